"""Generic implementation behind the remma_epiAA / remma_epiAD / remma_epiDD drop-in front-ends.

The reference keeps three near-identical copies of every routine (gmat/remma/remma_epi{AA,AD,DD}/);
here one parametrised implementation is specialised by an ``EpiKind``.  The host does what the reference
host does around the hot loop (phenotype parsing, the V / P preamble in numpy so that Z'Py and Z'PZ are the
same inputs both engines see, p-values with scipy for the few reported rows, file formats); every
O(pairs) computation runs in the CUDA engine.
"""
import gc
import logging
import os
import sys
import time
from functools import reduce

import numpy as np
import pandas as pd
import torch
from scipy import linalg
from scipy.stats import chi2

from .. import engine
from .._ffi import check, ffi, lib
from ..process_plink.process_plink import load_store
from ..uvlmm.design_matrix import design_matrix_wemai_multi_gmat
from . import parallel as par


class EpiKind(object):
    """What differs between the AA, AD and DD copies of the reference code."""

    def __init__(self, name, coding_u, coding_v, all_pairs):
        self.name = name
        self.coding_u, self.coding_v = coding_u, coding_v
        self.all_pairs = all_pairs          # AD exact: j in [0, m) incl. j == i (remma_epiAD.py:67-77)

    def default_snp_lst_0(self, num_snp):
        return range(num_snp) if self.all_pairs else range(num_snp - 1)

    def check_snp_lst_0(self, snp_lst_0, num_snp):
        # remma_epiAA.py:66 / remma_epiAD.py:70
        bad = (max(snp_lst_0) > num_snp - 1) if self.all_pairs else (max(snp_lst_0) >= num_snp - 1)
        if bad or min(snp_lst_0) < 0:
            logging.error('snp_lst_0 is out of range!')
            sys.exit()


AA = EpiKind('AA', engine.ADD, engine.ADD, False)
AD = EpiKind('AD', engine.ADD, engine.DOM, True)
DD = EpiKind('DD', engine.DOM, engine.DOM, False)


# ----------------------------------------------------------------------------------------------
# preamble (host numpy; restated from remma_epiAA.py:31-51 / remma_epiAA_pair.py:32-52)
# ----------------------------------------------------------------------------------------------
def preamble(y, xmat, zmat, gmat_lst, var_com, scipy_inv=True, need_pvp=True):
    """V = sigma_e^2 I + sum sigma_k^2 Z G_k Z';  P = V^-1 - V^-1 X (X'V^-1X)^-1 X'V^-1;
    returns (Z'Py [n,1], Z'PZ [n,n] or None).  ``scipy_inv`` selects scipy.linalg.inv (exact scan,
    remma_epiAA.py:40) or numpy.linalg.inv (pair / eff paths, remma_epiAA_pair.py:41)."""
    inv = linalg.inv if scipy_inv else np.linalg.inv
    y = np.array(y).reshape(-1, 1)
    n = y.shape[0]
    xmat = np.array(xmat).reshape(n, -1)
    vmat = np.diag([var_com[-1]] * n)
    for val in range(len(gmat_lst)):
        vmat += zmat.dot((zmat.dot(gmat_lst[val])).T) * var_com[val]
    vmat_inv = inv(vmat)
    vxmat = np.dot(vmat_inv, xmat)
    xvxmat = np.dot(xmat.T, vxmat)
    xvxmat = inv(xvxmat)
    pmat = reduce(np.dot, [vxmat, xvxmat, vxmat.T])
    pmat = vmat_inv - pmat
    pymat = zmat.T.dot(np.dot(pmat, y))
    pvpmat = zmat.T.dot((zmat.T.dot(pmat)).T) if need_pvp else None
    del vmat, vmat_inv, pmat
    gc.collect()
    return np.ascontiguousarray(pymat, dtype=np.float64), pvpmat


def preamble_device(y, xmat, zmat, gmat_lst, var_com, device, need_pvp=True):
    """The same preamble in float64 on the GPU: identical expressions, torch.linalg (cuSOLVER)
    for the two inverses and cuSPARSE for the products with the incidence matrix Z.  At N = 15,000 records the
    host version costs minutes (and under torchrun every rank repeats it on OMP_NUM_THREADS=1); this one takes
    a fraction of a second, and Z'Py / Z'PZ never visit the host.  Returns device tensors
    (Z'Py [n], Z'PZ [n,n] or None, bytes uploaded).  Agreement with the host version: ~1e-12 relative
    (tests/test_gpu_dropin.py::test_device_preamble_matches_host)."""
    f64 = torch.float64
    yv = torch.as_tensor(np.ascontiguousarray(np.array(y, dtype=np.float64).reshape(-1, 1))).to(device)
    n_rec = yv.shape[0]
    xm = torch.as_tensor(np.ascontiguousarray(np.array(xmat, dtype=np.float64).reshape(n_rec, -1))).to(device)
    zs, zbytes = engine.sparse_incidence(zmat, device)
    zt = zs.t().coalesce()
    h2d = yv.numel() * 8 + xm.numel() * 8 + zbytes
    vmat = torch.eye(n_rec, dtype=f64, device=device) * float(var_com[-1])
    for k in range(len(gmat_lst)):
        g = torch.as_tensor(np.ascontiguousarray(gmat_lst[k], dtype=np.float64)).to(device)
        h2d += g.numel() * 8
        vmat += torch.sparse.mm(zs, torch.sparse.mm(zs, g).t().contiguous()) * float(var_com[k])   # Z (Z G)'
        del g
    vinv = torch.linalg.inv(vmat)
    del vmat
    vx = vinv @ xm
    xvx = torch.linalg.inv(xm.t() @ vx)
    pmat = vinv - vx @ xvx @ vx.t()
    del vinv
    py = torch.sparse.mm(zt, pmat @ yv).reshape(-1).contiguous()
    pvp = None
    if need_pvp:
        pvp = torch.sparse.mm(zt, torch.sparse.mm(zt, pmat).t().contiguous()).contiguous()     # Z' (Z' P)'
    return py, pvp, h2d


def device_preamble_enabled():
    """GMAT_B200_PREAMBLE=host keeps the numpy preamble (bit-identical inputs for both engines: what the parity
    tests against the golden files use); default is the GPU preamble."""
    return os.environ.get("GMAT_B200_PREAMBLE", "device").lower() != "host"


class DeviceModel(object):
    """py = Z'Py and the packed Z'PZ resident on the device."""

    def __init__(self, pymat, pvpmat, device):
        self.py = torch.as_tensor(np.ascontiguousarray(pymat.reshape(-1))).to(device)
        self.pvp = torch.as_tensor(np.ascontiguousarray(pvpmat, dtype=np.float64)).to(device)
        self.packed = engine.pack_pvp(self.pvp)
        self.h2d_bytes = self.py.numel() * 8 + self.pvp.numel() * 8

    @classmethod
    def build(cls, y, xmat, zmat, gmat_lst, var_com, device, scipy_inv=True):
        """Preamble + upload: on the GPU by default, on the host with GMAT_B200_PREAMBLE=host."""
        if not device_preamble_enabled():
            pymat, pvpmat = preamble(y, xmat, zmat, gmat_lst, var_com, scipy_inv=scipy_inv)
            return cls(pymat, pvpmat, device)
        self = cls.__new__(cls)
        self.py, self.pvp, self.h2d_bytes = preamble_device(y, xmat, zmat, gmat_lst, var_com, device)
        self.packed = engine.pack_pvp(self.pvp)
        return self

    @classmethod
    def build_shared(cls, y, xmat, zmat, gmat_lst, var_com, device, scipy_inv=True):
        """One process per GPU: the preamble is the same on every rank, so rank 0 computes it and broadcasts Z'Py and
        Z'PZ (NCCL over NVLink: 200 MB at n = 5000) instead of every rank uploading the GRMs and inverting V
        (N = 15000 records in BASELINE configs[4]); the other ranks only re-tile Z'PZ for the FP64 pair kernel.
        All ranks then hold bit-identical model matrices.  One rank (or a gloo group): plain build."""
        import torch.distributed as dist
        rank, world = par.rank_world()
        if world == 1 or dist.get_backend() != "nccl":
            return cls.build(y, xmat, zmat, gmat_lst, var_com, device, scipy_inv=scipy_inv)
        n = zmat.shape[1]
        if rank == 0:
            self = cls.build(y, xmat, zmat, gmat_lst, var_com, device, scipy_inv=scipy_inv)
        else:
            self = cls.__new__(cls)
            self.py = torch.empty((n,), dtype=torch.float64, device=device)
            self.pvp = torch.empty((n, n), dtype=torch.float64, device=device)
            self.h2d_bytes = 0
        dist.broadcast(self.py, src=0)
        dist.broadcast(self.pvp, src=0)
        if rank != 0:
            self.packed = engine.pack_pvp(self.pvp)
        return self

    engine_note = ""

    def scanner(self, store, lut_u, lut_v, coding_v):
        """The exact-scan engine for this model and genotype store: ``scan(ilist, pair_space, chi_cut)``.
        GMAT_B200_EXACT_ENGINE = auto (default: the INT8 tensor-core engine) | i8 | f64 (the FP64 DMMA kernel)."""
        which = engine.exact_engine(store)
        if which == "i8":
            eng = engine.I8Exact(store, lut_u, lut_v, coding_v, self.py, self.pvp, packed=self.packed)
            self.engine_note = ("int8 tensor cores (tcgen05 kind::i8), %d base-256 digit planes = %d bits, weak pairs "
                                "re-checked by the FP64 pair kernel" % (eng.n_planes, 8 * eng.n_planes - 1))
            return eng.scan
        self.engine_note = "FP64 DMMA kernel (GMAT_B200_EXACT_ENGINE=f64)"
        return lambda ilist, pair_space, chi_cut: engine.epi_exact_scan(store, lut_u, lut_v, self.py, self.packed,
                                                                        ilist, pair_space, chi_cut)


# ----------------------------------------------------------------------------------------------
# exact scan (remma_epiAA.py:16-86 and the AD / DD copies)
# ----------------------------------------------------------------------------------------------
MAX_PAIRS_PER_LAUNCH = 1 << 28      # candidate pairs per kernel launch (bounds the hit buffer retries)


def _chunks(kind, snp_lst_0, num_snp):
    """Split the first-SNP list into launches of at most MAX_PAIRS_PER_LAUNCH candidate pairs."""
    out, cur, cur_pairs = [], [], 0
    for i in snp_lst_0:
        npair = num_snp if kind.all_pairs else (num_snp - 1 - i)
        if cur and cur_pairs + npair > MAX_PAIRS_PER_LAUNCH:
            out.append(cur)
            cur, cur_pairs = [], 0
        cur.append(i)
        cur_pairs += npair
    if cur:
        out.append(cur)
    return out


def _write_rows(out_file, cols):
    res = pd.DataFrame(cols)
    res.to_csv(out_file, sep=' ', header=False, index=False, mode='a')


REBALANCE_MIN_WORK = 1.5e15     # pairs per rank x n^2 from which the speed-aware deal is used (~3 s of scan per rank)
LAST_RATES = []                 # relative scan rates of the ranks measured by the last speed-aware deal (bench.py prints them)
PHASES = {}     # wall seconds of the phases of the last exact_scan call on this rank (bench.py prints them)


class _Phase(object):
    """with _Phase('name'): ... -- adds the wall time of the block (device work included) to PHASES[name]."""

    def __init__(self, name):
        self.name = name

    def __enter__(self):
        self.t0 = time.perf_counter()

    def __exit__(self, *exc):
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        PHASES[self.name] = PHASES.get(self.name, 0.0) + time.perf_counter() - self.t0
        return False


def exact_scan(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0=None, p_cut=1.0e-5,
               out_file='epi'):
    design_s = PHASES.get("design_matrix")
    PHASES.clear()
    if design_s is not None:
        PHASES["design_matrix"] = design_s
    logging.info("Calculate the phenotypic covariance matrix and inversion")
    device = engine.current_device()
    rank, world = par.rank_world()
    with _Phase("preamble"):
        model = DeviceModel.build_shared(y, xmat, zmat, gmat_lst, var_com, device, scipy_inv=True)
    logging.info("Read the SNP")
    if rank == 0:
        np.savetxt(out_file, ['snp_0 snp_1 eff chi p_val'], fmt='%s')     # header first (remma_epiAA.py:53)
    with _Phase("bed_read_h2d"):
        store = load_store(bed_file)
    num_snp = store.m
    logging.info('Test')
    if snp_lst_0 is None:
        snp_lst_0 = kind.default_snp_lst_0(num_snp)
    else:
        kind.check_snp_lst_0(snp_lst_0, num_snp)
    snp_lst_0 = np.fromiter((int(i) for i in snp_lst_0), dtype=np.int64)
    # position of every first SNP in the caller's list (first occurrence): the reference's row order
    order = np.full(num_snp, -1, dtype=np.int64)
    uniq, first_pos = np.unique(snp_lst_0, return_index=True)
    order[uniq] = first_pos
    lut_u, lut_v = store.lut(kind.coding_u), store.lut(kind.coding_v)
    # p < p_cut  <=>  chi > isf(p_cut); the device keeps a slightly larger set, the host re-applies the
    # reference's own comparison on scipy p-values (remma_epiAA.py:77,81)
    chi_cut = -1.0 if p_cut >= 1.0 else float(chi2.isf(p_cut, 1)) * (1.0 - 1e-9)
    pair_space = engine.PAIRS_ALL if kind.all_pairs else engine.PAIRS_UPPER
    # Sharding over the ranks: a static deal of first-SNP blocks by pair count; for a long scan (seconds per rank)
    # the deal is speed-aware -- a first share is scanned and timed, the rest dealt by the measured rates, because
    # the GPUs of one box run a few per cent apart under the power cap and everyone waits for the slowest
    # (GMAT_B200_REBALANCE=0 keeps the static deal)
    n_tests = float(num_snp) * len(snp_lst_0) if kind.all_pairs else float(np.maximum(num_snp - 1 - snp_lst_0, 0).sum())
    rebalance = (world > 1 and os.environ.get("GMAT_B200_REBALANCE", "1") != "0"
                 and n_tests / world * float(store.n) ** 2 >= REBALANCE_MIN_WORK and len(snp_lst_0) > 0)
    rest = None
    del LAST_RATES[:]
    if rebalance:
        my_list, head_pairs, rest = par.calibration_split(snp_lst_0, num_snp, kind.all_pairs, rank, world)
    else:
        my_list = par.shard_first_snps(snp_lst_0, num_snp, kind.all_pairs, rank, world)
    clock_t0 = time.perf_counter()
    cpu_t0 = time.process_time()

    def emit(hits):
        if len(hits) == 0:
            return
        hits = hits[np.lexsort((hits['snp_1'], order[hits['snp_0']]))]     # (i in list order, j ascending)
        p_vec = chi2.sf(hits['chi'], 1)
        keep = p_vec < p_cut                                   # the reference's comparison (remma_epiAA.py:81)
        _write_rows(out_file, {0: hits['snp_0'][keep].astype(np.int64), 1: hits['snp_1'][keep].astype(np.int64),
                               2: hits['eff'][keep], 3: hits['chi'][keep], 4: p_vec[keep]})

    mine = []
    with _Phase("engine_init"):
        scan = model.scanner(store, lut_u, lut_v, kind.coding_v)
        logging.info("Exact-scan engine: %s", model.engine_note)
    while True:
        t_scan0 = time.perf_counter()
        for chunk in _chunks(kind, my_list, num_snp):
            with _Phase("scan"):
                hits = scan(np.asarray(chunk, dtype=np.int64), pair_space, chi_cut)
            if world == 1:
                with _Phase("write"):
                    emit(hits)                                 # appended launch by launch (resume granularity)
            else:
                mine.append(hits)
        if rest is None:
            break
        # the calibration share is done: exchange the measured rates (pairs/s; one float per rank, no data-path
        # collective) and take this rank's part of what is left
        torch.cuda.synchronize()
        secs = max(time.perf_counter() - t_scan0, 1e-6)
        rates = par.all_gather_floats(head_pairs / secs if head_pairs > 0 else 0.0, store.device)
        good = [v for v in rates if v > 0]
        rates = [v if v > 0 else (min(good) if good else 1.0) for v in rates]
        LAST_RATES[:] = [round(v / max(rates), 4) for v in rates]
        my_list, rest = par.redeal(rest, rates, rank), None
    if world > 1:
        # the only collective of the scan: variable-length gather of hit records to rank 0
        with _Phase("gather"):
            local = np.concatenate(mine) if mine else np.zeros((0,), dtype=engine.HIT_DTYPE)
            hits = par.gather_hits(local, store.device)
        if rank == 0:
            with _Phase("write"):
                emit(hits)
    torch.cuda.synchronize()
    clock_t1 = time.perf_counter()
    cpu_t1 = time.process_time()
    logging.info("Running time: Clock time, {:.5f} sec; CPU time, {:.5f} sec.".format(clock_t1 - clock_t0,
                                                                                   cpu_t1 - cpu_t0))
    return 0


def exact_scan_parallel(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, parallel, p_cut, out_file):
    """remma_epiAA.py:109-140: part k of P, output ``out_file.k``."""
    logging.info("Parallel: " + str(parallel[0]) + ', ' + str(parallel[1]))
    num_snp = engine.count_lines(bed_file + '.bim')
    snp_list_0, pos = par.parallel_lists(num_snp, parallel, kind.all_pairs)
    logging.info('SNP position point: ' + ','.join(str(v) for v in pos))
    return exact_scan(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0=snp_list_0, p_cut=p_cut,
                      out_file=out_file + '.' + str(parallel[1]))


# ----------------------------------------------------------------------------------------------
# explicit pair list (remma_epiAA_pair.py:16-92 and copies)
# ----------------------------------------------------------------------------------------------
def pair_test(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, snp_pair_file, max_test_pair=50000,
              p_cut=1.0e-4, out_file='epi_pair'):
    logging.info("Calculate the phenotypic covariance matrix and inversion")
    model = DeviceModel.build(y, xmat, zmat, gmat_lst, var_com, engine.current_device(), scipy_inv=False)
    logging.info("Read the SNP")
    store = load_store(bed_file)
    num_snp = store.m
    lut_u, lut_v = store.lut(kind.coding_u), store.lut(kind.coding_v)
    logging.info('Test')
    np.savetxt(out_file, ['snp_0 snp_1 eff var chi p'], fmt='%s')
    clock_t0 = time.perf_counter()
    cpu_t0 = time.process_time()
    ipart = -1
    i8_pairs = None
    while True:
        ipart += 1
        skiprows = 1 + ipart * max_test_pair          # first line is a header (remma_epiAA_pair.py:69)
        try:
            snp_pair = pd.read_csv(snp_pair_file, header=None, sep=r'\s+', skiprows=skiprows,
                                   nrows=max_test_pair)
        except Exception as e:
            logging.info(e)
            break
        snp_pair = np.array(snp_pair.iloc[:, 0:2], dtype=int)
        if np.max(snp_pair) > num_snp - 1 or np.min(snp_pair) < 0:
            logging.error('snp_pair is out of range!')
            sys.exit()
        which = engine.pair_engine(store, len(snp_pair))
        if ipart == 0:
            logging.info("Pair-list engine: %s", "int8 tensor cores (tcgen05), FP64 re-check of weak pairs"
                         if which == "i8" else "FP64 DMMA pair kernel")
        if which == "i8":
            if i8_pairs is None:
                i8_pairs = engine.I8Pairs(store, kind.coding_u, kind.coding_v, lut_u, lut_v, model.py, model.pvp,
                                          packed=model.packed)
            res = i8_pairs.stats(snp_pair)
        else:
            res = engine.epi_exact_pairs(store, lut_u, lut_v, model.py, model.packed, snp_pair)
        p_vec = chi2.sf(res['chi'], 1)
        keep = p_vec < p_cut
        _write_rows(out_file, {0: snp_pair[keep, 0], 1: snp_pair[keep, 1], 2: res['eff'][keep],
                               3: res['var'][keep], 4: res['chi'][keep], 5: p_vec[keep]})
    torch.cuda.synchronize()
    clock_t1 = time.perf_counter()
    cpu_t1 = time.process_time()
    logging.info("Running time: Clock time, {:.5f} sec; CPU time, {:.5f} sec.".format(clock_t1 - clock_t0,
                                                                                   cpu_t1 - cpu_t0))
    return 0


# ----------------------------------------------------------------------------------------------
# approximate screen through the drop-in C symbols (remma_epiAA_eff.py:20-96 and copies)
# ----------------------------------------------------------------------------------------------
def _pymat_host(y, xmat, zmat, gmat_lst, var_com):
    """Z'Py as a host [n,1] array for the drop-in C symbols (host pointers), computed on the GPU by default."""
    if not device_preamble_enabled():
        return preamble(y, xmat, zmat, gmat_lst, var_com, scipy_inv=False, need_pvp=False)[0]
    py, _, _ = preamble_device(y, xmat, zmat, gmat_lst, var_com, engine.current_device(), need_pvp=False)
    return np.ascontiguousarray(py.cpu().numpy().reshape(-1, 1))


def _screen_defaults(kind, maf, snp_lst_0, num_snp):
    # quirk kept: AD _eff defaults to range(m) (remma_epiAD_eff.py:57), AD _maf_eff to range(m-1)
    # (remma_epiAD_maf_eff.py:59); both are equivalent (the last first-SNP has no j > i)
    if snp_lst_0 is None:
        return range(num_snp) if (kind.all_pairs and not maf) else range(num_snp - 1)
    upper_ok = (max(snp_lst_0) <= num_snp - 1) if (kind.all_pairs and not maf) else (max(snp_lst_0) < num_snp - 1)
    if not upper_ok or min(snp_lst_0) < 0:
        logging.error('snp_lst_0 is out of range!')
        sys.exit()
    return snp_lst_0


def _append_p_app(temp_file, out_file, deno_of):
    """The 'Add the approximate P values' step of the screen front ends (remma_epiAA_eff.py:85-95,
    remma_epiAA_maf_eff.py:88-101): every row of the C screen's text file gets chi_app = eff^2 / deno -- eff re-read
    from the %g-truncated text, like the reference -- and p_app = chi2.sf(chi_app, 1).  Same bytes as the reference's
    per-line loop (same IEEE operations, same str() of the same doubles), evaluated as arrays: the loop costs 70 us a
    row in scipy's scalar path, more than the whole GPU screen for a few hundred rows."""
    with open(temp_file) as fin:
        head_line = fin.readline().strip() + ' chi_app p_app\n'
        rows = [line.split() for line in fin]
    with open(out_file, 'w') as fout:
        fout.write(head_line)
        if rows:
            eff = np.array([float(r[-1]) for r in rows], dtype=np.float64)
            chi_app = eff * eff / deno_of(rows)
            p_app = chi2.sf(chi_app, 1)
            fout.write(''.join(' '.join(r + [str(c), str(q)]) + '\n'
                               for r, c, q in zip(rows, chi_app.tolist(), p_app)))
    os.remove(temp_file)


def eff_screen(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0=None, var_app=1.0, p_cut=1.0e-5,
               out_file='epi_eff'):
    logging.info("Calculate the phenotypic covariance matrix and inversion")
    pymat = _pymat_host(y, xmat, zmat, gmat_lst, var_com)
    num_snp = engine.count_lines(bed_file + '.bim')
    num_id = engine.count_lines(bed_file + '.fam')
    snp_lst_0 = _screen_defaults(kind, False, snp_lst_0, num_snp)
    engine.current_device()
    logging.info("Convert python variates to C type")
    pbed_file = ffi.new("char[]", bed_file.encode('ascii'))
    snp_lst_0 = np.array(list(snp_lst_0), dtype=np.longlong)
    psnp_lst_0 = ffi.cast("long long *", snp_lst_0.ctypes.data)
    ppymat = ffi.cast("double *", pymat.ctypes.data)
    chi_cut = chi2.isf(p_cut, 1)
    eff_cut = np.sqrt(chi_cut * var_app)
    temp_file = out_file + '.temp'
    pout_file = ffi.new("char[]", temp_file.encode('ascii'))
    logging.info('Test')
    clock_t0 = time.perf_counter()
    cpu_t0 = time.process_time()
    fn = getattr(lib, 'remma_epi%s_eff_cpu' % kind.name)
    check(fn(pbed_file, num_id, num_snp, psnp_lst_0, len(snp_lst_0), ppymat, float(eff_cut), pout_file),
          'remma_epi%s_eff_cpu' % kind.name)
    logging.info("Screen engine: %s", ffi.string(lib.gmat_b200_last_engine_note()).decode())
    clock_t1 = time.perf_counter()
    cpu_t1 = time.process_time()
    logging.info("Running time: Clock time, {:.5f} sec; CPU time, {:.5f} sec.".format(clock_t1 - clock_t0,
                                                                                   cpu_t1 - cpu_t0))
    logging.info('Add the approximate P values')
    _append_p_app(temp_file, out_file, lambda rows: var_app)
    return 0


def eff_screen_parallel(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, parallel, var_app, p_cut, out_file):
    logging.info("Parallel: " + str(parallel[0]) + ', ' + str(parallel[1]))
    num_snp = engine.count_lines(bed_file + '.bim')
    # last cut of part 1: num_snp - 1 in the AA / DD copies (remma_epiAA_eff.py:148), num_snp in the AD copy
    # (remma_epiAD_eff.py:147) -- the extra first SNP has no j > i, so only the logged cut points differ
    snp_list_0, pos = par.parallel_lists(num_snp, parallel, kind.all_pairs)
    logging.info('SNP position point: ' + ','.join(str(v) for v in pos))
    return eff_screen(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0=snp_list_0, var_app=var_app,
                      p_cut=p_cut, out_file=out_file + '.' + str(parallel[1]))


def maf_eff_screen_parallel(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, parallel, freq, freq2, freq_deno,
                            p_cut, out_file):
    """remma_epiAA_maf_eff.py:129-161 and the AD / DD copies: part k of P of the MAF-binned screen, output
    ``out_file.k``.  All three copies cut part 1 at num_snp - 1 (remma_epiAD_maf_eff.py:158 too)."""
    logging.info("Parallel: " + str(parallel[0]) + ', ' + str(parallel[1]))
    num_snp = engine.count_lines(bed_file + '.bim')
    snp_list_0, pos = par.parallel_lists(num_snp, parallel, False)
    logging.info('SNP position point: ' + ','.join(str(v) for v in pos))
    return maf_eff_screen(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0=snp_list_0, freq=freq,
                          freq2=freq2, freq_deno=freq_deno, p_cut=p_cut, out_file=out_file + '.' + str(parallel[1]))


def maf_eff_screen(kind, y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0=None, freq=None, freq2=None,
                   freq_deno=None, p_cut=1.0e-5, out_file='epi_maf_eff'):
    """remma_epiAA_maf_eff.py:20-104; ``freq2`` is the second bin vector of the AD variant (freqD)."""
    logging.info("Calculate the phenotypic covariance matrix and inversion")
    pymat = _pymat_host(y, xmat, zmat, gmat_lst, var_com)
    num_snp = engine.count_lines(bed_file + '.bim')
    num_id = engine.count_lines(bed_file + '.fam')
    snp_lst_0 = _screen_defaults(kind, True, snp_lst_0, num_snp)
    if freq is None:
        freq = np.zeros((num_snp,), dtype=np.longlong)
    if freq2 is None:
        freq2 = np.zeros((num_snp,), dtype=np.longlong) if kind.all_pairs else freq
    if freq_deno is None:
        freq_deno = np.ones(111)
    engine.current_device()
    freq = np.ascontiguousarray(freq, dtype=np.longlong)
    freq2 = np.ascontiguousarray(freq2, dtype=np.longlong)
    pbed_file = ffi.new("char[]", bed_file.encode('ascii'))
    snp_lst_0 = np.array(list(snp_lst_0), dtype=np.longlong)
    psnp_lst_0 = ffi.cast("long long *", snp_lst_0.ctypes.data)
    ppymat = ffi.cast("double *", pymat.ctypes.data)
    chi_cut = chi2.isf(p_cut, 1)
    eff_cut = np.ascontiguousarray(np.sqrt(chi_cut * np.asarray(freq_deno, dtype=np.float64)))
    if not kind.all_pairs:
        np.savetxt('eff_cut', eff_cut)                              # side effect kept (remma_epiAA_maf_eff.py:79)
    peff_cut = ffi.cast("double *", eff_cut.ctypes.data)
    temp_file = out_file + '.temp'
    pout_file = ffi.new("char[]", temp_file.encode('ascii'))
    logging.info('Test')
    fn = getattr(lib, 'remma_epi%s_maf_eff_cpu' % kind.name)
    pf = ffi.cast("long long *", freq.ctypes.data)
    if kind.all_pairs:
        rc = fn(pbed_file, num_id, num_snp, psnp_lst_0, len(snp_lst_0), ppymat, pf,
                ffi.cast("long long *", freq2.ctypes.data), peff_cut, pout_file)
    else:
        rc = fn(pbed_file, num_id, num_snp, psnp_lst_0, len(snp_lst_0), ppymat, pf, peff_cut, pout_file)
    check(rc, 'remma_epi%s_maf_eff_cpu' % kind.name)
    logging.info("Screen engine: %s", ffi.string(lib.gmat_b200_last_engine_note()).decode())
    logging.info('Add the approximate P values')
    freq_deno = np.asarray(freq_deno, dtype=np.float64)
    _append_p_app(temp_file, out_file,
                  lambda rows: freq_deno[freq[[int(r[0]) for r in rows]] * 10 + freq2[[int(r[1]) for r in rows]]])
    return 0


# ----------------------------------------------------------------------------------------------
# approximate pipelines (remma_epiAA_approx.py:10-53, remma_epiAA_maf_approx.py:11-90 and copies),
# run in memory on one resident genotype store / one preamble instead of three.
# ----------------------------------------------------------------------------------------------
class Prepared(object):
    """Everything the three pipeline stages share."""

    def __init__(self, kind, pheno_file, bed_file, gmat_lst, var_com):
        from ..engine import GenotypeStore
        self.kind = kind
        y, xmat, zmat = design_matrix_wemai_multi_gmat(pheno_file, bed_file)
        self.raw_store = GenotypeStore(bed_file)              # what the C screen sees (no imputation)
        self.store = self.raw_store
        if self.raw_store.n_missing() > 0:
            self.store = load_store(bed_file)                 # what the numpy paths see (imputed)
        self.model = DeviceModel.build(y, xmat, zmat, gmat_lst, var_com, self.store.device, scipy_inv=False)
        self.num_snp, self.num_id = self.store.m, self.store.n
        self._screen_ops = None                               # int8 screen operands, shared by every screen call
        self._i8_pairs = None                                 # int8 pair-list engine (planes of Z'PZ), built on first use

    def pair_stats(self, pairs):
        k = self.kind
        lut_u, lut_v = self.store.lut(k.coding_u), self.store.lut(k.coding_v)
        which = engine.pair_engine(self.store, len(pairs))
        logging.info("Pair-list engine: %s", "int8 tensor cores (tcgen05), FP64 re-check of weak pairs"
                     if which == "i8" else "FP64 DMMA pair kernel")
        if which == "i8":
            if self._i8_pairs is None:
                self._i8_pairs = engine.I8Pairs(self.store, k.coding_u, k.coding_v, lut_u, lut_v, self.model.py,
                                                self.model.pvp, packed=self.model.packed)
            return self._i8_pairs.stats(pairs)
        return engine.epi_exact_pairs(self.store, lut_u, lut_v, self.model.py, self.model.packed, pairs)

    def screen(self, snp_lst_0, eff_cut=0.0, cut_table=None, bin_i=None, bin_j=None):
        """Hits of the C screen in single-threaded C row order; returns list of (snp_0, snp_1, eff)."""
        k, st = self.kind, self.raw_store
        la = st.lut(engine.ADD, engine.FREQ_CLOOP)
        ld = st.lut(engine.DOM, engine.FREQ_CLOOP)
        ilist = np.asarray(list(snp_lst_0), dtype=np.int64)
        A, D = engine.ADD, engine.DOM
        if k.name == 'AD':
            passes = [(la, ld, engine.EFF_GE, (A, D)), (ld, la, engine.EFF_SWAP, (D, A))]
        elif k.name == 'AA':
            passes = [(la, la, 0, (A, A))]
        else:
            passes = [(ld, ld, 0, (D, D))]
        rows = []
        if self._screen_ops is None and engine.screen_engine(st) == "i8":
            self._screen_ops = engine.ScreenOperands(st, self.model.py)
        for d, (lu, lv, flags, codings) in enumerate(passes):
            h = engine.epi_eff(st, lu, lv, self.model.py, ilist, flags, eff_cut, cut_table, bin_i, bin_j,
                               codings=codings, ops=self._screen_ops)
            logging.info("Screen engine: %s %s", engine.LAST_SCREEN_ENGINE, engine.LAST_SCREEN_NOTE)
            swap = bool(flags & engine.EFF_SWAP)
            for s0, s1, e in zip(h['snp_0'].tolist(), h['snp_1'].tolist(), h['eff'].tolist()):
                rows.append(((s1, s0, d) if swap else (s0, s1, d), s0, s1, e))
        rows.sort(key=lambda r: r[0])
        return [(r[1], r[2], r[3]) for r in rows]


def random_pair(num_snp, out_file='random_pair', num_pair=100000, num_each_pair=5000, ordered=True):
    """Restated from gmat/remma/random_pair.py:6-65 (``ordered`` False = random_pairAD): global numpy RNG,
    string set, truncation to num_pair, written with a 'snp_0 snp_1' header and read back."""
    limit = num_snp * (num_snp - 1) / 2 if ordered else num_snp * (num_snp - 1)
    if num_pair > limit:
        logging.error('num_pair must be not greater than: ' + str(limit))
        sys.exit()
    if num_pair < num_each_pair:
        logging.error('num_pair must be greater than num_each_pair')
        sys.exit()
    res_lst = set()
    while True:
        arr = np.random.randint(num_snp, size=(num_each_pair, 2))
        arr = arr[arr[:, 0] < arr[:, 1], :] if ordered else arr[arr[:, 0] != arr[:, 1], :]
        # the same strings as str(a) + ' ' + str(b), entered through a temporary set like the reference does: the
        # iteration order of res_lst (hence WHICH pairs survive the truncation) depends on the insertion history
        res_lst.update(set(['%d %d' % (a, b) for a, b in arr.tolist()]))
        if len(res_lst) >= num_pair:
            break
    res_lst = list(res_lst)[:num_pair]
    # the bytes np.savetxt(out_file, np.array(res_lst), fmt="%s", header='snp_0 snp_1', comments='') writes, and the
    # array np.loadtxt(out_file, dtype=int, skiprows=1) reads back (random_pair.py:33-36 + the next stage's read),
    # without numpy's per-line text machinery (0.2 s of a 0.7 s pipeline at 100,000 pairs)
    with open(out_file, 'w') as fout:
        fout.write('snp_0 snp_1\n' + '\n'.join(res_lst) + '\n')
    return np.array(' '.join(res_lst).split(), dtype=int).reshape(-1, 2)


def random_pairAD(num_snp, out_file='random_pair', num_pair=100000, num_each_pair=5000):
    return random_pair(num_snp, out_file, num_pair, num_each_pair, ordered=False)


def _exact_rows_text(pairs, res):
    """The rows remma_epiXX_pair(p_cut=1) would have written for these pairs, as lists of strings."""
    import io
    p_vec = chi2.sf(res['chi'], 1)
    keep = p_vec < 1
    buf = io.StringIO()
    pd.DataFrame({0: pairs[keep, 0], 1: pairs[keep, 1], 2: res['eff'][keep], 3: res['var'][keep],
                  4: res['chi'][keep], 5: p_vec[keep]}).to_csv(buf, sep=' ', header=False, index=False)
    return [line.split() for line in buf.getvalue().splitlines()], res['var'][keep]


def _maf_bins(kind, store):
    """Per-SNP bin vectors (allele-frequency / heterozygosity * 20, truncated) and the arrays the reference
    saves next to the output (remma_epiAA_maf_approx.py:38-42, remma_epiAD_maf_approx.py:39-49,
    remma_epiDD_maf_approx.py:39-44).  Formed from exact device genotype counts."""
    counts = store.counts.cpu().numpy().astype(np.int64)
    num_id = store.n
    allele_sum = (counts[:, 2] + 2 * counts[:, 3]).astype(np.float64)
    saved = {}
    if kind.name == 'AA':
        freq = 1 - allele_sum / (2 * num_id)
        freq[freq > 0.5] = 1 - freq[freq > 0.5]
        saved['.freq'] = freq.copy()
        b = list(map(np.longlong, freq * 20))
        return b, b, saved
    het = counts[:, 2].astype(np.float64) / num_id
    het[het > 0.5] = 1 - het[het > 0.5]
    if kind.name == 'DD':
        saved['.heter'] = het.copy()
        b = np.array(het * 20, dtype=np.longlong)
        return b, b, saved
    freq_a = allele_sum / (2 * num_id)
    freq_a[freq_a > 0.5] = 1 - freq_a[freq_a > 0.5]
    saved['.maf'], saved['.heter'] = freq_a.copy(), het.copy()
    return np.array(freq_a * 20, dtype=np.longlong), np.array(het * 20, dtype=np.longlong), saved


def approx(kind, pheno_file, bed_file, gmat_lst, var_com, p_cut=1.0e-5, num_random_pair=100000,
           out_file='epi_approx', maf=False, parallel=None):
    sfx = '' if parallel is None else '.' + str(parallel[1])
    logging.info("\n\n#####Randomly select {:d} pairs, and test these SNP pairs#####".format(num_random_pair))
    num_snp = engine.count_lines(bed_file + '.bim')
    ad = kind.name == 'AD'
    rp_name = out_file + ('.random_pairAD' if (ad and maf) else '.random_pair') + sfx
    PHASES.clear()
    with _Phase("random_pairs"):
        pairs = (random_pairAD if ad else random_pair)(num_snp, out_file=rp_name, num_pair=num_random_pair)
        os.remove(rp_name)
        pairs = np.asarray(pairs, dtype=np.int64).reshape(-1, 2)
    with _Phase("model_and_store"):
        ctx = Prepared(kind, pheno_file, bed_file, gmat_lst, var_com)
    with _Phase("random_pairs_exact_test"):
        res = ctx.pair_stats(pairs)
        if maf:                                # rows as text: the per-bin table and out_file.random need them
            rand_rows, rand_var = _exact_rows_text(pairs, res)
        else:                                  # only the variances of the rows _pair(p_cut=1) keeps (p < 1)
            rand_rows, rand_var = None, res['var'][chi2.sf(res['chi'], 1) < 1]
    if parallel is None:
        snp_lst_0 = range(num_snp) if (ad and not maf) else range(num_snp - 1)
    else:
        snp_lst_0, _ = par.parallel_lists(num_snp, parallel, False)
    chi_cut = chi2.isf(p_cut, 1)
    logging.info("\n\n#####Screen the epistatic effects and select top SNP pairs based on approximate test#####")
    if not maf:
        var_median = np.median(rand_var)                                   # remma_epiAA_approx.py:30-31
        with _Phase("screen"):
            hits = ctx.screen(snp_lst_0, eff_cut=float(np.sqrt(chi_cut * var_median)))
        deno = lambda s0, s1: var_median                                    # noqa: E731
    else:
        # the _maf_ pipelines of the reference leave the random pairs' exact statistics on disk: they never remove
        # out_file.random[.k] (remma_epiAA_maf_approx.py:29-31,114-116; the plain _approx variants do, :33)
        with open(out_file + '.random' + sfx, 'w') as fout:
            fout.write('snp_0 snp_1 eff var chi p\n')
            for arr in rand_rows:
                fout.write(' '.join(arr) + '\n')
        bin_i, bin_j, saved = _maf_bins(kind, ctx.store)
        for ext, arr in saved.items():
            np.savetxt(out_file + ext + sfx, arr)
        sums, cnts = {}, {}
        for arr in rand_rows:                     # per-bin MEAN of the variances (named median in the reference)
            keys = [(bin_i[int(arr[0])], bin_j[int(arr[1])])]
            if not ad:
                keys.append((bin_i[int(arr[1])], bin_j[int(arr[0])]))
            for kv in keys:
                kv = ' '.join([str(kv[0]), str(kv[1])])
                cnts[kv] = cnts.get(kv, 0) + 1
                sums[kv] = sums.get(kv, 0.0) + float(arr[-3])
        all_sum, all_cnt = 0, 0
        for kv in cnts:
            all_sum += sums[kv]
            all_cnt += cnts[kv]
            sums[kv] = sums[kv] / cnts[kv]
        all_mean = all_sum / all_cnt
        freq_deno = np.ones(111)
        with open(out_file + '.freq_denominator' + sfx, 'w') as fout:
            for key1 in set(bin_i):
                for key2 in set(bin_j):
                    kv = ' '.join([str(key1), str(key2)])
                    if kv not in sums:
                        sums[kv] = all_mean
                    fout.write(kv + ' ' + str(sums[kv]) + '\n')
                    freq_deno[key1 * 10 + key2] = sums[kv]
        bi, bj = np.asarray(bin_i, dtype=np.int64), np.asarray(bin_j, dtype=np.int64)
        if not ad:
            np.savetxt('eff_cut', np.sqrt(chi_cut * freq_deno))   # written by the nested _maf_eff call (remma_epiAA_maf_eff.py:79)
        with _Phase("screen"):
            hits = ctx.screen(snp_lst_0, cut_table=np.sqrt(chi_cut * freq_deno), bin_i=bi, bin_j=bj)
        deno = lambda s0, s1: freq_deno[bi[s0] * 10 + bj[s1]]              # noqa: E731
    # approximate p from the %g-truncated effect, exactly like the text round trip of the reference
    # (as arrays: the same IEEE operations and the same str() of the same doubles as a per-row loop, without scipy's
    # 70 us scalar call per row)
    surv = [(s0, s1) for s0, s1, _ in hits]
    p_app_txt = None
    with _Phase("p_app"):
        if hits:
            e_trunc = np.array([float('%g' % e) for _, _, e in hits], dtype=np.float64)
            den = np.array([deno(s0, s1) for s0, s1 in surv], dtype=np.float64)
            p_app = chi2.sf(e_trunc * e_trunc / den, 1)
            p_app_txt = np.array([str(q) for q in p_app], dtype=object)      # the text the .approx_p file would hold
    logging.info("\n\n#####Calculate exact p values for top SNP pairs#####")
    with _Phase("survivors_exact_test_and_write"), open(out_file + sfx, 'w') as fout:
        # header of .exact_p with the token p_app spliced in (remma_epiAA_approx.py:41-50); the rows are those
        # remma_epiXX_pair(p_cut=1) keeps (p < 1) with the survivor's p_app inserted before p -- the reference looks it
        # up by the 'snp_0 snp_1' key, here it is the same row of the same list; written as one table (pandas formats
        # column by column: the same text per field as the row-by-row form)
        fout.write('snp_0 snp_1 eff var chi p_app p\n')
        if surv:
            sp = np.asarray(surv, dtype=np.int64)
            res = ctx.pair_stats(sp)
            p_vec = chi2.sf(res['chi'], 1)
            keep = p_vec < 1
            pd.DataFrame({0: sp[keep, 0], 1: sp[keep, 1], 2: res['eff'][keep], 3: res['var'][keep],
                          4: res['chi'][keep], 5: p_app_txt[keep], 6: p_vec[keep]}).to_csv(
                fout, sep=' ', header=False, index=False)
    return 0
