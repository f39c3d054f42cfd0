"""Device-side plumbing shared by the drop-in front-ends: torch owns every device buffer, the CUDA
engine (libgmat_b200.so, through cffi) does all the work.  Nothing here computes statistics on the
host; if CUDA or the library is unavailable every entry point raises.
"""
import os

import numpy as np
import torch

from ._ffi import CONST, check, ffi, lib

HIT_DTYPE = np.dtype([("snp_0", "<i4"), ("snp_1", "<i4"), ("eff", "<f8"), ("var", "<f8"), ("chi", "<f8")])
assert HIT_DTYPE.itemsize == 32

ADD, DOM, RAW = CONST["GMAT_B200_CODING_ADD"], CONST["GMAT_B200_CODING_DOM"], CONST["GMAT_B200_CODING_RAW"]
FREQ_NUMPY, FREQ_CLOOP = CONST["GMAT_B200_FREQ_NUMPY"], CONST["GMAT_B200_FREQ_CLOOP"]
PAIRS_UPPER, PAIRS_ALL = CONST["GMAT_B200_PAIRS_UPPER"], CONST["GMAT_B200_PAIRS_ALL"]
EFF_GE, EFF_SWAP = CONST["GMAT_B200_EFF_GE"], CONST["GMAT_B200_EFF_SWAP"]


def require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError("gmat_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")


def current_device():
    require_cuda()
    dev = torch.cuda.current_device()
    check(lib.gmat_b200_set_device(dev), "gmat_b200_set_device")
    return torch.device("cuda", dev)


def _p(t, ctype):
    return ffi.cast(ctype, t.data_ptr()) if t is not None else ffi.NULL


def _stream():
    return ffi.cast("void *", torch.cuda.current_stream().cuda_stream)


def sparse_incidence(zmat, device):
    """scipy CSR/COO incidence matrix Z (records x individuals) -> coalesced torch sparse COO float64 on the
    device, and the bytes uploaded."""
    import warnings
    zc = zmat.tocoo()
    idx = torch.as_tensor(np.vstack([zc.row, zc.col]).astype(np.int64)).to(device)
    val = torch.as_tensor(np.asarray(zc.data, dtype=np.float64)).to(device)
    with warnings.catch_warnings():
        warnings.filterwarnings("ignore", message="Sparse invariant checks")
        zs = torch.sparse_coo_tensor(idx, val, size=tuple(zc.shape), dtype=torch.float64, device=device,
                                     check_invariants=False).coalesce()
    return zs, idx.numel() * 8 + val.numel() * 8


def count_lines(path):
    """Row count of a .bim/.fam file (the reference counts rows with pd.read_csv(header=None),
    gmat/remma/remma_epiAA/remma_epiAA_eff.py:55-56)."""
    with open(path) as f:
        return sum(1 for line in f if line.strip())


def row_stride(n):
    """Bytes between SNP rows of the device genotype store: ceil(n/4) padded to 32 B (=128 ids)."""
    return ((n + 3) // 4 + 31) // 32 * 32


class GenotypeStore(object):
    """The PLINK .bed payload resident in HBM as 2-bit codes, plus per-SNP counts and coding LUTs.

    Replaces the reference's dense float64 ``snp_mat`` (``read_plink``; gmat/process_plink/process_plink.py:7-9)
    and its centred copies (remma_epiAA.py:59-61): a coded column is ``lut[snp][code]``.
    """

    def __init__(self, bed_file=None, n=None, m=None, payload=None, device=None):
        self.device = device or current_device()
        if bed_file is not None:
            n = count_lines(bed_file + ".fam") if n is None else n
            m = count_lines(bed_file + ".bim") if m is None else m
        self.n, self.m = int(n), int(m)
        self.row_bytes = row_stride(self.n)
        host = torch.empty((self.m, self.row_bytes), dtype=torch.uint8, pin_memory=True)
        if bed_file is not None:
            check(lib.gmat_b200_bed_read_host(bed_file.encode("ascii"), self.n, self.m, self.row_bytes,
                                              ffi.cast("unsigned char *", host.data_ptr())),
                  "gmat_b200_bed_read_host")
        else:
            src = np.frombuffer(payload, dtype=np.uint8).reshape(self.m, (self.n + 3) // 4)
            host.zero_()
            host.numpy()[:, :src.shape[1]] = src
        self.h2d_bytes = host.numel()
        self.bed = host.to(self.device, non_blocking=True)
        self._luts = {}
        self._freq = {}
        self.counts = torch.empty((self.m, 4), dtype=torch.int32, device=self.device)
        check(lib.gmat_b200_bed_stats(_p(self.bed, "unsigned char *"), self.n, self.m, self.row_bytes,
                                      _p(self.counts, "int *"), _stream()), "gmat_b200_bed_stats")

    # -- statistics ------------------------------------------------------------------------------
    def n_missing(self):
        if getattr(self, "_n_missing", None) is None:
            self._n_missing = int(self.counts[:, 1].sum().item())
        return self._n_missing

    def lut(self, coding, freq_mode=FREQ_NUMPY):
        key = (coding, freq_mode)
        if key not in self._luts:
            lut = torch.empty((self.m, 4), dtype=torch.float64, device=self.device)
            freq = torch.empty((self.m,), dtype=torch.float64, device=self.device)
            check(lib.gmat_b200_build_lut(_p(self.bed, "unsigned char *"), self.n, self.m, self.row_bytes,
                                          _p(self.counts, "int *"), coding, freq_mode,
                                          _p(lut, "double *"), _p(freq, "double *"), _stream()),
                  "gmat_b200_build_lut")
            self._luts[key] = lut
            self._freq[freq_mode] = freq
        return self._luts[key]

    def freq(self, freq_mode=FREQ_NUMPY):
        """p = allele frequency of the counted (A2) allele, numpy form (remma_epiAA.py:59)."""
        if freq_mode not in self._freq:
            self.lut(ADD, freq_mode)
        return self._freq[freq_mode]

    def decode(self, coding=None, freq_mode=FREQ_NUMPY, snp_begin=0, snp_end=None):
        """Dense float64 [snp][id] (device).  coding None = raw read_plink_bed values (code 01 -> 1/3)."""
        snp_end = self.m if snp_end is None else snp_end
        ms = snp_end - snp_begin
        out = torch.empty((ms, self.n), dtype=torch.float64, device=self.device)
        lut = None if coding is None else self.lut(coding, freq_mode)[snp_begin:snp_end]
        check(lib.gmat_b200_decode(_p(self.bed[snp_begin:snp_end], "unsigned char *"), self.n, ms, self.row_bytes,
                                   _p(lut, "double *"), _p(out, "double *"), _stream()), "gmat_b200_decode")
        return out

    @classmethod
    def from_device_payload(cls, bed_dev, n):
        """Wrap an existing device tensor uint8 [m][row_bytes] (row_bytes % 16 == 0) as a store."""
        self = cls.__new__(cls)
        self.device, self.n, self.m = bed_dev.device, int(n), int(bed_dev.shape[0])
        self.row_bytes = int(bed_dev.shape[1])
        self.h2d_bytes = 0
        self.bed = bed_dev
        self._luts, self._freq = {}, {}
        self.counts = torch.empty((self.m, 4), dtype=torch.int32, device=self.device)
        check(lib.gmat_b200_bed_stats(_p(self.bed, "unsigned char *"), self.n, self.m, self.row_bytes,
                                      _p(self.counts, "int *"), _stream()), "gmat_b200_bed_stats")
        return self

    @classmethod
    def from_dense(cls, geno_snp_major, device=None):
        """Pack a dense [snp][id] matrix of 0/1/2 (NaN -> missing) into a store (used after imputation)."""
        device = device or current_device()
        g = torch.as_tensor(np.ascontiguousarray(geno_snp_major), dtype=torch.float64).to(device)
        m, n = g.shape
        self = cls.__new__(cls)
        self.device, self.n, self.m = device, int(n), int(m)
        self.row_bytes = row_stride(self.n)
        self.h2d_bytes = g.numel() * 8
        self.bed = torch.zeros((self.m, self.row_bytes), dtype=torch.uint8, device=device)
        check(lib.gmat_b200_pack(_p(g, "double *"), self.n, self.m, self.row_bytes,
                                 _p(self.bed, "unsigned char *"), _stream()), "gmat_b200_pack")
        self._luts, self._freq = {}, {}
        self.counts = torch.empty((self.m, 4), dtype=torch.int32, device=device)
        check(lib.gmat_b200_bed_stats(_p(self.bed, "unsigned char *"), self.n, self.m, self.row_bytes,
                                      _p(self.counts, "int *"), _stream()), "gmat_b200_bed_stats")
        return self


# ----------------------------------------------------------------------------------------------
def pack_pvp(pvp_dev):
    """Re-tile Z'PZ (device, [n][n] float64) into the exact kernel's streaming order."""
    n = pvp_dev.shape[0]
    elems = lib.gmat_b200_pvp_packed_elems(n)
    packed = torch.empty((elems,), dtype=torch.float64, device=pvp_dev.device)
    check(lib.gmat_b200_pack_pvp(_p(pvp_dev, "double *"), n, _p(packed, "double *"), _stream()),
          "gmat_b200_pack_pvp")
    return packed


def hits_to_numpy(hits_dev, count):
    raw = hits_dev[:count].cpu().numpy()
    return raw.view(HIT_DTYPE).reshape(-1)


def alloc_hits(cap, device):
    return torch.empty((max(int(cap), 1), 32), dtype=torch.uint8, device=device)


def epi_exact_scan(store, lutU, lutV, py_dev, packed, ilist, pair_space, chi_cut, cap=1 << 20):
    """Run the exact scan for first-SNPs ``ilist`` (host int64 array); returns a HIT_DTYPE array."""
    dev = store.device
    il = torch.as_tensor(np.ascontiguousarray(ilist, dtype=np.int64)).to(dev)
    count = torch.zeros((1,), dtype=torch.int64, device=dev)
    while True:
        hits = alloc_hits(cap, dev)
        count.zero_()
        check(lib.gmat_b200_epi_exact_scan(
            _p(store.bed, "unsigned char *"), store.n, store.m, store.row_bytes,
            _p(lutU, "double *"), _p(lutV, "double *"), _p(py_dev, "double *"), _p(packed, "double *"),
            _p(il, "long long *"), il.numel(), pair_space, float(chi_cut),
            ffi.cast("gmat_b200_hit *", hits.data_ptr()), cap,
            ffi.cast("unsigned long long *", count.data_ptr()), _stream()), "gmat_b200_epi_exact_scan")
        cnt = int(count.item())
        if cnt <= cap:
            return hits_to_numpy(hits, cnt)
        cap = cnt               # the kernel counted every hit: one retry with the exact capacity


def exact_engine(store):
    """Which kernel runs the exact scan: "i8" (tcgen05 int8 digit planes; GMAT_B200_EXACT_ENGINE=auto or i8) or
    "f64" (DMMA)."""
    want = os.environ.get("GMAT_B200_EXACT_ENGINE", "auto").lower()
    if want not in ("auto", "i8", "f64"):
        raise ValueError("GMAT_B200_EXACT_ENGINE must be auto, i8 or f64")
    return "i8" if want == "auto" else want


def i8_planes():
    """Digit planes of the int8 engine (GMAT_B200_I8_PLANES, 2..7).  S base-256 planes carry 8S - 1 bits of
    max|diag(u) Z'PZ diag(u)|: 6 (default) = 47 bits, var within ~1e-12 of FP64 on real Z'PZ; 7 = 55 bits, finer
    than the FP64 rounding of the matrix itself.  Pairs whose variance is too small for the planes in use are
    re-evaluated in FP64 (I8Exact.RECHECK_TOL), so this is a speed knob."""
    return int(os.environ.get("GMAT_B200_I8_PLANES", "6"))


# The two digit-plane buffers of an I8Exact (4 GiB each by default) are kept for the next engine of the process: a
# fresh cudaMalloc of that size costs 20 - 400 ms (measured on the multi-GPU boxes, where every rank allocates at the
# same moment), which a scan of a second or two cannot hide; the front-ends build one engine per call (AA, then AD,
# then DD ...).  A buffer is zeroed when handed out (3 ms): the split kernel never writes the blocks right of the
# diagonal or the padding, and the geometry of the previous user may have been another.
_PLANE_FREE = {}


def _acquire_plane_buffer(dev, nbytes):
    key = (dev.index if dev.index is not None else torch.cuda.current_device(), int(nbytes))
    for other in [k for k in _PLANE_FREE if k[0] == key[0] and k != key]:
        del _PLANE_FREE[other]                         # another size is wanted now: give the old ones back to torch
    free = _PLANE_FREE.get(key)
    if free:
        return free.pop().zero_()
    return torch.zeros((int(nbytes),), dtype=torch.int8, device=dev)


def release_plane_buffers():
    """Give the kept digit-plane buffers (8 GiB per device by default) back to torch's allocator."""
    _PLANE_FREE.clear()


def _release_plane_buffers(dev, bufs):
    for b in bufs:
        key = (dev.index if dev.index is not None else torch.cuda.current_device(), int(b.numel()))
        lst = _PLANE_FREE.setdefault(key, [])
        if len(lst) < 2:
            lst.append(b)


class I8Exact(object):
    """INT8 tensor-core engine of the exact scan (gmat_b200/csrc/ozaki.cu): same statistics as
    ``epi_exact_scan`` (remma_epiAA.py:72-82), the n^2-per-pair quadratic form evaluated with tcgen05 int8 MMAs on
    an error-free digit-plane split of diag(u_i) Z'PZ diag(u_i); the two O(n)-per-pair terms ride in the same MMAs as
    four extra plane rows.  Every step is a kernel of libgmat_b200.so; torch owns the buffers, streams and events.
    Built once per scan (operands of the second SNPs stay resident: integer codes [m][ld] int8, o_j, sum a_j^2)."""

    # budget of ONE digit-plane buffer in bytes (first SNPs per launch follow from it); two exist.  GMAT_B200_I8_PLANE_GIB
    PLANE_BYTES = int(os.environ.get("GMAT_B200_I8_PLANE_GIB", "4")) << 30
    MAX_BATCH = 128                # first SNPs per launch at most
    RECHECK_TOL = 2.5e-10          # a pair is re-evaluated in FP64 when 10 sigma of the split error > this x |var|
    RECHECK_CAP = 1 << 20
    FIRST_BATCH = 16               # the first launch of a scan is small: its operands cannot be built behind a scan
    MINOR_ALLELE = True            # code additive second SNPs by their minor allele (False: A/B runs of scripts/i8_accuracy.py)

    def __init__(self, store, lutU, lutV, coding_v, py_dev, pvp_dev, n_planes=None, packed=None):
        self.store, self.lutU, self.lutV, self.py = store, lutU, lutV, py_dev.reshape(-1).contiguous()
        self.A = pvp_dev
        self._packed = packed          # Z'PZ in the FP64 pair kernel's order (built on first use: re-checks are rare)
        self.n_planes = int(i8_planes() if n_planes is None else n_planes)
        if not 2 <= self.n_planes <= 7:
            raise ValueError("GMAT_B200_I8_PLANES must be in [2, 7]")
        n, m, dev = store.n, store.m, store.device
        self.ld = int(lib.gmat_b200_i8_ld(n))
        self.g8 = torch.empty((m, self.ld), dtype=torch.int8, device=dev)
        self.g2 = torch.empty((m, self.ld // 4), dtype=torch.uint8, device=dev)
        # with the counts: additive SNPs are coded by their MINOR allele (a' = 2 - a where 11 outnumbers 00): same
        # var, eff sign restored by the scan, and the split error -- proportional to sum a^2 -- up to 35x smaller
        check(lib.gmat_b200_i8_dosage(_p(store.bed, "unsigned char *"), n, m, store.row_bytes, coding_v,
                                      _p(store.counts, "int *") if self.MINOR_ALLELE else ffi.NULL,
                                      _p(self.g8, "signed char *"), _p(self.g2, "unsigned char *"), self.ld,
                                      _stream()), "gmat_b200_i8_dosage")
        self.offs = torch.empty((m,), dtype=torch.float64, device=dev)
        self.asq = torch.empty((m,), dtype=torch.float64, device=dev)
        check(lib.gmat_b200_i8_offs(_p(store.counts, "int *"), _p(lutV, "double *"), m, coding_v,
                                    1 if self.MINOR_ALLELE else 0,
                                    _p(self.offs, "double *"), _p(self.asq, "double *"), _stream()),
              "gmat_b200_i8_offs")
        self.amax = torch.empty((1,), dtype=torch.float64, device=dev)
        check(lib.gmat_b200_absmax(_p(self.A, "double *"), self.A.numel(), _p(self.amax, "double *"), _stream()),
              "gmat_b200_absmax")
        per_first = self.n_planes * self.ld * self.ld
        self.batch = int(max(1, min(self.MAX_BATCH, self.PLANE_BYTES // per_first)))
        # two plane buffers (and two sets of per-batch scratch): the operands of batch k+1 are built on a side
        # stream while the scan of batch k is running
        self.planes = [_acquire_plane_buffer(dev, self.batch * per_first) for _ in range(2)]
        self.u = [torch.empty((self.batch, n), dtype=torch.float64, device=dev) for _ in range(2)]
        self.r = [torch.empty((self.batch, n), dtype=torch.float64, device=dev) for _ in range(2)]
        self.side = [torch.empty((self.batch, 8), dtype=torch.float64, device=dev) for _ in range(2)]
        self._prep_stream = torch.cuda.Stream(device=dev)
        # kernel variant: 3 = only the plane rows that can be non-zero are issued (default); 2 = full 256-row MMAs
        # (same bits, kept for A/B)
        self.wide = int(os.environ.get("GMAT_B200_I8_VARIANT", "3"))
        if self.wide not in (2, 3):
            raise ValueError("GMAT_B200_I8_VARIANT must be 2 or 3")
        self.kernel_events = None      # bench.py: list of (start, end) CUDA events around every i8_scan launch
        self.last_rechecked = 0        # pairs of the last scan that were re-evaluated in FP64

    def __del__(self):
        try:
            planes, self.planes = self.planes, []
            torch.cuda.current_stream(self.store.device).synchronize()      # nothing of this engine is still running
            self._prep_stream.synchronize()
            _release_plane_buffers(self.store.device, planes)
        except Exception:                                                    # interpreter shutdown
            pass

    def _work_items(self, ih, pair_space):
        """(item_prefix[n_i+1], jt0[n_i]) int32 host arrays: tiles of 256 second SNPs every first SNP needs."""
        m = self.store.m
        tj = 256                                       # second SNPs per work item (a CTA pair)
        n_jt = (m + tj - 1) // tj
        if pair_space == PAIRS_ALL:
            jt0 = np.zeros(len(ih), dtype=np.int64)
            cnt = np.full(len(ih), n_jt, dtype=np.int64)
        else:
            jt0 = (ih + 1) // tj
            cnt = np.where(ih + 1 < m, n_jt - jt0, 0)
        prefix = np.concatenate([[0], np.cumsum(cnt)])
        return prefix.astype(np.int32), jt0.astype(np.int32)

    def packed(self):
        if self._packed is None:
            self._packed = pack_pvp(self.A)
        return self._packed

    def scan(self, ilist, pair_space, chi_cut, cap=1 << 20, q_debug=False):
        """All launches of the scan are queued back to back (one hit buffer, one counter, no host synchronisation
        between first-SNP batches); the counters are read once at the end and the scan repeated with the exact
        capacity in the rare case a buffer was too small.  Per batch (plane-buffer budget) the operand kernels
        (first-SNP decode, Z'PZ u, side table, digit split) run on a side stream into the other plane buffer while
        the tensor-core scan of the previous batch is still running."""
        store, dev = self.store, self.store.device
        n, m = store.n, store.m
        ilist = np.ascontiguousarray(ilist, dtype=np.int64)
        if len(ilist) and (ilist.min() < 0 or ilist.max() >= m):
            raise ValueError("first SNP out of range")
        count = torch.zeros((2,), dtype=torch.int64, device=dev)          # [hits, re-checks]
        main = torch.cuda.current_stream()
        prep = self._prep_stream
        il_dev = torch.as_tensor(ilist).to(dev, non_blocking=True)
        # batches: a short one first (nothing hides its operand kernels), then full plane buffers
        bounds, b0 = [], 0
        while b0 < len(ilist):
            ni = min(self.FIRST_BATCH if not bounds else self.batch, self.batch, len(ilist) - b0)
            bounds.append((b0, ni))
            b0 += ni
        # work tables of every batch in one upload: [prefix (ni+1), jt0 (ni)] per batch, back to back
        tabs, offs, totals = [], [], []
        at = 0
        for b0, ni in bounds:
            prefix, jt0 = self._work_items(ilist[b0:b0 + ni], pair_space)
            tabs += [prefix, jt0]
            offs.append(at)
            totals.append(int(prefix[-1]))
            at += len(prefix) + len(jt0)
        tab_dev = torch.as_tensor(np.concatenate(tabs) if tabs else np.zeros(1, np.int32)).to(dev, non_blocking=True)
        rcap = self.RECHECK_CAP
        while True:
            hits = alloc_hits(cap, dev)
            recheck = torch.empty((max(rcap, 1), 2), dtype=torch.int32, device=dev)
            count.zero_()
            inputs_ready = main.record_event()
            dbg = []
            scan_done = [None, None]                                        # event per plane buffer
            k = 0
            for bi, (b0, ni) in enumerate(bounds):
                total = totals[bi]
                if total == 0:
                    continue
                slot = k & 1
                planes, u, r, side = self.planes[slot], self.u[slot], self.r[slot], self.side[slot]
                il_ptr = ffi.cast("long long *", il_dev.data_ptr() + 8 * b0)
                with torch.cuda.stream(prep):
                    prep.wait_event(inputs_ready)
                    if scan_done[slot] is not None:
                        prep.wait_event(scan_done[slot])                    # the scan that last read this buffer
                    check(lib.gmat_b200_i8_first(
                        _p(store.bed, "unsigned char *"), n, m, store.row_bytes, _p(self.lutU, "double *"),
                        _p(self.py, "double *"), _p(self.A, "double *"), _p(self.amax, "double *"), il_ptr, ni,
                        self.n_planes, _p(u, "double *"), _p(r, "double *"), _p(side, "double *"),
                        _p(planes, "signed char *"), _stream()), "gmat_b200_i8_first")
                    check(lib.gmat_b200_i8_split(_p(self.A, "double *"), n, _p(u, "double *"), ni,
                                                 _p(side, "double *"), self.n_planes,
                                                 _p(planes, "signed char *"), _stream()), "gmat_b200_i8_split")
                    split_done = prep.record_event()
                main.wait_event(split_done)
                qd = torch.zeros((ni, m), dtype=torch.float64, device=dev) if q_debug else None
                if self.kernel_events is not None:
                    ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                    ev[0].record()
                tab_ptr = tab_dev.data_ptr() + 4 * offs[bi]
                check(lib.gmat_b200_i8_scan(
                    _p(self.g8, "signed char *"), _p(self.g2, "unsigned char *"), n, m, _p(planes, "signed char *"),
                    self.n_planes, il_ptr, ni,
                    ffi.cast("int *", tab_ptr), ffi.cast("int *", tab_ptr + 4 * (ni + 1)), total,
                    _p(side, "double *"), _p(self.offs, "double *"), _p(self.asq, "double *"),
                    pair_space, float(chi_cut), ffi.cast("gmat_b200_hit *", hits.data_ptr()), cap,
                    ffi.cast("unsigned long long *", count.data_ptr()), _p(recheck, "int *"), rcap,
                    ffi.cast("unsigned long long *", count.data_ptr() + 8), float(self.RECHECK_TOL),
                    _p(qd, "double *"), int(self.wide), _stream()), "gmat_b200_i8_scan")
                if self.kernel_events is not None:
                    ev[1].record()
                    self.kernel_events.append(ev)
                scan_done[slot] = main.record_event()
                if q_debug:
                    dbg.append(qd)
                k += 1
            cnt, rcnt = (int(v) for v in count.tolist())                    # the one synchronisation of the scan
            prep.synchronize()
            if cnt <= cap and rcnt <= rcap:
                break
            cap, rcap = max(cap, cnt), max(rcap, rcnt)
        res = hits_to_numpy(hits, cnt)
        self.last_rechecked = rcnt
        if rcnt:
            # pairs the digit planes could not decide: FP64 DMMA pair kernel, then the same threshold
            pairs = recheck[:rcnt].cpu().numpy()
            ex = epi_exact_pairs(store, self.lutU, self.lutV, self.py, self.packed(), pairs)
            ex = ex[ex["chi"] > chi_cut] if chi_cut >= 0.0 else ex
            res = np.concatenate([res, ex])
        return (res, torch.cat(dbg)) if q_debug else res


def epi_exact_pairs(store, lutU, lutV, py_dev, packed, pairs):
    """Exact statistics for explicit pairs (host int array [k][2]); returns HIT_DTYPE array, list order."""
    dev = store.device
    pr = torch.as_tensor(np.ascontiguousarray(pairs, dtype=np.int32)).to(dev)
    k = pr.shape[0]
    out = alloc_hits(k, dev)
    check(lib.gmat_b200_epi_exact_pairs(
        _p(store.bed, "unsigned char *"), store.n, store.m, store.row_bytes,
        _p(lutU, "double *"), _p(lutV, "double *"), _p(py_dev, "double *"), _p(packed, "double *"),
        ffi.cast("int *", pr.data_ptr()), k, ffi.cast("gmat_b200_hit *", out.data_ptr()), _stream()),
        "gmat_b200_epi_exact_pairs")
    return hits_to_numpy(out, k)


def pair_engine(store, n_pairs):
    """Which kernel tests an explicit pair list: "i8" (int8 tensor cores, I8Pairs) or "f64" (DMMA pair kernel).
    GMAT_B200_PAIR_ENGINE = auto (default) | i8 | f64.  auto = i8 where it is measured faster: I8Pairs.MIN_IDS <= n <
    14000, at least I8Pairs.MIN_PAIRS pairs, no missing calls (5.3 operand rows per pair at S n^2 / 2 int8 MACs each
    against n^2 issued FP64 FMAs: 2.3x the DMMA kernel at n = 5000, 1.26x at n = 2000, scripts/pairs_speed.py)."""
    want = os.environ.get("GMAT_B200_PAIR_ENGINE", "auto").lower()
    if want not in ("auto", "i8", "f64"):
        raise ValueError("GMAT_B200_PAIR_ENGINE must be auto, i8 or f64")
    if want == "f64":
        return "f64"
    fits = store.n < 14000 and store.n_missing() == 0
    if want == "i8":
        if not fits:
            raise ValueError("the int8 pair engine needs n < 14000 and genotypes without missing calls")
        return "i8"
    return "i8" if (fits and store.n >= I8Pairs.MIN_IDS and n_pairs >= I8Pairs.MIN_PAIRS) else "f64"


class I8Pairs(object):
    """Exact test of explicit pair lists on the int8 tensor cores (gmat_b200/csrc/ozaki.cu, "pair lists";
    remma_epiAA_pair.py:66-87): digit planes of Z'PZ itself (built once per model), four integer operand rows per pair
    (q = a o b, q + a, q + b, a + b) plus one per distinct SNP, var / eff by polarisation.  Pairs the planes cannot
    resolve are re-evaluated by the FP64 pair kernel, like the scan's.  Same statistics as epi_exact_pairs."""

    MIN_PAIRS = 16384             # auto: below this the fixed costs (planes of Z'PZ, one launch) are not worth it
    MIN_IDS = 2000                # auto: below this n the FP64 DMMA pair kernel is as fast
    CHUNK = 65536                 # pairs per launch (operand rows: <= 6 x CHUNK x ld bytes = 2 GB at n = 5000)
    RECHECK_TOL = 2.5e-10
    # Seven planes (55 bits), not the scan's six: the operand rows are dense sums of two genotype vectors (sum x^2 up
    # to 36 n), and a re-check is dear here -- the FP64 pair kernel works in tiles of 128 pairs, so the 0.5 % of a
    # random list that six planes leave weak (231 of 50 000 pairs at n = 5000) took 13.9 ms on two SMs, more than the
    # whole tensor-core pass (11 ms).  With seven nothing but degenerate pairs is left: +17 % MMAs, -45 % time.
    PLANES = 7

    def __init__(self, store, coding_u, coding_v, lutU, lutV, py_dev, pvp_dev, packed=None, n_planes=None):
        self.store, self.cu, self.cv, self.lutU, self.lutV = store, int(coding_u), int(coding_v), lutU, lutV
        self.py = py_dev.reshape(-1).contiguous()
        self.A, self._packed = pvp_dev, packed
        self.n_planes = int(self.PLANES if n_planes is None else n_planes)
        n, dev = store.n, store.device
        self.ld = int(lib.gmat_b200_i8_ld(n))
        amax = torch.empty((1,), dtype=torch.float64, device=dev)
        check(lib.gmat_b200_absmax(_p(self.A, "double *"), self.A.numel(), _p(amax, "double *"), _stream()),
              "gmat_b200_absmax")
        ones = torch.ones((1, n), dtype=torch.float64, device=dev)
        self.t1 = torch.empty((1, n), dtype=torch.float64, device=dev)           # Z'PZ 1
        self.side = torch.empty((1, 8), dtype=torch.float64, device=dev)
        self.planes = torch.zeros((self.n_planes * self.ld * self.ld,), dtype=torch.int8, device=dev)
        check(lib.gmat_b200_i8_first(ffi.NULL, n, store.m, store.row_bytes, ffi.NULL, _p(self.py, "double *"),
                                     _p(self.A, "double *"), _p(amax, "double *"), ffi.NULL, 1, self.n_planes,
                                     _p(ones, "double *"), _p(self.t1, "double *"), _p(self.side, "double *"),
                                     _p(self.planes, "signed char *"), _stream()), "gmat_b200_i8_first")
        check(lib.gmat_b200_i8_split(_p(self.A, "double *"), n, _p(ones, "double *"), 1, _p(self.side, "double *"),
                                     self.n_planes, _p(self.planes, "signed char *"), _stream()), "gmat_b200_i8_split")
        self._ilist = torch.zeros((1,), dtype=torch.int64, device=dev)
        self._jt0 = torch.zeros((1,), dtype=torch.int32, device=dev)
        self._counters = torch.zeros((2,), dtype=torch.int64, device=dev)
        self.last_rechecked = 0

    def packed(self):
        if self._packed is None:
            self._packed = pack_pvp(self.A)
        return self._packed

    def stats(self, pairs):
        """HIT_DTYPE array (snp_0, snp_1, eff, var, chi) in list order."""
        pairs = np.ascontiguousarray(pairs, dtype=np.int32).reshape(-1, 2)
        st, dev, n, ld = self.store, self.store.device, self.store.n, self.ld
        outs, weak_all = [], []
        for c0 in range(0, len(pairs), self.CHUNK):
            ch = pairs[c0:c0 + self.CHUNK]
            k = len(ch)
            su, iu = np.unique(ch[:, 0], return_inverse=True)
            sv, iv = np.unique(ch[:, 1], return_inverse=True)
            rows = 4 * k + len(su) + len(sv)
            up = lambda a: torch.as_tensor(np.ascontiguousarray(a, dtype=np.int32)).to(dev)
            pr, su_d, sv_d, iu_d, iv_d = up(ch), up(su), up(sv), up(iu), up(iv)
            x8 = torch.empty((rows, ld), dtype=torch.int8, device=dev)
            x4 = torch.empty((rows, ld // 2), dtype=torch.uint8, device=dev)
            asq = torch.empty((rows,), dtype=torch.float64, device=dev)
            xt1 = torch.empty((rows,), dtype=torch.float64, device=dev)
            q = torch.empty((rows,), dtype=torch.float64, device=dev)
            e = torch.empty((rows,), dtype=torch.float64, device=dev)
            zeros = torch.zeros((rows,), dtype=torch.float64, device=dev)
            prefix = torch.tensor([0, (rows + 255) // 256], dtype=torch.int32, device=dev)
            check(lib.gmat_b200_i8_pair_rows(
                _p(st.bed, "unsigned char *"), n, st.row_bytes, self.cu, self.cv, _p(pr, "int *"), k,
                _p(su_d, "int *"), len(su), _p(sv_d, "int *"), len(sv), _p(self.t1, "double *"),
                _p(x8, "signed char *"), _p(x4, "unsigned char *"), _p(asq, "double *"), _p(xt1, "double *"),
                _stream()), "gmat_b200_i8_pair_rows")
            check(lib.gmat_b200_i8_rows_scan(
                _p(x8, "signed char *"), _p(x4, "unsigned char *"), n, rows, _p(self.planes, "signed char *"),
                self.n_planes, _p(self.side, "double *"), _p(zeros, "double *"), _p(asq, "double *"),
                _p(self._ilist, "long long *"), _p(prefix, "int *"), _p(self._jt0, "int *"), _p(q, "double *"),
                _p(e, "double *"), ffi.cast("unsigned long long *", self._counters.data_ptr()),
                ffi.cast("unsigned long long *", self._counters.data_ptr() + 8), _stream()), "gmat_b200_i8_rows_scan")
            out = alloc_hits(k, dev)
            weak = torch.empty((k,), dtype=torch.uint8, device=dev)
            check(lib.gmat_b200_i8_pairs_combine(
                _p(pr, "int *"), k, _p(iu_d, "int *"), len(su), _p(iv_d, "int *"), _p(self.lutU, "double *"),
                _p(self.lutV, "double *"), _p(q, "double *"), _p(e, "double *"), _p(xt1, "double *"),
                _p(asq, "double *"), _p(self.side, "double *"), float(self.RECHECK_TOL),
                ffi.cast("gmat_b200_hit *", out.data_ptr()), _p(weak, "unsigned char *"), _stream()),
                "gmat_b200_i8_pairs_combine")
            outs.append(out[:k])
            weak_all.append(weak)
        if not outs:
            return np.zeros((0,), dtype=HIT_DTYPE)
        res = torch.cat(outs).cpu().numpy().view(HIT_DTYPE).reshape(-1)
        wk = np.flatnonzero(torch.cat(weak_all).cpu().numpy())
        self.last_rechecked = int(len(wk))
        if len(wk):                               # rare variants: the FP64 pair kernel decides these
            fix = epi_exact_pairs(st, self.lutU, self.lutV, self.py, self.packed(), pairs[wk])
            res = res.copy()
            res[wk] = fix
        return res


def screen_engine(store):
    """Which engine runs the approximate screen: "i8" (int8 tensor-core prefilter + FP64 re-test of the survivors;
    stores with missing calls use the tripled integer coding, see ScreenOperands) or "f64" (DMMA).
    GMAT_B200_SCREEN_ENGINE = auto (default) | i8 | f64."""
    want = os.environ.get("GMAT_B200_SCREEN_ENGINE", "auto").lower()
    if want not in ("auto", "i8", "f64"):
        raise ValueError("GMAT_B200_SCREEN_ENGINE must be auto, i8 or f64")
    if want == "f64":
        return "f64"
    fits = store.n < (29000 if store.n_missing() > 0 else 60000)   # int32 plane sums must stay FP32-exact
    if want == "i8" and not fits:
        raise ValueError("the int8 screen needs n < 60000 (29000 with missing calls)")
    return "i8" if fits else "f64"


LAST_SCREEN_ENGINE = None      # "i8" / "f64": what the last epi_eff call actually ran (tests, bench)
LAST_SCREEN_NOTE = ""          # why, when the int8 path was asked for and declined


class ScreenOperands(object):
    """Operands of the int8 screen that depend only on (genotype store, Z'Py): the integer codes of every SNP per
    coding ([m][ld] int8) with t = a . Z'Py, the digits of Z'Py and its statistics, and the candidate buffer.  Built
    once and kept across screen calls -- the two directions of the AD screen, the stages of an _approx pipeline,
    repeated calls -- instead of re-coding the whole store on every call.  ``tripled``: the store has missing calls;
    the reference C screen does not impute (code 01 is the value 1/3, _remma_epi_eff_cpu.c:42), so the codes are 3a
    in {0,1,3,6} / {0,1,3,0} and Z'Py is cut into three base-32 digits instead of two base-128 ones."""

    CAND_CAP = 1 << 22

    def __init__(self, store, py_dev):
        self.store, self.py = store, py_dev.reshape(-1).contiguous()
        self.tripled = 1 if store.n_missing() > 0 else 0
        dev = store.device
        self.ld = int(lib.gmat_b200_screen_ld(store.n))
        self._ops = {}
        self.dig = torch.empty((3, self.ld), dtype=torch.int8, device=dev)
        self.stats = torch.empty((4,), dtype=torch.float64, device=dev)
        check(lib.gmat_b200_screen_py(_p(self.py, "double *"), store.n, self.tripled, _p(self.dig, "signed char *"),
                                      _p(self.stats, "double *"), _stream()), "gmat_b200_screen_py")
        self.cand = torch.empty((self.CAND_CAP, 2), dtype=torch.int32, device=dev)

    def operand(self, coding):
        if coding not in self._ops:
            st = self.store
            g8 = torch.empty((st.m, self.ld), dtype=torch.int8, device=st.device)
            t = torch.empty((st.m,), dtype=torch.float64, device=st.device)
            check(lib.gmat_b200_screen_operand(_p(st.bed, "unsigned char *"), st.n, st.m, st.row_bytes, int(coding),
                                               self.tripled, _p(self.py, "double *"), _p(g8, "signed char *"),
                                               _p(t, "double *"), _stream()), "gmat_b200_screen_operand")
            self._ops[coding] = (g8, t)
        return self._ops[coding]


def epi_eff(store, lutU, lutV, py_dev, ilist, flags, eff_cut, cut_table=None, binI=None, binJ=None,
            cap=1 << 20, codings=None, ops=None):
    """Approximate screen on device buffers; returns HIT_DTYPE array (unordered).  ``codings`` = (coding of lutU,
    coding of lutV): with it the int8 tensor-core engine runs (see screen_engine); without it the FP64 DMMA kernel.
    ``ops``: a ScreenOperands of (store, py_dev) to reuse across calls (else the operands are built for this call)."""
    global LAST_SCREEN_ENGINE, LAST_SCREEN_NOTE
    dev = store.device
    il = torch.as_tensor(np.ascontiguousarray(ilist, dtype=np.int64)).to(dev)
    count = torch.zeros((3,), dtype=torch.int64, device=dev)       # [hits, stage-1 survivors, status]
    tab = bI = bJ = None
    if cut_table is not None:
        tab = torch.as_tensor(np.ascontiguousarray(cut_table, dtype=np.float64)).to(dev)
        bI = torch.as_tensor(np.ascontiguousarray(binI, dtype=np.int32)).to(dev)
        bJ = torch.as_tensor(np.ascontiguousarray(binJ, dtype=np.int32)).to(dev)
    which = screen_engine(store) if codings is not None else "f64"
    LAST_SCREEN_NOTE = ""
    used = ffi.new("int *")
    while True:
        hits = alloc_hits(cap, dev)
        count.zero_()
        if which == "i8" and ops is not None:
            g8U, tU = ops.operand(codings[0])
            g8V, tV = ops.operand(codings[1])
            check(lib.gmat_b200_epi_eff_i8p(
                _p(store.bed, "unsigned char *"), store.n, store.m, store.row_bytes,
                _p(lutU, "double *"), _p(lutV, "double *"), _p(g8U, "signed char *"), _p(g8V, "signed char *"),
                _p(tU, "double *"), _p(tV, "double *"), _p(ops.dig, "signed char *"), _p(ops.stats, "double *"),
                ops.tripled, _p(py_dev, "double *"), _p(il, "long long *"), il.numel(), flags, float(eff_cut),
                _p(tab, "double *"), _p(bI, "int *"), _p(bJ, "int *"),
                ffi.cast("gmat_b200_hit *", hits.data_ptr()), cap,
                ffi.cast("unsigned long long *", count.data_ptr()), _p(ops.cand, "int *"), ops.CAND_CAP,
                ffi.cast("unsigned long long *", count.data_ptr() + 8), ffi.cast("int *", count.data_ptr() + 16),
                _stream()), "gmat_b200_epi_eff_i8p")
            cnt, _, status = (int(v) for v in count.tolist())       # the one synchronisation of the call
            if status != 0:
                LAST_SCREEN_NOTE = ("margin not small against the cut" if status == 1 else
                                    "more survivors than the candidate buffer") + ": FP64 DMMA kernel"
                which = "f64"
                continue
            LAST_SCREEN_ENGINE = "i8"
        elif which == "i8":
            check(lib.gmat_b200_epi_eff_i8(
                _p(store.bed, "unsigned char *"), store.n, store.m, store.row_bytes,
                _p(lutU, "double *"), _p(lutV, "double *"), int(codings[0]), int(codings[1]),
                1 if store.n_missing() > 0 else 0, _p(py_dev, "double *"),
                _p(il, "long long *"), il.numel(), flags, float(eff_cut),
                _p(tab, "double *"), _p(bI, "int *"), _p(bJ, "int *"),
                ffi.cast("gmat_b200_hit *", hits.data_ptr()), cap,
                ffi.cast("unsigned long long *", count.data_ptr()), used, _stream()), "gmat_b200_epi_eff_i8")
            LAST_SCREEN_ENGINE = "i8" if used[0] == 1 else "f64"
            cnt = int(count[0].item())
        else:
            check(lib.gmat_b200_epi_eff(
                _p(store.bed, "unsigned char *"), store.n, store.m, store.row_bytes,
                _p(lutU, "double *"), _p(lutV, "double *"), _p(py_dev, "double *"),
                _p(il, "long long *"), il.numel(), flags, float(eff_cut),
                _p(tab, "double *"), _p(bI, "int *"), _p(bJ, "int *"),
                ffi.cast("gmat_b200_hit *", hits.data_ptr()), cap,
                ffi.cast("unsigned long long *", count.data_ptr()), _stream()), "gmat_b200_epi_eff")
            LAST_SCREEN_ENGINE = "f64"
            cnt = int(count[0].item())
        if cnt <= cap:
            return hits_to_numpy(hits, cnt)
        cap = cnt


def grm_engine(store, snp_begin=0, snp_end=None):
    """Which kernel builds the GRM: "f4" (one exact FP4 tensor-core GEMM, tcgen05 kind::mxf4; needs a SNP range
    without missing genotypes, shorter than 2^21 SNPs), "i8" (the same as an int8 GEMM) or "f64" (DMMA).
    GMAT_B200_GRM_ENGINE = auto (default) | f4 | i8 | f64."""
    want = os.environ.get("GMAT_B200_GRM_ENGINE", "auto").lower()
    if want not in ("auto", "f4", "i8", "f64"):
        raise ValueError("GMAT_B200_GRM_ENGINE must be auto, f4, i8 or f64")
    if want == "f64":
        return "f64"
    snp_end = store.m if snp_end is None else snp_end
    clean = int(store.counts[snp_begin:snp_end, 1].sum().item()) == 0
    if want in ("i8", "f4") and not clean:
        raise ValueError("the tensor-core GRM kernels need genotypes without missing calls")
    if not clean:
        return "f64"
    short = snp_end - snp_begin < (1 << 21)
    if want == "f4" and not short:
        raise ValueError("the FP4 GRM kernel needs a SNP range shorter than 2^21")
    return "i8" if (want == "i8" or not short) else "f4"


def grm(store, coding, scale, small_val, snp_begin=0, snp_end=None, out=None, accumulate=False, which=None):
    """K = X X'/scale (+ diagonal inflation) on the device; returns the [n][n] float64 device tensor."""
    snp_end = store.m if snp_end is None else snp_end
    lut = store.lut(coding, FREQ_NUMPY)
    if out is None:
        out = torch.zeros((store.n, store.n), dtype=torch.float64, device=store.device) if accumulate \
            else torch.empty((store.n, store.n), dtype=torch.float64, device=store.device)
    which = grm_engine(store, snp_begin, snp_end) if which is None else which
    if which in ("i8", "f4"):
        fn = lib.gmat_b200_grm_i8 if which == "i8" else lib.gmat_b200_grm_f4
        check(fn(_p(store.bed, "unsigned char *"), store.n, store.m, store.row_bytes,
                 _p(lut, "double *"), coding, snp_begin, snp_end, float(scale), float(small_val),
                 1 if accumulate else 0, _p(out, "double *"), _stream()), "gmat_b200_grm_" + which)
        return out
    check(lib.gmat_b200_grm(_p(store.bed, "unsigned char *"), store.n, store.m, store.row_bytes,
                            _p(lut, "double *"), snp_begin, snp_end, float(scale), float(small_val),
                            1 if accumulate else 0, _p(out, "double *"), _stream()), "gmat_b200_grm")
    return out


def decode_into(store, coding, out, freq_mode=FREQ_NUMPY):
    """decode the whole store into a caller-owned [m][n] float64 device tensor (no allocation: timing)."""
    lut = None if coding is None else store.lut(coding, freq_mode)
    check(lib.gmat_b200_decode(_p(store.bed, "unsigned char *"), store.n, store.m, store.row_bytes,
                               _p(lut, "double *"), _p(out, "double *"), _stream()), "gmat_b200_decode")
    return out
