"""Drop-in for gmat/gmatrix/gmatrix.py: ``agmat`` / ``dgmat_as`` with the same signatures, return values
and output files (``bed_file.agrm0`` / ``.dgrm_as0`` ...), the GEMM done by the CUDA GRM kernel straight from
the 2-bit genotype store (no dense float64 genotype matrix is ever formed)."""
import logging
import sys
import time

import numpy as np
import pandas as pd
import torch

from .. import engine
from .._ffi import check, ffi, lib
from ..process_plink.process_plink import load_store
from ..remma import parallel as par


def output_mat(mat, id, out_file, out_fmt):
    """Same three formats as gmatrix.py:10-31."""
    if out_fmt == 'mat':
        # same bytes as np.savetxt(out_file + '0', mat), formatted by all host threads (include/gmat_b200.h)
        mat = np.ascontiguousarray(mat, dtype=np.float64)
        check(lib.gmat_b200_savetxt_host((out_file + '0').encode(), ffi.cast("double *", mat.ctypes.data),
                                         mat.shape[0], mat.shape[1], 0), "gmat_b200_savetxt_host")
    elif out_fmt in ('row_col_val', 'id_id_val'):
        # gmatrix.py:15-29: pandas to_csv of the lower triangle (np.tril_indices order).  Same bytes, written by
        # all host threads (gmat_b200_write_tril_host: repr(float) values, labels as pandas prints them).
        mat = np.ascontiguousarray(mat, dtype=np.float64)
        if out_fmt == 'row_col_val':
            ids, off, path = ffi.NULL, ffi.NULL, out_file + '1'
        else:
            labels = _csv_labels(id)
            blob = "".join(labels).encode()
            lens = np.array([len(x.encode()) for x in labels], dtype=np.int64)
            offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
            ids, off, path = ffi.from_buffer(blob), ffi.cast("long long *", offs.ctypes.data), out_file + '2'
        check(lib.gmat_b200_write_tril_host(path.encode(), ffi.cast("double *", mat.ctypes.data), mat.shape[0],
                                            ids, off, 0), "gmat_b200_write_tril_host")
    else:
        return 0
    return 1


def _csv_labels(id):
    """The n individual labels exactly as ``DataFrame.to_csv(sep=' ')`` prints them (ints, floats, strings, quoting
    of labels that hold the separator): pandas itself formats the n labels once, the n(n+1)/2 lines reuse them."""
    text = pd.DataFrame({"id": np.asarray(id)}).to_csv(sep=' ', index=False, header=False)
    labels = text.split("\n")[:-1]
    if len(labels) != len(id):                                   # a label with an embedded newline: not a .fam id
        raise ValueError("individual ids must not contain line breaks")
    return labels


def _read_ids(bed_file):
    fam_info = pd.read_csv(bed_file + '.fam', sep=r'\s+', header=None)
    return np.array(fam_info.iloc[:, 1])


def grm_device(store, coding, small_val=0.001):
    """(K on the device, scale).  coding ADD: gmatrix.py:53-66; coding DOM: gmatrix.py:116-130."""
    freq = store.freq().cpu().numpy().reshape(1, -1)
    if coding == engine.ADD:
        scale = np.sum(2 * freq * (1 - freq))                     # gmatrix.py:55-56
    else:
        scale_vec = 2 * freq * (1 - freq)
        scale = np.sum(scale_vec * (1 - scale_vec))               # gmatrix.py:118-119
    rank, world = par.rank_world()
    if world == 1:
        kin = engine.grm(store, coding, float(scale), small_val)
        return kin, float(scale)
    # one process per GPU: the K dimension (SNPs) is sharded in contiguous slices aligned to the kernel's
    # 64-SNP stages, every rank accumulates its partial X X' with no communication, one all_reduce (NCCL over
    # NVLink; 800 MB at n = 10,000) sums them, and every rank applies the same scale / diagonal inflation -- so
    # all ranks hold bit-identical copies, as the preamble of the scan that follows needs.
    import torch.distributed as dist
    per = (store.m + world - 1) // world
    per = (per + 63) // 64 * 64
    s0, s1 = min(store.m, rank * per), min(store.m, (rank + 1) * per)
    kin = engine.grm(store, coding, 1.0, 0.0, snp_begin=s0, snp_end=s1, accumulate=True)
    dist.all_reduce(kin)
    kin /= float(scale)
    d = torch.diagonal(kin)
    d += d * small_val                                            # gmatrix.py:64-66
    return kin, float(scale)


def _gmat(bed_file, coding, suffix, inv_suffix, inv, small_val, out_fmt, title):
    logging.info("{:#^80}".format("Read the SNP data"))
    store = load_store(bed_file, warn=logging.info)
    logging.info("There are {:d} individuals and {:d} SNPs.".format(store.n, store.m))
    logging.info("{:#^80}".format("Calculate the %s genomic relationship matrix" % title))
    clock_t0 = time.perf_counter()
    cpu_t0 = time.process_time()
    kin_dev, scale = grm_device(store, coding, small_val)
    torch.cuda.synchronize()
    clock_t1 = time.perf_counter()
    cpu_t1 = time.process_time()
    logging.info('The scaled factor is: {:.3f}'.format(scale))
    logging.info("Running time: Clock time, {:.5f} sec; CPU time, {:.5f} sec.".format(clock_t1 - clock_t0,
                                                                                   cpu_t1 - cpu_t0))
    kin = kin_dev.cpu().numpy()
    logging.info("{:#^80}".format("Output"))
    writer = par.rank_world()[0] == 0                             # under torchrun only rank 0 writes the files
    id = _read_ids(bed_file)
    out_file = bed_file + suffix
    logging.info("The output file is " + out_file)
    if out_fmt not in ('mat', 'row_col_val', 'id_id_val'):
        logging.error('Not Recognized output format: ' + out_fmt)
        sys.exit()
    if writer:
        output_mat(kin, id, out_file, out_fmt)
    kin_inv = None
    if inv:
        logging.info("{:#^80}".format("Calculate the inversion of kinship"))
        kin_inv = torch.linalg.inv(kin_dev).cpu().numpy()         # gmatrix.py:85 (scipy.linalg.inv)
        out_file = bed_file + inv_suffix
        logging.info("The output file is: " + out_file)
        if writer:
            output_mat(kin_inv, id, out_file, out_fmt)
    return kin, kin_inv


def agmat(bed_file, inv=False, small_val=0.001, out_fmt='mat'):
    """additive genomic relationship matrix and its inversion (reference: gmatrix.py:34-94)."""
    return _gmat(bed_file, engine.ADD, '.agrm', '.agiv', inv, small_val, out_fmt, 'additive')


def dgmat_as(bed_file, inv=False, small_val=0.001, out_fmt='mat'):
    """dominance genomic relationship matrix and its inversion (reference: gmatrix.py:97-159)."""
    return _gmat(bed_file, engine.DOM, '.dgrm_as', '.dgiv_as', inv, small_val, out_fmt, 'dominance')
