"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy + plain C) of GMAT's REMMAX hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module, and only as the checker or the timed CPU baseline.  Nothing under gmat_b200/ imports it.

Pinning: every function below is checked in tests/test_oracle.py against
  (a) tests/golden/*  -- outputs of the UNMODIFIED reference run in the build container by
      oracle/make_golden.py (incl. the reference's own golden vector var_com of
      examples/remma/remma_cpu.py:178), and
  (b) oracle/_ref/*.so -- the reference's own C (compiled from /root/reference by oracle/Makefile),
      which also travels to the GPU box.
Third-party arithmetic: ``pandas_plink`` (un-pinned in the reference's setup.py:21) does the .bed decode of
the numpy paths; it is absent here, its dtype is version dependent, and we pin float64 A2-counts = the
reference's own C reader (_read_plink_bed.c:36-37).
"""
import ctypes
import os
from functools import reduce

import numpy as np
from scipy import linalg
from scipy.stats import chi2

_HERE = os.path.dirname(os.path.abspath(__file__))
_PORT = os.path.join(_HERE, "_build", "libgmat_oracle.so")
_REF_EPI = os.path.join(_HERE, "_ref", "libgmat_ref_epi.so")
_LL, _D = ctypes.c_longlong, ctypes.c_double
_pd = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_pl = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")


def port_lib():
    if not os.path.exists(_PORT):
        raise RuntimeError("oracle C port not built: run `make -C oracle port`")
    lib = ctypes.CDLL(_PORT)
    lib.oracle_read_bed.argtypes = [ctypes.c_char_p, _LL, _LL, _pd]
    lib.oracle_read_bed.restype = ctypes.c_int
    lib.oracle_code.argtypes = [_pd, _LL, _LL, ctypes.c_int, ctypes.c_void_p]
    lib.oracle_code.restype = None
    lib.oracle_screen.argtypes = [_pd, _pd, _LL, _LL, _pl, _LL, _pd, _D, ctypes.c_void_p, ctypes.c_void_p,
                                  ctypes.c_void_p, ctypes.c_int, _pl, _pl, _pd, _LL]
    lib.oracle_screen.restype = _LL
    lib.oracle_exact_pairs.argtypes = [_pd, _pd, _LL, _pl, _pl, _LL, _pd, _pd, _pd, _pd]
    lib.oracle_exact_pairs.restype = None
    return lib


def count_lines(path):
    with open(path) as f:
        return sum(1 for line in f if line.strip())


# ---------------------------------------------------------------- decode + coding
def read_bed_raw(bed_prefix, n=None, m=None):
    """[m][n] float64, C map 0, 1/3, 1, 2 (gmat/process_plink/_read_plink_bed.c:19-46), pure numpy."""
    n = count_lines(bed_prefix + ".fam") if n is None else n
    m = count_lines(bed_prefix + ".bim") if m is None else m
    raw = np.fromfile(bed_prefix + ".bed", dtype=np.uint8)[3:]
    rows = raw[: m * ((n + 3) // 4)].reshape(m, (n + 3) // 4)
    codes = np.stack([(rows >> (2 * q)) & 3 for q in range(4)], axis=2).reshape(m, -1)[:, :n].astype(np.int64)
    return (codes * codes + codes) / 6.0


def read_plink(bed_prefix):
    """[n][m] float64, NaN for missing: the stand-in for pandas_plink (process_plink.py:7-9 + read_plink_bed.py:26)."""
    g = read_bed_raw(bed_prefix)
    g[np.abs(g - 1.0 / 3) < 0.0001] = np.nan
    return np.ascontiguousarray(g.T)


def code_additive(snp_mat):
    """remma_epiAA.py:59-61 (snp_mat is [n][m])."""
    n = snp_mat.shape[0]
    freq = (np.sum(snp_mat, axis=0) / (2 * n)).reshape(1, -1)
    return snp_mat - 2 * freq, freq


def code_dominance(snp_mat):
    """remma_epiAD.py:59-63 / remma_epiDD.py:59-65."""
    n = snp_mat.shape[0]
    freq = (np.sum(snp_mat, axis=0) / (2 * n)).reshape(1, -1)
    h = snp_mat.copy()
    h[h > 1.5] = 0.0
    return h - 2 * freq * (1 - freq), freq


def c_coded(bed_prefix, coding):
    """Coded [m][n] matrix exactly as the reference C scan builds it (p accumulated term by term)."""
    lib = port_lib()
    n, m = count_lines(bed_prefix + ".fam"), count_lines(bed_prefix + ".bim")
    x = np.empty((m, n), dtype=np.float64)
    assert lib.oracle_read_bed(bed_prefix.encode(), n, m, x) == 1
    freq = np.empty((m,), dtype=np.float64)
    lib.oracle_code(x, n, m, coding, freq.ctypes.data)
    return x, freq


# ---------------------------------------------------------------- GRM
def agmat(bed_prefix, small_val=0.001):
    """gmatrix.py:46-66 (no missing data assumed)."""
    snp_mat = read_plink(bed_prefix)
    x, freq = code_additive(snp_mat)
    scale = np.sum(2 * freq * (1 - freq))
    kin = np.dot(x, x.T) / scale
    d = np.diag(kin)
    np.fill_diagonal(kin, d + d * small_val)
    return kin, float(scale)


def dgmat_as(bed_prefix, small_val=0.001):
    """gmatrix.py:109-130."""
    snp_mat = read_plink(bed_prefix)
    n = snp_mat.shape[0]
    freq = (np.sum(snp_mat, axis=0) / (2 * n)).reshape(1, -1)
    scale_vec = 2 * freq * (1 - freq)
    scale = np.sum(scale_vec * (1 - scale_vec))
    snp_mat[snp_mat > 1.5] = 0.0
    x = snp_mat - scale_vec
    kin = np.dot(x, x.T) / scale
    d = np.diag(kin)
    np.fill_diagonal(kin, d + d * small_val)
    return kin, float(scale)


# ---------------------------------------------------------------- preamble + exact statistics
def preamble(y, xmat, zmat, gmat_lst, var_com, scipy_inv=True):
    """remma_epiAA.py:31-51."""
    inv = linalg.inv if scipy_inv else np.linalg.inv
    y = np.array(y).reshape(-1, 1)
    n = y.shape[0]
    xmat = np.array(xmat).reshape(n, -1)
    vmat = np.diag([var_com[-1]] * n)
    for val in range(len(gmat_lst)):
        vmat += zmat.dot((zmat.dot(gmat_lst[val])).T) * var_com[val]
    vmat_inv = inv(vmat)
    vxmat = np.dot(vmat_inv, xmat)
    xvxmat = inv(np.dot(xmat.T, vxmat))
    pmat = vmat_inv - reduce(np.dot, [vxmat, xvxmat, vxmat.T])
    pymat = zmat.T.dot(np.dot(pmat, y))
    pvpmat = zmat.T.dot((zmat.T.dot(pmat)).T)
    return pymat, pvpmat


def coded_pair(kind, snp_mat):
    """(U, V) coded [n][m] matrices for kind in {'AA','AD','DD'}."""
    if kind == "AA":
        a, _ = code_additive(snp_mat)
        return a, a
    d, _ = code_dominance(snp_mat)
    if kind == "DD":
        return d, d
    a, _ = code_additive(snp_mat)
    return a, d


def exact_first_snp(U, V, i, pymat, pvpmat, all_pairs):
    """One iteration of the reference hot loop (remma_epiAA.py:72-77 / remma_epiAD.py:77-82):
    returns (j indices, eff, var, chi, p)."""
    m = U.shape[1]
    if all_pairs:
        js = np.arange(m)
        epi_mat = U[:, i:(i + 1)] * V
    else:
        js = np.arange(i + 1, m)
        epi_mat = U[:, i:(i + 1)] * V[:, (i + 1):]
    eff_vec = np.dot(epi_mat.T, pymat)
    var_vec = np.sum(epi_mat * np.dot(pvpmat, epi_mat), axis=0).reshape(-1, 1)
    chi_vec = eff_vec * eff_vec / var_vec
    p_vec = chi2.sf(chi_vec, 1)
    return js, eff_vec[:, -1], var_vec[:, -1], chi_vec[:, -1], p_vec[:, -1]


def exact_pairs(U, V, pairs, pymat, pvpmat):
    """remma_epiAA_pair.py:78-84 for an explicit [k][2] pair array: (eff, var, chi, p)."""
    pairs = np.asarray(pairs)
    epi_mat = U[:, pairs[:, 0]] * V[:, pairs[:, 1]]
    eff_vec = np.dot(epi_mat.T, pymat)
    var_vec = np.sum(epi_mat * np.dot(pvpmat, epi_mat), axis=0).reshape(-1, 1)
    chi_vec = eff_vec * eff_vec / var_vec
    p_vec = chi2.sf(chi_vec, 1)
    return eff_vec[:, -1], var_vec[:, -1], chi_vec[:, -1], p_vec[:, -1]


def exact_scan(kind, snp_mat, snp_lst_0, pymat, pvpmat, p_cut):
    """Full exact scan, rows (i, j, eff, chi, p) with p < p_cut in reference order."""
    U, V = coded_pair(kind, snp_mat)
    rows = []
    for i in snp_lst_0:
        js, eff, var, chi, p = exact_first_snp(U, V, i, pymat, pvpmat, kind == "AD")
        for q in np.nonzero(p < p_cut)[0]:
            rows.append((int(i), int(js[q]), float(eff[q]), float(chi[q]), float(p[q])))
    return rows


# ---------------------------------------------------------------- approximate screen (C port)
def screen(bed_prefix, kind, snp_lst_0, pymat, eff_cut=0.0, table=None, bin_i=None, bin_j=None, cap=1 << 22):
    """The C screen (remma_epi{AA,AD,DD}[_maf]_eff_cpu) restated; returns sorted rows (snp_0, snp_1, eff)."""
    lib = port_lib()
    a = d = None
    if kind in ("AA", "AD"):
        a, _ = c_coded(bed_prefix, 1)
    if kind in ("DD", "AD"):
        d, _ = c_coded(bed_prefix, 2)
    m, n = (a if a is not None else d).shape
    il = np.ascontiguousarray(snp_lst_0, dtype=np.int64)
    py = np.ascontiguousarray(np.asarray(pymat, dtype=np.float64).reshape(-1))
    tab = None if table is None else np.ascontiguousarray(table, dtype=np.float64)
    bi = None if bin_i is None else np.ascontiguousarray(bin_i, dtype=np.int64)
    bj = None if bin_j is None else np.ascontiguousarray(bin_j, dtype=np.int64)

    def run(U, V, ge):
        hi, hj, he = np.empty(cap, np.int64), np.empty(cap, np.int64), np.empty(cap, np.float64)
        cnt = lib.oracle_screen(U, V, n, m, il, len(il), py, float(eff_cut),
                                None if tab is None else tab.ctypes.data,
                                None if bi is None else bi.ctypes.data,
                                None if bj is None else bj.ctypes.data, ge, hi, hj, he, cap)
        assert cnt <= cap
        return hi[:cnt], hj[:cnt], he[:cnt]

    rows = []
    if kind == "AA":
        passes = [(a, a, 0, False)]
    elif kind == "DD":
        passes = [(d, d, 0, False)]
    else:
        passes = [(a, d, 1, False), (d, a, 0, True)]
    for dirn, (U, V, ge, swap) in enumerate(passes):
        hi, hj, he = run(U, V, ge)
        for i, j, e in zip(hi.tolist(), hj.tolist(), he.tolist()):
            rows.append(((i, j, dirn), (j, i, e) if swap else (i, j, e)))
    rows.sort(key=lambda r: r[0])
    return [r[1] for r in rows]


def ref_screen_file(bed_prefix, kind, snp_lst_0, pymat, eff_cut, out_file, table=None, bin_i=None, bin_j=None):
    """Run the reference's OWN C (oracle/_ref/libgmat_ref_epi.so) and return its rows as a sorted list of
    (snp_0, snp_1, eff_text).  Works on the GPU box too (the .so travels)."""
    import cffi
    from .ref_harness import EPI_CDEF
    ffi = cffi.FFI()
    ffi.cdef(EPI_CDEF)
    lib = ffi.dlopen(_REF_EPI)
    n, m = count_lines(bed_prefix + ".fam"), count_lines(bed_prefix + ".bim")
    il = np.ascontiguousarray(snp_lst_0, dtype=np.int64)
    py = np.ascontiguousarray(np.asarray(pymat, dtype=np.float64).reshape(-1))
    args = [ffi.new("char[]", bed_prefix.encode()), n, m, ffi.cast("long long *", il.ctypes.data), len(il),
            ffi.cast("double *", py.ctypes.data)]
    keep = [il, py]
    if table is None:
        args.append(float(eff_cut))
        fn = getattr(lib, "remma_epi%s_eff_cpu" % kind)
    else:
        bi = np.ascontiguousarray(bin_i, dtype=np.int64)
        keep.append(bi)
        args.append(ffi.cast("long long *", bi.ctypes.data))
        if kind == "AD":
            bj = np.ascontiguousarray(bin_j, dtype=np.int64)
            keep.append(bj)
            args.append(ffi.cast("long long *", bj.ctypes.data))
        tab = np.ascontiguousarray(table, dtype=np.float64)
        keep.append(tab)
        args.append(ffi.cast("double *", tab.ctypes.data))
        fn = getattr(lib, "remma_epi%s_maf_eff_cpu" % kind)
    args.append(ffi.new("char[]", out_file.encode()))
    devnull = os.open(os.devnull, os.O_WRONLY)      # the reference prints a progress bar on stdout
    saved = os.dup(1)
    os.dup2(devnull, 1)
    try:
        rc = fn(*args)
    finally:
        os.dup2(saved, 1)
        os.close(devnull)
        os.close(saved)
    assert rc == 1
    rows = [l.split() for l in open(out_file).read().splitlines()[1:]]
    return sorted((int(r[0]), int(r[1]), r[2]) for r in rows)


def parallel_lists(num_snp, parallel, all_pairs=False):
    """remma_epiAA.py:125-137 / remma_epiAD.py:130-139."""
    s = int(num_snp / (2 * parallel[0]))
    p0, p1 = (parallel[1] - 1) * s, parallel[1] * s
    p2, p3 = (2 * parallel[0] - parallel[1]) * s, (2 * parallel[0] - parallel[1] + 1) * s
    if parallel[1] == 1:
        p3 = num_snp if all_pairs else num_snp - 1
    return list(range(p0, p1)) + list(range(p2, p3)), [p0, p1, p2, p3]


def random_pairs(num_snp, out_file, num_pair=100000, num_each_pair=5000, ordered=True):
    """gmat/remma/random_pair.py:21-34 (ordered: i < j, the AA / DD sampler) and :52-65 (not ordered: i != j, the AD
    sampler), step by step: draws from the global numpy RNG, a set of 'i j' strings filled through a temporary set
    per draw, truncation of the set's iteration order, np.savetxt, np.loadtxt of the file just written."""
    seen = set()
    while len(seen) < num_pair:
        draw = np.random.randint(num_snp, size=(num_each_pair, 2))
        draw = draw[draw[:, 0] < draw[:, 1]] if ordered else draw[draw[:, 0] != draw[:, 1]]
        text = np.array(draw, dtype=str)
        seen.update(set(map(lambda a, b: a + ' ' + b, text[:, 0], text[:, 1])))
    kept = list(seen)[:num_pair]
    np.savetxt(out_file, np.array(kept), fmt="%s", header='snp_0 snp_1', comments='')
    return np.loadtxt(out_file, dtype=int, skiprows=1)


# ---------------------------------------------------------------- REML (row f3), numpy restatement
def wemai_multi_gmat(y, xmat, zmat, gmat_lst, init=None, maxiter=200, cc_par=1.0e-8, cc_gra=1.0e-6):
    """gmat/uvlmm/uvlmm_varcom.py:8-104 restated: weighted EM + AI REML.  Pinned by tests/test_oracle.py against
    the reference's own golden vector (examples/remma/remma_cpu.py:178) and the fixtures of make_golden.py."""
    var_com = np.array([1.0] * (len(gmat_lst) + 1) if init is None else init[:], dtype=np.float64)
    y = np.array(y).reshape(-1, 1)
    n = y.shape[0]
    xmat = np.array(xmat).reshape(n, -1)
    zg = [zmat.dot((zmat.dot(g)).T) for g in gmat_lst]
    for _ in range(maxiter):
        vmat = np.diag([var_com[-1]] * n)
        for k in range(len(zg)):
            vmat += zg[k] * var_com[k]
        vmat = np.linalg.inv(vmat)
        vx = np.dot(vmat, xmat)
        xvx = np.linalg.inv(np.dot(xmat.T, vx))
        pmat = vmat - reduce(np.dot, [vx, xvx, vx.T])
        py = np.dot(pmat, y)
        fd, wv = [], []
        for k in range(len(zg)):
            fd.append(0.5 * np.sum(-np.trace(np.dot(pmat, zg[k])) + reduce(np.dot, [py.T, zg[k], py])))
            wv.append(np.dot(zg[k], py))
        fd.append(0.5 * np.sum(-np.trace(pmat) + np.dot(py.T, py)))
        fd = np.array(fd)
        wv.append(py)
        wv = np.concatenate(wv, axis=1)
        ai = 0.5 * reduce(np.dot, [wv.T, pmat, wv])
        em = np.diag([n / (v * v) for v in var_com])
        for j in range(0, 101):
            w = j * 0.01
            delta = np.dot(np.linalg.inv((1 - w) * ai + w * em), fd)
            new = var_com + delta
            if min(new) > 0:
                break
        cc_par_val = np.sqrt(np.sum(delta * delta) / np.sum(new * new))
        var_com = new[:]
        if np.sqrt(np.sum(fd * fd)) < cc_gra and cc_par_val < cc_par:
            break
    return var_com
