"""TEST INFRASTRUCTURE ONLY -- imports the UNMODIFIED reference (GMAT) from /root/reference.

Only usable in the build container (``/root/reference`` does not exist on the GPU box).  It is
used by ``oracle/make_golden.py`` to produce the fixtures under ``tests/golden/`` and by the
``not gpu`` tests (when the tree is present) to validate the restatement in ``oracle/oracle.py``
and ``oracle/gmat_oracle.c`` against the real thing.

Three shims are needed to import the 2020 code on this image:

1. ``pandas_plink`` is not installed.  ``gmat/process_plink/process_plink.py:7-9`` only uses
   ``read_plink1_bin(bed, bim, fam, verbose=False).values`` -> float64 ``[n_id][n_snp]``.  The
   stand-in decodes with the reference's OWN C reader (``oracle/_ref/libgmat_ref_bed.so`` built
   from ``gmat/process_plink/_read_plink_bed.c``) and applies the NaN mapping of
   ``gmat/process_plink/read_plink_bed.py:26-28``.  dtype pinned to float64.
2. ``np.int`` was removed from numpy (used at ``remma_epiAA_pair.py:75``, ``random_pair.py:33``).
3. the cffi API-mode extension modules ``_cremma_epi_eff_cpu`` / ``_cread_plink_bed`` are replaced
   by ABI-mode ``ffi.dlopen`` of the same C compiled by ``oracle/Makefile`` (same cdefs as
   ``gmat/remma/_build.py:8-31``).
"""
import os
import sys
import types
import warnings

import numpy as np

REF_ROOT = os.environ.get("GMAT_REFERENCE_ROOT", "/root/reference")
_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_SO_DIR = os.path.join(_HERE, "_ref")

EPI_CDEF = """
int remma_epiAA_maf_eff_cpu(char *bed_file, long long num_id, long long num_snp, long long *snp_lst_0,
    long long len_snp_lst_0, double *pymat, long long *freq, double *eff_cut, char* out_file);
int remma_epiAD_maf_eff_cpu(char *bed_file, long long num_id, long long num_snp, long long *snp_lst_0,
    long long len_snp_lst_0, double *pymat, long long *freqA, long long *freqD, double *eff_cut, char* out_file);
int remma_epiDD_maf_eff_cpu(char *bed_file, long long num_id, long long num_snp, long long *snp_lst_0,
    long long len_snp_lst_0, double *pymat, long long *freq, double *eff_cut, char* out_file);
int remma_epiAA_eff_cpu(char *bed_file, long long num_id, long long num_snp, long long *snp_lst_0,
    long long len_snp_lst_0, double *pymat, double eff_cut, char* out_file);
int remma_epiAD_eff_cpu(char *bed_file, long long num_id, long long num_snp,
    long long *snp_lst_0, long long len_snp_lst_0, double *pymat, double eff_cut, char* out_file);
int remma_epiDD_eff_cpu(char *bed_file, long long num_id, long long num_snp,
    long long *snp_lst_0, long long len_snp_lst_0, double *pymat, double eff_cut, char* out_file);
"""
BED_CDEF = "int read_plink_bed(char *bed_file, long long num_id, long long num_snp, double *marker_mat);"


def reference_available():
    return os.path.isdir(os.path.join(REF_ROOT, "gmat")) and \
        os.path.exists(os.path.join(_REF_SO_DIR, "libgmat_ref_epi.so"))


def ref_so_available():
    return os.path.exists(os.path.join(_REF_SO_DIR, "libgmat_ref_epi.so"))


def load_ref_epi():
    """cffi (ffi, lib) pair for the reference's scan C code compiled by oracle/Makefile."""
    import cffi
    ffi = cffi.FFI()
    ffi.cdef(EPI_CDEF)
    lib = ffi.dlopen(os.path.join(_REF_SO_DIR, "libgmat_ref_epi.so"))
    return ffi, lib


def load_ref_bed():
    import cffi
    ffi = cffi.FFI()
    ffi.cdef(BED_CDEF)
    lib = ffi.dlopen(os.path.join(_REF_SO_DIR, "libgmat_ref_bed.so"))
    return ffi, lib


def ref_read_bed(bed_prefix, num_id, num_snp):
    """The reference C reader, raw output: double[num_snp][num_id], code 01 -> 1/3."""
    ffi, lib = load_ref_bed()
    out = np.ones(num_id * num_snp, dtype=np.float64)
    lib.read_plink_bed(ffi.new("char[]", bed_prefix.encode("ascii")), num_id, num_snp,
                       ffi.cast("double *", ffi.from_buffer(out)))
    return out.reshape(num_snp, num_id)


def _count_lines(path):
    with open(path) as f:
        return sum(1 for _ in f)


class _Bin(object):
    def __init__(self, values):
        self.values = values


def _read_plink1_bin(bed, bim, fam, verbose=False):
    prefix = bed[:-4]
    n = _count_lines(fam)
    m = _count_lines(bim)
    mat = ref_read_bed(prefix, n, m)
    mat[np.abs(mat - 1.0 / 3) < 0.0001] = np.nan      # read_plink_bed.py:26
    return _Bin(np.ascontiguousarray(mat.T))           # [n_id][n_snp] float64


_installed = False


def install():
    """Install the shims and make ``import gmat`` resolve to the reference tree."""
    global _installed
    if _installed:
        return
    if not reference_available():
        raise RuntimeError("reference tree or oracle/_ref missing (run `make -C oracle ref` in the build container)")
    warnings.filterwarnings("ignore", category=SyntaxWarning)
    if not hasattr(np, "int"):
        np.int = int
    pp = types.ModuleType("pandas_plink")
    pp.read_plink1_bin = _read_plink1_bin
    sys.modules["pandas_plink"] = pp
    ffi, lib = load_ref_epi()
    m1 = types.ModuleType("_cremma_epi_eff_cpu")
    m1.ffi, m1.lib = ffi, lib
    sys.modules["_cremma_epi_eff_cpu"] = m1
    ffi2, lib2 = load_ref_bed()
    m2 = types.ModuleType("_cread_plink_bed")
    m2.ffi, m2.lib = ffi2, lib2
    sys.modules["_cread_plink_bed"] = m2
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    _installed = True


def ref():
    """Namespace with the reference functions on the hot path (imported lazily, unmodified)."""
    install()
    ns = types.SimpleNamespace()
    from gmat.gmatrix.gmatrix import agmat, dgmat_as
    from gmat.uvlmm.uvlmm_varcom import wemai_multi_gmat
    from gmat.uvlmm.design_matrix import design_matrix_wemai_multi_gmat
    from gmat.process_plink.process_plink import read_plink, impute_geno
    import gmat.remma.remma_epiAA as AA
    import gmat.remma.remma_epiAD as AD
    import gmat.remma.remma_epiDD as DD
    from gmat.remma.random_pair import random_pair, random_pairAD
    ns.agmat, ns.dgmat_as = agmat, dgmat_as
    ns.wemai_multi_gmat = wemai_multi_gmat
    ns.design_matrix_wemai_multi_gmat = design_matrix_wemai_multi_gmat
    ns.read_plink, ns.impute_geno = read_plink, impute_geno
    ns.AA, ns.AD, ns.DD = AA, AD, DD
    ns.random_pair, ns.random_pairAD = random_pair, random_pairAD
    return ns
