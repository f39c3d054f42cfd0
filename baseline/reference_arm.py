"""The reference arm of bench.py: the UNMODIFIED reference (yanjunzan/GMAT, pip-installed into baseline/_ref by
baseline/install_ref.sh) run through its own public functions on the host cores.  Nothing here imports gmat_b200.

Two things the 2020 code needs on this image (both outside the reference's files):
  * ``pandas_plink`` (install_requires, absent offline): gmat/process_plink/process_plink.py:7-9 only uses
    ``read_plink1_bin(bed, bim, fam, verbose=False).values`` -> float [n_id][n_snp] with NaN for missing.  The stand-in
    returns exactly what the reference's OWN reader returns (gmat.process_plink.read_plink_bed.Bed.read(): its C
    reader ``_cread_plink_bed`` + the NaN mapping of read_plink_bed.py:26-28).
  * ``np.int`` (removed from numpy; used by remma_epiAA_pair.py:75 and random_pair.py:33).
"""
import logging
import os
import re
import sys
import time
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def available():
    return os.path.isfile(os.path.join(REF_DIR, "gmat", "remma", "remma_epiAA", "remma_epiAA.py"))


_loaded = None


def load():
    """Import the installed reference; returns a namespace of the functions the arm calls."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError("baseline/_ref is missing: run baseline/install_ref.sh in the build container")
    import warnings
    warnings.filterwarnings("ignore", category=SyntaxWarning)
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    if not hasattr(np, "int"):
        np.int = int
    if "pandas_plink" not in sys.modules:
        pp = types.ModuleType("pandas_plink")

        class _Values(object):
            def __init__(self, values):
                self.values = values

        def read_plink1_bin(bed, bim, fam, verbose=False):
            from gmat.process_plink.read_plink_bed import Bed      # the reference's own C-backed reader
            return _Values(np.ascontiguousarray(Bed(bed[:-4]).read()))

        pp.read_plink1_bin = read_plink1_bin
        sys.modules["pandas_plink"] = pp
    ns = types.SimpleNamespace()
    from gmat.gmatrix.gmatrix import agmat
    from gmat.uvlmm.design_matrix import design_matrix_wemai_multi_gmat
    from gmat.remma.remma_epiAA.remma_epiAA import _remma_epiAA, remma_epiAA
    ns.agmat, ns.design_matrix, ns._remma_epiAA, ns.remma_epiAA = agmat, design_matrix_wemai_multi_gmat, _remma_epiAA, remma_epiAA
    _loaded = ns
    return ns


class _LoopClock(logging.Handler):
    """Catches the reference's own 'Running time: Clock time, X sec' record (remma_epiAA.py:83-85): the wall time
    of its scan loop, measured by the reference itself."""

    def __init__(self):
        super().__init__(level=logging.INFO)
        self.seconds = None

    def emit(self, record):
        m = re.search(r"Running time: Clock time, ([0-9.eE+-]+) sec", record.getMessage())
        if m:
            self.seconds = float(m.group(1))


class ExactArm(object):
    """remma_epiAA exact on the host cores.  Setup (outside the timed region, like the GPU arm's): the GRM through the
    reference's agmat(bed_file) and its design matrices.  sample(first_snps): ONE call of the reference's
    _remma_epiAA(y, xmat, zmat, gmat_lst, var_com, bed_file, snp_lst_0=first_snps, p_cut, out_file)
    (remma_epiAA.py:16-86: V, V^-1, P, Z'Py, Z'PZ, .bed read, then the per-first-SNP loop with its pandas filter /
    append).  Returns (pairs, loop seconds by the reference's own clock, whole-call seconds, rows written)."""

    def __init__(self, prefix, pheno, var_com, p_cut):
        self.ref = load()
        self.prefix, self.pheno, self.var_com, self.p_cut = prefix, pheno, var_com, p_cut
        ag = self.ref.agmat(prefix)[0]                       # writes prefix.agrm0 like every user's run
        self.gmat_lst = [ag, ag * ag]
        self.y, self.xmat, self.zmat = self.ref.design_matrix(pheno, prefix)
        self.m = sum(1 for _ in open(prefix + ".bim"))

    def sample(self, first_snps):
        out_file = self.prefix + ".ref_epiAA"
        clock = _LoopClock()
        root = logging.getLogger()
        old_level = root.level
        root.addHandler(clock)
        root.setLevel(logging.INFO)
        t0 = time.perf_counter()
        try:
            self.ref._remma_epiAA(self.y, self.xmat, self.zmat, self.gmat_lst, self.var_com, self.prefix,
                                  snp_lst_0=list(first_snps), p_cut=self.p_cut, out_file=out_file)
        finally:
            root.removeHandler(clock)
            root.setLevel(old_level)
        whole = time.perf_counter() - t0
        pairs = int(sum(self.m - 1 - i for i in first_snps))
        rows = sum(1 for _ in open(out_file)) - 1
        return pairs, (clock.seconds if clock.seconds else whole), whole, rows
