"""TEST INFRASTRUCTURE ONLY -- generate tests/golden/* by running the UNMODIFIED reference here.

Run once in the build container (needs /root/reference and `make -C oracle ref`):

    python oracle/make_golden.py            # ~5 minutes on 8 cores

Everything it writes is small text/JSON and is committed, so the GPU box (which has no
/root/reference) checks the CUDA path against outputs of the reference itself.
What is produced, and the reference entry point each file pins:

  mouse/plink.{bed,bim,fam}, mouse/pheno   copy of examples/data/mouse (input data, config C1)
  mouse/pheno_rep3                         3 records per individual (repeated measures, config C5 shape)
  mouse_golden.json                        agmat/dgmat_as scale + matrix samples (gmatrix.py:34,97),
                                           var_com (uvlmm_varcom.py:107; == examples/remma/remma_cpu.py:178),
                                           pymat/pvpmat samples (remma_epiAA.py:31-51), decode sample, partitions
  out/epiAA.txt epiDD.txt epiAD.txt        exact scans (remma_epiAA.py:89, remma_epiDD.py:93, remma_epiAD.py:94)
  out/epiXX_eff.txt                        C screen through remma_epiXX_eff (remma_epiAA_eff.py:99)
  out/epiXX_pair.txt + pairs_XX.txt        pair-list exact test, p_cut=1 (remma_epiAA_pair.py:95)
  out/epiXX_approx.txt + .random_pair      3-stage pipeline (remma_epiAA_approx.py:10)
  out/epiXX_maf_approx.txt (+ aux)         MAF-binned pipeline (remma_epiAA_maf_approx.py:11)
  out/epiAA_rep3.txt                       repeated-measures exact scan on a first-SNP subset
  out/epiAA_parallel.{1,2}                 _parallel parts (remma_epiAA.py:143)
  out/epiAD.anno_* epiAD.ld_table          annotation_snp_pos on out/epiAD.txt (annotation.py:22); made by the
                                           snippet in this file's annotation_golden()
"""
import json
import logging
import os
import shutil
import sys
import tempfile
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import ref_harness  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
SRC = os.path.join(ref_harness.REF_ROOT, "examples", "data", "mouse")


def flt(a):
    return [float(x) for x in np.asarray(a).ravel()]


def annotation_golden():
    """annotation_snp_pos fixtures from the unmodified reference (run after main(); needs out/epiAD.txt)."""
    ref_harness.install()
    from gmat.remma.annotation import annotation_snp_pos as ref_anno
    work = tempfile.mkdtemp()
    shutil.copyfile(os.path.join(GOLD, "mouse", "plink.bim"), os.path.join(work, "plink.bim"))
    res = os.path.join(work, "epiAD.txt")
    shutil.copyfile(os.path.join(GOLD, "out", "epiAD.txt"), res)
    rows = [line.split() for line in open(res).read().splitlines()[1:]]
    bim = [line.split() for line in open(os.path.join(work, "plink.bim"))]
    with open(os.path.join(work, "ld.txt"), "w") as f:       # plink --r2 style table: 5 linked, 5 weak pairs
        f.write(" CHR_A BP_A SNP_A CHR_B BP_B SNP_B R2\n")
        for k, r in enumerate(rows[:10]):
            a, b = bim[int(r[0])], bim[int(r[1])]
            f.write(" %s %s %s %s %s %s %s\n" % (a[0], a[3], a[1], b[0], b[3], b[1], "0.9" if k < 5 else "0.1"))
    ref_anno(res, os.path.join(work, "plink"), p_cut=1.0e-5, dis=0)
    shutil.copyfile(res + ".anno", os.path.join(GOLD, "out", "epiAD.anno_p1e-5"))
    ref_anno(res, os.path.join(work, "plink"), p_cut=1, dis=5.0e7, ld_file=os.path.join(work, "ld.txt"), r2=0.2)
    shutil.copyfile(res + ".anno", os.path.join(GOLD, "out", "epiAD.anno_dis5e7"))
    shutil.copyfile(res + ".anno.ld", os.path.join(GOLD, "out", "epiAD.anno_dis5e7.ld"))
    shutil.copyfile(os.path.join(work, "ld.txt"), os.path.join(GOLD, "out", "epiAD.ld_table"))


def main():
    logging.basicConfig(level=logging.WARNING)
    R = ref_harness.ref()
    os.makedirs(os.path.join(GOLD, "mouse"), exist_ok=True)
    os.makedirs(os.path.join(GOLD, "out"), exist_ok=True)
    for f in ("plink.bed", "plink.bim", "plink.fam", "pheno"):
        shutil.copyfile(os.path.join(SRC, f), os.path.join(GOLD, "mouse", f))
        os.chmod(os.path.join(GOLD, "mouse", f), 0o644)

    work = tempfile.mkdtemp(prefix="gmat_golden_")
    for f in ("plink.bed", "plink.bim", "plink.fam", "pheno"):
        shutil.copyfile(os.path.join(SRC, f), os.path.join(work, f))
    os.chdir(work)                      # the maf code writes 'eff_cut' into the cwd
    bed = os.path.join(work, "plink")
    pheno = os.path.join(work, "pheno")
    out = lambda name: os.path.join(GOLD, "out", name)  # noqa: E731
    G = {}
    t0 = time.time()

    # ---- repeated-measures phenotype: 3 records per individual, .fam order ------------------
    rng = np.random.Generator(np.random.PCG64(20201013))
    lines = [l.split() for l in open(pheno)]
    with open(os.path.join(GOLD, "mouse", "pheno_rep3"), "w") as f:
        for arr in lines:
            base = float(arr[-1])
            for r in range(3):
                val = base + 0.05 * rng.standard_normal()
                f.write(" ".join(arr[:-1] + ["%.6f" % val]) + "\n")
    pheno_rep = os.path.join(GOLD, "mouse", "pheno_rep3")

    # ---- decode + GRMs -------------------------------------------------------------------
    snp = R.read_plink(bed)
    n, m = snp.shape
    G["n"], G["m"] = int(n), int(m)
    G["any_missing"] = bool(np.any(np.isnan(snp)))
    G["geno_colsum"] = [int(x) for x in snp.sum(axis=0)]
    G["geno_rowsum_first16"] = [int(x) for x in snp.sum(axis=1)[:16]]
    G["geno_sample_rows0_8_cols0_16"] = [[int(x) for x in r[:16]] for r in snp[:8]]
    ag, _ = R.agmat(bed)
    dg, _ = R.dgmat_as(bed)
    freq = snp.sum(axis=0) / (2 * n)
    G["agmat_scale"] = float(np.sum(2 * freq * (1 - freq)))
    sv = 2 * freq * (1 - freq)
    G["dgmat_scale"] = float(np.sum(sv * (1 - sv)))
    ri, ci = list(range(0, n, 97)), list(range(0, n, 89))
    G["grm_sample_rows"], G["grm_sample_cols"] = ri, ci
    G["agmat_sample"] = [flt(ag[r, ci]) for r in ri]
    G["dgmat_sample"] = [flt(dg[r, ci]) for r in ri]
    G["agmat_diag_first16"], G["dgmat_diag_first16"] = flt(np.diag(ag)[:16]), flt(np.diag(dg)[:16])
    G["agmat_sum"], G["dgmat_sum"] = float(ag.sum()), float(dg.sum())
    G["agmat_fro"], G["dgmat_fro"] = float(np.linalg.norm(ag)), float(np.linalg.norm(dg))
    G["agmat_absmax"], G["dgmat_absmax"] = float(np.abs(ag).max()), float(np.abs(dg).max())
    # agrm0 text: first two lines verbatim pin the savetxt format (gmatrix.py:12)
    with open(bed + ".agrm0") as f:
        first = f.readline()
    G["agrm0_first_line_head"] = first[:200]
    print("GRMs done", time.time() - t0, flush=True)

    # ---- variance components (golden vector examples/remma/remma_cpu.py:178) ------------------
    vc_aa = R.wemai_multi_gmat(pheno, bed, [ag, ag * ag], out_file=os.path.join(work, "vc1"))
    vc_ad = R.wemai_multi_gmat(pheno, bed, [ag, dg, ag * ag], out_file=os.path.join(work, "vc2"))
    G["var_com_ag_agag"], G["var_com_ag_dg_agag"] = flt(vc_aa), flt(vc_ad)
    print("var_com", vc_aa, vc_ad, time.time() - t0, flush=True)
    gl_aa, gl_ad = [ag, ag * ag], [ag, dg, ag * ag]

    # ---- preamble samples (remma_epiAA.py:31-51) -------------------------------------------
    from functools import reduce
    from scipy import linalg
    y, xmat, zmat = R.design_matrix_wemai_multi_gmat(pheno, bed)
    N = y.shape[0]
    vmat = np.diag([vc_aa[-1]] * N)
    for k in range(2):
        vmat += zmat.dot((zmat.dot(gl_aa[k])).T) * vc_aa[k]
    vinv = linalg.inv(vmat)
    vx = np.dot(vinv, xmat)
    xvx = linalg.inv(np.dot(xmat.T, vx))
    pmat = vinv - reduce(np.dot, [vx, xvx, vx.T])
    pymat = zmat.T.dot(np.dot(pmat, y))
    pvpmat = zmat.T.dot((zmat.T.dot(pmat)).T)
    G["pymat_ag_agag"] = flt(pymat)
    G["pvpmat_sample_ag_agag"] = [flt(pvpmat[r, ci]) for r in ri]
    G["pvpmat_diag_first16"] = flt(np.diag(pvpmat)[:16])
    G["xmat_shape"], G["y_first8"] = list(xmat.shape), flt(y[:8])

    # ---- exact scans ---------------------------------------------------------------------
    R.AA.remma_epiAA(pheno, bed, gl_aa, vc_aa, p_cut=1e-5, out_file=out("epiAA.txt"))
    print("epiAA", time.time() - t0, flush=True)
    R.DD.remma_epiDD(pheno, bed, gl_ad, vc_ad, p_cut=1e-5, out_file=out("epiDD.txt"))
    print("epiDD", time.time() - t0, flush=True)
    R.AD.remma_epiAD(pheno, bed, gl_ad, vc_ad, p_cut=1e-5, out_file=out("epiAD.txt"))
    print("epiAD", time.time() - t0, flush=True)
    # looser threshold on a first-SNP subset: hundreds of rows, straddles the cut
    sub = list(range(0, m - 1, 29))
    G["snp_lst_0_subset"] = sub
    R.AA.remma_epiAA(pheno, bed, gl_aa, vc_aa, snp_lst_0=sub, p_cut=1e-3, out_file=out("epiAA_sub_p1e-3.txt"))
    R.DD.remma_epiDD(pheno, bed, gl_ad, vc_ad, snp_lst_0=sub, p_cut=1e-3, out_file=out("epiDD_sub_p1e-3.txt"))
    R.AD.remma_epiAD(pheno, bed, gl_ad, vc_ad, snp_lst_0=sub, p_cut=1e-3, out_file=out("epiAD_sub_p1e-3.txt"))
    # _parallel parts (remma_epiAA.py:143) with a loose cut so files are non-trivial
    for k in (1, 2):
        R.AA.remma_epiAA_parallel(pheno, bed, gl_aa, vc_aa, parallel=[24, k], p_cut=1e-3,
                                  out_file=out("epiAA_parallel"))
        R.AD.remma_epiAD_parallel(pheno, bed, gl_ad, vc_ad, parallel=[24, k], p_cut=1e-3,
                                  out_file=out("epiAD_parallel"))
    print("subset/parallel", time.time() - t0, flush=True)
    # repeated measures
    vc_rep = [0.05, 0.06, 0.02, 0.01]   # [ag, ag*ag, I(permanent env), residual]; fixed, not REML
    G["var_com_rep3"] = vc_rep
    R.AA.remma_epiAA(pheno_rep, bed, [ag, ag * ag, np.eye(n)], vc_rep, snp_lst_0=sub, p_cut=1e-3,
                     out_file=out("epiAA_rep3.txt"))
    print("rep3", time.time() - t0, flush=True)

    # ---- pair lists ----------------------------------------------------------------------
    a = rng.integers(0, m, size=(6000, 2))
    pa = a[a[:, 0] < a[:, 1]][:2500]
    np.savetxt(os.path.join(GOLD, "pairs_AA.txt"), pa, fmt="%d", header="snp_0 snp_1", comments="")
    pd_ = np.vstack([a[a[:, 0] != a[:, 1]][:2500], np.stack([np.arange(0, m, 97)] * 2, axis=1)])
    np.savetxt(os.path.join(GOLD, "pairs_AD.txt"), pd_, fmt="%d", header="snp_0 snp_1", comments="")
    R.AA.remma_epiAA_pair(pheno, bed, gl_aa, vc_aa, os.path.join(GOLD, "pairs_AA.txt"), p_cut=1,
                          out_file=out("epiAA_pair.txt"))
    R.DD.remma_epiDD_pair(pheno, bed, gl_ad, vc_ad, os.path.join(GOLD, "pairs_AA.txt"), p_cut=1,
                          out_file=out("epiDD_pair.txt"))
    R.AD.remma_epiAD_pair(pheno, bed, gl_ad, vc_ad, os.path.join(GOLD, "pairs_AD.txt"), p_cut=1,
                          out_file=out("epiAD_pair.txt"))
    # small chunk size exercises the chunk loop (remma_epiAA_pair.py:66-71)
    R.AA.remma_epiAA_pair(pheno, bed, gl_aa, vc_aa, os.path.join(GOLD, "pairs_AA.txt"), max_test_pair=700,
                          p_cut=0.05, out_file=out("epiAA_pair_chunk700_p05.txt"))
    import pandas as pd
    med = {}
    for kind in ("AA", "DD", "AD"):
        df = pd.read_csv(out("epi%s_pair.txt" % kind), sep=r"\s+")
        med[kind] = float(np.median(df["var"]))
    G["pair_var_median"] = med
    print("pairs", med, time.time() - t0, flush=True)

    # ---- C screens through the reference Python wrappers -------------------------------------
    R.AA.remma_epiAA_eff(pheno, bed, gl_aa, vc_aa, var_app=1575.0, p_cut=1e-5, out_file=out("epiAA_eff_p1e-5.txt"))
    R.AA.remma_epiAA_eff(pheno, bed, gl_aa, vc_aa, var_app=med["AA"], p_cut=1e-3, out_file=out("epiAA_eff.txt"))
    R.DD.remma_epiDD_eff(pheno, bed, gl_ad, vc_ad, var_app=med["DD"], p_cut=1e-3, out_file=out("epiDD_eff.txt"))
    R.AD.remma_epiAD_eff(pheno, bed, gl_ad, vc_ad, var_app=med["AD"], p_cut=1e-3, out_file=out("epiAD_eff.txt"))
    R.AA.remma_epiAA_eff(pheno, bed, gl_aa, vc_aa, snp_lst_0=sub, var_app=med["AA"], p_cut=1e-2,
                         out_file=out("epiAA_eff_sub.txt"))
    R.AA.remma_epiAA_eff_parallel(pheno, bed, gl_aa, vc_aa, parallel=[24, 2], var_app=med["AA"], p_cut=1e-3,
                                  out_file=out("epiAA_eff_parallel"))
    print("eff", time.time() - t0, flush=True)

    # ---- pipelines (random_pair output kept so the drop-in can be fed the same pairs) ---------
    def keep_pairs(modname, fn_name, dest):
        mod = sys.modules[modname]
        orig = getattr(mod, fn_name)

        def wrapped(num_snp, out_file="random_pair", num_pair=100000, num_each_pair=5000):
            res = orig(num_snp, out_file=out_file, num_pair=num_pair, num_each_pair=num_each_pair)
            shutil.copyfile(out_file, dest)
            return res
        setattr(mod, fn_name, wrapped)
        return lambda: setattr(mod, fn_name, orig)

    jobs = [
        ("AA", "remma_epiAA_approx", "random_pair", gl_aa, vc_aa),
        ("DD", "remma_epiDD_approx", "random_pair", gl_ad, vc_ad),
        ("AD", "remma_epiAD_approx", "random_pairAD", gl_ad, vc_ad),
        ("AA", "remma_epiAA_maf_approx", "random_pair", gl_aa, vc_aa),
        ("DD", "remma_epiDD_maf_approx", "random_pair", gl_ad, vc_ad),
        ("AD", "remma_epiAD_maf_approx", "random_pairAD", gl_ad, vc_ad),
    ]
    for kind, fn, rp, gl, vc in jobs:
        np.random.seed(20201013)
        modname = "gmat.remma.remma_epi%s.%s" % (kind, fn)
        stem = fn.replace("remma_", "")
        undo = keep_pairs(modname, rp, out(stem + ".random_pair"))
        tmp_out = os.path.join(work, stem)
        getattr(sys.modules[modname], fn)(pheno, bed, gl, vc, p_cut=1e-4, num_random_pair=20000, out_file=tmp_out)
        undo()
        shutil.copyfile(tmp_out, out(stem + ".txt"))
        for aux in (".freq_denominator", ".freq", ".maf", ".heter"):
            if os.path.exists(tmp_out + aux):
                shutil.copyfile(tmp_out + aux, out(stem + aux))
        print(fn, time.time() - t0, flush=True)

    # ---- _parallel partition arithmetic (remma_epiAA.py:125-137, remma_epiAD.py:130-139) -------
    parts = []
    for (mm, P) in ((1407, 3), (1407, 24), (50000, 8), (20000, 7), (101, 5)):
        for k in range(1, P + 1):
            s = int(mm / (2 * P))
            p0, p1, p2, p3 = (k - 1) * s, k * s, (2 * P - k) * s, (2 * P - k + 1) * s
            parts.append({"m": mm, "P": P, "k": k, "AA": [p0, p1, p2, (mm - 1) if k == 1 else p3],
                          "AD": [p0, p1, p2, mm if k == 1 else p3]})
    G["parallel_partitions"] = parts

    with open(os.path.join(GOLD, "mouse_golden.json"), "w") as f:
        json.dump(G, f, indent=0)
    shutil.rmtree(work, ignore_errors=True)
    print("done", time.time() - t0)


if __name__ == "__main__":
    main()
    annotation_golden()
