"""CPU tests: the oracle (numpy + C restatement) against the reference's own outputs (tests/golden, made by
oracle/make_golden.py from the unmodified reference) and against the reference C itself (oracle/_ref)."""
import os

import numpy as np
import pytest

from conftest import GOLD, gold_path, read_rows
from oracle import oracle, ref_harness


def test_decode_matches_golden_and_reference_c(golden, mouse_ro):
    bed, _ = mouse_ro
    g = oracle.read_bed_raw(bed)                       # [m][n]
    assert g.shape == (golden["m"], golden["n"])
    assert [int(x) for x in g.sum(axis=1)] == golden["geno_colsum"]
    assert [[int(x) for x in r[:16]] for r in g.T[:8]] == golden["geno_sample_rows0_8_cols0_16"]
    # C port == numpy restatement, bit for bit
    lib = oracle.port_lib()
    x = np.empty_like(g)
    assert lib.oracle_read_bed(bed.encode(), golden["n"], golden["m"], x) == 1
    assert np.array_equal(x, g)
    if ref_harness.ref_so_available():                 # the reference's own reader
        assert np.array_equal(ref_harness.ref_read_bed(bed, golden["n"], golden["m"]), g)


def test_grm_matches_reference(golden, mouse_model):
    ag, dg = mouse_model["ag"], mouse_model["dg"]
    ri, ci = golden["grm_sample_rows"], golden["grm_sample_cols"]
    np.testing.assert_allclose(ag[np.ix_(ri, ci)], np.array(golden["agmat_sample"]), rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(dg[np.ix_(ri, ci)], np.array(golden["dgmat_sample"]), rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(ag.sum(), golden["agmat_sum"], rtol=1e-9)
    np.testing.assert_allclose(np.linalg.norm(dg), golden["dgmat_fro"], rtol=1e-12)


def test_preamble_matches_reference(golden, mouse_model):
    M = mouse_model
    pymat, pvpmat = oracle.preamble(M["y"], M["xmat"], M["zmat"], M["gl_aa"], M["vc_aa"])
    np.testing.assert_allclose(pymat.ravel(), np.array(golden["pymat_ag_agag"]), rtol=1e-9, atol=1e-12)
    ri, ci = golden["grm_sample_rows"], golden["grm_sample_cols"]
    np.testing.assert_allclose(pvpmat[np.ix_(ri, ci)], np.array(golden["pvpmat_sample_ag_agag"]), rtol=1e-9,
                               atol=1e-13)
    # invariant: X has a column of ones => Z'Py sums to 0, rows of Z'PZ sum to 0
    assert abs(pymat.sum()) < 1e-8
    assert np.abs(pvpmat.sum(axis=1)).max() < 1e-8


@pytest.mark.parametrize("kind,model", [("AA", "aa"), ("DD", "ad"), ("AD", "ad")])
def test_exact_pairs_match_reference(kind, model, mouse_model, mouse_ro):
    M = mouse_model
    pymat, pvpmat = oracle.preamble(M["y"], M["xmat"], M["zmat"], M["gl_" + model], M["vc_" + model],
                                    scipy_inv=False)
    U, V = oracle.coded_pair(kind, oracle.read_plink(mouse_ro[0]))
    rows = read_rows(gold_path("epi%s_pair.txt" % kind))
    pairs = np.array([[int(r[0]), int(r[1])] for r in rows])
    eff, var, chi, p = oracle.exact_pairs(U, V, pairs, pymat, pvpmat)
    np.testing.assert_allclose(eff, [float(r[2]) for r in rows], rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(var, [float(r[3]) for r in rows], rtol=1e-9)
    np.testing.assert_allclose(p, [float(r[5]) for r in rows], rtol=1e-8)
    # C port of the same statistics
    lib = oracle.port_lib()
    Ut, Vt = np.ascontiguousarray(U.T), np.ascontiguousarray(V.T)
    e2, v2 = np.empty(200), np.empty(200)
    lib.oracle_exact_pairs(Ut, Vt, U.shape[0], np.ascontiguousarray(pairs[:200, 0]),
                           np.ascontiguousarray(pairs[:200, 1]), 200, np.ascontiguousarray(pymat.ravel()),
                           np.ascontiguousarray(pvpmat), e2, v2)
    np.testing.assert_allclose(e2, eff[:200], rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(v2, var[:200], rtol=1e-9)


def test_exact_scan_subset_matches_reference(golden, mouse_model, mouse_ro):
    M = mouse_model
    pymat, pvpmat = oracle.preamble(M["y"], M["xmat"], M["zmat"], M["gl_aa"], M["vc_aa"])
    sub = golden["snp_lst_0_subset"][:12]
    rows = oracle.exact_scan("AA", oracle.read_plink(mouse_ro[0]), sub, pymat, pvpmat, 1e-3)
    gold = [r for r in read_rows(gold_path("epiAA_sub_p1e-3.txt")) if int(r[0]) in sub]
    assert [(r[0], r[1]) for r in rows] == [(int(r[0]), int(r[1])) for r in gold]
    np.testing.assert_allclose([r[2] for r in rows], [float(r[2]) for r in gold], rtol=1e-9)
    np.testing.assert_allclose([r[3] for r in rows], [float(r[3]) for r in gold], rtol=1e-9)


@pytest.mark.parametrize("kind,model", [("AA", "aa"), ("DD", "ad"), ("AD", "ad")])
def test_screen_port_matches_reference(kind, model, golden, mouse_model, mouse_ro, tmp_path):
    """C port of the screen == the reference's own C (oracle/_ref) == golden file made through the reference
    Python wrapper; effects compared at the %g precision the reference writes."""
    M = mouse_model
    from scipy.stats import chi2
    pymat, _ = oracle.preamble(M["y"], M["xmat"], M["zmat"], M["gl_" + model], M["vc_" + model], scipy_inv=False)
    var_app = golden["pair_var_median"][kind]
    eff_cut = np.sqrt(chi2.isf(1e-3, 1) * var_app)
    m = golden["m"]
    lst = list(range(m)) if kind == "AD" else list(range(m - 1))
    rows = oracle.screen(mouse_ro[0], kind, lst, pymat, eff_cut=eff_cut)
    mine = sorted((i, j, "%g" % e) for i, j, e in rows)
    gold = sorted((int(r[0]), int(r[1]), r[2]) for r in read_rows(gold_path("epi%s_eff.txt" % kind)))
    assert mine == gold
    if ref_harness.ref_so_available():
        ref = oracle.ref_screen_file(mouse_ro[0], kind, lst, pymat, eff_cut, str(tmp_path / "ref.txt"))
        assert ref == mine


def test_parallel_partition_matches_reference(golden):
    for part in golden["parallel_partitions"]:
        lst, pos = oracle.parallel_lists(part["m"], [part["P"], part["k"]], False)
        assert pos == part["AA"]
        lst, pos = oracle.parallel_lists(part["m"], [part["P"], part["k"]], True)
        assert pos == part["AD"]


def test_reference_golden_vector(golden):
    """The reference's only own known answer: examples/remma/remma_cpu.py:178."""
    np.testing.assert_allclose(golden["var_com_ag_agag"], [0.06289206, 0.07641075, 0.08121168], atol=5e-9)


def test_reml_restatement_matches_reference(golden, mouse_model):
    """oracle.wemai_multi_gmat against the variance components the unmodified reference produced on mouse
    (== the reference's own golden vector, examples/remma/remma_cpu.py:178, to its 8 printed digits)."""
    M = mouse_model
    vc = oracle.wemai_multi_gmat(M["y"], M["xmat"], M["zmat"], M["gl_aa"])
    np.testing.assert_allclose(vc, golden["var_com_ag_agag"], rtol=1e-7)
    assert [round(float(v), 8) for v in vc] == [0.06289206, 0.07641075, 0.08121168]


def test_impute_geno_reproduces_the_reference_draws():
    """gmat_b200's impute_geno against the unmodified reference's (fixture by oracle/make_golden_impute.py): same
    imputed matrix under the same state of the global numpy RNG, and the same RNG state afterwards."""
    import json
    from gmat_b200.process_plink.process_plink import impute_geno
    g = json.load(open(os.path.join(GOLD, "impute_geno.json")))
    src = np.array(g["input"], dtype=np.float64)
    src[src < 0] = np.nan
    np.random.seed(g["seed"])
    out = impute_geno(src.copy())
    assert not np.isnan(out).any()
    assert np.array_equal(out, np.array(g["output"], dtype=np.float64))
    assert float(np.random.random()) == g["next_random"]
