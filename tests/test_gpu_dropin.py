"""GPU parity tests, drop-in level: the reference-named Python front-ends on the bundled mouse data
(BASELINE.json configs[0]) against tests/golden -- files written by the UNMODIFIED reference
(oracle/make_golden.py).  Gates (north_star): identical reported pair set at p_cut, eff/var/chi within
1e-9 relative, p within 1e-7 relative (scipy on both sides)."""
import os
import shutil

import numpy as np
import pytest

from conftest import GOLD, gold_path, read_rows

pytestmark = pytest.mark.gpu
RTOL = 1e-9


@pytest.fixture(scope="module")
def grms(tmp_path_factory, golden):
    """agmat / dgmat_as through the drop-in API on a writable copy of mouse; checks outputs on the way."""
    import torch
    assert torch.cuda.is_available()
    from gmat_b200.gmatrix import agmat, dgmat_as
    d = tmp_path_factory.mktemp("mouse")
    for f in os.listdir(os.path.join(GOLD, "mouse")):
        shutil.copyfile(os.path.join(GOLD, "mouse", f), str(d / f))
    bed = str(d / "plink")
    ag, inv = agmat(bed)
    assert inv is None and os.path.exists(bed + ".agrm0")
    dg, _ = dgmat_as(bed)
    assert os.path.exists(bed + ".dgrm_as0")
    return {"bed": bed, "pheno": str(d / "pheno"), "pheno_rep3": str(d / "pheno_rep3"), "ag": ag, "dg": dg,
            "dir": str(d)}


def _cmp_rows(got_file, gold_file, float_cols, rtol=RTOL, p_col=None):
    got, want = read_rows(got_file), read_rows(gold_file)
    assert open(got_file).readline() == open(gold_file).readline()             # header
    assert [(r[0], r[1]) for r in got] == [(r[0], r[1]) for r in want]          # identical pair set AND order
    for c in float_cols:
        g, w = np.array([float(r[c]) for r in got]), np.array([float(r[c]) for r in want])
        np.testing.assert_allclose(g, w, rtol=(1e-7 if c == p_col else rtol), atol=1e-300)
    return len(got)


def test_grm_matches_reference(grms, golden):
    ri, ci = golden["grm_sample_rows"], golden["grm_sample_cols"]
    for name, K in (("agmat", grms["ag"]), ("dgmat", grms["dg"])):
        mx = golden[name + "_absmax"]
        assert np.abs(K[np.ix_(ri, ci)] - np.array(golden[name + "_sample"])).max() <= RTOL * mx
        assert np.abs(np.diag(K)[:16] - np.array(golden[name + "_diag_first16"])).max() <= RTOL * mx
        np.testing.assert_allclose(K.sum(), golden[name + "_sum"], rtol=1e-8)
        np.testing.assert_allclose(np.linalg.norm(K), golden[name + "_fro"], rtol=1e-10)
    # the text file is np.savetxt('%.18e'): first row round-trips to the returned matrix
    first = np.array(open(grms["bed"] + ".agrm0").readline().split(), dtype=float)
    assert np.array_equal(first, grms["ag"][0])


def test_grm_other_formats_and_inverse(grms):
    from gmat_b200.gmatrix import agmat
    ag, aginv = agmat(grms["bed"], inv=True, out_fmt="row_col_val")
    rows = read_rows(grms["bed"] + ".agrm1", skip_header=False)
    assert rows[0][:2] == ["1", "1"] and rows[2][:2] == ["2", "2"] and len(rows) == 1304 * 1305 // 2
    assert abs(float(rows[1][2]) - ag[1, 0]) <= 1e-15
    np.testing.assert_allclose(aginv @ ag, np.eye(1304), atol=1e-7)
    assert os.path.exists(grms["bed"] + ".agiv1")


@pytest.mark.parametrize("kind", ["AA", "DD", "AD"])
def test_exact_scan_full_mouse(grms, golden, kind, tmp_path):
    """BASELINE configs[0] data, exact scans: 989,121 AA/DD pairs, 1,979,649 AD tests, p_cut = 1e-5."""
    import importlib
    mod = importlib.import_module("gmat_b200.remma.remma_epi" + kind)
    ag, dg = grms["ag"], grms["dg"]
    gl, vc = ([ag, ag * ag], golden["var_com_ag_agag"]) if kind == "AA" else \
        ([ag, dg, ag * ag], golden["var_com_ag_dg_agag"])
    out = str(tmp_path / ("epi" + kind))
    assert getattr(mod, "remma_epi" + kind)(grms["pheno"], grms["bed"], gl, vc, p_cut=1e-5, out_file=out) == 0
    n = _cmp_rows(out, gold_path("epi%s.txt" % kind), (2, 3, 4), p_col=4)
    assert n == {"AA": 8, "DD": 2, "AD": 42}[kind]


@pytest.mark.parametrize("kind", ["AA", "DD", "AD"])
def test_exact_scan_subset_loose_cut(grms, golden, kind, tmp_path):
    import importlib
    mod = importlib.import_module("gmat_b200.remma.remma_epi" + kind)
    ag, dg = grms["ag"], grms["dg"]
    gl, vc = ([ag, ag * ag], golden["var_com_ag_agag"]) if kind == "AA" else \
        ([ag, dg, ag * ag], golden["var_com_ag_dg_agag"])
    out = str(tmp_path / "sub")
    getattr(mod, "remma_epi" + kind)(grms["pheno"], grms["bed"], gl, vc, snp_lst_0=golden["snp_lst_0_subset"],
                                     p_cut=1e-3, out_file=out)
    assert _cmp_rows(out, gold_path("epi%s_sub_p1e-3.txt" % kind), (2, 3, 4), p_col=4) > 10


def test_exact_scan_parallel_parts(grms, golden, tmp_path):
    from gmat_b200.remma.remma_epiAA import remma_epiAA_parallel
    from gmat_b200.remma.remma_epiAD import remma_epiAD_parallel
    ag, dg = grms["ag"], grms["dg"]
    for k in (1, 2):
        out = str(tmp_path / "par")
        remma_epiAA_parallel(grms["pheno"], grms["bed"], [ag, ag * ag], golden["var_com_ag_agag"],
                             parallel=[24, k], p_cut=1e-3, out_file=out)
        _cmp_rows(out + ".%d" % k, gold_path("epiAA_parallel.%d" % k), (2, 3, 4), p_col=4)
        remma_epiAD_parallel(grms["pheno"], grms["bed"], [ag, dg, ag * ag], golden["var_com_ag_dg_agag"],
                             parallel=[24, k], p_cut=1e-3, out_file=out + "AD")
        _cmp_rows(out + "AD.%d" % k, gold_path("epiAD_parallel.%d" % k), (2, 3, 4), p_col=4)


def test_exact_scan_repeated_measures(grms, golden, tmp_path):
    """Config C5 shape: 3 records per individual, GRMs [A, A*A, I]; Z is a real N x n incidence matrix."""
    from gmat_b200.remma.remma_epiAA import remma_epiAA
    ag = grms["ag"]
    out = str(tmp_path / "rep3")
    remma_epiAA(grms["pheno_rep3"], grms["bed"], [ag, ag * ag, np.eye(ag.shape[0])], golden["var_com_rep3"],
                snp_lst_0=golden["snp_lst_0_subset"], p_cut=1e-3, out_file=out)
    assert _cmp_rows(out, gold_path("epiAA_rep3.txt"), (2, 3, 4), p_col=4) > 50


@pytest.mark.parametrize("kind", ["AA", "DD", "AD"])
def test_pair_list(grms, golden, kind, tmp_path):
    import importlib
    mod = importlib.import_module("gmat_b200.remma.remma_epi" + kind)
    ag, dg = grms["ag"], grms["dg"]
    gl, vc = ([ag, ag * ag], golden["var_com_ag_agag"]) if kind == "AA" else \
        ([ag, dg, ag * ag], golden["var_com_ag_dg_agag"])
    pairs = os.path.join(GOLD, "pairs_AD.txt" if kind == "AD" else "pairs_AA.txt")
    out = str(tmp_path / "pair")
    getattr(mod, "remma_epi%s_pair" % kind)(grms["pheno"], grms["bed"], gl, vc, pairs, p_cut=1, out_file=out)
    assert _cmp_rows(out, gold_path("epi%s_pair.txt" % kind), (2, 3, 4, 5), p_col=5) >= 2500
    if kind == "AA":                      # chunked reading + a real threshold
        getattr(mod, "remma_epiAA_pair")(grms["pheno"], grms["bed"], gl, vc, pairs, max_test_pair=700, p_cut=0.05,
                                         out_file=out + "c")
        _cmp_rows(out + "c", gold_path("epiAA_pair_chunk700_p05.txt"), (2, 3, 4, 5), p_col=5)


def _cmp_set(got_file, gold_file, float_cols, rtol):
    """Row order of the C screen is unspecified (OpenMP fprintf race): compare as sets keyed by the pair."""
    assert open(got_file).readline() == open(gold_file).readline()
    got = {(r[0], r[1]): r for r in read_rows(got_file)}
    want = {(r[0], r[1]): r for r in read_rows(gold_file)}
    assert set(got) == set(want)
    for k in want:
        for c in float_cols:
            assert abs(float(got[k][c]) - float(want[k][c])) <= rtol * abs(float(want[k][c])) + 1e-300, (k, c)
    return len(got)


@pytest.mark.parametrize("kind", ["AA", "DD", "AD"])
def test_eff_screen_wrappers(grms, golden, kind, tmp_path):
    """remma_epiXX_eff (CUDA screen behind the reference's cffi call shape) vs the reference C + wrapper."""
    import importlib
    mod = importlib.import_module("gmat_b200.remma.remma_epi" + kind)
    ag, dg = grms["ag"], grms["dg"]
    gl, vc = ([ag, ag * ag], golden["var_com_ag_agag"]) if kind == "AA" else \
        ([ag, dg, ag * ag], golden["var_com_ag_dg_agag"])
    out = str(tmp_path / "eff")
    getattr(mod, "remma_epi%s_eff" % kind)(grms["pheno"], grms["bed"], gl, vc, var_app=golden["pair_var_median"][kind],
                                           p_cut=1e-3, out_file=out)
    # eff is %g text (6 digits): allow one unit in the last printed digit; chi_app/p_app follow from it
    assert _cmp_set(out, gold_path("epi%s_eff.txt" % kind), (2,), 2e-6) > 100
    assert _cmp_set(out, gold_path("epi%s_eff.txt" % kind), (3, 4), 2e-4) > 100
    assert not os.path.exists(out + ".temp")
    if kind == "AA":
        getattr(mod, "remma_epiAA_eff")(grms["pheno"], grms["bed"], gl, vc, var_app=1575.0, p_cut=1e-5,
                                        out_file=out + "b")
        assert _cmp_set(out + "b", gold_path("epiAA_eff_p1e-5.txt"), (2,), 2e-6) == 7
        getattr(mod, "remma_epiAA_eff")(grms["pheno"], grms["bed"], gl, vc, snp_lst_0=golden["snp_lst_0_subset"],
                                        var_app=golden["pair_var_median"]["AA"], p_cut=1e-2, out_file=out + "c")
        _cmp_set(out + "c", gold_path("epiAA_eff_sub.txt"), (2,), 2e-6)
        getattr(mod, "remma_epiAA_eff_parallel")(grms["pheno"], grms["bed"], gl, vc, parallel=[24, 2],
                                                 var_app=golden["pair_var_median"]["AA"], p_cut=1e-3,
                                                 out_file=out + "p")
        _cmp_set(out + "p.2", gold_path("epiAA_eff_parallel.2"), (2,), 2e-6)


@pytest.mark.parametrize("kind", ["AA", "DD", "AD"])
@pytest.mark.parametrize("maf", [False, True])
def test_approx_pipelines(grms, golden, kind, maf, tmp_path, monkeypatch):
    """3-stage pipelines fed the very random pairs the reference drew (kept by make_golden.py)."""
    import importlib
    from gmat_b200.remma import _core
    mod = importlib.import_module("gmat_b200.remma.remma_epi" + kind)
    stem = "epi%s_%sapprox" % (kind, "maf_" if maf else "")

    def fake_random_pair(num_snp, out_file="random_pair", num_pair=100000, num_each_pair=5000, ordered=True):
        shutil.copyfile(gold_path(stem + ".random_pair"), out_file)
        return np.loadtxt(out_file, dtype=int, skiprows=1)
    monkeypatch.setattr(_core, "random_pair", fake_random_pair)
    monkeypatch.setattr(_core, "random_pairAD", fake_random_pair)
    ag, dg = grms["ag"], grms["dg"]
    gl, vc = ([ag, ag * ag], golden["var_com_ag_agag"]) if kind == "AA" else \
        ([ag, dg, ag * ag], golden["var_com_ag_dg_agag"])
    out = str(tmp_path / stem)
    monkeypatch.chdir(tmp_path)                  # the _maf_ pipelines write ./eff_cut, like the reference
    fn = getattr(mod, "remma_epi%s_%sapprox" % (kind, "maf_" if maf else ""))
    assert fn(grms["pheno"], grms["bed"], gl, vc, p_cut=1e-4, num_random_pair=20000, out_file=out) == 0
    # what is left on disk is what the reference leaves (remma_epiAA_maf_approx.py:29-31,88-89 / remma_epiAA_approx.py:32-33,51-52)
    for ext in (".random_pair", ".random_pairAD", ".approx_p", ".exact_p"):
        assert not os.path.exists(out + ext)
    assert os.path.exists(out + ".random") == maf
    if maf:
        # out_file.random = the reference's own _pair call on the random pairs with p_cut = 1
        getattr(mod, "remma_epi%s_pair" % kind)(grms["pheno"], grms["bed"], gl, vc,
                                                snp_pair_file=gold_path(stem + ".random_pair"), p_cut=1,
                                                out_file=out + ".random_by_pair")
        assert open(out + ".random").read() == open(out + ".random_by_pair").read()
        assert os.path.exists(str(tmp_path / "eff_cut")) == (kind != "AD")
        if kind != "AD":
            assert np.loadtxt(str(tmp_path / "eff_cut")).shape == (111,)
    # columns: snp_0 snp_1 eff var chi p_app p ; p_app derives from the %g-truncated effect
    n = _cmp_set(out, gold_path(stem + ".txt"), (2, 3, 4), RTOL)
    _cmp_set(out, gold_path(stem + ".txt"), (6,), 1e-7)
    _cmp_set(out, gold_path(stem + ".txt"), (5,), 2e-4)
    assert n > 20
    if maf:
        want = {tuple(r[:2]): float(r[2]) for r in read_rows(gold_path(stem + ".freq_denominator"), False)}
        got = {tuple(r[:2]): float(r[2]) for r in read_rows(out + ".freq_denominator", False)}
        assert set(got) == set(want)
        for k in want:
            assert abs(got[k] - want[k]) <= RTOL * abs(want[k])


def test_device_preamble_matches_host(grms, golden):
    """The GPU preamble (default of the front-ends) against the host numpy restatement of
    remma_epiAA.py:31-51 on the mouse model, without and with repeated records (3 per individual, extra I GRM)."""
    import torch
    from gmat_b200.remma import _core
    from gmat_b200.uvlmm.design_matrix import design_matrix_wemai_multi_gmat
    ag = grms["ag"]
    cases = [(grms["pheno"], [ag, ag * ag], golden["var_com_ag_agag"]),
             (grms["pheno_rep3"], [ag, ag * ag, np.eye(ag.shape[0])], golden["var_com_rep3"])]
    for pheno, gl, vc in cases:
        y, xmat, zmat = design_matrix_wemai_multi_gmat(pheno, grms["bed"])
        py_h, pvp_h = _core.preamble(y, xmat, zmat, gl, vc, scipy_inv=True)
        py_d, pvp_d, _ = _core.preamble_device(y, xmat, zmat, gl, vc, torch.device("cuda"))
        np.testing.assert_allclose(py_d.cpu().numpy(), py_h.ravel(), rtol=1e-10, atol=1e-12 * np.abs(py_h).max())
        np.testing.assert_allclose(pvp_d.cpu().numpy(), pvp_h, rtol=0, atol=1e-11 * np.abs(pvp_h).max())


def test_reml_wemai_multi_gmat_matches_reference(grms, golden, tmp_path):
    """gmat_b200.uvlmm.wemai_multi_gmat (GPU REML) against the reference's variance components
    on mouse for both model lists; the first equals the reference's own golden vector remma_cpu.py:178."""
    from gmat_b200.uvlmm import wemai_multi_gmat
    ag, dg = grms["ag"], grms["dg"]
    out = str(tmp_path / "var.txt")
    vc = wemai_multi_gmat(grms["pheno"], grms["bed"], [ag, ag * ag], out_file=out)
    np.testing.assert_allclose(vc, golden["var_com_ag_agag"], rtol=1e-6)
    assert [round(float(v), 8) for v in vc] == [0.06289206, 0.07641075, 0.08121168]
    np.testing.assert_allclose(np.loadtxt(out), vc, rtol=0, atol=0)
    vc2 = wemai_multi_gmat(grms["pheno"], grms["bed"], [ag, dg, ag * ag], out_file=out)
    np.testing.assert_allclose(vc2, golden["var_com_ag_dg_agag"], rtol=1e-5)
