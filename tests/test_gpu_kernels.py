"""GPU parity tests, kernel level: every CUDA entry point through the C-ABI against the oracle on seeded
inputs (bit-exact for integer/byte/table work, 1e-9 relative for FP64 statistics -- the tolerance
BASELINE.json's north_star states)."""
import os

import numpy as np
import pytest

from conftest import GOLD, gold_path, read_rows
from oracle import oracle

pytestmark = pytest.mark.gpu
RTOL = 1e-9


@pytest.fixture(scope="module")
def eng():
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from gmat_b200 import engine
    return engine


def _dataset(tmp_path, n, m, seed, missing=0.0):
    import synth
    prefix = str(tmp_path / ("syn_%d_%d" % (n, m)))
    return prefix, synth.write_dataset(prefix, n, m, seed=seed, missing_rate=missing)


# ------------------------------------------------------------------------------------------ decode
@pytest.mark.parametrize("n,m,missing", [(1304, 1407, 0), (5, 3, 0), (64, 7, 0.1), (65, 9, 0.1), (66, 5, 0),
                                         (67, 11, 0.2), (1, 4, 0), (257, 33, 0.05)])
def test_read_plink_bed_bit_exact(eng, tmp_path, mouse_ro, n, m, missing):
    """Drop-in symbol read_plink_bed == reference reader (_read_plink_bed.c), every n mod 4 tail, all 4 codes."""
    from gmat_b200.process_plink import read_plink_bed
    if (n, m) == (1304, 1407):
        prefix = mouse_ro[0]
    else:
        prefix, _ = _dataset(tmp_path, n, m, seed=n * 1000 + m, missing=missing)
    got = read_plink_bed(prefix, n, m)
    want = oracle.read_bed_raw(prefix, n, m)
    assert got.shape == want.shape and np.array_equal(got, want)
    if missing:
        assert np.any(np.abs(want - 1.0 / 3) < 1e-4)            # code 01 exercised


def test_read_plink_and_bed_class(eng, tmp_path, mouse_ro):
    from gmat_b200.process_plink import Bed, read_plink
    assert np.array_equal(read_plink(mouse_ro[0]), oracle.read_plink(mouse_ro[0]))
    prefix, _ = _dataset(tmp_path, 37, 21, seed=5, missing=0.15)
    a, b = read_plink(prefix), oracle.read_plink(prefix)
    assert np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(np.nan_to_num(a), np.nan_to_num(b))
    c = Bed(prefix).read()
    assert np.array_equal(np.nan_to_num(c), np.nan_to_num(b))


def test_counts_luts_and_coded_decode_bit_exact(eng, tmp_path, mouse_ro):
    for prefix, nm in ((mouse_ro[0], None), (_dataset(tmp_path, 131, 40, seed=9)[0], (131, 40))):
        store = eng.GenotypeStore(prefix)
        raw = oracle.read_bed_raw(prefix)                       # [m][n]
        cnt = store.counts.cpu().numpy()
        for c, val in enumerate((0.0, 1.0 / 3, 1.0, 2.0)):
            assert np.array_equal(cnt[:, c], np.sum(np.abs(raw - val) < 1e-6, axis=1))
        snp_mat = np.ascontiguousarray(raw.T)
        xa, freq = oracle.code_additive(snp_mat)
        xd, _ = oracle.code_dominance(snp_mat)
        assert np.array_equal(store.freq().cpu().numpy(), freq.ravel())
        assert np.array_equal(store.decode(eng.ADD).cpu().numpy(), xa.T)
        assert np.array_equal(store.decode(eng.DOM).cpu().numpy(), xd.T)
        # C-flavoured coding: p accumulated term by term (_remma_epi_eff_cpu.c:103-112)
        ca, cfreq = oracle.c_coded(prefix, 1)
        cd, _ = oracle.c_coded(prefix, 2)
        assert np.array_equal(store.freq(eng.FREQ_CLOOP).cpu().numpy(), cfreq)
        assert np.array_equal(store.decode(eng.ADD, eng.FREQ_CLOOP).cpu().numpy(), ca)
        assert np.array_equal(store.decode(eng.DOM, eng.FREQ_CLOOP).cpu().numpy(), cd)


def test_pack_roundtrip(eng, tmp_path):
    prefix, ds = _dataset(tmp_path, 203, 57, seed=3, missing=0.1)
    store = eng.GenotypeStore(prefix)
    g = ds["geno"].astype(np.float64)
    g[g < 0] = np.nan
    store2 = eng.GenotypeStore.from_dense(g)
    nb = (203 + 3) // 4
    a, b = store.bed.cpu().numpy()[:, :nb].copy(), store2.bed.cpu().numpy()[:, :nb].copy()
    a[:, -1] &= 0x3f                                             # 203 % 4 == 3: top pair is padding
    b[:, -1] &= 0x3f
    assert np.array_equal(a, b)
    assert np.array_equal(store.counts.cpu().numpy(), store2.counts.cpu().numpy())


# ------------------------------------------------------------------------------------------ approximate screen
def _assert_g_text(texts, values):
    """The file holds eff as C '%g' (6 significant digits, _remma_epi_eff_cpu.c:75): the text must be the
    '%g' rendering of the oracle value, or of a value within 1e-9 relative of it (a rounding-boundary case)."""
    for t, v in zip(texts, values):
        assert t in ("%g" % v, "%g" % (v * (1 + RTOL)), "%g" % (v * (1 - RTOL))), (t, v)


def _pymat(n, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    py = rng.normal(size=n)
    return py - py.mean()


@pytest.mark.parametrize("kind", ["AA", "DD", "AD"])
@pytest.mark.parametrize("n,m", [(131, 300), (64, 129), (500, 257)])
def test_screen_dropin_symbols_vs_oracle(eng, tmp_path, kind, n, m):
    """remma_epiXX_eff_cpu (CUDA) == C restatement: identical row set, effects within 1e-9 (file: %g)."""
    from gmat_b200._ffi import check, ffi, lib
    from scipy.stats import chi2
    prefix, _ = _dataset(tmp_path, n, m, seed=n + m, missing=0.02 if n == 131 else 0.0)
    py = _pymat(n, 1)
    lst = np.arange(m - 1, dtype=np.int64)
    full = oracle.screen(prefix, kind, lst, py, eff_cut=0.0)
    cut = float(np.quantile(np.abs([r[2] for r in full]), 0.98)) * (1 + 1e-7)
    want = [r for r in oracle.screen(prefix, kind, lst, py, eff_cut=cut)]
    out = str(tmp_path / "scr.txt")
    fn = getattr(lib, "remma_epi%s_eff_cpu" % kind)
    eng.current_device()
    check(fn(ffi.new("char[]", prefix.encode()), n, m, ffi.cast("long long *", lst.ctypes.data), len(lst),
             ffi.cast("double *", py.ctypes.data), cut, ffi.new("char[]", out.encode())))
    rows = read_rows(out)
    assert open(out).readline() == "snp_0 snp_1 eff\n"
    assert [(int(r[0]), int(r[1])) for r in rows] == [(r[0], r[1]) for r in want]     # same set, same order
    _assert_g_text([r[2] for r in rows], [r[2] for r in want])
    # full precision through the device API
    store = eng.GenotypeStore(prefix)
    import torch
    pyd = torch.as_tensor(py).cuda()
    la, ld = store.lut(eng.ADD, eng.FREQ_CLOOP), store.lut(eng.DOM, eng.FREQ_CLOOP)
    passes = {"AA": [(la, la, 0)], "DD": [(ld, ld, 0)],
              "AD": [(la, ld, eng.EFF_GE), (ld, la, eng.EFF_SWAP)]}[kind]
    got = {}
    for lu, lv, flags in passes:
        h = eng.epi_eff(store, lu, lv, pyd, lst, flags, cut)
        for a, b, e in zip(h["snp_0"], h["snp_1"], h["eff"]):
            assert (int(a), int(b)) not in got
            got[(int(a), int(b))] = float(e)
    ref = {(r[0], r[1]): r[2] for r in want}
    assert set(got) == set(ref)
    for k in got:
        assert abs(got[k] - ref[k]) <= RTOL * abs(ref[k]) + 1e-12


@pytest.mark.parametrize("kind", ["AA", "DD", "AD"])
def test_maf_screen_dropin_vs_oracle(eng, tmp_path, kind):
    from gmat_b200._ffi import check, ffi, lib
    n, m = 200, 260
    prefix, ds = _dataset(tmp_path, n, m, seed=77)
    py = _pymat(n, 2)
    lst = np.arange(0, m - 1, 2, dtype=np.int64)               # ragged, non-contiguous first-SNP list
    g = ds["geno"].astype(np.float64)
    maf = g.sum(axis=1) / (2 * n)
    maf = np.minimum(maf, 1 - maf)
    het = (g == 1).sum(axis=1) / n
    het = np.minimum(het, 1 - het)
    bi = np.array(maf * 20, dtype=np.int64) if kind != "DD" else np.array(het * 20, dtype=np.int64)
    bj = np.array(het * 20, dtype=np.int64) if kind == "AD" else bi
    rng = np.random.Generator(np.random.PCG64(4))
    base = np.quantile(np.abs([r[2] for r in oracle.screen(prefix, kind, lst, py)]), 0.97)
    table = base * rng.uniform(0.8, 1.2, size=111)
    want = oracle.screen(prefix, kind, lst, py, table=table, bin_i=bi, bin_j=bj)
    out = str(tmp_path / "maf.txt")
    eng.current_device()
    args = [ffi.new("char[]", prefix.encode()), n, m, ffi.cast("long long *", lst.ctypes.data), len(lst),
            ffi.cast("double *", py.ctypes.data), ffi.cast("long long *", bi.ctypes.data)]
    if kind == "AD":
        args.append(ffi.cast("long long *", bj.ctypes.data))
    args += [ffi.cast("double *", table.ctypes.data), ffi.new("char[]", out.encode())]
    check(getattr(lib, "remma_epi%s_maf_eff_cpu" % kind)(*args))
    rows = read_rows(out)
    assert len(want) > 20
    assert [(int(r[0]), int(r[1])) for r in rows] == [(r[0], r[1]) for r in want]
    _assert_g_text([r[2] for r in rows], [r[2] for r in want])


def test_screen_error_paths(eng, tmp_path):
    """Failures return a negative status + message instead of exit(1) (_remma_epi_eff_cpu.c:18-21)."""
    from gmat_b200._ffi import GmatB200Error, check, ffi, lib
    py = np.zeros(8)
    lst = np.zeros(1, dtype=np.int64)
    eng.current_device()
    rc = lib.remma_epiAA_eff_cpu(ffi.new("char[]", str(tmp_path / "nope").encode()), 8, 4,
                                 ffi.cast("long long *", lst.ctypes.data), 1,
                                 ffi.cast("double *", py.ctypes.data), 1.0,
                                 ffi.new("char[]", str(tmp_path / "o.txt").encode()))
    assert rc == -2 and b"nope.bed" in ffi.string(lib.gmat_b200_last_error())
    bad = tmp_path / "bad.bed"
    bad.write_bytes(b"\x00\x01\x02" + b"\x00" * 64)
    with pytest.raises(GmatB200Error):
        check(lib.read_plink_bed(ffi.new("char[]", str(tmp_path / "bad").encode()), 8, 4,
                                 ffi.cast("double *", np.zeros(32).ctypes.data)))


# ------------------------------------------------------------------------------------------ exact test
def _random_model(n, seed):
    """A symmetric PSD 'Z'PZ' with zero row sums and a 'Z'Py' with zero sum (SURVEY appendix B invariants)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    b = rng.normal(size=(n, n + 3))
    a = b @ b.T / n
    c = np.eye(n) - np.ones((n, n)) / n
    a = c @ a @ c
    py = c @ rng.normal(size=n)
    return py.reshape(-1, 1), 0.5 * (a + a.T)


@pytest.mark.parametrize("kind", ["AA", "AD", "DD"])
@pytest.mark.parametrize("n,m", [(128, 40), (131, 77), (300, 50), (17, 9)])
def test_exact_pairs_and_scan_vs_oracle_synthetic(eng, tmp_path, kind, n, m):
    """Ragged sizes (n not a multiple of 16/128, m not a multiple of 16): eff, var within 1e-9 of numpy."""
    import torch
    prefix, _ = _dataset(tmp_path, n, m, seed=n * 7 + m)
    pymat, pvp = _random_model(n, n + m)
    store = eng.GenotypeStore(prefix)
    packed = eng.pack_pvp(torch.as_tensor(pvp).cuda())
    pyd = torch.as_tensor(pymat.ravel().copy()).cuda()
    cu, cv = {"AA": (eng.ADD, eng.ADD), "AD": (eng.ADD, eng.DOM), "DD": (eng.DOM, eng.DOM)}[kind]
    lu, lv = store.lut(cu), store.lut(cv)
    U, V = oracle.coded_pair(kind, oracle.read_plink(prefix))
    rng = np.random.Generator(np.random.PCG64(1))
    pairs = rng.integers(0, m, size=(333, 2))
    res = eng.epi_exact_pairs(store, lu, lv, pyd, packed, pairs)
    eff, var, chi, _ = oracle.exact_pairs(U, V, pairs, pymat, pvp)
    assert np.array_equal(res["snp_0"], pairs[:, 0]) and np.array_equal(res["snp_1"], pairs[:, 1])
    np.testing.assert_allclose(res["eff"], eff, rtol=RTOL, atol=1e-11)
    np.testing.assert_allclose(res["var"], var, rtol=RTOL)
    np.testing.assert_allclose(res["chi"], chi, rtol=4 * RTOL, atol=1e-13)     # eff ~ 0 -> chi ~ 0
    # scan over a ragged first-SNP list, every pair reported (chi_cut < 0)
    all_pairs = kind == "AD"
    lst = [i for i in range(0, m - (0 if all_pairs else 1)) if i % 3 != 1]
    hits = eng.epi_exact_scan(store, lu, lv, pyd, packed, np.array(lst), eng.PAIRS_ALL if all_pairs else eng.PAIRS_UPPER,
                              -1.0)
    want = {}
    for i in lst:
        js, e, v, c, _ = oracle.exact_first_snp(U, V, i, pymat, pvp, all_pairs)
        for q in range(len(js)):
            want[(i, int(js[q]))] = (e[q], v[q])
    got = {(int(a), int(b)): (e, v) for a, b, e, v in zip(hits["snp_0"], hits["snp_1"], hits["eff"], hits["var"])}
    assert set(got) == set(want) and len(hits) == len(want)
    ge = np.array([got[k] for k in want])
    we = np.array([want[k] for k in want])
    np.testing.assert_allclose(ge[:, 0], we[:, 0], rtol=RTOL, atol=1e-11)
    np.testing.assert_allclose(ge[:, 1], we[:, 1], rtol=RTOL)


def test_exact_scan_hit_buffer_overflow_retry(eng, tmp_path):
    import torch
    n, m = 64, 120
    prefix, _ = _dataset(tmp_path, n, m, seed=11)
    pymat, pvp = _random_model(n, 5)
    store = eng.GenotypeStore(prefix)
    packed = eng.pack_pvp(torch.as_tensor(pvp).cuda())
    pyd = torch.as_tensor(pymat.ravel().copy()).cuda()
    lu = store.lut(eng.ADD)
    hits = eng.epi_exact_scan(store, lu, lu, pyd, packed, np.arange(m - 1), eng.PAIRS_UPPER, -1.0, cap=100)
    assert len(hits) == m * (m - 1) // 2
    assert len({(int(a), int(b)) for a, b in zip(hits["snp_0"], hits["snp_1"])}) == len(hits)


# ------------------------------------------------------------------------------------------ GRM
@pytest.mark.parametrize("which", ["f4", "i8", "f64"])
@pytest.mark.parametrize("n,m", [(131, 300), (128, 16), (260, 1000), (5, 3), (513, 130), (250, 513), (495, 257)])
def test_grm_vs_oracle_synthetic(eng, tmp_path, monkeypatch, n, m, which):
    """The three GRM kernels (FP4 and int8 tensor cores: gram_i8.cu; FP64 DMMA: grm.cu) against the numpy oracle;
    n = 250 / 495: the 18 digit rows of the FP4 kernel cross / open a 256-row tile, m = 513 / 257: ragged last
    256-SNP slab."""
    from gmat_b200.gmatrix.gmatrix import grm_device
    monkeypatch.setenv("GMAT_B200_GRM_ENGINE", which)
    prefix, _ = _dataset(tmp_path, n, m, seed=n + 3 * m)
    store = eng.GenotypeStore(prefix)
    for coding, fn in ((eng.ADD, oracle.agmat), (eng.DOM, oracle.dgmat_as)):
        want, scale = fn(prefix)
        kin, s = grm_device(store, coding)
        got = kin.cpu().numpy()
        assert s == scale
        assert np.abs(got - want).max() <= RTOL * np.abs(want).max()
        assert np.array_equal(got, got.T)
    # K-sliced accumulation (the multi-GPU path): partial sums over SNP ranges add up to X X'
    import torch
    acc = torch.zeros((n, n), dtype=torch.float64, device="cuda")
    cuts = [0, m // 3, m // 3, m]
    for a, b in zip(cuts[:-1], cuts[1:]):
        eng.grm(store, eng.ADD, 1.0, 0.0, snp_begin=a, snp_end=b, out=acc, accumulate=True)
    want, scale = oracle.agmat(prefix, small_val=0.0)
    assert np.abs(acc.cpu().numpy() / scale - want).max() <= RTOL * np.abs(want).max()


@pytest.mark.parametrize("which", ["f4", "i8", "f64"])
def test_grm_streamk_many_tiles_reproducible(eng, tmp_path, monkeypatch, which):
    """n large enough for more tiles than SMs (the equal-range stream-K split with atomically added partial tiles):
    parity with the oracle and bit-identical results run to run (FP64: a tile never has more than two parts;
    int8: the partial sums are integers)."""
    from gmat_b200.gmatrix.gmatrix import grm_device
    monkeypatch.setenv("GMAT_B200_GRM_ENGINE", which)
    n, m = 2300, 200                      # 18 blocks -> 171 tiles > 148 SMs; 4 stages of 64 SNPs (last one ragged)
    prefix, _ = _dataset(tmp_path, n, m, seed=99)
    store = eng.GenotypeStore(prefix)
    want, scale = oracle.agmat(prefix)
    kin, s = grm_device(store, eng.ADD)
    got = kin.cpu().numpy()
    assert s == scale
    assert np.abs(got - want).max() <= RTOL * np.abs(want).max()
    assert np.array_equal(got, got.T)
    for _ in range(3):
        again, _ = grm_device(store, eng.ADD)
        assert np.array_equal(again.cpu().numpy(), got)


@pytest.mark.parametrize("n,m", [(3300, 300), (4200, 140), (300, 9000)])
def test_grm_i8_wave_schedule_and_chunked_corrections(eng, monkeypatch, n, m):
    """The int8 GRM's schedule beyond one wave of tiles -- n=3300: 91 tiles of 256 x 256 = one lockstep wave of
    whole tiles (plain stores) + 17 tiles cut stream-K style (integer atomics); n=4200: two waves + 5 -- and its
    FP64 corrections over more than one 4096-SNP chunk (m=9000), against numpy on genotypes generated in HBM, and
    against the FP64 DMMA kernel; bit-identical run to run."""
    import torch
    import synth
    from gmat_b200.gmatrix.gmatrix import grm_device
    payload = synth.device_payload(n, m, 4242 + n, "cuda", eng.row_stride(n))
    store = eng.GenotypeStore.from_device_payload(payload, n)
    g = store.decode(None).t().contiguous().cpu().numpy()               # [n][m] dosages 0/1/2 (decode is bit-exact)
    assert not np.any(np.abs(g - 1.0 / 3) < 1e-3)
    freq = g.sum(axis=0) / (2.0 * n)
    x = g - 2.0 * freq                                                    # gmatrix.py:53-66
    scale = np.sum(2.0 * freq * (1.0 - freq))
    want = x @ x.T / scale
    want[np.diag_indices(n)] *= 1.001
    results = {}
    for which in ("i8", "f4"):
        monkeypatch.setenv("GMAT_B200_GRM_ENGINE", which)
        kin, s = grm_device(store, eng.ADD)
        got = kin.cpu().numpy()
        assert abs(s - scale) <= 1e-12 * scale
        assert np.abs(got - want).max() <= RTOL * np.abs(want).max(), which
        assert np.array_equal(got, got.T)
        again, _ = grm_device(store, eng.ADD)
        assert np.array_equal(again.cpu().numpy(), got)
        results[which] = got
    # both tensor-core kernels form the same exact integer sums; they differ only in how c_k is rounded
    assert np.abs(results["f4"] - results["i8"]).max() <= 1e-13 * np.abs(want).max()
    monkeypatch.setenv("GMAT_B200_GRM_ENGINE", "f64")
    ref, _ = grm_device(store, eng.ADD)
    assert np.abs(ref.cpu().numpy() - got).max() <= 1e-12 * np.abs(want).max()


def test_grm_engine_selection_and_missing_calls(eng, tmp_path, monkeypatch):
    """auto = FP4 tensor-core kernel for SNP ranges without a missing call, DMMA kernel otherwise (same result as
    forcing f64); forcing a tensor-core kernel on data with missing calls is refused."""
    from gmat_b200.gmatrix.gmatrix import grm_device
    prefix, _ = _dataset(tmp_path, 90, 150, seed=4, missing=0.02)
    store = eng.GenotypeStore(prefix)
    monkeypatch.delenv("GMAT_B200_GRM_ENGINE", raising=False)
    assert eng.grm_engine(store) == "f64"
    clean = [s for s in range(150) if int(store.counts[s, 1]) == 0]
    if clean:
        assert eng.grm_engine(store, clean[0], clean[0] + 1) == "f4"
    a, _ = grm_device(store, eng.ADD)
    monkeypatch.setenv("GMAT_B200_GRM_ENGINE", "f64")
    b, _ = grm_device(store, eng.ADD)
    assert np.array_equal(a.cpu().numpy(), b.cpu().numpy(), equal_nan=True)
    monkeypatch.setenv("GMAT_B200_GRM_ENGINE", "i8")
    with pytest.raises(ValueError):
        grm_device(store, eng.ADD)


def test_screen_chunked_operand_scratch(eng, tmp_path, monkeypatch):
    """The screen packs its A operand for a chunk of I-blocks at a time; force 1-block chunks (8 launches) and
    compare with the single-chunk result and the oracle."""
    n, m = 150, 1000
    prefix, _ = _dataset(tmp_path, n, m, seed=17)
    store = eng.GenotypeStore(prefix)
    rng = np.random.Generator(np.random.PCG64(2))
    pymat = rng.normal(size=(n, 1))
    import torch
    pyd = torch.as_tensor(pymat.ravel().copy()).cuda()
    la = store.lut(eng.ADD, eng.FREQ_CLOOP)
    lst = np.array([i for i in range(m - 1) if i % 5 != 2], dtype=np.int64)
    cut = 6.0
    one = eng.epi_eff(store, la, la, pyd, lst, 0, cut)
    monkeypatch.setenv("GMAT_B200_EFF_SCRATCH_BYTES", str(10 * 16 * 128 * 8))       # one I-block (10 slabs)
    many = eng.epi_eff(store, la, la, pyd, lst, 0, cut)
    key = lambda h: sorted(zip(h["snp_0"].tolist(), h["snp_1"].tolist(), h["eff"].tolist()))
    assert key(one) == key(many) and len(one) > 0
    ref = oracle.screen(prefix, "AA", lst, pymat, eff_cut=cut)
    assert [(a, b) for a, b, _ in key(one)] == [(a, b) for a, b, _ in ref]
    np.testing.assert_allclose([e for _, _, e in key(one)], [e for _, _, e in ref], rtol=RTOL, atol=1e-12)
