"""GPU parity tests of the INT8 tensor-core screen (gmat_b200/csrc/screen_i8.cu): stage 1 (two exact int8 GEMMs of a
two-digit split of Z'Py, FP32 combine, margin) must never lose a pair, stage 2 re-tests the survivors in FP64 -- so
the reported set must EQUAL the set of the FP64 DMMA engine and of the C restatement of the reference, and the
reported effects must be FP64 effects (1e-9 relative is the contract; 1e-12 is what two FP64 sums give)."""
import numpy as np
import pytest

from oracle import oracle

pytestmark = pytest.mark.gpu
RTOL = 1e-9


@pytest.fixture(scope="module")
def eng():
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from gmat_b200 import engine
    return engine


def _passes(eng, store, kind):
    la, ld = store.lut(eng.ADD, eng.FREQ_CLOOP), store.lut(eng.DOM, eng.FREQ_CLOOP)
    A, D = eng.ADD, eng.DOM
    return {"AA": [(la, la, 0, (A, A))], "DD": [(ld, ld, 0, (D, D))],
            "AD": [(la, ld, eng.EFF_GE, (A, D)), (ld, la, eng.EFF_SWAP, (D, A))]}[kind]


def _mid_gap_cut(abs_eff, quantile):
    """A threshold in the middle of a gap between two adjacent |eff| values: the hit set does not depend on the last
    bits of either engine's FP64 sum."""
    v = np.sort(np.asarray(abs_eff))
    k = int(len(v) * quantile)
    k = min(max(k, 1), len(v) - 1)
    while k < len(v) - 1 and v[k] - v[k - 1] <= 1e-9 * v[k]:
        k += 1
    return 0.5 * (v[k] + v[k - 1])


@pytest.mark.parametrize("kind", ["AA", "DD", "AD"])
@pytest.mark.parametrize("n,m", [(131, 77), (300, 600), (520, 300), (64, 1000), (1100, 260)])
def test_i8_screen_equals_f64_engine_and_oracle(eng, tmp_path, kind, n, m):
    import torch
    import synth
    prefix = str(tmp_path / "scr")
    synth.write_dataset(prefix, n, m, seed=3 * n + m)
    store = eng.GenotypeStore(prefix)
    assert store.n_missing() == 0
    rng = np.random.Generator(np.random.PCG64(n + m))
    py = rng.normal(size=n) * rng.choice([1.0, 25.0], size=n)             # a few large entries: scale >> typical
    py -= py.mean()
    pyd = torch.as_tensor(py).cuda()
    lst = np.array([i for i in range(m - 1) if i % 7 != 3], dtype=np.int64)
    for lu, lv, flags, codings in _passes(eng, store, kind):
        every = eng.epi_eff(store, lu, lv, pyd, lst, flags, 0.0)           # FP64 engine, every pair
        for quant in (0.97, 0.9995):
            cut = _mid_gap_cut(np.abs(every["eff"]), quant)
            ref = eng.epi_eff(store, lu, lv, pyd, lst, flags, cut)
            got = eng.epi_eff(store, lu, lv, pyd, lst, flags, cut, codings=codings)
            assert eng.LAST_SCREEN_ENGINE == "i8"
            key = lambda h: h["snp_0"].astype(np.int64) * m + h["snp_1"]
            ref, got = ref[np.argsort(key(ref))], got[np.argsort(key(got))]
            assert len(ref) > 0 and np.array_equal(key(ref), key(got)), (kind, quant, len(ref), len(got))
            np.testing.assert_allclose(got["eff"], ref["eff"], rtol=1e-12, atol=1e-13 * np.abs(ref["eff"]).max())
    # and the C restatement of the reference, direction 1 / the only direction
    want = oracle.screen(prefix, kind, lst, py, eff_cut=cut)
    rows = {}
    for lu, lv, flags, codings in _passes(eng, store, kind):
        h = eng.epi_eff(store, lu, lv, pyd, lst, flags, cut, codings=codings)
        for a, b, e in zip(h["snp_0"].tolist(), h["snp_1"].tolist(), h["eff"].tolist()):
            rows[(a, b)] = e
    assert set(rows) == {(r[0], r[1]) for r in want}
    for a, b, e in want:
        assert abs(rows[(a, b)] - e) <= RTOL * abs(e) + 1e-12


def test_i8_screen_maf_table_small_cuts_and_forced_engines(eng, tmp_path, monkeypatch):
    import torch
    import synth
    n, m = 260, 700
    prefix = str(tmp_path / "maf")
    ds = synth.write_dataset(prefix, n, m, seed=5)
    store = eng.GenotypeStore(prefix)
    rng = np.random.Generator(np.random.PCG64(9))
    py = rng.normal(size=n)
    py -= py.mean()
    pyd = torch.as_tensor(py).cuda()
    la = store.lut(eng.ADD, eng.FREQ_CLOOP)
    lst = np.arange(m - 1, dtype=np.int64)
    g = ds["geno"].astype(np.float64)
    maf = g.sum(axis=1) / (2 * n)
    bins = np.array(np.minimum(maf, 1 - maf) * 20, dtype=np.int32)
    every = eng.epi_eff(store, la, la, pyd, lst, 0, 0.0)
    base = float(np.quantile(np.abs(every["eff"]), 0.98))
    table = base * rng.uniform(0.8, 1.25, size=111)
    ref = eng.epi_eff(store, la, la, pyd, lst, 0, 0.0, table, bins, bins)
    got = eng.epi_eff(store, la, la, pyd, lst, 0, 0.0, table, bins, bins, codings=(eng.ADD, eng.ADD))
    assert eng.LAST_SCREEN_ENGINE == "i8"
    key = lambda h: sorted(zip(h["snp_0"].tolist(), h["snp_1"].tolist()))
    assert len(ref) > 50 and key(ref) == key(got)
    # a cut the margin is not small against: the library falls back to the FP64 kernel by itself
    zero = eng.epi_eff(store, la, la, pyd, lst, 0, 0.0, codings=(eng.ADD, eng.ADD))
    assert eng.LAST_SCREEN_ENGINE == "f64" and len(zero) == len(every)
    # hit buffer smaller than the number of hits: counted, retried with the exact capacity
    small = eng.epi_eff(store, la, la, pyd, lst, 0, base, cap=16, codings=(eng.ADD, eng.ADD))
    assert eng.LAST_SCREEN_ENGINE == "i8" and len(small) == int(np.sum(np.abs(every["eff"]) > base))
    # first SNPs in several chunks (plane-buffer budget of one 128-SNP block): same hits
    whole = eng.epi_eff(store, la, la, pyd, lst, 0, base, codings=(eng.ADD, eng.ADD))
    monkeypatch.setenv("GMAT_B200_SCREEN_SCRATCH_BYTES", "1")
    parts = eng.epi_eff(store, la, la, pyd, lst, 0, base, codings=(eng.ADD, eng.ADD))
    monkeypatch.delenv("GMAT_B200_SCREEN_SCRATCH_BYTES")
    assert eng.LAST_SCREEN_ENGINE == "i8" and key(whole) == key(parts) and len(whole) == len(small)
    assert sorted(whole["eff"].tolist()) == sorted(parts["eff"].tolist())
    monkeypatch.setenv("GMAT_B200_SCREEN_ENGINE", "f64")
    eng.epi_eff(store, la, la, pyd, lst, 0, base, codings=(eng.ADD, eng.ADD))
    assert eng.LAST_SCREEN_ENGINE == "f64"
    # missing calls: the int8 engine runs them too (tripled integer coding, three base-32 digit planes)
    monkeypatch.delenv("GMAT_B200_SCREEN_ENGINE")
    synth.write_dataset(prefix + "m", n, 90, seed=6, missing_rate=0.02)
    st2 = eng.GenotypeStore(prefix + "m")
    l2 = st2.lut(eng.ADD, eng.FREQ_CLOOP)
    r64 = eng.epi_eff(st2, l2, l2, pyd, np.arange(89), 0, base)
    r8 = eng.epi_eff(st2, l2, l2, pyd, np.arange(89), 0, base, codings=(eng.ADD, eng.ADD))
    assert eng.LAST_SCREEN_ENGINE == "i8" and key(r64) == key(r8)


@pytest.mark.parametrize("kind", ["AA", "DD", "AD"])
@pytest.mark.parametrize("n,m,rate", [(131, 77, 0.05), (300, 600, 0.01), (520, 300, 0.002), (1100, 260, 0.02)])
def test_i8_screen_with_missing_calls(eng, tmp_path, kind, n, m, rate):
    """Stores WITH missing calls (what real PLINK files look like): the reference C screen does not impute, code 01 is
    the value 1/3 (_remma_epi_eff_cpu.c:42).  The int8 engine carries 3a in {0,1,3,6} and three base-32 digit planes of
    Z'Py: same hit set as the FP64 DMMA engine and as the reference's own C (oracle/_ref), FP64 effects; one-shot and
    prepared (ScreenOperands, reused across the passes) forms agree."""
    import torch
    import synth
    from oracle import ref_harness
    prefix = str(tmp_path / "miss")
    synth.write_dataset(prefix, n, m, seed=5 * n + m, missing_rate=rate)
    store = eng.GenotypeStore(prefix)
    assert store.n_missing() > 0
    rng = np.random.Generator(np.random.PCG64(n + 3 * m))
    py = rng.normal(size=n) * rng.choice([1.0, 10.0], size=n)
    py -= py.mean()
    pyd = torch.as_tensor(py).cuda()
    lst = np.array([i for i in range(m - 1) if i % 5 != 2], dtype=np.int64)
    ops = eng.ScreenOperands(store, pyd)
    assert ops.tripled == 1
    key = lambda h: h["snp_0"].astype(np.int64) * m + h["snp_1"]
    rows = {}
    for lu, lv, flags, codings in _passes(eng, store, kind):
        every = eng.epi_eff(store, lu, lv, pyd, lst, flags, 0.0)
        cut = _mid_gap_cut(np.abs(every["eff"]), 0.98)
        ref = eng.epi_eff(store, lu, lv, pyd, lst, flags, cut)
        one = eng.epi_eff(store, lu, lv, pyd, lst, flags, cut, codings=codings)
        assert eng.LAST_SCREEN_ENGINE == "i8"
        pre = eng.epi_eff(store, lu, lv, pyd, lst, flags, cut, codings=codings, ops=ops)
        assert eng.LAST_SCREEN_ENGINE == "i8"
        ref, one, pre = ref[np.argsort(key(ref))], one[np.argsort(key(one))], pre[np.argsort(key(pre))]
        assert len(ref) > 0 and np.array_equal(key(ref), key(one)) and np.array_equal(key(ref), key(pre))
        np.testing.assert_allclose(one["eff"], ref["eff"], rtol=1e-12, atol=1e-13 * np.abs(ref["eff"]).max())
        assert np.array_equal(one["eff"], pre["eff"])
        rows[flags] = (cut, ref)
    if kind != "AD" and ref_harness.ref_so_available():
        # the reference's own C on the same files
        cut, ref = rows[0]
        out = str(tmp_path / "ref_screen.txt")
        oracle.ref_screen_file(prefix, kind, lst, py.reshape(-1, 1), cut, out)
        want = {(int(r[0]), int(r[1])) for r in (line.split() for line in open(out).read().splitlines()[1:])}
        assert want == set(zip(ref["snp_0"].tolist(), ref["snp_1"].tolist()))
