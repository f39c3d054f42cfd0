"""GPU parity tests of the pair-list form of the int8 tensor-core engine (gmat_b200/csrc/ozaki.cu "pair lists",
engine.I8Pairs; remma_epiAA_pair.py:66-87): digit planes of Z'PZ itself, four integer operand rows per pair, var / eff
by polarisation.  Same oracle and the same 1e-9 relative tolerance as the FP64 pair kernel it stands in for."""
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


def _model(n, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    b = rng.normal(size=(n, n + 3))
    a = b @ b.T / n
    c = np.eye(n) - np.ones((n, n)) / n
    a = c @ a @ c
    return (c @ rng.normal(size=n)).reshape(-1, 1), 0.5 * (a + a.T)


def _codings(eng, kind):
    return {"AA": (eng.ADD, eng.ADD), "AD": (eng.ADD, eng.DOM), "DD": (eng.DOM, eng.DOM)}[kind]


def _setup(eng, tmp_path, n, m, seed, kind, geno=None, centred=True):
    import torch
    import synth
    prefix = str(tmp_path / ("syn_%d_%d" % (n, m)))
    synth.write_dataset(prefix, n, m, seed=seed, missing_rate=0.0, geno=geno)
    pymat, pvp = _model(n, seed + 1)
    if not centred:                       # a model without an intercept: Z'PZ 1 != 0, the terms with 1 matter
        rng = np.random.Generator(np.random.PCG64(seed + 2))
        b = rng.normal(size=(n, n + 3))
        pvp = b @ b.T / n
        pvp = 0.5 * (pvp + pvp.T)
        pymat = rng.normal(size=(n, 1))
    store = eng.GenotypeStore(prefix)
    cu, cv = _codings(eng, kind)
    pyd = torch.as_tensor(pymat.ravel().copy()).cuda()
    pvd = torch.as_tensor(pvp).cuda().contiguous()
    U, V = oracle.coded_pair(kind, oracle.read_plink(prefix))
    return store, cu, cv, pyd, pvd, U, V, pymat, pvp


@pytest.mark.parametrize("kind", ["AA", "AD", "DD"])
@pytest.mark.parametrize("n,m,centred", [(131, 60, True), (300, 150, True), (257, 90, False), (520, 200, False)])
def test_i8_pairs_vs_oracle(eng, tmp_path, kind, n, m, centred):
    """Random pairs (both orders, repeated entries, i == j): eff, var within 1e-9 of the numpy oracle and of the FP64
    pair kernel, in list order; several chunks."""
    store, cu, cv, pyd, pvd, U, V, pymat, pvp = _setup(eng, tmp_path, n, m, 3 * n + m, kind, centred=centred)
    rng = np.random.Generator(np.random.PCG64(n))
    pairs = rng.integers(0, m, size=(1500, 2))
    pairs[:40, 1] = pairs[:40, 0]                                     # i == j
    pairs[40:80] = pairs[80:120]                                      # repeated entries
    eng_p = eng.I8Pairs(store, cu, cv, store.lut(cu), store.lut(cv), pyd, pvd)
    eng_p.CHUNK = 400                                                 # four launches
    got = eng_p.stats(pairs)
    assert np.array_equal(got["snp_0"], pairs[:, 0]) and np.array_equal(got["snp_1"], pairs[:, 1])
    eff, var, _, _ = oracle.exact_pairs(U, V, pairs, pymat, pvp)
    # (a pair whose variance is 1e-9 of the typical one -- i == j of a near-constant SNP -- carries the FP64 rounding
    # of the typical magnitude in the oracle as well: absolute floor 1e-13 max|var|, as in the scan's tests)
    np.testing.assert_allclose(got["var"], var, rtol=RTOL, atol=1e-13 * np.abs(var).max())
    np.testing.assert_allclose(got["eff"], eff, rtol=RTOL, atol=1e-12 * np.abs(eff).max())
    ref = eng.epi_exact_pairs(store, store.lut(cu), store.lut(cv), pyd, eng.pack_pvp(pvd), pairs)
    np.testing.assert_allclose(got["var"], ref["var"], rtol=RTOL, atol=1e-13 * np.abs(var).max())
    big = np.abs(var) > 1e-6 * np.abs(var).max()
    np.testing.assert_allclose(got["chi"][big], ref["chi"][big], rtol=4 * RTOL)
    assert eng_p.last_rechecked < len(pairs) // 10                    # the tensor cores did the work


@pytest.mark.parametrize("planes", [4, 6, 7])
def test_i8_pairs_rare_and_monomorphic_snps(eng, tmp_path, planes):
    """Rare variants (tiny var against max|Z'PZ| sum x^2) are re-evaluated in FP64 when the planes cannot resolve them;
    a monomorphic SNP gives eff = var = 0 and chi = NaN like the FP64 kernel and the reference (0/0)."""
    import synth
    n, m = 400, 120
    g = synth.random_genotypes(n, m, 5, 0.0)
    for snp in (2, 9, 50):
        g[snp, :] = 0
        g[snp, [snp, snp + 7, 2 * snp + 1]] = 1
    for snp in (30, 77):
        g[snp, :] = 2
        g[snp, [snp, snp + 5]] = 1
    g[11, :] = 1                                                      # monomorphic
    store, cu, cv, pyd, pvd, U, V, pymat, pvp = _setup(eng, tmp_path, n, m, 1, "AA", geno=g)
    pairs = np.array([(i, j) for i in range(m) for j in range(i + 1, m)][::3])
    eng_p = eng.I8Pairs(store, cu, cv, store.lut(cu), store.lut(cv), pyd, pvd, n_planes=planes)
    got = eng_p.stats(pairs)
    ref = eng.epi_exact_pairs(store, store.lut(cu), store.lut(cv), pyd, eng.pack_pvp(pvd), pairs)
    dead = (pairs == 11).any(axis=1)
    assert dead.sum() > 10
    assert np.all(got["eff"][dead] == 0) and np.all(got["var"][dead] == 0) and np.all(np.isnan(got["chi"][dead]))
    eff, var, _, _ = oracle.exact_pairs(U, V, pairs[~dead], pymat, pvp)
    np.testing.assert_allclose(got["var"][~dead], var, rtol=RTOL, atol=1e-13 * np.abs(var).max())
    np.testing.assert_allclose(got["eff"][~dead], eff, rtol=RTOL, atol=1e-12)
    np.testing.assert_allclose(got["var"][~dead], ref["var"][~dead], rtol=RTOL, atol=1e-13 * np.abs(var).max())
    assert eng_p.last_rechecked >= dead.sum()
    if planes <= 4:
        assert eng_p.last_rechecked > dead.sum()                      # the guard fired on rare pairs too


def test_pair_engine_selection(eng, tmp_path, monkeypatch):
    store, *_ = _setup(eng, tmp_path, 140, 50, 9, "AA")
    monkeypatch.delenv("GMAT_B200_PAIR_ENGINE", raising=False)
    assert eng.pair_engine(store, 10) == "f64" and eng.pair_engine(store, 100000) == "f64"      # n = 140: DMMA
    monkeypatch.setattr(eng.I8Pairs, "MIN_IDS", 100)                  # (2000 by default)
    assert eng.pair_engine(store, 10) == "f64" and eng.pair_engine(store, 100000) == "i8"
    monkeypatch.setenv("GMAT_B200_PAIR_ENGINE", "f64")
    assert eng.pair_engine(store, 100000) == "f64"
    monkeypatch.setenv("GMAT_B200_PAIR_ENGINE", "i8")
    assert eng.pair_engine(store, 10) == "i8"
    monkeypatch.setenv("GMAT_B200_PAIR_ENGINE", "bogus")
    with pytest.raises(ValueError):
        eng.pair_engine(store, 10)
