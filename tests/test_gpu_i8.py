"""GPU parity tests of the INT8 tensor-core engine of the exact scan (gmat_b200/csrc/ozaki.cu, tcgen05 kind::i8):
the same oracle and the same 1e-9 relative tolerance as the FP64 DMMA kernel, plus the properties the digit-plane
split itself must have (integer operands bit-exact, the planes reconstruct diag(u) Z'PZ diag(u) to the last bit of
the split, results do not depend on which engine ran)."""
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


def _dataset(tmp_path, n, m, seed, missing=0.0, geno=None):
    from gmat_b200 import synth
    prefix = str(tmp_path / ("syn_%d_%d" % (n, m)))
    return prefix, synth.write_dataset(prefix, n, m, seed=seed, missing_rate=missing, geno=geno)


def _random_model(n, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    b = rng.normal(size=(n, n + 3))
    a = b @ b.T / n
    c = np.eye(n) - np.ones((n, n)) / n
    a = c @ a @ c
    py = c @ rng.normal(size=n)
    return py.reshape(-1, 1), 0.5 * (a + a.T)


def _codings(eng, kind):
    return {"AA": (eng.ADD, eng.ADD), "AD": (eng.ADD, eng.DOM), "DD": (eng.DOM, eng.DOM)}[kind]


@pytest.mark.parametrize("kind", ["AA", "AD", "DD"])
@pytest.mark.parametrize("n,m", [(128, 40), (131, 77), (300, 150), (17, 9), (257, 129), (520, 260)])
def test_i8_scan_vs_oracle(eng, tmp_path, kind, n, m):
    """Every pair of a ragged first-SNP list (chi_cut < 0): eff, var within 1e-9 of the numpy oracle; n and m off
    every tile boundary (128 second SNPs, 256 plane rows, 128 ids per slab)."""
    import torch
    prefix, _ = _dataset(tmp_path, n, m, seed=n * 7 + m)
    pymat, pvp = _random_model(n, n + m)
    store = eng.GenotypeStore(prefix)
    cu, cv = _codings(eng, kind)
    lu, lv = store.lut(cu), store.lut(cv)
    pyd = torch.as_tensor(pymat.ravel().copy()).cuda()
    pvd = torch.as_tensor(pvp).cuda().contiguous()
    U, V = oracle.coded_pair(kind, oracle.read_plink(prefix))
    all_pairs = kind == "AD"
    lst = [i for i in range(0, m - (0 if all_pairs else 1)) if i % 3 != 1]
    scanner = eng.I8Exact(store, lu, lv, cv, pyd, pvd)
    hits = scanner.scan(np.array(lst), eng.PAIRS_ALL if all_pairs else eng.PAIRS_UPPER, -1.0)
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
    np.testing.assert_allclose(ge[:, 1], we[:, 1], rtol=RTOL, atol=1e-13 * np.abs(we[:, 1]).max())


def test_i8_operands_exact(eng, tmp_path):
    """Integer codes are the reference's dosage / heterozygote indicator bit for bit; the digit planes rebuild
    diag(u) A diag(u) (lower block-triangle, diagonal blocks halved) to 128^-S of its largest entry."""
    import torch
    from gmat_b200.engine import lib, _p, _stream, check
    n, m = 300, 90
    prefix, _ = _dataset(tmp_path, n, m, seed=3)
    pymat, pvp = _random_model(n, 8)
    store = eng.GenotypeStore(prefix)
    raw = oracle.read_bed_raw(prefix)                                     # [m][n] 0/1/2
    pyd = torch.as_tensor(pymat.ravel().copy()).cuda()
    pvd = torch.as_tensor(pvp).cuda().contiguous()
    for coding, want in ((eng.ADD, raw), (eng.DOM, np.where(raw > 1.5, 0.0, raw))):
        sc = eng.I8Exact(store, store.lut(coding), store.lut(coding), coding, pyd, pvd)
        g8 = sc.g8.cpu().numpy()
        assert np.array_equal(g8[:, :n].astype(np.float64), want) and not g8[:, n:].any()
    for planes in (3, 6, 8):
        sc = eng.I8Exact(store, store.lut(eng.ADD), store.lut(eng.ADD), eng.ADD, pyd, pvd, n_planes=planes)
        il = torch.as_tensor(np.array([1, m - 2], dtype=np.int64)).cuda()
        u = sc._coded_first(il)
        scale = (u.abs().amax(dim=1) ** 2 * (sc.amax / 0.99)).contiguous()
        check(lib.gmat_b200_i8_split(_p(pvd, "double *"), n, _p(u, "double *"), 2, _p(scale, "double *"), planes,
                                     _p(sc.planes[0], "signed char *"), _stream()), "gmat_b200_i8_split")
        pl = sc.planes[0][: 2 * planes * sc.ld * sc.ld].view(2, planes, sc.ld, sc.ld).cpu().numpy().astype(np.float64)
        rec = sum(pl[:, s] * 128.0 ** -(s + 1) for s in reversed(range(planes))) * scale.cpu().numpy()[:, None, None]
        B = np.zeros((2, sc.ld, sc.ld))
        uh = u.cpu().numpy()
        B[:, :n, :n] = uh[:, :, None] * pvp[None] * uh[:, None, :]
        blk = np.arange(sc.ld) // 128
        lower = (blk[:, None] > blk[None, :]) + 0.5 * (blk[:, None] == blk[None, :])
        err = np.abs(rec - B * lower).max() / scale.cpu().numpy().max()
        assert err <= 0.51 * 128.0 ** -planes + 1e-15
        assert np.abs(pl).max() <= 127


def test_i8_and_f64_engines_report_the_same_hits(eng, tmp_path):
    """Thresholded scan (what the drop-in runs): identical pair set, statistics within 1e-9 of the DMMA kernel."""
    import torch
    n, m = 700, 1500
    prefix, _ = _dataset(tmp_path, n, m, seed=21)
    pymat, pvp = _random_model(n, 4)
    store = eng.GenotypeStore(prefix)
    lu = store.lut(eng.ADD)
    pyd = torch.as_tensor(pymat.ravel().copy()).cuda()
    pvd = torch.as_tensor(pvp).cuda().contiguous()
    packed = eng.pack_pvp(pvd)
    first = np.arange(0, m - 1, 7)
    ref = eng.epi_exact_scan(store, lu, lu, pyd, packed, first, eng.PAIRS_UPPER, 6.0)
    got = eng.I8Exact(store, lu, lu, eng.ADD, pyd, pvd).scan(first, eng.PAIRS_UPPER, 6.0)
    key = lambda h: h["snp_0"].astype(np.int64) * m + h["snp_1"]
    ref, got = ref[np.argsort(key(ref))], got[np.argsort(key(got))]
    border = np.abs(ref["chi"] / 6.0 - 1) < 1e-7
    assert len(ref) > 1000
    assert set(key(ref)[~border]) <= set(key(got)) and len(got) <= len(ref) + border.sum()
    if len(ref) == len(got):
        assert np.array_equal(key(ref), key(got))
        np.testing.assert_allclose(got["eff"], ref["eff"], rtol=RTOL, atol=1e-12)
        np.testing.assert_allclose(got["var"], ref["var"], rtol=RTOL)
        np.testing.assert_allclose(got["chi"], ref["chi"], rtol=4 * RTOL)


def test_i8_monomorphic_snps_and_small_batches(eng, tmp_path, monkeypatch):
    """A second SNP whose code is the same for everybody has w == 0: eff = var = 0 and chi = NaN, as in the
    reference (0/0), never a hit; a first SNP of that kind too.  Also: plane buffer budget of one first SNP per
    launch gives the same result as one launch."""
    import torch
    from gmat_b200 import synth
    n, m = 140, 60
    g = synth.random_genotypes(n, m, 77, 0.0)
    g[5, :] = 0
    g[17, :] = 2
    g[30, :] = 1
    prefix, _ = _dataset(tmp_path, n, m, seed=1, geno=g)
    pymat, pvp = _random_model(n, 2)
    store = eng.GenotypeStore(prefix)
    lu = store.lut(eng.ADD)
    pyd = torch.as_tensor(pymat.ravel().copy()).cuda()
    pvd = torch.as_tensor(pvp).cuda().contiguous()
    packed = eng.pack_pvp(pvd)
    first = np.arange(m - 1)
    ref = eng.epi_exact_scan(store, lu, lu, pyd, packed, first, eng.PAIRS_UPPER, -1.0)
    sc = eng.I8Exact(store, lu, lu, eng.ADD, pyd, pvd)
    got = sc.scan(first, eng.PAIRS_UPPER, -1.0)
    monkeypatch.setattr(sc, "batch", 1)
    got1 = sc.scan(first, eng.PAIRS_UPPER, -1.0)
    key = lambda h: h["snp_0"].astype(np.int64) * m + h["snp_1"]
    ref, got, got1 = ref[np.argsort(key(ref))], got[np.argsort(key(got))], got1[np.argsort(key(got1))]
    assert np.array_equal(key(ref), key(got)) and np.array_equal(key(ref), key(got1))
    # (the two O(n) terms come from library DGEMMs whose summation order depends on the batch shape)
    np.testing.assert_allclose(got1["var"], got["var"], rtol=1e-12)
    np.testing.assert_allclose(got1["eff"], got["eff"], rtol=1e-12, atol=1e-13)
    dead = np.isin(ref["snp_0"], [5, 17, 30]) | np.isin(ref["snp_1"], [5, 17, 30])
    assert dead.sum() > 0
    assert np.all(got["eff"][dead] == 0) and np.all(got["var"][dead] == 0) and np.all(np.isnan(got["chi"][dead]))
    assert np.all(ref["var"][dead] == 0) and np.all(np.isnan(ref["chi"][dead]))
    np.testing.assert_allclose(got["eff"][~dead], ref["eff"][~dead], rtol=RTOL, atol=1e-12)
    np.testing.assert_allclose(got["var"][~dead], ref["var"][~dead], rtol=RTOL)
    assert len(sc.scan(first, eng.PAIRS_UPPER, 1e-3)) == int(np.sum(ref["chi"] > 1e-3))
    # hit buffer too small: the counter keeps counting, the scan is repeated once with the exact capacity
    small = sc.scan(first, eng.PAIRS_UPPER, -1.0, cap=100)
    assert len(small) == len(ref) and len(set(key(small).tolist())) == len(ref)
    # the single-CTA kernel variant gives the same numbers as the CTA-pair kernel
    sc0 = eng.I8Exact(store, lu, lu, eng.ADD, pyd, pvd)
    sc0.wide = 0
    got0 = sc0.scan(first, eng.PAIRS_UPPER, -1.0)
    got0 = got0[np.argsort(key(got0))]
    assert np.array_equal(key(got0), key(got))
    np.testing.assert_allclose(got0["var"], got["var"], rtol=1e-12)


def test_engine_selection(eng, tmp_path, monkeypatch):
    """auto = i8 = the int8 tensor-core engine, f64 = the DMMA kernel, anything else is an error."""
    p0, _ = _dataset(tmp_path, 40, 12, seed=5)
    s0 = eng.GenotypeStore(p0)
    monkeypatch.delenv("GMAT_B200_EXACT_ENGINE", raising=False)
    assert eng.exact_engine(s0) == "i8"
    monkeypatch.setenv("GMAT_B200_EXACT_ENGINE", "f64")
    assert eng.exact_engine(s0) == "f64"
    monkeypatch.setenv("GMAT_B200_EXACT_ENGINE", "bogus")
    with pytest.raises(ValueError):
        eng.exact_engine(s0)


@pytest.mark.parametrize("kind", ["AA", "AD"])
def test_i8_missing_calls_are_nan_like_the_reference(eng, tmp_path, kind):
    """A missing call makes the coded column NaN in the reference's numpy path, so every statistic of a pair that
    involves such a SNP is NaN (never reported); all other pairs are untouched.  Same NaN pattern and same numbers
    as the DMMA kernel and the oracle."""
    import torch
    from gmat_b200 import synth
    n, m = 150, 70
    g = synth.random_genotypes(n, m, 9, 0.0)
    for snp, ids in ((3, [0, 7]), (20, [149]), (41, [5, 6, 77]), (69, [10])):
        g[snp, ids] = -1
    prefix, _ = _dataset(tmp_path, n, m, seed=2, geno=g)
    pymat, pvp = _random_model(n, 6)
    store = eng.GenotypeStore(prefix)
    assert store.n_missing() == 7
    cu, cv = _codings(eng, kind)
    lu, lv = store.lut(cu), store.lut(cv)
    pyd = torch.as_tensor(pymat.ravel().copy()).cuda()
    pvd = torch.as_tensor(pvp).cuda().contiguous()
    packed = eng.pack_pvp(pvd)
    space = eng.PAIRS_ALL if kind == "AD" else eng.PAIRS_UPPER
    first = np.arange(m if kind == "AD" else m - 1)
    ref = eng.epi_exact_scan(store, lu, lv, pyd, packed, first, space, -1.0)
    got = eng.I8Exact(store, lu, lv, cv, pyd, pvd).scan(first, space, -1.0)
    key = lambda h: h["snp_0"].astype(np.int64) * m + h["snp_1"]
    ref, got = ref[np.argsort(key(ref))], got[np.argsort(key(got))]
    assert np.array_equal(key(ref), key(got))
    dead = np.isin(ref["snp_0"], [3, 20, 41, 69]) | np.isin(ref["snp_1"], [3, 20, 41, 69])
    for col in ("eff", "var", "chi"):
        assert np.all(np.isnan(ref[col][dead])) and np.all(np.isnan(got[col][dead]))
        assert not np.isnan(got[col][~dead]).any()
    np.testing.assert_allclose(got["eff"][~dead], ref["eff"][~dead], rtol=RTOL, atol=1e-12)
    np.testing.assert_allclose(got["var"][~dead], ref["var"][~dead], rtol=RTOL)
    # thresholded: no pair with a missing call is ever reported
    hits = eng.I8Exact(store, lu, lv, cv, pyd, pvd).scan(first, space, 0.5)
    assert len(hits) == int(np.nansum(ref["chi"] > 0.5)) and not (np.isin(hits["snp_0"], [3, 20, 41, 69]).any()
                                                                  or np.isin(hits["snp_1"], [3, 20, 41, 69]).any())
    U, V = oracle.coded_pair(kind, oracle.read_plink(prefix))
    js, e, v, c, _ = oracle.exact_first_snp(U, V, 5, pymat, pvp, kind == "AD")
    mine = got[got["snp_0"] == 5]
    assert np.array_equal(mine["snp_1"], js)
    assert np.array_equal(np.isnan(v), np.isnan(mine["var"]))
    ok = ~np.isnan(v)
    np.testing.assert_allclose(mine["var"][ok], v[ok], rtol=RTOL)
    np.testing.assert_allclose(mine["eff"][ok], e[ok], rtol=RTOL, atol=1e-11)


@pytest.mark.parametrize("n,m", [(257, 300), (400, 270), (520, 130), (900, 515), (1100, 64)])
def test_i8_kernel_variants_are_bit_identical(eng, tmp_path, n, m):
    """The CTA-pair kernels that issue only the plane rows that can be non-zero (last row tile trimmed to n, zero
    block right of the diagonal skipped) -- variant 4 (the default: two digit planes per genotype tile) and variant 3
    (one plane at a time) -- must give the SAME BITS as the full-tile CTA-pair kernel (2) and the single-CTA kernel
    (0): the integer sums are exact, the FP64 finish is the same expression in the same order.
    Shapes put n just past a 256-row tile (N = 16), past the half tile (half-slab N = 16) and in between."""
    import torch
    prefix, _ = _dataset(tmp_path, n, m, seed=n + m)
    pymat, pvp = _random_model(n, n * 3 + m)
    store = eng.GenotypeStore(prefix)
    lu = store.lut(eng.ADD)
    pyd = torch.as_tensor(pymat.ravel().copy()).cuda()
    pvd = torch.as_tensor(pvp).cuda().contiguous()
    first = np.arange(0, m - 1, 2)
    res = {}
    for variant in (4, 3, 2, 0):
        sc = eng.I8Exact(store, lu, lu, eng.ADD, pyd, pvd)
        sc.wide = variant
        h, q = sc.scan(first, eng.PAIRS_UPPER, -1.0, q_debug=True)
        key = h["snp_0"].astype(np.int64) * m + h["snp_1"]
        res[variant] = (h[np.argsort(key)], q.cpu().numpy())
    for variant in (4, 2, 0):
        assert np.array_equal(res[3][0]["snp_0"], res[variant][0]["snp_0"])
        assert np.array_equal(res[3][0]["snp_1"], res[variant][0]["snp_1"])
        assert np.array_equal(res[3][0]["var"], res[variant][0]["var"]), "variant %d: var differs" % variant
        assert np.array_equal(res[3][0]["eff"], res[variant][0]["eff"])
    # a_j' B_i a_j itself (only the entries of scanned pairs j > i are written by every variant)
    q4, q3, q2 = res[4][1], res[3][1], res[2][1]
    for r, i in enumerate(first):
        assert np.array_equal(q3[r, i + 1:], q2[r, i + 1:]) and np.array_equal(q4[r, i + 1:], q2[r, i + 1:])
    # and against the oracle, like every other engine test
    U, V = oracle.coded_pair("AA", oracle.read_plink(prefix))
    h = res[3][0]
    pairs = np.stack([h["snp_0"], h["snp_1"]], axis=1)
    eff, var, _, _ = oracle.exact_pairs(U, V, pairs, pymat, pvp)
    np.testing.assert_allclose(h["eff"], eff, rtol=RTOL, atol=1e-11)
    np.testing.assert_allclose(h["var"], var, rtol=RTOL, atol=1e-13 * np.abs(var).max())
