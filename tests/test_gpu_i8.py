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
    import synth
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
    diag(u) A diag(u) (lower block-triangle, diagonal blocks halved) to half a unit of the last base-256 digit; the
    four extra rows rebuild r_i = u o (A u) and u o py to FP64 accuracy; the side table holds the sums the expanded
    variance needs."""
    import torch
    from gmat_b200.engine import lib, _p, _stream, check
    n, m = 300, 90
    prefix, _ = _dataset(tmp_path, n, m, seed=3)
    pymat, pvp = _random_model(n, 8)
    store = eng.GenotypeStore(prefix)
    raw = oracle.read_bed_raw(prefix)                                     # [m][n] 0/1/2
    pyd = torch.as_tensor(pymat.ravel().copy()).cuda()
    pvd = torch.as_tensor(pvp).cuda().contiguous()
    # additive SNPs with more 11 than 00 homozygotes are coded by the other allele (2 - dosage) and marked by the sign
    # of asq; the dominance indicator is never flipped
    major = (raw == 2).sum(axis=1) > (raw == 0).sum(axis=1)
    assert 10 < major.sum() < m - 10
    for coding, want, flipped in ((eng.ADD, np.where(major[:, None], 2.0 - raw, raw), major),
                                  (eng.DOM, np.where(raw > 1.5, 0.0, raw), np.zeros(m, dtype=bool))):
        sc = eng.I8Exact(store, store.lut(coding), store.lut(coding), coding, pyd, pvd)
        g8 = sc.g8.cpu().numpy()
        assert np.array_equal(g8[:, :n].astype(np.float64), want) and not g8[:, n:].any()
        g2 = sc.g2.cpu().numpy()
        unpacked = np.stack([(g2 >> (2 * q)) & 3 for q in range(4)], axis=2).reshape(m, -1)
        assert np.array_equal(unpacked[:, :sc.ld], g8)
        asq = sc.asq.cpu().numpy()
        assert np.array_equal(np.abs(asq), (want ** 2).sum(axis=1)) and np.array_equal(np.signbit(asq), flipped)
        offs = sc.offs.cpu().numpy()
        o_ref = -store.lut(coding).cpu().numpy()[:, 0]
        assert np.array_equal(offs, np.where(flipped, 2.0 - o_ref, o_ref))
    U, _ = oracle.coded_pair("AA", oracle.read_plink(prefix))            # [n][m] coded columns
    for planes in (3, 6, 7):
        sc = eng.I8Exact(store, store.lut(eng.ADD), store.lut(eng.ADD), eng.ADD, pyd, pvd, n_planes=planes)
        assert float(sc.amax.item()) == np.abs(pvp).max()
        ilh = np.array([1, m - 2], dtype=np.int64)
        il = torch.as_tensor(ilh).cuda()
        u, r, side, pl_dev = sc.u[0], sc.r[0], sc.side[0], sc.planes[0]
        check(lib.gmat_b200_i8_first(_p(store.bed, "unsigned char *"), n, m, store.row_bytes,
                                     _p(sc.lutU, "double *"), _p(pyd, "double *"), _p(pvd, "double *"),
                                     _p(sc.amax, "double *"), _p(il, "long long *"), 2, planes, _p(u, "double *"),
                                     _p(r, "double *"), _p(side, "double *"), _p(pl_dev, "signed char *"), _stream()),
              "gmat_b200_i8_first")
        check(lib.gmat_b200_i8_split(_p(pvd, "double *"), n, _p(u, "double *"), 2, _p(side, "double *"), planes,
                                     _p(pl_dev, "signed char *"), _stream()), "gmat_b200_i8_split")
        uh = u[:2].cpu().numpy()
        assert np.array_equal(uh, U[:, ilh].T)                           # coded first SNPs, bit for bit
        sd = side[:2].cpu().numpy()
        rh = uh * (uh @ pvp)
        np.testing.assert_allclose(r[:2].cpu().numpy(), rh, rtol=1e-12, atol=1e-14 * np.abs(rh).max())
        np.testing.assert_allclose(sd[:, 1], rh.sum(axis=1), rtol=1e-10, atol=1e-13 * np.abs(rh).sum(axis=1).max())
        np.testing.assert_allclose(sd[:, 2], (uh * pymat.ravel()).sum(axis=1), rtol=1e-10, atol=1e-13)
        scale = sd[:, 0]
        np.testing.assert_allclose(scale, np.abs(uh).max(axis=1) ** 2 * np.abs(pvp).max() / 0.49, rtol=1e-14)
        pl = pl_dev[: 2 * planes * sc.ld * sc.ld].view(2, planes, sc.ld, sc.ld).cpu().numpy().astype(np.float64)
        assert np.abs(pl).max() <= 128
        rec = sum(pl[:, s] * 256.0 ** -(s + 1) for s in reversed(range(planes)))
        xrow = (n + 3) // 4 * 4
        B = np.zeros((2, sc.ld, sc.ld))
        B[:, :n, :n] = uh[:, :, None] * pvp[None] * uh[:, None, :]
        blk = np.arange(sc.ld) // 128
        lower = (blk[:, None] > blk[None, :]) + 0.5 * (blk[:, None] == blk[None, :])
        body = rec.copy()
        body[:, xrow:xrow + 4, :] = 0.0
        err = np.abs(body * scale[:, None, None] - B * lower).max() / scale.max()
        assert err <= 0.51 * 256.0 ** -planes + 1e-15
        # extra rows: value = scale * (hi + 2 * 256^-S * lo)
        wlo = 2.0 * 256.0 ** -planes
        r_rec = sd[:, 3, None] * (rec[:, xrow, :n] + wlo * rec[:, xrow + 1, :n])
        e_rec = sd[:, 4, None] * (rec[:, xrow + 2, :n] + wlo * rec[:, xrow + 3, :n])
        tol = 4e-16 + 256.0 ** (-2 * planes)                             # 2S digits: beyond FP64 from S = 4 on
        assert np.all(np.abs(r_rec - rh) <= tol * np.abs(rh).max(axis=1, keepdims=True) / 0.49)
        eh = uh * pymat.ravel()
        assert np.all(np.abs(e_rec - eh) <= tol * np.abs(eh).max(axis=1, keepdims=True) / 0.49)
        assert not rec[:, xrow:xrow + 4, n:].any() and not rec[:, xrow + 4:, :].any() and not rec[:, n:xrow, :].any()


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
    # values on the intersection of the two hit sets, unconditionally (the border band only excuses membership)
    common, ir, ig = np.intersect1d(key(ref), key(got), return_indices=True)
    assert len(common) >= len(ref) - border.sum()
    np.testing.assert_allclose(got["eff"][ig], ref["eff"][ir], rtol=RTOL, atol=1e-12)
    np.testing.assert_allclose(got["var"][ig], ref["var"][ir], rtol=RTOL)
    np.testing.assert_allclose(got["chi"][ig], ref["chi"][ir], rtol=4 * RTOL)


def test_i8_monomorphic_snps_and_small_batches(eng, tmp_path, monkeypatch):
    """A second SNP whose code is the same for everybody has w == 0: eff = var = 0 and chi = NaN, as in the
    reference (0/0), never a hit; a first SNP of that kind too.  Also: plane buffer budget of one first SNP per
    launch gives the same result as one launch."""
    import torch
    import synth
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
    # every operand kernel sums in an order that does not depend on the batch shape: same bits
    assert np.array_equal(got1["var"], got["var"]) and np.array_equal(got1["eff"], got["eff"])
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
    import synth
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


@pytest.mark.parametrize("n,m", [(257, 300), (400, 270), (520, 130), (900, 515), (1100, 64), (252, 70), (256, 70),
                                 (381, 40), (512, 300)])
def test_i8_kernel_variants_are_bit_identical(eng, tmp_path, n, m):
    """The kernel that issues only the plane rows that can be non-zero (variant 3, the default: last row tile trimmed
    to the extra rows, zero block right of the diagonal skipped) must give the SAME BITS as full 256-row MMAs
    (variant 2): the integer sums are exact, the FP64 finish is the same expression in the same order.
    Shapes put n just past a 256-row tile (N = 16), past the half tile (half-slab N = 16), in between, and at
    n = 252 / 256 / 512, where the four extra rows open a row tile of their own."""
    import torch
    prefix, _ = _dataset(tmp_path, n, m, seed=n + m)
    pymat, pvp = _random_model(n, n * 3 + m)
    store = eng.GenotypeStore(prefix)
    lu = store.lut(eng.ADD)
    pyd = torch.as_tensor(pymat.ravel().copy()).cuda()
    pvd = torch.as_tensor(pvp).cuda().contiguous()
    first = np.arange(0, m - 1, 2)
    res = {}
    for variant in (3, 2):
        sc = eng.I8Exact(store, lu, lu, eng.ADD, pyd, pvd)
        sc.wide = variant
        h, q = sc.scan(first, eng.PAIRS_UPPER, -1.0, q_debug=True)
        key = h["snp_0"].astype(np.int64) * m + h["snp_1"]
        res[variant] = (h[np.argsort(key)], q.cpu().numpy())
    for col in ("snp_0", "snp_1", "var", "eff"):
        assert np.array_equal(res[3][0][col], res[2][0][col]), col
    # a_j' B_i a_j itself (only the entries of scanned pairs j > i are written)
    q3, q2 = res[3][1], res[2][1]
    for r, i in enumerate(first):
        assert np.array_equal(q3[r, i + 1:], q2[r, i + 1:])
    # and against the oracle, like every other engine test
    U, V = oracle.coded_pair("AA", oracle.read_plink(prefix))
    h = res[3][0]
    pairs = np.stack([h["snp_0"], h["snp_1"]], axis=1)
    eff, var, _, _ = oracle.exact_pairs(U, V, pairs, pymat, pvp)
    np.testing.assert_allclose(h["eff"], eff, rtol=RTOL, atol=1e-11)
    np.testing.assert_allclose(h["var"], var, rtol=RTOL, atol=1e-13 * np.abs(var).max())


@pytest.mark.parametrize("planes", [4, 5, 6, 7])
def test_i8_weak_pairs_are_rechecked_in_fp64(eng, tmp_path, planes):
    """The a-posteriori guard: with few digit planes (or two very rare variants) the split cannot resolve a small
    variance; such pairs are re-evaluated by the FP64 pair kernel, so the result holds the 1e-9 tolerance for ANY
    number of planes, and every pair is still reported exactly once.  Rare first / second SNPs (3 carriers) make
    var tiny against max|u|^2 max|A|."""
    import torch
    import synth
    n, m = 400, 120
    g = synth.random_genotypes(n, m, 5, 0.0)
    for snp in (2, 9, 50, 51, 100):                                     # singletons / rare variants
        g[snp, :] = 0
        g[snp, [snp, snp + 7, 2 * snp + 1]] = 1
    for snp in (30, 31, 77):                                            # the same with the COUNTED allele the common
        g[snp, :] = 2                                                   # one: coded by the other allele in the engine
        g[snp, [snp, snp + 5, 3 * snp + 1]] = 1
        g[snp, snp + 11] = 0
    prefix, _ = _dataset(tmp_path, n, m, seed=1, geno=g)
    pymat, pvp = _random_model(n, 12)
    store = eng.GenotypeStore(prefix)
    lu = store.lut(eng.ADD)
    pyd = torch.as_tensor(pymat.ravel().copy()).cuda()
    pvd = torch.as_tensor(pvp).cuda().contiguous()
    first = np.arange(m - 1)
    sc = eng.I8Exact(store, lu, lu, eng.ADD, pyd, pvd, n_planes=planes)
    got = sc.scan(first, eng.PAIRS_UPPER, -1.0)
    assert len(got) == m * (m - 1) // 2
    key = got["snp_0"].astype(np.int64) * m + got["snp_1"]
    assert len(np.unique(key)) == len(got)
    got = got[np.argsort(key)]
    U, V = oracle.coded_pair("AA", oracle.read_plink(prefix))
    pairs = np.stack([got["snp_0"], got["snp_1"]], axis=1)
    eff, var, _, _ = oracle.exact_pairs(U, V, pairs, pymat, pvp)
    np.testing.assert_allclose(got["var"], var, rtol=RTOL, atol=0)
    np.testing.assert_allclose(got["eff"], eff, rtol=RTOL, atol=1e-12)
    if planes <= 5:
        assert sc.last_rechecked > 0                                    # the guard did fire
    # a tiny re-check buffer: the scan is repeated once with the exact capacity
    sc.RECHECK_CAP = 1
    again = sc.scan(first, eng.PAIRS_UPPER, -1.0)
    again = again[np.argsort(again["snp_0"].astype(np.int64) * m + again["snp_1"])]
    assert np.array_equal(again["var"], got["var"])
    # thresholded scan: same hit set as thresholding the full result
    cut = float(np.nanmedian(got["chi"]))
    hit = sc.scan(first, eng.PAIRS_UPPER, cut)
    assert len(hit) == int(np.sum(got["chi"] > cut))
