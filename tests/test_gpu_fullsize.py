"""Parity at BASELINE.json's full shapes, where a full CPU scan is infeasible (hours to days):
 * every statistic the GPU reports for a sample of pairs is recomputed by the numpy oracle on the same bytes
   (eff, var within 1e-9 relative -- the tolerance north_star states);
 * size-independent properties: scan and pair-list kernels agree on every reported hit, eff is linear in py and
   var does not depend on it, the screen's hit set on a first-SNP subset equals the oracle's, the GRM is
   symmetric and equals X X'/scale on sampled entries.
Genotypes are generated in HBM (synth.device_payload) and the same bytes are decoded on the host for the oracle.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-9          # north_star: effects and variances within 1e-9 relative


@pytest.fixture(scope="module")
def eng():
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from gmat_b200 import engine
    return engine


def _host_columns(payload_host, n, snps, coding):
    """Coded columns [n][len(snps)] of the selected SNPs from the raw 2-bit payload, numpy, restating
    _read_plink_bed.c:36-37 (code -> dosage) and remma_epiAA.py:59-61 / remma_epiDD.py:59-65 (coding)."""
    rows = payload_host[np.asarray(snps)][:, : (n + 3) // 4]
    codes = np.stack([(rows >> (2 * q)) & 3 for q in range(4)], axis=2).reshape(len(snps), -1)[:, :n]
    g = np.array([0.0, np.nan, 1.0, 2.0])[codes]                      # no missing codes in these data sets
    assert not np.isnan(g).any()
    freq = g.sum(axis=1, keepdims=True) / (2 * n)
    if coding == "A":
        x = g - 2 * freq
    else:
        h = np.where(g > 1.5, 0.0, g)
        x = h - 2 * freq * (1 - freq)
    return np.ascontiguousarray(x.T)


def _model(n, seed, torch, dev):
    gen = torch.Generator(device=dev)
    gen.manual_seed(seed)
    b = torch.randn((n, n + 8), dtype=torch.float64, device=dev, generator=gen)
    pvp = b @ b.T / n
    pvp = pvp - pvp.mean(dim=0, keepdim=True)
    pvp = pvp - pvp.mean(dim=1, keepdim=True)
    pvp = 0.5 * (pvp + pvp.T)
    py = torch.randn((n,), dtype=torch.float64, device=dev, generator=gen)
    py = py - py.mean()
    return pvp, py


def _oracle_pairs(payload_host, n, pairs, cu, cv, py, pvp):
    """eff = w'py, var = w' pvp w for explicit pairs (remma_epiAA_pair.py:78-84), numpy."""
    su = sorted(set(int(a) for a in pairs[:, 0]))
    sv = sorted(set(int(b) for b in pairs[:, 1]))
    U = _host_columns(payload_host, n, su, cu)
    V = _host_columns(payload_host, n, sv, cv)
    iu = {s: k for k, s in enumerate(su)}
    iv = {s: k for k, s in enumerate(sv)}
    W = U[:, [iu[int(a)] for a in pairs[:, 0]]] * V[:, [iv[int(b)] for b in pairs[:, 1]]]
    eff = W.T @ py
    var = np.sum(W * (pvp @ W), axis=0)
    return eff, var


@pytest.mark.parametrize("n,m,kind", [(2000, 20000, "AA"), (5000, 50000, "AA"), (5000, 50000, "DD"),
                                      (5000, 50000, "AD")])
def test_exact_scan_full_shape_sampled_against_oracle(eng, n, m, kind):
    import torch
    import synth
    dev = eng.current_device()
    payload = synth.device_payload(n, m, 20201013 + n, dev, eng.row_stride(n))
    store = eng.GenotypeStore.from_device_payload(payload, n)
    host = payload.cpu().numpy()
    pvp, py = _model(n, 5, torch, dev)
    packed = eng.pack_pvp(pvp)
    cu, cv = kind[0], kind[1]
    lu = store.lut(eng.ADD if cu == "A" else eng.DOM)
    lv = store.lut(eng.ADD if cv == "A" else eng.DOM)
    space = eng.PAIRS_ALL if kind == "AD" else eng.PAIRS_UPPER
    first = np.array([3, 4, 5, 6, 7, 8, 9, 10, m // 2, m // 2 + 1, m - 3, m - 2], dtype=np.int64)
    chi_cut = 9.0
    hits = eng.epi_exact_scan(store, lu, lv, py, packed, first, space, chi_cut)
    n_cand = sum((m if kind == "AD" else m - 1 - int(i)) for i in first)
    assert 0 < len(hits) < n_cand                                   # chi^2_1 > 9: ~0.27 % under the null
    assert np.all(hits["chi"] > chi_cut)
    assert set(hits["snp_0"].tolist()) <= set(first.tolist())
    if kind != "AD":
        assert np.all(hits["snp_1"] > hits["snp_0"])
    keys = set(zip(hits["snp_0"].tolist(), hits["snp_1"].tolist()))
    assert len(keys) == len(hits)                                   # no pair reported twice
    # (1) the pair-list kernel recomputes every reported hit: same numbers
    pairs = np.stack([hits["snp_0"], hits["snp_1"]], axis=1).astype(np.int64)
    again = eng.epi_exact_pairs(store, lu, lv, py, packed, pairs)
    np.testing.assert_allclose(again["eff"], hits["eff"], rtol=1e-12, atol=1e-13)
    np.testing.assert_allclose(again["var"], hits["var"], rtol=1e-12)
    # (2) oracle on a sample of the hits plus pairs the scan did NOT report (they must fall below the cut)
    rng = np.random.Generator(np.random.PCG64(n + m))
    pick = rng.choice(len(hits), size=min(48, len(hits)), replace=False)
    extra = np.array([[int(first[q % len(first)]), int(rng.integers(int(first[q % len(first)]) + 1, m))] for q in range(16)])
    sample = np.concatenate([pairs[pick], extra])
    res = eng.epi_exact_pairs(store, lu, lv, py, packed, sample)
    eff, var = _oracle_pairs(host, n, sample, cu, cv, py.cpu().numpy(), pvp.cpu().numpy())
    np.testing.assert_allclose(res["eff"], eff, rtol=RTOL, atol=1e-9 * np.abs(eff).max())
    np.testing.assert_allclose(res["var"], var, rtol=RTOL)
    chi = eff * eff / var
    for (a, b), c in zip(extra.tolist(), chi[len(pick):].tolist()):
        assert ((a, b) in keys) == (c > chi_cut) or abs(c / chi_cut - 1) < 1e-7
    # (3) linearity: eff(2 py) = 2 eff, var unchanged, chi x 4
    res2 = eng.epi_exact_pairs(store, lu, lv, 2.0 * py, packed, sample)
    np.testing.assert_allclose(res2["eff"], 2.0 * res["eff"], rtol=1e-14, atol=0)
    assert np.array_equal(res2["var"], res["var"])


def _real_model(eng, store, n, seed, torch, dev):
    """Z'PZ and Z'Py of a REAL mixed model on this very data (GRMs [A, A*A] from the GRM kernel, fixed variance
    components, one record per id, X = intercept): unlike the centred Wishart of _model its spectrum is what a scan
    sees (a few large eigenvalues from population structure in the GRM, a long tail near 1/sigma_e^2)."""
    from gmat_b200.gmatrix.gmatrix import grm_device
    ag = grm_device(store, eng.ADD)[0]
    gen = torch.Generator(device=dev)
    gen.manual_seed(seed)
    y = torch.randn((n, 1), dtype=torch.float64, device=dev, generator=gen)
    V = 0.35 * ag + 0.10 * ag * ag + 0.55 * torch.eye(n, dtype=torch.float64, device=dev)
    Vi = torch.linalg.inv(V)
    X = torch.ones((n, 1), dtype=torch.float64, device=dev)
    VX = Vi @ X
    P = Vi - VX @ torch.linalg.inv(X.T @ VX) @ VX.T
    return (0.5 * (P + P.T)).contiguous(), (P @ y).reshape(-1).contiguous()


def _plant_rare_snps(payload, n, snps, carriers, torch):
    """Overwrite the given SNP rows of a device payload with rare variants: ``carriers`` heterozygotes spread over
    the ids, everybody else homozygous for the other allele (code 00)."""
    for q, snp in enumerate(snps):
        payload[snp].zero_()
        ids = (torch.arange(carriers, device=payload.device) * (n // carriers) + 3 * q + 1) % n
        for k in ids.tolist():
            payload[snp, k // 4] |= 2 << (2 * (k % 4))                 # code 10 = heterozygote


@pytest.mark.parametrize("n,m,kind", [(2000, 20000, "AA"), (5000, 50000, "AA"), (5000, 50000, "DD"),
                                      (5000, 50000, "AD")])
def test_i8_scan_full_shape_sampled_against_oracle(eng, n, m, kind):
    """The int8 tensor-core engine at BASELINE's shapes on a REAL Z'PZ (GRMs of the data), first SNPs including
    planted RARE variants (5 and 12 carriers: where the digit planes lose most against max|u|^2 max|A|): same hit set
    as the DMMA kernel on the same first SNPs, values compared on the whole intersection, and EVERY hit plus 2000
    unreported pairs recomputed by the numpy oracle to 1e-9."""
    import torch
    import synth
    dev = eng.current_device()
    payload = synth.device_payload(n, m, 20201013 + n, dev, eng.row_stride(n))
    rare = [7, m // 2 + 1, 40, m - 3]
    _plant_rare_snps(payload, n, rare[:2], 5, torch)
    _plant_rare_snps(payload, n, rare[2:], 12, torch)
    store = eng.GenotypeStore.from_device_payload(payload, n)
    host = payload.cpu().numpy()
    pvp, py = _real_model(eng, store, n, 5, torch, dev)
    cu, cv = kind[0], kind[1]
    code = {"A": eng.ADD, "D": eng.DOM}
    lu, lv = store.lut(code[cu]), store.lut(code[cv])
    space = eng.PAIRS_ALL if kind == "AD" else eng.PAIRS_UPPER
    first = np.array([3, 4, 5, 6, 7, 8, 9, 10, m // 2, m // 2 + 1, m - 3, m - 2], dtype=np.int64)
    chi_cut = 9.0
    sc = eng.I8Exact(store, lu, lv, code[cv], py, pvp)
    hits = sc.scan(first, space, chi_cut)
    ref = eng.epi_exact_scan(store, lu, lv, py, eng.pack_pvp(pvp), first, space, chi_cut)
    key = lambda h: h["snp_0"].astype(np.int64) * m + h["snp_1"]
    hits, ref = hits[np.argsort(key(hits))], ref[np.argsort(key(ref))]
    assert len(np.unique(key(hits))) == len(hits)
    border = set(key(ref)[np.abs(ref["chi"] / chi_cut - 1) < 1e-7].tolist())
    assert len(ref) > 100 and (set(key(hits).tolist()) ^ set(key(ref).tolist())) <= border
    # values on the whole intersection, unconditionally (the border band only excuses set membership)
    _, ih, ir = np.intersect1d(key(hits), key(ref), return_indices=True)
    assert len(ih) >= len(ref) - len(border)
    np.testing.assert_allclose(hits["var"][ih], ref["var"][ir], rtol=RTOL)
    np.testing.assert_allclose(hits["eff"][ih], ref["eff"][ir], rtol=RTOL, atol=1e-9 * np.abs(ref["eff"]).max())
    # the oracle: every hit, in blocks (the host product is n^2 per pair)
    pyh, pvh = py.cpu().numpy(), pvp.cpu().numpy()
    allp = np.stack([hits["snp_0"], hits["snp_1"]], axis=1).astype(np.int64)
    for k0 in range(0, len(allp), 1024):
        blk = slice(k0, k0 + 1024)
        eff, var = _oracle_pairs(host, n, allp[blk], cu, cv, pyh, pvh)
        np.testing.assert_allclose(hits["eff"][blk], eff, rtol=RTOL, atol=1e-9 * np.abs(eff).max())
        np.testing.assert_allclose(hits["var"][blk], var, rtol=RTOL)
        assert np.all(eff * eff / var > chi_cut * (1 - 1e-7))
    # ... and 2000 pairs the scan did NOT report: the oracle's chi is below the cut
    rng = np.random.Generator(np.random.PCG64(n + m + 1))
    have = set(key(hits).tolist())
    miss = []
    while len(miss) < 2000:
        i = int(first[rng.integers(len(first))])
        j = int(rng.integers(0 if kind == "AD" else i + 1, m))
        if i * m + j not in have:
            miss.append((i, j))
    eff, var = _oracle_pairs(host, n, np.array(miss), cu, cv, pyh, pvh)
    chi = eff * eff / var
    assert np.all(~(chi > chi_cut * (1 + 1e-7)))                      # NaN (monomorphic partner) is not a hit either


def test_screen_full_c3_shape_subset_equals_oracle(eng):
    import torch
    import synth
    n, m = 2000, 20000
    dev = eng.current_device()
    payload = synth.device_payload(n, m, 77, dev, eng.row_stride(n))
    store = eng.GenotypeStore.from_device_payload(payload, n)
    host = payload.cpu().numpy()
    _, py = _model(n, 9, torch, dev)
    pyh = py.cpu().numpy()
    la = store.lut(eng.ADD)                         # numpy-style p (the C-loop p differs in the last ulp only)
    first = np.array(sorted(set(range(40, 56)) | set(range(m - 300, m - 284))), dtype=np.int64)
    # threshold at 3 sigma of the null effect distribution: ~0.27 % of the pairs
    X = _host_columns(host, n, first.tolist(), "A")
    sd = float(np.sqrt(np.mean((X * X).sum(axis=0)) * np.mean(pyh * pyh)) * np.sqrt(np.mean(X * X)))
    cut = 3.0 * sd
    hits = eng.epi_eff(store, la, la, py, first, 0, cut)
    got = {(int(a), int(b)): e for a, b, e in zip(hits["snp_0"], hits["snp_1"], hits["eff"])}
    assert len(got) == len(hits) and len(hits) > 0
    # oracle: effects of ALL pairs (i in first, j > i), numpy, in column blocks
    want = {}
    Xi = X * pyh[:, None]
    for j0 in range(0, m, 2000):
        cols = list(range(j0, min(m, j0 + 2000)))
        E = Xi.T @ _host_columns(host, n, cols, "A")
        for r, i in enumerate(first.tolist()):
            for c in np.nonzero(np.abs(E[r]) > cut * (1 - 1e-9))[0]:
                j = cols[c]
                if j > i:
                    want[(i, j)] = E[r, c]
    border = {k for k, e in want.items() if abs(abs(e) / cut - 1) < 1e-7}
    assert set(got) - border == set(k for k in want if abs(want[k]) > cut) - border
    for k, e in got.items():
        assert abs(e - want[k]) <= RTOL * abs(e) + 1e-9 * sd


def test_grm_c3_shape_sampled_entries(eng):
    import torch
    import synth
    from gmat_b200.gmatrix.gmatrix import grm_device
    n, m = 2000, 20000
    dev = eng.current_device()
    payload = synth.device_payload(n, m, 31, dev, eng.row_stride(n))
    store = eng.GenotypeStore.from_device_payload(payload, n)
    host = payload.cpu().numpy()
    for coding, tag in ((eng.ADD, "A"), (eng.DOM, "D")):
        kin, scale = grm_device(store, coding, small_val=0.001)
        K = kin.cpu().numpy()
        assert np.array_equal(K, K.T)
        X = _host_columns(host, n, list(range(m)), tag)                 # [n][m] float64, 320 MB
        rows = [0, 1, 127, 128, 1000, n - 1]
        want = X[rows] @ X.T
        freq = None
        if tag == "A":
            g = X + 0.0
            # scale = sum 2p(1-p) (gmatrix.py:53-58); recover p from the centred column: mean(g) = 2p
            rowsall = host[:, : (n + 3) // 4]
            codes = np.stack([(rowsall >> (2 * q)) & 3 for q in range(4)], axis=2).reshape(m, -1)[:, :n]
            freq = np.array([0.0, 0.0, 1.0, 2.0])[codes].sum(axis=1) / (2 * n)
            sc = float(np.sum(2 * freq * (1 - freq)))
        else:
            rowsall = host[:, : (n + 3) // 4]
            codes = np.stack([(rowsall >> (2 * q)) & 3 for q in range(4)], axis=2).reshape(m, -1)[:, :n]
            freq = np.array([0.0, 0.0, 1.0, 2.0])[codes].sum(axis=1) / (2 * n)
            sv = 2 * freq * (1 - freq)
            sc = float(np.sum(sv * (1 - sv)))                             # gmatrix.py:116-122
        assert abs(scale - sc) <= 1e-12 * sc
        want = want / sc
        for q, r in enumerate(rows):
            want[q, r] *= 1.001
        assert np.abs(K[rows] - want).max() <= RTOL * np.abs(want).max()
        del X
