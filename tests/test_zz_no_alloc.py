"""gmat_b200_epi_exact_scan_noalloc: the exact scan with its scratch (the tile table) in a file-scope static
__device__ array instead of a cudaMallocAsync block -- for callers that may not allocate (CUDA graph capture, serving
loops).  No counterpart in the reference (include/gmat_b200.h).

Checked against the existing entry point gmat_b200_epi_exact_scan through the C-ABI, on two calls in a row that use
DIFFERENT first-SNP lists and pair spaces (the array keeps the last call's contents: the second call must not see
them), with the 1e-9 relative tolerance of the other exact-scan tests.  This file is named to be collected last."""
import numpy as np
import pytest

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
    return c @ rng.normal(size=n), 0.5 * (a + a.T)


def _scan_noalloc(eng, store, lu, lv, pyd, packed, first, space, chi_cut, cap):
    """One call of the new entry point; every buffer is allocated by the caller BEFORE the call."""
    import torch
    from gmat_b200._ffi import ffi, lib, check
    dev = store.device
    il = torch.as_tensor(np.ascontiguousarray(first, dtype=np.int64)).to(dev)
    hits = eng.alloc_hits(cap, dev)
    count = torch.zeros((1,), dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    rc = lib.gmat_b200_epi_exact_scan_noalloc(
        ffi.cast("unsigned char *", store.bed.data_ptr()), store.n, store.m, store.row_bytes,
        ffi.cast("double *", lu.data_ptr()), ffi.cast("double *", lv.data_ptr()),
        ffi.cast("double *", pyd.data_ptr()), ffi.cast("double *", packed.data_ptr()),
        ffi.cast("long long *", il.data_ptr()), il.numel(), space, float(chi_cut),
        ffi.cast("gmat_b200_hit *", hits.data_ptr()), cap,
        ffi.cast("unsigned long long *", count.data_ptr()),
        ffi.cast("void *", torch.cuda.current_stream().cuda_stream))
    check(rc, "gmat_b200_epi_exact_scan_noalloc")
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0, "the call changed the device's free memory"
    cnt = int(count.item())
    assert cnt <= cap
    return eng.hits_to_numpy(hits, cnt)


def _same(ref, got, m):
    key = lambda h: h["snp_0"].astype(np.int64) * m + h["snp_1"]
    ref, got = ref[np.argsort(key(ref))], got[np.argsort(key(got))]
    assert len(ref) == len(got) and np.array_equal(key(ref), key(got))
    np.testing.assert_allclose(got["eff"], ref["eff"], rtol=RTOL, atol=1e-12)
    np.testing.assert_allclose(got["var"], ref["var"], rtol=RTOL)
    ok = ~np.isnan(ref["chi"])
    assert np.array_equal(np.isnan(got["chi"]), ~ok)
    np.testing.assert_allclose(got["chi"][ok], ref["chi"][ok], rtol=4 * RTOL)


def test_noalloc_scan_matches_the_allocating_entry_point_twice_in_a_row(eng, tmp_path):
    import torch
    import synth
    n, m = 300, 420
    prefix = str(tmp_path / "syn")
    synth.write_dataset(prefix, n, m, seed=97, missing_rate=0.0)
    py, pvp = _model(n, 5)
    store = eng.GenotypeStore(prefix)
    lu, lv = store.lut(eng.ADD), store.lut(eng.DOM)
    pyd = torch.as_tensor(py.copy()).cuda()
    packed = eng.pack_pvp(torch.as_tensor(pvp).cuda().contiguous())
    # call 1: a long ragged list, upper-triangle pair space, every pair reported (AA)
    first1 = np.array([i for i in range(m - 1) if i % 5 != 2])
    ref1 = eng.epi_exact_scan(store, lu, lu, pyd, packed, first1, eng.PAIRS_UPPER, -1.0)
    got1 = _scan_noalloc(eng, store, lu, lu, pyd, packed, first1, eng.PAIRS_UPPER, -1.0, cap=len(ref1) + 8)
    assert len(ref1) == int(sum(m - 1 - i for i in first1))
    _same(ref1, got1, m)
    # call 2, straight after: a SHORTER, different list in the other pair space (AD, all j), thresholded -- a tile
    # table left over from call 1 would give other pairs
    first2 = np.arange(7, m, 11)
    ref2 = eng.epi_exact_scan(store, lu, lv, pyd, packed, first2, eng.PAIRS_ALL, 1.5)
    got2 = _scan_noalloc(eng, store, lu, lv, pyd, packed, first2, eng.PAIRS_ALL, 1.5, cap=len(first2) * m)
    assert 0 < len(ref2) < len(first2) * m
    _same(ref2, got2, m)
    # and call 1 again: same result as the first time
    got3 = _scan_noalloc(eng, store, lu, lu, pyd, packed, first1, eng.PAIRS_UPPER, -1.0, cap=len(ref1) + 8)
    _same(ref1, got3, m)


def test_noalloc_scan_rejects_lists_beyond_the_static_scratch(eng, tmp_path):
    import torch
    import synth
    from gmat_b200._ffi import ffi, lib, CONST
    limit = CONST["GMAT_B200_NOALLOC_MAX_FIRST"]
    assert lib.gmat_b200_noalloc_max_first() == limit
    n, m = 64, 40
    prefix = str(tmp_path / "syn")
    synth.write_dataset(prefix, n, m, seed=3, missing_rate=0.0)
    py, pvp = _model(n, 1)
    store = eng.GenotypeStore(prefix)
    lu = store.lut(eng.ADD)
    pyd = torch.as_tensor(py.copy()).cuda()
    packed = eng.pack_pvp(torch.as_tensor(pvp).cuda().contiguous())
    il = torch.zeros((limit + 1,), dtype=torch.int64, device="cuda")
    count = torch.zeros((1,), dtype=torch.int64, device="cuda")
    hits = eng.alloc_hits(16, store.device)

    def call(n_i):
        return lib.gmat_b200_epi_exact_scan_noalloc(
            ffi.cast("unsigned char *", store.bed.data_ptr()), store.n, store.m, store.row_bytes,
            ffi.cast("double *", lu.data_ptr()), ffi.cast("double *", lu.data_ptr()),
            ffi.cast("double *", pyd.data_ptr()), ffi.cast("double *", packed.data_ptr()),
            ffi.cast("long long *", il.data_ptr()), n_i, eng.PAIRS_UPPER, 1e300,
            ffi.cast("gmat_b200_hit *", hits.data_ptr()), 16,
            ffi.cast("unsigned long long *", count.data_ptr()), ffi.NULL)

    assert call(limit + 1) == CONST["GMAT_B200_ERR_ARG"]
    msg = ffi.string(lib.gmat_b200_last_error()).decode()
    assert "GMAT_B200_NOALLOC_MAX_FIRST" in msg and str(limit) in msg
    torch.cuda.synchronize()
    assert int(count.item()) == 0
    # the largest list the scratch holds is accepted (first SNP 0 repeated: no pair passes chi > 1e300)
    assert call(limit) == CONST["GMAT_B200_OK"]
    torch.cuda.synchronize()
    assert int(count.item()) == 0
