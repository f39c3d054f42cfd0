"""CPU restatement in numpy of the arithmetic of the int8 GRM (gmat_b200/csrc/gram_i8.cu): X X' = G G' - c 1' - 1 c' + d
with G the integer code matrix (one exact integer GEMM), c_k = sum_s o_s a_ks and d = sum_s o_s^2 in FP64, finished
with the kernel's expression ((S - c_hi) - c_lo) + d -- against the reference formulation (gmatrix.py:53-66,
116-130) on the same genotypes.  No GPU, no library call."""
import numpy as np
import pytest


@pytest.mark.parametrize("coding", ["add", "dom"])
def test_integer_gram_plus_rank_one_corrections(coding):
    rng = np.random.Generator(np.random.PCG64(len(coding)))
    n, m = 150, 4000
    g = rng.binomial(2, rng.uniform(0.05, 0.95, size=m)[None, :], size=(n, m)).astype(np.float64)
    freq = g.sum(axis=0) / (2.0 * n)
    if coding == "add":
        a, o = g, 2.0 * freq                                           # gmatrix.py:53-56
        scale = np.sum(2.0 * freq * (1.0 - freq))
    else:
        a = np.where(g > 1.5, 0.0, g)                                  # gmatrix.py:116-122
        o = 2.0 * freq * (1.0 - freq)
        scale = np.sum(o * (1.0 - o))
    x = a - o[None, :]
    want = x @ x.T / scale
    want[np.diag_indices(n)] *= 1.001                                  # small_val, gmatrix.py:64-66
    s_int = a.astype(np.int64) @ a.astype(np.int64).T                  # the int8 tensor-core GEMM, exact
    c = a @ o
    d = float(np.sum(o * o))
    hi, lo = np.maximum.outer(np.arange(n), np.arange(n)), np.minimum.outer(np.arange(n), np.arange(n))
    k = ((s_int[hi, lo].astype(np.float64) - c[hi]) - c[lo] + d) / scale   # i8_gram_finish_kernel
    k[np.diag_indices(n)] += k[np.diag_indices(n)] * 0.001
    assert np.array_equal(k, k.T)
    assert np.abs(k - want).max() <= 1e-12 * np.abs(want).max()
