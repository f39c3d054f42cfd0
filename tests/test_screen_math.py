"""CPU restatement of stage 1 of the int8 screen (gmat_b200/csrc/screen_i8.cu) in numpy: the two-digit split of Z'Py,
the exact integer plane sums, the FP32 combine in the kernel's operation order -- and the property everything rests
on: the prefilter value never differs from the FP64 effect by more than the margin the library subtracts from the
cut, so stage 1 cannot lose a pair.  No GPU, no library call: this pins the arithmetic, the GPU tests pin the kernel."""
import numpy as np
import pytest


def _digits(py, scale):
    """sc_digits_kernel: balanced base-128 digits of py / scale, two planes."""
    t = np.rint(py / scale * 16384.0).astype(np.int64)
    d1 = ((t + 64) & 127) - 64
    d0 = (t - d1) >> 7
    return d0, d1


def _stage1(a_u, o_u, a_v, o_v, py):
    """ẽ[i][j] exactly as i8_screen_pair_kernel forms it (int32 plane sums -> FP32), and the margin."""
    n = len(py)
    scale = np.abs(py).max() / 0.49
    d0, d1 = _digits(py, scale)
    assert d0.min() >= -64 and d0.max() <= 63 and d1.min() >= -64 and d1.max() <= 63
    p0, p1 = a_u * d0[None, :], a_u * d1[None, :]                    # sc_planes_kernel: int8 operands
    assert p0.min() >= -128 and p0.max() <= 127 and p1.min() >= -128 and p1.max() <= 127
    y0, y1 = p0 @ a_v.T, p1 @ a_v.T                                   # exact int8 x int8 -> int32 GEMMs
    assert np.abs(y0).max() < 2 ** 24 and np.abs(y1).max() < 2 ** 24  # exact in FP32
    f = np.float32
    t_u = (a_u @ py).astype(f)                                        # sc_dot_kernel (FP64) -> FP32 copies
    t_v = a_v @ py
    c_v = (t_v - o_v * py.sum()).astype(f)                            # sc_second_kernel
    ou, ov = o_u.astype(f), o_v.astype(f)
    qscale = f(scale / 16384.0)
    t32 = y0.astype(f) * f(128.0) + y1.astype(f)                      # fmaf(v0, 128, v1): exact product, one rounding
    r = (-ov)[None, :] * t_u[:, None] + (-(ou[:, None] * c_v[None, :]))
    est = t32 * qscale + r
    margin = scale * n * (2.0 / 16384.0 + 4.0e-6)
    return est.astype(np.float64), margin, scale


@pytest.mark.parametrize("kind", ["AA", "AD", "DD"])
@pytest.mark.parametrize("n,m,outliers", [(300, 90, False), (1000, 60, True), (2000, 40, True)])
def test_stage1_never_misses_by_more_than_the_margin(kind, n, m, outliers):
    rng = np.random.Generator(np.random.PCG64(n + m + len(kind)))
    p = rng.uniform(0.03, 0.97, size=m)
    g = rng.binomial(2, p[:, None], size=(m, n)).astype(np.int64)
    freq = g.sum(axis=1) / (2.0 * n)
    add = (g, 2.0 * freq)                                             # a = dosage, o = 2p          (remma_epiAA.py:59-61)
    dom = (np.where(g == 2, 0, g), 2.0 * freq * (1.0 - freq))         # a = het indicator, o = 2pq  (remma_epiDD.py:59-65)
    (a_u, o_u), (a_v, o_v) = {"AA": (add, add), "AD": (add, dom), "DD": (dom, dom)}[kind]
    py = rng.normal(size=n)
    if outliers:
        py[rng.integers(0, n, size=5)] *= 40.0                        # scale is set by max |py|: the hard case
    py -= py.mean()
    est, margin, scale = _stage1(a_u, o_u, a_v, o_v, py)
    x_u, x_v = a_u - o_u[:, None], a_v - o_v[:, None]
    eff = (x_u * py[None, :]) @ x_v.T                                 # FP64 effects of every pair (_remma_epi_eff_cpu.c:69-77)
    err = np.abs(est - eff).max()
    assert err <= margin, (err, margin)
    # the bound is not vacuous: without outliers in Z'Py the margin is a small fraction of a p = 1e-5 cut on these
    # effects; with extreme outliers it can reach cut / 4, which is where the library hands the call to the FP64 kernel
    cut = 4.42 * np.sqrt(np.median(eff ** 2) / 0.4549)
    if not outliers:
        assert margin < 0.05 * cut
    if not margin < 0.25 * cut:
        return
    # and no pair that passes the cut in FP64 fails the prefilter test |e~| >= cut - margin
    passing = np.abs(eff) > cut
    assert np.all(np.abs(est)[passing] >= cut - margin)
