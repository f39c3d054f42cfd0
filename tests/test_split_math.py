"""CPU restatement in numpy of the arithmetic of the int8 exact engine (gmat_b200/csrc/ozaki.cu): the error-free split
of B_i = diag(u_i) Z'PZ diag(u_i) into signed base-128 digit planes, the exact integer plane products with the integer
genotype vector, and the FP64 recombination in the kernel's order.  Property: the quadratic form a' B a comes out within
a few ulp of a long-double evaluation for 8 planes (56 bits), and degrades by 128 per dropped plane -- the numbers
DESIGN.md 3.0 quotes.  No GPU, no library call."""
import numpy as np
import pytest


def _split(b, scale, planes):
    """i8_split_kernel: t = rn(b / scale * 128^S) as int64, balanced digits in [-64, 63], top digit |.| <= 127."""
    t = np.rint(b * (128.0 ** planes / scale)).astype(np.int64)
    digs = []
    for _ in range(planes - 1):
        d = ((t + 64) & 127) - 64
        t = (t - d) >> 7
        digs.append(d)
    assert np.abs(t).max() <= 127
    digs.append(t)
    return digs[::-1]                                                  # plane 0 = most significant


@pytest.mark.parametrize("planes,tol", [(8, 2e-14), (7, 2e-12), (6, 3e-10)])
def test_digit_planes_reproduce_the_quadratic_form(planes, tol):
    rng = np.random.Generator(np.random.PCG64(planes))
    n, m = 260, 40
    c = np.eye(n) - 1.0 / n
    z = rng.normal(size=(n, n + 5))
    a_mat = c @ (z @ z.T / n) @ c
    a_mat = 0.5 * (a_mat + a_mat.T)                                    # a stand-in for Z'PZ: symmetric PSD, centred
    g = rng.binomial(2, rng.uniform(0.05, 0.95, size=m)[:, None], size=(m, n)).astype(np.int64)
    u = g[0] - g[0].mean()                                             # coded first SNP
    b = (u[:, None] * a_mat) * u[None, :]
    scale = np.abs(u).max() ** 2 * np.abs(a_mat).max() / 0.99          # engine.I8Exact.scan
    digs = _split(b, scale, planes)
    for d in digs:
        assert d.min() >= -128 and d.max() <= 127                      # int8 operands
    recon = sum(d.astype(np.float64) * 128.0 ** -(s + 1) for s, d in enumerate(digs))
    # the split itself: half a unit of the last digit, plus the rounding of this FP64 reconstruction
    assert np.abs(recon * scale - b).max() <= scale * (128.0 ** -planes + 4e-16)
    worst = 0.0
    for j in range(1, m):
        aj = g[j]
        qacc, wgt = 0.0, 1.0
        for d in digs:                                                 # kernel order: planes 0, 1, 2, ... one FP64 add each
            wgt *= 0.0078125
            qacc += float(int(aj @ d @ aj)) * wgt                      # exact integer plane sum (int64)
        quad = scale * qacc
        want = float(np.longdouble(1) * (aj.astype(np.longdouble) @ b.astype(np.longdouble) @ aj.astype(np.longdouble)))
        # relative to the size of the terms (the quadratic form of a non-centred integer vector is far from cancelling)
        worst = max(worst, abs(quad - want) / abs(want))
    assert worst <= tol, worst
