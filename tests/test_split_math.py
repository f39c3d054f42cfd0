"""CPU restatement in numpy of the arithmetic of the int8 exact engine (gmat_b200/csrc/ozaki.cu): the error-free split
of B_i = diag(u_i) Z'PZ diag(u_i) into signed base-256 digit planes, the exact integer plane products with the integer
genotype vector, and the FP64 recombination in the kernel's order.  Property: the quadratic form a' B a comes out within
a few ulp of a long-double evaluation for 7 planes (55 bits), and degrades by 256 per dropped plane -- the numbers
DESIGN.md 3.0 quotes.  No GPU, no library call."""
import numpy as np
import pytest


def _split(b, scale, planes):
    """i8_split_kernel: t = rn(b / scale * 256^S) as int64, balanced digits in [-128, 127], top digit |.| <= 127."""
    t = np.rint(b * (256.0 ** planes / scale)).astype(np.int64)
    digs = []
    for _ in range(planes - 1):
        d = ((t + 128) & 255) - 128
        t = (t - d) >> 8
        digs.append(d)
    assert np.abs(t).max() <= 127
    digs.append(t)
    return digs[::-1]                                                  # plane 0 = most significant


@pytest.mark.parametrize("planes,tol", [(7, 2e-14), (6, 6e-12), (5, 2e-9)])
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
    scale = np.abs(u).max() ** 2 * np.abs(a_mat).max() / 0.49          # i8_side_kernel
    digs = _split(b, scale, planes)
    for d in digs:
        assert d.min() >= -128 and d.max() <= 127                      # int8 operands
    recon = sum(d.astype(np.float64) * 256.0 ** -(s + 1) for s, d in enumerate(digs))
    # the split itself: half a unit of the last digit, plus the rounding of this FP64 reconstruction
    assert np.abs(recon * scale - b).max() <= scale * (0.5 * 256.0 ** -planes + 4e-16)
    worst = 0.0
    for j in range(1, m):
        aj = g[j]
        qacc, wgt = 0.0, 1.0
        for d in digs:                                                 # kernel order: planes 0, 1, 2, ... one FP64 add each
            wgt *= 0.00390625
            qacc += float(int(aj @ d @ aj)) * wgt                      # exact integer plane sum (int64)
        quad = scale * qacc
        want = float(np.longdouble(1) * (aj.astype(np.longdouble) @ b.astype(np.longdouble) @ aj.astype(np.longdouble)))
        # relative to the size of the terms (the quadratic form of a non-centred integer vector is far from cancelling)
        worst = max(worst, abs(quad - want) / abs(want))
    assert worst <= tol, worst
