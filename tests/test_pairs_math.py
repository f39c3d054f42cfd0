"""The arithmetic the int8 PAIR-LIST engine rests on (gmat_b200/csrc/ozaki.cu "pair lists", engine.I8Pairs), restated
in numpy: w = (a - o_i) o (b - o_j) = q - o_j a - o_i b + o_i o_j 1 with q = a o b, the bilinear forms of q, a, b by
polarisation from QUADRATIC forms of integer vectors (rows q, q+a, q+b, a+b and one row per SNP), the terms with 1 as
dot products with A 1, and the coefficients the a-posteriori guard weights the rows' split errors with."""
import numpy as np


def _combine(Q, E, T, asq, oi, oj, s11, sp):
    """i8_pairs_combine_kernel, line by line: rows 0..3 = q, q+a, q+b, a+b; 4 = a; 5 = b."""
    c2, c3, c4 = -oj, -oi, oi * oj
    Bqa = 0.5 * ((Q[1] - Q[0]) - Q[4])
    Bqb = 0.5 * ((Q[2] - Q[0]) - Q[5])
    Bab = 0.5 * ((Q[3] - Q[4]) - Q[5])
    var = (Q[0] + c2 * c2 * Q[4] + c3 * c3 * Q[5] + 2.0 * (c2 * Bqa + c3 * Bqb + c2 * c3 * Bab)
           + 2.0 * c4 * (T[0] + c2 * T[4] + c3 * T[5]) + c4 * c4 * s11)
    eff = E[0] + c2 * E[4] + c3 * E[5] + c4 * sp
    coef = np.array([1.0 - c2 - c3, c2, c3, c2 * c3, c2 * c2 - c2 - c2 * c3, c3 * c3 - c3 - c2 * c3])
    return var, eff, coef


def test_polarisation_identity_and_guard_coefficients():
    rng = np.random.Generator(np.random.PCG64(4))
    n = 257
    for trial in range(40):
        b = rng.normal(size=(n, n + 5))
        A = b @ b.T / n
        if trial % 2 == 0:                            # with an intercept in the model A 1 = 0, otherwise it is not
            c = np.eye(n) - np.ones((n, n)) / n
            A = c @ A @ c
        A = 0.5 * (A + A.T)
        py = rng.normal(size=n)
        kind = ("AA", "AD", "DD")[trial % 3]
        ga, gb = rng.binomial(2, rng.uniform(0.03, 0.97), size=n), rng.binomial(2, rng.uniform(0.03, 0.97), size=n)
        a = ga if kind[0] == "A" else (ga == 1).astype(int)
        bb = gb if kind[1] == "A" else (gb == 1).astype(int)
        pa, pb = ga.sum() / (2.0 * n), gb.sum() / (2.0 * n)
        oi = 2 * pa if kind[0] == "A" else 2 * pa * (1 - pa)     # -lut[.][0]: remma_epiAA.py:59-61, remma_epiDD.py:59-65
        oj = 2 * pb if kind[1] == "A" else 2 * pb * (1 - pb)
        w = (a - oi) * (bb - oj)
        want_var, want_eff = w @ A @ w, w @ py
        q = a * bb
        rows = np.stack([q, q + a, q + bb, a + bb, a, bb]).astype(np.float64)
        assert rows.max() <= 6 and rows.min() >= 0               # int8 operand, 4-bit epilogue codes
        t1 = A @ np.ones(n)
        Q = np.einsum("rk,kl,rl->r", rows, A, rows)
        E, T, asq = rows @ py, rows @ t1, (rows ** 2).sum(axis=1)
        var, eff, coef = _combine(Q, E, T, asq, oi, oj, t1.sum(), py.sum())
        scale = np.abs(np.einsum("k,kl,l->", np.abs(w), np.abs(A), np.abs(w))) + 1e-300
        assert abs(var - want_var) <= 1e-11 * max(abs(want_var), 1e-3 * np.abs(Q).max())
        assert abs(eff - want_eff) <= 1e-11 * max(abs(want_eff), 1e-3 * np.abs(E).max())
        # the guard's weights ARE d var / d Q_r: perturb every row's quadratic form and watch var move
        for r in range(6):
            Qp = Q.copy()
            Qp[r] += 1.0
            vp, _, _ = _combine(Qp, E, T, asq, oi, oj, t1.sum(), py.sum())
            assert abs((vp - var) - coef[r]) <= 1e-9 * max(1.0, abs(coef[r]))
        assert scale > 0


def test_split_error_of_a_quadratic_form_scales_with_sum_x2():
    """What the guard assumes: truncating A / scale to S base-256 digits moves x'Ax by at most ~ 256^-S scale sum x^2
    (half a unit per element, random signs: the standard deviation is 0.29 of that bound, 4x is the guard's unit)."""
    rng = np.random.Generator(np.random.PCG64(9))
    n, S = 300, 3
    b = rng.normal(size=(n, n + 5))
    A = b @ b.T / n
    A = 0.5 * (A + A.T)
    scale = np.abs(A).max() / 0.49
    At = np.rint(A / scale * 256.0 ** S) / 256.0 ** S * scale
    worst = 0.0
    for _ in range(200):
        x = rng.integers(0, 7, size=n).astype(np.float64)
        err = abs(x @ At @ x - x @ A @ x)
        worst = max(worst, err / (256.0 ** -S * scale * (x ** 2).sum()))
    assert worst < 4.0
