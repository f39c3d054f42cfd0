#!/usr/bin/env python
"""Measured accuracy of the int8 exact engine as a function of the number of base-256 digit planes (run on a B200).

Data: synthetic genotypes of the bench shape, a REAL Z'PZ (GRMs [A, A*A] of this very data, fixed variance components,
one record per id, X = intercept).  For S in --planes: every pair (i, j > i) of a sample of first SNPs -- evenly spread
plus the --rare lowest-MAF SNPs of the data -- is computed by the int8 engine and by the FP64 DMMA kernel; a few pairs
per first SNP are also evaluated in long double on the host.  Reports max relative error of var / eff and how many
pairs the a-posteriori guard sent to the FP64 pair kernel.  Output: gpurun_out/i8_accuracy.json (-> profiles/)."""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ids", dest="n", type=int, default=2000)
    ap.add_argument("--snps", dest="m", type=int, default=20000)
    ap.add_argument("--first", type=int, default=24)
    ap.add_argument("--rare", type=int, default=8)
    ap.add_argument("--planes", default="5,6,7")
    ap.add_argument("--min-maf", type=float, default=0.002, help="lowest MAF of the synthetic data")
    ap.add_argument("--major-share", type=float, default=0.5,
                    help="share of the SNPs whose COUNTED (.bim A2) allele is the major one -- allele labels are "
                         "arbitrary in real files; the engine codes such SNPs by the other allele (I8Exact.MINOR_ALLELE)")
    args = ap.parse_args()
    import torch
    from gmat_b200 import engine
    import synth
    from gmat_b200.gmatrix.gmatrix import grm_device
    dev = engine.current_device()
    n, m = args.n, args.m
    # synthetic genotypes with a long low-MAF tail (the regime where var is small against max|u|^2 max|A|)
    rng = np.random.Generator(np.random.PCG64(11))
    maf = np.concatenate([rng.uniform(args.min_maf, 0.02, size=m // 10), rng.uniform(0.02, 0.5, size=m - m // 10)])
    rng.shuffle(maf)
    g = rng.binomial(2, maf[:, None], size=(m, n)).astype(np.int8)
    for j in np.where(g.max(axis=1) == g.min(axis=1))[0]:
        g[j, rng.integers(n)] = 1                                # no monomorphic SNP
    swap = rng.random(m) < args.major_share
    g[swap] = 2 - g[swap]                                        # the counted allele is the common one there
    store = engine.GenotypeStore.from_dense(g.astype(np.float64))
    lut = store.lut(engine.ADD)
    ag = grm_device(store, engine.ADD)[0]
    gen = torch.Generator(device=dev)
    gen.manual_seed(3)
    y = torch.randn((n, 1), dtype=torch.float64, device=dev, generator=gen)
    V = 0.35 * ag + 0.10 * ag * ag + 0.55 * torch.eye(n, dtype=torch.float64, device=dev)
    Vi = torch.linalg.inv(V)
    X = torch.ones((n, 1), dtype=torch.float64, device=dev)
    VX = Vi @ X
    P = Vi - VX @ torch.linalg.inv(X.T @ VX) @ VX.T
    A = (0.5 * (P + P.T)).contiguous()
    py = (P @ y).reshape(-1).contiguous()
    packed = engine.pack_pvp(A)
    freq = g.sum(axis=1) / (2.0 * n)
    mafs = np.minimum(freq, 1 - freq)
    rare = np.argsort(mafs[: m - 1])[: args.rare]
    first = np.unique(np.concatenate([np.linspace(0, m - 2, args.first).astype(np.int64), rare]))
    ref = engine.epi_exact_scan(store, lut, lut, py, packed, first, engine.PAIRS_UPPER, -1.0, cap=len(first) * m)
    key = lambda h: h["snp_0"].astype(np.int64) * m + h["snp_1"]
    ref = ref[np.argsort(key(ref))]
    Ah = A.cpu().numpy().astype(np.longdouble)
    lut_h = lut.cpu().numpy()
    codes = np.array([0, 0, 1, 2])                               # no missing calls here
    pick = np.linspace(0, len(ref) - 1, 48).astype(np.int64)
    exact = []
    raw = np.where(g < 0, 0, g)
    for q in pick:
        i, j = int(ref["snp_0"][q]), int(ref["snp_1"][q])
        w = ((raw[i] - 2 * freq[i]) * (raw[j] - 2 * freq[j])).astype(np.longdouble)
        exact.append(float(w @ (Ah @ w)))
    exact = np.array(exact)
    out = {"n": n, "m": m, "first_snps": len(first), "pairs": int(len(ref)), "lowest_maf": float(mafs.min()),
           "snps_counting_the_major_allele": int(swap.sum()),
           "dmma_vs_longdouble_var": float(np.max(np.abs(ref["var"][pick] - exact) / np.abs(exact))),
           "planes": {}, "planes_plain_dosage": {}}
    for S, minor in [(int(v), mn) for mn in (True, False) for v in args.planes.split(",")]:
        engine.I8Exact.MINOR_ALLELE = minor
        eng = engine.I8Exact(store, lut, lut, engine.ADD, py, A, n_planes=S, packed=packed)
        got = eng.scan(first, engine.PAIRS_UPPER, -1.0, cap=len(first) * m)
        got = got[np.argsort(key(got))]
        assert np.array_equal(key(got), key(ref))
        rechecked = int(eng.last_rechecked)
        rv = np.abs(got["var"] - ref["var"]) / np.abs(ref["var"])
        re = np.abs(got["eff"] - ref["eff"]) / np.abs(ref["eff"]).max()
        # the same without the guard: what the planes alone deliver
        eng.RECHECK_TOL = float("inf")
        raw_got = eng.scan(first, engine.PAIRS_UPPER, -1.0, cap=len(first) * m)
        raw_got = raw_got[np.argsort(key(raw_got))]
        rv0 = np.abs(raw_got["var"] - ref["var"]) / np.abs(ref["var"])
        out["planes" if minor else "planes_plain_dosage"][str(S)] = {
            "bits": 8 * S - 1, "var_max_rel_vs_dmma": float(rv.max()), "var_p999_rel_vs_dmma": float(np.quantile(rv, 0.999)),
            "eff_max_rel_to_max_vs_dmma": float(re.max()),
            "var_max_rel_vs_longdouble": float(np.max(np.abs(got["var"][pick] - exact) / np.abs(exact))),
            "rechecked_pairs": rechecked, "rechecked_frac": rechecked / float(len(ref)),
            "without_guard_var_max_rel_vs_dmma": float(rv0.max())}
        print(S, "minor-allele coding" if minor else "plain dosage",
              json.dumps(out["planes" if minor else "planes_plain_dosage"][str(S)]), flush=True)
    engine.I8Exact.MINOR_ALLELE = True
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/i8_accuracy.json", "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
