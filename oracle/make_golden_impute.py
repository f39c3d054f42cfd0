#!/usr/bin/env python
"""Golden vector for impute_geno: the UNMODIFIED reference function (gmat/process_plink/process_plink.py:12-25,
imported from /root/reference through oracle/ref_harness.py) on a small seeded genotype matrix with holes, under a
fixed state of the global numpy RNG.  Writes tests/golden/impute_geno.json (input and output, so the test needs no
reference at run time).  Run in the build container:  python oracle/make_golden_impute.py"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_harness  # noqa: E402


def main():
    ref = ref_harness.ref()
    rng = np.random.Generator(np.random.PCG64(20201013))
    n, m = 60, 37
    p = rng.uniform(0.05, 0.95, size=m)
    g = rng.binomial(2, p, size=(n, m)).astype(np.float64)
    g[rng.random((n, m)) < 0.07] = np.nan
    g[:, 5] = np.where(np.isnan(g[:, 5]), np.nan, 2.0)          # a column whose observed calls are all 2
    src = g.copy()
    np.random.seed(424242)
    out = ref.impute_geno(g)
    after = float(np.random.random())                            # RNG state after the call: same number of draws
    code = lambda a: [[-1 if np.isnan(v) else int(v) for v in row] for row in a]
    path = os.path.join(ROOT, "tests", "golden", "impute_geno.json")
    with open(path, "w") as f:
        json.dump({"seed": 424242, "input": code(src), "output": code(out), "next_random": after}, f)
    print("wrote", path, "holes:", int(np.isnan(src).sum()))


if __name__ == "__main__":
    main()
