#!/usr/bin/env python
"""Pair-list engines side by side at the bench shapes: 100,000 random pairs, FP64 DMMA pair kernel vs int8 (I8Pairs)."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from gmat_b200 import engine
import synth
dev = engine.current_device()
out = {}
for n, m in ((2000, 20000), (5000, 50000)):
    store = engine.GenotypeStore.from_device_payload(synth.device_payload(n, m, 11, dev, engine.row_stride(n)), n)
    gen = torch.Generator(device=dev); gen.manual_seed(5)
    b = torch.randn((n, n + 8), dtype=torch.float64, device=dev, generator=gen)
    A = (b @ b.T / n).contiguous()
    A -= A.mean(dim=0, keepdim=True); A -= A.mean(dim=1, keepdim=True); A = (0.5 * (A + A.T)).contiguous()
    py = torch.randn((n,), dtype=torch.float64, device=dev, generator=gen); py = (py - py.mean()).contiguous()
    lu = store.lut(engine.ADD)
    rng = np.random.Generator(np.random.PCG64(3))
    pairs = rng.integers(0, m, size=(100000, 2)); pairs = pairs[pairs[:, 0] < pairs[:, 1]]
    packed = engine.pack_pvp(A)
    def t(f, reps=3):
        f(); torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(reps): r = f()
        torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps, r
    t64, ref = t(lambda: engine.epi_exact_pairs(store, lu, lu, py, packed, pairs))
    t0 = time.perf_counter(); ip = engine.I8Pairs(store, engine.ADD, engine.ADD, lu, lu, py, A, packed=packed); torch.cuda.synchronize()
    tinit = time.perf_counter() - t0
    t8, got = t(lambda: ip.stats(pairs))
    rel = float(np.max(np.abs(got["var"] - ref["var"]) / np.abs(ref["var"])))
    out["n%d" % n] = {"pairs": int(len(pairs)), "f64_s": t64, "i8_s": t8, "i8_init_s": tinit, "speedup": t64 / t8,
                      "var_max_rel": rel, "rechecked": ip.last_rechecked}
    print(n, out["n%d" % n], flush=True)
print(json.dumps(out))
