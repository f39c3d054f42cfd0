#!/usr/bin/env python
"""Two calls of the int8 pair-list engine at n=5000 x m=50000 on 50,000 random pairs (for an ncu capture of
i8_scan_rows_kernel; run once without ncu first)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from gmat_b200 import engine
import synth
dev = engine.current_device()
n, m = 5000, 50000
store = engine.GenotypeStore.from_device_payload(synth.device_payload(n, m, 11, dev, engine.row_stride(n)), n)
gen = torch.Generator(device=dev); gen.manual_seed(5)
b = torch.randn((n, n + 8), dtype=torch.float64, device=dev, generator=gen)
A = (b @ b.T / n).contiguous(); A -= A.mean(dim=0, keepdim=True); A -= A.mean(dim=1, keepdim=True)
A = (0.5 * (A + A.T)).contiguous()
py = torch.randn((n,), dtype=torch.float64, device=dev, generator=gen); py = (py - py.mean()).contiguous()
lu = store.lut(engine.ADD)
rng = np.random.Generator(np.random.PCG64(3))
pairs = rng.integers(0, m, size=(100000, 2)); pairs = pairs[pairs[:, 0] < pairs[:, 1]]
ip = engine.I8Pairs(store, engine.ADD, engine.ADD, lu, lu, py, A)
for _ in range(2):
    h = ip.stats(pairs)
torch.cuda.synchronize()
print(len(h), ip.last_rechecked)
