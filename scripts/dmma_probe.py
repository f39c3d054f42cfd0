#!/usr/bin/env python
"""DMMA (mma.sync m8n8k4 f64) issue model on this GPU: TFLOP/s vs resident warps per SM and register-tile shape."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gmat_b200 import engine
from probe import ffi, lib


def check(rc):
    assert rc >= 1, ffi.string(lib.gmat_b200_last_error())
engine.current_device()
out = ffi.new("double *")
res = []
for (ia, ib) in [(1, 1), (2, 2), (4, 2), (4, 4), (8, 4)]:
    for warps, ctas in [(4, 1), (8, 1), (12, 1), (16, 1), (20, 1), (24, 1), (32, 1), (16, 2)]:
        if (ia * ib == 32 and warps > 8) or (ia * ib == 16 and warps > 16):
            continue
        check(lib.gmat_b200_dmma_probe(warps, ia, ib, ctas, 4000, out))
        res.append({"tile": [ia, ib], "warps_per_sm": warps * ctas, "tflops": round(out[0], 2)})
        print(res[-1], flush=True)
json.dump(res, open("gpurun_out/dmma_probe.json", "w"))
