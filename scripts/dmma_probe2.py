#!/usr/bin/env python
"""DMMA + LDS issue mix probe (see gmat_b200_dmma_lds_probe)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gmat_b200 import engine
from probe import ffi, lib


def check(rc):
    assert rc >= 1, ffi.string(lib.gmat_b200_last_error())
engine.current_device()
out = ffi.new("double *")
res = []
names = {0: "4x4 B=LDS.64", 1: "4x4 B=LDS.128", 2: "8x4 B=LDS.64", 3: "8x4 B=LDS.128", 4: "4x2 B=LDS.128"}
for shape in range(5):
    for warps, ctas in [(4, 1), (8, 1), (16, 1), (4, 4), (8, 2), (4, 2)]:
        if shape in (2, 3) and warps * ctas > 8:
            continue
        check(lib.gmat_b200_dmma_lds_probe(warps, shape, ctas, 20000, out))
        res.append({"shape": names[shape], "warps": warps, "ctas_per_sm": ctas, "tflops": round(out[0], 2)})
        print(res[-1], flush=True)
json.dump(res, open("gpurun_out/dmma_lds_probe.json", "w"))
