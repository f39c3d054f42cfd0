#!/usr/bin/env python
"""DMMA + DMUL mix probe (see gmat_b200_dmma_dmul_probe): does a DMUL in the DMMA stream cost more than its slot?"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gmat_b200 import engine
from probe import ffi, lib


def check(rc):
    assert rc >= 1, ffi.string(lib.gmat_b200_last_error())
engine.current_device()
out = ffi.new("double *")
res = []
for nmul, mode in [(0, 0), (1, 0), (4, 0), (4, 1), (4, 2), (8, 0), (8, 1)]:
    for warps in (4, 8):
        check(lib.gmat_b200_dmma_dmul_probe(warps, nmul, mode, 20000, out))
        res.append({"dmul_per_32_dmma": nmul, "mode": ["grouped", "spread", "grouped+dependent"][mode], "warps_per_sm": warps,
                    "dmma_tflops": round(out[0], 2)})
        print(res[-1], flush=True)
json.dump(res, open("gpurun_out/dmma_dmul_probe.json", "w"))
