#!/usr/bin/env python
"""int8 MMA issue rate: a cold burst, then the same kernel back to back for a few seconds (power-capped rate)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import probe
torch.cuda.set_device(0)
out = {"burst_tops": probe.i8_peak(20000)}
t0 = time.time()
rates = []
while time.time() - t0 < 4.0:
    rates.append(probe.i8_peak(200000))
out["sustained_tops_series"] = [round(r, 1) for r in rates]
out["sustained_tops"] = sum(rates[-5:]) / len(rates[-5:])
print(json.dumps(out))
