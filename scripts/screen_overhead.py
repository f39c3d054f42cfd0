#!/usr/bin/env python
"""Where does the time of one prepared int8 screen call go at the C3 shape (host timeline, not a bench value)."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from gmat_b200 import engine
from gmat_b200.engine import lib, ffi, _p, _stream, check
import synth
dev = engine.current_device()
n, m = 2000, 20000
store = engine.GenotypeStore.from_device_payload(synth.device_payload(n, m, 11, dev, engine.row_stride(n)), n)
gen = torch.Generator(device=dev); gen.manual_seed(5)
py = torch.randn((n,), dtype=torch.float64, device=dev, generator=gen); py = (py - py.mean()).contiguous()
la = store.lut(engine.ADD, engine.FREQ_CLOOP)
full = np.arange(m - 1, dtype=np.int64)
smp = engine.epi_eff(store, la, la, py, np.arange(0, m - 1, 97, dtype=np.int64), 0, 0.0)
out = {}
for tag, q in (("few_hits", 1 - 1e-5), ("many_hits", 1 - 5e-4)):
    cut = float(np.quantile(np.abs(smp["eff"]), q))
    ops = engine.ScreenOperands(store, py)
    for _ in range(3):
        h = engine.epi_eff(store, la, la, py, full, 0, cut, codings=(engine.ADD, engine.ADD), ops=ops)
    torch.cuda.synchronize()
    # phase timing
    T = {}
    t0 = time.perf_counter()
    il = torch.as_tensor(full).to(dev)
    count = torch.zeros((3,), dtype=torch.int64, device=dev)
    hits = engine.alloc_hits(1 << 20, dev)
    g8, t = ops.operand(engine.ADD)
    torch.cuda.synchronize(); T["upload_alloc"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    check(lib.gmat_b200_epi_eff_i8p(
        _p(store.bed, "unsigned char *"), n, m, store.row_bytes, _p(la, "double *"), _p(la, "double *"),
        _p(g8, "signed char *"), _p(g8, "signed char *"), _p(t, "double *"), _p(t, "double *"),
        _p(ops.dig, "signed char *"), _p(ops.stats, "double *"), ops.tripled, _p(py, "double *"),
        _p(il, "long long *"), il.numel(), 0, cut, ffi.NULL, ffi.NULL, ffi.NULL,
        ffi.cast("gmat_b200_hit *", hits.data_ptr()), 1 << 20, ffi.cast("unsigned long long *", count.data_ptr()),
        _p(ops.cand, "int *"), ops.CAND_CAP, ffi.cast("unsigned long long *", count.data_ptr() + 8),
        ffi.cast("int *", count.data_ptr() + 16), _stream()), "i8p")
    T["c_call_host_side"] = time.perf_counter() - t0
    torch.cuda.synchronize(); T["c_call_until_gpu_done"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    cnt = count.tolist(); T["count_read"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    res = engine.hits_to_numpy(hits, cnt[0]); T["hits_d2h"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    h = engine.epi_eff(store, la, la, py, full, 0, cut, codings=(engine.ADD, engine.ADD), ops=ops)
    torch.cuda.synchronize(); T["whole_epi_eff_call"] = time.perf_counter() - t0
    out[tag] = {"hits": cnt[0], "candidates": cnt[1], **{k: round(v * 1e3, 4) for k, v in T.items()}}
print(json.dumps(out))
import cProfile, pstats, io
cut = float(np.quantile(np.abs(smp["eff"]), 1 - 1e-5))
ops = engine.ScreenOperands(store, py)
ts = []
for _ in range(30):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    engine.epi_eff(store, la, la, py, full, 0, cut, codings=(engine.ADD, engine.ADD), ops=ops)
    torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
print("epi_eff ms: min %.3f median %.3f max %.3f" % (min(ts), sorted(ts)[15], max(ts)))
pr = cProfile.Profile(); pr.enable()
for _ in range(50):
    engine.epi_eff(store, la, la, py, full, 0, cut, codings=(engine.ADD, engine.ADD), ops=ops)
torch.cuda.synchronize()
pr.disable()
sio = io.StringIO(); pstats.Stats(pr, stream=sio).sort_stats("tottime").print_stats(14); print(sio.getvalue()[:3500])
