#!/usr/bin/env python
"""Staged check of the INT8 tensor-core exact engine against the FP64 DMMA kernel (run on a B200):
operands (integer codes, digit planes), the quadratic form, the emitted statistics, and the rate."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ids", dest="n", type=int, default=300)
    ap.add_argument("--snps", dest="m", type=int, default=700)
    ap.add_argument("--first", type=int, default=5)
    ap.add_argument("--planes", type=int, default=8)
    ap.add_argument("--kind", default="AA")
    ap.add_argument("--operands", action="store_true", help="also verify codes and digit planes in torch")
    ap.add_argument("--time", action="store_true")
    args = ap.parse_args()
    import torch
    from gmat_b200 import engine, synth
    from gmat_b200.engine import lib, _p, _stream, check
    dev = engine.current_device()
    n, m = args.n, args.m
    payload = synth.device_payload(n, m, 20201013 + n, dev, engine.row_stride(n))
    store = engine.GenotypeStore.from_device_payload(payload, n)
    cu = engine.ADD if args.kind[0] == "A" else engine.DOM
    cv = engine.ADD if args.kind[1] == "A" else engine.DOM
    lu, lv = store.lut(cu), store.lut(cv)
    gen = torch.Generator(device=dev)
    gen.manual_seed(5)
    b = torch.randn((n, n + 8), dtype=torch.float64, device=dev, generator=gen)
    A = b @ b.T / n
    A = A - A.mean(dim=0, keepdim=True)
    A = A - A.mean(dim=1, keepdim=True)
    A = (0.5 * (A + A.T)).contiguous()
    py = torch.randn((n,), dtype=torch.float64, device=dev, generator=gen)
    py = (py - py.mean()).contiguous()
    packed = engine.pack_pvp(A)
    space = engine.PAIRS_ALL if args.kind == "AD" else engine.PAIRS_UPPER
    first = np.unique(np.linspace(0, m - 2, args.first).astype(np.int64))
    out = {"n": n, "m": m, "first": len(first), "planes": args.planes, "kind": args.kind}

    eng = engine.I8Exact(store, lu, lv, cv, py, A, n_planes=args.planes)
    if args.operands:
        raw = store.decode(None)
        want = raw if cv == engine.ADD else torch.where(raw > 1.5, torch.zeros_like(raw), raw)
        out["codes_ok"] = bool(torch.equal(eng.g8[:, :n].double(), want) and int(eng.g8[:, n:].abs().sum()) == 0)
        il = torch.as_tensor(first[:2]).to(dev)
        u = eng._coded_first(il)
        scale = (u.abs().amax(dim=1) ** 2 * (eng.amax / 0.99)).contiguous()
        check(lib.gmat_b200_i8_split(_p(A, "double *"), n, _p(u, "double *"), 2, _p(scale, "double *"), args.planes,
                                     _p(eng.planes[0], "signed char *"), _stream()), "split")
        pl = eng.planes[0][: 2 * args.planes * eng.ld * eng.ld].view(2, args.planes, eng.ld, eng.ld).double()
        rec = torch.zeros((2, eng.ld, eng.ld), dtype=torch.float64, device=dev)
        for s in reversed(range(args.planes)):
            rec += pl[:, s] * 128.0 ** -(s + 1)
        rec = rec * scale[:, None, None]
        B = torch.zeros((2, eng.ld, eng.ld), dtype=torch.float64, device=dev)
        B[:, :n, :n] = u[:, :, None] * A[None] * u[:, None, :]
        blk = torch.arange(eng.ld, device=dev) // 128
        lower = (blk[:, None] > blk[None, :]).double() + 0.5 * (blk[:, None] == blk[None, :]).double()
        out["planes_max_err_rel"] = float((rec - B * lower).abs().max() / B.abs().max())

    t0 = time.perf_counter()
    hits, qd = eng.scan(first, space, -1.0, cap=len(first) * m + 16, q_debug=True)
    torch.cuda.synchronize()
    out["scan_s_first_call"] = time.perf_counter() - t0
    ref = engine.epi_exact_scan(store, lu, lv, py, packed, first, space, -1.0, cap=len(first) * m + 16)
    # quadratic form against torch FP64
    il = torch.as_tensor(first).to(dev)
    u = eng._coded_first(il)
    gd = eng.g8[:, :n].double()
    q_ref = torch.stack([((gd * u[k]) @ A * (gd * u[k])).sum(dim=1) for k in range(min(len(first), 3))])
    qq = qd[: q_ref.shape[0]]
    mask = torch.ones_like(q_ref, dtype=torch.bool)
    if space == engine.PAIRS_UPPER:      # tiles left of the first needed tile are never computed
        for k in range(q_ref.shape[0]):
            mask[k, : (int(first[k]) + 1) // 128 * 128] = False
    out["quad_max_rel_err"] = float(((qq - q_ref).abs() / q_ref.abs().clamp_min(1e-300))[mask].max())
    key = lambda h: (h["snp_0"].astype(np.int64) * m + h["snp_1"])
    a, r = hits[np.argsort(key(hits))], ref[np.argsort(key(ref))]
    out["pairs"] = int(len(r))
    out["same_pair_set"] = bool(len(a) == len(r) and np.array_equal(key(a), key(r)))
    if out["same_pair_set"]:
        out["eff_max_rel"] = float(np.max(np.abs(a["eff"] - r["eff"]) / np.abs(r["eff"]).max()))
        out["var_max_rel"] = float(np.max(np.abs(a["var"] - r["var"]) / np.abs(r["var"])))
        out["chi_max_rel"] = float(np.max(np.abs(a["chi"] - r["chi"]) / np.maximum(np.abs(r["chi"]), 1e-12)))
    if args.time:
        for name, fn in (("i8", lambda: eng.scan(first, space, 20.0)),
                         ("f64", lambda: engine.epi_exact_scan(store, lu, lv, py, packed, first, space, 20.0))):
            fn()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            h = fn()
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            out[name + "_pairs_per_s"] = len(r) / dt
            out[name + "_hits"] = int(len(h))
    print(json.dumps(out))


if __name__ == "__main__":
    main()
