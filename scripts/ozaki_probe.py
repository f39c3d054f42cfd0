#!/usr/bin/env python
"""Experiment (not product code): can the n^2-per-pair part of the exact test run on the INT8 tensor cores
without giving up FP64-level results?

    w_ij = u_i o v_j,  v_j = a_j - o_j   (a_j in {0,1,2} integer dosage, o_j = 2 p_j; no missing in j)
    var_ij = w' A w = a_j' B_i a_j - 2 o_j a_j.r_i + o_j^2 s_i
             B_i = diag(u_i) A diag(u_i),  r_i = u_i o (A u_i),  s_i = sum(r_i)

Only a_j' B_i a_j costs n^2 per pair, and a_j is an exact small integer vector.  B_i is split error-free into S
signed base-128 digit planes (int8); each plane times the integer genotype matrix is an exact int8 x int8 -> int32
GEMM; the planes are recombined in FP64.  This script measures (a) the relative error of var against the FP64
DMMA kernel and against a float128 host evaluation as a function of S, (b) the int8 GEMM rate cuBLASLt reaches on
this GPU for the shapes involved (torch._int_mm), next to the FP64 kernel's pairs/s.
"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def digit_planes(B, S):
    """Error-free split of B/scale into S signed base-128 digits: B ~ scale * sum_s d_s 128^-(s+1)."""
    import torch
    scale = float(B.abs().max()) / 0.99
    r = B / scale
    planes = []
    for _ in range(S):
        r = r * 128.0
        d = torch.round(r)
        r = r - d
        planes.append(d.to(torch.int8))
    return planes, scale


def main():
    import torch
    from gmat_b200 import engine, synth
    from gmat_b200.gmatrix.gmatrix import grm_device
    from gmat_b200.remma import _core
    out = {}
    dev = engine.current_device()
    n, m = int(os.environ.get("OZ_N", 2000)), int(os.environ.get("OZ_M", 20000))
    payload = synth.device_payload(n, m, 20201013 + n, dev, engine.row_stride(n))
    store = engine.GenotypeStore.from_device_payload(payload, n)
    lut = store.lut(engine.ADD)
    # a realistic Z'PZ: GRMs of this very data set, fixed variance components, one record per id, X = intercept
    ag = grm_device(store, engine.ADD)[0]
    gen = torch.Generator(device=dev)
    gen.manual_seed(3)
    y = torch.randn((n, 1), dtype=torch.float64, device=dev, generator=gen)
    V = 0.35 * ag + 0.10 * ag * ag + 0.55 * torch.eye(n, dtype=torch.float64, device=dev)
    Vi = torch.linalg.inv(V)
    X = torch.ones((n, 1), dtype=torch.float64, device=dev)
    VX = Vi @ X
    P = Vi - VX @ torch.linalg.inv(X.T @ VX) @ VX.T
    A = 0.5 * (P + P.T)
    py = (P @ y).reshape(-1).contiguous()
    packed = engine.pack_pvp(A)

    G = store.decode(None)                               # [m][n] raw dosage 0/1/2 float64
    G8 = G.to(torch.int8).contiguous()                   # [m][n], K(=id)-major: the TN operand cuBLASLt wants
    del G
    o = lut[:, 0].neg().contiguous()                     # lut[j][code 0] = 0 - 2p  -> o_j = 2 p_j
    firsts = [5, m // 3, m - 4000]
    mj = 2048                                            # second SNPs checked per first SNP
    res = []
    for i in firsts:
        js = torch.arange(i + 1, min(m, i + 1 + mj), device=dev)
        ui = G8[i].double() - o[i]
        Bi = ui[:, None] * A * ui[None, :]
        ri = ui * (A @ ui)
        si = ri.sum()
        Gj = G8[js]                                      # [mj][n] int8
        T1 = Gj.double() @ ri
        oj = o[js]
        pairs = np.stack([np.full(len(js), i, dtype=np.int64), js.cpu().numpy()], axis=1)
        ref = engine.epi_exact_pairs(store, lut, lut, py, packed, pairs)      # FP64 DMMA kernel
        # float128 host evaluation of a few pairs
        Ah = A.cpu().numpy().astype(np.longdouble)
        uh = ui.cpu().numpy().astype(np.longdouble)
        sel = list(range(0, len(js), max(1, len(js) // 12)))[:12]
        exact = []
        for q in sel:
            vj = (Gj[q].cpu().numpy().astype(np.longdouble) - np.longdouble(float(oj[q])))
            w = uh * vj
            exact.append(float(w @ (Ah @ w)))
        exact = np.array(exact)
        row = {"first": i, "err_dmma_vs_f128": float(np.max(np.abs(ref["var"][sel] - exact) / np.abs(exact)))}
        for S in (6, 7, 8):
            planes, scale = digit_planes(Bi, S)
            Q = torch.zeros(len(js), dtype=torch.float64, device=dev)
            for s in reversed(range(S)):                 # smallest plane first
                Y = torch._int_mm(planes[s], Gj.t())     # [n][mj] int32, exact
                qs = (Y.long() * Gj.t().long()).sum(dim=0)           # sum_k a_jk Y[k][j], exact int64
                Q += qs.double() * (scale * 128.0 ** -(s + 1))
            var = (Q - 2.0 * oj * T1 + oj * oj * si).cpu().numpy()
            row["S%d_vs_dmma" % S] = float(np.max(np.abs(var - ref["var"]) / np.abs(ref["var"])))
            row["S%d_vs_f128" % S] = float(np.max(np.abs(var[sel] - exact) / np.abs(exact)))
        res.append(row)
        print(json.dumps(row), flush=True)
    out["accuracy"] = res

    # ---- attainable int8 GEMM rate (cuBLASLt through torch._int_mm) for [S*n x n] @ [n x m] -------------
    rates = []
    for S, mm in ((8, 8192), (8, m // 8 * 8)):
        a = torch.randint(-127, 127, (S * n, n), dtype=torch.int8, device=dev)
        b = G8[:mm].t()
        for _ in range(2):
            torch._int_mm(a, b)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 5
        e0.record()
        for _ in range(reps):
            torch._int_mm(a, b)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        macs = float(S * n) * n * mm
        rates.append({"M": S * n, "K": n, "N": mm, "ms": ms, "int8_TOPS": 2 * macs / ms / 1e9,
                      "pairs_per_s_if_S_planes": mm / (ms * 1e-3)})
        print(json.dumps(rates[-1]), flush=True)
        del a
    out["int8_gemm"] = rates
    # bf16 for comparison of the library's tensor-core rate on the same shape
    a = torch.randn((8 * n, n), dtype=torch.bfloat16, device=dev)
    b = torch.randn((n, 8192), dtype=torch.bfloat16, device=dev)
    for _ in range(2):
        a @ b
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        a @ b
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    out["bf16_gemm_TFLOPS"] = 2.0 * 8 * n * n * 8192 / ms / 1e9
    print(json.dumps(out))
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/ozaki_probe.json", "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
