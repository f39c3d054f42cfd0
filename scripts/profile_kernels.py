#!/usr/bin/env python
"""One launch of every hot kernel at a moderate, roofline-relevant size -- the command that is run first
plain and then under ``ncu --set full`` (see scripts/gpu_profile.sh).  Prints CUDA-event times so the plain
run doubles as a sanity check; numbers printed under ncu are never bench values.

    python scripts/profile_kernels.py [--only exact,eff,grm,decode,stats] [--n 2000] [--m 20000]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="i8,exact,eff,grm,decode,stats")
    ap.add_argument("--n", type=int, default=2000)
    ap.add_argument("--m", type=int, default=20000)
    ap.add_argument("--first", type=int, default=64, help="first-SNPs of the FP64 exact-scan launch")
    ap.add_argument("--i8-first", type=int, default=144, help="first-SNPs of the int8 scan (16 + 128: two launches)")
    ap.add_argument("--missing", type=float, default=0.005, help="missing-call rate of the second screen store")
    ap.add_argument("--grm-n", type=int, default=4096)
    ap.add_argument("--grm-m", type=int, default=65536)
    ap.add_argument("--dec-n", type=int, default=5000)
    ap.add_argument("--dec-m", type=int, default=50000)
    ap.add_argument("--reps", type=int, default=1)
    args = ap.parse_args()
    only = set(args.only.split(","))
    import torch
    from gmat_b200 import engine
    import synth
    dev = engine.current_device()
    out = {}

    def timed(fn):
        best = 1e30
        for _ in range(args.reps + 1):             # first call is the warm-up
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return best * 1e-3

    n, m = args.n, args.m
    rb = engine.row_stride(n)
    store = engine.GenotypeStore.from_device_payload(synth.device_payload(n, m, 11, dev, rb), n)
    if "stats" in only:
        t = timed(lambda: engine.GenotypeStore.from_device_payload(store.bed, n))
        out["bed_stats"] = {"seconds": t, "gb_per_s": m * rb / t / 1e9}
    if only & {"exact", "eff", "i8"}:
        gen = torch.Generator(device=dev)
        gen.manual_seed(5)
        b = torch.randn((n, n + 8), dtype=torch.float64, device=dev, generator=gen)
        pvp = b @ b.T / n
        pvp = pvp - pvp.mean(dim=0, keepdim=True)
        pvp = pvp - pvp.mean(dim=1, keepdim=True)
        py = torch.randn((n,), dtype=torch.float64, device=dev, generator=gen)
        py = py - py.mean()
    if "exact" in only:
        packed = engine.pack_pvp(pvp)
        lut = store.lut(engine.ADD)
        # contiguous blocks of 8 first-SNPs (one I-block of the scan tiling), blocks spread over the range --
        # what a full scan looks like to the kernel; isolated first-SNPs would leave 7/8 of a tile's rows masked
        nblk = max(1, args.first // 8)
        starts = np.linspace(0, m - 10, nblk).astype(np.int64)
        il = np.unique((starts[:, None] + np.arange(8)[None, :]).ravel())
        pairs = int(np.sum(m - 1 - il))
        t = timed(lambda: engine.epi_exact_scan(store, lut, lut, py, packed, il, engine.PAIRS_UPPER, 30.0))
        out["exact"] = {"seconds": t, "pairs": pairs, "pairs_per_s": pairs / t,
                        "tflops_algorithmic": pairs * (2.0 * n * n + 5 * n) / t / 1e12}
    if "i8" in only:
        lut = store.lut(engine.ADD)
        pvp = pvp.contiguous()
        sc = engine.I8Exact(store, lut, lut, engine.ADD, py, pvp)
        nblk = max(1, args.i8_first // 8)
        il = np.unique((np.linspace(0, m - 10, nblk).astype(np.int64)[:, None] + np.arange(8)[None, :]).ravel())
        pairs = int(np.sum(m - 1 - il))
        t = timed(lambda: sc.scan(il, engine.PAIRS_UPPER, 30.0))
        out["i8"] = {"seconds": t, "pairs": pairs, "pairs_per_s": pairs / t, "planes": sc.n_planes,
                     "first_snps": len(il), "rechecked": sc.last_rechecked,
                     "int8_tops_algorithmic": pairs * sc.n_planes * float(n) * (n + 1) / t / 1e12}
        del sc
    if "eff" in only:
        la = store.lut(engine.ADD, engine.FREQ_CLOOP)
        full = np.arange(m - 1, dtype=np.int64)
        pairs = m * (m - 1) // 2
        t = timed(lambda: engine.epi_eff(store, la, la, py, full, 0, 1e30))
        out["eff"] = {"seconds": t, "pairs": pairs, "pairs_per_s": pairs / t,
                      "tflops_algorithmic": pairs * 2.0 * n / t / 1e12}
        # the int8 tensor-core screen (prefilter + FP64 re-test) at a cut that keeps ~1e-5 of the pairs, on prepared
        # operands; then the same store with missing calls (tripled coding, three base-32 planes)
        smp = engine.epi_eff(store, la, la, py, np.arange(0, m - 1, 97, dtype=np.int64), 0, 0.0)
        cut = float(np.quantile(np.abs(smp["eff"]), 1.0 - 1e-5))
        ops = engine.ScreenOperands(store, py)
        t = timed(lambda: engine.epi_eff(store, la, la, py, full, 0, cut, codings=(engine.ADD, engine.ADD), ops=ops))
        out["eff_i8"] = {"seconds": t, "pairs": pairs, "pairs_per_s": pairs / t, "engine": engine.LAST_SCREEN_ENGINE,
                         "int8_tops_algorithmic": pairs * 4.0 * n / t / 1e12}
        del ops
        if args.missing > 0:
            bed = store.bed.clone()
            gen2 = torch.Generator(device=dev)
            gen2.manual_seed(9)
            hole = torch.rand(bed.shape, generator=gen2, device=dev) < args.missing * 4
            slot = torch.randint(0, 4, bed.shape, generator=gen2, device=dev, dtype=torch.uint8)
            clear = (~(torch.tensor(3, dtype=torch.uint8, device=dev) << (2 * slot))) & 0xFF
            bed = torch.where(hole, (bed & clear) | (torch.tensor(1, dtype=torch.uint8, device=dev) << (2 * slot)), bed)
            stm = engine.GenotypeStore.from_device_payload(bed, n)
            lm = stm.lut(engine.ADD, engine.FREQ_CLOOP)
            opm = engine.ScreenOperands(stm, py)
            t = timed(lambda: engine.epi_eff(stm, lm, lm, py, full, 0, cut, codings=(engine.ADD, engine.ADD), ops=opm))
            out["eff_i8_missing"] = {"seconds": t, "pairs": pairs, "pairs_per_s": pairs / t,
                                     "engine": engine.LAST_SCREEN_ENGINE, "missing_calls": stm.n_missing(),
                                     "int8_tops_algorithmic": pairs * 6.0 * n / t / 1e12}
            del opm, stm
    if "decode" in only:
        n2, m2 = args.dec_n, args.dec_m
        st2 = engine.GenotypeStore.from_device_payload(synth.device_payload(n2, m2, 12, dev, engine.row_stride(n2)), n2)
        nbytes = m2 * ((n2 + 3) // 4) + 8.0 * m2 * n2
        dec = torch.empty((m2, n2), dtype=torch.float64, device=dev)
        t = timed(lambda: engine.decode_into(st2, None, dec))
        out["decode_raw"] = {"seconds": t, "gb_per_s": nbytes / t / 1e9}
        t = timed(lambda: engine.decode_into(st2, engine.ADD, dec))
        out["decode_add"] = {"seconds": t, "gb_per_s": nbytes / t / 1e9}
        del st2, dec
    if "grm" in only:
        n2, m2 = args.grm_n, args.grm_m
        st2 = engine.GenotypeStore.from_device_payload(synth.device_payload(n2, m2, 13, dev, engine.row_stride(n2)), n2)
        st2.lut(engine.ADD)
        kin = torch.empty((n2, n2), dtype=torch.float64, device=dev)
        for which in ("f4", "i8", "f64"):                # FP4 (default) and int8 tensor-core GEMMs, FP64 DMMA kernel
            t = timed(lambda: engine.grm(st2, engine.ADD, 1.0, 0.001, out=kin, which=which))
            out["grm_" + which] = {"seconds": t, "tflops_algorithmic_n(n+1)m": float(n2) * (n2 + 1) * m2 / t / 1e12}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
