#!/usr/bin/env python
"""int8 tensor-core GRM against the FP64 DMMA GRM kernel: max difference and time (run on a B200)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    from gmat_b200 import engine
    import synth
    n, m = int(sys.argv[1]), int(sys.argv[2])
    dev = engine.current_device()
    store = engine.GenotypeStore.from_device_payload(synth.device_payload(n, m, 13, dev, engine.row_stride(n)), n)
    out = {"n": n, "m": m}
    for coding, tag in ((engine.ADD, "add"), (engine.DOM, "dom")):
        store.lut(coding)
        res = {}
        for which in ("i8", "f64"):
            kin = torch.empty((n, n), dtype=torch.float64, device=dev)
            best = 1e30
            for it in range(3):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                engine.grm(store, coding, 1234.5, 0.001, out=kin, which=which)
                e1.record()
                torch.cuda.synchronize()
                if it:
                    best = min(best, e0.elapsed_time(e1))
            res[which] = kin
            out["%s_%s_ms" % (tag, which)] = best
        out[tag + "_max_rel_diff"] = float((res["i8"] - res["f64"]).abs().max() / res["f64"].abs().max())
        out[tag + "_symmetric"] = bool(torch.equal(res["i8"], res["i8"].T))
        part = engine.grm(store, coding, 1.0, 0.0, snp_begin=64, snp_end=m - 7, accumulate=True, which="i8")
        ref = engine.grm(store, coding, 1.0, 0.0, snp_begin=64, snp_end=m - 7, accumulate=True, which="f64")
        out[tag + "_partial_max_rel_diff"] = float((part - ref).abs().max() / ref.abs().max())
    print(json.dumps(out))


if __name__ == "__main__":
    main()
