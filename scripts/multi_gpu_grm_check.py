#!/usr/bin/env python
"""Sharded (one process per GPU) GRM against the single-GPU GRM: launched under torchrun by
tests/test_gpu_multi.py and by hand.  Every rank computes both; prints one JSON line on rank 0."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import torch.distributed as dist
    from gmat_b200 import engine
    import synth
    from gmat_b200.gmatrix.gmatrix import grm_device
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    n, m = int(sys.argv[1]), int(sys.argv[2])
    dev = engine.current_device()
    store = engine.GenotypeStore.from_device_payload(synth.device_payload(n, m, 5, dev, engine.row_stride(n)), n)
    single_a, sa = grm_device(store, engine.ADD)           # no process group yet: the single-GPU path
    single_d, sd = grm_device(store, engine.DOM)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    shard_a, sa2 = grm_device(store, engine.ADD)
    shard_d, sd2 = grm_device(store, engine.DOM)
    out = {"world": dist.get_world_size(), "n": n, "m": m, "scale_equal": bool(sa == sa2 and sd == sd2)}
    for name, a, b in (("add", single_a, shard_a), ("dom", single_d, shard_d)):
        out[name + "_max_rel"] = float((a - b).abs().max() / a.abs().max())
        out[name + "_symmetric"] = bool(torch.equal(b, b.t()))
        ref = [torch.empty_like(b) for _ in range(dist.get_world_size())]
        dist.all_gather(ref, b)
        out[name + "_identical_on_all_ranks"] = all(bool(torch.equal(ref[0], r)) for r in ref)
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
