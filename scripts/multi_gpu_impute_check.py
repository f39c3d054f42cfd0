#!/usr/bin/env python
"""One imputation for all ranks: under torchrun, load_store() of a .bed WITH missing calls must give every
rank the same imputed 2-bit store (rank 0 draws, broadcasts) -- although every rank seeds the global numpy RNG
differently -- and the sharded GRM built from it must be identical on all ranks.  Prints one JSON line on rank 0."""
import json
import os
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import torch.distributed as dist
    from gmat_b200 import engine
    import synth
    from gmat_b200.gmatrix.gmatrix import grm_device
    from gmat_b200.process_plink.process_plink import load_store
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    n, m = int(sys.argv[1]), int(sys.argv[2])
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    box = [None]
    if rank == 0:
        d = tempfile.mkdtemp(prefix="gmat_impute_")
        synth.write_dataset(os.path.join(d, "x"), n, m, seed=8, missing_rate=0.01)
        box[0] = os.path.join(d, "x")
    dist.broadcast_object_list(box, src=0)
    np.random.seed(100 + rank)                       # different RNG state per rank: only rank 0's draws may count
    store = load_store(box[0], warn=lambda *_: None)
    world = dist.get_world_size()
    beds = [torch.empty_like(store.bed) for _ in range(world)]
    dist.all_gather(beds, store.bed)
    kin, scale = grm_device(store, engine.ADD)
    kins = [torch.empty_like(kin) for _ in range(world)]
    dist.all_gather(kins, kin)
    scales = [None] * world
    dist.all_gather_object(scales, scale)
    out = {"world": world, "missing_in_file": int(engine.GenotypeStore(box[0]).n_missing()),
           "missing_after_impute": int(store.n_missing()),
           "store_identical_on_all_ranks": all(bool(torch.equal(beds[0], b)) for b in beds),
           "grm_identical_on_all_ranks": all(bool(torch.equal(kins[0], k)) for k in kins),
           "scale_identical": len(set(scales)) == 1}
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
