"""Sharding of the pair space and the final hit gather.

Two layers, as in SURVEY section 8e:
 * ``parallel_lists``: the reference's manual ``parallel=[P, k]`` partition, arithmetic reproduced verbatim
   (remma_epiAA.py:125-137; AD variant remma_epiAD.py:130-139) -- file-level drop-in (``out_file.k``).
 * ``shard_first_snps`` / ``gather_hits``: one process per GPU (torch.distributed).  First-SNP blocks are dealt
   to the least-loaded rank by pair count so the triangular work is balanced; the scan itself has no
   collective; hit records (32 bytes each) are gathered to rank 0 once at the end (NCCL on GPUs, gloo in
   CPU tests).
"""
import numpy as np
import torch
import torch.distributed as dist

HIT_BYTES = 32


_SINGLE = [False]


class single_rank(object):
    """Context manager: inside it the front-ends behave as a one-process job even though a process group exists
    (bench.py uses it on rank 0 to compare a sharded scan with the same scan on one GPU)."""

    def __enter__(self):
        _SINGLE[0] = True
        return self

    def __exit__(self, *exc):
        _SINGLE[0] = False
        return False


def rank_world():
    if not _SINGLE[0] and dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def parallel_lists(num_snp, parallel, all_pairs=False):
    """First-SNP list of part ``parallel[1]`` (1-based) of ``parallel[0]`` parts, and the four cut points."""
    num_snp_part = int(num_snp / (2 * parallel[0]))
    snp_pos_0 = (parallel[1] - 1) * num_snp_part
    snp_pos_1 = parallel[1] * num_snp_part
    snp_pos_2 = (2 * parallel[0] - parallel[1]) * num_snp_part
    snp_pos_3 = (2 * parallel[0] - parallel[1] + 1) * num_snp_part
    if parallel[1] == 1:
        snp_pos_3 = num_snp if all_pairs else num_snp - 1
    snp_list_0 = list(range(snp_pos_0, snp_pos_1)) + list(range(snp_pos_2, snp_pos_3))
    return snp_list_0, [snp_pos_0, snp_pos_1, snp_pos_2, snp_pos_3]


SHARD_BLOCK = 32        # first-SNPs per dealt block


def shard_first_snps(snp_lst_0, num_snp, all_pairs, rank, world):
    """This rank's share of the first-SNP list.  Blocks of SHARD_BLOCK consecutive list entries are dealt, in list
    order, to the rank that has the fewest PAIRS so far (a first SNP i has num_snp - 1 - i of them in the triangular
    scans, num_snp in the AD scan; ties go to the lowest rank), so the ranks' pair counts differ by less than one
    block whatever the list looks like.  Every rank computes the same deal; list order is kept within a rank."""
    snp_lst_0 = np.asarray(list(snp_lst_0), dtype=np.int64)
    if world == 1:
        return snp_lst_0.tolist()
    weight = np.full(len(snp_lst_0), num_snp, dtype=np.int64) if all_pairs else np.maximum(num_snp - 1 - snp_lst_0, 0)
    load = [0] * world
    mine = []
    for b0 in range(0, len(snp_lst_0), SHARD_BLOCK):
        owner = min(range(world), key=lambda r: (load[r], r))
        load[owner] += int(weight[b0:b0 + SHARD_BLOCK].sum())
        if owner == rank:
            mine.append(snp_lst_0[b0:b0 + SHARD_BLOCK])
    return np.concatenate(mine).tolist() if mine else []


def _blocks(snp_lst_0, num_snp, all_pairs):
    snp_lst_0 = np.asarray(list(snp_lst_0), dtype=np.int64)
    weight = np.full(len(snp_lst_0), num_snp, dtype=np.int64) if all_pairs else np.maximum(num_snp - 1 - snp_lst_0, 0)
    starts = np.arange(0, len(snp_lst_0), SHARD_BLOCK)
    wblk = np.add.reduceat(weight, starts) if len(starts) else np.zeros(0, dtype=np.int64)
    return snp_lst_0, starts, wblk


def deal_blocks(wblk, rates, load=None):
    """Owner of every block: in order, to the rank that would FINISH it first, (load + w) / rate (ties: lowest rank).
    With equal rates this is the deal of shard_first_snps."""
    world = len(rates)
    load = [0.0] * world if load is None else [float(v) for v in load]
    owner = np.empty(len(wblk), dtype=np.int64)
    for b, w in enumerate(wblk.tolist()):
        r = min(range(world), key=lambda q: ((load[q] + w) / rates[q], q))
        owner[b] = r
        load[r] += w
    return owner


CALIBRATION_SHARE = 0.08


def calibration_split(snp_lst_0, num_snp, all_pairs, rank, world, share=CALIBRATION_SHARE):
    """Speed-aware sharding, step 1.  GPUs of one box do not run at the same clock under the power cap (a few per cent
    apart), and a static deal by pair count makes every rank wait for the slowest one.  So: the static deal's first
    blocks of every rank, ``share`` of its pairs, are scanned first and timed (``head``, this rank's first SNPs);
    the rest of ALL ranks' blocks (``rest``: (snp_lst_0 array, block starts, block weights) in list order) is dealt again
    by ``redeal`` once the ranks have exchanged their measured rates.  Every rank computes the same split."""
    lst, starts, wblk = _blocks(snp_lst_0, num_snp, all_pairs)
    owner = deal_blocks(wblk, [1.0] * world)
    total = np.array([wblk[owner == r].sum() for r in range(world)], dtype=np.float64)
    in_head = np.zeros(len(wblk), dtype=bool)
    acc = [0.0] * world
    for b in range(len(wblk)):
        r = int(owner[b])
        if acc[r] < share * total[r]:
            in_head[b] = True
            acc[r] += float(wblk[b])
    head = [lst[starts[b]:starts[b] + SHARD_BLOCK] for b in range(len(wblk)) if in_head[b] and owner[b] == rank]
    head = np.concatenate(head).tolist() if head else []
    rest_blocks = np.flatnonzero(~in_head)
    return head, float(acc[rank]), (lst, starts[rest_blocks], wblk[rest_blocks])


def redeal(rest, rates, rank):
    """Speed-aware sharding, step 2: this rank's first SNPs of the remaining blocks, dealt by the measured rates
    (pairs per second of every rank's calibration scan)."""
    lst, starts, wblk = rest
    owner = deal_blocks(wblk, [float(v) for v in rates])
    mine = [lst[s:s + SHARD_BLOCK] for s, o in zip(starts.tolist(), owner.tolist()) if o == rank]
    return np.concatenate(mine).tolist() if mine else []


def all_gather_floats(value, device):
    """One float of every rank, as a list in rank order (all ranks get the same list)."""
    rank, world = rank_world()
    if world == 1:
        return [float(value)]
    dev = device if dist.get_backend() == "nccl" else torch.device("cpu")
    t = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    out = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    return [float(v.item()) for v in out]


def gather_hits(local, device):
    """Gather variable-length hit arrays (engine.HIT_DTYPE) to rank 0.  Returns the concatenation on rank 0,
    None elsewhere.  One all_gather of counts + one padded all_gather of records."""
    rank, world = rank_world()
    if world == 1:
        return local
    backend = dist.get_backend()
    dev = device if backend == "nccl" else torch.device("cpu")
    cnt = torch.tensor([len(local)], dtype=torch.int64, device=dev)
    counts = [torch.zeros_like(cnt) for _ in range(world)]
    dist.all_gather(counts, cnt)
    counts = [int(c.item()) for c in counts]
    mx = max(counts)
    if mx == 0:
        return local if rank == 0 else None
    buf = torch.zeros((mx, HIT_BYTES), dtype=torch.uint8, device=dev)
    if len(local):
        raw = np.frombuffer(np.ascontiguousarray(local).tobytes(), dtype=np.uint8).reshape(-1, HIT_BYTES)
        buf[:len(local)] = torch.from_numpy(raw.copy()).to(dev)
    bufs = [torch.zeros_like(buf) for _ in range(world)]
    dist.all_gather(bufs, buf)
    if rank != 0:
        return None
    parts = [bufs[r][:counts[r]].cpu().numpy().view(local.dtype).reshape(-1) for r in range(world)]
    return np.concatenate(parts)
