"""Drop-in for gmat/process_plink/process_plink.py (read_plink, impute_geno).

``read_plink`` in the reference is ``pandas_plink.read_plink1_bin(...).values`` (process_plink.py:7-9):
a dense float matrix [n_id][n_snp] holding the count of the second .bim allele, NaN for missing.
Here the 2-bit payload is uploaded once and decoded by the CUDA decode kernel; dtype is float64
(the pin recorded in DESIGN.md: equals the reference's own C reader, _read_plink_bed.c:36-37).
"""
import logging

import numpy as np
import torch

from ..engine import GenotypeStore


def read_plink(bed_file):
    """PLINK .bed -> float64 ndarray [n_id][n_snp], NaN for missing (reference: process_plink.py:7-9)."""
    store = GenotypeStore(bed_file)
    dense = store.decode(None)                              # [snp][id], code 01 -> 1/3
    miss = (store.counts[:, 1] > 0).any().item()
    out = dense.t().contiguous().cpu().numpy()
    if miss:
        out[np.abs(out - 1.0 / 3) < 0.0001] = np.nan        # read_plink_bed.py:26
    return out


GENOTYPES = [0.0, 1.0, 2.0]


def impute_geno(snp_mat):
    """Random imputation by observed genotype frequency, in place (reference: process_plink.py:12-25).  Reproduces
    the reference's draws for the same state of the global numpy RNG: columns are visited in the iteration order
    of ``set(column indices of the NaNs)`` (row-major NaN order, as ``np.where`` lists them), and each column makes
    ONE ``np.random.choice`` call for all of its holes with p = the observed frequencies of 0 / 1 / 2."""
    holes = np.isnan(snp_mat)
    for col in set(np.nonzero(holes)[1]):
        gap = holes[:, col]
        seen = snp_mat[~gap, col]
        tally = np.array([np.count_nonzero(np.abs(seen - g) < 1e-10) for g in GENOTYPES])
        snp_mat[gap, col] = np.random.choice(GENOTYPES, int(gap.sum()), p=tally / tally.sum())
    return snp_mat


def load_store(bed_file, warn=logging.warning):
    """Device genotype store for the numpy-flavoured paths: if the file has missing calls they are
    imputed exactly like the reference does before any statistic is formed (remma_epiAA.py:56-58),
    then re-packed to 2 bits on the device.  One process per GPU: the reference imputes ONCE per call, so rank 0
    draws the genotypes and broadcasts the re-packed 2-bit store -- every rank then tests against the same matrix
    (per-rank draws from the unseeded global RNG would give every shard its own genotypes)."""
    from ..remma import parallel as par
    store = GenotypeStore(bed_file)
    if store.n_missing() > 0:
        warn('Missing genotypes are imputed with random genotypes.')
        rank, world = par.rank_world()
        if rank == 0:
            dense = store.decode(None).t().contiguous().cpu().numpy()
            dense[np.abs(dense - 1.0 / 3) < 0.0001] = np.nan
            dense = impute_geno(dense)
            store = GenotypeStore.from_dense(np.ascontiguousarray(dense.T), device=store.device)
        if world > 1:
            import torch.distributed as dist
            bed = store.bed if rank == 0 else torch.empty_like(store.bed)
            if dist.get_backend() == "nccl":
                dist.broadcast(bed, src=0)
            else:
                host = bed.cpu()
                dist.broadcast(host, src=0)
                bed = host.to(store.device)
            if rank != 0:
                store = GenotypeStore.from_device_payload(bed, store.n)
    return store
