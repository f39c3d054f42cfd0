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


def impute_geno(snp_mat):
    """Random imputation by observed genotype frequency, column by column, with the global numpy RNG
    (restated from process_plink.py:12-25 -- same column order, same draws for the same RNG state)."""
    ind_na = np.where(np.isnan(snp_mat))
    col_na = set(ind_na[1])
    for i in col_na:
        snpi = snp_mat[:, i]
        code0 = np.sum(np.absolute(snpi - 0.0) < 1e-10)
        code1 = np.sum(np.absolute(snpi - 1.0) < 1e-10)
        code2 = np.sum(np.absolute(snpi - 2.0) < 1e-10)
        code_count = code0 + code1 + code2
        p_lst = [code0 / code_count, code1 / code_count, code2 / code_count]
        icol_na = np.where(np.isnan(snpi))
        snpi[icol_na] = np.random.choice([0.0, 1.0, 2.0], len(icol_na[0]), p=p_lst)
        snp_mat[:, i] = snpi
    return snp_mat


def load_store(bed_file, warn=logging.warning):
    """Device genotype store for the numpy-flavoured paths: if the file has missing calls they are
    imputed exactly like the reference does before any statistic is formed (remma_epiAA.py:56-58),
    then re-packed to 2 bits on the device."""
    store = GenotypeStore(bed_file)
    if store.n_missing() > 0:
        warn('Missing genotypes are imputed with random genotypes.')
        dense = store.decode(None).t().contiguous().cpu().numpy()
        dense[np.abs(dense - 1.0 / 3) < 0.0001] = np.nan
        dense = impute_geno(dense)
        store = GenotypeStore.from_dense(np.ascontiguousarray(dense.T), device=store.device)
    return store
