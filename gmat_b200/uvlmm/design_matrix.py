"""Phenotype file + .fam -> (y, X, Z) for the mixed model: drop-in for
``gmat.uvlmm.design_matrix.design_matrix_wemai_multi_gmat`` (gmat/uvlmm/design_matrix.py:7-57), same three return
values for the same files.

Caller-side glue of the hot path: text parsing on the host, done as array operations.
Contract reproduced from the reference:
  * a phenotype row is "<fam> <id> <covariates ...> <value>"; rows whose value is NA / NaN / nan / na are dropped;
  * records are emitted in .fam order, the records of one genotyped individual contiguously in file order;
  * every genotyped individual needs at least one record (else the reference's error message and ``sys.exit()``);
    phenotyped individuals that are not genotyped are ignored;
  * a .fam row listed twice emits its records twice (the reference iterates over the .fam lines);
  * Z (N records x n individuals, CSR, ones) gets one column per distinct INDIVIDUAL id -- the second .fam column
    alone -- numbered in order of first appearance.
"""
import logging
import sys

import numpy as np
from scipy.sparse import csr_matrix

MISSING_VALUE_TOKENS = frozenset(('NA', 'NaN', 'nan', 'na'))


def _rows(path):
    with open(path) as handle:
        return [fields for fields in (line.split() for line in handle) if fields]


def design_matrix_wemai_multi_gmat(pheno_file, bed_file):
    fam_keys = np.array([f[0] + " " + f[1] for f in _rows(bed_file + '.fam')], dtype=object)
    pheno = [f for f in _rows(pheno_file) if f[-1] not in MISSING_VALUE_TOKENS]
    rec_keys = np.array([f[0] + " " + f[1] for f in pheno], dtype=object)
    # every record gets the code of its (fam, id) key; genotyped keys are looked up in the same code book
    book, rec_code = np.unique(rec_keys, return_inverse=True) if len(pheno) else (np.array([], dtype=object), np.array([], dtype=np.int64))
    slot = np.searchsorted(book, fam_keys)
    slot_ok = slot < len(book)
    known = np.zeros(len(fam_keys), dtype=bool)
    known[slot_ok] = book[slot[slot_ok]] == fam_keys[slot_ok]
    if not known.all():
        absent = list(dict.fromkeys(fam_keys[~known].tolist()))
        logging.error('The below genotyped id is not in the phenotype file:\n {}'.format('\n'.join(absent)))
        sys.exit()
    # records grouped by key, file order kept inside a group (stable sort), then concatenated in .fam order
    by_key = np.argsort(rec_code, kind='stable')
    starts = np.searchsorted(rec_code[by_key], np.arange(len(book)), side='left')
    stops = np.searchsorted(rec_code[by_key], np.arange(len(book)), side='right')
    take = np.concatenate([by_key[starts[s]:stops[s]] for s in slot]) if len(slot) else np.array([], dtype=np.int64)
    n_rec = len(take)
    y = np.array([float(pheno[r][-1]) for r in take], dtype=np.float64).reshape(-1, 1)
    xmat = np.array([pheno[r][2:-1] for r in take], dtype=float).reshape(n_rec, -1)
    person = np.array([pheno[r][1] for r in take], dtype=object)
    # column of an individual = rank of its first appearance among the emitted records
    names, first_at, which = np.unique(person, return_index=True, return_inverse=True)
    column_of_name = np.empty(len(names), dtype=np.int64)
    column_of_name[np.argsort(first_at, kind='stable')] = np.arange(len(names))
    zmat = csr_matrix((np.ones(n_rec), (np.arange(n_rec), column_of_name[which])))
    return y, xmat, zmat
