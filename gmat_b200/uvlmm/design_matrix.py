"""Host restatement of gmat/uvlmm/design_matrix.py:7-57 (phenotype file + .fam -> y, X, Z).

Caller-side glue of the hot path: pure text parsing, kept on the host.
Records of one individual are emitted contiguously in .fam order; Z is the N x n incidence matrix.
"""
import logging
import sys

import numpy as np
from scipy.sparse import csr_matrix

_NA = ('NA', 'NaN', 'nan', 'na')


def design_matrix_wemai_multi_gmat(pheno_file, bed_file):
    id_bed_lst = []
    with open(bed_file + '.fam') as fin:
        for line in fin:
            arr = line.split()
            id_bed_lst.append(" ".join([arr[0], arr[1]]))
    records = {}
    with open(pheno_file) as fin:
        for line in fin:
            arr = line.split()
            if arr[-1] not in _NA:
                records.setdefault(" ".join([arr[0], arr[1]]), []).append(arr)
    id_not_pheno = set(id_bed_lst) - set(records.keys())
    if len(id_not_pheno) > 0:
        logging.error('The below genotyped id is not in the phenotype file:\n {}'.format(
            '\n'.join(list(id_not_pheno))))
        sys.exit()
    y, xmat, id_lst = [], [], []
    for key in id_bed_lst:
        for arr in records[key]:
            y.append(float(arr[-1]))
            xmat.append(arr[2:-1])
            id_lst.append(arr[1])
    y = np.array(y).reshape(-1, 1)
    xmat = np.array(xmat, dtype=float).reshape(y.shape[0], -1)
    id_dct, row, col = {}, [], []
    for i, ind in enumerate(id_lst):
        row.append(i)
        if ind not in id_dct:
            id_dct[ind] = len(id_dct)
        col.append(id_dct[ind])
    zmat = csr_matrix(([1.0] * len(row), (row, col)))
    return y, xmat, zmat
