"""Drop-in for gmat/process_plink/read_plink_bed.py: class ``Bed`` around the C symbol
``read_plink_bed`` -- here the CUDA library's symbol of the same name and signature."""
import numpy as np

from .._ffi import check, ffi, lib
from ..engine import count_lines, current_device


def read_plink_bed(bed_file, num_id, num_snp):
    """Raw reader: float64 [num_snp][num_id] with the C map 0, 1/3 (missing), 1, 2
    (gmat/process_plink/_read_plink_bed.c:36-37)."""
    current_device()
    marker_mat = np.ones(num_id * num_snp, dtype=np.float64)
    check(lib.read_plink_bed(ffi.new("char[]", bed_file.encode('ascii')), num_id, num_snp,
                             ffi.cast("double *", ffi.from_buffer(marker_mat))), "read_plink_bed")
    return marker_mat.reshape(num_snp, num_id)


class Bed(object):
    def __init__(self, bed_file):
        self.bed_file = bed_file
        self.num_snp = count_lines(bed_file + '.bim')
        self.num_id = count_lines(bed_file + '.fam')

    def read(self):
        marker_mat = read_plink_bed(self.bed_file, self.num_id, self.num_snp)
        marker_mat[np.abs(marker_mat - 1.0 / 3) < 0.0001] = np.nan      # read_plink_bed.py:26
        return marker_mat.T
