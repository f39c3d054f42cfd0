from ._core import random_pair, random_pairAD
from .annotation import annotation_snp_pos
