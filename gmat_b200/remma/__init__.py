from ._core import random_pair, random_pairAD
