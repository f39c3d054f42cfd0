from .process_plink import read_plink, impute_geno, load_store
from .read_plink_bed import Bed, read_plink_bed
