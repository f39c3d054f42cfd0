#!/usr/bin/env python
"""cProfile of the end-to-end drop-in call bench.py times (second call, warm): where does the host time go?"""
import cProfile
import os
import pstats
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    import torch
    from gmat_b200 import engine
    from gmat_b200.gmatrix.gmatrix import grm_device
    from gmat_b200.remma.remma_epiAA import remma_epiAA
    d = tempfile.mkdtemp()
    prefix, pheno = bench.make_dataset(d)
    store = engine.GenotypeStore(prefix)
    ag = grm_device(store, engine.ADD)[0].cpu().numpy()
    gl = [ag, ag * ag]
    lst = bench.batch_first_snps(bench.N_SNP, 0, 1, bench.E2E_FIRST)
    out = os.path.join(d, "o")
    remma_epiAA(pheno, prefix, gl, bench.var_com_for(pheno), snp_lst_0=lst, p_cut=1e-5, out_file=out)
    torch.cuda.synchronize()
    pr = cProfile.Profile()
    t0 = time.perf_counter()
    pr.enable()
    remma_epiAA(pheno, prefix, gl, bench.var_com_for(pheno), snp_lst_0=lst, p_cut=1e-5, out_file=out)
    torch.cuda.synchronize()
    pr.disable()
    print("total %.3f s for %d pairs" % (time.perf_counter() - t0, bench.n_pairs(lst, bench.N_SNP)))
    pstats.Stats(pr).sort_stats("cumulative").print_stats(28)


if __name__ == "__main__":
    main()
