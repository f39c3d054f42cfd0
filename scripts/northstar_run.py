#!/usr/bin/env python
"""North-star run: the exhaustive EXACT remma_epiAA scan of a synthetic n x m data set through the public
drop-in API, sharded over the GPUs of one box (launch with torchrun), followed by a verification of the output
file against the numpy oracle.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 \
        scripts/northstar_run.py --ids 5000 --snps 50000 --out gpurun_out/northstar.json

What is checked (rank 0, after the scan):
  * every row of the output file is recomputed by the oracle (coded columns from the same .bed bytes, the host
    numpy preamble): eff, chi within 1e-9 relative, and p < p_cut holds;
  * every planted pair whose oracle p-value is below p_cut is in the file (and none above is);
  * a random sample of pairs that are NOT in the file has oracle p >= p_cut.
"""
import argparse
import json
import os
import sys
import time

if int(os.environ.get("LOCAL_WORLD_SIZE", "1")) > 1 and os.environ.get("OMP_NUM_THREADS", "1") == "1":
    _thr = str(max(1, (os.cpu_count() or 1) // int(os.environ["LOCAL_WORLD_SIZE"])))
    for _k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[_k] = _thr

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ids", dest="n", type=int, default=5000)
    ap.add_argument("--snps", dest="m", type=int, default=50000)
    ap.add_argument("--p-cut", type=float, default=1e-7)
    ap.add_argument("--workdir", default="/tmp/gmat_northstar")
    ap.add_argument("--out", default="gpurun_out/northstar.json")
    ap.add_argument("--kind", default="AA", choices=["AA", "DD", "AD"])
    ap.add_argument("--records", type=int, default=1,
                    help="records per individual (> 1: repeated measures, BASELINE configs[4]: GRM list gets an "
                         "identity for the permanent environment and Z is a real N x n incidence matrix)")
    ap.add_argument("--first", type=int, default=0,
                    help="scan only this many first SNPs (evenly spread, plus the first SNP of every planted pair) "
                         "through snp_lst_0 -- a single-GPU sample of the full shape; 0 = exhaustive")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from scipy.stats import chi2
    from gmat_b200 import engine
    import synth
    from gmat_b200.gmatrix.gmatrix import grm_device
    from gmat_b200 import remma
    import importlib
    mod = importlib.import_module("gmat_b200.remma.remma_epi" + args.kind)
    scan = getattr(mod, "remma_epi" + args.kind)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    n, m = args.n, args.m
    prefix = os.path.join(args.workdir, "ns_n%d_m%d_r%d_%s" % (n, m, args.records, args.kind))
    t0 = time.perf_counter()
    planted = None
    if rank == 0:
        os.makedirs(args.workdir, exist_ok=True)
        ds = synth.write_dataset(prefix, n, m, seed=20201013 + 4, n_planted=10, planted_kind=args.kind,
                                 records_per_id=args.records)
        planted = ds["planted"]
        del ds
    if world > 1:
        box = [planted]
        dist.broadcast_object_list(box, src=0)
        planted = box[0]
    barrier()
    t_data = time.perf_counter() - t0

    # GRMs on the GPU (every rank: they are inputs of its own preamble), variance components fixed
    t0 = time.perf_counter()
    store = engine.GenotypeStore(prefix)
    ag = grm_device(store, engine.ADD)[0].cpu().numpy()
    # variance components: fixed proportions scaled to the phenotypic variance (REML, wemai_multi_gmat, is out of
    # scope; unscaled components would make every chi-square anti-conservative and flood the output file)
    s2 = float(np.var([float(line.split()[-1]) for line in open(prefix + ".pheno")]))
    if args.kind == "AA":
        gmat_lst, var_com = [ag, ag * ag], [0.35 * s2, 0.10 * s2, 0.55 * s2]
    else:
        dg = grm_device(store, engine.DOM)[0].cpu().numpy()
        gmat_lst, var_com = [ag, dg, ag * ag], [0.30 * s2, 0.05 * s2, 0.10 * s2, 0.55 * s2]
    if args.records > 1:                                      # permanent-environment component (SURVEY config C5)
        gmat_lst = gmat_lst + [np.eye(n)]
        var_com = [v * 0.8 for v in var_com[:-1]] + [0.15 * s2, var_com[-1] * 0.8]
    snp_lst_0 = None
    if args.first > 0:
        last = m if args.kind == "AD" else m - 1
        pick = set(np.linspace(0, last - 1, args.first).astype(np.int64).tolist()) | {int(a) for a, _ in planted}
        snp_lst_0 = sorted(pick)
    del store
    barrier()
    t_grm = time.perf_counter() - t0

    out_file = os.path.join(args.workdir, "epi%s_full" % args.kind)
    barrier()
    t0 = time.perf_counter()
    scan(prefix + ".pheno", prefix, gmat_lst, var_com, snp_lst_0=snp_lst_0, p_cut=args.p_cut, out_file=out_file)
    barrier()
    t_scan = time.perf_counter() - t0
    if snp_lst_0 is None:
        n_tests = m * m if args.kind == "AD" else m * (m - 1) // 2
    else:
        n_tests = len(snp_lst_0) * m if args.kind == "AD" else int(sum(m - 1 - i for i in snp_lst_0))

    if rank == 0:
        # ------------------------------ verification against the oracle (host numpy) ------------------------
        from oracle import oracle
        from gmat_b200.uvlmm.design_matrix import design_matrix_wemai_multi_gmat
        t0 = time.perf_counter()
        rows = [line.split() for line in open(out_file).read().splitlines()[1:]]
        got = {(int(r[0]), int(r[1])): (float(r[2]), float(r[3]), float(r[4])) for r in rows}
        y, xmat, zmat = design_matrix_wemai_multi_gmat(prefix + ".pheno", prefix)
        pymat, pvpmat = oracle.preamble(y, xmat, zmat, gmat_lst, var_com)
        raw = np.fromfile(prefix + ".bed", dtype=np.uint8)[3:].reshape(m, (n + 3) // 4)

        def cols(snps, coding):
            r = raw[np.asarray(snps)]
            c = np.stack([(r >> (2 * q)) & 3 for q in range(4)], axis=2).reshape(len(snps), -1)[:, :n]
            g = np.array([0.0, np.nan, 1.0, 2.0])[c]
            f = g.sum(axis=1, keepdims=True) / (2 * n)
            x = g - 2 * f if coding == "A" else np.where(g > 1.5, 0.0, g) - 2 * f * (1 - f)
            return x.T

        def stats(pairs):
            pairs = np.asarray(pairs, dtype=np.int64).reshape(-1, 2)
            su, sv = sorted(set(pairs[:, 0].tolist())), sorted(set(pairs[:, 1].tolist()))
            U, V = cols(su, args.kind[0]), cols(sv, args.kind[1])
            iu, iv = {s: k for k, s in enumerate(su)}, {s: k for k, s in enumerate(sv)}
            W = U[:, [iu[a] for a in pairs[:, 0].tolist()]] * V[:, [iv[b] for b in pairs[:, 1].tolist()]]
            eff = (W.T @ pymat).ravel()
            var = np.sum(W * (pvpmat @ W), axis=0)
            chi = eff * eff / var
            return eff, chi, chi2.sf(chi, 1)

        ok = True
        msgs = []
        keys = sorted(got)
        max_rel = 0.0
        for k0 in range(0, len(keys), 512):
            part = keys[k0:k0 + 512]
            eff, chi, p = stats(part)
            for q, k in enumerate(part):
                ge, gc, gp = got[k]
                rel = max(abs(ge - eff[q]) / abs(eff[q]), abs(gc - chi[q]) / abs(chi[q]))
                max_rel = max(max_rel, rel)
                if rel > 1e-9 or not (p[q] < args.p_cut):
                    ok = False
                    msgs.append("row %s: rel %.2e, oracle p %.3e" % (k, rel, p[q]))
        pl_eff, pl_chi, pl_p = stats(planted)
        planted_sig = [tuple(pr) for pr, pv in zip(planted, pl_p) if pv < args.p_cut]
        for pr, pv in zip(planted, pl_p):
            if (tuple(pr) in got) != (pv < args.p_cut):
                ok = False
                msgs.append("planted %s: oracle p %.3e, in file: %s" % (pr, pv, tuple(pr) in got))
        rng = np.random.Generator(np.random.PCG64(1))
        rnd = []
        firsts = np.asarray(snp_lst_0 if snp_lst_0 is not None else np.arange(m))
        while len(rnd) < 2000:
            a, b = int(firsts[rng.integers(0, len(firsts))]), int(rng.integers(0, m))
            if args.kind != "AD" and b <= a:
                continue
            if (a, b) not in got:
                rnd.append((a, b))
        _, _, rp = stats(rnd)
        if not np.all(rp >= args.p_cut):
            ok = False
            msgs.append("unreported random pairs below p_cut: %d" % int(np.sum(rp < args.p_cut)))
        t_verify = time.perf_counter() - t0
        res = {
            "what": "remma_epi%s exact, %s, synthetic n=%d x m=%d, %d record(s) per individual, %d GRMs, p_cut=%g, "
                    "through the drop-in API (files in, file out), %d x B200"
                    % (args.kind, "exhaustive" if snp_lst_0 is None else "%d first SNPs (snp_lst_0)" % len(snp_lst_0),
                       n, m, args.records, len(gmat_lst), args.p_cut, world),
            "records": n * args.records,
            "n_gpus": world, "tests": n_tests, "seconds_scan_call": t_scan, "pairs_per_s_e2e": n_tests / t_scan,
            "seconds_dataset": t_data, "seconds_grm": t_grm, "seconds_verify": t_verify,
            "rows_in_file": len(got), "var_com": var_com, "host_cores": os.cpu_count(), "planted": planted, "planted_significant": planted_sig,
            "max_rel_err_vs_oracle": max_rel, "random_unreported_checked": len(rnd),
            "verified": ok, "messages": msgs[:20],
            "algorithmic_tflops_per_gpu": n_tests * (2.0 * n * n + 5 * n) / t_scan / world / 1e12,
        }
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as f:
            json.dump(res, f, indent=1)
        print(json.dumps(res))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
