#!/usr/bin/env python
"""bench.py -- headline benchmark of the REMMAX hot path on B200 (contract: see DESIGN.md section 6).

    python bench.py --gpus N --steps K --warmup W                    # CUDA engine (this repo)
    python bench.py --impl reference --gpus N --steps K --warmup W   # the UNMODIFIED reference on the host cores

Headline workload (config.workload): BASELINE.json configs[2] -- remma_epiAA EXACT test, synthetic n=2,000 x m=20,000.
One step = one pass of the hot path over one batch of 1024 first-SNPs (all j > i: ~1.0e7 pairs); under torchrun
every rank takes its own batch (weak scaling), value = pairs of all ranks / max-over-ranks device time.

Also in the same JSON line:
  e2e              the public drop-in call remma_epiAA(files in, file out) on the same data
  extra (N=1)      approximate screen, agmat (configs[1]), decode, the FP64 DMMA engine
  extra.northstar  the north-star shape n=5,000 x m=50,000 through the drop-in API, every reported row re-derived by the
                   numpy oracle in the run: a FIXED sub-problem (every 8th first SNP of the exact AA scan) at every N
                   (strong scaling), remma_epiAD / remma_epiDD (configs[3]) and the 15,000-record repeated-measures
                   model (configs[4]) -- exhaustive at N=8 (or --northstar-full), on first-SNP subsets below that
  extra.multi_gpu_parity (N>=2)   sharded output file == single-GPU output file, byte for byte (data with missing calls)
"""
import argparse
import json
import os
import sys
import tempfile
import threading
import time

# torchrun exports OMP_NUM_THREADS=1 for every rank when it is unset; the host-side parts of the end-to-end call
# (numpy/BLAS in the design-matrix and file stages) then crawl.  Give each rank its share of the host cores.
if int(os.environ.get("LOCAL_WORLD_SIZE", "1")) > 1 and os.environ.get("OMP_NUM_THREADS", "1") == "1":
    # (the reference arm runs on rank 0 alone and may use every core)
    _share = 1 if "reference" in sys.argv else int(os.environ["LOCAL_WORLD_SIZE"])
    _thr = str(max(1, (os.cpu_count() or 1) // _share))
    for _k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[_k] = _thr

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_ID, N_SNP, SEED = 2000, 20000, 20201013 + 2
VAR_PROP = [0.35, 0.10, 0.55]           # [A, A*A, residual] proportions of the phenotypic variance; fixed
P_CUT = 1.0e-5
BATCH = 1024                            # first-SNPs per step
E2E_FIRST = 4096                        # first-SNPs per GPU of the end-to-end call (~4.1e7 pairs)
REF_FIRST = 4                           # first-SNPs per step of the reference arm (~5.0e4 pairs, ~1 s of numpy/BLAS)
NS_N, NS_M, NS_P_CUT = 5000, 50000, 1.0e-7     # the north-star shape (BASELINE configs[3]/[4])
METRIC = "snp_pairs_per_sec_remma_epiAA_exact"
# dram__bytes_read.sum + dram__bytes_write.sum of one i8_scan_pair_kernel launch (128 first SNPs of this shape), from the
# committed ncu --set full capture (profiles/r02_kernels.md: 2.515 GB read + 4.4 MB written; the compulsory bytes are
# the digit planes once -- 1.58 GB written by i8_split_kernel just before -- and the int8 / 2-bit codes, 51 MB)
I8_TRAFFIC_BYTES = 2.52e9
CONFIG = {"workload": "remma_epiAA exact test, synthetic n=%d x m=%d, GRMs [A, A*A] (BASELINE configs[2])"
                      % (N_ID, N_SNP),
          "step": "one batch of %d first-SNPs (evenly spread blocks of 8), all j>i (~1.0e7 pairs) per rank" % BATCH,
          "p_cut": P_CUT, "l2": "flushed (256 MB write) between timed iterations"}


def var_com_for(pheno, props=VAR_PROP):
    """Fixed variance components (REML is not on the timed path): proportions scaled to the phenotypic variance of the
    data set, so that the test statistics are calibrated and the output file holds the handful of rows a real scan
    yields (unscaled components made every chi-square anti-conservative: ~0.3 % of all pairs were written)."""
    yv = np.array([float(line.split()[-1]) for line in open(pheno)])
    return [float(v * np.var(yv)) for v in props]


def i8_ops_per_pair(n, planes, variant=3):
    """int8 tensor-core ops (2 per MAC) per pair: (algorithmic, issued).  Algorithmic = the lower triangle of the
    symmetric B_i once per digit plane, planes * n(n+1); issued = what the kernel's tiling launches: per plane and per
    256-row tile kt the id slabs 0..min(2kt+2, ceil(n/128))-1, each N x 128 MACs per pair with N = 256 plane rows,
    or (variant 3, ozaki.cu::i8p_shape) N = the tile's rows up to the four extra rows (n rounded up to 4, + 4),
    rounded up to 16, minus 128 for slab 2kt+1."""
    rows_all = (n + 3) // 4 * 4 + 4
    n_kt, n_slab = (rows_all + 255) // 256, (n + 127) // 128
    macs = 0.0
    for kt in range(n_kt):
        rows = 256 if variant < 3 else min(256, (rows_all - 256 * kt + 15) // 16 * 16)
        for l in range(min(2 * kt + 2, n_slab)):
            macs += (rows - 128 if variant >= 3 and l == 2 * kt + 1 else rows) * 128.0
    return planes * float(n) * (n + 1), planes * macs * 2.0


def f_exact(n):
    """Algorithmic flop per pair of the reference formulation: GEMM 2n^2 + 5n."""
    return 2.0 * n * n + 5.0 * n


def exact_issued_flop_per_pair(n, bm=128):
    """DMMA flop the exact kernel issues per pair (gmat_b200/csrc/exact.cu): per 128-row block and per warp row
    (64 rows) the slabs left of / inside the warp's diagonal sub-block, the upper-right quarter of the diagonal
    block skipped, slabs past n skipped, and in the last block only the m-tile pairs that hold a row below n."""
    nb = (n + bm - 1) // bm
    kslabs = (n + 15) // 16
    flop = 0.0
    for mb in range(nb):
        diag0, nks = mb * (bm // 16), (mb + 1) * (bm // 16)
        for wm in (0, 1):
            rows_left = n - (mb * bm + wm * 64)
            mtv = 8 if rows_left >= 64 else (0 if rows_left <= 0 else (rows_left + 15) // 16 * 2)
            mma_end = min(nks if wm else diag0 + bm // 32, kslabs)
            flop += 2.0 * (mtv * 8) * 16 * mma_end
    return flop


def batch_first_snps(m, rank, world, batch=BATCH):
    """``batch`` first-SNPs as contiguous blocks of 8 (one I-block of the scan tiling -- what a full scan
    hands the kernel), the blocks spread evenly over [0, m-1) and offset per rank so ranks take disjoint batches."""
    blk = 8 if batch >= 8 else 1
    nblk = max(1, batch // blk)
    stride = max(blk, (m - 1 - blk) // nblk)
    off = (rank * (stride // max(world, 1))) // blk * blk
    lst = []
    for q in range(nblk):
        s0 = (q * stride + off) % (m - 1 - blk)
        lst.extend(range(s0, s0 + blk))
    return sorted(set(lst))


def n_pairs(lst, m):
    return int(sum(m - 1 - i for i in lst))


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons with NVML during the timed region."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag = index, [], set(), False
        self.max_mhz = None

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
                     0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown"}
            while not self.stop_flag:
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
                time.sleep(0.1)
        except Exception as e:                      # never fail the benchmark because of NVML
            self.reasons.add("nvml_unavailable:%s" % type(e).__name__)

    def result(self):
        self.stop_flag = True
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def make_dataset(workdir, n=N_ID, m=N_SNP):
    import synth
    prefix = os.path.join(workdir, "c3_n%d_m%d" % (n, m))
    ds = synth.write_dataset(prefix, n, m, seed=SEED, n_planted=10, planted_kind="AA")
    return prefix, ds["pheno"]


# ----------------------------------------------------------------------------------------------------------------
# reference arm: the unmodified reference on the host cores (baseline/reference_arm.py)
# ----------------------------------------------------------------------------------------------------------------
REF_SAMPLE = ("one call of the unmodified reference's _remma_epiAA(y, xmat, zmat, gmat_lst, var_com, bed_file, "
              "snp_lst_0=%d evenly spread first-SNPs, p_cut, out_file) (gmat/remma/remma_epiAA/remma_epiAA.py:16-86, "
              "pip-installed into baseline/_ref) on the same synthetic files, all host threads (numpy/BLAS); timed: its "
              "scan loop by its own clock ('Running time: Clock time', remma_epiAA.py:83-85), i.e. without its preamble "
              "and .bed read" % REF_FIRST)


def reference_sample(prefix, pheno, arm=None):
    """(pairs, loop seconds, whole-call seconds, arm) of one bounded sample of the reference's exact scan."""
    from baseline import reference_arm
    if arm is None:
        arm = reference_arm.ExactArm(prefix, pheno, var_com_for(pheno), P_CUT)
    lst = batch_first_snps(N_SNP, 0, 1, REF_FIRST)
    pairs, loop_s, whole_s, rows = arm.sample(lst)
    return pairs, loop_s, whole_s, arm


def run_reference(args):
    """--impl reference: the UNMODIFIED reference (baseline/_ref) on the host cores, bounded sample per step."""
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    from baseline import reference_arm
    if not reference_arm.available():
        print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref is missing (run baseline/install_ref.sh "
                                                              "where /root/reference exists)"}))
        return 0
    with tempfile.TemporaryDirectory() as d:
        prefix, pheno = make_dataset(d)
        arm = None
        times, whole, pairs = [], [], 0
        for it in range(args.warmup + args.steps):
            pairs, loop_s, whole_s, arm = reference_sample(prefix, pheno, arm)
            if it >= args.warmup:
                times.append(loop_s)
                whole.append(whole_s)
    t = float(np.mean(times))
    value = pairs / t
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "pairs/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": CONFIG,
        "cpu_baseline": {"value": value, "unit": "pairs/s", "cores": os.cpu_count(), "kind": "reference",
                         "sample": REF_SAMPLE + "; %d pairs per step; whole call incl. preamble and file read: %.2f s"
                                   % (pairs, float(np.mean(whole)))},
        "e2e": {"value": value, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0


def measure_fp64_peaks(torch):
    """Roofline denominators for FP64 (absent from MEASURED_PEAKS.json): the DMMA issue-rate microbenchmark of
    probe/ and cuBLAS DGEMM 4096^3 (torch.matmul float64), best of a few runs, CUDA events."""
    import probe
    dmma = probe.fp64_peak(0, 20000)
    dfma = probe.fp64_peak(1, 20000)
    a = torch.randn((4096, 4096), dtype=torch.float64, device="cuda")
    b = torch.randn((4096, 4096), dtype=torch.float64, device="cuda")
    best = 1e30
    for it in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b)
        e1.record()
        torch.cuda.synchronize()
        if it:
            best = min(best, e0.elapsed_time(e1))
    dgemm = 2.0 * 4096 ** 3 / (best * 1e-3) / 1e12
    return {"dmma_tflops": dmma, "dfma_tflops": dfma, "cublas_dgemm_tflops": dgemm}


def i8_peak_sustained(seconds=1.5):
    """The int8 MMA issue-rate microbenchmark back to back for ``seconds``: the power-capped rate of the part
    (mean of the second half of the window)."""
    import probe
    t0, rates = time.time(), []
    while time.time() - t0 < seconds:
        rates.append(probe.i8_peak(200000))
    tail = rates[len(rates) // 2:]
    return float(sum(tail) / len(tail))


def run_gpu(args):
    import torch
    import torch.distributed as dist
    from scipy.stats import chi2
    import synth
    import probe
    from gmat_b200 import engine
    from gmat_b200._ffi import lib
    from gmat_b200.gmatrix.gmatrix import grm_device
    from gmat_b200.remma import _core
    from gmat_b200.remma.remma_epiAA import remma_epiAA
    from gmat_b200.uvlmm.design_matrix import design_matrix_wemai_multi_gmat

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        # NCCL prints its version banner to stdout when the communicator is created (box-level NCCL_DEBUG): keep
        # stdout to the one JSON line by pointing fd 1 at stderr while the communicator comes up
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    dev = engine.current_device()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(v, op):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=op)
        return float(t.item())

    MAX = dist.ReduceOp.MAX if world > 1 else None
    SUM = dist.ReduceOp.SUM if world > 1 else None

    # ONE int8 denominator for every tensor-roofline fraction of this run: the issue-rate microbenchmark on a cold,
    # unthrottled part, measured before anything else (the sustained, power-capped rate is measured after the timed
    # steps and reported beside it)
    i8_peak_burst = probe.i8_peak(20000) if rank == 0 else None
    f4_peak_burst = probe.f4_peak(20000) if rank == 0 else None     # same microbenchmark with the GRM's FP4 MMA

    workdir = tempfile.mkdtemp(prefix="gmat_bench_r%d_" % rank)
    prefix, pheno = make_dataset(workdir)            # every rank writes its own identical copy (seeded)
    # ------------- setup, outside the timed region: GRM on the GPU, preamble on the host, upload -------------
    store = engine.GenotypeStore(prefix)
    with _core.par.single_rank():                    # every rank needs the full GRM of its own copy here
        ag_dev, _ = grm_device(store, engine.ADD)
    ag = ag_dev.cpu().numpy()
    gmat_lst = [ag, ag * ag]
    y, xmat, zmat = design_matrix_wemai_multi_gmat(pheno, prefix)
    var_com = var_com_for(pheno)
    pymat, pvpmat = _core.preamble(y, xmat, zmat, gmat_lst, var_com, scipy_inv=True)
    model = _core.DeviceModel(pymat, pvpmat, dev)
    lut = store.lut(engine.ADD)
    chi_cut = float(chi2.isf(P_CUT, 1)) * (1 - 1e-9)
    lst = batch_first_snps(N_SNP, rank, world)
    pairs_rank = n_pairs(lst, N_SNP)
    il = np.asarray(lst, dtype=np.int64)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)       # > 126 MB L2
    # the engine the drop-in runs for this data (GMAT_B200_EXACT_ENGINE, default auto -> int8 tensor cores)
    which = engine.exact_engine(store)
    i8 = engine.I8Exact(store, lut, lut, engine.ADD, model.py, model.pvp, packed=model.packed) if which == "i8" else None

    def step():
        if i8 is not None:
            return i8.scan(il, engine.PAIRS_UPPER, chi_cut)
        return engine.epi_exact_scan(store, lut, lut, model.py, model.packed, il, engine.PAIRS_UPPER, chi_cut)

    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    launches0 = lib.gmat_b200_kernel_launches()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    if i8 is not None:
        i8.kernel_events = []
    hits = None
    rechecked = 0
    for k in range(args.steps):
        flush.zero_()                                # L2 flush between timed iterations
        ev[k][0].record()
        hits = step()
        ev[k][1].record()
        rechecked += i8.last_rechecked if i8 is not None else 0
    barrier()
    launches = lib.gmat_b200_kernel_launches() - launches0
    i8_peak_load = i8_peak_sustained() if rank == 0 else None          # right after the load, still power-capped
    clocks = sampler.result()
    ms_local = sum(a.elapsed_time(b) for a, b in ev) / args.steps
    kern_ms = kern_n = None
    if i8 is not None:
        kern_n = len(i8.kernel_events)
        kern_ms = sum(a.elapsed_time(b) for a, b in i8.kernel_events)     # all i8_scan launches of the timed steps
        i8.kernel_events = None
    print("[bench] rank %d: %.1f ms/step, per-step %s" % (rank, ms_local, [round(a.elapsed_time(b), 1) for a, b in ev]),
          file=sys.stderr)
    ms = reduce(ms_local, MAX)
    total_pairs = reduce(float(pairs_rank), SUM)
    value = total_pairs / (ms * 1e-3)

    # ------------- e2e: the public drop-in call, HOST buffers, files in / file out, copies inside -------------
    # every rank makes the SAME call (E2E_FIRST first-SNPs per GPU); the front-end shards the list over the ranks
    # and gathers the hits to rank 0 (gmat_b200/remma/parallel.py), which writes the file
    e2e_lst = batch_first_snps(N_SNP, 0, 1, min(E2E_FIRST * world, (N_SNP - 1) // 8 * 8))
    e2e_pairs = n_pairs(e2e_lst, N_SNP)
    out_file = os.path.join(workdir, "epiAA_e2e")
    e2e_t = []
    for it in range(3):
        barrier()
        t0 = time.perf_counter()
        remma_epiAA(pheno, prefix, gmat_lst, var_com, snp_lst_0=e2e_lst, p_cut=P_CUT, out_file=out_file)
        barrier()
        if it:
            e2e_t.append(time.perf_counter() - t0)
    e2e_phases = {k: round(v, 4) for k, v in _core.PHASES.items()}
    e2e_s = reduce(float(np.mean(e2e_t)), MAX)
    e2e_value = float(e2e_pairs) / e2e_s
    e2e_rows = sum(1 for _ in open(out_file)) - 1 if rank == 0 else 0
    # rank 0 uploads the GRMs, y and X (the preamble is computed once and broadcast); every rank its .bed payload
    h2d = world * store.h2d_bytes + sum(g.nbytes for g in gmat_lst) + y.nbytes + xmat.nbytes + len(e2e_lst) * 8
    d2h = 16 + 32 * e2e_rows

    extra = {}
    if world > 1 and not args.no_parity:
        extra["multi_gpu_parity"] = multi_gpu_parity(torch, dist, engine, synth, remma_epiAA, rank, world, workdir,
                                                     barrier)
    if not args.no_northstar:
        extra["northstar"] = northstar(args, torch, dist, engine, _core, synth, rank, world, barrier, reduce, MAX)

    line = None
    if rank == 0:
        n = N_ID
        mp = {}
        try:
            mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        i8_peaks = {"i8_mma_issue_tops": i8_peak_burst, "i8_mma_issue_tops_sustained": i8_peak_load,
                    "int8_from_measured_bf16_burst_x2": 2.0 * mp["bf16_tflops"] if mp else None,
                    "int8_from_measured_bf16_sustained_x2": 2.0 * mp["bf16_tflops_sustained"] if mp else None}
        line = {
            "metric": METRIC, "value": value, "unit": "pairs/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": CONFIG,
            "clocks": clocks, "gpu_launches": int(launches),
            "e2e": {"value": e2e_value, "unit": "pairs/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "seconds": e2e_s, "pairs": e2e_pairs, "rows_in_file": e2e_rows,
                    "phases_rank0_s": e2e_phases,
                    "what": "gmat_b200.remma.remma_epiAA(pheno_file, bed_file, gmat_lst, var_com, snp_lst_0=%d "
                            "first-SNPs): phenotype parse + GRM upload + GPU preamble + .bed read + H2D + scan + D2H "
                            "+ file" % len(e2e_lst)},
        }
        if i8 is not None:
            planes = i8.n_planes
            alg, iss = i8_ops_per_pair(n, planes, i8.wide)
            launches_per_step = kern_n / float(args.steps)
            kern_s = kern_ms * 1e-3 / args.steps                              # scan-kernel seconds per step
            achieved = pairs_rank * alg / kern_s / 1e12
            issued = pairs_rank * iss / kern_s / 1e12
            peak = i8_peak_burst
            line["dtype"] = ("f64 results; n^2 term as exact int8 x int8 -> int32 tensor-core products of an error-free "
                             "%d-plane base-256 split (%d bits), weak pairs re-checked in FP64, rest f64"
                             % (planes, 8 * planes - 1))
            line["roofline"] = {
                "bound": "tensor", "achieved": achieved, "unit": "TOP/s", "peak": peak, "frac": achieved / peak,
                "frac_of_sustained_peak": achieved / i8_peak_load,
                "traffic": I8_TRAFFIC_BYTES,
                "kernel": "i8_scan_pair_kernel (tcgen05.mma cta_group::2 kind::i8, 256xNx32 per CTA pair, N = the plane "
                          "rows that can be non-zero)" if i8.wide == 3 else
                          "i8_scan_pair_kernel (tcgen05.mma cta_group::2 kind::i8, 256x256x32 per CTA pair)",
                "kernel_ms_per_launch": kern_ms / max(kern_n, 1), "kernel_launches_per_step": launches_per_step,
                "kernel_share_of_step": kern_ms / args.steps / ms_local,
                "ops_per_pair_algorithmic": alg, "ops_per_pair_issued": iss, "issued_tops": issued,
                "frac_issued": issued / peak, "frac_issued_of_sustained_peak": issued / i8_peak_load,
                "pairs_rechecked_in_fp64_per_step": rechecked / float(args.steps),
                "fp64_equivalent_tflops": pairs_rank * f_exact(n) / (ms * 1e-3) / 1e12,
                "peak_source": "ONE denominator for every tensor-roofline fraction of this line: the issue rate of "
                               "back-to-back tcgen05.mma kind::i8 128x256x32 on resident tiles, all SMs "
                               "(probe/libgmat_b200_probe.so), measured on rank 0 at the start of the run on the cold "
                               "part; i8_mma_issue_tops_sustained = the same microbenchmark for 1.5 s right after the "
                               "timed steps (the power-capped rate, what *_of_sustained_peak divides by); "
                               "MEASURED_PEAKS.json has no int8 figure, 2 x its dense bf16 numbers are listed beside",
                **i8_peaks,
                "note": "achieved = planes x n(n+1) int8 ops per pair (lower triangle of the symmetric quadratic form, "
                        "once per digit plane) / the scan kernel's own launch time (the operand kernels of the next "
                        "batch run beside it on spare SM resources); issued adds the tiling's padding and the zero "
                        "half of the diagonal 256x256 blocks"}
        else:
            achieved = (pairs_rank * f_exact(n)) / (ms * 1e-3) / 1e12          # per GPU, algorithmic flops
            from gmat_b200._ffi import ffi
            cfg = ffi.new("int[3]")
            lib.gmat_b200_exact_config(cfg, cfg + 1, cfg + 2)
            bm, bn = int(cfg[0]), int(cfg[1])
            issued = pairs_rank * exact_issued_flop_per_pair(n, bm) / (ms * 1e-3) / 1e12
            peaks = measure_fp64_peaks(torch)
            peak = max(peaks["dmma_tflops"], peaks["cublas_dgemm_tflops"])
            line["roofline"] = {"bound": "tensor", "achieved": achieved, "unit": "TFLOP/s", "peak": peak,
                                "frac": achieved / peak, "frac_issued": issued / peak, "traffic": None,
                                "kernel": "epi_exact_rb_kernel<%d,..,scan>" % bn,
                                "flops_per_pair_algorithmic": f_exact(n), "issued_tflops": issued,
                                "note": "achieved = pairs/s x (2n^2+5n), the reference formulation's flop; the kernel "
                                        "exploits the symmetry of Z'PZ and issues ~51.5 % of them, so frac > 1 is "
                                        "expected: read frac_issued as pipe utilisation", **peaks}
        if world == 1 and not args.no_extra:
            e2e_f64 = None
            if i8 is not None:
                # the same drop-in call on the FP64 DMMA engine (one timed call after one warm call)
                os.environ["GMAT_B200_EXACT_ENGINE"] = "f64"
                try:
                    for it in range(2):
                        torch.cuda.synchronize()
                        t0 = time.perf_counter()
                        remma_epiAA(pheno, prefix, gmat_lst, var_com, snp_lst_0=e2e_lst, p_cut=P_CUT,
                                    out_file=out_file + ".f64")
                        torch.cuda.synchronize()
                        e2e_f64 = float(e2e_pairs) / (time.perf_counter() - t0)
                finally:
                    del os.environ["GMAT_B200_EXACT_ENGINE"]
            extra.update(extras(engine, torch, synth, store, model, lut, args, prefix, pymat, i8_peak_burst, e2e_f64,
                                f4_peak=f4_peak_burst))
        if extra:
            line["extra"] = extra
        if world == 1 and not args.no_cpu:
            cp, loop_s, whole_s, _ = reference_sample(prefix, pheno)
            line["cpu_baseline"] = {"value": cp / loop_s, "unit": "pairs/s", "cores": os.cpu_count(),
                                    "kind": "reference",
                                    "sample": REF_SAMPLE + "; %d pairs, loop %.2f s, whole call incl. preamble and file "
                                                           "read %.2f s" % (cp, loop_s, whole_s)}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


# ----------------------------------------------------------------------------------------------------------------
# multi-GPU parity inside the run: sharded file == single-GPU file
# ----------------------------------------------------------------------------------------------------------------
def multi_gpu_parity(torch, dist, engine, synth, remma_epiAA, rank, world, workdir, barrier):
    """The same remma_epiAA call (data WITH missing calls: the imputed genotypes must be the same on every rank)
    sharded over all ranks and on rank 0 alone; the two output files must be equal byte for byte."""
    from gmat_b200.gmatrix.gmatrix import grm_device
    from gmat_b200.process_plink.process_plink import load_store
    from gmat_b200.remma import parallel as par
    n, m = 600, 3000
    box = [None]
    if rank == 0:
        d = tempfile.mkdtemp(prefix="gmat_parity_")
        pre = os.path.join(d, "par")
        synth.write_dataset(pre, n, m, seed=SEED + 7, missing_rate=0.002, n_planted=6)
        box[0] = pre
    dist.broadcast_object_list(box, src=0)
    pre = box[0]
    barrier()
    with par.single_rank():
        np.random.seed(11)
        st = engine.GenotypeStore(pre)
        ag = grm_device(load_store(pre, warn=lambda *_: None), engine.ADD)[0].cpu().numpy()
    vc = var_com_for(pre + ".pheno")
    first = list(range(0, m - 1, 2))
    np.random.seed(12)                                       # rank 0's draws decide the imputation of both runs
    remma_epiAA(pre + ".pheno", pre, [ag, ag * ag], vc, snp_lst_0=first, p_cut=1e-3, out_file=pre + ".sharded")
    barrier()
    # once more with the speed-aware deal of long scans forced on this small one (calibration share, exchange of the
    # measured rates, re-deal of the rest: _core.exact_scan)
    from gmat_b200.remma import _core
    keep_min, _core.REBALANCE_MIN_WORK = _core.REBALANCE_MIN_WORK, 0.0
    np.random.seed(12)
    remma_epiAA(pre + ".pheno", pre, [ag, ag * ag], vc, snp_lst_0=first, p_cut=1e-3, out_file=pre + ".sharded2")
    _core.REBALANCE_MIN_WORK = keep_min
    rates = list(_core.LAST_RATES)
    barrier()
    res = None
    if rank == 0:
        np.random.seed(12)
        with par.single_rank():
            remma_epiAA(pre + ".pheno", pre, [ag, ag * ag], vc, snp_lst_0=first, p_cut=1e-3, out_file=pre + ".single")
        a, b = open(pre + ".sharded", "rb").read(), open(pre + ".single", "rb").read()
        a2 = open(pre + ".sharded2", "rb").read()
        res = {"identical": a == b and a2 == b, "identical_static_deal": a == b, "identical_speed_aware_deal": a2 == b,
               "speed_aware_deal_used": len(rates) == world,
               "rows": a.count(b"\n") - 1, "bytes": len(a), "missing_calls": int(st.n_missing()),
               "what": "remma_epiAA on synthetic n=%d x m=%d with %.1f %% missing calls, %d first-SNPs, p_cut=1e-3: "
                       "output file of the %d-rank sharded run (rank 0 imputes and broadcasts the 2-bit store, hits "
                       "gathered over NCCL), with the static deal and with the speed-aware deal forced, vs the same "
                       "call on rank 0 alone" % (n, m, 0.2, len(first), world)}
    barrier()
    return res


# ----------------------------------------------------------------------------------------------------------------
# north-star shape through the drop-in API, verified by the oracle inside the run
# ----------------------------------------------------------------------------------------------------------------
def northstar(args, torch, dist, engine, _core, synth, rank, world, barrier, reduce, MAX):
    import importlib
    from scipy.stats import chi2
    from gmat_b200.gmatrix.gmatrix import grm_device
    from gmat_b200.uvlmm.design_matrix import design_matrix_wemai_multi_gmat
    n, m = args.ns_n, args.ns_m
    full = args.northstar_full or world >= 8
    out = {"shape": "synthetic n=%d x m=%d, planted AA/AD/DD pairs, p_cut=%g" % (n, m, NS_P_CUT), "n_gpus": world}
    t0 = time.perf_counter()
    box = [None]
    if rank == 0:
        d = tempfile.mkdtemp(prefix="gmat_northstar_")
        pre = os.path.join(d, "ns")
        ds = synth.write_dataset(pre, n, m, seed=20201013 + 4, n_planted=12, planted_kind="mix", also_records=(3,))
        box[0] = (pre, ds["planted"])
        del ds
    if world > 1:
        dist.broadcast_object_list(box, src=0)
    pre, planted = box[0]
    barrier()
    out["seconds_dataset"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    store = engine.GenotypeStore(pre)
    ag = grm_device(store, engine.ADD)[0].cpu().numpy()            # sharded over the ranks + all_reduce when N > 1
    dg = grm_device(store, engine.DOM)[0].cpu().numpy()
    del store
    barrier()
    out["seconds_grms"] = time.perf_counter() - t0
    s2 = float(np.var([float(line.split()[-1]) for line in open(pre + ".pheno")]))
    s2r = float(np.var([float(line.split()[-1]) for line in open(pre + ".rep3.pheno")]))
    models = {
        "AA": (pre + ".pheno", [ag, ag * ag], [0.35 * s2, 0.10 * s2, 0.55 * s2]),
        "ADD": (pre + ".pheno", [ag, dg, ag * ag], [0.30 * s2, 0.05 * s2, 0.10 * s2, 0.55 * s2]),
        "REP": (pre + ".rep3.pheno", [ag, ag * ag, np.eye(n)], [0.28 * s2r, 0.08 * s2r, 0.15 * s2r, 0.44 * s2r]),
    }
    oracle_models = {}

    def run_block(name, kind, model_key, first, what):
        pheno, gmat_lst, var_com = models[model_key]
        mod = importlib.import_module("gmat_b200.remma.remma_epi" + kind)
        scan = getattr(mod, "remma_epi" + kind)
        out_file = pre + "." + name
        tests = (len(first) * m if kind == "AD" else n_pairs(first, m)) if first is not None else \
            (m * m if kind == "AD" else m * (m - 1) // 2)
        barrier()
        t0 = time.perf_counter()
        scan(pheno, pre, gmat_lst, var_com, snp_lst_0=first, p_cut=NS_P_CUT, out_file=out_file)
        barrier()
        secs = reduce(time.perf_counter() - t0, MAX)
        phases = dict(_core.PHASES)
        blk = None
        if world > 1:                                              # slowest rank per phase
            keys = sorted(set(phases) | {"gather", "write", "design_matrix", "preamble", "bed_read_h2d", "engine_init",
                                         "scan"})
            t = torch.tensor([phases.get(k, 0.0) for k in keys], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            phases = dict(zip(keys, t.tolist()))
        if rank == 0:
            blk = {"what": what, "kind": kind, "tests": int(tests), "seconds": secs, "pairs_per_s": tests / secs,
                   "first_snps": len(first) if first is not None else (m if kind == "AD" else m - 1),
                   "grms": len(gmat_lst), "records": sum(1 for _ in open(pheno)),
                   "phases_max_over_ranks_s": {k: round(v, 3) for k, v in phases.items()}}
            if _core.LAST_RATES:           # the speed-aware deal was used: relative scan rates of the ranks it measured
                blk["rank_rates_rel"] = list(_core.LAST_RATES)
            if not args.no_verify:
                blk.update(verify_file(out_file, pre, pheno, kind, model_key, gmat_lst, var_com, first, planted, n, m,
                                       oracle_models, design_matrix_wemai_multi_gmat, chi2))
        barrier()
        return blk

    def run_approx(exact_name, exact_scope):
        """The three-stage approximate pipeline (remma_epiAA_approx.py:10-53: 100,000 random pairs -> median null
        variance -> screen of ALL pairs -> exact re-test of the survivors) through the public call, files in, file
        out, on rank 0's GPU alone (the call has no sharded form: the reference's _approx_parallel parts draw their
        own random pairs and cuts)."""
        from gmat_b200.remma.remma_epiAA import remma_epiAA_approx
        pheno, gmat_lst, var_com = models["AA"]
        out_file = pre + ".AAapprox"
        blk = None
        barrier()
        if rank == 0:
            np.random.seed(20201013)
            with _core.par.single_rank():
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                remma_epiAA_approx(pheno, pre, gmat_lst, var_com, p_cut=NS_P_CUT, num_random_pair=100000,
                                   out_file=out_file)
                torch.cuda.synchronize()
                secs = time.perf_counter() - t0
            tests = m * (m - 1) // 2
            blk = {"what": "remma_epiAA_approx, ALL pairs screened (approximate test at p_cut on the median null "
                           "variance of 100,000 random pairs), survivors re-tested exactly; public call on files, one GPU",
                   "kind": "AA", "tests": int(tests), "seconds": secs, "pairs_per_s": tests / secs, "n_gpus_used": 1,
                   "screen_engine": engine.LAST_SCREEN_ENGINE,
                   "phases_s": {k: round(v, 3) for k, v in _core.PHASES.items()}}
            if not args.no_verify:
                if "AA" not in oracle_models:
                    y, xmat, zmat = design_matrix_wemai_multi_gmat(pheno, pre)
                    from oracle import oracle
                    oracle_models["AA"] = oracle.preamble(y, xmat, zmat, gmat_lst, var_com)
                exact_rows = None
                if os.path.exists(pre + "." + exact_name):
                    exact_rows = {(int(r.split()[0]), int(r.split()[1]))
                                  for r in open(pre + "." + exact_name).read().splitlines()[1:]}
                blk.update(verify_approx_file(out_file, pre, n, m, oracle_models["AA"][0], oracle_models["AA"][1],
                                              exact_rows, exact_scope, chi2))
        barrier()
        return blk

    sub8 = list(range(0, m - 1, 8))
    blocks = {}
    blocks["strong_AA"] = run_block(
        "strongAA", "AA", "AA", sub8,
        "FIXED problem at every N (strong scaling): remma_epiAA exact, every 8th first SNP, all j>i (1/8 of the "
        "exhaustive scan)")
    if full:
        blocks["AA"] = run_block("AA", "AA", "AA", None, "remma_epiAA exact, EXHAUSTIVE, GRMs [A, A*A]")
        blocks["AD"] = run_block("AD", "AD", "ADD", None, "remma_epiAD exact, EXHAUSTIVE (every ordered pair incl. "
                                                          "j=i), GRMs [A, D, A*A] (BASELINE configs[3])")
        blocks["DD"] = run_block("DD", "DD", "ADD", None, "remma_epiDD exact, EXHAUSTIVE, GRMs [A, D, A*A] "
                                                          "(BASELINE configs[3])")
        blocks["AA_15000_records"] = run_block(
            "REP", "AA", "REP", None, "remma_epiAA exact, EXHAUSTIVE, 3 records per individual (15,000 records), GRMs "
                                      "[A, A*A, I] (BASELINE configs[4])")
    else:
        blocks["AD"] = run_block("AD", "AD", "ADD", list(range(0, m, 16)),
                                 "remma_epiAD exact, every 16th first SNP x all m second SNPs, GRMs [A, D, A*A] "
                                 "(BASELINE configs[3]; exhaustive at N=8)")
        blocks["DD"] = run_block("DD", "DD", "ADD", sub8, "remma_epiDD exact, every 8th first SNP, GRMs [A, D, A*A] "
                                                          "(BASELINE configs[3]; exhaustive at N=8)")
        blocks["AA_15000_records"] = run_block(
            "REP", "AA", "REP", sub8, "remma_epiAA exact, every 8th first SNP, 3 records per individual (15,000 "
                                      "records), GRMs [A, A*A, I] (BASELINE configs[4]; exhaustive at N=8)")
    blocks["AA_approx"] = run_approx("AA" if full else "strongAA", None if full else set(sub8))
    out["exhaustive"] = bool(full)
    out["blocks"] = blocks
    if rank == 0:
        out["verified"] = all(b.get("verified", args.no_verify) for b in blocks.values())
    return out


def oracle_pair_stats(pre, n, m, kind, pymat, pvpmat, pairs, chi2):
    """eff, var, chi, p of explicit pairs by the numpy oracle expressions (remma_epiAA.py:59-61,72-77) on the files."""
    raw = np.fromfile(pre + ".bed", dtype=np.uint8)[3:].reshape(m, (n + 3) // 4)

    def cols(snps, coding):
        r = raw[np.asarray(snps)]
        c = np.stack([(r >> (2 * q)) & 3 for q in range(4)], axis=2).reshape(len(snps), -1)[:, :n]
        g = np.array([0.0, np.nan, 1.0, 2.0])[c]
        f = g.sum(axis=1, keepdims=True) / (2 * n)
        x = g - 2 * f if coding == "A" else np.where(g > 1.5, 0.0, g) - 2 * f * (1 - f)
        return x.T

    pairs = np.asarray(pairs, dtype=np.int64).reshape(-1, 2)
    effs, vars_ = [], []
    for k0 in range(0, len(pairs), 512):
        part = pairs[k0:k0 + 512]
        su, sv = sorted(set(part[:, 0].tolist())), sorted(set(part[:, 1].tolist()))
        U, V = cols(su, kind[0]), cols(sv, kind[1])
        iu, iv = {s_: k for k, s_ in enumerate(su)}, {s_: k for k, s_ in enumerate(sv)}
        W = U[:, [iu[a] for a in part[:, 0].tolist()]] * V[:, [iv[b] for b in part[:, 1].tolist()]]
        effs.append((W.T @ pymat).ravel())
        vars_.append(np.sum(W * (pvpmat @ W), axis=0))
    eff = np.concatenate(effs) if effs else np.zeros(0)
    var = np.concatenate(vars_) if vars_ else np.zeros(0)
    chi = eff * eff / var
    return eff, var, chi, chi2.sf(chi, 1)


def verify_approx_file(out_file, pre, n, m, pymat, pvpmat, exact_rows, exact_scope, chi2):
    """The final file of remma_epiAA_approx (snp_0 snp_1 eff var chi p_app p) against the numpy oracle: every row with
    p < 1e-5 and 2000 random other rows re-derived (eff, var, chi within 1e-9, p within 1e-7); p_app consistent with
    the %g-truncated effect (one null variance for the whole file); and against the EXACT scan of the same run: every
    approx row with p < p_cut whose first SNP the exact scan covered must be in the exact file (the FP64 pair kernel
    and the int8 scan kernel decide the same pairs), recall = the share of the exact scan's rows the screen kept."""
    t0 = time.perf_counter()
    lines = open(out_file).read().splitlines()
    ok, msgs = lines[0] == "snp_0 snp_1 eff var chi p_app p", []
    rows = [ln.split() for ln in lines[1:]]
    key = np.array([[int(r[0]), int(r[1])] for r in rows], dtype=np.int64).reshape(-1, 2)
    val = np.array([[float(v) for v in r[2:]] for r in rows], dtype=np.float64).reshape(-1, 5)
    rng = np.random.Generator(np.random.PCG64(2))
    strong = np.flatnonzero(val[:, 4] < 1e-5)
    rest = np.setdiff1d(np.arange(len(rows)), strong)
    pick = np.concatenate([strong, rng.choice(rest, size=min(2000, len(rest)), replace=False)]).astype(np.int64)
    eff, var, chi, pv = oracle_pair_stats(pre, n, m, "AA", pymat, pvpmat, key[pick], chi2)
    rel = np.max(np.abs(np.stack([val[pick, 0] / eff, val[pick, 1] / var, val[pick, 2] / chi]) - 1.0)) if len(pick) else 0.0
    relp = np.max(np.abs(val[pick, 4] / pv - 1.0)) if len(pick) else 0.0
    if rel > 1e-9 or relp > 1e-7:
        ok = False
        msgs.append("rows vs oracle: rel %.2e, p rel %.2e" % (rel, relp))
    # p_app = chi2.sf(float('%g' % eff)^2 / var_median): one denominator must explain every row
    e_tr = np.array([float("%g" % e) for e in val[:, 0]])
    den = e_tr * e_tr / chi2.isf(val[:, 3], 1)
    spread = float(np.max(np.abs(den / np.median(den) - 1.0))) if len(den) else 0.0
    if spread > 1e-6:
        ok = False
        msgs.append("p_app not consistent with one null variance: spread %.2e" % spread)
    sig = {(int(a), int(b)) for (a, b), q in zip(key.tolist(), val[:, 4].tolist()) if q < NS_P_CUT}
    res = {"rows_in_file": len(rows), "rows_exact_p_below_cut": len(sig), "rows_checked_by_oracle": int(len(pick)),
           "max_rel_err_vs_oracle": float(rel), "null_variance": float(np.median(den)) if len(den) else None}
    if exact_rows is not None:
        in_scope = {k for k in sig if exact_scope is None or k[0] in exact_scope}
        missing = sorted(in_scope - exact_rows)
        if missing:
            ok = False
            msgs.append("significant in the approx file but not in the exact scan: %s" % missing[:5])
        res["exact_scan_rows"] = len(exact_rows)
        res["recall_of_exact_scan_rows"] = len(in_scope & exact_rows) / len(exact_rows) if exact_rows else None
    res["verified"] = ok
    res["seconds_verify"] = time.perf_counter() - t0
    if msgs:
        res["messages"] = msgs[:10]
    return res


def verify_file(out_file, pre, pheno, kind, model_key, gmat_lst, var_com, first, planted, n, m, cache,
                design_matrix, chi2):
    """Re-derive the output file with the numpy oracle (host preamble on the same files): every reported row (eff, chi
    within 1e-9, p < p_cut), every planted pair in scope (reported iff its oracle p < p_cut), and 2000 random
    unreported pairs (oracle p >= p_cut)."""
    from oracle import oracle
    t0 = time.perf_counter()
    rows = [line.split() for line in open(out_file).read().splitlines()[1:]]
    got = {(int(r[0]), int(r[1])): (float(r[2]), float(r[3]), float(r[4])) for r in rows}
    if model_key not in cache:
        y, xmat, zmat = design_matrix(pheno, pre)
        cache[model_key] = oracle.preamble(y, xmat, zmat, gmat_lst, var_com)
    pymat, pvpmat = cache[model_key]
    raw = np.fromfile(pre + ".bed", dtype=np.uint8)[3:].reshape(m, (n + 3) // 4)

    def cols(snps, coding):
        r = raw[np.asarray(snps)]
        c = np.stack([(r >> (2 * q)) & 3 for q in range(4)], axis=2).reshape(len(snps), -1)[:, :n]
        g = np.array([0.0, np.nan, 1.0, 2.0])[c]
        f = g.sum(axis=1, keepdims=True) / (2 * n)
        x = g - 2 * f if coding == "A" else np.where(g > 1.5, 0.0, g) - 2 * f * (1 - f)
        return x.T

    def stats(pairs):
        pairs = np.asarray(pairs, dtype=np.int64).reshape(-1, 2)
        if len(pairs) == 0:
            return np.zeros(0), np.zeros(0), np.zeros(0)
        su, sv = sorted(set(pairs[:, 0].tolist())), sorted(set(pairs[:, 1].tolist()))
        U, V = cols(su, kind[0]), cols(sv, kind[1])
        iu, iv = {s: k for k, s in enumerate(su)}, {s: k for k, s in enumerate(sv)}
        W = U[:, [iu[a] for a in pairs[:, 0].tolist()]] * V[:, [iv[b] for b in pairs[:, 1].tolist()]]
        eff = (W.T @ pymat).ravel()
        var = np.sum(W * (pvpmat @ W), axis=0)
        chi = eff * eff / var
        return eff, chi, chi2.sf(chi, 1)

    ok, msgs, max_rel = True, [], 0.0
    keys = sorted(got)
    for k0 in range(0, len(keys), 512):
        part = keys[k0:k0 + 512]
        eff, chi, p = stats(part)
        for q, k in enumerate(part):
            ge, gc, gp = got[k]
            rel = max(abs(ge - eff[q]) / abs(eff[q]), abs(gc - chi[q]) / abs(chi[q]))
            max_rel = max(max_rel, rel)
            if rel > 1e-9 or not (p[q] < NS_P_CUT):
                ok = False
                msgs.append("row %s: rel %.2e, oracle p %.3e" % (k, rel, p[q]))
    in_first = set(first) if first is not None else None
    scope = []
    for a, b in planted:
        cands = [(a, b), (b, a)] if kind == "AD" else [(a, b)]
        scope += [c for c in cands if in_first is None or c[0] in in_first]
    _, _, pl_p = stats(scope)
    planted_sig = 0
    for pr, pv in zip(scope, pl_p):
        planted_sig += int(pv < NS_P_CUT)
        if (tuple(pr) in got) != bool(pv < NS_P_CUT):
            ok = False
            msgs.append("planted %s: oracle p %.3e, in file: %s" % (pr, pv, tuple(pr) in got))
    rng = np.random.Generator(np.random.PCG64(1))
    firsts = np.asarray(first if first is not None else np.arange(m if kind == "AD" else m - 1))
    rnd = []
    while len(rnd) < 2000:
        a, b = int(firsts[rng.integers(0, len(firsts))]), int(rng.integers(0, m))
        if kind != "AD" and b <= a:
            continue
        if (a, b) not in got:
            rnd.append((a, b))
    _, _, rp = stats(rnd)
    if not np.all(rp >= NS_P_CUT):
        ok = False
        msgs.append("unreported random pairs below p_cut: %d" % int(np.sum(rp < NS_P_CUT)))
    res = {"verified": ok, "rows_in_file": len(got), "max_rel_err_vs_oracle": max_rel,
           "planted_in_scope": len(scope), "planted_significant_and_reported": planted_sig,
           "random_unreported_checked": len(rnd), "seconds_verify": time.perf_counter() - t0}
    if msgs:
        res["messages"] = msgs[:10]
    return res


def extras(engine, torch, synth, store, model, lut, args, prefix, pymat, i8_peak, e2e_f64=None, f4_peak=None):
    """Secondary numbers of BASELINE.json's metric, same process, N=1: approximate screen pairs/s on the same
    data, agmat TFLOP/s on configs[1] (n=10,000 x m=500,000, generated in HBM), decode GB/s.  Every int8 fraction
    divides by the same ``i8_peak`` as the headline's."""
    out = {}

    def timed(fn, reps=3):
        best = 1e30
        for it in range(reps + 1):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            if it:
                best = min(best, e0.elapsed_time(e1))
        return best * 1e-3

    # the FP64 DMMA engine on a 256-first-SNP batch of the same data: the mechanism north_star names for the
    # quadratic forms (GMAT_B200_EXACT_ENGINE=f64), with its own roofline against the measured FP64 peak
    sub = np.asarray(batch_first_snps(N_SNP, 0, 1, 256), dtype=np.int64)
    t = timed(lambda: engine.epi_exact_scan(store, lut, lut, model.py, model.packed, sub, engine.PAIRS_UPPER, 19.5),
              reps=2)
    from gmat_b200._ffi import ffi
    cfg = ffi.new("int[3]")
    engine.lib.gmat_b200_exact_config(cfg, cfg + 1, cfg + 2)
    peaks = measure_fp64_peaks(torch)
    peak = max(peaks["dmma_tflops"], peaks["cublas_dgemm_tflops"])
    issued = n_pairs(sub, N_SNP) * exact_issued_flop_per_pair(N_ID, int(cfg[0])) / t / 1e12
    out["exact_f64_engine"] = {"pairs_per_s": n_pairs(sub, N_SNP) / t, "seconds": t,
                               "tflops_algorithmic": n_pairs(sub, N_SNP) * f_exact(N_ID) / t / 1e12,
                               "tflops_issued": issued, "fp64_peak_tflops": peak, "frac_issued": issued / peak,
                               "workload": "epi_exact_rb_kernel (FP64 DMMA, Z'PZ staged by TMA), 256 first-SNPs, all j>i",
                               **peaks}
    if e2e_f64 is not None:
        out["exact_f64_engine"]["e2e_pairs_per_s"] = e2e_f64
    # approximate screen: full scan of the C3 data (1.9999e8 pairs).  Default engine: int8 tensor-core prefilter
    # (2 planes x 2n int8 ops per pair) + FP64 re-test of the survivors (screen_i8.cu); the FP64 DMMA kernel
    # (F_approx = 2n flop per pair, eff.cu) beside it.  The cut is a real one (p_cut = 1e-5 on a variance of the size
    # random pairs show), so the survivors, the re-test and the hit records are in the time.
    la = store.lut(engine.ADD, engine.FREQ_CLOOP)
    full = np.arange(N_SNP - 1, dtype=np.int64)
    pairs = N_SNP * (N_SNP - 1) // 2
    sample = engine.epi_eff(store, la, la, model.py, np.arange(0, N_SNP - 1, 97, dtype=np.int64), 0, 0.0)
    from scipy.stats import chi2
    cut = float(np.sqrt(chi2.isf(P_CUT, 1) * np.median(sample["eff"] ** 2) / chi2.ppf(0.5, 1)))
    box = {}
    ops = engine.ScreenOperands(store, model.py)        # int8 codes, a.Z'Py, digits of Z'Py: built once, kept

    def run_screen(codings, prepared=True):
        box["hits"] = engine.epi_eff(store, la, la, model.py, full, 0, cut, codings=codings,
                                     ops=ops if (prepared and codings is not None) else None)
    t = timed(lambda: run_screen((engine.ADD, engine.ADD)))
    which = engine.LAST_SCREEN_ENGINE
    n_hits = len(box["hits"])
    out["approx_screen"] = {"engine": which, "pairs_per_s": pairs / t, "seconds": t, "pairs": pairs, "hits": n_hits,
                            "eff_cut": cut,
                            "fp64_equivalent_tflops": pairs * 2.0 * N_ID / t / 1e12,
                            "workload": "remma_epiAA_eff_cpu formulation, full scan n=%d x m=%d, whole call on "
                                        "prepared operands (first-SNP upload, planes of the first SNPs, prefilter, "
                                        "FP64 re-test, hit records to the host; the int8 codes of the store and the "
                                        "digits of Z'Py are kept across calls)" % (N_ID, N_SNP)}
    if which == "i8":
        tops = pairs * 4.0 * N_ID / t / 1e12
        out["approx_screen"]["roofline"] = {"bound": "tensor", "unit": "TOP/s", "achieved": tops, "peak": i8_peak,
                                            "frac": tops / i8_peak,
                                            "note": "4n int8 ops per pair (two digit planes of Z'Py) / the whole call"}
        t1 = timed(lambda: run_screen((engine.ADD, engine.ADD), prepared=False))
        out["approx_screen"]["one_shot_call_seconds"] = t1
    t64 = timed(lambda: run_screen(None))
    out["approx_screen"]["f64_engine"] = {"pairs_per_s": pairs / t64, "seconds": t64, "hits": len(box["hits"]),
                                          "same_hit_count": len(box["hits"]) == n_hits,
                                          "tflops_algorithmic": pairs * 2.0 * N_ID / t64 / 1e12,
                                          "kernel": "epi_eff_kernel (FP64 DMMA)"}
    del ops
    if not args.no_cpu:
        # the reference's own C screen (oracle/_ref, compiled from gmat/remma/_remma_epi_eff_cpu.c) on the same
        # files with all host threads; bounded sample of first-SNPs.  It re-reads and re-codes the whole .bed into
        # double[m*n] on every call, which is part of what its users pay (remma_epiAA_eff.py:78-83 times it too).
        from oracle import oracle
        sub = batch_first_snps(N_SNP, 0, 1, 2048)
        t0 = time.perf_counter()
        oracle.ref_screen_file(prefix, "AA", sub, pymat, 1e30, prefix + ".ref_screen.tmp")
        dt = time.perf_counter() - t0
        out["approx_screen"]["cpu_reference"] = {
            "pairs_per_s": n_pairs(sub, N_SNP) / dt, "seconds": dt, "cores": os.cpu_count(), "kind": "reference",
            "sample": "oracle/_ref remma_epiAA_eff_cpu (OpenMP, all threads) on 2048 first-SNPs in blocks of 8 "
                      "spread over the range = %d pairs, incl. its .bed read" % n_pairs(sub, N_SNP)}
    # decode: .bed -> float64 [m][n] (what read_plink_bed materialises), HBM-bound; C4 shape (2 GB out > L2)
    n3, m3 = 5000, 50000
    st3 = engine.GenotypeStore.from_device_payload(
        synth.device_payload(n3, m3, SEED + 3, store.device, engine.row_stride(n3)), n3)
    dec = torch.empty((m3, n3), dtype=torch.float64, device=store.device)
    t = timed(lambda: engine.decode_into(st3, None, dec))
    nbytes = m3 * ((n3 + 3) // 4) + 8.0 * m3 * n3
    mp = {}
    try:
        mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = mp.get("hbm_gbs")
    out["decode"] = {"gb_per_s": nbytes / t / 1e9, "seconds": t, "bytes": nbytes,
                     "roofline": {"bound": "hbm", "unit": "GB/s", "achieved": nbytes / t / 1e9, "peak": hbm,
                                  "frac": nbytes / t / 1e9 / hbm if hbm else None,
                                  "peak_source": "MEASURED_PEAKS.json hbm_gbs (copy bandwidth)"},
                     "workload": "gmat_b200_decode n=%d x m=%d (0.25 B in + 8 B out per genotype)" % (n3, m3)}
    del dec
    # approximate screen at the north-star shape (1.25e9 pairs), clean store and the same store with 0.5 % missing
    # calls (the C path's code 01 = 1/3: tripled integer coding, three base-32 planes)
    gen = torch.Generator(device=store.device)
    gen.manual_seed(5)
    py3 = torch.randn((n3,), dtype=torch.float64, device=store.device, generator=gen)
    py3 = (py3 - py3.mean()).contiguous()
    full3 = np.arange(m3 - 1, dtype=np.int64)
    pairs3 = m3 * (m3 - 1) // 2
    out["approx_screen_northstar"] = {}
    for tag, st in (("clean", st3), ("missing_0.5pct", None)):
        if st is None:
            bed = st3.bed.clone()
            hole = torch.rand(bed.shape, generator=gen, device=store.device) < 0.005 * 4
            slot = torch.randint(0, 4, bed.shape, generator=gen, device=store.device, dtype=torch.uint8)
            clear = (~(torch.tensor(3, dtype=torch.uint8, device=store.device) << (2 * slot))) & 0xFF
            setb = torch.tensor(1, dtype=torch.uint8, device=store.device) << (2 * slot)
            bed = torch.where(hole, (bed & clear) | setb, bed)             # code 01 in one slot of ~2 % of the bytes
            st = engine.GenotypeStore.from_device_payload(bed, n3)
        l3 = st.lut(engine.ADD, engine.FREQ_CLOOP)
        smp = engine.epi_eff(st, l3, l3, py3, np.arange(0, m3 - 1, 997, dtype=np.int64), 0, 0.0)
        cut3 = float(np.sqrt(chi2.isf(1e-7, 1) * np.median(smp["eff"] ** 2) / chi2.ppf(0.5, 1)))
        ops3 = engine.ScreenOperands(st, py3)
        hbox = {}

        def go():
            hbox["h"] = engine.epi_eff(st, l3, l3, py3, full3, 0, cut3, codings=(engine.ADD, engine.ADD), ops=ops3)
        t3 = timed(go, reps=2)
        planes3 = 3 if ops3.tripled else 2
        tops3 = pairs3 * 2.0 * planes3 * n3 / t3 / 1e12
        out["approx_screen_northstar"][tag] = {
            "engine": engine.LAST_SCREEN_ENGINE, "note": engine.LAST_SCREEN_NOTE, "missing_calls": int(st.n_missing()),
            "pairs": pairs3, "seconds": t3, "pairs_per_s": pairs3 / t3, "hits": len(hbox["h"]), "eff_cut": cut3,
            "digit_planes": planes3,
            "roofline": {"bound": "tensor", "unit": "TOP/s", "achieved": tops3, "peak": i8_peak, "frac": tops3 / i8_peak,
                         "note": "2 n int8 ops per pair and digit plane / the whole call"},
            "workload": "full AA screen n=%d x m=%d on prepared operands, cut at p = 1e-7 on the median null variance"
                        % (n3, m3)}
        del ops3
    del st3
    if not args.no_grm:
        n2, m2 = args.grm_n, args.grm_m
        rb = engine.row_stride(n2)
        payload = synth.device_payload(n2, m2, SEED + 1, store.device, rb)
        st2 = engine.GenotypeStore.from_device_payload(payload, n2)
        st2.lut(engine.ADD)
        kin = torch.empty((n2, n2), dtype=torch.float64, device=store.device)
        # the engine agmat() selects for this data (no missing calls -> the int8 tensor-core GEMM G G' with FP64
        # rank-one corrections, gram_i8.cu), then the FP64 DMMA kernel (grm.cu) north_star names, same output buffer
        which = engine.grm_engine(st2)
        t = timed(lambda: engine.grm(st2, engine.ADD, 1.0, 0.001, out=kin, which=which), reps=2)
        ops_n = float(n2) * (n2 + 1) * m2
        out["agmat"] = {"engine": which,
                        "tflops_full_gemm_convention_2n2m": 2.0 * n2 * n2 * m2 / t / 1e12,
                        "tflops_algorithmic_n(n+1)m": ops_n / t / 1e12,
                        "seconds": t, "workload": "agmat GRM, synthetic n=%d x m=%d generated in HBM "
                                                  "(BASELINE configs[1])" % (n2, m2)}
        if which in ("f4", "i8"):
            # the FP4 GRM's denominator is the issue rate of ITS MMA (tcgen05 kind::mxf4.block_scale 128x256x64, unit
            # scale factors), measured by the same microbenchmark as the int8 one at the start of the run on the cold
            # part (probe/: f4_peak); 2 x the int8 rate, the nominal ratio, is listed beside it
            pk = (f4_peak if f4_peak else 2.0 * i8_peak) if which == "f4" else i8_peak
            out["agmat"]["roofline"] = {
                "bound": "tensor", "unit": "TOP/s", "achieved": ops_n / t / 1e12, "peak": pk, "frac": ops_n / t / 1e12 / pk,
                "frac_of_int8_peak": ops_n / t / 1e12 / i8_peak,
                "f4_mma_issue_tops": f4_peak, "i8_mma_issue_tops_x2": 2.0 * i8_peak,
                "note": "n(n+1)m ops (lower triangle of G G', 2 ops per MAC) / the whole call (2-bit -> packed operand "
                        "transpose, the GEMM kernel, corrections); peak = "
                        + ("the measured issue rate of tcgen05.mma kind::mxf4.block_scale (E2M1 operands) on "
                           "resident tiles, all SMs, cold part" if which == "f4" else
                           "the int8 issue rate of the headline")}
            if which == "f4":
                t8 = timed(lambda: engine.grm(st2, engine.ADD, 1.0, 0.001, out=kin, which="i8"), reps=1)
                out["agmat"]["i8_engine"] = {"seconds": t8, "tflops_algorithmic_n(n+1)m": ops_n / t8 / 1e12,
                                             "frac_of_int8_peak": ops_n / t8 / 1e12 / i8_peak,
                                             "kernel": "i8_gram_pair_kernel (tcgen05 kind::i8)"}
            t64 = timed(lambda: engine.grm(st2, engine.ADD, 1.0, 0.001, out=kin, which="f64"), reps=1)
            out["agmat"]["f64_engine"] = {"seconds": t64,
                                          "tflops_algorithmic_n(n+1)m": ops_n / t64 / 1e12,
                                          "fp64_peak_tflops": peak,
                                          "frac": ops_n / t64 / 1e12 / peak,
                                          "kernel": "grm_kernel (FP64 DMMA, 2-bit codes staged by TMA)"}
        if args.agmat_e2e:
            # the public call on files: agmat(bed_file) = .bed/.bim/.fam read + H2D + counts + GRM kernel + D2H +
            # the 2.5 GB '%.18e' text file the reference format demands (written by all host threads)
            from gmat_b200.gmatrix import agmat
            d = tempfile.mkdtemp(prefix="gmat_agmat_e2e_")
            pre = os.path.join(d, "c2")
            with open(pre + ".bed", "wb") as f:
                f.write(bytes([0x6c, 0x1b, 0x01]))
                f.write(payload[:, :(n2 + 3) // 4].contiguous().cpu().numpy().tobytes())
            with open(pre + ".bim", "w") as f:
                f.write("".join("1 snp%d 0 %d A T\n" % (j, j + 1) for j in range(m2)))
            with open(pre + ".fam", "w") as f:
                f.write("".join("fam%d id%d 0 0 0 -9\n" % (k, k) for k in range(n2)))
            t0 = time.perf_counter()
            agmat(pre)
            dt = time.perf_counter() - t0
            out["agmat"]["e2e"] = {"seconds": dt, "tflops_algorithmic_n(n+1)m": float(n2) * (n2 + 1) * m2 / dt / 1e12,
                                   "text_bytes": os.path.getsize(pre + ".agrm0"),
                                   "what": "gmat_b200.gmatrix.agmat(bed_file): file read + H2D + GRM + D2H + .agrm0 text"}
            for ext in (".bed", ".bim", ".fam", ".agrm0"):
                os.remove(pre + ext)
        del kin, st2, payload
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="gmat_b200", choices=["gmat_b200", "reference"])
    ap.add_argument("--no-extra", action="store_true", help="skip the approx/GRM/decode secondary numbers")
    ap.add_argument("--no-grm", action="store_true")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-northstar", action="store_true", help="skip the n=5000 x m=50000 blocks")
    ap.add_argument("--northstar-full", action="store_true", help="exhaustive north-star scans at any N (default: N>=8)")
    ap.add_argument("--no-verify", action="store_true", help="skip the oracle re-derivation of the north-star files")
    ap.add_argument("--no-parity", action="store_true", help="skip the sharded-vs-single-GPU file comparison (N>=2)")
    ap.add_argument("--ns-n", type=int, default=NS_N)
    ap.add_argument("--ns-m", type=int, default=NS_M)
    ap.add_argument("--agmat-e2e", action="store_true", help="also time agmat(bed_file) on files (writes ~4 GB)")
    ap.add_argument("--grm-n", type=int, default=10000)
    ap.add_argument("--grm-m", type=int, default=500000)
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
