#!/usr/bin/env python
"""Turn the ncu artefacts a gpurun call brought back into the small, tracked summaries under profiles/.

    python scripts/ncu_summary.py --rep gpurun_out/prof_kernels.ncu-rep --out profiles/r01_kernels_v1.md
    python scripts/ncu_summary.py --launches gpurun_out/launches.csv --out profiles/r01_launches_v1.md

(ncu itself is only ever run on the GPU box, after the same command exited 0 without it; this script runs
here, on the CPU box, on the imported report.)
"""
import argparse
import collections
import csv
import io
import re
import subprocess

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs/thread"),
    ("launch__shared_mem_per_block_dynamic", "dyn smem/block"),
    ("sm__ops_path_tensor_src_fp64.sum", "FP64 tensor-path ops (flop)"),
    ("sm__ops_path_tensor_src_fp64.avg.pct_of_peak_sustained_elapsed", "FP64 tensor path % of peak (elapsed)"),
    ("sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active", "DMMA sub-pipe active %"),
    ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "FP64 (DFMA/DMUL) pipe active %"),
    ("sm__ops_path_tensor_op_utcimma_src_int8_sparsity_off.avg.pct_of_peak_sustained_elapsed", "int8 tcgen05 (UTCIMMA) ops % of peak (elapsed)"),
    ("sm__pipe_tensor_subpipe_imma_cycles_active.avg.pct_of_peak_sustained_active", "IMMA sub-pipe active %"),
    ("sm__pipe_shared_cycles_active.avg.pct_of_peak_sustained_elapsed", "shared pipe active % (elapsed)"),
    ("sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor-memory path active % (elapsed)"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 throughput % of peak"),
    ("l1tex__m_xbar2l1tex_read_bytes_mem_global_op_tma_ld.sum", "TMA load bytes (L2 -> SM)"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput % of peak"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("lts__t_bytes.sum", "L2 bytes"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem wavefronts"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank-conflict wavefronts"),
    ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "stall math_pipe_throttle"),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall wait"),
    ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stall barrier"),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall short_scoreboard"),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long_scoreboard"),
    ("smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "stall mio_throttle"),
    ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "stall lg_throttle"),
    ("smsp__average_warps_issue_stalled_membar_per_issue_active.ratio", "stall membar"),
    ("smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio", "stall sleeping"),
]


def summarise_report(rep, out, title):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    with open(out, "w") as f:
        f.write("# %s\n\nSource: `ncu --set full --clock-control none --import-source on` capture `%s` "
                "(read with `ncu -i ... --page raw --csv`). One launch per section; per-launch values.\n" % (title, rep))
        for r in rows[2:]:
            name = re.sub(r"\(.*", "", r[ki])
            f.write("\n## %s\n\n| metric | value |\n|---|---|\n" % name)
            for key, label in KEYS:
                if key in hdr:
                    i = hdr.index(key)
                    f.write("| %s (`%s`) | %s %s |\n" % (label, key, r[i], units[i]))
    print("wrote", out)


def summarise_launches(path, out, title):
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "msecond": 1.0, "s": 1e3, "usecond": 1e-3, "nsecond": 1e-6}
    agg = collections.OrderedDict()
    for r in rows[1:]:
        v = float(r[vi].replace(",", "")) * scale[r[ui]]
        name = re.sub(r"\(.*", "", r[ki])
        name = name if len(name) < 90 else name[:87] + "..."
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    ours = sum(a[1] for k, a in agg.items() if "gb::" in k)
    with open(out, "w") as f:
        f.write("# %s\n\nSource: `ncu --metrics gpu__time_duration.sum --clock-control none --csv` over `%s` "
                "(cold-cache, serialised per-launch times: compare SHARES, not absolutes).\n\n" % (title, path))
        f.write("Launches: %d; total device time %.1f ms; kernels of this library (`gb::*`): %.1f ms (%.1f %%).\n\n"
                % (len(rows) - 1, tot, ours, 100 * ours / tot))
        f.write("| kernel | launches | total ms | share |\n|---|---:|---:|---:|\n")
        for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
            if a[1] / tot < 2e-4 and "gb::" not in k:
                continue
            f.write("| `%s` | %d | %.3f | %.2f %% |\n" % (k, a[0], a[1], 100 * a[1] / tot))
    print("wrote", out)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--rep")
    ap.add_argument("--launches")
    ap.add_argument("--out", required=True)
    ap.add_argument("--title", default="ncu summary")
    a = ap.parse_args()
    if a.rep:
        summarise_report(a.rep, a.out, a.title)
    if a.launches:
        summarise_launches(a.launches, a.out, a.title)
