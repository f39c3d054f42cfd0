#!/bin/bash
# One gpurun call: GPU parity tests, then (each only after its plain run exited 0) the ncu launch list of
# bench.py and one `ncu --set full` capture of every hot kernel.  Logs -> gpurun_out/.
mkdir -p gpurun_out
if [ "${SKIP_TESTS:-0}" != "1" ]; then
timeout 1500 python -m pytest tests -m gpu -q --maxfail=40 --timeout=300 > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
fi
PK="python scripts/profile_kernels.py --reps 0 ${PROFILE_ARGS:-}"
timeout 600 $PK > gpurun_out/prof_plain.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on \
    -k regex:'i8_scan|i8_split|epi_exact|epi_eff_kernel|grm_kernel|decode_kernel' -c 14 \
    -f -o gpurun_out/prof_kernels $PK > gpurun_out/prof_ncu.log 2>&1
echo "profile rc=$?"; tail -2 gpurun_out/prof_plain.log; tail -3 gpurun_out/prof_ncu.log
BN="python bench.py --steps 2 --warmup 1 --no-cpu ${BENCH_ARGS:-}"
timeout 900 $BN > gpurun_out/bench_plain.log 2>&1 &&
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/launches.csv $BN > gpurun_out/bench_ncu.log 2>&1
echo "launch list rc=$?"; tail -1 gpurun_out/bench_plain.log | cut -c1-400; wc -l gpurun_out/launches.csv
