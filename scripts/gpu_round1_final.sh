#!/bin/bash
# One gpurun call for the end of round 1: checks the default int8 kernel variant on the GPU first (3, the
# trimmed CTA-pair kernel, must pass the int8 parity tests, else everything below runs with variant 2), then bench, the full GPU
# parity suite, one `ncu --set full` capture of the int8 scan and GRM kernels and the ncu launch list of bench.py
# (each ncu pass only after the same command exited 0 without ncu).  Logs -> gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 180 python -m pytest tests/test_gpu_i8.py -m gpu -q -x --timeout=120 > gpurun_out/pytest_i8.log 2>&1
rc=$?; echo "i8 tests (default variant) rc=$rc"; tail -25 gpurun_out/pytest_i8.log | cut -c1-300
if [ $rc -ne 0 ]; then
    export GMAT_B200_I8_VARIANT=2
    echo "FALLBACK to variant 2" | tee gpurun_out/variant.txt
    timeout 420 python -m pytest tests/test_gpu_i8.py -m gpu -q --timeout=200 \
        --deselect tests/test_gpu_i8.py::test_i8_kernel_variants_are_bit_identical > gpurun_out/pytest_i8_v2.log 2>&1
    echo "i8 tests (variant 2) rc=$?"; tail -4 gpurun_out/pytest_i8_v2.log
else
    echo "variant 3" > gpurun_out/variant.txt
fi
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench rc=$?"; cut -c1-600 gpurun_out/bench.json; tail -3 gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
echo "bench ref rc=$?"; cut -c1-300 gpurun_out/bench_ref.json
timeout 900 python -m pytest tests -m gpu -q --maxfail=20 --timeout=300 --ignore=tests/test_gpu_i8.py \
    > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
PK="python scripts/profile_kernels.py --reps 0 --only i8,grm,eff"
timeout 300 $PK > gpurun_out/prof_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on \
    -k regex:'i8_scan_pair|i8_gram_pair|i8_dosage_t|i8_split|i8_screen_pair|sc_planes|screen_exact' -c 10 \
    -f -o gpurun_out/prof_kernels $PK > gpurun_out/prof_ncu.log 2>&1
echo "profile rc=$?"; tail -2 gpurun_out/prof_plain.log; tail -2 gpurun_out/prof_ncu.log
BN="python bench.py --steps 2 --warmup 1 --no-cpu --no-extra"
timeout 400 $BN > gpurun_out/bench_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/launches.csv $BN > gpurun_out/bench_ncu.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/launches.csv
if [ "$(cat gpurun_out/variant.txt)" = "variant 3" ]; then
    GMAT_B200_I8_VARIANT=4 timeout 300 python bench.py --no-extra --no-cpu > gpurun_out/bench_variant4.json 2> /dev/null
    echo "bench variant 4 rc=$?"; cut -c1-200 gpurun_out/bench_variant4.json
fi
