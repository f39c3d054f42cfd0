#!/bin/bash
# round 2, step A: the rewritten int8 exact engine (base-256 planes, side rows, guard) on the GPU
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_i8.py -m gpu -q -x --timeout=300 > gpurun_out/pytest_i8.log 2>&1
echo "i8 tests rc=$?"; tail -5 gpurun_out/pytest_i8.log | cut -c1-300
timeout 600 python scripts/i8_accuracy.py > gpurun_out/i8_accuracy.log 2>&1
echo "accuracy rc=$?"; tail -5 gpurun_out/i8_accuracy.log | cut -c1-400
GMAT_B200_I8_PLANES=6 timeout 600 python bench.py --steps 5 --warmup 3 --no-extra --no-cpu > gpurun_out/bench_S6.json 2> gpurun_out/bench_S6.err
echo "bench S=6 rc=$?"; cut -c1-300 gpurun_out/bench_S6.json; tail -3 gpurun_out/bench_S6.err
timeout 2400 python -m pytest tests -m gpu -q --maxfail=40 --timeout=600 > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -40 gpurun_out/pytest_gpu.log | cut -c1-300
