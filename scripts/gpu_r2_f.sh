#!/bin/bash
# round 2, step F: the FP4 GRM kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x --timeout=300 -k "grm" > gpurun_out/pytest_grm.log 2>&1
echo "grm tests rc=$?"; tail -25 gpurun_out/pytest_grm.log | cut -c1-300
timeout 600 python scripts/profile_kernels.py --reps 2 --only grm --grm-n 10000 --grm-m 500000 > gpurun_out/prof_grm_plain.log 2>&1
echo "grm c2 rc=$?"; tail -2 gpurun_out/prof_grm_plain.log | cut -c1-600
