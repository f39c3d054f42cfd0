#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_screen_i8.py tests/test_gpu_dropin.py tests/test_gpu_kernels.py -m gpu -q -x --timeout=300 -k "screen or eff or approx or dropin" > gpurun_out/pytest_screen.log 2>&1
echo "screen tests rc=$?"; tail -5 gpurun_out/pytest_screen.log | cut -c1-300
python scripts/screen_overhead.py 2>&1 | grep -E "few_hits|epi_eff ms"
