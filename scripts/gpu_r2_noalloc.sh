#!/bin/bash
# the allocation-free exact scan: alone in a fresh process, then after every other GPU test, then smoke()
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/noalloc_alone.log 2>&1; echo "alone rc=$?" | tee -a gpurun_out/noalloc_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/gpu_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
