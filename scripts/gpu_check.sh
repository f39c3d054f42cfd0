#!/bin/bash
# One gpurun call: peaks, smoke, GPU parity tests, short bench.  Logs -> gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 300 python - > gpurun_out/peaks.log 2>&1 <<'PY'
import torch, json
from gmat_b200 import engine
import bench
print(json.dumps(bench.measure_fp64_peaks(engine, torch)))
PY
echo "peaks rc=$?"; cat gpurun_out/peaks.log | tail -3
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "smoke rc=$?"; tail -5 gpurun_out/smoke.log
timeout 2400 python -m pytest tests -m gpu -q --maxfail=40 -x --timeout=600 > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -40 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err
echo "bench rc=$?"; tail -3 gpurun_out/bench.log; tail -5 gpurun_out/bench.err
