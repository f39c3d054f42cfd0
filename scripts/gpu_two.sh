#!/bin/bash
# Two-GPU check (gpurun --gpus 2): the multi-GPU parity tests, then the 2-rank bench line.  Logs -> gpurun_out/.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q --timeout=300 > gpurun_out/pytest_multi.log 2>&1
echo "multi tests rc=$?"; tail -5 gpurun_out/pytest_multi.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err
echo "bench n2 rc=$?"; cut -c1-400 gpurun_out/bench_n2.json; tail -2 gpurun_out/bench_n2.err
