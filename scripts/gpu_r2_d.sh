#!/bin/bash
# multi-GPU step: the 2-GPU parity tests, then the bench line under torchrun (N = $1)
N=${1:-2}
mkdir -p gpurun_out
if [ "$N" = "2" ]; then timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q --timeout=600 > gpurun_out/pytest_multi.log 2>&1; fi
echo "multi tests rc=$?"; tail -5 gpurun_out/pytest_multi.log | cut -c1-300
( time timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err ) 2> gpurun_out/bench_n$N.time
echo "bench n$N rc=$?"; tail -3 gpurun_out/bench_n$N.time; grep -v "INFO\|it/s\|^$" gpurun_out/bench_n$N.err | tail -8 | cut -c1-300
python - $N <<'PY'
import json, sys
N=sys.argv[1]
d=json.loads([l for l in open('gpurun_out/bench_n%s.json'%N) if l.startswith('{')][-1])
print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['kernel_share_of_step'], d['e2e']['phases_rank0_s'])
print(d['extra'].get('multi_gpu_parity'))
ns=d['extra']['northstar']
print(ns['seconds_dataset'], ns['seconds_grms'], ns.get('verified'))
for k,b in ns['blocks'].items():
    print(k, {x:b[x] for x in ('tests','seconds','pairs_per_s','verified','rows_in_file','max_rel_err_vs_oracle','seconds_verify','phases_max_over_ranks_s') if x in b}, b.get('messages'))
PY
