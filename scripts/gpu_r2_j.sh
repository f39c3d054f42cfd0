#!/bin/bash
# the approximate pipeline at the north-star shape inside the bench line; drop-in tests of the pipelines
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_dropin.py -m gpu -q > gpurun_out/pytest_dropin.log 2>&1; echo "dropin tests rc=$?"; tail -n 3 gpurun_out/pytest_dropin.log | cut -c1-300
( time timeout 1200 python bench.py --steps 5 --warmup 3 --no-cpu --no-extra > gpurun_out/bench_j.json 2> gpurun_out/bench_j.err ) 2>&1 | tail -n 3
echo "bench rc=$?"; grep -v "INFO\|it/s\|^$" gpurun_out/bench_j.err | tail -n 6 | cut -c1-300
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_j.json') if l.startswith('{')][-1])
ns=d['extra']['northstar']
print(ns.get('verified'))
for k,b in ns['blocks'].items():
    print(k, {x:b[x] for x in b if x not in ('what','phases_max_over_ranks_s')})
PY
