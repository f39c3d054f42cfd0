#!/bin/bash
# int8 pair-list engine: its tests, every drop-in test, the bench line (AA_approx uses it at n = 5000), speeds
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_i8_pairs.py tests/test_gpu_dropin.py tests/test_gpu_i8.py -m gpu -x -q > gpurun_out/pytest_l.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/pytest_l.log | cut -c1-300
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu --no-extra > gpurun_out/bench_l.json 2> gpurun_out/bench_l.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_l.json') if l.startswith('{')][-1])
print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['kernel_share_of_step'])
b=d['extra']['northstar']['blocks']
print({k:(round(v['seconds'],3), v['verified']) for k,v in b.items()})
print(b['AA_approx'])
PY
timeout 300 python scripts/pairs_speed.py > gpurun_out/pairs_speed.json 2> gpurun_out/pairs_speed.err; tail -n 1 gpurun_out/pairs_speed.json | cut -c1-600
