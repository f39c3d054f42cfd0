#!/bin/bash
# FP4 issue-rate probe, the _maf_approx side files, the bench line with the measured FP4 denominator
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_dropin.py -m gpu -q -k approx > gpurun_out/pytest_approx.log 2>&1; echo "approx tests rc=$?"; tail -n 3 gpurun_out/pytest_approx.log | cut -c1-300
timeout 300 python - <<'PY' 2>&1 | tail -n 5
import probe, torch
torch.cuda.init()
for k in range(3):
    print("i8 %.1f  f4 %.1f TOP/s" % (probe.i8_peak(20000), probe.f4_peak(20000)))
print("f4 long", [round(probe.f4_peak(200000), 1) for _ in range(6)])
print("i8 long", [round(probe.i8_peak(200000), 1) for _ in range(6)])
PY
timeout 900 python bench.py --steps 10 --warmup 3 --no-northstar --no-cpu > gpurun_out/bench_i.json 2> gpurun_out/bench_i.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_i.json') if l.startswith('{')][-1])
print(d['value'], d['e2e']['value'], d['roofline']['frac'])
print(json.dumps(d['extra']['agmat']['roofline']))
PY
