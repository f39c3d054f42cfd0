#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_i8.py -m gpu -q -x --timeout=300 > gpurun_out/pytest_i8.log 2>&1
echo "i8 tests rc=$?"; tail -3 gpurun_out/pytest_i8.log | cut -c1-300
timeout 100 python scripts/peak_sustained.py
timeout 600 python bench.py --steps 5 --warmup 3 --no-extra --no-cpu > gpurun_out/bench_S6.json 2> gpurun_out/bench_S6.err
echo "bench S=6 rc=$?"; tail -3 gpurun_out/bench_S6.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_S6.json'))
r=d['roofline']
print(d['value'], d['ms_per_step'], {k:r[k] for k in ('achieved','peak','frac','kernel_ms_per_launch','kernel_launches_per_step','kernel_share_of_step','issued_tops','frac_issued')}, d['e2e']['value'], d['clocks'])
PY
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-extra --no-cpu > gpurun_out/bench_ncu.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/launches.csv
grep -E "i8_split|i8_au|i8_side" gpurun_out/launches.csv | tail -4 | cut -c1-200
