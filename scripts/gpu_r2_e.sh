#!/bin/bash
# round 2, step E: full GPU suite, bench N=1, ncu captures (scan + screen at C3, GRM at configs[1]), launch list
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q --maxfail=40 --timeout=600 > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -25 gpurun_out/pytest_gpu.log | cut -c1-400
( time timeout 1500 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err ) 2> gpurun_out/bench_n1.time
echo "bench rc=$?"; tail -3 gpurun_out/bench_n1.time; grep -v "INFO\|it/s\|^$" gpurun_out/bench_n1.err | tail -5 | cut -c1-300
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_n1.json') if l.startswith('{')][-1])
r=d['roofline']
print(d['value'], d['e2e']['value'], r['frac'], r['frac_of_sustained_peak'], r['kernel_share_of_step'], r['i8_mma_issue_tops'], r['i8_mma_issue_tops_sustained'])
print(d['extra']['northstar'].get('verified'))
for k in ('approx_screen','approx_screen_northstar','agmat','decode'):
    print(k, json.dumps(d['extra'][k])[:1500])
PY
PK="python scripts/profile_kernels.py --reps 0 --only i8,eff"
timeout 600 $PK > gpurun_out/prof_plain.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on \
    -k regex:'i8_scan_pair|i8_split|i8_au|i8_side|i8_screen_pair|screen_exact|sc_planes' -c 14 \
    -f -o gpurun_out/prof_kernels $PK > gpurun_out/prof_ncu.log 2>&1
echo "profile rc=$?"; tail -2 gpurun_out/prof_plain.log | cut -c1-600; tail -3 gpurun_out/prof_ncu.log
PG="python scripts/profile_kernels.py --reps 0 --only grm --grm-n 10000 --grm-m 500000"
timeout 600 $PG > gpurun_out/prof_grm_plain.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'i8_gram_pair|i8_dosage_t' -c 2 \
    -f -o gpurun_out/prof_grm $PG > gpurun_out/prof_grm_ncu.log 2>&1
echo "grm profile rc=$?"; tail -1 gpurun_out/prof_grm_plain.log | cut -c1-400
BN="python bench.py --steps 2 --warmup 1 --no-cpu --no-extra --no-northstar"
timeout 900 $BN > gpurun_out/bench_plain.log 2>&1 &&
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/launches.csv $BN > gpurun_out/bench_ncu.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/launches.csv
