#!/bin/bash
# round 2, step G: full GPU suite, bench N=1, ncu of the FP4 GRM kernel at configs[1]
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q --maxfail=40 --timeout=600 > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log | cut -c1-400
( time timeout 1500 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err ) 2> gpurun_out/bench_n1.time
echo "bench rc=$?"; tail -3 gpurun_out/bench_n1.time; grep -v "INFO\|it/s\|^$" gpurun_out/bench_n1.err | tail -5 | cut -c1-300
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_n1.json') if l.startswith('{')][-1])
r=d['roofline']
print(d['value'], d['e2e']['value'], r['frac'], r['frac_of_sustained_peak'], r['kernel_share_of_step'], r['i8_mma_issue_tops'], r['i8_mma_issue_tops_sustained'])
print(d['extra']['northstar'].get('verified'), d.get('cpu_baseline',{}).get('value'))
for k in ('approx_screen','agmat'):
    print(k, json.dumps(d['extra'][k])[:1800])
PY
PG="python scripts/profile_kernels.py --reps 0 --only grm --grm-n 10000 --grm-m 500000"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'f4_gram_pair|f4_dosage_t|f4_offs_rows|f4_c_kernel|i8_gram_finish' -c 5 \
    -f -o gpurun_out/prof_grm_f4 $PG > gpurun_out/prof_grm_f4_ncu.log 2>&1
echo "grm f4 profile rc=$?"; tail -2 gpurun_out/prof_grm_f4_ncu.log
