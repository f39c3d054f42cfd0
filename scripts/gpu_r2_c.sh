#!/bin/bash
# round 2, step C: the full bench line at N=1 (timed), then the reference arm
mkdir -p gpurun_out
( time timeout 1500 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err ) 2> gpurun_out/bench_n1.time
echo "bench rc=$?"; tail -3 gpurun_out/bench_n1.time; tail -5 gpurun_out/bench_n1.err | cut -c1-300
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_n1.json'))
print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['kernel_share_of_step'], d['e2e']['phases_rank0_s'])
ns=d['extra']['northstar']
print(ns['seconds_dataset'], ns['seconds_grms'], ns.get('verified'))
for k,b in ns['blocks'].items():
    print(k, {x:b[x] for x in ('tests','seconds','pairs_per_s','verified','rows_in_file','max_rel_err_vs_oracle','seconds_verify','phases_max_over_ranks_s') if x in b}, b.get('messages'))
print(d.get('cpu_baseline'))
for k in ('approx_screen','agmat','decode'):
    print(k, json.dumps(d['extra'][k])[:600])
PY
( time timeout 900 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err ) 2>&1 | tail -3
cut -c1-300 gpurun_out/bench_ref.json
