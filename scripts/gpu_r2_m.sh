#!/bin/bash
# last single-GPU pass of round 2 on the final tree: full GPU suite, smoke, bench line, reference arm, pair engines
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 2 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 1500 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --steps 10 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 300 python scripts/pairs_speed.py > gpurun_out/pairs_speed.json 2> gpurun_out/pairs_speed.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_n1.json') if l.startswith('{')][-1])
ref=json.loads([l for l in open('gpurun_out/bench_ref.json') if l.startswith('{')][-1])
r=d['roofline']
print('value %.4g e2e %.4g frac %.3f frac_sust %.3f share %.3f' % (d['value'], d['e2e']['value'], r['frac'], r['frac_of_sustained_peak'], r['kernel_share_of_step']))
print('reference %.4g ratio e2e %.0f same_config %s' % (ref['value'], d['e2e']['value']/ref['value'], ref['config']==d['config']))
x=d['extra']; b=x['northstar']['blocks']
print({k:(round(v['seconds'],3), v['verified']) for k,v in b.items()}, b['AA_approx']['phases_s'], b['AA_approx']['max_rel_err_vs_oracle'])
print('screen C3 %.3f ms; ns %.2f / %.2f ms; agmat %.2f ms frac %.2f' % (x['approx_screen']['seconds']*1e3, x['approx_screen_northstar']['clean']['seconds']*1e3, x['approx_screen_northstar']['missing_0.5pct']['seconds']*1e3, x['agmat']['seconds']*1e3, x['agmat']['roofline']['frac']))
print(open('gpurun_out/pairs_speed.json').read().splitlines()[-1][:500])
PY
