#!/bin/bash
# round 2, final single-GPU evidence: GPU suite, smoke, bench line, reference arm, launch list, ncu of the hot kernels
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log | cut -c1-300
timeout 2400 python -m pytest tests -m gpu -q --maxfail=40 --timeout=600 > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log | cut -c1-300
( time timeout 1500 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err ) 2> gpurun_out/bench_n1.time
echo "bench rc=$?"; tail -3 gpurun_out/bench_n1.time
( time timeout 900 python bench.py --impl reference --steps 10 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err ) 2>&1 | tail -3
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_n1.json') if l.startswith('{')][-1])
ref=json.loads([l for l in open('gpurun_out/bench_ref.json') if l.startswith('{')][-1])
r=d['roofline']
print('value %.4g e2e %.4g frac %.3f frac_sust %.3f share %.3f peak %.0f sust %.0f' % (d['value'], d['e2e']['value'], r['frac'], r['frac_of_sustained_peak'], r['kernel_share_of_step'], r['i8_mma_issue_tops'], r['i8_mma_issue_tops_sustained']))
print('reference %.4g ratio e2e %.0f same_config %s' % (ref['value'], d['e2e']['value']/ref['value'], ref['config']==d['config']))
x=d['extra']
print('northstar verified', x['northstar'].get('verified'), {k:(round(b['seconds'],2), b['verified']) for k,b in x['northstar']['blocks'].items()})
print('screen C3 %.3f ms frac %.2f; northstar clean %.2f ms frac %.2f; missing %.2f ms frac %.2f' % (x['approx_screen']['seconds']*1e3, x['approx_screen']['roofline']['frac'], x['approx_screen_northstar']['clean']['seconds']*1e3, x['approx_screen_northstar']['clean']['roofline']['frac'], x['approx_screen_northstar']['missing_0.5pct']['seconds']*1e3, x['approx_screen_northstar']['missing_0.5pct']['roofline']['frac']))
print('agmat %s %.2f ms frac %.2f (of the measured FP4 issue rate) i8 %.2f ms' % (x['agmat']['engine'], x['agmat']['seconds']*1e3, x['agmat']['roofline']['frac'], x['agmat']['i8_engine']['seconds']*1e3))
print('decode %.0f GB/s' % x['decode']['gb_per_s'])
b=x['northstar']['blocks']['AA_approx']
print('AA_approx %.3f s %.3g pairs/s verified %s recall %s phases %s' % (b['seconds'], b['pairs_per_s'], b['verified'], b.get('recall_of_exact_scan_rows'), b['phases_s']))
PY
BN="python bench.py --steps 2 --warmup 1 --no-cpu --no-extra --no-northstar"
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/launches.csv $BN > gpurun_out/bench_ncu.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/launches.csv
PK="python scripts/profile_kernels.py --reps 0 --only i8,eff"
timeout 1500 ncu --set full --clock-control none --import-source on \
    -k regex:'i8_scan_pair|i8_split|i8_au|i8_side|i8_screen_pair|screen_exact|sc_planes' -c 16 \
    -f -o gpurun_out/prof_kernels $PK > gpurun_out/prof_ncu.log 2>&1
echo "profile rc=$?"; tail -2 gpurun_out/prof_ncu.log
PN="python scripts/profile_kernels.py --reps 0 --only i8 --n 5000 --m 50000 --i8-first 43"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'i8_scan_pair' -s 1 -c 1 \
    -f -o gpurun_out/prof_scan_n5000 $PN > gpurun_out/prof_n5000_ncu.log 2>&1
echo "n5000 profile rc=$?"; tail -2 gpurun_out/prof_n5000_ncu.log
