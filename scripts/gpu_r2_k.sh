#!/bin/bash
# minor-allele coding of the second SNPs: int8 engine tests, full-size tests, accuracy study, bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_i8.py tests/test_gpu_fullsize.py tests/test_gpu_dropin.py -m gpu -x -q > gpurun_out/pytest_k.log 2>&1; echo "tests rc=$?"; tail -n 4 gpurun_out/pytest_k.log | cut -c1-400
timeout 900 python scripts/i8_accuracy.py > gpurun_out/i8_accuracy.log 2>&1; echo "accuracy rc=$?"; grep -v "^{" gpurun_out/i8_accuracy.log | tail -n 8 | cut -c1-420
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-extra --no-northstar > gpurun_out/bench_k.json 2> gpurun_out/bench_k.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_k.json') if l.startswith('{')][-1])
print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['pairs_rechecked_in_fp64_per_step'])
PY
