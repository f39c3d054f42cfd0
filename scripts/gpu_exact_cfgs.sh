#!/bin/bash
# compare the exact-kernel tile configurations (parity first, then speed)
mkdir -p gpurun_out
for c in ${CFGS:-0 1}; do
  export GMAT_B200_EXACT_CFG=$c
  timeout 600 python -m pytest tests -m gpu -q -k "exact or pair_list or smoke" --timeout=300 > gpurun_out/pytest_cfg$c.log 2>&1
  echo "cfg $c pytest rc=$?"; tail -2 gpurun_out/pytest_cfg$c.log
  timeout 300 python scripts/profile_kernels.py --only exact --reps 2 --first 256 > gpurun_out/exact_cfg$c.log 2>&1
  echo "cfg $c: $(tail -1 gpurun_out/exact_cfg$c.log)"
done
