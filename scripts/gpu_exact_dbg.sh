#!/bin/bash
mkdir -p gpurun_out
export GMAT_B200_EXACT_CFG=${CFG:-2}
for d in ${DBGS:-0 1 2 4 6 7 16 23}; do
  export GMAT_B200_EXACT_DBG=$d
  echo "dbg $d: $(timeout 300 python scripts/profile_kernels.py --only exact --reps 1 --first 128 2>&1 | tail -1)"
done
