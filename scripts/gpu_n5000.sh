#!/bin/bash
# The int8 exact scan at the north-star shape (n=5000 x m=50000) on one GPU: rate of a 64-first-SNP scan, then one
# ncu --set full capture of its kernel (after the plain run exited 0).  Logs -> gpurun_out/.
mkdir -p gpurun_out
PK="python scripts/profile_kernels.py --only i8 --n 5000 --m 50000 --first 64"
timeout 300 $PK --reps 0 > gpurun_out/n5000_plain.log 2>&1
echo "plain rc=$?"; tail -1 gpurun_out/n5000_plain.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'i8_scan_pair' -s 1 -c 1 \
    -f -o gpurun_out/prof_n5000 $PK --reps 0 > gpurun_out/n5000_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/n5000_ncu.log
