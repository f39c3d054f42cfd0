#!/bin/bash
# BASELINE configs[3]/[4] at their full shape (n=5000 x m=50000) on ONE GPU, sampled through snp_lst_0: repeated
# measures (15,000 records, GRMs [A, A*A, I]) for AA, and DD / AD with GRMs [A, D, A*A]; every reported row, the
# planted pairs and 2000 unreported pairs re-derived by the numpy oracle.  Logs -> gpurun_out/.
mkdir -p gpurun_out
run() { timeout $1 python scripts/northstar_run.py --workdir /tmp/ns --p-cut 1e-6 "${@:3}" --out gpurun_out/$2.json \
        > gpurun_out/$2.log 2>&1; echo "$2 rc=$?"; tail -1 gpurun_out/$2.log | cut -c1-700; }
run 230 c5_rep3_AA --kind AA --records 3 --first 256
run 150 c4_DD --kind DD --first 192
run 150 c4_AD --kind AD --first 96
