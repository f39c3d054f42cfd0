"""Multi-GPU paths on a box with at least two GPUs (skipped otherwise): one process per GPU under torchrun,
rendezvous on 127.0.0.1."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(nproc, script, *args, port=29577):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc),
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "scripts", script)] + list(args)
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert res.returncode == 0, res.stderr[-3000:]
    return json.loads([line for line in res.stdout.splitlines() if line.startswith("{")][-1])


def test_sharded_grm_equals_single_gpu():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    out = _torchrun(2, "multi_gpu_grm_check.py", "700", "3000")
    assert out["world"] == 2 and out["scale_equal"]
    for k in ("add", "dom"):
        assert out[k + "_max_rel"] <= 1e-13 and out[k + "_symmetric"] and out[k + "_identical_on_all_ranks"]


def test_sharded_exact_scan_file_equals_oracle():
    """The public remma_epiAA call under torchrun (first-SNP blocks sharded over two GPUs, hits gathered with NCCL,
    rank 0 writes the file), every row re-derived by the numpy oracle (scripts/northstar_run.py)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    out_json = os.path.join(ROOT, "gpurun_out", "northstar_test.json")
    out = _torchrun(2, "northstar_run.py", "--ids", "500", "--snps", "2500", "--p-cut", "1e-4", "--out", out_json,
                    "--workdir", "/tmp/gmat_northstar_test", port=29578)
    assert out["verified"] and out["n_gpus"] == 2 and out["rows_in_file"] > 0
    assert out["max_rel_err_vs_oracle"] <= 1e-9


def test_one_imputation_for_all_ranks():
    """Missing calls under torchrun: rank 0 imputes and broadcasts the 2-bit store (the reference imputes once per
    call); every rank then holds the same genotypes and the same GRM although their numpy RNG states differ."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    out = _torchrun(2, "multi_gpu_impute_check.py", "300", "900", port=29579)
    assert out["world"] == 2 and out["missing_in_file"] > 0 and out["missing_after_impute"] == 0
    assert out["store_identical_on_all_ranks"] and out["grm_identical_on_all_ranks"] and out["scale_identical"]
