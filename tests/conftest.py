import json
import os
import shutil
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(GOLD, "mouse_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session", autouse=True)
def _build_oracle():
    """Compile the oracle's C port (and oracle/_ref when the reference tree is present)."""
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "all"], check=True,
                   stdout=subprocess.DEVNULL)


@pytest.fixture()
def mouse(tmp_path):
    """Writable copy of the mouse data set; returns (bed_prefix, pheno_file, tmp dir)."""
    d = tmp_path / "mouse"
    shutil.copytree(os.path.join(GOLD, "mouse"), d)
    return str(d / "plink"), str(d / "pheno"), str(d)


@pytest.fixture(scope="session")
def mouse_ro():
    return os.path.join(GOLD, "mouse", "plink"), os.path.join(GOLD, "mouse", "pheno")


def read_rows(path, skip_header=True):
    with open(path) as f:
        lines = f.read().splitlines()
    return [l.split() for l in (lines[1:] if skip_header else lines)]


def gold_path(name):
    return os.path.join(GOLD, "out", name)


@pytest.fixture(scope="session")
def mouse_model(golden, mouse_ro):
    """Oracle-side model on mouse: GRMs, preambles for [ag, ag*ag] and [ag, dg, ag*ag] (cached per session)."""
    from oracle import oracle
    from gmat_b200.uvlmm.design_matrix import design_matrix_wemai_multi_gmat
    bed, pheno = mouse_ro
    ag, _ = oracle.agmat(bed)
    dg, _ = oracle.dgmat_as(bed)
    y, xmat, zmat = design_matrix_wemai_multi_gmat(pheno, bed)
    out = {"ag": ag, "dg": dg, "y": y, "xmat": xmat, "zmat": zmat,
           "gl_aa": [ag, ag * ag], "vc_aa": golden["var_com_ag_agag"],
           "gl_ad": [ag, dg, ag * ag], "vc_ad": golden["var_com_ag_dg_agag"]}
    return out
