"""Measurement scaffolding (NOT the product): cffi binding of probe/libgmat_b200_probe.so, the issue-rate
microbenchmarks bench.py and scripts/ use as roofline denominators.  Nothing under gmat_b200/ imports this."""
import os
import re

import cffi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgmat_b200_probe.so")

ffi = cffi.FFI()
_src = re.sub(r"/\*.*?\*/", "", open(os.path.join(_HERE, "gmat_b200_probe.h")).read(), flags=re.S)
ffi.cdef("\n".join(l for l in _src.splitlines()
                   if not l.strip().startswith("#") and not l.strip().startswith('extern "C"') and l.strip() != "}"))
lib = ffi.dlopen(LIB_PATH)


def _call(fn, *args):
    import torch
    out = ffi.new("double *")
    lib.gmat_b200_set_device(torch.cuda.current_device())
    rc = fn(*args, out)
    if rc < 1:
        raise RuntimeError(ffi.string(lib.gmat_b200_last_error()).decode("utf-8", "replace"))
    return float(out[0])


def i8_peak(iters=20000):
    """Measured issue rate of the int8 tensor-core MMA on this GPU, TOPS (2 ops per MAC)."""
    return _call(lib.gmat_b200_i8_peak, iters)


def f4_peak(iters=20000):
    """Measured issue rate of the FP4 (E2M1, block-scaled) tensor-core MMA on this GPU, TOPS (2 ops per MAC)."""
    return _call(lib.gmat_b200_f4_peak, iters)


def fp64_peak(kind=0, iters=20000):
    """kind 0: DMMA (mma.sync m8n8k4 f64), kind 1: DFMA; TFLOP/s."""
    return _call(lib.gmat_b200_fp64_peak, kind, iters)
