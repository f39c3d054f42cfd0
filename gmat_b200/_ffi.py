"""cffi (ABI mode) binding of ``libgmat_b200.so`` -- the only door from Python into the CUDA engine.

The reference binds its C scan the same way (cffi; ``from _cremma_epi_eff_cpu import ffi, lib``,
gmat/remma/remma_epiAA/remma_epiAA_eff.py:17), so ``ffi`` / ``lib`` here are drop-in replacements
for that pair: ``lib.remma_epiAA_eff_cpu(...)`` has the identical signature.

There is NO CPU fallback: if the shared library is missing or does not load, importing this module
raises, and every product function fails loudly.
"""
import os
import re

import cffi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgmat_b200.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "gmat_b200.h")


def _cdef_from_header(path):
    """Strip preprocessor lines / extern "C" so cffi can parse the public header verbatim."""
    src = open(path).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    consts = dict(re.findall(r"^#define\s+(GMAT_B200_\w+)\s+(-?\d+)", src, flags=re.M))
    lines = []
    for line in src.splitlines():
        s = line.strip()
        if s.startswith("#") or s.startswith('extern "C"') or s == "}":
            continue
        lines.append(line)
    return "\n".join(lines), {k: int(v) for k, v in consts.items()}


ffi = cffi.FFI()
_CDEF, CONST = _cdef_from_header(HEADER_PATH)
ffi.cdef(_CDEF)

if not os.path.exists(LIB_PATH):
    raise ImportError(
        "gmat_b200: %s not found. Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(or `make -C gmat_b200/csrc`). There is no CPU fallback." % LIB_PATH)
lib = ffi.dlopen(LIB_PATH)


class GmatB200Error(RuntimeError):
    pass


def check(rc, what=""):
    """Raise on a negative status from any C-ABI call."""
    if rc is None or rc >= 1:
        return rc
    msg = ffi.string(lib.gmat_b200_last_error()).decode("utf-8", "replace")
    raise GmatB200Error("%s failed (status %d): %s" % (what or "gmat_b200 call", rc, msg))


def exported_symbols():
    """Names of every function the public header declares (used by the symbol-export test)."""
    return sorted(set(re.findall(r"\b(gmat_b200_\w+|remma_epi\w+_cpu|read_plink_bed)\s*\(", _CDEF)))
