"""TEST INFRASTRUCTURE: writes tests/golden/frontend_signatures.json -- name, parameter names and defaults of every public
function of the reference's gmat/remma/remma_epi{AA,AD,DD} packages (imported unmodified through oracle/ref_harness.py).
Run in the build container (needs /root/reference)."""
import inspect
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import ref_harness  # noqa: E402

ref_harness.install()
import importlib  # noqa: E402

out = {}
for kind in ("AA", "AD", "DD"):
    pkg = importlib.import_module("gmat.remma.remma_epi" + kind)
    table = {}
    for sub in ("", "_pair", "_eff", "_maf_eff", "_approx", "_maf_approx"):
        mod = importlib.import_module("gmat.remma.remma_epi%s.remma_epi%s%s" % (kind, kind, sub))
        for name, fn in inspect.getmembers(mod, inspect.isfunction):
            if fn.__module__ != mod.__name__ or "remma_epi" not in name:
                continue
            sig = inspect.signature(fn)
            table[name] = [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
                           for p in sig.parameters.values()]
    out[kind] = table
path = os.path.join(os.path.dirname(HERE), "tests", "golden", "frontend_signatures.json")
with open(path, "w") as f:
    json.dump(out, f, indent=1, sort_keys=True)
print(path, {k: len(v) for k, v in out.items()})
