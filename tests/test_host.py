"""CPU tests of the host side: the C-ABI library loads and exports every symbol include/gmat_b200.h declares
(no compute calls without a GPU), host logic (partitions, design matrix, formats), and the world_size-2
gloo path of the hit gather."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from conftest import GOLD, ROOT, read_rows


def test_library_loads_and_exports_every_declared_symbol():
    from gmat_b200 import _ffi
    names = _ffi.exported_symbols()
    assert len(names) >= 24
    for required in ("read_plink_bed", "remma_epiAA_eff_cpu", "remma_epiAD_eff_cpu", "remma_epiDD_eff_cpu",
                     "remma_epiAA_maf_eff_cpu", "remma_epiAD_maf_eff_cpu", "remma_epiDD_maf_eff_cpu",
                     "gmat_b200_epi_exact_scan", "gmat_b200_epi_exact_pairs", "gmat_b200_epi_eff", "gmat_b200_grm",
                     "gmat_b200_decode", "gmat_b200_bed_stats", "gmat_b200_build_lut", "gmat_b200_pack_pvp"):
        assert required in names
    for n in names:
        assert hasattr(_ffi.lib, n), n
    out = subprocess.run(["nm", "-D", "--defined-only", _ffi.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (\w+)", out))
    assert set(names) <= exported
    assert _ffi.lib.gmat_b200_version() == 100
    assert _ffi.ffi.sizeof("gmat_b200_hit") == 32


def test_reference_signatures_match_the_reference_cdef():
    """The seven drop-in symbols have exactly the reference's cffi cdef signatures (gmat/remma/_build.py:8-31)."""
    from gmat_b200 import _ffi
    from oracle.ref_harness import BED_CDEF, EPI_CDEF

    def norm(sig):
        return re.sub(r"\s+", " ", sig.replace("char* ", "char *")).strip()
    theirs = {re.search(r"(\w+)\s*\(", s).group(1): norm(s) for s in (EPI_CDEF + BED_CDEF).split(";") if "(" in s}
    ours = {}
    for s in _ffi._CDEF.split(";"):
        mm = re.search(r"\b(remma_epi\w+_cpu|read_plink_bed)\s*\(", s)
        if mm:
            ours[mm.group(1)] = norm(s[s.index("int "):])
    assert set(ours) == set(theirs) and len(ours) == 7
    for k in theirs:
        a = re.sub(r"\b\w+(?=[,)])", "", theirs[k]).replace(" ", "")     # drop parameter names
        b = re.sub(r"\b\w+(?=[,)])", "", ours[k]).replace(" ", "")
        assert a == b, (k, a, b)


def test_host_bed_reader_without_gpu(tmp_path):
    from gmat_b200._ffi import ffi, lib
    bed = os.path.join(GOLD, "mouse", "plink")
    n, m = 1304, 1407
    rb = 352
    buf = np.zeros(m * rb, dtype=np.uint8)
    assert lib.gmat_b200_bed_read_host(bed.encode(), n, m, rb, ffi.cast("unsigned char *", buf.ctypes.data)) == 1
    raw = np.fromfile(bed + ".bed", dtype=np.uint8)[3:].reshape(m, 326)
    assert np.array_equal(buf.reshape(m, rb)[:, :326], raw) and not buf.reshape(m, rb)[:, 326:].any()
    assert lib.gmat_b200_bed_read_host(str(tmp_path / "missing").encode(), n, m, rb, ffi.NULL) == -2
    assert b"missing.bed" in ffi.string(lib.gmat_b200_last_error())
    (tmp_path / "bad.bed").write_bytes(b"abc")
    assert lib.gmat_b200_bed_read_host(str(tmp_path / "bad").encode(), 4, 1, 16, ffi.NULL) == -2
    assert lib.gmat_b200_pvp_packed_elems(1304) == 8 * 11 * 12 // 2 * 2048


def test_no_cuda_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from gmat_b200 import engine
    with pytest.raises(RuntimeError):
        engine.current_device()
    from gmat_b200.gmatrix import agmat
    with pytest.raises(RuntimeError):
        agmat(os.path.join(GOLD, "mouse", "plink"))


def test_product_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "gmat_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f
                assert "oracle/" not in src and "/root/reference" not in src, f


def test_parallel_partition_matches_reference(golden):
    from gmat_b200.remma.parallel import parallel_lists, shard_first_snps
    for part in golden["parallel_partitions"]:
        lst, pos = parallel_lists(part["m"], [part["P"], part["k"]], False)
        assert pos == part["AA"] and lst == list(range(pos[0], pos[1])) + list(range(pos[2], pos[3]))
        assert parallel_lists(part["m"], [part["P"], part["k"]], True)[1] == part["AD"]
    # the reference partition covers [0, m-1) exactly once
    for m, P in ((1407, 3), (50000, 8), (101, 5)):
        allv = sorted(sum((parallel_lists(m, [P, k])[0] for k in range(1, P + 1)), []))
        assert allv == list(range(m - 1))
    # GPU sharding: disjoint cover, triangular work balanced
    m, R = 50000, 8
    full = list(range(m - 1))
    shards = [shard_first_snps(full, m, False, r, R) for r in range(R)]
    assert sorted(sum(shards, [])) == full
    work = [sum(m - 1 - i for i in s) for s in shards]
    assert max(work) / min(work) < 1.002
    # any list (a strided subset, descending order, the AD scan's m pairs per first SNP): still a disjoint cover,
    # pair counts within one dealt block of each other
    for lst, allp in ((list(range(0, m - 1, 8)), False), (list(range(m - 2, -1, -3)), False), (list(range(0, m, 16)), True)):
        sh = [shard_first_snps(lst, m, allp, r, R) for r in range(R)]
        assert sorted(sum(sh, [])) == sorted(lst)
        wk = [sum((m if allp else m - 1 - i) for i in s) for s in sh]
        assert max(wk) - min(wk) <= 32 * m


def test_design_matrix_matches_reference_shapes(golden):
    from gmat_b200.uvlmm.design_matrix import design_matrix_wemai_multi_gmat
    bed = os.path.join(GOLD, "mouse", "plink")
    y, xmat, zmat = design_matrix_wemai_multi_gmat(os.path.join(GOLD, "mouse", "pheno"), bed)
    assert list(xmat.shape) == golden["xmat_shape"] and np.allclose(y[:8, 0], golden["y_first8"])
    assert zmat.shape == (1304, 1304) and np.array_equal(zmat.toarray(), np.eye(1304))
    y3, x3, z3 = design_matrix_wemai_multi_gmat(os.path.join(GOLD, "mouse", "pheno_rep3"), bed)
    assert z3.shape == (3912, 1304) and np.array_equal(np.asarray(z3.sum(axis=0)).ravel(), np.full(1304, 3.0))
    assert np.array_equal(z3.indices[:6], [0, 0, 0, 1, 1, 1])


def test_host_preamble_matches_reference(golden, mouse_model):
    from gmat_b200.remma._core import preamble
    M = mouse_model
    pymat, pvpmat = preamble(M["y"], M["xmat"], M["zmat"], M["gl_aa"], M["vc_aa"], scipy_inv=True)
    np.testing.assert_allclose(pymat.ravel(), golden["pymat_ag_agag"], rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(np.diag(pvpmat)[:16], golden["pvpmat_diag_first16"], rtol=1e-9)


def test_random_pair_restatement(tmp_path):
    from gmat_b200.remma import random_pair, random_pairAD
    np.random.seed(3)
    a = random_pair(500, out_file=str(tmp_path / "rp"), num_pair=6000)
    assert a.shape == (6000, 2) and (a[:, 0] < a[:, 1]).all() and len({tuple(r) for r in a}) == 6000
    assert open(str(tmp_path / "rp")).readline() == "snp_0 snp_1\n"
    assert np.array_equal(a, np.loadtxt(str(tmp_path / "rp"), dtype=int, skiprows=1))   # what the reference reads back
    b = random_pairAD(500, out_file=str(tmp_path / "rq"), num_pair=6000)
    assert (b[:, 0] != b[:, 1]).all() and (b[:, 0] > b[:, 1]).any()
    # same RNG state, same process (same string hashes): the same pairs in the same order and the same file bytes as
    # the step-by-step restatement of random_pair.py (the product builds the strings and the file without numpy's
    # text machinery)
    from oracle import oracle
    for fn, ordered, shape in ((random_pair, True, (1407, 20000)), (random_pairAD, False, (1407, 20000)),
                               (random_pair, True, (50000, 100000))):
        np.random.seed(11)
        got = fn(shape[0], out_file=str(tmp_path / "g"), num_pair=shape[1])
        state = np.random.get_state()[1].copy()
        np.random.seed(11)
        want = oracle.random_pairs(shape[0], str(tmp_path / "w"), num_pair=shape[1], ordered=ordered)
        assert np.array_equal(got, want) and got.dtype == want.dtype
        assert open(str(tmp_path / "g"), "rb").read() == open(str(tmp_path / "w"), "rb").read()
        assert np.array_equal(state, np.random.get_state()[1])        # and the RNG is left where the reference leaves it


def test_synth_bed_roundtrip(tmp_path):
    import synth
    from oracle import oracle
    prefix = str(tmp_path / "s")
    ds = synth.write_dataset(prefix, 37, 19, seed=1, missing_rate=0.1)
    raw = oracle.read_bed_raw(prefix, 37, 19)
    g = ds["geno"].astype(float)
    want = np.where(g < 0, 1.0 / 3, g)
    assert np.allclose(raw, want)


_GLOO_WORKER = r'''
import os, sys
sys.path.insert(0, %r)
import numpy as np, torch, torch.distributed as dist
from gmat_b200.engine import HIT_DTYPE
from gmat_b200.remma import parallel as par
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%%s" %% sys.argv[2], rank=int(sys.argv[1]), world_size=2)
rank = dist.get_rank()
m = 1000
lst = par.shard_first_snps(list(range(m - 1)), m, False, rank, 2)
hits = np.zeros(len(lst) // (7 + rank), dtype=HIT_DTYPE)       # different counts per rank, rank-tagged payload
hits["snp_0"] = lst[:len(hits)]
hits["snp_1"] = rank
hits["eff"] = np.arange(len(hits)) + 0.5 * rank
out = par.gather_hits(hits, torch.device("cpu"))
empty = par.gather_hits(np.zeros(0, dtype=HIT_DTYPE), torch.device("cpu"))
# the speed-aware deal's exchange of measured rates, and the deal itself on two ranks: every first SNP once
assert par.all_gather_floats(rank + 0.5, torch.device("cpu")) == [0.5, 1.5]
head, head_pairs, rest = par.calibration_split(list(range(m - 1)), m, False, rank, 2)
rates = par.all_gather_floats([1.0, 0.8][rank], torch.device("cpu"))
tail = par.redeal(rest, rates, rank)
cover = par.gather_hits(np.array([(i, 0, 0.0, 0.0, 0.0) for i in head + tail], dtype=HIT_DTYPE), torch.device("cpu"))
if rank == 0:
    assert np.array_equal(np.sort(cover["snp_0"]), np.arange(m - 1))
if rank == 0:
    n0 = len(par.shard_first_snps(list(range(m - 1)), m, False, 0, 2)) // 7
    n1 = len(par.shard_first_snps(list(range(m - 1)), m, False, 1, 2)) // 8
    assert len(out) == n0 + n1 and (out["snp_1"][:n0] == 0).all() and (out["snp_1"][n0:] == 1).all()
    assert np.allclose(out["eff"][n0:], np.arange(n1) + 0.5) and len(empty) == 0
    print("GLOO_OK")
else:
    assert out is None
dist.destroy_process_group()
'''


def test_two_rank_hit_gather_gloo(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_GLOO_WORKER % ROOT)
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    procs = [subprocess.Popen([sys.executable, str(script), str(r), str(port)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert "GLOO_OK" in outs[0]


def test_savetxt_host_is_byte_identical_to_numpy(tmp_path):
    """gmat_b200_savetxt_host == np.savetxt(path, mat) (the 'mat' output of agmat, gmatrix.py:10-14), no GPU."""
    from gmat_b200._ffi import ffi, lib
    rng = np.random.Generator(np.random.PCG64(4))
    mats = [rng.normal(size=(37, 53)) * 10.0 ** rng.integers(-30, 30, size=(37, 53)),
            np.array([[0.0, -0.0, 1.0, -1.0, 1e-310, 1.7976931348623157e308, 5e-324, 0.1, 1 / 3]]),
            np.array([[np.nan, -np.nan, np.inf, -np.inf, np.copysign(np.nan, -1.0)]]),
            rng.integers(0, 2 ** 64, size=(400, 25), dtype=np.uint64).view(np.float64),   # any bit pattern
            rng.normal(size=(1300, 7)), np.zeros((0, 4)), rng.normal(size=(5, 1))]
    for k, mat in enumerate(mats):
        mat = np.ascontiguousarray(mat)
        a, b = str(tmp_path / ("np%d.txt" % k)), str(tmp_path / ("c%d.txt" % k))
        np.savetxt(a, mat)
        for threads in (1, 0):
            assert lib.gmat_b200_savetxt_host(b.encode(), ffi.cast("double *", mat.ctypes.data), mat.shape[0],
                                              mat.shape[1], threads) == 1
            assert open(a, "rb").read() == open(b, "rb").read()
    assert lib.gmat_b200_savetxt_host(str(tmp_path / "no" / "dir.txt").encode(), ffi.NULL, 0, 0, 1) == -2


def _reference_output_mat(mat, id, out_file, out_fmt):
    """The two list branches of the reference's output_mat, restated verbatim in behaviour (gmatrix.py:15-29)."""
    import pandas as pd
    ind = np.tril_indices_from(mat)
    if out_fmt == 'row_col_val':
        df = pd.DataFrame({"row": ind[0] + 1, "col": ind[1] + 1, "val": mat[ind]})
        df.to_csv(out_file + '1', sep=' ', index=False, header=False, columns=["row", "col", "val"])
    else:
        df = pd.DataFrame({"id0": id[ind[0]], "id1": id[ind[1]], "val": mat[ind]})
        df.to_csv(out_file + '2', sep=' ', index=False, header=False, columns=["id0", "id1", "val"])


def test_list_format_writers_are_byte_identical_to_pandas(tmp_path):
    """output_mat(..., 'row_col_val' / 'id_id_val') == the reference's pandas to_csv of the lower triangle
    (gmatrix.py:15-29): repr(float) values incl. exponent forms, integers, zeros, NaN as the empty string; integer,
    string and float individual ids; one thread and all threads; no GPU."""
    from gmat_b200._ffi import ffi, lib
    from gmat_b200.gmatrix.gmatrix import output_mat
    rng = np.random.Generator(np.random.PCG64(8))
    n = 483                                                    # 116,886 lines: more than one chunk per thread
    mat = rng.normal(size=(n, n)) * 10.0 ** rng.integers(-12, 20, size=(n, n))
    mat[0, 0], mat[1, 0], mat[1, 1], mat[2, 0], mat[2, 1] = 0.0, -0.0, 1.0, 1e16, 1e-5
    mat[3, 0], mat[3, 1], mat[3, 2] = np.nan, np.inf, 123456.0
    id_sets = [np.arange(1000, 1000 + n), np.array(["id%d_x" % k for k in range(n)], dtype=object),
               np.arange(n) + 0.5]
    for k, ids in enumerate(id_sets):
        a, b = str(tmp_path / ("ref%d" % k)), str(tmp_path / ("got%d" % k))
        for fmt, ext in (("row_col_val", "1"), ("id_id_val", "2")):
            _reference_output_mat(mat, ids, a, fmt)
            assert output_mat(mat, ids, b, fmt) == 1
            assert open(a + ext, "rb").read() == open(b + ext, "rb").read(), (k, fmt)
    small = np.ascontiguousarray(mat[:5, :5])
    p1 = str(tmp_path / "t1")
    assert lib.gmat_b200_write_tril_host(p1.encode(), ffi.cast("double *", small.ctypes.data), 5, ffi.NULL, ffi.NULL,
                                         1) == 1
    _reference_output_mat(small, None, str(tmp_path / "t1ref"), "row_col_val")
    assert open(p1, "rb").read() == open(str(tmp_path / "t1ref") + "1", "rb").read()
    assert lib.gmat_b200_write_tril_host(str(tmp_path / "no" / "dir").encode(), ffi.NULL, 0, ffi.NULL, ffi.NULL, 1) == -2
    buf = ffi.new("char[40]")
    for v in [0.1, 1e-4, 9.9e-5, 1e15, 1e16, 2.0 ** 53, 5e-324, 1.7976931348623157e308, -1 / 3, 1e22, 1e23]:
        lib.gmat_b200_repr_host(v, buf)
        assert ffi.string(buf).decode() == repr(v)


def test_annotation_snp_pos_matches_reference(tmp_path):
    """annotation_snp_pos (README step 4, gmat/remma/annotation.py:22-74) against files written by the unmodified
    reference (fixtures made in the build container: p_cut filter, distance filter, LD filter)."""
    import shutil
    from gmat_b200.remma.annotation import annotation_snp_pos
    out = os.path.join(GOLD, "out")
    res = str(tmp_path / "epiAD.txt")
    shutil.copyfile(os.path.join(out, "epiAD.txt"), res)
    bed = os.path.join(GOLD, "mouse", "plink")
    assert annotation_snp_pos(res, bed, p_cut=1.0e-5, dis=0) == 0
    assert open(res + ".anno").read() == open(os.path.join(out, "epiAD.anno_p1e-5")).read()
    assert annotation_snp_pos(res, bed, p_cut=1, dis=5.0e7, ld_file=os.path.join(out, "epiAD.ld_table"), r2=0.2) == 0
    assert open(res + ".anno").read() == open(os.path.join(out, "epiAD.anno_dis5e7")).read()
    assert open(res + ".anno.ld").read() == open(os.path.join(out, "epiAD.anno_dis5e7.ld")).read()


def test_p_app_step_is_byte_identical_to_the_reference_loop(tmp_path):
    """The 'Add the approximate P values' step of the screen front ends, evaluated as arrays, writes the bytes the
    reference's per-line loop writes (remma_epiAA_eff.py:85-95 scalar var_app; remma_epiAA_maf_eff.py:88-101 table)."""
    from scipy.stats import chi2
    from gmat_b200.remma import _core
    rng = np.random.Generator(np.random.PCG64(3))
    m, k = 300, 2500
    rows = [(int(a), int(b), "%g" % e) for a, b, e in
            zip(rng.integers(0, m, k), rng.integers(0, m, k), rng.normal(size=k) * 10.0 ** rng.integers(-4, 5, k))]
    freq, freq2 = rng.integers(0, 11, m).astype(np.longlong), rng.integers(0, 11, m).astype(np.longlong)
    table = rng.uniform(0.5, 5.0, 111)
    for mode in ("scalar", "table", "empty"):
        tmp, out = str(tmp_path / "t.temp"), str(tmp_path / "t.out")
        use = [] if mode == "empty" else rows
        with open(tmp, "w") as f:
            f.write("snp_0 snp_1 eff\n")
            f.writelines("%d %d %s\n" % r for r in use)
        want = ["snp_0 snp_1 eff chi_app p_app\n"]
        for a, b, e in use:                                              # the reference's loop, verbatim in behaviour
            deno = 2.37 if mode == "scalar" else table[freq[a] * 10 + freq2[b]]
            chi_app = float(e) * float(e) / deno
            want.append(' '.join([str(a), str(b), e, str(chi_app), str(chi2.sf(chi_app, 1))]) + '\n')
        if mode == "table":
            _core._append_p_app(tmp, out, lambda rr: table[freq[[int(r[0]) for r in rr]] * 10
                                                           + freq2[[int(r[1]) for r in rr]]])
        else:
            _core._append_p_app(tmp, out, lambda rr: 2.37)
        assert open(out).read() == ''.join(want) and not os.path.exists(tmp)


def test_bench_reference_arm_contract():
    """`bench.py --impl reference` (the CPU arm the driver runs beside the GPU arm) works without a GPU, runs the
    UNMODIFIED reference from baseline/_ref (kind "reference"), prints one JSON line with the keys of the measurement
    contract and the SAME config as the GPU arm, and maps nothing of this repo's product (no gmat_b200 module, no
    libgmat_b200.so)."""
    import json
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if not os.path.isdir(os.path.join(root, "baseline", "_ref", "gmat")):
        subprocess.run(["bash", os.path.join(root, "baseline", "install_ref.sh")], check=True, capture_output=True)
    if not os.path.isdir(os.path.join(root, "baseline", "_ref", "gmat")):
        pytest.skip("baseline/_ref not installed and /root/reference absent")
    probe = ("import runpy, sys, json\n"
             "sys.argv = ['bench.py', '--impl', 'reference', '--steps', '1', '--warmup', '1']\n"
             "try:\n    runpy.run_path(%r, run_name='__main__')\nexcept SystemExit:\n    pass\n"
             "maps = open('/proc/self/maps').read()\n"
             "print(json.dumps({'mods': [m for m in sys.modules if m.startswith('gmat_b200')],"
             " 'so': 'libgmat_b200' in maps}), file=sys.stderr)\n" % os.path.join(root, "bench.py"))
    out = subprocess.run([sys.executable, "-c", probe], capture_output=True, text=True, timeout=900, cwd=root)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip().startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "snp_pairs_per_sec_remma_epiAA_exact" and d["unit"] == "pairs/s"
    assert d["value"] > 0 and d["higher_is_better"] is True and d["steps"] == 1 and d["n_gpus"] == 1
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["cores"] >= 1
    assert d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    sys.path.insert(0, root)
    import bench
    assert d["config"] == json.loads(json.dumps(bench.CONFIG))            # byte-identical config in both arms
    tail = json.loads(out.stderr.strip().splitlines()[-1])
    assert tail["mods"] == [] and tail["so"] is False, tail


def test_every_reference_front_end_exists_with_the_same_signature():
    """Every public function of the reference's remma_epi{AA,AD,DD} packages (fixture written from the unmodified
    reference by oracle/make_golden_signatures.py) exists in gmat_b200.remma.remma_epiXX with the same parameter names,
    order and defaults."""
    import importlib
    import inspect
    import json
    want = json.load(open(os.path.join(GOLD, "frontend_signatures.json")))
    for kind, table in want.items():
        mod = importlib.import_module("gmat_b200.remma.remma_epi" + kind)
        assert len(table) >= 18
        for name, params in table.items():
            assert hasattr(mod, name), name
            got = [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
                   for p in inspect.signature(getattr(mod, name)).parameters.values()]
            assert got == params, (name, got, params)


def test_bench_approx_verifier_accepts_a_correct_file_and_rejects_a_wrong_one(tmp_path):
    """bench.py's checker of the approximate pipeline's file (extra.northstar.blocks.AA_approx): a file built from the
    oracle's own numbers passes, one corrupted value / one significant row the exact scan does not hold fails."""
    import importlib.util
    from scipy.stats import chi2
    from oracle import oracle
    import synth
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    n, m = 90, 50
    pre = str(tmp_path / "d")
    synth.write_dataset(pre, n, m, seed=5, n_planted=3)
    from gmat_b200.uvlmm.design_matrix import design_matrix_wemai_multi_gmat
    y, xmat, zmat = design_matrix_wemai_multi_gmat(pre + ".pheno", pre)
    snp = oracle.read_plink(pre)
    x = snp - 2 * snp.mean(axis=0) / 2
    ag = x @ x.T / m
    pymat, pvpmat = oracle.preamble(y, xmat, zmat, [ag, ag * ag], [0.3, 0.1, 0.6])
    pairs = np.array([(i, j) for i in range(m - 1) for j in range(i + 1, m)][::7])
    eff, var, chi, p = bench.oracle_pair_stats(pre, n, m, "AA", pymat, pvpmat, pairs, chi2)
    # the same numbers by the oracle's own routine
    U, V = oracle.coded_pair("AA", snp)
    w = U[:, pairs[:, 0]] * V[:, pairs[:, 1]]
    np.testing.assert_allclose(eff, (w.T @ pymat).ravel(), rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(var, np.sum(w * (pvpmat @ w), axis=0), rtol=1e-12)
    var_med = float(np.median(var))

    def write(path, eff_rows):
        with open(path, "w") as f:
            f.write("snp_0 snp_1 eff var chi p_app p\n")
            for (a, b), e, v, c, q in zip(pairs.tolist(), eff_rows, var, chi, p):
                e_tr = float("%g" % e)
                f.write(" ".join([str(a), str(b), repr(float(e)), repr(float(v)), repr(float(c)),
                                  str(chi2.sf(e_tr * e_tr / var_med, 1)), repr(float(q))]) + "\n")
    bench.NS_P_CUT = cut = 0.05              # the tiny data set has nothing at the north-star cut
    exact = {(int(a), int(b)) for (a, b), q in zip(pairs.tolist(), p) if q < cut}
    good = str(tmp_path / "good")
    write(good, eff)
    res = bench.verify_approx_file(good, pre, n, m, pymat, pvpmat, exact, None, chi2)
    assert res["verified"], res
    assert res["rows_in_file"] == len(pairs) and abs(res["null_variance"] / var_med - 1) < 1e-6
    assert res["recall_of_exact_scan_rows"] == 1.0
    bad_eff = eff.copy()
    bad_eff[np.argmin(p)] *= 1 + 1e-6
    bad = str(tmp_path / "bad")
    write(bad, bad_eff)
    assert not bench.verify_approx_file(bad, pre, n, m, pymat, pvpmat, exact, None, chi2)["verified"]
    assert len(exact) > 3
    fewer = set(sorted(exact)[1:])
    assert not bench.verify_approx_file(good, pre, n, m, pymat, pvpmat, fewer, None, chi2)["verified"]


def test_speed_aware_deal_covers_the_list_once_and_follows_the_rates():
    """parallel.calibration_split + redeal (the speed-aware sharding of long multi-GPU scans): whatever the measured
    rates, every first SNP is scanned by exactly one rank; with equal rates the pair counts are balanced like the
    static deal; a rank measured 10 % slower gets ~10 % fewer pairs of the re-dealt part."""
    from gmat_b200.remma import parallel as par
    m = 50000
    for all_pairs, lst in ((False, np.arange(m - 1)), (True, np.arange(m)), (False, np.arange(0, m - 1, 8))):
        w = (lambda a: np.full(len(a), m) if all_pairs else m - 1 - np.asarray(a))
        for world, rates in ((2, [1.0, 1.0]), (4, [1.0, 0.9, 1.0, 1.05]), (8, [1.0] * 7 + [0.9])):
            heads, tails, head_pairs = [], [], []
            for r in range(world):
                head, hp, rest = par.calibration_split(lst, m, all_pairs, r, world)
                heads.append(head)
                head_pairs.append(hp)
                tails.append(par.redeal(rest, rates, r))
                assert hp == float(np.sum(w(head)))
            everything = np.concatenate([np.asarray(h + t, dtype=np.int64) for h, t in zip(heads, tails)])
            assert len(everything) == len(lst) and np.array_equal(np.sort(everything), lst)
            total = float(np.sum(w(lst)))
            blk = par.SHARD_BLOCK * m
            for r in range(world):                  # the calibration share: 8 % of a rank's pairs, to within a block
                assert 0.08 * (total / world - blk) <= head_pairs[r] <= 0.08 * (total / world + blk) + blk
            tail_pairs = np.array([float(np.sum(w(t))) for t in tails])
            want = np.asarray(rates) / np.sum(rates) * tail_pairs.sum()
            assert np.max(np.abs(tail_pairs - want)) <= 2 * par.SHARD_BLOCK * m
    # equal rates = the static deal
    lst = np.arange(m - 1)
    _, starts, wblk = par._blocks(lst, m, False)
    owner = par.deal_blocks(wblk, [1.0] * 4)
    for r in range(4):
        mine = np.concatenate([lst[s:s + par.SHARD_BLOCK] for s, o in zip(starts, owner) if o == r])
        assert np.array_equal(mine, par.shard_first_snps(lst, m, False, r, 4))
