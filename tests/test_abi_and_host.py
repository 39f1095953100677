"""CPU tests: the C-ABI library loads and exports every symbol include/tpb.h
declares; the Python host layer fails loudly without a GPU (no fallback); the
workload generators are deterministic and shard-consistent."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "tpb.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(tpb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(pkg):
    if not os.path.exists(pkg.lib_path()):
        from __graft_entry__ import build
        build()
    lib = ctypes.CDLL(pkg.lib_path())
    syms = _declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), f"libtpb.so does not export {s}"
    assert pkg.lib().tpb_abi_version() == 4


def test_header_compiles_as_c():
    import subprocess
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        f = os.path.join(d, "t.c")
        open(f, "w").write('#include "tpb.h"\nint main(void){tpb_config c; (void)c; return 0;}\n')
        subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                        "-c", f, "-o", os.path.join(d, "t.o")], check=True)


def test_no_cpu_fallback(pkg):
    """Without a CUDA device the engine refuses to exist (kFailedPrecondition)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.TpbError) as ei:
        pkg.Engine(max_paths=4, num_samples=16, num_dofs=3, max_waypoints=3)
    assert ei.value.code == 9
    with pytest.raises(pkg.TpbError):
        knots, ctrl = pkg.workloads.dense_splines(1, 3, 2, 8)
        pkg.spline_eval_host(knots, ctrl, 3, 2, u=np.zeros((1, 4)))


def test_product_does_not_touch_the_oracle():
    """The product (package + csrc + host) must not import/link/execute oracle/."""
    pkg_dir = os.path.join(ROOT, "x-edr-trajectory-planning_b200")
    for dp, _, files in os.walk(pkg_dir):
        if os.sep + "lib" in dp:
            continue
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cc", ".cpp")) or fn == "Makefile":
                txt = open(os.path.join(dp, fn), errors="ignore").read()
                assert "pyoracle" not in txt and "tp_oracle" not in txt and "oracle/" not in txt, \
                    os.path.join(dp, fn)


def test_workloads_deterministic_and_shardable(pkg):
    W = pkg.workloads
    a = W.joint_paths(64, 10, 7, seed=1000)
    b = W.joint_paths(64, 10, 7, seed=1000)
    for x, y in zip(a, b):
        np.testing.assert_array_equal(x, y)
    big = W.joint_paths(256, 10, 7, seed=1000)
    for x, y in zip(a, big):
        np.testing.assert_array_equal(x, y[:64])  # shard == rows of the full batch
    wp, vmax, amax = a
    assert np.abs(wp).max() <= np.pi and vmax.min() >= 0.5 and amax.max() <= 4.0
    k = W.clamped_uniform_knots(5, 3)
    assert list(k) == [0, 0, 0, 0, 0.5, 1, 1, 1, 1]
