"""tpb_resample_uniform_noalloc (include/tpb.h): the allocation-free variant of
tpb_resample whose scratch (the per-path row counts) lives in a file-scope
`static __device__` array.  No counterpart in the reference.
Called through the C ABI and compared with the existing entry
point (tpb_resample on an engine) bit for bit -- both run the same resampling kernels --
on two calls in a row that use different time steps, so the second call sees the first
call's stale scratch.  Collected last on purpose."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()
pytestmark = pytest.mark.gpu

B, W, D, N = 24, 6, 7, 256
ROWS = ("R_TIME", "R_S", "R_SD", "R_SDD", "R_POS", "R_VEL", "R_ACC")


def test_noalloc_resample_equals_engine_resample_twice():
    import torch

    dev = torch.device("cuda", 0)
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=77)
    cap = B * 8192
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=cap)
    eng.set_joint_waypoints(wp)
    eng.set_joint_limits(vmax, amax)
    eng.sample_paths()
    eng.optimize()
    eng.sync()
    status = eng.get(pkg.ARR.STATUS)
    assert (status == 0).sum() >= B - 2

    # inputs: the engine's own device arrays + the limits uploaded by the test
    d_amax = torch.from_numpy(amax).to(dev)
    names = {"time": pkg.ARR.TIME, "s": pkg.ARR.S, "sd": pkg.ARR.SD, "sdd": pkg.ARR.SDD,
             "q": pkg.ARR.Q, "dq": pkg.ARR.DQ, "ddq": pkg.ARR.DDQ}
    inputs = {k: eng.devptr(v)[0] for k, v in names.items()}
    inputs["max_acceleration"] = d_amax.data_ptr()
    # caller-provided outputs (allocated here, by the test -- the entry point allocates nothing)
    out = {"r_offset": torch.full((B + 1,), -1, dtype=torch.int64, device=dev)}
    for k in ("r_time", "r_s", "r_sd", "r_sdd"):
        out[k] = torch.full((cap,), float("nan"), dtype=torch.float64, device=dev)
    for k in ("r_pos", "r_vel", "r_acc"):
        out[k] = torch.full((cap, D), float("nan"), dtype=torch.float64, device=dev)
    outputs = {k: v.data_ptr() for k, v in out.items()}
    d_status = torch.from_numpy(status.astype(np.int32)).to(dev)

    # two calls in a row; the second one has other row counts than the scratch still holds,
    # and reads the path status from a caller array instead of treating every path as timed
    for call, (start, dt, st_ptr) in enumerate(((0.0, 0.004, None), (0.1, 0.01, d_status.data_ptr()))):
        total = eng.resample(start, dt)  # the existing entry point
        eng.sync()
        want_off = eng.get(pkg.ARR.R_OFFSET).copy()
        want = {k: eng.get(getattr(pkg.ARR, k)).copy() for k in ROWS}
        assert total == want_off[-1] and 0 < total <= cap
        for v in out.values():
            v.fill_(-1 if v.dtype == torch.int64 else float("nan"))
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(dev)[0]
        rc, msg = pkg.resample_uniform_noalloc(0, B, N, D, dict(inputs, status=st_ptr), outputs,
                                               start, dt, cap, stream=None)
        torch.cuda.synchronize()
        free_after = torch.cuda.mem_get_info(dev)[0]
        assert rc == 0, msg
        if call == 1:  # (the first launch may load the kernels' module)
            assert free_after == free_before, "the call changed the free device memory"
        got_off = out["r_offset"].cpu().numpy()
        ok = status == 0
        if st_ptr is None and not ok.all():
            # without a status array failed paths are resampled too: compare the others' counts
            assert np.array_equal(np.diff(got_off)[ok], np.diff(want_off)[ok])
            continue
        assert np.array_equal(got_off, want_off), call
        for k in ROWS:
            got = out[k.lower()].cpu().numpy()[:total]
            assert got.tobytes() == want[k][:total].tobytes(), (call, k)
        # rows beyond the batch stay untouched
        assert np.isnan(out["r_time"].cpu().numpy()[total:]).all()
    eng.close()


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    rc, msg = pkg.resample_uniform_noalloc(0, 65537, N, D, {}, {}, 0.0, 0.004, 16)
    assert rc == 11 and "65536" in msg  # TPB_OUT_OF_RANGE
    rc, msg = pkg.resample_uniform_noalloc(0, 16, N, D, {}, {}, 0.0, 0.004, 16)
    assert rc == 3 and "null" in msg    # TPB_INVALID_ARGUMENT
