"""The multi-GPU front-end (tpb_multi, csrc/tpb_multi.cu): a batch sharded in contiguous
blocks over logical shards of one GPU and over all GPUs present gives, bit for bit, the
single-engine result -- durations, status, profiles, offsets and packed rows -- and
tpb_shard_range is the partition of SURVEY.md 8(e)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()
pytestmark = pytest.mark.gpu

B, W, D, N = 203, 5, 7, 160
ROWS = ("r_pos", "r_vel", "r_acc", "r_s", "r_sd", "r_sdd", "r_time")


def single_engine(wp, vmax, amax, nw):
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=B * 8192)
    eng.set_joint_waypoints(wp, nw)
    eng.set_joint_limits(vmax, amax)
    eng.sample_paths()
    eng.optimize()
    total = eng.resample(0.0, 0.004)
    A = pkg.ARR
    res = {k: eng.get(getattr(A, k)) for k in ("DURATION", "STATUS", "TIME", "S", "SD", "SDD", "Q",
                                               "R_OFFSET", "R_POS", "R_VEL", "R_ACC", "R_S",
                                               "R_SD", "R_SDD", "R_TIME")}
    res["total"] = total
    eng.close()
    return res


def device_lists():
    import torch
    n = torch.cuda.device_count()
    lists = [[0, 0], [0, 0, 0], [0] * 7]          # logical shards of one GPU
    if n > 1:
        lists.append(list(range(n)))                # all GPUs present
        lists.append(list(range(n)) * 2)            # two logical shards on each
    return lists


def test_multi_plan_joint_host_equals_single_engine():
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=21)
    nw = np.full(B, W, np.int32)
    nw[::9] = 2
    nw[4::13] = 1
    ref = single_engine(wp, vmax, amax, nw)
    cap = int(ref["total"])
    for devices in device_lists():
        me = pkg.MultiEngine(B, N, D, W, devices=devices, max_time_samples_total=B * 8192)
        out = {"duration": np.full(B, np.nan), "status": np.full(B, -1, np.int32),
               "r_offset": np.full(B + 1, -1, np.int64)}
        out.update({k: np.full((cap, D), np.nan) for k in ("r_pos", "r_vel", "r_acc")})
        out.update({k: np.full(cap, np.nan) for k in ("r_s", "r_sd", "r_sdd", "r_time")})
        out.update({k: np.full((B, N), np.nan) for k in ("profile_time", "profile_s",
                                                         "profile_sd", "profile_sdd")})
        for _ in range(2):
            total = me.plan_joint_host(wp, vmax, amax, 0.0, 0.004, num_waypoints=nw, out=out)
        assert total == cap, devices
        assert out["duration"].tobytes() == ref["DURATION"].tobytes(), devices
        np.testing.assert_array_equal(out["status"], ref["STATUS"])
        np.testing.assert_array_equal(out["r_offset"], ref["R_OFFSET"])
        for k in ROWS:
            assert out[k].tobytes() == ref[k.upper()][:cap].tobytes(), (devices, k)
        for k, a in (("profile_time", "TIME"), ("profile_s", "S"), ("profile_sd", "SD"),
                     ("profile_sdd", "SDD")):
            assert out[k].tobytes() == ref[a].tobytes(), (devices, k)
        # the shards tile the batch and the rows
        G = len(devices)
        path, row = 0, 0
        for g in range(G):
            s = me.shard(g)
            assert s["device"] == devices[g]
            assert (s["first_path"], s["num_paths"]) == pkg.shard_range(B, G, g)
            assert s["first_path"] == path and s["first_row"] == row
            path += s["num_paths"]
            row += s["num_rows"]
        assert path == B and row == cap
        # too small a host capacity is reported, not overrun
        small = dict(out, r_pos=np.zeros((cap // 2, D)))
        with pytest.raises(pkg.TpbError) as ei:
            me.plan_joint_host(wp, vmax, amax, 0.0, 0.004, num_waypoints=nw, out=small)
        assert ei.value.code == 11
        me.close()


def test_multi_device_resident_plan_and_gather():
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=22)
    ref = single_engine(wp, vmax, amax, None)
    for devices in device_lists()[:2] + device_lists()[3:4]:
        me = pkg.MultiEngine(B, N, D, W, devices=devices, max_time_samples_total=B * 8192)
        me.set_joint_paths(wp, vmax, amax)
        assert me.plan_device(0.0, 0.004) == ref["total"]
        assert me.gather(pkg.ARR.DURATION).tobytes() == ref["DURATION"].tobytes()
        np.testing.assert_array_equal(me.gather(pkg.ARR.STATUS, dtype=np.int32), ref["STATUS"])
        assert me.gather(pkg.ARR.TIME, (N,)).tobytes() == ref["TIME"].tobytes()
        assert me.gather(pkg.ARR.Q, (N, D)).tobytes() == ref["Q"].tobytes()
        assert me.kernel_launches > 0
        # fewer paths than shards: trailing shards stay empty
        few = 2
        me.set_joint_paths(wp[:few], vmax[:few], amax[:few])
        me.plan_device(0.0, 0.0)
        assert me.gather(pkg.ARR.DURATION).tobytes() == ref["DURATION"][:few].tobytes()
        with pytest.raises(pkg.TpbError) as ei:
            me.set_joint_paths(np.zeros((B + 1, W, D)), np.ones((B + 1, D)), np.ones((B + 1, D)))
        assert ei.value.code == 11
        me.close()
