"""Device-resident TrajectoryBuffers (tpb_tbuf_*) against the CPU oracle, bit for
bit: the scenarios of trajectory_buffer_test.cc, seeded random operation sequences
(one buffer, and many buffers running different sequences through the same batched
calls), the golden observations written from the reference's object code, and
size-independent properties on the resampled trajectories of a timed batch (device
pointers end to end)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import pyoracle as orc  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402
import tbuf_cases as TC  # noqa: E402
import test_oracle_tbuf as T  # noqa: E402

pkg = load_package()
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("scenario", [
    T.test_insert_segment_works, T.test_append_sample_works,
    T.test_insert_segment_uses_timestep_tolerance, T.test_clear_and_discard_work,
    T.test_get_positions_up_to_time_and_at_time, T.test_stop_at_index_and_before_time])
def test_reference_test_scenarios(scenario):
    scenario("gpu")


@pytest.mark.parametrize("seed", range(6))
def test_one_buffer_random_sequences_are_bit_identical_to_the_oracle(seed):
    spec = TC.ops_spec(seed, D=[7, 3, 9, 1][seed % 4], num_ops=60, tol=[1e-6, 1e-5, 1e-3][seed % 3])
    ops = TC.generate_ops(spec, orc.TrajectoryBuffer(spec["D"], spec["tol"]))
    want = TC.run_ops(orc.TrajectoryBuffer(spec["D"], spec["tol"]), ops)
    got = TC.run_ops(pkg.TrajectoryBufferBatch(3, spec["D"], 1024, spec["tol"]).view(2), ops)
    for k, (x, y) in enumerate(zip(got, want)):
        assert x == y, (seed, k, ops[k][0], {f: (x[f], y[f]) for f in ("result", "n", "seq")})


def test_golden_observations_from_the_reference_object_code():
    g = np.load(T.GOLDEN)
    for c in range(int(g["num_cases"])):
        spec = dict(seed=int(g[f"seed_{c}"]), D=int(g[f"D_{c}"]), num_ops=int(g[f"num_ops_{c}"]),
                    dt=float(g[f"dt_{c}"]), tol=float(g[f"tol_{c}"]))
        ops = TC.generate_ops(spec, orc.TrajectoryBuffer(spec["D"], spec["tol"]))
        obs = TC.run_ops(pkg.TrajectoryBufferBatch(2, spec["D"], 1024, spec["tol"]).view(0), ops)
        assert TC.digest(obs) == str(g[f"digest_{c}"])


def run_batched(batch, ops_per_buffer):
    """Step k applies operation k of every buffer: the buffers whose operation is of
    the same kind share one batched call, the others are skipped by it."""
    B, D = batch.B, batch.D
    results = [[None] * len(ops_per_buffer[0]) for _ in range(B)]
    for k in range(len(ops_per_buffer[0])):
        kinds = {}
        for b in range(B):
            kinds.setdefault(ops_per_buffer[b][k][0], []).append(b)
        for kind, members in sorted(kinds.items()):
            arg = lambda b, i: ops_per_buffer[b][k][1][i]
            sel = np.zeros(B, np.int32)
            sel[members] = 1
            if kind == "clear":
                batch.clear(sel)
                st = None
            elif kind == "insert_segment":
                length = np.full(B, -1, np.int32)
                begin = np.zeros(B, np.int64)
                row = 0
                for b in members:
                    begin[b], length[b] = row, len(arg(b, 0))
                    row += length[b]
                cat = lambda i: (np.concatenate([np.asarray(arg(b, i), np.float64).reshape(
                    length[b], D if i else 1) for b in members]) if row else np.zeros((0, D)))
                st = batch.insert_segments(begin, length, cat(0).reshape(-1), cat(1), cat(2), cat(3))
            elif kind == "append_sample":
                t = np.zeros(B)
                rows = [np.zeros((B, D)) for _ in range(3)]
                for b in members:
                    t[b] = arg(b, 0)
                    for i in range(3):
                        rows[i][b] = arg(b, i + 1)
                st = batch.append_samples(t, *rows, select=sel)
            elif kind in ("discard_before_sec", "discard_before_ns"):
                t = np.full(B, np.nan)
                for b in members:
                    t[b] = arg(b, 0) if kind.endswith("sec") else pkg.time_to_sec(arg(b, 0))
                batch.discard_before(t)
                st = None
            elif kind == "positions_up_to_time_ns":
                t = np.full(B, np.nan)
                for b in members:
                    t[b] = pkg.time_to_sec(arg(b, 0))
                st = batch.positions_up_to_time(t)
            elif kind == "get_at_time":
                t = np.full((B, 1), np.nan)
                for b in members:
                    t[b, 0] = arg(b, 1)
                s4 = batch.sample(t)
                st = [(int(s4[0][b, 0]), s4[1 + arg(b, 0)][b, 0].copy()) if b in members else None
                      for b in range(B)]
            elif kind == "stop_at_index":
                idx = np.full(B, pkg.SKIP_INDEX, np.int32)
                amax = np.ones((B, D))
                steps = {arg(b, 2) for b in members}
                st = np.zeros(B, np.int32)
                for step in sorted(steps):  # time_step is one scalar per call
                    idx[:] = pkg.SKIP_INDEX
                    for b in members:
                        if arg(b, 2) == step:
                            idx[b], amax[b] = arg(b, 0), arg(b, 1)
                    r = batch.stop_at_index(idx, amax, step)
                    for b in members:
                        if arg(b, 2) == step:
                            st[b] = r[b]
            elif kind == "stop_before_time":
                t = np.full(B, np.nan)
                amax = np.ones((B, D))
                for b in members:
                    t[b], amax[b] = arg(b, 0), arg(b, 1)
                st = batch.stop_before_time(t, amax, arg(members[0], 2))
            elif kind == "add_offset":
                off = np.zeros(B)
                for b in members:
                    off[b] = arg(b, 0)
                batch.add_offset(off)
                st = None
            else:
                raise AssertionError(kind)
            for b in members:
                results[b][k] = None if st is None else st[b]
    return results


def test_many_buffers_with_different_sequences_through_the_same_batched_calls():
    B, D, tol = 24, 7, 1e-6
    specs = [TC.ops_spec(500 + b, D=D, num_ops=50, tol=tol) for b in range(B)]
    ops = [TC.generate_ops(s, orc.TrajectoryBuffer(D, tol)) for s in specs]
    want = [TC.run_ops(orc.TrajectoryBuffer(D, tol), o) for o in ops]
    batch = pkg.TrajectoryBufferBatch(B, D, 1024, tol)
    results = run_batched(batch, ops)
    state = batch.state()
    for b in range(B):
        for k, (r, w) in enumerate(zip(results[b], want[b])):
            wr = w["result"]
            if isinstance(wr, tuple):
                assert int(r[0]) == wr[0] and (wr[0] != 0 or r[1].tobytes() == wr[1]), (b, k)
            else:
                assert (None if r is None else int(r)) == wr, (b, k, ops[b][k][0])
        last = want[b][-1]
        assert state["num_samples"][b] == last["n"] and state["sequence_number"][b] == last["seq"]
        t, p, v, a = batch.rows(b)
        assert (t.tobytes(), p.tobytes(), v.tobytes(), a.tobytes()) == (
            last["t"], last["p"], last["v"], last["a"]), b


def test_capacity_is_reported_and_rows_move_to_the_front_when_needed():
    D, cap = 3, 64
    batch = pkg.TrajectoryBufferBatch(2, D, cap)
    buf, ref = batch.view(0), orc.TrajectoryBuffer(D)
    rng = np.random.default_rng(1)
    seg = TC.smooth_segment(rng, 0.0, 60, D, 0.01)
    assert buf.insert_segment(*seg) == ref.insert_segment(*seg) == 0
    # discard 50 rows in front (head moves), then insert 40 behind the 10 kept rows:
    # 50 + 10 + 40 > 64 slots, so the kept rows move to slot 0
    for b in (buf, ref):
        b.discard_before_sec(seg[0][50])
    seg2 = TC.smooth_segment(rng, float(seg[0][-1]), 40, D, 0.01)
    assert buf.insert_segment(*seg2) == ref.insert_segment(*seg2) == 0
    for x, y in zip(buf.rows(), ref.rows()):
        assert x.tobytes() == y.tobytes()
    # more rows than slots: kResourceExhausted, buffer and sequence number unchanged
    before, seq = buf.rows(), buf.sequence_number
    seg3 = TC.smooth_segment(rng, float(seg2[0][5]), 70, D, 0.01)
    assert buf.insert_segment(*seg3) == 8
    assert buf.sequence_number == seq
    for x, y in zip(buf.rows(), before):
        assert x.tobytes() == y.tobytes()
    # appending into a full window behind a moved head
    small = pkg.TrajectoryBufferBatch(1, D, 8).view(0)
    ref2 = orc.TrajectoryBuffer(D)
    for k in range(8):
        row = (0.1 * k, np.full(D, k), np.full(D, 2.0 * k), np.full(D, 3.0 * k))
        assert small.append_sample(*row) == ref2.append_sample(*row) == 0
    assert small.append_sample(1.0, np.zeros(D), np.zeros(D), np.zeros(D)) == 8
    for b in (small, ref2):
        b.discard_before_sec(0.35)
    row = (1.0, np.full(D, -1.0), np.full(D, -2.0), np.full(D, -3.0))
    assert small.append_sample(*row) == ref2.append_sample(*row) == 0
    for x, y in zip(small.rows(), ref2.rows()):
        assert x.tobytes() == y.tobytes()


class DevArray:
    """A raw device pointer with the two attributes the binding looks at."""
    is_cuda = True

    def __init__(self, ptr, shape):
        self.ptr, self.shape = ptr, shape

    def data_ptr(self):
        return self.ptr

    def is_contiguous(self):
        return True


def test_timed_batch_goes_into_buffers_without_leaving_the_device():
    """The engine's packed resampled trajectories (TPB_ARR_R_*) are inserted by device
    pointer; sampling every buffer at its own time stamps returns its rows bit for bit,
    discarding + stopping match the oracle on a few buffers."""
    import torch
    B, W, D, N, dt = 96, 8, 7, 512, 0.004
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=77)
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=B * 8192)
    eng.set_joint_waypoints(wp)
    eng.set_joint_limits(vmax, amax)
    eng.sample_paths()
    eng.optimize()
    total = eng.resample(0.0, dt)
    off = eng.get(pkg.ARR.R_OFFSET)
    r_t, r_p = eng.get(pkg.ARR.R_TIME), eng.get(pkg.ARR.R_POS)
    r_v, r_a = eng.get(pkg.ARR.R_VEL), eng.get(pkg.ARR.R_ACC)
    status = eng.get(pkg.ARR.STATUS)
    assert (status == 0).all()
    counts = np.diff(off).astype(np.int32)
    cap = int(counts.max()) + 8
    batch = pkg.TrajectoryBufferBatch(B, D, cap)
    dev = lambda what, shape: DevArray(eng.devptr(what)[0], shape)
    length = torch.tensor(counts, dtype=torch.int32, device="cuda")
    st = torch.full((B,), -1, dtype=torch.int32, device="cuda")
    batch.insert_segments(dev(pkg.ARR.R_OFFSET, (B + 1,)), length, dev(pkg.ARR.R_TIME, (total,)),
                          dev(pkg.ARR.R_POS, (total, D)), dev(pkg.ARR.R_VEL, (total, D)),
                          dev(pkg.ARR.R_ACC, (total, D)), status=st)
    batch.sync()
    assert (st.cpu().numpy() == 0).all()
    state = batch.state()
    assert (state["num_samples"] == counts).all()
    assert (state["front_time"] == r_t[off[:-1]]).all() and (state["back_time"] == r_t[off[1:] - 1]).all()
    # sampling at the buffers' own time stamps returns the rows themselves
    Q = int(counts.min())
    tq = np.stack([r_t[off[b]:off[b] + Q] for b in range(B)])
    s, p, v, a = batch.sample(tq)
    assert (s == 0).all()
    for b in range(B):
        sl = slice(off[b], off[b] + Q)
        assert p[b].tobytes() == r_p[sl].tobytes() and v[b].tobytes() == r_v[sl].tobytes()
        assert a[b].tobytes() == r_a[sl].tobytes()
    # between the time stamps: the oracle's interpolation, and the documented tolerance
    tm = tq[:, :-1] + 0.37 * np.diff(tq, axis=1)
    s, p, v, a = batch.sample(tm)
    assert (s == 0).all()
    for b in (0, 17, B - 1):
        ref = orc.TrajectoryBuffer(D)
        sl = slice(off[b], off[b + 1])
        assert ref.insert_segment(r_t[sl], r_p[sl], r_v[sl], r_a[sl]) == 0
        for q in (0, 5, Q - 2):
            for which, got in enumerate((p, v, a)):
                assert got[b, q].tobytes() == ref.get_at_time(which, tm[b, q])[1].tobytes()
    # discard the first third and stop two thirds in, all buffers at once
    t_cut = np.array([r_t[off[b] + counts[b] // 3] + 0.001 for b in range(B)])
    t_stop = np.array([r_t[off[b] + 2 * counts[b] // 3] for b in range(B)])
    batch.discard_before(t_cut)
    st = batch.stop_before_time(t_stop, amax, dt)
    state = batch.state()
    for b in range(B):
        ref = orc.TrajectoryBuffer(D)
        sl = slice(off[b], off[b + 1])
        ref.insert_segment(r_t[sl], r_p[sl], r_v[sl], r_a[sl])
        ref.discard_before_sec(t_cut[b])
        assert st[b] == ref.stop_before_time(t_stop[b], amax[b], dt), b
        assert state["num_samples"][b] == ref.num_samples
        if b % 8 == 0:
            for x, y in zip(batch.rows(b), ref.rows()):
                assert x.tobytes() == y.tobytes()
    assert (st == 0).sum() > B // 2
    eng.close()
    batch.close()
