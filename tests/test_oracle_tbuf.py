"""TrajectoryBuffer: the CPU restatement (oracle/tp_oracle_tbuf.c) against the
reference's own object code (oracle/_ref, trajectory_buffer.cc compiled verbatim),
against golden observations written from that object code
(tests/golden/trajectory_buffer.npz), and on the scenarios of
trajectory_buffer_test.cc.  CPU only."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import pyoracle as orc  # noqa: E402
import tbuf_cases as TC  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden", "trajectory_buffer.npz")
OK, INVALID, NOT_FOUND, FAILED_PRECONDITION, OUT_OF_RANGE = 0, 3, 5, 9, 11
D = TC.K_NUM_JOINTS
DT = TC.K_SAMPLE_TIME


def ns(sec):
    return orc.lib().orc_time_from_sec(float(sec))


def make(impl="port", tol=1e-6, dofs=D):
    if impl == "gpu":  # one buffer of a device batch (tests/test_gpu_tbuf.py)
        from __graft_entry__ import load_package
        return load_package().TrajectoryBufferBatch(3, dofs, 2048, tol).view(1)
    b = orc.TrajectoryBuffer(dofs, tol, impl=impl)
    assert b.create_status == OK
    return b


IMPLS = ["port"] + (["ref"] if orc.have_ref() else [])


def test_create_checks_options():
    """trajectory_buffer_test.cc:99-103"""
    for impl in IMPLS:
        assert orc.TrajectoryBuffer(D, -1.0, impl=impl).create_status == FAILED_PRECONDITION
    from __graft_entry__ import load_package
    pkg = load_package()
    with pytest.raises(pkg.TpbError) as ei:  # checked before any device is touched
        pkg.TrajectoryBufferBatch(2, D, 64, timestep_tolerance=-1.0)
    assert ei.value.code == FAILED_PRECONDITION


@pytest.mark.parametrize("impl", IMPLS)
def test_insert_segment_works(impl):
    """trajectory_buffer_test.cc:105-195"""
    b = make(impl)
    assert (b.num_samples, b.start_ns, b.sequence_number) == (0, 0, 0)
    e = np.zeros((0, D))
    assert b.insert_segment(np.zeros(0), e, e, e) == OK
    assert (b.start_ns, b.end_ns, b.sequence_number, b.num_samples) == (0, 0, 1, 0)
    b.clear()
    tr = TC.test_trajectory(1.0, 10)
    assert b.insert_segment(*tr) == OK
    assert (b.start_ns, b.end_ns) == (ns(tr[0][0]), ns(tr[0][-1]))
    assert (b.sequence_number, b.num_samples) == (0, 10)
    for got, want in zip(b.rows(), tr):
        assert got.tobytes() == want.tobytes()
    tr2 = TC.test_trajectory(1.0 + 3 * DT, 5)
    assert b.insert_segment(*tr2) == OK
    assert (b.start_ns, b.end_ns) == (ns(1.0), ns(tr2[0][-1]))
    assert (b.sequence_number, b.num_samples) == (1, 8)
    t, p, v, a = b.rows()
    assert list(t) == [1.0 + k * DT for k in range(8)]
    seq = np.array([0, 1, 2, 0, 1, 2, 3, 4.0])
    assert np.all(p == seq[:, None]) and np.all(v == 10 * seq[:, None]) and np.all(a == 100 * seq[:, None])


@pytest.mark.parametrize("impl", IMPLS)
def test_append_sample_works(impl):
    """trajectory_buffer_test.cc:197-228"""
    b = make(impl)
    c = lambda x: np.full(D, x)
    assert b.append_sample(1.0, c(1), c(2), c(3)) == OK
    assert b.append_sample(1.0, c(1), c(2), c(3)) == INVALID
    assert b.append_sample(-1.0, c(1), c(2), c(3)) == INVALID
    assert b.append_sample(1.1, c(1.1), c(2.1), c(3.1)) == OK
    t, p, v, a = b.rows()
    assert b.num_samples == 2 and list(t) == [1.0, 1.1]
    assert np.all(p == np.array([[1.0], [1.1]])) and np.all(v == np.array([[2.0], [2.1]]))
    assert np.all(a == np.array([[3.0], [3.1]]))


@pytest.mark.parametrize("impl", IMPLS)
def test_insert_segment_uses_timestep_tolerance(impl):
    """trajectory_buffer_test.cc:230-288"""
    tol = 1e-5
    for sign in (-0.5, 0.5):
        b = make(impl, tol)
        assert b.insert_segment(*TC.test_trajectory(1.0, 10)) == OK
        t2, p2, v2, a2 = TC.test_trajectory(1.0 + 3 * DT, 5)
        t2[0] += sign * tol
        assert b.insert_segment(t2, p2, v2, a2) == OK
        want = [1.0 + k * DT for k in range(8)]
        want[3] = (1.0 + 3 * DT) + sign * tol
        assert list(b.rows()[0]) == want
    # appending at the final time step works (:313-322)
    b = make(impl)
    tr = TC.test_trajectory(1.0, 10)
    assert b.insert_segment(*tr) == OK
    t3 = orc.lib().orc_time_to_sec(b.end_ns) + DT * np.arange(10)
    assert b.insert_segment(t3, *tr[1:]) == OK


@pytest.mark.parametrize("impl", IMPLS)
def test_clear_and_discard_work(impl):
    """trajectory_buffer_test.cc:325-471"""
    b = make(impl)
    b.discard_before_ns(ns(10))
    b.discard_before_ns(ns(-10))
    assert (b.sequence_number, b.num_samples) == (0, 0)
    tr = TC.test_trajectory(1.0, 10)
    t = tr[0]
    assert b.insert_segment(*tr) == OK
    b.discard_before_ns(b.start_ns)
    assert (b.start_ns, b.end_ns, b.num_samples) == (ns(t[0]), ns(t[-1]), 10)
    b.discard_before_ns(ns(t[4]))
    assert b.start_ns == ns(t[4]) and b.rows()[0][0] == t[4]
    assert (b.end_ns, b.num_samples) == (ns(t[-1]), 6)
    b.discard_before_ns(ns(t[-1] + 1.0 / 1e9))
    assert (b.start_ns, b.end_ns, b.num_samples) == (0, 0, 0)
    eps, pos_eps = 1e-10, 1e-8

    def fresh():
        b.clear()
        assert b.insert_segment(*tr) == OK
        assert (b.sequence_number, b.start_ns) == (0, ns(t[0]))

    fresh()
    b.discard_before_sec(t[3])
    bt, bp, _, _ = b.rows()
    assert abs(bt[0] - t[3]) <= eps and np.all(np.abs(bp[0] - tr[1][3]) <= pos_eps)
    assert np.all(np.abs(np.diff(bt) - DT) <= eps)
    for tq in (np.nextafter(t[6], -1e99), np.nextafter(t[6], 1e99)):
        fresh()
        b.discard_before_sec(tq)
        bt, bp, _, _ = b.rows()
        assert abs(bt[0] - tq) <= eps and np.all(np.abs(bp[0] - tr[1][6]) <= pos_eps)
        assert np.all(np.abs(np.diff(bt) - DT) <= eps)
    fresh()
    tq = 0.5 * (t[5] + t[6])
    b.discard_before_sec(tq)
    bt, bp, _, _ = b.rows()
    assert abs(bt[0] - tq) <= eps
    assert not np.all(np.abs(bp[0] - tr[1][6]) <= eps) and not np.all(np.abs(bp[0] - tr[1][5]) <= eps)
    assert np.all(np.diff(bt) <= DT + 2 * eps)


@pytest.mark.parametrize("impl", IMPLS)
def test_get_positions_up_to_time_and_at_time(impl):
    """trajectory_buffer_test.cc:473-528"""
    b = make(impl)
    tr = TC.test_trajectory(1.0, 5)
    t = tr[0]
    assert b.insert_segment(*tr) == OK
    assert b.positions_up_to_time_ns(ns(10)) == 0 and b.positions_up_to_time_ns(ns(-1)) == 0
    assert b.positions_up_to_time_ns(ns(t[2])) == 2
    assert b.positions_up_to_time_ns(ns(np.nextafter(t[2], t[1]))) == 1
    assert b.positions_up_to_time_ns(ns(np.nextafter(t[2], t[3]))) == 2
    to_sec = orc.lib().orc_time_to_sec
    for i in range(5):
        st, pos = b.get_at_time(0, to_sec(ns(t[i])))
        assert st == OK and np.all(np.abs(pos - tr[1][i]) <= 1e-10)
        st, vel = b.get_at_time(1, to_sec(ns(t[i])))
        assert st == OK and np.all(np.abs(vel - tr[2][i]) <= 1e-8)
    assert b.get_at_time(0, t[-1] + 1.0)[0] == OUT_OF_RANGE
    assert make(impl).get_at_time(0, 0.0)[0] == FAILED_PRECONDITION


def check_fd_consistency(t, p, v, tol):
    fd = np.diff(p, axis=0) / np.diff(t)[:, None]
    assert np.max(np.abs(fd - 0.5 * (v[1:] + v[:-1]))) <= tol


@pytest.mark.parametrize("impl", IMPLS)
def test_stop_at_index_and_before_time(impl):
    """trajectory_buffer_test.cc:530-716"""
    dt, n = 1e-3, 1001
    tr = TC.quadratic_test_trajectory(n, dt)
    amax = np.full(D, 5.0)
    k = n // 2
    b = make(impl)
    assert b.insert_segment(*tr) == OK
    assert b.stop_at_index(k, amax, dt) == OK
    t, p, v, a = b.rows()
    assert b.num_samples >= k + 1 and np.all(np.abs(v[-1]) <= 1e-12)
    assert np.all(np.abs(a) <= amax + 1e-8)
    check_fd_consistency(t, p, v, 5e-2)
    # errors (:570-591)
    e = make(impl)
    assert e.stop_at_index(-1, np.full(D, 4.0), DT) == OUT_OF_RANGE
    assert e.stop_at_index(0, np.full(D, 4.0), DT) == OUT_OF_RANGE
    assert e.insert_segment(*TC.test_trajectory(1.0, 5)) == OK
    assert e.stop_at_index(5, np.full(D, 4.0), DT) == OUT_OF_RANGE
    assert e.stop_at_index(2, np.full(D, 4.0), -0.1) == INVALID
    # infeasible (:593-632)
    f = make(impl)
    assert f.insert_segment(*tr) == OK
    before = f.rows()
    assert f.stop_at_index(2, np.ones(D), dt) == NOT_FOUND
    assert f.stop_before_time(dt * 2, np.ones(D), dt) == NOT_FOUND
    assert f.stop_before_time(dt * 3.1415, np.ones(D), dt) == NOT_FOUND
    for x, y in zip(before, f.rows()):
        assert x.tobytes() == y.tobytes()
    # StopBeforeTime at and between samples (:634-685)
    g = make(impl)
    assert g.insert_segment(*tr) == OK
    assert g.stop_before_time(tr[0][k], amax, dt) == OK
    t, p, v, a = g.rows()
    assert g.num_samples >= k + 1 and np.all(np.abs(v[-1]) <= 1e-12)
    assert np.all(np.abs(a) <= amax + 1e-8)
    check_fd_consistency(t, p, v, 5e-2)
    g.clear()
    assert g.insert_segment(*tr) == OK
    assert g.stop_before_time(0.5 * (tr[0][k] + tr[0][k + 1]), amax, dt) == OK
    assert g.num_samples >= k + 1
    # cut-off beyond the final time step (:687-713)
    h = make(impl, dofs=3)
    cnt, step = 50, 0.01
    tt = np.array([i * step for i in range(cnt)])
    pp = np.array([[1.0 * step * i] * 3 for i in range(cnt)])
    assert h.insert_segment(tt, pp, np.ones((cnt, 3)), np.zeros((cnt, 3))) == OK
    assert h.stop_before_time(tt[-1] + 12.0, np.full(3, 2.0), step) == OK
    assert np.all(np.abs(h.rows()[2][-1]) <= 1e-12)
    # an empty buffer has nothing to stop (:369-371)
    assert make(impl).stop_before_time(1.0, amax, dt) == OK


@pytest.mark.skipif(not orc.have_ref(), reason="oracle/_ref not built")
def test_time_conversions_match_reference_object_code():
    rng = np.random.default_rng(3)
    for x in list(rng.uniform(-1e4, 1e4, 200)) + [0.0, 1.016, -0.3, 1e-9, 123456.789012345]:
        n = orc.ref().tpref_time_from_sec_ns(float(x))
        assert orc.lib().orc_time_from_sec(float(x)) == n
        assert orc.lib().orc_time_to_sec(n) == orc.ref().tpref_time_to_sec(n)


@pytest.mark.skipif(not orc.have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("seed", range(24))
def test_port_is_bit_identical_to_reference_object_code(seed):
    spec = TC.ops_spec(seed, D=[7, 3, 9, 1][seed % 4], num_ops=60,
                       tol=[1e-6, 1e-5, 1e-3][seed % 3])
    ops = TC.generate_ops(spec, make("port", spec["tol"], spec["D"]))
    a = TC.run_ops(make("port", spec["tol"], spec["D"]), ops)
    b = TC.run_ops(make("ref", spec["tol"], spec["D"]), ops)
    for k, (x, y) in enumerate(zip(a, b)):
        assert x == y, (seed, k, ops[k][0], {f: (x[f], y[f]) for f in ("result", "n", "seq")})
    kinds = {op[0] for op in ops}
    assert {"insert_segment", "discard_before_sec", "get_at_time"} <= kinds


def test_port_matches_committed_golden_observations():
    """tests/golden/trajectory_buffer.npz was written by
    tests/golden/make_trajectory_buffer.py from the reference object code."""
    g = np.load(GOLDEN)
    for c in range(int(g["num_cases"])):
        spec = dict(seed=int(g[f"seed_{c}"]), D=int(g[f"D_{c}"]), num_ops=int(g[f"num_ops_{c}"]),
                    dt=float(g[f"dt_{c}"]), tol=float(g[f"tol_{c}"]))
        ops = TC.generate_ops(spec, make("port", spec["tol"], spec["D"]))
        obs = TC.run_ops(make("port", spec["tol"], spec["D"]), ops)
        assert [o["n"] for o in obs] == list(g[f"n_{c}"])
        assert [o["seq"] for o in obs] == list(g[f"seq_{c}"])
        assert [o["start"] for o in obs] == list(g[f"start_{c}"])
        assert [-1 if o["result"] is None else o["result"] if isinstance(o["result"], int)
                else o["result"][0] for o in obs] == list(g[f"status_{c}"])
        t, p, v, a = (np.frombuffer(obs[-1][k]) for k in "tpva")
        assert t.tobytes() == g[f"t_{c}"].tobytes() and p.tobytes() == g[f"p_{c}"].tobytes()
        assert v.tobytes() == g[f"v_{c}"].tobytes() and a.tobytes() == g[f"a_{c}"].tobytes()
        assert TC.digest(obs) == str(g[f"digest_{c}"])
