"""Shared TrajectoryBuffer test material: the reference test's trajectories
(trajectory_buffer_test.cc:60-97), seeded random operation sequences, and a
runner that applies a sequence to any implementation with the pyoracle
TrajectoryBuffer interface (port, reference object code, or one buffer of the
device batch) and records everything observable after every operation."""
import numpy as np

K_NUM_JOINTS = 9
K_SAMPLE_TIME = 8e-3


def quadratic_test_trajectory(count, dt=K_SAMPLE_TIME, D=K_NUM_JOINTS):
    """x = (t - T)^2 / 2 (trajectory_buffer_test.cc:60-77)"""
    final_time = (count - 1) * dt
    t = np.array([dt * i for i in range(count)])
    trel = t - final_time
    pos = np.repeat((0.5 * trel * trel)[:, None], D, axis=1)
    vel = np.repeat(trel[:, None], D, axis=1)
    acc = np.ones((count, D))
    return t, pos, vel, acc


def test_trajectory(offset, count, dt=K_SAMPLE_TIME, D=K_NUM_JOINTS):
    """trajectory_buffer_test.cc:80-97"""
    i = np.arange(count, dtype=np.float64)
    t = np.array([k * dt + offset for k in range(count)])
    one = np.ones((1, D))
    return t, i[:, None] * one, 10 * i[:, None] * one, 100 * i[:, None] * one


test_trajectory.__test__ = False  # not a pytest test


def smooth_segment(rng, t0, n, D, dt):
    """A smooth random segment of n rows starting at t0 (slightly irregular grid)."""
    t = t0 + np.concatenate([[0.0], np.cumsum(dt * rng.uniform(0.9, 1.1, n - 1))]) if n else np.zeros(0)
    vel = np.cumsum(rng.normal(0, 0.02, (n, D)), axis=0) + rng.normal(0, 0.6, D)
    acc = rng.normal(0, 0.5, (n, D))
    pos = np.cumsum(vel, axis=0) * dt + rng.normal(0, 1, D)
    return t, pos, vel, acc


def ops_spec(seed, D, num_ops=40, dt=4e-3, tol=1e-6):
    """Parameters of a seeded operation sequence (materialised by generate_ops: the
    operations' times depend on the buffer's content at that point)."""
    return dict(seed=seed, D=D, num_ops=num_ops, dt=dt, tol=tol)


def generate_ops(spec, shadow):
    """Materialise the operation list of `spec` by running it on `shadow` (any
    implementation; the port in practice).  Returns a list of (name, args)."""
    rng = np.random.default_rng(spec["seed"])
    D, dt, tol = spec["D"], spec["dt"], spec["tol"]
    ops = []

    def times():
        return shadow.rows()[0]

    for k in range(spec["num_ops"]):
        t = times()
        n = len(t)
        kind = rng.choice(["insert", "insert", "append", "discard", "discard_ns", "get", "upto",
                           "stop_index", "stop_time", "offset", "clear", "insert_tol"],
                          p=[0.2, 0.1, 0.08, 0.12, 0.05, 0.12, 0.06, 0.1, 0.06, 0.04, 0.02, 0.05])
        if n == 0 and kind not in ("insert", "append", "get", "discard", "stop_time", "upto"):
            kind = "insert"
        if kind in ("insert", "insert_tol"):
            m = int(rng.integers(0, 120))
            if n == 0:
                t0 = float(rng.uniform(-2, 5))
            else:
                where = rng.choice(["inside", "end", "before", "beyond", "at_sample"])
                if where == "inside":
                    t0 = float(rng.uniform(t[0], t[-1]))
                elif where == "end":
                    t0 = float(t[-1])
                elif where == "before":
                    t0 = float(t[0] - rng.uniform(0, 0.1))
                elif where == "beyond":
                    t0 = float(t[-1] + rng.uniform(0, 0.1))
                else:
                    t0 = float(t[int(rng.integers(0, n))])
                if kind == "insert_tol":
                    t0 = float(t[int(rng.integers(0, n))] + rng.choice([-0.5, 0.5, 0.999, 1.001, 2.0]) * tol)
            seg = smooth_segment(rng, t0, m, D, dt)
            op = ("insert_segment", seg)
        elif kind == "append":
            last = t[-1] if n else 0.0
            tt = float(last + rng.choice([dt, 0.0, -dt, 1e-9]))
            op = ("append_sample", (tt, rng.normal(0, 1, D), rng.normal(0, 1, D), rng.normal(0, 1, D)))
        elif kind in ("discard", "discard_ns"):
            if n == 0:
                tt = float(rng.uniform(-1, 1))
            else:
                i = int(rng.integers(0, n))
                tt = float(rng.choice([t[i], np.nextafter(t[i], -1e99), np.nextafter(t[i], 1e99),
                                       t[i] + 0.5 * tol, t[i] - 0.5 * tol, t[i] + 1.5 * tol,
                                       rng.uniform(t[0] - 0.01, t[-1] + 0.01)]))
                if rng.uniform() < 0.05:
                    tt = float(t[-1] + 1e-9)
            if kind == "discard":
                op = ("discard_before_sec", (tt,))
            else:
                op = ("discard_before_ns", (int(tt * 1e9),))
        elif kind == "get":
            if n == 0:
                tt = 0.0
            else:
                i = int(rng.integers(0, n))
                tt = float(rng.choice([t[i], t[-1], t[0], rng.uniform(t[0] - 0.01, t[-1] + 0.01),
                                       np.nextafter(t[i], 1e99)]))
            op = ("get_at_time", (int(rng.integers(0, 3)), tt))
        elif kind == "upto":
            tt = float(rng.uniform(t[0] - 0.01, t[-1] + 0.01)) if n else 0.5
            op = ("positions_up_to_time_ns", (int(tt * 1e9),))
        elif kind == "stop_index":
            idx = int(rng.integers(-1, n + 2)) if rng.uniform() < 0.3 else int(rng.integers(1, max(n, 2)))
            amax = rng.uniform(0.5, 6.0, D)
            if rng.uniform() < 0.05:
                amax[int(rng.integers(0, D))] = 0.0
            step = dt if rng.uniform() > 0.05 else -dt
            op = ("stop_at_index", (idx, amax, step))
        elif kind == "stop_time":
            tt = float(rng.uniform(t[0] - 0.01, t[-1] + 0.05)) if n else 1.0
            if n and rng.uniform() < 0.3:
                tt = float(t[int(rng.integers(0, n))])
            op = ("stop_before_time", (tt, rng.uniform(0.5, 6.0, D), dt))
        elif kind == "offset":
            op = ("add_offset", (float(rng.uniform(-1, 1)),))
        else:
            op = ("clear", ())
        if op[0] == "stop_at_index" and would_abort_in_reference(shadow, op[1][0]):
            op = ("get_at_time", (0, 0.0))
        if op[0] == "stop_before_time" and n:
            lower = int(np.searchsorted(t, op[1][0], side="left"))
            if would_abort_in_reference(shadow, min(lower + 1, n - 1)):
                op = ("get_at_time", (0, 0.0))
        apply_op(shadow, op)
        ops.append(op)
    return ops


def would_abort_in_reference(buf, index):
    """StopAtIndex CHECK-aborts (trajectory_buffer.cc:338) when the sample at `index`
    stands still but is not the last one, and reads out of bounds on an empty buffer
    with index > 0; such operations are left out of the comparisons with the reference."""
    t, _, v, _ = buf.rows()
    n = len(t)
    if n == 0:
        return index > 0
    if index <= 0 or index > n - 1:
        return False
    still = np.max(np.abs(v[index])) < 1e-8
    return bool(still and index < n - 1)


def apply_op(buf, op):
    name, args = op
    return getattr(buf, name)(*args)


def observe(buf, result):
    t, p, v, a = buf.rows()
    if isinstance(result, tuple):  # get_at_time -> (status, row); the row only counts when ok
        st, row = result
        res = (int(st), row.tobytes() if st == 0 else b"")
    else:
        res = None if result is None else int(result)
    return dict(result=res, n=buf.num_samples, seq=buf.sequence_number, start=buf.start_ns,
                end=buf.end_ns, t=t.tobytes(), p=p.tobytes(), v=v.tobytes(), a=a.tobytes())


def run_ops(buf, ops):
    """Apply `ops` to `buf`; returns the list of observations after every operation."""
    return [observe(buf, apply_op(buf, op)) for op in ops]


def digest(observations):
    """SHA-256 over everything observed after every operation."""
    import hashlib
    h = hashlib.sha256()
    for o in observations:
        r = o["result"]
        h.update(repr(r if not isinstance(r, tuple) else (r[0],)).encode())
        if isinstance(r, tuple):
            h.update(r[1])
        h.update(repr((o["n"], o["seq"], o["start"], o["end"])).encode())
        for k in "tpva":
            h.update(o[k])
    return h.hexdigest()
