"""GPU parity: joint-space paths through the C ABI vs the CPU oracle.

Tolerances from BASELINE.json: spline positions/derivatives 1e-12 absolute,
durations 1e-6 relative, resampled joint positions 1e-9 rad.  Because the CUDA
build uses the oracle's operation order without FMA contraction, the
intermediate arrays are in fact expected to be bit-identical; that stronger
property is asserted where it holds by construction.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL_SPLINE = 1e-12
TOL_DURATION_REL = 1e-6
TOL_RESAMPLED = 1e-9


def _run_engine(pkg, wp, vmax, amax, N, dt=0.004, num_waypoints=None, delta=None,
                options=None):
    B, W, D = wp.shape
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=B * 16384)
    if options:
        eng.set_path_options(*options)
    if delta is not None:
        eng.set_joint_waypoints(wp, num_waypoints)
        eng.set_sampling(delta)
    else:
        eng.set_joint_waypoints(wp, num_waypoints)
    eng.set_joint_limits(vmax, amax)
    eng.sample_paths()
    eng.optimize()
    total = eng.resample(0.0, dt)
    A = pkg.ARR
    res = {k: eng.get(getattr(A, k)) for k in
           ("CTRL", "KNOTS", "NUM_POINTS", "Q", "DQ", "DDQ", "TIME", "S", "SD", "SDD", "SD2",
            "B_SD2_MAX", "B_SDD_MAX", "B_SDD_MIN", "B_SD2_ZERO", "B_TYPE", "STATUS",
            "DURATION", "LAST_EXTREMAL", "KNOT_END", "DELTA", "R_OFFSET", "R_TIME", "R_POS",
            "R_VEL", "R_ACC", "R_S", "R_SD", "R_SDD")}
    res["total"] = total
    eng.close()
    return res


def _check_against_oracle(po, res, wp, vmax, amax, N, dt=0.004, nw=None, rounding=0.2,
                          safety=0.8, delta=None):
    B = wp.shape[0]
    exact = 0
    for b in range(B):
        W = wp.shape[1] if nw is None else int(nw[b])
        o = po.plan_joint(wp[b, :W], vmax[b], amax[b], N, rounding=rounding, safety=safety,
                          delta=-1.0 if delta is None else float(delta[b]), dt=dt,
                          M_cap=16384)
        n = res["NUM_POINTS"][b]
        ctrl, knots = po.joint_fit(wp[b, :W], rounding)
        assert n == len(ctrl)
        assert np.max(np.abs(res["CTRL"][b, :n] - ctrl)) <= TOL_SPLINE
        assert np.max(np.abs(res["KNOTS"][b, :n + 3] - knots)) <= TOL_SPLINE
        for k, ok in (("Q", "q"), ("DQ", "dq"), ("DDQ", "ddq")):
            err = np.max(np.abs(res[k][b] - o[ok]))
            assert err <= TOL_SPLINE, (b, k, err)
        assert (res["STATUS"][b] == 0) == o["ok"], (b, res["STATUS"][b], o["status"])
        if not o["ok"]:
            assert np.isnan(res["DURATION"][b])
            continue
        assert abs(res["DURATION"][b] - o["duration"]) <= TOL_DURATION_REL * o["duration"]
        assert res["LAST_EXTREMAL"][b] == o["last_extremal_index"]
        for k, ok in (("TIME", "time"), ("SD", "sd"), ("SDD", "sdd"), ("S", "s")):
            np.testing.assert_allclose(res[k][b], o[ok], rtol=1e-9, atol=1e-9, err_msg=f"{b} {k}")
        lo, hi = res["R_OFFSET"][b], res["R_OFFSET"][b + 1]
        assert hi - lo == o["M"], (b, hi - lo, o["M"])
        assert np.max(np.abs(res["R_POS"][lo:hi] - o["r_pos"])) <= TOL_RESAMPLED
        assert np.max(np.abs(res["R_VEL"][lo:hi] - o["r_vel"])) <= 1e-8
        assert np.max(np.abs(res["R_ACC"][lo:hi] - o["r_acc"])) <= 1e-7
        assert np.max(np.abs(res["R_TIME"][lo:hi] - o["r_t"])) == 0.0
        bit = all(np.array_equal(res[k][b], o[ok]) for k, ok in
                  (("Q", "q"), ("DQ", "dq"), ("DDQ", "ddq"), ("TIME", "time"), ("SD", "sd"),
                   ("SDD", "sdd"))) and np.array_equal(res["R_POS"][lo:hi], o["r_pos"])
        exact += bit
    return exact


def test_cfg2_shape_small_batch(pkg, po):
    """cfg2 shape (D=7, W=10, N=2048), 48 paths."""
    wp, vmax, amax = pkg.workloads.joint_paths(48, 10, 7, seed=1000)
    res = _run_engine(pkg, wp, vmax, amax, 2048)
    exact = _check_against_oracle(po, res, wp, vmax, amax, 2048)
    assert exact == 48, f"only {exact}/48 paths bit-identical to the oracle"


def test_cfg1_single_path(pkg, po):
    wp, vmax, amax = pkg.workloads.single_path_cfg1()
    res = _run_engine(pkg, wp, vmax, amax, 1000)
    assert _check_against_oracle(po, res, wp, vmax, amax, 1000) == 1


@pytest.mark.parametrize("D,W,N", [(6, 4, 257), (3, 3, 1000), (7, 2, 64), (1, 5, 100),
                                   (16, 3, 128)])
def test_other_shapes(pkg, po, D, W, N):
    wp, vmax, amax = pkg.workloads.joint_paths(16, W, D, seed=5 + D)
    res = _run_engine(pkg, wp, vmax, amax, N)
    assert _check_against_oracle(po, res, wp, vmax, amax, N) == 16


def test_ragged_waypoint_counts_and_single_waypoint(pkg, po):
    """Ragged batches; W == 1 is PolyLineToBspline3Waypoints' special case
    (spline_utils.cc:55-60): a point path, zero duration."""
    B, W, D, N = 12, 8, 7, 300
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=77)
    nw = np.array([8, 2, 3, 1, 5, 8, 4, 2, 7, 6, 1, 3], np.int32)
    res = _run_engine(pkg, wp, vmax, amax, N, num_waypoints=nw)
    _check_against_oracle(po, res, wp, vmax, amax, N, nw=nw)


def test_short_segments_and_duplicate_waypoints(pkg, po):
    """Segments shorter than 4*rounding (offset = norm/4) and zero-length
    segments (norm <= kMinNorm), spline_utils.cc:28-43."""
    B, W, D, N = 6, 5, 7, 400
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=9)
    wp[0, 2] = wp[0, 1]                     # duplicate waypoint
    wp[1, 1:] = wp[1, :1] + 0.01 * wp[1, 1:]  # tiny path
    wp[2, 3] = wp[2, 2] + 1e-8              # below kMinNorm
    res = _run_engine(pkg, wp, vmax, amax, N)
    _check_against_oracle(po, res, wp, vmax, amax, N)


def test_explicit_delta_with_tail_beyond_path_end(pkg, po):
    """delta larger than knots.back()/(N-1): samples beyond the path end take the
    tail branch (timeable_path_joint_spline.cc:309-313), zero rows -> kMaxSd2."""
    B, W, D, N = 8, 6, 7, 500
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=21)
    ctrl_len = np.array([po.joint_fit(wp[b], 0.2)[1][-1] for b in range(B)])
    delta = ctrl_len / (N - 1) * np.linspace(1.0, 1.3, B)
    res = _run_engine(pkg, wp, vmax, amax, N, delta=delta)
    _check_against_oracle(po, res, wp, vmax, amax, N, delta=delta)


def test_options_safety_rounding(pkg, po):
    B, W, D, N = 8, 6, 7, 300
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=33)
    res = _run_engine(pkg, wp, vmax, amax, N, options=(0.5, 0.05))
    _check_against_oracle(po, res, wp, vmax, amax, N, rounding=0.05, safety=0.5)


def test_plan_joint_host_matches_staged_calls(pkg, po):
    B, W, D, N = 16, 10, 7, 512
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=4)
    res = _run_engine(pkg, wp, vmax, amax, N)
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=B * 16384)
    cap = int(res["total"])
    out = {"duration": np.zeros(B), "status": np.zeros(B, np.int32),
           "r_offset": np.zeros(B + 1, np.int64), "r_pos": np.zeros((cap, D)),
           "r_vel": np.zeros((cap, D)), "r_acc": np.zeros((cap, D)), "r_s": np.zeros(cap),
           "r_sd": np.zeros(cap), "r_sdd": np.zeros(cap)}
    total = eng.plan_joint_host(wp, vmax, amax, 0.0, 0.004, out=out)
    eng.close()
    assert total == res["total"]
    np.testing.assert_array_equal(out["duration"], res["DURATION"])
    np.testing.assert_array_equal(out["r_offset"], res["R_OFFSET"])
    np.testing.assert_array_equal(out["r_pos"], res["R_POS"])
    np.testing.assert_array_equal(out["r_acc"], res["R_ACC"])
    np.testing.assert_array_equal(out["r_sdd"], res["R_SDD"])


def _limit_curve_oracle(po, wp_b, vmax_b, amax_b, N, rounding=0.2, safety=0.8):
    ctrl, knots = po.joint_fit(wp_b, rounding)
    delta = knots[-1] / (N - 1)
    q, dq, ddq = po.joint_sample(knots, ctrl, N, 0.0, delta)
    cons = po.joint_constraints(dq, ddq, vmax_b, amax_b, safety)
    return po.solve(cons, 0.0, delta * (N - 1), max_loops=max(100, 10 * N)), cons


@pytest.mark.parametrize("D", [2, 4, 7, 8])
def test_tied_rows_duplicate_and_mirrored_dofs(pkg, po, D):
    """Dofs that are copies / mirror images of dof 0 with the same limits make
    two or more constraint rows active within kTiny in FindMaxSd2Simplex
    (time_optimal_path_timing.cc:1187-1260): the tie handling (active-set order,
    strict '<' slope choice, pairwise IsOptimal) must match the reference's
    list semantics.  The limit curve itself is compared bit for bit."""
    B, W, N = 6, 5, 300
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=123 + D)
    for b in range(B):
        for d in range(1, D):
            kind = (b + d) % 3
            if kind == 0:
                wp[b, :, d] = wp[b, :, 0]
            elif kind == 1:
                wp[b, :, d] = -wp[b, :, 0]
            if kind != 2:
                vmax[b, d] = vmax[b, 0]
                amax[b, d] = amax[b, 0]
    res = _run_engine(pkg, wp, vmax, amax, N)
    assert _check_against_oracle(po, res, wp, vmax, amax, N) == B
    for b in range(B):
        o, _ = _limit_curve_oracle(po, wp[b], vmax[b], amax[b], N)
        for k, ok in (("B_SD2_MAX", "b_sd2_max"), ("B_SDD_MAX", "b_sdd_max"),
                      ("B_SDD_MIN", "b_sdd_min"), ("B_SD2_ZERO", "b_sd2_zero"),
                      ("B_TYPE", "b_type")):
            assert np.array_equal(res[k][b], o[ok]), (b, k)


def test_limit_curve_bitwise_cfg2_shape(pkg, po):
    """The register-only joint LP (tpb_joint_limit.cuh) against the oracle's
    FindMaxSd2Simplex / FindSddMax / FindSddMin on cfg2-shaped paths: limit
    curve after CalculateBoundary's passes, boundary types and sd^2, bit for bit."""
    B, W, D, N = 8, 10, 7, 1024
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=2024)
    res = _run_engine(pkg, wp, vmax, amax, N)
    for b in range(B):
        o, _ = _limit_curve_oracle(po, wp[b], vmax[b], amax[b], N)
        for k, ok in (("B_SD2_MAX", "b_sd2_max"), ("B_SDD_MAX", "b_sdd_max"),
                      ("B_SDD_MIN", "b_sdd_min"), ("B_SD2_ZERO", "b_sd2_zero"),
                      ("B_TYPE", "b_type"), ("SD2", "sd2")):
            assert np.array_equal(res[k][b], o[ok], equal_nan=True), (b, k)


def test_plan_joint_device_matches_staged_calls(pkg):
    """tpb_plan_joint_device (one call for the resident batch; one chunk, or several
    chunks on overlapping streams) returns exactly what sample_paths + optimize +
    resample return."""
    B, W, D, N = 700, 5, 7, 256
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=12)
    res = _run_engine(pkg, wp, vmax, amax, N)
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=B * 16384)
    eng.set_joint_waypoints(wp)
    eng.set_joint_limits(vmax, amax)
    for chunks in (1, 3, 1, 5):
        total = eng.plan_joint_device(0.0, 0.004, chunks)
        assert total == res["total"]
        for k in ("DURATION", "STATUS", "R_OFFSET", "R_POS", "R_VEL", "R_ACC", "R_S", "TIME", "SD"):
            np.testing.assert_array_equal(eng.get(getattr(pkg.ARR, k)), res[k], err_msg=f"{chunks} {k}")
    eng.close()


def test_plan_joint_host_chunked_pipeline_matches_staged_calls(pkg):
    """Batches of >= 1024 paths go through tpb_plan_joint_host in chunks whose
    device-to-host copies overlap the next chunk's kernels; every output must be
    identical to the staged whole-batch calls."""
    B, W, D, N = 1500, 4, 7, 128
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=11)
    nw = np.full(B, W, np.int32)
    nw[::7] = 2
    nw[5::11] = 1
    res = _run_engine(pkg, wp, vmax, amax, N, num_waypoints=nw)
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=B * 16384)
    cap = int(res["total"])
    out = {"duration": np.zeros(B), "status": np.zeros(B, np.int32),
           "r_offset": np.zeros(B + 1, np.int64), "r_pos": np.zeros((cap, D)),
           "r_vel": np.zeros((cap, D)), "r_acc": np.zeros((cap, D)), "r_s": np.zeros(cap),
           "r_sd": np.zeros(cap), "r_sdd": np.zeros(cap)}
    for _ in range(2):  # a second call reuses the pipeline's stream / events
        total = eng.plan_joint_host(wp, vmax, amax, 0.0, 0.004, num_waypoints=nw, out=out)
    assert total == res["total"]
    np.testing.assert_array_equal(out["duration"], res["DURATION"])
    np.testing.assert_array_equal(out["status"], res["STATUS"])
    np.testing.assert_array_equal(out["r_offset"], res["R_OFFSET"])
    for k, a in (("r_pos", "R_POS"), ("r_vel", "R_VEL"), ("r_acc", "R_ACC"), ("r_s", "R_S"),
                 ("r_sd", "R_SD"), ("r_sdd", "R_SDD")):
        np.testing.assert_array_equal(out[k], res[a], err_msg=k)
    # staged getters still work on the engine after the pipelined call
    np.testing.assert_array_equal(eng.get(pkg.ARR.R_POS), res["R_POS"])
    np.testing.assert_array_equal(eng.get(pkg.ARR.TIME), res["TIME"])
    # too small a host capacity is reported, not overrun
    small = dict(out)
    for k in ("r_pos", "r_vel", "r_acc"):
        small[k] = np.zeros((cap // 2, D))
    for k in ("r_s", "r_sd", "r_sdd"):
        small[k] = np.zeros(cap // 2)
    with pytest.raises(pkg.TpbError) as ei:
        eng.plan_joint_host(wp, vmax, amax, 0.0, 0.004, num_waypoints=nw, out=small)
    assert ei.value.code == 11
    eng.close()


@pytest.mark.parametrize("B", [96, 1500])
def test_plan_joint_host_selective_outputs(pkg, B):
    """tpb_plan_io copies only the arrays a caller passes (ABI 4): positions only,
    the profile only (no resampling at all), and time stamps + profile together, for
    the unchunked and the chunked pipeline; values equal the staged calls'."""
    W, D, N = 4, 7, 128
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=12)
    res = _run_engine(pkg, wp, vmax, amax, N)
    cap = int(res["total"])
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=B * 16384)
    # positions only
    out = {"r_offset": np.zeros(B + 1, np.int64), "r_pos": np.full((cap, D), np.nan)}
    assert eng.plan_joint_host(wp, vmax, amax, 0.0, 0.004, out=out) == cap
    np.testing.assert_array_equal(out["r_offset"], res["R_OFFSET"])
    np.testing.assert_array_equal(out["r_pos"], res["R_POS"])
    # the profile only: time step 0 = no resampling
    prof = {k: np.full((B, N), np.nan) for k in ("profile_time", "profile_s", "profile_sd",
                                                   "profile_sdd")}
    prof["duration"] = np.zeros(B)
    assert eng.plan_joint_host(wp, vmax, amax, 0.0, 0.0, out=prof) == 0
    np.testing.assert_array_equal(prof["duration"], res["DURATION"])
    for k, a in (("profile_time", "TIME"), ("profile_s", "S"), ("profile_sd", "SD"),
                 ("profile_sdd", "SDD")):
        np.testing.assert_array_equal(prof[k], res[a], err_msg=k)
    # time stamps + one profile array with the rows
    both = {"r_time": np.full(cap, np.nan), "r_sd": np.full(cap, np.nan),
            "profile_time": np.full((B, N), np.nan)}
    assert eng.plan_joint_host(wp, vmax, amax, 0.0, 0.004, out=both) == cap
    np.testing.assert_array_equal(both["r_time"], res["R_TIME"])
    np.testing.assert_array_equal(both["r_sd"], res["R_SD"])
    np.testing.assert_array_equal(both["profile_time"], res["TIME"])
    eng.close()


def test_resample_skipping_samples_closer_than_time_step_bitwise(pkg, po):
    """kSkipSamplesCloserThanTimeStep (path_timing_trajectory.cc:785-836): every packed
    array bit-identical to the oracle, for time steps below, around and above the
    spacing of the path samples, and a start time inside the trajectory."""
    B, W, D, N = 12, 6, 7, 300
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=99)
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=B * (N + 1))
    eng.set_joint_waypoints(wp)
    eng.set_joint_limits(vmax, amax)
    eng.sample_paths()
    eng.optimize()
    for dt, start in ((0.004, 0.0), (0.02, 0.0), (0.1, 0.0), (0.02, 0.37)):
        total = eng.resample(start, dt, skip=True)
        off = eng.get(pkg.ARR.R_OFFSET)
        assert total == off[-1]
        got = {k: eng.get(getattr(pkg.ARR, k)) for k in
               ("R_TIME", "R_S", "R_SD", "R_SDD", "R_POS", "R_VEL", "R_ACC")}
        for b in range(B):
            o = po.plan_joint(wp[b], vmax[b], amax[b], N, start_sec=0.0, dt=dt, M_cap=N + 1,
                              skip=True) if start == 0.0 else None
            if o is None:
                # a later start: resample the oracle's path samples directly
                full = po.plan_joint(wp[b], vmax[b], amax[b], N, dt=0.0)
                o = po.resample_skip(full, amax[b], start, dt)
            lo, hi = off[b], off[b + 1]
            assert hi - lo == o["M"]
            assert np.array_equal(got["R_TIME"][lo:hi], o["r_t"])
            assert np.array_equal(got["R_S"][lo:hi], o["r_s"])
            assert np.array_equal(got["R_SD"][lo:hi], o["r_sd"])
            assert np.array_equal(got["R_SDD"][lo:hi], o["r_sdd"])
            assert np.array_equal(got["R_POS"].reshape(-1, D)[lo:hi], o["r_pos"])
            assert np.array_equal(got["R_VEL"].reshape(-1, D)[lo:hi], o["r_vel"])
            assert np.array_equal(got["R_ACC"].reshape(-1, D)[lo:hi], o["r_acc"])
            # the reference's own rule, restated on the result: consecutive kept
            # samples are at least 0.95 dt apart
            assert np.all(np.diff(o["r_t"]) >= 0.95 * dt - 1e-15)
    eng.close()


def test_cfg2_full_size_properties(pkg, po):
    """BASELINE cfg2 at full size (4,096 paths x 2,048 samples, D = 7, W = 10, the bench's
    batch): what holds whatever the size -- every path timed, limits kept, trajectories
    start and end on their waypoints at rest, time and path parameter monotone, the
    packed layout consistent -- plus independence of the batch: paths taken from all over
    the batch and run as a small batch of their own give the same bits, and the oracle's."""
    B, W, D, N, dt = 4096, 10, 7, 2048, 0.004
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=1000)
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=B * 6144)
    eng.set_joint_waypoints(wp)
    eng.set_joint_limits(vmax, amax)
    eng.sample_paths()
    eng.optimize()
    total = eng.resample(0.0, dt)
    A = pkg.ARR
    res = {k: eng.get(getattr(A, k)) for k in
           ("STATUS", "DURATION", "TIME", "S", "SD", "KNOT_END", "R_OFFSET", "R_TIME", "R_POS",
            "R_VEL", "R_ACC", "R_S")}
    eng.close()
    off = res["R_OFFSET"]
    assert (res["STATUS"] == 0).all()
    dur = res["DURATION"]
    assert np.isfinite(dur).all() and (dur > 0).all()
    assert off[0] == 0 and off[-1] == total and (np.diff(off) >= 2).all()
    # ResampleEquidistantlyInTime (:755-757): rows 0, dt, 2 dt, ... and the end of the path
    rows = np.diff(off)
    assert (np.abs(rows - (np.ceil(dur / dt) + 1)) <= 1).all()
    assert np.array_equal(res["TIME"][:, 0], np.zeros(B))
    assert np.array_equal(res["TIME"][:, -1], dur)
    assert (np.diff(res["TIME"], axis=1) >= 0).all()
    assert (np.diff(res["S"], axis=1) > 0).all()
    assert (res["SD"] >= 0).all() and (res["SD"][:, 0] == 0).all() and (res["SD"][:, -1] == 0).all()
    first, last = off[:-1], off[1:] - 1
    assert np.array_equal(res["R_TIME"][first], np.zeros(B))
    assert np.max(np.abs(res["R_POS"][first] - wp[:, 0])) <= TOL_RESAMPLED
    assert np.max(np.abs(res["R_POS"][last] - wp[:, -1])) <= TOL_RESAMPLED
    assert np.array_equal(res["R_VEL"][last], np.zeros((B, D)))
    assert np.array_equal(res["R_ACC"][last], np.zeros((B, D)))
    # per-row limits: ConstraintSetup scales them by constraint_safety = 0.8, accelerations
    # are clamped to the unscaled limit (path_timing_trajectory.cc:458-472)
    owner = np.repeat(np.arange(B), rows)
    assert (np.abs(res["R_VEL"]) <= 0.8 * vmax[owner] * (1 + 1e-6) + 1e-9).all()
    assert (np.abs(res["R_ACC"]) <= amax[owner]).all()
    within = np.ones(total - 1, bool)   # differences between rows of the same path ...
    within[last[:-1]] = False
    assert (np.diff(res["R_S"])[within] >= 0).all()
    within[last - 1] = False            # ... and not onto the final row (the path's end time)
    assert np.max(np.abs(np.diff(res["R_TIME"])[within] - dt)) <= 1e-12 * float(np.max(dur))
    # batch independence and the oracle on a sample of the batch
    pick = np.array([0, 1, 777, 2048, 3000, 4095])
    sub = _run_engine(pkg, wp[pick], vmax[pick], amax[pick], N, dt=dt)
    for i, b in enumerate(pick):
        lo, hi = off[b], off[b + 1]
        slo, shi = sub["R_OFFSET"][i], sub["R_OFFSET"][i + 1]
        assert hi - lo == shi - slo
        assert res["DURATION"][b] == sub["DURATION"][i]
        for k in ("R_TIME", "R_POS", "R_VEL", "R_ACC", "R_S"):
            assert np.array_equal(res[k][lo:hi], sub[k][slo:shi]), (b, k)
        assert np.array_equal(res["TIME"][b], sub["TIME"][i])
    exact = _check_against_oracle(po, sub, wp[pick], vmax[pick], amax[pick], N, dt=dt)
    assert exact == len(pick)
    # ALL 4,096 paths against the reference's own object code (TimeableJointSplinePath +
    # PathTimingTrajectory compiled verbatim into oracle/_ref, all host threads): status,
    # duration, row count, the profile's time stamps and every resampled position, bit for bit
    if po.have_ref():
        ref = po.plan_joint_batch_ref(wp, vmax, amax, N, dt=dt, threads=os.cpu_count() or 1,
                                      r_offset=off, want_time=True)
        assert np.array_equal(ref["status"] == 0, res["STATUS"] == 0)
        assert ref["duration"].tobytes() == res["DURATION"].tobytes()
        assert np.array_equal(ref["rows"], rows)
        assert ref["time"].tobytes() == res["TIME"].tobytes()
        assert ref["r_pos"].tobytes() == res["R_POS"][:total].tobytes()


def test_resample_with_more_path_samples_than_shared_memory_holds(pkg, po):
    """N = 32,768 path samples: the time stamps of a path (256 KB) do not fit shared memory,
    so the resampling searches them in global memory (k_resample_fill<false, .>); every
    array still equals the oracle bit for bit."""
    B, W, D, N = 2, 5, 7, 32768
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=41)
    res = _run_engine(pkg, wp, vmax, amax, N)
    _check_against_oracle(po, res, wp, vmax, amax, N)
    # and with the skipping time sampling
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=B * (N + 1))
    eng.set_joint_waypoints(wp)
    eng.set_joint_limits(vmax, amax)
    eng.sample_paths()
    eng.optimize()
    total = eng.resample(0.0, 0.004, skip=True)
    off = eng.get(pkg.ARR.R_OFFSET)
    pos, t = eng.get(pkg.ARR.R_POS), eng.get(pkg.ARR.R_TIME)
    for b in range(B):
        o = po.plan_joint(wp[b], vmax[b], amax[b], N, dt=0.004, M_cap=N + 1, skip=True)
        assert o["ok"] and off[b + 1] - off[b] == o["M"]
        assert pos[off[b]:off[b + 1]].tobytes() == o["r_pos"].tobytes()
        assert t[off[b]:off[b + 1]].tobytes() == o["r_t"].tobytes()
    assert total == off[-1]
    eng.close()
