"""GPU parity: Cartesian timeable paths (BASELINE config 5) through the C ABI vs
the CPU oracle (oracle/tp_oracle_cart.c).

Two modes, like the reference's API allows:
 * host callbacks (user PathIKFunc / JacobianFunc): the path-IK positions and the
   Jacobians are produced on the host (here: by the oracle's restatement of the
   reference test's fake kinematics) and uploaded; forward differences,
   constraint rows, limit curve, timing and resampling run on the GPU and must be
   BIT-IDENTICAL to the oracle given the same positions;
 * built-in device kinematics (TPB_KIN_FAKE7): everything on the GPU.  The
   quaternion spline and the kinematics use sin/cos/atan2/log/exp, where CUDA's
   libm and glibc differ by an ulp, so this mode is compared with the tolerances
   BASELINE.json states (1e-12 spline values, 1e-6 relative durations, 1e-9 rad
   resampled positions) -- the forward differences amplify an ulp by 1/delta^2.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

N = 1000
DT = 0.004


def _engine(pkg, B, W, cap=40000):
    return pkg.Engine(max_paths=B, num_samples=N, num_dofs=7, max_waypoints=W,
                      max_time_samples_total=B * cap, cartesian=True)


def _oracle(po, pw, jw, vm, am, delta=-1.0, **kw):
    return po.plan_cart_fake(pw, jw, vm, am, N, delta=delta, dt=DT, M_cap=40000, **kw)


def test_fit_and_targets_match_oracle(pkg, po):
    B, W = 12, 5
    pw, jw, vm, am = pkg.workloads.cart_paths(B, W, seed=5000)
    eng = _engine(pkg, B, W)
    eng.set_cart_options(0.05, 0.5, 0.25)
    eng.set_cart_waypoints(pw, jw)
    eng.cart_sample_targets()
    A = pkg.ARR
    pose_ctrl, knots, npts = eng.get(A.POSE_CTRL), eng.get(A.KNOTS), eng.get(A.NUM_POINTS)
    ctrl, pt, jt = eng.get(A.CTRL), eng.get(A.POSE_TARGETS), eng.get(A.JOINT_TARGETS)
    delta = eng.get(A.DELTA)
    eng.close()
    for b in range(B):
        pc, jc, kn = po.cart_fit(pw[b], jw[b], 0.05, 0.2)
        n = npts[b]
        assert n == len(pc) == 3 * W - 2
        assert np.max(np.abs(pose_ctrl[b, :n] - pc)) <= 1e-12
        assert np.max(np.abs(ctrl[b, :n] - jc)) <= 1e-12
        assert np.max(np.abs(knots[b, :n + 3] - kn)) <= 1e-12 * kn[-1]
        d = kn[-1] / (N - 1)
        assert abs(delta[b] - d) <= 1e-15
        opt, ojt = po.cart_sample_targets(kn, pc, jc, N, d)
        assert np.max(np.abs(pt[b] - opt)) <= 1e-12
        assert np.max(np.abs(jt[b] - ojt)) <= 1e-12
        # unit quaternions with non-negative real part (bsplineq.cc:99-109)
        assert np.max(np.abs(np.linalg.norm(pt[b][:, 3:], axis=1) - 1.0)) <= 1e-12
        assert (pt[b][:, 6] >= 0).all()
        # IK index -1 and 0 are both parameter 0 (.cc:478-482)
        assert np.array_equal(pt[b][0], pt[b][1])


@pytest.mark.parametrize("explicit_delta", [False, True])
def test_host_callback_mode_is_bit_identical(pkg, po, explicit_delta):
    """Positions and Jacobians uploaded (what a user PathIKFunc / JacobianFunc
    would return): derivatives, rows, limit curve, timing, resampling on the GPU."""
    B, W = 8, 5
    pw, jw, vm, am = pkg.workloads.cart_paths(B, W, seed=5100)
    eng = _engine(pkg, B, W)
    eng.set_cart_options(0.05, 0.5, 0.25)
    eng.set_cart_waypoints(pw, jw)
    if explicit_delta:
        # CartesianPathOptions' default delta_parameter 0.005: the first planning
        # window s in [0, 4.995] of a longer path
        eng.set_sampling(np.full(B, 0.005))
    eng.set_joint_limits(vm, am)
    ora = [_oracle(po, pw[b], jw[b], vm[b], am[b], delta=0.005 if explicit_delta else -1.0)
           for b in range(B)]
    q = np.stack([o["q"] for o in ora])
    jac = np.stack([np.stack([po.fake_jacobian(q[b, i]) for i in range(N)]) for b in range(B)])
    if not explicit_delta:
        # the engine's automatic delta must be the oracle's, bit for bit, for the
        # rest to be comparable at that level; it depends on the device-side pose
        # control polygon (sin/cos), so upload the oracle's value when it differs
        d_gpu = eng.get(pkg.ARR.DELTA)
        d_orc = np.array([o["delta"] for o in ora])
        assert np.max(np.abs(d_gpu - d_orc)) <= 1e-15
        if not np.array_equal(d_gpu, d_orc):
            eng.set_sampling(d_orc)
    eng.cart_set_ik_positions(q)
    eng.cart_setup_constraints(model=0, jacobians=jac)
    eng.optimize()
    total = eng.resample(0.0, DT)
    A = pkg.ARR
    res = {k: eng.get(getattr(A, k)) for k in
           ("DQ", "DDQ", "CONS", "TIME", "S", "SD", "SDD", "STATUS", "DURATION",
            "LAST_EXTREMAL", "R_OFFSET", "R_POS", "R_VEL", "R_ACC", "R_TIME")}
    eng.close()
    assert total == res["R_OFFSET"][-1]
    for b, o in enumerate(ora):
        assert np.array_equal(res["DQ"][b], o["dq"])
        assert np.array_equal(res["DDQ"][b], o["ddq"])
        assert np.array_equal(res["CONS"][b], o["cons"])
        assert (res["STATUS"][b] == 0) == o["ok"], (b, res["STATUS"][b], o["status"])
        if not o["ok"]:
            continue
        for k, ok in (("TIME", "time"), ("S", "s"), ("SD", "sd"), ("SDD", "sdd")):
            assert np.array_equal(res[k][b], o[ok]), (b, k)
        assert res["DURATION"][b] == o["duration"]
        assert res["LAST_EXTREMAL"][b] == o["last_extremal_index"]
        lo, hi = res["R_OFFSET"][b], res["R_OFFSET"][b + 1]
        assert hi - lo == o["M"]
        assert np.array_equal(res["R_POS"][lo:hi], o["r_pos"])
        assert np.array_equal(res["R_VEL"][lo:hi], o["r_vel"])
        assert np.array_equal(res["R_ACC"][lo:hi], o["r_acc"])


def test_builtin_kinematics_pipeline(pkg, po):
    """Everything on the device with TPB_KIN_FAKE7, BASELINE tolerances."""
    B, W = 16, 5
    pw, jw, vm, am = pkg.workloads.cart_paths(B, W, seed=5200)
    eng = _engine(pkg, B, W)
    eng.set_cart_options(0.05, 0.5, 0.25)
    eng.set_cart_waypoints(pw, jw)
    eng.set_joint_limits(vm, am)
    eng.cart_sample_targets()
    eng.cart_solve_ik(1)
    eng.cart_setup_constraints(1)
    eng.optimize()
    eng.resample(0.0, DT)
    A = pkg.ARR
    res = {k: eng.get(getattr(A, k)) for k in
           ("Q", "DQ", "STATUS", "DURATION", "R_OFFSET", "R_POS", "TIME")}
    eng.close()
    n_ok = 0
    for b in range(B):
        o = _oracle(po, pw[b], jw[b], vm[b], am[b])
        assert np.max(np.abs(res["Q"][b] - o["q"])) <= 1e-12
        # forward differences amplify by 1/delta (delta ~ 0.03-0.05 here)
        assert np.max(np.abs(res["DQ"][b] - o["dq"])) <= 1e-10
        assert (res["STATUS"][b] == 0) == o["ok"]
        if not o["ok"]:
            continue
        n_ok += 1
        assert abs(res["DURATION"][b] - o["duration"]) <= 1e-6 * o["duration"]
        lo, hi = res["R_OFFSET"][b], res["R_OFFSET"][b + 1]
        assert abs((hi - lo) - o["M"]) <= 1
        m = min(hi - lo, o["M"]) - 1  # the last row is the end position in both
        assert np.max(np.abs(res["R_POS"][lo:lo + m] - o["r_pos"][:m])) <= 1e-9
        # path ends at the last waypoint (fake IK of the last pose)
        assert np.max(np.abs(res["R_POS"][hi - 1] - jw[b, -1])) <= 1e-9
    assert n_ok >= B // 2


def test_cfg5_full_size_against_reference_object_code(pkg, po):
    """BASELINE config 5 at its full size -- 4,096 Cartesian paths (5 pose waypoints, N = 1000),
    the batch bench.py times -- everything on the device, against the reference's own
    TimeableCartesianSplinePath + PathTimingTrajectory object code (oracle/_ref, all host
    threads, the same fake robot as its callbacks): status of every path, total duration
    within 1e-6 relative (BASELINE.json), the profile's time stamps within 1e-6 of the
    duration, row counts of a 20 ms resampling within one row; size-independent
    properties on all rows (uniform grid, rest to rest, end on the last joint waypoint)."""
    import os
    if not po.have_ref():
        pytest.skip("oracle/_ref not built")
    B, W = 4096, 5
    dt = 0.02  # ~2,100 rows per path: 9 M rows in all
    pw, jw, vm, am = pkg.workloads.cart_paths(B, W, seed=5000)
    eng = _engine(pkg, B, W, cap=4096)
    eng.set_cart_options(0.05, 0.5, 0.25)
    eng.set_cart_waypoints(pw, jw)
    eng.set_joint_limits(vm, am)
    eng.cart_sample_targets()
    eng.cart_solve_ik(1)
    eng.cart_setup_constraints(1)
    eng.optimize()
    eng.resample(0.0, dt)
    A = pkg.ARR
    st, dur, tim = eng.get(A.STATUS), eng.get(A.DURATION), eng.get(A.TIME)
    off, r_t, r_pos, r_vel = eng.get(A.R_OFFSET), eng.get(A.R_TIME), eng.get(A.R_POS), eng.get(A.R_VEL)
    eng.close()
    ref = po.plan_cart_batch_ref(pw, jw, vm, am, N, dt=dt, threads=os.cpu_count(), want_time=True)
    assert ((st == 0) == (ref["status"] == 0)).all()
    ok = st == 0
    assert ok.sum() >= 0.99 * B
    rel = np.zeros(B)
    rel[ok] = np.abs(dur[ok] - ref["duration"][ok]) / ref["duration"][ok]
    # With the whole path in one window (delta = knots.back() / (N - 1)) SamplePath's test
    # "parameter < knots.back() - delta" (timeable_path_cartesian_spline.cc:488) compares two
    # numbers that are equal in exact arithmetic at IK index N - 2; knots.back() is a sum of
    # control-polygon lengths that went through sin / cos / atan2, so the device's and glibc's
    # last bit can differ and the second-to-last sample is then the spline value on one side
    # and the end pose on the other.  Such a path (1 of these 4,096) agrees up to its last
    # few samples; every other path is within BASELINE.json's 1e-6.
    edge = rel > 1e-6
    assert edge.sum() <= B // 1000, int(edge.sum())
    good = ok & ~edge
    assert np.max(np.abs(tim[good] - ref["time"][good]) / ref["duration"][good, None]) <= 1e-6
    for b in np.flatnonzero(edge):
        d = np.abs(tim[b] - ref["time"][b]) / ref["duration"][b]
        assert d[:N - 8].max() <= 1e-6 and rel[b] <= 2e-3, (b, rel[b])
    rows = np.diff(off)
    assert np.max(np.abs(rows[good] - ref["rows"][good])) <= 1
    first, last = off[:-1][ok], off[1:][ok] - 1
    assert (r_t[first] == 0.0).all()
    inner = np.ones(len(r_t), bool)
    inner[off[1:] - 1] = False  # the last row of a path is its end time, off the grid
    steps = np.diff(r_t)[inner[:-1] & inner[1:]]
    assert np.max(np.abs(steps - dt)) <= 1e-9
    assert (r_vel[first] == 0).all() and (r_vel[last] == 0).all()
    assert np.max(np.abs(r_pos[last] - jw[ok, -1])) <= 1e-9
    assert np.max(np.abs(r_vel)) <= 0.8 * vm.max() + 1e-9


def _limit_curve_run(pkg, pw, jw, vm, am, jac_from=None, variants=()):
    B, W = pw.shape[0], pw.shape[1]
    eng = _engine(pkg, B, W)
    for v in variants:  # cross-check kernels (tpb_set_kernel_variant)
        eng.set_kernel_variant(v)
    eng.set_cart_options(0.05, 0.5, 0.25)
    eng.set_cart_waypoints(pw, jw)
    eng.set_joint_limits(vm, am)
    eng.cart_sample_targets()
    eng.cart_solve_ik(1)
    n0 = eng.kernel_launches
    if jac_from is None:
        eng.cart_setup_constraints(1)
    else:
        q = eng.get(pkg.ARR.Q)
        jac = np.stack([np.stack([jac_from(q[b, i]) for i in range(0, N)]) for b in range(B)])
        eng.cart_setup_constraints(model=0, jacobians=jac)
    launches = eng.kernel_launches - n0
    eng.optimize()
    eng.resample(0.0, DT)
    A = pkg.ARR
    res = {k: eng.get(getattr(A, k)) for k in
           ("DQ", "DDQ", "CONS", "B_SD2_MAX", "B_SDD_MAX", "B_SDD_MIN", "B_SD2_ZERO", "B_TYPE",
            "TIME", "SD", "SDD", "SD2", "STATUS", "DURATION", "LAST_EXTREMAL", "R_OFFSET", "R_POS",
            "R_VEL", "R_ACC", "R_S", "R_SD", "R_SDD")}
    eng.close()
    return res, launches


@pytest.mark.parametrize("uploaded_jacobians", [False, True])
def test_fused_rows_and_limit_curve_equal_the_two_kernel_path(pkg, po, monkeypatch, uploaded_jacobians):
    """k_cart_constraints_limit (rows + LP on the registers that built them) against
    k_cart_constraints followed by k_generic_limit reading the stored rows: every
    array bit for bit, on paths that include failing ones."""
    B, W = (6, 5) if uploaded_jacobians else (48, 5)
    pw, jw, vm, am = pkg.workloads.cart_paths(B, W, seed=5300)
    jf = po.fake_jacobian if uploaded_jacobians else None
    fused, n_fused = _limit_curve_run(pkg, pw, jw, vm, am, jf)
    plain, n_plain = _limit_curve_run(pkg, pw, jw, vm, am, jf, variants=("cart_unfused",))
    assert (n_fused, n_plain) == (1, 2)
    assert (fused["STATUS"] == 0).sum() >= B // 2
    for k in fused:
        assert fused[k].tobytes() == plain[k].tobytes(), k


@pytest.mark.parametrize("uploaded_jacobians", [False, True])
def test_joint_structured_sweeps_equal_the_generic_solver(pkg, po, monkeypatch, uploaded_jacobians):
    """Cartesian rows are the joint rows plus two speed rows: the sweeps run on
    Joint8Policy<7, true> / k_optimize_pair<7, true> (rows from q', q'' and the two stored
    coefficients) and must give the bits of GenericPolicy reading all 16 stored rows."""
    B, W = (6, 5) if uploaded_jacobians else (64, 5)
    pw, jw, vm, am = pkg.workloads.cart_paths(B, W, seed=5400)
    if not uploaded_jacobians:
        # some unusual paths: tight / loose limits, a repeated waypoint
        vm[3] *= 0.5
        am[5] *= 100.0
        am[6] *= 0.3
        pw[7, 2] = pw[7, 1]
        jw[7, 2] = jw[7, 1]
    jf = po.fake_jacobian if uploaded_jacobians else None
    fast, _ = _limit_curve_run(pkg, pw, jw, vm, am, jf)
    generic, _ = _limit_curve_run(pkg, pw, jw, vm, am, jf, variants=("cart_generic",))
    assert (fast["STATUS"] == 0).sum() >= B // 2
    for k in fast:
        assert fast[k].tobytes() == generic[k].tobytes(), k


def _six_dof_run(pkg, pw, jw6, vm, am, q, jac, variants=()):
    B, W = pw.shape[0], pw.shape[1]
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=6, max_waypoints=W,
                     max_time_samples_total=B * 60000, cartesian=True)
    for v in variants:
        eng.set_kernel_variant(v)
    eng.set_cart_options(0.05, 0.5, 0.25)
    eng.set_cart_waypoints(pw, jw6)
    eng.set_joint_limits(vm, am)
    eng.cart_set_ik_positions(q)
    eng.cart_setup_constraints(model=0, jacobians=jac)
    eng.optimize()
    eng.resample(0.0, DT)
    A = pkg.ARR
    res = {k: eng.get(getattr(A, k)) for k in
           ("DQ", "DDQ", "CONS", "B_SD2_MAX", "B_SDD_MAX", "B_SDD_MIN", "B_SD2_ZERO", "B_TYPE",
            "TIME", "SD", "SDD", "SD2", "STATUS", "DURATION", "LAST_EXTREMAL", "R_OFFSET", "R_POS",
            "R_VEL", "R_ACC", "R_S", "R_SD", "R_SDD")}
    eng.close()
    return res


def test_six_dof_arm_fast_kernels_equal_the_generic_ones(pkg, monkeypatch):
    """A 6-dof arm with user kinematics (positions and Jacobians uploaded, here smooth
    random ones): the fused rows/limit kernel and the joint-structured sweeps for kD = 6
    against k_cart_constraints + k_generic_limit + GenericPolicy, bit for bit."""
    B, W = 24, 4
    rng = np.random.default_rng(5500)
    pw, jw, _, _ = pkg.workloads.cart_paths(B, W, seed=5500)
    jw6 = np.ascontiguousarray(jw[..., :6])
    vm = rng.uniform(0.5, 2.0, (B, 6))
    am = rng.uniform(1.0, 4.0, (B, 6))
    # smooth joint motion along the path and a smooth, well-conditioned Jacobian field
    u = np.linspace(0.0, 1.0, N)[None, :, None]
    q = (rng.uniform(-1, 1, (B, 1, 6)) + u * rng.uniform(-2, 2, (B, 1, 6))
         + 0.3 * np.sin(2 * np.pi * u * rng.uniform(0.5, 2.0, (B, 1, 6))))
    jac = (rng.normal(0, 0.3, (B, 1, 6, 6)) + np.eye(6)[None, None]
           + 0.2 * np.sin(3.0 * u[..., None] + rng.uniform(0, 6, (B, 1, 6, 6))))
    q, jac = np.ascontiguousarray(q), np.ascontiguousarray(jac)
    fast = _six_dof_run(pkg, pw, jw6, vm, am, q, jac)
    generic = _six_dof_run(pkg, pw, jw6, vm, am, q, jac,
                           variants=("cart_unfused", "cart_generic"))
    assert (fast["STATUS"] == 0).sum() >= B // 2
    for k in fast:
        assert fast[k].tobytes() == generic[k].tobytes(), k


def test_cart_argument_errors(pkg):
    eng = _engine(pkg, 2, 3, cap=100)
    with pytest.raises(pkg.TpbError) as ei:
        eng.set_cart_options(0.0, 0.5, 0.25)   # SetTranslationRounding(0) -> kInvalidArgument
    assert ei.value.code == 3
    with pytest.raises(pkg.TpbError) as ei:
        eng.set_cart_options(0.05, -1.0, 0.25)  # SetMaxCartesianVelocity -> kInvalidArgument
    assert ei.value.code == 3
    with pytest.raises(pkg.TpbError) as ei:
        eng.cart_sample_targets()               # no path set
    assert ei.value.code == 9
    eng.close()
    joint_only = pkg.Engine(max_paths=2, num_samples=N, num_dofs=7, max_waypoints=3)
    with pytest.raises(pkg.TpbError) as ei:
        joint_only.set_cart_options(0.05, 0.5, 0.25)
    assert ei.value.code == 9
    joint_only.close()
