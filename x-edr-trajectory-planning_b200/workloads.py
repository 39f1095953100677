"""Seeded synthetic inputs for the BASELINE.json configs.

The same generators feed the GPU engine, the CPU oracle and the timed CPU
baseline, so that all three see identical bits.  numpy only.
"""
from __future__ import annotations

import numpy as np


def joint_paths(num_paths, num_waypoints=10, num_dofs=7, seed=1000):
    """cfg2 / cfg4: random-walk joint waypoints with per-path limits.

    q_0 ~ U(-pi, pi)^D, q_{w+1} = clip(q_w + U(-1, 1)^D, +-pi) (keeps segment
    lengths O(1)), v_max ~ U(0.5, 2)^D rad/s, a_max ~ U(1, 4)^D rad/s^2.
    Path b's numbers depend only on (seed, b): generation is row-wise from one
    stream, so a shard [lo, hi) equals rows lo..hi of the full batch.
    Returns (waypoints [B][W][D], vmax [B][D], amax [B][D]).
    """
    rng = np.random.default_rng(seed)
    B, W, D = num_paths, num_waypoints, num_dofs
    per = rng.random((B, W * D + 2 * D))
    q0 = (per[:, :D] * 2.0 - 1.0) * np.pi
    steps = (per[:, D:W * D] * 2.0 - 1.0).reshape(B, W - 1, D)
    wp = np.empty((B, W, D))
    wp[:, 0] = q0
    for w in range(1, W):
        wp[:, w] = np.clip(wp[:, w - 1] + steps[:, w - 1], -np.pi, np.pi)
    vmax = 0.5 + 1.5 * per[:, W * D:W * D + D]
    amax = 1.0 + 3.0 * per[:, W * D + D:]
    return np.ascontiguousarray(wp), np.ascontiguousarray(vmax), np.ascontiguousarray(amax)


def single_path_cfg1(seed=1):
    """cfg1: one 7-dof path, 10 waypoints ~ U(-pi, pi)^7, v_max = 1, a_max = 2
    (the reference test's limits, path_timing_trajectory_test.cc:120-123)."""
    rng = np.random.default_rng(seed)
    wp = rng.uniform(-np.pi, np.pi, (1, 10, 7))
    return wp, np.full((1, 7), 1.0), np.full((1, 7), 2.0)


def clamped_uniform_knots(num_points, degree, scale=1.0):
    """Clamped uniform knot vector by running sum, like
    BSplineBase::MakeUniformKnotVector (splines/bspline_base.cc:356-381)."""
    nk = num_points + degree + 1
    k = np.zeros(nk)
    spacing = 1.0 / (nk - 2.0 * (degree + 1.0) + 1.0)
    for i in range(degree + 1, nk - degree - 1):
        k[i] = k[i - 1] + spacing
    k[nk - degree - 1:] = 1.0
    return k * scale


def dense_splines(num_splines, degree, dim, num_points=32, knot_scale=8.0, seed=7):
    """cfg3: control points ~ U(-pi, pi), clamped uniform knots on [0, knot_scale]."""
    rng = np.random.default_rng(seed + 100 * degree + dim)
    ctrl = rng.uniform(-np.pi, np.pi, (num_splines, num_points, dim))
    knots = np.tile(clamped_uniform_knots(num_points, degree, knot_scale), (num_splines, 1))
    return np.ascontiguousarray(knots), np.ascontiguousarray(ctrl)


# ---- reference test scenarios (time_optimal_path_timing_test.cc) -----------
def circle_constraints(N, s0=0.0, s1=np.pi, R=2.0, vmax=1.2, amax=1.0):
    """MakeCircleExampleConstraints, time_optimal_path_timing_test.cc:125-158."""
    ds = (s1 - s0) / (N - 1)
    c = np.zeros((N, 4, 4))
    for i in range(N):
        s = i * ds + s0
        c[i, 0] = [-R * np.sin(s), -R * np.cos(s), -amax, amax]
        c[i, 1] = [R * np.cos(s), -R * np.sin(s), -amax, amax]
        c[i, 2] = [0.0, (R * np.sin(s)) ** 2, 0.0, vmax ** 2]
        c[i, 3] = [0.0, (R * np.cos(s)) ** 2, 0.0, vmax ** 2]
    return c


def sine_constraints(N, s0, s1, R=2.0, vmax=1.2, amax=1.0):
    """MakeSineExampleConstraints, time_optimal_path_timing_test.cc:49-74."""
    ds = (s1 - s0) / (N - 1)
    c = np.zeros((N, 2, 4))
    for i in range(N):
        s = i * ds + s0
        c[i, 0] = [-R * np.sin(s), -R * np.cos(s), -amax, amax]
        c[i, 1] = [0.0, (R * np.sin(s)) ** 2, 0.0, vmax ** 2]
    return c


def line_constraints(N, slope=2.0, vmax=1.0, amax=1.0):
    """MakeLineExampleConstraints, time_optimal_path_timing_test.cc:216-247."""
    c = np.zeros((N, 4, 4))
    c[:, 0] = [1.0, 0.0, -amax, amax]
    c[:, 1] = [slope, 0.0, -amax, amax]
    c[:, 2] = [0.0, slope ** 2, 0.0, vmax ** 2]
    c[:, 3] = [0.0, 1.0, 0.0, vmax ** 2]
    return c


def scalar_straight_constraints(N, vmax=0.5, amax=1.0):
    """MakeScalarStraightExampleConstraints, :299-322."""
    c = np.zeros((N, 2, 4))
    c[:, 0] = [1.0, 0.0, -amax, amax]
    c[:, 1] = [0.0, 1.0, 0.0, vmax ** 2]
    return c


def scalar_curved_constraints(N, s0=-3.0, s1=1.0, m=(1.0, 1.0, 2.0), vmax=1.0, amax=0.2):
    """MakeScalarCurvedExampleConstraints, :362-389 (upper(1) = vmax, sic)."""
    ds = (s1 - s0) / (N - 1)
    c = np.zeros((N, 2, 4))
    for i in range(N):
        s = i * ds + s0
        d1 = 3.0 * m[0] * s * s + 2.0 * m[1] * s + m[2]
        c[i, 0] = [d1, 6.0 * m[0] * s + 2.0 * m[1], -amax, amax]
        c[i, 1] = [0.0, d1 ** 2, 0.0, vmax]
    return c


def random_constraints(rng, N, C):
    """Smooth random rows: A, B follow low-order trig polynomials of s so that
    the limit curve has interior critical points; symmetric accel-like bounds
    plus positive velocity-like rows."""
    s = np.linspace(0.0, 1.0, N)
    c = np.zeros((N, C, 4))
    half = max(C // 2, 1)
    for r in range(C):
        a = rng.uniform(-2, 2, 3)
        b = rng.uniform(-3, 3, 3)
        ph = rng.uniform(0, 2 * np.pi, 2)
        A = a[0] + a[1] * np.sin(2 * np.pi * s + ph[0]) + a[2] * np.cos(4 * np.pi * s)
        Bc = b[0] + b[1] * np.cos(2 * np.pi * s + ph[1]) + b[2] * np.sin(6 * np.pi * s)
        if r < half:
            lim = rng.uniform(0.5, 3.0)
            c[:, r, 0], c[:, r, 1], c[:, r, 2], c[:, r, 3] = A, Bc, -lim, lim
        else:
            c[:, r, 0], c[:, r, 1], c[:, r, 2], c[:, r, 3] = 0.0, A * A, 0.0, rng.uniform(0.5, 4.0)
    return c


def _exp_so3(v):
    """Rotation vector -> quaternion (x, y, z, w)."""
    th = np.linalg.norm(v, axis=-1, keepdims=True)
    k = np.where(th < 1e-8, 0.5 - th * th / 48.0, np.sin(0.5 * th) / np.where(th == 0, 1.0, th))
    return np.concatenate([k * v, np.cos(0.5 * th)], axis=-1)


def cart_paths(num_paths, num_waypoints=5, seed=5000):
    """cfg5: Cartesian paths for the reference test's fake 7-dof kinematics
    (path_timing_trajectory_test.cc:548-587).  Waypoint translations
    ~ U(-0.5, 0.5)^3 m, rotations ExpSO3(U(-1, 1)^3); joint waypoints are the
    fake IK of the poses (q[0:3] = translation, q[3:6] = rotation vector) plus
    q[6] ~ U(-1, 1); v_max = 1 rad/s, a_max = 2 rad/s^2 (the reference test's
    limits).  Returns (pose_wp [B][W][7] rows (tx,ty,tz,qx,qy,qz,qw),
    joint_wp [B][W][7], vmax [B][7], amax [B][7])."""
    rng = np.random.default_rng(seed)
    B, W = num_paths, num_waypoints
    per = rng.random((B, W * 7))
    per = per.reshape(B, W, 7)
    trans = per[..., 0:3] - 0.5
    rotvec = per[..., 3:6] * 2.0 - 1.0
    q6 = per[..., 6:7] * 2.0 - 1.0
    pose = np.concatenate([trans, _exp_so3(rotvec)], axis=-1)
    joint = np.concatenate([trans, rotvec, q6], axis=-1)
    vmax = np.full((B, 7), 1.0)
    amax = np.full((B, 7), 2.0)
    return (np.ascontiguousarray(pose), np.ascontiguousarray(joint), vmax, amax)
