#!/usr/bin/env python
"""Regenerate the golden fixtures in this directory.  Run it in the build
container, where /root/reference exists; the fixtures are committed because the
GPU box has no /root/reference.

  bspline_mathematica.json   the Mathematica table of the reference's
                             BSplineGoldenTest (splines/bspline_test.cc:97-727):
                             degree 3, knots {0,0,0,0,.5,1,1,1,1}, control points
                             (1,1),(2,3),(3,-1),(4,1),(5,0), 101 samples, values +
                             3 derivatives; tolerance 5e-14 (:70-75).
  quat_exp_mathematica.json  QuatExp inputs/outputs of QuatExpAndLogTest
                             (splines/bsplineq_test.cc:99-198), order w,x,y,z.
  lp_regression.json         the 5 regression LPs of FindMaxSd2Regression
                             (time_optimal_path_timing_test.cc:744-1088) with the
                             outputs of the reference's own FindMaxSd2Simplex and
                             FindMaxSd2BruteForce (oracle/_ref).
  ref_solver_scenarios.json  outputs of the reference's own
                             TimeOptimalPathProfile (oracle/_ref, compiled verbatim)
                             on its test scenarios (:440-701): durations, last
                             extremal index, and full sd/sdd/time arrays.
  ref_joint_paths.json       whole-pipeline outputs (oracle stages A/B/D + reference
                             stage C) for a few seeded joint paths: durations and
                             checksums.
"""
import json
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference/trajectory_planning"

NUM = r"[-+]?(?:\d+\.\d*|\.\d+|\d+)(?:[eE][-+]?\d+)?"


def parse_bspline_tables():
    src = open(os.path.join(REF, "splines/bspline_test.cc")).read()
    out = {"degree": 3, "knots": [0, 0, 0, 0, 0.5, 1, 1, 1, 1],
           "control_points": [[1, 1], [2, 3], [3, -1], [4, 1], [5, 0]],
           "num_samples": 101, "tolerance": 5e-14,
           "source": "splines/bspline_test.cc:97-727 (Mathematica BSplineFunction)"}
    for name in ("X", "Y", "Xp", "Yp", "Xpp", "Ypp", "Xppp", "Yppp"):
        m = re.search(r"kGoldenData%s\[GetNumSamples\(\)\]\s*=\s*\{(.*?)\};" % name, src, re.S)
        vals = [float(x) for x in re.findall(NUM, m.group(1))]
        assert len(vals) == 101, (name, len(vals))
        out[name] = vals
    return out


def parse_quat_exp():
    src = open(os.path.join(REF, "splines/bsplineq_test.cc")).read()
    body = src[src.index("TEST(QuatExpAndLogTest, ResultsMatchGoldenValues)"):]
    body = body[:body.index("\n}\n")]
    quats = re.findall(r"Quaterniond\(\s*(%s),\s*(%s),\s*(%s),\s*(%s)\)" % (NUM, NUM, NUM, NUM),
                       body)
    q = [[float(x) for x in t] for t in quats]
    assert len(q) == 30, len(q)
    return {"order": "w,x,y,z", "inputs": q[:15], "outputs": q[15:],
            "source": "splines/bsplineq_test.cc:99-198 (Mathematica)"}


def parse_lp_regression():
    src = open(os.path.join(REF, "time_optimal_path_timing_test.cc")).read()
    a = src.index("std::vector<LPInfo> lpinfo = {")
    b = src.index("for (const auto &lp : lpinfo)")
    toks = re.findall(NUM, src[a + len("std::vector<LPInfo> lpinfo = {"):b])
    lps, i = [], 0
    while i < len(toks):
        sz = int(toks[i])
        vals = [float(x) for x in toks[i + 1:i + 1 + 4 * sz]]
        assert len(vals) == 4 * sz
        lps.append({"sz": sz, "a": vals[:sz], "b": vals[sz:2 * sz],
                    "lower": vals[2 * sz:3 * sz], "upper": vals[3 * sz:]})
        i += 1 + 4 * sz
    assert len(lps) == 5, len(lps)
    return lps


def main():
    from __graft_entry__ import load_oracle, load_package
    po = load_oracle()
    po.build()
    pkg = load_package()
    W = pkg.workloads
    assert po.have_ref(), "oracle/_ref missing: run make -C oracle"

    json.dump(parse_bspline_tables(), open(os.path.join(HERE, "bspline_mathematica.json"), "w"))
    json.dump(parse_quat_exp(), open(os.path.join(HERE, "quat_exp_mathematica.json"), "w"))

    lps = parse_lp_regression()
    for lp in lps:
        rows = np.stack([lp["a"], lp["b"], lp["lower"], lp["upper"]], 1)
        lp["ref_simplex"] = [float(x) for x in po.find_max_sd2(rows, "ref_simplex")]
        lp["ref_brute"] = [float(x) for x in po.find_max_sd2(rows, "ref_brute")]
    json.dump({"source": "time_optimal_path_timing_test.cc:744-1088", "lps": lps},
              open(os.path.join(HERE, "lp_regression.json"), "w"))

    scen = []

    def add(name, cons, s0, s1, sd0=0.0):
        r = po.solve(cons, s0, s1, sd0, impl="ref")
        assert r["ok"], name
        scen.append({"name": name, "N": int(cons.shape[0]), "s_start": s0, "s_end": s1,
                     "sd_start": sd0, "duration": float(r["time"][-1] - r["time"][0]),
                     "last_extremal_index": int(r["last_extremal_index"]),
                     "satisfies": int(r["extra"]),
                     "time": [float(x) for x in r["time"]],
                     "sd": [float(x) for x in r["sd"]],
                     "sdd": [float(x) for x in r["sdd"]]})

    for N in (50, 51):
        for sd0 in (0.0, 0.1):
            add(f"circle_N{N}_sd{sd0}", W.circle_constraints(N), 0.0, float(np.pi), sd0)
    for off, nm in ((0.0, "0"), (np.pi / 2, "pi2"), (np.pi / 4, "pi4")):
        for N in (30, 31, 100, 111):
            add(f"sine_off{nm}_N{N}", W.sine_constraints(N, off, np.pi + off), float(off),
                float(np.pi + off))
    add("line_N30", W.line_constraints(30), 0.0, float(np.pi))
    add("scalar_straight_N30", W.scalar_straight_constraints(30), 0.0, 1.0)
    add("scalar_curved_N100", W.scalar_curved_constraints(100), -3.0, 1.0)
    json.dump({"source": "reference TimeOptimalPathProfile compiled verbatim (oracle/_ref) on "
                         "time_optimal_path_timing_test.cc:440-701 scenarios",
               "scenarios": scen}, open(os.path.join(HERE, "ref_solver_scenarios.json"), "w"))

    wp, vmax, amax = W.joint_paths(16, 10, 7, seed=1000)
    paths = []
    for b in range(16):
        r = po.plan_joint(wp[b], vmax[b], amax[b], 2048, dt=0.004, M_cap=16384, impl="ref")
        assert r["ok"]
        paths.append({"index": b, "duration": float(r["duration"]), "M": int(r["M"]),
                      "knot_end": float(r["knot_end"]),
                      "last_extremal_index": int(r["last_extremal_index"]),
                      "sum_time": float(np.sum(r["time"])), "sum_sd": float(np.sum(r["sd"])),
                      "sum_q": float(np.sum(r["q"])), "sum_r_pos": float(np.sum(r["r_pos"]))})
    json.dump({"source": "workloads.joint_paths(16, 10, 7, seed=1000), N=2048, dt=4 ms; stage C "
                         "= reference .cc verbatim, stages A/B/D = oracle restatement",
               "paths": paths}, open(os.path.join(HERE, "ref_joint_paths.json"), "w"))
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
