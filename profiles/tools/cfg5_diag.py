import os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
from __graft_entry__ import load_package
from oracle import pyoracle as po
pkg = load_package()
B, W, N = 4096, 5, 1000
pw, jw, vm, am = pkg.workloads.cart_paths(B, W, seed=5000)
eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=7, max_waypoints=W, max_time_samples_total=B * 64, cartesian=True)
eng.set_cart_options(0.05, 0.5, 0.25)
eng.set_cart_waypoints(pw, jw); eng.set_joint_limits(vm, am)
eng.cart_sample_targets(); eng.cart_solve_ik(1); eng.cart_setup_constraints(1); eng.optimize()
A = pkg.ARR
st, dur, tim, q = eng.get(A.STATUS), eng.get(A.DURATION), eng.get(A.TIME), eng.get(A.Q)
ref = po.plan_cart_batch_ref(pw, jw, vm, am, N, dt=0.02, threads=os.cpu_count(), want_time=True)
ok = (st == 0) & (ref["status"] == 0)
rel = np.where(ok, np.abs(dur - ref["duration"]) / ref["duration"], 0)
idx = np.argsort(-rel)[:12]
print("status mismatch", int(((st == 0) != (ref["status"] == 0)).sum()), "ok", int(ok.sum()))
print("rel > 1e-6:", int((rel > 1e-6).sum()), "> 1e-9:", int((rel > 1e-9).sum()), "> 1e-12:", int((rel > 1e-12).sum()))
for b in idx: print(b, rel[b], dur[b], ref["duration"][b])
b = idx[0]
# host-callback mode for the worst path: positions from the oracle's fake IK of the device targets
o = po.plan_cart_fake(pw[b], jw[b], vm[b], am[b], N, dt=0.0)
print("port duration", o["duration"], "max |q - q_port|", np.max(np.abs(q[b] - o["q"])))
dt_ = np.abs(tim[b] - ref["time"][b]); print("first sample where time differs > 1e-9:", np.argmax(dt_ > 1e-9), dt_.max())
