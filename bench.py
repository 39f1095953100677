#!/usr/bin/env python
"""Benchmark of the batched path-timing hot path (BASELINE.json metric:
"7-DoF paths timed/sec at 1/2/4/8 B200; B-spline evals/sec vs HBM roofline").

    python bench.py --gpus N --steps K --warmup W            # our arm (CUDA, C ABI)
    python bench.py --impl reference --gpus N --steps K ...  # reference CPU arm

One step = one pass of the whole hot path (fit -> SamplePath+ConstraintSetup ->
limit curve -> Pfeiffer-Johanni phase-plane integration -> trajectory assembly
and uniform-in-time resampling) over one batch of BASELINE config 2:
4,096 synthetic 7-DoF joint-spline paths x 2,048 samples per GPU.  With N > 1
(torchrun, one rank per GPU) every rank times its own shard of independent
paths (weak scaling, no data-path collective; a final all_gather of the
durations happens outside the timed region).

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for every key.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_oracle, load_package  # noqa: E402

METRIC = "7-DoF paths timed/sec"
UNIT = "paths/s"
PATHS_PER_GPU = 4096
NUM_SAMPLES = 2048
NUM_WAYPOINTS = 10
NUM_DOFS = 7
TIME_STEP = 0.004  # planner time step of the reference tests (4 ms)
WORKLOAD = ("cfg2: 4096 synthetic 7-DoF joint-spline paths x 2048 samples per GPU, "
            "time-optimal timing + 4 ms uniform resampling")


#: the sources of the kernels whose ncu counters bench.py reads (profiles/tools/ops_from_raw.py
#: hashes the same list)
HOT_KERNEL_SOURCES = ("tpb_kernels.cu", "tpb_integrate.cuh", "tpb_integrate_pair.cuh",
                      "tpb_joint_limit.cuh", "tpb_row_limit.cuh", "tpb_solver.cuh", "tpb_device.cuh",
                      "tpb_bulk.cuh", "tpb_fit.cuh")


def _csrc_hash():
    """Hash of the CUDA sources; profiles/fp64_ops.json and traffic.json carry the hash of
    the build they were captured from ("_csrc_hash"), so stale counters are visible."""
    import hashlib
    h = hashlib.sha256()
    for name in HOT_KERNEL_SOURCES:
        with open(os.path.join(ROOT, "x-edr-trajectory-planning_b200", "csrc", name), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def _stale(path):
    try:
        with open(path) as f:
            return json.load(f).get("_csrc_hash") != _csrc_hash()
    except (OSError, ValueError):
        return True


def _traffic(kernel):
    """DRAM bytes per launch of `kernel` from the committed `ncu --set full` capture
    (profiles/traffic.json: dram__bytes_read.sum + dram__bytes_write.sum), or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(p) as f:
            return json.load(f).get(kernel, {}).get("dram_bytes_per_launch")
    except (OSError, ValueError):
        return None


def _fp64_report(pkg, device, kernel_ms):
    """Achieved FP64 FLOP/s of the compute-bound kernels: thread-level
    dadd + dmul + 2 dfma per launch from the committed ncu capture
    (profiles/fp64_ops.json; the workload is deterministic) over the kernel's live
    CUDA-event time, against the FP64 FMA rate measured on this device now."""
    p = os.path.join(ROOT, "profiles", "fp64_ops.json")
    try:
        with open(p) as f:
            ops = json.load(f)
    except (OSError, ValueError):
        return None
    peak = pkg.measure_fp64_fma_peak(device)
    out = {"peak_tflops": peak, "peak_source": "tpb_measure_fp64_fma_peak (8 FMA chains/thread)",
           "ops_source": "profiles/fp64_ops.json (ncu --set full, same workload)",
           "stale": _stale(p), "kernels": {}}
    for k, ms in kernel_ms.items():
        if k in ops and ms > 0:
            tf = ops[k]["flops_per_launch"] / (ms / 1e3) / 1e12
            out["kernels"][k] = {"flops_per_launch": ops[k]["flops_per_launch"], "ms": ms,
                                 "tflops": tf, "frac_of_fma_peak": tf / peak,
                                 "fp64_pipe_pct_ncu": ops[k]["fp64_pipe_pct"]}
    return out


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# --------------------------------------------------------------------------
# clocks
# --------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100", "-i", str(index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        self.p.wait()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx = float(parts[1])
            except ValueError:
                continue
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        self.f.close()
        os.unlink(self.f.name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx,
                "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------
# reference arm (CPU): oracle/_ref when compiled, else the oracle port
# --------------------------------------------------------------------------
def _cpu_paths_per_sec(po, wp, vmax, amax, threads, impl, r_offset=None):
    """Time the reference's CPU algorithm on `threads` host threads.  impl "ref": every
    path goes through the reference's own object code (TimeableJointSplinePath +
    PathTimingTrajectory, oracle/_ref) inside ONE C call that spawns the threads
    (tpref_plan_joint_batch: no Python, no per-path allocation of result arrays in the
    timed loop).  impl "port": the C restatement, one ctypes call per path from Python
    threads (only when oracle/_ref could not be built).  Returns paths/s, seconds, the
    number of paths timed successfully and the per-path results."""
    B = wp.shape[0]
    if impl == "ref":
        t0 = time.perf_counter()
        res = po.plan_joint_batch_ref(wp, vmax, amax, NUM_SAMPLES, dt=TIME_STEP, threads=threads,
                                      r_offset=r_offset)
        dt = time.perf_counter() - t0
        return B / dt, dt, int((res["status"] == 0).sum()), res

    from concurrent.futures import ThreadPoolExecutor

    dur = np.full(B, np.nan)
    rows = np.zeros(B, np.int32)

    def work(lo_hi):
        lo, hi = lo_hi
        ok = 0
        for b in range(lo, hi):
            r = po.plan_joint(wp[b], vmax[b], amax[b], NUM_SAMPLES, dt=TIME_STEP,
                              M_cap=16384, impl=impl)
            ok += bool(r["ok"])
            dur[b], rows[b] = r["duration"], r["M"]
        return ok

    chunks = [(i * B // threads, (i + 1) * B // threads) for i in range(threads)]
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        ok = sum(ex.map(work, chunks))
    dt = time.perf_counter() - t0
    return B / dt, dt, ok, {"duration": dur, "rows": rows,
                            "status": np.where(np.isnan(dur), 13, 0).astype(np.int32)}


def _cpu_setup():
    po = load_oracle()
    po.build()
    impl = "ref" if po.have_ref() else "port"
    kind = "reference" if impl == "ref" else "port"
    cores = os.cpu_count() or 1
    return po, impl, kind, cores


def _config():
    """The workload description BOTH arms print (identical keys and values)."""
    return {"workload": WORKLOAD, "paths_per_gpu_per_step": PATHS_PER_GPU,
            "num_samples": NUM_SAMPLES, "num_dofs": NUM_DOFS, "num_waypoints": NUM_WAYPOINTS,
            "time_step_s": TIME_STEP,
            "l2": "per-step working set ~2 GB (> 126 MB L2), no flush needed"}


def _quiet_stderr():
    """The reference logs "Negative sd2" / "Unbounded problem" hints to stderr."""
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(2)
    os.dup2(devnull, 2)
    return saved


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    pkg = load_package()
    po, impl, kind, cores = _cpu_setup()
    # one step = the 4,096 paths of the cfg2 batch once on all host threads (~1 s of wall
    # time on 16 threads); the host does not grow with --gpus, so neither does the sample
    sample = PATHS_PER_GPU
    wp, vmax, amax = pkg.workloads.joint_paths(PATHS_PER_GPU, NUM_WAYPOINTS, NUM_DOFS, 1000)
    saved = _quiet_stderr()
    try:
        for _ in range(args.warmup):
            _cpu_paths_per_sec(po, wp, vmax, amax, cores, impl)
        times = []
        for _ in range(args.steps):
            _, dt, ok, _ = _cpu_paths_per_sec(po, wp, vmax, amax, cores, impl)
            times.append(dt)
        # the single-thread figure on a bounded sample (the reference is single-threaded)
        one = min(sample, 128)
        v1, dt1, _, _ = _cpu_paths_per_sec(po, wp[:one], vmax[:one], amax[:one], 1, impl)
    finally:
        os.dup2(saved, 2)
    ms = 1e3 * float(np.mean(times))
    value = sample / (ms / 1e3)
    desc = (f"the {sample} paths of the cfg2 batch per step; every stage is the reference's own "
            f"object code (TimeableJointSplinePath + PathTimingTrajectory compiled verbatim, "
            f"oracle/_ref), one path per task, threads spawned inside one C call"
            if kind == "reference" else
            f"the {sample} paths of the cfg2 batch per step; C restatement of all stages")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": _config(),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": desc, "paths_ok": ok,
                         "single_thread": {"value": v1, "unit": UNIT, "cores": 1,
                                           "sample": f"first {one} paths, {dt1:.1f} s"},
                         "per_thread": value / cores},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "host_threads": cores,
    }))


# --------------------------------------------------------------------------
# our arm (CUDA through the C ABI)
# --------------------------------------------------------------------------
def _bytes_per_sample_model(D):
    """Algorithmic HBM bytes per path sample of each stage kernel."""
    return {
        # writes q,q',q'' + (sd2_max, sd2_zero, sdd_max, sdd_min) + flag
        "sample_limit": (3 * D + 4) * 8 + 1,
        # reads raw limit curve (3x8) + sd2_zero (8) + flag (1) + q',q'' (2D x 8);
        # writes working limit curve (3x8) + type (1) + s, sd2, sdd, sd, time (5x8)
        "integrate": 3 * 8 + 8 + 1 + 2 * D * 8 + 3 * 8 + 1 + 5 * 8,
        # the sweeps alone (k_optimize_pair): read q',q'' (2D x 8), sd2_max (8), type (1);
        # read-modify-write sd2 (2 x 8), write sdd (8)
        "optimize": 2 * D * 8 + 8 + 1 + 2 * 8 + 8,
    }


def run_ours(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the engine has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    pkg = load_package()
    B, N, D, W = PATHS_PER_GPU, NUM_SAMPLES, NUM_DOFS, NUM_WAYPOINTS
    wp, vmax, amax = pkg.workloads.joint_paths(B * world, W, D, seed=1000)
    sl = slice(rank * B, (rank + 1) * B)
    wp, vmax, amax = wp[sl].copy(), vmax[sl].copy(), amax[sl].copy()
    d_wp = torch.from_numpy(wp).to(dev)
    d_vmax = torch.from_numpy(vmax).to(dev)
    d_amax = torch.from_numpy(amax).to(dev)

    r_cap = B * 6144
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=r_cap, device=local_rank)
    # a dedicated (non-default) stream shared by torch's events and the engine:
    # tpb_set_stream(NULL) would mean "engine-owned stream", and torch.cuda.Event
    # only sees work on the stream it is recorded on
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    eng.set_stream(stream.cuda_stream)

    def stage_fit():
        eng.set_joint_waypoints(d_wp)
        eng.set_joint_limits(d_vmax, d_amax)

    def step_device():
        stage_fit()
        eng.sample_paths()
        eng.optimize()
        eng.resample(0.0, TIME_STEP, want_total=False)

    # ---- value: inputs resident in HBM, CUDA events on the launching stream ----
    for _ in range(args.warmup):
        step_device()
    torch.cuda.synchronize()
    barrier()
    clocks = ClockSampler(local_rank) if rank == 0 else None
    l0 = eng.kernel_launches
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    ev0.record(stream)
    for _ in range(args.steps):
        step_device()
    ev1.record(stream)
    torch.cuda.synchronize()
    barrier()
    elapsed_ms = max_over_ranks(ev0.elapsed_time(ev1))
    launches = eng.kernel_launches - l0
    clock_info = clocks.stop() if clocks else None
    ms_per_step = elapsed_ms / args.steps
    value = B * world / (ms_per_step / 1e3)

    status = eng.get(pkg.ARR.STATUS)
    durations = eng.get(pkg.ARR.DURATION)
    n_ok = int((status == 0).sum())
    total_rows = int(eng.get(pkg.ARR.R_OFFSET)[-1])
    if total_rows > r_cap:
        raise SystemExit(f"resampled rows {total_rows} exceed capacity {r_cap}")

    # ---- per-stage device times (same work, events around each C-ABI stage) ----
    stage_ms = {"fit": 0.0, "sample_limit": 0.0, "integrate": 0.0, "resample": 0.0}
    phase_ms = [0.0, 0.0, 0.0]  # setup, sweeps, finalize kernels of the integrate stage
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
    reps = max(3, args.steps)
    eng.time_phases(True)
    for _ in range(reps):
        evs[0].record(stream)
        stage_fit()
        evs[1].record(stream)
        eng.sample_paths()
        evs[2].record(stream)
        eng.optimize()
        evs[3].record(stream)
        eng.resample(0.0, TIME_STEP, want_total=False)
        evs[4].record(stream)
        torch.cuda.synchronize()
        for i, k in enumerate(stage_ms):
            stage_ms[k] += evs[i].elapsed_time(evs[i + 1]) / reps
        for i, ms in enumerate(eng.last_phase_ms()):
            phase_ms[i] += ms / reps
    eng.time_phases(False)
    split = phase_ms[0] > 0.0  # three kernels (joint paths, D <= 8) or one
    peak, peak_src = _peaks()
    model = _bytes_per_sample_model(D)
    # the dominant KERNEL: the integrate stage is three launches on this path, of
    # which the sweeps (k_optimize_pair) are timed by the engine's own events
    kernel_ms = {"sample_limit": stage_ms["sample_limit"], "resample": stage_ms["resample"]}
    if split:
        kernel_ms["optimize"] = phase_ms[1]
    else:
        kernel_ms["integrate"] = stage_ms["integrate"]
    dom = max(kernel_ms, key=lambda k: kernel_ms[k])
    if dom == "resample":
        alg_bytes = total_rows * (3 * D + 4) * 8 + B * N * (3 * D + 4) * 8
    else:
        alg_bytes = B * N * model[dom]
    achieved = alg_bytes / (kernel_ms[dom] / 1e3) / 1e9
    kernel_names = {"sample_limit": "k_joint_sample_limit", "integrate": "k_integrate",
                    "optimize": "k_optimize_pair", "resample": "k_resample_fill"}
    if split:
        stage_ms["integrate_phases"] = {"setup": phase_ms[0], "sweeps": phase_ms[1],
                                        "finalize": phase_ms[2]}
    fp64 = _fp64_report(pkg, local_rank, {"k_joint_sample_limit": stage_ms["sample_limit"],
                                           "k_integrate_setup": phase_ms[0],
                                           "k_optimize_pair": phase_ms[1],
                                           "k_integrate_finalize": phase_ms[2],
                                           "k_resample_fill": stage_ms["resample"]} if split else {})
    hbm = {"achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
           "algorithmic_bytes_per_launch": alg_bytes, "peak_source": peak_src}
    traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
    fk = (fp64 or {}).get("kernels", {}).get(kernel_names[dom])
    if dom == "resample" or fk is None:
        # a streaming kernel: HBM is its roof
        roofline = dict(hbm, bound="hbm", kernel=kernel_names[dom])
    else:
        # limit curve / phase-plane sweeps: FP64 arithmetic on a sequential dependence
        # chain.  The roof SURVEY 8(d) names is the FP64 pipe; HBM is shown beside it.
        roofline = {"bound": "fp64/latency", "kernel": kernel_names[dom],
                    "achieved": fk["tflops"], "peak": fp64["peak_tflops"], "unit": "TFLOP/s",
                    "frac": fk["frac_of_fma_peak"], "flops_per_launch": fk["flops_per_launch"],
                    "fp64_pipe_pct_ncu": fk["fp64_pipe_pct_ncu"],
                    "peak_source": fp64["peak_source"], "hbm": hbm,
                    "note": "sequential FP64 phase-plane integration: bound by the dependent-"
                            "instruction chain of one extremal step and FP64 issue, not by HBM "
                            "(DESIGN.md 4)"}
    roofline.update(traffic=_traffic(kernel_names[dom]), avg_launch_ms=kernel_ms[dom],
                    stale=_stale(traffic_file) or bool(fp64 and fp64["stale"]))

    # ---- stopping trajectories on the resident resampled batch ----
    stop = run_stop_leg(pkg, torch, dev, stream, eng, d_amax, B, D, max_over_ranks, barrier,
                        world, args)

    # ---- trajectory buffers fed from the resident resampled batch ----
    tbuf = run_tbuf_leg(pkg, torch, dev, stream, local_rank, eng, d_amax, B, D, peak,
                        max_over_ranks, barrier, world, args)

    # ---- e2e: the call a C++ user makes, PathTimingTrajectoryBatch::Plan (host class ->
    # tpb_multi_plan_joint_host), host buffers in, page-locked result arrays out ----
    h_wp = torch.from_numpy(wp).pin_memory()
    h_vmax = torch.from_numpy(vmax).pin_memory()
    h_amax = torch.from_numpy(amax).pin_memory()
    rows = total_rows
    hb = pkg.HostBatch(D, N, int(round(TIME_STEP * 1e9)), B, W, r_cap, devices=(local_rank,))
    h2d = wp.nbytes + vmax.nbytes + amax.nbytes
    PCIE_GBS = 55.0  # what one Gen5 x16 link sustains device-to-host (profiles/r1_d2h_scaling)

    def e2e_leg(want, d2h_bytes):
        for _ in range(min(args.warmup, 3)):
            hb.plan(h_wp, h_vmax, h_amax, want=want)
        torch.cuda.synchronize()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            got = hb.plan(h_wp, h_vmax, h_amax, want=want)
        sec = max_over_ranks(time.perf_counter() - t0) / args.steps
        barrier()
        return got, {"value": B * world / sec, "unit": UNIT, "ms_per_step": 1e3 * sec,
                     "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h_bytes),
                     "pcie_frac": d2h_bytes / sec / 1e9 / PCIE_GBS}

    fixed = B * 8 + B * 4 + (B + 1) * 8  # duration, status, offsets
    got, e2e = e2e_leg("trajectory", fixed + rows * (3 * D + 3) * 8)
    assert got == rows
    e2e["api"] = ("PathTimingTrajectoryBatch::Plan (host C++ class, libtpb_hostapi.so): whole "
                  "trajectory (positions, velocities, accelerations, s, sd, sdd)")
    # keep copies of what the parity check below compares
    out = {"duration": torch.from_numpy(hb.array("duration").copy()),
           "status": torch.from_numpy(hb.array("status").copy()),
           "r_offset": torch.from_numpy(hb.array("offset").copy()),
           "r_pos": torch.from_numpy(hb.array("positions").reshape(-1, D).copy())}
    same = bool(np.array_equal(out["duration"].numpy(), durations, equal_nan=True))
    # fewer bytes: a GetPositions() consumer, and a TimeOptimalPathProfile consumer
    got, e2e_pos = e2e_leg("positions", fixed + rows * D * 8)
    assert got == rows
    same = same and np.array_equal(hb.array("positions").reshape(-1, D), out["r_pos"].numpy())
    got, e2e_prof = e2e_leg("profile", B * 12 + 4 * B * N * 8)
    same = same and bool(np.array_equal(hb.array("duration"), durations, equal_nan=True))
    e2e["positions_only"] = e2e_pos
    e2e["timing_only"] = e2e_prof
    hb.close()

    # ---- dense spline evaluation sweep (BASELINE config 3) ----
    spline = run_spline_sweep(pkg, torch, dev, stream, peak, max_over_ranks, barrier, world,
                              args)

    # ---- Cartesian paths (BASELINE config 5) ----
    cartesian = run_cartesian_leg(pkg, torch, dev, stream, local_rank, max_over_ranks, barrier,
                                  world, rank, args)

    # ---- moving-horizon re-planning ----
    planner = run_planner_leg(pkg, torch, dev, stream, local_rank, d_wp, d_vmax, d_amax, B, D, W,
                              max_over_ranks, barrier, world, args)

    # ---- BASELINE config 1: one path, latency ----
    cfg1 = run_cfg1_leg(pkg, local_rank, rank, world, args)

    # ---- BASELINE config 4: 65,536 paths in total, strong-scaled over the ranks ----
    cfg4 = run_cfg4_leg(pkg, torch, dev, stream, local_rank, rank, world, max_over_ranks,
                        barrier, args)

    # ---- final gather of the per-path durations (outside any timed region) ----
    if world > 1:
        g = [torch.empty(B, dtype=torch.float64, device=dev) for _ in range(world)]
        dist.all_gather(g, torch.from_numpy(durations).to(dev))
        all_dur = torch.cat(g).cpu().numpy()
    else:
        all_dur = durations

    cpu_baseline, parity = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        po, impl, kind, cores = _cpu_setup()
        # bounded sample: the 4,096 paths of this batch once on all host threads (10-30
        # core-seconds) + the first 128 on one thread.  The results are not thrown away:
        # every path's duration, row count and resampled positions are compared with the
        # GPU's (the e2e leg's host buffers).
        gpu_off = out["r_offset"].numpy()
        saved = _quiet_stderr()
        try:
            _cpu_paths_per_sec(po, wp[:cores], vmax[:cores], amax[:cores], cores, impl)
            v, dt, ok, res = _cpu_paths_per_sec(po, wp, vmax, amax, cores, impl,
                                                r_offset=gpu_off if impl == "ref" else None)
            one = min(B, 128)
            v1, dt1, _, _ = _cpu_paths_per_sec(po, wp[:one], vmax[:one], amax[:one], 1, impl)
        finally:
            os.dup2(saved, 2)
        what = ("every stage = the reference's own object code, oracle/_ref"
                if kind == "reference" else "oracle port")
        cpu_baseline = {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
                        "sample": f"the {B} paths of this batch, {dt:.1f} s wall = "
                                  f"{dt * cores:.0f} core-seconds on {cores} host threads ({what})",
                        "per_thread": v / cores,
                        "single_thread": {"value": v1, "unit": UNIT, "cores": 1,
                                          "sample": f"first {one} paths, {dt1:.1f} s"}}
        gpu_dur = out["duration"].numpy()
        gpu_ok = out["status"].numpy() == 0
        cpu_ok = res["status"] == 0
        both = gpu_ok & cpu_ok
        rel = np.abs(gpu_dur[both] - res["duration"][both]) / res["duration"][both]
        gpu_rows = np.diff(gpu_off)
        parity = {"paths": int(B), "status_equal": bool(np.array_equal(gpu_ok, cpu_ok)),
                  "paths_ok": int(both.sum()),
                  "max_rel_duration_err": float(rel.max()) if rel.size else None,
                "paths_beyond_tolerance": int((rel > 1e-6).sum()),
                  "durations_bit_equal": int((gpu_dur[both] == res["duration"][both]).sum()),
                  "rows_equal": bool(np.array_equal(gpu_rows[both], res["rows"][both])),
                  "tolerances": "duration 1e-6 relative, positions 1e-9 rad (BASELINE.json)",
                  "against": what}
        if "r_pos" in res:
            d = np.abs(out["r_pos"].numpy() - res["r_pos"])
            parity["max_abs_position_err_rad"] = float(np.nanmax(d))
            parity["positions_compared"] = int(np.isfinite(d).sum())
            parity["positions_bit_equal"] = int((out["r_pos"].numpy() == res["r_pos"]).sum())

    if rank == 0:
        # the second half of the metric ("B-spline evals/sec vs HBM roofline"), compact:
        # the headline configuration (degree 3, 7 dofs, two derivatives) and the worst one
        spline_compact = None
        if spline:
            poly = {k: v for k, v in spline.items()
                    if isinstance(v, dict) and "frac_of_peak" in v and k.count("_") == 1}
            head = spline["p3_d7"]
            worst = min(poly, key=lambda k: poly[k]["frac_of_peak"])
            spline_compact = {
                "kernel": "k_spline_eval_poly", "config": "cfg3 p3_d7: degree 3, 7 dofs, 2 derivatives, "
                "4096 splines x 24576 parameters", "evals_per_s": head["evals_per_s"],
                "achieved": head["hbm_gbs_per_gpu"], "peak": peak, "unit": "GB/s",
                "frac": head["frac_of_peak"], "worst_config": worst,
                "worst_frac": poly[worst]["frac_of_peak"],
                "shuffled_u_frac": spline.get("p3_d7_shuffled", {}).get("frac_of_peak"),
                "exact_mode_frac": spline.get("p3_d7_exact", {}).get("frac_of_peak"),
                "cpu_baseline": spline.get("cpu_baseline")}
        cfg5 = None
        if cartesian:
            cb = cartesian.get("cpu_baseline") or {}
            cfg5 = {"workload": "4096 Cartesian paths (W = 5 poses, N = 1000) per GPU, device "
                                "kinematics, no resampling",
                    "paths_per_s": cartesian["paths_per_s"], "ms": cartesian["ms"],
                    "cpu_single_thread_paths_per_s": cb.get("single_thread_paths_per_s"),
                    "cpu_all_cores_paths_per_s": cb.get("all_cores_paths_per_s"),
                    "cpu_cores": cb.get("cores"), "cpu_kind": cb.get("kind"),
                    "max_rel_duration_err": cb.get("max_rel_duration_err")}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": _config(),
            "spline_eval": spline_compact,
            "roofline": roofline, "cpu_baseline": cpu_baseline, "parity": parity, "e2e": e2e,
            "cfg4": cfg4, "cfg1": cfg1, "cfg5": cfg5,
            "gpu_launches": int(launches), "clocks": clock_info,
            "stage_ms": stage_ms, "paths_ok": n_ok, "resampled_rows_per_gpu": total_rows,
            "mean_duration_s": float(np.nanmean(all_dur)),
            "e2e_matches_device_run": same,
            "parallelism": f"independent-path sharding x{world}, no data-path collective",
        }
        # everything else (secondary legs) goes to a file, so that the line above survives
        # the tail of any log: gpurun_out/bench_full.json
        full = dict(line, spline_eval_full=spline, fp64=fp64, cartesian=cartesian, stop=stop,
                    planner=planner, trajectory_buffer=tbuf)
        try:
            os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
            with open(os.path.join(ROOT, "gpurun_out", f"bench_full_n{world}.json"), "w") as f:
                json.dump(full, f)
            line["full"] = f"gpurun_out/bench_full_n{world}.json"
        except OSError:
            line["full"] = None
        print(json.dumps(line))
    eng.close()
    if world > 1:
        dist.destroy_process_group()


def run_cfg1_leg(pkg, local_rank, rank, world, args):
    """BASELINE config 1, the reference's own test-sized workload: ONE 7-DoF joint-space path
    (10 waypoints ~ U(-pi, pi)^7, degree-2 spline as the reference fits it, v_max 1, a_max 2),
    N = 1000 path samples, 4 ms resampling, whole path in one window.  Latency of one call:
    the host C++ class with a batch of one (PathTimingTrajectoryBatch::Plan on the GPU,
    host buffers in and out) next to the reference's own object code on one host thread."""
    if rank != 0 or world != 1:
        return None
    rng = np.random.default_rng(1)
    D, W, N = NUM_DOFS, 10, 1000
    wp = rng.uniform(-np.pi, np.pi, (1, W, D))
    vmax, amax = np.full((1, D), 1.0), np.full((1, D), 2.0)
    hb = pkg.HostBatch(D, N, int(round(TIME_STEP * 1e9)), 1, W, 65536, devices=(local_rank,))
    lat = []
    for i in range(25):
        t0 = time.perf_counter()
        rows = hb.plan(wp, vmax, amax, want="trajectory")
        lat.append(time.perf_counter() - t0)
    lat = np.array(lat[5:]) * 1e6
    dur = float(hb.array("duration")[0])
    hb.close()
    out = {"workload": "cfg1: 1 path, 7 dof, 10 waypoints, N = 1000, 4 ms resampling",
           "gpu_us_best": float(lat.min()), "gpu_us_median": float(np.median(lat)),
           "rows": int(rows), "duration_s": dur,
           "api": "PathTimingTrajectoryBatch::Plan, batch of one, host buffers in / out"}
    if not args.no_cpu_baseline:
        po, impl, kind, _ = _cpu_setup()
        if impl == "ref":
            saved = _quiet_stderr()
            try:
                cl = []
                for i in range(25):
                    t0 = time.perf_counter()
                    r = po.plan_joint_batch_ref(wp, vmax, amax, N, dt=TIME_STEP, threads=1)
                    cl.append(time.perf_counter() - t0)
            finally:
                os.dup2(saved, 2)
            cl = np.array(cl[5:]) * 1e6
            out.update(cpu_us_best=float(cl.min()), cpu_us_median=float(np.median(cl)),
                       cpu_kind="reference (object code, 1 thread)",
                       duration_equal=bool(r["duration"][0] == dur),
                       rows_equal=bool(r["rows"][0] == rows))
    return out


CFG4_PATHS = 65536


def run_cfg4_leg(pkg, torch, dev, stream, local_rank, rank, world, max_over_ranks, barrier, args):
    """BASELINE config 4 as written: 65,536 synthetic 7-DoF paths x 2,048 samples IN TOTAL,
    sharded over the ranks in the contiguous blocks of tpb_shard_range (strong scaling: at
    N = 1 one GPU times all 65,536), independent-path partitioning, no data-path
    collective; the only gather is durations + status to the host (timed in `e2e`).
    Device-timed with CUDA events on the launching stream, max over ranks."""
    if args.no_cfg4:
        return None
    N, D, W = NUM_SAMPLES, NUM_DOFS, NUM_WAYPOINTS
    first, count = pkg.shard_range(CFG4_PATHS, world, rank)
    wp, vmax, amax = pkg.workloads.joint_paths(CFG4_PATHS, W, D, seed=1000)
    sl = slice(first, first + count)
    wp, vmax, amax = wp[sl].copy(), vmax[sl].copy(), amax[sl].copy()
    d_wp, d_vm, d_am = (torch.from_numpy(x).to(dev) for x in (wp, vmax, amax))
    eng = pkg.Engine(max_paths=count, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=count * 4608, device=local_rank)
    eng.set_stream(stream.cuda_stream)
    eng.time_phases(True)
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
    stage = np.zeros(4)
    phase = np.zeros(3)
    reps = max(3, min(args.steps, 5))

    def step():
        evs[0].record(stream)
        eng.set_joint_waypoints(d_wp)
        eng.set_joint_limits(d_vm, d_am)
        evs[1].record(stream)
        eng.sample_paths()
        evs[2].record(stream)
        eng.optimize()
        evs[3].record(stream)
        eng.resample(0.0, TIME_STEP, want_total=False)
        evs[4].record(stream)

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    barrier()
    total_ms = 0.0
    for _ in range(reps):
        step()
        torch.cuda.synchronize()
        total_ms += evs[0].elapsed_time(evs[4]) / reps
        stage += [evs[i].elapsed_time(evs[i + 1]) / reps for i in range(4)]
        phase += np.array(eng.last_phase_ms()) / reps
    ms = max_over_ranks(total_ms)
    eng.time_phases(False)
    status = eng.get(pkg.ARR.STATUS)
    rows = int(eng.get(pkg.ARR.R_OFFSET)[-1])
    # e2e: host waypoints in, durations + status out (SURVEY 8e's optional final gather);
    # the trajectories stay on their GPUs for device-resident consumers
    h = [torch.from_numpy(x).pin_memory() for x in (wp, vmax, amax)]
    out = {"duration": torch.empty(count, dtype=torch.float64).pin_memory(),
           "status": torch.empty(count, dtype=torch.int32).pin_memory()}
    for _ in range(2):
        eng.plan_joint_host(h[0], h[1], h[2], 0.0, TIME_STEP, out=out)
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        eng.plan_joint_host(h[0], h[1], h[2], 0.0, TIME_STEP, out=out)
    e2e_s = max_over_ranks(time.perf_counter() - t0) / reps
    ok = int((status == 0).sum())
    eng.close()
    del d_wp, d_vm, d_am
    torch.cuda.empty_cache()
    return {"paths_total": CFG4_PATHS, "paths_this_rank": count, "scaling": "strong",
            "ms_per_step": ms, "paths_per_s": CFG4_PATHS / (ms / 1e3),
            "stage_ms_rank0": {"fit": stage[0], "sample_limit": stage[1], "integrate": stage[2],
                               "resample": stage[3], "setup": phase[0], "sweeps": phase[1],
                               "finalize": phase[2]},
            "paths_ok_rank0": ok, "rows_rank0": rows,
            "e2e": {"value": CFG4_PATHS / e2e_s, "unit": UNIT, "ms_per_step": 1e3 * e2e_s,
                    "h2d_bytes_per_step": int(wp.nbytes + vmax.nbytes + amax.nbytes),
                    "d2h_bytes_per_step": int(count * 12),
                    "api": "tpb_plan_joint_host, durations + status out (rows stay in HBM)"},
            "partition": "tpb_shard_range: contiguous blocks of ceil(65536 / N) paths"}


def run_planner_leg(pkg, torch, dev, stream, local_rank, d_wp, d_vmax, d_amax, B, D, W,
                    max_over_ranks, barrier, world, args):
    """Moving-horizon re-planning of the cfg2 batch: every path is planned
    in windows of 512 path samples that cover a quarter of the longest path, Plan() is called
    again every 200 ms with a 750 ms horizon (the reference test's cadence,
    path_timing_trajectory_test.cc:64-65) until every path is at its end.  All planner state
    stays in HBM; per call the host reads one 4-byte flag per planning window.  One untimed
    pass collects the call / window counts, the timed pass repeats it (CUDA events on the
    launching stream around the whole loop)."""
    if args.no_planner:
        return None
    N = 512
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=B * 4096, device=local_rank)
    eng.set_stream(stream.cuda_stream)
    eng.set_joint_waypoints(d_wp)
    eng.set_joint_limits(d_vmax, d_amax)
    knot_end = float(eng.get(pkg.ARR.KNOT_END).max())
    delta = knot_end / (N - 1) / 4.0
    eng.set_sampling(np.full(B, delta))
    horizon, replan, dt = 750_000_000, 200_000_000, TIME_STEP

    def one_pass(collect):
        eng.planner_reset(8 * N, dt)
        start, calls, windows, erased, rows = 0, 0, 0, 0, 0
        while calls < 400:
            total = eng.planner_plan(start, horizon, want_total=collect)
            calls += 1
            if collect:
                st = eng.get(pkg.ARR.PLAN_STATUS)
                windows += int(eng.get(pkg.ARR.PLAN_LOOPS).sum())
                erased += int((eng.get(pkg.ARR.PLAN_ERASED) == 1).sum())
                rows = max(rows, total)
                done = (eng.get(pkg.ARR.PLAN_TARGET_REACHED) == 1) | (st != 0)
                if done.all():
                    break
            elif calls == one_pass.calls:
                break
            start += replan
        return calls, windows, erased, rows

    calls, windows, erased, rows = one_pass(True)
    one_pass.calls = calls
    one_pass(False)
    torch.cuda.synchronize()
    barrier()
    l0 = eng.kernel_launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    one_pass(False)
    e1.record(stream)
    torch.cuda.synchronize()
    ms = max_over_ranks(e0.elapsed_time(e1))
    launches = eng.kernel_launches - l0
    reached = int((eng.get(pkg.ARR.PLAN_TARGET_REACHED) == 1).sum())
    eng.close()
    return {"path_plans_per_s": world * B * calls / (ms / 1e3), "ms_total": ms,
            "ms_per_plan_call": ms / calls, "plan_calls": calls, "paths_per_gpu": B,
            "windows_planned": windows, "paths_only_erased": erased,
            "window_paths_per_s": world * windows / (ms / 1e3), "paths_at_end": reached,
            "max_rows_per_gpu": int(rows), "gpu_launches": int(launches),
            "config": {"window_samples": N, "windows_per_longest_path": 4, "horizon_ms": 750,
                       "replan_ms": 200, "time_step_s": dt}}


class _DevPtr:
    """A raw device address with the data_ptr() the binding expects."""
    is_cuda = True

    def __init__(self, ptr, shape=None):
        self._p = ptr
        self.shape = shape

    def data_ptr(self):
        return self._p


def run_tbuf_leg(pkg, torch, dev, stream, local_rank, eng, d_amax, B, D, peak, max_over_ranks,
                 barrier, world, args):
    """Device-resident TrajectoryBuffers as the consumer of the timed batch: (1) every
    resampled trajectory (the engine's packed TPB_ARR_R_* arrays) is inserted into its own
    buffer without leaving HBM, (2) every buffer is sampled (position, velocity,
    acceleration) at 2,048 times spread over its span, (3) every buffer is stopped before
    60 % of its span (StopBeforeTime).  CUDA events on the launching stream."""
    if args.no_tbuf:
        return None
    off = eng.get(pkg.ARR.R_OFFSET).astype(np.int64)
    r_t = eng.get(pkg.ARR.R_TIME)
    rows = int(off[-1])
    counts = (off[1:] - off[:-1]).astype(np.int32)
    cap = int(counts.max()) + 64
    batch = pkg.TrajectoryBufferBatch(B, D, cap, device=local_rank)
    batch.set_stream(stream.cuda_stream)
    dp = lambda what, shape: _DevPtr(eng.devptr(what)[0], shape)
    seg = (dp(pkg.ARR.R_OFFSET, (B + 1,)), torch.from_numpy(counts).to(dev),
           dp(pkg.ARR.R_TIME, (rows,)), dp(pkg.ARR.R_POS, (rows, D)), dp(pkg.ARR.R_VEL, (rows, D)),
           dp(pkg.ARR.R_ACC, (rows, D)))
    st = torch.zeros(B, dtype=torch.int32, device=dev)
    Q = 2048
    t0, t1 = r_t[off[:-1]], r_t[off[1:] - 1]
    tq = t0[:, None] + (t1 - t0)[:, None] * ((np.arange(Q) + 0.5) / Q)[None, :]
    d_tq = torch.from_numpy(np.ascontiguousarray(tq)).to(dev)
    outs = tuple(torch.empty((B, Q, D), dtype=torch.float64, device=dev) for _ in range(3))
    d_qst = torch.zeros((B, Q), dtype=torch.int32, device=dev)
    d_tstop = torch.from_numpy(t0 + 0.6 * (t1 - t0)).to(dev)
    d_sst = torch.zeros(B, dtype=torch.int32, device=dev)
    amax = _DevPtr(d_amax.data_ptr(), (B, D))
    sel = torch.ones(B, dtype=torch.int32, device=dev)

    def insert():
        batch.clear(sel)
        batch.insert_segments(*seg, status=st)

    def sample():
        batch.sample(d_tq, out=outs + (d_qst,))

    def stop():
        lib = pkg.lib()
        rc = lib.tpb_tbuf_stop_before_time(batch._h, d_tstop.data_ptr(), amax.data_ptr(), 0.004,
                                           d_sst.data_ptr(), 1)
        assert rc == 0

    def timed(fn, reps, before=None):
        for _ in range(3):
            if before:
                before()
            fn()
        torch.cuda.synchronize()
        barrier()
        total = 0.0
        for _ in range(reps):
            if before:
                before()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            fn()
            e1.record(stream)
            torch.cuda.synchronize()
            total += e0.elapsed_time(e1)
        return max_over_ranks(total) / reps

    reps = max(3, args.steps)
    launches0 = batch.kernel_launches
    ms_insert = timed(insert, reps)
    assert (st.cpu().numpy() == 0).all()
    ms_sample = timed(sample, reps)
    assert (d_qst.cpu().numpy() == 0).all()
    ms_stop = timed(stop, reps, before=insert)
    stop_ok = int((d_sst.cpu().numpy() == 0).sum())
    launches = batch.kernel_launches - launches0
    row_bytes = (3 * D + 1) * 8
    ins_gbs = 2.0 * rows * row_bytes / (ms_insert / 1e3) / 1e9
    query_bytes = 8 + 2 * 3 * D * 8 + 3 * D * 8 + 4
    smp_gbs = B * Q * query_bytes / (ms_sample / 1e3) / 1e9
    batch.close()
    return {"insert": {"ms": ms_insert, "rows_per_s": world * rows / (ms_insert / 1e3),
                       "hbm_gbs_per_gpu": ins_gbs, "frac_of_peak": ins_gbs / peak,
                       "bytes_per_row": f"{row_bytes} read + {row_bytes} written",
                       "kernel": "k_tbuf_clear + k_tbuf_insert"},
            "sample": {"ms": ms_sample, "queries_per_s": world * B * Q / (ms_sample / 1e3),
                       "hbm_gbs_per_gpu": smp_gbs, "frac_of_peak": smp_gbs / peak,
                       "queries_per_buffer": Q,
                       "bytes_per_query": f"{query_bytes}: query time 8 + two bracketing rows of "
                                          f"three arrays {2 * 3 * D * 8} read + {3 * D * 8} "
                                          "written + status 4",
                       "kernel": "k_tbuf_sample"},
            "stop_before_time": {"ms": ms_stop, "buffers_per_s": world * B / (ms_stop / 1e3),
                                 "ok": stop_ok,
                                 "kernels": "k_tbuf_stop_prepare + k_rescale_to_stop + "
                                            "k_tbuf_stop_commit"},
            "buffers_per_gpu": B, "rows_per_gpu": rows, "capacity_rows": cap,
            "gpu_launches": int(launches)}


def run_stop_leg(pkg, torch, dev, stream, eng, d_amax, B, D, max_over_ranks, barrier, world,
                 args):
    """Batched RescaleTrajectoryBackwardToStop: every resampled trajectory of the batch is cut
    at 60 % of its samples (where it still moves) and the fastest stop tracing the tail of what
    is left is generated, all trajectories in one launch.  Inputs are the engine's packed
    TPB_ARR_R_* arrays in HBM; CUDA events on the launching stream."""
    if args.no_stop:
        return None
    off = eng.get(pkg.ARR.R_OFFSET).astype(np.int64)
    rows = int(off[-1])
    length = np.maximum(2, ((off[1:] - off[:-1]) * 0.6).astype(np.int64)).astype(np.int32)
    d_begin = torch.from_numpy(off[:-1].copy()).to(dev)
    d_len = torch.from_numpy(length).to(dev)
    t_in = _DevPtr(eng.devptr(pkg.ARR.R_TIME)[0])
    v_in = _DevPtr(eng.devptr(pkg.ARR.R_VEL)[0])
    a_in = _DevPtr(eng.devptr(pkg.ARR.R_ACC)[0])
    cnt = torch.zeros(B, dtype=torch.int32, device=dev)
    sw = torch.zeros(B, dtype=torch.int32, device=dev)
    st = torch.zeros(B, dtype=torch.int32, device=dev)
    o_t = torch.empty(rows, dtype=torch.float64, device=dev)
    o_v = torch.empty(rows * D, dtype=torch.float64, device=dev)
    o_a = torch.empty(rows * D, dtype=torch.float64, device=dev)

    class _Shape:  # the binding reads .shape of these two
        def __init__(self, t, shape):
            self._t, self.shape = t, shape

        def data_ptr(self):
            return self._t.data_ptr()

    amax = _Shape(d_amax, (B, D))
    times = _Shape(t_in, (rows,))

    def run():
        pkg.rescale_to_stop_device(amax, d_begin, d_len, times, v_in, a_in, cnt, sw, st, o_t, o_v,
                                   o_a, stream=stream.cuda_stream)

    for _ in range(3):
        run()
    torch.cuda.synchronize()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = max(3, args.steps)
    e0.record(stream)
    for _ in range(reps):
        run()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = max_over_ranks(e0.elapsed_time(e1)) / reps
    c = cnt.cpu().numpy()
    return {"trajectories_per_s": world * B / (ms / 1e3), "ms": ms, "trajectories_per_gpu": B,
            "ok": int((st.cpu().numpy() == 0).sum()), "mean_input_samples": float(length.mean()),
            "mean_stop_samples": float(c.mean()), "kernel": "k_rescale_to_stop",
            "note": "tail of each 4 ms resampled trajectory cut at 60 % of its samples"}


def run_cartesian_leg(pkg, torch, dev, stream, local_rank, max_over_ranks, barrier, world, rank,
                      args):
    """BASELINE config 5: 4,096 Cartesian paths (position + quaternion spline, W = 5 pose
    waypoints, N = 1000 samples) per GPU: fit, SamplePath (quaternion / translation / joint
    spline targets, built-in fake-robot IK, forward-difference derivatives), ConstraintSetup
    (2D+2 = 16 rows), limit curve and time-optimal timing.  The whole path is one planning
    window (delta = knots.back() / (N - 1)).  Inputs resident in HBM, CUDA events."""
    if args.no_cartesian:
        return None
    B, W, N = 4096, 5, 1000
    pw, jw, vm, am = pkg.workloads.cart_paths(B * world, W, seed=5000)
    sl = slice(rank * B, (rank + 1) * B)
    d_pw = torch.from_numpy(pw[sl].copy()).to(dev)
    d_jw = torch.from_numpy(jw[sl].copy()).to(dev)
    d_vm = torch.from_numpy(vm[sl].copy()).to(dev)
    d_am = torch.from_numpy(am[sl].copy()).to(dev)
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=7, max_waypoints=W, cartesian=True,
                     device=local_rank)
    eng.set_stream(stream.cuda_stream)
    eng.set_cart_options(0.05, 0.5, 0.25)

    def step():
        eng.set_cart_waypoints(d_pw, d_jw)
        eng.set_joint_limits(d_vm, d_am)
        eng.cart_sample_targets()
        eng.cart_solve_ik(1)
        eng.cart_setup_constraints(1)
        eng.optimize()

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = max(3, args.steps)
    e0.record(stream)
    for _ in range(reps):
        step()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = max_over_ranks(e0.elapsed_time(e1)) / reps
    status = eng.get(pkg.ARR.STATUS)
    dur = eng.get(pkg.ARR.DURATION)
    eng.close()
    out = {"paths_per_s": world * B / (ms / 1e3), "ms": ms, "paths_per_gpu": B,
           "num_samples": N, "num_waypoints": W, "rows_per_sample": 16,
           "kinematics": "TPB_KIN_FAKE7 (reference test's fake robot) on the device",
           "paths_ok": int((status == 0).sum()), "mean_duration_s": float(np.nanmean(dur))}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # the reference's own TimeableCartesianSplinePath + PathTimingTrajectory object code
        # (oracle/_ref) on a bounded sample of the same paths; its PathIKFunc / JacobianFunc
        # are the same fake robot as host callbacks
        po, impl, kind, cores = _cpu_setup()
        if impl == "ref":
            n1, nall = 64, min(B, 64 * cores)
            t0 = time.perf_counter()
            r1 = po.plan_cart_batch_ref(pw[:n1], jw[:n1], vm[:n1], am[:n1], N, dt=TIME_STEP,
                                        threads=1)
            t1 = time.perf_counter()
            ra = po.plan_cart_batch_ref(pw[:nall], jw[:nall], vm[:nall], am[:nall], N,
                                        dt=TIME_STEP, threads=cores)
            t2 = time.perf_counter()
            both = (ra["status"] == 0) & (status[:nall] == 0)
            rel = np.abs(dur[:nall][both] - ra["duration"][both]) / ra["duration"][both]
            out["cpu_baseline"] = {
                "kind": kind, "cores": cores,
                "single_thread_paths_per_s": n1 / (t1 - t0),
                "all_cores_paths_per_s": nall / (t2 - t1),
                "sample": f"{n1} paths on one thread, {nall} on {cores} (reference object code "
                          "of bsplineq.cc / timeable_path_cartesian_spline.cc / "
                          "path_timing_trajectory.cc incl. resampling)",
                "paths_compared": int(both.sum()),
                "status_equal": bool(((ra["status"] == 0) == (status[:nall] == 0)).all()),
                "max_rel_duration_err": float(rel.max()) if rel.size else None,
                "tolerance": "1e-6 relative on T_total (BASELINE.json); device kinematics "
                             "differ from the host callbacks by libm rounding",
                "single_thread_status_ok": int((r1["status"] == 0).sum())}
    return out


def run_spline_sweep(pkg, torch, dev, stream, peak, max_over_ranks, barrier, world, args):
    """BASELINE config 3: 4,096 splines x 24,576 parameters = 1.007e8 evaluations
    with 1st/2nd derivatives per (degree, dim); output [S][E][3][dim] in HBM."""
    if args.no_spline:
        return None
    S, E, npts = 4096, 24576, 32
    res = {}
    out = torch.empty(S * E * 3 * 7, dtype=torch.float64, device=dev)
    for P in (3, 4, 5):
        for D in (7, 6):
            knots, ctrl = pkg.workloads.dense_splines(S, P, D, npts, 8.0)
            d_k = torch.from_numpy(knots).to(dev)
            d_c = torch.from_numpy(ctrl).to(dev)
            u0 = torch.zeros(S, dtype=torch.float64, device=dev)
            du = torch.full((S,), 8.0 / (E - 1), dtype=torch.float64, device=dev)
            o = out[:S * E * 3 * D].view(S, E, 3, D)

            for exact in (0, 1, -1):
                def run():
                    pkg.spline_eval_device(d_k, d_c, P, 2, o, u0=u0, du=du, evals=E,
                                           exact=exact, stream=stream.cuda_stream)
                for _ in range(3):
                    run()
                torch.cuda.synchronize()
                barrier()
                e0, e1 = (torch.cuda.Event(enable_timing=True),
                          torch.cuda.Event(enable_timing=True))
                reps = 5
                e0.record(stream)
                for _ in range(reps):
                    run()
                e1.record(stream)
                torch.cuda.synchronize()
                ms = max_over_ranks(e0.elapsed_time(e1)) / reps
                gbs = S * E * 3 * D * 8 / (ms / 1e3) / 1e9
                key = f"p{P}_d{D}" + {0: "", 1: "_exact", -1: "_threadstores"}[exact]
                res[key] = {"evals_per_s": world * S * E / (ms / 1e3), "ms": ms,
                            "hbm_gbs_per_gpu": gbs, "frac_of_peak": gbs / peak}
            # shuffled-u variant: the same parameters in random order per spline,
            # read from an explicit array (8 B per evaluation more)
            if (P, D) in ((3, 7), (5, 6)):
                g = torch.Generator(device="cpu").manual_seed(1234 + P)
                grid = torch.linspace(0.0, 8.0, E, dtype=torch.float64)
                perm = torch.stack([grid[torch.randperm(E, generator=g)] for _ in range(8)])
                d_u = perm.repeat(S // 8, 1).contiguous().to(dev)

                def run_u():
                    pkg.spline_eval_device(d_k, d_c, P, 2, o, u=d_u, exact=False,
                                           stream=stream.cuda_stream)
                for _ in range(3):
                    run_u()
                torch.cuda.synchronize()
                barrier()
                e0, e1 = (torch.cuda.Event(enable_timing=True),
                          torch.cuda.Event(enable_timing=True))
                e0.record(stream)
                for _ in range(5):
                    run_u()
                e1.record(stream)
                torch.cuda.synchronize()
                ms = max_over_ranks(e0.elapsed_time(e1)) / 5
                gbs = S * E * (3 * D * 8 + 8) / (ms / 1e3) / 1e9
                res[f"p{P}_d{D}_shuffled"] = {"evals_per_s": world * S * E / (ms / 1e3), "ms": ms,
                                              "hbm_gbs_per_gpu": gbs, "frac_of_peak": gbs / peak,
                                              "bytes_per_eval": 3 * D * 8 + 8}
                del d_u
    if world == 1 and not args.no_cpu_baseline:
        # CPU baseline of cfg3: the reference's own BSplineXd::EvalCurveAndDerivatives object
        # code, one call per parameter, degree 3, 7 dofs, 2 derivatives, the same splines
        po, impl, kind, cores = _cpu_setup()
        if impl == "ref":
            knots, ctrl = pkg.workloads.dense_splines(8 * cores, 3, 7, npts, 8.0)
            u = np.tile(np.linspace(0.0, 8.0, E), (8 * cores, 1))
            t0 = time.perf_counter()
            ref1, _ = po.spline_eval_batch_ref(knots[:8], ctrl[:8], 3, 2, u[:8], threads=1)
            t1 = time.perf_counter()
            po.spline_eval_batch_ref(knots, ctrl, 3, 2, u, threads=cores)
            t2 = time.perf_counter()
            # parity of the device evaluation with it, on the first 8 splines
            d_k = torch.from_numpy(knots[:8]).to(dev)
            d_c = torch.from_numpy(ctrl[:8]).to(dev)
            d_u = torch.from_numpy(u[:8]).to(dev)
            o8 = torch.empty((8, E, 3, 7), dtype=torch.float64, device=dev)
            err = {}
            for mode, name in ((1, "exact"), (0, "polynomial")):
                pkg.spline_eval_device(d_k, d_c, 3, 2, o8, u=d_u, exact=mode,
                                       stream=stream.cuda_stream)
                torch.cuda.synchronize()
                err[name] = float(np.max(np.abs(o8.cpu().numpy() - ref1)))
            res["cpu_baseline"] = {
                "kind": "reference (BSplineXd::EvalCurveAndDerivatives object code)",
                "config": "p3_d7, 2 derivatives", "cores": cores,
                "single_thread_evals_per_s": 8 * E / (t1 - t0),
                "all_cores_evals_per_s": 8 * cores * E / (t2 - t1),
                "max_abs_err_vs_reference": err,
                "tolerance": "1e-12 absolute (BASELINE.json); exact mode is bit-identical"}
    res["bytes_per_eval"] = "3*dim*8 written (implicit parameter grid, tables in smem)"
    res["modes"] = ("pP_dD: per-span polynomial form (<= 1e-12 of the reference), tiles leave shared "
                    "memory as bulk async copies (UBLKCP); pP_dD_threadstores: the same with per-"
                    "thread 16-byte stores (A/B); pP_dD_exact: "
                    "NURBS A2.3 in the reference's operation order (bit-identical, "
                    "FP64-division bound)")
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-spline", action="store_true")
    ap.add_argument("--no-cartesian", action="store_true")
    ap.add_argument("--no-stop", action="store_true")
    ap.add_argument("--no-planner", action="store_true")
    ap.add_argument("--no-tbuf", action="store_true")
    ap.add_argument("--no-cfg4", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
