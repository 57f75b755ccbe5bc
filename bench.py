#!/usr/bin/env python
"""bench.py — throughput of the hot path (reprojection residual + Jacobian + Schur elimination) on a B200.

One "step" = one pass of eval + Schur over the synthetic graph of the named BASELINE.json config (default:
config 4, the 2k-keyframe / 200k-landmark / ~2M-observation global refinement). Prints ONE JSON line.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config 1..5] [--scale S] [--impl reference]
"""
import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

CONFIG_NAMES = {1: "C1 VIO window 10kf/300lm", 2: "C2 LVIO window 20kf/1k lm", 3: "C3 submap refinement 200kf/20k lm",
                4: "C4 global map refinement 2k kf/200k lm/~2M obs", 5: "C5 BA stress 10k kf/1M lm/~10M obs"}
METRIC = "reproj_obs_per_s (resid+Jac+Schur)"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 7 for n, v in zip(names, r[3:7]) if v.lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def snapshot_path(config, scale):
    import tempfile
    return os.path.join(tempfile.gettempdir(), f"bsnls_snapshot_c{config}_s{scale}.bsnls")


def build_problem(config, scale):
    """Synthetic graph of the config + its flattened problem; the problem is also left as a snapshot file so that the
    other arm of the same box works on the identical arrays without regenerating them."""
    from beam_slam_b200 import host
    t0 = time.time()
    g = host.HashGraph.synthetic(config, scale=scale)
    p = g.flatten()
    try:
        if int(os.environ.get("RANK", "0")) == 0 and not os.path.exists(snapshot_path(config, scale)):
            p.save(snapshot_path(config, scale))
    except OSError:
        pass
    return g, p, time.time() - t0


def load_problem_for_reference(config, scale):
    """The reference arm reads the snapshot with numpy alone when the other arm already left one (then it loads none of the
    product's native libraries); otherwise it generates the graph."""
    from beam_slam_b200 import cabi
    path = snapshot_path(config, scale)
    if os.path.exists(path):
        p = cabi.FlatProblem.load_numpy(path)
        nobs = p.count(cabi.FACTOR_REPROJECTION)
        info = {"num_observations": nobs, "num_landmarks": int((p.block_group == 0).sum()), "num_keyframes": None}
        return p, info, "snapshot " + path
    g, p, _ = build_problem(config, scale)
    return p, g.info, "generated"


def algorithmic_bytes(info, nf_dim, dense):
    """184 B/obs eval + 168 B/obs Schur, + 120 B/landmark + 56 B/keyframe + reduced-system write."""
    nobs, nlm, nkf = info["num_observations"], info["num_landmarks"], info["num_keyframes"]
    eval_b = 184 * nobs + 56 * nkf + 24 * nlm
    schur_b = 168 * nobs + 96 * nlm + 8 * nf_dim * (nf_dim + 1) // 2 * (1 if dense else 0) + 8 * nf_dim
    return eval_b, schur_b


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def load_traffic():
    """Per-launch DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) of the hot kernels, from the committed
    `ncu --set full` capture (profiles/): {"capture": ..., "kernels": {name: bytes}, "stages": {stage: [names]}}."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    return json.load(open(p)) if os.path.exists(p) else None


def cpu_pass(p, steps, warmup):
    """Oracle (CPU restatement of the reference's Ceres path) on the SAME graph, every host core: seconds per
    eval + Schur pass (residuals, Jacobians, landmark elimination into the block-sparse reduced system)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as orc
    cores = host_cores()
    orc.set_num_threads(cores)     # torchrun exports OMP_NUM_THREADS=1
    for _ in range(warmup):
        orc.time_eval_schur(p, 1e4, 1)
    ts = []
    for _ in range(steps):
        sec, threads = orc.time_eval_schur(p, 1e4, 1)
        ts.append(sec)
    return ts, threads


def run_reference(args):
    """The reference arm: the reference's own (CPU) implementation of the path — Ceres is not installable here, so its
    restatement under oracle/ — on the GPU arm's workload (same graph, same size), all host cores, rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    t0 = time.time()
    p, info, source = load_problem_for_reference(args.config, args.scale)
    ts, threads = cpu_pass(p, args.steps, min(args.warmup, 1))
    nobs = info["num_observations"]
    v = nobs * len(ts) / sum(ts)
    sample = f"config {args.config} at scale {args.scale} ({source}): {info['num_landmarks']} lm, {nobs} obs per step (one eval+Schur pass per step)"
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": "obs/s", "n_gpus": args.gpus, "steps": args.steps,
                      "warmup": min(args.warmup, 1), "ms_per_step": 1e3 * sum(ts) / len(ts), "higher_is_better": True, "scaling": "strong",
                      "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                      "config": {"workload": CONFIG_NAMES[args.config], "scale": args.scale, "observations": nobs, "sample": sample},
                      "cpu_baseline": {"value": v, "unit": "obs/s", "cores": threads, "host_cores": host_cores(), "kind": "port", "sample": sample},
                      "e2e": {"value": v, "unit": "obs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                      "wall_s": time.time() - t0}))


def smoother_cycles(host, opts, cycles=24):
    """The fixed-lag smoother's call pattern on a 10-keyframe VIO window (fixed_lag_smoother.cpp:220-281): every cycle a
    keyframe enters, the oldest is pseudo-marginalised, optimize() runs on the changed graph. Wall clock of optimize()
    (flatten + bsnls_set_problem + solve + write-back), median over the cycles after two warm-ups."""
    r = host.SmootherReplay(cycles + 12, 10)
    walls, parts, dev, its_n, upd = [], [], [], [], []
    for _ in range(cycles):
        upd.append(r.advance())
        t0 = time.perf_counter()
        s, its, timing = r.optimize(opts)
        walls.append(time.perf_counter() - t0)
        parts.append(timing)
        dev.append(s.device_time_in_seconds)
        its_n.append(s.num_iterations)
    k = slice(2, None)
    med = lambda v: float(np.median(v[k]))
    return {"pattern": "one keyframe added + oldest pseudo-marginalised per cycle, structure rebuilt every call",
            "optimize_wall_ms_median": 1e3 * med(walls), "optimize_wall_ms_max": 1e3 * float(np.max(walls[k])),
            "flatten_ms": 1e3 * med([q["flatten_s"] for q in parts]), "set_problem_ms": 1e3 * med([q["upload_s"] for q in parts]),
            "solve_ms": 1e3 * med([q["solve_s"] for q in parts]), "writeback_ms": 1e3 * med([q["download_s"] for q in parts]),
            "solve_device_ms": 1e3 * med(dev), "graph_update_ms (caller side)": 1e3 * med(upd),
            "iterations_median": float(np.median(its_n[k])), "observations": r.observations(), "cycles": cycles}


def marginalization_cycles(host, opts, poses=40, cycles=12):
    """The smoother's other branch (fixed_lag_smoother.cpp:269-272: fuse_constraints::marginalizeVariables) on a pose-graph
    window (odometry + longer links, no landmarks): every cycle the oldest pose is marginalised for real -- its constraints
    are linearised on the device (bsnls_linearize), eliminated on the host, a MarginalConstraint stays on its neighbours --
    and optimize() runs on the changed graph."""
    import ctypes as C
    lib = host.load_library()

    def pose(k):
        out = []
        for kind in (b"fuse_variables::Position3DStamped", b"fuse_variables::Orientation3DStamped"):
            buf = C.create_string_buffer(16)
            lib.bsfuse_uuid_stamped(kind, 10 + k, 0, buf)
            out.append(buf.raw)
        return out
    g = host.HashGraph.pose_chain(poses, seed=2)
    g.optimize(opts)
    marg, opt, dev = [], [], []
    for k in range(cycles):
        t0 = time.perf_counter()
        g.marginalize(pose(k))
        t1 = time.perf_counter()
        s, _ = g.optimize(opts)
        t2 = time.perf_counter()
        marg.append(t1 - t0); opt.append(t2 - t1); dev.append(s.device_time_in_seconds)
    med = lambda v: 1e3 * float(np.median(v[2:]))
    return {"pattern": f"pose graph of {poses} poses: the oldest pose truly marginalised per cycle (linearise on the device, QR on the host, MarginalConstraint), then optimize()",
            "marginalize_wall_ms_median": med(marg), "optimize_wall_ms_median": med(opt), "solve_device_ms_median": med(dev), "cycles": cycles}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from beam_slam_b200 import cabi, host, sharding
    from beam_slam_b200.solver import Solver
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU path)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    peak, peak_src = load_peaks()
    g, p, t_build = build_problem(args.config, args.scale)
    info = g.info
    nobs = info["num_observations"]
    opts = host.HashGraph.synthetic_options(args.config)
    opts.device = local
    s = sharding.create_sharded_solver(opts) if world > 1 else Solver(opts)   # landmark shard per rank, NCCL exchange
    t0 = time.time()
    s.set_problem(p)
    t_setup = time.time() - t0
    _, fb = s.ordering()
    nf_dim = 3 * len(fb)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(args.warmup):
        s.run_pass(1e4)
    launches0 = s.launch_count()
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    ev_ms, sc_ms, parts = [], [], []
    wall0 = time.time()
    for _ in range(args.steps):
        s.run_pass(1e4)             # one complete linearisation: residuals + Jacobians + landmark elimination + reduced system
        a, b = s.last_kernel_ms()   # CUDA events on the library's stream, around the pass (one CUDA-graph launch)
        ev_ms.append(a)
        sc_ms.append(b)
    barrier()
    wall = time.time() - wall0
    launches = s.launch_count() - launches0
    dev_ms = max_over_ranks(sum(ev_ms) + sum(sc_ms))
    if wall < 0.6:
        # the timed region is shorter than nvidia-smi's sampling period: keep the same kernels running (untimed) so
        # that the clock / throttle record is taken under the same load. Same count on every rank (collectives inside).
        for _ in range(min(4000, int(0.8 / max(dev_ms / args.steps * 1e-3, 1e-5)) + 1)):
            s.run_pass(1e4)
    # per-kernel split: the same pass launched kernel by kernel with events in between (outside the timed region)
    s.enable_breakdown(True)
    for _ in range(3):
        s.run_pass(1e4)
        parts.append(s.last_breakdown_ms())
    s.enable_breakdown(False)
    clocks = sampler.stop()
    clocks["sampled"] = "during the timed region" if wall >= 0.6 else "same kernels looped for 0.8 s right after the timed region (region too short to sample)"
    value = nobs * args.steps / (dev_ms * 1e-3)

    # pass level, through the C-ABI with host buffers: H2D of the step's parameter values, eval + Schur, D2H of the cost
    x_host = torch.from_numpy(p.block_values.copy()).pin_memory().numpy()
    barrier()
    t0 = time.time()
    for _ in range(args.steps):
        s.set_values(x_host)
        s.run_pass(1e4)
    barrier()
    pass_e2e = nobs * args.steps / max_over_ranks(time.time() - t0)

    # whole-solve numbers (LM iterations/s) on the same graph
    s.set_values(p.block_values)
    summ, its = s.solve()
    s.set_values(p.block_values)
    summ, its = s.solve()          # second solve of the structure: CUDA-graph replay where graphs are used
    lm_iters = max(1, summ.num_iterations - 1)
    lm_iters_per_s = lm_iters / summ.device_time_in_seconds if summ.device_time_in_seconds > 0 else None
    lm = {"iterations": summ.num_iterations, "lm_iters_per_s": lm_iters_per_s, "solve_device_ms": summ.device_time_in_seconds * 1e3,
          "termination": summ.message.decode(), "initial_cost": summ.initial_cost, "final_cost": summ.final_cost}

    # The plug-in call a fuse user makes: GpuHashGraph::optimize with host buffers in and out (Variable::data()):
    # flatten or value gather, bsnls_set_problem or bsnls_set_values, bsnls_solve, bsnls_get_values, scatter.
    # One GPU only (the graph-level drop-in is the single-GPU path; the sharded solve is driven through the C-ABI).
    plug = None
    if world == 1:
        x0 = g.current_values()
        first = None
        walls, splits, iters = [], [], []
        for i in range(args.warmup + args.steps):
            g.set_values(x0)                      # the caller's new estimates; not part of optimize()
            t0 = time.perf_counter()
            sg, itg = g.optimize(opts)
            w = time.perf_counter() - t0
            if i == 0:
                first = {"wall_ms": 1e3 * w, **{k: (1e3 * v if k.endswith("_s") else v) for k, v in g.last_timing().items()}}
            if i >= args.warmup:
                walls.append(w)
                splits.append(g.last_timing())
                iters.append(max(1, sg.num_iterations - 1))
        tot = sum(walls)
        plug = {"call": "GpuHashGraph::optimize (flatten | set_problem or set_values | solve | get_values + scatter), host buffers in and out",
                "value": nobs * sum(iters) / tot, "unit": "obs/s", "optimize_wall_ms": 1e3 * tot / len(walls), "lm_steps_per_call": float(np.mean(iters)),
                "flatten_or_gather_ms": 1e3 * float(np.mean([q["flatten_s"] for q in splits])), "upload_ms": 1e3 * float(np.mean([q["upload_s"] for q in splits])),
                "solve_ms": 1e3 * float(np.mean([q["solve_s"] for q in splits])), "writeback_ms": 1e3 * float(np.mean([q["download_s"] for q in splits])),
                "structure_reused": bool(splits[-1]["structure_reused"]), "first_call": first,
                "h2d_bytes_per_step": int(8 * p.num_values), "d2h_bytes_per_step": int(8 * p.num_values + 88 * sg.num_iterations),
                "note": "value = observations x LM steps per call / wall time of the call; every LM step passes every observation through residual + Jacobian + Schur"}
        assert abs(sg.final_cost - summ.final_cost) <= 1e-9 * summ.final_cost

    # N > 1: the sharded solve against a single-GPU solve of the same graph on rank 0 (parity of the NCCL path where the
    # driver runs N ranks)
    if world > 1:
        if rank == 0:
            o1 = host.HashGraph.synthetic_options(args.config)
            o1.device = local
            s1 = Solver(o1)
            s1.set_problem(p)
            sm1, it1 = s1.solve()
            lm["rel_diff_vs_single_gpu"] = {
                "iterations_equal": len(it1) == len(its),
                "max_cost_rel_diff_per_iteration": max(abs(a.cost - b.cost) / abs(b.cost) for a, b in zip(its, it1)),
                "final_cost_rel_diff": abs(summ.final_cost - sm1.final_cost) / abs(sm1.final_cost),
                "first_step_norm_rel_diff": abs(its[1].step_norm - it1[1].step_norm) / it1[1].step_norm if len(its) > 1 and len(it1) > 1 else None,
                "single_gpu_solve_device_ms": sm1.device_time_in_seconds * 1e3}
            s1.close()
        barrier()

    mean = lambda k: float(np.mean([q[k] for q in parts]))
    br = {k: mean(k) for k in Solver.BREAKDOWN}
    dense = nf_dim <= 160   # sliding windows keep a dense reduced system; everything larger is 3x3-block sparse
    eval_b, schur_b = algorithmic_bytes(info, nf_dim, dense)
    if not dense:
        schur_b += 72 * (2 * summ.reduced_system_nnz_blocks - nf_dim // 3)   # block-sparse S write (full symmetric storage)
    eval_b, schur_b = eval_b // world, schur_b // world   # per-rank share of the algorithmic bytes
    fused = br["landmark"] == 0.0 and nobs > 0   # chunk-major fused pass: Jacobians, elimination and chunk tiles in one kernel
    if fused:
        stages = {"fused eval+elimination+SYRK (chunk_fused_kernel) + pair_reduce + rows": (br["eval_reproj"] + br["pair"] + br["generic_rows"] + br["clear_S"], eval_b + schur_b)}
    else:
        stages = {"eval_reproj": (br["eval_reproj"], eval_b), "schur(landmark+pair+rows)": (br["landmark"] + br["pair"] + br["generic_rows"] + br["clear_S"], schur_b)}
    dom = max(stages, key=lambda k: stages[k][0])
    dom_ms, dom_bytes = stages[dom]
    achieved = dom_bytes / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else 0.0
    tr = load_traffic()
    traffic, traffic_detail = None, None
    if tr and world == 1 and args.config == tr.get("config") and args.scale == tr.get("scale"):
        names = tr["stages"].get(dom, [])
        traffic = int(sum(tr["kernels"][n] for n in names)) if names else None
        traffic_detail = {"capture": tr["capture"], "dram_bytes_per_launch": tr["kernels"],
                          "dram_over_algorithmic": {st: sum(tr["kernels"][n] for n in ns) / stages[st][1] for st, ns in tr["stages"].items() if st in stages}}
    e2e = {"value": plug["value"], "unit": "obs/s", "h2d_bytes_per_step": plug["h2d_bytes_per_step"], "d2h_bytes_per_step": plug["d2h_bytes_per_step"],
           "call": plug["call"], "optimize_wall_ms": plug["optimize_wall_ms"]} if plug else \
          {"value": pass_e2e, "unit": "obs/s", "h2d_bytes_per_step": int(8 * p.num_values), "d2h_bytes_per_step": 8 * 16,
           "call": "bsnls_set_values + bsnls_run_pass per step (sharded solve: driven through the C-ABI)"}
    out = {"metric": METRIC, "value": value, "unit": "obs/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
           "data": "synthetic",
           "config": {"workload": CONFIG_NAMES[args.config], "scale": args.scale, "keyframes": info["num_keyframes"],
                      "landmarks": info["num_landmarks"], "observations": nobs, "reduced_system_dim": nf_dim,
                      "generator": "analytic smooth trajectory (closed-form position, Euler-angle attitude), not the basalt quintic B-spline of SURVEY 8(d)",
                      "l2": "inputs larger than L2 (Jacobian + elimination records %.0f MB)" % (272e-6 * nobs) if nobs * 272 > 130e6 else "inputs fit in L2 (latency-bound config)",
                      "multi_gpu": "single GPU" if world == 1 else f"landmarks sharded over {world} GPUs, NCCL all-reduce of the reduced camera system per step"},
           "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                        "traffic": traffic, "traffic_detail": traffic_detail, "peak_source": peak_src, "algorithmic_bytes_per_launch": dom_bytes, "kernel_ms": dom_ms,
                        "whole_pass_frac": (eval_b + schur_b) / (dev_ms / args.steps * 1e-3) / 1e9 / peak, "per": "GPU"},
           "kernel_ms": br, "kernel_ms_note": "the timed steps are one CUDA-graph launch each; this split comes from 3 extra passes launched kernel by kernel with events in between",
           "e2e": e2e,
           "e2e_pass": {"value": pass_e2e, "unit": "obs/s", "call": "bsnls_set_values + bsnls_run_pass per step", "h2d_bytes_per_step": int(8 * p.num_values), "d2h_bytes_per_step": 8 * 16},
           "plugin_optimize": plug,
           "gpu_launches": launches, "clocks": clocks, "lm": lm,
           "setup_s": {"generate+flatten": t_build, "set_problem": t_setup}, "wall_s_timed": wall}
    if rank == 0:
        if world == 1:
            try:
                ts, threads = cpu_pass(p, 3, 1)
                out["cpu_baseline"] = {"value": nobs * len(ts) / sum(ts), "unit": "obs/s", "cores": threads, "host_cores": host_cores(), "kind": "port",
                                       "sample": f"the same graph at full size ({nobs} obs): 3 eval+Schur passes after 1 warm-up, {1e3 * sum(ts) / len(ts):.0f} ms each"}
            except Exception as e:  # the baseline is reported, never required
                out["cpu_baseline"] = {"value": None, "unit": "obs/s", "cores": 0, "kind": "port", "sample": f"failed: {e}"}
            try:  # the latency-bound sliding-window target of the north star: the 10-keyframe VIO window
                o1 = host.HashGraph.synthetic_options(1)
                o1.device = local
                g1 = host.HashGraph.synthetic(1)
                x1 = g1.current_values() if g1.flatten() is not None else None
                walls, dev = [], []
                for _ in range(10):
                    g1.set_values(x1)
                    t0 = time.perf_counter()
                    sm1, _ = g1.optimize(o1)
                    walls.append(time.perf_counter() - t0)
                    dev.append(sm1.device_time_in_seconds * 1e3)
                out["vio_window"] = {"config": CONFIG_NAMES[1], "iterations": sm1.num_iterations, "observations": g1.info["num_observations"],
                                     "unchanged_graph": {"optimize_wall_ms_median": 1e3 * float(np.median(walls[3:])), "solve_device_ms_median": float(np.median(dev[3:]))},
                                     "smoother_cycles": smoother_cycles(host, o1)}
            except Exception as e:
                out["vio_window"] = {"error": str(e)}
            try:
                o3 = host.HashGraph.synthetic_options(3)
                o3.device = local
                out["marginalization"] = marginalization_cycles(host, o3)
            except Exception as e:
                out["marginalization"] = {"error": str(e)}
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", type=int, default=4)
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--impl", default="ours")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
