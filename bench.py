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


def build_problem(config, scale):
    from beam_slam_b200 import host
    t0 = time.time()
    g = host.HashGraph.synthetic(config, scale=scale)
    p = g.flatten()
    return g, p, time.time() - t0


def algorithmic_bytes(info, nf_dim, dense):
    """184 B/obs eval + 168 B/obs Schur, + 120 B/landmark + 56 B/keyframe + reduced-system write."""
    nobs, nlm, nkf = info["num_observations"], info["num_landmarks"], info["num_keyframes"]
    eval_b = 184 * nobs + 56 * nkf + 24 * nlm
    schur_b = 168 * nobs + 96 * nlm + 8 * nf_dim * (nf_dim + 1) // 2 * (1 if dense else 0) + 8 * nf_dim
    return eval_b, schur_b


def cpu_sample(config, scale, steps=2):
    """Oracle (CPU port of the reference's Ceres path) on a bounded sample of the same workload."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as orc
    sample_scale = min(scale, {1: 1.0, 2: 1.0, 3: 1.0, 4: 0.1, 5: 0.02}[config])
    g, p, _ = build_problem(config, sample_scale)
    sec, threads = orc.time_eval_schur(p, 1e4, steps)
    return g.info["num_observations"] / sec, threads, f"config {config} at scale {sample_scale}: {g.info['num_keyframes']} kf, " \
        f"{g.info['num_landmarks']} lm, {g.info['num_observations']} obs; best of {steps} eval+Schur passes"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    t0 = time.time()
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as orc
    sample_scale = min(args.scale, {1: 1.0, 2: 1.0, 3: 1.0, 4: 0.1, 5: 0.02}[args.config])
    g, p, _ = build_problem(args.config, sample_scale)
    for _ in range(args.warmup):
        orc.time_eval_schur(p, 1e4, 1)
    ts = []
    for _ in range(args.steps):
        sec, threads = orc.time_eval_schur(p, 1e4, 1)
        ts.append(sec)
    nobs = g.info["num_observations"]
    v = nobs * len(ts) / sum(ts)
    sample = f"config {args.config} at scale {sample_scale}: {g.info['num_keyframes']} kf, {g.info['num_landmarks']} lm, {nobs} obs per step"
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": "obs/s", "n_gpus": args.gpus, "steps": args.steps,
                      "warmup": args.warmup, "ms_per_step": 1e3 * sum(ts) / len(ts), "higher_is_better": True, "scaling": "strong",
                      "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                      "config": {"workload": CONFIG_NAMES[args.config], "sample": sample},
                      "cpu_baseline": {"value": v, "unit": "obs/s", "cores": threads, "kind": "port", "sample": sample},
                      "e2e": {"value": v, "unit": "obs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                      "wall_s": time.time() - t0}))


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

    for _ in range(args.warmup):
        s.run_eval()
        s.run_schur(1e4)
    launches0 = s.launch_count()
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    ev_ms, sc_ms, parts = [], [], []
    wall0 = time.time()
    for _ in range(args.steps):
        s.run_eval()
        s.run_schur(1e4)
        a, b = s.last_kernel_ms()   # CUDA events on the library's stream
        ev_ms.append(a)
        sc_ms.append(b)
        parts.append(s.last_breakdown_ms())
    barrier()
    wall = time.time() - wall0
    launches = s.launch_count() - launches0
    dev_ms = sum(ev_ms) + sum(sc_ms)
    if world > 1:
        t = torch.tensor([dev_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms = float(t.item())
    if wall < 0.6:
        # the timed region is shorter than nvidia-smi's sampling period: keep the same kernels running (untimed) so
        # that the clock / throttle record is taken under the same load. Same count on every rank (collectives inside).
        for _ in range(min(4000, int(0.8 / max(dev_ms / args.steps * 1e-3, 1e-5)) + 1)):
            s.run_eval()
            s.run_schur(1e4)
    clocks = sampler.stop()
    clocks["sampled"] = "during the timed region" if wall >= 0.6 else "same kernels looped for 0.8 s right after the timed region (region too short to sample)" 
    value = nobs * args.steps / (dev_ms * 1e-3)

    # end to end through the C-ABI with host buffers: H2D of the step's parameter values, eval + Schur, D2H of the cost
    x_host = torch.from_numpy(p.block_values.copy()).pin_memory().numpy()
    barrier()
    t0 = time.time()
    for _ in range(args.steps):
        s.set_values(x_host)
        s.run_eval()
        s.run_schur(1e4)
    barrier()
    e2e_s = time.time() - t0
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e = nobs * args.steps / e2e_s

    # whole-solve numbers (LM iterations/s) on the same graph
    s.set_values(p.block_values)
    summ, its = s.solve()
    lm_iters = max(1, summ.num_iterations - 1)
    lm_iters_per_s = lm_iters / summ.device_time_in_seconds if summ.device_time_in_seconds > 0 else None

    mean = lambda k: float(np.mean([q[k] for q in parts]))
    br = {k: mean(k) for k in Solver.BREAKDOWN}
    dense = nf_dim <= 160   # sliding windows keep a dense reduced system; everything larger is 3x3-block sparse
    eval_b, schur_b = algorithmic_bytes(info, nf_dim, dense)
    if not dense:
        schur_b += 72 * (2 * summ.reduced_system_nnz_blocks - nf_dim // 3)   # block-sparse S write (full symmetric storage)
    eval_b, schur_b = eval_b // world, schur_b // world   # per-rank share of the algorithmic bytes
    stages = {"eval_reproj": (br["eval_reproj"], eval_b), "schur(landmark+pair+rows)": (br["landmark"] + br["pair"] + br["generic_rows"] + br["clear_S"], schur_b)}
    dom = max(stages, key=lambda k: stages[k][0])
    dom_ms, dom_bytes = stages[dom]
    achieved = dom_bytes / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else 0.0
    out = {"metric": METRIC, "value": value, "unit": "obs/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
           "data": "synthetic",
           "config": {"workload": CONFIG_NAMES[args.config], "scale": args.scale, "keyframes": info["num_keyframes"],
                      "landmarks": info["num_landmarks"], "observations": nobs, "reduced_system_dim": nf_dim,
                      "l2": "inputs larger than L2 (Jacobian + elimination records %.0f MB)" % (272e-6 * nobs) if nobs * 272 > 130e6 else "inputs fit in L2 (latency-bound config)",
                      "multi_gpu": "single GPU" if world == 1 else f"landmarks sharded over {world} GPUs, NCCL all-reduce of the reduced camera system per step"},
           "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                        "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": dom_bytes, "kernel_ms": dom_ms,
                        "whole_pass_frac": (eval_b + schur_b) / (dev_ms / args.steps * 1e-3) / 1e9 / peak, "per": "GPU"},
           "kernel_ms": br,
           "e2e": {"value": e2e, "unit": "obs/s", "h2d_bytes_per_step": int(8 * p.num_values), "d2h_bytes_per_step": 8 * 16},
           "gpu_launches": launches, "clocks": clocks,
           "lm": {"iterations": summ.num_iterations, "lm_iters_per_s": lm_iters_per_s, "solve_device_ms": summ.device_time_in_seconds * 1e3,
                  "termination": summ.message.decode(), "initial_cost": summ.initial_cost, "final_cost": summ.final_cost},
           "setup_s": {"generate+flatten": t_build, "set_problem": t_setup}, "wall_s_timed": wall}
    if rank == 0:
        if world == 1:
            try:
                v, threads, sample = cpu_sample(args.config, args.scale)
                out["cpu_baseline"] = {"value": v, "unit": "obs/s", "cores": threads, "kind": "port", "sample": sample}
            except Exception as e:  # the baseline is reported, never required
                out["cpu_baseline"] = {"value": None, "unit": "obs/s", "cores": 0, "kind": "port", "sample": f"failed: {e}"}
            try:  # the latency-bound sliding-window target of the north star: whole solve of the 10-keyframe VIO window
                g1 = host.HashGraph.synthetic(1)
                p1 = g1.flatten()
                o1 = host.HashGraph.synthetic_options(1)
                o1.device = local
                s1 = Solver(o1)
                s1.set_problem(p1)
                dev = []
                for _ in range(8):
                    s1.set_values(p1.block_values)
                    sm1, _ = s1.solve()
                    dev.append(sm1.device_time_in_seconds * 1e3)
                out["vio_window"] = {"config": CONFIG_NAMES[1], "solve_device_ms_median": float(np.median(dev[2:])),
                                     "iterations": sm1.num_iterations, "observations": g1.info["num_observations"]}
            except Exception as e:
                out["vio_window"] = {"error": str(e)}
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
