#!/usr/bin/env python
"""bench.py — tracker step of the dense-SLAM backend hot path on B200 (BASELINE.json configs[1]).

One "step" = one frame-to-keyframe tracker step on a synthetic 512x384 pair with 24-d
descriptors: iter_proj + refine_matches (b = 1, n = 196608 query pixels) followed by
gauss_newton_rays over the 2 directed edges of the pair (N = 2 poses, 10 iterations,
delta_thresh = 0 so all 10 run).  `value` = matched pixels per second over the whole step with
inputs resident in HBM; `e2e` = the same through the public driver API (matching.match +
gauss_newton_rays) from pinned HOST buffers, H2D and D2H inside the timed region.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload tracker|graph]

--gpus N > 1 (under torchrun): every rank runs its own independent tracker steps (pairs are
independent units: no data-path collective, weak scaling).  --workload graph runs the
edge-sharded global optimisation with one NCCL allreduce per iteration instead.
--impl reference times the CPU oracle (the reference ships no CPU path for this code; its
CUDA-only algorithm restated in C, oracle/*.c) on the host cores on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

# NCCL prints its version banner on stdout at VERSION level; the contract is ONE JSON line on stdout
if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
    os.environ["NCCL_DEBUG"] = "WARN"

import torch

# ... and torch's NCCL process group prints the banner anyway on some builds: keep the real stdout for the
# one JSON line and send everything else written to fd 1 (by any library, any thread) to stderr.
if __name__ == "__main__" and "_B200SLAM_STDOUT_FD" not in os.environ:
    sys.stdout.flush()
    os.environ["_B200SLAM_STDOUT_FD"] = str(os.dup(1))
    os.dup2(2, 1)


def emit(line):
    """The bench contract's ONE JSON line, on the process's original stdout."""
    data = (json.dumps(line) + "\n").encode()
    fd = os.environ.get("_B200SLAM_STDOUT_FD")
    if fd is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(int(fd), data)


ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H, W, FDIM = 384, 512, 24
N_PIX = H * W
GN = dict(sigma_ray=0.003, sigma_dist=10.0, C_thresh=0.0, Q_thresh=1.5, iters=10)
# algorithmic bytes per unit (SURVEY.md section 8d / DESIGN.md)
B_ITER_PROJ, B_REFINE, B_GN_POINT, B_GN_EDGE = 65, 128, 45, 840


# the ONE config dict of both arms (the driver compares them)
CONFIG = {"workload": "tracker step 512x384 24-d: iter_proj+refine_matches (b=1) + gauss_newton_rays N=2 E=2 10 iters",
          "l2": "6 rotating input sets of ~56 MB (> 126 MB L2)",
          "ranks": "headline: independent pairs per rank, no collective; `graph` block: configs[3] edge-sharded over "
                   "the ranks with one all-gather per GN iteration"}


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag = index, [], set(), False
        self.max_mhz = None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i",
                                      str(self.index)], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                self.samples.append(float(f[0]))
                self.max_mhz = float(f[1])
                for nm, v in zip(names, f[2:6]):
                    if v.lower().startswith("active"):
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.05)

    def result(self):
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def make_sets(n_sets, seed0):
    """n_sets independent tracker-step inputs (CPU tensors)."""
    from splatt3r_slam_b200 import synthetic
    sets = []
    for s in range(n_sets):
        X11, X21, D11, D21 = synthetic.make_pair(seed=seed0 + s, h=H, w=W, fdim=FDIM, b=1)
        g = synthetic.make_graph(seed=seed0 + s, N=2, E_undirected=1, h=H, w=W)
        sets.append(dict(X11=X11, X21=X21, D11=D11, D21=D21, **{k: g[k] for k in
                    ("Twc", "Xs", "Cs", "ii", "jj", "idx_ii2jj", "valid_match", "Q")}))
    return sets


def run_reference(args):
    """CPU arm: the oracle on the host cores, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle
    from splatt3r_slam_b200 import matching
    cores = os.cpu_count()
    torch.set_num_threads(cores)
    s = make_sets(1, 1000)[0]
    rays, pts, p_init = matching.prep_for_iter_proj(s["X11"], s["X21"], None)
    D11h, D21h = s["D11"].half(), s["D21"].view(1, N_PIX, -1).half().contiguous()
    cfg = matching.DEFAULT_CFG

    def step():  # one FULL tracker step: all 196608 query pixels, all 10 GN iterations — nothing extrapolated
        p_new, _ = oracle.iter_proj(rays, pts, p_init, cfg["max_iter"], cfg["lambda_init"], cfg["convergence_thresh"])
        oracle.refine_matches(D11h, D21h, p_new.long(), cfg["radius"], cfg["dilation_max"])
        T = s["Twc"].clone()
        oracle.gauss_newton_rays(T, s["Xs"], s["Cs"], s["ii"], s["jj"], s["idx_ii2jj"], s["valid_match"], s["Q"],
                                 GN["sigma_ray"], GN["sigma_dist"], GN["C_thresh"], GN["Q_thresh"], GN["iters"], 0.0)

    # a step is ~3 s of CPU work on 16 threads: the driver's 20 steps end in about a minute
    steps = max(1, args.steps)
    step()  # one warm-up step (pages, OpenMP pool)
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    t_step = (time.perf_counter() - t0) / steps
    val = N_PIX / t_step
    sample = (f"{steps} full tracker steps (one 512x384 pair: iter_proj + refine_matches over all 196608 pixels + 10 GN "
              f"iterations), wall clock, OpenMP over all host threads; 1 warm-up step")
    line = {"impl": "reference", "metric": "matched_px_per_s", "value": val, "unit": "px/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_step * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": CONFIG, "timed_steps": steps,
            "cpu_baseline": {"value": val, "unit": "px/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": "px/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="tracker")
    ap.add_argument("--kfs", type=int, default=32)
    ap.add_argument("--edges", type=int, default=128, help="directed edges for --workload graph")
    ap.add_argument("--kind", default="rays", choices=["rays", "calib"], help="residual model for --workload graph")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="skip the configs[2]/[3] graph sub-blocks of the tracker line")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    from bench_impl import run_tracker, run_graph
    if args.workload == "graph":
        return run_graph(args)
    return run_tracker(args)


if __name__ == "__main__":
    main()
