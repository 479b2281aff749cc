"""Per-call device times (CUDA events, rotating cold inputs) of every backend entry on the bench
workloads, plus the split of one graph iteration into build / solve.  Tuning aid; prints JSON.
  python tools/kernel_timing.py [--graph N E]"""
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import GN, N_PIX, make_sets  # noqa: E402
from splatt3r_slam_b200 import matching, synthetic  # noqa: E402
import splatt3r_slam_b200.mast3r_slam_backends as be  # noqa: E402

dev = torch.device("cuda:0")
cfg = matching.DEFAULT_CFG


def timeit(fn, n_sets, reps=30):
    for i in range(3):
        fn(i % n_sets)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(reps):
        fn(i % n_sets)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3


out = {"env": {k: v for k, v in os.environ.items() if k.startswith("B200SLAM")}}
n_sets = 6
sets = []
for s in make_sets(n_sets, 100):
    d = {k: v.to(dev) for k, v in s.items()}
    rays, pts, p_init = matching.prep_for_iter_proj(d["X11"], d["X21"], None)
    d.update(rays=rays, pts=pts, p_init=p_init, D11h=d["D11"].half().contiguous(),
             D21h=d["D21"].view(1, N_PIX, -1).half().contiguous())
    p_new, _ = be.iter_proj(rays, pts, p_init, cfg["max_iter"], cfg["lambda_init"], cfg["convergence_thresh"])
    d["p1"] = p_new.long()
    d["T"] = d["Twc"].clone()
    sets.append(d)
out["iter_proj_us"] = timeit(lambda i: be.iter_proj(sets[i]["rays"], sets[i]["pts"], sets[i]["p_init"], cfg["max_iter"],
                                                    cfg["lambda_init"], cfg["convergence_thresh"]), n_sets)
out["refine_us"] = timeit(lambda i: be.refine_matches(sets[i]["D11h"], sets[i]["D21h"], sets[i]["p1"], cfg["radius"],
                                                      cfg["dilation_max"]), n_sets)
# whole match() call the way tracker.py:31 makes it: reference glue (torch ops) + 2 backend calls vs the fused entry
for name, fused in (("match_unfused_us", False), ("match_fused_us", True)):
    out[name] = timeit(lambda i, fused=fused: matching.match(sets[i]["X11"], sets[i]["X21"], sets[i]["D11h"], sets[i]["D21h"],
                                                             fused=fused), n_sets)
out["match_fused_f32desc_us"] = timeit(lambda i: matching.match(sets[i]["X11"], sets[i]["X21"], sets[i]["D11"], sets[i]["D21"],
                                                                fused=True), n_sets)
out["match_unfused_f32desc_us"] = timeit(lambda i: matching.match(sets[i]["X11"], sets[i]["X21"], sets[i]["D11"], sets[i]["D21"]),
                                         n_sets)
for iters in (1, 10):
    def gn(i, iters=iters):
        d = sets[i]
        d["T"].copy_(d["Twc"])
        be.gauss_newton_rays(d["T"], d["Xs"], d["Cs"], d["ii"], d["jj"], d["idx_ii2jj"], d["valid_match"], d["Q"],
                             GN["sigma_ray"], GN["sigma_dist"], GN["C_thresh"], GN["Q_thresh"], iters, 0.0)
    out[f"gn_tracker_{iters}it_us"] = timeit(gn, n_sets)
ws = be._workspaces[("cuda", 0)]
st = ws[:48].cpu().view(torch.int64)[2:6].tolist()
out["gn_tracker_tail_ns"] = {"finalize": st[1] - st[0], "assemble_chol": st[2] - st[1], "finish": st[3] - st[2]}
pf = ws[88:136].cpu().view(torch.int64).tolist()  # State::prof[0..2] of the persistent kernel: start, points staged, loop done
out["gn_tracker_kernel_ns"] = {"prologue": pf[1] - pf[0], "iterations": pf[2] - pf[1]}
# on-device frame tracker (8f rank 2) on a 512x384-point problem; reference restatement on the host beside it
import time  # noqa: E402
from oracle import tracker_ref  # noqa: E402
from splatt3r_slam_b200 import tracker  # noqa: E402
TP = synthetic.make_track_problem(seed=3, n=N_PIX)
TD = {k: v.to(dev) for k, v in TP.items()}
c = tracker.DEFAULT_CFG
targs = (c["sigma_ray"], c["sigma_dist"], c["huber"], c["max_iters"], c["rel_error"], c["delta_norm"])
_, _, info = be.track_ray_dist(TD["Xf"], TD["Xk"], TD["T_WCf"], TD["T_WCk"], TD["Qk"], TD["valid"], *targs)
out["track_ray_dist"] = {"iters": info["iters"], "us": timeit(lambda i: be.track_ray_dist(
    TD["Xf"], TD["Xk"], TD["T_WCf"], TD["T_WCk"], TD["Qk"], TD["valid"], *targs, check=False), 1)}
t0 = time.perf_counter()
_, _, steps = tracker_ref.opt_pose_ray_dist_sim3(TP["Xf"], TP["Xk"], TP["T_WCf"], TP["T_WCk"], TP["Qk"], TP["valid"])
out["track_ray_dist"]["cpu_restatement_us"] = (time.perf_counter() - t0) * 1e6
out["track_ray_dist"]["cpu_threads"] = torch.get_num_threads()
del sets
torch.cuda.empty_cache()

if "--no-graph" in sys.argv:
    print(json.dumps(out))
    sys.exit(0)
# ---- graph iteration split ------------------------------------------------------------------
N, E = 32, 128
if "--graph" in sys.argv:
    k = sys.argv.index("--graph")
    N, E = int(sys.argv[k + 1]), int(sys.argv[k + 2])
g = synthetic.make_graph(seed=7, N=N, E_undirected=E // 2, h=384, w=512)
G = {k: g[k].to(dev) for k in ("Twc", "Xs", "Cs", "ii", "jj", "idx_ii2jj", "valid_match", "Q")}
n = G["Xs"].shape[1]
prm = be.make_params(0, 0.003, 10.0, 0.0, 1.5)
ii_row, jj_row = be.edge_rows(G["ii"], G["jj"])
nbytes = be._lib.b200slam_gn_workspace_bytes(N, E, n)
wsg = torch.empty(int(nbytes), dtype=torch.uint8, device=dev)
eb = torch.zeros((E, 56), dtype=torch.float64, device=dev)
dx = torch.zeros((N - 1, 7), dtype=torch.float32, device=dev)
T = G["Twc"].clone()
stream = lambda: torch.cuda.current_stream(dev).cuda_stream
be._check(be._lib.b200slam_gn_reset(wsg.data_ptr(), nbytes, stream()), "reset")


def build(_):
    be._check(be._lib.b200slam_gn_edge_blocks(
        ctypes.byref(prm), T.data_ptr(), G["Xs"].data_ptr(), G["Cs"].data_ptr(), None, ii_row.data_ptr(),
        jj_row.data_ptr(), G["idx_ii2jj"].data_ptr(), G["valid_match"].data_ptr(), G["Q"].data_ptr(), N, E, n, 0, E,
        eb.data_ptr(), wsg.data_ptr(), nbytes, stream()), "edge_blocks")


def solve(_):
    T.copy_(G["Twc"])
    be._check(be._lib.b200slam_gn_solve_retract(T.data_ptr(), ii_row.data_ptr(), jj_row.data_ptr(), eb.data_ptr(), N, E,
                                                n, 0.0, dx.data_ptr(), wsg.data_ptr(), nbytes, stream()), "solve")


def full(_):
    T.copy_(G["Twc"])
    be.gauss_newton_rays(T, G["Xs"], G["Cs"], G["ii"], G["jj"], G["idx_ii2jj"], G["valid_match"], G["Q"],
                         GN["sigma_ray"], GN["sigma_dist"], GN["C_thresh"], GN["Q_thresh"], 10, 0.0)


tb = timeit(build, 1, 10)
tf = timeit(full, 1, 5)
out[f"graph_N{N}_E{E}"] = {"build_us": tb, "build_GBps": E * (45.0 * n + 840) / tb / 1e3,
                           "gn_10it_us": tf, "solve_us": timeit(solve, 1, 10)}
stg = wsg[:48].cpu().view(torch.int64)[2:6].tolist()
out[f"graph_N{N}_E{E}"]["solve_kernel_ns"] = {"load": stg[1] - stg[0], "cholesky": stg[2] - stg[1], "finish": stg[3] - stg[2]}
print(json.dumps(out))
