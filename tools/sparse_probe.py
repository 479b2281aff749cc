"""Timing of the large-graph solvers on ring-like graphs (odometry ring + loop closures): block-sparse Cholesky vs
the dense tiled Cholesky.  python tools/sparse_probe.py"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from splatt3r_slam_b200 import synthetic  # noqa: E402
import splatt3r_slam_b200.mast3r_slam_backends as be  # noqa: E402
from splatt3r_slam_b200.dist_gn import CudaHooks  # noqa: E402

dev = torch.device("cuda:0")
for (N, Eu) in ((130, 130), (300, 330), (512, 560), (1024, 1024), (1024, 1100), (1024, 1200), (1024, 1400)):
    g = synthetic.make_graph_device(3, N, Eu, 8, 12, dev)
    prm = be.make_params(0, 0.003, 10.0, 0.0, 1.5)
    E, n = 2 * Eu, 96
    rows = be.edge_rows(g["ii"], g["jj"])
    for solver in ("auto", "dense"):
        be.gn_set_solver(solver)
        hooks = CudaHooks(prm, N, E, n, dev)
        eb = torch.zeros((E, 56), dtype=torch.float64, device=dev)
        hooks.edge_blocks(g["Twc"], g["Xs"], g["Cs"], None, rows[0], rows[1], g["idx_ii2jj"], g["valid_match"], g["Q"], 0, E, eb)
        T = g["Twc"].clone()
        dx = torch.zeros((N - 1, 7), device=dev)
        t0 = time.perf_counter(); hooks.analyze(rows[0], rows[1]); torch.cuda.synchronize(); ta = time.perf_counter() - t0
        t0 = time.perf_counter(); hooks.analyze(rows[0], rows[1]); torch.cuda.synchronize(); tb = time.perf_counter() - t0
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(3):
            hooks.solve_retract(T, rows[0], rows[1], eb, 0.0, dx)
        a.record()
        for _ in range(10):
            hooks.solve_retract(T, rows[0], rows[1], eb, 0.0, dx)
        b.record()
        torch.cuda.synchronize()
        st = be.gn_solver_status(hooks.ws)
        print(N, Eu, solver, "solve_ms", round(a.elapsed_time(b) / 10, 3), "analyze_ms first/cached", round(ta * 1e3, 2),
              round(tb * 1e3, 2), {k: st[k] for k in ("sparse_runs", "dense_runs", "cg_iters", "pcg_bad", "solve_us")})
    be.gn_set_solver("default")
