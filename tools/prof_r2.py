"""Workloads for the round-2 ncu captures:  python tools/prof_r2.py graph32 | graph256 | solver1024
  graph32    2 GN iterations, N=32 / E=128 directed edges, 512x384, rays then calib  (configs[2] shape)
  graph256   2 GN iterations, N=256 / E=2048, 512x384: pointmaps 0.8 GB >> L2 (what configs[3] looks like to the build kernel)
  solver1024 the large-graph solver on configs[3]'s graph (N=1024, E=8192) at tiny n (the solver does not depend on n)"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from splatt3r_slam_b200 import synthetic  # noqa: E402
import splatt3r_slam_b200.mast3r_slam_backends as be  # noqa: E402

dev = torch.device("cuda:0")
what = sys.argv[1] if len(sys.argv) > 1 else "graph32"
N, E, h, w = {"graph32": (32, 128, 384, 512), "graph256": (256, 2048, 384, 512), "solver1024": (1024, 8192, 8, 12)}[what]
G = synthetic.make_graph_device(7, N, E // 2, h, w, dev)
K = G["K"].to(dev)
for kind in (("rays", "calib") if what != "solver1024" else ("rays",)):
    for _ in range(2):
        T = G["Twc"].clone()
        if kind == "rays":
            be.gauss_newton_rays(T, G["Xs"], G["Cs"], G["ii"], G["jj"], G["idx_ii2jj"], G["valid_match"], G["Q"], 0.003, 10.0,
                                 0.0, 1.5, 2, 0.0)
        else:
            be.gauss_newton_calib(T, G["Xs"], G["Cs"], K, G["ii"], G["jj"], G["idx_ii2jj"], G["valid_match"], G["Q"], h, w, -10,
                                  1e-6, 1.0, 10.0, 0.0, 1.5, 2, 0.0)
torch.cuda.synchronize()
print("ok", what, be.gn_solver_status(dev) if N > 32 else "")
