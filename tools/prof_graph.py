"""Runs two GN iterations on the graph workload (N=32 keyframes, E=128 directed edges, 512x384) so
that ncu can capture the non-fused build_kernel / finalize / assemble / solve kernels."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from splatt3r_slam_b200 import synthetic
import splatt3r_slam_b200.mast3r_slam_backends as be
dev = torch.device("cuda:0")
N, E = 32, 128
g = synthetic.make_graph(seed=7, N=N, E_undirected=E // 2, h=384, w=512)
G = {k: g[k].to(dev) for k in ("Twc", "Xs", "Cs", "ii", "jj", "idx_ii2jj", "valid_match", "Q")}
for _ in range(2):
    T = G["Twc"].clone()
    be.gauss_newton_rays(T, G["Xs"], G["Cs"], G["ii"], G["jj"], G["idx_ii2jj"], G["valid_match"], G["Q"],
                         0.003, 10.0, 0.0, 1.5, 2, 0.0)
torch.cuda.synchronize()
print("ok")
