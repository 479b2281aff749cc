"""Tuning aid: PCG iteration counts / residuals / times on a few graph topologies (small images)."""
import ctypes, os, sys, struct
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle
from splatt3r_slam_b200 import synthetic
import splatt3r_slam_b200.mast3r_slam_backends as be
dev = torch.device("cuda:0")
for (N, Eu, h, w) in [(130, 130, 16, 24), (64, 128, 16, 24), (256, 1024, 16, 24), (1024, 4096, 8, 12)]:
    g = synthetic.make_graph(seed=N + Eu, N=N, E_undirected=Eu, h=h, w=w)
    G = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in g.items()}
    T = G["Twc"].clone()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    be.gauss_newton_rays(T, G["Xs"], G["Cs"], G["ii"], G["jj"], G["idx_ii2jj"], G["valid_match"], G["Q"], 0.003, 10.0, 0.0, 1.5, 1, 0.0)
    b.record(); torch.cuda.synchronize()
    ws = be._workspaces[("cuda", 0)]
    raw = bytes(ws[:72].cpu().numpy())
    cg_iters = struct.unpack_from("i", raw, 52)[0]; cg_rel = struct.unpack_from("d", raw, 56)[0]
    To = g["Twc"].clone()
    if N <= 256:
        oracle.gauss_newton_rays(To, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"], 0.003, 10.0, 0.0, 1.5, 1, 0.0)
        err = (T.cpu() - To).abs().max().item()
    else:
        err = float("nan")
    tt = struct.unpack_from("4q", raw, 16)
    print("   dense cycles: phase1, phase2, phase3, backward =", tt)
    print(f"N={N} Eu={Eu}: 1 GN iteration {a.elapsed_time(b)*1e3:.0f} us, cg_iters={cg_iters}, rel={cg_rel:.2e}, pose err vs oracle {err:.2e}, status {be.gn_status(dev)}")
