"""Runs each hot kernel a few times on the bench workload (for ncu captures)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import make_sets, GN, N_PIX
from splatt3r_slam_b200 import matching
import splatt3r_slam_b200.mast3r_slam_backends as be
dev = torch.device("cuda:0")
cfg = matching.DEFAULT_CFG
s = {k: v.to(dev) for k, v in make_sets(1, 100)[0].items()}
rays, pts, p_init = matching.prep_for_iter_proj(s["X11"], s["X21"], None)
D11h, D21h = s["D11"].half().contiguous(), s["D21"].view(1, N_PIX, -1).half().contiguous()
for _ in range(3):
    p_new, conv = be.iter_proj(rays, pts, p_init, cfg["max_iter"], cfg["lambda_init"], cfg["convergence_thresh"])
    p1 = p_new.long()
    be.refine_matches(D11h, D21h, p1, cfg["radius"], cfg["dilation_max"])
    T = s["Twc"].clone()
    be.gauss_newton_rays(T, s["Xs"], s["Cs"], s["ii"], s["jj"], s["idx_ii2jj"], s["valid_match"], s["Q"],
                         GN["sigma_ray"], GN["sigma_dist"], GN["C_thresh"], GN["Q_thresh"], 3, 0.0)
    # the two entries beyond the reference module
    matching.match(s["X11"], s["X21"], D11h, D21h, fused=True)
from splatt3r_slam_b200 import synthetic, tracker
TP = {k: v.to(dev) for k, v in synthetic.make_track_problem(seed=3, n=N_PIX).items()}
for _ in range(2):
    tracker.opt_pose_ray_dist_sim3(TP["Xf"], TP["Xk"], TP["T_WCf"], TP["T_WCk"], TP["Qk"], TP["valid"])
torch.cuda.synchronize()
print("ok")
