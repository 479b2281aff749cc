"""Debug aid: time the fused GN path and print the tail's %globaltimer stamps."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import make_sets, GN
import splatt3r_slam_b200.mast3r_slam_backends as be
dev = torch.device("cuda:0")
s = {k: v.to(dev) for k, v in make_sets(1, 100)[0].items()}
def run(iters):
    T = s["Twc"].clone()
    be.gauss_newton_rays(T, s["Xs"], s["Cs"], s["ii"], s["jj"], s["idx_ii2jj"], s["valid_match"], s["Q"],
                         GN["sigma_ray"], GN["sigma_dist"], GN["C_thresh"], GN["Q_thresh"], iters, 0.0)
for _ in range(3): run(10)
torch.cuda.synchronize()
for iters in (1, 10):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20): run(iters)
    b.record(); torch.cuda.synchronize()
    print(f"iters={iters}: {a.elapsed_time(b)/20*1e3:.1f} us per call")
ws = be._workspaces[("cuda", 0)]
st = ws[:48].cpu().view(torch.int64)
t = st[2:6].tolist()
print("tail ns: finalize", t[1]-t[0], "assemble+chol", t[2]-t[1], "finish", t[3]-t[2])
