import os, sys, json, torch
sys.path.insert(0, "/root/repo")
from bench import make_sets, GN
import splatt3r_slam_b200.mast3r_slam_backends as be
dev = torch.device("cuda:0")
sets = [{k: v.to(dev) for k, v in s.items()} for s in make_sets(6, 100)]
for d in sets: d["T"] = d["Twc"].clone()
def timeit(fn, reps=40):
    for i in range(5): fn(i % 6)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(reps): fn(i % 6)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3
out = {}
for iters in (1, 2, 5, 10, 20, 40):
    def gn(i, iters=iters):
        d = sets[i]
        d["T"].copy_(d["Twc"])
        be.gauss_newton_rays(d["T"], d["Xs"], d["Cs"], d["ii"], d["jj"], d["idx_ii2jj"], d["valid_match"], d["Q"],
                             GN["sigma_ray"], GN["sigma_dist"], GN["C_thresh"], GN["Q_thresh"], iters, 0.0)
    out[iters] = round(timeit(gn), 1)
print(json.dumps(out))
# same, 4 calls queued back to back per timing unit -> host overhead amortised?
