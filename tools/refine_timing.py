"""refine_matches / fused match device time for the library selected by B200SLAM_LIB (A/B aid); prints JSON."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import N_PIX, make_sets  # noqa: E402
from splatt3r_slam_b200 import matching  # noqa: E402
import splatt3r_slam_b200.mast3r_slam_backends as be  # noqa: E402

dev = torch.device("cuda:0")
cfg = matching.DEFAULT_CFG


def timeit(fn, n_sets, reps=60):
    for i in range(5):
        fn(i % n_sets)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(reps):
        fn(i % n_sets)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3


n_sets = 6
sets = []
for s in make_sets(n_sets, 100):
    d = {k: v.to(dev) for k, v in s.items() if k in ("X11", "X21", "D11", "D21")}
    rays, pts, p_init = matching.prep_for_iter_proj(d["X11"], d["X21"], None)
    d.update(D11h=d["D11"].half().contiguous(), D21h=d["D21"].view(1, N_PIX, -1).half().contiguous())
    p_new, _ = be.iter_proj(rays, pts, p_init, cfg["max_iter"], cfg["lambda_init"], cfg["convergence_thresh"])
    d["p1"] = p_new.long()
    sets.append(d)
out = {"lib": os.environ.get("B200SLAM_LIB", "default")}
out["refine_us"] = timeit(lambda i: be.refine_matches(sets[i]["D11h"], sets[i]["D21h"], sets[i]["p1"], cfg["radius"],
                                                      cfg["dilation_max"]), n_sets)
out["match_fused_us"] = timeit(lambda i: matching.match(sets[i]["X11"], sets[i]["X21"], sets[i]["D11h"], sets[i]["D21h"],
                                                        fused=True), n_sets)
# checksum of the refined pixels: identical across builds when the result is bit-identical
(p,) = be.refine_matches(sets[0]["D11h"], sets[0]["D21h"], sets[0]["p1"], cfg["radius"], cfg["dilation_max"])
out["checksum"] = int((p.view(-1) * torch.arange(1, p.numel() + 1, device=dev)).sum().item())
if hasattr(be._lib, "b200slam_internal_refine_stats"):  # -DRT_STATS=1 builds: counters of ONE call on set 0
    import ctypes
    buf = (ctypes.c_ulonglong * 32)()
    be._lib.b200slam_internal_refine_stats(buf)
    be.refine_matches(sets[0]["D11h"], sets[0]["D21h"], sets[0]["p1"], cfg["radius"], cfg["dilation_max"])
    torch.cuda.synchronize()
    be._lib.b200slam_internal_refine_stats(buf)
    out["stats_by_dilation"] = {d: {"live": buf[4 * d], "out_of_box": buf[4 * d + 1], "tiles_overflow": buf[4 * d + 2],
                                    "tiles_merged": buf[4 * d + 3]} for d in range(1, 8) if buf[4 * d] or buf[4 * d + 3]}
print(json.dumps(out))
