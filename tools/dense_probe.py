"""Stage-3 solver in isolation on synthetic SPD edge blocks: time of b200slam_gn_solve_retract and the
cycle split the dense kernel leaves in the workspace head.  python tools/dense_probe.py N E [N E ...]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import splatt3r_slam_b200.mast3r_slam_backends as be  # noqa: E402
from splatt3r_slam_b200.dist_gn import CudaHooks  # noqa: E402

dev = torch.device("cuda:0")
args = [int(a) for a in sys.argv[1:]] or [256, 2048, 1024, 8192]
for N, E in zip(args[::2], args[1::2]):
    gen = torch.Generator().manual_seed(N + E)
    ii = torch.arange(E // 2) % N
    jj = (ii + 1 + (torch.randint(0, N - 2, (E // 2,), generator=gen) * (torch.arange(E // 2) >= N))) % N
    ii, jj = torch.cat((ii, jj)).to(dev), torch.cat((jj, ii)).to(dev)
    R = torch.randn(E, 7, 7, generator=gen, dtype=torch.float64)
    Bm = R @ R.transpose(1, 2) + 7.0 * torch.eye(7, dtype=torch.float64)
    eb = torch.cat((Bm.reshape(E, 49), torch.randn(E, 7, generator=gen, dtype=torch.float64)), dim=1).contiguous().to(dev)
    prm = be.make_params(0, 0.003, 10.0, 0.0, 1.5)
    ii_row, jj_row = be.edge_rows(ii, jj)
    hooks = CudaHooks(prm, N, E, 64, dev)
    T0 = torch.zeros(N, 8, device=dev)
    T0[:, 6] = 1.0
    T0[:, 7] = 1.0
    T = T0.clone()
    dx = torch.zeros((N - 1, 7), dtype=torch.float32, device=dev)
    for _ in range(2):
        T.copy_(T0)
        hooks.solve_retract(T, ii_row, jj_row, eb, 0.0, dx)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    a.record()
    for _ in range(reps):
        T.copy_(T0)
        hooks.solve_retract(T, ii_row, jj_row, eb, 0.0, dx)
    b.record()
    torch.cuda.synchronize()
    st = hooks.ws[:48].cpu().view(torch.int64)[2:6].tolist()
    print(json.dumps({"N": N, "E": E, "D": 7 * (N - 1), "solve_ms": a.elapsed_time(b) / reps, "chol_fail": hooks.done()["chol_fail"],
                      "dx_absmax": float(dx.abs().max()), "cycles": {"diag_cta0": st[0], "panel": st[1], "trail": st[2],
                                                                    "kernel": st[3]}}))
