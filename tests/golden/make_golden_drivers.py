"""Generates tests/golden/drivers_64x48.npz by running the REFERENCE's own Python drivers (byte-compiled from
/root/reference by oracle/build_lietorch.py, with the reference's vendored lietorch, CPU backend) on seeded inputs:

  * FrameTracker.opt_pose_ray_dist_sim3 / opt_pose_calib_sim3 (tracker.py:156-266)  -> T_WCf, T_CkCf
  * matching.match (matching.py:8-90) with the CPU oracle as `mast3r_slam_backends`   -> idx_1_to_2, valid
  * FactorGraph.solve_GN_rays / solve_GN_calib (global_opt.py:121-213) with the CPU oracle as backend -> poses

Run here (where /root/reference exists):  python tests/golden/make_golden_drivers.py
The fixture pins oracle/tracker_ref.py (the restatement) and the host mirrors in splatt3r_slam_b200/ on every box.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle  # noqa: E402
import refdrivers  # noqa: E402
from oracle import tracker_ref  # noqa: E402
from splatt3r_slam_b200 import synthetic  # noqa: E402

H, W = 48, 64


def main():
    ns = refdrivers.load()
    Sim3 = ns.lietorch.Sim3
    out = {}
    # ---- tracker, ray + distance ------------------------------------------------------------------------
    for seed in (1, 2):
        P = synthetic.make_track_problem(seed=seed, n=H * W)
        tr = ns.tracker.FrameTracker(None, None, "cpu")
        Tw, Tr = tr.opt_pose_ray_dist_sim3(P["Xf"], P["Xk"], Sim3(P["T_WCf"].view(1, 8)), Sim3(P["T_WCk"].view(1, 8)),
                                           P["Qk"], P["valid"])
        for k in ("Xf", "Xk", "T_WCf", "T_WCk", "Qk", "valid"):
            out[f"track{seed}_{k}"] = P[k].numpy()
        out[f"track{seed}_out_T_WCf"] = Tw.data.view(8).numpy()
        out[f"track{seed}_out_T_CkCf"] = Tr.data.view(8).numpy()
    # ---- tracker, calibrated ----------------------------------------------------------------------------
    P = synthetic.make_track_problem_calib(seed=1, h=H, w=W)
    meas_k, valid_meas_k = tracker_ref.calib_measurements(P["Xk"], P["img_size"])
    tr = ns.tracker.FrameTracker(None, None, "cpu")
    # the reference builds meas_k itself in get_points_poses (tracker.py:141-152); cross-check the restatement's
    uv = ns.geometry.get_pixel_coords(1, P["img_size"], device="cpu", dtype=torch.float32).view(-1, 2)
    mk = torch.cat((uv, torch.log(P["Xk"][..., 2:3])), dim=-1)
    vk = P["Xk"][..., 2:3] > 1e-6
    mk[~vk.repeat(1, 3)] = 0.0
    assert torch.equal(mk, meas_k) and torch.equal(vk, valid_meas_k)
    Tw, Tr = tr.opt_pose_calib_sim3(P["Xf"], P["Xk"], Sim3(P["T_WCf"].view(1, 8)), Sim3(P["T_WCk"].view(1, 8)),
                                    P["Qk"], P["valid"], meas_k, valid_meas_k, P["K"], P["img_size"])
    for k in ("Xf", "Xk", "T_WCf", "T_WCk", "Qk", "valid", "K"):
        out[f"calib_{k}"] = P[k].numpy()
    out["calib_out_T_WCf"] = Tw.data.view(8).numpy()
    out["calib_out_T_CkCf"] = Tr.data.view(8).numpy()
    # ---- matching driver on the oracle backend ----------------------------------------------------------------
    refdrivers.bind_backend(ns, oracle)
    X11, X21, D11, D21 = synthetic.make_pair(seed=5, h=H, w=W, fdim=24, b=2)
    idx, valid = ns.matching.match(X11, X21, D11, D21)
    idx2, valid2 = ns.matching.match(X11, X21, D11, D21, idx_1_to_2_init=idx)
    out.update(match_X11=X11.numpy(), match_X21=X21.numpy(), match_D11=D11.numpy(), match_D21=D21.numpy(),
               match_idx=idx.numpy(), match_valid=valid.numpy(), match_idx2=idx2.numpy(), match_valid2=valid2.numpy())
    # ---- global optimisation driver on the oracle backend ---------------------------------------------------------
    g = synthetic.make_graph(seed=4, N=4, E_undirected=5, h=H, w=W)
    Eu = g["ii"].numel() // 2
    for name in ("rays", "calib"):
        store = refdrivers.KeyframeStore(g["Xs"].clone(), g["Cs"], g["Twc"], H, W)
        fg = ns.global_opt.FactorGraph(None, store, K=g["K"], device="cpu")
        # the synthetic graph lists both directions: first half = (ii -> jj), second half = the reverse edges
        fg.ii, fg.jj = g["ii"][:Eu], g["jj"][:Eu]
        fg.idx_ii2jj, fg.idx_jj2ii = g["idx_ii2jj"][:Eu], g["idx_ii2jj"][Eu:]
        fg.valid_match_j, fg.valid_match_i = g["valid_match"][:Eu], g["valid_match"][Eu:]
        fg.Q_ii2jj, fg.Q_jj2ii = g["Q"][:Eu], g["Q"][Eu:]
        (fg.solve_GN_rays if name == "rays" else fg.solve_GN_calib)()
        out[f"gopt_{name}_T"] = store.T_WC.view(-1, 8).numpy().copy()
    for k in ("Twc", "Xs", "Cs", "K", "ii", "jj", "idx_ii2jj", "valid_match", "Q"):
        out[f"gopt_{k}"] = g[k].numpy()
    path = os.path.join(HERE, "drivers_64x48.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes;", len(out), "arrays")


if __name__ == "__main__":
    main()
