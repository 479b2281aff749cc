"""GPU tier — SURVEY.md 8 (a10): the reference's OWN drivers, unchanged, on top of this repo's
`mast3r_slam_backends` shim (byte-compiled from /root/reference into oracle/_ref by oracle/build_lietorch.py; the
tests skip where that build did not travel), compared with the same drivers on the reference's CUDA build:

  * matching.match (matching.py:8-90)                     shim == reference build, torch.equal, at 384x512;
                                                          the fused entry b200slam_match gives the same bits
  * FrameTracker.opt_pose_ray_dist_sim3 / _calib_sim3      reference (lietorch + torch on CUDA) vs the on-device
    (tracker.py:156-266)                                   tracker kernel, and vs the CPU fixtures
  * FactorGraph.solve_GN_rays / solve_GN_calib             shim vs reference build (global_opt.py:121-213), and
                                                          the zero-copy two-way mirror vs the driver fixture
Pose tolerance 1e-5 relative (north star); where the reference's own fp32 arithmetic is looser it is said so."""
import os
from types import SimpleNamespace

import numpy as np
import pytest
import torch

import refdrivers
from conftest import GOLDEN
from splatt3r_slam_b200 import global_opt, matching, synthetic, tracker

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
needs_drivers = pytest.mark.skipif(not refdrivers.available(), reason="oracle/_ref has no byte-compiled reference drivers")


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLDEN, "drivers_64x48.npz"))


# ---------------------------------------------------------------------------------------------- fixtures (always)
@pytest.mark.parametrize("seed", [1, 2])
def test_track_kernel_vs_reference_tracker_fixture(backend, gold, seed):
    P = {k: T(gold[f"track{seed}_{k}"]) for k in ("Xf", "Xk", "T_WCf", "T_WCk", "Qk", "valid")}
    Tw, Tr = tracker.opt_pose_ray_dist_sim3(P["Xf"], P["Xk"], P["T_WCf"], P["T_WCk"], P["Qk"], P["valid"])
    assert (Tw - T(gold[f"track{seed}_out_T_WCf"])).abs().max() < 1e-5
    assert (Tr - T(gold[f"track{seed}_out_T_CkCf"])).abs().max() < 1e-5


def test_track_calib_kernel_vs_reference_tracker_fixture(backend, gold):
    from oracle import tracker_ref
    P = {k: T(gold[f"calib_{k}"]) for k in ("Xf", "Xk", "T_WCf", "T_WCk", "Qk", "valid", "K")}
    meas_k, valid_meas_k = tracker_ref.calib_measurements(P["Xk"].cpu(), (48, 64))
    Tw, Tr = tracker.opt_pose_calib_sim3(P["Xf"], P["Xk"], P["T_WCf"], P["T_WCk"], P["Qk"], P["valid"], meas_k.to(DEV),
                                         valid_meas_k.to(DEV), P["K"], (48, 64))
    assert (Tw - T(gold["calib_out_T_WCf"])).abs().max() < 1e-5
    assert (Tr - T(gold["calib_out_T_CkCf"])).abs().max() < 1e-5


def test_match_vs_reference_driver_fixture(backend, gold):
    """The fixture's indices come from the reference driver on CPU torch glue + the C oracle; on the GPU the glue is
    torch's CUDA kernels, whose normalize / conv rounding differs in the last bit for a few pixels — the bit-exact
    comparison is the live one below (same CUDA glue on both sides); here at most 1e-3 of the pixels may differ."""
    X11, X21, D11, D21 = (T(gold[f"match_{k}"]) for k in ("X11", "X21", "D11", "D21"))
    idx, valid = matching.match(X11, X21, D11, D21, backend=backend)
    n = idx.numel()
    assert (idx != T(gold["match_idx"])).sum().item() <= 1e-3 * n
    assert (valid != T(gold["match_valid"])).sum().item() <= 1e-3 * n


@pytest.mark.parametrize("name", ["rays", "calib"])
def test_global_opt_mirror_vs_reference_driver_fixture(backend, gold, name):
    g = {k: T(gold[f"gopt_{k}"]) for k in ("Twc", "Xs", "Cs", "K", "ii", "jj", "idx_ii2jj", "valid_match", "Q")}
    Eu = g["ii"].numel() // 2
    c = lambda t: t.contiguous()
    graph = SimpleNamespace(ii=c(g["ii"][:Eu]), jj=c(g["jj"][:Eu]), idx_ii2jj=c(g["idx_ii2jj"][:Eu]),
                            idx_jj2ii=c(g["idx_ii2jj"][Eu:]), valid_match_j=c(g["valid_match"][:Eu]),
                            valid_match_i=c(g["valid_match"][Eu:]), Q_ii2jj=c(g["Q"][:Eu]), Q_jj2ii=c(g["Q"][Eu:]))
    T_buf = g["Twc"].clone()
    if name == "rays":
        global_opt.solve_GN_rays(graph, g["Xs"], g["Cs"], T_buf, backend=backend)
    else:
        uv = torch.stack(torch.meshgrid(torch.arange(64), torch.arange(48), indexing="xy"), -1).reshape(1, -1, 2).to(DEV).float()
        K = g["K"]  # constrain_points_to_ray (geometry.py:36-42, global_opt.py:172)
        z = g["Xs"][..., 2:3]
        Xc = torch.cat((((uv[..., 0:1] - K[0, 2]) / K[0, 0]) * z, ((uv[..., 1:2] - K[1, 2]) / K[1, 1]) * z, z), -1).contiguous()
        global_opt.solve_GN_calib(graph, Xc, g["Cs"], T_buf, K, (48, 64), backend=backend)
    ref = T(gold[f"gopt_{name}_T"])
    assert ((T_buf - ref).abs().max() / ref.abs().max()).item() < 1e-5


# ------------------------------------------------------------------------------- live reference drivers (when built)
@needs_drivers
@pytest.mark.parametrize("h,w,b,seed", [(96, 128, 2, 0), (384, 512, 1, 1)])
def test_reference_matching_driver_shim_equals_reference_build(backend, ref_backend, h, w, b, seed):
    """matching.py of the reference, unchanged, once on this repo's shim and once on the reference's CUDA build: the
    two must agree bit for bit (same torch CUDA glue on both sides); so must the fused entry."""
    ns = refdrivers.load()
    X11, X21, D11, D21 = (t.to(DEV) for t in synthetic.make_pair(seed=seed, h=h, w=w, fdim=24, b=b))
    refdrivers.bind_backend(ns, backend)
    idx_s, val_s = ns.matching.match(X11, X21, D11, D21)
    idx_s2, val_s2 = ns.matching.match(X11, X21, D11, D21, idx_1_to_2_init=idx_s)
    idx_f, val_f = matching.match(X11, X21, D11, D21, backend=backend, fused=True)
    idx_f2, val_f2 = matching.match(X11, X21, D11, D21, idx_1_to_2_init=idx_s, backend=backend, fused=True)
    idx_m, val_m = matching.match(X11, X21, D11, D21, backend=backend)
    assert torch.equal(idx_m, idx_s) and torch.equal(val_m, val_s)        # the mirror IS the reference driver
    assert torch.equal(idx_f, idx_s) and torch.equal(val_f, val_s)        # fused entry, cold start
    assert torch.equal(idx_f2, idx_s2) and torch.equal(val_f2, val_s2)    # fused entry, warm start
    if ref_backend is None:
        pytest.skip("reference CUDA build not on this box: shim-vs-mirror-vs-fused checked only")
    refdrivers.bind_backend(ns, ref_backend)
    idx_r, val_r = ns.matching.match(X11, X21, D11, D21)
    idx_r2, val_r2 = ns.matching.match(X11, X21, D11, D21, idx_1_to_2_init=idx_r)
    assert torch.equal(idx_s, idx_r) and torch.equal(val_s, val_r)
    assert torch.equal(idx_s2, idx_r2) and torch.equal(val_s2, val_r2)


@needs_drivers
@pytest.mark.parametrize("n,seed", [(3000, 4), (196608, 5)])
def test_reference_tracker_on_cuda_vs_track_kernel(backend, n, seed):
    """FrameTracker.opt_pose_ray_dist_sim3 (torch + lietorch on CUDA tensors) vs b200slam_track_ray_dist.  The
    reference accumulates A^T A over 4n rows and factors it in fp32 (condition ~1e7): its own result moves by ~1e-5
    between CPU and CUDA summation orders, so 2e-5 absolute is what the reference supports here."""
    ns = refdrivers.load()
    Sim3 = ns.lietorch.Sim3
    P = {k: v.to(DEV) for k, v in synthetic.make_track_problem(seed=seed, n=n).items()}
    tr = ns.tracker.FrameTracker(None, None, DEV)
    Tw_r, Tr_r = tr.opt_pose_ray_dist_sim3(P["Xf"], P["Xk"], Sim3(P["T_WCf"].view(1, 8)), Sim3(P["T_WCk"].view(1, 8)),
                                           P["Qk"], P["valid"])
    Tw, Tr = tracker.opt_pose_ray_dist_sim3(P["Xf"], P["Xk"], P["T_WCf"], P["T_WCk"], P["Qk"], P["valid"])
    assert (Tw - Tw_r.data.view(-1)).abs().max() < 2e-5 and (Tr - Tr_r.data.view(-1)).abs().max() < 2e-5


@needs_drivers
def test_reference_tracker_calib_on_cuda_vs_track_kernel(backend):
    from oracle import tracker_ref
    ns = refdrivers.load()
    Sim3 = ns.lietorch.Sim3
    P = synthetic.make_track_problem_calib(seed=3, h=96, w=128)
    meas_k, valid_meas_k = tracker_ref.calib_measurements(P["Xk"], P["img_size"])
    g = lambda t: t.to(DEV)
    tr = ns.tracker.FrameTracker(None, None, DEV)
    Tw_r, Tr_r = tr.opt_pose_calib_sim3(g(P["Xf"]), g(P["Xk"]), Sim3(g(P["T_WCf"]).view(1, 8)), Sim3(g(P["T_WCk"]).view(1, 8)),
                                        g(P["Qk"]), g(P["valid"]), g(meas_k), g(valid_meas_k), g(P["K"]), P["img_size"])
    Tw, Tr = tracker.opt_pose_calib_sim3(g(P["Xf"]), g(P["Xk"]), g(P["T_WCf"]), g(P["T_WCk"]), g(P["Qk"]), g(P["valid"]),
                                         g(meas_k), g(valid_meas_k), g(P["K"]), P["img_size"])
    assert (Tw - Tw_r.data.view(-1)).abs().max() < 2e-5 and (Tr - Tr_r.data.view(-1)).abs().max() < 2e-5


@needs_drivers
@pytest.mark.parametrize("name,N,Eu,h,w", [("rays", 4, 5, 96, 128), ("calib", 4, 5, 96, 128), ("rays", 8, 12, 384, 512),
                                           ("calib", 8, 12, 384, 512)])
def test_reference_factor_graph_shim_vs_reference_build(backend, ref_backend, name, N, Eu, h, w):
    """FactorGraph.solve_GN_rays / solve_GN_calib of the reference, unchanged (stack of pointmaps, two-way cat,
    constrain_points_to_ray, in-place pose view, update_T_WCs), on the shim and on the reference's CUDA build."""
    ns = refdrivers.load()
    g = synthetic.make_graph(seed=N + Eu, N=N, E_undirected=Eu, h=h, w=w)
    G = {k: (v.to(DEV) if torch.is_tensor(v) else v) for k, v in g.items()}

    def run(be_mod):
        refdrivers.bind_backend(ns, be_mod)
        store = refdrivers.KeyframeStore(G["Xs"].clone(), G["Cs"], G["Twc"], h, w)
        fg = ns.global_opt.FactorGraph(None, store, K=G["K"], device=DEV)
        fg.ii, fg.jj = G["ii"][:Eu], G["jj"][:Eu]
        fg.idx_ii2jj, fg.idx_jj2ii = G["idx_ii2jj"][:Eu], G["idx_ii2jj"][Eu:]
        fg.valid_match_j, fg.valid_match_i = G["valid_match"][:Eu], G["valid_match"][Eu:]
        fg.Q_ii2jj, fg.Q_jj2ii = G["Q"][:Eu], G["Q"][Eu:]
        (fg.solve_GN_rays if name == "rays" else fg.solve_GN_calib)()
        return store.T_WC.view(-1, 8).clone()

    T_s = run(backend)
    assert torch.equal(T_s[0], G["Twc"][0]) and (T_s - G["Twc"]).abs().max() > 1e-4  # pinned pose kept, others moved
    if ref_backend is None:
        pytest.skip("reference CUDA build not on this box")
    T_r = run(ref_backend)
    assert ((T_s - T_r).abs().max() / T_r.abs().max()).item() < 1e-5
