"""CPU tier: the C oracle against the golden fixtures produced by the reference CUDA build
(tests/golden/make_golden.py, run on a B200).  Bit-exact for matching; 1e-5 relative for GN."""
import os

import numpy as np
import pytest
import torch

import oracle
from conftest import GOLDEN

CFG = dict(max_iter=10, lambda_init=1e-8, convergence_thresh=1e-6, radius=3, dilation_max=5)
GN = dict(sigma_ray=0.003, sigma_dist=10.0, sigma_pixel=1.0, sigma_depth=10.0, sigma_point=0.05,
          C_thresh=0.0, Q_thresh=1.5, pixel_border=-10, z_eps=1e-6)
H, W = 48, 64
RTOL = 1e-5  # BASELINE.json north_star: Hessians / pose updates within 1e-5 relative (fp32)


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a))


def rel(a, b):
    return ((a - b).abs().max() / b.abs().max()).item()


@pytest.fixture(scope="module")
def gm():
    return np.load(os.path.join(GOLDEN, "matching_64x48.npz"))


@pytest.fixture(scope="module")
def gg():
    return np.load(os.path.join(GOLDEN, "gn_3kf_64x48.npz"))


def test_iter_proj_bit_exact(gm):
    p_new, conv = oracle.iter_proj(T(gm["rays"]), T(gm["pts"]), T(gm["p_init"]), CFG["max_iter"],
                                   CFG["lambda_init"], CFG["convergence_thresh"])
    assert np.array_equal(p_new.numpy().view(np.uint32), gm["p_new"].view(np.uint32))
    assert np.array_equal(conv.numpy(), gm["converged"])
    assert 0.2 < conv.float().mean() < 1.0  # the fixture exercises both outcomes


def test_iter_proj_zero_iterations_clamps_only(gm):
    p_new, conv = oracle.iter_proj(T(gm["rays"]), T(gm["pts"]), T(gm["p_init"]), 0, 1e-8, 1e-6)
    p = T(gm["p_init"])
    exp = torch.stack([p[..., 0].clamp(1, W - 2), p[..., 1].clamp(1, H - 2)], -1)
    assert torch.equal(p_new, exp) and not conv.any()


def test_refine_f16_bit_exact(gm):
    (p1,) = oracle.refine_matches(T(gm["D11"]), T(gm["D21"]), T(gm["p1"]), CFG["radius"], CFG["dilation_max"])
    assert np.array_equal(p1.numpy(), gm["p1_refined"])
    assert (p1.numpy() != gm["p1"]).any()


def test_refine_f32_bit_exact(gm):
    (p1,) = oracle.refine_matches(T(gm["D11"]).float(), T(gm["D21"]).float(), T(gm["p1"]), CFG["radius"],
                                  CFG["dilation_max"])
    assert np.array_equal(p1.numpy(), gm["p1_refined_f32"])


def test_refine_radius_zero_is_identity(gm):
    (p1,) = oracle.refine_matches(T(gm["D11"]), T(gm["D21"]), T(gm["p1"]), 3, 0)
    assert np.array_equal(p1.numpy(), gm["p1"])


def _graph(gg):
    return [T(gg[k]) for k in ("Twc", "Xs", "Cs", "K", "ii", "jj", "idx_ii2jj", "valid_match", "Q")]


@pytest.mark.parametrize("name,kind", [("rays", 0), ("calib", 1), ("points", 2)])
def test_gn_blocks_vs_reference(gg, name, kind):
    Twc, Xs, Cs, K, ii, jj, idx, valid, Q = _graph(gg)
    sa, sb = {"rays": (GN["sigma_ray"], GN["sigma_dist"]), "calib": (GN["sigma_pixel"], GN["sigma_depth"]),
              "points": (GN["sigma_point"], 1.0)}[name]
    Hs, gs = oracle.gn_build(kind, Twc, Xs, Cs, K if kind == 1 else None, ii, jj, idx, valid, Q, sa, sb,
                             GN["C_thresh"], GN["Q_thresh"], H, W, GN["pixel_border"], GN["z_eps"])
    Hr, gr = T(gg[f"{name}_Hs"]), T(gg[f"{name}_gs"])
    for blk in range(4):
        assert rel(Hs[blk], Hr[blk]) < RTOL
    for blk in range(2):
        assert rel(gs[blk], gr[blk]) < RTOL
    # structure the B200 kernels rely on: H_ii = H_jj = -H_ij = -H_ji^T, g_i = -g_j (gn_kernels.cu:998-999)
    assert rel(Hr[0], Hr[3]) < 1e-6 and rel(Hr[1], -Hr[3]) < 1e-5 and rel(Hr[2], -Hr[3]) < 1e-5
    assert rel(gr[0], -gr[1]) < 1e-6


@pytest.mark.parametrize("iters", [1, 4])
def test_gn_loop_vs_reference(gg, iters):
    Twc, Xs, Cs, K, ii, jj, idx, valid, Q = _graph(gg)
    T1 = Twc.clone()
    (dx,) = oracle.gauss_newton_rays(T1, Xs, Cs, ii, jj, idx, valid, Q, GN["sigma_ray"], GN["sigma_dist"],
                                     GN["C_thresh"], GN["Q_thresh"], iters, 0.0)
    assert rel(T1, T(gg[f"rays_T{iters}"])) < RTOL
    T2 = Twc.clone()
    oracle.gauss_newton_calib(T2, Xs, Cs, K, ii, jj, idx, valid, Q, H, W, GN["pixel_border"], GN["z_eps"],
                              GN["sigma_pixel"], GN["sigma_depth"], GN["C_thresh"], GN["Q_thresh"], iters, 0.0)
    assert rel(T2, T(gg[f"calib_T{iters}"])) < RTOL
    T3 = Twc.clone()
    oracle.gauss_newton_points(T3, Xs, Cs, ii, jj, idx, valid, Q, GN["sigma_point"], GN["C_thresh"],
                               GN["Q_thresh"], iters, 0.0)
    assert rel(T3, T(gg[f"points_T{iters}"])) < RTOL
    if iters == 1:
        assert rel(dx, T(gg["rays_dx1"])) < 1e-4  # dx itself is small; poses carry the contract
        assert torch.equal(T1[0], Twc[0])  # pose 0 is the gauge (num_fix = 1)


def test_gn_early_exit_and_failed_factorisation(gg):
    Twc, Xs, Cs, K, ii, jj, idx, valid, Q = _graph(gg)
    T1 = Twc.clone()
    oracle.gauss_newton_rays(T1, Xs, Cs, ii, jj, idx, valid, Q, 0.003, 10.0, 0.0, 1.5, 10, 1e9)
    assert oracle._gauss_newton.last_iters == 1  # ||dx|| < huge threshold after the first update
    # no valid residual at all -> H = 0 -> not SPD -> dx = 0, poses untouched (gn_kernels.cu:141-149)
    T2 = Twc.clone()
    (dx,) = oracle.gauss_newton_rays(T2, Xs, Cs, ii, jj, idx, torch.zeros_like(valid), Q, 0.003, 10.0, 0.0,
                                     1.5, 3, 1e-8)
    assert torch.equal(T2, Twc) and not dx.any() and oracle._gauss_newton.last_iters == 1


# ---- tracker restatement (oracle/tracker_ref.py; parity unpinned, see its header) -------------------
def test_tracker_restatement_recovers_known_pose():
    from oracle import tracker_ref as tr
    from splatt3r_slam_b200 import synthetic
    P = synthetic.make_track_problem(seed=1, n=3000)
    T_WCf, T_CkCf, steps = tr.opt_pose_ray_dist_sim3(P["Xf"], P["Xk"], P["T_WCf"], P["T_WCk"], P["Qk"], P["valid"])
    assert 2 <= steps < 50
    assert (T_CkCf - P["T_CkCf_gt"]).abs().max() < 5e-3      # scale and depth translation are weakly observed (sigma_dist = 10)
    assert torch.allclose(T_WCf, tr.sim3_mul(P["T_WCk"], T_CkCf))
    # Sim3 helpers against the pinned C oracle: Exp(0) retraction is the identity, inv * T = identity
    ident = tr.sim3_mul(tr.sim3_inv(P["T_WCk"]), P["T_WCk"])
    assert torch.allclose(ident, torch.tensor([0, 0, 0, 0, 0, 0, 1.0, 1.0]), atol=1e-6)
    assert torch.allclose(tr.sim3_retr(P["T_WCk"], torch.zeros(7)), P["T_WCk"], atol=1e-7)


def test_tracker_restatement_raises_on_degenerate_system():
    from oracle import tracker_ref as tr
    from splatt3r_slam_b200 import synthetic
    P = synthetic.make_track_problem(seed=2, n=64)
    with pytest.raises(RuntimeError):  # no valid point: H = 0, torch.linalg.cholesky raises (tracker.py:88-90)
        tr.opt_pose_ray_dist_sim3(P["Xf"], P["Xk"], P["T_WCf"], P["T_WCk"], P["Qk"], torch.zeros_like(P["valid"]))
