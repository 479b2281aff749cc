"""CPU tier: tests/golden/drivers_64x48.npz — outputs of the reference's OWN Python drivers (tracker.py,
matching.py, global_opt.py run unchanged with the reference's lietorch; tests/golden/make_golden_drivers.py) —
pin the test infrastructure that stands in for them where they cannot travel:

  * oracle/tracker_ref.py (restatement of FrameTracker.opt_pose_*_sim3)      vs the reference tracker's poses
  * splatt3r_slam_b200/matching.py (mirror of matching.py) on the CPU oracle  vs the reference driver's indices
and, when oracle/_ref holds the byte-compiled drivers (this container, and any box the build travelled to),
the same comparisons live on fresh seeds.  Tolerance for poses: 1e-5 absolute on Sim3 data of magnitude <= 1
(measured 7e-7 .. 1.7e-6; both sides are fp32 with an fp32 Cholesky)."""
import os

import numpy as np
import pytest
import torch

import oracle
import refdrivers
from conftest import GOLDEN
from oracle import tracker_ref
from splatt3r_slam_b200 import matching, synthetic

TOL = 1e-5


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLDEN, "drivers_64x48.npz"))


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a))


@pytest.mark.parametrize("seed", [1, 2])
def test_tracker_restatement_vs_reference_fixture(gold, seed):
    P = {k: T(gold[f"track{seed}_{k}"]) for k in ("Xf", "Xk", "T_WCf", "T_WCk", "Qk", "valid")}
    Tw, Tr, steps = tracker_ref.opt_pose_ray_dist_sim3(P["Xf"], P["Xk"], P["T_WCf"], P["T_WCk"], P["Qk"], P["valid"])
    assert (Tw.view(-1) - T(gold[f"track{seed}_out_T_WCf"])).abs().max() < TOL
    assert (Tr.view(-1) - T(gold[f"track{seed}_out_T_CkCf"])).abs().max() < TOL
    assert 1 <= steps < 50


def test_tracker_calib_restatement_vs_reference_fixture(gold):
    P = {k: T(gold[f"calib_{k}"]) for k in ("Xf", "Xk", "T_WCf", "T_WCk", "Qk", "valid", "K")}
    img_size = (48, 64)
    meas_k, valid_meas_k = tracker_ref.calib_measurements(P["Xk"], img_size)
    Tw, Tr, steps = tracker_ref.opt_pose_calib_sim3(P["Xf"], P["Xk"], P["T_WCf"], P["T_WCk"], P["Qk"], P["valid"], meas_k,
                                                    valid_meas_k, P["K"], img_size)
    assert (Tw.view(-1) - T(gold["calib_out_T_WCf"])).abs().max() < TOL
    assert (Tr.view(-1) - T(gold["calib_out_T_CkCf"])).abs().max() < TOL


def test_matching_mirror_vs_reference_driver_fixture(gold):
    X11, X21, D11, D21 = (T(gold[f"match_{k}"]) for k in ("X11", "X21", "D11", "D21"))
    idx, valid = matching.match(X11, X21, D11, D21, backend=oracle)
    assert torch.equal(idx, T(gold["match_idx"])) and torch.equal(valid, T(gold["match_valid"]))
    idx2, valid2 = matching.match(X11, X21, D11, D21, idx_1_to_2_init=idx, backend=oracle)
    assert torch.equal(idx2, T(gold["match_idx2"])) and torch.equal(valid2, T(gold["match_valid2"]))


def test_global_opt_fixture_is_the_oracle_loop_on_the_drivers_inputs(gold):
    """The reference FactorGraph.solve_GN_* (prep_two_way_edges, get_poses_points, constrain_points_to_ray, pin)
    fed the oracle: the same numbers must come out of calling the oracle on the synthetic two-way arrays."""
    g = {k: T(gold[f"gopt_{k}"]) for k in ("Twc", "Xs", "Cs", "K", "ii", "jj", "idx_ii2jj", "valid_match", "Q")}
    Tw = g["Twc"].clone()
    oracle.gauss_newton_rays(Tw, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"], 0.003, 10.0,
                             0.0, 1.5, 10, 1e-8)
    assert torch.equal(Tw, T(gold["gopt_rays_T"]))
    assert (Tw - g["Twc"]).abs().max() > 1e-4  # the solve moved the poses


needs_drivers = pytest.mark.skipif(not refdrivers.available(), reason="oracle/_ref has no byte-compiled reference drivers")


@needs_drivers
@pytest.mark.parametrize("seed", [11, 12, 13])
def test_live_reference_tracker_vs_restatement(seed):
    ns = refdrivers.load()
    Sim3 = ns.lietorch.Sim3
    P = synthetic.make_track_problem(seed=seed, n=3000)
    tr = ns.tracker.FrameTracker(None, None, "cpu")
    Tw_r, Tr_r = tr.opt_pose_ray_dist_sim3(P["Xf"], P["Xk"], Sim3(P["T_WCf"].view(1, 8)), Sim3(P["T_WCk"].view(1, 8)),
                                           P["Qk"], P["valid"])
    Tw, Tr, _ = tracker_ref.opt_pose_ray_dist_sim3(P["Xf"], P["Xk"], P["T_WCf"], P["T_WCk"], P["Qk"], P["valid"])
    assert (Tw.view(-1) - Tw_r.data.view(-1)).abs().max() < TOL and (Tr.view(-1) - Tr_r.data.view(-1)).abs().max() < TOL


@needs_drivers
def test_live_reference_matching_driver_on_the_oracle_equals_mirror():
    ns = refdrivers.load()
    refdrivers.bind_backend(ns, oracle)
    X11, X21, D11, D21 = synthetic.make_pair(seed=21, h=40, w=56, fdim=24, b=1)
    idx_r, val_r = ns.matching.match(X11, X21, D11, D21)
    idx_m, val_m = matching.match(X11, X21, D11, D21, backend=oracle)
    assert torch.equal(idx_r, idx_m) and torch.equal(val_r, val_m)
