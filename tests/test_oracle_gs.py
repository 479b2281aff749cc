"""CPU tier: the numpy oracle of the 3DGS forward pass (oracle/gs_ref.py) against the golden vectors produced by the
reference's own CUDA build (tests/golden/make_golden_gs.py), and its basic invariants.  Same floating-point bar as
tests/test_gpu_gs.py."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import gs_ref
from splatt3r_slam_b200 import synthetic


def run(S, use_cov, use_sh, deg=3):
    n = lambda k: S[k].numpy()
    return gs_ref.rasterize(n("means"), n("cov") if use_cov else None, None if use_cov else n("scales"),
                            None if use_cov else n("rotations"), 1.0, n("opacities"), None if use_sh else n("colors"),
                            n("shs") if use_sh else None, deg, n("campos"), n("view"), n("proj"), S["tanfovx"], S["tanfovy"],
                            n("bg"), S["H"], S["W"])


def test_oracle_vs_reference_golden():
    path = os.path.join(GOLDEN, "gs_96x128.npz")
    if not os.path.exists(path):
        pytest.skip("golden fixture not generated yet")
    g = np.load(path)
    S = synthetic.make_splat_scene(seed=5, P=3000, H=96, W=128)
    for (use_cov, use_sh, tag) in ((True, False, "cov_rgb"), (False, True, "sr_sh3")):
        c, r = run(S, use_cov, use_sh)
        gr, gc = g["radii_" + tag].astype(np.int64), g["color_" + tag].astype(np.float64)
        assert (r == gr).mean() >= 0.999 and np.abs(r - gr).max() <= 1
        d = np.abs(c - gc)
        assert d.max() < 6e-3 and (d < 2e-5).mean() >= 0.999, (d.max(), (d < 2e-5).mean())


def test_oracle_invariants():
    S = synthetic.make_splat_scene(seed=1, P=300, H=32, W=48)
    c, r = run(S, True, False)
    assert c.shape == (3, 32, 48) and r.shape == (300,) and (r >= 0).all() and (r[:15] == 0).all()  # first 5 %: behind the camera
    # the two parameterisations of the covariance describe the same scene
    c2, r2 = run(S, False, False)
    assert np.abs(c - c2).max() < 6e-3 and (r == r2).mean() > 0.99
    # zero opacity -> background
    S0 = dict(S); S0["opacities"] = S["opacities"] * 0
    c0, _ = run(S0, True, False)
    assert np.allclose(c0, S["bg"].numpy()[:, None, None])


def test_rasterizer_mirror_argument_checks_without_a_gpu():
    """The mirror of the reference's GaussianRasterizer raises the reference's own exceptions for bad argument
    combinations (diff_gaussian_rasterization/__init__.py:190-194) and refuses CPU tensors (no CPU path)."""
    import torch
    from splatt3r_slam_b200.gaussian_rasterizer import GaussianRasterizationSettings, GaussianRasterizer
    S = synthetic.make_splat_scene(seed=1, P=16, H=16, W=16)
    rs = GaussianRasterizationSettings(16, 16, S["tanfovx"], S["tanfovy"], S["bg"], 1.0, S["view"], S["proj"], 0, S["campos"],
                                       False, False)
    R = GaussianRasterizer(rs)
    with pytest.raises(Exception, match="excatly one of either SHs"):
        R(S["means"], None, S["opacities"], shs=S["shs"], colors_precomp=S["colors"], cov3D_precomp=S["cov"])
    with pytest.raises(Exception, match="exactly one of either scale/rotation"):
        R(S["means"], None, S["opacities"], colors_precomp=S["colors"])
    with pytest.raises(RuntimeError, match="no CPU path"):
        R(S["means"], None, S["opacities"], colors_precomp=S["colors"], cov3D_precomp=S["cov"])
    assert torch.is_tensor(S["cov"]) and S["cov"].shape == (16, 6)
