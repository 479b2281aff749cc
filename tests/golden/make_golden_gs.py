"""Golden vectors for the 3DGS forward rasteriser, produced by the REFERENCE's own CUDA build
(oracle/build_gs_ref.py -> oracle/_ref/diffgauss_ref.so) on a GPU box:
    python tests/golden/make_golden_gs.py gpurun_out/gs_96x128.npz      (then copy to tests/golden/)
Two calls on one synthetic scene (splatt3r_slam_b200.synthetic.make_splat_scene, seed 5): precomputed covariance +
precomputed colours (the SLAM viewer's call, visualization.py:563-596) and scales / rotations + degree-3 SH
(cuda_splatting.py:96-124)."""
import importlib.util
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from splatt3r_slam_b200 import synthetic  # noqa: E402


def load_ref():
    so = os.path.join(ROOT, "oracle", "_ref", "diffgauss_ref.so")
    spec = importlib.util.spec_from_file_location("diffgauss_ref", so)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def ref_forward(ref, S, dev, use_cov, use_sh, sh_degree=3):
    e = torch.empty(0, device=dev)
    d = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in S.items()}
    out = ref.rasterize_gaussians(d["bg"], d["means"], e if use_sh else d["colors"], d["opacities"],
                                  e if use_cov else d["scales"], e if use_cov else d["rotations"], 1.0,
                                  d["cov"] if use_cov else e, d["view"], d["proj"], S["tanfovx"], S["tanfovy"], S["H"], S["W"],
                                  d["shs"] if use_sh else e, sh_degree, d["campos"], False, False)
    return out[1], out[2]  # color, radii


if __name__ == "__main__":
    dev = torch.device("cuda:0")
    ref = load_ref()
    S = synthetic.make_splat_scene(seed=5, P=3000, H=96, W=128)
    c1, r1 = ref_forward(ref, S, dev, True, False)
    c2, r2 = ref_forward(ref, S, dev, False, True)
    np.savez_compressed(sys.argv[1], color_cov_rgb=c1.cpu().numpy(), radii_cov_rgb=r1.cpu().numpy(),
                        color_sr_sh3=c2.cpu().numpy(), radii_sr_sh3=r2.cpu().numpy())
    print("wrote", sys.argv[1], float(c1.mean()), int((r1 > 0).sum()))
