"""Generate the golden fixtures in tests/golden/ by running the REFERENCE backend itself
(oracle/_ref/mast3r_slam_backends_ref.so, built by oracle/build_ref.py from the unmodified
device code of /root/reference) on a B200.  Run on the GPU box:

    python tests/golden/make_golden.py gpurun_out/golden      # then copy *.npz to tests/golden/

Inputs are the seeded synthetic tensors of splatt3r_slam_b200.synthetic; the exact kernel inputs
are stored next to the reference outputs so the fixtures do not depend on the platform's conv /
normalise implementation.
"""
import importlib.util
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
REF_SO = os.path.join(ROOT, "oracle", "_ref", "mast3r_slam_backends_ref.so")


def load_ref():
    spec = importlib.util.spec_from_file_location("mast3r_slam_backends_ref", REF_SO)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)  # also registers torch.ops.refshim.*
    return mod


GN_CFG = dict(sigma_ray=0.003, sigma_dist=10.0, sigma_pixel=1.0, sigma_depth=10.0, sigma_point=0.05,
              C_thresh=0.0, Q_thresh=1.5, pixel_border=-10, z_eps=1e-6)


def main(out_dir):
    from splatt3r_slam_b200 import matching, synthetic
    ref = load_ref()
    dev = torch.device("cuda:0")
    os.makedirs(out_dir, exist_ok=True)
    cfg = matching.DEFAULT_CFG

    # ---- matching: b=2 pairs of 64x48 ------------------------------------------------------
    h, w = 48, 64
    X11, X21, D11, D21 = [t.to(dev) for t in synthetic.make_pair(seed=11, h=h, w=w, fdim=24, b=2, noise=0.005)]
    rays, pts, p_init = matching.prep_for_iter_proj(X11, X21, None)
    # second batch row starts from a shifted, partly out-of-range initial guess (tests the clamp)
    p_init[1] = p_init[1] + torch.tensor([7.5, -3.25], device=dev)
    p_new, conv = ref.iter_proj(rays, pts, p_init.contiguous(), cfg["max_iter"], cfg["lambda_init"],
                                cfg["convergence_thresh"])
    p1 = p_new.long()
    D11h, D21h = D11.half(), D21.view(2, h * w, -1).half().contiguous()
    (p1_ref,) = ref.refine_matches(D11h, D21h, p1, cfg["radius"], cfg["dilation_max"])
    (p1_ref_f32,) = ref.refine_matches(D11h.float(), D21h.float(), p1,  # fp32 kernel on the fp16-rounded values
                                       cfg["radius"], cfg["dilation_max"])
    torch.cuda.synchronize()
    np.savez_compressed(
        os.path.join(out_dir, "matching_64x48.npz"),
        rays=rays.cpu().numpy(), pts=pts.cpu().numpy(), p_init=p_init.cpu().numpy(),
        p_new=p_new.cpu().numpy(), converged=conv.cpu().numpy(),
        D11=D11h.cpu().numpy(), D21=D21h.cpu().numpy(), p1=p1.cpu().numpy(),
        p1_refined=p1_ref.cpu().numpy(), p1_refined_f32=p1_ref_f32.cpu().numpy())

    # ---- GN: N=3 keyframes, 3 undirected = 6 directed edges, 64x48 -------------------------
    g = synthetic.make_graph(seed=5, N=3, E_undirected=3, h=h, w=w, noise=0.005)
    G = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in g.items()}
    c = GN_CFG
    out = {k: g[k].numpy() for k in ("Twc", "Xs", "Cs", "K", "ii", "jj", "idx_ii2jj", "valid_match", "Q")}
    Hs, gs = torch.ops.refshim.ray_align(G["Twc"], G["Xs"], G["Cs"], G["ii"], G["jj"], G["idx_ii2jj"],
                                         G["valid_match"], G["Q"], c["sigma_ray"], c["sigma_dist"],
                                         c["C_thresh"], c["Q_thresh"])
    out["rays_Hs"], out["rays_gs"] = Hs.cpu().numpy(), gs.cpu().numpy()
    Hs, gs = torch.ops.refshim.calib_proj(G["Twc"], G["Xs"], G["Cs"], G["K"], G["ii"], G["jj"],
                                          G["idx_ii2jj"], G["valid_match"], G["Q"], h, w, c["pixel_border"],
                                          c["z_eps"], c["sigma_pixel"], c["sigma_depth"], c["C_thresh"],
                                          c["Q_thresh"])
    out["calib_Hs"], out["calib_gs"] = Hs.cpu().numpy(), gs.cpu().numpy()
    Hs, gs = torch.ops.refshim.point_align(G["Twc"], G["Xs"], G["Cs"], G["ii"], G["jj"], G["idx_ii2jj"],
                                           G["valid_match"], G["Q"], c["sigma_point"], c["C_thresh"],
                                           c["Q_thresh"])
    out["points_Hs"], out["points_gs"] = Hs.cpu().numpy(), gs.cpu().numpy()
    for iters in (1, 4):
        T = G["Twc"].clone()
        (dx,) = ref.gauss_newton_rays(T, G["Xs"], G["Cs"], G["ii"], G["jj"], G["idx_ii2jj"], G["valid_match"],
                                      G["Q"], c["sigma_ray"], c["sigma_dist"], c["C_thresh"], c["Q_thresh"],
                                      iters, 0.0)
        out[f"rays_T{iters}"], out[f"rays_dx{iters}"] = T.cpu().numpy(), dx.cpu().numpy()
        T = G["Twc"].clone()
        (dx,) = ref.gauss_newton_calib(T, G["Xs"], G["Cs"], G["K"], G["ii"], G["jj"], G["idx_ii2jj"],
                                       G["valid_match"], G["Q"], h, w, c["pixel_border"], c["z_eps"],
                                       c["sigma_pixel"], c["sigma_depth"], c["C_thresh"], c["Q_thresh"],
                                       iters, 0.0)
        out[f"calib_T{iters}"], out[f"calib_dx{iters}"] = T.cpu().numpy(), dx.cpu().numpy()
        T = G["Twc"].clone()
        (dx,) = ref.gauss_newton_points(T, G["Xs"], G["Cs"], G["ii"], G["jj"], G["idx_ii2jj"],
                                        G["valid_match"], G["Q"], c["sigma_point"], c["C_thresh"],
                                        c["Q_thresh"], iters, 0.0)
        out[f"points_T{iters}"], out[f"points_dx{iters}"] = T.cpu().numpy(), dx.cpu().numpy()
    torch.cuda.synchronize()
    np.savez_compressed(os.path.join(out_dir, "gn_3kf_64x48.npz"), **out)
    print("golden fixtures written to", out_dir)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
