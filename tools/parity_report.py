"""GPU-box report: B200 kernels vs the reference CUDA build (oracle/_ref) vs the CPU oracle at
full MASt3R size, plus CUDA-event timings of both GPU implementations.  Test infrastructure.

    python tools/parity_report.py gpurun_out/parity.json [--small]
"""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))


def timeit(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3  # us


def rel(a, b):
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def main(out_path, small=False):
    import oracle
    from make_golden import GN_CFG, load_ref
    from splatt3r_slam_b200 import matching, synthetic
    import splatt3r_slam_b200.mast3r_slam_backends as be
    try:
        ref = load_ref()
    except Exception as exc:  # noqa
        ref = None
        print("reference build not loadable:", exc)
    dev = torch.device("cuda:0")
    h, w = (96, 128) if small else (384, 512)
    cfg = matching.DEFAULT_CFG
    rep = {"h": h, "w": w, "gpu": torch.cuda.get_device_name(0), "ref_loaded": ref is not None}

    # ---------------- matching ----------------
    X11, X21, D11, D21 = synthetic.make_pair(seed=0, h=h, w=w, fdim=24, b=2)
    Xg = [t.to(dev) for t in (X11, X21, D11, D21)]
    rays, pts, p_init = matching.prep_for_iter_proj(Xg[0], Xg[1], None)
    args = (cfg["max_iter"], cfg["lambda_init"], cfg["convergence_thresh"])
    p_b, c_b = be.iter_proj(rays, pts, p_init, *args)
    t0 = time.time()
    p_o, c_o = oracle.iter_proj(rays.cpu(), pts.cpu(), p_init.cpu(), *args)
    rep["iter_proj_oracle_cpu_s"] = time.time() - t0
    rep["iter_proj_vs_oracle_p_mismatch"] = int((p_b.cpu() != p_o).sum())
    rep["iter_proj_vs_oracle_conv_mismatch"] = int((c_b.cpu() != c_o).sum())
    rep["iter_proj_us"] = timeit(lambda: be.iter_proj(rays, pts, p_init, *args))
    if ref:
        p_r, c_r = ref.iter_proj(rays, pts, p_init, *args)
        rep["iter_proj_vs_ref_p_mismatch"] = int((p_b != p_r).sum())
        rep["iter_proj_vs_ref_conv_mismatch"] = int((c_b != c_r).sum())
        rep["iter_proj_ref_vs_oracle_p_mismatch"] = int((p_r.cpu() != p_o).sum())
        rep["iter_proj_ref_us"] = timeit(lambda: ref.iter_proj(rays, pts, p_init, *args))
    p1 = p_b.long()
    D11h, D21h = Xg[2].half(), Xg[3].view(2, h * w, -1).half().contiguous()
    (r_b,) = be.refine_matches(D11h, D21h, p1, cfg["radius"], cfg["dilation_max"])
    t0 = time.time()
    (r_o,) = oracle.refine_matches(D11h.cpu(), D21h.cpu(), p1.cpu(), cfg["radius"], cfg["dilation_max"])
    rep["refine_oracle_cpu_s"] = time.time() - t0
    rep["refine_vs_oracle_mismatch"] = int((r_b.cpu() != r_o).sum())
    rep["refine_moved_frac"] = float((r_b != p1).any(-1).float().mean())
    rep["refine_us"] = timeit(lambda: be.refine_matches(D11h, D21h, p1, cfg["radius"], cfg["dilation_max"]))
    if ref:
        (r_r,) = ref.refine_matches(D11h, D21h, p1, cfg["radius"], cfg["dilation_max"])
        rep["refine_vs_ref_mismatch"] = int((r_b != r_r).sum())
        rep["refine_ref_us"] = timeit(lambda: ref.refine_matches(D11h, D21h, p1, cfg["radius"], cfg["dilation_max"]), iters=5)
        D11f, D21f = Xg[2].float(), Xg[3].view(2, h * w, -1).float().contiguous()
        (rf_b,) = be.refine_matches(D11f, D21f, p1, cfg["radius"], cfg["dilation_max"])
        (rf_r,) = ref.refine_matches(D11f, D21f, p1, cfg["radius"], cfg["dilation_max"])
        rep["refine_f32_vs_ref_mismatch"] = int((rf_b != rf_r).sum())
    rep["pixels_per_call"] = 2 * h * w

    # ---------------- GN ----------------
    c = GN_CFG
    for (N, Eu) in ((2, 1), (8, 16)):
        g = synthetic.make_graph(seed=2, N=N, E_undirected=Eu, h=h, w=w)
        G = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in g.items()}
        tag = f"N{N}_E{2 * Eu}"
        kinds = {
            "rays": (be.make_params(0, c["sigma_ray"], c["sigma_dist"], c["C_thresh"], c["Q_thresh"]), None),
            "calib": (be.make_params(1, c["sigma_pixel"], c["sigma_depth"], c["C_thresh"], c["Q_thresh"], h, w,
                                     c["pixel_border"], c["z_eps"]), G["K"]),
            "points": (be.make_params(2, c["sigma_point"], 1.0, c["C_thresh"], c["Q_thresh"]), None),
        }
        for name, (prm, K) in kinds.items():
            Hs, gs = be.gn_build_blocks(prm, G["Twc"], G["Xs"], G["Cs"], K, G["ii"], G["jj"], G["idx_ii2jj"],
                                        G["valid_match"], G["Q"])
            Ho, go = oracle.gn_build(prm.kind, g["Twc"], g["Xs"], g["Cs"], g["K"] if K is not None else None,
                                     g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"], prm.sigma_a,
                                     prm.sigma_b, prm.C_thresh, prm.Q_thresh, prm.height, prm.width,
                                     prm.pixel_border, prm.z_eps)
            rep[f"gn_{name}_{tag}_Hs_rel_vs_oracle"] = rel(Hs.cpu(), Ho)
            rep[f"gn_{name}_{tag}_gs_rel_vs_oracle"] = rel(gs.cpu(), go)
            if ref:
                if name == "rays":
                    Hr, gr = torch.ops.refshim.ray_align(G["Twc"], G["Xs"], G["Cs"], G["ii"], G["jj"], G["idx_ii2jj"],
                                                         G["valid_match"], G["Q"], prm.sigma_a, prm.sigma_b,
                                                         prm.C_thresh, prm.Q_thresh)
                elif name == "calib":
                    Hr, gr = torch.ops.refshim.calib_proj(G["Twc"], G["Xs"], G["Cs"], G["K"], G["ii"], G["jj"],
                                                          G["idx_ii2jj"], G["valid_match"], G["Q"], h, w,
                                                          prm.pixel_border, prm.z_eps, prm.sigma_a, prm.sigma_b,
                                                          prm.C_thresh, prm.Q_thresh)
                else:
                    Hr, gr = torch.ops.refshim.point_align(G["Twc"], G["Xs"], G["Cs"], G["ii"], G["jj"],
                                                           G["idx_ii2jj"], G["valid_match"], G["Q"], prm.sigma_a,
                                                           prm.C_thresh, prm.Q_thresh)
                rep[f"gn_{name}_{tag}_Hs_rel_vs_ref"] = rel(Hs, Hr)
                rep[f"gn_{name}_{tag}_gs_rel_vs_ref"] = rel(gs, gr)
                rep[f"gn_{name}_{tag}_oracle_Hs_rel_vs_ref"] = rel(Ho, Hr.cpu())
        # full loop, 10 iterations forced
        def run_b():
            T = G["Twc"].clone()
            be.gauss_newton_rays(T, G["Xs"], G["Cs"], G["ii"], G["jj"], G["idx_ii2jj"], G["valid_match"], G["Q"],
                                 c["sigma_ray"], c["sigma_dist"], c["C_thresh"], c["Q_thresh"], 10, 0.0)
            return T
        Tb = run_b()
        rep[f"gn_rays_{tag}_10it_us"] = timeit(run_b, iters=10)
        To = g["Twc"].clone()
        t0 = time.time()
        oracle.gauss_newton_rays(To, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"],
                                 c["sigma_ray"], c["sigma_dist"], c["C_thresh"], c["Q_thresh"], 10, 0.0)
        rep[f"gn_rays_{tag}_oracle_cpu_s"] = time.time() - t0
        rep[f"gn_rays_{tag}_pose_abs_vs_oracle"] = (Tb.cpu() - To).abs().max().item()
        if ref:
            def run_r():
                T = G["Twc"].clone()
                ref.gauss_newton_rays(T, G["Xs"], G["Cs"], G["ii"], G["jj"], G["idx_ii2jj"], G["valid_match"],
                                      G["Q"], c["sigma_ray"], c["sigma_dist"], c["C_thresh"], c["Q_thresh"], 10, 0.0)
                return T
            Tr = run_r()
            rep[f"gn_rays_{tag}_pose_abs_vs_ref"] = (Tb - Tr).abs().max().item()
            rep[f"gn_rays_{tag}_ref_10it_us"] = timeit(run_r, iters=5)
    os.makedirs(os.path.dirname(os.path.abspath(out_path)), exist_ok=True)
    json.dump(rep, open(out_path, "w"), indent=1)
    print(json.dumps(rep, indent=1))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/parity.json", small="--small" in sys.argv)
