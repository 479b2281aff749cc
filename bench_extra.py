"""Sub-blocks of the bench line beyond the headline tracker step (all device-timed with CUDA events on the
launching stream, warm, inputs larger than L2 or rotating):
  * batched_matching_block  — iter_proj / refine_matches at b = 8 and b = 32 pairs per call (the
    roofline figure of the matcher is quoted on b >= 8; global_opt.py:40-51 produces b = 2 * new edges)
  * ref_cuda_block          — the reference's own CUDA build (oracle/_ref, when it travelled) on the same
    rotating inputs, timed OUTSIDE the product arm's timed regions
  * bind_numa               — pin the rank's process to the CPU cores of its GPU's NUMA node
"""
import importlib.util
import os

import torch

from bench import B_ITER_PROJ, B_REFINE, GN, N_PIX, peaks

ROOT = os.path.dirname(os.path.abspath(__file__))


def bind_numa(local):
    """Bind this process to the cores of the NUMA node its GPU hangs off, so that the rank's pinned arena and the
    thread that drives the copies are local to the GPU's PCIe root.  Returns a description (or why not)."""
    try:
        p = torch.cuda.get_device_properties(local)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read().strip())
        if node < 0:
            return {"bound": False, "why": "numa_node = -1 (single node or not exposed)", "pci": bdf}
        cpus = []
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.extend(range(int(a), int(b or a) + 1))
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if not allowed:
            return {"bound": False, "why": "no allowed core on the GPU's node", "node": node, "pci": bdf}
        os.sched_setaffinity(0, allowed)
        return {"bound": True, "node": node, "cores": len(allowed), "pci": bdf}
    except Exception as exc:  # noqa: BLE001 - reporting only
        return {"bound": False, "why": f"{type(exc).__name__}: {exc}"}


def _time_us(fn, reps, n_rot):
    for i in range(3):
        fn(i % n_rot)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(reps):
        fn(i % n_rot)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3


def batched_matching_block(be, cfg, sets, reps=10):
    """iter_proj and refine_matches at b = 8 and b = 32 pairs per call.  Batches are built from the bench's distinct
    512x384 input sets (cycled; every batch row is its own memory, b = 8 is 0.3 GB and b = 32 1.2 GB of inputs,
    both far beyond the 126 MB L2; two batches alternate)."""
    peak, _ = peaks()
    out = {}
    for b in (8, 32):
        rot = []
        for r in range(2):
            pick = [sets[(r + k) % len(sets)] for k in range(b)]
            d = {k: torch.cat([s[k] for s in pick], 0).contiguous() for k in ("rays", "pts", "p_init", "D11h", "D21h")}
            p_new, _ = be.iter_proj(d["rays"], d["pts"], d["p_init"], cfg["max_iter"], cfg["lambda_init"],
                                    cfg["convergence_thresh"])
            d["p1"] = p_new.long()
            rot.append(d)
        t_ip = _time_us(lambda i: be.iter_proj(rot[i]["rays"], rot[i]["pts"], rot[i]["p_init"], cfg["max_iter"],
                                               cfg["lambda_init"], cfg["convergence_thresh"]), reps, 2)
        t_rf = _time_us(lambda i: be.refine_matches(rot[i]["D11h"], rot[i]["D21h"], rot[i]["p1"], cfg["radius"],
                                                    cfg["dilation_max"]), reps, 2)
        px = b * N_PIX
        out[f"b{b}"] = {
            "iter_proj_us": t_ip, "refine_us": t_rf, "match_px_per_s": px / ((t_ip + t_rf) * 1e-6),
            "iter_proj_GBps": B_ITER_PROJ * px / (t_ip * 1e-6) / 1e9,
            "iter_proj_frac_hbm": B_ITER_PROJ * px / (t_ip * 1e-6) / 1e9 / peak,
            "refine_GBps": B_REFINE * px / (t_rf * 1e-6) / 1e9,
            "refine_frac_hbm": B_REFINE * px / (t_rf * 1e-6) / 1e9 / peak,
        }
        del rot
        torch.cuda.empty_cache()
    return out


def load_ref_cuda():
    """The reference's CUDA extension built by oracle/build_ref.py (None when it did not travel)."""
    so = os.path.join(ROOT, "oracle", "_ref", "mast3r_slam_backends_ref.so")
    if not os.path.exists(so):
        return None
    try:
        spec = importlib.util.spec_from_file_location("mast3r_slam_backends_ref", so)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod
    except Exception:  # noqa: BLE001
        return None


def ref_cuda_block(cfg, sets, reps=10):
    """The reference's own CUDA kernels (matching_kernels.cu / gn_kernels.cu compiled for sm_100a, device code
    untouched) on the bench's rotating tracker inputs, same box, same events.  Context for the product numbers; it
    is never inside a timed region of the product arm."""
    ref = load_ref_cuda()
    if ref is None:
        return {"unavailable": "oracle/_ref/mast3r_slam_backends_ref.so not present on this box"}
    n = len(sets)
    p1s = []
    for d in sets:
        p_new, _ = ref.iter_proj(d["rays"], d["pts"], d["p_init"], cfg["max_iter"], cfg["lambda_init"],
                                 cfg["convergence_thresh"])
        p1s.append(p_new.long())
    Ts = [d["Twc"].clone() for d in sets]
    t_ip = _time_us(lambda i: ref.iter_proj(sets[i]["rays"], sets[i]["pts"], sets[i]["p_init"], cfg["max_iter"],
                                            cfg["lambda_init"], cfg["convergence_thresh"]), reps, n)
    t_rf = _time_us(lambda i: ref.refine_matches(sets[i]["D11h"], sets[i]["D21h"], p1s[i], cfg["radius"],
                                                 cfg["dilation_max"]), reps, n)

    def gn(i):
        Ts[i].copy_(sets[i]["Twc"])
        ref.gauss_newton_rays(Ts[i], sets[i]["Xs"], sets[i]["Cs"], sets[i]["ii"], sets[i]["jj"], sets[i]["idx_ii2jj"],
                              sets[i]["valid_match"], sets[i]["Q"], GN["sigma_ray"], GN["sigma_dist"], GN["C_thresh"],
                              GN["Q_thresh"], GN["iters"], 0.0)
    t_gn = _time_us(gn, max(3, reps // 2), n)
    step_ms = (t_ip + t_rf + t_gn) * 1e-3
    return {"iter_proj_us": t_ip, "refine_us": t_rf, "gn_10it_us": t_gn, "step_ms": step_ms,
            "matched_px_per_s": N_PIX / (step_ms * 1e-3),
            "what": "reference CUDA extension (oracle/_ref: /root/reference sources, 4 host-line edits for torch 2.11, "
                    "device code untouched) timed per call on the same rotating inputs"}


def gs_render_block(reps=10, keyframes=4, H=384, W=512):
    """SURVEY.md 8f rank 4 (last item): the forward 3DGS rasteriser on a scene of the shape the SLAM viewer renders — one
    Gaussian per pixel of `keyframes` 512x384 keyframes (visualization.py:470-596: all accumulated Gaussians, precomputed
    covariances and colours) — through the mirror of the reference's GaussianRasterizer, and the reference's own CUDA
    build (oracle/_ref/diffgauss_ref.so, when it travelled) on the same tensors."""
    import torch
    from splatt3r_slam_b200 import synthetic
    from splatt3r_slam_b200.gaussian_rasterizer import GaussianRasterizationSettings, GaussianRasterizer
    dev = torch.device("cuda", torch.cuda.current_device())
    P = keyframes * H * W
    S = synthetic.make_splat_scene(seed=9, P=P, H=H, W=W)
    S["scales"] = S["scales"] * 0.08   # pixel-sized splats, as one Gaussian per pointmap pixel gives
    S["cov"] = S["cov"] * 0.08 ** 2
    d = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in S.items()}
    rs = GaussianRasterizationSettings(H, W, S["tanfovx"], S["tanfovy"], d["bg"], 1.0, d["view"], d["proj"], 0, d["campos"],
                                       False, False)
    R = GaussianRasterizer(rs)
    mg = torch.zeros_like(d["means"])
    ours = lambda i: R(d["means"], mg, d["opacities"], colors_precomp=d["colors"], cov3D_precomp=d["cov"])
    color, radii = ours(0)
    t_ours = _time_us(ours, reps, 1)
    vis = int((radii > 0).sum())
    out = {"workload": f"{P} splats ({keyframes} keyframes x {H}x{W}), {H}x{W} image, precomputed covariance + colour",
           "visible_splats": vis, "us": t_ours, "splats_per_s": P / (t_ours * 1e-6),
           # bytes the forward pass has to touch at least once: inputs (3+6+1+3 floats per splat) + image
           "algorithmic_GBps": (P * 13 * 4 + 3 * H * W * 4) / (t_ours * 1e-6) / 1e9}
    so = os.path.join(ROOT, "oracle", "_ref", "diffgauss_ref.so")
    if os.path.exists(so):
        try:
            spec = importlib.util.spec_from_file_location("diffgauss_ref", so)
            ref = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(ref)
            e = torch.empty(0, device=dev)
            fn = lambda i: ref.rasterize_gaussians(d["bg"], d["means"], d["colors"], d["opacities"], e, e, 1.0, d["cov"], d["view"],
                                                   d["proj"], S["tanfovx"], S["tanfovy"], H, W, e, 0, d["campos"], False, False)
            rc = fn(0)[1]
            out["ref_cuda_us"] = _time_us(fn, reps, 1)
            out["max_abs_diff_vs_ref"] = float((rc - color).abs().max())
        except Exception as exc:  # noqa: BLE001
            out["ref_cuda_us"] = None
            out["ref_error"] = str(exc)[:200]
    return out
