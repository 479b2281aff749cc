"""Host-side matching driver — mirror of the reference's `splatt3r_slam/matching.py:8-90`
(same function names, argument meaning and outputs) so parity tests and the bench can drive
the backend exactly the way `tracker.py:31` / `global_opt.py:40-51` do.

The reference reads its scalars from the global YAML config (`config["matching"]`,
config/base.yaml:8-14); here they are the `cfg` dict with the same keys and defaults.
`backend` is the module exporting `iter_proj` / `refine_matches` — the CUDA backend by default;
tests pass the CPU oracle to run the identical glue on CPU tensors.
"""
import torch
import torch.nn.functional as F

DEFAULT_CFG = dict(max_iter=10, lambda_init=1e-8, convergence_thresh=1e-6, dist_thresh=1e-1,
                   radius=3, dilation_max=5)


def _default_backend():
    from . import mast3r_slam_backends
    return mast3r_slam_backends


def pixel_to_lin(p1, w):
    return p1[..., 0] + (w * p1[..., 1])


def lin_to_pixel(idx_1_to_2, w):
    return torch.stack((idx_1_to_2 % w, idx_1_to_2 // w), dim=-1)


def img_gradient(img):
    """Scharr-like 3x3 gradients with reflect padding (reference image.py:5-38). img (b,c,h,w)."""
    c = img.shape[1]
    k = torch.tensor([[-3.0, 0.0, 3.0], [-10.0, 0.0, 10.0], [-3.0, 0.0, 3.0]],
                     device=img.device, dtype=img.dtype)
    gx_k = ((1.0 / 32.0) * k).repeat(c, 1, 1, 1)
    gy_k = ((1.0 / 32.0) * k.t().contiguous()).repeat(c, 1, 1, 1)
    pad = F.pad(img, (1, 1, 1, 1), mode="reflect")
    return F.conv2d(pad, gx_k, groups=c), F.conv2d(pad, gy_k, groups=c)


def prep_for_iter_proj(X11, X21, idx_1_to_2_init):
    b, h, w, _ = X11.shape
    rays = F.normalize(X11, dim=-1).permute(0, 3, 1, 2)
    gx, gy = img_gradient(rays)
    rays_with_grad_img = torch.cat((rays, gx, gy), dim=1).permute(0, 2, 3, 1).contiguous()
    pts3d_norm = F.normalize(X21.view(b, -1, 3), dim=-1)
    if idx_1_to_2_init is None:
        idx_1_to_2_init = torch.arange(h * w, device=X11.device)[None, :].repeat(b, 1)
    p_init = lin_to_pixel(idx_1_to_2_init, w).float()
    return rays_with_grad_img, pts3d_norm, p_init


def match_iterative_proj(X11, X21, D11, D21, idx_1_to_2_init=None, cfg=None, backend=None):
    cfg = DEFAULT_CFG if cfg is None else cfg
    be = _default_backend() if backend is None else backend
    b, h, w = X21.shape[:3]
    rays_with_grad_img, pts3d_norm, p_init = prep_for_iter_proj(X11, X21, idx_1_to_2_init)
    p1, valid_proj2 = be.iter_proj(rays_with_grad_img, pts3d_norm, p_init, cfg["max_iter"],
                                   cfg["lambda_init"], cfg["convergence_thresh"])
    p1 = p1.long()
    # occlusion test on 3-D distance (matching.py:70-76)
    batch_inds = torch.arange(b, device=X11.device)[:, None].repeat(1, h * w)
    dists2 = torch.linalg.norm(
        X11[batch_inds, p1[..., 1], p1[..., 0], :].reshape(b, h, w, 3) - X21, dim=-1)
    valid_proj2 = valid_proj2 & (dists2 < cfg["dist_thresh"]).view(b, -1)
    if cfg["radius"] > 0:
        (p1,) = be.refine_matches(D11.half(), D21.view(b, h * w, -1).half(), p1, cfg["radius"],
                                  cfg["dilation_max"])
    return pixel_to_lin(p1, w), valid_proj2.unsqueeze(-1)


def match_fused(X11, X21, D11, D21, idx_1_to_2_init=None, cfg=None, backend=None):
    """Same inputs and outputs as `match_iterative_proj`, through the backend's single fused entry
    (`b200slam_match`: 4 launches instead of ~12 torch kernels + 3 backend launches)."""
    cfg = DEFAULT_CFG if cfg is None else cfg
    be = _default_backend() if backend is None else backend
    idx, valid = be.match(X11, X21, D11, D21, idx_1_to_2_init, cfg["max_iter"], cfg["lambda_init"],
                          cfg["convergence_thresh"], cfg["dist_thresh"], cfg["radius"], cfg["dilation_max"])
    return idx, valid


def match(X11, X21, D11, D21, idx_1_to_2_init=None, cfg=None, backend=None, fused=False):
    if fused:
        return match_fused(X11, X21, D11, D21, idx_1_to_2_init, cfg, backend)
    return match_iterative_proj(X11, X21, D11, D21, idx_1_to_2_init, cfg, backend)
