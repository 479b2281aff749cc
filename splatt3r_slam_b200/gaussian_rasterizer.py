"""Drop-in for the FORWARD interface of the reference's vendored rasteriser
(`thirdparty/diff-gaussian-rasterization-modified/diff_gaussian_rasterization/__init__.py:172-221`:
`GaussianRasterizationSettings`, `GaussianRasterizer.forward`) on top of `b200slam_gs_render` (C ABI, sm_100a).

Same names, argument meaning and error behaviour (exactly one of shs / colors_precomp, exactly one of
(scales, rotations) / cov3D_precomp); returns `(color (3,H,W), radii (P,))`.  Rendering only: the result carries no
autograd graph (`means2D`, the reference's gradient holder, is accepted and ignored).  No CPU path.
"""
import ctypes
from typing import NamedTuple

import torch

from . import mast3r_slam_backends as _be

_lib = _be._lib
_vp, _i, _f = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
_lib.b200slam_gs_workspace_bytes.argtypes = [_i, _i, _i, ctypes.c_longlong]
_lib.b200slam_gs_workspace_bytes.restype = ctypes.c_size_t
_lib.b200slam_gs_render.argtypes = ([_vp] * 4 + [_f] + [_vp] * 3 + [_i, _i] + [_vp] * 3 + [_f, _f, _vp, _i, _i, _i, _vp, _vp,
                                    ctypes.c_longlong, ctypes.POINTER(ctypes.c_longlong), _vp, ctypes.c_size_t, _vp])
_lib.b200slam_gs_render.restype = _i
EWORKSPACE = -2
_workspaces = {}  # (device index, stream) -> (uint8 tensor, instance capacity): calls on different streams must not share it


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def _f32(t, name):
    if t is None or t.numel() == 0:  # (the reference passes empty tensors for the unused alternatives)
        return None
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (this rasteriser has no CPU path)")
    if t.dtype == torch.float32 and t.is_contiguous():
        return t  # (no new tensor object: eleven arguments per call, each conversion costs microseconds of host time
        #          during which the GPU idles - the call synchronises once in the middle)
    return t.detach().to(torch.float32).contiguous()


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, raster_settings):
    rs = raster_settings
    dev = means3D.device
    P, H, W = int(means3D.shape[0]), int(rs.image_height), int(rs.image_width)
    t = dict(means=_f32(means3D, "means3D"), op=_f32(opacities, "opacities"), sh=_f32(sh, "shs"),
             col=_f32(colors_precomp, "colors_precomp"), sc=_f32(scales, "scales"), rot=_f32(rotations, "rotations"),
             cov=_f32(cov3Ds_precomp, "cov3D_precomp"), view=_f32(rs.viewmatrix, "viewmatrix"),
             proj=_f32(rs.projmatrix, "projmatrix"), bg=_f32(rs.bg, "bg"), cam=_f32(rs.campos, "campos"))
    ptr = lambda x: None if x is None or x.numel() == 0 else x.data_ptr()
    sh_coeffs = int(t["sh"].shape[1]) if t["sh"] is not None and t["sh"].numel() else 0
    with torch.cuda.device(dev):
        color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
        radii = torch.zeros((P,), dtype=torch.int32, device=dev)
        stream = torch.cuda.current_stream(dev).cuda_stream
        key = (dev.index, stream)
        ws, cap = _workspaces.get(key, (None, 0))
        need = ctypes.c_longlong(0)
        cap = max(cap, 8 * P + 1024)
        for _ in range(3):
            nbytes = _lib.b200slam_gs_workspace_bytes(P, H, W, cap)
            if ws is None or ws.numel() < nbytes:
                ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            _workspaces[key] = (ws, cap)
            rc = _lib.b200slam_gs_render(ptr(t["means"]), ptr(t["cov"]), ptr(t["sc"]), ptr(t["rot"]), float(rs.scale_modifier),
                                         ptr(t["op"]), ptr(t["col"]), ptr(t["sh"]), int(rs.sh_degree), sh_coeffs, ptr(t["cam"]),
                                         ptr(t["view"]), ptr(t["proj"]), float(rs.tanfovx), float(rs.tanfovy), ptr(t["bg"]), P, H,
                                         W, color.data_ptr(), radii.data_ptr(), cap, ctypes.byref(need), ws.data_ptr(),
                                         ws.numel(), stream)
            if rc == EWORKSPACE and need.value > cap:  # more (splat, tile) instances than the workspace holds: grow, retry
                cap = int(need.value * 1.25) + 1024
                continue
            break
        if rc != 0:
            raise RuntimeError(f"b200slam gs_render failed with code {rc}")
    return color, radii


class GaussianRasterizer(torch.nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception("Please provide excatly one of either SHs or precomputed colors!")
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (
                (scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!")
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                                   self.raster_settings)
