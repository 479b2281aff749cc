"""Which fp32 operation order does torch's CUDA path use for the matching glue (F.normalize over a
3-vector, the grouped 3x3 conv of image.py:5-38)?  Emulates candidate orders with exact fp64
products rounded to fp32 step by step and prints the fraction of bit-identical outputs, so the fused
prep kernel can follow the order the reference's glue actually produces on this GPU/torch build."""
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from splatt3r_slam_b200 import matching  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)
X = (torch.randn(1, 96, 128, 3, device=dev) * torch.tensor([1.0, 0.7, 3.0], device=dev)).contiguous()


def r32(x64):
    return x64.float().double()


def eq(a, b):
    return float((a == b).float().mean())


out = {}
x, y, z = [X[..., k].double() for k in range(3)]
ref_n = torch.linalg.norm(X, dim=-1)
cands = {
    "fma_seq_xyz": r32(z * z + r32(y * y + r32(x * x))),
    "fma_seq_zyx": r32(x * x + r32(y * y + r32(z * z))),
    "unfused_seq": r32(r32(r32(x * x) + r32(y * y)) + r32(z * z)),
    "fma_0_start": r32(z * z + r32(y * y + r32(x * x + 0.0))),
    "tree_xz_y": r32(r32(z * z + r32(x * x)) + r32(y * y)),
    "tree_xy_z_unf": r32(r32(x * x + r32(y * y)) + r32(z * z)),
    "unf_tree_xz_y": r32(r32(r32(x * x) + r32(z * z)) + r32(y * y)),
    "unf_tree_yz_x": r32(r32(r32(y * y) + r32(z * z)) + r32(x * x)),
    "unf_y_then_xz": r32(r32(y * y) + r32(r32(x * x) + r32(z * z))),
    "exact_sum": r32(x * x + y * y + z * z),
}
sq = (X.double() ** 2).sum(-1)
out["norm_vs_exact_ulp_hist"] = torch.bincount(((ref_n.view(torch.int32) - torch.sqrt(sq).float().view(torch.int32)).flatten() + 4).clamp(0, 8), minlength=9).tolist()
out["norm"] = {k: eq(torch.sqrt(v.float()), ref_n) for k, v in cands.items()}
ref_r = F.normalize(X, dim=-1)
out["normalize_div_given_norm"] = eq(X / ref_n.clamp_min(1e-12)[..., None], ref_r)
out["normalize_mul_rcp"] = eq(X * (1.0 / ref_n.clamp_min(1e-12))[..., None], ref_r)

rays = ref_r.permute(0, 3, 1, 2).contiguous()
gx, gy = matching.img_gradient(rays)
pad = F.pad(rays, (1, 1, 1, 1), mode="reflect").double()
h, w = rays.shape[2:]
kx = [[-3.0, 0.0, 3.0], [-10.0, 0.0, 10.0], [-3.0, 0.0, 3.0]]


def conv_seq(k, order, skip_zero):
    acc = torch.zeros_like(rays, dtype=torch.float64)
    for (a, b) in order:
        wgt = k[a][b] / 32.0
        if skip_zero and wgt == 0.0:
            continue
        acc = r32(acc + wgt * pad[:, :, a:a + h, b:b + w])
    return acc.float()


row_major = [(a, b) for a in range(3) for b in range(3)]
col_major = [(a, b) for b in range(3) for a in range(3)]
ky = [[kx[b][a] for b in range(3)] for a in range(3)]
out["gx"] = {"row_major_fma": eq(conv_seq(kx, row_major, True), gx), "col_major_fma": eq(conv_seq(kx, col_major, True), gx),
             "row_major_rev": eq(conv_seq(kx, row_major[::-1], True), gx)}
out["gy"] = {"row_major_fma": eq(conv_seq(ky, row_major, True), gy), "col_major_fma": eq(conv_seq(ky, col_major, True), gy),
             "row_major_rev": eq(conv_seq(ky, row_major[::-1], True), gy)}
out["gx_maxdiff_row_major"] = float((conv_seq(kx, row_major, True) - gx).abs().max())
out["cudnn_enabled"] = torch.backends.cudnn.enabled
print(json.dumps(out, indent=1))
