"""GPU tier: the fused match entry (b200slam_match) against the unfused path
it replaces — the reference's own PyTorch glue (matching.py:25-90, restated in
splatt3r_slam_b200/matching.py) around iter_proj / refine_matches.  The fused kernels follow the float
operation order torch's CUDA kernels use for that glue, so idx and valid must be IDENTICAL.  Against the
CPU oracle (same glue on CPU tensors, where torch's norm reduction sums in another order) the
match indices may differ on a handful of boundary pixels: tolerance 1e-3 of the pixels, stated below."""
import pytest
import torch

import oracle
from splatt3r_slam_b200 import matching, synthetic

pytestmark = pytest.mark.gpu


def dev():
    return torch.device("cuda:0")


def pair(h, w, b, seed):
    return [t.to(dev()) for t in synthetic.make_pair(seed=seed, h=h, w=w, fdim=24, b=b)]


@pytest.mark.parametrize("h,w,b", [(48, 64, 1), (37, 53, 2), (96, 128, 3), (384, 512, 1)])
@pytest.mark.parametrize("desc", ["f16", "f32"])
def test_fused_equals_unfused(backend, h, w, b, desc):
    X11, X21, D11, D21 = pair(h, w, b, seed=h + w)
    if desc == "f16":
        D11, D21 = D11.half(), D21.half()
    idx_u, val_u = matching.match(X11, X21, D11, D21, backend=backend)
    idx_f, val_f = matching.match(X11, X21, D11, D21, backend=backend, fused=True)
    assert idx_f.dtype == torch.int64 and val_f.dtype == torch.bool and val_f.shape == val_u.shape
    assert torch.equal(idx_f, idx_u) and torch.equal(val_f, val_u)
    assert 0.2 < val_f.float().mean() < 1.0
    # second round seeded with the first result (tracker.py / global_opt.py pass idx back in)
    init = idx_u.roll(3, dims=1).contiguous()
    idx_u2, val_u2 = matching.match(X11, X21, D11, D21, init, backend=backend)
    idx_f2, val_f2 = matching.match(X11, X21, D11, D21, init, backend=backend, fused=True)
    assert torch.equal(idx_f2, idx_u2) and torch.equal(val_f2, val_u2)


def test_fused_without_refinement(backend):
    X11, X21, D11, D21 = pair(48, 64, 2, seed=5)
    cfg = dict(matching.DEFAULT_CFG, radius=0)
    idx_u, val_u = matching.match(X11, X21, D11, D21, cfg=cfg, backend=backend)
    idx_f, val_f = matching.match(X11, X21, D11, D21, cfg=cfg, backend=backend, fused=True)
    assert torch.equal(idx_f, idx_u) and torch.equal(val_f, val_u)


def test_fused_degenerate_points(backend):
    """Zero-length points (F.normalize's eps clamp), points far away (occlusion test fails) and a
    non-tiled refine shape (w too small for the staged box -> per-thread kernel + pixel_to_lin)."""
    X11, X21, D11, D21 = pair(24, 40, 1, seed=9)
    X21 = X21.clone()
    X21[0, 3, 5] = 0.0
    X21[0, 7, 9] *= 50.0
    X11 = X11.clone()
    X11[0, 10, 11] = 0.0
    idx_u, val_u = matching.match(X11, X21, D11, D21, backend=backend)
    idx_f, val_f = matching.match(X11, X21, D11, D21, backend=backend, fused=True)
    assert torch.equal(idx_f, idx_u) and torch.equal(val_f, val_u)
    assert not val_f[0, 7 * 40 + 9, 0]


def test_fused_vs_cpu_oracle(backend):
    h, w = 96, 128
    X11, X21, D11, D21 = pair(h, w, 1, seed=11)
    idx_f, val_f = matching.match(X11, X21, D11, D21, backend=backend, fused=True)
    idx_o, val_o = matching.match(X11.cpu(), X21.cpu(), D11.cpu(), D21.cpu(), backend=oracle)
    frac_idx = (idx_f.cpu() != idx_o).float().mean().item()
    frac_val = (val_f.cpu() != val_o).float().mean().item()
    assert frac_idx <= 1e-3 and frac_val <= 1e-3, (frac_idx, frac_val)


def test_fused_argument_errors(backend):
    X11, X21, D11, D21 = pair(24, 40, 1, seed=2)
    with pytest.raises(RuntimeError):
        backend.match(X11.permute(0, 2, 1, 3), X21, D11, D21, None, 10, 1e-8, 1e-6, 0.1, 3, 5)
    with pytest.raises(RuntimeError):
        backend.match(X11, X21, D11.double(), D21.double(), None, 10, 1e-8, 1e-6, 0.1, 3, 5)
    with pytest.raises(RuntimeError):
        backend.match(X11.cpu(), X21, D11, D21, None, 10, 1e-8, 1e-6, 0.1, 3, 5)
