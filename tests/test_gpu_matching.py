"""GPU tier: iter_proj / refine_matches through the drop-in module (-> C ABI -> sm_100a kernels)
against (1) the golden fixtures of the reference CUDA build, (2) the C oracle on seeded inputs,
(3) the reference CUDA build itself when oracle/_ref travelled to this box.  Bit-exact."""
import os

import numpy as np
import pytest
import torch

import oracle
from conftest import GOLDEN
from splatt3r_slam_b200 import matching, synthetic

pytestmark = pytest.mark.gpu
CFG = matching.DEFAULT_CFG
IP = (CFG["max_iter"], CFG["lambda_init"], CFG["convergence_thresh"])
RF = (CFG["radius"], CFG["dilation_max"])


def dev():
    return torch.device("cuda:0")


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev())


def bits(t):
    return t.contiguous().view(torch.int32)


def test_golden_matching(backend):
    g = np.load(os.path.join(GOLDEN, "matching_64x48.npz"))
    p_new, conv = backend.iter_proj(T(g["rays"]), T(g["pts"]), T(g["p_init"]), *IP)
    assert torch.equal(bits(p_new), bits(T(g["p_new"]))) and torch.equal(conv, T(g["converged"]))
    (p1,) = backend.refine_matches(T(g["D11"]), T(g["D21"]), T(g["p1"]), *RF)
    assert torch.equal(p1, T(g["p1_refined"]))
    (p1f,) = backend.refine_matches(T(g["D11"]).float(), T(g["D21"]).float(), T(g["p1"]), *RF)
    assert torch.equal(p1f, T(g["p1_refined_f32"]))


@pytest.mark.parametrize("h,w,b", [(48, 64, 1), (37, 53, 2), (384, 512, 1)])
def test_matching_vs_oracle(backend, ref_backend, h, w, b):
    """Whole driver (prep + iter_proj + occlusion test + refine) on seeded synthetic pairs;
    37x53 gives n = 1961, not a multiple of the CTA size (ragged tail; the reference would
    read out of bounds there, matching_kernels.cu:35,130)."""
    X11, X21, D11, D21 = [t.to(dev()) for t in synthetic.make_pair(seed=h + w, h=h, w=w, fdim=24, b=b)]
    rays, pts, p_init = matching.prep_for_iter_proj(X11, X21, None)
    p_new, conv = backend.iter_proj(rays, pts, p_init, *IP)
    p_o, c_o = oracle.iter_proj(rays.cpu(), pts.cpu(), p_init.cpu(), *IP)
    assert torch.equal(bits(p_new).cpu(), bits(p_o)) and torch.equal(conv.cpu(), c_o)
    p1 = p_new.long()
    D11h, D21h = D11.half(), D21.view(b, h * w, -1).half().contiguous()
    (r,) = backend.refine_matches(D11h, D21h, p1, *RF)
    (r_o,) = oracle.refine_matches(D11h.cpu(), D21h.cpu(), p1.cpu(), *RF)
    assert torch.equal(r.cpu(), r_o)
    assert (r != p1).any()
    if ref_backend is not None and (h * w) % 16 == 0:
        p_r, c_r = ref_backend.iter_proj(rays, pts, p_init, *IP)
        assert torch.equal(bits(p_new), bits(p_r)) and torch.equal(conv, c_r)
        (r_r,) = ref_backend.refine_matches(D11h, D21h, p1, *RF)
        assert torch.equal(r, r_r)
    # end-to-end driver: same indices and masks as the identical glue run on the oracle
    idx_g, val_g = matching.match(X11, X21, D11, D21, backend=backend)
    idx_o, val_o = matching.match(X11.cpu(), X21.cpu(), D11.cpu(), D21.cpu(), backend=oracle)
    if torch.equal(rays.cpu(), matching.prep_for_iter_proj(X11.cpu(), X21.cpu(), None)[0]):
        assert torch.equal(idx_g.cpu(), idx_o) and torch.equal(val_g.cpu(), val_o)
    assert idx_g.min() >= 0 and idx_g.max() < h * w


def test_iter_proj_with_previous_matches_as_init(backend):
    """tracker.py:31 passes the previous frame's matches as the initial guess."""
    h, w = 96, 128
    X11, X21, D11, D21 = [t.to(dev()) for t in synthetic.make_pair(seed=21, h=h, w=w)]
    idx0, _ = matching.match(X11, X21, D11, D21, backend=backend)
    rays, pts, p_init = matching.prep_for_iter_proj(X11, X21, idx0)
    p_new, conv = backend.iter_proj(rays, pts, p_init, *IP)
    p_o, c_o = oracle.iter_proj(rays.cpu(), pts.cpu(), p_init.cpu(), *IP)
    assert torch.equal(bits(p_new).cpu(), bits(p_o)) and torch.equal(conv.cpu(), c_o)


def test_iter_proj_edge_cases(backend):
    h, w = 24, 40
    g = torch.Generator().manual_seed(0)
    rays = torch.randn(1, h, w, 9, generator=g)
    pts = torch.nn.functional.normalize(torch.randn(1, 300, 3, generator=g), dim=-1)
    # wild initial guesses: far outside, negative, exactly on the clamp borders
    p_init = torch.rand(1, 300, 2, generator=g) * 200 - 80
    p_init[0, :4] = torch.tensor([[1.0, 1.0], [w - 2.0, h - 2.0], [0.0, 0.0], [1e9, -1e9]])
    for it in (0, 1, 10):
        p_new, conv = backend.iter_proj(rays.to(dev()), pts.to(dev()), p_init.to(dev()), it, 1e-8, 1e-6)
        p_o, c_o = oracle.iter_proj(rays, pts, p_init, it, 1e-8, 1e-6)
        assert torch.equal(bits(p_new).cpu(), bits(p_o)) and torch.equal(conv.cpu(), c_o)
        assert p_new[..., 0].min() >= 1 and p_new[..., 0].max() <= w - 2
    # empty inputs
    e = backend.iter_proj(rays.to(dev()), pts[:, :0].contiguous().to(dev()), p_init[:, :0].contiguous().to(dev()), 10, 1e-8, 1e-6)
    assert e[0].shape == (1, 0, 2) and e[1].shape == (1, 0)


def test_refine_edge_cases(backend):
    h, w, fdim = 20, 28, 24
    g = torch.Generator().manual_seed(1)
    D11 = torch.nn.functional.normalize(torch.randn(1, h, w, fdim, generator=g), dim=-1).half()
    D21 = torch.nn.functional.normalize(torch.randn(1, 200, fdim, generator=g), dim=-1).half()
    p1 = torch.stack([torch.randint(0, w, (1, 200), generator=g), torch.randint(0, h, (1, 200), generator=g)], -1)
    p1[0, :4] = torch.tensor([[0, 0], [w - 1, h - 1], [0, h - 1], [w - 1, 0]])  # corners: windows leave the image
    for radius, dil in ((3, 5), (1, 1), (2, 3), (3, 0), (0, 2)):
        (r,) = backend.refine_matches(D11.to(dev()), D21.to(dev()), p1.to(dev()), radius, dil)
        (r_o,) = oracle.refine_matches(D11, D21, p1, radius, dil)
        assert torch.equal(r.cpu(), r_o), (radius, dil)
    # ties + all-non-positive scores: constant descriptors (every candidate ties -> first wins),
    # and negated queries (no score > 0 -> the input pixel is kept)
    Dc = torch.full((1, h, w, fdim), 0.2041).half()
    for q in (Dc[:, 0, :1].expand(1, 200, fdim).contiguous(), -Dc[:, 0, :1].expand(1, 200, fdim).contiguous()):
        (r,) = backend.refine_matches(Dc.to(dev()), q.to(dev()), p1.to(dev()), 3, 5)
        (r_o,) = oracle.refine_matches(Dc, q, p1, 3, 5)
        assert torch.equal(r.cpu(), r_o)
    assert torch.equal(r.cpu(), p1)
    # other descriptor widths: vector path (16, 32) and generic path (7)
    for fd in (16, 32, 7):
        A = torch.randn(1, h, w, fd, generator=g).half()
        B = torch.randn(1, 200, fd, generator=g).half()
        (r,) = backend.refine_matches(A.to(dev()), B.to(dev()), p1.to(dev()), 3, 5)
        (r_o,) = oracle.refine_matches(A, B, p1, 3, 5)
        assert torch.equal(r.cpu(), r_o), fd


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_refine_float_and_double_instantiations(backend, ref_backend, dtype):
    """The reference dispatches refine_matches over double, float and Half (matching_kernels.cu:102); the drivers only
    use Half, but a drop-in must take all three: same indices as the oracle and as the reference build."""
    h, w, fdim, n = 48, 64, 24, 48 * 64
    g = torch.Generator().manual_seed(5)
    D11 = torch.nn.functional.normalize(torch.randn(2, h, w, fdim, generator=g, dtype=torch.float64), dim=-1).to(dtype)
    D21 = torch.nn.functional.normalize(torch.randn(2, n, fdim, generator=g, dtype=torch.float64), dim=-1).to(dtype)
    p1 = torch.stack([torch.randint(0, w, (2, n), generator=g), torch.randint(0, h, (2, n), generator=g)], -1)
    (r,) = backend.refine_matches(D11.to(dev()), D21.to(dev()), p1.to(dev()), *RF)
    (r_o,) = oracle.refine_matches(D11, D21, p1, *RF)
    assert torch.equal(r.cpu(), r_o) and (r.cpu() != p1).any()
    if ref_backend is not None:
        (r_r,) = ref_backend.refine_matches(D11.to(dev()), D21.to(dev()), p1.to(dev()), *RF)
        assert torch.equal(r, r_r)


def test_error_convention(backend):
    x = torch.zeros(1, 8, 8, 18, device=dev())[..., ::2]
    with pytest.raises(RuntimeError, match="must be contiguous"):  # gn.h:5 CHECK_CONTIGUOUS
        backend.iter_proj(x, torch.zeros(1, 64, 3, device=dev()), torch.zeros(1, 64, 2, device=dev()), 10, 1e-8, 1e-6)
    with pytest.raises(RuntimeError):
        backend.refine_matches(torch.zeros(1, 8, 8, 24, device=dev(), dtype=torch.bfloat16),
                               torch.zeros(1, 64, 24, device=dev(), dtype=torch.bfloat16),
                               torch.zeros(1, 64, 2, device=dev(), dtype=torch.int64), 3, 5)


@pytest.mark.parametrize("h,w", [(40, 56), (45, 72), (50, 96)])
def test_refine_tiled_kernel_incoherent_and_out_of_range(backend, h, w):
    """n == h*w selects the tiled (bulk-copy staged) kernel when w >= 32 + 6*dilation_max + 2 (w = 56: dilation
    1 and 3 only; 72 and 96: all three, with a ragged last tile column and row bands of unequal height).
    Random centres make every tile incoherent (global-memory path per query, lanes re-dealt at every
    level), a coherent half exercises the staged path, and centres far outside the image / beyond int32
    must be returned untouched."""
    fdim = 24
    n = h * w
    g = torch.Generator().manual_seed(5)
    D11 = torch.nn.functional.normalize(torch.randn(2, h, w, fdim, generator=g), dim=-1).half()
    D21 = torch.nn.functional.normalize(torch.randn(2, n, fdim, generator=g), dim=-1).half()
    p1 = torch.stack([torch.randint(0, w, (2, n), generator=g), torch.randint(0, h, (2, n), generator=g)], -1)
    ident = torch.stack([torch.arange(n) % w, torch.arange(n) // w], -1)
    p1[1] = (ident + torch.tensor([3, -2])).clamp_min(0)  # coherent shift (second batch row)
    p1[1, :, 0].clamp_(max=w - 1)
    p1[1, :, 1].clamp_(max=h - 1)
    p1[0, :6] = torch.tensor([[-1, 5], [w, 5], [5, -16], [-100, -100], [2 ** 40, 3], [7, -(2 ** 35)]])
    for dil in (5, 1, 3):
        (r,) = backend.refine_matches(D11.to(dev()), D21.to(dev()), p1.to(dev()), 3, dil)
        (r_o,) = oracle.refine_matches(D11, D21, p1, 3, dil)
        assert torch.equal(r.cpu(), r_o), dil
    assert torch.equal(r.cpu()[0, 3:6], p1[0, 3:6])
