"""GPU tier, collected last: the allocation-free entry point `b200slam_iter_proj_static` (no counterpart in the
reference; its planar scratch is a static device array of the library) called through the C ABI and compared,
bit for bit (the bar `tests/test_gpu_matching.py` uses for iter_proj), with `b200slam_iter_proj` on two calls in a
row that share the scratch."""
import ctypes

import pytest
import torch

from splatt3r_slam_b200 import matching, synthetic

pytestmark = pytest.mark.gpu
CFG = matching.DEFAULT_CFG
IP = (CFG["max_iter"], CFG["lambda_init"], CFG["convergence_thresh"])
ESTATIC_LIMIT = -3  # B200SLAM_ESTATIC_LIMIT


def _inputs(seed, h, w, b):
    dev = torch.device("cuda:0")
    X11, X21, _, _ = [t.to(dev) for t in synthetic.make_pair(seed=seed, h=h, w=w, fdim=24, b=b)]
    return matching.prep_for_iter_proj(X11, X21, None)


def _call_static(lib, rays, pts, p_init, p_new, conv):
    b, h, w, _ = rays.shape
    n = p_init.shape[1]
    vp = ctypes.c_void_p
    return lib.b200slam_iter_proj_static(
        vp(rays.data_ptr()), vp(pts.data_ptr()), vp(p_init.data_ptr()), vp(p_new.data_ptr()), vp(conv.data_ptr()),
        b, h, w, n, int(IP[0]), float(IP[1]), float(IP[2]), vp(torch.cuda.current_stream().cuda_stream))


def test_iter_proj_static_equals_iter_proj_on_two_calls_in_a_row(backend):
    lib = backend._lib
    # call 1 fills the whole static array (one 384x512 image = the stated maximum); call 2 is a smaller, batched
    # problem that finds call 1's planes in the scratch and must not see them
    cases = [_inputs(11, 384, 512, 1), _inputs(12, 48, 64, 2)]
    expected = [backend.iter_proj(r, p, q, *IP) for (r, p, q) in cases]
    outs = []
    for rays, pts, p_init in cases:
        b, n = p_init.shape[:2]
        outs.append((torch.full((b, n, 2), -7.0, device=rays.device), torch.zeros((b, n), dtype=torch.bool, device=rays.device)))
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for (rays, pts, p_init), (p_new, conv) in zip(cases, outs):
        assert _call_static(lib, rays, pts, p_init, p_new, conv) == 0
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    for (p_new, conv), (p_ref, c_ref) in zip(outs, expected):
        assert torch.equal(p_new.view(torch.int32), p_ref.view(torch.int32))
        assert torch.equal(conv, c_ref)
        assert conv.any()
    # and once more through the Python wrapper, outputs passed by the caller
    rays, pts, p_init = cases[0]
    p_new, conv = outs[0]
    p_new.fill_(0)
    backend.iter_proj_static(rays, pts, p_init, *IP, p_new, conv)
    assert torch.equal(p_new.view(torch.int32), expected[0][0].view(torch.int32)) and torch.equal(conv, expected[0][1])


def test_iter_proj_static_rejects_sizes_beyond_its_scratch(backend):
    rays, pts, p_init = _inputs(13, 384, 512, 1)
    rays2, pts2, p2 = [torch.cat([t, t]).contiguous() for t in (rays, pts, p_init)]  # b = 2: twice the maximum
    p_new = torch.full((2, p2.shape[1], 2), -7.0, device=rays.device)
    conv = torch.zeros((2, p2.shape[1]), dtype=torch.bool, device=rays.device)
    assert _call_static(backend._lib, rays2, pts2, p2, p_new, conv) == ESTATIC_LIMIT
    torch.cuda.synchronize()
    assert (p_new == -7.0).all() and not conv.any()  # nothing was launched
    with pytest.raises(RuntimeError, match="exceed the built-in"):
        backend.iter_proj_static(rays2, pts2, p2, *IP, p_new, conv)
