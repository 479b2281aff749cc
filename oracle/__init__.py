"""CPU oracle (TEST INFRASTRUCTURE ONLY) — ctypes front-end to oracle/liboracle.so.

Mirrors the five `mast3r_slam_backends` entry points (gn.cpp:116-123 of the reference) on
CPU torch tensors so the parity tests read like calls into the backend.  May be imported
only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
The product package (splatt3r_slam_b200) never imports this module.
"""
import ctypes
import os
import subprocess

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build():
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("oracle_matching.c", "oracle_gn.c", "oracle_gn_solve.c")]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build())
        _LIB.oracle_gauss_newton.restype = ctypes.c_int
        _LIB.oracle_gn_solve.restype = ctypes.c_int
    return _LIB


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def _cpu(t, dtype):
    assert t.device.type == "cpu" and t.dtype == dtype and t.is_contiguous(), (t.device, t.dtype)
    return t


def iter_proj(rays_img_with_grad, pts_3d_norm, p_init, max_iter, lambda_init, cost_thresh):
    b, h, w, c = rays_img_with_grad.shape
    assert c == 9
    n = p_init.shape[1]
    _cpu(rays_img_with_grad, torch.float32), _cpu(pts_3d_norm, torch.float32), _cpu(p_init, torch.float32)
    p_new = torch.zeros(b, n, 2, dtype=torch.float32)
    conv = torch.zeros(b, n, dtype=torch.bool)
    lib().oracle_iter_proj(_p(rays_img_with_grad), _p(pts_3d_norm), _p(p_init), _p(p_new), _p(conv),
                           b, h, w, n, int(max_iter), ctypes.c_float(lambda_init),
                           ctypes.c_float(cost_thresh))
    return [p_new, conv]


def refine_matches(D11, D21, p1, radius, dilation_max):
    b, h, w, fdim = D11.shape
    n = p1.shape[1]
    _cpu(p1, torch.int64)
    p1_new = torch.zeros(b, n, 2, dtype=torch.int64)
    if D11.dtype == torch.float16:
        fn = lib().oracle_refine_matches_f16
    elif D11.dtype == torch.float32:
        fn = lib().oracle_refine_matches_f32
    elif D11.dtype == torch.float64:
        fn = lib().oracle_refine_matches_f64
    else:
        raise TypeError(D11.dtype)
    assert D21.dtype == D11.dtype and D11.is_contiguous() and D21.is_contiguous()
    fn(_p(D11), _p(D21), _p(p1), _p(p1_new), b, h, w, n, fdim, int(radius), int(dilation_max))
    return [p1_new]


def _rows(ii, jj):
    uniq = torch.unique(torch.cat([ii, jj]), sorted=True)
    return torch.searchsorted(uniq, ii).contiguous(), torch.searchsorted(uniq, jj).contiguous()


def gn_build(kind, Twc, Xs, Cs, K, ii, jj, idx, valid, Q, sigma_a, sigma_b, C_thresh, Q_thresh,
             height=0, width=0, pixel_border=0, z_eps=0.0):
    """One Hessian build: returns Hs (4,E,7,7), gs (2,E,7) like the reference's private buffers
    (gn_kernels.cu:1172-1173).  kind: 0 rays, 1 calib, 2 points.  ii/jj are keyframe ids."""
    E, n = idx.shape[0], Xs.shape[1]
    ir, jr = _rows(ii, jj)
    Hs = torch.zeros(4, E, 7, 7, dtype=torch.float32)
    gs = torch.zeros(2, E, 7, dtype=torch.float32)
    vb = valid.view(torch.uint8) if valid.dtype == torch.bool else valid
    lib().oracle_gn_build(int(kind), _p(_cpu(Twc, torch.float32)), _p(_cpu(Xs, torch.float32)),
                          _p(_cpu(Cs, torch.float32)), _p(K), _p(ir), _p(jr),
                          _p(_cpu(idx, torch.int64)), _p(vb.contiguous()), _p(_cpu(Q, torch.float32)),
                          E, n, ctypes.c_float(sigma_a), ctypes.c_float(sigma_b),
                          ctypes.c_float(C_thresh), ctypes.c_float(Q_thresh), int(height), int(width),
                          int(pixel_border), ctypes.c_float(z_eps), _p(Hs), _p(gs))
    return Hs, gs


def _gauss_newton(kind, Twc, Xs, Cs, K, ii, jj, idx, valid, Q, sigma_a, sigma_b, C_thresh, Q_thresh,
                  height, width, pixel_border, z_eps, max_iter, delta_thresh):
    N, n = Xs.shape[0], Xs.shape[1]
    E = ii.shape[0]
    dx = torch.zeros(N - 1, 7, dtype=torch.float32)
    vb = valid.view(torch.uint8) if valid.dtype == torch.bool else valid
    assert Twc.is_contiguous()
    iters = lib().oracle_gauss_newton(
        int(kind), _p(_cpu(Twc, torch.float32)), _p(_cpu(Xs, torch.float32)), _p(_cpu(Cs, torch.float32)),
        _p(K), _p(_cpu(ii, torch.int64)), _p(_cpu(jj, torch.int64)), _p(_cpu(idx, torch.int64)),
        _p(vb.contiguous()), _p(_cpu(Q, torch.float32)), N, E, n, ctypes.c_float(sigma_a),
        ctypes.c_float(sigma_b), ctypes.c_float(C_thresh), ctypes.c_float(Q_thresh), int(height),
        int(width), int(pixel_border), ctypes.c_float(z_eps), int(max_iter),
        ctypes.c_float(delta_thresh), _p(dx))
    _gauss_newton.last_iters = iters
    return [dx]


def gauss_newton_rays(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_ray, sigma_dist,
                      C_thresh, Q_thresh, max_iter, delta_thresh):
    return _gauss_newton(0, Twc, Xs, Cs, None, ii, jj, idx_ii2jj, valid_match, Q, sigma_ray,
                         sigma_dist, C_thresh, Q_thresh, 0, 0, 0, 0.0, max_iter, delta_thresh)


def gauss_newton_calib(Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, height, width,
                       pixel_border, z_eps, sigma_pixel, sigma_depth, C_thresh, Q_thresh, max_iter,
                       delta_thresh):
    return _gauss_newton(1, Twc, Xs, Cs, _cpu(K, torch.float32), ii, jj, idx_ii2jj, valid_match, Q,
                         sigma_pixel, sigma_depth, C_thresh, Q_thresh, height, width, pixel_border,
                         z_eps, max_iter, delta_thresh)


def gauss_newton_points(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_point, C_thresh,
                        Q_thresh, max_iter, delta_thresh):
    return _gauss_newton(2, Twc, Xs, Cs, None, ii, jj, idx_ii2jj, valid_match, Q, sigma_point, 1.0,
                         C_thresh, Q_thresh, 0, 0, 0, 0.0, max_iter, delta_thresh)


def gn_solve(Hs, gs, ii_opt, jj_opt, M):
    """SparseBlock assembly + LLT solve (gn_kernels.cu:56-158): dx (M,7) f32, ok flag."""
    E = ii_opt.shape[0]
    dx = torch.zeros(M, 7, dtype=torch.float32)
    ok = lib().oracle_gn_solve(_p(_cpu(Hs, torch.float32)), _p(_cpu(gs, torch.float32)),
                               _p(_cpu(ii_opt, torch.int64)), _p(_cpu(jj_opt, torch.int64)), E, M, _p(dx))
    return dx, bool(ok)


def pose_retr(Twc, dx, num_fix=1):
    """pose_retr_kernel (gn_kernels.cu:414-452), in place."""
    lib().oracle_pose_retr(_p(_cpu(Twc, torch.float32)), _p(_cpu(dx, torch.float32)), Twc.shape[0], num_fix)
