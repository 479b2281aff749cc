"""Drop-in replacement for the reference's `mast3r_slam_backends` extension module.

Same five functions, positional signatures, return lists, dtypes/layouts and in-place
semantics as the reference's pybind module (splatt3r_slam/backend/src/gn.cpp:116-123,
include/gn.h:22-117), so `matching.py:60,79` and `global_opt.py:140,190` of the reference run
unchanged on top of it.  This layer only checks arguments, allocates outputs, remaps
keyframe ids to rows (gn_kernels.cu:160-169) and forwards raw device pointers + the current
CUDA stream to the C ABI in libb200slam.so (include/b200slam.h).  There is NO CPU fallback:
importing works without a GPU (so the symbol table can be checked), calling does not.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.environ.get("B200SLAM_LIB") or os.path.join(_HERE, "libb200slam.so")  # override: A/B builds

GN_RAYS, GN_CALIB, GN_POINTS = 0, 1, 2


class GNParams(ctypes.Structure):
    _fields_ = [("kind", ctypes.c_int), ("sigma_a", ctypes.c_float), ("sigma_b", ctypes.c_float),
                ("C_thresh", ctypes.c_float), ("Q_thresh", ctypes.c_float), ("height", ctypes.c_int),
                ("width", ctypes.c_int), ("pixel_border", ctypes.c_int), ("z_eps", ctypes.c_float),
                ("kf_slots", ctypes.c_void_p)]  # optional device (N) i64: row k reads Xs / Cs row kf_slots[k]


def _load():
    if not os.path.exists(_LIB_PATH):
        raise ImportError(
            f"{_LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  There is no CPU fallback for this backend.")
    lib = ctypes.CDLL(_LIB_PATH)
    vp, i, f, sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_size_t
    P = ctypes.POINTER(GNParams)
    lib.b200slam_version.restype = i
    lib.b200slam_iter_proj.argtypes = [vp] * 5 + [i] * 5 + [f, f, vp, sz, vp]
    lib.b200slam_iter_proj_static.argtypes = [vp] * 5 + [i] * 5 + [f, f, vp]
    lib.b200slam_iter_proj_workspace_bytes.argtypes = [i, i, i]
    lib.b200slam_iter_proj_workspace_bytes.restype = sz
    lib.b200slam_refine_matches_f16.argtypes = [vp] * 4 + [i] * 7 + [vp]
    lib.b200slam_refine_matches_f32.argtypes = [vp] * 4 + [i] * 7 + [vp]
    lib.b200slam_refine_matches_f64.argtypes = [vp] * 4 + [i] * 7 + [vp]
    lib.b200slam_gn_workspace_bytes.argtypes = [i, i, i]
    lib.b200slam_gn_workspace_bytes.restype = sz
    lib.b200slam_gn_build_blocks.argtypes = [P] + [vp] * 9 + [i] * 5 + [vp, vp, vp, sz, vp]
    lib.b200slam_gn_edge_blocks.argtypes = [P] + [vp] * 9 + [i] * 5 + [vp, vp, sz, vp]
    lib.b200slam_gn_edge_blocks_push.argtypes = [P] + [vp] * 9 + [i] * 5 + [vp, vp, i, i, ctypes.c_uint, vp, sz, vp]
    lib.b200slam_gn_wait_peers.argtypes = [vp, i, ctypes.c_uint, vp]
    lib.b200slam_gn_solve_retract.argtypes = [vp] * 4 + [i, i, i, f, vp, vp, sz, vp]
    lib.b200slam_gn_reset.argtypes = [vp, sz, vp]
    lib.b200slam_gn_status.argtypes = [vp, ctypes.POINTER(i), ctypes.POINTER(i), ctypes.POINTER(i), vp]
    D = ctypes.POINTER(ctypes.c_double)
    lib.b200slam_gn_solver_status.argtypes = [vp] + [ctypes.POINTER(i)] * 4 + [D, D, ctypes.POINTER(ctypes.c_ulonglong), vp]
    lib.b200slam_gn_set_solver.argtypes = [i]
    lib.b200slam_gn_analyze.argtypes = [vp, vp, i, i, i, vp, sz, vp]
    lib.b200slam_gn_sparse_host_solve.argtypes = [vp, vp, vp, i, i, vp, vp]
    lib.b200slam_gauss_newton.argtypes = [P] + [vp] * 9 + [i] * 4 + [f, vp, vp, sz, vp]
    lib.b200slam_gauss_newton_ids.argtypes = [P] + [vp] * 9 + [i] * 4 + [f, vp, vp, sz, vp]
    lib.b200slam_edge_rows.argtypes = [vp, vp, i, vp, vp, sz, vp]
    lib.b200slam_gauss_newton_twoway.argtypes = [P] + [vp] * 12 + [i] * 4 + [f, vp, vp, sz, vp]
    lib.b200slam_track_workspace_bytes.argtypes = []
    lib.b200slam_track_workspace_bytes.restype = sz
    lib.b200slam_track_ray_dist.argtypes = [vp] * 8 + [i, f, f, f, i, f, f, vp, sz, vp]
    lib.b200slam_track_calib.argtypes = [vp] * 10 + [i, i, i, f, f, f, i, f, i, f, f, vp, sz, vp]
    lib.b200slam_track_status.argtypes = [vp, ctypes.POINTER(i), ctypes.POINTER(i), ctypes.POINTER(i),
                                          ctypes.POINTER(ctypes.c_double), vp]
    lib.b200slam_match_workspace_bytes.argtypes = [i] * 5
    lib.b200slam_match_workspace_bytes.restype = sz
    lib.b200slam_match.argtypes = [vp] * 4 + [i] + [vp] * 3 + [i] * 5 + [f, f, f, i, i, vp, sz, vp]
    for name in ("iter_proj", "iter_proj_static", "refine_matches_f16", "refine_matches_f32", "refine_matches_f64", "gn_build_blocks",
                 "gn_edge_blocks", "gn_edge_blocks_push", "gn_wait_peers", "gn_solve_retract", "gn_reset", "gn_status", "gn_solver_status", "gn_set_solver", "gn_analyze", "gn_sparse_host_solve", "gauss_newton",
                 "gauss_newton_ids", "gauss_newton_twoway", "edge_rows", "match", "track_ray_dist", "track_calib", "track_status"):
        getattr(lib, "b200slam_" + name).restype = i
    return lib


_lib = _load()
EXPORTS = ("gauss_newton_points", "gauss_newton_rays", "gauss_newton_calib", "iter_proj", "refine_matches")


def _check(rc, what):
    if rc != 0:
        raise RuntimeError(f"b200slam {what} failed with code {rc}")


def _contig(**tensors):
    # gn.h:5 CHECK_CONTIGUOUS -> RuntimeError("<name> must be contiguous")
    for name, t in tensors.items():
        if not t.is_contiguous():
            raise RuntimeError(f"{name} must be contiguous")


def _cuda(t, dtype, name):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (this backend has no CPU path)")
    if t.dtype != dtype:
        raise RuntimeError(f"expected scalar type {dtype} for {name} but found {t.dtype}")
    return t.data_ptr()


def _stream(t):
    return torch.cuda.current_stream(t.device).cuda_stream


class _NoGuard:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


_NOGUARD = _NoGuard()


def _guard(t):
    """Device guard only when needed (torch.cuda.device costs ~10 us per call otherwise)."""
    return _NOGUARD if t.device.index == torch.cuda.current_device() else torch.cuda.device(t.device)


def iter_proj(rays_img_with_grad, pts_3d_norm, p_init, max_iter, lambda_init, cost_thresh):
    _contig(rays_img_with_grad=rays_img_with_grad, pts_3d_norm=pts_3d_norm, p_init=p_init)
    b, h, w, c = rays_img_with_grad.shape
    if c != 9:
        raise RuntimeError("rays_img_with_grad must have 9 channels (ray, d/du, d/dv)")
    n = p_init.shape[1]
    ptrs = (_cuda(rays_img_with_grad, torch.float32, "rays_img_with_grad"),
            _cuda(pts_3d_norm, torch.float32, "pts_3d_norm"), _cuda(p_init, torch.float32, "p_init"))
    with _guard(p_init):
        p_new = torch.empty((b, n, 2), dtype=torch.float32, device=p_init.device)
        conv = torch.empty((b, n), dtype=torch.bool, device=p_init.device)
        nbytes = _lib.b200slam_iter_proj_workspace_bytes(b, h, w)
        ws = torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=p_init.device)
        _check(_lib.b200slam_iter_proj(
            *ptrs, p_new.data_ptr(), conv.data_ptr(), b, h, w, n, int(max_iter), float(lambda_init),
            float(cost_thresh), ws.data_ptr(), nbytes, _stream(p_init)), "iter_proj")
    return [p_new, conv]


ITER_PROJ_STATIC_MAX_BYTES = 196608 * 9 * 4  # B200SLAM_ITER_PROJ_STATIC_MAX_BYTES, include/b200slam.h


def iter_proj_static(rays_img_with_grad, pts_3d_norm, p_init, max_iter, lambda_init, cost_thresh, p_new, converged):
    """Allocation-free `iter_proj` (`b200slam_iter_proj_static`; no counterpart in the reference): the caller passes
    the outputs `p_new` (b,n,2) f32 and `converged` (b,n) bool, the planar scratch is a static device array of the
    library (at most ITER_PROJ_STATIC_MAX_BYTES = one 512x384 image).  Neither this wrapper nor the entry point
    allocates; calls that overlap on the GPU share the scratch and race."""
    _contig(rays_img_with_grad=rays_img_with_grad, pts_3d_norm=pts_3d_norm, p_init=p_init, p_new=p_new,
            converged=converged)
    b, h, w, c = rays_img_with_grad.shape
    if c != 9:
        raise RuntimeError("rays_img_with_grad must have 9 channels (ray, d/du, d/dv)")
    n = p_init.shape[1]
    if tuple(p_new.shape) != (b, n, 2) or tuple(converged.shape) != (b, n):
        raise RuntimeError("iter_proj_static: p_new must be (b,n,2) and converged (b,n)")
    if b * h * w * 36 > ITER_PROJ_STATIC_MAX_BYTES:
        raise RuntimeError(f"iter_proj_static: b*h*w*36 = {b * h * w * 36} bytes of scratch exceed the built-in "
                           f"{ITER_PROJ_STATIC_MAX_BYTES} (one 512x384 image); use iter_proj")
    ptrs = (_cuda(rays_img_with_grad, torch.float32, "rays_img_with_grad"),
            _cuda(pts_3d_norm, torch.float32, "pts_3d_norm"), _cuda(p_init, torch.float32, "p_init"),
            _cuda(p_new, torch.float32, "p_new"), _cuda(converged, torch.bool, "converged"))
    with _guard(p_init):
        _check(_lib.b200slam_iter_proj_static(*ptrs, b, h, w, n, int(max_iter), float(lambda_init),
                                              float(cost_thresh), _stream(p_init)), "iter_proj_static")
    return [p_new, converged]


def refine_matches(D11, D21, p1, radius, dilation_max):
    _contig(D11=D11, D21=D21, p1=p1)
    b, h, w, fdim = D11.shape
    n = p1.shape[1]
    if D11.dtype == torch.float16:
        fn = _lib.b200slam_refine_matches_f16
    elif D11.dtype == torch.float32:
        fn = _lib.b200slam_refine_matches_f32
    elif D11.dtype == torch.float64:
        fn = _lib.b200slam_refine_matches_f64
    else:
        raise RuntimeError(f"refine_matches: unsupported descriptor dtype {D11.dtype}")
    ptrs = (_cuda(D11, D11.dtype, "D11"), _cuda(D21, D11.dtype, "D21"), _cuda(p1, torch.int64, "p1"))
    with _guard(p1):
        p1_new = torch.empty((b, n, 2), dtype=torch.int64, device=p1.device)
        _check(fn(*ptrs, p1_new.data_ptr(), b, h, w, n, fdim, int(radius), int(dilation_max), _stream(p1)),
               "refine_matches")
    return [p1_new]


def match(X11, X21, D11, D21, idx_1_to_2_init, max_iter, lambda_init, cost_thresh, dist_thresh, radius,
          dilation_max):
    """Fused `match_iterative_proj` (reference matching.py:52-90 including prep_for_iter_proj :25-49
    and the occlusion test :66-76): X11, X21 (b,h,w,3) f32; D11 (b,h,w,fdim), D21 (b,h,w,fdim) or
    (b,h*w,fdim), f16 or f32 (f32 is rounded to f16 like the reference's `.half()`);
    idx_1_to_2_init (b,h*w) int64 or None.  Returns [idx_1_to_2 (b,h*w) int64, valid (b,h*w,1) bool] —
    bit-identical to the unfused glue + iter_proj + refine_matches (an entry
    point the reference does not have; `splatt3r_slam_b200.matching.match(..., fused=True)` uses it)."""
    _contig(X11=X11, X21=X21, D11=D11, D21=D21)
    b, h, w, _ = X11.shape
    fdim = D11.shape[-1]
    if D11.dtype not in (torch.float16, torch.float32) or D21.dtype != D11.dtype:
        raise RuntimeError(f"match: unsupported descriptor dtypes {D11.dtype}, {D21.dtype}")
    ptrs = (_cuda(X11, torch.float32, "X11"), _cuda(X21, torch.float32, "X21"), _cuda(D11, D11.dtype, "D11"),
            _cuda(D21, D11.dtype, "D21"))
    f32 = int(D11.dtype == torch.float32)
    init = None
    if idx_1_to_2_init is not None:
        _contig(idx_1_to_2_init=idx_1_to_2_init)
        init = _cuda(idx_1_to_2_init, torch.int64, "idx_1_to_2_init")
    with _guard(X11):
        idx = torch.empty((b, h * w), dtype=torch.int64, device=X11.device)
        valid = torch.empty((b, h * w, 1), dtype=torch.bool, device=X11.device)
        nbytes = _lib.b200slam_match_workspace_bytes(b, h, w, fdim, f32)
        key = ("match", X11.device.index, _stream(X11))  # per (device, stream): two streams must not share scratch
        ws = _workspaces.get(key)
        if ws is None or ws.numel() < nbytes:
            ws = _workspaces[key] = torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=X11.device)
        _check(_lib.b200slam_match(*ptrs, f32, init, idx.data_ptr(), valid.data_ptr(), b, h, w, fdim, int(max_iter),
                                   float(lambda_init), float(cost_thresh), float(dist_thresh), int(radius),
                                   int(dilation_max), ws.data_ptr(), nbytes, _stream(X11)), "match")
    return [idx, valid]


def track_ray_dist(Xf, Xk, T_WCf, T_WCk, Qk, valid, sigma_ray, sigma_dist, huber, max_iters, rel_error, delta_norm,
                   check=True):
    """On-device `FrameTracker.opt_pose_ray_dist_sim3` (reference tracker.py:174-214):
    Xf (n,3) matched frame points, Xk (n,3) keyframe points, T_WCf / T_WCk Sim3 data (8,) or (1,8)
    [t, q_xyzw, s], Qk (n,1) or (n,), valid (n,1) or (n,) bool.  Returns
    [T_WCf_new (8,), T_CkCf (8,), info] with info = dict(iters, converged, cost) when `check`
    (one 64-byte device->host read; the reference syncs once per iteration) else None.  A non-SPD
    system raises RuntimeError like the reference's torch.linalg.cholesky (tracker.py:88-90)."""
    _contig(Xf=Xf, Xk=Xk, T_WCf=T_WCf, T_WCk=T_WCk, Qk=Qk, valid=valid)
    n = Xk.shape[0]
    if Xf.shape != (n, 3) or Xk.shape != (n, 3) or Qk.numel() != n or valid.numel() != n:
        raise RuntimeError("track_ray_dist: Xf, Xk must be (n,3) and Qk, valid must hold n values")
    if T_WCf.numel() != 8 or T_WCk.numel() != 8:
        raise RuntimeError("track_ray_dist: poses must be Sim3 data vectors of 8 floats")
    ptrs = (_cuda(Xf, torch.float32, "Xf"), _cuda(Xk, torch.float32, "Xk"), _cuda(Qk, torch.float32, "Qk"),
            _cuda(valid, torch.bool, "valid"), _cuda(T_WCf, torch.float32, "T_WCf"),
            _cuda(T_WCk, torch.float32, "T_WCk"))
    with _guard(Xk):
        out = torch.empty((2, 8), dtype=torch.float32, device=Xk.device)
        ws, nbytes = _track_workspace(Xk.device)
        st = _stream(Xk)
        _check(_lib.b200slam_track_ray_dist(*ptrs, out[0].data_ptr(), out[1].data_ptr(), n, float(sigma_ray),
                                            float(sigma_dist), float(huber), int(max_iters), float(rel_error),
                                            float(delta_norm), ws.data_ptr(), nbytes, st), "track_ray_dist")
        info = _track_info(ws, st, "track_ray_dist") if check else None
    return [out[0], out[1], info]


def track_calib(Xf, Xk, T_WCf, T_WCk, Qk, valid, meas_k, valid_meas_k, K, img_size, sigma_pixel, sigma_depth, huber,
                pixel_border, depth_eps, max_iters, rel_error, delta_norm, check=True):
    """On-device `FrameTracker.opt_pose_calib_sim3` (reference tracker.py:216-266): arguments in the
    reference's order (Xk is unused there too: the keyframe enters through meas_k (n,3) = [u, v, log z]
    and valid_meas_k (n,1) bool, tracker.py:143-152); K (3,3) f32 on the device, img_size = (h, w).
    Returns [T_WCf_new, T_CkCf, info] like `track_ray_dist`."""
    _contig(Xf=Xf, T_WCf=T_WCf, T_WCk=T_WCk, Qk=Qk, valid=valid, meas_k=meas_k, valid_meas_k=valid_meas_k, K=K)
    n = Xf.shape[0]
    if Xf.shape != (n, 3) or meas_k.shape != (n, 3) or Qk.numel() != n or valid.numel() != n or valid_meas_k.numel() != n:
        raise RuntimeError("track_calib: Xf, meas_k must be (n,3) and Qk, valid, valid_meas_k must hold n values")
    if T_WCf.numel() != 8 or T_WCk.numel() != 8 or K.shape != (3, 3):
        raise RuntimeError("track_calib: poses must be 8 floats and K (3,3)")
    ptrs = (_cuda(Xf, torch.float32, "Xf"), _cuda(meas_k, torch.float32, "meas_k"), _cuda(Qk, torch.float32, "Qk"),
            _cuda(valid, torch.bool, "valid"), _cuda(valid_meas_k, torch.bool, "valid_meas_k"),
            _cuda(K, torch.float32, "K"), _cuda(T_WCf, torch.float32, "T_WCf"), _cuda(T_WCk, torch.float32, "T_WCk"))
    with _guard(Xf):
        out = torch.empty((2, 8), dtype=torch.float32, device=Xf.device)
        ws, nbytes = _track_workspace(Xf.device)
        st = _stream(Xf)
        _check(_lib.b200slam_track_calib(*ptrs, out[0].data_ptr(), out[1].data_ptr(), n, int(img_size[0]),
                                         int(img_size[1]), float(sigma_pixel), float(sigma_depth), float(huber),
                                         int(pixel_border), float(depth_eps), int(max_iters), float(rel_error),
                                         float(delta_norm), ws.data_ptr(), nbytes, st), "track_calib")
        info = _track_info(ws, st, "track_calib") if check else None
    return [out[0], out[1], info]


def _track_workspace(device):
    nbytes = _lib.b200slam_track_workspace_bytes()
    key = ("track", device.index, torch.cuda.current_stream(device).cuda_stream)
    ws = _workspaces.get(key)
    if ws is None:
        ws = _workspaces[key] = torch.empty(int(nbytes), dtype=torch.uint8, device=device)
    return ws, nbytes


def _track_info(ws, st, what):
    it, cv, cf, cost = ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_double()
    _check(_lib.b200slam_track_status(ws.data_ptr(), ctypes.byref(it), ctypes.byref(cv), ctypes.byref(cf),
                                      ctypes.byref(cost), st), "track_status")
    if cf.value:
        raise RuntimeError(f"{what}: linalg.cholesky: the normal equations are not positive-definite")
    return dict(iters=it.value, converged=bool(cv.value), cost=cost.value)


def edge_rows(ii, jj):
    """Keyframe ids -> rows of Twc/Xs/Cs: dense rank among the sorted unique ids
    (reference: torch::_unique + searchsorted, gn_kernels.cu:160-169).  No host sync.
    CUDA tensors go through the library's kernel; CPU tensors (tests, gloo ranks) through torch."""
    E = ii.numel()
    if ii.is_cuda and E > 0:
        rows = torch.empty(2 * E, dtype=torch.int64, device=ii.device)
        scratch = torch.empty(2 * E, dtype=torch.uint8, device=ii.device)
        with _guard(ii):
            _check(_lib.b200slam_edge_rows(_cuda(ii.contiguous(), torch.int64, "ii"),
                                           _cuda(jj.contiguous(), torch.int64, "jj"), E, rows.data_ptr(),
                                           scratch.data_ptr(), 2 * E, _stream(ii)), "edge_rows")
        return rows[:E], rows[E:]
    s, order = torch.sort(torch.cat([ii, jj]))
    rank_sorted = torch.cumsum(torch.cat([s.new_zeros(1), (s[1:] != s[:-1]).to(s.dtype)]), 0)
    rows = torch.empty_like(rank_sorted)
    rows[order] = rank_sorted
    return rows[:E].contiguous(), rows[E:].contiguous()


_workspaces = {}


def _workspace(device, nbytes):
    """GN workspace, cached per (device, current stream): calls on different streams of one device must not share
    the State / partials / ticket buffers (the reference allocates Hs/gs per call).  A workspace that has to grow
    is replaced; the old tensor was only ever used on this same stream, so handing it back to the caching
    allocator is ordered.  The last workspace used on a device is remembered for gn_status()."""
    index = device.index if device.index is not None else torch.cuda.current_device()
    key = (device.type, index, torch.cuda.current_stream(device).cuda_stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(int(nbytes), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    _workspaces[(device.type, index)] = ws
    return ws


def _gn_ptrs(Twc, Xs, Cs, K, ii_row, jj_row, idx, valid, Q):
    return (_cuda(Twc, torch.float32, "Twc"), _cuda(Xs, torch.float32, "Xs"), _cuda(Cs, torch.float32, "Cs"),
            (_cuda(K, torch.float32, "K") if K is not None else None), ii_row.data_ptr(), jj_row.data_ptr(),
            _cuda(idx, torch.int64, "idx_ii2jj"), _cuda(valid, torch.bool, "valid_match"),
            _cuda(Q, torch.float32, "Q"))


def _gauss_newton(prm, Twc, Xs, Cs, K, ii, jj, idx, valid, Q, max_iter, delta_thresh):
    _contig(Twc=Twc, Xs=Xs, Cs=Cs, ii=ii, jj=jj, idx_ii2jj=idx, valid_match=valid, Q=Q)
    if K is not None:
        _contig(K=K)
    N, n = Xs.shape[0], Xs.shape[1]
    E = ii.shape[0]
    ptrs = _gn_ptrs(Twc, Xs, Cs, K, ii, jj, idx, valid, Q)
    _cuda(ii, torch.int64, "ii"), _cuda(jj, torch.int64, "jj")
    with _guard(Twc):
        dx = torch.zeros((max(N - 1, 0), 7), dtype=torch.float32, device=Twc.device)
        if max_iter <= 0 or N < 2 or E == 0:
            return [dx]
        nbytes = _lib.b200slam_gn_workspace_bytes(N, E, n)
        ws = _workspace(Twc.device, nbytes)
        _check(_lib.b200slam_gauss_newton_ids(
            ctypes.byref(prm), *ptrs, N, E, n, int(max_iter), float(delta_thresh), dx.data_ptr(),
            ws.data_ptr(), nbytes, _stream(Twc)), "gauss_newton")
    return [dx]


def gauss_newton_rays(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_ray, sigma_dist, C_thresh,
                      Q_thresh, max_iter, delta_thresh):
    prm = GNParams(GN_RAYS, sigma_ray, sigma_dist, C_thresh, Q_thresh, 0, 0, 0, 0.0)
    return _gauss_newton(prm, Twc, Xs, Cs, None, ii, jj, idx_ii2jj, valid_match, Q, max_iter, delta_thresh)


def gauss_newton_calib(Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, height, width, pixel_border,
                       z_eps, sigma_pixel, sigma_depth, C_thresh, Q_thresh, max_iter, delta_thresh):
    prm = GNParams(GN_CALIB, sigma_pixel, sigma_depth, C_thresh, Q_thresh, int(height), int(width),
                   int(pixel_border), z_eps)
    return _gauss_newton(prm, Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, max_iter, delta_thresh)


def gauss_newton_points(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_point, C_thresh, Q_thresh,
                        max_iter, delta_thresh):
    prm = GNParams(GN_POINTS, sigma_point, 1.0, C_thresh, Q_thresh, 0, 0, 0, 0.0)
    return _gauss_newton(prm, Twc, Xs, Cs, None, ii, jj, idx_ii2jj, valid_match, Q, max_iter, delta_thresh)


def gauss_newton_twoway(prm, Twc, Xs, Cs, K, ii, jj, idx_ii2jj, idx_jj2ii, valid_match_j, valid_match_i, Q_ii2jj,
                        Q_jj2ii, max_iter, delta_thresh):
    """The GN loop on the factor graph's ONE-WAY edge lists (reference FactorGraph fields, global_opt.py:21-29):
    same result as `gauss_newton_*` on `prep_two_way_edges()`'s concatenations (global_opt.py:104-110), without
    making them.  prm: `make_params(...)`.  Updates Twc in place; returns [dx]."""
    _contig(Twc=Twc, Xs=Xs, Cs=Cs, ii=ii, jj=jj, idx_ii2jj=idx_ii2jj, idx_jj2ii=idx_jj2ii, valid_match_j=valid_match_j,
            valid_match_i=valid_match_i, Q_ii2jj=Q_ii2jj, Q_jj2ii=Q_jj2ii)
    if K is not None:
        _contig(K=K)
    N, n, E1 = Twc.shape[0], Xs.shape[1], ii.shape[0]  # (with prm.kf_slots, Xs / Cs are whole keyframe buffers)
    ptrs = (_cuda(Twc, torch.float32, "Twc"), _cuda(Xs, torch.float32, "Xs"), _cuda(Cs, torch.float32, "Cs"),
            (_cuda(K, torch.float32, "K") if K is not None else None), _cuda(ii, torch.int64, "ii"),
            _cuda(jj, torch.int64, "jj"), _cuda(idx_ii2jj, torch.int64, "idx_ii2jj"),
            _cuda(idx_jj2ii, torch.int64, "idx_jj2ii"), _cuda(valid_match_j, torch.bool, "valid_match_j"),
            _cuda(valid_match_i, torch.bool, "valid_match_i"), _cuda(Q_ii2jj, torch.float32, "Q_ii2jj"),
            _cuda(Q_jj2ii, torch.float32, "Q_jj2ii"))
    with _guard(Twc):
        dx = torch.zeros((max(N - 1, 0), 7), dtype=torch.float32, device=Twc.device)
        if max_iter <= 0 or N < 2 or E1 == 0:
            return [dx]
        nbytes = _lib.b200slam_gn_workspace_bytes(N, 2 * E1, n)
        ws = _workspace(Twc.device, nbytes)
        _check(_lib.b200slam_gauss_newton_twoway(
            ctypes.byref(prm), *ptrs, N, E1, n, int(max_iter), float(delta_thresh), dx.data_ptr(), ws.data_ptr(), nbytes,
            _stream(Twc)), "gauss_newton_twoway")
    return [dx]


# ---- extras beyond the reference module (used by tests, bench and the multi-GPU driver) ----------
def make_params(kind, sigma_a, sigma_b, C_thresh, Q_thresh, height=0, width=0, pixel_border=0, z_eps=0.0, kf_slots=None):
    """`kf_slots` (optional (N,) int64 CUDA tensor): row k of the solve reads its pointmap / confidence from row
    kf_slots[k] of Xs / Cs (keyframe buffers used in place); the tensor must outlive the calls made with the params."""
    prm = GNParams(kind, sigma_a, sigma_b, C_thresh, Q_thresh, int(height), int(width), int(pixel_border), z_eps, None)
    if kf_slots is not None:
        if not (kf_slots.is_cuda and kf_slots.dtype == torch.int64 and kf_slots.is_contiguous()):
            raise RuntimeError("kf_slots must be a contiguous int64 CUDA tensor")
        prm.kf_slots = kf_slots.data_ptr()
        prm._keep = kf_slots
    return prm


def gn_build_blocks(prm, Twc, Xs, Cs, K, ii, jj, idx, valid, Q):
    """One Hessian build in the reference's private layout: Hs (4,E,7,7), gs (2,E,7) fp32."""
    N, n, E = Xs.shape[0], Xs.shape[1], ii.shape[0]
    with _guard(Twc):
        ii_row, jj_row = edge_rows(ii, jj)
        Hs = torch.zeros((4, E, 7, 7), dtype=torch.float32, device=Twc.device)
        gs = torch.zeros((2, E, 7), dtype=torch.float32, device=Twc.device)
        nbytes = _lib.b200slam_gn_workspace_bytes(N, E, n)
        ws = _workspace(Twc.device, nbytes)
        _check(_lib.b200slam_gn_reset(ws.data_ptr(), nbytes, _stream(Twc)), "gn_reset")
        _check(_lib.b200slam_gn_build_blocks(
            ctypes.byref(prm), *_gn_ptrs(Twc, Xs, Cs, K, ii_row, jj_row, idx, valid, Q), N, E, n, 0, E,
            Hs.data_ptr(), gs.data_ptr(), ws.data_ptr(), nbytes, _stream(Twc)), "gn_build_blocks")
    return Hs, gs


def _ws_of(device):
    device = torch.device(device)
    index = device.index if device.index is not None else torch.cuda.current_device()
    return _workspaces[(device.type, index)]


def gn_status(device):
    return gn_status_ws(_ws_of(device))


SOLVERS = {"default": -1, "auto": 0, "dense": 1, "pcg1": 2}


def gn_set_solver(name):
    """Large-graph solver selection (process-wide): 'default' | 'auto' | 'dense' | 'pcg1'."""
    _check(_lib.b200slam_gn_set_solver(SOLVERS[name]), "gn_set_solver")


def gn_solver_status(ws):
    """Large-graph solver diagnostics of the last GN call on workspace `ws` (a device, or the uint8 tensor)."""
    if not isinstance(ws, torch.Tensor):
        ws = _ws_of(ws)
    v = [ctypes.c_int() for _ in range(4)]
    rel, true_rel = ctypes.c_double(), ctypes.c_double()
    stamps = (ctypes.c_ulonglong * 12)()
    _check(_lib.b200slam_gn_solver_status(ws.data_ptr(), *[ctypes.byref(x) for x in v], ctypes.byref(rel),
                                          ctypes.byref(true_rel), stamps,
                                          torch.cuda.current_stream(ws.device).cuda_stream), "gn_solver_status")
    return dict(cg_iters=v[0].value, cg_total=v[1].value, pcg_bad=v[2].value, dense_runs=v[3].value & 0xffff,
                sparse_runs=v[3].value >> 16,
                cg_rel=rel.value, cg_true_rel=true_rel.value,
                # last solver kernel: microseconds of {set-up, iterations, tail} (device %globaltimer)
                solve_us=[(stamps[k + 1] - stamps[k]) * 1e-3 if stamps[k + 1] >= stamps[k] else None for k in range(3)],
                cg_loop_cycles=dict(publish=stamps[4], wait=stamps[5], update=stamps[6], local=stamps[7], scalars=stamps[8],
                                    matvec=stamps[9], busy_max_over_ctas=stamps[10], busy_min_over_ctas=stamps[11]))


def gn_status_ws(ws):
    device = ws.device
    it, dn, cf = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    _check(_lib.b200slam_gn_status(ws.data_ptr(), ctypes.byref(it), ctypes.byref(dn), ctypes.byref(cf),
                                   torch.cuda.current_stream(device).cuda_stream), "gn_status")
    return dict(iters=it.value, done=dn.value, chol_fail=cf.value)
