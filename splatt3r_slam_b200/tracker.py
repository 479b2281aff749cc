"""Host-side mirror of the pose optimisation in the reference's `splatt3r_slam/tracker.py:174-214`
(`FrameTracker.opt_pose_ray_dist_sim3`): same argument order and meaning, with the `self.cfg` scalars of
config/base.yaml:16-31 as the `cfg` dict.  Poses are lietorch Sim3 `.data` vectors (8,) = [t, q_xyzw, s]
instead of `lietorch.Sim3` objects (lietorch is not a dependency of this package): the caller wraps
the result as `lietorch.Sim3(T[None])`.  `backend` is the module exporting `track_ray_dist` — the
CUDA backend by default; tests pass `oracle.tracker_ref` wrapped to the same call on CPU tensors."""
DEFAULT_CFG = dict(max_iters=50, rel_error=1e-3, delta_norm=1e-3, huber=1.345, sigma_ray=0.003, sigma_dist=10.0,
                   sigma_pixel=1.0, sigma_depth=10.0, pixel_border=-10, depth_eps=1e-6, C_conf=0.0, Q_conf=1.5)


def opt_pose_ray_dist_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, cfg=None, backend=None):
    cfg = DEFAULT_CFG if cfg is None else cfg
    if backend is None:
        from . import mast3r_slam_backends as backend
    T_WCf_new, T_CkCf, _ = backend.track_ray_dist(Xf, Xk, T_WCf, T_WCk, Qk, valid, cfg["sigma_ray"], cfg["sigma_dist"],
                                                  cfg["huber"], cfg["max_iters"], cfg["rel_error"], cfg["delta_norm"])
    return T_WCf_new, T_CkCf


def opt_pose_calib_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, meas_k, valid_meas_k, K, img_size, cfg=None, backend=None):
    """tracker.py:216-266, same argument order."""
    cfg = DEFAULT_CFG if cfg is None else cfg
    if backend is None:
        from . import mast3r_slam_backends as backend
    T_WCf_new, T_CkCf, _ = backend.track_calib(Xf, Xk, T_WCf, T_WCk, Qk, valid, meas_k, valid_meas_k, K, img_size,
                                               cfg["sigma_pixel"], cfg["sigma_depth"], cfg["huber"], cfg["pixel_border"],
                                               cfg["depth_eps"], cfg["max_iters"], cfg["rel_error"], cfg["delta_norm"])
    return T_WCf_new, T_CkCf
