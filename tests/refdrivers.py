"""TEST INFRASTRUCTURE: import the reference's OWN Python drivers — `splatt3r_slam/matching.py`, `tracker.py`,
`global_opt.py`, `geometry.py`, `nonlinear_optimizer.py` — and its vendored lietorch, from the sourceless
byte-code + extension that `oracle/build_lietorch.py` compiled out of /root/reference into `oracle/_ref/`
(git-ignored; travels to the GPU box where /root/reference does not exist).

The drivers are executed UNCHANGED; only what they import but the hot path never touches is stubbed:
  * `splatt3r_slam.splatt3r_utils` (network inference: splatt3r_match_*; not called by the methods used here)
  * `splatt3r_slam.frame` (shared-memory keyframe store): replaced by `KeyframeStore` below, which offers what
    `FactorGraph.get_poses_points / solve_GN_*` read (global_opt.py:112-119,158: X_canon, T_WC,
    get_average_conf(), img, update_T_WCs)
  * `mast3r_slam_backends`: whatever module the caller binds — this repo's shim, the reference CUDA build
    (oracle/_ref), or the CPU oracle — via `bind_backend`.
`config` is filled with the values of the reference's config/base.yaml (cited per block).
"""
import importlib
import importlib.abc
import importlib.machinery
import importlib.util
import os
import sys
import types

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
REFPY = os.path.join(REF_DIR, "refpy")

# config/base.yaml of the reference (line numbers of that file)
BASE_CFG = {
    "use_calib": False,  # :1
    "single_thread": False,
    "matching": {"max_iter": 10, "lambda_init": 1e-8, "convergence_thresh": 1e-6, "dist_thresh": 1e-1, "radius": 3,
                 "dilation_max": 5},  # :8-14
    "tracking": {"min_match_frac": 0.05, "max_iters": 50, "C_conf": 0.0, "Q_conf": 1.5, "rel_error": 1e-3,
                 "delta_norm": 1e-3, "huber": 1.345, "match_frac_thresh": 0.333, "sigma_ray": 0.003, "sigma_dist": 1e+1,
                 "sigma_pixel": 1.0, "sigma_depth": 1e+1, "sigma_point": 0.05, "pixel_border": -10, "depth_eps": 1e-6,
                 "filtering_mode": "weighted_pointmap", "filtering_score": "median"},  # :16-33
    "local_opt": {"pin": 1, "window_size": 1e+6, "C_conf": 0.0, "Q_conf": 1.5, "min_match_frac": 0.1, "pixel_border": -10,
                  "depth_eps": 1e-6, "max_iters": 10, "sigma_ray": 0.003, "sigma_dist": 1e+1, "sigma_pixel": 1.0,
                  "sigma_depth": 1e+1, "sigma_point": 0.05, "delta_norm": 1e-8, "use_cuda": True},  # :35-50
}


def available():
    return (os.path.exists(os.path.join(REF_DIR, "lietorch_backends.so"))
            and os.path.exists(os.path.join(REFPY, "splatt3r_slam", "tracker.refbc")))


class KeyframeStore:
    """Stand-in for SharedKeyframes (frame.py:247-345) holding N keyframes: pointmaps X (N,n,3), confidences
    C (N,n,1), poses T_WC (N,1,8) Sim3 data, image size (h, w)."""

    class _KF:
        def __init__(self, store, k):
            import lietorch
            self.X_canon = store.X[k]
            self.T_WC = lietorch.Sim3(store.T_WC[k])
            self._C = store.C[k]
            self.img = torch.empty((3, store.h, store.w), device=store.X.device)

        def get_average_conf(self):
            return self._C

    def __init__(self, X, C, T_WC, h, w):
        self.X, self.C, self.T_WC, self.h, self.w = X, C, T_WC.view(-1, 1, 8).clone(), h, w

    def __getitem__(self, k):
        return KeyframeStore._KF(self, int(k))

    def __len__(self):
        return self.X.shape[0]

    def update_T_WCs(self, T_WCs, idx):  # frame.py:336-338
        self.T_WC[idx] = T_WCs.data


class _ByteCodeFinder(importlib.abc.MetaPathFinder):
    """Resolves `lietorch[.x]` and `splatt3r_slam.x` to the sourceless byte-code files under oracle/_ref/refpy."""

    def find_spec(self, fullname, path=None, target=None):
        parts = fullname.split(".")
        if parts[0] not in ("lietorch", "splatt3r_slam") or len(parts) > 2:
            return None
        pkg_dir = os.path.join(REFPY, parts[0])
        if len(parts) == 1:
            init = os.path.join(pkg_dir, "__init__.refbc")
            if os.path.exists(init):
                loader = importlib.machinery.SourcelessFileLoader(fullname, init)
                return importlib.util.spec_from_file_location(fullname, init, loader=loader,
                                                              submodule_search_locations=[pkg_dir])
            spec = importlib.machinery.ModuleSpec(fullname, None, is_package=True)  # namespace package
            spec.submodule_search_locations = [pkg_dir]
            return spec
        f = os.path.join(pkg_dir, parts[1] + ".refbc")
        if not os.path.exists(f):
            return None
        return importlib.util.spec_from_file_location(fullname, f, loader=importlib.machinery.SourcelessFileLoader(fullname, f))


_loaded = None


def load():
    """Returns a namespace with the reference modules: lietorch, config, matching, geometry, tracker, global_opt."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError("oracle/_ref has no reference drivers: run oracle/build_lietorch.py where /root/reference exists")
    if "lietorch_backends" not in sys.modules:
        spec = importlib.util.spec_from_file_location("lietorch_backends", os.path.join(REF_DIR, "lietorch_backends.so"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        sys.modules["lietorch_backends"] = mod
    if not any(isinstance(f, _ByteCodeFinder) for f in sys.meta_path):
        sys.meta_path.insert(0, _ByteCodeFinder())
    importlib.invalidate_caches()
    utils = types.ModuleType("splatt3r_slam.splatt3r_utils")

    def _not_on_path(*a, **k):
        raise RuntimeError("network inference is outside the hot path under test")
    utils.splatt3r_match_asymmetric = utils.splatt3r_match_symmetric = utils.resize_img = _not_on_path
    sys.modules["splatt3r_slam.splatt3r_utils"] = utils
    frame = types.ModuleType("splatt3r_slam.frame")
    frame.Frame, frame.SharedKeyframes = object, KeyframeStore
    sys.modules["splatt3r_slam.frame"] = frame
    if "mast3r_slam_backends" not in sys.modules:  # bound per call through bind_backend
        sys.modules["mast3r_slam_backends"] = types.ModuleType("mast3r_slam_backends")
    ns = types.SimpleNamespace()
    ns.lietorch = importlib.import_module("lietorch")
    ns.config = importlib.import_module("splatt3r_slam.config")
    ns.config.config.clear()
    ns.config.config.update({k: (dict(v) if isinstance(v, dict) else v) for k, v in BASE_CFG.items()})
    for name in ("matching", "geometry", "nonlinear_optimizer", "tracker", "global_opt"):
        setattr(ns, name, importlib.import_module("splatt3r_slam." + name))
    _loaded = ns
    return ns


def bind_backend(ns, backend):
    """Make the reference drivers call `backend` as their `mast3r_slam_backends`."""
    ns.matching.mast3r_slam_backends = backend
    ns.global_opt.mast3r_slam_backends = backend
