"""Import shim: `import mast3r_slam_backends` (the name the reference's matching.py:5 and
global_opt.py:9 import) resolves to the B200 backend when the repo root is on sys.path."""
from splatt3r_slam_b200.mast3r_slam_backends import *  # noqa: F401,F403
from splatt3r_slam_b200.mast3r_slam_backends import (  # noqa: F401
    gauss_newton_calib, gauss_newton_points, gauss_newton_rays, iter_proj, refine_matches)
