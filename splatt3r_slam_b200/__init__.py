"""splatt3r_slam_b200 — B200-native (sm_100a) dense-SLAM backend hot path.

Contents: csrc/ (CUDA kernels + the C ABI of include/b200slam.h, built into libb200slam.so),
`mast3r_slam_backends` (drop-in mirror of the reference's extension module), `matching`
(mirror of the reference's matching driver), `dist_gn` (edge-sharded multi-GPU Gauss-Newton),
`synthetic` (seeded MASt3R-shaped inputs).  Importing the package does not load the library.
"""
__version__ = "0.1.0"
