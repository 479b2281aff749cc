/* b200slam — C ABI of the B200-native dense-SLAM backend hot path.
 *
 * Drop-in boundary for the reference's `mast3r_slam_backends` extension
 * (Looong01/Splatt3R-SLAM, splatt3r_slam/backend).  Every entry point takes plain device
 * pointers + extents + a CUDA stream (passed as void*, i.e. a cudaStream_t) and returns 0 on
 * success or a negative B200SLAM_E* / positive cudaError_t code.  No torch types.  All tensors
 * are row-major contiguous, exactly the layouts the reference's pybind functions receive
 * (include/gn.h:22-117, src/gn.cpp:116-123).
 *
 * The library is compiled for sm_100a only; there is no CPU fallback.
 */
#ifndef B200SLAM_H
#define B200SLAM_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200SLAM_OK 0
#define B200SLAM_EINVAL (-1)    /* bad argument (null pointer, negative extent, n not supported) */
#define B200SLAM_EWORKSPACE (-2) /* workspace too small */
#define B200SLAM_ESTATIC_LIMIT (-3) /* input larger than the built-in scratch of an allocation-free entry point */

/* Residual model of the Gauss-Newton build. */
#define B200SLAM_GN_RAYS 0   /* ray + distance, gn_kernels.cu:812-1137 (ray_align_kernel)   */
#define B200SLAM_GN_CALIB 1  /* pixel + log-depth, gn_kernels.cu:1230-1542 (calib_proj_kernel) */
#define B200SLAM_GN_POINTS 2 /* 3-D point, gn_kernels.cu:454-722 (point_align_kernel)        */

int b200slam_version(void);

/* Replaces iter_proj_cuda / iter_proj_kernel, matching_kernels.cu:118-315 (pybind gn.cpp:84-99).
 * rays_img (b,h,w,9) f32; pts_3d_norm (b,n,3) f32; p_init (b,n,2) f32 (u,v)
 * -> p_new (b,n,2) f32; converged (b,n) u8 (torch.bool storage).  Bit-exact with the reference.
 * workspace: b200slam_iter_proj_workspace_bytes(b,h,w) bytes of device scratch (planar copy of
 * the ray image: the kernel re-lays the 9 interleaved channels out as planes so that a warp's
 * bilinear taps coalesce). */
size_t b200slam_iter_proj_workspace_bytes(int b, int h, int w);
int b200slam_iter_proj(const float *rays_img, const float *pts_3d_norm, const float *p_init,
                       float *p_new, uint8_t *converged, int b, int h, int w, int n, int max_iter,
                       float lambda_init, float cost_thresh, void *workspace, size_t workspace_bytes,
                       void *stream);

/* Allocation-free variant of b200slam_iter_proj.  It has NO counterpart in the reference and was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it): same inputs, outputs and bits as b200slam_iter_proj, but the planar
 * copy of the ray image lives in a file-scope `static __device__` array of the library instead of a caller
 * workspace, and the call performs no cudaMalloc / cudaFree / allocating library routine.
 * Limit: b*h*w*9*4 bytes <= B200SLAM_ITER_PROJ_STATIC_MAX_BYTES (one 512x384 image, or any b*h*w <= 196 608);
 * larger inputs return B200SLAM_ESTATIC_LIMIT and launch nothing.  Every GPU that loads the library holds the array
 * until exit.  All calls on a device share it (nothing needs zeroing: each call overwrites what it reads), so calls
 * that OVERLAP on the GPU - two streams - race; serialise them.  The first call on each device looks the array's
 * address up (cudaGetSymbolAddress): make it before a graph capture starts. */
#define B200SLAM_ITER_PROJ_STATIC_MAX_BYTES (196608 * 9 * 4)
int b200slam_iter_proj_static(const float *rays_img, const float *pts_3d_norm, const float *p_init,
                              float *p_new, uint8_t *converged, int b, int h, int w, int n, int max_iter,
                              float lambda_init, float cost_thresh, void *stream);

/* Replaces refine_matches_cuda / refine_matches_kernel<c10::Half>, matching_kernels.cu:24-115
 * (pybind gn.cpp:101-114).  D11 (b,h,w,fdim) f16; D21 (b,n,fdim) f16; p1 (b,n,2) i64 (u,v)
 * -> p1_new (b,n,2) i64.  Bit-exact with the reference (sequential fp16-rounded MACs). */
int b200slam_refine_matches_f16(const void *D11, const void *D21, const int64_t *p1,
                                int64_t *p1_new, int b, int h, int w, int n, int fdim, int radius,
                                int dilation_max, void *stream);
/* float instantiation of the same kernel (AT_DISPATCH at matching_kernels.cu:102). */
int b200slam_refine_matches_f32(const float *D11, const float *D21, const int64_t *p1,
                                int64_t *p1_new, int b, int h, int w, int n, int fdim, int radius,
                                int dilation_max, void *stream);
/* double instantiation (the third type AT_DISPATCH_FLOATING_TYPES_AND_HALF covers, matching_kernels.cu:102). */
int b200slam_refine_matches_f64(const double *D11, const double *D21, const int64_t *p1,
                                int64_t *p1_new, int b, int h, int w, int n, int fdim, int radius,
                                int dilation_max, void *stream);

/* Fused match_iterative_proj: replaces the PyTorch glue of splatt3r_slam/matching.py:25-49
 * (prep_for_iter_proj: F.normalize, image.py:5-38 gradients, cat/permute, lin_to_pixel), :52-65
 * (iter_proj call), :66-76 (occlusion distance test), :78-85 (refine_matches on .half() descriptors)
 * and :87-90 (pixel_to_lin) by four kernel launches.  NOT an entry of the reference module: the
 * reference's matching.py would call it instead of its own glue.
 * X11, X21 (b,h,w,3) f32 pointmaps; D11, D21 (b,h,w,fdim) descriptors, f16 (desc_f32 = 0) or f32
 * (desc_f32 = 1: rounded to f16 first, as `.half()` does); idx_init (b,h*w) i64 or NULL (identity map)
 * -> idx_1_to_2 (b,h*w) i64 linear index u + w*v; valid (b,h*w) u8 = converged & (distance < dist_thresh).
 * radius == 0 skips the descriptor refinement (matching.py:78).  Same float operation order as
 * torch's CUDA kernels for that glue, so results equal the unfused path bit for bit. */
size_t b200slam_match_workspace_bytes(int b, int h, int w, int fdim, int desc_f32);
int b200slam_match(const float *X11, const float *X21, const void *D11, const void *D21,
                   int desc_f32, const int64_t *idx_init, int64_t *idx_1_to_2, uint8_t *valid,
                   int b, int h, int w, int fdim, int max_iter, float lambda_init, float cost_thresh,
                   float dist_thresh, int radius, int dilation_max, void *workspace,
                   size_t workspace_bytes, void *stream);

/* On-device frame-to-keyframe tracker: replaces FrameTracker.opt_pose_ray_dist_sim3 + solve +
 * check_convergence, splatt3r_slam/tracker.py:156-214 and nonlinear_optimizer.py:5-33 (PyTorch there:
 * ~30 kernels and one host sync per Gauss-Newton step).  NOT an entry of the reference module
 * (SURVEY.md section 8f rank 2): tracker.py would call it instead of its own loop.
 * Xf (n,3) frame points already gathered by the match index (tracker.py:150), Xk (n,3) keyframe points,
 * Qk (n) match confidence, valid (n) u8; T_WCf, T_WCk: Sim3 data [t(3), q_xyzw(4), s] as lietorch stores
 * it -> T_WCf_new = T_WCk * T_CkCf and T_CkCf (8 floats each).  One cooperative launch for the whole
 * loop (<= max_iters steps; stops when |cost decrease| / old cost < rel_error or |tau| < delta_norm).
 * n <= 5632 * (number of SMs).  b200slam_track_status blocks on `stream` and reports the last call's
 * step count, convergence flag, failed-factorisation flag (the reference raises there) and final cost. */
size_t b200slam_track_workspace_bytes(void);
int b200slam_track_ray_dist(const float *Xf, const float *Xk, const float *Qk, const uint8_t *valid,
                            const float *T_WCf, const float *T_WCk, float *T_WCf_new, float *T_CkCf,
                            int n, float sigma_ray, float sigma_dist, float huber_k, int max_iters,
                            float rel_error, float delta_norm, void *workspace, size_t workspace_bytes,
                            void *stream);
/* Same loop for FrameTracker.opt_pose_calib_sim3, tracker.py:216-266 (residual [u, v, log z] through
 * project_calib, geometry.py:63-105): meas_k (n,3) = [u_k, v_k, log z_k] and valid_meas_k (n) u8 as
 * tracker.py:143-152 builds them; K device (3,3) row-major intrinsics (read by the kernel, no host sync);
 * a projected point counts when pixel_border < u < width-1-pixel_border (same for v) and z > z_eps. */
int b200slam_track_calib(const float *Xf, const float *meas_k, const float *Qk, const uint8_t *valid,
                         const uint8_t *valid_meas_k, const float *K, const float *T_WCf,
                         const float *T_WCk, float *T_WCf_new, float *T_CkCf, int n, int height,
                         int width, float sigma_pixel, float sigma_depth, float huber_k,
                         int pixel_border, float z_eps, int max_iters, float rel_error,
                         float delta_norm, void *workspace, size_t workspace_bytes, void *stream);
int b200slam_track_status(const void *workspace, int *iters, int *converged, int *chol_fail,
                          double *cost, void *stream);

/* Scalar parameters of the three GN residual models (union of the reference's scalar args,
 * gn.h:22-117).  sigma_a/sigma_b: rays = (sigma_ray, sigma_dist); calib = (sigma_pixel,
 * sigma_depth); points = (sigma_point, unused). */
typedef struct {
  int kind; /* B200SLAM_GN_* */
  float sigma_a, sigma_b;
  float C_thresh, Q_thresh;
  int height, width, pixel_border; /* calib only */
  float z_eps;                     /* calib only */
  /* Optional (NULL = identity): device array (N) i64, row k of the solve reads its pointmap / confidence from row
   * kf_slots[k] of Xs / Cs.  Lets a caller hand the kernels its keyframe BUFFERS (reference frame.py:266-268
   * SharedKeyframes.X / .C, indexed by keyframe id) in place for ANY set of keyframe ids, where the reference
   * stacks fresh (N,n,3) / (N,n,1) tensors before every solve (global_opt.py:111-119).  Twc stays the compact
   * (N,8) array (32 bytes per keyframe). */
  const int64_t *kf_slots;
} b200slam_gn_params;

/* Bytes of device workspace the GN entry points need for N poses, E directed edges, n points. */
size_t b200slam_gn_workspace_bytes(int N, int E, int n);

/* One Hessian build in the reference's private buffer layout (gn_kernels.cu:1172-1194):
 * Hs (4,E,7,7) f32 = [H_ii, H_ij, H_ji, H_jj] per edge, gs (2,E,7) f32 = [g_i, g_j].
 * Twc (N,8) f32 [t, q xyzw, s]; Xs (N,n,3); Cs (N,n); K (3,3) or NULL; ii_row/jj_row (E) i64 are
 * ROWS into Twc/Xs/Cs (the reference's ii_edge/jj_edge, :1161-1163); idx (E,n) i64;
 * valid (E,n) u8; Q (E,n) f32.  Edges [e0, e0+Ecount) only (edge sharding); Hs/gs are indexed
 * by the global edge number, extent E. */
int b200slam_gn_build_blocks(const b200slam_gn_params *prm, const float *Twc, const float *Xs,
                             const float *Cs, const float *K, const int64_t *ii_row,
                             const int64_t *jj_row, const int64_t *idx, const uint8_t *valid,
                             const float *Q, int N, int E, int n, int e0, int Ecount, float *Hs,
                             float *gs, void *workspace, size_t workspace_bytes, void *stream);

/* Stage 1+2 of one iteration for edges [e0,e0+Ecount): per-edge reduced blocks in fp64,
 * edge_blocks (E,56) f64 = [H_jj (49, row-major) | g_j (7)] (H_ii = H_jj, H_ij = H_ji = -H_jj,
 * g_i = -g_j; gn_kernels.cu:998-999).  Rows outside the shard are left untouched, so a
 * zero-initialised buffer can be summed across ranks with one allreduce. */
/* Multi-GPU (no counterpart in the reference, which is single-GPU): the same stage 1+2 with the exchange FUSED
 * into the finalizing kernel — every edge block of the shard is stored straight into every rank's (E,56) array
 * through peer pointers (host arrays peer_edge_blocks[world], peer_flags[world] of device pointers into NVLink-
 * mapped symmetric memory), then flags[rank] = seq is released at system scope in every rank's flag array.
 * b200slam_gn_wait_peers blocks the stream until all `world` flags of THIS rank's array have reached seq.
 * Buffers must be double-buffered by the caller (ranks run at most one iteration apart). */
int b200slam_gn_edge_blocks_push(const b200slam_gn_params *prm, const float *Twc, const float *Xs,
                                 const float *Cs, const float *K, const int64_t *ii_row,
                                 const int64_t *jj_row, const int64_t *idx, const uint8_t *valid,
                                 const float *Q, int N, int E, int n, int e0, int Ecount,
                                 double *const *peer_edge_blocks, unsigned *const *peer_flags, int world,
                                 int rank, unsigned seq, void *workspace, size_t workspace_bytes, void *stream);
int b200slam_gn_wait_peers(const unsigned *flags, int world, unsigned seq, void *stream);
int b200slam_gn_edge_blocks(const b200slam_gn_params *prm, const float *Twc, const float *Xs,
                            const float *Cs, const float *K, const int64_t *ii_row,
                            const int64_t *jj_row, const int64_t *idx, const uint8_t *valid,
                            const float *Q, int N, int E, int n, int e0, int Ecount,
                            double *edge_blocks, void *workspace, size_t workspace_bytes,
                            void *stream);

/* Stage 3: assemble the 7(N-1) normal equations from edge_blocks (num_fix = 1, fixed rows
 * skipped, duplicates summed: SparseBlock::update_lhs/rhs, gn_kernels.cu:70-112), fp64
 * Cholesky solve (zero update when not SPD, :141-149), dx = -x (:1208), in-place Sim3
 * retraction of Twc rows >= 1 (pose_retr_kernel, :414-452), and the ||dx|| < delta test
 * (:1218-1221) recorded in the workspace so later iterations become no-ops.  dx (N-1,7) f32. */
int b200slam_gn_solve_retract(float *Twc, const int64_t *ii_row, const int64_t *jj_row,
                              const double *edge_blocks, int N, int E, int n, float delta_thresh,
                              float *dx, void *workspace, size_t workspace_bytes, void *stream);

/* Reset the iteration state (done flag, iteration counter) kept at the head of the workspace. */
int b200slam_gn_reset(void *workspace, size_t workspace_bytes, void *stream);
/* Read back {iterations executed, done flag, cholesky failures} (synchronises the stream). */
int b200slam_gn_status(const void *workspace, int *iters, int *done, int *chol_fail, void *stream);
/* Large graphs (7(N-1) > 217): decide, for THIS call, between the block-sparse Cholesky (graphs whose factor stays
 * sparse under a minimum-degree ordering: odometry chains with few loop closures) and the pipelined PCG (graphs that
 * fill in).  The analysis half of the reference's SimplicialLLT (ordering + symbolic factorisation,
 * gn_kernels.cu:131-152, there re-done on the host every iteration) runs here once per topology: the 2E edge rows
 * are copied to the host (ONE stream synchronisation), hashed, and only a changed topology is re-analysed; the plan
 * lives in the workspace.  Call after b200slam_gn_reset and before the solves; b200slam_gauss_newton* do it
 * themselves.  Without this call the PCG (with its exact dense fallback) is used. */
int b200slam_gn_analyze(const int64_t *ii_row, const int64_t *jj_row, int N, int E, int n, void *workspace,
                        size_t workspace_bytes, void *stream);
/* The same plan executed on the HOST (no GPU): test aid that pins the ordering / symbolic factorisation / schedule.
 * edge_blocks (E,56) and the rows are host pointers; x (7(N-1)) solves H x = b; returns 1 when H is not SPD.
 * stats[4] = {off-diagonal blocks of L, block products, elimination-tree levels, 1 if within the device caps}. */
int b200slam_gn_sparse_host_solve(const int64_t *ii_row, const int64_t *jj_row, const double *edge_blocks, int N, int E,
                                  double *x, long long *stats);
/* Large-graph (7(N-1) > 217) solver selection, process-wide: -1 = default (environment B200SLAM_GN_SOLVER,
 * else auto), 0 = auto (pipelined PCG, exact tiled Cholesky as gated fallback), 1 = exact tiled Cholesky
 * only, 2 = first-generation PCG.  A/B and test aid; all three solve the same system (the reference's
 * SparseBlock::solve, gn_kernels.cu:131-152). */
int b200slam_gn_set_solver(int mode);
/* Large-graph solver diagnostics of the last call on this workspace (synchronises the stream): the
 * reference's counterpart is SparseBlock::solve's solver.info() (gn_kernels.cu:141-149), which it does not
 * expose.  cg_iters / cg_rel / cg_true_rel: iterations and relative (r, D^-1 r) — recursive and true — of
 * the last pipelined-PCG solve; cg_total: PCG iterations summed over the call; pcg_bad: the PCG gave up and
 * the exact tiled Cholesky took over; dense_runs: low 16 bits = exact dense solves of the call, high bits =
 * block-sparse Cholesky solves; stamps_ns[12]
 * (optional): %globaltimer at {start, set-up done, iterations done, end} of the last solver kernel, then its CG
 * loop's cycles in {publish, wait, recurrences, local dot products + D^-1 w, scalars, mat-vec} on one warp of CTA 0, then the maximum and minimum over the CTAs of their non-waiting cycles. */
int b200slam_gn_solver_status(const void *workspace, int *cg_iters, int *cg_total, int *pcg_bad,
                              int *dense_runs, double *cg_rel, double *cg_true_rel,
                              unsigned long long *stamps_ns, void *stream);

/* Whole driver: replaces gauss_newton_{rays,calib,points}_cuda (gn_kernels.cu:1139-1227,
 * :1545-1637, :724-810) for a single GPU.  Enqueues max_iter iterations with no host
 * synchronisation; iterations after convergence are skipped on the device.  Twc is updated in
 * place; dx (N-1,7) f32 receives the last update. */
int b200slam_gauss_newton(const b200slam_gn_params *prm, float *Twc, const float *Xs,
                          const float *Cs, const float *K, const int64_t *ii_row,
                          const int64_t *jj_row, const int64_t *idx, const uint8_t *valid,
                          const float *Q, int N, int E, int n, int max_iter, float delta_thresh,
                          float *dx, void *workspace, size_t workspace_bytes, void *stream);

/* Keyframe ids -> rows of Twc/Xs/Cs: rows[0..E) for ii, rows[E..2E) for jj; row = rank of the id
 * among the sorted distinct ids of cat(ii, jj).  Replaces get_unique_kf_idx / create_inds
 * (torch::_unique + searchsorted, gn_kernels.cu:160-169).  scratch: 2*E bytes. */
int b200slam_edge_rows(const int64_t *ii, const int64_t *jj, int E, int64_t *rows, void *scratch,
                       size_t scratch_bytes, void *stream);

/* Same as b200slam_gauss_newton but takes the KEYFRAME IDS the reference's pybind functions take
 * (gn.cpp:28-82) and derives the rows on the device; nothing in the call synchronises the host
 * (except the 36-byte read of K for the calib model). */
int b200slam_gauss_newton_ids(const b200slam_gn_params *prm, float *Twc, const float *Xs,
                              const float *Cs, const float *K, const int64_t *ii, const int64_t *jj,
                              const int64_t *idx, const uint8_t *valid, const float *Q, int N, int E,
                              int n, int max_iter, float delta_thresh, float *dx, void *workspace,
                              size_t workspace_bytes, void *stream);

/* Same loop on the factor graph's ONE-WAY edge lists: the reference doubles every per-edge array with
 * torch.cat before each solve (FactorGraph.prep_two_way_edges, global_opt.py:104-110: 2.56 MB per edge at
 * 512x384); here directed edge e < E1 is (ii[e] -> jj[e]) with idx_ii2jj / valid_match_j / Q_ii2jj and edge
 * E1 + e is (jj[e] -> ii[e]) with idx_jj2ii / valid_match_i / Q_jj2ii, read in place.  Results are
 * bit-identical to b200slam_gauss_newton_ids on the concatenated arrays.  Workspace:
 * b200slam_gn_workspace_bytes(N, 2 * E1, n).   */
int b200slam_gauss_newton_twoway(const b200slam_gn_params *prm, float *Twc, const float *Xs,
                                 const float *Cs, const float *K, const int64_t *ii, const int64_t *jj,
                                 const int64_t *idx_ii2jj, const int64_t *idx_jj2ii,
                                 const uint8_t *valid_match_j, const uint8_t *valid_match_i,
                                 const float *Q_ii2jj, const float *Q_jj2ii, int N, int E1, int n,
                                 int max_iter, float delta_thresh, float *dx, void *workspace,
                                 size_t workspace_bytes, void *stream);

/* Forward 3D-Gaussian-splat tile rasteriser (last item).  Replaces the FORWARD pass of
 * the reference's vendored rasteriser, thirdparty/diff-gaussian-rasterization-modified: RasterizeGaussiansCUDA
 * (rasterize_points.cu:39-130) -> CudaRasterizer::Rasterizer::forward (cuda_rasterizer/rasterizer_impl.cu:195-330:
 * preprocessCUDA forward.cu:157-260, duplicateWithKeys / sort / identifyTileRanges, renderCUDA forward.cu:262-370),
 * as the SLAM system calls it for rendering (splatt3r_slam/visualization.py:563-596,
 * splatt3r_core/src/pixelsplat_src/cuda_splatting.py:96-124).  The backward pass is not provided.
 * All pointers are device pointers unless noted.  means3D (P,3); cov3D_precomp (P,6) upper triangle, or NULL with
 * scales (P,3) + rotations (P,4) quaternion (r,x,y,z); opacities (P); colors_precomp (P,3), or NULL with shs
 * (P,sh_coeffs,3) of degree sh_degree <= 3 and campos (3); viewmatrix / projmatrix (16) in the reference's layout
 * (the transposed 4x4 matrices GaussianRasterizationSettings carries); bg (3) -> out_color (3,H,W) f32, radii (P) i32.
 * The number of (splat, tile) instances is data dependent: the workspace is sized for max_instances
 * (b200slam_gs_workspace_bytes); the call reads the actual count on the host (one stream synchronisation, as the
 * reference does at rasterizer_impl.cu:289), stores it in *num_instances (host pointer, may be NULL) and returns
 * B200SLAM_EWORKSPACE without rendering when it exceeds max_instances — call again with a larger workspace.
 * Limits: H, W <= 65535 * 16; fewer than 2^31 instances per call (a scene beyond that must be rendered in parts). */
size_t b200slam_gs_workspace_bytes(int P, int H, int W, long long max_instances);
int b200slam_gs_render(const float *means3D, const float *cov3D_precomp, const float *scales,
                       const float *rotations, float scale_modifier, const float *opacities,
                       const float *colors_precomp, const float *shs, int sh_degree, int sh_coeffs,
                       const float *campos, const float *viewmatrix, const float *projmatrix, float tan_fovx,
                       float tan_fovy, const float *bg, int P, int H, int W, float *out_color, int *radii,
                       long long max_instances, long long *num_instances, void *workspace,
                       size_t workspace_bytes, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* B200SLAM_H */
