// b200slam — C entry points of the on-device frame tracker (gn_track.cuh).  Included by gn.cu.
#pragma once

extern "C" size_t b200slam_track_workspace_bytes(void) {
  return 256 + (size_t)1024 * b200gn::TRACK_NACC * sizeof(double);  // state + one partial per CTA (<= 1024 SMs)
}

namespace {
template <int KIND>
int launch_track(const float *Xf, const float *B, const float *Qk, const uint8_t *valid, const uint8_t *valid2,
                 const float *T_WCf, const float *T_WCk, float *T_WCf_new, float *T_CkCf, int n,
                 b200gn::TrackConsts C, int max_iters, void *workspace, size_t workspace_bytes, void *stream) {
  using namespace b200gn;
  if (n < 0 || max_iters < 0) return B200SLAM_EINVAL;
  if (!T_WCf || !T_WCk || !T_WCf_new || !T_CkCf) return B200SLAM_EINVAL;
  if (n > 0 && (!Xf || !B || !Qk || !valid || (KIND == 1 && !valid2))) return B200SLAM_EINVAL;
  if (!workspace || workspace_bytes < b200slam_track_workspace_bytes()) return B200SLAM_EWORKSPACE;
  const b200dev::Info dev = b200dev::info();  // per device, not per process
  if (!dev.coop) return B200SLAM_EINVAL;
  const int num_sms = dev.num_sms;
  int S = num_sms > 1024 ? 1024 : num_sms;
  int gper = (n + S - 1) / S;
  gper = (gper + 31) / 32 * 32;
  if (gper < 32) gper = 32;
  if (gper > PERSIST_MAX_POINTS) return B200SLAM_EINVAL;  // > 833 k points on 148 SMs
  static b200dev::Once attr;
  if (attr.first()) {
    cudaError_t err = cudaFuncSetAttribute(track_kernel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           PERSIST_WORDS * PERSIST_MAX_POINTS * (int)sizeof(float));
    if (err != cudaSuccess) { attr.retry(); return (int)err; }
  }
  cudaStream_t st = (cudaStream_t)stream;
  TrackState *state = (TrackState *)workspace;
  double *partials = (double *)((char *)workspace + 256);
  cudaError_t err = cudaMemsetAsync(state, 0, 256, st);
  if (err != cudaSuccess) return (int)err;
  void *args[] = {&Xf, &B, &Qk, &valid, &valid2, &T_WCf, &T_WCk, &T_WCf_new, &T_CkCf, &n, &gper, &C,
                  &max_iters, &partials, &state};
  const size_t smem = (size_t)PERSIST_WORDS * gper * sizeof(float);
  return (int)cudaLaunchCooperativeKernel((const void *)track_kernel<KIND>, dim3(S), dim3(PERSIST_THREADS), args, smem,
                                          st);
}
}  // namespace

extern "C" int b200slam_track_ray_dist(const float *Xf, const float *Xk, const float *Qk, const uint8_t *valid,
                                       const float *T_WCf, const float *T_WCk, float *T_WCf_new, float *T_CkCf,
                                       int n, float sigma_ray, float sigma_dist, float huber_k, int max_iters,
                                       float rel_error, float delta_norm, void *workspace, size_t workspace_bytes,
                                       void *stream) {
  if (!(sigma_ray > 0.0f) || !(sigma_dist > 0.0f)) return B200SLAM_EINVAL;
  b200gn::TrackConsts C{1.0f / sigma_ray, 1.0f / sigma_dist, huber_k, rel_error, delta_norm, 0, 0, 0, 0, 0, nullptr};
  return launch_track<0>(Xf, Xk, Qk, valid, nullptr, T_WCf, T_WCk, T_WCf_new, T_CkCf, n, C, max_iters, workspace,
                         workspace_bytes, stream);
}

extern "C" int b200slam_track_calib(const float *Xf, const float *meas_k, const float *Qk, const uint8_t *valid,
                                    const uint8_t *valid_meas_k, const float *K, const float *T_WCf,
                                    const float *T_WCk, float *T_WCf_new, float *T_CkCf, int n, int height,
                                    int width, float sigma_pixel, float sigma_depth, float huber_k,
                                    int pixel_border, float z_eps, int max_iters, float rel_error,
                                    float delta_norm, void *workspace, size_t workspace_bytes, void *stream) {
  if (!(sigma_pixel > 0.0f) || !(sigma_depth > 0.0f) || !K || height < 1 || width < 1) return B200SLAM_EINVAL;
  b200gn::TrackConsts C{1.0f / sigma_pixel, 1.0f / sigma_depth, huber_k, rel_error, delta_norm,
                        (float)pixel_border, (float)(width - 1 - pixel_border), (float)pixel_border,
                        (float)(height - 1 - pixel_border), z_eps, K};
  return launch_track<1>(Xf, meas_k, Qk, valid, valid_meas_k, T_WCf, T_WCk, T_WCf_new, T_CkCf, n, C, max_iters,
                         workspace, workspace_bytes, stream);
}

// Blocking read of the last call's outcome (synchronises `stream`).
extern "C" int b200slam_track_status(const void *workspace, int *iters, int *converged, int *chol_fail,
                                     double *cost, void *stream) {
  if (!workspace) return B200SLAM_EINVAL;
  b200gn::TrackState h;
  cudaError_t err = cudaMemcpyAsync(&h, workspace, sizeof(h), cudaMemcpyDeviceToHost, (cudaStream_t)stream);
  if (err != cudaSuccess) return (int)err;
  err = cudaStreamSynchronize((cudaStream_t)stream);
  if (err != cudaSuccess) return (int)err;
  if (iters) *iters = h.iters;
  if (converged) *converged = h.converged;
  if (chol_fail) *chol_fail = h.chol_fail;
  if (cost) *cost = h.cost;
  return B200SLAM_OK;
}
