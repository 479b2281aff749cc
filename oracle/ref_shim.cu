// TEST INFRASTRUCTURE ONLY (GPU oracle).  Built by oracle/build_ref.py into oracle/_ref/.
//
// Pulls the staged copy of the reference's gn_kernels.cu into this translation unit
// (so the reference device code is compiled byte-for-byte) and additionally exposes
// ONE launch of each per-edge Hessian kernel as torch.ops.refshim.{ray,calib,point}_align,
// returning the raw Hs (4,E,7,7) / gs (2,E,7) that the reference driver keeps private
// (gn_kernels.cu:1172-1194).  This lets the parity tests compare Hessian blocks directly.
#include "gn_kernels.cu"
#include <torch/library.h>

namespace {

std::vector<torch::Tensor> edge_rows(torch::Tensor ii, torch::Tensor jj) {
  torch::Tensor uniq = get_unique_kf_idx(ii, jj);
  return create_inds(uniq, 0, ii, jj);
}

std::tuple<torch::Tensor, torch::Tensor> shim_ray_align(
    torch::Tensor Twc, torch::Tensor Xs, torch::Tensor Cs, torch::Tensor ii, torch::Tensor jj,
    torch::Tensor idx, torch::Tensor valid, torch::Tensor Q,
    double sigma_ray, double sigma_dist, double C_thresh, double Q_thresh) {
  auto rows = edge_rows(ii, jj);
  const int E = ii.size(0);
  torch::Tensor Hs = torch::zeros({4, E, 7, 7}, Twc.options());
  torch::Tensor gs = torch::zeros({2, E, 7}, Twc.options());
  ray_align_kernel<<<E, THREADS>>>(
      Twc.packed_accessor32<float,2,torch::RestrictPtrTraits>(),
      Xs.packed_accessor32<float,3,torch::RestrictPtrTraits>(),
      Cs.packed_accessor32<float,3,torch::RestrictPtrTraits>(),
      rows[0].packed_accessor32<long,1,torch::RestrictPtrTraits>(),
      rows[1].packed_accessor32<long,1,torch::RestrictPtrTraits>(),
      idx.packed_accessor32<long,2,torch::RestrictPtrTraits>(),
      valid.packed_accessor32<bool,3,torch::RestrictPtrTraits>(),
      Q.packed_accessor32<float,3,torch::RestrictPtrTraits>(),
      Hs.packed_accessor32<float,4,torch::RestrictPtrTraits>(),
      gs.packed_accessor32<float,3,torch::RestrictPtrTraits>(),
      (float)sigma_ray, (float)sigma_dist, (float)C_thresh, (float)Q_thresh);
  return {Hs, gs};
}

std::tuple<torch::Tensor, torch::Tensor> shim_point_align(
    torch::Tensor Twc, torch::Tensor Xs, torch::Tensor Cs, torch::Tensor ii, torch::Tensor jj,
    torch::Tensor idx, torch::Tensor valid, torch::Tensor Q,
    double sigma_point, double C_thresh, double Q_thresh) {
  auto rows = edge_rows(ii, jj);
  const int E = ii.size(0);
  torch::Tensor Hs = torch::zeros({4, E, 7, 7}, Twc.options());
  torch::Tensor gs = torch::zeros({2, E, 7}, Twc.options());
  point_align_kernel<<<E, THREADS>>>(
      Twc.packed_accessor32<float,2,torch::RestrictPtrTraits>(),
      Xs.packed_accessor32<float,3,torch::RestrictPtrTraits>(),
      Cs.packed_accessor32<float,3,torch::RestrictPtrTraits>(),
      rows[0].packed_accessor32<long,1,torch::RestrictPtrTraits>(),
      rows[1].packed_accessor32<long,1,torch::RestrictPtrTraits>(),
      idx.packed_accessor32<long,2,torch::RestrictPtrTraits>(),
      valid.packed_accessor32<bool,3,torch::RestrictPtrTraits>(),
      Q.packed_accessor32<float,3,torch::RestrictPtrTraits>(),
      Hs.packed_accessor32<float,4,torch::RestrictPtrTraits>(),
      gs.packed_accessor32<float,3,torch::RestrictPtrTraits>(),
      (float)sigma_point, (float)C_thresh, (float)Q_thresh);
  return {Hs, gs};
}

std::tuple<torch::Tensor, torch::Tensor> shim_calib_proj(
    torch::Tensor Twc, torch::Tensor Xs, torch::Tensor Cs, torch::Tensor K,
    torch::Tensor ii, torch::Tensor jj,
    torch::Tensor idx, torch::Tensor valid, torch::Tensor Q,
    int64_t height, int64_t width, int64_t pixel_border, double z_eps,
    double sigma_pixel, double sigma_depth, double C_thresh, double Q_thresh) {
  auto rows = edge_rows(ii, jj);
  const int E = ii.size(0);
  torch::Tensor Hs = torch::zeros({4, E, 7, 7}, Twc.options());
  torch::Tensor gs = torch::zeros({2, E, 7}, Twc.options());
  calib_proj_kernel<<<E, THREADS>>>(
      Twc.packed_accessor32<float,2,torch::RestrictPtrTraits>(),
      Xs.packed_accessor32<float,3,torch::RestrictPtrTraits>(),
      Cs.packed_accessor32<float,3,torch::RestrictPtrTraits>(),
      K.packed_accessor32<float,2,torch::RestrictPtrTraits>(),
      rows[0].packed_accessor32<long,1,torch::RestrictPtrTraits>(),
      rows[1].packed_accessor32<long,1,torch::RestrictPtrTraits>(),
      idx.packed_accessor32<long,2,torch::RestrictPtrTraits>(),
      valid.packed_accessor32<bool,3,torch::RestrictPtrTraits>(),
      Q.packed_accessor32<float,3,torch::RestrictPtrTraits>(),
      Hs.packed_accessor32<float,4,torch::RestrictPtrTraits>(),
      gs.packed_accessor32<float,3,torch::RestrictPtrTraits>(),
      (int)height, (int)width, (int)pixel_border, (float)z_eps,
      (float)sigma_pixel, (float)sigma_depth, (float)C_thresh, (float)Q_thresh);
  return {Hs, gs};
}

}  // namespace

TORCH_LIBRARY(refshim, m) {
  m.def("ray_align", &shim_ray_align);
  m.def("point_align", &shim_point_align);
  m.def("calib_proj", &shim_calib_proj);
}
