// `droid_backends` as a compiled torch extension over the libdba C ABI (include/dba.h): what the reference's
// src/droid.cpp becomes when its CUDA is replaced by libdba (reference build: setup.py:7-31, CUDAExtension
// 'droid_backends'; call sites: depth_video.pyc cuda_ba@232 / distance@195-205, hislam2/util/droid_visualization.py:
// 97-101).  Argument lists and the CHECK_INPUT convention follow the reference's extension (tensors must be contiguous
// CUDA tensors); every function forwards pointers and sizes, nothing is computed here.
// Built by cut3r_slam_b200/build.py into cut3r_slam_b200/ext/droid_backends.<abi>.so (rpath $ORIGIN/.. -> libdba.so);
// the ctypes module cut3r_slam_b200/droid_backends.py stays the package's default binding.
#include <torch/extension.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>

#include <mutex>
#include <vector>

#include "dba.h"

#define CHECK_INPUT(x) TORCH_CHECK((x).is_cuda() && (x).is_contiguous(), #x " must be a contiguous CUDA tensor")
#define CHECK_F32(x) TORCH_CHECK((x).scalar_type() == torch::kFloat32, #x " must be float32")
#define CHECK_I64(x) TORCH_CHECK((x).scalar_type() == torch::kInt64, #x " must be int64")

namespace {

dba_ctx* ctx_for(int dev) {
  static std::mutex mu;
  static dba_ctx* ctxs[64] = {};
  std::lock_guard<std::mutex> lock(mu);
  TORCH_CHECK(dev >= 0 && dev < 64, "device index out of range");
  if (!ctxs[dev]) ctxs[dev] = dba_create(dev);
  TORCH_CHECK(ctxs[dev] != nullptr, "dba_create failed (no CUDA device " + std::to_string(dev) + "?)");
  return ctxs[dev];
}

void check_rc(dba_ctx* c, int rc, const char* what) {
  TORCH_CHECK(rc == 0, "libdba ", what, " error ", rc, ": ", dba_last_error(c));
}

void* stream_of(const torch::Tensor& t) { return at::cuda::getCurrentCUDAStream(t.get_device()).stream(); }

const float* fptr_or_null(const torch::Tensor& t) { return t.defined() && t.numel() > 0 ? t.data_ptr<float>() : nullptr; }

// ba(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1, iterations, lm, ep, motion_only)
std::vector<torch::Tensor> ba(torch::Tensor poses, torch::Tensor disps, torch::Tensor intrinsics, torch::Tensor disps_sens,
                              torch::Tensor targets, torch::Tensor weights, torch::Tensor eta, torch::Tensor ii,
                              torch::Tensor jj, int64_t t0, int64_t t1, int64_t iterations, double lm, double ep,
                              bool motion_only) {
  CHECK_INPUT(poses); CHECK_INPUT(disps); CHECK_INPUT(intrinsics); CHECK_INPUT(targets); CHECK_INPUT(weights);
  CHECK_INPUT(ii); CHECK_INPUT(jj);
  CHECK_F32(poses); CHECK_F32(disps); CHECK_F32(intrinsics); CHECK_F32(targets); CHECK_F32(weights);
  CHECK_I64(ii); CHECK_I64(jj);
  TORCH_CHECK(disps.dim() == 3 && poses.dim() == 2 && poses.size(1) == 7, "poses [N,7], disps [N,ht,wd]");
  if (!motion_only) { CHECK_INPUT(eta); CHECK_F32(eta); }
  if (disps_sens.defined() && disps_sens.numel() > 0) { CHECK_INPUT(disps_sens); CHECK_F32(disps_sens); }
  const c10::cuda::CUDAGuard guard(poses.device());
  const int64_t N = disps.size(0), E = ii.numel(), M = motion_only ? 0 : eta.size(0);
  const int ht = int(disps.size(1)), wd = int(disps.size(2));
  TORCH_CHECK(jj.numel() == E, "ii and jj must have the same length");
  auto dx = torch::zeros({std::max<int64_t>(t1 - t0, 0), 6}, poses.options());
  auto dz = torch::zeros({M, ht, wd}, poses.options());
  dba_ctx* c = ctx_for(poses.get_device());
  const int rc = dba_ba(c, poses.data_ptr<float>(), disps.data_ptr<float>(), intrinsics.data_ptr<float>(),
                        fptr_or_null(disps_sens), E ? targets.data_ptr<float>() : nullptr,
                        E ? weights.data_ptr<float>() : nullptr, motion_only ? nullptr : eta.data_ptr<float>(),
                        E ? ii.data_ptr<int64_t>() : nullptr, E ? jj.data_ptr<int64_t>() : nullptr, E, N, ht, wd, M, int(t0),
                        int(t1), int(iterations), float(lm), float(ep), motion_only ? 1 : 0, dx.data_ptr<float>(),
                        M ? dz.data_ptr<float>() : nullptr, stream_of(poses));
  check_rc(c, rc, "ba");
  return {dx, dz};
}

// frame_distance(poses, disps, intrinsics, ii, jj, beta) -> [K]
torch::Tensor frame_distance(torch::Tensor poses, torch::Tensor disps, torch::Tensor intrinsics, torch::Tensor ii,
                             torch::Tensor jj, double beta) {
  CHECK_INPUT(poses); CHECK_INPUT(disps); CHECK_INPUT(intrinsics); CHECK_INPUT(ii); CHECK_INPUT(jj);
  CHECK_F32(poses); CHECK_F32(disps); CHECK_F32(intrinsics); CHECK_I64(ii); CHECK_I64(jj);
  TORCH_CHECK(ii.numel() == jj.numel(), "ii and jj must have the same length");
  const c10::cuda::CUDAGuard guard(poses.device());
  auto out = torch::empty({ii.numel()}, poses.options());
  dba_ctx* c = ctx_for(poses.get_device());
  check_rc(c, dba_frame_distance(c, poses.data_ptr<float>(), disps.data_ptr<float>(), intrinsics.data_ptr<float>(),
                                 ii.data_ptr<int64_t>(), jj.data_ptr<int64_t>(), ii.numel(), disps.size(0),
                                 int(disps.size(1)), int(disps.size(2)), float(beta), out.data_ptr<float>(),
                                 stream_of(poses)), "frame_distance");
  return out;
}

// covis_distance(poses, disps, intrinsics, ii) -> [K]
torch::Tensor covis_distance(torch::Tensor poses, torch::Tensor disps, torch::Tensor intrinsics, torch::Tensor ii) {
  CHECK_INPUT(poses); CHECK_INPUT(disps); CHECK_INPUT(intrinsics); CHECK_INPUT(ii);
  CHECK_F32(poses); CHECK_F32(disps); CHECK_F32(intrinsics); CHECK_I64(ii);
  const c10::cuda::CUDAGuard guard(poses.device());
  auto out = torch::empty({ii.numel()}, poses.options());
  dba_ctx* c = ctx_for(poses.get_device());
  check_rc(c, dba_covis_distance(c, poses.data_ptr<float>(), disps.data_ptr<float>(), intrinsics.data_ptr<float>(),
                                 ii.data_ptr<int64_t>(), ii.numel(), disps.size(0), int(disps.size(1)), int(disps.size(2)),
                                 out.data_ptr<float>(), stream_of(poses)), "covis_distance");
  return out;
}

// iproj(poses_inv, disps, intrinsics) -> [K,ht,wd,3]
torch::Tensor iproj(torch::Tensor poses, torch::Tensor disps, torch::Tensor intrinsics) {
  CHECK_INPUT(poses); CHECK_INPUT(disps); CHECK_INPUT(intrinsics);
  CHECK_F32(poses); CHECK_F32(disps); CHECK_F32(intrinsics);
  TORCH_CHECK(poses.dim() == 2 && poses.size(0) == disps.size(0) && poses.size(1) == 7, "poses must be [K,7]");
  const c10::cuda::CUDAGuard guard(poses.device());
  auto out = torch::empty({disps.size(0), disps.size(1), disps.size(2), 3}, poses.options());
  dba_ctx* c = ctx_for(poses.get_device());
  check_rc(c, dba_iproj(c, poses.data_ptr<float>(), disps.data_ptr<float>(), intrinsics.data_ptr<float>(), disps.size(0),
                        int(disps.size(1)), int(disps.size(2)), out.data_ptr<float>(), stream_of(poses)), "iproj");
  return out;
}

// depth_filter(poses, disps, intrinsics, ix, thresh) -> [K,ht,wd]
torch::Tensor depth_filter(torch::Tensor poses, torch::Tensor disps, torch::Tensor intrinsics, torch::Tensor ix,
                           torch::Tensor thresh) {
  CHECK_INPUT(poses); CHECK_INPUT(disps); CHECK_INPUT(intrinsics); CHECK_INPUT(ix); CHECK_INPUT(thresh);
  CHECK_F32(poses); CHECK_F32(disps); CHECK_F32(intrinsics); CHECK_I64(ix); CHECK_F32(thresh);
  TORCH_CHECK(thresh.numel() == ix.numel(), "thresh must hold one threshold per entry of ix");
  const c10::cuda::CUDAGuard guard(poses.device());
  auto out = torch::empty({ix.numel(), disps.size(1), disps.size(2)}, poses.options());
  dba_ctx* c = ctx_for(poses.get_device());
  check_rc(c, dba_depth_filter(c, poses.data_ptr<float>(), disps.data_ptr<float>(), intrinsics.data_ptr<float>(),
                               ix.data_ptr<int64_t>(), thresh.data_ptr<float>(), ix.numel(), disps.size(0),
                               int(disps.size(1)), int(disps.size(2)), out.data_ptr<float>(), stream_of(poses)),
           "depth_filter");
  return out;
}

// pgba(poses [N,8], disps, intrinsics, targets, weights, eta, ii, jj, rel_ii, rel_jj, rel_poses, rel_infos, t0, t1,
//      iterations, lm, ep, pw, motion_only)  -- 19 arguments like the call at depth_video.pyc cuda_pgba@307-309
std::vector<torch::Tensor> pgba(torch::Tensor poses, torch::Tensor disps, torch::Tensor intrinsics, torch::Tensor targets,
                                torch::Tensor weights, torch::Tensor eta, torch::Tensor ii, torch::Tensor jj,
                                torch::Tensor rel_ii, torch::Tensor rel_jj, torch::Tensor rel_poses, torch::Tensor rel_infos,
                                int64_t t0, int64_t t1, int64_t iterations, double lm, double ep, double pw,
                                bool motion_only) {
  TORCH_CHECK(!motion_only, "pgba: motion_only is not supported");
  CHECK_INPUT(poses); CHECK_INPUT(disps); CHECK_INPUT(intrinsics); CHECK_INPUT(targets); CHECK_INPUT(weights);
  CHECK_INPUT(eta); CHECK_INPUT(ii); CHECK_INPUT(jj);
  CHECK_F32(poses); CHECK_F32(disps); CHECK_F32(targets); CHECK_F32(weights); CHECK_F32(eta); CHECK_I64(ii); CHECK_I64(jj);
  TORCH_CHECK(poses.dim() == 2 && poses.size(1) == 8, "poses must be [N,8] Sim3 data");
  const int64_t R = rel_ii.defined() ? rel_ii.numel() : 0;
  if (R) {
    CHECK_INPUT(rel_ii); CHECK_INPUT(rel_jj); CHECK_INPUT(rel_poses); CHECK_INPUT(rel_infos);
    CHECK_I64(rel_ii); CHECK_I64(rel_jj); CHECK_F32(rel_poses); CHECK_F32(rel_infos);
  }
  const c10::cuda::CUDAGuard guard(poses.device());
  const int64_t N = disps.size(0), E = ii.numel(), M = eta.size(0);
  const int ht = int(disps.size(1)), wd = int(disps.size(2));
  auto dx = torch::zeros({std::max<int64_t>(t1 - t0, 0), 7}, poses.options());
  auto dz = torch::zeros({M, ht, wd}, poses.options());
  dba_ctx* c = ctx_for(poses.get_device());
  check_rc(c, dba_pgba(c, poses.data_ptr<float>(), disps.data_ptr<float>(), intrinsics.data_ptr<float>(),
                       targets.data_ptr<float>(), weights.data_ptr<float>(), eta.data_ptr<float>(), ii.data_ptr<int64_t>(),
                       jj.data_ptr<int64_t>(), R ? rel_ii.data_ptr<int64_t>() : nullptr,
                       R ? rel_jj.data_ptr<int64_t>() : nullptr, R ? rel_poses.data_ptr<float>() : nullptr,
                       R ? rel_infos.data_ptr<float>() : nullptr, E, R, N, ht, wd, M, int(t0), int(t1), int(iterations),
                       float(lm), float(ep), float(pw), dx.data_ptr<float>(), dz.data_ptr<float>(), stream_of(poses)),
           "pgba");
  return {dx, dz};
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "droid_backends over libdba (B200-native dense bundle adjustment), compiled binding";
  m.def("ba", &ba, "dense bundle adjustment, in place on poses / disps; returns [dx, dz]");
  m.def("pgba", &pgba, "Sim3 pose-graph bundle adjustment; returns [dx, dz]");
  m.def("frame_distance", &frame_distance, "frame distance");
  m.def("covis_distance", &covis_distance, "covisibility distance to the temporal successor");
  m.def("iproj", &iproj, "back-project depth maps");
  m.def("depth_filter", &depth_filter, "depth filter");
  m.def("version", []() { return std::string(dba_version()); });
}
