// Companion ops of the `droid_backends` interface: frame_distance, iproj, depth_filter.
//
// The reference's CUDA for these (src/droid_kernels.cu) is absent from its tree, so their exact
// constants are PARITY UNPINNED; the definitions implemented here are the ones documented in
// DESIGN.md ("Companion ops") and restated on the CPU in oracle/dba.py, built from
// hislam2/geom/projective_ops.py:44-62 and the call sites depth_video.pyc distance@195-205,
// hislam2/util/droid_visualization.py:97,100-101.
#include "dba_kernels.h"
#include "dba_math.cuh"

namespace dba {

namespace {

constexpr float kFrameDistanceSentinel = 1000.0f;
constexpr int kDepthFilterRadius = 3;

__global__ void __launch_bounds__(256)
k_frame_distance(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr,
                 const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, int N, int HW, int wd, float beta,
                 float* __restrict__ out) {
  __shared__ float sRt[12];
  __shared__ float sred[2][8];
  const int pair = blockIdx.x;
  const int64_t i = ii[pair], j = jj[pair];
  const bool ok = i >= 0 && i < N && j >= 0 && j < N;
  if (threadIdx.x == 0 && ok) relative_pose_Rt(poses, int(i), int(j), sRt);
  __syncthreads();
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  float num = 0.f, den = 0.f;
  if (ok) {
    float Rt[12];
#pragma unroll
    for (int t = 0; t < 12; ++t) Rt[t] = sRt[t];
    for (int p = threadIdx.x; p < HW; p += blockDim.x) {
      const float d = disps[i * HW + p];
      const int py = p / wd, px = p - py * wd;
      const float xf = float(px), yf = float(py);
      const float x0 = (xf - cx) / fx, y0 = (yf - cy) / fy;
      // full relative motion
      float X = Rt[0] * x0 + Rt[1] * y0 + Rt[2] + Rt[9] * d;
      float Y = Rt[3] * x0 + Rt[4] * y0 + Rt[5] + Rt[10] * d;
      float Z = Rt[6] * x0 + Rt[7] * y0 + Rt[8] + Rt[11] * d;
      float v1 = Z > kMinDepth ? 1.f : 0.f;
      float Zs = Z > kMinDepth ? Z : 1.f;
      float du = fx * (X / Zs) + cx - xf, dv = fy * (Y / Zs) + cy - yf;
      const float f1 = sqrtf(du * du + dv * dv) * v1;
      // translation only
      X = x0 + Rt[9] * d; Y = y0 + Rt[10] * d; Z = 1.f + Rt[11] * d;
      float v2 = Z > kMinDepth ? 1.f : 0.f;
      Zs = Z > kMinDepth ? Z : 1.f;
      du = fx * (X / Zs) + cx - xf; dv = fy * (Y / Zs) + cy - yf;
      const float f2 = sqrtf(du * du + dv * dv) * v2;
      num += beta * f1 + (1.f - beta) * f2;
      den += beta * v1 + (1.f - beta) * v2;
    }
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    num += __shfl_xor_sync(0xffffffffu, num, o);
    den += __shfl_xor_sync(0xffffffffu, den, o);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { sred[0][warp] = num; sred[1][warp] = den; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float n = 0.f, dn = 0.f;
    for (int w = 0; w < 8; ++w) { n += sred[0][w]; dn += sred[1][w]; }
    out[pair] = (!ok || dn < 1.0f) ? kFrameDistanceSentinel : n / fmaxf(dn, 1e-12f);
  }
}

// covis_distance(poses, disps, intrinsics, ii) -> [K]  (call shape depth_video.pyc distance_covis@217; PARITY UNPINNED).
// Definition used: for keyframe i = ii[k] and its temporal successor j = i + 1 (i - 1 for the last frame of the
// buffer): the mean induced-flow magnitude |proj(G_ij X_i) - x_i| in pixels over the COVISIBLE pixels of frame i --
// those whose transformed depth exceeds MIN_DEPTH and whose projection falls inside the image of frame j;
// sentinel 1000 when no pixel is covisible.  The caller rescales by 1 / median(disps[ii]) (@218).
__global__ void __launch_bounds__(256)
k_covis_distance(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr,
                 const int64_t* __restrict__ ii, int N, int HW, int wd, int ht, float* __restrict__ out) {
  __shared__ float sRt[12];
  __shared__ float sred[2][8];
  const int64_t i = ii[blockIdx.x];
  const int64_t j = (i + 1 < N) ? i + 1 : i - 1;
  const bool ok = i >= 0 && i < N && j >= 0 && j < N;
  if (threadIdx.x == 0 && ok) relative_pose_Rt(poses, int(i), int(j), sRt);
  __syncthreads();
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  float num = 0.f, den = 0.f;
  if (ok) {
    float Rt[12];
#pragma unroll
    for (int t = 0; t < 12; ++t) Rt[t] = sRt[t];
    for (int p = threadIdx.x; p < HW; p += blockDim.x) {
      const float d = disps[i * HW + p];
      const int py = p / wd, px = p - py * wd;
      const float xf = float(px), yf = float(py);
      const float x0 = (xf - cx) / fx, y0 = (yf - cy) / fy;
      const float X = Rt[0] * x0 + Rt[1] * y0 + Rt[2] + Rt[9] * d;
      const float Y = Rt[3] * x0 + Rt[4] * y0 + Rt[5] + Rt[10] * d;
      const float Z = Rt[6] * x0 + Rt[7] * y0 + Rt[8] + Rt[11] * d;
      const float Zs = Z > kMinDepth ? Z : 1.f;
      const float u = fx * (X / Zs) + cx, v = fy * (Y / Zs) + cy;
      const bool vis = Z > kMinDepth && u >= 0.f && u <= float(wd - 1) && v >= 0.f && v <= float(ht - 1);
      const float du = u - xf, dv = v - yf;
      if (vis) { num += sqrtf(du * du + dv * dv); den += 1.f; }
    }
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    num += __shfl_xor_sync(0xffffffffu, num, o);
    den += __shfl_xor_sync(0xffffffffu, den, o);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { sred[0][warp] = num; sred[1][warp] = den; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float n = 0.f, dn = 0.f;
    for (int w = 0; w < 8; ++w) { n += sred[0][w]; dn += sred[1][w]; }
    out[blockIdx.x] = (!ok || dn < 1.0f) ? kFrameDistanceSentinel : n / dn;
  }
}

__global__ void __launch_bounds__(256)
k_iproj(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr, int HW,
        int wd, float* __restrict__ out) {
  const int k = blockIdx.y;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= HW) return;
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const SE3 g = se3_load(poses + 7 * k);
  const float d = disps[int64_t(k) * HW + p];
  const int py = p / wd, px = p - py * wd;
  V3 pt{((float(px) - cx) / fx) / d, ((float(py) - cy) / fy) / d, 1.0f / d};
  V3 r = qrot(g.q, pt);
  float* o = out + (int64_t(k) * HW + p) * 3;
  o[0] = r.x + g.t.x; o[1] = r.y + g.t.y; o[2] = r.z + g.t.z;
}

__global__ void __launch_bounds__(256)
k_depth_filter(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr,
               const int64_t* __restrict__ ix, const float* __restrict__ thresh, int N, int ht, int wd,
               float* __restrict__ out) {
  __shared__ float sRt[2 * kDepthFilterRadius][12];
  __shared__ int sj[2 * kDepthFilterRadius];
  const int k = blockIdx.y;
  const int HW = ht * wd;
  const int64_t i = ix[k];
  const bool ok = i >= 0 && i < N;
  if (threadIdx.x < 2 * kDepthFilterRadius) {
    const int t = threadIdx.x;
    const int off = (t < kDepthFilterRadius) ? t - kDepthFilterRadius : t - kDepthFilterRadius + 1;
    const int64_t j = i + off;
    const bool in = ok && j >= 0 && j < N;
    sj[t] = in ? int(j) : -1;
    if (in) relative_pose_Rt(poses, int(i), int(j), sRt[t]);
  }
  __syncthreads();
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= HW) return;
  float cnt = 0.f;
  if (ok) {
    const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
    const float th = thresh[k];
    const float d = disps[i * HW + p];
    const int py = p / wd, px = p - py * wd;
    const float x0 = (float(px) - cx) / fx, y0 = (float(py) - cy) / fy;
    for (int t = 0; t < 2 * kDepthFilterRadius; ++t) {
      const int j = sj[t];
      if (j < 0) continue;
      const float* Rt = sRt[t];
      const float X = Rt[0] * x0 + Rt[1] * y0 + Rt[2] + Rt[9] * d;
      const float Y = Rt[3] * x0 + Rt[4] * y0 + Rt[5] + Rt[10] * d;
      const float Z = Rt[6] * x0 + Rt[7] * y0 + Rt[8] + Rt[11] * d;
      const bool front = Z > kMinDepth;
      const float Zs = front ? Z : 1.f;
      const float uj = fx * (X / Zs) + cx, vj = fy * (Y / Zs) + cy;
      const float dj = d / Zs;
      const float u0 = floorf(uj), v0 = floorf(vj);
      const bool inb = front && u0 >= 0.f && v0 >= 0.f && u0 < float(wd - 1) && v0 < float(ht - 1);
      if (!inb) continue;
      const int iu = int(u0), iv = int(v0);
      const float wx = uj - u0, wy = vj - v0;
      const float* D = disps + int64_t(j) * HW;
      const float d00 = D[iv * wd + iu], d01 = D[iv * wd + iu + 1];
      const float d10 = D[(iv + 1) * wd + iu], d11 = D[(iv + 1) * wd + iu + 1];
      const float dhat = (1.f - wy) * ((1.f - wx) * d00 + wx * d01) + wy * ((1.f - wx) * d10 + wx * d11);
      const float err = fabsf(1.0f / fmaxf(dj, 1e-12f) - 1.0f / fmaxf(dhat, 1e-12f));
      if (dj > 0.f && err < th) cnt += 1.f;
    }
  }
  out[int64_t(k) * HW + p] = cnt;
}

}  // namespace

// Objective of the reprojection problem at the current state (debug / convergence monitor, option "chi2"):
//   chi2 = sum over edges and pixels of 0.001 * valid * (w_x r_x^2 + w_y r_y^2),  r = target - coords, the weighting of
//   hislam2/geom/ba.py:44-45.  One CTA per edge, per-CTA partial sums in fp64 (deterministic slot per edge; the host
//   adds the E slots).  Not on the optimisation path.
namespace {
__global__ void __launch_bounds__(256)
k_chi2(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr,
       const float* __restrict__ targets, const float* __restrict__ weights, const int64_t* __restrict__ ii,
       const int64_t* __restrict__ jj, int N, int HW, int wd, double* __restrict__ out) {
  __shared__ float sRt[12];
  __shared__ double sred[8];
  const int e = blockIdx.x;
  const int64_t i = ii[e], j = jj[e];
  const bool ok = i >= 0 && i < N && j >= 0 && j < N;
  if (threadIdx.x == 0 && ok) relative_pose_Rt(poses, int(i), int(j), sRt);
  __syncthreads();
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  double acc = 0.0;
  if (ok) {
    float Rt[12];
#pragma unroll
    for (int t = 0; t < 12; ++t) Rt[t] = sRt[t];
    const float* tg = targets + int64_t(e) * 2 * HW;
    const float* wg = weights + int64_t(e) * 2 * HW;
    for (int p = threadIdx.x; p < HW; p += blockDim.x) {
      const int py = p / wd, px = p - py * wd;
      const Proj pr = project(Rt, (float(px) - cx) / fx, (float(py) - cy) / fy, disps[i * HW + p], fx, fy, cx, cy);
      const float r0 = tg[p] - pr.u, r1 = tg[HW + p] - pr.v;
      acc += double(kWeightScale * pr.valid * (wg[p] * r0 * r0 + wg[HW + p] * r1 * r1));
    }
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) sred[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t += sred[w];
    out[e] = t;
  }
}
}  // namespace

cudaError_t launch_chi2(const float* poses, const float* disps, const float* intr, const float* targets,
                        const float* weights, const int64_t* ii, const int64_t* jj, int64_t E, int N, int HW, int wd,
                        double* out, cudaStream_t st) {
  if (E <= 0) return cudaSuccess;
  k_chi2<<<unsigned(E), 256, 0, st>>>(poses, disps, intr, targets, weights, ii, jj, N, HW, wd, out);
  g_launch_count++;
  return cudaGetLastError();
}

cudaError_t launch_frame_distance(const float* poses, const float* disps, const float* intr, const int64_t* ii,
                                  const int64_t* jj, int64_t K, int N, int ht, int wd, float beta, float* out,
                                  cudaStream_t st) {
  k_frame_distance<<<unsigned(K), 256, 0, st>>>(poses, disps, intr, ii, jj, N, ht * wd, wd, beta, out);
  g_launch_count++;
  return cudaGetLastError();
}

cudaError_t launch_covis_distance(const float* poses, const float* disps, const float* intr, const int64_t* ii,
                                  int64_t K, int N, int ht, int wd, float* out, cudaStream_t st) {
  k_covis_distance<<<unsigned(K), 256, 0, st>>>(poses, disps, intr, ii, N, ht * wd, wd, ht, out);
  g_launch_count++;
  return cudaGetLastError();
}

cudaError_t launch_iproj(const float* poses, const float* disps, const float* intr, int64_t K, int ht, int wd,
                         float* out, cudaStream_t st) {
  dim3 grid((ht * wd + 255) / 256, unsigned(K));
  k_iproj<<<grid, 256, 0, st>>>(poses, disps, intr, ht * wd, wd, out);
  g_launch_count++;
  return cudaGetLastError();
}

cudaError_t launch_depth_filter(const float* poses, const float* disps, const float* intr, const int64_t* ix,
                                const float* thresh, int64_t K, int N, int ht, int wd, float* out,
                                cudaStream_t st) {
  dim3 grid((ht * wd + 255) / 256, unsigned(K));
  k_depth_filter<<<grid, 256, 0, st>>>(poses, disps, intr, ix, thresh, N, ht, wd, out);
  g_launch_count++;
  return cudaGetLastError();
}

}  // namespace dba
