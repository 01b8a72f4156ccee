// The two elementwise steps either side of `ba` inside FactorGraph.update:
//   reproject(ii, jj) -> coords [E,ht,wd,2], valid [E,ht,wd,1]
//       hislam2/depth_video.pyc reproject@170-178 = geom/projective_ops.py:44-74 without Jacobians
//   cvx_upsample(data, mask) -> [B,8ht,8wd,dim]
//       hislam2/modules/droid_net.py:21-38: per fine pixel a softmax over the 3x3 coarse neighbourhood (logits
//       mask[B,9,8,8,ht,wd]; neighbours outside the image get -inf) and the convex combination of the neighbours
// Both are HBM-bound streaming kernels: every input element is read once with full 128-byte warp transactions,
// every output written once (the upsampler transposes its 8x8 sub-pixel block through shared memory so that the
// rows of the fine image leave as contiguous runs).
#include "dba_kernels.h"
#include "dba_math.cuh"

namespace dba {

namespace {

constexpr int RP_T = 256;

__global__ void __launch_bounds__(RP_T)
k_reproject(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr,
            const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, int HW, int wd, float2* __restrict__ coords,
            float* __restrict__ valid) {
  __shared__ float sRt[12];
  const int e = blockIdx.y;
  const int i = int(ii[e]), j = int(jj[e]);
  if (threadIdx.x == 0) relative_pose_Rt(poses, i, j, sRt);
  __syncthreads();
  const int p = blockIdx.x * RP_T + threadIdx.x;
  if (p >= HW) return;
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const int py = p / wd, px = p - py * wd;
  const float d = disps[int64_t(i) * HW + p];
  const Proj pr = project(sRt, (float(px) - cx) / fx, (float(py) - cy) / fy, d, fx, fy, cx, cy);
  __stcs(coords + int64_t(e) * HW + p, make_float2(pr.u, pr.v));
  __stcs(valid + int64_t(e) * HW + p, pr.valid);
}

constexpr int UP_T = 256;          // 8 sub-columns j x 32 coarse pixels
constexpr int UP_C = 4;            // channels per pass

__global__ void __launch_bounds__(UP_T)
k_cvx_upsample(const float* __restrict__ data, const float* __restrict__ mask, int ht, int wd, int dim,
               float* __restrict__ up) {
  __shared__ float srow[8][32 * 8 * UP_C];
  const int b = blockIdx.z, y = blockIdx.y, x0 = blockIdx.x * 32;
  const int tid = threadIdx.x, j = tid >> 5, xx = tid & 31, x = x0 + xx;
  const int nx = min(32, wd - x0);
  const bool act = xx < nx;
  const int64_t plane = int64_t(ht) * wd;
  bool inside[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    const int yy = y + k / 3 - 1, xk = x + k % 3 - 1;
    inside[k] = act && yy >= 0 && yy < ht && xk >= 0 && xk < wd;
  }
  for (int c0 = 0; c0 < dim; c0 += UP_C) {
    const int nc = min(UP_C, dim - c0);
    float nb[9][UP_C];
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const int yy = y + k / 3 - 1, xk = x + k % 3 - 1;
#pragma unroll
      for (int c = 0; c < UP_C; ++c)
        nb[k][c] = (inside[k] && c < nc) ? data[((int64_t(b) * ht + yy) * wd + xk) * dim + c0 + c] : 0.f;
    }
#pragma unroll 2
    for (int i = 0; i < 8; ++i) {
      float l[9];
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        // mask viewed as [B,9,8,8,ht,wd] (droid_net.py:24)
        const int64_t off = (((int64_t(b) * 9 + k) * 8 + i) * 8 + j) * plane + int64_t(y) * wd + x;
        l[k] = inside[k] ? __ldcs(mask + off) : -INFINITY;
      }
      float mx = l[4];                                     // the centre is always inside
#pragma unroll
      for (int k = 0; k < 9; ++k) mx = fmaxf(mx, l[k]);
      float den = 0.f, acc[UP_C] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        const float ev = inside[k] ? expf(l[k] - mx) : 0.f;
        den += ev;
#pragma unroll
        for (int c = 0; c < UP_C; ++c) acc[c] = fmaf(ev, nb[k][c], acc[c]);
      }
      const float inv = 1.0f / den;
      if (act) {
#pragma unroll
        for (int c = 0; c < UP_C; ++c)
          if (c < nc) srow[i][(xx * 8 + j) * nc + c] = acc[c] * inv;
      }
    }
    __syncthreads();
    // fine rows 8y .. 8y+7, columns 8x0 .. 8x0 + 8nx - 1: contiguous runs when one pass covers all channels
    const int run = nx * 8 * nc;
    for (int i = 0; i < 8; ++i) {
      float* dst = up + ((int64_t(b) * 8 * ht + 8 * y + i) * 8 * wd + 8 * x0) * dim;
      for (int t = tid; t < run; t += UP_T) {
        const int col = t / nc, c = t - col * nc;
        __stcs(dst + int64_t(col) * dim + c0 + c, srow[i][t]);
      }
    }
    __syncthreads();
  }
}

}  // namespace

cudaError_t launch_reproject(const float* poses, const float* disps, const float* intr, const int64_t* ii,
                             const int64_t* jj, int E, int HW, int wd, float* coords, float* valid, cudaStream_t st) {
  if (E <= 0) return cudaSuccess;
  dim3 grid((HW + RP_T - 1) / RP_T, E);
  k_reproject<<<grid, RP_T, 0, st>>>(poses, disps, intr, ii, jj, HW, wd, reinterpret_cast<float2*>(coords), valid);
  g_launch_count++;
  return cudaGetLastError();
}

cudaError_t launch_cvx_upsample(const float* data, const float* mask, int B, int ht, int wd, int dim, float* up,
                                cudaStream_t st) {
  if (B <= 0) return cudaSuccess;
  dim3 grid((wd + 31) / 32, ht, B);
  k_cvx_upsample<<<grid, UP_T, 0, st>>>(data, mask, ht, wd, dim, up);
  g_launch_count++;
  return cudaGetLastError();
}

}  // namespace dba
