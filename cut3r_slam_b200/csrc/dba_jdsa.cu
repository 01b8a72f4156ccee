// JDSA path of HI-SLAM2 (hislam2/geom/ba.py:161-241, chol.py:80-108): the three
// `droid_backends` ops / solves it needs.
//   proj_trans(poses, disps, intrinsics, target, weight, ii, jj) -> (C, w)   [ba.py:200]
//       per source keyframe and pixel: C = sum_edges sum_rows w Jz^2, w = sum w r Jz -- the depth block and
//       depth gradient of ba.py:62-68,83-90 WITHOUT eta (JDSA adds its own terms, ba.py:228-229)
//   bi_inter(scales, grid) -> (vals, J)                                        [ba.py:167]
//       bilinear interpolation of a coarse scale grid at fractional positions + the dense Jacobian w.r.t.
//       the grid nodes (4 non-zeros per pixel)
//   mono_prior_solve: schur_solve_mono_prior (chol.py:80-108) for the block structure JDSA builds: Hs, Es are
//       scattered at (kx, kx) only (ba.py:222-223), so the system is M independent D x D problems.
// The reference CUDA of proj_trans / bi_inter is absent: their definitions here are pinned to the torch
// statements around the call sites (oracle/jdsa.py); the solve is pinned to the reference's own
// schur_solve_mono_prior (tests/golden/jdsa_*.npz).
#include "dba_kernels.h"
#include "dba_math.cuh"

namespace dba {

namespace {

constexpr int PT_T = 128;
constexpr int PT_CHUNK = 64;

// one thread per pixel of (source keyframe slot m, 128-pixel tile): loop over the keyframe's edges
__global__ void __launch_bounds__(PT_T)
k_proj_trans(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr,
             const float* __restrict__ targets, const float* __restrict__ weights, const int32_t* __restrict__ kx,
             const int32_t* __restrict__ koff, const int32_t* __restrict__ cedge, const int32_t* __restrict__ cj,
             int HW, int wd, int ntiles, float* __restrict__ Cout, float* __restrict__ wout) {
  __shared__ float sRt[PT_CHUNK * 12];
  __shared__ int sedge[PT_CHUNK];
  const int m = blockIdx.x / ntiles, tile = blockIdx.x % ntiles;
  const int tid = threadIdx.x;
  const int k = kx[m];
  const int e0 = koff[m], deg = koff[m + 1] - e0;
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const int p = tile * PT_T + tid;
  const bool inb = p < HW;
  const float dd = inb ? disps[int64_t(k) * HW + p] : 0.f;
  const int py = p / wd, px = p - py * wd;
  const float xx = (float(px) - cx) / fx, yy = (float(py) - cy) / fy;
  float C = 0.f, w = 0.f;
  for (int c0 = 0; c0 < deg; c0 += PT_CHUNK) {
    const int nc = min(PT_CHUNK, deg - c0);
    __syncthreads();
    for (int le = tid; le < nc; le += PT_T) {
      relative_pose_Rt(poses, k, cj[e0 + c0 + le], sRt + 12 * le);
      sedge[le] = cedge[e0 + c0 + le];
    }
    __syncthreads();
    if (inb) {
      for (int le = 0; le < nc; ++le) {
        const float* tg = targets + int64_t(sedge[le]) * 2 * HW;
        const float* wg = weights + int64_t(sedge[le]) * 2 * HW;
        const float tx = __ldcs(tg + p), ty = __ldcs(tg + HW + p);
        const float wx = __ldcs(wg + p), wy = __ldcs(wg + HW + p);
        const Proj pr = project(sRt + 12 * le, xx, yy, dd, fx, fy, cx, cy);
        const float r0 = tx - pr.u, r1 = ty - pr.v;
        const float a0 = kWeightScale * (pr.valid * wx) * pr.jz0, a1 = kWeightScale * (pr.valid * wy) * pr.jz1;
        C += a0 * pr.jz0 + a1 * pr.jz1;          // ba.py:68
        w += a0 * r0 + a1 * r1;                  // ba.py:67
      }
    }
  }
  if (inb) {
    Cout[int64_t(m) * HW + p] = C;
    wout[int64_t(m) * HW + p] = w;
  }
}

// vals[m,p] = bilinear(scales[m], grid[m,p]) ; J[m,p,:] = d vals / d scales[m,:] (hs*ws entries, 4 non-zero).
// grid[...,0] runs along ws (x), grid[...,1] along hs (y), both in node units (ba.py:164-166).
__global__ void __launch_bounds__(256)
k_bi_inter(const float* __restrict__ scales, const float* __restrict__ grid, int HW, int hs, int ws,
           float* __restrict__ vals, float* __restrict__ J) {
  const int m = blockIdx.y;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= HW) return;
  const int D = hs * ws;
  const float gx = grid[(int64_t(m) * HW + p) * 2 + 0], gy = grid[(int64_t(m) * HW + p) * 2 + 1];
  int x0 = int(floorf(gx)), y0 = int(floorf(gy));
  x0 = max(0, min(x0, ws - 2 >= 0 ? ws - 2 : 0));
  y0 = max(0, min(y0, hs - 2 >= 0 ? hs - 2 : 0));
  const int x1 = min(x0 + 1, ws - 1), y1 = min(y0 + 1, hs - 1);
  const float wx = gx - float(x0), wy = gy - float(y0);
  const float w00 = (1.f - wy) * (1.f - wx), w01 = (1.f - wy) * wx, w10 = wy * (1.f - wx), w11 = wy * wx;
  const float* s = scales + int64_t(m) * D;
  vals[int64_t(m) * HW + p] = w00 * s[y0 * ws + x0] + w01 * s[y0 * ws + x1] + w10 * s[y1 * ws + x0] + w11 * s[y1 * ws + x1];
  float* Jp = J + (int64_t(m) * HW + p) * D;
  for (int i = 0; i < D; ++i) Jp[i] = 0.f;
  Jp[y0 * ws + x0] += w00;
  Jp[y0 * ws + x1] += w01;
  Jp[y1 * ws + x0] += w10;
  Jp[y1 * ws + x1] += w11;
}

}  // namespace

cudaError_t launch_proj_trans(const float* poses, const float* disps, const float* intr, const float* targets,
                              const float* weights, const int32_t* kx, const int32_t* koff, const int32_t* cedge,
                              const int32_t* cj, int M, int HW, int wd, float* Cout, float* wout, cudaStream_t st) {
  if (M == 0) return cudaSuccess;
  const int ntiles = (HW + PT_T - 1) / PT_T;
  k_proj_trans<<<M * ntiles, PT_T, 0, st>>>(poses, disps, intr, targets, weights, kx, koff, cedge, cj, HW, wd, ntiles,
                                            Cout, wout);
  g_launch_count++;
  return cudaGetLastError();
}

cudaError_t launch_bi_inter(const float* scales, const float* grid, int M, int HW, int hs, int ws, float* vals,
                            float* J, cudaStream_t st) {
  if (M == 0) return cudaSuccess;
  dim3 g((HW + 255) / 256, unsigned(M));
  k_bi_inter<<<g, 256, 0, st>>>(scales, grid, HW, hs, ws, vals, J);
  g_launch_count++;
  return cudaGetLastError();
}

}  // namespace dba
