// Relative-pose estimate + diagonal covariance per loop-closure edge: PGOBuffer.add_rel_poses
// (hislam2/pgo_buffer.py:85-123).  Per edge (i -> j):
//     H = sum w Jj^T Jj + 1e-4 I ,  v = sum w Jj^T r ,  dx = H^-1 v          (w = .001 valid weight, r = target - coords)
//     rel = Exp(dx)^4 (G_j G_i^-1)       -- the reference relinearises at the unchanged `poses` four times (:99-110),
//                                           so the four retractions apply the same dx
//     cov = diag( (sum w (Jj dx - r)^2) H^-1 ) ,  valid = mean(weight > 0.1)
// One CTA per edge, two passes over the pixels (the Jacobian rows are recomputed, nothing is stored), fixed-order
// reductions, the 6x6 solve in fp64 by one thread.  targets / weights in the op layout [E,2,ht,wd].
#include "dba_kernels.h"
#include "dba_math.cuh"

namespace dba {

namespace {

constexpr int RPT = 256;
constexpr int NRED = 29;          // 21 (H upper) + 6 (v) + 1 (count) + 1 (pass 2: sigma^2)

struct JRows { float j0[6], j1[6], r0, r1, w0, w1; };

__device__ __forceinline__ JRows jrows(const float* Rt, float xx, float yy, float d, float fx, float fy, float cx,
                                       float cy, float tx, float ty, float wx, float wy) {
  const Proj pr = project(Rt, xx, yy, d, fx, fy, cx, cy);
  JRows o;
  // Jj_r = [d g_r, X1 x g_r], g0 = [a0, 0, c0], g1 = [0, b1, c1]   (dba_math.cuh)
  const float X = pr.X1.x, Y = pr.X1.y, Z = pr.X1.z;
  o.j0[0] = d * pr.a0; o.j0[1] = 0.f;       o.j0[2] = d * pr.c0;
  o.j0[3] = Y * pr.c0; o.j0[4] = Z * pr.a0 - X * pr.c0; o.j0[5] = -Y * pr.a0;
  o.j1[0] = 0.f;       o.j1[1] = d * pr.b1; o.j1[2] = d * pr.c1;
  o.j1[3] = Y * pr.c1 - Z * pr.b1; o.j1[4] = -X * pr.c1; o.j1[5] = X * pr.b1;
  o.r0 = tx - pr.u; o.r1 = ty - pr.v;
  o.w0 = kWeightScale * (pr.valid * wx); o.w1 = kWeightScale * (pr.valid * wy);
  return o;
}

// fixed-order block reduction of n values per thread -> sred[0..n) (valid for every thread after the call)
template <int N>
__device__ __forceinline__ void block_reduce(float (&v)[N], float* swarp, float* sred, int tid) {
  const int lane = tid & 31, warp = tid >> 5;
#pragma unroll
  for (int i = 0; i < N; ++i)
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
  if (lane == 0)
#pragma unroll
    for (int i = 0; i < N; ++i) swarp[warp * NRED + i] = v[i];
  __syncthreads();
  if (tid < N) {
    float s = 0.f;
    for (int w = 0; w < RPT / 32; ++w) s += swarp[w * NRED + tid];
    sred[tid] = s;
  }
  __syncthreads();
}

__global__ void __launch_bounds__(RPT)
k_rel_pose(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr,
           const float* __restrict__ targets, const float* __restrict__ weights, const int64_t* __restrict__ ii,
           const int64_t* __restrict__ jj, int N, int HW, int wd, float* __restrict__ rel_out, float* __restrict__ cov_out,
           float* __restrict__ valid_out) {
  __shared__ float sRt[12], swarp[(RPT / 32) * NRED], sred[NRED], sdx[6], sHd[6];
  const int e = blockIdx.x, tid = threadIdx.x;
  const int64_t i64 = ii[e], j64 = jj[e];
  if (i64 < 0 || i64 >= N || j64 < 0 || j64 >= N) {        // outside the frame buffer: the edge is reported as masked out
    if (tid < 7) rel_out[int64_t(e) * 7 + tid] = (tid == 6) ? 1.f : 0.f;
    if (tid < 6) cov_out[int64_t(e) * 6 + tid] = 0.f;
    if (tid == 0) valid_out[e] = 0.f;
    return;
  }
  const int i = int(i64), j = int(j64);
  if (tid == 0) relative_pose_Rt(poses, i, j, sRt);
  __syncthreads();
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const float* dsrc = disps + int64_t(i) * HW;
  const float* tg = targets + int64_t(e) * 2 * HW;
  const float* wg = weights + int64_t(e) * 2 * HW;
  float acc[28];
#pragma unroll
  for (int q = 0; q < 28; ++q) acc[q] = 0.f;
  for (int p = tid; p < HW; p += RPT) {
    const int py = p / wd, px = p - py * wd;
    const float wx = __ldg(wg + p), wy = __ldg(wg + HW + p);
    const JRows J = jrows(sRt, (float(px) - cx) / fx, (float(py) - cy) / fy, __ldg(dsrc + p), fx, fy, cx, cy, __ldg(tg + p),
                          __ldg(tg + HW + p), wx, wy);
    int q = 0;
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      const float a0 = J.w0 * J.j0[r], a1 = J.w1 * J.j1[r];
#pragma unroll
      for (int c = r; c < 6; ++c) { acc[q] = fmaf(a0, J.j0[c], fmaf(a1, J.j1[c], acc[q])); ++q; }
      acc[21 + r] = fmaf(a0, J.r0, fmaf(a1, J.r1, acc[21 + r]));
    }
    acc[27] += (wx > 0.1f ? 1.f : 0.f) + (wy > 0.1f ? 1.f : 0.f);
  }
  block_reduce<28>(acc, swarp, sred, tid);
  if (tid == 0) {
    // (H + 1e-4 I)^-1 and dx in fp64: Cholesky, L^-1, H^-1 = L^-T L^-1
    double H[6][6], L[6][6], W[6][6];
    int q = 0;
    for (int r = 0; r < 6; ++r)
      for (int c = r; c < 6; ++c) { H[r][c] = H[c][r] = double(sred[q]) + (r == c ? 1e-4 : 0.0); ++q; }
    for (int c = 0; c < 6; ++c)
      for (int r = 0; r < 6; ++r) { L[r][c] = 0.0; W[r][c] = 0.0; }
    for (int c = 0; c < 6; ++c) {
      double d = H[c][c];
      for (int k = 0; k < c; ++k) d -= L[c][k] * L[c][k];
      const double l = sqrt(d > 0.0 ? d : 1e-300);
      L[c][c] = l;
      for (int r = c + 1; r < 6; ++r) {
        double s = H[r][c];
        for (int k = 0; k < c; ++k) s -= L[r][k] * L[c][k];
        L[r][c] = s / l;
      }
    }
    for (int c = 0; c < 6; ++c)
      for (int r = c; r < 6; ++r) {
        double s = (r == c) ? 1.0 : 0.0;
        for (int k = c; k < r; ++k) s -= L[r][k] * W[k][c];
        W[r][c] = s / L[r][r];
      }
    for (int r = 0; r < 6; ++r) {
      double dx = 0.0;
      for (int c = 0; c < 6; ++c) {
        double h = 0.0;
        for (int k = (r > c ? r : c); k < 6; ++k) h += W[k][r] * W[k][c];
        dx += h * double(sred[21 + c]);
        if (c == r) sHd[r] = float(h);
      }
      sdx[r] = float(dx);
    }
  }
  __syncthreads();
  float dx[6];
#pragma unroll
  for (int r = 0; r < 6; ++r) dx[r] = sdx[r];
  float sg[1] = {0.f};
  for (int p = tid; p < HW; p += RPT) {
    const int py = p / wd, px = p - py * wd;
    const JRows J = jrows(sRt, (float(px) - cx) / fx, (float(py) - cy) / fy, __ldg(dsrc + p), fx, fy, cx, cy, __ldg(tg + p),
                          __ldg(tg + HW + p), __ldg(wg + p), __ldg(wg + HW + p));
    float v0 = -J.r0, v1 = -J.r1;
#pragma unroll
    for (int r = 0; r < 6; ++r) { v0 = fmaf(J.j0[r], dx[r], v0); v1 = fmaf(J.j1[r], dx[r], v1); }
    sg[0] = fmaf(J.w0 * v0, v0, fmaf(J.w1 * v1, v1, sg[0]));
  }
  const float cnt = sred[27];
  __syncthreads();                                   // sred is about to be reused
  block_reduce<1>(sg, swarp, sred, tid);
  if (tid == 0) {
    float rel[7];
    {
      SE3 gi = se3_load(poses + 7 * i), gj = se3_load(poses + 7 * j);
      SE3 g = se3_mul(gj, se3_inv(gi));
      rel[0] = g.t.x; rel[1] = g.t.y; rel[2] = g.t.z; rel[3] = g.q.x; rel[4] = g.q.y; rel[5] = g.q.z; rel[6] = g.q.w;
    }
    for (int it = 0; it < 4; ++it) se3_retract(rel, dx);
    for (int c = 0; c < 7; ++c) rel_out[int64_t(e) * 7 + c] = rel[c];
    for (int c = 0; c < 6; ++c) cov_out[int64_t(e) * 6 + c] = sred[0] * sHd[c];
    valid_out[e] = cnt / float(2 * HW);
  }
}

}  // namespace

cudaError_t launch_rel_pose(const float* poses, const float* disps, const float* intr, const float* targets,
                            const float* weights, const int64_t* ii, const int64_t* jj, int E, int N, int HW, int wd,
                            float* rel_out, float* cov_out, float* valid_out, cudaStream_t st) {
  if (E <= 0) return cudaSuccess;
  k_rel_pose<<<E, RPT, 0, st>>>(poses, disps, intr, targets, weights, ii, jj, N, HW, wd, rel_out, cov_out, valid_out);
  g_launch_count++;
  return cudaGetLastError();
}

}  // namespace dba
