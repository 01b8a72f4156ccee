// pgba: Sim3 pose-graph bundle adjustment (depth_video.pyc cuda_pgba@244-314 calls an absent
// CUDA export with 19 arguments).  Mathematics: the reference's `BA` (hislam2/geom/ba.py:32-107) with lietorch.Sim3
// poses -- 7-dof tangent [tau, phi, sigma], Sim3 column of the point Jacobian (geom/projective_ops.py:32-38), Sim3
// adjoint / Exp (thirdparty/lietorch/lietorch/include/sim3.h:91-104,156-166, rxso3.h:168-233) -- plus the relative-pose
// factor blocks of hislam2/pgo_buffer.py:19-62 (numeric Jacobians of log(Gij Gi Gj^-1), fp64, eps = 1e-8).
// Both parts are pinned to the unmodified reference (tests/golden/pgba.npz); how the absent op adds the factor blocks
// to the reprojection system is bytecode-only: restated in oracle/pgba.py (H += Hsp, v += vsp before the damping).
//
// This is the loop-closure event path (a few calls per run), built for correctness first: plain SIMT kernels, the
// per-edge E rows are materialised ([E,14,HW] floats: 2 GB at 512 keyframes / 12k edges), the reduced system is
// gathered in the fixed order of the same host plan the 6-dof path uses (dba_plan.cpp works in block units), the
// factorisation / solves are the 6-dof path's (n = 7 P').
#include "dba_kernels.h"
#include "dba_math.cuh"
#include "dba_plan.h"

namespace dba {

namespace {

constexpr int D7 = 7;

// ---- Sim3 on [t(3), q(xyzw), s] ------------------------------------------------------------------------------
template <class T> struct S3 { T t[3]; T q[4]; T s; };

template <class T> __device__ __forceinline__ S3<T> s3_load(const float* __restrict__ p) {
  S3<T> g;
  g.t[0] = T(p[0]); g.t[1] = T(p[1]); g.t[2] = T(p[2]);
  T x = T(p[3]), y = T(p[4]), z = T(p[5]), w = T(p[6]);
  const T n = T(1) / sqrt(x * x + y * y + z * z + w * w);          // so3.h:31-37: normalised on every load
  g.q[0] = x * n; g.q[1] = y * n; g.q[2] = z * n; g.q[3] = w * n;
  g.s = T(p[7]);
  return g;
}
template <class T> __device__ __forceinline__ void q_rot(const T* q, const T* p, T* o) {   // so3.h:55-60
  T uv[3] = {T(2) * (q[1] * p[2] - q[2] * p[1]), T(2) * (q[2] * p[0] - q[0] * p[2]), T(2) * (q[0] * p[1] - q[1] * p[0])};
  o[0] = p[0] + q[3] * uv[0] + (q[1] * uv[2] - q[2] * uv[1]);
  o[1] = p[1] + q[3] * uv[1] + (q[2] * uv[0] - q[0] * uv[2]);
  o[2] = p[2] + q[3] * uv[2] + (q[0] * uv[1] - q[1] * uv[0]);
}
template <class T> __device__ __forceinline__ void q_mul(const T* a, const T* b, T* o) {
  T r[4] = {a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1], a[3] * b[1] + a[1] * b[3] + a[2] * b[0] - a[0] * b[2],
            a[3] * b[2] + a[2] * b[3] + a[0] * b[1] - a[1] * b[0], a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2]};
  const T n = T(1) / sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2] + r[3] * r[3]);
  o[0] = r[0] * n; o[1] = r[1] * n; o[2] = r[2] * n; o[3] = r[3] * n;
}
template <class T> __device__ __forceinline__ S3<T> s3_mul(const S3<T>& a, const S3<T>& b) {   // sim3.h:51-53
  S3<T> r;
  T rt[3];
  q_rot(a.q, b.t, rt);
  for (int i = 0; i < 3; ++i) r.t[i] = a.t[i] + a.s * rt[i];
  q_mul(a.q, b.q, r.q);
  r.s = a.s * b.s;
  return r;
}
template <class T> __device__ __forceinline__ S3<T> s3_inv(const S3<T>& g) {                   // sim3.h:42-44
  S3<T> r;
  r.q[0] = -g.q[0]; r.q[1] = -g.q[1]; r.q[2] = -g.q[2]; r.q[3] = g.q[3];
  T rt[3];
  q_rot(r.q, g.t, rt);
  r.s = T(1) / g.s;
  for (int i = 0; i < 3; ++i) r.t[i] = -rt[i] * r.s;
  return r;
}
template <class T> __device__ __forceinline__ void q_to_R(const T* q, T* R) {                  // row-major
  const T tx = T(2) * q[0], ty = T(2) * q[1], tz = T(2) * q[2];
  const T twx = tx * q[3], twy = ty * q[3], twz = tz * q[3];
  const T txx = tx * q[0], txy = ty * q[0], txz = tz * q[0], tyy = ty * q[1], tyz = tz * q[1], tzz = tz * q[2];
  R[0] = T(1) - (tyy + tzz); R[1] = txy - twz;          R[2] = txz + twy;
  R[3] = txy + twz;          R[4] = T(1) - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy;          R[7] = tyz + twx;          R[8] = T(1) - (txx + tyy);
}
// W = A Phi + B Phi^2 + C I  (rxso3.h:190-233)
template <class T> __device__ __forceinline__ void s3_calcW(const T* phi, T sigma, T* W) {
  const T eps = T(1e-6);
  const T th2 = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  const T th = sqrt(th2);
  const T sc = exp(sigma);
  T A, B, C;
  if (fabs(sigma) < eps) {
    C = T(1);
    if (fabs(th) < eps) { A = T(0.5); B = T(1.0 / 6.0); }
    else { A = (T(1) - cos(th)) / th2; B = (th - sin(th)) / (th2 * th); }
  } else {
    C = (sc - T(1)) / sigma;
    const T s2 = sigma * sigma;
    if (fabs(th) < eps) {
      A = ((sigma - T(1)) * sc + T(1)) / s2;
      B = (sc * T(0.5) * s2 + sc - T(1) - sigma * sc) / (s2 * sigma);
    } else {
      const T a = sc * sin(th), b = sc * cos(th), c = th2 + s2;
      A = (a * sigma + (T(1) - b) * th) / (th * c);
      B = (C - ((b - T(1)) * sigma + a * th) / c) / th2;
    }
  }
  const T P[9] = {T(0), -phi[2], phi[1], phi[2], T(0), -phi[0], -phi[1], phi[0], T(0)};
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      T p2 = T(0);
      for (int k = 0; k < 3; ++k) p2 += P[i * 3 + k] * P[k * 3 + j];
      W[i * 3 + j] = A * P[i * 3 + j] + B * p2 + (i == j ? C : T(0));
    }
}
template <class T> __device__ __forceinline__ S3<T> s3_exp(const T* xi) {                      // sim3.h:156-166
  S3<T> g;
  const T* phi = xi + 3;
  const T th2 = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  const T th = sqrt(th2);
  T im, re;
  if (th < T(1e-6)) {
    const T th4 = th2 * th2;
    im = T(0.5) - th2 / T(48) + th4 / T(3840);
    re = T(1) - th2 / T(8) + th4 / T(384);
  } else { im = sin(T(0.5) * th) / th; re = cos(T(0.5) * th); }
  T q[4] = {im * phi[0], im * phi[1], im * phi[2], re};
  const T n = T(1) / sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  for (int i = 0; i < 4; ++i) g.q[i] = q[i] * n;
  g.s = exp(xi[6]);
  T W[9];
  s3_calcW(phi, xi[6], W);
  for (int i = 0; i < 3; ++i) g.t[i] = W[i * 3] * xi[0] + W[i * 3 + 1] * xi[1] + W[i * 3 + 2] * xi[2];
  return g;
}
// Log (sim3.h:145-154, rxso3.h:131-166): double only (relative-pose residuals)
__device__ __forceinline__ void s3_log(const S3<double>& g, double* xi) {
  const double n2 = g.q[0] * g.q[0] + g.q[1] * g.q[1] + g.q[2] * g.q[2], w = g.q[3];
  double f;
  if (n2 < 1e-12) f = 2.0 / w - (2.0 / 3.0) * n2 / (w * w * w);
  else {
    const double n = sqrt(n2);
    if (fabs(w) < 1e-6) f = (w > 0 ? 3.14159265358979323846 : -3.14159265358979323846) / n;
    else f = 2.0 * atan(n / w) / n;
  }
  xi[3] = f * g.q[0]; xi[4] = f * g.q[1]; xi[5] = f * g.q[2]; xi[6] = log(g.s);
  double W[9];
  s3_calcW(xi + 3, xi[6], W);
  // tau = W^-1 t  (3 x 3 by cofactors)
  const double det = W[0] * (W[4] * W[8] - W[5] * W[7]) - W[1] * (W[3] * W[8] - W[5] * W[6]) + W[2] * (W[3] * W[7] - W[4] * W[6]);
  const double id = 1.0 / det;
  const double* t = g.t;
  xi[0] = id * (t[0] * (W[4] * W[8] - W[5] * W[7]) - W[1] * (t[1] * W[8] - W[5] * t[2]) + W[2] * (t[1] * W[7] - W[4] * t[2]));
  xi[1] = id * (W[0] * (t[1] * W[8] - W[5] * t[2]) - t[0] * (W[3] * W[8] - W[5] * W[6]) + W[2] * (W[3] * t[2] - t[1] * W[6]));
  xi[2] = id * (W[0] * (W[4] * t[2] - t[1] * W[7]) - W[1] * (W[3] * t[2] - t[1] * W[6]) + t[0] * (W[3] * W[7] - W[4] * W[6]));
}

}  // namespace

namespace {

constexpr int PG_T = 256;
constexpr int NG = 105 + 14;           // lower triangle of the 14 x 14 Gram of [Ji | Jj] + the 14 gradient sums

// One CTA per edge (CSR position c): 7-dof Jacobians per pixel (projective_ops.py:15-74 with Sim3), the per-edge
// blocks Hii, Hij, Hji, Hjj, vi, vj (ba.py:54-59) by a block reduction, and the per-pixel rows Ei, Ej, Ck, wk
// (ba.py:62-68) to global memory: Ep [E][14][HW] (rows 0-6 Ei, 7-13 Ej), CW [E][2][HW].
__global__ void __launch_bounds__(PG_T)
k_pg_edge(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr,
          const float* __restrict__ targets, const float* __restrict__ weights, const int32_t* __restrict__ ci,
          const int32_t* __restrict__ cj, const int32_t* __restrict__ cedge, int HW, int wd,
          float* __restrict__ Hblk, float* __restrict__ vblk, float* __restrict__ Ep, float* __restrict__ CW) {
  __shared__ float sRt[12];            // s R (row-major) | t of the relative pose
  __shared__ float sAd[49];
  __shared__ float sred[PG_T / 32][NG];
  const int c = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int i = ci[c], j = cj[c];
  if (tid == 0) {
    const S3<float> gi = s3_load<float>(poses + 8 * i), gj = s3_load<float>(poses + 8 * j);
    const S3<float> g = s3_mul(gj, s3_inv(gi));                  // projective_ops.py:51
    float R[9];
    q_to_R(g.q, R);
    for (int k = 0; k < 9; ++k) sRt[k] = g.s * R[k];
    for (int k = 0; k < 3; ++k) sRt[9 + k] = g.t[k];
    // Ad = [[sR, [t]x R, -t], [0, R, 0], [0, 0, 1]]  (sim3.h:91-104)
    for (int k = 0; k < 49; ++k) sAd[k] = 0.f;
    const float tx = g.t[0], ty = g.t[1], tz = g.t[2];
    for (int cc = 0; cc < 3; ++cc) {
      const float r0 = R[cc], r1 = R[3 + cc], r2 = R[6 + cc];
      sAd[0 * 7 + cc] = g.s * r0; sAd[1 * 7 + cc] = g.s * r1; sAd[2 * 7 + cc] = g.s * r2;
      sAd[0 * 7 + 3 + cc] = ty * r2 - tz * r1;
      sAd[1 * 7 + 3 + cc] = tz * r0 - tx * r2;
      sAd[2 * 7 + 3 + cc] = tx * r1 - ty * r0;
      sAd[3 * 7 + 3 + cc] = r0; sAd[4 * 7 + 3 + cc] = r1; sAd[5 * 7 + 3 + cc] = r2;
    }
    sAd[0 * 7 + 6] = -tx; sAd[1 * 7 + 6] = -ty; sAd[2 * 7 + 6] = -tz;
    sAd[48] = 1.f;
  }
  __syncthreads();
  float Rt[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) Rt[k] = sRt[k];
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const float* tg = targets + int64_t(cedge[c]) * 2 * HW;
  const float* wg = weights + int64_t(cedge[c]) * 2 * HW;
  const float* dsrc = disps + int64_t(i) * HW;
  // Reduced form (the one the 6-dof kernels use, tests/test_kernel_algebra.py): Ji = -Jj Ad is never formed.  Per pixel
  // only the 7 x 7 Gram of Jj, vj and Ej are accumulated (35 sums instead of 119, 133 flops instead of 364); the
  // source-pose row is Ei = -Ad^T Ej, and the blocks follow once per edge: Hji = -Hjj Ad, Hij = Hji^T,
  // Hii = Ad^T Hjj Ad = -Ad^T Hji, vi = -Ad^T vj.
  constexpr int NR = 28 + 7;
  float acc[NR];
#pragma unroll
  for (int k = 0; k < NR; ++k) acc[k] = 0.f;
  for (int p = tid; p < HW; p += PG_T) {
    const int py = p / wd, px = p - py * wd;
    const float dd = dsrc[p];
    const Proj pr = project(Rt, (float(px) - cx) / fx, (float(py) - cy) / fy, dd, fx, fy, cx, cy);
    const float X = pr.X1.x, Y = pr.X1.y, Z = pr.X1.z;
    const float rr[2] = {tg[p] - pr.u, tg[HW + p] - pr.v};
    const float ww[2] = {kWeightScale * (pr.valid * wg[p]), kWeightScale * (pr.valid * wg[HW + p])};
    const float jz[2] = {pr.jz0, pr.jz1};
    float J[2][7];
    // Jj = Jp Ja: [d g, X1 x g, X1 . g]  with g0 = [a0, 0, c0], g1 = [0, b1, c1]
    J[0][0] = dd * pr.a0; J[0][1] = 0.f; J[0][2] = dd * pr.c0;
    J[0][3] = Y * pr.c0; J[0][4] = Z * pr.a0 - X * pr.c0; J[0][5] = -Y * pr.a0; J[0][6] = pr.a0 * X + pr.c0 * Z;
    J[1][0] = 0.f; J[1][1] = dd * pr.b1; J[1][2] = dd * pr.c1;
    J[1][3] = Y * pr.c1 - Z * pr.b1; J[1][4] = -X * pr.c1; J[1][5] = X * pr.b1; J[1][6] = pr.b1 * Y + pr.c1 * Z;
    float ej[7], ck = 0.f, wk = 0.f;
#pragma unroll
    for (int k = 0; k < 7; ++k) ej[k] = 0.f;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const float w = ww[r];
      int q = 0;
#pragma unroll
      for (int a = 0; a < 7; ++a) {
        const float wa = w * J[r][a];
#pragma unroll
        for (int bb = 0; bb <= a; ++bb) acc[q++] += wa * J[r][bb];
        acc[28 + a] += wa * rr[r];
        ej[a] += wa * jz[r];
      }
      ck += w * jz[r] * jz[r];
      wk += w * rr[r] * jz[r];
    }
#pragma unroll
    for (int cc = 0; cc < 7; ++cc) {                             // Ei = -Ad^T Ej
      float a = 0.f;
#pragma unroll
      for (int k = 0; k < 7; ++k) a += ej[k] * sAd[k * 7 + cc];
      Ep[(int64_t(c) * 14 + cc) * HW + p] = -a;
    }
#pragma unroll
    for (int k = 0; k < 7; ++k) Ep[(int64_t(c) * 14 + 7 + k) * HW + p] = ej[k];
    CW[(int64_t(c) * 2) * HW + p] = ck;
    CW[(int64_t(c) * 2 + 1) * HW + p] = wk;
  }
#pragma unroll
  for (int k = 0; k < NR; ++k) {
    float v = acc[k];
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) sred[warp][k] = v;
  }
  __syncthreads();
  __shared__ float sHjj[49], sHji[49], svj[7];
  if (tid < NR) {
    float v = 0.f;
    for (int w = 0; w < PG_T / 32; ++w) v += sred[w][tid];
    if (tid < 28) {
      int a = 0;
      while ((a + 1) * (a + 2) / 2 <= tid) ++a;
      const int bb = tid - a * (a + 1) / 2;
      sHjj[a * 7 + bb] = v; sHjj[bb * 7 + a] = v;
    } else {
      svj[tid - 28] = v;
    }
  }
  __syncthreads();
  float* out = Hblk + int64_t(c) * 4 * 49;
  if (tid < 49) {                                                // Hji = -Hjj Ad
    const int r = tid / 7, cc = tid - r * 7;
    float a = 0.f;
#pragma unroll
    for (int k = 0; k < 7; ++k) a += sHjj[r * 7 + k] * sAd[k * 7 + cc];
    sHji[tid] = -a;
    out[2 * 49 + tid] = -a;                                      // Hji
    out[1 * 49 + cc * 7 + r] = -a;                               // Hij = Hji^T
    out[3 * 49 + tid] = sHjj[tid];                               // Hjj
  } else if (tid < 56) {                                         // vi = -Ad^T vj
    const int cc = tid - 49;
    float a = 0.f;
#pragma unroll
    for (int k = 0; k < 7; ++k) a += sAd[k * 7 + cc] * svj[k];
    vblk[int64_t(c) * 14 + cc] = -a;
    vblk[int64_t(c) * 14 + 7 + cc] = svj[cc];
  }
  __syncthreads();
  if (tid < 49) {                                                // Hii = -Ad^T Hji
    const int r = tid / 7, cc = tid - r * 7;
    float a = 0.f;
#pragma unroll
    for (int k = 0; k < 7; ++k) a += sAd[k * 7 + r] * sHji[k * 7 + cc];
    out[0 * 49 + tid] = -a;
  }
}

// per source keyframe pixel: C = sum Ck + eta + 1e-7, w = sum wk, Q = 1 / C; the source-pose row sum_e Ei
__global__ void __launch_bounds__(256)
k_pg_cq(const float* __restrict__ CW, const float* __restrict__ Ep, const float* __restrict__ eta,
        const int32_t* __restrict__ koff, int HW, float* __restrict__ Q, float* __restrict__ Wv, float* __restrict__ Eis) {
  const int m = blockIdx.y, p = blockIdx.x * 256 + threadIdx.x;
  if (p >= HW) return;
  float C = 0.f, w = 0.f, ei[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int c = koff[m]; c < koff[m + 1]; ++c) {
    C += CW[(int64_t(c) * 2) * HW + p];
    w += CW[(int64_t(c) * 2 + 1) * HW + p];
#pragma unroll
    for (int k = 0; k < 7; ++k) ei[k] += Ep[(int64_t(c) * 14 + k) * HW + p];
  }
  C += eta[int64_t(m) * HW + p] + kCEps;
  Q[int64_t(m) * HW + p] = 1.0f / C;
  Wv[int64_t(m) * HW + p] = w;
#pragma unroll
  for (int k = 0; k < 7; ++k) Eis[(int64_t(m) * 7 + k) * HW + p] = ei[k];
}

}  // namespace

namespace {

// Schur products of one source keyframe m and one panel-block pair (A >= B): X = sum_p Q E_A E_B^T (7 x 7) into
// Spart [pairbase[m] + tri(A,B)][49]; the diagonal pairs also write E_A (Q w) into vEpart [blkbase[m] + A][7].
// Panel block 0 = the source pose (Eis), block 1 + le = target pose of the keyframe's edge le (rows 7-13 of Ep).
__global__ void __launch_bounds__(PG_T)
k_pg_schur(const float* __restrict__ Ep, const float* __restrict__ Eis, const float* __restrict__ Q,
           const float* __restrict__ Wv, const int32_t* __restrict__ koff, const int32_t* __restrict__ pairbase,
           const int32_t* __restrict__ blkbase, int HW, float* __restrict__ Spart, float* __restrict__ vEpart) {
  __shared__ float sred[PG_T / 32][56];
  const int m = blockIdx.y, pi = blockIdx.x;
  const int e0 = koff[m], nblk = koff[m + 1] - e0 + 1;
  if (pi >= nblk * (nblk + 1) / 2) return;
  int A = 0;
  while ((A + 1) * (A + 2) / 2 <= pi) ++A;
  const int B = pi - A * (A + 1) / 2;
  const float* ea = (A == 0) ? Eis + int64_t(m) * 7 * HW : Ep + (int64_t(e0 + A - 1) * 14 + 7) * HW;
  const float* eb = (B == 0) ? Eis + int64_t(m) * 7 * HW : Ep + (int64_t(e0 + B - 1) * 14 + 7) * HW;
  const float* q = Q + int64_t(m) * HW;
  const float* wv = Wv + int64_t(m) * HW;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float acc[56];
#pragma unroll
  for (int k = 0; k < 56; ++k) acc[k] = 0.f;
  for (int p = tid; p < HW; p += PG_T) {
    const float qq = q[p], qw = qq * wv[p];
    float a[7], b[7];
#pragma unroll
    for (int k = 0; k < 7; ++k) { a[k] = ea[int64_t(k) * HW + p]; b[k] = eb[int64_t(k) * HW + p]; }
#pragma unroll
    for (int r = 0; r < 7; ++r) {
      const float qa = qq * a[r];
#pragma unroll
      for (int c = 0; c < 7; ++c) acc[r * 7 + c] += qa * b[c];
      acc[49 + r] += qw * a[r];
    }
  }
#pragma unroll
  for (int k = 0; k < 56; ++k) {
    float v = acc[k];
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) sred[warp][k] = v;
  }
  __syncthreads();
  if (tid < 56) {
    float v = 0.f;
    for (int w = 0; w < PG_T / 32; ++w) v += sred[w][tid];
    if (tid < 49) Spart[(int64_t(pairbase[m]) + pi) * 49 + tid] = v;
    else if (A == B) vEpart[(int64_t(blkbase[m]) + A) * 7 + (tid - 49)] = v;
  }
}

// The same products with one CTA per (keyframe, panel block A): the seven rows of E_A are staged once in shared memory
// (7 HW floats, 93 KB at 43 x 77: two CTAs per SM) and the CTA walks B = 0 .. A, so only the E_B rows stream from L2 --
// half the traffic of the pair-per-CTA kernel (3.5 GB per iteration on the backend graph, the L2 bound of pgba).
// Each thread still adds its pixels in ascending order: bit-identical slabs.  Largest A first.
__global__ void __launch_bounds__(PG_T)
k_pg_schur_rows(const float* __restrict__ Ep, const float* __restrict__ Eis, const float* __restrict__ Q,
                const float* __restrict__ Wv, const int32_t* __restrict__ koff, const int32_t* __restrict__ pairbase,
                const int32_t* __restrict__ blkbase, int HW, int nblkmax, float* __restrict__ Spart,
                float* __restrict__ vEpart) {
  extern __shared__ __align__(16) float sEa[];              // [7][HW]
  __shared__ float sred[PG_T / 32][56];
  const int m = blockIdx.y, A = nblkmax - 1 - int(blockIdx.x);
  const int e0 = koff[m], nblk = koff[m + 1] - e0 + 1;
  if (A >= nblk) return;
  const float* ea = (A == 0) ? Eis + int64_t(m) * 7 * HW : Ep + (int64_t(e0 + A - 1) * 14 + 7) * HW;
  const float* q = Q + int64_t(m) * HW;
  const float* wv = Wv + int64_t(m) * HW;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < 7 * HW; i += PG_T) sEa[i] = ea[i];
  __syncthreads();
  for (int B = 0; B <= A; ++B) {
    const float* eb = (B == 0) ? Eis + int64_t(m) * 7 * HW : Ep + (int64_t(e0 + B - 1) * 14 + 7) * HW;
    float acc[56];
#pragma unroll
    for (int k = 0; k < 56; ++k) acc[k] = 0.f;
    // two pixels per pass: sixteen L2 loads in flight per thread (two CTAs per SM leave only 16 warps to hide them)
    for (int p = tid; p < HW; p += 2 * PG_T) {
      const int p1 = p + PG_T;
      const bool two = p1 < HW;
      float b[7], b1[7];
#pragma unroll
      for (int k = 0; k < 7; ++k) { b[k] = eb[int64_t(k) * HW + p]; b1[k] = two ? eb[int64_t(k) * HW + p1] : 0.f; }
      const float qq = q[p], qw = qq * wv[p];
      const float qq1 = two ? q[p1] : 0.f, qw1 = two ? qq1 * wv[p1] : 0.f;
#pragma unroll
      for (int r = 0; r < 7; ++r) {
        const float ar = sEa[r * HW + p];
        const float qa = qq * ar;
#pragma unroll
        for (int c = 0; c < 7; ++c) acc[r * 7 + c] += qa * b[c];
        acc[49 + r] += qw * ar;
      }
      if (two) {
#pragma unroll
        for (int r = 0; r < 7; ++r) {
          const float ar = sEa[r * HW + p1];
          const float qa = qq1 * ar;
#pragma unroll
          for (int c = 0; c < 7; ++c) acc[r * 7 + c] += qa * b1[c];
          acc[49 + r] += qw1 * ar;
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 56; ++k) {
      float v = acc[k];
#pragma unroll
      for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) sred[warp][k] = v;
    }
    __syncthreads();
    if (tid < 56) {
      float v = 0.f;
      for (int w = 0; w < PG_T / 32; ++w) v += sred[w][tid];
      const int pi = A * (A + 1) / 2 + B;
      if (tid < 49) Spart[(int64_t(pairbase[m]) + pi) * 49 + tid] = v;
      else if (A == B) vEpart[(int64_t(blkbase[m]) + A) * 7 + (tid - 49)] = v;
    }
    __syncthreads();
  }
}

// reduced system: block ob of the lower block triangle (CTAs [0, nSblk)) or the rhs of pose a (CTAs nSblk ..),
// gathered in the plan's fixed order with fp64 sums; chol.py:55-57 damping on the true diagonal
__global__ void __launch_bounds__(64)
k_pg_assemble(const int32_t* __restrict__ soff, const Item* __restrict__ sitems, const int32_t* __restrict__ roff,
              const Item* __restrict__ ritems, const float* __restrict__ Hblk, const float* __restrict__ vblk,
              const float* __restrict__ Spart, const float* __restrict__ vEpart, int nSblk, int Pp, float lm, float ep,
              float* __restrict__ Amat, int ld) {
  const int t = threadIdx.x;
  if (int(blockIdx.x) >= nSblk) {
    const int a = blockIdx.x - nSblk;
    if (t >= 7) return;
    double vs = 0.0, es = 0.0;
    for (int i = roff[a]; i < roff[a + 1]; ++i) {
      const Item it = ritems[i];
      if ((it.flags & 1) == 0) vs += double(vblk[int64_t(it.base) * 7 + t]);
      else es += double(vEpart[int64_t(it.base) * 7 + t]);
    }
    Amat[int64_t(7 * Pp) * ld + a * 7 + t] = float(vs - es);
    return;
  }
  if (t >= 49) return;
  const int ob = blockIdx.x;
  int a = 0;
  while ((a + 1) * (a + 2) / 2 <= ob) ++a;
  const int b = ob - a * (a + 1) / 2;
  const int r = t / 7, c = t - r * 7, tT = c * 7 + r;
  double hs = 0.0, ss = 0.0;
  for (int i = soff[ob]; i < soff[ob + 1]; ++i) {
    const Item it = sitems[i];
    const int e = (it.flags & 2) ? tT : t;
    if ((it.flags & 1) == 0) hs += double(Hblk[int64_t(it.base) * 49 + e]);
    else ss += double(Spart[int64_t(it.base) * 49 + e]);
  }
  // the relative-pose blocks are added later (k_pg_rel_scatter) with the same (1 + lm) on the true diagonal
  double val = (a == b && r == c) ? hs + (double(ep) + double(lm) * hs) : hs;
  Amat[int64_t(7 * a + r) * ld + 7 * b + c] = float(val - ss);
}

// Relative-pose factor blocks (pgo_buffer.py:19-62): one CTA per factor, thread k < 29 evaluates
// f = log(Gij G_i' G_j'^-1) for one perturbed pair (14 columns x +-eps, and the unperturbed residual), fp64.
__global__ void __launch_bounds__(64)
k_pg_rel_blocks(const float* __restrict__ poses, const int64_t* __restrict__ rii, const int64_t* __restrict__ rjj,
                const float* __restrict__ rel, const float* __restrict__ infos, float pw, int N,
                float* __restrict__ Hsp, float* __restrict__ vsp) {
  __shared__ double sf[29][7];
  __shared__ float sJ[2][7][7];          // [which pose][residual row][tangent column]
  __shared__ float swJT[2][7][7];        // (pw J)^T info
  __shared__ float sr[7];
  const int f = blockIdx.x, t = threadIdx.x;
  const int64_t i = rii[f], j = rjj[f];
  if (i < 0 || i >= N || j < 0 || j >= N) return;
  if (t < 29) {
    S3<double> gi = s3_load<double>(poses + 8 * i), gj = s3_load<double>(poses + 8 * j);
    float r8[8];
    for (int k = 0; k < 7; ++k) r8[k] = rel[int64_t(f) * 7 + k];
    r8[7] = 1.f;
    const S3<double> gij = s3_load<double>(r8);
    if (t < 28) {
      const int col = (t >> 1) % 7, first = t < 14;
      double d[7] = {0, 0, 0, 0, 0, 0, 0};
      d[col] = (t & 1) ? -1e-8 : 1e-8;
      const S3<double> e = s3_exp(d);
      if (first) gi = s3_mul(e, gi); else gj = s3_mul(e, gj);
    }
    double xi[7];
    s3_log(s3_mul(s3_mul(gij, gi), s3_inv(gj)), xi);
    for (int k = 0; k < 7; ++k) sf[t][k] = xi[k];
  }
  __syncthreads();
  if (t < 49) {                          // J[which][row][col] = (f+ - f-) / 2 / eps, cast to fp32 as the reference does
    const int row = t / 7, col = t - row * 7;
    for (int which = 0; which < 2; ++which) {
      const int base = which * 14 + 2 * col;
      sJ[which][row][col] = float((sf[base][row] - sf[base + 1][row]) / 2.0 / 1e-8);
    }
    if (col == 0) sr[row] = float(sf[28][row]);
  }
  __syncthreads();
  if (t < 49) {                          // wJT = ((pw J)^T info) in fp64, cast to fp32
    const int a = t / 7, c = t - a * 7;
    for (int which = 0; which < 2; ++which) {
      double s = 0.0;
      for (int k = 0; k < 7; ++k) s += double(pw) * double(sJ[which][k][a]) * double(infos[(int64_t(f) * 7 + k) * 7 + c]);
      swJT[which][a][c] = float(s);
    }
  }
  __syncthreads();
  if (t < 49) {
    const int a = t / 7, c = t - a * 7;
    for (int wa = 0; wa < 2; ++wa)
      for (int wb = 0; wb < 2; ++wb) {   // Hsp[2 wa + wb] = wJ_a^T J_b
        float s = 0.f;
        for (int k = 0; k < 7; ++k) s += swJT[wa][a][k] * sJ[wb][k][c];
        Hsp[((int64_t(2 * wa + wb) * gridDim.x + f) * 7 + a) * 7 + c] = s;
      }
    if (c == 0)
      for (int wa = 0; wa < 2; ++wa) {
        float s = 0.f;
        for (int k = 0; k < 7; ++k) s += swJT[wa][a][k] * sr[k];
        vsp[(int64_t(wa) * gridDim.x + f) * 7 + a] = -s;
      }
  }
}

// factor blocks into the (lower) reduced system and its rhs row, one factor after the other (fixed order)
__global__ void __launch_bounds__(64)
k_pg_rel_scatter(const float* __restrict__ Hsp, const float* __restrict__ vsp, const int64_t* __restrict__ rii,
                 const int64_t* __restrict__ rjj, int R, int t0, int Pp, float lm, float* __restrict__ Amat, int ld) {
  const int t = threadIdx.x;
  const int r = t / 7, c = t - r * 7;
  for (int f = 0; f < R; ++f) {
    const int a = int(rii[f]) - t0, b = int(rjj[f]) - t0;
    const bool ia = a >= 0 && a < Pp, ib = b >= 0 && b < Pp;
    if (t < 49) {
      if (ia) Amat[int64_t(7 * a + r) * ld + 7 * a + c] += Hsp[((int64_t(0) * R + f) * 7 + r) * 7 + c] * (r == c ? 1.f + lm : 1.f);
      if (ib) Amat[int64_t(7 * b + r) * ld + 7 * b + c] += Hsp[((int64_t(3) * R + f) * 7 + r) * 7 + c] * (r == c ? 1.f + lm : 1.f);
      if (ia && ib) {
        if (a > b) Amat[int64_t(7 * a + r) * ld + 7 * b + c] += Hsp[((int64_t(1) * R + f) * 7 + r) * 7 + c];
        else if (b > a) Amat[int64_t(7 * b + r) * ld + 7 * a + c] += Hsp[((int64_t(2) * R + f) * 7 + r) * 7 + c];
      }
    } else if (t < 56) {
      const int k = t - 49;
      if (ia) Amat[int64_t(7 * Pp) * ld + 7 * a + k] += vsp[(int64_t(0) * R + f) * 7 + k];
      if (ib) Amat[int64_t(7 * Pp) * ld + 7 * b + k] += vsp[(int64_t(1) * R + f) * 7 + k];
    }
    // no barrier between factors: thread t owns entry (r, c) of EVERY block it touches, so two factors that meet in
    // one block are added by the same thread, in factor order
  }
}

}  // namespace

namespace {

// dz = Q (w - E^T dx) (chol.py:70), disparity retraction + guards of the source keyframes (ba.py:102-105)
__global__ void __launch_bounds__(256)
k_pg_backsub(const float* __restrict__ Ep, const float* __restrict__ Eis, const float* __restrict__ Q,
             const float* __restrict__ Wv, const float* __restrict__ dx, const int32_t* __restrict__ kx,
             const int32_t* __restrict__ koff, const int32_t* __restrict__ cj, int HW, int t0, int Pp,
             float* __restrict__ disps, float* __restrict__ dz) {
  const int m = blockIdx.y, p = blockIdx.x * 256 + threadIdx.x;
  if (p >= HW) return;
  const int k = kx[m];
  float s = Wv[int64_t(m) * HW + p];
  const int a = k - t0;
  if (a >= 0 && a < Pp) {
#pragma unroll
    for (int r = 0; r < 7; ++r) s -= Eis[(int64_t(m) * 7 + r) * HW + p] * dx[a * 7 + r];
  }
  for (int c = koff[m]; c < koff[m + 1]; ++c) {
    const int b = cj[c] - t0;
    if (b < 0 || b >= Pp) continue;
#pragma unroll
    for (int r = 0; r < 7; ++r) s -= Ep[(int64_t(c) * 14 + 7 + r) * HW + p] * dx[b * 7 + r];
  }
  const float d = Q[int64_t(m) * HW + p] * s;
  dz[int64_t(m) * HW + p] = d;
  float nd = disps[int64_t(k) * HW + p] + d;
  if (nd > 10.f) nd = 0.f;
  disps[int64_t(k) * HW + p] = fmaxf(nd, 0.001f);
}

// guards over the whole slice disps[:t1] (ba.py:104-105), applied once per call
__global__ void __launch_bounds__(256)
k_pg_guard(float* __restrict__ disps, int64_t n) {
  const int64_t i = int64_t(blockIdx.x) * 256 + threadIdx.x;
  if (i >= n) return;
  float d = disps[i];
  if (d > 10.f) d = 0.f;
  disps[i] = fmaxf(d, 0.001f);
}

// pose <- Exp(dx) * pose for the frames [t0, t1) (groups.py:153-156 with Sim3)
__global__ void __launch_bounds__(128)
k_pg_retract(float* __restrict__ poses, const float* __restrict__ dx, int t0, int Pp, float* __restrict__ dx_out) {
  const int a = blockIdx.x * 128 + threadIdx.x;
  if (a >= Pp) return;
  // Exp in fp64: the scale updates of a converging solve are 1e-6 .. 1e-5, where W's coefficient C = (e^sigma - 1) / sigma
  // (rxso3.h:190-233, general branch above |sigma| = 1e-6) keeps two digits in fp32 and B none -- the translation of
  // Exp(dx) is then off by ~1 % of |tau| (3e-4 on a 2e-2 step; tools/sim3_state_diag.py: the state after one iteration
  // was 2.8e-4 from the fp64 result, and the NEXT update 1.7e-2 off, although this update was exact to 6e-5).  The
  // reference evaluates the same formulas in fp32 and carries the same noise (its own fp32 run: 3.6e-3 on that update);
  // one thread per pose makes fp64 free here.
  double xi[7];
  for (int k = 0; k < 7; ++k) { xi[k] = double(dx[a * 7 + k]); if (dx_out) dx_out[a * 7 + k] = dx[a * 7 + k]; }
  float* p = poses + 8 * (t0 + a);
  const S3<double> g = s3_mul(s3_exp(xi), s3_load<double>(p));
  p[0] = float(g.t[0]); p[1] = float(g.t[1]); p[2] = float(g.t[2]);
  p[3] = float(g.q[0]); p[4] = float(g.q[1]); p[5] = float(g.q[2]); p[6] = float(g.q[3]);
  p[7] = float(g.s);
}

}  // namespace

// ---- one linearisation + assembly (everything before the factorisation) and the two steps after the solve ------
cudaError_t launch_pgba_linearize(const PgbaArgs& a, cudaStream_t st) {
  const int HW = a.HW;
  k_pg_edge<<<a.E, PG_T, 0, st>>>(a.poses, a.disps, a.intr, a.targets, a.weights, a.ci, a.cj, a.cedge, HW, a.wd, a.Hblk,
                                  a.vblk, a.Ep, a.CW);
  k_pg_cq<<<dim3((HW + 255) / 256, a.M), 256, 0, st>>>(a.CW, a.Ep, a.eta, a.koff, HW, a.Q, a.Wv, a.Eis);
  const size_t smem_rows = size_t(7) * HW * sizeof(float);
  if (a.nblkmax > 0 && smem_rows <= size_t(100) * 1024) {     // two CTAs per SM; larger images keep the pair-per-CTA kernel
    cudaError_t e = cudaFuncSetAttribute(k_pg_schur_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem_rows));
    if (e != cudaSuccess) return e;
    k_pg_schur_rows<<<dim3(a.nblkmax, a.M), PG_T, smem_rows, st>>>(a.Ep, a.Eis, a.Q, a.Wv, a.koff, a.pairbase, a.blkbase, HW,
                                                                   a.nblkmax, a.Spart, a.vEpart);
  } else {
    k_pg_schur<<<dim3(a.npairs_max, a.M), PG_T, 0, st>>>(a.Ep, a.Eis, a.Q, a.Wv, a.koff, a.pairbase, a.blkbase, HW, a.Spart,
                                                         a.vEpart);
  }
  k_pg_assemble<<<a.nSblk + a.Pp, 64, 0, st>>>(a.sblk_off, a.sblk_items, a.rhs_off, a.rhs_items, a.Hblk, a.vblk, a.Spart,
                                               a.vEpart, a.nSblk, a.Pp, a.lm, a.ep, a.A, a.ldA);
  g_launch_count += 4;
  if (a.R > 0) {
    k_pg_rel_blocks<<<a.R, 64, 0, st>>>(a.poses, a.rel_ii, a.rel_jj, a.rel_poses, a.rel_infos, a.pw, a.N, a.Hsp, a.vsp);
    k_pg_rel_scatter<<<1, 64, 0, st>>>(a.Hsp, a.vsp, a.rel_ii, a.rel_jj, a.R, a.t0, a.Pp, a.lm, a.A, a.ldA);
    g_launch_count += 2;
  }
  return cudaGetLastError();
}

cudaError_t launch_pgba_update(const PgbaArgs& a, float* dx_out, cudaStream_t st) {
  k_pg_backsub<<<dim3((a.HW + 255) / 256, a.M), 256, 0, st>>>(a.Ep, a.Eis, a.Q, a.Wv, a.dx, a.kx, a.koff, a.cj, a.HW, a.t0,
                                                            a.Pp, a.disps, a.dz);
  k_pg_retract<<<(a.Pp + 127) / 128, 128, 0, st>>>(a.poses, a.dx, a.t0, a.Pp, dx_out);
  g_launch_count += 2;
  return cudaGetLastError();
}

cudaError_t launch_pgba_guard(float* disps, int64_t n, cudaStream_t st) {
  k_pg_guard<<<unsigned((n + 255) / 256), 256, 0, st>>>(disps, n);
  g_launch_count++;
  return cudaGetLastError();
}

}  // namespace dba
