// Device-side building blocks shared by the two linearisation kernels (SIMT Schur products in
// dba_linearize.cu, tcgen05 Schur products in dba_linearize_tc.cu): per-edge residuals / Jacobians
// of one pixel tile (phase 1) and the per-pixel inverse-depth block (phase 2).
//
// Maths follows hislam2/geom/ba.py:39-68,83-92 + chol.py:53-64 of the reference in the reduced
// form proven in tests/test_kernel_algebra.py (Ji/Hii/Hij/Ei are never formed per pixel).
#pragma once
#include "dba_kernels.h"
#include "dba_math.cuh"

namespace dba {

struct LinParams {
  const float* poses; const float* disps; const float* intr; const float* disps_sens;
  const float* targets; const float* weights; const float* eta;
  const int32_t *kx, *koff, *cedge, *cj, *pairbase, *blkbase;
  const int32_t* kf_list;     // source-keyframe slots handled by this launch (null: all, in order)
  float *Hpart, *Spart, *vEpart, *Q, *w;
  int HW, wd, S, tiles_per_split, ntiles, degmax;
};

// value i of every lane summed onto lane i (tests/test_kernel_algebra.py::test_butterfly_index_map)
__device__ __forceinline__ float warp_reduce_scatter32(float (&v)[32], int lane) {
#pragma unroll
  for (int mask = 16, cnt = 32; mask >= 1; mask >>= 1, cnt >>= 1) {
    const int half = cnt >> 1;
    const bool up = (lane & mask) != 0;
#pragma unroll
    for (int i = 0; i < half; ++i) {
      float send = up ? v[i] : v[i + half];
      float keep = up ? v[i + half] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, mask);
    }
  }
  return v[0];
}

__device__ __forceinline__ void decode_pair(int pi, int& A, int& B) {
  int a = int((sqrtf(8.0f * float(pi) + 1.0f) - 1.0f) * 0.5f);
  while ((a + 1) * (a + 2) / 2 <= pi) ++a;
  while (a * (a + 1) / 2 > pi) --a;
  A = a;
  B = pi - a * (a + 1) / 2;
}

// Phase 1, one warp, one edge, one tile of T = 32*PPL pixels starting at p0 (lanes stride over the
// pixels).  Adds the 21+6 sums of Hjj / vj of the tile onto hacc[0..26] (shared, owned by this warp)
// and, unless MOTION_ONLY, writes the 8 panel rows of the edge to pe[row*T + pixel]:
// rows 0-2 h, rows 3-5 X1 x h, row 6 ck, row 7 wk (phase 2 scales rows 0-2 by the disparity).
// the tile's streamed inputs of one edge (read once: evict-first): v = {tx, ty, wx, wy} x PPL
template <int PPL>
__device__ __forceinline__ void phase1_load(const float* __restrict__ tg, const float* __restrict__ wg, int HW, int p0,
                                            int lane, float (&v)[4 * PPL]) {
#pragma unroll
  for (int q = 0; q < PPL; ++q) {
    const int p = p0 + lane + 32 * q;
    const bool inb = p < HW;
    v[4 * q + 0] = inb ? __ldcs(tg + p) : 0.f;
    v[4 * q + 1] = inb ? __ldcs(tg + HW + p) : 0.f;
    v[4 * q + 2] = inb ? __ldcs(wg + p) : 0.f;
    v[4 * q + 3] = inb ? __ldcs(wg + HW + p) : 0.f;
  }
}

template <int PPL, bool MOTION_ONLY>
__device__ __forceinline__ void phase1_edge(const float (&v)[4 * PPL], int HW, int p0,
                                            int lane, const float* sRt_e, const float (&d)[PPL],
                                            const float (&x0)[PPL], const float (&y0)[PPL], float fx, float fy,
                                            float cx, float cy, float* pe, float* hacc) {
  constexpr int T = 32 * PPL;
  float Rt[12];
#pragma unroll
  for (int i = 0; i < 12; ++i) Rt[i] = sRt_e[i];
  float hs[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) hs[i] = 0.f;
#pragma unroll
  for (int q = 0; q < PPL; ++q) {
    const int pl = lane + 32 * q;
    const int p = p0 + pl;
    const bool inb = p < HW;
    const Proj pr = project(Rt, x0[q], y0[q], d[q], fx, fy, cx, cy);
    const float r0 = v[4 * q + 0] - pr.u, r1 = v[4 * q + 1] - pr.v;
    const float w0 = kWeightScale * (pr.valid * v[4 * q + 2]), w1 = kWeightScale * (pr.valid * v[4 * q + 3]);
    const float X = pr.X1.x, Y = pr.X1.y, Z = pr.X1.z, dd = d[q];
    // Jj rows: [d g, X1 x g]
    float J0[6] = {dd * pr.a0, 0.f, dd * pr.c0, Y * pr.c0, Z * pr.a0 - X * pr.c0, -Y * pr.a0};
    float J1[6] = {0.f, dd * pr.b1, dd * pr.c1, Y * pr.c1 - Z * pr.b1, -X * pr.c1, X * pr.b1};
    if (inb) {
      float wJ0[6], wJ1[6];
#pragma unroll
      for (int r = 0; r < 6; ++r) { wJ0[r] = w0 * J0[r]; wJ1[r] = w1 * J1[r]; }
#pragma unroll
      for (int r = 0; r < 6; ++r) {
#pragma unroll
        for (int c = 0; c <= r; ++c) {
          if (r != 1 && c != 1) hs[r * (r + 1) / 2 + c] += wJ0[r] * J0[c];
          if (r != 0 && c != 0) hs[r * (r + 1) / 2 + c] += wJ1[r] * J1[c];
        }
        if (r != 1) hs[21 + r] += wJ0[r] * r0;
        if (r != 0) hs[21 + r] += wJ1[r] * r1;
      }
    }
    if (!MOTION_ONLY) {
      const float a0 = w0 * pr.jz0, a1 = w1 * pr.jz1;
      float hx = a0 * pr.a0, hy = a1 * pr.b1, hz = a0 * pr.c0 + a1 * pr.c1;
      float ck = a0 * pr.jz0 + a1 * pr.jz1;
      float wk = a0 * r0 + a1 * r1;
      if (!inb) { hx = hy = hz = ck = wk = 0.f; }
      pe[0 * T + pl] = hx;                  // rows 0-2 hold h until phase 2 scales them by d
      pe[1 * T + pl] = hy;
      pe[2 * T + pl] = hz;
      pe[3 * T + pl] = Y * hz - Z * hy;     // X1 x h
      pe[4 * T + pl] = Z * hx - X * hz;
      pe[5 * T + pl] = X * hy - Y * hx;
      pe[6 * T + pl] = ck;
      pe[7 * T + pl] = wk;
    }
  }
  const float tot = warp_reduce_scatter32(hs, lane);
  if (lane < 27) hacc[lane] += tot;
}

// Phase 2, one thread, one pixel of the tile (local index pl): C = sum_e ck + eta + 1e-7 (+ the
// sensor-depth prior), w = sum_e wk, Q = 1/C, rows 0-2 of every edge block scaled by d, and the
// source-pose block E_ii = -[d G, X0 x G] with G = sum_e R_e^T h_e written to panel block 0.
// Q and w go to global (the back-substitution needs them); returns Q and w.
template <int T>
__device__ __forceinline__ void phase2_pixel(const LinParams& P, int m, int k, int deg, int p0, int pl, const float* dk,
                                             const float* sRt, float* spanel, int BS, float fx, float fy, float cx,
                                             float cy, float& Qv, float& wv) {
  const int p = p0 + pl;
  const bool inb = p < P.HW;
  const float dd = inb ? __ldg(dk + p) : 0.f;
  const int py = p / P.wd, px = p - py * P.wd;
  const float xx = (float(px) - cx) / fx, yy = (float(py) - cy) / fy;
  float C = 0.f, wsum = 0.f, Gx = 0.f, Gy = 0.f, Gz = 0.f;
  for (int le = 0; le < deg; ++le) {
    float* pe = spanel + (1 + le) * BS;
    const float* R = sRt + 12 * le;
    const float hx = pe[pl], hy = pe[T + pl], hz = pe[2 * T + pl];
    C += pe[6 * T + pl];
    wsum += pe[7 * T + pl];
    Gx += R[0] * hx + R[3] * hy + R[6] * hz;    // R^T h
    Gy += R[1] * hx + R[4] * hy + R[7] * hz;
    Gz += R[2] * hx + R[5] * hy + R[8] * hz;
    pe[pl] = dd * hx; pe[T + pl] = dd * hy; pe[2 * T + pl] = dd * hz;
  }
  Qv = 0.f; wv = 0.f;
  if (inb) {
    const float et = __ldg(P.eta + int64_t(m) * P.HW + p);
    if (P.disps_sens != nullptr) {
      const float ds = __ldg(P.disps_sens + int64_t(k) * P.HW + p);
      const float ms = ds > 0.f ? 1.f : 0.f;
      C = C + ms * kSensAlpha + (1.f - ms) * et + kCEps;
      wsum = wsum - ms * kSensAlpha * (dd - ds);
    } else {
      C = C + et + kCEps;
    }
    Qv = 1.0f / C;
    wv = wsum;
    P.Q[int64_t(m) * P.HW + p] = Qv;
    P.w[int64_t(m) * P.HW + p] = wv;
  }
  // E_ii = -[d G, X0 x G],  X0 = [xx, yy, 1]
  spanel[0 * T + pl] = -dd * Gx;
  spanel[1 * T + pl] = -dd * Gy;
  spanel[2 * T + pl] = -dd * Gz;
  spanel[3 * T + pl] = -(yy * Gz - Gy);
  spanel[4 * T + pl] = -(Gx - xx * Gz);
  spanel[5 * T + pl] = -(xx * Gy - yy * Gx);
}

}  // namespace dba
