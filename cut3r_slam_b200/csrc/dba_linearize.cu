// Kernel 1-3 of the DBA iteration, fused: per source keyframe, per pixel tile
//   phase 1  residuals + Jacobians of every edge leaving the keyframe (one warp per edge, lanes
//            stride over the tile's pixels): the 21+6 per-edge sums of Hjj / vj are reduced with a
//            warp reduce-scatter (no atomics); the depth-coupling panel E (6 rows per edge) is
//            staged in shared memory and never leaves the SM;
//   phase 2  per pixel: C = sum_e Ck + eta + 1e-7, w = sum_e wk, Q = 1/C, the source-pose panel
//            block E_ii = -[d G, X0 x G] with G = sum_e R_e^T h_e;
//   phase 3  Schur products  X_AB = sum_p Q_p E_A[:,p] E_B[:,p]^T  for every block pair of the
//            keyframe's panel and  E_A (Q w), accumulated in registers across the CTA's tiles.
//
// Maths follows hislam2/geom/ba.py:39-68,83-92 + chol.py:53-64 of the reference in the reduced
// form proven in tests/test_kernel_algebra.py (Ji/Hii/Hij/Ei are never formed per pixel).
#include "dba_kernels.h"
#include "dba_math.cuh"

namespace dba {

namespace {

constexpr int NT = 256;   // threads per CTA
constexpr int NW = NT / 32;

struct LinParams {
  const float* poses; const float* disps; const float* intr; const float* disps_sens;
  const float* targets; const float* weights; const float* eta;
  const int32_t *kx, *koff, *cedge, *cj, *pairbase, *blkbase;
  float *Hpart, *Spart, *vEpart, *Q, *w;
  int HW, wd, S, tiles_per_split, ntiles, degmax;
};

__host__ __device__ inline int round4(int x) { return (x + 3) & ~3; }

struct SmemLayout {
  int Rt, Hacc, q, qw, panel, BS, total;
  __host__ __device__ SmemLayout(int degmax, int T, bool motion_only) {
    Rt = 0;
    Hacc = Rt + degmax * 12;
    q = Hacc + degmax * 32;
    if (motion_only) { qw = q; panel = q; BS = 0; total = q; return; }
    qw = q + T;
    panel = qw + T;
    BS = 8 * T + 4;      // +4 floats: consecutive panel blocks start 16 B apart in bank space
    total = panel + (degmax + 1) * BS;
  }
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

template <int PPL, int PPT, bool MOTION_ONLY>
__global__ void __launch_bounds__(NT, (PPT == 1 ? 2 : 1))
k_linearize(const LinParams P) {
  constexpr int T = 32 * PPL;
  extern __shared__ __align__(16) float smem[];
  const SmemLayout L(P.degmax, T, MOTION_ONLY);
  float* sRt = smem + L.Rt;
  float* sHacc = smem + L.Hacc;
  float* sq = smem + L.q;
  float* sqw = smem + L.qw;
  float* spanel = smem + L.panel;
  const int BS = L.BS;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int m = blockIdx.x / P.S, s = blockIdx.x % P.S;
  const int k = P.kx[m];
  const int e0 = P.koff[m];
  const int deg = P.koff[m + 1] - e0;
  const int nblk = deg + 1;
  const int npairs = nblk * (nblk + 1) / 2;
  const int HW = P.HW;
  const float fx = P.intr[0], fy = P.intr[1], cx = P.intr[2], cy = P.intr[3];

  for (int le = tid; le < deg; le += NT) relative_pose_Rt(P.poses, k, P.cj[e0 + le], sRt + 12 * le);
  for (int i = tid; i < deg * 32; i += NT) sHacc[i] = 0.f;

  // Schur accumulators (register resident across tiles when every pair fits)
  const bool resident = MOTION_ONLY || (npairs <= PPT * NT);
  float acc[PPT][36];
  float vacc[PPT][6];
#pragma unroll
  for (int u = 0; u < PPT; ++u) {
#pragma unroll
    for (int i = 0; i < 36; ++i) acc[u][i] = 0.f;
#pragma unroll
    for (int i = 0; i < 6; ++i) vacc[u][i] = 0.f;
  }
  const int64_t slab0 = (int64_t(P.pairbase[m]) * P.S + int64_t(s) * npairs) * 36;
  const int64_t vE0 = (int64_t(P.blkbase[m]) * P.S + int64_t(s) * nblk) * 6;

  const int tile_lo = s * P.tiles_per_split;
  const int tile_hi = min(tile_lo + P.tiles_per_split, (P.ntiles + PPL - 1) / PPL);
  const float* dk = P.disps + int64_t(k) * HW;
  __syncthreads();

  for (int tile = tile_lo; tile < tile_hi; ++tile) {
    const int p0 = tile * T;
    // ---------------- phase 1: one warp per edge ------------------------------------------
    float d[PPL], x0[PPL], y0[PPL];
#pragma unroll
    for (int q = 0; q < PPL; ++q) {
      int p = p0 + lane + 32 * q;
      bool inb = p < HW;
      d[q] = inb ? __ldg(dk + p) : 0.f;
      int py = p / P.wd, px = p - py * P.wd;
      x0[q] = (float(px) - cx) / fx;
      y0[q] = (float(py) - cy) / fy;
    }
    for (int le = warp; le < deg; le += NW) {
      float Rt[12];
#pragma unroll
      for (int i = 0; i < 12; ++i) Rt[i] = sRt[12 * le + i];
      const int e = P.cedge[e0 + le];
      const float* tg = P.targets + int64_t(e) * 2 * HW;
      const float* wg = P.weights + int64_t(e) * 2 * HW;
      float hs[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) hs[i] = 0.f;
      float* pe = spanel + (1 + le) * BS;
#pragma unroll
      for (int q = 0; q < PPL; ++q) {
        const int pl = lane + 32 * q;
        const int p = p0 + pl;
        const bool inb = p < HW;
        float tx = 0.f, ty = 0.f, wx = 0.f, wy = 0.f;
        if (inb) {
          tx = __ldcs(tg + p); ty = __ldcs(tg + HW + p);
          wx = __ldcs(wg + p); wy = __ldcs(wg + HW + p);
        }
        const Proj pr = project(Rt, x0[q], y0[q], d[q], fx, fy, cx, cy);
        const float r0 = tx - pr.u, r1 = ty - pr.v;
        const float w0 = kWeightScale * (pr.valid * wx), w1 = kWeightScale * (pr.valid * wy);
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
      float tot = warp_reduce_scatter32(hs, lane);
      if (lane < 27) sHacc[le * 32 + lane] += tot;
    }
    if (MOTION_ONLY) continue;   // edge `le` always belongs to warp le % NW: no cross-warp hazard
    __syncthreads();
    // ---------------- phase 2: one thread per pixel ---------------------------------------
    if (tid < T) {
      const int p = p0 + tid;
      const bool inb = p < HW;
      const float dd = inb ? __ldg(dk + p) : 0.f;
      const int py = p / P.wd, px = p - py * P.wd;
      const float xx = (float(px) - cx) / fx, yy = (float(py) - cy) / fy;
      float C = 0.f, wsum = 0.f, Gx = 0.f, Gy = 0.f, Gz = 0.f;
      for (int le = 0; le < deg; ++le) {
        float* pe = spanel + (1 + le) * BS;
        const float* R = sRt + 12 * le;
        float hx = pe[tid], hy = pe[T + tid], hz = pe[2 * T + tid];
        C += pe[6 * T + tid];
        wsum += pe[7 * T + tid];
        Gx += R[0] * hx + R[3] * hy + R[6] * hz;    // R^T h
        Gy += R[1] * hx + R[4] * hy + R[7] * hz;
        Gz += R[2] * hx + R[5] * hy + R[8] * hz;
        pe[tid] = dd * hx; pe[T + tid] = dd * hy; pe[2 * T + tid] = dd * hz;
      }
      float Qv = 0.f, wv = 0.f;
      if (inb) {
        float et = __ldg(P.eta + int64_t(m) * HW + p);
        if (P.disps_sens != nullptr) {
          float ds = __ldg(P.disps_sens + int64_t(k) * HW + p);
          float ms = ds > 0.f ? 1.f : 0.f;
          C = C + ms * kSensAlpha + (1.f - ms) * et + kCEps;
          wsum = wsum - ms * kSensAlpha * (dd - ds);
        } else {
          C = C + et + kCEps;
        }
        Qv = 1.0f / C;
        wv = wsum;
        P.Q[int64_t(m) * HW + p] = Qv;
        P.w[int64_t(m) * HW + p] = wv;
      }
      sq[tid] = Qv;
      sqw[tid] = Qv * wv;
      // E_ii = -[d G, X0 x G],  X0 = [xx, yy, 1]
      spanel[0 * T + tid] = -dd * Gx;
      spanel[1 * T + tid] = -dd * Gy;
      spanel[2 * T + tid] = -dd * Gz;
      spanel[3 * T + tid] = -(yy * Gz - Gy);
      spanel[4 * T + tid] = -(Gx - xx * Gz);
      spanel[5 * T + tid] = -(xx * Gy - yy * Gx);
    }
    __syncthreads();
    // ---------------- phase 3: Schur products of the panel --------------------------------
    for (int base = 0; base < npairs; base += PPT * NT) {
      if (!resident) {
#pragma unroll
        for (int u = 0; u < PPT; ++u) {
#pragma unroll
          for (int i = 0; i < 36; ++i) acc[u][i] = 0.f;
#pragma unroll
          for (int i = 0; i < 6; ++i) vacc[u][i] = 0.f;
        }
      }
#pragma unroll
      for (int u = 0; u < PPT; ++u) {
        const int pi = base + tid + u * NT;
        if (pi < npairs) {
          int A, B;
          decode_pair(pi, A, B);
          const float* pa = spanel + A * BS;
          const float* pb = spanel + B * BS;
          for (int p4 = 0; p4 < T; p4 += 4) {
            const float4 q4 = *reinterpret_cast<const float4*>(sq + p4);
            float4 a4[6], b4[6];
#pragma unroll
            for (int r = 0; r < 6; ++r) {
              a4[r] = *reinterpret_cast<const float4*>(pa + r * T + p4);
              b4[r] = *reinterpret_cast<const float4*>(pb + r * T + p4);
            }
            if (B == 0) {
              const float4 w4 = *reinterpret_cast<const float4*>(sqw + p4);
#pragma unroll
              for (int r = 0; r < 6; ++r)
                vacc[u][r] += a4[r].x * w4.x + a4[r].y * w4.y + a4[r].z * w4.z + a4[r].w * w4.w;
            }
#pragma unroll
            for (int r = 0; r < 6; ++r) {
              const float qx = q4.x * a4[r].x, qy = q4.y * a4[r].y, qz = q4.z * a4[r].z, qw_ = q4.w * a4[r].w;
#pragma unroll
              for (int c = 0; c < 6; ++c)
                acc[u][r * 6 + c] += qx * b4[c].x + qy * b4[c].y + qz * b4[c].z + qw_ * b4[c].w;
            }
          }
          if (!resident) {   // CTA-owned slab, read-modify-write in tile order (no atomics)
            float* dst = P.Spart + slab0 + int64_t(pi) * 36;
            const bool first = (tile == tile_lo);
#pragma unroll
            for (int i = 0; i < 36; ++i) dst[i] = first ? acc[u][i] : dst[i] + acc[u][i];
            if (B == 0) {
              float* dv = P.vEpart + vE0 + A * 6;
#pragma unroll
              for (int i = 0; i < 6; ++i) dv[i] = first ? vacc[u][i] : dv[i] + vacc[u][i];
            }
          }
        }
      }
    }
    __syncthreads();   // panel is overwritten by the next tile
  }

  // ---------------- write the CTA's partial sums ---------------------------------------------
  for (int i = tid; i < deg * 32; i += NT) {
    int le = i >> 5, l = i & 31;
    P.Hpart[(int64_t(e0 + le) * P.S + s) * 32 + l] = sHacc[i];
  }
  if (!MOTION_ONLY && resident) {
#pragma unroll
    for (int u = 0; u < PPT; ++u) {
      const int pi = tid + u * NT;
      if (pi < npairs) {
        float4* dst = reinterpret_cast<float4*>(P.Spart + slab0 + int64_t(pi) * 36);
#pragma unroll
        for (int i = 0; i < 9; ++i)
          dst[i] = make_float4(acc[u][4 * i], acc[u][4 * i + 1], acc[u][4 * i + 2], acc[u][4 * i + 3]);
        int A, B;
        decode_pair(pi, A, B);
        if (B == 0) {
          float* dv = P.vEpart + vE0 + A * 6;
#pragma unroll
          for (int i = 0; i < 6; ++i) dv[i] = vacc[u][i];
        }
      }
    }
  }
}

template <int PPL, int PPT, bool MO>
cudaError_t launch_one(const LinParams& P, int grid, size_t smem, cudaStream_t st) {
  auto kern = k_linearize<PPL, PPT, MO>;
  cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
  if (err != cudaSuccess) return err;
  kern<<<grid, NT, smem, st>>>(P);
  g_launch_count++;
  return cudaGetLastError();
}

template <int PPL>
cudaError_t launch_ppl(const LinParams& P, int grid, size_t smem, bool motion_only, int npairs_max,
                       cudaStream_t st) {
  if (motion_only) return launch_one<PPL, 1, true>(P, grid, smem, st);
  if (npairs_max <= NT) return launch_one<PPL, 1, false>(P, grid, smem, st);
  if (npairs_max <= 2 * NT) return launch_one<PPL, 2, false>(P, grid, smem, st);
  return launch_one<PPL, 4, false>(P, grid, smem, st);   // beyond 4*NT pairs: slab read-modify-write
}

}  // namespace

size_t linearize_smem_bytes(int degmax, int ppl, bool motion_only) {
  return size_t(SmemLayout(degmax, 32 * ppl, motion_only).total) * sizeof(float);
}

int linearize_pick_ppl(int degmax, int nblkmax, bool motion_only, int HW, size_t smem_cap) {
  (void)nblkmax;
  if (motion_only) return 4;
  // widest tile whose panel leaves room for two CTAs per SM; else the widest that fits at all
  const int cand[4] = {8, 4, 2, 1};
  for (int c : cand)
    if (32 * c <= ((HW + 31) / 32) * 32 && linearize_smem_bytes(degmax, c, false) * 2 + 2048 <= smem_cap) return c;
  for (int c : cand)
    if (linearize_smem_bytes(degmax, c, false) <= smem_cap) return c;
  return 0;
}

cudaError_t launch_linearize(const DeviceGraph& g, const BAInputs& in, const Workspace& ws, bool motion_only,
                             int tiles_per_split, int ppl, cudaStream_t st) {
  LinParams P;
  P.poses = in.poses; P.disps = in.disps; P.intr = in.intrinsics; P.disps_sens = in.disps_sens;
  P.targets = in.targets; P.weights = in.weights; P.eta = in.eta;
  P.kx = g.kx; P.koff = g.koff; P.cedge = g.cedge; P.cj = g.cj; P.pairbase = g.pairbase; P.blkbase = g.blkbase;
  P.Hpart = ws.Hpart; P.Spart = ws.Spart; P.vEpart = ws.vEpart; P.Q = ws.Q; P.w = ws.w;
  P.HW = g.HW; P.wd = g.wd; P.S = g.S; P.tiles_per_split = tiles_per_split;
  P.ntiles = (g.HW + 31) / 32;   // in units of 32 pixels; the kernel divides by its PPL
  P.degmax = g.degmax;
  const int grid = g.M * g.S;
  if (grid == 0) return cudaSuccess;
  const size_t smem = linearize_smem_bytes(g.degmax, ppl, motion_only);
  switch (ppl) {
    case 1: return launch_ppl<1>(P, grid, smem, motion_only, g.npairs_max, st);
    case 2: return launch_ppl<2>(P, grid, smem, motion_only, g.npairs_max, st);
    case 4: return launch_ppl<4>(P, grid, smem, motion_only, g.npairs_max, st);
    case 8: return launch_ppl<8>(P, grid, smem, motion_only, g.npairs_max, st);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace dba
