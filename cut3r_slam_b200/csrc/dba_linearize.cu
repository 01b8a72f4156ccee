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
#include "dba_linearize.cuh"

namespace dba {

namespace {

constexpr int NT = 256;   // threads per CTA
constexpr int NW = NT / 32;

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
  const int mi = blockIdx.x / P.S, s = blockIdx.x % P.S;
  const int m = (P.kf_list != nullptr) ? P.kf_list[mi] : mi;
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
      const int e = P.cedge[e0 + le];
      float v[4 * PPL];
      phase1_load<PPL>(P.targets + int64_t(e) * 2 * HW, P.weights + int64_t(e) * 2 * HW, HW, p0, lane, v);
      phase1_edge<PPL, MOTION_ONLY>(v, HW, p0, lane, sRt + 12 * le, d, x0, y0, fx, fy, cx, cy,
                                    spanel + (1 + le) * BS, sHacc + le * 32);
    }
    if (MOTION_ONLY) continue;   // edge `le` always belongs to warp le % NW: no cross-warp hazard
    __syncthreads();
    // ---------------- phase 2: one thread per pixel ---------------------------------------
    if (tid < T) {
      float Qv, wv;
      phase2_pixel<T>(P, m, k, deg, p0, tid, dk, sRt, spanel, BS, fx, fy, cx, cy, Qv, wv);
      sq[tid] = Qv;
      sqw[tid] = Qv * wv;
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
                             int tiles_per_split, int ppl, const int32_t* kf_list, int nkf, cudaStream_t st) {
  LinParams P;
  P.poses = in.poses; P.disps = in.disps; P.intr = in.intrinsics; P.disps_sens = in.disps_sens;
  P.targets = in.targets; P.weights = in.weights; P.eta = in.eta;
  P.kx = g.kx; P.koff = g.koff; P.cedge = g.cedge; P.cj = g.cj; P.pairbase = g.pairbase; P.blkbase = g.blkbase;
  P.Hpart = ws.Hpart; P.Spart = ws.Spart; P.vEpart = ws.vEpart; P.Q = ws.Q; P.w = ws.w;
  P.kf_list = kf_list;
  P.HW = g.HW; P.wd = g.wd; P.S = g.S; P.tiles_per_split = tiles_per_split;
  P.ntiles = (g.HW + 31) / 32;   // in units of 32 pixels; the kernel divides by its PPL
  P.degmax = g.degmax;
  const int grid = (kf_list != nullptr ? nkf : g.M) * g.S;
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
