// Linearisation with the Schur products on the 5th-generation tensor cores (tcgen05, sm_100a).
//
// Same phases 1 and 2 as dba_linearize.cu (shared code in dba_linearize.cuh).  Phase 3, the Schur
// products of the keyframe's panel  X = sum_p Q_p E[:,p] E[:,p]^T  and  E (Q w), is a symmetric rank-K
// update with K = pixels: the panel rows scaled by sqrt(Q_p), plus one extra row sqrt(Q_p) w_p, form a
// K-major operand P' [N x K] and X = P' P'^T.  Per 64-pixel tile the CTA
//   * converts the fp32 panel to two tf32 operands hi = tf32(v), lo = v - hi, written by its own
//     threads (st.shared) straight into the SWIZZLE_128B layout the tensor core reads;
//   * one thread issues 24 tcgen05.mma (kind::tf32, M = 128, N = 16*ceil((R+1)/16), fp32 accumulate
//     in TMEM): hi hi^T into accumulator 1, hi lo^T + lo hi^T into accumulator 2 ("3xTF32": ~22
//     mantissa bits; the small cross terms get their own accumulator);
//   * the MMAs run while the SIMT pipes do phases 1-2 of the next tile.
// TMEM accumulation is NOT round-to-nearest (measured: a 3328-pixel chain loses 4e-5 relative), so
// every tile starts fresh accumulators and the per-tile results are added into fp32 registers
// (tcgen05.ld + FADD, round-to-nearest): tools/umma_test.cu measures 5e-7, the level of a
// sequential fp32 FMA chain.
// Limits: R = 6 (deg + 1) <= 126, i.e. out-degree <= 20; larger keyframes take the SIMT kernel.
#include "dba_linearize.cuh"
#include "dba_umma.cuh"

namespace dba {

namespace {

using namespace umma;

constexpr int NT = 256;
constexpr int NW = NT / 32;
constexpr int PPL = 2;
constexpr int T = 32 * PPL;            // pixels per tile = K per flush
constexpr int KB = T / 32;             // 32-pixel K-blocks per tile
constexpr int BS = 8 * T + 8;          // floats per staged panel block (6 E rows + ck + wk; +8: phase 2 bank spread)
constexpr uint32_t TILE = TILE_BYTES_128;
constexpr int ACC2 = 128;              // TMEM column of the cross-term accumulator
constexpr int TMEM_COLS = 256;

__host__ __device__ inline size_t tc_float_region(int degmax) {
  // sRt [deg,12] | sHacc [deg,32] | ssq [T] | ssw [T] | staging [(deg+1), BS]
  return size_t(degmax) * 44 + 2 * T + size_t(degmax + 1) * BS;
}

__global__ void __launch_bounds__(NT, 2)
k_linearize_tc(const LinParams P) {
  // dynamic shared memory only (no static __shared__), so the window starts 1024-byte aligned as the
  // SWIZZLE_128B operand tiles require; all pointers stay in the shared address space (LDS / STS)
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* hi = smem_raw;                   // [KB][128 rows][32] swizzled
  uint8_t* lo = smem_raw + KB * TILE;
  float* sRt = reinterpret_cast<float*>(smem_raw + 2 * KB * TILE);
  float* sHacc = sRt + P.degmax * 12;
  float* ssq = sHacc + P.degmax * 32;
  float* ssw = ssq + T;
  float* spanel = ssw + T;
  uint64_t* mbar_p = reinterpret_cast<uint64_t*>(spanel + (P.degmax + 1) * BS + ((P.degmax + 1) * BS & 1));
  uint32_t* tmem_p = reinterpret_cast<uint32_t*>(mbar_p + 1);
#define mbar (*mbar_p)
#define tmem_s (*tmem_p)

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int mi = blockIdx.x / P.S, s = blockIdx.x % P.S;
  const int m = (P.kf_list != nullptr) ? P.kf_list[mi] : mi;
  const int k = P.kx[m];
  const int e0 = P.koff[m];
  const int deg = P.koff[m + 1] - e0;
  const int nblk = deg + 1;
  const int npairs = nblk * (nblk + 1) / 2;
  const int R = 6 * nblk;                   // panel rows; row R carries w
  const int N = (R + 1 + 15) & ~15;
  const int HW = P.HW;
  const float fx = P.intr[0], fy = P.intr[1], cx = P.intr[2], cy = P.intr[3];

  for (int le = tid; le < deg; le += NT) relative_pose_Rt(P.poses, k, P.cj[e0 + le], sRt + 12 * le);
  for (int i = tid; i < deg * 32; i += NT) sHacc[i] = 0.f;
  if ((smem_u32(smem_raw) & 1023u) != 0u) __trap();
  {  // rows the panel never writes must not hold NaN patterns: zero both operand tiles once
    float4* z = reinterpret_cast<float4*>(smem_raw);
    for (int i = tid; i < int(2 * KB * TILE / 16); i += NT) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  if (warp == 0) tmem_alloc(&tmem_s, TMEM_COLS);
  if (tid == 0) mbar_init(&mbar, 1);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = tmem_s;

  // fp32 running sums of the tile results: thread = (row of D, 16-column groups g = 2*gi + h)
  const int qd = warp & 3, h = warp >> 2;
  const int row = 32 * qd + lane;
  const uint32_t tlane = uint32_t(32 * qd) << 16;
  const int ngroups = N >> 4;
  const int maxcol = 6 * ((32 * qd + 31) / 6) + 5;      // last column any row of this warp needs
  const int gw = R >> 4;                                // group of the w column
  float tot[4][16];
#pragma unroll
  for (int gi = 0; gi < 4; ++gi)
#pragma unroll
    for (int j = 0; j < 16; ++j) tot[gi][j] = 0.f;

  auto flush = [&]() {
#pragma unroll
    for (int gi = 0; gi < 4; ++gi) {
      const int g = 2 * gi + h;
      if (g < ngroups && (16 * g <= maxcol || g == gw)) {       // warp-uniform
        uint32_t v1[16], v2[16];
        tmem_ld16(tmem + tlane + 16 * g, v1);
        tmem_ld16(tmem + tlane + ACC2 + 16 * g, v2);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) tot[gi][j] += __uint_as_float(v1[j]) + __uint_as_float(v2[j]);
      }
    }
  };

  const int64_t slab0 = (int64_t(P.pairbase[m]) * P.S + int64_t(s) * npairs) * 36;
  const int64_t vE0 = (int64_t(P.blkbase[m]) * P.S + int64_t(s) * nblk) * 6;
  const int tile_lo = s * P.tiles_per_split;
  const int tile_hi = min(tile_lo + P.tiles_per_split, (P.ntiles + PPL - 1) / PPL);
  const float* dk = P.disps + int64_t(k) * HW;
  uint32_t phase = 0;
  bool pending = false;
  float cur[4 * PPL];                        // streamed inputs of the warp's next (edge, tile)
#pragma unroll
  for (int i = 0; i < 4 * PPL; ++i) cur[i] = 0.f;
  if (warp < deg && tile_lo < tile_hi) {
    const int e = P.cedge[e0 + warp];
    phase1_load<PPL>(P.targets + int64_t(e) * 2 * HW, P.weights + int64_t(e) * 2 * HW, HW, tile_lo * T, lane, cur);
  }

  for (int tile = tile_lo; tile < tile_hi; ++tile) {
    const int p0 = tile * T;
    // ---------------- phase 1: one warp per edge (overlaps the previous tile's MMAs) ------------
    float d[PPL], x0[PPL], y0[PPL];
#pragma unroll
    for (int q = 0; q < PPL; ++q) {
      const int p = p0 + lane + 32 * q;
      const bool inb = p < HW;
      d[q] = inb ? __ldg(dk + p) : 0.f;
      const int py = p / P.wd, px = p - py * P.wd;
      x0[q] = (float(px) - cx) / fx;
      y0[q] = (float(py) - cy) / fy;
    }
    for (int le = warp; le < deg; le += NW) {
      // rolling prefetch: the next edge of this tile, else this warp's first edge of the next tile
      float nxt[4 * PPL];
      {
        int nle = le + NW, np0 = p0;
        if (nle >= deg) { nle = warp; np0 = p0 + T; }
        if (np0 < HW && (tile + 1 < tile_hi || np0 == p0)) {
          const int ne = P.cedge[e0 + nle];
          phase1_load<PPL>(P.targets + int64_t(ne) * 2 * HW, P.weights + int64_t(ne) * 2 * HW, HW, np0, lane, nxt);
        }
      }
      phase1_edge<PPL, false>(cur, HW, p0, lane, sRt + 12 * le, d, x0, y0, fx, fy, cx, cy, spanel + (1 + le) * BS,
                              sHacc + le * 32);
#pragma unroll
      for (int i = 0; i < 4 * PPL; ++i) cur[i] = nxt[i];
    }
    __syncthreads();
    // ---------------- phase 2: four threads per pixel (a quarter of the edges each) ---------------
    {
      const int pl = warp * 8 + (lane >> 2), part = lane & 3;
      const int p = p0 + pl;
      const bool inb = p < HW;
      const float dd = inb ? __ldg(dk + p) : 0.f;
      float C = 0.f, wsum = 0.f, Gx = 0.f, Gy = 0.f, Gz = 0.f;
      for (int le = part; le < deg; le += 4) {
        float* pe = spanel + (1 + le) * BS;
        const float* Rm = sRt + 12 * le;
        const float hx = pe[pl], hy = pe[T + pl], hz = pe[2 * T + pl];
        C += pe[6 * T + pl];
        wsum += pe[7 * T + pl];
        Gx += Rm[0] * hx + Rm[3] * hy + Rm[6] * hz;    // R^T h
        Gy += Rm[1] * hx + Rm[4] * hy + Rm[7] * hz;
        Gz += Rm[2] * hx + Rm[5] * hy + Rm[8] * hz;
        pe[pl] = dd * hx; pe[T + pl] = dd * hy; pe[2 * T + pl] = dd * hz;
      }
#pragma unroll
      for (int mk = 1; mk <= 2; mk <<= 1) {
        C += __shfl_xor_sync(0xffffffffu, C, mk);
        wsum += __shfl_xor_sync(0xffffffffu, wsum, mk);
        Gx += __shfl_xor_sync(0xffffffffu, Gx, mk);
        Gy += __shfl_xor_sync(0xffffffffu, Gy, mk);
        Gz += __shfl_xor_sync(0xffffffffu, Gz, mk);
      }
      if (part == 0) {
        float Qv = 0.f, wv = 0.f;
        if (inb) {
          const float et = __ldg(P.eta + int64_t(m) * HW + p);
          if (P.disps_sens != nullptr) {
            const float ds = __ldg(P.disps_sens + int64_t(k) * HW + p);
            const float ms = ds > 0.f ? 1.f : 0.f;
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
        ssq[pl] = sqrtf(Qv);
        ssw[pl] = wv;
        const int py = p / P.wd, px = p - py * P.wd;
        const float xx = (float(px) - cx) / fx, yy = (float(py) - cy) / fy;
        // E_ii = -[d G, X0 x G],  X0 = [xx, yy, 1]
        spanel[0 * T + pl] = -dd * Gx;
        spanel[1 * T + pl] = -dd * Gy;
        spanel[2 * T + pl] = -dd * Gz;
        spanel[3 * T + pl] = -(yy * Gz - Gy);
        spanel[4 * T + pl] = -(Gx - xx * Gz);
        spanel[5 * T + pl] = -(xx * Gy - yy * Gx);
      }
    }
    __syncthreads();
    // ---------------- phase 3: tensor cores ------------------------------------------------------
    if (pending) {   // previous tile: MMAs done -> accumulators readable, operand tiles reusable
      mbar_wait(&mbar, phase);
      phase ^= 1;
      tc_fence_after_sync();
      flush();
      tc_fence_before_sync();
    }
    for (int r = warp; r <= R; r += NW) {                       // one panel row per warp, lane = pixel
      const int blk = r / 6, rr = r - blk * 6;
      const float* src = (r < R) ? (spanel + blk * BS + rr * T) : ssw;
      const uint32_t off = sw128_offset(r, lane);
#pragma unroll
      for (int kb = 0; kb < KB; ++kb) {
        const int pl = kb * 32 + lane;
        const float v = src[pl] * ssq[pl];
        float vh, vl;
        split_tf32(v, vh, vl);
        *reinterpret_cast<float*>(hi + kb * TILE + off) = vh;
        *reinterpret_cast<float*>(lo + kb * TILE + off) = vl;
      }
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after_sync();
      const uint32_t idesc = make_idesc_tf32(128, N);
      const uint32_t ah = smem_u32(hi), al = smem_u32(lo);
#pragma unroll
      for (int kb = 0; kb < KB; ++kb)
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t dh = make_desc_sw128(ah + kb * TILE + ks * 32);
          const uint64_t dl = make_desc_sw128(al + kb * TILE + ks * 32);
          const bool first = (kb == 0 && ks == 0);
          mma_tf32_ss(tmem, dh, dh, idesc, !first);
          mma_tf32_ss(tmem + ACC2, dh, dl, idesc, !first);
          mma_tf32_ss(tmem + ACC2, dl, dh, idesc, true);
        }
      mma_commit(&mbar);
    }
    pending = true;
  }
  if (pending) {
    mbar_wait(&mbar, phase);
    tc_fence_after_sync();
    flush();
    tc_fence_before_sync();
  }

  // ---------------- write the CTA's partial sums -------------------------------------------------
  for (int i = tid; i < deg * 32; i += NT) {
    const int le = i >> 5, l = i & 31;
    P.Hpart[(int64_t(e0 + le) * P.S + s) * 32 + l] = sHacc[i];
  }
  if (row < R) {
    const int A = row / 6, ri = row - A * 6;
    float* Sdst = P.Spart + slab0 + int64_t(A) * (A + 1) / 2 * 36 + ri * 6;
#pragma unroll
    for (int gi = 0; gi < 4; ++gi) {
      const int g = 2 * gi + h;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int c = 16 * g + j;
        if (c < R) {
          const int B = c / 6, cj = c - B * 6;
          if (B <= A) Sdst[B * 36 + cj] = tot[gi][j];
        } else if (c == R) {
          P.vEpart[vE0 + A * 6 + ri] = tot[gi][j];
        }
      }
    }
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, TMEM_COLS);
#undef mbar
#undef tmem_s
}

}  // namespace

size_t linearize_tc_smem_bytes(int degmax) { return 2 * KB * TILE + tc_float_region(degmax) * sizeof(float) + 32; }
int linearize_tc_max_degree() { return 20; }
int linearize_tc_ppl() { return PPL; }

cudaError_t launch_linearize_tc(const DeviceGraph& g, const BAInputs& in, const Workspace& ws, int tiles_per_split,
                                const int32_t* kf_list, int nkf, int degmax, cudaStream_t st) {
  if (nkf == 0) return cudaSuccess;
  LinParams P;
  P.poses = in.poses; P.disps = in.disps; P.intr = in.intrinsics; P.disps_sens = in.disps_sens;
  P.targets = in.targets; P.weights = in.weights; P.eta = in.eta;
  P.kx = g.kx; P.koff = g.koff; P.cedge = g.cedge; P.cj = g.cj; P.pairbase = g.pairbase; P.blkbase = g.blkbase;
  P.kf_list = kf_list;
  P.Hpart = ws.Hpart; P.Spart = ws.Spart; P.vEpart = ws.vEpart; P.Q = ws.Q; P.w = ws.w;
  P.HW = g.HW; P.wd = g.wd; P.S = g.S; P.tiles_per_split = tiles_per_split;
  P.ntiles = (g.HW + 31) / 32;
  P.degmax = degmax;
  const size_t smem = linearize_tc_smem_bytes(degmax);
  cudaError_t err = cudaFuncSetAttribute(k_linearize_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
  if (err != cudaSuccess) return err;
  k_linearize_tc<<<nkf * g.S, NT, smem, st>>>(P);
  g_launch_count++;
  return cudaGetLastError();
}

}  // namespace dba
