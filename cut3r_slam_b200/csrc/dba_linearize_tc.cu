// Linearisation with the Schur products on the 5th-generation tensor cores (tcgen05, sm_100a).
//
// Same phases 1 and 2 as dba_linearize.cu (shared code in dba_linearize.cuh).  Phase 3, the Schur
// products of the keyframe's panel  X = sum_p Q_p E[:,p] E[:,p]^T  and  E (Q w), is a symmetric rank-K
// update with K = pixels: the panel rows scaled by sqrt(Q_p), plus one extra row sqrt(Q_p) w_p, form a
// K-major operand P' [N x K] and X = P' P'^T.  Per 64-pixel tile the CTA
//   * converts the fp32 panel to two tf32 operands hi = tf32(v), lo = v - hi, written by its own
//     threads (st.shared) straight into the SWIZZLE_128B layout the tensor core reads;
//   * one thread issues 24 tcgen05.mma (kind::tf32, M = 128, N = 16*ceil((R+1)/16), fp32 accumulate
//     in TMEM) into one accumulator: the sixteen cross terms hi lo^T + lo hi^T first, then the eight
//     hi hi^T steps ("3xTF32": ~22 mantissa bits; the truncating TMEM adds of the small terms act on a
//     small accumulator, the main chain is as long as with a separate accumulator);
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
constexpr int BS = 6 * T;              // floats per staged panel block (6 E rows)
constexpr int PS = 5 * T + 8;          // floats per warp of the per-pixel partial sums (ck, wk, G): +8 bank spread
constexpr int TMEM_COLS = 128;         // one fp32 accumulator, N <= 128 columns
// WIDE variant (out-degree 21..25, R + 1 <= 160 panel rows): operand tiles of 256 rows, a second product
// D2 = P'[128..255] P'[0..N2)^T (N2 <= 160) in TMEM columns 128.., whose 32 meaningful lanes belong to the
// warps of lane quarter 0; one CTA per SM (80 more accumulator registers in those warps).
constexpr int WIDE_MAXDEG = 25;
// WIDE = 2 (out-degree 26..41, R + 1 <= 253): the second product has all 128 rows and up to 256 columns, every
// warp holds up to 8 more column groups of its lane quarter (what does not fit 255 registers lives in local memory).
constexpr int WIDE2_MAXDEG = 41;
constexpr int TMEM_COLS_WIDE = 512;
constexpr int D2COL = 128;

// float region after the two operand tiles:
//   sRt [deg,12] | sHacc [deg,32] | ssq [T] | ssw [T] | spx [2][3][T] | spart [NW][PS] | staging [(deg+1), BS] |
//   sEoff [deg] (int64: element offset of the edge's targets / weights planes)
__host__ __device__ inline size_t tc_float_region(int degmax) {
  return size_t(degmax) * 44 + 2 * T + 6 * T + size_t(NW) * PS + size_t(degmax + 1) * BS + 2 * size_t(degmax);
}

// Phase 1 of one (edge, tile): lanes own pixels lane and lane+32 of the tile.  Adds the 21+6 sums of
// Hjj / vj onto hacc (shared, owned by this warp), writes the 6 panel rows of the edge
// [d h, X1 x h] to pe[row*T + pixel] and accumulates the per-pixel sums over the warp's edges
// acc[q] = {ck, wk, G = R^T h} in registers (hislam2/geom/ba.py:39-68 in the reduced form of
// tests/test_kernel_algebra.py).
__device__ __forceinline__ void p1_edge(const float (&v)[4 * PPL], bool inb0, bool inb1, int lane,
                                        const float* sRt_e, const float (&d)[PPL], const float (&x0)[PPL],
                                        const float (&y0)[PPL], float fx, float fy, float cx, float cy, float* pe,
                                        float* hacc, float (&acc)[PPL][5]) {
  float Rt[12];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const float4 r4 = *reinterpret_cast<const float4*>(sRt_e + 4 * i);
    Rt[4 * i] = r4.x; Rt[4 * i + 1] = r4.y; Rt[4 * i + 2] = r4.z; Rt[4 * i + 3] = r4.w;
  }
  float hs[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) hs[i] = 0.f;
#pragma unroll
  for (int q = 0; q < PPL; ++q) {
    const int pl = lane + 32 * q;
    const bool inb = q == 0 ? inb0 : inb1;
    const Proj pr = project(Rt, x0[q], y0[q], d[q], fx, fy, cx, cy);
    const float r0 = v[4 * q + 0] - pr.u, r1 = v[4 * q + 1] - pr.v;
    // out-of-range pixels carry zero weights (the loads return 0), so every sum below is exact
    const float w0 = kWeightScale * (pr.valid * v[4 * q + 2]), w1 = kWeightScale * (pr.valid * v[4 * q + 3]);
    const float X = pr.X1.x, Y = pr.X1.y, Z = pr.X1.z, dd = d[q];
    const float J0[6] = {dd * pr.a0, 0.f, dd * pr.c0, Y * pr.c0, Z * pr.a0 - X * pr.c0, -Y * pr.a0};
    const float J1[6] = {0.f, dd * pr.b1, dd * pr.c1, Y * pr.c1 - Z * pr.b1, -X * pr.c1, X * pr.b1};
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
    const float a0 = w0 * pr.jz0, a1 = w1 * pr.jz1;
    float hx = a0 * pr.a0, hy = a1 * pr.b1, hz = a0 * pr.c0 + a1 * pr.c1;
    float ck = a0 * pr.jz0 + a1 * pr.jz1;
    float wk = a0 * r0 + a1 * r1;
    if (!inb) { hx = hy = hz = ck = wk = 0.f; }
    pe[0 * T + pl] = dd * hx;
    pe[1 * T + pl] = dd * hy;
    pe[2 * T + pl] = dd * hz;
    pe[3 * T + pl] = Y * hz - Z * hy;     // X1 x h
    pe[4 * T + pl] = Z * hx - X * hz;
    pe[5 * T + pl] = X * hy - Y * hx;
    acc[q][0] += ck;
    acc[q][1] += wk;
    acc[q][2] += Rt[0] * hx + Rt[3] * hy + Rt[6] * hz;     // R^T h
    acc[q][3] += Rt[1] * hx + Rt[4] * hy + Rt[7] * hz;
    acc[q][4] += Rt[2] * hx + Rt[5] * hy + Rt[8] * hz;
  }
  const float tot = warp_reduce_scatter32(hs, lane);
  if (lane < 27) hacc[lane] += tot;
}

// the tile's streamed inputs of one edge: v = {tx, ty, wx, wy} x PPL.  STREAM: evict-first loads (the
// factor graph does not fit the L2); otherwise normal loads, so that the 2nd..nth LM iteration of a
// call finds targets / weights in the 126 MB L2.
template <bool STREAM>
__device__ __forceinline__ float ld_in(const float* p) { return STREAM ? __ldcs(p) : __ldg(p); }
template <bool STREAM>
__device__ __forceinline__ void tc_load(const float* __restrict__ tg, const float* __restrict__ wg, int HW, int p0,
                                        int lane, float (&v)[4 * PPL]) {
#pragma unroll
  for (int q = 0; q < PPL; ++q) {
    const int p = p0 + lane + 32 * q;
    const bool inb = p < HW;
    v[4 * q + 0] = inb ? ld_in<STREAM>(tg + p) : 0.f;
    v[4 * q + 1] = inb ? ld_in<STREAM>(tg + HW + p) : 0.f;
    v[4 * q + 2] = inb ? ld_in<STREAM>(wg + p) : 0.f;
    v[4 * q + 3] = inb ? ld_in<STREAM>(wg + HW + p) : 0.f;
  }
}

template <bool STREAM, int WIDE>
__global__ void __launch_bounds__(NT, WIDE ? 1 : 2)
k_linearize_tc(const LinParams P) {
  constexpr uint32_t TILE = WIDE ? TILE_BYTES_256 : TILE_BYTES_128;
  // dynamic shared memory only (no static __shared__), so the window starts 1024-byte aligned as the
  // SWIZZLE_128B operand tiles require; all pointers stay in the shared address space (LDS / STS)
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* hi = smem_raw;                   // [KB][128 rows][32] swizzled
  uint8_t* lo = smem_raw + KB * TILE;
  float* sRt = reinterpret_cast<float*>(smem_raw + 2 * KB * TILE);
  float* sHacc = sRt + P.degmax * 12;
  float* ssq = sHacc + P.degmax * 32;
  float* ssw = ssq + T;
  float* spx = ssw + T;                     // [2][3][T]: x0, y0, d of the tile's pixels (double-buffered)
  float* spart = spx + 6 * T;               // [NW][PS]: per-warp partial sums of ck, wk, G
  float* spanel = spart + NW * PS;          // [(deg+1)][6][T]
  int64_t* sEoff = reinterpret_cast<int64_t*>(spanel + (P.degmax + 1) * BS);
  uint64_t* mbar_p = reinterpret_cast<uint64_t*>(sEoff + P.degmax);
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
  const int N2 = (R + 1 + 15) & ~15;         // columns that hold data (incl. the w row)
  const int N = WIDE ? 128 : N2;             // columns of the first product
  const int HW = P.HW;
  const float fx = P.intr[0], fy = P.intr[1], cx = P.intr[2], cy = P.intr[3];
  const int tile_lo = s * P.tiles_per_split;
  const int tile_hi = min(tile_lo + P.tiles_per_split, (P.ntiles + PPL - 1) / PPL);
  const float* dk = P.disps + int64_t(k) * HW;

  // x0, y0, d of one tile's pixels (pinhole.py:6-25), computed once per tile by 64 threads
  auto fill_px = [&](int tile, int pl, float dval) {
    float* b = spx + (tile & 1) * 3 * T;
    const int p = tile * T + pl;
    const int py = p / P.wd, px = p - py * P.wd;
    b[pl] = (float(px) - cx) / fx;
    b[T + pl] = (float(py) - cy) / fy;
    b[2 * T + pl] = dval;
  };

  for (int le = tid; le < deg; le += NT) relative_pose_Rt(P.poses, k, P.cj[e0 + le], sRt + 12 * le);
  for (int i = tid; i < deg * 32; i += NT) sHacc[i] = 0.f;
  for (int le = tid; le < deg; le += NT) sEoff[le] = int64_t(P.cedge[e0 + le]) * 2 * HW;
  if (tid < T && tile_lo < tile_hi) fill_px(tile_lo, tid, (tile_lo * T + tid < HW) ? __ldg(dk + tile_lo * T + tid) : 0.f);
  if ((smem_u32(smem_raw) & 1023u) != 0u) __trap();
  {  // rows the panel never writes must not hold NaN patterns: zero both operand tiles once
    float4* z = reinterpret_cast<float4*>(smem_raw);
    for (int i = tid; i < int(2 * KB * TILE / 16); i += NT) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  if (warp == 0) tmem_alloc(&tmem_s, WIDE ? TMEM_COLS_WIDE : TMEM_COLS);
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
  // WIDE: rows 128 + lane of the second product, column groups g = 2*gi + h (warps 0 and 4 only)
  constexpr int NG2 = WIDE == 2 ? 8 : (WIDE == 1 ? 5 : 1);
  const bool has2 = WIDE == 2 || (WIDE == 1 && qd == 0);       // warp-uniform: this warp reads the second product
  const int maxcol2 = 6 * ((128 + 32 * qd + 31) / 6) + 5;
  float tot2[NG2][16];
#pragma unroll
  for (int gi = 0; gi < NG2; ++gi)
#pragma unroll
    for (int j = 0; j < 16; ++j) tot2[gi][j] = 0.f;

  auto flush = [&]() {
#pragma unroll
    for (int gi = 0; gi < 4; ++gi) {
      const int g = 2 * gi + h;
      if (g < ngroups && (16 * g <= maxcol || g == gw)) {       // warp-uniform
        uint32_t v1[16];
        tmem_ld16(tmem + tlane + 16 * g, v1);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) tot[gi][j] += __uint_as_float(v1[j]);
      }
    }
    if (WIDE && has2) {
#pragma unroll
      for (int gi = 0; gi < NG2; ++gi) {
        const int g = 2 * gi + h;
        if (16 * g < N2 && 16 * g <= maxcol2) {                   // warp-uniform
          uint32_t v1[16];
          tmem_ld16(tmem + tlane + D2COL + 16 * g, v1);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) tot2[gi][j] += __uint_as_float(v1[j]);
        }
      }
    }
  };

  const int64_t slab0 = (int64_t(P.pairbase[m]) * P.S + int64_t(s) * npairs) * 36;
  const int64_t vE0 = (int64_t(P.blkbase[m]) * P.S + int64_t(s) * nblk) * 6;
  uint32_t phase = 0;
  bool pending = false;
  float cur[4 * PPL];                        // streamed inputs of the warp's next (edge, tile)
#pragma unroll
  for (int i = 0; i < 4 * PPL; ++i) cur[i] = 0.f;
  if (warp < deg && tile_lo < tile_hi) {
    const int64_t eo = sEoff[warp];
    tc_load<STREAM>(P.targets + eo, P.weights + eo, HW, tile_lo * T, lane, cur);
  }
  const int nwp = min(deg, NW);              // warps that own at least one edge
  // per-pixel inputs of phase 2 are fetched one tile ahead (thread = (pixel tid>>2, part tid&3):
  // part 0 carries eta, part 1 the disparity of the next tile's pixel)
  const float* etak = P.eta + int64_t(m) * HW;
  float eta_c = 0.f;
  if ((tid & 3) == 0 && tile_lo < tile_hi && tile_lo * T + (tid >> 2) < HW) eta_c = __ldg(etak + tile_lo * T + (tid >> 2));

  for (int tile = tile_lo; tile < tile_hi; ++tile) {
    const int p0 = tile * T;
    const float* px = spx + (tile & 1) * 3 * T;
    float pre_n = 0.f;                       // part 0: eta, part 1: d of pixel (tile+1, tid>>2)
    if (tile + 1 < tile_hi && (tid & 3) < 2) {
      const int pn = p0 + T + (tid >> 2);
      if (pn < HW) pre_n = __ldg(((tid & 3) == 0 ? etak : dk) + pn);
    }
    // ---------------- phase 1: one warp per edge (overlaps the previous tile's MMAs) ------------
    {
      float d[PPL], x0[PPL], y0[PPL], acc[PPL][5];
#pragma unroll
      for (int q = 0; q < PPL; ++q) {
        x0[q] = px[lane + 32 * q];
        y0[q] = px[T + lane + 32 * q];
        d[q] = px[2 * T + lane + 32 * q];
#pragma unroll
        for (int i = 0; i < 5; ++i) acc[q][i] = 0.f;
      }
      const bool inb0 = p0 + lane < HW, inb1 = p0 + lane + 32 < HW;
      for (int le = warp; le < deg; le += NW) {
        // rolling prefetch: the next edge of this tile, else this warp's first edge of the next tile
        float nxt[4 * PPL];
        {
          int nle = le + NW, np0 = p0;
          if (nle >= deg) { nle = warp; np0 = p0 + T; }
          if (np0 < HW && (tile + 1 < tile_hi || np0 == p0)) {
            const int64_t eo = sEoff[nle];
            tc_load<STREAM>(P.targets + eo, P.weights + eo, HW, np0, lane, nxt);
          }
        }
        p1_edge(cur, inb0, inb1, lane, sRt + 12 * le, d, x0, y0, fx, fy, cx, cy, spanel + (1 + le) * BS,
                sHacc + le * 32, acc);
#pragma unroll
        for (int i = 0; i < 4 * PPL; ++i) cur[i] = nxt[i];
      }
      if (warp < nwp) {
        float* sp = spart + warp * PS;
#pragma unroll
        for (int q = 0; q < PPL; ++q)
#pragma unroll
          for (int i = 0; i < 5; ++i) sp[i * T + lane + 32 * q] = acc[q][i];
      }
    }
    __syncthreads();
    // ---------------- phase 2: four threads per pixel sum the warps' partials in a fixed order -----
    {
      const int pl = tid >> 2, part = tid & 3;
      float sum[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
      for (int w = part; w < nwp; w += 4) {
        const float* sp = spart + w * PS + pl;
#pragma unroll
        for (int i = 0; i < 5; ++i) sum[i] += sp[i * T];
      }
#pragma unroll
      for (int mk = 1; mk <= 2; mk <<= 1)
#pragma unroll
        for (int i = 0; i < 5; ++i) sum[i] += __shfl_xor_sync(0xffffffffu, sum[i], mk);
      if (part == 1 && tile + 1 < tile_hi) fill_px(tile + 1, pl, pre_n);
      if (part == 0) {
        const int p = p0 + pl;
        const bool inb = p < HW;
        const float dd = px[2 * T + pl], xx = px[pl], yy = px[T + pl];
        float C = sum[0], wsum = sum[1];
        const float Gx = sum[2], Gy = sum[3], Gz = sum[4];
        float Qv = 0.f, wv = 0.f;
        if (inb) {
          const float et = eta_c;
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
        // E_ii = -[d G, X0 x G],  X0 = [xx, yy, 1]
        spanel[0 * T + pl] = -dd * Gx;
        spanel[1 * T + pl] = -dd * Gy;
        spanel[2 * T + pl] = -dd * Gz;
        spanel[3 * T + pl] = -(yy * Gz - Gy);
        spanel[4 * T + pl] = -(Gx - xx * Gz);
        spanel[5 * T + pl] = -(xx * Gy - yy * Gx);
        eta_c = pre_n;
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
    {
      // operand conversion: half a warp per panel row, lane = 4 consecutive pixels (one 16-byte chunk of
      // the swizzled K-major tile): v * sqrt(Q) -> hi = tf32(v), lo = v - hi
      const int c16 = lane & 15;                           // chunk of the 64-pixel row
      const int kb = c16 >> 3, ch = c16 & 7;
      const float4 sq4 = *reinterpret_cast<const float4*>(ssq + 4 * c16);
      for (int r = 2 * warp + (lane >> 4); r <= R; r += 2 * NW) {
        const int blk = r / 6, rr = r - blk * 6;
        const float* src = (r < R) ? (spanel + blk * BS + rr * T) : ssw;
        const float4 v4 = *reinterpret_cast<const float4*>(src + 4 * c16);
        float4 vh, vl;
        split_tf32(v4.x * sq4.x, vh.x, vl.x);
        split_tf32(v4.y * sq4.y, vh.y, vl.y);
        split_tf32(v4.z * sq4.z, vh.z, vl.z);
        split_tf32(v4.w * sq4.w, vh.w, vl.w);
        const uint32_t off = uint32_t(kb) * TILE + uint32_t(r >> 3) * 1024u + uint32_t(r & 7) * 128u +
                             ((uint32_t(ch) ^ uint32_t(r & 7)) << 4);
        *reinterpret_cast<float4*>(hi + off) = vh;
        *reinterpret_cast<float4*>(lo + off) = vl;
      }
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after_sync();
      const uint32_t idesc = make_idesc_tf32(128, N);
      const uint32_t ah = smem_u32(hi), al = smem_u32(lo);
      // the small cross terms first, so that the truncating TMEM adds act on a small accumulator;
      // the eight hi hi^T steps follow (the same chain length as a separate main accumulator)
#pragma unroll
      for (int kb = 0; kb < KB; ++kb)
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t dh = make_desc_sw128(ah + kb * TILE + ks * 32);
          const uint64_t dl = make_desc_sw128(al + kb * TILE + ks * 32);
          mma_tf32_ss(tmem, dh, dl, idesc, !(kb == 0 && ks == 0));
          mma_tf32_ss(tmem, dl, dh, idesc, true);
        }
#pragma unroll
      for (int kb = 0; kb < KB; ++kb)
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t dh = make_desc_sw128(ah + kb * TILE + ks * 32);
          mma_tf32_ss(tmem, dh, dh, idesc, true);
        }
      if (WIDE) {   // rows 128.. against all rows: A = the tile's second 128 rows (16 KB further), B = rows [0, N2)
        const uint32_t idesc2 = make_idesc_tf32(128, N2);
#pragma unroll
        for (int kb = 0; kb < KB; ++kb)
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            const uint64_t dh = make_desc_sw128(ah + kb * TILE + ks * 32);
            const uint64_t dl = make_desc_sw128(al + kb * TILE + ks * 32);
            const uint64_t dh2 = make_desc_sw128(ah + kb * TILE + TILE_BYTES_128 + ks * 32);
            const uint64_t dl2 = make_desc_sw128(al + kb * TILE + TILE_BYTES_128 + ks * 32);
            mma_tf32_ss(tmem + D2COL, dh2, dl, idesc2, !(kb == 0 && ks == 0));
            mma_tf32_ss(tmem + D2COL, dl2, dh, idesc2, true);
          }
#pragma unroll
        for (int kb = 0; kb < KB; ++kb)
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            const uint64_t dh = make_desc_sw128(ah + kb * TILE + ks * 32);
            const uint64_t dh2 = make_desc_sw128(ah + kb * TILE + TILE_BYTES_128 + ks * 32);
            mma_tf32_ss(tmem + D2COL, dh2, dh, idesc2, true);
          }
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
  if (WIDE && has2) {
    const int row2 = 128 + 32 * qd + lane;
    const int A = row2 / 6, ri = row2 - A * 6;
    float* Sdst = P.Spart + slab0 + int64_t(A) * (A + 1) / 2 * 36 + ri * 6;
#pragma unroll
    for (int gi = 0; gi < NG2; ++gi) {
      const int g = 2 * gi + h;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int c = 16 * g + j;
        if (c < R) {
          if (row2 < R) {
            const int B = c / 6, cj = c - B * 6;
            if (B <= A) Sdst[B * 36 + cj] = tot2[gi][j];
          } else if (row2 == R) {
            P.vEpart[vE0 + c] = tot2[gi][j];       // row R = sum_p Q w E[c,p] (the product is symmetric)
          }
        }
      }
    }
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, WIDE ? TMEM_COLS_WIDE : TMEM_COLS);
#undef mbar
#undef tmem_s
}

}  // namespace

int g_tc_stream_mode = -1;   // -1 auto, 0 L2-resident loads, 1 evict-first loads ("stream_inputs" option)

size_t linearize_tc_smem_bytes(int degmax, int wide) {
  return 2 * KB * (wide ? TILE_BYTES_256 : TILE_BYTES_128) + tc_float_region(degmax) * sizeof(float) + 32;
}
int linearize_tc_max_degree() { return 20; }
int linearize_tc_wide_max_degree(int level) { return level == 2 ? WIDE2_MAXDEG : WIDE_MAXDEG; }
int linearize_tc_ppl() { return PPL; }

cudaError_t launch_linearize_tc(const DeviceGraph& g, const BAInputs& in, const Workspace& ws, int tiles_per_split,
                                const int32_t* kf_list, int nkf, int degmax, int wide, cudaStream_t st) {
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
  const size_t smem = linearize_tc_smem_bytes(degmax, wide);
  // inputs that fit the L2 with room to spare stay resident across the LM iterations of a call
  const bool stream = g_tc_stream_mode < 0 ? size_t(g.E) * g.HW * 16 > size_t(100) << 20 : g_tc_stream_mode != 0;
  auto kern = wide == 2 ? (stream ? k_linearize_tc<true, 2> : k_linearize_tc<false, 2>)
            : wide == 1 ? (stream ? k_linearize_tc<true, 1> : k_linearize_tc<false, 1>)
                        : (stream ? k_linearize_tc<true, 0> : k_linearize_tc<false, 0>);
  cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
  if (err != cudaSuccess) return err;
  kern<<<nkf * g.S, NT, smem, st>>>(P);
  g_launch_count++;
  return cudaGetLastError();
}

}  // namespace dba
