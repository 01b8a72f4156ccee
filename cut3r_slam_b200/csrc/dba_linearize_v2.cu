// Linearisation, second generation (sm_100a): ONE WARP PER EDGE, Schur products on tcgen05.
//
// Same mathematics and the same outputs as dba_linearize_tc.cu (hislam2/geom/ba.py:39-92 + chol.py:53-64 in the
// reduced form of tests/test_kernel_algebra.py); what changed is where the instructions go.  The first-generation
// kernel spent 546 thread-instructions per edge-pixel (ncu, profiles/r1c_*): 14 % in the 31-shuffle reduce-scatter
// of the 27 Hjj / vj sums that every warp ran per (edge, 64-pixel tile), 12 % converting the panel to tf32 operand
// pairs with a division per row, 19 % of its stall samples at a barrier because 8 warps shared `deg` edges unevenly.
// Here a CTA = (source keyframe, pixel split) has one warp per outgoing edge (NW = 16 or 20 warps, one CTA per SM):
//   * the 27 sums of Hjj / vj stay in the warp's REGISTERS across all tiles of the CTA; one reduce-scatter per CTA;
//   * every warp carries the same load (one edge), so the three barriers per tile close quickly;
//   * the panel is staged as plain rows [6 (deg + 1) + 1][64] (row r of the operand = row r of the staging area),
//     each half-warp converts one row per pass with a constant address stride: no division, no block decode;
//   * the per-tile tensor-core results are added (round-to-nearest fp32, as before: TMEM accumulation truncates)
//     into a shared-memory accumulator [128][132] with 128-bit LDS / FADD x4 / STS instead of 64 registers per
//     thread -- that is what makes room for the Hjj registers -- and two TMEM accumulators alternate, so the
//     flush of tile t-1 runs while the MMAs of tile t execute.
// Limits: out-degree <= 20 (6 (deg + 1) + 1 <= 127 operand rows).  Wider keyframes: dba_linearize_tc.cu (WIDE).
#include "dba_linearize.cuh"
#include "dba_umma.cuh"

// -DDBA_V2_TRACE: thread 0 (an owner, reducer and tail thread) and thread 64 (a flush warp) of CTA 0 print the clock
// cycles they spent in every phase of the tile loop (developer experiment, profiles/r2_linearize_v2.md)
#ifdef DBA_V2_TRACE
#include <cstdio>
#define TR(i) do { if (blockIdx.x == 0 && (tid == 0 || tid == 64)) { const long long t_ = clock64(); tr[i] += t_ - tlast; tlast = t_; } } while (0)
#else
#define TR(i) do { } while (0)
#endif

// -DDBA_V2_ABLATE=bits (developer timing experiment through tools/build_v2_variant.sh, results are WRONG): 1 no MMAs,
// 2 only the hi x hi product, 4 no flush, 8 no operand conversion, 16 no reduce / tail, 32 no phase-1 arithmetic.
// Reading the deltas: switching a consumer off also removes the producer code that feeds it (dead-code elimination),
// so the deltas over-attribute; profiles/r2c_linearize_experiments.md.
#ifndef DBA_V2_ABLATE
#define DBA_V2_ABLATE 0
#endif


namespace dba {

namespace {

using namespace umma;

constexpr int T = 64;                  // pixels per tile = K per MMA batch
constexpr int KB = T / 32;             // 32-pixel K-blocks
constexpr int PS = 5 * T + 8;          // floats per warp of the per-pixel partial sums (ck, wk, G)
constexpr int ACS = 132;               // row stride of the fp32 accumulator (132 mod 32 = 4: conflict-free LDS.128)
constexpr int ACC_COLS = 128;          // TMEM columns per accumulator

template <int NW>
__host__ __device__ constexpr size_t v2_smem_bytes() {
  return size_t(2 * KB) * TILE_BYTES_128 +
         sizeof(float) * (size_t(128) * ACS + NW * 12 + T + 5 * T + 6 * T + size_t(NW) * PS + (size_t(6) * (NW + 1) + 1) * T) +
         2 * sizeof(uint64_t) + 16;
}

template <bool STREAM>
__device__ __forceinline__ float ld_in(const float* p) { return STREAM ? __ldcs(p) : __ldg(p); }

template <bool STREAM>
__device__ __forceinline__ void load_inputs(const float* __restrict__ tg, const float* __restrict__ wg, int HW, int p0,
                                            int lane, float (&v)[8]) {
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const int p = p0 + lane + 32 * q;
    const bool inb = p < HW;
    v[4 * q + 0] = inb ? ld_in<STREAM>(tg + p) : 0.f;
    v[4 * q + 1] = inb ? ld_in<STREAM>(tg + HW + p) : 0.f;
    v[4 * q + 2] = inb ? ld_in<STREAM>(wg + p) : 0.f;
    v[4 * q + 3] = inb ? ld_in<STREAM>(wg + HW + p) : 0.f;
  }
}

template <int NW, bool STREAM>
__global__ void __launch_bounds__(32 * NW, 1)
k_linearize_v2(const LinParams P) {
  constexpr int NT = 32 * NW;
  constexpr uint32_t TILE = TILE_BYTES_128;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* hi = smem_raw;                                   // [KB][128 rows][32 fp32] swizzled
  uint8_t* lo = smem_raw + KB * TILE;
  float* sAcc = reinterpret_cast<float*>(smem_raw + 2 * KB * TILE);   // [128][ACS] running sums of the tile products
  float* sRt = sAcc + 128 * ACS;                            // [NW][12]
  float* ssq = sRt + NW * 12;                               // [T] sqrt(Q)
  float* ssum = ssq + T;                                    // [5][T]: ck, wk, G summed over the keyframe's edges
  float* spx = ssum + 5 * T;                                // [2][3][T]: x0, y0, d of the tile's pixels (double-buffered)
  float* spart = spx + 6 * T;                               // [NW][PS]: the edge's ck, wk, G per pixel
  float* spanel = spart + NW * PS;                          // [6 (deg+1) + 1][T]: rows 0-5 E_ii, 6 (1+e).. edge e, last w
  uint64_t* mbar = reinterpret_cast<uint64_t*>(spanel + (6 * (NW + 1) + 1) * T);   // [2]
  uint32_t* tmem_p = reinterpret_cast<uint32_t*>(mbar + 2);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int mi = blockIdx.x / P.S, s = blockIdx.x % P.S;
  const int m = (P.kf_list != nullptr) ? P.kf_list[mi] : mi;
  const int k = P.kx[m];
  const int e0 = P.koff[m];
  const int deg = P.koff[m + 1] - e0;
  const int nblk = deg + 1;
  const int npairs = nblk * (nblk + 1) / 2;
  const int R = 6 * nblk;                                   // panel rows; row R carries w
  const int N = (R + 1 + 15) & ~15;                         // MMA N: columns that hold data
  const int HW = P.HW;
  const float fx = P.intr[0], fy = P.intr[1], cx = P.intr[2], cy = P.intr[3];
  const int tile_lo = s * P.tiles_per_split;
  const int tile_hi = min(tile_lo + P.tiles_per_split, (HW + T - 1) / T);
  const float* dk = P.disps + int64_t(k) * HW;
  const bool owner = warp < deg;                            // this warp owns edge `warp` of the keyframe

  const float inv_wd = 1.0f / float(P.wd);
  auto fill_px = [&](int tile, int pl, float dval) {        // pinhole.py:6-25
    float* b = spx + (tile & 1) * 3 * T;
    const int p = tile * T + pl;
    const int py = __float2int_rd((float(p) + 0.5f) * inv_wd), px = p - py * P.wd;   // exact for p < 2^22
    b[pl] = (float(px) - cx) / fx;
    b[T + pl] = (float(py) - cy) / fy;
    b[2 * T + pl] = dval;
  };

  // ---- the edge's streamed inputs, one tile ahead; the 27 sums of Hjj / vj in registers for the whole CTA ------
  const float* tg = P.targets;
  const float* wg = P.weights;
  float cur[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) cur[i] = 0.f;
  if (owner) {
    const int64_t eo = int64_t(P.cedge[e0 + warp]) * 2 * HW;
    tg += eo; wg += eo;
    if (tile_lo < tile_hi) load_inputs<STREAM>(tg, wg, HW, tile_lo * T, lane, cur);
  }
  if (owner && lane == 0) relative_pose_Rt(P.poses, k, P.cj[e0 + warp], sRt + 12 * warp);
  if (tid < T && tile_lo < tile_hi) fill_px(tile_lo, tid, (tile_lo * T + tid < HW) ? __ldg(dk + tile_lo * T + tid) : 0.f);
  if ((smem_u32(smem_raw) & 1023u) != 0u) __trap();
  {  // operand rows the panel never writes must be zero (no NaN patterns); the accumulator starts at zero
    float4* z = reinterpret_cast<float4*>(smem_raw);
    constexpr int n16 = int((2 * KB * TILE + 128 * ACS * sizeof(float)) / 16);
    for (int i = tid; i < n16; i += NT) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  if (warp == 0) tmem_alloc(tmem_p, 2 * ACC_COLS);
  if (tid == 0) { mbar_init(&mbar[0], 1); mbar_init(&mbar[1], 1); }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = *tmem_p;

  float hs[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) hs[i] = 0.f;

  // phase 2 mapping: warps 0-1, one thread per pixel of the tile (the other warps flush the previous tile meanwhile)
  const int pl2 = tid;
  const float* etak = P.eta + int64_t(m) * HW;
  float eta_c = 0.f;
  if (tid < T && tile_lo < tile_hi && tile_lo * T + pl2 < HW) eta_c = __ldg(etak + tile_lo * T + pl2);

  // flush mapping: warps 2.. ; quarter q = warp & 3 (TMEM lanes 32 q ..), the quarter's 16-column groups round-robin
  // over its flush warps (quarters 0 and 1 lend warps 0 and 1 to phase 2)
  const int fq = warp & 3;
  const int fj = (warp >> 2) - (fq < 2 ? 1 : 0);
  const int fnq = NW / 4 - (fq < 2 ? 1 : 0);
  const int frow = 32 * fq + lane;
  const uint32_t tlane = uint32_t(32 * fq) << 16;
  const int fmaxg = (6 * ((32 * fq + 31) / 6) + 5) >> 4;    // last 16-column group any row of this quarter needs
  const int gw = R >> 4;                                    // group of the w column
  const int ngroups = N >> 4;
  auto flush = [&](int buf) {
    if (DBA_V2_ABLATE & 4) return;
    if (warp < 2 || 32 * fq >= R) return;                   // warp-uniform
    for (int g = fj; g < ngroups; g += fnq) {
      if (g > fmaxg && g != gw) continue;                   // warp-uniform
      uint32_t v[16];
      tmem_ld16(tmem + tlane + uint32_t(buf * ACC_COLS + 16 * g), v);
      float4* a = reinterpret_cast<float4*>(sAcc + frow * ACS + 16 * g);
      float4 t4[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) t4[c] = a[c];
      tmem_ld_wait();
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        t4[c].x += __uint_as_float(v[4 * c + 0]); t4[c].y += __uint_as_float(v[4 * c + 1]);
        t4[c].z += __uint_as_float(v[4 * c + 2]); t4[c].w += __uint_as_float(v[4 * c + 3]);
      }
#pragma unroll
      for (int c = 0; c < 4; ++c) a[c] = t4[c];
    }
  };

#ifdef DBA_V2_TRACE
  long long tr[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  long long tlast = clock64(), trw = 0;
#endif
#pragma unroll 1
  for (int tile = tile_lo; tile < tile_hi; ++tile) {
    TR(9);
    const int p0 = tile * T;
    const float* px = spx + (tile & 1) * 3 * T;
    float eta_n = 0.f, d_n = 0.f;                           // warps 0-1: eta and disparity of pixel (tile + 1, pl2)
    if (warp < 2 && tile + 1 < tile_hi) {
      const int pn = p0 + T + pl2;
      if (pn < HW) { eta_n = __ldg(etak + pn); d_n = __ldg(dk + pn); }
    }
    // ---------------- phase 1: the warp's edge, lanes own pixels lane and lane + 32 ---------------------------
    if (owner && !(DBA_V2_ABLATE & 32)) {
      float Rt[12];
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        const float4 r4 = *reinterpret_cast<const float4*>(sRt + 12 * warp + 4 * i);
        Rt[4 * i] = r4.x; Rt[4 * i + 1] = r4.y; Rt[4 * i + 2] = r4.z; Rt[4 * i + 3] = r4.w;
      }
      float* pe = spanel + 6 * (1 + warp) * T;
      float* sp = spart + warp * PS;
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int pl = lane + 32 * q;
        const bool inb = p0 + pl < HW;
        const float dd = px[2 * T + pl];
        const Proj pr = project(Rt, px[pl], px[T + pl], dd, fx, fy, cx, cy);
        const float r0 = cur[4 * q + 0] - pr.u, r1 = cur[4 * q + 1] - pr.v;
        // out-of-range pixels carry zero weights (the loads return 0), so every sum below is exact
        const float w0 = kWeightScale * (pr.valid * cur[4 * q + 2]), w1 = kWeightScale * (pr.valid * cur[4 * q + 3]);
        const float X = pr.X1.x, Y = pr.X1.y, Z = pr.X1.z;
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
        pe[0 * T + pl] = dd * hx;                           // E_j = [d h, X1 x h]
        pe[1 * T + pl] = dd * hy;
        pe[2 * T + pl] = dd * hz;
        pe[3 * T + pl] = Y * hz - Z * hy;
        pe[4 * T + pl] = Z * hx - X * hz;
        pe[5 * T + pl] = X * hy - Y * hx;
        sp[0 * T + pl] = ck;
        sp[1 * T + pl] = wk;
        sp[2 * T + pl] = Rt[0] * hx + Rt[3] * hy + Rt[6] * hz;   // G = R^T h
        sp[3 * T + pl] = Rt[1] * hx + Rt[4] * hy + Rt[7] * hz;
        sp[4 * T + pl] = Rt[2] * hx + Rt[5] * hy + Rt[8] * hz;
      }
      // the next tile's inputs are requested AFTER this tile's were consumed: the scoreboards are counters, a wait
      // for `cur` issued after newer loads would wait for those too (ncu r2h: 7 % of the samples stalled there)
      asm volatile("" ::: "memory");
#pragma unroll
      for (int i = 0; i < 8; ++i) cur[i] = 0.f;
      if (tile + 1 < tile_hi) load_inputs<STREAM>(tg, wg, HW, p0 + T, lane, cur);
    }
    TR(0);
    __syncthreads();
    TR(1);
    // ---------------- phase 2 (warps 0-1, one pixel each): the edges' partials summed in a fixed order; the other
    // warps meanwhile add the PREVIOUS tile's tensor-core products into the shared-memory accumulator ------------
#ifdef DBA_V2_TRACE
    if (blockIdx.x == 0 && (tid == 0 || tid == 64)) { const long long t_ = clock64(); trw += t_ - tlast; tlast = t_; }
#endif
    if (warp < 8 && !(DBA_V2_ABLATE & 16)) {         // four threads per pixel sum the edges' partials (fixed order), lane 0 of the four stores
      const int plr = tid >> 2, part = tid & 3;
      float sum[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
      for (int w = part; w < deg; w += 4) {
        const float* spw = spart + w * PS + plr;
#pragma unroll
        for (int i = 0; i < 5; ++i) sum[i] += spw[i * T];
      }
#pragma unroll
      for (int mk = 1; mk <= 2; mk <<= 1)
#pragma unroll
        for (int i = 0; i < 5; ++i) sum[i] += __shfl_xor_sync(0xffffffffu, sum[i], mk);
      if (part == 0) {
#pragma unroll
        for (int i = 0; i < 5; ++i) ssum[i * T + plr] = sum[i];
      }
    }
    TR(2);
    __syncthreads();
    TR(3);
    if (warp < 2 && !(DBA_V2_ABLATE & 16)) {         // one thread per pixel: Q, w, the source-pose rows
      float sum[5];
#pragma unroll
      for (int i = 0; i < 5; ++i) sum[i] = ssum[i * T + pl2];
      const int p = p0 + pl2;
      const bool inb = p < HW;
      const float dd = px[2 * T + pl2], xx = px[pl2], yy = px[T + pl2];
      float C = sum[0], wsum = sum[1];
      const float Gx = sum[2], Gy = sum[3], Gz = sum[4];
      float Qv = 0.f, wv = 0.f, sq = 0.f;
      if (inb) {
        const float et = eta_c;
        if (P.disps_sens != nullptr) {
          const float ds = __ldg(P.disps_sens + int64_t(k) * HW + p);
          const float ms = ds > 0.f ? 1.f : 0.f;
          C = C + ms * kSensAlpha + (1.f - ms) * et + kCEps;
          wsum = wsum - ms * kSensAlpha * (dd - ds);
        } else {
          C = C + et + kCEps;                               // ba.py:92
        }
        float rs;                                           // sqrt(Q) = C^-1/2: MUFU.RSQ + one Newton step (~0.5 ulp)
        asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(rs) : "f"(C));
        rs = rs * fmaf(-0.5f * C, rs * rs, 1.5f);
        sq = rs;
        Qv = rs * rs;
        wv = wsum;
        P.Q[int64_t(m) * HW + p] = Qv;
        P.w[int64_t(m) * HW + p] = wv;
      }
      ssq[pl2] = sq;
      spanel[R * T + pl2] = wv;                             // operand row R
      spanel[0 * T + pl2] = -dd * Gx;                       // E_ii = -[d G, X0 x G],  X0 = [xx, yy, 1]
      spanel[1 * T + pl2] = -dd * Gy;
      spanel[2 * T + pl2] = -dd * Gz;
      spanel[3 * T + pl2] = -(yy * Gz - Gy);
      spanel[4 * T + pl2] = -(Gx - xx * Gz);
      spanel[5 * T + pl2] = -(xx * Gy - yy * Gx);
      eta_c = eta_n;
      if (tile + 1 < tile_hi) fill_px(tile + 1, pl2, d_n);
    } else if (tile > tile_lo) {
      // The wait for the previous tile's MMAs sits HERE, on the flush warps only: the reduce above needs neither the
      // accumulator nor the operand tiles, so it runs while the tensor pipe finishes (the wait used to be in front of
      // it, for all warps).  Warps 0-1 reach the operand tiles only behind the next barrier, i.e. after these waits.
      mbar_wait(&mbar[(tile - tile_lo - 1) & 1], uint32_t((tile - tile_lo - 1) >> 1) & 1u);
      tc_fence_after_sync();
      flush(((tile - tile_lo) & 1) ^ 1);
      tc_fence_before_sync();
    }
    TR(4);
    __syncthreads();
    TR(5);
    // ---------------- phase 3: operands for the tensor core ----------------------------------------------------
    if (!(DBA_V2_ABLATE & 8)) {
      // half a warp per operand row, lane = 4 consecutive pixels = one 16-byte chunk of the swizzled K-major tile:
      // v * sqrt(Q) -> hi = tf32(v), lo = v - hi.  Rows advance by 2 NW (a multiple of 8): constant byte stride.
      const int c16 = lane & 15;
      const int r0 = 2 * warp + (lane >> 4);
      const float4 sq4 = *reinterpret_cast<const float4*>(ssq + 4 * c16);
      uint32_t off = uint32_t(c16 >> 3) * TILE + uint32_t(r0 >> 3) * 1024u + uint32_t(r0 & 7) * 128u +
                     ((uint32_t(c16 & 7) ^ uint32_t(r0 & 7)) << 4);
      const float* src = spanel + r0 * T + 4 * c16;
#pragma unroll 2
      for (int r = r0; r <= R; r += 2 * NW) {
        const float4 v4 = *reinterpret_cast<const float4*>(src);
        float4 vh, vl;
        split_tf32(v4.x * sq4.x, vh.x, vl.x);
        split_tf32(v4.y * sq4.y, vh.y, vl.y);
        split_tf32(v4.z * sq4.z, vh.z, vl.z);
        split_tf32(v4.w * sq4.w, vh.w, vl.w);
        *reinterpret_cast<float4*>(hi + off) = vh;
        *reinterpret_cast<float4*>(lo + off) = vl;
        src += 2 * NW * T;
        off += uint32_t(2 * NW / 8) * 1024u;
      }
    }
    fence_proxy_async_smem();
    TR(6);
    __syncthreads();
    TR(7);
    const int buf = (tile - tile_lo) & 1;
    if (tid == NT - 32) {                                   // the last warp owns an edge only when deg == NW
      tc_fence_after_sync();
      const uint32_t idesc = make_idesc_tf32(128, N);
      const uint32_t ah = smem_u32(hi), al = smem_u32(lo);
      const uint32_t acc = tmem + uint32_t(buf * ACC_COLS);
      // the small cross terms first, so that the truncating TMEM adds act on a small accumulator
#pragma unroll
      for (int kb = 0; kb < ((DBA_V2_ABLATE & 3) ? 0 : KB); ++kb)
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t dh = make_desc_sw128(ah + kb * TILE + ks * 32);
          const uint64_t dl = make_desc_sw128(al + kb * TILE + ks * 32);
          mma_tf32_ss(acc, dh, dl, idesc, !(kb == 0 && ks == 0));
          mma_tf32_ss(acc, dl, dh, idesc, true);
        }
#pragma unroll
      for (int kb = 0; kb < KB; ++kb)
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t dh = make_desc_sw128(ah + kb * TILE + ks * 32);
          if (!(DBA_V2_ABLATE & 1)) mma_tf32_ss(acc, dh, dh, idesc, (DBA_V2_ABLATE & 2) ? !(kb == 0 && ks == 0) : true);
        }
      mma_commit(&mbar[buf]);
    }
    TR(8);
  }
#ifdef DBA_V2_TRACE
  if (blockIdx.x == 0 && (tid == 0 || tid == 64))
    printf("v2 trace tid %d deg %d tiles %d: P1 %lld syncA %lld reduce %lld syncA2 %lld tail/flush %lld syncB %lld convert %lld syncC %lld mma %lld looptop %lld mbarwait %lld\n",
           tid, deg, tile_hi - tile_lo, tr[0], tr[1], tr[2], tr[3], tr[4], tr[5], tr[6], tr[7], tr[8], tr[9], trw);
#endif
  if (tile_lo < tile_hi) {
    const int last = tile_hi - tile_lo - 1;
    mbar_wait(&mbar[last & 1], uint32_t(last >> 1) & 1u);
    tc_fence_after_sync();
    flush(last & 1);
    tc_fence_before_sync();
  }

  // ---------------- the CTA's partial sums -------------------------------------------------------------------
  if (owner) {
    const float tot = warp_reduce_scatter32(hs, lane);
    P.Hpart[(int64_t(e0 + warp) * P.S + s) * 32 + lane] = (lane < 27) ? tot : 0.f;
  }
  __syncthreads();                                          // sAcc complete
  {
    const int64_t slab0 = (int64_t(P.pairbase[m]) * P.S + int64_t(s) * npairs) * 36;
    const int64_t vE0 = (int64_t(P.blkbase[m]) * P.S + int64_t(s) * nblk) * 6;
    // (row, block B <= A): six consecutive floats of the slab [pair (A,B)][ri][0..5]
    for (int i = tid; i < R * nblk; i += NT) {
      const int row = i / nblk, B = i - row * nblk;
      const int A = row / 6, ri = row - A * 6;
      if (B <= A) {
        float* dst = P.Spart + slab0 + (int64_t(A) * (A + 1) / 2 + B) * 36 + ri * 6;
        const float* src = sAcc + row * ACS + 6 * B;
#pragma unroll
        for (int c = 0; c < 6; ++c) dst[c] = src[c];
      }
    }
    for (int row = tid; row < R; row += NT) P.vEpart[vE0 + row] = sAcc[row * ACS + R];
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 2 * ACC_COLS);
}

}  // namespace

int linearize_v2_max_degree() { return 20; }

size_t linearize_v2_smem_bytes(int degmax) { return degmax <= 16 ? v2_smem_bytes<16>() : v2_smem_bytes<20>(); }

cudaError_t launch_linearize_v2(const DeviceGraph& g, const BAInputs& in, const Workspace& ws, int tiles_per_split,
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
  // inputs that fit the L2 with room to spare stay resident across the LM iterations of a call
  const bool stream = g_tc_stream_mode < 0 ? size_t(g.E) * g.HW * 16 > size_t(100) << 20 : g_tc_stream_mode != 0;
  cudaError_t err;
  if (degmax <= 16) {
    auto kern = stream ? k_linearize_v2<16, true> : k_linearize_v2<16, false>;
    constexpr size_t smem = v2_smem_bytes<16>();
    err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (err != cudaSuccess) return err;
    kern<<<nkf * g.S, 32 * 16, smem, st>>>(P);
  } else {
    auto kern = stream ? k_linearize_v2<20, true> : k_linearize_v2<20, false>;
    constexpr size_t smem = v2_smem_bytes<20>();
    err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (err != cudaSuccess) return err;
    kern<<<nkf * g.S, 32 * 20, smem, st>>>(P);
  }
  g_launch_count++;
  return cudaGetLastError();
}

}  // namespace dba
