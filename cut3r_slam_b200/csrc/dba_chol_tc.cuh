// Large reduced systems (n > ~1000: loop-closure global BA, n = 3060): right-looking Cholesky with 128-column
// panels and the trailing update on the 5th-generation tensor cores.  Included by dba_chol_fused.cu (it re-uses
// potrf32_warp / solve_row / the Ldiag layout of that file).
//
// Why: the cooperative kernel above walks 32-column steps, so the trailing matrix (37 MB at n = 3060) is read and
// written 96 times and updated with SIMT FMAs: 1.49 ms per factorisation, 5.7x its fp32 FMA floor (r1 profiles).
// Here a step is
//   k_potrf128 : one CTA factors the 128 x 128 diagonal block in shared memory (four 32-column sub-steps: register
//                potrf in one warp, one thread per row for the rows below, SIMT update of the rest of the block) and
//                publishes L_kk plus the per-32-block L^T, 1/diag, L^-1 the triangular solves use;
//   k_syrk_tc<1>: X = A[rows below, panel] L_kk^-T as a tensor-core product with the explicit inverse of the diagonal
//                block (the rhs row n rides along as a matrix row: the forward solve falls out of the factorisation);
//                a first version with one thread per row doing the substitution took 80 us per panel
//   k_syrk_tc<0>: C -= X_i X_j^T on 128 x 128 tiles: the fp32 panel rows are split into tf32 (hi, lo) pairs while
//                they are copied into the SWIZZLE_128B K-major operand tiles, tcgen05.mma accumulates
//                hi lo^T + lo hi^T + hi hi^T over the 128 panel columns in TMEM ("3xTF32", two K-halves of 64),
//                the result is subtracted from the tile in global memory (L2 resident).
// 24 steps instead of 96; 9.5 GFLOP of trailing update become ~1.1 k tensor-core tiles.
#pragma once

namespace {

using namespace umma;

constexpr int PW = 128;              // panel width
constexpr int PLD = PW + 4;          // row stride of the diagonal block in shared memory

// dynamic shared memory of k_potrf128: sA [PW][PLD] | LT [32][32] | rinv [32] | Pp [32][PW] | col [2][32] | sLi [PW][PLD]
__host__ __device__ inline size_t potrf128_smem_floats() { return 2 * size_t(PW) * PLD + 32 * 32 + 32 + 32 * PW + 64; }

__global__ void __launch_bounds__(256)
k_potrf128(const float* __restrict__ A, float* __restrict__ Lm, float* __restrict__ Ldiag, float* __restrict__ LinvW,
           int ld, int n, int k0, int* __restrict__ fail) {
  extern __shared__ __align__(16) float psm[];
  float* sA = psm;                                   // the diagonal block, lower triangle
  float (*LT)[NB] = reinterpret_cast<float (*)[NB]>(sA + PW * PLD);   // LT[j][c] = L[c][j] of the current 32-block
  float* rinv = sA + PW * PLD + NB * NB;
  float* Pp = rinv + NB;                             // [32][PW]: solved rows below the 32-block, k-major
  float* col = Pp + NB * PW;
  float* sLi = col + 64;                             // [PW][PLD]: L_kk^-1 (lower), built from the 32-block inverses
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int w = min(PW, n - k0);
  for (int e = tid; e < PW * PLD; e += 256) sLi[e] = 0.f;
  for (int e = tid; e < PW * (PW / 4); e += 256) {   // 128 rows x 32 float4 chunks
    const int i = e >> 5, c4 = (e & 31) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < w && c4 <= i) v = __ldcg(reinterpret_cast<const float4*>(A + int64_t(k0 + i) * ld + k0 + c4));
    *reinterpret_cast<float4*>(sA + i * PLD + c4) = v;
  }
  __syncthreads();
  const int nsb = (w + NB - 1) / NB;
  for (int sb = 0; sb < nsb; ++sb) {
    const int s0 = sb * NB, nb = min(NB, w - s0), r0 = s0 + nb;
    float* Lg = Ldiag + int64_t(k0 / NB + sb) * LD_STRIDE;
    if (warp == 0) {                                 // ---- 32 x 32 diagonal block in registers
      float a[NB];
#pragma unroll
      for (int c = 0; c < NB; ++c) a[c] = (lane < nb && c < nb) ? sA[(s0 + lane) * PLD + s0 + c] : ((c == lane) ? 1.f : 0.f);
      const bool bad = potrf32_warp(a, lane, col);
#pragma unroll
      for (int c = 0; c < NB; ++c) {
        const float v = (c <= lane) ? a[c] : 0.f;
        LT[c][lane] = v;
        Lg[LD_LT + c * NB + lane] = v;
        if (lane < nb && c < nb && c <= lane) sA[(s0 + lane) * PLD + s0 + c] = v;
      }
      __syncwarp();
      const float ri = 1.0f / LT[lane][lane];
      rinv[lane] = ri;
      Lg[LD_RINV + lane] = ri;
      if (bad && lane == 0) *fail = 1;
    }
    __syncthreads();
    const int nrow = w - r0;                         // rows of the block below this 32-block (<= 96)
    if (tid < nrow) {                                // ---- x L^T = b, one thread per row
      float b[NB];
      load_row(b, sA + (r0 + tid) * PLD + s0);
      solve_row(b, LT, rinv);
#pragma unroll
      for (int j = 0; j < NB; ++j) {
        Pp[j * PW + tid] = (j < nb) ? b[j] : 0.f;
        if (j < nb) sA[(r0 + tid) * PLD + s0 + j] = b[j];
      }
    } else if (warp == 7) {                          // ---- L_kk^-1 of the 32-block (column `lane` solves L v = e_lane)
      float b[NB];
#pragma unroll
      for (int c = 0; c < NB; ++c) b[c] = (c == lane) ? 1.f : 0.f;
      solve_row(b, LT, rinv);
#pragma unroll
      for (int j = 0; j < NB; ++j) { Lg[LD_LINV + j * NB + lane] = b[j]; sLi[(s0 + j) * PLD + s0 + lane] = b[j]; }
    }
    __syncthreads();
    {  // ---- rest of the block (lower part), 4 x 4 register tiles: tile (ty, tx <= ty) of the nrow x nrow remainder
      const int nt4 = (nrow + 3) >> 2;
      for (int t = tid; t < nt4 * nt4; t += 256) {
        const int ty = t / nt4, tx = t - ty * nt4;
        if (tx > ty) continue;
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
#pragma unroll 8
        for (int m = 0; m < NB; ++m) {
          const float4 a4 = *reinterpret_cast<const float4*>(Pp + m * PW + 4 * ty);
          const float4 b4 = *reinterpret_cast<const float4*>(Pp + m * PW + 4 * tx);
          const float av[4] = {a4.x, a4.y, a4.z, a4.w}, bv[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int ri = 4 * ty + i, rc = 4 * tx + j;
            if (ri < nrow && rc <= ri) sA[(r0 + ri) * PLD + r0 + rc] -= acc[i][j];
          }
      }
    }
    __syncthreads();
  }
  // ---- the full inverse of the 128 x 128 factor from its 32-block inverses (k_trsm as a tensor-core product):
  // column block j, rows i > j in turn:  Linv_ij = -Linv_ii (sum_{k=j}^{i-1} L_ik Linv_kj);  Pp holds the sum
  for (int j = 0; j + 1 < nsb; ++j)
    for (int i = j + 1; i < nsb; ++i) {
      {
        const int r = tid >> 3, c4 = (tid & 7) * 4;   // thread -> row r of the 32 x 32 block, 4 columns
        float t4[4] = {0.f, 0.f, 0.f, 0.f};
        for (int k = j; k < i; ++k)
#pragma unroll 8
          for (int m = 0; m < NB; ++m) {
            const float l = sA[(i * NB + r) * PLD + k * NB + m];
            const float4 v = *reinterpret_cast<const float4*>(sLi + (k * NB + m) * PLD + j * NB + c4);
            t4[0] = fmaf(l, v.x, t4[0]); t4[1] = fmaf(l, v.y, t4[1]); t4[2] = fmaf(l, v.z, t4[2]); t4[3] = fmaf(l, v.w, t4[3]);
          }
        *reinterpret_cast<float4*>(Pp + r * PW + c4) = make_float4(t4[0], t4[1], t4[2], t4[3]);
      }
      __syncthreads();
      {
        const int r = tid >> 3, c4 = (tid & 7) * 4;
        float t4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 8
        for (int m = 0; m < NB; ++m) {
          const float l = sLi[(i * NB + r) * PLD + i * NB + m];
          const float4 v = *reinterpret_cast<const float4*>(Pp + m * PW + c4);
          t4[0] = fmaf(l, v.x, t4[0]); t4[1] = fmaf(l, v.y, t4[1]); t4[2] = fmaf(l, v.z, t4[2]); t4[3] = fmaf(l, v.w, t4[3]);
        }
        *reinterpret_cast<float4*>(sLi + (i * NB + r) * PLD + j * NB + c4) = make_float4(-t4[0], -t4[1], -t4[2], -t4[3]);
      }
      __syncthreads();
    }
  for (int e = tid; e < PW * (PW / 4); e += 256) {   // L_kk^-1, rows / columns >= w zero
    const int i = e >> 5, c4 = (e & 31) * 4;
    float4 v = *reinterpret_cast<const float4*>(sLi + i * PLD + c4);
    if (i >= w) v = make_float4(0.f, 0.f, 0.f, 0.f);
    *reinterpret_cast<float4*>(LinvW + i * PW + c4) = v;
  }
  for (int e = tid; e < w * (PW / 4); e += 256) {    // the factor's diagonal block (lower; the rest is never read)
    const int i = e >> 5, c4 = (e & 31) * 4;
    if (c4 <= i) *reinterpret_cast<float4*>(Lm + int64_t(k0 + i) * ld + k0 + c4) = *reinterpret_cast<const float4*>(sA + i * PLD + c4);
  }
}

}  // namespace

namespace {

constexpr int SY_T = 256;
constexpr uint32_t SY_TILE = TILE_BYTES_128;         // one K-block: 128 rows x 32 fp32, swizzled
// dynamic shared memory of k_syrk_tc: A hi | A lo | B hi | B lo, two K-blocks (64 panel columns) each
constexpr size_t SY_SMEM = 4 * 2 * SY_TILE + 64;

// rows [row0, row0 + 128) x panel columns [kc, kc + 64) of X (in Lm) -> tf32 (hi, lo) operand tiles
__device__ __forceinline__ void syrk_stage(const float* __restrict__ Lm, int ld, int row0, int row_max, int k0, int kc,
                                           int w, uint8_t* hi, uint8_t* lo, int tid) {
#pragma unroll 8
  for (int e = tid; e < 128 * 16; e += SY_T) {       // 128 rows x 16 chunks of 4 columns
    const int r = e >> 4, c16 = e & 15;
    const int row = row0 + r, kk = kc + 4 * c16;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row <= row_max && kk < w) v = __ldcg(reinterpret_cast<const float4*>(Lm + int64_t(row) * ld + k0 + kk));
    if (kk + 1 >= w) v.y = 0.f;
    if (kk + 2 >= w) v.z = 0.f;
    if (kk + 3 >= w) v.w = 0.f;
    float4 vh, vl;
    split_tf32(v.x, vh.x, vl.x); split_tf32(v.y, vh.y, vl.y); split_tf32(v.z, vh.z, vl.z); split_tf32(v.w, vh.w, vl.w);
    const uint32_t off = uint32_t(c16 >> 3) * SY_TILE + uint32_t(r >> 3) * 1024u + uint32_t(r & 7) * 128u +
                         ((uint32_t(c16 & 7) ^ uint32_t(r & 7)) << 4);
    *reinterpret_cast<float4*>(hi + off) = vh;
    *reinterpret_cast<float4*>(lo + off) = vl;
  }
}

// MODE 0: C[ti, tj] -= X_i X_j^T for the lower tiles of the trailing matrix (rows r0..n incl. the rhs row, columns
//         r0..n-1); X = the solved panel in Lm.
// MODE 1: the panel solve as a product with the inverse of the diagonal block: X[rows r0..n] = A[rows, panel] L_kk^-T,
//         written to Lm (one CTA per 128 rows; LinvW [128][128] from k_potrf128).
template <int MODE>
__global__ void __launch_bounds__(SY_T, 1)
k_syrk_tc(float* __restrict__ A, float* __restrict__ Lm, const float* __restrict__ LinvW, int ld, int n, int k0) {
  extern __shared__ __align__(1024) uint8_t ssm[];
  uint8_t* Ahi = ssm; uint8_t* Alo = ssm + 2 * SY_TILE; uint8_t* Bhi = ssm + 4 * SY_TILE; uint8_t* Blo = ssm + 6 * SY_TILE;
  uint64_t* mbar = reinterpret_cast<uint64_t*>(ssm + 8 * SY_TILE);
  uint32_t* tmem_p = reinterpret_cast<uint32_t*>(mbar + 1);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int w = min(PW, n - k0), r0 = k0 + w;
  int ti = blockIdx.x, tj = 0;
  if (MODE == 0) tri_decode(blockIdx.x, ti, tj);
  const int rowi = r0 + ti * PW, rowj = (MODE == 0) ? r0 + tj * PW : 0;
  const bool diag = MODE == 0 && ti == tj;
  if ((smem_u32(ssm) & 1023u) != 0u) __trap();
  if (warp == 0) tmem_alloc(tmem_p, 128);
  if (tid == 0) mbar_init(mbar, 1);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = *tmem_p;
  const uint32_t idesc = make_idesc_tf32(128, 128);
  uint32_t phase = 0;
  for (int kc = 0; kc < w; kc += 64) {
    if (kc > 0) {                                    // the first half's MMAs still read the operand tiles
      mbar_wait(mbar, phase);
      phase ^= 1;
      tc_fence_after_sync();
    }
    if (MODE == 0) {
      syrk_stage(Lm, ld, rowi, n, k0, kc, w, Ahi, Alo, tid);
      if (!diag) syrk_stage(Lm, ld, rowj, n - 1, k0, kc, w, Bhi, Blo, tid);
    } else {
      syrk_stage(A, ld, rowi, n, k0, kc, w, Ahi, Alo, tid);
      syrk_stage(LinvW, PW, 0, w - 1, 0, kc, w, Bhi, Blo, tid);
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after_sync();
      const uint32_t ah = smem_u32(Ahi), al = smem_u32(Alo);
      const uint32_t bh = diag ? ah : smem_u32(Bhi), bl = diag ? al : smem_u32(Blo);
#pragma unroll
      for (int kb = 0; kb < 2; ++kb)
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint32_t o = kb * SY_TILE + ks * 32;
          mma_tf32_ss(tmem, make_desc_sw128(ah + o), make_desc_sw128(bl + o), idesc, !(kc == 0 && kb == 0 && ks == 0));
          mma_tf32_ss(tmem, make_desc_sw128(al + o), make_desc_sw128(bh + o), idesc, true);
        }
#pragma unroll
      for (int kb = 0; kb < 2; ++kb)
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint32_t o = kb * SY_TILE + ks * 32;
          mma_tf32_ss(tmem, make_desc_sw128(ah + o), make_desc_sw128(bh + o), idesc, true);
        }
      mma_commit(mbar);
    }
  }
  mbar_wait(mbar, phase);
  tc_fence_after_sync();
  {  // C -= D: warp = (lane quarter q, column half h); a lane owns one row, 16 consecutive columns per TMEM load
    const int q = warp & 3, h = warp >> 2;
    const int gi = rowi + 32 * q + lane;
    const uint32_t tlane = uint32_t(32 * q) << 16;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const int cbase = 64 * h + 16 * g;
      uint32_t v[16];
      tmem_ld16(tmem + tlane + uint32_t(cbase), v);
      tmem_ld_wait();
      const int gj = rowj + cbase;
      if (gi <= n) {
#pragma unroll
        for (int c4 = 0; c4 < 16; c4 += 4) {
          const int col = gj + c4;
          if (MODE == 0) {
            // columns past the diagonal of a diagonal tile hold symmetric values nobody reads; columns >= n do not exist
            if (col < n && col <= gi) {
              float4* p = reinterpret_cast<float4*>(A + int64_t(gi) * ld + col);
              float4 c = __ldcg(p);
              c.x -= __uint_as_float(v[c4]); c.y -= __uint_as_float(v[c4 + 1]);
              c.z -= __uint_as_float(v[c4 + 2]); c.w -= __uint_as_float(v[c4 + 3]);
              __stcg(p, c);
            }
          } else if (col < w) {
            __stcg(reinterpret_cast<float4*>(Lm + int64_t(gi) * ld + k0 + col),
                   make_float4(__uint_as_float(v[c4]), __uint_as_float(v[c4 + 1]), __uint_as_float(v[c4 + 2]),
                               __uint_as_float(v[c4 + 3])));
          }
        }
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 128);
}

}  // namespace
