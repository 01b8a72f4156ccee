// Cholesky + solve of the reduced pose system, latency-oriented (n = 6 P' is 60 .. ~3000: the
// factorisation is a chain of n/32 dependent block steps, so the cost is per-step latency).
//
// Replaces torch.linalg.cholesky + cholesky_solve of the reference (hislam2/geom/chol.py:5-20,66).
// Storage: A (n+1) x ld row-major holds the running Schur complement (lower) with the rhs as row n
// (the forward solve falls out of the factorisation); the factor goes to a SEPARATE matrix Lm of
// the same shape, so the pre-solve panel entries of A stay readable by every CTA during a step.
// Ldiag holds, per 32x32 diagonal block: L^T, 1/diag(L) and L^-1.
//
// k_chol_fused -- one co-resident grid, ONE grid barrier per 32-column block step k:
//   * CTA 0 runs one step ahead (look-ahead): it solves the 32 panel rows of block row k+1, updates
//     the next diagonal block, factors it in one warp (registers + shuffles, rsqrt, the next pivot
//     issued before the rank-1 tail) and publishes L^T, 1/diag, L^-1 for step k+1;
//   * every other CTA owns 64x64 trailing tiles: the diagonal factor, the tile's panel rows
//     (cp.async, no register staging) and the tile itself (registers) arrive in one L2 round trip;
//     the panel rows are solved by forward substitution, one thread per row, against the factor
//     broadcast from shared memory (redundant per tile instead of a second grid barrier); owners of
//     diagonal tiles publish their rows of the factor; the tile is updated in registers and stored.
// k_backsolve (1 CTA): L^T x = y by 32-column blocks: x_k = L_kk^-T y_k (a 32x32 product in one warp),
// then y -= L_k.^T x_k with the rows of L fetched while the warp works.
// A non-positive / NaN pivot raises the flag -> dx = 0 (chol.py:14-18 behaviour).
#include "dba_kernels.h"
#include "dba_umma.cuh"

namespace dba {

#ifdef DBA_CHOL_TRACE
void chol_trace_read(unsigned long long* out, int n);
#endif

namespace {

constexpr int NB = 32;        // block step
constexpr int TT = 64;        // trailing tile
constexpr int NTH = 256;
constexpr int SBS = NB + 4;   // row stride of the staged panel / diagonal block (16-byte aligned rows)
constexpr int SPS = TT + 4;   // row stride of the solved panel (k-major)
constexpr unsigned FULL = 0xffffffffu;
// Ldiag, per diagonal block: LT[32][32] (LT[j][c] = L[c][j]) | rinv[32] | Linv[32][32] (row-major)
constexpr int LD_LT = 0, LD_RINV = NB * NB, LD_LINV = NB * NB + NB, LD_STRIDE = 2 * NB * NB + NB;

__device__ __forceinline__ void tri_decode(int t, int& ti, int& tj) {
  int a = int((sqrtf(8.0f * float(t) + 1.0f) - 1.0f) * 0.5f);
  while ((a + 1) * (a + 2) / 2 <= t) ++a;
  while (a * (a + 1) / 2 > t) --a;
  ti = a;
  tj = t - a * (a + 1) / 2;
}

__device__ __forceinline__ void cp_async16(float* smem_dst, const float* gsrc, bool valid) {
  const unsigned d = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
  const int sz = valid ? 16 : 0;       // src-size 0: the 16 bytes are zero-filled, nothing is read
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(gsrc), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }

// Grid barrier over co-resident CTAs (cooperative launch): one atomic per CTA + an acquire spin.
// All data shared between CTAs is read with .cg (L2) loads, so no L1 invalidation is needed.
__device__ __forceinline__ void grid_barrier(unsigned* bar, unsigned target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(bar, 1u);
    unsigned v;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];\n" : "=r"(v) : "l"(bar) : "memory");
    } while (v < target);
  }
  __syncthreads();
}

#ifdef DBA_CHOL_TRACE
__device__ unsigned long long g_trace[16 * 512];
__device__ __forceinline__ unsigned long long gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define TRACE(cta, slot) do { if (blockIdx.x == (cta) && threadIdx.x == 0 && kb >= 0 && kb < 512) g_trace[kb * 16 + (slot)] = gtimer(); } while (0)
#else
#define TRACE(cta, slot) do { } while (0)
#endif

// (a0, a1) -= s * (x0, x1) as ONE packed instruction (sm_100 FFMA2 with a scalar broadcast operand; each component
// rounds like fmaf(-s, x, a): bit-identical to two FFMAs).  The factorisation's critical path is single-warp code that
// issues one instruction every ~3.6 cycles, so halving the FMA count is what shortens it.
__device__ __forceinline__ void fms2(float& a0, float& a1, float s, float x0, float x1) {
  const float2 r = __ffma2_rn(make_float2(-s, -s), make_float2(x0, x1), make_float2(a0, a1));
  a0 = r.x; a1 = r.y;
}

struct __align__(16) CholSmem {
  float LT[NB][NB];          // LT[j][c] = L[c][j]  (column j of the factor, contiguous)
  float rinv[NB];            // 1 / L[j][j]          (LT and rinv are contiguous: one flat copy from Ldiag)
  float D[NB][SBS];          // CTA 0: the next diagonal block
  float B[2][TT][SBS];       // staged pre-solve panel rows of the tile's two row groups
  float P[2][NB][SPS];       // solved panel rows, k-major: P[g][c][row]
};

// One warp, lane = row of the block held in a[0..31] (lower part valid, identity padding).
// On exit a[] holds row `lane` of L.  Returns true when a pivot was not positive.
// Critical chain per column: FMUL -> FFMA -> SHFL -> MUFU.RSQ; the rank-1 tail fills the bubbles.
__device__ __forceinline__ float rsqrt_fast(float x) {
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
#ifndef DBA_CHOL_NO_NEWTON
  r = r * fmaf(-0.5f * x, r * r, 1.5f);     // one Newton step: the 2-ulp approximation -> ~0.5 ulp
#endif
  return r;
}

__device__ __forceinline__ bool potrf32_warp(float (&a)[NB], int lane, float* col /* smem [2][32] scratch */) {
  // A non-positive pivot turns into NaN (rsqrt of a negative number, 0 * inf) and poisons every later
  // pivot, so positivity is tested once on the finished diagonal instead of inside the chain.
  // Only two shuffles per column are on the chain (next pivot, next column's multiplier); the rest of
  // column j reaches the other lanes through shared memory (broadcast float4 reads).
  float d = __shfl_sync(FULL, a[0], 0);
  float dg = 1.f;
#pragma unroll
  for (int j = 0; j < NB; ++j) {
    const float r = rsqrt_fast(d);
    const float l = a[j] * r;            // lane j: d * rsqrt(d) = sqrt(d); lanes < j: unused upper part
    a[j] = l;
    if (lane == j) dg = l;
    if (j + 1 < NB) {
      d = __shfl_sync(FULL, fmaf(-l, l, a[j + 1]), j + 1);       // lane j+1 owns a[j+1] and its own l
      const float l1 = __shfl_sync(FULL, l, j + 1);
      a[j + 1] = fmaf(-l, l1, a[j + 1]);
    }
    if (j + 2 < NB) {
      float* cj = col + (j & 1) * NB;
      cj[lane] = l;
      __syncwarp();
#pragma unroll
      for (int c4 = ((j + 2) & ~3); c4 < NB; c4 += 4) {
        const float4 l4 = *reinterpret_cast<const float4*>(cj + c4);
        // columns >= j + 2 (meaningful for lane >= c), in aligned pairs where both are
        if (c4 + 0 >= j + 2) fms2(a[c4 + 0], a[c4 + 1], l, l4.x, l4.y);
        else if (c4 + 1 >= j + 2) a[c4 + 1] = fmaf(-l, l4.y, a[c4 + 1]);
        if (c4 + 2 >= j + 2) fms2(a[c4 + 2], a[c4 + 3], l, l4.z, l4.w);
        else if (c4 + 3 >= j + 2) a[c4 + 3] = fmaf(-l, l4.w, a[c4 + 3]);
      }
    }
  }
  const bool ok = (dg > 0.f) && (dg < 3.0e38f);
  return !__all_sync(FULL, ok);
}

// x L^T = b for one row b (forward substitution, column-oriented); b is overwritten by x.
__device__ __forceinline__ void solve_row(float (&b)[NB], const float (*LT)[NB], const float* rinv) {
  // column j+1 of L and its reciprocal pivot are fetched (broadcast float4 reads) before the
  // multiply-adds of column j are issued, so the shared-memory latency stays off the chain
  float4 cur[NB / 4], nxt[NB / 4];
  float ri = rinv[0], rn = 0.f;
#pragma unroll
  for (int g = 0; g < NB / 4; ++g) cur[g] = *reinterpret_cast<const float4*>(&LT[0][g * 4]);
#pragma unroll
  for (int j = 0; j < NB; ++j) {
    if (j + 1 < NB) {
      rn = rinv[j + 1];
#pragma unroll
      for (int g = (j + 2) / 4; g < NB / 4; ++g) nxt[g] = *reinterpret_cast<const float4*>(&LT[j + 1][g * 4]);
    }
    const float x = b[j] * ri;
    b[j] = x;
#pragma unroll
    for (int g = (j + 1) / 4; g < NB / 4; ++g) {
      const int c4 = g * 4;
      if (c4 + 0 > j) fms2(b[c4 + 0], b[c4 + 1], x, cur[g].x, cur[g].y);
      else if (c4 + 1 > j) b[c4 + 1] = fmaf(-x, cur[g].y, b[c4 + 1]);
      if (c4 + 2 > j) fms2(b[c4 + 2], b[c4 + 3], x, cur[g].z, cur[g].w);
      else if (c4 + 3 > j) b[c4 + 3] = fmaf(-x, cur[g].w, b[c4 + 3]);
    }
    ri = rn;
#pragma unroll
    for (int g = 0; g < NB / 4; ++g) cur[g] = nxt[g];
  }
}

__device__ __forceinline__ void load_row(float (&b)[NB], const float* src) {
#pragma unroll
  for (int c4 = 0; c4 < NB; c4 += 4) {
    const float4 v = *reinterpret_cast<const float4*>(src + c4);
    b[c4] = v.x; b[c4 + 1] = v.y; b[c4 + 2] = v.z; b[c4 + 3] = v.w;
  }
}

// L^-1 of the block held in S.LT / S.rinv (needed by the backward solve only): column `lane` solves L v = e_lane
__device__ __forceinline__ void inverse_stage(CholSmem& S, float* Lg, int lane) {
  float b[NB];
#pragma unroll
  for (int c = 0; c < NB; ++c) b[c] = (c == lane) ? 1.f : 0.f;
  solve_row(b, S.LT, S.rinv);
#pragma unroll
  for (int j = 0; j < NB; ++j) Lg[LD_LINV + j * NB + lane] = b[j];
}

// CTA 0: factor diagonal block kn (rows q0..q0+nbn-1).  With `panel`, S.LT / S.rinv hold block kn-1
// (columns p0..p0+31) and the block row is first brought up to date: X = A[q, p] L_pp^-T, D -= X X^T.
__device__ __forceinline__ void diag_stage(CholSmem& S, float* A, float* Lm, float* Ldiag, int ld, int n, int kn,
                                           bool panel, int* fail) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int kb = kn - 1;   // trace slot
  const int q0 = kn * NB, nbn = min(NB, n - q0), p0 = q0 - NB;
  {  // 32 rows x 8 chunks: the diagonal block, and the raw panel rows of the block row
    const int r = tid >> 3, ch = tid & 7;
    cp_async16(&S.D[r][ch * 4], A + int64_t(q0 + r) * ld + q0 + ch * 4, (r < nbn) && (q0 + ch * 4 < ld));
    if (panel) cp_async16(&S.B[0][r][ch * 4], A + int64_t(q0 + r) * ld + p0 + ch * 4, r < nbn);
  }
  cp_async_wait_all();
  __syncthreads();
  TRACE(0, 8);
  if (panel) {
    if (warp == 0) {
      float b[NB];
      load_row(b, &S.B[0][lane][0]);
      solve_row(b, S.LT, S.rinv);
#pragma unroll
      for (int j = 0; j < NB; ++j) S.P[0][j][lane] = b[j];
    } else if (warp == 1) {
      inverse_stage(S, Ldiag + int64_t(kn - 1) * LD_STRIDE, lane);   // off the critical path
    }
    __syncthreads();
    TRACE(0, 9);
    {  // D -= X X^T : thread -> row r, columns c4..c4+3
      const int r = tid >> 3, c4 = (tid & 7) * 4;
      float4 d = *reinterpret_cast<const float4*>(&S.D[r][c4]);
      if (c4 <= r) {                       // only the lower triangle of the block is read afterwards
        // four independent chains (m mod 4), added once at the end: the sum over m is what the factor needs, not its
        // order (the panel solve that produced X is rounded already); 8 dependent FFMA2 per chain instead of 32
        float4 e1 = make_float4(0.f, 0.f, 0.f, 0.f), e2 = e1, e3 = e1;
#pragma unroll
        for (int m = 0; m < NB; m += 4) {
          const float x0 = S.P[0][m][r], x1 = S.P[0][m + 1][r], x2 = S.P[0][m + 2][r], x3 = S.P[0][m + 3][r];
          const float4 c0 = *reinterpret_cast<const float4*>(&S.P[0][m][c4]);
          const float4 c1 = *reinterpret_cast<const float4*>(&S.P[0][m + 1][c4]);
          const float4 c2 = *reinterpret_cast<const float4*>(&S.P[0][m + 2][c4]);
          const float4 c3 = *reinterpret_cast<const float4*>(&S.P[0][m + 3][c4]);
          fms2(d.x, d.y, x0, c0.x, c0.y);   fms2(d.z, d.w, x0, c0.z, c0.w);
          fms2(e1.x, e1.y, x1, c1.x, c1.y); fms2(e1.z, e1.w, x1, c1.z, c1.w);
          fms2(e2.x, e2.y, x2, c2.x, c2.y); fms2(e2.z, e2.w, x2, c2.z, c2.w);
          fms2(e3.x, e3.y, x3, c3.x, c3.y); fms2(e3.z, e3.w, x3, c3.z, c3.w);
        }
        d.x += (e1.x + e2.x) + e3.x; d.y += (e1.y + e2.y) + e3.y; d.z += (e1.z + e2.z) + e3.z; d.w += (e1.w + e2.w) + e3.w;
      }
      *reinterpret_cast<float4*>(&S.D[r][c4]) = d;
    }
    __syncthreads();
    TRACE(0, 10);
  }
  float* Lg = Ldiag + int64_t(kn) * LD_STRIDE;
  if (warp == 0) {
    float a[NB];
    load_row(a, &S.D[lane][0]);
#pragma unroll
    for (int c = 0; c < NB; ++c)
      if (!(lane < nbn && c < nbn)) a[c] = (c == lane) ? 1.f : 0.f;       // identity padding
    const bool bad = potrf32_warp(a, lane, &S.P[1][0][0]);
#pragma unroll
    for (int c = 0; c < NB; ++c) {
      const float v = (c <= lane) ? a[c] : 0.f;
      S.LT[c][lane] = v;
      Lg[LD_LT + c * NB + lane] = v;
      if (lane < nbn && c < nbn && c <= lane) Lm[int64_t(q0 + lane) * ld + q0 + c] = v;
    }
    __syncwarp();
    const float ri = 1.0f / S.LT[lane][lane];
    S.rinv[lane] = ri;
    Lg[LD_RINV + lane] = ri;
    if (bad && lane == 0) *fail = 1;
  }
  __syncthreads();
  TRACE(0, 11);
}

// MINB = 2 (n >= 1024): 128 registers with ~170 bytes of spills so that two CTAs share an SM and cover each other's tile
// loads (n = 3060: 1452 -> 1374 us); MINB = 1 (200 registers, no spills) where the look-ahead CTA's chain is the step
// (n = 756: the two-CTA build costs 9 us per factorisation in the pipeline)
template <int MINB>
__global__ void __launch_bounds__(NTH, MINB)
k_chol_fused(float* A, float* Lm, float* Ldiag, int ld, int n, int* fail, unsigned* bar) {
  __shared__ CholSmem S;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nkb = (n + NB - 1) / NB;
  const int nworkers = int(gridDim.x) - 1;
  unsigned target = 0;

  if (blockIdx.x == 0) diag_stage(S, A, Lm, Ldiag, ld, n, 0, false, fail);
  target += gridDim.x;
  grid_barrier(bar, target);

  for (int kb = 0; kb < nkb; ++kb) {
    const int k0 = kb * NB, nb = min(NB, n - k0), r0 = k0 + nb;
    const int ntj = (n - r0 + TT - 1) / TT;            // row groups with matrix rows
    const int nti = (n + 1 - r0 + TT - 1) / TT;        // ... including the rhs row n
    const int ntiles = ntj * (ntj + 1) / 2 + (nti > ntj ? ntj : 0);
    TRACE(1, 0);
    if (blockIdx.x == 0) {
      TRACE(0, 6);
      if (ntj == 0) {
        // last block: only the rhs row is left below the diagonal block: y_k = L_kk^-1 rhs_k
        if (warp == 0) {
          float b = (lane < nb) ? __ldcg(A + int64_t(n) * ld + k0 + lane) : 0.f;
          float x = 0.f;
#pragma unroll
          for (int j = 0; j < NB; ++j) {
            const float xj = __shfl_sync(FULL, b, j) * S.rinv[j];
            if (lane == j) x = xj;
            b = fmaf(-xj, S.LT[j][lane], b);           // LT[j][lane] = L[lane][j], zero for lane < j
          }
          if (lane < nb) Lm[int64_t(n) * ld + k0 + lane] = x;
        } else if (warp == 1) {
          inverse_stage(S, Ldiag + int64_t(kb) * LD_STRIDE, lane);
        }
      } else {
        diag_stage(S, A, Lm, Ldiag, ld, n, kb + 1, true, fail);
      }
      TRACE(0, 7);
    } else if (int(blockIdx.x) - 1 < ntiles) {
      // ---- issue every load of the first tile: factor of the diagonal block, panel rows, the tile ----
      {
        const float* Lg = Ldiag + int64_t(kb) * LD_STRIDE;
        float* dst = &S.LT[0][0];
        for (int ch = tid; ch < (NB * NB + NB) / 4; ch += NTH) cp_async16(dst + ch * 4, Lg + ch * 4, true);
      }
      int ti = 0, tj = 0;
      float4 cacc[4];
      auto issue_tile = [&](int tt) {
        tri_decode(tt, ti, tj);
        const int rowi = r0 + ti * TT, rowj = r0 + tj * TT;
#pragma unroll
        for (int q = 0; q < 4; ++q) {                    // 2 groups x 64 rows x 8 chunks = 1024 chunks
          const int idx = tid + q * NTH;
          const int g = idx >> 9, rr = (idx >> 3) & 63, ch = idx & 7;
          const int gr = (g == 0 ? rowi : rowj) + rr;
          cp_async16(&S.B[g][rr][ch * 4], A + int64_t(gr) * ld + k0 + ch * 4, gr <= n);
        }
        const int ty = tid >> 4, tx = tid & 15;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int gi = rowi + ty * 4 + i, gj = rowj + tx * 4;
          cacc[i] = (gi <= n && gj < ld) ? __ldcg(reinterpret_cast<const float4*>(A + int64_t(gi) * ld + gj))
                                         : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      };
      bool first = true;
      for (int t = int(blockIdx.x) - 1; t < ntiles; t += nworkers) {
        issue_tile(t);
        cp_async_wait_all();
        __syncthreads();
        if (first) TRACE(1, 1);
        const int rowi = r0 + ti * TT, rowj = r0 + tj * TT;
        // ---- forward substitution, one thread per panel row: x L^T = b ---------------------------
        if (tid < 2 * TT) {
          const int g = tid >> 6, rr = tid & 63;
          float b[NB];
          load_row(b, &S.B[g][rr][0]);
          solve_row(b, S.LT, S.rinv);
#pragma unroll
          for (int j = 0; j < NB; ++j) S.P[g][j][rr] = b[j];
        }
        __syncthreads();
        if (first) TRACE(1, 2);
        // ---- publish the factor's panel rows: diagonal tile owner; the rhs-only row group: tj == 0
        if ((tj == ti) || (ti >= ntj && tj == 0)) {
          for (int idx = tid; idx < TT * NB; idx += NTH) {
            const int rr = idx >> 5, c = idx & 31;       // coalesced rows of Lm
            const int gr = rowi + rr;
            if (gr <= n) Lm[int64_t(gr) * ld + k0 + c] = S.P[0][c][rr];
          }
        }
        {  // ---- C -= X_i X_j^T in registers, write back ------------------------------------------
          const int ty = tid >> 4, tx = tid & 15;
          float acc[4][4];
#pragma unroll
          for (int i = 0; i < 4; ++i) { acc[i][0] = cacc[i].x; acc[i][1] = cacc[i].y; acc[i][2] = cacc[i].z; acc[i][3] = cacc[i].w; }
#pragma unroll 8
          for (int c = 0; c < NB; ++c) {
            const float4 av = *reinterpret_cast<const float4*>(&S.P[0][c][ty * 4]);
            const float4 bv = *reinterpret_cast<const float4*>(&S.P[1][c][tx * 4]);
            const float a4[4] = {av.x, av.y, av.z, av.w}, b4[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              fms2(acc[i][0], acc[i][1], a4[i], b4[0], b4[1]);
              fms2(acc[i][2], acc[i][3], a4[i], b4[2], b4[3]);
            }
          }
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int gi = rowi + ty * 4 + i, gj = rowj + tx * 4;
            // columns past the diagonal of a diagonal tile hold symmetric values nobody reads
            if (gi <= n && gj < ld && gj <= gi)
              __stcg(reinterpret_cast<float4*>(A + int64_t(gi) * ld + gj), make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]));
          }
        }
        __syncthreads();
        if (first) TRACE(1, 3);
        first = false;
      }
    }
    if (kb + 1 < nkb) {
      target += gridDim.x;
      grid_barrier(bar, target);
    }
    TRACE(1, 4);
  }
}

// ---- backward solve L^T x = y (y = row n of Lm), one CTA ------------------------------------------
constexpr int BSN = 512;
constexpr int BCPT = 2;                  // columns per thread held in registers per pass

__global__ void __launch_bounds__(BSN)
k_backsolve(const float* __restrict__ Lm, const float* __restrict__ Ldiag, int ld, int n, int* fail,
            unsigned* bar, float* __restrict__ dx, int mode) {
  extern __shared__ __align__(16) float bsm[];
  float* sLi = bsm;                      // [2][32][32] L_kk^-1 of the current / next block
  float* xs = bsm + 2 * NB * NB;         // [n + 32] running right-hand side -> solution
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nkb = (n + NB - 1) / NB;
  for (int i = tid; i < n + NB; i += BSN) xs[i] = (i < n) ? __ldcg(Lm + int64_t(n) * ld + i) : 0.f;
  if (tid < NB * NB / 4)
    cp_async16(sLi + ((nkb - 1) & 1) * NB * NB + tid * 4, Ldiag + int64_t(nkb - 1) * LD_STRIDE + LD_LINV + tid * 4, true);
  for (int kb = nkb - 1; kb >= 0; --kb) {
    const int k0 = kb * NB, nb = min(NB, n - k0);
    // rows k0..k0+nb-1 of L at the thread's columns c < k0: in flight while warp 0 solves the block
    float lr[BCPT][NB];
    const int npass = (k0 + BSN * BCPT - 1) / (BSN * BCPT);
    auto load_rows = [&](int pass) {
#pragma unroll
      for (int q = 0; q < BCPT; ++q) {
        const int c = (pass * BCPT + q) * BSN + tid;
#pragma unroll
        for (int r = 0; r < NB; ++r) lr[q][r] = (c < k0 && r < nb) ? __ldcg(Lm + int64_t(k0 + r) * ld + c) : 0.f;
      }
    };
    if (npass > 0) load_rows(0);
    cp_async_wait_all();
    __syncthreads();                     // L_kk^-1 landed; xs of the previous step complete
    if (kb > 0 && tid < NB * NB / 4)
      cp_async16(sLi + ((kb - 1) & 1) * NB * NB + tid * 4, Ldiag + int64_t(kb - 1) * LD_STRIDE + LD_LINV + tid * 4, true);
    if (warp == 0) {
      // x_c = sum_r Linv[r][c] y_r   (L^-T y; Linv is lower triangular, identity padded)
      const float* Li = sLi + (kb & 1) * NB * NB;
      float x = 0.f;
#pragma unroll
      for (int r = 0; r < NB; ++r) x = fmaf(Li[r * NB + lane], xs[k0 + r], x);
      __syncwarp();
      if (lane < nb) xs[k0 + lane] = x;
    }
    __syncthreads();
    for (int pass = 0; pass < npass; ++pass) {
      if (pass > 0) load_rows(pass);
#pragma unroll
      for (int q = 0; q < BCPT; ++q) {
        const int c = (pass * BCPT + q) * BSN + tid;
        if (c < k0) {
          float s = 0.f;
#pragma unroll
          for (int r = 0; r < NB; ++r) s = fmaf(lr[q][r], xs[k0 + r], s);
          xs[c] -= s;
        }
      }
    }
  }
  __syncthreads();
  const int failed = *reinterpret_cast<volatile int*>(fail);
  __syncthreads();
  // mode bit 1: dx += x (the correction of an iterative-refinement pass); bit 0: last solve of the iteration
  for (int i = tid; i < n; i += BSN) dx[i] = failed ? 0.f : ((mode & 2) ? dx[i] + xs[i] : xs[i]);
  if (tid == 0 && (mode & 1)) { *fail = 0; *bar = 0u; }
}


// ---- backward solve, shared-memory variant (n <= ~830): the 32-row slab of L that step k multiplies
// with x_k does not depend on x, so the slab of step k-1 streams into the other half of a double
// buffer (cp.async) while step k computes: no L2 latency on the chain.
constexpr int BSW_MAX = 832;             // widest slab (columns) that fits twice in shared memory

__host__ __device__ inline size_t backsolve_smem_floats(int n, int W) {
  return size_t(2 * NB * NB) + size_t(n + NB) + size_t(2) * NB * W;
}

__global__ void __launch_bounds__(BSN)
k_backsolve_smem(const float* __restrict__ Lm, const float* __restrict__ Ldiag, int ld, int n, int W, int* fail,
                 unsigned* bar, float* __restrict__ dx, int mode) {
  extern __shared__ __align__(16) float bsm[];
  float* sLi = bsm;                      // [2][32][32]
  float* slab = bsm + 2 * NB * NB;       // [2][32][W]
  float* xs = slab + 2 * NB * W;         // [n + 32]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nkb = (n + NB - 1) / NB;
  auto prefetch = [&](int kb) {          // slab rows k0..k0+31 (columns < k0) and L_kk^-1 of block kb
    const int k0 = kb * NB, nb = min(NB, n - k0);
    float* dst = slab + (kb & 1) * NB * W;
    // thread = (row tid / 16, chunk lane tid % 16): no index division in the copy loop -- this kernel is bound by the
    // instructions its one CTA issues (ncu: issue slots 52 % busy, 45 % of the instructions were this loop's)
    const int cpr = k0 >> 2;             // 16-byte chunks per row
    const int r = tid >> 4;
    const float* src = Lm + int64_t(k0 + r) * ld;
    float* drow = dst + r * W;
    const bool rv = r < nb;
    for (int ch = tid & 15; ch < cpr; ch += 16) cp_async16(drow + ch * 4, src + ch * 4, rv);
    if (tid < NB * NB / 4)
      cp_async16(sLi + (kb & 1) * NB * NB + tid * 4, Ldiag + int64_t(kb) * LD_STRIDE + LD_LINV + tid * 4, true);
    asm volatile("cp.async.commit_group;\n" ::: "memory");
  };
  prefetch(nkb - 1);
  for (int i = tid; i < n + NB; i += BSN) xs[i] = (i < n) ? __ldcg(Lm + int64_t(n) * ld + i) : 0.f;
  for (int kb = nkb - 1; kb >= 0; --kb) {
    const int k0 = kb * NB, nb = min(NB, n - k0);
    if (kb > 0) {
      prefetch(kb - 1);
      asm volatile("cp.async.wait_group 1;\n" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    }
    __syncthreads();                     // slab / L_kk^-1 of step kb landed; xs of the previous step complete
    if (warp == 0) {
      const float* Li = sLi + (kb & 1) * NB * NB;
      float x = 0.f;
#pragma unroll
      for (int r = 0; r < NB; ++r) x = fmaf(Li[r * NB + lane], xs[k0 + r], x);
      __syncwarp();
      if (lane < nb) xs[k0 + lane] = x;
    }
    __syncthreads();
    const float* sl = slab + (kb & 1) * NB * W;
    float xk[NB];
#pragma unroll
    for (int r4 = 0; r4 < NB; r4 += 4) {
      const float4 v = *reinterpret_cast<const float4*>(xs + k0 + r4);
      xk[r4] = v.x; xk[r4 + 1] = v.y; xk[r4 + 2] = v.z; xk[r4 + 3] = v.w;
    }
    for (int c = tid; c < k0; c += BSN) {
      float s0 = 0.f, s1 = 0.f;
#pragma unroll
      for (int r = 0; r < NB; r += 2) {
        s0 = fmaf(sl[r * W + c], xk[r], s0);
        s1 = fmaf(sl[(r + 1) * W + c], xk[r + 1], s1);
      }
      xs[c] -= s0 + s1;
    }
    __syncthreads();                     // slab (kb & 1) is free for the prefetch of step kb - 2
  }
  __syncthreads();
  const int failed = *reinterpret_cast<volatile int*>(fail);
  __syncthreads();
  // mode bit 1: dx += x (the correction of an iterative-refinement pass); bit 0: last solve of the iteration
  for (int i = tid; i < n; i += BSN) dx[i] = failed ? 0.f : ((mode & 2) ? dx[i] + xs[i] : xs[i]);
  if (tid == 0 && (mode & 1)) { *fail = 0; *bar = 0u; }
}


// ---- small systems (n <= SMALL_N: the frontend's sliding window): the whole factorisation and both
// solves in ONE CTA, the matrix resident in shared memory, __syncthreads instead of grid barriers.  Same
// building blocks as above: register potrf of the 32x32 diagonal block in one warp, one thread per panel
// row for the triangular solve, L_kk^-1 (a second warp, off the chain) for the backward solve.
constexpr int SMALL_N = 96;
constexpr int SMALL_T = 256;

__global__ void __launch_bounds__(SMALL_T)
k_chol_small(const float* __restrict__ A, int ldA, int n, float* __restrict__ dx, int* __restrict__ fail) {
  __shared__ CholSmem S;                       // LT / rinv of the current block; P[0] = solved panel (k-major)
  extern __shared__ __align__(16) float csm[];
  const int ld = ((n + 3) & ~3) + 4;           // 16-byte aligned rows
  float* sA = csm;                             // [(n+1)][ld] lower triangle + rhs row n
  float* sLi = sA + (n + 1) * ld;              // [nkb][32][32] inverses of the diagonal blocks
  float* xs = sLi + 3 * NB * NB;               // [n + 32] solution
  __shared__ int sbad;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nkb = (n + NB - 1) / NB;
  if (tid == 0) sbad = 0;
  {  // lower triangle + rhs row, every thread's loads independent (one L2 round trip, not one per row)
    const int tot = (n + 1) * ld;
#pragma unroll 4
    for (int e = tid; e < tot; e += SMALL_T) {
      const int i = e / ld, k = e - i * ld;
      sA[e] = (k < n && (k <= i || i == n)) ? __ldcg(A + int64_t(i) * ldA + k) : 0.f;
    }
  }
  __syncthreads();
  for (int kb = 0; kb < nkb; ++kb) {
    const int k0 = kb * NB, nb = min(NB, n - k0), r0 = k0 + nb;
    if (warp == 0) {                           // ---- diagonal block
      float a[NB];
#pragma unroll
      for (int c = 0; c < NB; ++c) a[c] = (lane < nb && c < nb) ? sA[(k0 + lane) * ld + k0 + c] : ((c == lane) ? 1.f : 0.f);
      const bool bad = potrf32_warp(a, lane, &S.P[1][0][0]);
#pragma unroll
      for (int c = 0; c < NB; ++c) {
        const float v = (c <= lane) ? a[c] : 0.f;
        S.LT[c][lane] = v;
        if (lane < nb && c < nb && c <= lane) sA[(k0 + lane) * ld + k0 + c] = v;
      }
      __syncwarp();
      S.rinv[lane] = 1.0f / S.LT[lane][lane];
      if (bad && lane == 0) sbad = 1;
    }
    __syncthreads();
    const int nrow = n + 1 - r0;               // rows below the block, the rhs row included
    if (tid < nrow) {                          // ---- panel rows: x L^T = b, one thread per row (<= 65 rows)
      float b[NB];
      load_row(b, sA + (r0 + tid) * ld + k0);
      solve_row(b, S.LT, S.rinv);
#pragma unroll
      for (int j = 0; j < NB; ++j) {               // columns past n (last, partial block) are padding
        S.P[0][j][tid] = (j < nb) ? b[j] : 0.f;
        if (j < nb) sA[(r0 + tid) * ld + k0 + j] = b[j];
      }
    } else if (warp == SMALL_T / 32 - 1) {     // ---- L_kk^-1, column `lane` solves L v = e_lane (off the chain)
      float b[NB];
#pragma unroll
      for (int c = 0; c < NB; ++c) b[c] = (c == lane) ? 1.f : 0.f;
      solve_row(b, S.LT, S.rinv);
#pragma unroll
      for (int j = 0; j < NB; ++j) sLi[kb * NB * NB + j * NB + lane] = b[j];
    }
    __syncthreads();
    // ---- trailing update: rows i >= r0 (rhs row included), columns r0 <= c <= min(i, n-1)
    for (int e = tid; e < nrow * (n - r0); e += SMALL_T) {
      const int ri = e / (n - r0), rc = e - ri * (n - r0);
      if (rc <= ri) {
        float acc = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;      // four chains: this single-CTA kernel is latency-bound
#pragma unroll
        for (int m = 0; m < NB; m += 4) {
          acc = fmaf(S.P[0][m][ri], S.P[0][m][rc], acc);
          acc1 = fmaf(S.P[0][m + 1][ri], S.P[0][m + 1][rc], acc1);
          acc2 = fmaf(S.P[0][m + 2][ri], S.P[0][m + 2][rc], acc2);
          acc3 = fmaf(S.P[0][m + 3][ri], S.P[0][m + 3][rc], acc3);
        }
        sA[(r0 + ri) * ld + r0 + rc] -= (acc + acc1) + (acc2 + acc3);
      }
    }
    __syncthreads();
  }
  // ---- backward solve L^T x = y (y = row n), block by block, one warp
  if (warp == 0) {
    for (int i = lane; i < n + NB; i += 32) xs[i] = (i < n) ? sA[n * ld + i] : 0.f;
    __syncwarp();
    for (int kb = nkb - 1; kb >= 0; --kb) {
      const int k0 = kb * NB, nb = min(NB, n - k0), r0 = k0 + nb;
      float t = (lane < nb) ? xs[k0 + lane] : 0.f;
      if (lane < nb)
        for (int r = r0; r < n; ++r) t = fmaf(-sA[r * ld + k0 + lane], xs[r], t);   // column k0+lane of L below the block
      __syncwarp();
      xs[k0 + lane] = (lane < nb) ? t : 0.f;
      __syncwarp();
      const float* Li = sLi + kb * NB * NB;
      float x = 0.f, x1 = 0.f, x2 = 0.f, x3 = 0.f;
#pragma unroll
      for (int r = 0; r < NB; r += 4) {
        x = fmaf(Li[r * NB + lane], xs[k0 + r], x);
        x1 = fmaf(Li[(r + 1) * NB + lane], xs[k0 + r + 1], x1);
        x2 = fmaf(Li[(r + 2) * NB + lane], xs[k0 + r + 2], x2);
        x3 = fmaf(Li[(r + 3) * NB + lane], xs[k0 + r + 3], x3);
      }
      x = (x + x1) + (x2 + x3);
      __syncwarp();
      if (lane < nb) xs[k0 + lane] = x;
      __syncwarp();
    }
    const bool bad = sbad != 0;
    for (int i = lane; i < n; i += 32) dx[i] = bad ? 0.f : xs[i];      // chol.py:14-18: failed factorisation -> zeros
    if (lane == 0) *fail = 0;
  }
}

}  // namespace

#include "dba_chol_tc.cuh"

// Factorisation + forward solve of the rhs row with 128-column panels (dba_chol_tc.cuh), then the backward solve.
// 3 launches per panel; all on `st`.
cudaError_t launch_cholesky_tc(const Workspace& ws, int n, cudaStream_t st, bool finalize) {
  static bool attr_done = false;
  const size_t smem_p = potrf128_smem_floats() * sizeof(float);
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(k_potrf128, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem_p));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_syrk_tc<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(SY_SMEM));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_syrk_tc<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(SY_SMEM));
    if (e != cudaSuccess) return e;
    attr_done = true;
  }
  float* LinvW = ws.Ldiag + size_t((n + NB - 1) / NB) * LD_STRIDE;      // [128][128] behind the per-block entries
  for (int k0 = 0; k0 < n; k0 += PW) {
    const int w = (n - k0 < PW) ? n - k0 : PW, r0 = k0 + w;
    k_potrf128<<<1, 256, smem_p, st>>>(ws.A, ws.Lm, ws.Ldiag, LinvW, ws.ldA, n, k0, ws.fail);
    k_syrk_tc<1><<<(n + 1 - r0 + PW - 1) / PW, SY_T, SY_SMEM, st>>>(ws.A, ws.Lm, LinvW, ws.ldA, n, k0);
    g_launch_count += 2;
    if (r0 < n) {
      const int ntj = (n - r0 + PW - 1) / PW, nti = (n + 1 - r0 + PW - 1) / PW;
      const int ntiles = ntj * (ntj + 1) / 2 + (nti > ntj ? ntj : 0);
      k_syrk_tc<0><<<ntiles, SY_T, SY_SMEM, st>>>(ws.A, ws.Lm, LinvW, ws.ldA, n, k0);
      g_launch_count++;
    }
  }
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return err;
  return launch_backsolve(ws, n, finalize ? 1 : 0, st);
}

int cholesky_tc_min_n() { return 1024; }

#ifdef DBA_CHOL_TRACE
void chol_trace_read(unsigned long long* out, int n) { cudaMemcpyFromSymbol(out, g_trace, sizeof(unsigned long long) * n); }
#endif

cudaError_t launch_cholesky_coop(const Workspace& ws, int n, int max_ctas, cudaStream_t st, bool finalize) {
  const int r0 = (n < NB) ? n : NB;
  const int ntj = (n - r0 + TT - 1) / TT, nti = (n + 1 - r0 + TT - 1) / TT;
  const int ntiles = ntj * (ntj + 1) / 2 + (nti > ntj ? ntj : 0);
  int grid = ntiles + 1;                       // CTA 0 is the look-ahead (diagonal) CTA
  if (grid > max_ctas) grid = max_ctas;
  float* A = ws.A; float* Lm = ws.Lm; float* Ldiag = ws.Ldiag; int ld = ws.ldA; int* fail = ws.fail;
  unsigned* bar = reinterpret_cast<unsigned*>(ws.fail + 1);
  float* dx = ws.dx;
  void* args[] = {&A, &Lm, &Ldiag, &ld, &n, &fail, &bar};
  void* kern = (n >= 1024) ? reinterpret_cast<void*>(k_chol_fused<2>) : reinterpret_cast<void*>(k_chol_fused<1>);
  cudaError_t err = cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(NTH), args, 0, st);
  g_launch_count++;
  if (err != cudaSuccess) return err;
  return launch_backsolve(ws, n, finalize ? 1 : 0, st);
}

// L^T x = y with y = row n of Lm; mode: bit 0 = last solve of the iteration (resets the flags), bit 1 = dx += x
cudaError_t launch_backsolve(const Workspace& ws, int n, int mode, cudaStream_t st) {
  const float* Lm = ws.Lm; const float* Ldiag = ws.Ldiag; int ld = ws.ldA; int* fail = ws.fail;
  unsigned* bar = reinterpret_cast<unsigned*>(ws.fail + 1);
  float* dx = ws.dx;
  {  // the factor resident in a cluster's shared memory where it fits (n <= 768): 28 -> ~9 us at n = 756
    const cudaError_t ce = launch_trisolve_cluster(ws, n, false, mode, st);
    if (ce == cudaSuccess) return ce;
    if (ce != cudaErrorNotSupported) return ce;
  }
  const int nkb = (n + NB - 1) / NB;
  const int W = (nkb - 1) * NB > 0 ? (nkb - 1) * NB : 4;
  cudaError_t err;
  if (W <= BSW_MAX) {
    const size_t smem = backsolve_smem_floats(n, W) * sizeof(float);
    err = cudaFuncSetAttribute(k_backsolve_smem, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (err != cudaSuccess) return err;
    k_backsolve_smem<<<1, BSN, smem, st>>>(Lm, Ldiag, ld, n, W, fail, bar, dx, mode);
  } else {
    const size_t smem = size_t(2 * NB * NB + n + NB) * sizeof(float);
    k_backsolve<<<1, BSN, smem, st>>>(Lm, Ldiag, ld, n, fail, bar, dx, mode);
  }
  g_launch_count++;
  return cudaGetLastError();
}

int cholesky_small_max_n() { return SMALL_N; }

cudaError_t launch_cholesky_small(const Workspace& ws, int n, cudaStream_t st) {
  const int ld = ((n + 3) & ~3) + 4;
  const size_t smem = (size_t(n + 1) * ld + 3 * NB * NB + n + NB) * sizeof(float);
  cudaError_t err = cudaFuncSetAttribute(k_chol_small, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
  if (err != cudaSuccess) return err;
  k_chol_small<<<1, SMALL_T, smem, st>>>(ws.A, ws.ldA, n, ws.dx, ws.fail);
  g_launch_count++;
  return cudaGetLastError();
}

size_t cholesky_ldiag_floats(int n) { return size_t((n + NB - 1) / NB) * LD_STRIDE + size_t(PW) * PW; }
void cholesky_ldiag_layout(int* stride, int* linv_offset) { *stride = LD_STRIDE; *linv_offset = LD_LINV; }

int cholesky_coop_max_ctas(int n) {
  int dev = 0, sms = 0, per = 0, coop = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
  if (!coop || n > 9000) return 0;      // k_backsolve keeps x in shared memory
  const cudaError_t oe = (n >= 1024) ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_chol_fused<2>, NTH, 0)
                                     : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_chol_fused<1>, NTH, 0);
  if (oe != cudaSuccess || per < 1) return 0;
  return sms * (per > 2 ? 2 : per);
}

}  // namespace dba
