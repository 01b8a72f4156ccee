// Single-launch Cholesky + solve of the reduced pose system: one cooperative (co-resident) grid,
// ONE grid barrier per 32-column block step, the backward solve in the same kernel.
//
// Replaces torch.linalg.cholesky + cholesky_solve of the reference (hislam2/geom/chol.py:5-20,66).
// Storage: A (n+1) x ld row-major holds the running Schur complement (lower) with the rhs as row n;
// the factor goes to a SEPARATE matrix Lm (same shape) so that the pre-solve panel entries stay
// readable by every CTA (no intra-step race), inverse diagonal blocks to Linv [nkb,32,32].
//
// Step k (right-looking, redundancy instead of a second barrier):
//   every CTA with work: warp 0 factors the 32x32 diagonal block in registers (shuffles, no block
//   barrier) and inverts it; the panel rows a CTA's 64x64 trailing tile needs are solved as a small
//   GEMM X = B * Linv^T by all threads (redundantly per tile), then the tile is updated
//   C -= X_i X_j^T.  The owner of a diagonal tile publishes its panel rows into Lm.
// A non-positive / NaN pivot raises the flag -> dx = 0 (chol.py:14-18 behaviour).
#include <cooperative_groups.h>

#include "dba_kernels.h"

namespace cg = cooperative_groups;

namespace dba {

namespace {

constexpr int NB = 32;
constexpr int TT = 64;
constexpr int NTH = 256;
constexpr int kCoopMaxN = 2 * TT * (NB + 1) + 2 * NB * (TT + 4);   // xs aliases the staging buffers

__device__ __forceinline__ void tri_decode(int t, int& ti, int& tj) {
  int a = int((sqrtf(8.0f * float(t) + 1.0f) - 1.0f) * 0.5f);
  while ((a + 1) * (a + 2) / 2 <= t) ++a;
  while (a * (a + 1) / 2 > t) --a;
  ti = a;
  tj = t - a * (a + 1) / 2;
}

// warp-level: factor the diagonal block at (k0,k0) into sL (lower, identity padded) and its inverse
// into sLi.  Returns true when a pivot was not positive.
__device__ __forceinline__ bool potrf_inv_warp(const float* __restrict__ A, int ld, int n, int k0, int nb, int lane,
                                               float (*sL)[NB + 1], float (*sLi)[NB + 1]) {
  float a[NB];
#pragma unroll
  for (int c = 0; c < NB; ++c) {
    float v = (c == lane) ? 1.f : 0.f;
    if (lane < nb && c < nb && c <= lane) v = __ldcg(A + int64_t(k0 + lane) * ld + k0 + c);
    a[c] = v;
  }
  bool bad = false;
#pragma unroll
  for (int j = 0; j < NB; ++j) {
    float ajj = __shfl_sync(0xffffffffu, a[j], j);
    if (!(ajj > 0.f)) { bad = true; ajj = 1.f; }
    const float djj = sqrtf(ajj);
    float lij = a[j] / djj;
    if (lane == j) lij = djj;
    if (lane < j) lij = 0.f;
    a[j] = lij;
#pragma unroll
    for (int c = j + 1; c < NB; ++c) {
      const float lcj = __shfl_sync(0xffffffffu, lij, c);
      if (lane >= c) a[c] -= lij * lcj;
    }
  }
#pragma unroll
  for (int c = 0; c < NB; ++c) sL[lane][c] = a[c];
  __syncwarp();
  // inverse: lane = column c of Linv; forward substitution L x = e_c
  float x[NB];
#pragma unroll
  for (int j = 0; j < NB; ++j) {
    float s = (j == lane) ? 1.f : 0.f;
#pragma unroll
    for (int i = 0; i < j; ++i) s -= sL[j][i] * x[i];
    x[j] = s / sL[j][j];
  }
#pragma unroll
  for (int j = 0; j < NB; ++j) sLi[j][lane] = (j >= lane) ? x[j] : 0.f;
  __syncwarp();
  return bad;
}

__global__ void __launch_bounds__(NTH)
k_chol_coop(float* A, float* Lm, int ld, int n, float* Linv, int* fail, float* __restrict__ dx) {
  cg::grid_group grid = cg::this_grid();
  __shared__ float sL[NB][NB + 1];
  __shared__ float sLi[NB][NB + 1];
  __shared__ __align__(16) float sBP[2 * TT * (NB + 1) + 2 * NB * (TT + 4)];
  float (*sB)[TT][NB + 1] = reinterpret_cast<float (*)[TT][NB + 1]>(sBP);
  float (*sP)[NB][TT + 4] = reinterpret_cast<float (*)[NB][TT + 4]>(sBP + 2 * TT * (NB + 1));
  float* xs = sBP;                  // [n] for the backward solve (CTA 0), aliases the tile staging
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nkb = (n + NB - 1) / NB;

  for (int kb = 0; kb < nkb; ++kb) {
    const int k0 = kb * NB, nb = min(NB, n - k0), r0 = k0 + nb;
    const int ntj = (n - r0 + TT - 1) / TT;
    const int nti = (n + 1 - r0 + TT - 1) / TT;
    const int ntiles = ntj * (ntj + 1) / 2 + (nti > ntj ? ntj : 0);
    const bool work = (int(blockIdx.x) < ntiles) || blockIdx.x == 0;
    if (work) {
      if (warp == 0) {
        const bool bad = potrf_inv_warp(A, ld, n, k0, nb, lane, sL, sLi);
        if (blockIdx.x == 0) {
          float* dst = Linv + int64_t(kb) * NB * NB;
#pragma unroll
          for (int j = 0; j < NB; ++j) dst[j * NB + lane] = sLi[j][lane];
          if (bad && lane == 0) *fail = 1;
        }
      }
      __syncthreads();
      if (ntj == 0) {
        // last block: only the rhs row is left below the diagonal block
        if (blockIdx.x == 0 && tid < NB) {
          float s = 0.f;
          for (int c = 0; c < nb; ++c) s += __ldcg(A + int64_t(n) * ld + k0 + c) * sLi[tid][c];
          if (tid < nb) Lm[int64_t(n) * ld + k0 + tid] = s;
        }
      }
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
        int ti, tj;
        tri_decode(t, ti, tj);
        const int rowi = r0 + ti * TT, rowj = r0 + tj * TT;
        // stage the pre-solve panel rows of both row groups
        for (int idx = tid; idx < 2 * TT * NB; idx += NTH) {
          const int g = idx / (TT * NB), rem = idx % (TT * NB), rr = rem / NB, c = rem % NB;
          const int gr = (g == 0 ? rowi : rowj) + rr;
          sB[g][rr][c] = (gr <= n && c < nb) ? __ldcg(A + int64_t(gr) * ld + k0 + c) : 0.f;
        }
        __syncthreads();
        {  // X = B Linv^T : thread -> (group, row, 16 columns)
          const int g = tid >> 7, u = tid & 127, rr = u & 63, jh = u >> 6;
          float b[NB];
#pragma unroll
          for (int c = 0; c < NB; ++c) b[c] = sB[g][rr][c];
#pragma unroll
          for (int jj = 0; jj < 16; ++jj) {
            const int j = jh * 16 + jj;
            float s = 0.f;
#pragma unroll
            for (int c = 0; c < NB; ++c) s += b[c] * sLi[j][c];     // Linv is lower: zeros above
            sP[g][j][rr] = s;
          }
        }
        __syncthreads();
        // publish the factor's panel rows: diagonal tile owner; the rhs-only row group has none
        const bool writer = (tj == ti) || (ti >= ntj && tj == 0);
        if (writer) {
          for (int idx = tid; idx < TT * NB; idx += NTH) {
            const int rr = idx / NB, c = idx % NB;
            const int gr = rowi + rr;
            if (gr <= n && c < nb) Lm[int64_t(gr) * ld + k0 + c] = sP[0][c][rr];
          }
        }
        {  // C -= X_i X_j^T
          const int ty = tid / 16, tx = tid % 16;
          float acc[4][4];
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
#pragma unroll 8
          for (int c = 0; c < NB; ++c) {
            const float4 av = *reinterpret_cast<const float4*>(&sP[0][c][ty * 4]);
            const float4 bv = *reinterpret_cast<const float4*>(&sP[1][c][tx * 4]);
            const float a4[4] = {av.x, av.y, av.z, av.w}, b4[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
              for (int j = 0; j < 4; ++j) acc[i][j] += a4[i] * b4[j];
          }
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int gi = rowi + ty * 4 + i;
            if (gi > n) continue;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int gj = rowj + tx * 4 + j;
              if (gj < n && gj <= gi) { float* pc = A + int64_t(gi) * ld + gj; *pc = __ldcg(pc) - acc[i][j]; }
            }
          }
        }
        __syncthreads();
      }
    }
    __threadfence();
    grid.sync();
  }

  if (blockIdx.x != 0) return;
  // ---- backward solve L^T x = y on CTA 0 (y = row n of Lm) --------------------------------------
  for (int i = tid; i < n; i += NTH) xs[i] = __ldcg(Lm + int64_t(n) * ld + i);
  __syncthreads();
  for (int kb = nkb - 1; kb >= 0; --kb) {
    const int k0 = kb * NB, nb = min(NB, n - k0);
    if (warp == 0) {
      const float* Li = Linv + int64_t(kb) * NB * NB;
      const float r = (lane < nb) ? xs[k0 + lane] : 0.f;
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < NB; ++i) s += __ldcg(Li + i * NB + lane) * __shfl_sync(0xffffffffu, r, i);   // (Linv^T r)_lane
      if (lane < nb) xs[k0 + lane] = s;
    }
    __syncthreads();
    for (int c = tid; c < k0; c += NTH) {
      float s = 0.f;
#pragma unroll 8
      for (int r = 0; r < nb; ++r) s += __ldcg(Lm + int64_t(k0 + r) * ld + c) * xs[k0 + r];
      xs[c] -= s;
    }
    __syncthreads();
  }
  const int failed = *reinterpret_cast<volatile int*>(fail);
  __syncthreads();
  for (int i = tid; i < n; i += NTH) dx[i] = failed ? 0.f : xs[i];
  if (tid == 0) *fail = 0;
}

}  // namespace

// Returns cudaErrorNotSupported-like failure through the normal error path; the caller falls back to the
// multi-launch path when cooperative launch is unavailable.
cudaError_t launch_cholesky_coop(const Workspace& ws, int n, int max_ctas, cudaStream_t st) {
  const int r0 = (n < NB) ? n : NB;
  const int ntj = (n - r0 + TT - 1) / TT, nti = (n + 1 - r0 + TT - 1) / TT;
  int ntiles = ntj * (ntj + 1) / 2 + (nti > ntj ? ntj : 0);
  int grid = ntiles < 1 ? 1 : ntiles;
  if (grid > max_ctas) grid = max_ctas;
  const size_t smem = 0;
  float* A = ws.A; float* Lm = ws.Lm; int ld = ws.ldA; float* Linv = ws.Ldiag; int* fail = ws.fail; float* dx = ws.dx;
  void* args[] = {&A, &Lm, &ld, &n, &Linv, &fail, &dx};
  cudaError_t err = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(k_chol_coop), dim3(grid), dim3(NTH), args, smem, st);
  g_launch_count++;
  return err;
}

int cholesky_coop_max_ctas(int n) {
  int dev = 0, sms = 0, per = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const size_t smem = 0;
  if (n > kCoopMaxN) return 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_chol_coop, NTH, smem) != cudaSuccess) return 0;
  int coop = 0;
  cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
  if (!coop) return 0;
  return sms * (per > 2 ? 2 : per);
}

}  // namespace dba
