// Dense blocked Cholesky of the reduced pose system on the device, with the right-hand side
// carried as an extra row (so the forward solve falls out of the factorisation), a backward solve
// and the failure flag.  Replaces the reference's torch.linalg.cholesky + cholesky_solve
// (hislam2/geom/chol.py:5-20,66): a non-positive / NaN pivot yields dx = 0 like the reference's
// blanket `except` (chol.py:14-18).
//
// Storage: A is (n+1) x ld row-major, lower triangle of S in rows 0..n-1, rhs in row n.
// Right-looking, NB = 32:   for every diagonal block
//   k_potrf_trsm   every CTA refactors the 32x32 diagonal block in one warp (registers + shuffles,
//                  no barriers) and solves its 128 rows of the panel; CTA 0 publishes the factor
//                  into Ldiag (A's diagonal blocks keep their pre-factor values)
//   k_trailing     64x64 tiles of the trailing lower triangle: C -= P_i P_j^T
#include "dba_kernels.h"

namespace dba {

namespace {

constexpr int NB = 32;
constexpr int TR = 128;   // rows per CTA in the panel solve

__global__ void __launch_bounds__(TR)
k_potrf_trsm(float* __restrict__ A, int ld, int n, int k0, float* __restrict__ Ldiag_kb, int* __restrict__ fail) {
  __shared__ float sL[NB][NB + 1];
  __shared__ float sInv[NB];
  const int tid = threadIdx.x;
  const int nb = min(NB, n - k0);
  if (tid < 32) {
    const int lane = tid;
    float a[NB];
    // row `lane` of the diagonal block (lower part), identity padding beyond nb
#pragma unroll
    for (int c = 0; c < NB; ++c) {
      float v = (c == lane) ? 1.f : 0.f;
      if (lane < nb && c < nb && c <= lane) v = A[int64_t(k0 + lane) * ld + k0 + c];
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
    __syncwarp();   // a[lane] would need a dynamic register index: read the diagonal back from smem
    sInv[lane] = 1.0f / sL[lane][lane];
    if (blockIdx.x == 0) {
#pragma unroll
      for (int c = 0; c < NB; ++c) Ldiag_kb[lane * NB + c] = (c <= lane) ? a[c] : 0.f;
      if (bad && lane == 0) *fail = 1;
    }
  }
  __syncthreads();
  const int row = k0 + nb + blockIdx.x * TR + tid;
  if (row > n) return;      // rows k0+nb .. n (row n = rhs)
  float* ar = A + int64_t(row) * ld + k0;
  float x[NB];
#pragma unroll
  for (int c4 = 0; c4 < NB; c4 += 4) {
    // k0 is a multiple of 32 and ld of 4: 16-byte aligned; columns beyond nb are never used
    if (c4 < nb) {
      float4 v = *reinterpret_cast<const float4*>(ar + c4);
      x[c4] = v.x; x[c4 + 1] = v.y; x[c4 + 2] = v.z; x[c4 + 3] = v.w;
    } else {
      x[c4] = x[c4 + 1] = x[c4 + 2] = x[c4 + 3] = 0.f;
    }
  }
#pragma unroll
  for (int j = 0; j < NB; ++j) {
    float s = x[j];
#pragma unroll
    for (int c = 0; c < j; ++c) s -= x[c] * sL[j][c];
    x[j] = s * sInv[j];
  }
#pragma unroll
  for (int c4 = 0; c4 < NB; c4 += 4) {
    if (c4 + 3 < nb) {
      *reinterpret_cast<float4*>(ar + c4) = make_float4(x[c4], x[c4 + 1], x[c4 + 2], x[c4 + 3]);
    } else {
#pragma unroll
      for (int t = 0; t < 4; ++t)
        if (c4 + t < nb) ar[c4 + t] = x[c4 + t];
    }
  }
}

constexpr int TT = 64;   // trailing tile

__global__ void __launch_bounds__(256)
k_trailing(float* __restrict__ A, int ld, int n, int k0, int nb, int r0) {
  // tile (ti >= tj) of the trailing matrix starting at row/col r0 = k0 + nb
  __shared__ __align__(16) float sPi[NB][TT + 4];
  __shared__ __align__(16) float sPj[NB][TT + 4];
  int t = blockIdx.x;
  int ti = int((sqrtf(8.0f * float(t) + 1.0f) - 1.0f) * 0.5f);
  while ((ti + 1) * (ti + 2) / 2 <= t) ++ti;
  while (ti * (ti + 1) / 2 > t) --ti;
  const int tj = t - ti * (ti + 1) / 2;
  const int rowi = r0 + ti * TT, rowj = r0 + tj * TT;
  const int tid = threadIdx.x;
  // load the two 64 x nb panel pieces, transposed to [c][row]
  for (int idx = tid; idx < TT * NB; idx += 256) {
    const int rr = idx / NB, c = idx % NB;
    const int gi = rowi + rr, gj = rowj + rr;
    sPi[c][rr] = (gi <= n && c < nb) ? A[int64_t(gi) * ld + k0 + c] : 0.f;
    sPj[c][rr] = (gj < n && c < nb) ? A[int64_t(gj) * ld + k0 + c] : 0.f;
  }
  __syncthreads();
  const int ty = tid / 16, tx = tid % 16;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
#pragma unroll 8
  for (int c = 0; c < NB; ++c) {
    const float4 a = *reinterpret_cast<const float4*>(&sPi[c][ty * 4]);
    const float4 b = *reinterpret_cast<const float4*>(&sPj[c][tx * 4]);
    const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] += av[i] * bv[j];
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gi = rowi + ty * 4 + i;
    if (gi > n) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gj = rowj + tx * 4 + j;
      if (gj < n && gj <= gi) A[int64_t(gi) * ld + gj] -= acc[i][j];
    }
  }
}

// L^T x = y (y = row n of A after the factorisation), one CTA; writes dx (zeros on failure)
constexpr int BT = 1024;

__global__ void __launch_bounds__(BT)
k_backsolve(const float* __restrict__ A, int ld, int n, const float* __restrict__ Ldiag, int* __restrict__ fail,
            float* __restrict__ dx) {
  extern __shared__ float xs[];          // [n] solution, then [32][33] partials
  float* part = xs + ((n + 3) & ~3);
  const int tid = threadIdx.x, lane = tid & 31, grp = tid >> 5;
  const int nkb = (n + NB - 1) / NB;
  for (int kb = nkb - 1; kb >= 0; --kb) {
    const int k0 = kb * NB, nb = min(NB, n - k0);
    float s = 0.f;
    if (lane < nb)
      for (int i = k0 + nb + grp; i < n; i += 32) s += A[int64_t(i) * ld + k0 + lane] * xs[i];
    part[grp * 33 + lane] = s;
    __syncthreads();
    if (tid < 32) {
      float tot = 0.f;
      for (int g = 0; g < 32; ++g) tot += part[g * 33 + lane];
      float val = (lane < nb) ? A[int64_t(n) * ld + k0 + lane] - tot : 0.f;
      const float* L = Ldiag + int64_t(kb) * NB * NB;
      for (int j = nb - 1; j >= 0; --j) {
        const float xj = __shfl_sync(0xffffffffu, val, j) / L[j * NB + j];
        if (lane == j) val = xj;
        if (lane < j) val -= L[j * NB + lane] * xj;
      }
      if (lane < nb) xs[k0 + lane] = val;
    }
    __syncthreads();
  }
  const int failed = *fail;
  __syncthreads();
  for (int i = tid; i < n; i += BT) dx[i] = failed ? 0.f : xs[i];
  if (tid == 0) *fail = 0;
}

}  // namespace

cudaError_t launch_cholesky_solve(const Workspace& ws, int n, cudaStream_t st) {
  const int nkb = (n + NB - 1) / NB;
  for (int kb = 0; kb < nkb; ++kb) {
    const int k0 = kb * NB, nb = (n - k0 < NB) ? n - k0 : NB;
    const int rows = n + 1 - (k0 + nb);             // includes the rhs row
    k_potrf_trsm<<<(rows + TR - 1) / TR, TR, 0, st>>>(ws.A, ws.ldA, n, k0, ws.Ldiag + int64_t(kb) * NB * NB, ws.fail);
    g_launch_count++;
    const int r0 = k0 + nb;
    if (r0 < n) {
      const int nti = (n + 1 - r0 + TT - 1) / TT;   // row tiles (incl. rhs row)
      const int ntj = (n - r0 + TT - 1) / TT;       // column tiles
      // tiles with tj <= ti, tj < ntj; nti is ntj or ntj+1
      const int ntiles = ntj * (ntj + 1) / 2 + (nti > ntj ? ntj : 0);
      k_trailing<<<ntiles, 256, 0, st>>>(ws.A, ws.ldA, n, k0, nb, r0);
      g_launch_count++;
    }
  }
  const size_t smem = (size_t((n + 3) & ~3) + 32 * 33) * sizeof(float);
  cudaError_t err = cudaSuccess;
  if (smem > 48 * 1024) {
    err = cudaFuncSetAttribute(k_backsolve, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (err != cudaSuccess) return err;
  }
  k_backsolve<<<1, BT, smem, st>>>(ws.A, ws.ldA, n, ws.Ldiag, ws.fail, ws.dx);
  g_launch_count++;
  return cudaGetLastError();
}

}  // namespace dba
