// One step of iterative refinement of the reduced pose system (sm_100a).
//
// The reference solves S dx = v with torch.linalg.cholesky + cholesky_solve in fp32 (hislam2/geom/chol.py:5-20,66).
// The device factorisation (dba_chol_fused.cu) is as accurate as LAPACK's fp32 at n = 756 but 12x worse at
// n = 3060 (profiles/r2_parity_margins.md): 96 right-looking block steps, each rounding the trailing matrix.
// cond(S) ~ 1e5 turns that into 6e-3 on dx.  Classic mixed-precision refinement repairs it with the factor we
// already have:   r = v - S dx  (fp64 accumulation, S and dx in fp32),   L y = r,   L^T d = y,   dx += d.
// One step multiplies the solver's error by ~cond(S) eps_fp32 ~ 1e-2; what is left is the rounding of S itself.
//   k_residual      : one warp per row of S (lower triangle stored: row part coalesced, column part strided, the
//                     matrix is L2-resident), fixed-order shuffle tree -> bit-reproducible
//   k_forwardsolve  : one CTA, 32-column block steps: y_k = L_kk^-1 r_k (the inverse blocks the factorisation stored
//                     for the backward solve), then r -= L[:, k] y_k with one matrix row per thread (128 contiguous
//                     bytes); the next step's rows are in flight while warp 0 solves the block
//   backward solve  : launch_backsolve(mode = accumulate | finalize) of dba_chol_fused.cu
#include "dba_kernels.h"

namespace dba {

namespace {

constexpr int NB = 32;
constexpr int FSN = 512;             // threads of the forward solve
constexpr int FRT = 2;               // matrix rows per thread per pass

// r = rhs - S x, S symmetric with its lower triangle in Sorig [(n+1), ld], rhs = row n; written to Lm row n
__global__ void __launch_bounds__(256)
k_residual(const float* __restrict__ Sorig, const float* __restrict__ x, int ld, int n, float* __restrict__ r_out) {
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (i >= n) return;
  double acc = 0.0;
  const float* row = Sorig + int64_t(i) * ld;
  for (int j = lane; j <= i; j += 32) acc += double(__ldcg(row + j)) * double(__ldcg(x + j));
  for (int j = i + 1 + lane; j < n; j += 32) acc += double(__ldcg(Sorig + int64_t(j) * ld + i)) * double(__ldcg(x + j));
#pragma unroll
  for (int mk = 16; mk >= 1; mk >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, mk);
  if (lane == 0) r_out[i] = float(double(__ldcg(Sorig + int64_t(n) * ld + i)) - acc);
}

// L y = r in place on yr (= row n of Lm, where the backward solve expects y)
__global__ void __launch_bounds__(FSN)
k_forwardsolve(const float* __restrict__ Lm, const float* __restrict__ Ldiag, int ld_stride, int linv_off, int ld, int n,
               float* __restrict__ yr) {
  extern __shared__ __align__(16) float fsm[];
  float* xs = fsm;                         // [n + 32] running right-hand side -> solution
  float* sLi = fsm + ((n + NB + 3) & ~3);  // [32][33] L_kk^-1 (padded: row reads by lane)
  float* yk = sLi + NB * (NB + 1);         // [32]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nkb = (n + NB - 1) / NB;
  for (int i = tid; i < n + NB; i += FSN) xs[i] = (i < n) ? __ldcg(yr + i) : 0.f;
  for (int kb = 0; kb < nkb; ++kb) {
    const int k0 = kb * NB, nb = min(NB, n - k0), r0 = k0 + nb;
    // the rows below the block at columns k0..k0+31 do not depend on the running vector: request the first pass now
    float4 lr[FRT][NB / 4];
    auto load_rows = [&](int pass) {
#pragma unroll
      for (int q = 0; q < FRT; ++q) {
        const int c = r0 + (pass * FRT + q) * FSN + tid;
#pragma unroll
        for (int g = 0; g < NB / 4; ++g)
          lr[q][g] = (c < n) ? __ldcg(reinterpret_cast<const float4*>(Lm + int64_t(c) * ld + k0) + g)
                             : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    };
    const int npass = (n - r0 + FSN * FRT - 1) / (FSN * FRT);
    if (npass > 0) load_rows(0);
    for (int e = tid; e < NB * NB; e += FSN)
      sLi[(e >> 5) * (NB + 1) + (e & 31)] = __ldcg(Ldiag + int64_t(kb) * ld_stride + linv_off + e);
    __syncthreads();                       // L_kk^-1 staged; xs of the previous step complete
    if (warp == 0) {                       // y_k = L_kk^-1 r_k (identity padded beyond nb)
      float y = 0.f;
#pragma unroll
      for (int c = 0; c < NB; ++c) y = fmaf(sLi[lane * (NB + 1) + c], xs[k0 + c], y);
      yk[lane] = (lane < nb) ? y : 0.f;
      __syncwarp();
      if (lane < nb) xs[k0 + lane] = y;
    }
    __syncthreads();
    float yv[NB];
#pragma unroll
    for (int c4 = 0; c4 < NB; c4 += 4) {
      const float4 v = *reinterpret_cast<const float4*>(yk + c4);
      yv[c4] = v.x; yv[c4 + 1] = v.y; yv[c4 + 2] = v.z; yv[c4 + 3] = v.w;
    }
    for (int pass = 0; pass < npass; ++pass) {
      if (pass > 0) load_rows(pass);
#pragma unroll
      for (int q = 0; q < FRT; ++q) {
        const int c = r0 + (pass * FRT + q) * FSN + tid;
        if (c < n) {
          float s0 = 0.f, s1 = 0.f;
#pragma unroll
          for (int g = 0; g < NB / 4; ++g) {
            s0 = fmaf(lr[q][g].x, yv[4 * g], s0);     s1 = fmaf(lr[q][g].y, yv[4 * g + 1], s1);
            s0 = fmaf(lr[q][g].z, yv[4 * g + 2], s0); s1 = fmaf(lr[q][g].w, yv[4 * g + 3], s1);
          }
          xs[c] -= s0 + s1;
        }
      }
    }
  }
  __syncthreads();
  for (int i = tid; i < n; i += FSN) yr[i] = xs[i];
}

}  // namespace

cudaError_t launch_refine(const Workspace& ws, const float* Sorig, int n, cudaStream_t st) {
  float* yr = ws.Lm + int64_t(n) * ws.ldA;
  k_residual<<<(n + 7) / 8, 256, 0, st>>>(Sorig, ws.dx, ws.ldA, n, yr);
  g_launch_count++;
  {
    const cudaError_t ce = launch_trisolve_cluster(ws, n, true, 0, st);
    if (ce == cudaSuccess) return launch_backsolve(ws, n, 3, st);
    if (ce != cudaErrorNotSupported) return ce;
  }
  int stride = 0, linv = 0;
  cholesky_ldiag_layout(&stride, &linv);
  const size_t smem = (size_t((n + NB + 3) & ~3) + NB * (NB + 1) + NB) * sizeof(float);
  cudaError_t err = cudaFuncSetAttribute(k_forwardsolve, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
  if (err != cudaSuccess) return err;
  k_forwardsolve<<<1, FSN, smem, st>>>(ws.Lm, ws.Ldiag, stride, linv, ws.ldA, n, yr);
  g_launch_count++;
  err = cudaGetLastError();
  if (err != cudaSuccess) return err;
  return launch_backsolve(ws, n, 3, st);
}

}  // namespace dba
