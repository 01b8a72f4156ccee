// Depth covariance of the Schur solve (hislam2/geom/chol.py:74-76, the third output of BA, geom/ba.py:98-107;
// SURVEY.md section 8f rank 4):
//     F = L^-1 (E Q) ; dzcov = colsum(F^2) + Q        with S = L L^T the damped reduced pose system
// i.e. per pixel (k,p):  dzcov = Q^2 u^T S^-1 u + Q,  u = column (k,p) of E -- non-zero only in the 6-blocks of pose k
// and of the poses k has edges to.  The reference forms F densely ([6P, M*HW]: 1.3 GB at 128 keyframes); here
//   k_linv    W = L^-1 by 32-column blocks (one CTA per block column, forward substitution by blocks)
//   k_wtw     S^-1 = W^T W (lower tiles, mirrored), summed and STORED in fp64
//   k_dzcov   per (source keyframe, pixel tile): the columns of E are recomputed from the inputs (never stored),
//             kept in shared memory, and the quadratic form runs over the (deg+1)^2 6x6 blocks of S^-1 it touches,
//             in fp64.  Why fp64 behind an fp32 W: u^T (W^T W) u cancels where |W u|^2 (the reference's sum of squares)
//             does not -- with fp32 sums the rounding error is n eps rho^2 against the reference's 2 n eps rho,
//             rho = || |W| |u| || / || W u || (measured 3.4x .. 4x the reference's own fp32 error on ill-conditioned
//             ragged graphs, tests/test_ba_ragged_gpu.py); exact products of the fp32 entries summed in fp64 leave
//             only W's own rounding, which enters both forms alike.
#include "dba_kernels.h"
#include "dba_math.cuh"

namespace dba {

namespace {

constexpr int CB = 32;

__device__ __forceinline__ float L_at(const float* __restrict__ L, int ld, int n, int r, int c) {
  // the factor padded with an identity to a multiple of 32
  return (r < n && c < n) ? __ldg(L + int64_t(r) * ld + c) : (r == c ? 1.f : 0.f);
}

// W [nkb*32, ldw] (zero-initialised): block column c of L^-1 by block forward substitution,
//   W[c][c] = L_cc^-1 ;  W[r][c] = -L_rr^-1 sum_{m=c}^{r-1} L[r][m] W[m][c]
// with the inverses of the diagonal blocks taken from the factorisation (Ldiag, identity-padded like L_at).
__global__ void __launch_bounds__(256)
k_linv(const float* __restrict__ L, int ld, int n, const float* __restrict__ Ldiag, int ld_stride, int ld_linv, float* W,
       int ldw) {
  __shared__ __align__(16) float sT[8][CB][CB + 4];         // rows 16-byte aligned: broadcast reads as float4
  __shared__ float sX[CB][CB + 1], sLi[CB][CB + 1];
  const int nkb = (n + CB - 1) / CB;
  const int c = blockIdx.x, c0 = c * CB;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int e = tid; e < CB * CB; e += 256)
    W[int64_t(c0 + (e >> 5)) * ldw + c0 + (e & 31)] = __ldg(Ldiag + int64_t(c) * ld_stride + ld_linv + e);
  __syncthreads();
  for (int r = c + 1; r < nkb; ++r) {
    const int r0 = r * CB;
    for (int e = tid; e < CB * CB; e += 256) sLi[e >> 5][e & 31] = __ldg(Ldiag + int64_t(r) * ld_stride + ld_linv + e);
    // T = sum_{m=c}^{r-1} L[r][m] W[m][c], the m range dealt over the warps; lane = column of the product
    float acc[CB];
#pragma unroll
    for (int i = 0; i < CB; ++i) acc[i] = 0.f;
    for (int m = c + warp; m < r; m += 8) {
      const int m0 = m * CB;
      float wk[CB];
#pragma unroll
      for (int k = 0; k < CB; ++k) wk[k] = __ldcg(W + int64_t(m0 + k) * ldw + c0 + lane);
      // the L block goes through this warp's slice of sT (free until the partial sums are parked there)
      __syncwarp();
#pragma unroll
      for (int i = 0; i < CB; ++i) sT[warp][i][lane] = L_at(L, ld, n, r0 + i, m0 + lane);
      __syncwarp();
#pragma unroll
      for (int i = 0; i < CB; ++i) {
        float s0 = 0.f, s1 = 0.f;
#pragma unroll
        for (int k = 0; k < CB; k += 4) {
          const float4 l4 = *reinterpret_cast<const float4*>(&sT[warp][i][k]);
          s0 = fmaf(l4.x, wk[k], s0);
          s1 = fmaf(l4.y, wk[k + 1], s1);
          s0 = fmaf(l4.z, wk[k + 2], s0);
          s1 = fmaf(l4.w, wk[k + 3], s1);
        }
        acc[i] += s0 + s1;
      }
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < CB; ++i) sT[warp][i][lane] = acc[i];
    __syncthreads();
    for (int e = tid; e < CB * CB; e += 256) {
      const int i = e >> 5, j = e & 31;
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) s += sT[w][i][j];
      sX[i][j] = s;
    }
    __syncthreads();
    for (int e = tid; e < CB * CB; e += 256) {                 // W[r][c] = -L_rr^-1 T
      const int i = e >> 5, j = e & 31;
      float s = 0.f;
#pragma unroll
      for (int k = 0; k < CB; ++k) s = fmaf(sLi[i][k], sX[k][j], s);
      W[int64_t(r0 + i) * ldw + c0 + j] = -s;
    }
    __syncthreads();                                          // block-wide visibility of W[r][c] for the next steps
  }
}

// Right-looking variant for n <= 32 * LINV_RL_MAX: the running sums T_r = sum_m L[r][m] W[m][c] of ALL later block
// rows live in shared memory and receive one term per step, so the chain per step is only
//   warp 0: W[m][c] = -L_mm^-1 T_m  ->  barrier  ->  warp 0: T_{m+1} += L[m+1][m] W[m][c]
// (the other warps update the rows further down meanwhile; L blocks and the next L_mm^-1 are independent of the chain).
constexpr int LINV_RL_MAX = 40;                              // block rows whose T fit next to the staging buffers
constexpr int RL_T = 512, RL_W = RL_T / 32;

__global__ void __launch_bounds__(RL_T, 1)
k_linv_rl(const float* __restrict__ L, int ld, int n, const float* __restrict__ Ldiag, int ld_stride, int ld_linv,
          float* __restrict__ W, int ldw) {
  extern __shared__ __align__(16) float rsm[];
  const int nkb = (n + CB - 1) / CB;
  const int c = blockIdx.x, c0 = c * CB;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nrow = nkb - c - 1;                              // block rows below the diagonal block
  float* sWm = rsm;                                          // [2][32][32]   W[m][c], k-major
  float* sLi = sWm + 2 * CB * CB;                            // [3][32][36]   L_mm^-1, fetched two steps ahead
  float* sLw = sLi + 3 * CB * (CB + 4);                      // [RL_W][32][36] the warp's L[r][m] block
  float* sT = sLw + RL_W * CB * (CB + 4);                       // [nrow][32][32]
  for (int e = tid; e < nrow * CB * CB; e += RL_T) sT[e] = 0.f;
  auto load_linv = [&](int m, int buf) {                     // by one warp: 32 rows x 32 floats
    for (int i = 0; i < CB; ++i)
      sLi[(buf * CB + i) * (CB + 4) + lane] = __ldg(Ldiag + int64_t(m) * ld_stride + ld_linv + i * CB + lane);
  };
  if (warp == 1) load_linv(c, 0);
  if (warp == 2 && c + 1 < nkb) load_linv(c + 1, 1);
  __syncthreads();
  for (int m = c; m < nkb; ++m) {
    const int buf = (m - c) & 1;
    float* Wm = sWm + buf * CB * CB;
    if (warp == 0) {
      // W[m][c]: lane = column j; the diagonal step takes L_cc^-1 itself
      const float* Li = sLi + ((m - c) % 3) * CB * (CB + 4);
      if (m == c) {
#pragma unroll 8
        for (int i = 0; i < CB; ++i) {
          const float v = Li[i * (CB + 4) + lane];
          Wm[i * CB + lane] = v;
          W[int64_t(c0 + i) * ldw + c0 + lane] = v;
        }
      } else {
        const float* T = sT + size_t(m - c - 1) * CB * CB;
        float tk[CB];
#pragma unroll
        for (int k = 0; k < CB; ++k) tk[k] = T[k * CB + lane];
        // k outermost: 32 independent accumulators, so the broadcast loads of consecutive rows pipeline
        float acc[CB];
#pragma unroll
        for (int i = 0; i < CB; ++i) acc[i] = 0.f;
#pragma unroll
        for (int k = 0; k < CB; k += 4) {
#pragma unroll
          for (int i = 0; i < CB; ++i) {
            const float4 l4 = *reinterpret_cast<const float4*>(Li + i * (CB + 4) + k);
            acc[i] = fmaf(l4.x, tk[k], acc[i]); acc[i] = fmaf(l4.y, tk[k + 1], acc[i]);
            acc[i] = fmaf(l4.z, tk[k + 2], acc[i]); acc[i] = fmaf(l4.w, tk[k + 3], acc[i]);
          }
        }
#pragma unroll
        for (int i = 0; i < CB; ++i) {
          Wm[i * CB + lane] = -acc[i];
          W[int64_t(m * CB + i) * ldw + c0 + lane] = -acc[i];
        }
      }
    }
    __syncthreads();                                         // W[m][c] visible; the other sWm / sLi buffers are free
    if (warp == RL_W - 1 && m + 2 < nkb) load_linv(m + 2, (m + 2 - c) % 3);   // read after the barrier of step m + 1
    // T_r += L[r][m] W[m][c] for the rows below, dealt over the warps (warp 0 takes r = m + 1: the chain)
    float wk[CB];
#pragma unroll
    for (int k = 0; k < CB; ++k) wk[k] = Wm[k * CB + lane];
    float* Lw = sLw + warp * CB * (CB + 4);
    for (int r = m + 1 + warp; r < nkb; r += RL_W) {
      __syncwarp();
#pragma unroll
      for (int i = 0; i < CB; ++i) Lw[i * (CB + 4) + lane] = L_at(L, ld, n, r * CB + i, m * CB + lane);
      __syncwarp();
      float* T = sT + size_t(r - c - 1) * CB * CB;
      float acc[CB];
#pragma unroll
      for (int i = 0; i < CB; ++i) acc[i] = T[i * CB + lane];
#pragma unroll
      for (int k = 0; k < CB; k += 4) {
#pragma unroll
        for (int i = 0; i < CB; ++i) {
          const float4 l4 = *reinterpret_cast<const float4*>(Lw + i * (CB + 4) + k);
          acc[i] = fmaf(l4.x, wk[k], acc[i]); acc[i] = fmaf(l4.y, wk[k + 1], acc[i]);
          acc[i] = fmaf(l4.z, wk[k + 2], acc[i]); acc[i] = fmaf(l4.w, wk[k + 3], acc[i]);
        }
      }
#pragma unroll
      for (int i = 0; i < CB; ++i) T[i * CB + lane] = acc[i];
    }
    // no barrier here: T_{m+1} belongs to warp 0 itself, everything else is protected by the barrier of step m + 1
  }
}

size_t linv_rl_smem(int nkb) {
  return (size_t(2) * CB * CB + 3 * CB * (CB + 4) + RL_W * CB * (CB + 4) + size_t(nkb > 1 ? nkb - 1 : 0) * CB * CB) * sizeof(float);
}

// Sinv = W^T W, lower tiles (a >= b) mirrored.  W is lower triangular by blocks: rows >= a only.
__global__ void __launch_bounds__(256)
k_wtw(const float* __restrict__ W, int ldw, int nkb, double* __restrict__ Sinv) {
  __shared__ float sA[CB][CB + 1], sB[CB][CB + 1];
  int a = int((sqrtf(8.0f * float(blockIdx.x) + 1.0f) - 1.0f) * 0.5f);
  while ((a + 1) * (a + 2) / 2 <= int(blockIdx.x)) ++a;
  while (a * (a + 1) / 2 > int(blockIdx.x)) --a;
  const int b = int(blockIdx.x) - a * (a + 1) / 2;
  const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;     // outputs (i = ty + 8q, j = tx)
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  for (int r = a; r < nkb; ++r) {
    for (int e = tid; e < CB * CB; e += 256) {
      const int i = e >> 5, j = e & 31;
      sA[i][j] = W[int64_t(r * CB + i) * ldw + a * CB + j];
      sB[i][j] = W[int64_t(r * CB + i) * ldw + b * CB + j];
    }
    __syncthreads();
#pragma unroll 8
    for (int k = 0; k < CB; ++k) {
      const double bv = double(sB[k][tx]);
#pragma unroll
      for (int q = 0; q < 4; ++q) acc[q] = fma(double(sA[k][ty + 8 * q]), bv, acc[q]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int i = a * CB + ty + 8 * q, j = b * CB + tx;
    Sinv[int64_t(i) * ldw + j] = acc[q];
    Sinv[int64_t(j) * ldw + i] = acc[q];
  }
}

constexpr int CV_CHUNK = 32;

// one thread per pixel of (source keyframe slot m, PT-pixel tile); dynamic smem: u [(deg+1)*6][PT] | Sinv blocks
__global__ void
k_dzcov(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr,
        const float* __restrict__ targets, const float* __restrict__ weights, const int32_t* __restrict__ kx,
        const int32_t* __restrict__ koff, const int32_t* __restrict__ cedge, const int32_t* __restrict__ cj, int HW, int wd,
        int ntiles, int t0, int Pp, const double* __restrict__ Sinv, int lds, const float* __restrict__ Q,
        float* __restrict__ dzcov) {
  extern __shared__ __align__(16) float smem[];
  __shared__ float sRt[CV_CHUNK * 12], sAd[CV_CHUNK * 36];
  __shared__ int sedge[CV_CHUNK];
  const int PT = blockDim.x;
  const int m = blockIdx.x / ntiles, tile = blockIdx.x % ntiles;
  const int tid = threadIdx.x;
  const int k = kx[m];
  const int e0 = koff[m], deg = koff[m + 1] - e0;
  float* su = smem;                                         // [(1 + deg) * 6][PT]
  double* sS = reinterpret_cast<double*>(smem + size_t(1 + deg) * 6 * PT);   // [2][(1 + deg)][36], 8-byte aligned (PT % 32 == 0)
  int* sslot = reinterpret_cast<int*>(sS + size_t(1 + deg) * 72);
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const int p = tile * PT + tid;
  const bool inb = p < HW;
  const float dd = inb ? disps[int64_t(k) * HW + p] : 0.f;
  const int py = p / wd, px = p - py * wd;
  const float xx = (float(px) - cx) / fx, yy = (float(py) - cy) / fy;
  // entry 0 = pose k itself, entry 1 + e = the target pose of edge e; slot < 0: fixed pose, no row in S
  for (int e = tid; e <= deg; e += PT) {
    const int f = (e == 0) ? k : cj[e0 + e - 1];
    const int s = f - t0;
    sslot[e] = (s >= 0 && s < Pp) ? s : -1;
  }
  float ui[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int c0 = 0; c0 < deg; c0 += CV_CHUNK) {
    const int nc = min(CV_CHUNK, deg - c0);
    __syncthreads();
    for (int le = tid; le < nc; le += PT) {
      float Rt[12], Ad[36];
      relative_pose_Rt(poses, k, cj[e0 + c0 + le], Rt);
      adjoint_from_Rt(Rt, Ad);
      for (int i = 0; i < 12; ++i) sRt[12 * le + i] = Rt[i];
      for (int i = 0; i < 36; ++i) sAd[36 * le + i] = Ad[i];
      sedge[le] = cedge[e0 + c0 + le];
    }
    __syncthreads();
    float nwx = 0.f, nwy = 0.f;                             // weights of the next edge, one edge ahead
    if (inb && nc > 0) {
      const float* wg = weights + int64_t(sedge[0]) * 2 * HW;
      nwx = __ldcs(wg + p); nwy = __ldcs(wg + HW + p);
    }
    for (int le = 0; le < nc; ++le) {
      float Ej[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      const float wx = nwx, wy = nwy;
      if (inb && le + 1 < nc) {
        const float* wg = weights + int64_t(sedge[le + 1]) * 2 * HW;
        nwx = __ldcs(wg + p); nwy = __ldcs(wg + HW + p);
      }
      if (inb) {
        const Proj pr = project(sRt + 12 * le, xx, yy, dd, fx, fy, cx, cy);
        // h = sum_rows w jz g (3-vector); Ej = [d h, X1 x h]   (Jj_r = [d g_r, X1 x g_r], dba_math.cuh)
        const float s0 = kWeightScale * (pr.valid * wx) * pr.jz0, s1 = kWeightScale * (pr.valid * wy) * pr.jz1;
        const float hx = s0 * pr.a0, hy = s1 * pr.b1, hz = s0 * pr.c0 + s1 * pr.c1;
        Ej[0] = dd * hx; Ej[1] = dd * hy; Ej[2] = dd * hz;
        Ej[3] = pr.X1.y * hz - pr.X1.z * hy;
        Ej[4] = pr.X1.z * hx - pr.X1.x * hz;
        Ej[5] = pr.X1.x * hy - pr.X1.y * hx;
        const float* Ad = sAd + 36 * le;
#pragma unroll
        for (int r = 0; r < 6; ++r) {
          float s = 0.f;
#pragma unroll
          for (int t = 0; t < 6; ++t) s = fmaf(Ad[t * 6 + r], Ej[t], s);
          ui[r] -= s;                                      // Ei = -Ad^T Ej (projective_ops.py:66)
        }
      }
#pragma unroll
      for (int c = 0; c < 6; ++c) su[(size_t(1 + c0 + le) * 6 + c) * PT + tid] = Ej[c];
    }
  }
#pragma unroll
  for (int c = 0; c < 6; ++c) su[size_t(c) * PT + tid] = ui[c];
  __syncthreads();
  double quad = 0.0;
  // rows of S^-1 blocks for entry a: [b <= a][6][6], double-buffered -- the next row is in flight while this one is used
  const int rowlen = (1 + deg) * 36;
  double* sSb[2] = {sS, sS + rowlen};
  auto fetch = [&](int a, double* dst) {                     // cp.async: global -> shared without registers
    const int sa = sslot[a];
    if (sa < 0) return;
    for (int e = tid; e < (a + 1) * 36; e += PT) {
      const int b = e / 36, rc = e - b * 36, r = rc / 6, c = rc - r * 6;
      const int sb = sslot[b];
      const double* src = (sb >= 0) ? Sinv + int64_t(6 * sa + r) * lds + 6 * sb + c : Sinv;
      const unsigned d = static_cast<unsigned>(__cvta_generic_to_shared(dst + e));
      const int sz = (sb >= 0) ? 8 : 0;                     // src-size 0: zero fill
      asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(d), "l"(src), "r"(sz) : "memory");
    }
  };
  fetch(0, sSb[0]);
  for (int a = 0; a <= deg; ++a) {
    double* cur = sSb[a & 1];
    asm volatile("cp.async.wait_all;\n" ::: "memory");
    __syncthreads();                                        // row a landed; everyone is done with the other buffer
    if (a + 1 <= deg) fetch(a + 1, sSb[(a + 1) & 1]);
    if (sslot[a] < 0) continue;                             // block-uniform: a fixed pose has no row in S
    double ua[6];
#pragma unroll
    for (int c = 0; c < 6; ++c) ua[c] = double(su[(size_t(a) * 6 + c) * PT + tid]);
    // t = sum_{b<a} S_ab u_b + S_aa u_a / 2 ; quad += 2 u_a . t   (S symmetric: each off-diagonal pair once)
    double t[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    for (int b = 0; b <= a; ++b) {
      const double2* Sb = reinterpret_cast<const double2*>(cur + b * 36);     // 16-byte aligned: 36 doubles per block
      const double sc = (b == a) ? 0.5 : 1.0;
      double ub[6];
#pragma unroll
      for (int c = 0; c < 6; ++c) ub[c] = sc * double(su[(size_t(b) * 6 + c) * PT + tid]);
#pragma unroll
      for (int r = 0; r < 6; ++r) {
#pragma unroll
        for (int c2 = 0; c2 < 3; ++c2) {
          const double2 v = Sb[r * 3 + c2];
          t[r] = fma(v.x, ub[2 * c2], t[r]);
          t[r] = fma(v.y, ub[2 * c2 + 1], t[r]);
        }
      }
    }
    double dq = 0.0;
#pragma unroll
    for (int r = 0; r < 6; ++r) dq = fma(ua[r], t[r], dq);
    quad = fma(2.0, dq, quad);
  }
  if (inb) {
    const double q = double(Q[int64_t(m) * HW + p]);
    dzcov[int64_t(m) * HW + p] = float(q * q * quad + q);
  }
}

}  // namespace

size_t cov_scratch_floats(int n) {
  const size_t npad = size_t((n + CB - 1) / CB) * CB;
  return 3 * npad * npad;             // W (fp32) + S^-1 (fp64)
}

cudaError_t launch_depth_cov(const DeviceGraph& g, const BAInputs& in, const Workspace& ws, float* scratch,
                             size_t smem_cap, float* dzcov, cudaStream_t st) {
  const int n = g.n6;
  const int nkb = (n + CB - 1) / CB, npad = nkb * CB;
  float* W = scratch;
  double* Sinv = reinterpret_cast<double*>(scratch + size_t(npad) * npad);   // npad % 32 == 0: 8-byte aligned
  cudaError_t err = cudaMemsetAsync(W, 0, size_t(npad) * npad * sizeof(float), st);
  if (err != cudaSuccess) return err;
  if (n > 0) {
    int ld_stride = 0, ld_linv = 0;
    cholesky_ldiag_layout(&ld_stride, &ld_linv);
    if (nkb <= LINV_RL_MAX && linv_rl_smem(nkb) + 1024 <= smem_cap) {
      err = cudaFuncSetAttribute(k_linv_rl, cudaFuncAttributeMaxDynamicSharedMemorySize, int(linv_rl_smem(nkb)));
      if (err != cudaSuccess) return err;
      k_linv_rl<<<nkb, RL_T, linv_rl_smem(nkb), st>>>(ws.Lm, ws.ldA, n, ws.Ldiag, ld_stride, ld_linv, W, npad);
    } else {
      k_linv<<<nkb, 256, 0, st>>>(ws.Lm, ws.ldA, n, ws.Ldiag, ld_stride, ld_linv, W, npad);
    }
    k_wtw<<<nkb * (nkb + 1) / 2, 256, 0, st>>>(W, npad, nkb, Sinv);
    g_launch_count += 2;
  }
  if (g.M > 0) {
    int PT = 128;
    auto bytes = [&](int pt) { return (size_t(1 + g.degmax) * 6 * pt + size_t(1 + g.degmax) * 145 + 8) * sizeof(float); };
    while (PT > 32 && bytes(PT) > (smem_cap > 12288 ? smem_cap - 12288 : 0)) PT >>= 1;
    if (bytes(PT) + 12288 > smem_cap) return cudaErrorInvalidConfiguration;
    err = cudaFuncSetAttribute(k_dzcov, cudaFuncAttributeMaxDynamicSharedMemorySize, int(bytes(PT)));
    if (err != cudaSuccess) return err;
    const int ntiles = (g.HW + PT - 1) / PT;
    k_dzcov<<<g.M * ntiles, PT, bytes(PT), st>>>(in.poses, in.disps, in.intrinsics, in.targets, in.weights, g.kx, g.koff,
                                                  g.cedge, g.cj, g.HW, g.wd, ntiles, g.t0, g.Pp, Sinv, npad, ws.Q, dzcov);
    g_launch_count++;
  }
  return cudaGetLastError();
}

}  // namespace dba
