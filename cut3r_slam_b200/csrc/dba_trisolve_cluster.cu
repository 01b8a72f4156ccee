// Triangular solves of the reduced pose system on ONE THREAD-BLOCK CLUSTER (sm_100a), n <= 768.
//
// The reference calls torch.cholesky_solve (hislam2/geom/chol.py:66); here the backward solve L^T x = y runs after
// every factorisation and the forward solve L y = r in the refinement step (dba_refine.cu).  The single-CTA kernels
// (k_backsolve_smem, k_forwardsolve) stream the factor through one SM: 1.1 MB at n = 756 in 32-column slabs, each
// step waiting for its slab's L2 round trip -- 28 us and 77 us for what is a chain of 24 small products.
// Here the factor is RESIDENT: a cluster of NC = ceil(n / 64) CTAs (12 at n = 756; sizes above 8 are the opt-in
// non-portable kind); CTA c owns the pose columns C_c = [64 c, 64 c + 64) and keeps in its shared memory the 64 x 64
// blocks it multiplies (<= 11 x 18 KB, staged with cp.async by all CTAs at once).  The dependency chain runs through the
// cluster, one hop per panel:
//   forward  (L y = r),   CTA 0 first:     CTA c solves y_c, then sends each later CTA d the product L[C_d, C_c] y_c;
//   backward (L^T x = y), last CTA first:  CTA c solves x_c, then sends each earlier CTA d the product L[C_c, C_d]^T x_c;
//   (nearest destination first); a receiver subtracts the vectors in arrival order and then solves its own 64 x 64
//   triangle with the stored inverses of the two diagonal blocks (three 32-term products on all eight warps).
// A product's 64 floats go straight into the receiver's shared memory with st.async ... mbarrier::complete_tx::bytes:
// the store itself credits the receiver's barrier (armed with expect_tx = 256 bytes), no release / arrive round trip,
// no CTA barrier on the sender (each warp stages, multiplies and sends its own eight outputs).
// Measured (B200, n = 756, globaltimer trace -DDBA_CS_TRACE): 0.73 us per hop (solve 0.35, product 0.25, DSMEM 0.13),
// 12.7 us per solve against 28.4 us (backward) / 77 us (forward) for the slab kernels.  Sums are split in a fixed
// order: bit-reproducible, not bit-identical to the slab kernels.
// Every wait is bounded (2 s of %globaltimer): a lost signal raises the failure flag (dx = 0) instead of hanging.
#include "dba_kernels.h"
#ifdef DBA_CS_TRACE
#include <cstdio>
#define CST(i) do { if (tid == 0) tst[i] = cs_timer(); } while (0)
#else
#define CST(i) do { } while (0)
#endif

namespace dba {

int g_cluster_solve_mode = 1;      // option "cluster_solve": 0 off, 1 on where it applies

namespace {

constexpr int CS_T = 256;          // threads per CTA
constexpr int CS_W = 64;           // columns per CTA
constexpr int CS_MAXC = 16;        // largest cluster
constexpr int CS_NMAX = 768;
constexpr int NB = 32;
constexpr int LIS = NB + 1;        // row stride of a staged inverse block

__device__ __forceinline__ uint32_t cs_smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ uint32_t cs_mapa(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
// one float into another CTA's shared memory; the store itself credits 4 bytes to the receiver's mbarrier (st.async
// with complete_tx: no separate release / arrive round trip, the receiver's acquire-wait sees the data)
__device__ __forceinline__ void cs_st_signal(uint32_t addr, float v, uint32_t bar_addr) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];"
               ::"r"(addr), "r"(__float_as_uint(v)), "r"(bar_addr) : "memory");
}
__device__ __forceinline__ void cs_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ unsigned long long cs_timer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// phase 0 of a barrier that is used once; false after 2 s
__device__ __forceinline__ bool cs_wait(uint64_t* bar) {
  const uint32_t a = cs_smem_u32(bar);
  unsigned long long t0 = 0;
  for (unsigned it = 0;; ++it) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(done) : "r"(a), "r"(0u) : "memory");
    if (done) return true;
    if ((it & 1023u) == 1023u) {
      const unsigned long long t = cs_timer();
      if (t0 == 0) t0 = t;
      else if (t - t0 > 2000000000ull) return false;
    }
  }
}
__device__ __forceinline__ void cs_cp16(float* dst, const float* src, bool valid) {
  const unsigned d = cs_smem_u32(dst);
  const int sz = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(src), "r"(sz) : "memory");
}

constexpr int BLD = CS_W + 8;      // row stride of a staged 64 x 64 block: 72 = 8 banks per row step, see the backward product
constexpr int CLD = NB + 4;        // row stride of the 32 x 32 coupling block of the CTA's own triangle

__device__ __forceinline__ void cs_cp4(float* dst, const float* src, bool valid) {
  const unsigned d = cs_smem_u32(dst);
  const int sz = valid ? 4 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(d), "l"(src), "r"(sz) : "memory");
}

__host__ __device__ inline size_t cs_smem_floats(int n) {
  const int nc = (n + CS_W - 1) / CS_W;
  return size_t(nc > 1 ? nc - 1 : 1) * CS_W * BLD + NB * CLD + 2 * NB * LIS + CS_MAXC * CS_W + 2 * CS_W + 64;
}

__device__ __forceinline__ void cs_wait_pending(int npend) {     // cp.async.wait_group takes an immediate
  switch (npend) {
    case 0: asm volatile("cp.async.wait_group 0;\n" ::: "memory"); break;
    case 1: asm volatile("cp.async.wait_group 1;\n" ::: "memory"); break;
    case 2: asm volatile("cp.async.wait_group 2;\n" ::: "memory"); break;
    case 3: asm volatile("cp.async.wait_group 3;\n" ::: "memory"); break;
    case 4: asm volatile("cp.async.wait_group 4;\n" ::: "memory"); break;
    case 5: asm volatile("cp.async.wait_group 5;\n" ::: "memory"); break;
    case 6: asm volatile("cp.async.wait_group 6;\n" ::: "memory"); break;
    case 7: asm volatile("cp.async.wait_group 7;\n" ::: "memory"); break;
    case 8: asm volatile("cp.async.wait_group 8;\n" ::: "memory"); break;
    case 9: asm volatile("cp.async.wait_group 9;\n" ::: "memory"); break;
    case 10: asm volatile("cp.async.wait_group 10;\n" ::: "memory"); break;
    case 11: asm volatile("cp.async.wait_group 11;\n" ::: "memory"); break;
    default: asm volatile("cp.async.wait_group 0;\n" ::: "memory"); break;
  }
}

// One cluster = the whole grid.  CTA c owns the pose columns C_c = [64 c, 64 c + 64).  FWD (L y = r): CTA 0 first, after
// y_c every later CTA d gets L[C_d, C_c] y_c; BWD (L^T x = y): the last CTA first, after x_c every earlier CTA d gets
// L[C_c, C_d]^T x_c.  The 64 x 64 blocks those products need are staged per WARP (each warp loads exactly the eight
// rows / columns it multiplies, so a product needs no CTA barrier), nearest destination first, one cp.async group
// per destination; a destination's barrier counts the eight warps of the sender.
template <bool FWD>
__global__ void __launch_bounds__(CS_T, 1)
k_trisolve_cluster(const float* __restrict__ Lm, const float* __restrict__ Ldiag, int ld_stride, int linv_off, int ld, int n,
                   float* __restrict__ yr, int* fail, unsigned* bar, float* __restrict__ dx, int mode) {
  extern __shared__ __align__(16) float sm[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int NC = int(gridDim.x), c = int(blockIdx.x);
  const int c0 = c * CS_W, w = min(CS_W, n - c0);
  const int ndest = FWD ? NC - 1 - c : c;
  float* Bd = sm;                                            // [ndest][64][BLD]
  float* Cb = Bd + size_t(NC > 1 ? NC - 1 : 1) * CS_W * BLD; // [32][CLD]: L[c0 + 32 + j][c0 + i]
  float* Li = Cb + NB * CLD;                                 // [2][32][33]: inverses of the two diagonal blocks
  float* msg = Li + 2 * NB * LIS;                            // [NC][64]: vectors written by the other CTAs
  float* rhs = msg + CS_MAXC * CS_W;                         // [64]
  float* xs = rhs + CS_W;                                    // [64]
  float* tmp = xs + CS_W;                                    // [64]
  __shared__ uint64_t mb[CS_MAXC];                           // one-shot barriers, one per sender
  const int failed_in = *reinterpret_cast<volatile int*>(fail);
#ifdef DBA_CS_TRACE
  unsigned long long tst[8] = {};
#endif
  CST(0);
  // ---- stage, all asynchronous: the CTA's own triangle (group 0), then the destination blocks (warp-private parts) --
  for (int e = tid; e < NB * NB; e += CS_T) {                // coupling block (rows 32.., columns ..31 of the triangle)
    const int j = e >> 5, i = e & 31;
    cs_cp4(Cb + j * CLD + i, Lm + int64_t(c0 + NB + j) * ld + c0 + i, c0 + NB + j < n);
  }
  for (int e = tid; e < 2 * NB * NB; e += CS_T) {            // L_kk^-1 of the two diagonal blocks
    const int b = e >> 10, r = (e >> 5) & 31, cc = e & 31;
    const int kb = 2 * c + b;
    cs_cp4(Li + b * NB * LIS + r * LIS + cc, Ldiag + int64_t(kb) * ld_stride + linv_off + r * NB + cc, kb * NB < n);
  }
  asm volatile("cp.async.commit_group;\n" ::: "memory");
  for (int b = 0; b < ndest; ++b) {
    const int d = FWD ? c + 1 + b : c - 1 - b;
    float* blk = Bd + size_t(b) * CS_W * BLD;
    if (FWD) {              // rows 8 warp .. 8 warp + 7 of L[C_d, C_c]: 16 chunks of 16 bytes each
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int q = lane + 32 * t;
        const int i = 8 * warp + (q >> 4), ch = q & 15;
        cs_cp16(blk + i * BLD + ch * 4, Lm + int64_t(d * CS_W + i) * ld + c0 + ch * 4, d * CS_W + i < n);
      }
    } else {                // columns 8 warp .. 8 warp + 7 of L[C_c, C_d]: two 16-byte chunks per row
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int q = lane + 32 * t;
        const int j = q >> 1, half = q & 1;
        cs_cp16(blk + j * BLD + 8 * warp + half * 4, Lm + int64_t(c0 + j) * ld + d * CS_W + 8 * warp + half * 4, c0 + j < n);
      }
    }
    asm volatile("cp.async.commit_group;\n" ::: "memory");
  }
  CST(7);
  if (tid < CS_W) {
    rhs[tid] = (tid < w) ? __ldcg(yr + c0 + tid) : 0.f;
    xs[tid] = 0.f;
  }
  if (tid < NC) {
    const uint32_t a = cs_smem_u32(&mb[tid]);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(a) : "memory");
    // the one pending arrival is this one; the phase completes when the sender's 64 floats have landed
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(a), "r"(uint32_t(CS_W * sizeof(float))) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  cs_cluster_sync();                                         // every CTA's barriers exist before anyone signals
  bool ok = true;
  CST(1);
  cs_wait_pending(ndest > 0 ? ndest - 1 : 0);                // the triangle's blocks and the nearest destination's
  if (2 * c * NB + NB >= n && tid < NB) Li[NB * LIS + tid * LIS + tid] = 1.f;   // no second diagonal block: identity
  // ---- the other panels' contributions, in arrival order (the nearest sender's comes last) -----------------------
  for (int s = 0; s < (FWD ? c : NC - 1 - c); ++s) {
    const int cp = FWD ? s : NC - 1 - s;
    if (ok && !cs_wait(&mb[cp])) ok = false;
    if (tid < CS_W) rhs[tid] -= msg[cp * CS_W + tid];
  }
  __syncthreads();
  CST(2);
  // ---- the CTA's own 64 x 64 triangle: three 32-term products on all eight warps (thread -> output o, terms 4 q ..) --
  {
    const int o = tid >> 3, q = tid & 7;
    auto dot32 = [&](const float* B, bool transpose, const float* v) {
      float s = 0.f;
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int k = 4 * q + t;
        s = fmaf(transpose ? B[k * LIS + o] : B[o * LIS + k], v[k], s);
      }
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      s += __shfl_xor_sync(0xffffffffu, s, 4);
      return s;
    };
    auto couple = [&](const float* v, bool transpose) {     // sum_i C[o][i] v[i]  or  sum_j C[j][o] v[j]
      float s = 0.f;
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int k = 4 * q + t;
        s = fmaf(transpose ? Cb[k * CLD + o] : Cb[o * CLD + k], v[k], s);
      }
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      s += __shfl_xor_sync(0xffffffffu, s, 4);
      return s;
    };
    if (FWD) {
      const float ylo = dot32(Li, false, rhs);                                   // y_lo = L_ll^-1 rhs_lo
      if (q == 0) xs[o] = ylo;
      __syncthreads();
      const float cp_ = couple(xs, false);                                       // rhs_hi -= L[hi, lo] y_lo
      if (q == 0) tmp[o] = rhs[NB + o] - cp_;
      __syncthreads();
      const float yhi = dot32(Li + NB * LIS, false, tmp);                        // y_hi = L_hh^-1 rhs_hi
      if (q == 0) xs[NB + o] = yhi;
    } else {
      const float xhi = dot32(Li + NB * LIS, true, rhs + NB);                    // x_hi = L_hh^-T rhs_hi
      if (q == 0) xs[NB + o] = xhi;
      __syncthreads();
      const float cp_ = couple(xs + NB, true);                                   // rhs_lo -= L[hi, lo]^T x_hi
      if (q == 0) tmp[o] = rhs[o] - cp_;
      __syncthreads();
      const float xlo = dot32(Li, true, tmp);                                    // x_lo = L_ll^-T rhs_lo
      if (q == 0) xs[o] = xlo;
    }
    __syncthreads();
  }
  CST(3);
  // ---- products for the destinations, nearest first; each warp: its eight outputs, stores, one arrive ------------
  for (int b = 0; b < ndest; ++b) {
    const int d = FWD ? c + 1 + b : c - 1 - b;
    if (b > 0) {
      cs_wait_pending(ndest - 1 - b);
      __syncwarp();
    }
    if (b == 0) CST(4);
    const float* blk = Bd + size_t(b) * CS_W * BLD;
    const int il = lane >> 2, kq = lane & 3, i = 8 * warp + il;
    float s = 0.f;
    float s1 = 0.f;
    if (FWD) {              // out[i] = sum_k blk[i][k] y[k]: the lane's quarter of row i
      const float4* row = reinterpret_cast<const float4*>(blk + i * BLD + kq * 16);
      const float4* v4 = reinterpret_cast<const float4*>(xs + kq * 16);
#pragma unroll
      for (int g = 0; g < 4; g += 2) {
        const float4 a = row[g], v = v4[g], a2 = row[g + 1], v2 = v4[g + 1];
        s = fmaf(a.x, v.x, s); s = fmaf(a.y, v.y, s); s = fmaf(a.z, v.z, s); s = fmaf(a.w, v.w, s);
        s1 = fmaf(a2.x, v2.x, s1); s1 = fmaf(a2.y, v2.y, s1); s1 = fmaf(a2.z, v2.z, s1); s1 = fmaf(a2.w, v2.w, s1);
      }
    } else {                // out[i] = sum_j blk[j][i] x[j], rows j = 4 t + kq: the four lanes of an output read rows
                            // 72 floats = 8 banks apart, the eight outputs of the warp fill the gaps: no bank conflict
#pragma unroll
      for (int t = 0; t < 16; t += 2) {
        const int j = 4 * t + kq;
        s = fmaf(blk[j * BLD + i], xs[j], s);
        s1 = fmaf(blk[(j + 4) * BLD + i], xs[j + 4], s1);
      }
    }
    s += s1;
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    if (b == 0) CST(5);
    if (kq == 0) cs_st_signal(cs_mapa(cs_smem_u32(msg + c * CS_W + i), uint32_t(d)), s, cs_mapa(cs_smem_u32(&mb[c]), uint32_t(d)));
    if (b == 0) CST(6);
  }
  if (!ok && tid == 0) atomicExch(fail, 4);
  if (FWD) {
    if (tid < w) yr[c0 + tid] = xs[tid];
  } else {
    const bool failed = failed_in != 0 || !ok;
    // mode bit 1: dx += x (the correction of an iterative-refinement pass); bit 0: last solve of the iteration
    if (tid < w) dx[c0 + tid] = failed ? 0.f : ((mode & 2) ? dx[c0 + tid] + xs[tid] : xs[tid]);
  }
  cs_cluster_sync();                               // nobody leaves while its shared memory can still be written
#ifdef DBA_CS_TRACE
  if (tid == 0) printf("cs trace %s cta %2d: copies issued +%llu staged +%llu received +%llu solved +%llu | block ready +%llu product +%llu stored +%llu\n", FWD ? "fwd" : "bwd", c, tst[7] - tst[0], tst[1] - tst[0], tst[2] - tst[0], tst[3] - tst[0], tst[4] - tst[3], tst[5] - tst[3], tst[6] - tst[3]);
#endif
  // last solve of the iteration: the flags start the next one clean (a timed-out wait has zeroed THIS update)
  if (!FWD && (mode & 1) && c == 0 && tid == 0) { *fail = 0; *bar = 0u; }
}

// ---------------------------------------------------------------------------------------------------------------
// The same chain for systems that do not fit a cluster (768 < n <= ~9000: loop-closure graphs), on a co-resident GRID:
// NC = ceil(n / 64) CTAs (cooperative launch), messages through global memory -- slot [dst][src] of 64 floats and a flag
// that holds the launch's epoch (st.release.gpu after a CTA barrier; the receiver's thread 0 polls with ld.acquire.gpu).
// A CTA cannot keep its <= 140 blocks resident: they stream through a ring of GR_RING warp-private slots (each warp
// restages its slice of a slot right after its own product, so the ring needs no CTA barrier either).
// The epoch lives in device memory and is bumped by the chain's LAST CTA, so a captured CUDA graph replays correctly.
// n = 3060: ~48 hops of ~1.8 us against 370 us (backward) / ~1 ms (forward) for the single-CTA slab kernels.
constexpr int GR_RING = 8;

__host__ __device__ inline size_t gr_smem_floats() {
  return size_t(GR_RING) * CS_W * BLD + NB * CLD + 2 * NB * LIS + 3 * CS_W + 64;
}

__device__ __forceinline__ unsigned gr_ld_acquire(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void gr_st_release(unsigned* p, unsigned v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <bool FWD>
__global__ void __launch_bounds__(CS_T, 1)
k_trisolve_grid(const float* __restrict__ Lm, const float* __restrict__ Ldiag, int ld_stride, int linv_off, int ld, int n,
                float* __restrict__ yr, int* fail, unsigned* bar, float* __restrict__ dx, int mode,
                float* __restrict__ gmsg, unsigned* __restrict__ gflag) {
  extern __shared__ __align__(16) float sm[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int NC = int(gridDim.x), c = int(blockIdx.x);
  const int c0 = c * CS_W, w = min(CS_W, n - c0);
  const int ndest = FWD ? NC - 1 - c : c;
  const int nsend = FWD ? c : NC - 1 - c;
  float* Bd = sm;                                            // [GR_RING][64][BLD]
  float* Cb = Bd + size_t(GR_RING) * CS_W * BLD;             // [32][CLD]
  float* Li = Cb + NB * CLD;                                 // [2][32][33]
  float* rhs = Li + 2 * NB * LIS;                            // [64]
  float* xs = rhs + CS_W;                                    // [64]
  float* tmp = xs + CS_W;                                    // [64]
  __shared__ int s_ok;
  const int failed_in = *reinterpret_cast<volatile int*>(fail);
  unsigned* ectr = gflag + size_t(NC) * NC;                  // launch epoch - 1 (bumped by the chain's last CTA)
  const unsigned epoch = gr_ld_acquire(ectr) + 1u;
  if (tid == 0) s_ok = 1;

  for (int e = tid; e < NB * NB; e += CS_T) {
    const int j = e >> 5, i = e & 31;
    cs_cp4(Cb + j * CLD + i, Lm + int64_t(c0 + NB + j) * ld + c0 + i, c0 + NB + j < n);
  }
  for (int e = tid; e < 2 * NB * NB; e += CS_T) {
    const int b = e >> 10, r = (e >> 5) & 31, cc = e & 31;
    const int kb = 2 * c + b;
    cs_cp4(Li + b * NB * LIS + r * LIS + cc, Ldiag + int64_t(kb) * ld_stride + linv_off + r * NB + cc, kb * NB < n);
  }
  asm volatile("cp.async.commit_group;\n" ::: "memory");
  int ncommitted = 0;
  auto stage_dest = [&](int b) {                             // the warp's slice of destination block b into slot b % RING
    const int d = FWD ? c + 1 + b : c - 1 - b;
    float* blk = Bd + size_t(b % GR_RING) * CS_W * BLD;
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int q = lane + 32 * t;
      if (FWD) {
        const int i = 8 * warp + (q >> 4), ch = q & 15;
        cs_cp16(blk + i * BLD + ch * 4, Lm + int64_t(d * CS_W + i) * ld + c0 + ch * 4, d * CS_W + i < n);
      } else {
        const int j = q >> 1, half = q & 1;
        cs_cp16(blk + j * BLD + 8 * warp + half * 4, Lm + int64_t(c0 + j) * ld + d * CS_W + 8 * warp + half * 4, c0 + j < n);
      }
    }
    asm volatile("cp.async.commit_group;\n" ::: "memory");
    ++ncommitted;
  };
  for (int b = 0; b < min(ndest, GR_RING); ++b) stage_dest(b);
  if (tid < CS_W) {
    rhs[tid] = (tid < w) ? __ldcg(yr + c0 + tid) : 0.f;
    xs[tid] = 0.f;
  }
  cs_wait_pending(ncommitted > 0 ? ncommitted - 1 : 0);      // the triangle's blocks and the nearest destination's
  if (2 * c * NB + NB >= n && tid < NB) Li[NB * LIS + tid * LIS + tid] = 1.f;
  __syncthreads();
  // ---- the other panels' contributions, in arrival order ---------------------------------------------------------
  for (int s = 0; s < nsend; ++s) {
    const int cp = FWD ? s : NC - 1 - s;
    if (tid == 0) {
      const unsigned* f = gflag + size_t(c) * NC + cp;
      unsigned long long t0 = 0;
      for (unsigned it = 0; gr_ld_acquire(f) != epoch; ++it) {
        if ((it & 255u) == 255u) {
          const unsigned long long t = cs_timer();
          if (t0 == 0) t0 = t;
          else if (t - t0 > 2000000000ull) { s_ok = 0; break; }
        }
      }
    }
    __syncthreads();
    if (tid < CS_W) rhs[tid] -= __ldcg(gmsg + (size_t(c) * NC + cp) * CS_W + tid);
  }
  __syncthreads();
  {  // ---- the CTA's own 64 x 64 triangle (as in the cluster kernel) --------------------------------------------------
    const int o = tid >> 3, q = tid & 7;
    auto dot32 = [&](const float* B, int sb, int so, const float* v) {
      float s = 0.f;
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int k = 4 * q + t;
        s = fmaf(B[k * sb + o * so], v[k], s);
      }
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      s += __shfl_xor_sync(0xffffffffu, s, 4);
      return s;
    };
    if (FWD) {
      const float ylo = dot32(Li, 1, LIS, rhs);
      if (q == 0) xs[o] = ylo;
      __syncthreads();
      const float cp_ = dot32(Cb, 1, CLD, xs);
      if (q == 0) tmp[o] = rhs[NB + o] - cp_;
      __syncthreads();
      const float yhi = dot32(Li + NB * LIS, 1, LIS, tmp);
      if (q == 0) xs[NB + o] = yhi;
    } else {
      const float xhi = dot32(Li + NB * LIS, LIS, 1, rhs + NB);
      if (q == 0) xs[NB + o] = xhi;
      __syncthreads();
      const float cp_ = dot32(Cb, CLD, 1, xs + NB);
      if (q == 0) tmp[o] = rhs[o] - cp_;
      __syncthreads();
      const float xlo = dot32(Li, LIS, 1, tmp);
      if (q == 0) xs[o] = xlo;
    }
    __syncthreads();
  }
  // ---- products for the destinations, nearest first --------------------------------------------------------------
  for (int b = 0; b < ndest; ++b) {
    const int d = FWD ? c + 1 + b : c - 1 - b;
    if (b > 0) {
      cs_wait_pending(ncommitted - 1 - b);
      __syncwarp();
    }
    const float* blk = Bd + size_t(b % GR_RING) * CS_W * BLD;
    const int il = lane >> 2, kq = lane & 3, i = 8 * warp + il;
    float s = 0.f, s1 = 0.f;
    if (FWD) {
      const float4* row = reinterpret_cast<const float4*>(blk + i * BLD + kq * 16);
      const float4* v4 = reinterpret_cast<const float4*>(xs + kq * 16);
#pragma unroll
      for (int g = 0; g < 4; g += 2) {
        const float4 a = row[g], v = v4[g], a2 = row[g + 1], v2 = v4[g + 1];
        s = fmaf(a.x, v.x, s); s = fmaf(a.y, v.y, s); s = fmaf(a.z, v.z, s); s = fmaf(a.w, v.w, s);
        s1 = fmaf(a2.x, v2.x, s1); s1 = fmaf(a2.y, v2.y, s1); s1 = fmaf(a2.z, v2.z, s1); s1 = fmaf(a2.w, v2.w, s1);
      }
    } else {
#pragma unroll
      for (int t = 0; t < 16; t += 2) {
        const int j = 4 * t + kq;
        s = fmaf(blk[j * BLD + i], xs[j], s);
        s1 = fmaf(blk[(j + 4) * BLD + i], xs[j + 4], s1);
      }
    }
    s += s1;
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    if (kq == 0) __stcg(gmsg + (size_t(d) * NC + c) * CS_W + i, s);
    __syncwarp();
    if (b + GR_RING < ndest) stage_dest(b + GR_RING);        // the warp's slice of this slot is free again
    __syncthreads();                                         // all eight warps' outputs are written
    if (tid == 0) {
      __threadfence();
      gr_st_release(gflag + size_t(d) * NC + c, epoch);
    }
  }
  const bool ok = s_ok != 0;
  if (!ok && tid == 0) atomicExch(fail, 4);
  if (FWD) {
    if (tid < w) yr[c0 + tid] = xs[tid];
  } else {
    const bool failed = failed_in != 0 || !ok;
    if (tid < w) dx[c0 + tid] = failed ? 0.f : ((mode & 2) ? dx[c0 + tid] + xs[tid] : xs[tid]);
  }
  // the chain's last CTA has heard from every other one (each of them read the epoch before it sent anything)
  if (ndest == 0 && tid == 0) {
    gr_st_release(ectr, epoch);
    if (!FWD && (mode & 1)) { *fail = 0; *bar = 0u; }
  }
}

}  // namespace

cudaError_t launch_trisolve_grid(const Workspace& ws, int n, bool forward, int mode, cudaStream_t st);

// cudaErrorNotSupported: the caller falls back to the single-CTA kernels
cudaError_t launch_trisolve_cluster(const Workspace& ws, int n, bool forward, int mode, cudaStream_t st) {
  if (n > CS_NMAX) return launch_trisolve_grid(ws, n, forward, mode, st);
  if (g_cluster_solve_mode == 0 || n < 2 * CS_W) return cudaErrorNotSupported;
  const int NC = (n + CS_W - 1) / CS_W;
  const size_t smem = cs_smem_floats(n) * sizeof(float);
  static int ok_for[CS_MAXC + 1][2] = {};          // 0 unknown, 1 usable, -1 not
  void* kern = forward ? reinterpret_cast<void*>(k_trisolve_cluster<true>) : reinterpret_cast<void*>(k_trisolve_cluster<false>);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(NC);
  cfg.blockDim = dim3(CS_T);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = NC; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int& state = ok_for[NC][forward ? 1 : 0];
  if (state == 0) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(cs_smem_floats(CS_NMAX) * sizeof(float)));
    if (e == cudaSuccess && NC > 8) e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    int nclusters = 0;
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveClusters(&nclusters, kern, &cfg);
    state = (e == cudaSuccess && nclusters >= 1) ? 1 : -1;
    if (e != cudaSuccess) cudaGetLastError();
  }
  if (state < 0) return cudaErrorNotSupported;
  const float* Lm = ws.Lm; const float* Ldiag = ws.Ldiag; int ld = ws.ldA; int* fail = ws.fail;
  unsigned* bar = reinterpret_cast<unsigned*>(ws.fail + 1);
  float* dx = ws.dx;
  float* yr = ws.Lm + int64_t(n) * ws.ldA;
  int stride = 0, linv = 0;
  cholesky_ldiag_layout(&stride, &linv);
  void* args[] = {&Lm, &Ldiag, &stride, &linv, &ld, &n, &yr, &fail, &bar, &dx, &mode};
  cudaError_t err = cudaLaunchKernelExC(&cfg, kern, args);
  if (err != cudaSuccess) return err;
  g_launch_count++;
  return cudaGetLastError();
}

// 768 < n: the chain on a cooperative grid with global-memory messages (ws.solve_msg / ws.solve_flag from the context)
cudaError_t launch_trisolve_grid(const Workspace& ws, int n, bool forward, int mode, cudaStream_t st) {
  const int NC = (n + CS_W - 1) / CS_W;
  if (g_cluster_solve_mode == 0 || ws.solve_msg == nullptr || ws.solve_flag == nullptr || n <= CS_NMAX) return cudaErrorNotSupported;
  void* kern = forward ? reinterpret_cast<void*>(k_trisolve_grid<true>) : reinterpret_cast<void*>(k_trisolve_grid<false>);
  const size_t smem = gr_smem_floats() * sizeof(float);
  static int max_ctas[2] = {0, 0};
  int& mc = max_ctas[forward ? 1 : 0];
  if (mc == 0) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    int dev = 0, sms = 0, per = 0, coop = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, kern, CS_T, smem);
    mc = (e == cudaSuccess && coop && per >= 1) ? sms * per : -1;
    if (e != cudaSuccess) cudaGetLastError();
  }
  if (mc < NC) return cudaErrorNotSupported;
  const float* Lm = ws.Lm; const float* Ldiag = ws.Ldiag; int ld = ws.ldA; int* fail = ws.fail;
  unsigned* bar = reinterpret_cast<unsigned*>(ws.fail + 1);
  float* dx = ws.dx;
  float* yr = ws.Lm + int64_t(n) * ws.ldA;
  float* gmsg = ws.solve_msg; unsigned* gflag = ws.solve_flag;
  int stride = 0, linv = 0;
  cholesky_ldiag_layout(&stride, &linv);
  void* args[] = {&Lm, &Ldiag, &stride, &linv, &ld, &n, &yr, &fail, &bar, &dx, &mode, &gmsg, &gflag};
  cudaError_t err = cudaLaunchCooperativeKernel(kern, dim3(NC), dim3(CS_T), args, smem, st);
  if (err != cudaSuccess) return err;
  g_launch_count++;
  return cudaGetLastError();
}

size_t trisolve_grid_msg_floats(int n) { const size_t nc = (n + CS_W - 1) / CS_W; return n > CS_NMAX ? nc * nc * CS_W : 0; }
size_t trisolve_grid_flag_words(int n) { const size_t nc = (n + CS_W - 1) / CS_W; return n > CS_NMAX ? nc * nc + 4 : 0; }

}  // namespace dba
