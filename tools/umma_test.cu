// Developer harness: validates the hand-written tcgen05 path used by the Schur products
// (SWIZZLE_128B K-major fp32 tiles written with st.shared, tf32 UMMA with a hi/lo split,
// mbarrier commit, tcgen05.ld epilogue) against a CPU fp64 product.   D = P P^T, P is [R x K].
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cstdint>
#include <cstring>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

#include "../cut3r_slam_b200/csrc/dba_umma.cuh"

using namespace dba::umma;

// P: [R][K] row-major fp32 in global. One CTA, 128 threads. K multiple of CH (pixels per chunk), R <= 128.
// mode 0: plain tf32, accumulate everything in TMEM; 1: 3xTF32, everything in TMEM;
// mode 2: 3xTF32, TMEM accumulation restarted every chunk and flushed into fp32 registers (RN adds)
template <int CH>
__global__ void __launch_bounds__(128) k_test(const float* __restrict__ P, int R, int K, int N, float* __restrict__ D, int mode, int kahan) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t mbar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  constexpr int nkb = CH / 32;
  uint8_t* hi = base;                                   // [nkb][128 rows][32] swizzled
  uint8_t* lo = base + size_t(nkb) * TILE_BYTES_128;
  if (warp == 0) tmem_alloc(&tmem_base_s, 256);
  if (tid == 0) mbar_init(&mbar, 1);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = tmem_base_s;
  float racc[128], rcmp[128];
  for (int j = 0; j < 128; ++j) { racc[j] = 0.f; rcmp[j] = 0.f; }
  uint32_t phase = 0;
  for (int c = 0; c < K / CH; ++c) {
    for (int idx = tid; idx < 128 * CH; idx += 128) {
      const int r = idx / CH, k = idx % CH;
      const float v = (r < R) ? P[r * K + c * CH + k] : 0.f;
      float h, l;
      split_tf32(v, h, l);
      const uint32_t off = (k >> 5) * TILE_BYTES_128 + sw128_offset(r, k & 31);
      *reinterpret_cast<float*>(hi + off) = h;
      *reinterpret_cast<float*>(lo + off) = l;
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after_sync();
      const uint32_t idesc = make_idesc_tf32(128, N);
      for (int kb = 0; kb < nkb; ++kb)
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t dh = make_desc_sw128(smem_u32(hi) + kb * TILE_BYTES_128 + ks * 32);
          const uint64_t dl = make_desc_sw128(smem_u32(lo) + kb * TILE_BYTES_128 + ks * 32);
          const bool fresh = (kb == 0 && ks == 0) && (mode >= 2 || c == 0);
          mma_tf32_ss(tmem, dh, dh, idesc, !fresh);
          if (mode == 1 || mode == 2) { mma_tf32_ss(tmem, dh, dl, idesc, true); mma_tf32_ss(tmem, dl, dh, idesc, true); }
          if (mode == 3) { mma_tf32_ss(tmem + 128, dh, dl, idesc, !fresh); mma_tf32_ss(tmem + 128, dl, dh, idesc, true); }
        }
      mma_commit(&mbar);
    }
    mbar_wait(&mbar, phase);
    phase ^= 1;
    tc_fence_after_sync();
    if (mode >= 2) {
      for (int c0 = 0; c0 < N; c0 += 16) {
        uint32_t v[16], u[16];
        tmem_ld16(tmem + (uint32_t(warp * 32) << 16) + c0, v);
        if (mode == 3) tmem_ld16(tmem + (uint32_t(warp * 32) << 16) + 128 + c0, u);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float x = __uint_as_float(v[j]);
          if (!kahan) racc[c0 + j] += (mode == 3) ? (x + __uint_as_float(u[j])) : x;
          else {   // two-sum accumulation: racc + rcmp carries the exact running sum of the flushed chunks
            const float a = racc[c0 + j];
            const float sum = a + x;
            const float bb = sum - a;
            const float e = (a - (sum - bb)) + (x - bb);
            racc[c0 + j] = sum;
            rcmp[c0 + j] += e + ((mode == 3) ? __uint_as_float(u[j]) : 0.f);
          }
        }
      }
      tc_fence_before_sync();
    }
    __syncthreads();
  }
  for (int c0 = 0; c0 < N; c0 += 16) {
    uint32_t v[16];
    tmem_ld16(tmem + (uint32_t(warp * 32) << 16) + c0, v);
    tmem_ld_wait();
    for (int j = 0; j < 16; ++j) D[(warp * 32 + (tid & 31)) * N + c0 + j] = (mode >= 2) ? (racc[c0 + j] + rcmp[c0 + j]) : __uint_as_float(v[j]);
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

// fp32 SIMT reference with sequential fp32 accumulation (what the current SIMT Schur kernel does)
static float simt_dot(const float* a, const float* b, int K) { float s = 0.f; for (int k = 0; k < K; ++k) s = fmaf(a[k], b[k], s); return s; }

int main() {
  const int R = 102, K = 3328, N = 112;
  for (int positive = 0; positive < 3; ++positive) {
    std::vector<float> P(size_t(R) * K);
    srand(7);
    for (auto& v : P) { v = float(rand() / double(RAND_MAX) - (positive ? 0.0 : 0.5)) * 3.f; if (positive == 2) { uint32_t u; memcpy(&u, &v, 4); u &= 0xFFFFE000u; memcpy(&v, &u, 4); } }
    float *dP, *dD;
    CK(cudaMalloc(&dP, P.size() * 4)); CK(cudaMalloc(&dD, 128 * N * 4));
    CK(cudaMemcpy(dP, P.data(), P.size() * 4, cudaMemcpyHostToDevice));
    const size_t smem = 2 * size_t(64 / 32) * TILE_BYTES_128 + 1024;
    CK(cudaFuncSetAttribute(k_test<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    CK(cudaFuncSetAttribute(k_test<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    std::vector<double> ref(size_t(R) * R);
    double nrm = 0, e_simt = 0;
    for (int r = 0; r < R; ++r)
      for (int c = 0; c < R; ++c) {
        double s = 0;
        for (int k = 0; k < K; ++k) s += double(P[size_t(r) * K + k]) * double(P[size_t(c) * K + k]);
        ref[r * R + c] = s; nrm = fmax(nrm, fabs(s));
        e_simt = fmax(e_simt, fabs(s - simt_dot(&P[size_t(r) * K], &P[size_t(c) * K], K)));
      }
    printf("data %s, K=%d: fp32 sequential fma rel err %.3e\n", positive == 2 ? "positive, tf32-exact" : positive ? "positive" : "zero-mean", K, e_simt / nrm);
    for (int kahan = 0; kahan < 2; ++kahan)
    for (int mode = 2; mode < 6; ++mode) {
      CK(cudaMemset(dD, 0, 128 * N * 4));
      if (mode < 4) k_test<64><<<1, 128, smem>>>(dP, R, K, N, dD, mode, kahan);
      else k_test<32><<<1, 128, smem>>>(dP, R, K, N, dD, mode - 2, kahan);
      CK(cudaDeviceSynchronize());
      std::vector<float> D(128 * N);
      CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
      double err = 0;
      for (int r = 0; r < R; ++r) for (int c = 0; c < R; ++c) err = fmax(err, fabs(ref[r * R + c] - D[r * N + c]));
      printf("  kahan=%d mode=%d (%s)  max abs err %.3e  rel %.3e\n", kahan, mode, mode==0?"tf32":mode==1?"3x, all in TMEM":mode==2?"3x, flush 64":mode==3?"3x, 2 acc, flush 64":mode==4?"3x, flush 32":"3x, 2 acc, flush 32", err, err / nrm);
    }
    cudaFree(dP); cudaFree(dD);
  }
  return 0;
}
