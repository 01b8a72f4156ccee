// Hand-written tcgen05 (5th-gen tensor core) helpers for sm_100a: shared-memory matrix descriptors,
// the kind::tf32 instruction descriptor, TMEM allocation / loads, mbarrier plumbing and the
// SWIZZLE_128B K-major tile layout our own threads write with st.shared (no TMA: the operands are
// produced on chip by the linearisation, they never exist in global memory).
//
// fp32-faithful products on the tf32 pipe ("3xTF32"): every operand value v is stored as
// hi = tf32(v) and lo = v - hi; D += hi*hi' + hi*lo' + lo*hi' recovers ~22 mantissa bits with fp32
// accumulation in TMEM (the dropped lo*lo' term is 2^-22 relative).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace dba {
namespace umma {

// One K-block of an operand tile: ROWS rows x 32 fp32 (128 bytes per row), 8-row groups of 1024 B,
// 16-byte chunks XOR-swizzled with the row index inside the group (Swizzle<3,4,3>).
constexpr uint32_t TILE_BYTES_128 = 128 * 128;   // 128 rows
constexpr uint32_t TILE_BYTES_256 = 256 * 128;   // 256 rows

// byte offset of element (row r, k in [0,32)) inside one K-block (base must be 1024-byte aligned)
__host__ __device__ __forceinline__ uint32_t sw128_offset(int r, int k) {
  return uint32_t(r >> 3) * 1024u + uint32_t(r & 7) * 128u + ((uint32_t(k >> 2) ^ uint32_t(r & 7)) << 4) + uint32_t(k & 3) * 4u;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

// tf32 split: hi = round-to-nearest tf32 of v (low 13 mantissa bits zero), lo = v - hi (exact in fp32)
__device__ __forceinline__ void split_tf32(float v, float& hi, float& lo) {
  uint32_t h;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(v));
  hi = __uint_as_float(h);
  lo = v - hi;
}

// Shared-memory matrix descriptor, K-major, SWIZZLE_128B: start address >> 4 in bits [0,14),
// leading byte offset (unused for swizzled K-major; 1) in [16,30), stride byte offset = 1024 B between
// 8-row groups in [32,46), descriptor version 1 (sm_100) in [46,48), layout type 2 in [61,64).
__device__ __forceinline__ uint64_t make_desc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= uint64_t((saddr & 0x3FFFFu) >> 4);
  d |= uint64_t(1) << 16;
  d |= uint64_t(1024 >> 4) << 32;
  d |= uint64_t(1) << 46;
  d |= uint64_t(2) << 61;
  return d;
}

// Instruction descriptor for kind::tf32, fp32 accumulate, A and B K-major:
// c_format F32 (1) at [4,6), a_format / b_format TF32 (2) at [7,10) / [10,13), N>>3 at [17,23), M>>4 at [24,29)
__host__ __device__ __forceinline__ uint32_t make_idesc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(N >> 3) << 17) | (uint32_t(M >> 4) << 24);
}

// D[tmem] (+)= A[smem] * B[smem]^T ; issued by ONE thread
__device__ __forceinline__ void mma_tf32_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, bool accumulate) {
  const uint32_t acc = accumulate ? 1u : 0u;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(acc) : "memory");
}

// all previously issued MMAs of this thread arrive on the mbarrier when they complete
// (implies tcgen05.fence::before_thread_sync)
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t a = smem_u32(bar);
  uint32_t done;
  do {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}\n" : "=r"(done) : "r"(a), "r"(parity) : "memory");
  } while (!done);
}

// generic-proxy writes to shared memory -> visible to the async proxy (tensor core operand fetch)
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// TMEM allocation: one full warp; ncols a power of two in [32,512]; the base address lands in *dst (smem)
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

// 32 lanes x 16 consecutive columns: thread i of the warp gets lane (base_lane + i); the lane field
// (bits 31:16 of taddr) must be the warp's quarter: 32 * (warp_id % 4)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

}  // namespace umma
}  // namespace dba
