// One-shot all-reduce of the reduced pose system over NVLink peer memory (sharded global BA).
//
// Every rank assembles its share of S (lower triangle + rhs row) into a slab of a buffer the other ranks of the box
// have mapped (cudaIpc).  ONE kernel per rank then (1) tells every peer "my slab of epoch e is written" (a flag store
// into the peer's memory), (2) waits until all peers said so, (3) reads the lower triangle of all N slabs and sums
// them in rank order 0..N-1 -- the same order on every rank, so the sum is bit-identical everywhere -- into the
// rank's private system matrix.  Against ncclAllReduce on the full square (r1: 54..66 us at 8 GPUs for 2.3 MB): half
// the bytes (only the triangle is read), one launch, no ring / tree steps; the latency is one flag round trip plus
// one remote read.  Slabs alternate by epoch: a rank can be at most one epoch ahead of the slowest reader.
#include "dba_kernels.h"

namespace dba {

namespace {

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer(const float4* p) {     // bypass L1: the line may be a stale copy of an older epoch
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

__global__ void __launch_bounds__(256)
k_p2p_allreduce(P2PPeers peers, int nranks, int rank, size_t slab_floats, size_t flags_off_floats, int parity,
                uint32_t epoch, int ld, int n, float* __restrict__ out, int* __restrict__ err, int debug) {
  __shared__ int s_bad;
  const int tid = threadIdx.x, lane = tid & 31;
  if (tid == 0) s_bad = 0;
  if (blockIdx.x == 0 && tid < nranks) {            // the previous kernel on this stream wrote my slab
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(peers.base[tid] + flags_off_floats) + rank, epoch);
  }
  __syncthreads();
  if (tid < nranks && !(debug & 1)) {                // every CTA waits for all ranks' flags in MY buffer
    const uint32_t* f = reinterpret_cast<const uint32_t*>(peers.base[rank] + flags_off_floats) + tid;
    const long long t0 = clock64();
    while (int32_t(*reinterpret_cast<const volatile uint32_t*>(f) - epoch) < 0) {    // polite polling: ~100 CTAs watch the line
      __nanosleep(200);
      if (clock64() - t0 > 6000000000LL) { s_bad = 1; break; }       // ~3 s: a peer died; do not hang the GPU
    }
    (void)ld_acquire_sys(f);                                         // order the slab reads after the flag
  }
  __syncthreads();
  if (s_bad && tid == 0) *err = 1;
  const int gw = blockIdx.x * 8 + (tid >> 5), nw = gridDim.x * 8;
  const size_t soff = size_t(parity) * slab_floats;
  if (debug & 2) return;
  for (int i = n - gw; i >= 0; i -= nw) {            // long rows first; row n = rhs (n columns), row i < n: i + 1 columns
    const int ncol = (i < n) ? i + 1 : n;
    const size_t roff = soff + size_t(i) * ld;
    for (int c4 = lane * 4; c4 < ncol; c4 += 256) {  // two chunks per pass, all ranks' loads in flight before the adds
      const bool two = c4 + 128 < ncol;
      float4 v0[8], v1[8];
#pragma unroll
      for (int r = 0; r < 8; ++r)
        if (r < nranks) {
          v0[r] = ld_peer(reinterpret_cast<const float4*>(peers.base[r] + roff + c4));
          if (two) v1[r] = ld_peer(reinterpret_cast<const float4*>(peers.base[r] + roff + c4 + 128));
        }
      float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
#pragma unroll
      for (int r = 0; r < 8; ++r)
        if (r < nranks) {
          a0.x += v0[r].x; a0.y += v0[r].y; a0.z += v0[r].z; a0.w += v0[r].w;
          if (two) { a1.x += v1[r].x; a1.y += v1[r].y; a1.z += v1[r].z; a1.w += v1[r].w; }
        }
      *reinterpret_cast<float4*>(out + size_t(i) * ld + c4) = a0;
      if (two) *reinterpret_cast<float4*>(out + size_t(i) * ld + c4 + 128) = a1;
    }
  }
}

// ---- two-shot all-reduce for the LARGE systems (n > p2p_max_n: loop-closure graphs, 18.8 MB of triangle at n = 3060).
// The one-shot kernel makes every rank read all N slabs (N x 18.8 MB inbound: slower than NCCL's ring).  Here the rows
// are cut into N ranges of equal element count; (A) rank r sums ITS rows over all slabs, in rank order, into the `red`
// region of its own exchange buffer; (B) every rank copies the N reduced ranges into its private system matrix.
// Inbound per rank: 2 (N - 1) / N of the triangle (33 MB at N = 8) and two flag round trips; the sum of a row is formed
// on ONE rank, so all ranks hold the same bits.  Ordering: `red` is rewritten by A(e + 1) only after every peer's slab
// flag e + 1 was seen, which a peer raises after its B(e) finished reading; slabs alternate by epoch as above.
struct P2PRows { int b[9]; };            // rows [b[r], b[r + 1]) belong to rank r (row n = the rhs row)

__device__ __forceinline__ bool p2p_wait_flags(const uint32_t* f, uint32_t epoch) {
  const long long t0 = clock64();
  while (int32_t(*reinterpret_cast<const volatile uint32_t*>(f) - epoch) < 0) {
    __nanosleep(200);
    if (clock64() - t0 > 6000000000LL) return false;
  }
  (void)ld_acquire_sys(f);
  return true;
}

__global__ void __launch_bounds__(256)
k_p2p_reduce_scatter(P2PPeers peers, int nranks, int rank, size_t slab_floats, size_t flags_off_floats, int parity,
                     uint32_t epoch, int ld, int n, P2PRows rows, int* __restrict__ err) {
  __shared__ int s_bad;
  const int tid = threadIdx.x, lane = tid & 31;
  if (tid == 0) s_bad = 0;
  if (blockIdx.x == 0 && tid < nranks) {            // the previous kernel on this stream wrote my slab
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(peers.base[tid] + flags_off_floats) + rank, epoch);
  }
  __syncthreads();
  if (tid < nranks && !p2p_wait_flags(reinterpret_cast<const uint32_t*>(peers.base[rank] + flags_off_floats) + tid, epoch)) s_bad = 1;
  __syncthreads();
  if (s_bad && tid == 0) *err = 1;
  const int gw = blockIdx.x * 8 + (tid >> 5), nw = gridDim.x * 8;
  const size_t soff = size_t(parity) * slab_floats;
  float* red = peers.base[rank] + 2 * slab_floats;
  for (int i = rows.b[rank + 1] - 1 - gw; i >= rows.b[rank]; i -= nw) {
    const int ncol = (i < n) ? i + 1 : n;
    const size_t roff = size_t(i) * ld;
    for (int c4 = lane * 4; c4 < ncol; c4 += 128) {
      float4 v[8];
#pragma unroll
      for (int r = 0; r < 8; ++r)
        if (r < nranks) v[r] = ld_peer(reinterpret_cast<const float4*>(peers.base[r] + soff + roff + c4));
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int r = 0; r < 8; ++r)
        if (r < nranks) { a.x += v[r].x; a.y += v[r].y; a.z += v[r].z; a.w += v[r].w; }
      *reinterpret_cast<float4*>(red + roff + c4) = a;
    }
  }
}

__global__ void __launch_bounds__(256)
k_p2p_allgather(P2PPeers peers, int nranks, int rank, size_t slab_floats, size_t flags_off_floats, uint32_t epoch, int ld,
                int n, P2PRows rows, float* __restrict__ out, int* __restrict__ err) {
  __shared__ int s_bad;
  const int tid = threadIdx.x, lane = tid & 31;
  if (tid == 0) s_bad = 0;
  if (blockIdx.x == 0 && tid < nranks) {            // the previous kernel on this stream wrote my reduced rows
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(peers.base[tid] + flags_off_floats) + 16 + rank, epoch);
  }
  __syncthreads();
  if (tid < nranks && !p2p_wait_flags(reinterpret_cast<const uint32_t*>(peers.base[rank] + flags_off_floats) + 16 + tid, epoch)) s_bad = 1;
  __syncthreads();
  if (s_bad && tid == 0) *err = 1;
  const int gw = blockIdx.x * 8 + (tid >> 5), nw = gridDim.x * 8;
  for (int i = n - gw; i >= 0; i -= nw) {
    int q = 0;
    while (q + 1 < nranks && i >= rows.b[q + 1]) ++q;
    const int ncol = (i < n) ? i + 1 : n;
    const float* src = peers.base[q] + 2 * slab_floats + size_t(i) * ld;
    float* dst = out + size_t(i) * ld;
    for (int c4 = lane * 4; c4 < ncol; c4 += 256) {
      const bool two = c4 + 128 < ncol;
      const float4 a0 = ld_peer(reinterpret_cast<const float4*>(src + c4));
      float4 a1 = a0;
      if (two) a1 = ld_peer(reinterpret_cast<const float4*>(src + c4 + 128));
      *reinterpret_cast<float4*>(dst + c4) = a0;
      if (two) *reinterpret_cast<float4*>(dst + c4 + 128) = a1;
    }
  }
}

// NCCL path of the large systems: only the lower triangle + rhs row carry data, so they are packed into a contiguous
// buffer (row i at i (i + 1) / 2, the rhs row behind the triangle), all-reduced and unpacked: half the bytes on NVLink.
template <bool PACK>
__global__ void __launch_bounds__(256)
k_tri_pack(float* __restrict__ A, float* __restrict__ buf, int ld, int n) {
  const int lane = threadIdx.x & 31;
  const int gw = blockIdx.x * 8 + (threadIdx.x >> 5), nw = gridDim.x * 8;
  for (int i = gw; i <= n; i += nw) {
    const int ncol = (i < n) ? i + 1 : n;
    float* row = A + size_t(i) * ld;
    float* dst = buf + (size_t(i) * (i + 1)) / 2;          // i = n: n (n + 1) / 2 = the rhs row's slot
    for (int c = lane; c < ncol; c += 32) {
      if (PACK) dst[c] = row[c]; else row[c] = dst[c];
    }
  }
}

}  // namespace

cudaError_t launch_p2p_allreduce(const P2PPeers& peers, int nranks, int rank, size_t slab_floats, size_t flags_off_floats,
                                 int parity, uint32_t epoch, int ld, int n, float* out, int* err, int debug, cudaStream_t st) {
  int grid = (n + 1 + 7) / 8;
  if (grid > 296) grid = 296;
  k_p2p_allreduce<<<grid, 256, 0, st>>>(peers, nranks, rank, slab_floats, flags_off_floats, parity, epoch, ld, n, out, err, debug);
  g_launch_count++;
  return cudaGetLastError();
}

// rows cut into nranks ranges of (nearly) equal element count; two launches on `st`
cudaError_t launch_p2p_allreduce_two_shot(const P2PPeers& peers, int nranks, int rank, size_t slab_floats,
                                          size_t flags_off_floats, int parity, uint32_t epoch, int ld, int n, float* out,
                                          int* err, cudaStream_t st) {
  P2PRows rows;
  const double total = double(n) * (n + 1) / 2 + n;
  int q = 1;
  double acc = 0.0;
  rows.b[0] = 0;
  for (int i = 0; i <= n && q < nranks; ++i) {
    acc += (i < n) ? double(i + 1) : double(n);
    while (q < nranks && acc >= total * q / nranks) rows.b[q++] = i + 1;
  }
  for (; q <= 8; ++q) rows.b[q] = n + 1;
  for (int r = nranks; r <= 8; ++r) rows.b[r] = n + 1;
  int grid = (n + 1 + 7) / 8;
  if (grid > 296) grid = 296;
  int gridA = ((n + 1) / nranks + 7) / 8 + 1;
  if (gridA > 296) gridA = 296;
  k_p2p_reduce_scatter<<<gridA, 256, 0, st>>>(peers, nranks, rank, slab_floats, flags_off_floats, parity, epoch, ld, n, rows, err);
  g_launch_count++;
  k_p2p_allgather<<<grid, 256, 0, st>>>(peers, nranks, rank, slab_floats, flags_off_floats, epoch, ld, n, rows, out, err);
  g_launch_count++;
  return cudaGetLastError();
}

}  // namespace dba

namespace dba {
size_t tri_pack_floats(int n) { return size_t(n) * (n + 1) / 2 + size_t(n); }
cudaError_t launch_tri_pack(float* A, float* buf, int ld, int n, bool pack, cudaStream_t st) {
  int grid = (n + 1 + 7) / 8;
  if (grid > 592) grid = 592;
  if (pack) k_tri_pack<true><<<grid, 256, 0, st>>>(A, buf, ld, n);
  else k_tri_pack<false><<<grid, 256, 0, st>>>(A, buf, ld, n);
  g_launch_count++;
  return cudaGetLastError();
}
}  // namespace dba
