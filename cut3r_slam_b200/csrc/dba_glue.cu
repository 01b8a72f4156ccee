// The two elementwise steps either side of `ba` inside FactorGraph.update:
//   reproject(ii, jj) -> coords [E,ht,wd,2], valid [E,ht,wd,1]
//       hislam2/depth_video.pyc reproject@170-178 = geom/projective_ops.py:44-74 without Jacobians
//   cvx_upsample(data, mask) -> [B,8ht,8wd,dim]
//       hislam2/modules/droid_net.py:21-38: per fine pixel a softmax over the 3x3 coarse neighbourhood (logits
//       mask[B,9,8,8,ht,wd]; neighbours outside the image get -inf) and the convex combination of the neighbours
// Both are HBM-bound streaming kernels: every input element is read once with full 128-byte warp transactions,
// every output written once (the upsampler transposes its 8x8 sub-pixel block through shared memory so that the
// rows of the fine image leave as contiguous runs).
#include "../../include/dba.h"

#include <climits>
#include <mutex>

#include "dba_kernels.h"
#include "dba_math.cuh"

namespace dba {

namespace {

constexpr int RP_T = 256;

// grid (nsplit, E): one relative pose per CTA, the CTA strides over its share of the pixels
__global__ void __launch_bounds__(RP_T)
k_reproject(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr,
            const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, int N, int HW, int wd,
            float2* __restrict__ coords, float* __restrict__ valid) {
  __shared__ float sRt[12];
  const int e = blockIdx.y;
  const int64_t i64 = ii[e], j64 = jj[e];
  if (i64 < 0 || i64 >= N || j64 < 0 || j64 >= N) {        // an index outside the frame buffer: defined output, no read
    for (int p = blockIdx.x * RP_T + threadIdx.x; p < HW; p += gridDim.x * RP_T) {
      coords[int64_t(e) * HW + p] = make_float2(0.f, 0.f);
      valid[int64_t(e) * HW + p] = 0.f;
    }
    return;
  }
  const int i = int(i64), j = int(j64);
  if (threadIdx.x == 0) relative_pose_Rt(poses, i, j, sRt);
  __syncthreads();
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const float* dsrc = disps + int64_t(i) * HW;
  for (int p = blockIdx.x * RP_T + threadIdx.x; p < HW; p += gridDim.x * RP_T) {
    const int py = p / wd, px = p - py * wd;
    const Proj pr = project(sRt, (float(px) - cx) / fx, (float(py) - cy) / fy, __ldg(dsrc + p), fx, fy, cx, cy);
    __stcs(coords + int64_t(e) * HW + p, make_float2(pr.u, pr.v));
    __stcs(valid + int64_t(e) * HW + p, pr.valid);
  }
}

// reproject + motion features in one pass (FactorGraph.update, factor_graph.pyc update@195-197):
//   coords1 = reproject(ii, jj);  motn = cat([coords1 - coords0, target - coords1], -1).permute(0,1,4,2,3).clamp(-64, 64)
// coords0 is the pixel grid.  Outputs: coords [E,HW] float2, valid [E,HW], motn [E,4,HW] (four coalesced planes).
__global__ void __launch_bounds__(RP_T)
k_motion_features(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr,
                  const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, const float2* __restrict__ target,
                  int N, int HW, int wd, float2* __restrict__ coords, float* __restrict__ valid,
                  float* __restrict__ motn) {
  __shared__ float sRt[12];
  const int e = blockIdx.y;
  const int64_t i64 = ii[e], j64 = jj[e];
  const bool ok = i64 >= 0 && i64 < N && j64 >= 0 && j64 < N;
  if (ok && threadIdx.x == 0) relative_pose_Rt(poses, int(i64), int(j64), sRt);
  __syncthreads();
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const float* dsrc = disps + (ok ? i64 : 0) * HW;
  for (int p = blockIdx.x * RP_T + threadIdx.x; p < HW; p += gridDim.x * RP_T) {
    const int py = p / wd, px = p - py * wd;
    float u = 0.f, v = 0.f, val = 0.f;
    if (ok) {
      const Proj pr = project(sRt, (float(px) - cx) / fx, (float(py) - cy) / fy, __ldg(dsrc + p), fx, fy, cx, cy);
      u = pr.u; v = pr.v; val = pr.valid;
    }
    const float2 tg = __ldcs(target + int64_t(e) * HW + p);
    __stcs(coords + int64_t(e) * HW + p, make_float2(u, v));
    __stcs(valid + int64_t(e) * HW + p, val);
    float* mo = motn + int64_t(e) * 4 * HW + p;
    __stcs(mo, fminf(fmaxf(u - float(px), -64.f), 64.f));
    __stcs(mo + HW, fminf(fmaxf(v - float(py), -64.f), 64.f));
    __stcs(mo + 2 * HW, fminf(fmaxf(tg.x - u, -64.f), 64.f));
    __stcs(mo + 3 * HW, fminf(fmaxf(tg.y - v, -64.f), 64.f));
  }
}

// damping of the edges' source keyframes (factor_graph.pyc update@214,227):
//   eta = 0.2 * damping[unique(ii)] + 1e-7, rows in sorted(unique(ii)) order.
// One CTA marks the frames that occur in ii and ranks them (slot[f] = row of frame f, -1 if absent) ...
__global__ void __launch_bounds__(1024)
k_source_slots(const int64_t* __restrict__ ii, int E, int N, int32_t* __restrict__ slot) {
  extern __shared__ int sflag[];                  // [N] flags, then an inclusive scan in place
  for (int f = threadIdx.x; f < N; f += blockDim.x) sflag[f] = 0;
  __syncthreads();
  for (int e = threadIdx.x; e < E; e += blockDim.x) {
    const int64_t f = ii[e];
    if (f >= 0 && f < N) sflag[f] = 1;
  }
  __syncthreads();
  if (threadIdx.x == 0) {                         // N is a few hundred to a few thousand: a serial scan is microseconds
    int run = 0;
    for (int f = 0; f < N; ++f) { const int fl = sflag[f]; slot[f] = fl ? run : -1; run += fl; }
  }
}
// ... and every marked frame's map is scaled into its row
__global__ void __launch_bounds__(256)
k_damping_gather(const float* __restrict__ damping, const int32_t* __restrict__ slot, int HW, int M,
                 float* __restrict__ eta) {
  const int f = blockIdx.y;
  const int r = slot[f];
  if (r < 0 || r >= M) return;
  const int p = blockIdx.x * 256 + threadIdx.x;
  if (p < HW) eta[int64_t(r) * HW + p] = 0.2f * __ldg(damping + int64_t(f) * HW + p) + 1e-7f;
}

constexpr int UP_T = 256;          // 8 sub-columns j x 32 coarse pixels

// NC channels per pass (1, 2 or 4); channels [c0, c0 + nc) of `dim`
template <int NC>
__global__ void __launch_bounds__(UP_T, 3)
k_cvx_upsample(const float* __restrict__ data, const float* __restrict__ mask, int ht, int wd, int dim, int c0, int nc,
               float* __restrict__ up) {
  __shared__ float srow[8][32 * 8 * NC + 4];
  const int b = blockIdx.z, y = blockIdx.y, x0 = blockIdx.x * 32;
  const int tid = threadIdx.x, j = tid >> 5, xx = tid & 31, x = x0 + xx;
  const int nx = min(32, wd - x0);
  const bool act = xx < nx;
  const int64_t plane = int64_t(ht) * wd;
  unsigned inside = 0;
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    const int yy = y + k / 3 - 1, xk = x + k % 3 - 1;
    if (act && yy >= 0 && yy < ht && xk >= 0 && xk < wd) inside |= 1u << k;
  }
  float nb[9][NC];
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    const int yy = y + k / 3 - 1, xk = x + k % 3 - 1;
#pragma unroll
    for (int c = 0; c < NC; ++c)
      nb[k][c] = ((inside >> k & 1u) && c < nc) ? __ldg(data + ((int64_t(b) * ht + yy) * wd + xk) * dim + c0 + c) : 0.f;
  }
  // mask viewed as [B,9,8,8,ht,wd] (droid_net.py:24): element (b,k,i,j,y,x)
  const float* mb = mask + (int64_t(b) * 9 * 64 + j) * plane + int64_t(y) * wd + x;
#pragma unroll 2
  for (int i = 0; i < 8; ++i) {
    float l[9];
#pragma unroll
    for (int k = 0; k < 9; ++k)
      l[k] = (inside >> k & 1u) ? __ldcs(mb + int64_t((k * 8 + i) * 8) * plane) : -INFINITY;
    float mx = l[4];                                     // the centre is always inside
#pragma unroll
    for (int k = 0; k < 9; ++k) mx = fmaxf(mx, l[k]);
    float den = 0.f, acc[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) acc[c] = 0.f;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      // ex2.approx of a non-positive argument: relative error <= 3e-6 on terms that are themselves <= 1 (the largest
      // term is exactly 1), i.e. <= 3e-7 of the convex combination -- inside the 2e-6 parity tolerance
      const float ev = (inside >> k & 1u) ? __expf(l[k] - mx) : 0.f;
      den += ev;
#pragma unroll
      for (int c = 0; c < NC; ++c) acc[c] = fmaf(ev, nb[k][c], acc[c]);
    }
    const float inv = 1.0f / den;
    if (act) {
#pragma unroll
      for (int c = 0; c < NC; ++c)
        if (c < nc) srow[i][(xx * 8 + j) * nc + c] = acc[c] * inv;
    }
  }
  __syncthreads();
  // fine rows 8y .. 8y+7, columns 8x0 .. 8x0 + 8nx - 1: contiguous runs when one pass covers all channels
  const int run = nx * 8 * nc;
  for (int i = 0; i < 8; ++i) {
    float* dst = up + ((int64_t(b) * 8 * ht + 8 * y + i) * 8 * wd + 8 * x0) * dim;
    for (int t = tid; t < run; t += UP_T) {
      const int col = t / nc, c = t - col * nc;
      __stcs(dst + int64_t(col) * dim + c0 + c, srow[i][t]);
    }
  }
}

// dim == 1 (disparities, the case FactorGraph.update runs): one thread per coarse pixel of the FLATTENED image, so
// that a CTA reads one contiguous 1 KB run from each of the 576 logit planes (no partially filled tiles at the row
// ends) with 36 independent loads in flight per thread, and writes each fine row as 16-byte pieces that
// interleave to contiguous runs across the warp -- no shared memory, no barriers.
__global__ void __launch_bounds__(256, 3)
k_cvx_upsample_flat(const float* __restrict__ data, const float* __restrict__ mask, int ht, int wd,
                    float* __restrict__ up) {
  const int b = blockIdx.y;
  const int plane = ht * wd;
  const int p = blockIdx.x * 256 + threadIdx.x;
  if (p >= plane) return;
  const int y = p / wd, x = p - y * wd;
  unsigned inside = 0;
  float nb[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    const int yy = y + k / 3 - 1, xk = x + k % 3 - 1;
    const bool in = yy >= 0 && yy < ht && xk >= 0 && xk < wd;
    if (in) inside |= 1u << k;
    nb[k] = in ? __ldg(data + (int64_t(b) * ht + yy) * wd + xk) : 0.f;
  }
  const float* mb = mask + int64_t(b) * 576 * plane + p;           // element (b,k,i,j,y,x) of [B,9,8,8,ht,wd]
  float* ub = up + (int64_t(b) * 8 * ht + 8 * y) * 8 * wd + 8 * x;
#pragma unroll 1
  for (int ih = 0; ih < 16; ++ih) {                        // (fine row i, half of its 8 sub-columns)
    const int i = ih >> 1, j0 = (ih & 1) * 4;
    float l[9][4];
#pragma unroll
    for (int k = 0; k < 9; ++k)
#pragma unroll
      for (int j = 0; j < 4; ++j)
        l[k][j] = (inside >> k & 1u) ? __ldcs(mb + int64_t((k * 8 + i) * 8 + j0 + j) * plane) : -INFINITY;
    float o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float mx = l[4][j];                                 // the centre is always inside
#pragma unroll
      for (int k = 0; k < 9; ++k) mx = fmaxf(mx, l[k][j]);
      float den = 0.f, acc = 0.f;
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        const float ev = (inside >> k & 1u) ? __expf(l[k][j] - mx) : 0.f;   // see k_cvx_upsample on the accuracy
        den += ev;
        acc = fmaf(ev, nb[k], acc);
      }
      o[j] = acc / den;
    }
    // 8x + j0 floats into fine row 8y + i: 16-byte aligned; the warp's pieces of a row are contiguous
    __stcs(reinterpret_cast<float4*>(ub + int64_t(i) * 8 * wd + j0), make_float4(o[0], o[1], o[2], o[3]));
  }
}

}  // namespace

cudaError_t launch_reproject(const float* poses, const float* disps, const float* intr, const int64_t* ii,
                             const int64_t* jj, int E, int N, int HW, int wd, float* coords, float* valid, cudaStream_t st) {
  if (E <= 0) return cudaSuccess;
  // about 16 CTAs per SM in total; an edge is never split finer than one pixel per thread
  int nsplit = (148 * 16 + E - 1) / E;
  const int maxsplit = (HW + RP_T - 1) / RP_T;
  if (nsplit > maxsplit) nsplit = maxsplit;
  if (nsplit < 1) nsplit = 1;
  dim3 grid(nsplit, E);
  k_reproject<<<grid, RP_T, 0, st>>>(poses, disps, intr, ii, jj, N, HW, wd, reinterpret_cast<float2*>(coords), valid);
  g_launch_count++;
  return cudaGetLastError();
}

cudaError_t launch_motion_features(const float* poses, const float* disps, const float* intr, const int64_t* ii,
                                   const int64_t* jj, const float* target, int E, int N, int HW, int wd, float* coords,
                                   float* valid, float* motn, cudaStream_t st) {
  int nsplit = (HW + RP_T * 4 - 1) / (RP_T * 4);
  if (nsplit < 1) nsplit = 1;
  k_motion_features<<<dim3(nsplit, E), RP_T, 0, st>>>(poses, disps, intr, ii, jj, reinterpret_cast<const float2*>(target),
                                                      N, HW, wd, reinterpret_cast<float2*>(coords), valid, motn);
  g_launch_count++;
  return cudaGetLastError();
}

cudaError_t launch_damping_gather(const float* damping, const int64_t* ii, int E, int N, int HW, int M, int32_t* slot,
                                  float* eta, cudaStream_t st) {
  k_source_slots<<<1, 1024, size_t(N) * sizeof(int), st>>>(ii, E, N, slot);
  g_launch_count++;
  k_damping_gather<<<dim3((HW + 255) / 256, N), 256, 0, st>>>(damping, slot, HW, M, eta);
  g_launch_count++;
  return cudaGetLastError();
}

cudaError_t launch_cvx_upsample(const float* data, const float* mask, int B, int ht, int wd, int dim, float* up,
                                cudaStream_t st) {
  if (B <= 0) return cudaSuccess;
  if (dim == 1 && (reinterpret_cast<uintptr_t>(up) & 15u) == 0) {
    dim3 gridf((ht * wd + 255) / 256, B);
    k_cvx_upsample_flat<<<gridf, 256, 0, st>>>(data, mask, ht, wd, up);
    g_launch_count++;
    return cudaGetLastError();
  }
  dim3 grid((wd + 31) / 32, ht, B);
  for (int c0 = 0; c0 < dim; c0 += 4) {
    const int nc = dim - c0 < 4 ? dim - c0 : 4;
    if (nc == 1) k_cvx_upsample<1><<<grid, UP_T, 0, st>>>(data, mask, ht, wd, dim, c0, nc, up);
    else if (nc == 2) k_cvx_upsample<2><<<grid, UP_T, 0, st>>>(data, mask, ht, wd, dim, c0, nc, up);
    else k_cvx_upsample<4><<<grid, UP_T, 0, st>>>(data, mask, ht, wd, dim, c0, nc, up);
    g_launch_count++;
  }
  return cudaGetLastError();
}

}  // namespace dba

// ---- dba_damping_gather_noalloc: the damping gather for callers that must not allocate (code under CUDA graph
// capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Same two kernels as dba_damping_gather; the frame -> row map (the only scratch of this op)
// lives in a file-scope device array instead of the context's growable buffer.
//
// kNoAllocMaxFrames * 4 bytes = 48 000 bytes are held by every GPU that loads the library, until the process exits.
// Every call on a device shares the array: k_source_slots rewrites entries [0, N) in full before k_damping_gather
// reads them, so nothing has to start at zero and stale contents of an earlier call are never read -- but two calls
// that overlap on the GPU (different streams, no ordering between them) race on it.
namespace dba {
namespace {
constexpr int kNoAllocMaxFrames = 12000;          // = the frame limit of dba_damping_gather
constexpr int kNoAllocMaxDevices = 64;
static __device__ __align__(256) int32_t g_noalloc_slot[kNoAllocMaxFrames];

// In host code `g_noalloc_slot` is only a stand-in for the symbol, not a device address: the address is looked up
// once per device, with that device current, and cached.
int32_t* g_noalloc_slot_addr[kNoAllocMaxDevices] = {};
std::mutex g_noalloc_mutex;

cudaError_t noalloc_slot_address(int device, int32_t** out) {
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (g_noalloc_slot_addr[device] == nullptr) {
    int32_t* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_slot);
    if (e != cudaSuccess) return e;
    if (p == nullptr) return cudaErrorInvalidSymbol;
    g_noalloc_slot_addr[device] = p;
  }
  *out = g_noalloc_slot_addr[device];
  return cudaSuccess;
}
}  // namespace
}  // namespace dba

extern "C" int dba_damping_gather_noalloc(dba_ctx* ctx, const float* damping, const int64_t* ii, int64_t E, int64_t N,
                                          int ht, int wd, int64_t M, float* eta_out, void* stream) {
  using namespace dba;
  if (!ctx) return -1;
  if (E < 0 || N <= 0 || ht <= 0 || wd <= 0 || M < 0) return ctx_fail_arg(ctx, "negative or zero size");
  if (N > kNoAllocMaxFrames)
    return ctx_fail_arg(ctx, "dba_damping_gather_noalloc: frame buffer larger than the static scratch (12000 frames)");
  if (E > INT32_MAX || int64_t(ht) * wd > INT32_MAX || M > INT32_MAX)
    return ctx_fail_arg(ctx, "dba_damping_gather_noalloc: size beyond int32");
  if (E == 0 || M == 0) return 0;
  if (!damping || !ii || !eta_out) return ctx_fail_arg(ctx, "NULL argument");
  const int device = ctx_device(ctx);
  if (device < 0 || device >= kNoAllocMaxDevices) return ctx_fail_arg(ctx, "dba_damping_gather_noalloc: device index beyond 63");
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return ctx_fail_cuda(ctx, e, "cudaSetDevice");
  int32_t* slot = nullptr;
  e = noalloc_slot_address(device, &slot);
  if (e != cudaSuccess) return ctx_fail_cuda(ctx, e, "cudaGetSymbolAddress(g_noalloc_slot)");
  g_launch_count = 0;
  e = launch_damping_gather(damping, ii, int(E), int(N), ht * wd, int(M), slot, eta_out,
                            reinterpret_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return ctx_fail_cuda(ctx, e, "damping_gather_noalloc");
  ctx_set_launches(ctx, g_launch_count);
  return 0;
}
