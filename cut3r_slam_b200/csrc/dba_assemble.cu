// Assembly of the reduced pose system from the per-CTA partial sums (fixed-order gathers, no
// atomics), back-substitution of the disparity update and the retractions.
//
// Reference maths: hislam2/geom/ba.py:54-59 (Hii,Hij,Hji,Hjj,vi,vj), :70-92 (scatter into H,v;
// C += eta + 1e-7), chol.py:55-57 (true-diagonal LM damping), :63-64 (S = H - E Q E^T,
// v' = v - E Q w), :70 (dz = Q (w - E^T dx)), ba.py:101-105 (retractions + disparity guards).
#include "dba_kernels.h"
#include "dba_math.cuh"
#include "dba_plan.h"

namespace dba {

namespace {

// ---------------------------------------------------------------------------------------------
// per edge: sum the split partials of Hjj / vj, expand to the four 6x6 blocks with the adjoint
// ---------------------------------------------------------------------------------------------
constexpr int EF_WARPS = 8;
__device__ __forceinline__ double sum_splits(const float* __restrict__ src, int64_t st, int S);

// one warp per edge: lanes 0..26 sum the split partials (fp64: a handful of adds, removes one layer of
// fp32 rounding), the 4 x 36 + 2 x 6 outputs are spread over the lanes
// one warp finalises edge c: sH / sHij / sAd [36], sv [8] are the warp's shared scratch
__device__ __forceinline__ void edge_finalize_warp(const float* __restrict__ poses, const int32_t* __restrict__ ci,
                                                   const int32_t* __restrict__ cj, const float* __restrict__ Hpart, int S,
                                                   float* __restrict__ Hblk, float* __restrict__ vblk, int c, int lane,
                                                   float* sH, float* sHij, float* sAd, float* sv) {
  {
    const float hv = float(sum_splits(Hpart + int64_t(c) * S * 32 + lane, 32, S));
    // lane l < 21 holds H[r][cc] with l = r(r+1)/2 + cc (cc <= r); lanes 21..26 hold vj
    if (lane < 21) {
      int r = 0;
      while ((r + 1) * (r + 2) / 2 <= lane) ++r;
      const int cc = lane - r * (r + 1) / 2;
      sH[r * 6 + cc] = hv;
      sH[cc * 6 + r] = hv;
    } else if (lane < 27) {
      sv[lane - 21] = hv;
    }
  }
  if (lane == 31) {      // an otherwise idle lane builds the adjoint of the edge's relative pose
    float AdR[36], Rt[12];
    relative_pose_Rt(poses, ci[c], cj[c], Rt);
    adjoint_from_Rt(Rt, AdR);
#pragma unroll
    for (int i = 0; i < 36; ++i) sAd[i] = AdR[i];
  }
  __syncwarp();
  const float* H = sH;
  const float* Ad = sAd;
  float* out = Hblk + int64_t(c) * 144;
  for (int e = lane; e < 36; e += 32) {          // Hij = -Ad^T Hjj ; Hji = -Hjj Ad ; Hjj
    const int r = e / 6, cc = e - r * 6;
    float a = 0.f, hji = 0.f;
#pragma unroll
    for (int t = 0; t < 6; ++t) {
      a += Ad[t * 6 + r] * H[t * 6 + cc];
      hji += H[r * 6 + t] * Ad[t * 6 + cc];
    }
    sHij[e] = -a;
    out[1 * 36 + e] = -a;
    out[2 * 36 + e] = -hji;
    out[3 * 36 + e] = H[e];
  }
  __syncwarp();
  for (int e = lane; e < 36; e += 32) {          // Hii = Ad^T Hjj Ad = -(Hij Ad)
    const int r = e / 6, cc = e - r * 6;
    float hii = 0.f;
#pragma unroll
    for (int t = 0; t < 6; ++t) hii += sHij[r * 6 + t] * Ad[t * 6 + cc];
    out[e] = -hii;
  }
  if (lane < 6) {
    float a = 0.f;
#pragma unroll
    for (int t = 0; t < 6; ++t) a += Ad[t * 6 + lane] * sv[t];
    vblk[int64_t(c) * 12 + lane] = -a;                   // vi = -Ad^T vj
    vblk[int64_t(c) * 12 + 6 + lane] = sv[lane];   // vj
  }
}

__global__ void __launch_bounds__(32 * EF_WARPS)
k_edge_finalize(const float* __restrict__ poses, const int32_t* __restrict__ ci, const int32_t* __restrict__ cj,
                const float* __restrict__ Hpart, int E, int S, float* __restrict__ Hblk, float* __restrict__ vblk) {
  __shared__ float sH[EF_WARPS][36], sHij[EF_WARPS][36], sAd[EF_WARPS][36], sv[EF_WARPS][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * EF_WARPS + warp;
  if (c >= E) return;
  edge_finalize_warp(poses, ci, cj, Hpart, S, Hblk, vblk, c, lane, sH[warp], sHij[warp], sAd[warp], sv[warp]);
}

// ---------------------------------------------------------------------------------------------
// S (lower block triangle) = damp(sum H) - sum Schur.  One CTA per 6x6 block: 8 item lanes x 36
// entries; item lane l gathers items l, l+8, ... and the 8 partials are summed in lane order
// (fixed order -> reproducible).  Diagonal blocks carry ~50 items, off-diagonal ones a handful.
// ---------------------------------------------------------------------------------------------
constexpr int ASM_LANES = 8;

// sum of the S split partials of one entry, eight loads in flight, added in split order (fp64)
__device__ __forceinline__ double sum_splits(const float* __restrict__ src, int64_t st, int S) {
  double t = 0.0;
  int s = 0;
  for (; s + 8 <= S; s += 8) {
    float v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = __ldcg(src + (s + u) * st);
#pragma unroll
    for (int u = 0; u < 8; ++u) t += double(v[u]);
  }
  for (; s < S; ++s) t += double(__ldcg(src + s * st));
  return t;
}

// gathers of one 6x6 block entry: items i0+il, i0+il+nl, ... (fp64 sums: S = H - E Q E^T cancels several
// digits, so the only roundings kept are those of the fp32 partials themselves and the final store)
__device__ __forceinline__ void gather_block(const Item* __restrict__ items, int i0, int i1, int il, int nl, int ent,
                                             int entT, const float* __restrict__ Hblk,
                                             const float* __restrict__ Spart, int S, double& hsum, double& ssum) {
  for (int i = i0 + il; i < i1; i += nl) {
    const Item it = items[i];
    const int e = (it.flags & 2) ? entT : ent;
    if ((it.flags & 1) == 0) {
      hsum += double(__ldcg(Hblk + int64_t(it.base) * 36 + e));
    } else {
      const float* src = Spart + int64_t(it.base) * 36 + e;
      const int64_t st = int64_t(it.stride) * 36;
      ssum += sum_splits(src, st, S);
    }
  }
}

// rhs row of the augmented system: v - E Q w for pose a.  16 item lanes x 6 entries (threads >= 96 of the
// CTA only take part in the barrier); runs as the trailing CTAs of k_assemble_S.
constexpr int RHS_LANES = 16;

__device__ __forceinline__ void assemble_rhs_cta(int a, const int32_t* __restrict__ off, const Item* __restrict__ items,
                                                 const float* __restrict__ vblk, const float* __restrict__ vEpart,
                                                 int S, int Pp, float* __restrict__ A, int ldA) {
  __shared__ double sv[RHS_LANES][6], se[RHS_LANES][6];
  const int il = threadIdx.x / 6, r = threadIdx.x % 6;
  const bool active = il < RHS_LANES;
  double vsum = 0.0, esum = 0.0;
  if (active) {
    for (int i = off[a] + il; i < off[a + 1]; i += RHS_LANES) {
      const Item it = items[i];
      if ((it.flags & 1) == 0) {
        vsum += double(__ldcg(vblk + int64_t(it.base) * 6 + r));
      } else {
        esum += sum_splits(vEpart + int64_t(it.base) * 6 + r, int64_t(it.stride) * 6, S);
      }
    }
    sv[il][r] = vsum;
    se[il][r] = esum;
  }
  __syncthreads();
  if (il != 0) return;
#pragma unroll
  for (int l = 1; l < RHS_LANES; ++l) { vsum += sv[l][r]; esum += se[l][r]; }
  A[int64_t(6 * Pp) * ldA + a * 6 + r] = float(vsum - esum);
}

// CTAs [0, nlightCTA): 8 light blocks (<= 8 items) per CTA, one thread per entry; the remaining CTAs: one
// heavy block (diagonal blocks carry ~50 items) with 8 item lanes per entry, summed in lane order.
__global__ void __launch_bounds__(36 * ASM_LANES)
k_assemble_S(const int32_t* __restrict__ off, const Item* __restrict__ items, const int32_t* __restrict__ blist,
             int nlight, int nlightCTA, const float* __restrict__ Hblk, const float* __restrict__ Spart, int S,
             float lm, float ep, bool block_diag_quirk, float* __restrict__ A, int ldA, int nSctas,
             const int32_t* __restrict__ roff, const Item* __restrict__ ritems, const float* __restrict__ vblk,
             const float* __restrict__ vEpart, int Pp) {
  if (int(blockIdx.x) >= nSctas) {               // trailing CTAs: one pose of the rhs row each
    assemble_rhs_cta(blockIdx.x - nSctas, roff, ritems, vblk, vEpart, S, Pp, A, ldA);
    return;
  }
  __shared__ double sh[ASM_LANES][36], ss[ASM_LANES][36];
  const bool light = int(blockIdx.x) < nlightCTA;
  const int grp = threadIdx.x / 36, ent = threadIdx.x % 36;
  const int r = ent / 6, c = ent % 6;
  int ob;
  if (light) {
    const int idx = blockIdx.x * ASM_LANES + grp;
    if (idx >= nlight) return;
    ob = blist[idx];
  } else {
    ob = blist[nlight + (blockIdx.x - nlightCTA)];
  }
  // ob = a(a+1)/2 + b
  int a = int((sqrtf(8.0f * float(ob) + 1.0f) - 1.0f) * 0.5f);
  while ((a + 1) * (a + 2) / 2 <= ob) ++a;
  while (a * (a + 1) / 2 > ob) --a;
  const int b = ob - a * (a + 1) / 2;
  double hsum = 0.0, ssum = 0.0;
  const int i0 = off[ob], i1 = off[ob + 1];
  if (light) {
    gather_block(items, i0, i1, 0, 1, ent, c * 6 + r, Hblk, Spart, S, hsum, ssum);
  } else {
    gather_block(items, i0, i1, grp, ASM_LANES, ent, c * 6 + r, Hblk, Spart, S, hsum, ssum);
    sh[grp][ent] = hsum;
    ss[grp][ent] = ssum;
    __syncthreads();
    if (grp != 0) return;
#pragma unroll
    for (int l = 1; l < ASM_LANES; ++l) { hsum += sh[l][ent]; ssum += ss[l][ent]; }
  }
  // chol.py:55-57: H + (ep + lm H) on the true diagonal; chol.py:35-37 (block_solve as written):
  // on the diagonal of EVERY 6x6 block when the quirk is requested (motion_only parity switch)
  const bool damp = (r == c) && (a == b || block_diag_quirk);
  double val = damp ? hsum + (double(ep) + double(lm) * hsum) : hsum;
  val -= ssum;
  A[int64_t(6 * a + r) * ldA + 6 * b + c] = float(val);
}

// ---------------------------------------------------------------------------------------------
// back-substitution: dz = Q (w - E^T dx), disparity retraction + guards (one thread per pixel)
//   E_j . dx_j = h . (d tau_j + phi_j x X1) ;  E_i . dx_k = -h . R (d tau_k + phi_k x X0)
// ---------------------------------------------------------------------------------------------
constexpr int BS_T = 128;
constexpr int BS_CHUNK = 64;

__global__ void __launch_bounds__(BS_T)
k_backsub(const float* __restrict__ poses, float* __restrict__ disps, const float* __restrict__ intr,
          const float* __restrict__ weights, const int32_t* __restrict__ kx, const int32_t* __restrict__ koff,
          const int32_t* __restrict__ cedge, const int32_t* __restrict__ cj, const float* __restrict__ Q,
          const float* __restrict__ wv, const float* __restrict__ dx, float* __restrict__ dz, int HW, int wd,
          int ntiles, int t0, int Pp) {
  __shared__ float sRt[BS_CHUNK * 12];
  __shared__ float sMc[BS_CHUNK * 12];   // per edge: M (3x3) and c (3) of u = M X0 + d c, see below
  __shared__ int sedge[BS_CHUNK];
  const int m = blockIdx.x / ntiles, tile = blockIdx.x % ntiles;
  const int tid = threadIdx.x;
  const int k = kx[m];
  const int e0 = koff[m], deg = koff[m + 1] - e0;
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const int p = tile * BS_T + tid;
  const bool inb = p < HW;
  const float dd = inb ? disps[int64_t(k) * HW + p] : 0.f;
  const int py = p / wd, px = p - py * wd;
  const float xx = (float(px) - cx) / fx, yy = (float(py) - cy) / fy;
  float tk[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (k - t0 >= 0 && k - t0 < Pp) {
#pragma unroll
    for (int i = 0; i < 6; ++i) tk[i] = dx[(k - t0) * 6 + i];
  }
  // E^T dx per edge-pixel is h . u with u = d tau_j + phi_j x X1 - R z, z = d tau_k + phi_k x X0, X1 = R X0 + d t.
  // Everything but h is linear in (X0, d) with per-edge coefficients:
  //   u = M X0 + d c ,  M = [phi_j]x R - R [phi_k]x ,  c = tau_j + phi_j x t - R tau_k
  // built once per edge (12 floats, like Rt) instead of ~30 FMAs per edge-pixel.
  float etdx = 0.f;
  for (int c0 = 0; c0 < deg; c0 += BS_CHUNK) {
    const int nc = min(BS_CHUNK, deg - c0);
    __syncthreads();
    for (int le = tid; le < nc; le += BS_T) {
      const int j = cj[e0 + c0 + le];
      float Rt[12];
      relative_pose_Rt(poses, k, j, Rt);
      const bool opt = (j - t0 >= 0 && j - t0 < Pp);
      float tj[6];
#pragma unroll
      for (int i = 0; i < 6; ++i) tj[i] = opt ? dx[(j - t0) * 6 + i] : 0.f;
      float* Mc = sMc + 12 * le;
#pragma unroll
      for (int col = 0; col < 3; ++col) {
        // column `col` of [phi_j]x R = phi_j x R[:,col] ; column `col` of R [phi_k]x = R (phi_k x e_col)
        const float r0 = Rt[col], r1 = Rt[3 + col], r2 = Rt[6 + col];
        const float ax = tj[4] * r2 - tj[5] * r1, ay = tj[5] * r0 - tj[3] * r2, az = tj[3] * r1 - tj[4] * r0;
        // phi_k x e_col
        const float px_ = (col == 0) ? 0.f : (col == 1 ? -tk[5] : tk[4]);
        const float py_ = (col == 0) ? tk[5] : (col == 1 ? 0.f : -tk[3]);
        const float pz_ = (col == 0) ? -tk[4] : (col == 1 ? tk[3] : 0.f);
        Mc[0 + col] = ax - (Rt[0] * px_ + Rt[1] * py_ + Rt[2] * pz_);
        Mc[3 + col] = ay - (Rt[3] * px_ + Rt[4] * py_ + Rt[5] * pz_);
        Mc[6 + col] = az - (Rt[6] * px_ + Rt[7] * py_ + Rt[8] * pz_);
      }
      Mc[9] = tj[0] + (tj[4] * Rt[11] - tj[5] * Rt[10]) - (Rt[0] * tk[0] + Rt[1] * tk[1] + Rt[2] * tk[2]);
      Mc[10] = tj[1] + (tj[5] * Rt[9] - tj[3] * Rt[11]) - (Rt[3] * tk[0] + Rt[4] * tk[1] + Rt[5] * tk[2]);
      Mc[11] = tj[2] + (tj[3] * Rt[10] - tj[4] * Rt[9]) - (Rt[6] * tk[0] + Rt[7] * tk[1] + Rt[8] * tk[2]);
#pragma unroll
      for (int i = 0; i < 12; ++i) sRt[12 * le + i] = Rt[i];
      sedge[le] = cedge[e0 + c0 + le];
    }
    __syncthreads();
    if (inb) {
      for (int le = 0; le < nc; ++le) {
        const float* Rt = sRt + 12 * le;
        const float* Mc = sMc + 12 * le;
        const float* wg = weights + int64_t(sedge[le]) * 2 * HW;
        const float wx = __ldcs(wg + p), wy = __ldcs(wg + HW + p);
        const Proj pr = project(Rt, xx, yy, dd, fx, fy, cx, cy);
        const float a0 = kWeightScale * (pr.valid * wx) * pr.jz0, a1 = kWeightScale * (pr.valid * wy) * pr.jz1;
        const float hx = a0 * pr.a0, hy = a1 * pr.b1, hz = a0 * pr.c0 + a1 * pr.c1;
        const float ux = Mc[0] * xx + Mc[1] * yy + Mc[2] + Mc[9] * dd;
        const float uy = Mc[3] * xx + Mc[4] * yy + Mc[5] + Mc[10] * dd;
        const float uz = Mc[6] * xx + Mc[7] * yy + Mc[8] + Mc[11] * dd;
        etdx += hx * ux + hy * uy + hz * uz;
      }
    }
  }
  if (inb) {
    const float q = Q[int64_t(m) * HW + p], w = wv[int64_t(m) * HW + p];
    const float dzv = q * (w - etdx);
    dz[int64_t(m) * HW + p] = dzv;
    float nd = dd + dzv;                        // ba.py:102
    nd = (nd > 10.f) ? 0.f : nd;                // ba.py:104
    nd = fmaxf(nd, 0.001f);                     // ba.py:105  (NaN -> 0.001, as torch.clamp does not: see DESIGN)
    disps[int64_t(k) * HW + p] = nd;
  }
}

// disparity guards on the frames of [0,t1) that are not a source keyframe (ba.py:104-105 run over
// the whole slice)
__global__ void __launch_bounds__(256)
k_guard(float* __restrict__ disps, const int32_t* __restrict__ frames, int HW) {
  const int f = frames[blockIdx.y];
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += gridDim.x * blockDim.x) {
    float v = disps[int64_t(f) * HW + p];
    float nd = (v > 10.f) ? 0.f : v;
    nd = fmaxf(nd, 0.001f);
    if (nd != v) disps[int64_t(f) * HW + p] = nd;
  }
}

__global__ void __launch_bounds__(128)
k_retract(float* __restrict__ poses, const float* __restrict__ dx, int t0, int Pp, float* __restrict__ dx_out) {
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= Pp) return;
  float xi[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) xi[i] = dx[a * 6 + i];
  se3_retract(poses + 7 * (t0 + a), xi);
  if (dx_out != nullptr) {
#pragma unroll
    for (int i = 0; i < 6; ++i) dx_out[a * 6 + i] = xi[i];
  }
}

}  // namespace

cudaError_t launch_assemble(const DeviceGraph& g, const BAInputs& in, const Workspace& ws, float lm, float ep,
                            float /*hscale_lm*/, bool motion_only, bool block_diag_quirk, cudaStream_t st) {
  (void)motion_only;
  if (g.E > 0) {
    k_edge_finalize<<<(g.E + EF_WARPS - 1) / EF_WARPS, 32 * EF_WARPS, 0, st>>>(in.poses, g.ci, g.cj, ws.Hpart, g.E, g.S, ws.Hblk, ws.vblk);
    g_launch_count++;
  }
  const int nlightCTA = (g.n_light + ASM_LANES - 1) / ASM_LANES;
  const int nSctas = nlightCTA + (g.nSblk - g.n_light);
  k_assemble_S<<<nSctas + g.Pp, 36 * ASM_LANES, 0, st>>>(
      g.sblk_off, g.sblk_items, g.sblk_list, g.n_light, nlightCTA, ws.Hblk, ws.Spart, g.S, lm, ep, block_diag_quirk,
      ws.A, ws.ldA, nSctas, g.rhs_off, g.rhs_items, ws.vblk, ws.vEpart, g.Pp);
  g_launch_count += 1;
  return cudaGetLastError();
}

cudaError_t launch_backsub(const DeviceGraph& g, const BAInputs& in, const Workspace& ws, cudaStream_t st) {
  const int ntiles = (g.HW + BS_T - 1) / BS_T;
  if (g.M > 0) {
    g_launch_count++;
    k_backsub<<<g.M * ntiles, BS_T, 0, st>>>(in.poses, in.disps, in.intrinsics, in.weights, g.kx, g.koff, g.cedge,
                                             g.cj, ws.Q, ws.w, ws.dx, ws.dz, g.HW, g.wd, ntiles, g.t0, g.Pp);
  }
  return cudaGetLastError();
}

cudaError_t launch_guard_nonsrc(const DeviceGraph& g, const BAInputs& in, cudaStream_t st) {
  if (g.n_nonsrc > 0) {
    dim3 grid((g.HW + 255) / 256, g.n_nonsrc);
    k_guard<<<grid, 256, 0, st>>>(in.disps, g.nonsrc, g.HW);
    g_launch_count++;
  }
  return cudaGetLastError();
}

cudaError_t launch_retract(const DeviceGraph& g, const BAInputs& in, const Workspace& ws, float* dx_out,
                           cudaStream_t st) {
  k_retract<<<(g.Pp + 127) / 128, 128, 0, st>>>(in.poses, ws.dx, g.t0, g.Pp, dx_out);
  g_launch_count++;
  return cudaGetLastError();
}

}  // namespace dba
