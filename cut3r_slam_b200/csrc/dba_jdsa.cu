// JDSA path of HI-SLAM2 (hislam2/geom/ba.py:161-241, chol.py:80-108): the three
// `droid_backends` ops / solves it needs.
//   proj_trans(poses, disps, intrinsics, target, weight, ii, jj) -> (C, w)   [ba.py:200]
//       per source keyframe and pixel: C = sum_edges sum_rows w Jz^2, w = sum w r Jz -- the depth block and
//       depth gradient of ba.py:62-68,83-90 WITHOUT eta (JDSA adds its own terms, ba.py:228-229)
//   bi_inter(scales, grid) -> (vals, J)                                        [ba.py:167]
//       bilinear interpolation of a coarse scale grid at fractional positions + the dense Jacobian w.r.t.
//       the grid nodes (4 non-zeros per pixel)
//   mono_prior_solve: schur_solve_mono_prior (chol.py:80-108) for the block structure JDSA builds: Hs, Es are
//       scattered at (kx, kx) only (ba.py:222-223), so the system is M independent D x D problems.
// The reference CUDA of proj_trans / bi_inter is absent: their definitions here are pinned to the torch
// statements around the call sites (oracle/jdsa.py); the solve is pinned to the reference's own
// schur_solve_mono_prior (tests/golden/jdsa_*.npz).
#include "dba_kernels.h"
#include "dba_math.cuh"

namespace dba {

namespace {

constexpr int PT_T = 128;
constexpr int PT_CHUNK = 64;

// one thread per pixel of (source keyframe slot m, 128-pixel tile): loop over the keyframe's edges
__global__ void __launch_bounds__(PT_T)
k_proj_trans(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr,
             const float* __restrict__ targets, const float* __restrict__ weights, const int32_t* __restrict__ kx,
             const int32_t* __restrict__ koff, const int32_t* __restrict__ cedge, const int32_t* __restrict__ cj,
             int HW, int wd, int ntiles, float* __restrict__ Cout, float* __restrict__ wout) {
  __shared__ float sRt[PT_CHUNK * 12];
  __shared__ int sedge[PT_CHUNK];
  const int m = blockIdx.x / ntiles, tile = blockIdx.x % ntiles;
  const int tid = threadIdx.x;
  const int k = kx[m];
  const int e0 = koff[m], deg = koff[m + 1] - e0;
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const int p = tile * PT_T + tid;
  const bool inb = p < HW;
  const float dd = inb ? disps[int64_t(k) * HW + p] : 0.f;
  const int py = p / wd, px = p - py * wd;
  const float xx = (float(px) - cx) / fx, yy = (float(py) - cy) / fy;
  float C = 0.f, w = 0.f;
  for (int c0 = 0; c0 < deg; c0 += PT_CHUNK) {
    const int nc = min(PT_CHUNK, deg - c0);
    __syncthreads();
    for (int le = tid; le < nc; le += PT_T) {
      relative_pose_Rt(poses, k, cj[e0 + c0 + le], sRt + 12 * le);
      sedge[le] = cedge[e0 + c0 + le];
    }
    __syncthreads();
    if (inb) {
#pragma unroll 4
      for (int le = 0; le < nc; ++le) {                 // four edges' loads in flight per thread
        const float* tg = targets + int64_t(sedge[le]) * 2 * HW;
        const float* wg = weights + int64_t(sedge[le]) * 2 * HW;
        const float tx = __ldcs(tg + p), ty = __ldcs(tg + HW + p);
        const float wx = __ldcs(wg + p), wy = __ldcs(wg + HW + p);
        const Proj pr = project(sRt + 12 * le, xx, yy, dd, fx, fy, cx, cy);
        const float r0 = tx - pr.u, r1 = ty - pr.v;
        const float a0 = kWeightScale * (pr.valid * wx) * pr.jz0, a1 = kWeightScale * (pr.valid * wy) * pr.jz1;
        C += a0 * pr.jz0 + a1 * pr.jz1;          // ba.py:68
        w += a0 * r0 + a1 * r1;                  // ba.py:67
      }
    }
  }
  if (inb) {
    Cout[int64_t(m) * HW + p] = C;
    wout[int64_t(m) * HW + p] = w;
  }
}

// vals[m,p] = bilinear(scales[m], grid[m,p]) ; J[m,p,:] = d vals / d scales[m,:] (hs*ws entries, 4 non-zero).
// grid[...,0] runs along ws (x), grid[...,1] along hs (y), both in node units (ba.py:164-166).
// A warp owns 32 consecutive pixels: their 32 x D block of J is one contiguous run, written lane-strided with the
// value of each element computed from the owning pixel's cell (fetched by shuffle) -- every store is a full line.
__global__ void __launch_bounds__(256)
k_bi_inter(const float* __restrict__ scales, const float* __restrict__ grid, int HW, int hs, int ws,
           float* __restrict__ vals, float* __restrict__ J) {
  const int m = blockIdx.y;
  const int lane = threadIdx.x & 31;
  const int pw = (blockIdx.x * blockDim.x + threadIdx.x) & ~31;       // first pixel of this warp
  if (pw >= HW) return;
  const int p = pw + lane;
  const bool inb = p < HW;
  const int D = hs * ws;
  float gx = 0.f, gy = 0.f;
  if (inb) {
    const float2 g2 = *reinterpret_cast<const float2*>(grid + (int64_t(m) * HW + p) * 2);
    gx = g2.x; gy = g2.y;
  }
  int x0 = int(floorf(gx)), y0 = int(floorf(gy));
  x0 = max(0, min(x0, ws - 2 >= 0 ? ws - 2 : 0));
  y0 = max(0, min(y0, hs - 2 >= 0 ? hs - 2 : 0));
  const int x1 = min(x0 + 1, ws - 1), y1 = min(y0 + 1, hs - 1);
  const float wx = gx - float(x0), wy = gy - float(y0);
  const float w00 = (1.f - wy) * (1.f - wx), w01 = (1.f - wy) * wx, w10 = wy * (1.f - wx), w11 = wy * wx;
  const float* s = scales + int64_t(m) * D;
  if (inb)
    vals[int64_t(m) * HW + p] = w00 * s[y0 * ws + x0] + w01 * s[y0 * ws + x1] + w10 * s[y1 * ws + x0] + w11 * s[y1 * ws + x1];
  const int npx = min(32, HW - pw);
  float* Jw = J + (int64_t(m) * HW + pw) * D;
  for (int idx = lane; idx < 32 * D; idx += 32) {                      // warp-uniform trip count (shuffles inside)
    const int q = idx / D, a = idx - q * D;
    const int qx0 = __shfl_sync(0xffffffffu, x0, q), qy0 = __shfl_sync(0xffffffffu, y0, q);
    const float qwx = __shfl_sync(0xffffffffu, wx, q), qwy = __shfl_sync(0xffffffffu, wy, q);
    const int qx1 = min(qx0 + 1, ws - 1), qy1 = min(qy0 + 1, hs - 1);
    const int ay = a / ws, ax = a - ay * ws;
    const float hy = (ay == qy0 ? 1.f - qwy : 0.f) + (ay == qy1 ? qwy : 0.f);
    const float hx = (ax == qx0 ? 1.f - qwx : 0.f) + (ax == qx1 ? qwx : 0.f);
    if (q < npx) __stcs(Jw + idx, hy * hx);
  }
}


// ---------------------------------------------------------------------------------------------
// One JDSA step per source keyframe (ba.py:203-239 + chol.py:80-108), one CTA each.  Because Hs, Es, vs are
// scattered at (kx, kx) only, the reduced system is M independent D x D problems (D = hs*ws <= 64):
//   g = -m prior ;  Jso_a = g h_a(p)  (h_a = bilinear weight of node a at pixel p, 4 non-zero per pixel)
//   C' = C + m alpha + (1-m) eta ; w' = w - m alpha rd ; rd = disp - prior * bilinear(scales) ; Q = 1/C'
//   S = alpha T1 (1 + lm I) + ep I - alpha^2 T2,  T1 = sum g^2 h_a h_b, T2 = sum Q g^2 h_a h_b
//   v = -alpha sum g h_a rd - alpha sum g h_a Q w'
//   dso = S^-1 v ; dz = Q (w' - alpha sum_a Jso_a dso_a) ; dzcov = u^T S^-1 u + Q, u_a = alpha Jso_a Q
// All sums are lane-strided partials + a fixed shuffle tree (deterministic).
// ---------------------------------------------------------------------------------------------
constexpr int JD_T = 1024;      // one CTA per keyframe: every phase is latency-bound, so as many warps as fit
constexpr int JD_DMAX = 64;

struct PixNodes { int a[4]; float h[4]; };

// dynamic shared memory of k_jdsa: per-pixel g, Q, w', rd [4][HW] | per-row y0, wy [2][ht] | per-column x0, wx [2][wd]
size_t jdsa_smem_bytes(int ht, int wd) { return (size_t(4) * ht * wd + 2 * size_t(ht) + 2 * size_t(wd)) * sizeof(float); }

__global__ void __launch_bounds__(JD_T)
k_jdsa(float* disps, const float* disps_in, const float* __restrict__ Cin, const float* __restrict__ win,
       const float* __restrict__ eta, const float* __restrict__ prior, float* scales, const float* scales_in,
       const int32_t* __restrict__ kx, int M, int ht, int wd, int hs, int ws, float alpha, float ep, float lm,
       float* __restrict__ dzcov, float* __restrict__ dz_out, float* __restrict__ dso_out) {
  extern __shared__ float jsm[];
  __shared__ float sS[JD_DMAX][JD_DMAX + 1], sW[JD_DMAX][JD_DMAX + 1], sv[JD_DMAX], sv2[JD_DMAX], sdso[JD_DMAX], ssc[JD_DMAX],
      ssc_src[JD_DMAX], s_r;
  __shared__ int s_found, srlo[JD_DMAX], srhi[JD_DMAX], sclo[JD_DMAX], schi[JD_DMAX];
  const int m = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int D = hs * ws, HW = ht * wd;
  const int k = kx[m];
  float* sg = jsm;                  // g = -m prior of the frame whose prior terms land in this slot
  float* sQ = sg + HW;              // 1 / C'
  float* swp = sQ + HW;             // w'
  float* srd = swp + HW;            // prior residual of that frame (rhs of the scale block)
  float* swy = srd + HW;            // per image row: interpolation weight and (as int) the upper node row
  int* sy0 = reinterpret_cast<int*>(swy + ht);
  float* swx = reinterpret_cast<float*>(sy0 + ht);
  int* sx0 = reinterpret_cast<int*>(swx + wd);
  // ba.py:218-225: the prior blocks Hs, Es, vs of keyframe kx[q] are scattered at slot kx[q] - kx[0], not at q, and
  // slots >= M are dropped.  Slot m therefore takes its prior terms from frame ks = kx[0] + m IF that frame is a
  // source keyframe (ks == k whenever unique(ii) is a contiguous range -- the normal case), else it has none;
  // C, w, eta, the retraction and dso -> dscales[kx[m]] stay with slot m's own keyframe.
  const int ks = kx[0] + m;
  if (tid == 0) s_found = 0;
  __syncthreads();
  for (int q = tid; q < M; q += JD_T) if (kx[q] == ks) s_found = 1;
  for (int i = tid; i < D; i += JD_T) ssc[i] = scales_in[int64_t(k) * D + i];
  // ba.py:163: torch.linspace(0, hs-1-1e-6, ht) / (0, ws-1-1e-6, wd) in fp32, torch's symmetric evaluation
  {
    const float endy = float(hs - 1) - 1e-6f, endx = float(ws - 1) - 1e-6f;
    const float stepy = ht > 1 ? endy / float(ht - 1) : 0.f, stepx = wd > 1 ? endx / float(wd - 1) : 0.f;
    for (int r = tid; r < ht; r += JD_T) {
      const float gy = r < ht / 2 ? stepy * float(r) : endy - stepy * float(ht - 1 - r);
      int y0 = int(floorf(gy));
      y0 = max(0, min(y0, hs - 2 >= 0 ? hs - 2 : 0));
      sy0[r] = y0; swy[r] = gy - float(y0);
    }
    for (int c = tid; c < wd; c += JD_T) {
      const float gx = c < wd / 2 ? stepx * float(c) : endx - stepx * float(wd - 1 - c);
      int x0 = int(floorf(gx));
      x0 = max(0, min(x0, ws - 2 >= 0 ? ws - 2 : 0));
      sx0[c] = x0; swx[c] = gx - float(x0);
    }
  }
  __syncthreads();
  const bool has_prior = s_found != 0;
  for (int i = tid; i < D; i += JD_T) ssc_src[i] = has_prior ? scales_in[int64_t(ks) * D + i] : 0.f;
  for (int i = tid; i < D * (JD_DMAX + 1); i += JD_T) { (&sS[0][0])[i] = 0.f; (&sW[0][0])[i] = 0.f; }
  // image rows / columns that see node row y / node column x (y0 is non-decreasing along the image)
  for (int y = tid; y < hs; y += JD_T) {
    int lo = ht, hi = -1;
    for (int r = 0; r < ht; ++r) {
      const int y0 = sy0[r], y1 = min(y0 + 1, hs - 1);
      if (y0 == y || y1 == y) { lo = min(lo, r); hi = max(hi, r); }
    }
    srlo[y] = lo; srhi[y] = hi;
  }
  for (int x = tid; x < ws; x += JD_T) {
    int lo = wd, hi = -1;
    for (int c = 0; c < wd; ++c) {
      const int x0 = sx0[c], x1 = min(x0 + 1, ws - 1);
      if (x0 == x || x1 == x) { lo = min(lo, c); hi = max(hi, c); }
    }
    sclo[x] = lo; schi[x] = hi;
  }
  __syncthreads();

  auto nodes_of = [&](int py, int px) {
    PixNodes n;
    const int y0 = sy0[py], x0 = sx0[px];
    const int y1 = min(y0 + 1, hs - 1), x1 = min(x0 + 1, ws - 1);
    const float wy = swy[py], wx = swx[px];
    n.a[0] = y0 * ws + x0; n.a[1] = y0 * ws + x1; n.a[2] = y1 * ws + x0; n.a[3] = y1 * ws + x1;
    n.h[0] = (1.f - wy) * (1.f - wx); n.h[1] = (1.f - wy) * wx; n.h[2] = wy * (1.f - wx); n.h[3] = wy * wx;
    return n;
  };
  // separable weight of node (ya, xa) at pixel (py, px)
  auto hy_of = [&](int ya, int py) {
    const int y0 = sy0[py], y1 = min(y0 + 1, hs - 1);
    const float wy = swy[py];
    return (y0 == ya ? 1.f - wy : 0.f) + (y1 == ya ? wy : 0.f);
  };
  auto hx_of = [&](int xa, int px) {
    const int x0 = sx0[px], x1 = min(x0 + 1, ws - 1);
    const float wx = swx[px];
    return (x0 == xa ? 1.f - wx : 0.f) + (x1 == xa ? wx : 0.f);
  };

  // ---- per-pixel terms, once (ba.py:205-229) ----
  {
    const float* dk = disps_in + int64_t(k) * HW;
    const float* pk = prior + int64_t(k) * HW;
    const float* dsrc = disps_in + int64_t(ks) * HW;
    const float* psrc = prior + int64_t(ks) * HW;
    for (int p = tid; p < HW; p += JD_T) {
      const int py = p / wd, px = p - py * wd;
      const PixNodes n = nodes_of(py, px);
      const float pr = pk[p];
      const float mk = pr > 0.f ? 1.f : 0.f;
      const float bil = n.h[0] * ssc[n.a[0]] + n.h[1] * ssc[n.a[1]] + n.h[2] * ssc[n.a[2]] + n.h[3] * ssc[n.a[3]];
      float rd = dk[p] - pr * bil;
      const float Cp = Cin[int64_t(m) * HW + p] + mk * alpha + (1.f - mk) * eta[int64_t(m) * HW + p];
      swp[p] = win[int64_t(m) * HW + p] - mk * alpha * rd;
      sQ[p] = 1.0f / Cp;
      float g = -mk * pr;
      if (ks != k) {                                       // prior terms of the frame scattered into this slot
        if (has_prior) {
          const float ps = psrc[p];
          const float bs = n.h[0] * ssc_src[n.a[0]] + n.h[1] * ssc_src[n.a[1]] + n.h[2] * ssc_src[n.a[2]] +
                           n.h[3] * ssc_src[n.a[3]];
          g = ps > 0.f ? -ps : 0.f;
          rd = dsrc[p] - ps * bs;
        } else {
          g = 0.f;
        }
      }
      sg[p] = g;
      srd[p] = rd;
    }
  }
  __syncthreads();

  // ---- reduced system: one warp per entry (a, b <= a) with adjacent nodes, and per rhs entry; the sums run over
  //      the pixels both nodes see, lane-strided + a fixed shuffle tree ----
  if (has_prior) {
    // tasks (a, o): o < 5 -> the entry (a, b) with b = a + offset o (the five neighbours with b <= a), o == 5 -> rhs
    const int ntask = 6 * D;
    for (int t = warp; t < ntask; t += JD_T / 32) {
      const int a = t / 6, o = t - 6 * a;
      const bool is_v = o == 5;
      const int ya = a / ws, xa = a - ya * ws;
      const int dy = (o >= 2 && o <= 4) ? -1 : 0;
      const int dx = (o == 1 || o == 2) ? -1 : (o == 4 ? 1 : 0);
      const int yb = ya + dy, xb = xa + dx;
      if (yb < 0 || xb < 0 || xb >= ws) continue;
      const int b = yb * ws + xb;
      const int r0 = max(srlo[ya], srlo[yb]), r1 = min(srhi[ya], srhi[yb]);
      const int c0 = max(sclo[xa], sclo[xb]), c1 = min(schi[xa], schi[xb]);
      const int rw = c1 - c0 + 1, rh = r1 - r0 + 1;
      float s1 = 0.f, s2 = 0.f;
      if (rw > 0 && rh > 0) {
        for (int idx = lane; idx < rw * rh; idx += 32) {
          const int py = r0 + idx / rw, px = c0 + idx % rw;
          const int p = py * wd + px;
          const float g = sg[p];
          const float ha = hy_of(ya, py) * hx_of(xa, px);
          if (is_v) {
            s1 = fmaf(g * ha, srd[p], s1);
            s2 = fmaf(g * ha, sQ[p] * swp[p], s2);
          } else {
            const float q = g * g * ha * (hy_of(yb, py) * hx_of(xb, px));
            s1 += q;
            s2 = fmaf(sQ[p], q, s2);
          }
        }
      }
#pragma unroll
      for (int o = 16; o >= 1; o >>= 1) { s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
      if (lane == 0) {
        if (is_v) sv[a] = -alpha * s1 - alpha * s2;
        else {
          const float dmp = (a == b) ? 1.f : 0.f;
          const float val = alpha * s1 + dmp * (ep + lm * alpha * s1) - alpha * alpha * s2;
          sS[a][b] = val; sS[b][a] = val;
        }
      }
    }
  } else {
    for (int i = tid; i < D; i += JD_T) { sS[i][i] = ep; sv[i] = 0.f; }
  }
  __syncthreads();
  // ---- Cholesky of S (lower, in place): one barrier per column.  Column j stays unscaled while its rank-1 update is
  //      applied (the scale 1/sqrt(d_j) goes in through registers); all columns are scaled at the end ----
  for (int j = 0; j < D; ++j) {
    const float d = sS[j][j];
    const float r = d > 0.f ? 1.0f / sqrtf(d) : 0.f;
    if (tid == 0) sv2[j] = r;
    const int ti = tid >> 5, tc = tid & 31;
    for (int i = j + 1 + ti; i < D; i += JD_T / 32) {
      const float lij = sS[i][j] * r;
      for (int c = j + 1 + tc; c <= i; c += 32) sS[i][c] = fmaf(-lij, sS[c][j] * r, sS[i][c]);
    }
    __syncthreads();
  }
  for (int e = tid; e < D * D; e += JD_T) {
    const int i = e / D, j = e - i * D;
    if (j <= i) sS[i][j] *= sv2[j];                         // L[j][j] = d_j / sqrt(d_j)
  }
  __syncthreads();
  // ---- W = L^-1: thread c owns column c; row i uses rows < i of the same column.  The sums run from q = 0 (W is
  //      zero above its diagonal) so that every lane has the same trip count and the loads pipeline ----
  if (tid < D) {
    const int c = tid;
    for (int i = 0; i < D; ++i) {
      float s0 = (i == c) ? 1.f : 0.f, s1 = 0.f;
      int q = 0;
      for (; q + 1 < i; q += 2) {
        s0 = fmaf(-sS[i][q], sW[q][c], s0);
        s1 = fmaf(-sS[i][q + 1], sW[q + 1][c], s1);
      }
      if (q < i) s0 = fmaf(-sS[i][q], sW[q][c], s0);
      if (i >= c) sW[i][c] = (s0 + s1) / sS[i][i];
    }
  }
  __syncthreads();
  // ---- S^-1 = W^T W into sS (full, symmetric) ; dso = S^-1 v ----
  {
    float keep[(JD_DMAX * JD_DMAX + JD_T - 1) / JD_T];
    int cnt = 0;
    for (int e = tid; e < D * D; e += JD_T, ++cnt) {
      const int a = e / D, b = e % D;
      float s = 0.f;
      for (int r = max(a, b); r < D; ++r) s = fmaf(sW[r][a], sW[r][b], s);
      keep[cnt] = s;
    }
    __syncthreads();
    cnt = 0;
    for (int e = tid; e < D * D; e += JD_T, ++cnt) sS[e / D][e % D] = keep[cnt];
  }
  __syncthreads();
  for (int a = tid; a < D; a += JD_T) {
    float s = 0.f;
    for (int b = 0; b < D; ++b) s = fmaf(sS[a][b], sv[b], s);
    sdso[a] = s;
  }
  __syncthreads();
  // ---- per pixel: dz, dzcov, disparity retraction + guards (ba.py:231-239) ----
  {
    const float* dk = disps_in + int64_t(k) * HW;
    for (int p = tid; p < HW; p += JD_T) {
      const int py = p / wd, px = p - py * wd;
      const PixNodes n = nodes_of(py, px);
      const float g = sg[p], Q = sQ[p];
      float et = 0.f, quad = 0.f;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        et += g * n.h[u] * sdso[n.a[u]];
        float row = 0.f;
#pragma unroll
        for (int v = 0; v < 4; ++v) row += n.h[v] * sS[n.a[u]][n.a[v]];
        quad += n.h[u] * row;
      }
      const float dzv = Q * (swp[p] - alpha * et);
      const float ua = alpha * g * Q;
      dzcov[int64_t(m) * HW + p] = ua * ua * quad + Q;
      if (dz_out != nullptr) dz_out[int64_t(m) * HW + p] = dzv;
      float nd = dk[p] + dzv;
      nd = (nd > 10.f) ? 0.f : nd;
      nd = fmaxf(nd, 0.001f);
      disps[int64_t(k) * HW + p] = nd;
    }
  }
  for (int i = tid; i < D; i += JD_T) {
    scales[int64_t(k) * D + i] = ssc[i] + sdso[i];
    if (dso_out != nullptr) dso_out[int64_t(m) * D + i] = sdso[i];
  }
}


__global__ void __launch_bounds__(256)
k_guard_all(float* __restrict__ disps, int64_t n) {
  for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
    const float v = disps[i];
    float nd = (v > 10.f) ? 0.f : v;
    nd = fmaxf(nd, 0.001f);
    if (nd != v) disps[i] = nd;
  }
}

}  // namespace

cudaError_t launch_proj_trans(const float* poses, const float* disps, const float* intr, const float* targets,
                              const float* weights, const int32_t* kx, const int32_t* koff, const int32_t* cedge,
                              const int32_t* cj, int M, int HW, int wd, float* Cout, float* wout, cudaStream_t st) {
  if (M == 0) return cudaSuccess;
  const int ntiles = (HW + PT_T - 1) / PT_T;
  k_proj_trans<<<M * ntiles, PT_T, 0, st>>>(poses, disps, intr, targets, weights, kx, koff, cedge, cj, HW, wd, ntiles,
                                            Cout, wout);
  g_launch_count++;
  return cudaGetLastError();
}

cudaError_t launch_bi_inter(const float* scales, const float* grid, int M, int HW, int hs, int ws, float* vals,
                            float* J, cudaStream_t st) {
  if (M == 0) return cudaSuccess;
  dim3 g((HW + 255) / 256, unsigned(M));
  k_bi_inter<<<g, 256, 0, st>>>(scales, grid, HW, hs, ws, vals, J);
  g_launch_count++;
  return cudaGetLastError();
}

int jdsa_max_nodes() { return JD_DMAX; }
size_t jdsa_dynamic_smem(int ht, int wd) { return jdsa_smem_bytes(ht, wd) + 36 * 1024; }   // + the static part

cudaError_t launch_jdsa(float* disps, const float* Cin, const float* win, const float* eta, const float* prior,
                        float* scales, const float* disps_in, const float* scales_in, const int32_t* kx, int M, int N, int ht, int wd, int hs, int ws, float alpha,
                        float ep, float lm, float* dzcov, float* dz_out, float* dso_out, cudaStream_t st) {
  const int HW = ht * wd;
  if (M > 0) {
    const size_t smem = jdsa_smem_bytes(ht, wd);
    cudaError_t err = cudaFuncSetAttribute(k_jdsa, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (err != cudaSuccess) return err;
    k_jdsa<<<M, JD_T, smem, st>>>(disps, disps_in, Cin, win, eta, prior, scales, scales_in, kx, M, ht, wd, hs, ws, alpha, ep, lm, dzcov, dz_out,
                               dso_out);
    g_launch_count++;
  }
  const int64_t n = int64_t(N) * HW;                       // ba.py:238-239 run over the whole disparity tensor
  k_guard_all<<<unsigned((n + 256 * 8 - 1) / (256 * 8)), 256, 0, st>>>(disps, n);
  g_launch_count++;
  return cudaGetLastError();
}

}  // namespace dba
