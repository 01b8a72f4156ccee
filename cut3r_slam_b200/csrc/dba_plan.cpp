// Host-side plan construction (see dba_plan.h).  Pure C++, no CUDA.
#include "dba_plan.h"

#include <algorithm>
#include <numeric>

namespace dba {

namespace {
inline int tri(int a, int b) { return a * (a + 1) / 2 + b; }   // a >= b

// back to the default state without giving the vectors' memory back
void reset(Plan& P) {
  P.E = P.N = P.t0 = P.t1 = P.Pp = P.M = P.n6 = 0;
  P.S = 1;
  P.motion_only = false;
  P.degmax = P.nblkmax = P.npairs_max = 0;
  P.slab_blocks = P.vE_blocks = 0;
  P.nSblk = 0;
  for (auto* v : {&P.kx, &P.koff, &P.cedge, &P.ci, &P.cj, &P.pairbase, &P.blkbase, &P.slot_of_frame, &P.sblk_off, &P.rhs_off})
    v->clear();
  P.sblk_items.clear(); P.rhs_items.clear();
  P.key_ii.clear(); P.key_jj.clear();
  P.error.clear();
}
}  // namespace

bool build_plan(Plan& P, const int64_t* ii, const int64_t* jj, int64_t E64, int N, int t0, int t1,
                bool motion_only, int64_t M_expected, int splits) {
  if (!build_plan_csr(P, ii, jj, E64, N, t0, t1, motion_only, M_expected, splits)) return false;
  build_plan_gather(P);
  return true;
}

bool build_plan_csr(Plan& P, const int64_t* ii, const int64_t* jj, int64_t E64, int N, int t0, int t1,
                    bool motion_only, int64_t M_expected, int splits) {
  reset(P);
  if (E64 < 0 || E64 > (int64_t(1) << 24)) { P.error = "edge count out of range"; return false; }
  const int E = int(E64);
  if (N <= 0 || t0 < 0 || t1 > N || t1 - t0 <= 0) {
    P.error = "need 0 <= t0 < t1 <= N (number of frames in the pose/disparity buffers)";
    return false;
  }
  for (int e = 0; e < E; ++e) {
    // the reference slices poses[:t1], disps[:t1] (oracle adapter; ba.py indexes with ii/jj) --
    // an index outside the slice is an IndexError there and an invalid argument here
    if (ii[e] < 0 || jj[e] < 0 || ii[e] >= t1 || jj[e] >= t1) {
      P.error = "edge index outside [0, t1)";
      return false;
    }
  }
  P.E = E; P.N = N; P.t0 = t0; P.t1 = t1; P.Pp = t1 - t0; P.n6 = 6 * P.Pp; P.S = std::max(1, splits);
  P.motion_only = motion_only;
  P.key_ii.assign(ii, ii + E);
  P.key_jj.assign(jj, jj + E);

  // kx = sorted unique(ii)  (ba.py:70)
  P.slot_of_frame.assign(N, -1);
  std::vector<int32_t>& deg_of_frame = P.scratch_deg;
  deg_of_frame.assign(N, 0);
  for (int e = 0; e < E; ++e) deg_of_frame[ii[e]]++;
  for (int f = 0; f < N; ++f)
    if (deg_of_frame[f] > 0) { P.slot_of_frame[f] = int32_t(P.kx.size()); P.kx.push_back(f); }
  P.M = int(P.kx.size());
  if (!motion_only && M_expected >= 0 && M_expected != P.M) {
    P.error = "eta must hold one map per unique source keyframe (rows of eta != |unique(ii)|)";
    return false;
  }
  // CSR: edges of slot m in ascending edge id
  P.koff.assign(P.M + 1, 0);
  for (int m = 0; m < P.M; ++m) P.koff[m + 1] = P.koff[m] + deg_of_frame[P.kx[m]];
  P.cedge.resize(E); P.ci.resize(E); P.cj.resize(E);
  {
    std::vector<int32_t>& cur = P.scratch_cur;
    cur.assign(P.koff.begin(), P.koff.end() - 1);
    for (int e = 0; e < E; ++e) {
      int m = P.slot_of_frame[ii[e]];
      int c = cur[m]++;
      P.cedge[c] = e; P.ci[c] = int32_t(ii[e]); P.cj[c] = int32_t(jj[e]);
    }
  }
  P.pairbase.assign(P.M + 1, 0);
  P.blkbase.assign(P.M + 1, 0);
  for (int m = 0; m < P.M; ++m) {
    int deg = P.koff[m + 1] - P.koff[m];
    int nblk = deg + 1;
    P.degmax = std::max(P.degmax, deg);
    P.nblkmax = std::max(P.nblkmax, nblk);
    P.npairs_max = std::max(P.npairs_max, nblk * (nblk + 1) / 2);
    int64_t pb = int64_t(P.pairbase[m]) + int64_t(nblk) * (nblk + 1) / 2;
    if (pb * P.S * 36 > (int64_t(1) << 31) - 1) { P.error = "Schur slab exceeds 2^31 floats"; return false; }
    P.pairbase[m + 1] = int32_t(pb);
    P.blkbase[m + 1] = P.blkbase[m] + nblk;
  }
  P.slab_blocks = motion_only ? 0 : int64_t(P.pairbase[P.M]) * P.S;
  P.vE_blocks = motion_only ? 0 : int64_t(P.blkbase[P.M]) * P.S;
  P.nSblk = P.Pp * (P.Pp + 1) / 2;

  return true;
}

void build_plan_gather(Plan& P) {
  const int E = P.E, t0 = P.t0;
  const bool motion_only = P.motion_only;
  // ---- gather lists --------------------------------------------------------------------------
  // Two walks over the same contributions in the same order: the first counts per destination block, the second
  // places every item at its final position (a stable counting sort without the intermediate list).
  const int Pp = P.Pp, S = P.S;
  const int32_t* ci = P.ci.data();
  const int32_t* cj = P.cj.data();
  auto walk = [&](auto&& to_S, auto&& to_rhs) {
    // per-edge blocks (ba.py:73-82,86-88): which = 0 Hii, 1 Hij, 2 Hji, 3 Hjj ; v: 0 vi, 1 vj
    for (int c = 0; c < E; ++c) {
      const int i = ci[c] - t0, j = cj[c] - t0;
      const bool oi = i >= 0 && i < Pp, oj = j >= 0 && j < Pp;
      if (oi) {
        to_S(tri(i, i), Item{0, c * 4 + 0, 0, 0});
        to_rhs(i, Item{0, c * 2 + 0, 0, 0});
      }
      if (oj) {
        to_S(tri(j, j), Item{0, c * 4 + 3, 0, 0});
        to_rhs(j, Item{0, c * 2 + 1, 0, 0});
      }
      if (oi && oj) {
        if (i >= j) to_S(tri(i, j), Item{0, c * 4 + 1, 0, 0});   // Hij at (i,j)
        if (j >= i) to_S(tri(j, i), Item{0, c * 4 + 2, 0, 0});   // Hji at (j,i)
      }
    }
    if (motion_only) return;
    // Schur products (chol.py:63-64): keyframe slot m, panel blocks A >= B
    for (int m = 0; m < P.M; ++m) {
      const int e0 = P.koff[m], deg = P.koff[m + 1] - e0, nblk = deg + 1;
      const int npairs = nblk * (nblk + 1) / 2;
      const int k0 = P.kx[m] - t0;
      const int64_t pb = int64_t(P.pairbase[m]) * S, bb = int64_t(P.blkbase[m]) * S;
      for (int A = 0; A < nblk; ++A) {
        const int pA = (A == 0) ? k0 : cj[e0 + A - 1] - t0;
        if (pA < 0 || pA >= Pp) continue;
        for (int B = 0; B <= A; ++B) {
          const int pB = (B == 0) ? k0 : cj[e0 + B - 1] - t0;
          if (pB < 0 || pB >= Pp) continue;
          const int32_t base = int32_t(pb + tri(A, B));
          // X = E_A Q E_B^T sits at (pA,pB); X^T at (pB,pA) when A != B
          if (pA >= pB) to_S(tri(pA, pB), Item{1, base, npairs, 0});
          if (A != B && pB >= pA) to_S(tri(pB, pA), Item{1 | 2, base, npairs, 0});
        }
        to_rhs(pA, Item{1, int32_t(bb + A), nblk, 0});
      }
    }
  };
  P.sblk_off.assign(P.nSblk + 1, 0);
  P.rhs_off.assign(Pp + 1, 0);
  {
    int32_t* so = P.sblk_off.data() + 1;
    int32_t* ro = P.rhs_off.data() + 1;
    walk([&](int dst, const Item&) { so[dst]++; }, [&](int dst, const Item&) { ro[dst]++; });
  }
  for (int i = 0; i < P.nSblk; ++i) P.sblk_off[i + 1] += P.sblk_off[i];
  for (int i = 0; i < Pp; ++i) P.rhs_off[i + 1] += P.rhs_off[i];
  P.sblk_items.resize(size_t(P.sblk_off[P.nSblk]));
  P.rhs_items.resize(size_t(P.rhs_off[Pp]));
  {
    std::vector<int32_t>& cs = P.scratch_cur;
    std::vector<int32_t>& cr = P.scratch_deg;
    cs.assign(P.sblk_off.begin(), P.sblk_off.end() - 1);
    cr.assign(P.rhs_off.begin(), P.rhs_off.end() - 1);
    Item* si = P.sblk_items.data();
    Item* ri = P.rhs_items.data();
    int32_t* csp = cs.data();
    int32_t* crp = cr.data();
    walk([&](int dst, const Item& it) { si[csp[dst]++] = it; }, [&](int dst, const Item& it) { ri[crp[dst]++] = it; });
  }
}

}  // namespace dba
