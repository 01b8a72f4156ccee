// Edge bookkeeping of the factor graph on host int64 arrays (include/dba.h, "graph" section).
// Bit-exact restatement of the reference's index helpers -- order of appended edges is preserved:
//   hislam2/geom/graph_utils.py:20-29        meshgrid / neighbourhood_graph
//   hislam2/factor_graph.py:29-39            __filter_repeated_edges
//   hislam2/factor_graph.py:59-82            add_factors (duplicate filter, max_factors eviction by age)
//   hislam2/factor_graph.py:86-106           rm_factors / rm_keyframe
//   hislam2/factor_graph.py:109-117          add_neighborhood_factors
//   factor_graph.pyc add_proximity_factors@354-424 (stale bytecode; SURVEY.md section 8a restatement)
// Pure C++, no CUDA: these run between `ba` calls on a few thousand edges.
#include "../../include/dba.h"

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <limits>
#include <numeric>
#include <unordered_set>
#include <vector>

namespace {
inline uint64_t key(int64_t i, int64_t j) { return (uint64_t(uint32_t(i)) << 32) | uint64_t(uint32_t(j)); }
}

extern "C" {

// graph_utils.py:25-29: all (i,j) of the n x n grid in row-major order with 1 <= |i-j| <= r.
// Returns the number of edges; writes at most cap of them.
int64_t dba_graph_neighbourhood(int64_t t0, int64_t t1, int64_t r, int64_t* ii, int64_t* jj, int64_t cap) {
  int64_t n = 0;
  for (int64_t i = t0; i < t1; ++i)
    for (int64_t j = t0; j < t1; ++j) {
      const int64_t d = std::llabs(i - j);
      if (d >= 1 && d <= r) {
        if (n < cap && ii && jj) { ii[n] = i; jj[n] = j; }
        ++n;
      }
    }
  return n;
}

// factor_graph.py:29-39: keep[k] = (new_ii[k], new_jj[k]) is not an edge of the CURRENT graph.
// (Duplicates inside the new list itself are kept, as in the reference.)  Returns the number kept.
int64_t dba_graph_filter_repeated(const int64_t* cur_ii, const int64_t* cur_jj, int64_t n_cur, const int64_t* new_ii,
                                  const int64_t* new_jj, int64_t n_new, uint8_t* keep) {
  std::unordered_set<uint64_t> eset;
  eset.reserve(size_t(n_cur) * 2 + 16);
  for (int64_t k = 0; k < n_cur; ++k) eset.insert(key(cur_ii[k], cur_jj[k]));
  int64_t kept = 0;
  for (int64_t k = 0; k < n_new; ++k) {
    keep[k] = eset.count(key(new_ii[k], new_jj[k])) ? 0 : 1;
    kept += keep[k];
  }
  return kept;
}

// factor_graph.py:75-77: when the graph would exceed max_factors, edges are dropped by
//   ix = arange(n)[argsort(age)] ; mask = ix >= max_factors - n_new
// i.e. edge k is removed when the index of the k-th youngest edge is >= the threshold (as written in
// the reference: the mask is applied positionally).  argsort is taken stable here (torch.argsort is
// not; ties in age make the reference implementation-defined).  remove[k] = 1 for dropped edges.
int64_t dba_graph_evict_mask(const int64_t* age, int64_t n, int64_t max_factors, int64_t n_new, uint8_t* remove) {
  std::vector<int64_t> ix(static_cast<size_t>(n), 0);
  std::iota(ix.begin(), ix.end(), int64_t(0));
  std::stable_sort(ix.begin(), ix.end(), [&](int64_t a, int64_t b) { return age[a] < age[b]; });
  int64_t cnt = 0;
  for (int64_t k = 0; k < n; ++k) {
    remove[k] = ix[size_t(k)] >= max_factors - n_new ? 1 : 0;
    cnt += remove[k];
  }
  return cnt;
}

// factor_graph.py:93-106 (edge part): mask m = (ii == ix) | (jj == ix) on the ORIGINAL indices, then
// every index >= ix is decremented, then masked edges are dropped.  In place; returns the new count.
int64_t dba_graph_rm_keyframe(int64_t* ii, int64_t* jj, int64_t* age, int64_t n, int64_t ix) {
  int64_t o = 0;
  for (int64_t k = 0; k < n; ++k) {
    const bool m = ii[k] == ix || jj[k] == ix;
    const int64_t i = ii[k] >= ix ? ii[k] - 1 : ii[k];
    const int64_t j = jj[k] >= ix ? jj[k] - 1 : jj[k];
    if (!m) {
      ii[o] = i; jj[o] = j;
      if (age) age[o] = age[k];
      ++o;
    }
  }
  return o;
}

}  // extern "C"

// ---- proximity selection (factor_graph.pyc add_proximity_factors@354-424) -------------------------
// d: [(t-t0) * (t-t1)] frame distances of the grid ii = arange(t0,t) x jj = arange(t1,t) ('ij' order),
// modified in place exactly like the reference's tensor.  ex_ii/ex_jj: the existing active + inactive
// edges.  Writes the selected edges (both directions, forced temporal neighbours first) to out_ii/out_jj
// (at most cap) and returns their number.  The caller then passes them to add_factors.
//   @366  d[ii - rad < jj] = inf ; @367 d[d > 100] = inf
//   @371-379 suppression around every existing edge ; @382-387 temporal edges i, j in [max(i-rad-1,0), i)
//   @389-397 the backend Euler-angle filter assigns into a temporary (no effect) -- not restated
//   @399-421 greedy selection in ascending distance (stable sort; the reference's argsort is not) with NMS
namespace {
struct Prox {
  float* d; int64_t t0, t1, t, nd;
  void kill(int64_t i, int64_t j) {            // d[(i-t0)*(t-t1) + (j-t1)] = inf with Python index semantics
    int64_t idx = (i - t0) * (t - t1) + (j - t1);
    if (idx < 0) idx += nd;
    if (idx >= 0 && idx < nd) d[idx] = std::numeric_limits<float>::infinity();
  }
  void suppress(int64_t i, int64_t j, int64_t nms) {
    const int64_t lim = std::max<int64_t>(std::min<int64_t>(std::llabs(i - j) - 2, nms), 0);
    for (int64_t di = -nms; di <= nms; ++di)
      for (int64_t dj = -nms; dj <= nms; ++dj)
        if (std::llabs(di) + std::llabs(dj) <= lim) {
          const int64_t i1 = i + di, j1 = j + dj;
          if (t0 <= i1 && i1 < t && t1 <= j1 && j1 < t) kill(i1, j1);
        }
  }
};
}  // namespace

extern "C" int64_t dba_graph_proximity(float* d, int64_t t0, int64_t t1, int64_t t, int64_t rad, int64_t nms,
                                       float thresh, int64_t max_factors, const int64_t* ex_ii, const int64_t* ex_jj,
                                       int64_t n_ex, int64_t* out_ii, int64_t* out_jj, int64_t cap) {
  const int64_t ni = t - t0, nj = t - t1;
  if (ni <= 0 || nj <= 0) return 0;
  Prox P{d, t0, t1, t, ni * nj};
  const float inf = std::numeric_limits<float>::infinity();
  for (int64_t a = 0; a < ni; ++a)
    for (int64_t b = 0; b < nj; ++b) {
      float& v = d[a * nj + b];
      if ((t0 + a) - rad < (t1 + b)) v = inf;
      if (v > 100.f || v != v) v = inf;      // NaN (never produced by frame_distance) is treated as 'no overlap'
    }
  for (int64_t k = 0; k < n_ex; ++k) P.suppress(ex_ii[k], ex_jj[k], nms);
  int64_t n = 0;
  auto emit = [&](int64_t i, int64_t j) {
    if (n < cap) { out_ii[n] = i; out_jj[n] = j; }
    ++n;
  };
  for (int64_t i = t0; i < t; ++i)
    for (int64_t j = std::max<int64_t>(i - rad - 1, 0); j < i; ++j) {
      emit(i, j); emit(j, i);
      P.kill(i, j);
    }
  std::vector<int64_t> order(static_cast<size_t>(P.nd), 0);
  std::iota(order.begin(), order.end(), int64_t(0));
  std::stable_sort(order.begin(), order.end(), [&](int64_t a, int64_t b) { return d[a] < d[b]; });
  for (int64_t k : order) {
    if (!(d[k] <= thresh)) continue;                        // @401 (written `continue`, not `break`)
    if (n > max_factors) break;                             // @404: len(es) > max_factors
    const int64_t i = t0 + k / nj, j = t1 + k % nj;
    emit(i, j); emit(j, i);
    P.suppress(i, j, nms);
  }
  return n;
}
