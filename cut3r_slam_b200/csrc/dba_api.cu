// libdba C ABI (include/dba.h): context, plan cache, workspace, the LM iteration loop and the
// optional NCCL all-reduce of the reduced system for sharded global BA.
#include "../../include/dba.h"

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nvtx3/nvToolsExt.h>

#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <chrono>
#include <cstring>
#include <functional>
#include <queue>
#include <string>
#include <vector>

#include "dba_kernels.h"
#include "dba_plan.h"

namespace dba {
thread_local int64_t g_launch_count = 0;
}

using namespace dba;

namespace {

uint64_t g_devbuf_allocs = 0;        // bumped by every (re)allocation: captured CUDA graphs hold the old addresses

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes, bool zero = false) {
    if (bytes <= cap) return cudaSuccess;
    ++g_devbuf_allocs;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) return e;
    cap = want;
    if (zero) e = cudaMemset(p, 0, want);
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

// ---- NCCL through dlopen: the single-GPU path has no NCCL dependency ---------------------------
struct Id128 { char b[128]; };
struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(void*) = nullptr;
  int (*CommInitRank)(void**, int, /*ncclUniqueId by value*/ Id128, int) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
NcclApi g_nccl;

bool load_nccl(std::string& err) {
  if (g_nccl.lib) return true;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (g_nccl.lib) break;
  }
  if (!g_nccl.lib) { err = "libnccl.so.2 not found (import torch first, or add it to LD_LIBRARY_PATH)"; return false; }
  g_nccl.GetUniqueId = reinterpret_cast<decltype(g_nccl.GetUniqueId)>(dlsym(g_nccl.lib, "ncclGetUniqueId"));
  g_nccl.CommInitRank = reinterpret_cast<decltype(g_nccl.CommInitRank)>(dlsym(g_nccl.lib, "ncclCommInitRank"));
  g_nccl.AllReduce = reinterpret_cast<decltype(g_nccl.AllReduce)>(dlsym(g_nccl.lib, "ncclAllReduce"));
  g_nccl.CommDestroy = reinterpret_cast<decltype(g_nccl.CommDestroy)>(dlsym(g_nccl.lib, "ncclCommDestroy"));
  g_nccl.GetErrorString = reinterpret_cast<decltype(g_nccl.GetErrorString)>(dlsym(g_nccl.lib, "ncclGetErrorString"));
  if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllReduce || !g_nccl.CommDestroy) {
    err = "NCCL symbols missing";
    return false;
  }
  return true;
}

}  // namespace

struct dba_ctx {
  int device = 0;
  std::string err;
  // options
  bool block_diag_quirk = false;
  bool keep_system = false;
  bool profile = false;            // CUDA events around every stage (adds a sync at the end of the call)
  double stage_ms[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // linearize, assemble, allreduce, cholesky, backsub, retract, -, launches
  std::vector<cudaEvent_t> ev;
  cudaStream_t side = nullptr;       // the few high-degree keyframes (SIMT kernel) run beside the tcgen05 launch
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_join2 = nullptr;
  cudaStream_t side2 = nullptr;      // the WIDE tcgen05 keyframes when the SIMT kernel already occupies `side`
  int opt_splits = 0;
  int opt_ppl = 0;
  // plan cache
  Plan plan;
  bool plan_valid = false;
  int plan_ppl = 0, plan_tps = 1, plan_HW = 0, plan_ht = 0, plan_wd = 0;
  uint64_t plan_key = 0;             // caller's identity of the edge lists the plan was built from (0: none)
  DevBuf d_plan_i32, d_items;
  DeviceGraph g{};
  // second plan slot: dba_depth_cov linearises on the fp32 SIMT kernels (see there) and keeps its plan here, so that
  // ba and depth_cov calls alternating on one graph rebuild neither plan (swap_plan_slot)
  struct PlanSlot {
    Plan plan;
    bool valid = false, v2 = false, pending = false;
    int ppl = 0, tps = 1, HW = 0, ht = 0, wd = 0;
    uint64_t key = 0;
    DevBuf i32, i32b, items;
    DeviceGraph g{};
  } cov_slot;
  int32_t last_dims[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // of the plan the last call ran on (debug_read "dims")
  // workspace
  DevBuf Hpart, Spart, vEpart, Q, w, Hblk, vblk, A, Lm, Ldiag, dx, dz, fail, Akeep, solve_msg, solve_flag;
  int coop_ctas = -1, coop_n = -1;   // cooperative Cholesky: co-resident CTA budget for the cached n6
  bool opt_multilaunch_chol = false;
  bool opt_tc = true;
  bool opt_v2 = true;                // narrow keyframes on the one-warp-per-edge kernel (dba_linearize_v2.cu)
  bool plan_v2 = false;              // ... and the cached plan was split for it
  int opt_wide_tc = 2;               // 0: SIMT above out-degree 20; 1: WIDE tcgen05 variant up to 25; 2: up to 41          // out-degree 21..25 on the WIDE tcgen05 variant instead of the SIMT kernel
  int opt_chol_tc = 0;               // 0 off (default: measured slower, profiles/r2_cholesky_tc.md); 1: n >= cholesky_tc_min_n() on 128-column panels + tcgen05 trailing update;
                                     // > 1: from that n on (tests)
  int opt_refine = 1;                // iterative refinement of the pose solve: 0 off, 1 auto, 2 every iteration
  bool opt_small = true;             // single-CTA solve for n6 <= small_system_max_n()                // Schur products on tcgen05 where the out-degree allows
  // host staging
  int64_t* h_idx = nullptr;
  size_t h_idx_cap = 0;
  char* h_plan[2] = {nullptr, nullptr};   // pinned staging of the plan upload, one region per half (pageable copies of
  size_t h_plan_cap[2] = {0, 0};          // the loop-closure plan's 3.7 MB of items took 1.4 - 3.4 ms)
  cudaEvent_t ev_plan = nullptr;          // behind the last upload: the staging regions are free again
  std::vector<int32_t> h_pack;            // packing scratch of the uploads
  bool plan_gather_pending = false;       // the plan's CSR half is on the device, its gather lists are not built yet
  DevBuf d_plan_i32b;                     // offsets / block list of the gather half (d_plan_i32 holds the CSR half)
  // host-entry staging (dba_ba_host)
  DevBuf s_poses, s_disps, s_intr, s_sens, s_targets, s_weights, s_eta, s_ii, s_jj, s_dx, s_dz;
  // proj_trans: CSR of the edges by source keyframe (cached by the edge lists)
  std::vector<int64_t> pt_ii, pt_jj;
  DevBuf d_pt;
  DevBuf d_jd;                 // proj_trans outputs inside dba_jdsa
  DevBuf d_cov;                // L^-1 and S^-1 of dba_depth_cov
  DevBuf d_slot;               // dba_damping_gather: frame -> eta row
  DevBuf d_pack;               // packed lower triangle for the NCCL all-reduce
  // dba_pgba (Sim3): its own plan (S = 1) and workspace
  Plan pg_plan;
  DevBuf pg_i32, pg_items, pg_H, pg_v, pg_Ep, pg_CW, pg_Q, pg_W, pg_Eis, pg_S, pg_vE, pg_Hsp, pg_vsp, pg_A, pg_Lm, pg_Ldiag,
      pg_dx, pg_dz, pg_fail, pg_keep, pg_smsg, pg_sflag;
  int pt_M = 0, pt_N = 0;
  bool pt_contig = true;       // unique(ii) is a contiguous range of frames
  // NCCL
  void* comm = nullptr;
  int nranks = 1, rank = 0;
  // one-shot all-reduce over peer memory (dba_p2p.cu): exchange buffer = 2 slabs + flags, mapped by every rank
  float* p2p_buf = nullptr;
  size_t p2p_slab_floats = 0;
  P2PPeers p2p_peers{};
  bool p2p_ok = false;
  bool opt_p2p = true;
  bool opt_nvtx = false;             // NVTX ranges around the call and its stages
  bool opt_chi2 = false;             // objective before every iteration and after the last one (debug_read "chi2")
  const int64_t* cur_ii_dev = nullptr;   // device edge lists of the call in flight (set by the entry points)
  const int64_t* cur_jj_dev = nullptr;
  DevBuf d_chi2;
  int chi2_rows = 0;
  int64_t chi2_E = 0;
  int opt_p2p_two_shot = 1;          // n > p2p_max_n: reduce-scatter + all-gather over peer memory instead of NCCL:
                                     // 0 never, 1 for two ranks (where it measured faster), 2 always
  int opt_p2p_max_n = 1536;          // larger systems are bandwidth-bound: NCCL's reduce-scatter + all-gather moves less
  uint32_t p2p_epoch = 0;
  int opt_p2p_debug = 0;             // timing experiments only: 1 skip the flag wait, 2 skip the sums
  int64_t last_launches = 0;
  size_t smem_cap = 0;
  // CUDA graphs of whole dba_ba_cached calls (plan hit by key, same buffers and scalars as an earlier call): the
  // second identical call is captured on `gstream`, later ones are one cudaGraphLaunch on the caller's stream
  struct BaGraph {
    const void* ptr[9] = {};         // poses, disps, intr, disps_sens, targets, weights, eta, dx_out, dz_out
    int64_t dims[8] = {};            // E, N, ht, wd, M, t0, t1, iterations
    float lm = 0.f, ep = 0.f;
    bool motion_only = false;
    uint64_t graph_key = 0, epoch = 0, allocs = 0, last_use = 0;
    cudaGraphExec_t exec = nullptr;
    int64_t launches = 0;
    bool failed = false;             // capture or instantiation failed once: this call shape stays on the plain path
  };
  std::vector<BaGraph> graphs;
  uint64_t graph_epoch = 0;          // bumped when the plan or an option changes: every cached graph is stale
  uint64_t graph_clock = 0;
  int64_t graph_replays = 0;
  bool opt_graphs = true;
  cudaStream_t gstream = nullptr;
};

namespace {

int fail_arg(dba_ctx* c, const std::string& m) { c->err = m; return -1; }
int fail_cuda(dba_ctx* c, cudaError_t e, const char* where) {
  c->err = std::string(where) + ": " + cudaGetErrorString(e);
  return int(e) > 0 ? int(e) : 999;
}
#define CK(call, where)                                      \
  do {                                                       \
    cudaError_t e__ = (call);                                \
    if (e__ != cudaSuccess) return fail_cuda(ctx, e__, where); \
  } while (0)

void swap_plan_slot(dba_ctx* c) {
  auto& a = c->cov_slot;
  std::swap(c->plan, a.plan); std::swap(c->plan_valid, a.valid); std::swap(c->plan_v2, a.v2);
  std::swap(c->plan_gather_pending, a.pending);
  std::swap(c->plan_ppl, a.ppl); std::swap(c->plan_tps, a.tps); std::swap(c->plan_HW, a.HW);
  std::swap(c->plan_ht, a.ht); std::swap(c->plan_wd, a.wd); std::swap(c->plan_key, a.key);
  std::swap(c->d_plan_i32, a.i32); std::swap(c->d_plan_i32b, a.i32b); std::swap(c->d_items, a.items); std::swap(c->g, a.g);
}

cudaError_t plan_staging(dba_ctx* c, size_t bytes, int region) {
  if (bytes <= c->h_plan_cap[region]) return cudaSuccess;
  if (c->h_plan[region]) cudaFreeHost(c->h_plan[region]);
  c->h_plan[region] = nullptr; c->h_plan_cap[region] = 0;
  const size_t want = bytes + bytes / 4;
  cudaError_t e = cudaMallocHost(&c->h_plan[region], want);
  if (e == cudaSuccess) c->h_plan_cap[region] = want;
  return e;
}

int out_degree_max(const int64_t* ii, int64_t E, int N, int* M_out) {
  std::vector<int> deg(N, 0);
  int mx = 0, M = 0;
  for (int64_t e = 0; e < E; ++e)
    if (ii[e] >= 0 && ii[e] < N) { if (deg[ii[e]]++ == 0) ++M; if (deg[ii[e]] > mx) mx = deg[ii[e]]; }
  *M_out = M;
  return mx;
}

// plan-cache hit by the caller's graph key alone: no look at the edge lists (they may still be on the device)
bool plan_hit_by_key(dba_ctx* ctx, uint64_t key, int64_t E, int N, int ht, int wd, int64_t M, int t0, int t1,
                     bool motion_only) {
  const Plan& P = ctx->plan;
  return key != 0 && ctx->plan_valid && ctx->plan_key == key && P.E == E && P.N == N && P.t0 == t0 && P.t1 == t1 &&
         P.motion_only == motion_only && ctx->plan_ht == ht && ctx->plan_wd == wd && (motion_only || M < 0 || M == P.M);
}

// (re)build the plan for this graph if it differs from the cached one, upload it
int prepare_plan(dba_ctx* ctx, const int64_t* ii, const int64_t* jj, int64_t E, int N, int ht, int wd, int64_t M,
                 int t0, int t1, bool motion_only, cudaStream_t st) {
  const int HW = ht * wd;
  Plan& P = ctx->plan;
  bool same = ctx->plan_valid && P.E == E && P.N == N && P.t0 == t0 && P.t1 == t1 && P.motion_only == motion_only &&
              ctx->plan_ht == ht && ctx->plan_wd == wd && int64_t(P.key_ii.size()) == E &&
              (E == 0 || (std::memcmp(P.key_ii.data(), ii, E * sizeof(int64_t)) == 0 &&
                          std::memcmp(P.key_jj.data(), jj, E * sizeof(int64_t)) == 0));
  if (same) {
    if (!motion_only && M >= 0 && M != P.M) return fail_arg(ctx, "eta rows != |unique(ii)|");
    return 0;
  }
  ctx->plan_valid = false;
  ++ctx->graph_epoch;
  // DBA_PLAN_TRACE=1: host time of the rebuild's sections on stderr (developer diagnostic for the cold-plan path)
  static const bool trace = std::getenv("DBA_PLAN_TRACE") != nullptr;
  auto now_us = [] { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  double tr_t[5] = {now_us(), 0, 0, 0, 0};
  if (E > 0) {
    for (int64_t e = 0; e < E; ++e)
      if (ii[e] < 0 || ii[e] >= N) return fail_arg(ctx, "edge index outside the frame buffer");
  }
  int Mq = 0;
  const int degmax = out_degree_max(ii, E, N, &Mq);
  // tensor-core path: keyframes with out-degree <= 20; both kernels then share the 64-pixel tiling
  int n_tc_kf = 0, degmax_tc = 0;
  if (ctx->opt_tc && !motion_only) {
    std::vector<int> dg(N, 0);
    for (int64_t e = 0; e < E; ++e) dg[ii[e]]++;
    for (int f = 0; f < N; ++f)
      if (dg[f] > 0 && dg[f] <= linearize_tc_max_degree()) { ++n_tc_kf; if (dg[f] > degmax_tc) degmax_tc = dg[f]; }
    if (n_tc_kf > 0 && linearize_tc_smem_bytes(degmax_tc) > ctx->smem_cap) n_tc_kf = 0;
  }
  bool use_tc = n_tc_kf > 0;
  if (use_tc) {
    // keyframes beyond the widest tcgen05 variant take the SIMT kernel, whose panel must fit at the tcgen05 tile
    // width (ppl = 2: out-degree <~ 105); if it does not, the whole graph takes the SIMT kernel with a narrower tile
    // (out-degree <~ 200) instead of being rejected
    const int tc_cap = ctx->opt_wide_tc >= 2 ? linearize_tc_wide_max_degree(2)
                     : ctx->opt_wide_tc >= 1 ? linearize_tc_wide_max_degree(1) : linearize_tc_max_degree();
    if (degmax > tc_cap && linearize_smem_bytes(degmax, linearize_tc_ppl(), motion_only) > ctx->smem_cap) {
      use_tc = false; n_tc_kf = 0; degmax_tc = 0;
    }
  }
  const bool use_v2 = use_tc && ctx->opt_v2 && linearize_v2_smem_bytes(degmax_tc) <= ctx->smem_cap;
  int ppl = use_tc ? linearize_tc_ppl()
                   : (ctx->opt_ppl > 0 ? ctx->opt_ppl : linearize_pick_ppl(degmax, degmax + 1, motion_only, HW, ctx->smem_cap));
  if (ppl == 0 || linearize_smem_bytes(degmax, ppl, motion_only) > ctx->smem_cap)
    return fail_arg(ctx, "a source keyframe has too many outgoing edges for the shared-memory panel (limit ~200)");
  const int T = 32 * ppl;
  const int ntilesT = (HW + T - 1) / T;
  int S = ctx->opt_splits;
  if (S <= 0) {
    int sm = 148;
    cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, ctx->device);
    if (use_tc) {
      // CTAs are dispatched in block order onto 2 slots per SM: simulate that list schedule for every
      // split count (jobs = keyframes sorted by out-degree x pixel splits) and keep the shortest
      std::vector<int> dg(N, 0), degs, wide;
      for (int64_t e = 0; e < E; ++e) dg[ii[e]]++;
      for (int f = 0; f < N; ++f) {
        if (dg[f] > 0 && dg[f] <= linearize_tc_max_degree()) degs.push_back(dg[f]);
        else if (dg[f] > 0) wide.push_back(dg[f]);
      }
      std::sort(degs.begin(), degs.end(), std::greater<int>());
      std::sort(wide.begin(), wide.end(), std::greater<int>());
      const int slots = use_v2 ? sm : 2 * sm;          // v2: one CTA per SM
      double best = 1e300;
      // measured (backend graph, 1 GPU): flat optimum for 2..5 waves of CTAs, clearly slower below; so only
      // split counts that give >= 3 waves compete (capped at 2 tiles per CTA), the simulation picks among them
      const int njobs = int(degs.size() + 2 * wide.size());
      int cand_lo = 1;
      while (cand_lo < ntilesT && (ntilesT + cand_lo - 1) / cand_lo > 2 && njobs * cand_lo < 3 * slots) ++cand_lo;
      // small windows (frontend: 12 keyframes): when the whole stage fits ONE wave of CTAs with a handful of tiles each,
      // that beats the multi-wave splits above (measured, frontend: S = 26 -> 0.162 ms per call, S = 12 = one wave ->
      // 0.130 ms, S = 13 = 156 CTAs -> 0.148 ms); the simulation decides
      int one_wave = 0;
      if (use_v2 && njobs > 0 && njobs * cand_lo > slots && slots / njobs >= 1) {
        const int s1 = std::min(slots / njobs, ntilesT);
        const int tps1 = (ntilesT + s1 - 1) / s1;
        if (tps1 <= 6) one_wave = (ntilesT + tps1 - 1) / tps1;
      }
      for (int cand = (one_wave > 0 ? cand_lo - 1 : cand_lo); cand <= 52 && cand <= ntilesT; ++cand) {
        const int cand_s = (cand < cand_lo) ? one_wave : cand;       // the extra first round tries the one-wave split
        const int tps_c = (ntilesT + cand_s - 1) / cand_s, S_c = (ntilesT + tps_c - 1) / tps_c;
        if (S_c != cand_s) { continue; }
        {
          // no schedule beats total work / slots (nor its longest job): a candidate whose bound already loses is not
          // simulated (same choice as without the test, checked on 3000 random degree sets; the simulation was 0.5 - 0.7 ms of
          // host time per new graph, 0.14 ms with it)
          double work = 0.0, longest = 0.0;
          const int nt_last = ntilesT - (S_c - 1) * tps_c;
          auto job = [&](int dgk, bool is_wide, int nt) {
            if (use_v2) return 1.4 + nt * (is_wide ? 0.1 * (dgk + 1.0) : 1.0 + 0.03 * dgk);
            return 4.0 + nt * (is_wide ? 0.064 * (dgk + 1.0) * (dgk + 1.0) : dgk + 6.0);
          };
          for (int dgk : wide) {
            work += ((S_c - 1) * job(dgk, true, tps_c) + job(dgk, true, nt_last)) * (use_v2 ? 1.0 : 2.0);
            longest = std::max(longest, job(dgk, true, tps_c));
          }
          for (int dgk : degs) {
            work += (S_c - 1) * job(dgk, false, tps_c) + job(dgk, false, nt_last);
            longest = std::max(longest, job(dgk, false, tps_c));
          }
          const double bound = std::max(work / slots, longest) * (1.0 - 1e-9);
          if (bound * (1.0 + 0.01 * S_c) >= best) continue;
        }
        std::priority_queue<double, std::vector<double>, std::greater<double>> free_at;
        for (int i = 0; i < slots; ++i) free_at.push(0.0);
        double makespan = 0.0;
        // the SIMT CTAs of the wide keyframes start first (side stream) and take a whole SM = two slots;
        // measured: a 64-pixel tile costs ~0.064 (deg+1)^2 of the unit in which a tcgen05 tile costs deg + 6
        if (use_v2) {
          // one CTA per SM for every kernel.  Unit ~ one v2 tile (one warp per edge: the tile time hardly depends on
          // the out-degree); a WIDE tcgen05 tile of out-degree d costs ~0.1 (d + 1) of it; a CTA costs ~1.4 tiles of
          // prologue + epilogue (split sweep, profiles/r2_linearize_v2.md)
          for (int dgk : wide)
            for (int sp = 0; sp < S_c; ++sp) {
              const int nt = std::min(tps_c, ntilesT - sp * tps_c);
              const double t0s = free_at.top(); free_at.pop();
              const double t1s = t0s + 1.4 + nt * 0.1 * (dgk + 1.0);
              free_at.push(t1s);
              if (t1s > makespan) makespan = t1s;
            }
          for (int dgk : degs)
            for (int sp = 0; sp < S_c; ++sp) {
              const int nt = std::min(tps_c, ntilesT - sp * tps_c);
              const double t0s = free_at.top(); free_at.pop();
              const double t1s = t0s + 1.4 + nt * (1.0 + 0.03 * dgk);      // measured: ~6 us per CTA, ~4.3 us per tile
              free_at.push(t1s);
              if (t1s > makespan) makespan = t1s;
            }
        } else {
        // the SIMT CTAs of the wide keyframes start first (side stream) and take a whole SM = two slots;
        // measured: a 64-pixel tile costs ~0.064 (deg+1)^2 of the unit in which a tcgen05 tile costs deg + 6
        for (int dgk : wide)
          for (int sp = 0; sp < S_c; ++sp) {
            const int nt = std::min(tps_c, ntilesT - sp * tps_c);
            const double a = free_at.top(); free_at.pop();
            const double b = free_at.top(); free_at.pop();
            const double t1s = std::max(a, b) + 4.0 + nt * 0.064 * (dgk + 1.0) * (dgk + 1.0);
            free_at.push(t1s); free_at.push(t1s);
            if (t1s > makespan) makespan = t1s;
          }
        for (int dgk : degs)
          for (int sp = 0; sp < S_c; ++sp) {
            const int nt = std::min(tps_c, ntilesT - sp * tps_c);
            const double t0s = free_at.top();
            free_at.pop();
            const double t1s = t0s + 4.0 + nt * (dgk + 6.0);      // fixed prologue/epilogue + per-tile work
            free_at.push(t1s);
            if (t1s > makespan) makespan = t1s;
          }
        }
        const double score = makespan * (1.0 + 0.01 * S_c);         // partial slabs cost assembly traffic
        if (score < best) { best = score; S = S_c; }
      }
    } else {
      S = (4 * sm + (Mq > 0 ? Mq : 1) - 1) / (Mq > 0 ? Mq : 1);
    }
  }
  if (S > ntilesT) S = ntilesT;
  if (S < 1) S = 1;
  const int tps = (ntilesT + S - 1) / S;
  S = (ntilesT + tps - 1) / tps;     // no empty split
  tr_t[1] = now_us();
  // first half of the plan: what the linearisation needs.  The gather lists of the assembly (finish_plan) are built
  // while the first linearisation of the new graph runs
  if (!build_plan_csr(P, ii, jj, E, N, t0, t1, motion_only, M, S)) return fail_arg(ctx, P.error);
  tr_t[2] = now_us();
  ctx->plan_v2 = use_v2;
  ctx->plan_ppl = ppl; ctx->plan_tps = tps; ctx->plan_HW = HW; ctx->plan_ht = ht; ctx->plan_wd = wd;

  // kernel assignment of the source keyframes
  std::vector<int32_t> tc_list, simt_list, wide_list, wide2_list;
  int degmax_wide = 0, degmax_wide2 = 0;
  const bool use_wide = use_tc && ctx->opt_wide_tc >= 1 &&
                        linearize_tc_smem_bytes(linearize_tc_wide_max_degree(1), 1) <= ctx->smem_cap;
  const bool use_wide2 = use_wide && ctx->opt_wide_tc >= 2 &&
                         linearize_tc_smem_bytes(linearize_tc_wide_max_degree(2), 2) <= ctx->smem_cap;
  for (int m = 0; m < P.M; ++m) {
    const int dgm = P.koff[m + 1] - P.koff[m];
    if (use_tc && dgm <= linearize_tc_max_degree()) tc_list.push_back(m);
    else if (use_wide && dgm <= linearize_tc_wide_max_degree(1)) { wide_list.push_back(m); degmax_wide = std::max(degmax_wide, dgm); }
    else if (use_wide2 && dgm <= linearize_tc_wide_max_degree(2)) { wide2_list.push_back(m); degmax_wide2 = std::max(degmax_wide2, dgm); }
    else simt_list.push_back(m);
  }
  auto by_degree = [&](int32_t a, int32_t b) { return (P.koff[a + 1] - P.koff[a]) > (P.koff[b + 1] - P.koff[b]); };
  std::stable_sort(tc_list.begin(), tc_list.end(), by_degree);       // heavy keyframes are dispatched first
  std::stable_sort(simt_list.begin(), simt_list.end(), by_degree);
  std::stable_sort(wide_list.begin(), wide_list.end(), by_degree);
  std::stable_sort(wide2_list.begin(), wide2_list.end(), by_degree);
  // non-source frames of [0,t1)
  std::vector<int32_t> nonsrc;
  for (int f = 0; f < t1; ++f) if (P.slot_of_frame[f] < 0) nonsrc.push_back(f);

  // pack the int32 arrays of this half into one upload
  std::vector<int32_t>& pack = ctx->h_pack;
  pack.clear();
  auto put = [&](const std::vector<int32_t>& v) {
    size_t off = pack.size();
    pack.insert(pack.end(), v.begin(), v.end());
    while (pack.size() % 4) pack.push_back(0);
    return off;
  };
  size_t o_kx = put(P.kx), o_koff = put(P.koff), o_cedge = put(P.cedge), o_ci = put(P.ci), o_cj = put(P.cj);
  size_t o_pb = put(P.pairbase), o_bb = put(P.blkbase), o_ns = put(nonsrc);
  size_t o_tc = put(tc_list), o_sl = put(simt_list), o_wl = put(wide_list), o_wl2 = put(wide2_list);
  tr_t[3] = now_us();
  CK(ctx->d_plan_i32.ensure(pack.size() * 4 + 16), "plan alloc");
  const size_t b_pack = pack.size() * 4;
  // pinned staging, two regions: [this half | the gather half]; the previous rebuild's copies have left it
  if (ctx->ev_plan) CK(cudaEventSynchronize(ctx->ev_plan), "plan upload");
  CK(plan_staging(ctx, b_pack + 64, 0), "pinned staging");
  std::memcpy(ctx->h_plan[0], pack.data(), b_pack);
  if (b_pack) CK(cudaMemcpyAsync(ctx->d_plan_i32.p, ctx->h_plan[0], b_pack, cudaMemcpyHostToDevice, st), "plan upload");
  if (!ctx->ev_plan) CK(cudaEventCreateWithFlags(&ctx->ev_plan, cudaEventDisableTiming), "plan upload");
  CK(cudaEventRecord(ctx->ev_plan, st), "plan upload");       // re-recorded behind the gather half's copies

  DeviceGraph& g = ctx->g;
  const int32_t* base = ctx->d_plan_i32.as<int32_t>();
  g.E = P.E; g.N = N; g.t0 = t0; g.t1 = t1; g.Pp = P.Pp; g.M = P.M; g.n6 = P.n6; g.S = P.S;
  g.degmax = P.degmax; g.nblkmax = P.nblkmax; g.npairs_max = P.npairs_max;
  g.HW = HW; g.ht = ht; g.wd = wd;
  g.kx = base + o_kx; g.koff = base + o_koff; g.cedge = base + o_cedge; g.ci = base + o_ci; g.cj = base + o_cj;
  g.pairbase = base + o_pb; g.blkbase = base + o_bb; g.nonsrc = base + o_ns; g.n_nonsrc = int(nonsrc.size());
  g.tc_list = base + o_tc; g.simt_list = base + o_sl; g.n_tc = int(tc_list.size()); g.n_simt = int(simt_list.size());
  g.degmax_tc = degmax_tc;
  g.wide_list = base + o_wl; g.n_wide = int(wide_list.size()); g.degmax_wide = degmax_wide;
  g.wide2_list = base + o_wl2; g.n_wide2 = int(wide2_list.size()); g.degmax_wide2 = degmax_wide2;
  g.nSblk = P.nSblk;
  g.sblk_off = nullptr; g.rhs_off = nullptr; g.sblk_items = nullptr; g.rhs_items = nullptr; g.sblk_list = nullptr; g.n_light = 0;
  ctx->plan_gather_pending = true;     // plan_valid stays false until finish_plan has run
  tr_t[4] = now_us();
  if (trace)
    std::fprintf(stderr, "[dba plan] E=%lld S=%d: split choice %.0f us, CSR half %.0f us, kernel lists + pack %.0f us, upload %.0f us\n",
                 (long long)E, S, tr_t[1] - tr_t[0], tr_t[2] - tr_t[1], tr_t[3] - tr_t[2], tr_t[4] - tr_t[3]);
  return 0;
}

// second half of a rebuilt plan: the gather lists of the assembly, uploaded on `st` (run_ba calls it behind the launches
// of the new graph's first linearisation, so the host work overlaps that kernel)
int finish_plan(dba_ctx* ctx, cudaStream_t st) {
  if (!ctx->plan_gather_pending) return 0;
  static const bool trace = std::getenv("DBA_PLAN_TRACE") != nullptr;
  auto now_us = [] { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  const double t_a = now_us();
  Plan& P = ctx->plan;
  build_plan_gather(P);
  const double t_b = now_us();
  std::vector<int32_t>& pack = ctx->h_pack;
  pack.clear();
  auto put = [&](const std::vector<int32_t>& v) {
    size_t off = pack.size();
    pack.insert(pack.end(), v.begin(), v.end());
    while (pack.size() % 4) pack.push_back(0);
    return off;
  };
  size_t o_so = put(P.sblk_off), o_ro = put(P.rhs_off);
  std::vector<int32_t> blist, heavy;
  // small systems: every block gets its own CTA (8 item lanes); large ones: blocks with <= 8 items share a CTA
  const bool all_heavy = P.nSblk <= 1184;
  for (int ob = 0; ob < P.nSblk; ++ob)
    (!all_heavy && P.sblk_off[ob + 1] - P.sblk_off[ob] <= 8 ? blist : heavy).push_back(ob);
  const int n_light = int(blist.size());
  blist.insert(blist.end(), heavy.begin(), heavy.end());
  size_t o_bl = put(blist);
  const size_t n_s = P.sblk_items.size(), n_r = P.rhs_items.size();
  const size_t b_pack = pack.size() * 4, b_s = n_s * sizeof(Item), b_r = n_r * sizeof(Item);
  CK(ctx->d_plan_i32b.ensure(b_pack + 16), "plan alloc");
  CK(ctx->d_items.ensure(b_s + b_r + 16), "plan alloc");
  CK(plan_staging(ctx, b_pack + b_s + b_r + 64, 1), "pinned staging");
  char* h = ctx->h_plan[1];
  std::memcpy(h, pack.data(), b_pack);
  if (b_s) std::memcpy(h + b_pack, P.sblk_items.data(), b_s);
  if (b_r) std::memcpy(h + b_pack + b_s, P.rhs_items.data(), b_r);
  if (b_pack) CK(cudaMemcpyAsync(ctx->d_plan_i32b.p, h, b_pack, cudaMemcpyHostToDevice, st), "plan upload");
  if (b_s + b_r) CK(cudaMemcpyAsync(ctx->d_items.p, h + b_pack, b_s + b_r, cudaMemcpyHostToDevice, st), "plan upload");
  CK(cudaEventRecord(ctx->ev_plan, st), "plan upload");       // the staging regions are free again behind this
  DeviceGraph& g = ctx->g;
  const int32_t* base = ctx->d_plan_i32b.as<int32_t>();
  g.sblk_off = base + o_so; g.rhs_off = base + o_ro;
  g.sblk_items = ctx->d_items.as<Item>();
  g.rhs_items = ctx->d_items.as<Item>() + n_s;
  g.sblk_list = base + o_bl; g.n_light = n_light;
  ctx->plan_gather_pending = false;
  ctx->plan_valid = true;
  if (trace)
    std::fprintf(stderr, "[dba plan]   gather half (behind the first linearisation): lists %.0f us, pack + upload %.0f us\n",
                 t_b - t_a, now_us() - t_b);
  return 0;
}

int ensure_workspace(dba_ctx* ctx, Workspace& ws) {
  const Plan& P = ctx->plan;
  const size_t HW = size_t(ctx->plan_HW);
  const size_t f = sizeof(float);
  CK(ctx->Hpart.ensure(size_t(P.E) * P.S * 32 * f + 16), "workspace");
  CK(ctx->Spart.ensure(size_t(P.slab_blocks) * 36 * f + 16), "workspace");
  CK(ctx->vEpart.ensure(size_t(P.vE_blocks) * 6 * f + 16), "workspace");
  CK(ctx->Q.ensure(size_t(P.M) * HW * f + 16), "workspace");
  CK(ctx->w.ensure(size_t(P.M) * HW * f + 16), "workspace");
  CK(ctx->dz.ensure(size_t(P.M) * HW * f + 16), "workspace");
  CK(ctx->Hblk.ensure(size_t(P.E) * 144 * f + 16), "workspace");
  CK(ctx->vblk.ensure(size_t(P.E) * 12 * f + 16), "workspace");
  const int ld = (P.n6 + 1 + 3) & ~3;
  CK(ctx->A.ensure(size_t(P.n6 + 1) * ld * f, true), "workspace");
  CK(ctx->Ldiag.ensure(cholesky_ldiag_floats(P.n6) * f + 16), "workspace");
  CK(ctx->Lm.ensure(size_t(P.n6 + 1) * ld * f), "workspace");
  if (ctx->coop_n != P.n6) { ctx->coop_ctas = cholesky_coop_max_ctas(P.n6); ctx->coop_n = P.n6; }
  CK(ctx->dx.ensure(size_t(P.n6) * f + 16), "workspace");
  CK(ctx->fail.ensure(16, true), "workspace");
  if (ctx->keep_system || ctx->opt_refine) CK(ctx->Akeep.ensure(size_t(P.n6 + 1) * ld * f), "workspace");
  ws.Hpart = ctx->Hpart.as<float>(); ws.Spart = ctx->Spart.as<float>(); ws.vEpart = ctx->vEpart.as<float>();
  ws.Q = ctx->Q.as<float>(); ws.w = ctx->w.as<float>(); ws.Hblk = ctx->Hblk.as<float>(); ws.vblk = ctx->vblk.as<float>();
  ws.A = ctx->A.as<float>(); ws.Lm = ctx->Lm.as<float>(); ws.ldA = ld; ws.Ldiag = ctx->Ldiag.as<float>(); ws.dx = ctx->dx.as<float>();
  ws.dz = ctx->dz.as<float>(); ws.fail = ctx->fail.as<int>();
  if (trisolve_grid_msg_floats(P.n6) > 0) {          // message slots of the grid-chain triangular solves (n > 768)
    CK(ctx->solve_msg.ensure(trisolve_grid_msg_floats(P.n6) * f), "workspace");
    CK(ctx->solve_flag.ensure(trisolve_grid_flag_words(P.n6) * sizeof(unsigned), true), "workspace");
    ws.solve_msg = ctx->solve_msg.as<float>(); ws.solve_flag = ctx->solve_flag.as<unsigned>();
  }
  return 0;
}

// The iteration loop on device pointers; ii/jj here are HOST copies of the edge lists.
int run_ba(dba_ctx* ctx, float* poses, float* disps, const float* intr, const float* disps_sens,
           const float* targets, const float* weights, const float* eta, const int64_t* ii_h, const int64_t* jj_h,
           int64_t E, int64_t N, int ht, int wd, int64_t M, int t0, int t1, int iterations, float lm, float ep,
           bool motion_only, float* dx_out, float* dz_out, cudaStream_t st, float* cov_out = nullptr,
           uint64_t graph_key = 0, bool capturing = false) {
  g_launch_count = 0;
  int rc = 0;
  // ii_h == nullptr: the caller's graph key matched the cached plan (plan_hit_by_key), nothing to look at
  if (ii_h != nullptr || E == 0) {
    rc = prepare_plan(ctx, ii_h, jj_h, E, int(N), ht, wd, motion_only ? -1 : M, t0, t1, motion_only, st);
    if (rc) return rc;
    ctx->plan_key = graph_key;
  }
  {
    const int32_t d[8] = {ctx->plan.n6, (ctx->plan.n6 + 1 + 3) & ~3, ctx->plan.Pp, ctx->plan.M, ctx->plan.S, ctx->plan_ppl,
                          ctx->plan.degmax, ctx->plan.npairs_max};
    std::memcpy(ctx->last_dims, d, sizeof(d));
  }
  Workspace ws{};
  rc = ensure_workspace(ctx, ws);
  if (rc) return rc;
  if (ctx->plan_gather_pending && iterations <= 0) { rc = finish_plan(ctx, st); if (rc) return rc; }
  // ---- whole-call CUDA graph: a call that repeats an earlier one exactly (plan hit by the caller's key, same
  // buffers, same scalars) is captured once and then replayed with ONE launch.  What this buys is launch latency:
  // the frontend / motion-only windows are 12-13 dependent kernels of 5-40 us per LM iteration.
  if (!capturing && ctx->opt_graphs && ii_h == nullptr && E > 0 && graph_key != 0 && cov_out == nullptr &&
      iterations > 0 && !ctx->profile && !ctx->keep_system && !ctx->opt_chi2 && !(ctx->comm != nullptr && ctx->nranks > 1)) {
    const void* ptr[9] = {poses, disps, intr, disps_sens, targets, weights, eta, dx_out, dz_out};
    const int64_t dims[8] = {E, N, ht, wd, M, t0, t1, iterations};
    dba_ctx::BaGraph* hit = nullptr;
    for (auto it = ctx->graphs.begin(); it != ctx->graphs.end();) {
      if (it->epoch != ctx->graph_epoch || it->allocs != g_devbuf_allocs) {        // stale: plan / option / workspace changed
        if (it->exec) cudaGraphExecDestroy(it->exec);
        it = ctx->graphs.erase(it);
        continue;
      }
      if (it->graph_key == graph_key && it->lm == lm && it->ep == ep && it->motion_only == motion_only &&
          std::memcmp(it->ptr, ptr, sizeof(ptr)) == 0 && std::memcmp(it->dims, dims, sizeof(dims)) == 0)
        hit = &*it;
      ++it;
    }
    if (hit != nullptr && hit->exec != nullptr) {
      hit->last_use = ++ctx->graph_clock;
      CK(cudaGraphLaunch(hit->exec, st), "graph launch");
      ++ctx->graph_replays;
      ctx->last_launches = hit->launches;
      return 0;
    }
    if (hit != nullptr && !hit->failed) {            // second identical call: capture it
      if (ctx->gstream == nullptr) CK(cudaStreamCreateWithFlags(&ctx->gstream, cudaStreamNonBlocking), "graph stream");
      cudaGraph_t graph = nullptr;
      cudaError_t ce = cudaStreamBeginCapture(ctx->gstream, cudaStreamCaptureModeThreadLocal);
      int crc = 0;
      if (ce == cudaSuccess) {
        crc = run_ba(ctx, poses, disps, intr, disps_sens, targets, weights, eta, nullptr, nullptr, E, N, ht, wd, M, t0, t1,
                     iterations, lm, ep, motion_only, dx_out, dz_out, ctx->gstream, nullptr, graph_key, true);
        ce = cudaStreamEndCapture(ctx->gstream, &graph);
      }
      cudaGraphExec_t exec = nullptr;
      if (ce == cudaSuccess && crc == 0 && graph != nullptr) ce = cudaGraphInstantiate(&exec, graph, 0);
      if (graph != nullptr) cudaGraphDestroy(graph);
      if (ce == cudaSuccess && crc == 0 && exec != nullptr) {
        hit->exec = exec;
        hit->launches = ctx->last_launches;
        hit->last_use = ++ctx->graph_clock;
        CK(cudaGraphLaunch(exec, st), "graph launch");
        ++ctx->graph_replays;
        return 0;
      }
      if (exec != nullptr) cudaGraphExecDestroy(exec);
      cudaGetLastError();                            // the capture failed: this call shape stays on the plain path
      hit->failed = true;
      g_launch_count = 0;
    } else if (hit == nullptr) {                     // first sighting: remember it, run it plainly
      if (ctx->graphs.size() >= 8) {
        size_t lru = 0;
        for (size_t i = 1; i < ctx->graphs.size(); ++i) if (ctx->graphs[i].last_use < ctx->graphs[lru].last_use) lru = i;
        if (ctx->graphs[lru].exec) cudaGraphExecDestroy(ctx->graphs[lru].exec);
        ctx->graphs.erase(ctx->graphs.begin() + lru);
      }
      dba_ctx::BaGraph ge;
      std::memcpy(ge.ptr, ptr, sizeof(ptr));
      std::memcpy(ge.dims, dims, sizeof(dims));
      ge.lm = lm; ge.ep = ep; ge.motion_only = motion_only; ge.graph_key = graph_key;
      ge.epoch = ctx->graph_epoch; ge.allocs = g_devbuf_allocs; ge.last_use = ++ctx->graph_clock;
      ctx->graphs.push_back(ge);
    }
  }
  const DeviceGraph& g = ctx->g;
  BAInputs in{poses, disps, intr, disps_sens, targets, weights, eta};
  // sharded: every rank damps its share of H; the constant ep is added once (rank 0)
  const float ep_eff = (ctx->rank == 0) ? ep : 0.f;
  const bool prof = ctx->profile;
  size_t nev = 0;
  if (prof) {
    while (ctx->ev.size() < size_t(iterations) * 7 + 1) {
      cudaEvent_t e;
      CK(cudaEventCreate(&e), "event");
      ctx->ev.push_back(e);
    }
  }
  // option "nvtx": host-side ranges around the call and the stage launches (Nsight Systems timelines; no cost when off)
  static const char* const kStage[6] = {"dba linearize", "dba assemble", "dba all-reduce", "dba cholesky+solve", "dba backsub", "dba retract"};
  const bool nvtx = ctx->opt_nvtx;
  int nvtx_stage = 0;
  if (nvtx) nvtxRangePushA(capturing ? "dba_ba (capture)" : "dba_ba");
  auto mark = [&]() {
    if (prof) cudaEventRecord(ctx->ev[nev++], st);
    if (nvtx) {
      if (nvtx_stage > 0) nvtxRangePop();
      if (nvtx_stage < 6) nvtxRangePushA(kStage[nvtx_stage]);
      nvtx_stage = (nvtx_stage + 1) % 7;
    }
  };
  struct NvtxCloser {          // every early return of the loop below leaves the ranges balanced
    bool on; int* stage;
    ~NvtxCloser() { if (on) { if (*stage > 0) nvtxRangePop(); nvtxRangePop(); } }
  } nvtx_closer{nvtx, &nvtx_stage};
  const bool chi2 = ctx->opt_chi2 && !capturing && cov_out == nullptr && g.E > 0 && ctx->cur_ii_dev != nullptr && iterations > 0;
  if (chi2) {
    CK(ctx->d_chi2.ensure(size_t(iterations + 1) * g.E * sizeof(double) + 16), "chi2 buffer");
    ctx->chi2_rows = iterations + 1;
    ctx->chi2_E = g.E;
  } else if (!capturing) {
    ctx->chi2_rows = 0;
  }
  auto chi2_at = [&](int row) {
    return launch_chi2(poses, disps, intr, targets, weights, ctx->cur_ii_dev, ctx->cur_jj_dev, g.E, g.N, g.HW, g.wd,
                       ctx->d_chi2.as<double>() + size_t(row) * g.E, st);
  };
  for (int it = 0; it < iterations; ++it) {
    if (chi2) CK(chi2_at(it), "chi2");
    mark();
    if (g.E > 0) {
      const bool fork = g.n_tc > 0 && (g.n_simt > 0 || g.n_wide > 0 || g.n_wide2 > 0) && ctx->side != nullptr;
      cudaStream_t s2 = fork ? ctx->side : st;
      // the WIDE launch gets its own stream when the SIMT kernel is on `side`, so that the two run side by side
      cudaStream_t sw = (fork && (g.n_simt > 0 || g.n_wide2 > 0) && g.n_wide > 0 && ctx->side2 != nullptr) ? ctx->side2 : s2;
      if (fork) {   // long one-per-SM CTAs first (side streams) so that they do not form the tail of the stage
        CK(cudaEventRecord(ctx->ev_fork, st), "fork");
        CK(cudaStreamWaitEvent(ctx->side, ctx->ev_fork, 0), "fork");
        if (sw != s2) CK(cudaStreamWaitEvent(sw, ctx->ev_fork, 0), "fork");
      }
      if (fork && g.n_simt > 0)
        CK(launch_linearize(g, in, ws, motion_only, ctx->plan_tps, ctx->plan_ppl, g.simt_list, g.n_simt, s2), "linearize");
      if (fork && g.n_wide2 > 0)
        CK(launch_linearize_tc(g, in, ws, ctx->plan_tps, g.wide2_list, g.n_wide2, g.degmax_wide2, 2, s2), "linearize (tcgen05, wide 2)");
      if (fork && g.n_wide > 0)
        CK(launch_linearize_tc(g, in, ws, ctx->plan_tps, g.wide_list, g.n_wide, g.degmax_wide, 1, sw), "linearize (tcgen05, wide)");
      if (fork) {
        CK(cudaEventRecord(ctx->ev_join, ctx->side), "join");
        if (sw != s2) CK(cudaEventRecord(ctx->ev_join2, sw), "join");
      }
      if (g.n_tc > 0 && ctx->plan_v2)
        CK(launch_linearize_v2(g, in, ws, ctx->plan_tps, g.tc_list, g.n_tc, g.degmax_tc, st), "linearize (v2)");
      else if (g.n_tc > 0)
        CK(launch_linearize_tc(g, in, ws, ctx->plan_tps, g.tc_list, g.n_tc, g.degmax_tc, 0, st), "linearize (tcgen05)");
      if (fork) {
        CK(cudaStreamWaitEvent(st, ctx->ev_join, 0), "join");
        if (sw != s2) CK(cudaStreamWaitEvent(st, ctx->ev_join2, 0), "join");
      } else {
        if (g.n_wide2 > 0)
          CK(launch_linearize_tc(g, in, ws, ctx->plan_tps, g.wide2_list, g.n_wide2, g.degmax_wide2, 2, st), "linearize (tcgen05, wide 2)");
        if (g.n_wide > 0)
          CK(launch_linearize_tc(g, in, ws, ctx->plan_tps, g.wide_list, g.n_wide, g.degmax_wide, 1, st), "linearize (tcgen05, wide)");
        if (g.n_simt > 0)
          CK(launch_linearize(g, in, ws, motion_only, ctx->plan_tps, ctx->plan_ppl, g.simt_list, g.n_simt, st), "linearize");
      }
    }
    mark();
    // a rebuilt plan: its gather lists are made now, on the host, while the linearisation above runs
    if (ctx->plan_gather_pending) { rc = finish_plan(ctx, st); if (rc) return rc; }
    const bool sharded = ctx->comm != nullptr && ctx->nranks > 1;
    const bool p2p_fits = sharded && ctx->p2p_ok && ctx->opt_p2p && size_t(g.n6 + 1) * ws.ldA <= ctx->p2p_slab_floats;
    // measured on loopclosure (n = 3060): 2 GPUs 17.69 ms against 18.05 with the packed ncclAllReduce; 4 GPUs 14.53
    // against 14.34, 8 GPUs 13.64 against 13.32 (NCCL reduces inside the switch there): on by default for two ranks
    const bool two_shot = p2p_fits && g.n6 > ctx->opt_p2p_max_n &&
                          (ctx->opt_p2p_two_shot >= 2 || (ctx->opt_p2p_two_shot == 1 && ctx->nranks <= 2));
    const bool use_p2p = (p2p_fits && g.n6 <= ctx->opt_p2p_max_n) || two_shot;
    if (use_p2p) {
      // the rank's share goes into its exchange slab; ONE kernel signals, waits and sums all ranks' lower triangles
      // (large systems: reduce-scatter + all-gather over the same slabs, dba_p2p.cu)
      ++ctx->p2p_epoch;
      const int parity = int(ctx->p2p_epoch & 1u);
      Workspace wsx = ws;
      wsx.A = ctx->p2p_buf + size_t(parity) * ctx->p2p_slab_floats;
      CK(launch_assemble(g, in, wsx, lm, ep_eff, 0.f, motion_only, ctx->block_diag_quirk && motion_only, st), "assemble");
      mark();
      if (two_shot)
        CK(launch_p2p_allreduce_two_shot(ctx->p2p_peers, ctx->nranks, ctx->rank, ctx->p2p_slab_floats, 3 * ctx->p2p_slab_floats,
                                         parity, ctx->p2p_epoch, ws.ldA, g.n6, ws.A, ws.fail + 2, st), "all-reduce (peer memory, two-shot)");
      else
        CK(launch_p2p_allreduce(ctx->p2p_peers, ctx->nranks, ctx->rank, ctx->p2p_slab_floats, 3 * ctx->p2p_slab_floats,
                                parity, ctx->p2p_epoch, ws.ldA, g.n6, ws.A, ws.fail + 2, ctx->opt_p2p_debug, st), "all-reduce (peer memory)");
    } else {
    CK(launch_assemble(g, in, ws, lm, ep_eff, 0.f, motion_only, ctx->block_diag_quirk && motion_only, st), "assemble");
    mark();
    }
    if (sharded && !use_p2p) {
      // lower triangle + rhs row packed: half the bytes of the square on NVLink (18.8 MB instead of 37.6 at n = 3060)
      const size_t cnt = tri_pack_floats(g.n6);
      CK(ctx->d_pack.ensure(cnt * sizeof(float) + 16), "all-reduce buffer");
      CK(launch_tri_pack(ws.A, ctx->d_pack.as<float>(), ws.ldA, g.n6, true, st), "pack");
      int r = g_nccl.AllReduce(ctx->d_pack.p, ctx->d_pack.p, cnt, /*ncclFloat32*/ 7, /*ncclSum*/ 0, ctx->comm, st);
      if (r == 0) CK(launch_tri_pack(ws.A, ctx->d_pack.as<float>(), ws.ldA, g.n6, false, st), "unpack");
      if (r != 0) {
        ctx->err = std::string("ncclAllReduce: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
        return 1000 + r;
      }
    }
    // iterative refinement (dba_refine.cu) of the cooperative solve on the LAST TWO iterations of the call: an earlier
    // iteration's solver error (4e-4 of dx at n = 756, 6e-3 at n = 3060) is corrected by the relinearisations that
    // follow, the last one's would stay in the result -- and so does most of the one before it along weakly observed
    // directions, where one more iteration contracts little (a chain-end keyframe on a 2-iteration call: disps 3.7e-4
    // from the fp64 result against 2.5e-5 with both iterations refined, the reference's own fp32 run 7.7e-5;
    // profiles/r2_parity_margins.md; motion-only calls have no disparity update to amplify it and keep refining the
    // last iteration only).  "refine" = 2: every iteration (+39 us each at n = 756: the copy of the system, the fp64
    // residual, two chain solves)
    const bool coop_path = cov_out == nullptr && !(ctx->opt_small && g.n6 <= cholesky_small_max_n()) &&
                           ctx->coop_ctas > 0 && !ctx->opt_multilaunch_chol;
    const bool refine = coop_path && ctx->opt_refine != 0 && (ctx->opt_refine >= 2 || it >= iterations - (motion_only ? 1 : 2));
    if (ctx->keep_system || refine)
      CK(cudaMemcpyAsync(ctx->Akeep.p, ws.A, size_t(g.n6 + 1) * ws.ldA * sizeof(float), cudaMemcpyDeviceToDevice, st),
         "keep_system");
    mark();
    if (cov_out != nullptr) {
      // depth covariance of THIS linearisation (chol.py:74-76): factor, then S^-1 and the per-pixel quadratic forms;
      // the state is left untouched
      if (ctx->coop_ctas <= 0) return fail_arg(ctx, "depth covariance needs the cooperative factorisation (system too large)");
      CK(launch_cholesky_coop(ws, g.n6, ctx->coop_ctas, st), "cholesky (cooperative)");
      CK(ctx->d_cov.ensure(cov_scratch_floats(g.n6) * sizeof(float) + 16), "covariance scratch");
      CK(launch_depth_cov(g, in, ws, ctx->d_cov.as<float>(), ctx->smem_cap, cov_out, st), "depth covariance");
      break;
    }
    if (ctx->opt_small && g.n6 <= cholesky_small_max_n())
      CK(launch_cholesky_small(ws, g.n6, st), "cholesky (single CTA)");
    else if (ctx->coop_ctas > 0 && !ctx->opt_multilaunch_chol) {
      if (ctx->opt_chol_tc != 0 && g.n6 >= (ctx->opt_chol_tc > 1 ? ctx->opt_chol_tc : cholesky_tc_min_n()))
        CK(launch_cholesky_tc(ws, g.n6, st, !refine), "cholesky (tcgen05 panels)");
      else
        CK(launch_cholesky_coop(ws, g.n6, ctx->coop_ctas, st, !refine), "cholesky (cooperative)");
      if (refine) CK(launch_refine(ws, ctx->Akeep.as<float>(), g.n6, st), "refinement");
    }
    else
      CK(launch_cholesky_solve(ws, g.n6, st), "cholesky");
    mark();
    if (!motion_only && g.E > 0) CK(launch_backsub(g, in, ws, st), "backsub");
    mark();
    CK(launch_retract(g, in, ws, (it == iterations - 1) ? dx_out : nullptr, st), "retract");
    mark();
  }
  if (cov_out != nullptr) { ctx->last_launches = g_launch_count; return 0; }
  if (chi2) CK(chi2_at(iterations), "chi2");
  if (prof && nev > 0) {
    CK(cudaEventSynchronize(ctx->ev[nev - 1]), "profile");
    for (int it = 0; it < iterations; ++it)
      for (int k = 0; k < 6; ++k) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, ctx->ev[it * 7 + k], ctx->ev[it * 7 + k + 1]);
        ctx->stage_ms[k] += ms;
      }
    ctx->stage_ms[7] += iterations;
  }
  if (!motion_only && iterations > 0) {
    CK(launch_guard_nonsrc(g, in, st), "guard");
    if (dz_out != nullptr && g.M > 0)
      CK(cudaMemcpyAsync(dz_out, ws.dz, size_t(g.M) * g.HW * sizeof(float), cudaMemcpyDeviceToDevice, st), "dz_out");
  }
  ctx->last_launches = g_launch_count;
  return 0;
}

int check_common(dba_ctx* ctx, const void* poses, const void* disps, const void* intr, const void* targets,
                 const void* weights, const void* eta, const void* ii, const void* jj, int64_t E, int64_t N, int ht,
                 int wd, int64_t M, int iterations, int motion_only) {
  if (!ctx) return -1;
  if (!poses || !disps || !intr) return fail_arg(ctx, "poses / disps / intrinsics must not be NULL");
  if (E < 0 || N <= 0 || ht <= 0 || wd <= 0 || iterations < 0) return fail_arg(ctx, "negative or zero size");
  if (int64_t(ht) * wd > (int64_t(1) << 24)) return fail_arg(ctx, "ht*wd too large");
  if (E > 0 && (!targets || !weights || !ii || !jj)) return fail_arg(ctx, "targets / weights / ii / jj must not be NULL");
  if (!motion_only && E > 0 && (!eta || M <= 0)) return fail_arg(ctx, "eta must not be NULL unless motion_only");
  return 0;
}

}  // namespace

// accessors for entry points defined beside their kernels in another translation unit (dba_glue.cu:
// dba_damping_gather_noalloc); dba_ctx itself is private to this file
namespace dba {
int ctx_device(dba_ctx* c) { return c->device; }
int ctx_fail_arg(dba_ctx* c, const char* m) { return fail_arg(c, m); }
int ctx_fail_cuda(dba_ctx* c, cudaError_t e, const char* where) { return fail_cuda(c, e, where); }
void ctx_set_launches(dba_ctx* c, int64_t n) { c->last_launches = n; }
}  // namespace dba

extern "C" {

const char* dba_version(void) { return "libdba 0.1 (sm_100a)"; }

dba_ctx* dba_create(int device) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) return nullptr;
  if (cudaSetDevice(device) != cudaSuccess) return nullptr;
  dba_ctx* c = new dba_ctx();
  c->device = device;
  int cap = 0;
  cudaDeviceGetAttribute(&cap, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
  c->smem_cap = size_t(cap);
  int lo_pri = 0, hi_pri = 0;
  cudaDeviceGetStreamPriorityRange(&lo_pri, &hi_pri);
  if (cudaStreamCreateWithPriority(&c->side, cudaStreamNonBlocking, hi_pri) != cudaSuccess) c->side = nullptr;
  if (cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming) != cudaSuccess) {
    if (c->side) cudaStreamDestroy(c->side);
    c->side = nullptr;
  }
  if (c->side != nullptr &&
      (cudaStreamCreateWithPriority(&c->side2, cudaStreamNonBlocking, hi_pri) != cudaSuccess ||
       cudaEventCreateWithFlags(&c->ev_join2, cudaEventDisableTiming) != cudaSuccess)) {
    if (c->side2) cudaStreamDestroy(c->side2);
    c->side2 = nullptr;
  }
  return c;
}

void dba_destroy(dba_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
  DevBuf* bufs[] = {&c->d_plan_i32, &c->d_plan_i32b, &c->d_items, &c->cov_slot.i32, &c->cov_slot.i32b, &c->cov_slot.items, &c->Hpart, &c->Spart, &c->vEpart, &c->Q, &c->w, &c->Hblk, &c->vblk,
                    &c->A, &c->Lm, &c->Ldiag, &c->dx, &c->dz, &c->fail, &c->Akeep, &c->solve_msg, &c->solve_flag, &c->d_chi2, &c->s_poses, &c->s_disps, &c->s_intr,
                    &c->s_sens, &c->s_targets, &c->s_weights, &c->s_eta, &c->s_ii, &c->s_jj, &c->s_dx, &c->s_dz, &c->d_pt, &c->d_jd, &c->d_cov, &c->d_slot, &c->d_pack, &c->pg_i32, &c->pg_items, &c->pg_H, &c->pg_v, &c->pg_Ep,
                    &c->pg_CW, &c->pg_Q, &c->pg_W, &c->pg_Eis, &c->pg_S, &c->pg_vE, &c->pg_Hsp, &c->pg_vsp, &c->pg_A, &c->pg_Lm,
                    &c->pg_Ldiag, &c->pg_dx, &c->pg_dz, &c->pg_fail, &c->pg_keep, &c->pg_smsg, &c->pg_sflag};
  for (DevBuf* b : bufs) b->release();
  for (cudaEvent_t e : c->ev) cudaEventDestroy(e);
  for (auto& ge : c->graphs) if (ge.exec) cudaGraphExecDestroy(ge.exec);
  if (c->gstream) cudaStreamDestroy(c->gstream);
  if (c->side) cudaStreamDestroy(c->side);
  if (c->side2) cudaStreamDestroy(c->side2);
  if (c->ev_join2) cudaEventDestroy(c->ev_join2);
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  if (c->ev_join) cudaEventDestroy(c->ev_join);
  if (c->h_idx) cudaFreeHost(c->h_idx);
  for (char* hp : c->h_plan) if (hp) cudaFreeHost(hp);
  if (c->ev_plan) cudaEventDestroy(c->ev_plan);
  delete c;
}

const char* dba_last_error(dba_ctx* c) { return c ? c->err.c_str() : "null context"; }

int dba_set_option(dba_ctx* ctx, const char* name, double value) {
  if (!ctx || !name) return -1;
  std::string n(name);
  ++ctx->graph_epoch;                // any option may change what a call launches: cached call graphs are stale
  if (n == "cuda_graphs") ctx->opt_graphs = value != 0;
  else if (n == "block_diag_quirk") ctx->block_diag_quirk = value != 0;
  else if (n == "keep_system") ctx->keep_system = value != 0;
  else if (n == "profile") { ctx->profile = value != 0; for (double& v : ctx->stage_ms) v = 0; }
  else if (n == "multilaunch_cholesky") ctx->opt_multilaunch_chol = value != 0;
  else if (n == "tensor_cores") { ctx->opt_tc = value != 0; ctx->plan_valid = false; }
  else if (n == "linearize_v2") { ctx->opt_v2 = value != 0; ctx->plan_valid = false; }
  else if (n == "wide_tensor_cores") { ctx->opt_wide_tc = int(value); ctx->plan_valid = false; }
  else if (n == "small_system") ctx->opt_small = value != 0;
  else if (n == "refine") ctx->opt_refine = int(value);
  else if (n == "p2p_allreduce") ctx->opt_p2p = value != 0;
  else if (n == "p2p_max_n") ctx->opt_p2p_max_n = int(value);
  else if (n == "p2p_two_shot") ctx->opt_p2p_two_shot = int(value);
  else if (n == "nvtx") ctx->opt_nvtx = value != 0;
  else if (n == "chi2") ctx->opt_chi2 = value != 0;
  else if (n == "p2p_debug") ctx->opt_p2p_debug = int(value);
  else if (n == "cholesky_tc") ctx->opt_chol_tc = int(value);
  else if (n == "stream_inputs") g_tc_stream_mode = int(value);
  else if (n == "cluster_solve") g_cluster_solve_mode = int(value);
  else if (n == "splits") { ctx->opt_splits = int(value); ctx->plan_valid = false; }
  else if (n == "ppl") {
    int v = int(value);
    if (!(v == 0 || v == 1 || v == 2 || v == 4 || v == 8)) return fail_arg(ctx, "ppl must be 0,1,2,4,8");
    ctx->opt_ppl = v; ctx->plan_valid = false;
  } else return fail_arg(ctx, "unknown option " + n);
  ctx->cov_slot.valid = false;
  return 0;
}

// device edge lists -> pinned host staging (one synchronisation of `st`)
static int fetch_edge_lists(dba_ctx* ctx, const int64_t* ii, const int64_t* jj, int64_t E, cudaStream_t st) {
  if (size_t(2 * E) > ctx->h_idx_cap) {
    if (ctx->h_idx) cudaFreeHost(ctx->h_idx);
    ctx->h_idx = nullptr; ctx->h_idx_cap = 0;
    CK(cudaMallocHost(&ctx->h_idx, size_t(2 * E + 64) * sizeof(int64_t)), "pinned staging");
    ctx->h_idx_cap = size_t(2 * E + 64);
  }
  if (E > 0) {
    CK(cudaMemcpyAsync(ctx->h_idx, ii, E * sizeof(int64_t), cudaMemcpyDeviceToHost, st), "ii -> host");
    CK(cudaMemcpyAsync(ctx->h_idx + E, jj, E * sizeof(int64_t), cudaMemcpyDeviceToHost, st), "jj -> host");
    CK(cudaStreamSynchronize(st), "ii/jj -> host");
  }
  return 0;
}

int dba_ba_cached(dba_ctx* ctx, float* poses, float* disps, const float* intrinsics, const float* disps_sens,
                  const float* targets, const float* weights, const float* eta, const int64_t* ii, const int64_t* jj,
                  int64_t E, int64_t N, int ht, int wd, int64_t M, int t0, int t1, int iterations, float lm, float ep,
                  int motion_only, float* dx_out, float* dz_out, void* stream, uint64_t graph_key,
                  const int64_t* ii_host, const int64_t* jj_host) {
  int rc = check_common(ctx, poses, disps, intrinsics, targets, weights, eta, ii, jj, E, N, ht, wd, M, iterations,
                        motion_only);
  if (rc) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  const int64_t* ih = nullptr;
  const int64_t* jh = nullptr;
  if (E > 0 && !plan_hit_by_key(ctx, graph_key, E, int(N), ht, wd, motion_only ? -1 : M, t0, t1, motion_only != 0)) {
    if (ii_host != nullptr && jj_host != nullptr) { ih = ii_host; jh = jj_host; }
    else {
      rc = fetch_edge_lists(ctx, ii, jj, E, st);      // the only host synchronisation of the call
      if (rc) return rc;
      ih = ctx->h_idx; jh = ctx->h_idx + E;
    }
  }
  ctx->cur_ii_dev = ii; ctx->cur_jj_dev = jj;
  rc = run_ba(ctx, poses, disps, intrinsics, disps_sens, targets, weights, eta, ih, jh, E, N, ht, wd, M, t0, t1,
              iterations, lm, ep, motion_only != 0, dx_out, dz_out, st, nullptr, graph_key);
  ctx->cur_ii_dev = nullptr; ctx->cur_jj_dev = nullptr;
  return rc;
}

int dba_ba(dba_ctx* ctx, float* poses, float* disps, const float* intrinsics, const float* disps_sens,
           const float* targets, const float* weights, const float* eta, const int64_t* ii, const int64_t* jj,
           int64_t E, int64_t N, int ht, int wd, int64_t M, int t0, int t1, int iterations, float lm, float ep,
           int motion_only, float* dx_out, float* dz_out, void* stream) {
  return dba_ba_cached(ctx, poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, E, N, ht, wd, M, t0,
                       t1, iterations, lm, ep, motion_only, dx_out, dz_out, stream, 0, nullptr, nullptr);
}

int dba_pgba(dba_ctx* ctx, float* poses, float* disps, const float* intrinsics, const float* targets, const float* weights,
             const float* eta, const int64_t* ii, const int64_t* jj, const int64_t* rel_ii, const int64_t* rel_jj,
             const float* rel_poses, const float* rel_infos, int64_t E, int64_t R, int64_t N, int ht, int wd, int64_t M,
             int t0, int t1, int iterations, float lm, float ep, float pw, float* dx_out, float* dz_out, void* stream) {
  int rc = check_common(ctx, poses, disps, intrinsics, targets, weights, eta, ii, jj, E, N, ht, wd, M, iterations, 0);
  if (rc) return rc;
  if (E <= 0) return fail_arg(ctx, "pgba needs reprojection edges");
  if (R < 0 || (R > 0 && (!rel_ii || !rel_jj || !rel_poses || !rel_infos))) return fail_arg(ctx, "relative-pose factor arrays");
  if (t0 < 0 || t1 <= t0 || t1 > N) return fail_arg(ctx, "need 0 <= t0 < t1 <= N");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  rc = fetch_edge_lists(ctx, ii, jj, E, st);
  if (rc) return rc;
  const int64_t* ih = ctx->h_idx; const int64_t* jh = ctx->h_idx + E;
  for (int64_t e = 0; e < E; ++e)
    if (jh[e] < 0 || jh[e] >= t1 || ih[e] >= t1) return fail_arg(ctx, "edge index outside [0, t1)");
  Plan& P = ctx->pg_plan;
  const bool same = P.E == E && P.N == N && P.t0 == t0 && P.t1 == t1 && int64_t(P.key_ii.size()) == E &&
                    std::memcmp(P.key_ii.data(), ih, E * sizeof(int64_t)) == 0 &&
                    std::memcmp(P.key_jj.data(), jh, E * sizeof(int64_t)) == 0 && ctx->pg_i32.p != nullptr;
  const size_t HW = size_t(ht) * wd, f = sizeof(float);
  auto r4 = [](size_t n) { return (n + 3) & ~size_t(3); };
  if (!same) {
    P = Plan();
    if (!build_plan(P, ih, jh, E, int(N), t0, t1, false, M, 1)) return fail_arg(ctx, P.error);
    std::vector<int32_t> pack;
    auto put = [&](const std::vector<int32_t>& v) { pack.insert(pack.end(), v.begin(), v.end()); while (pack.size() % 4) pack.push_back(0); };
    put(P.kx); put(P.koff); put(P.cedge); put(P.ci); put(P.cj); put(P.pairbase); put(P.blkbase); put(P.sblk_off); put(P.rhs_off);
    CK(ctx->pg_i32.ensure(pack.size() * 4 + 16), "pgba plan");
    CK(cudaMemcpyAsync(ctx->pg_i32.p, pack.data(), pack.size() * 4, cudaMemcpyHostToDevice, st), "pgba plan");
    std::vector<Item> items(P.sblk_items);
    items.insert(items.end(), P.rhs_items.begin(), P.rhs_items.end());
    CK(ctx->pg_items.ensure(items.size() * sizeof(Item) + 16), "pgba plan");
    if (!items.empty()) CK(cudaMemcpyAsync(ctx->pg_items.p, items.data(), items.size() * sizeof(Item), cudaMemcpyHostToDevice, st), "pgba plan");
    CK(cudaStreamSynchronize(st), "pgba plan");
  } else if (M != P.M) return fail_arg(ctx, "eta rows != |unique(ii)|");
  const int Pp = P.Pp, n7 = 7 * Pp, ld = (n7 + 1 + 3) & ~3;
  CK(ctx->pg_H.ensure(size_t(E) * 4 * 49 * f), "pgba workspace");  CK(ctx->pg_v.ensure(size_t(E) * 14 * f), "pgba workspace");
  CK(ctx->pg_Ep.ensure(size_t(E) * 14 * HW * f), "pgba workspace"); CK(ctx->pg_CW.ensure(size_t(E) * 2 * HW * f), "pgba workspace");
  CK(ctx->pg_Q.ensure(size_t(P.M) * HW * f), "pgba workspace");     CK(ctx->pg_W.ensure(size_t(P.M) * HW * f), "pgba workspace");
  CK(ctx->pg_Eis.ensure(size_t(P.M) * 7 * HW * f), "pgba workspace"); CK(ctx->pg_dz.ensure(size_t(P.M) * HW * f), "pgba workspace");
  CK(ctx->pg_S.ensure(size_t(P.pairbase[P.M]) * 49 * f + 16), "pgba workspace");
  CK(ctx->pg_vE.ensure(size_t(P.blkbase[P.M]) * 7 * f + 16), "pgba workspace");
  CK(ctx->pg_Hsp.ensure(size_t(R) * 4 * 49 * f + 16), "pgba workspace"); CK(ctx->pg_vsp.ensure(size_t(R) * 14 * f + 16), "pgba workspace");
  CK(ctx->pg_A.ensure(size_t(n7 + 1) * ld * f, true), "pgba workspace"); CK(ctx->pg_Lm.ensure(size_t(n7 + 1) * ld * f), "pgba workspace");
  CK(ctx->pg_keep.ensure(size_t(n7 + 1) * ld * f), "pgba workspace");
  CK(ctx->pg_Ldiag.ensure(cholesky_ldiag_floats(n7) * f + 16), "pgba workspace");
  CK(ctx->pg_dx.ensure(size_t(n7) * f + 16), "pgba workspace"); CK(ctx->pg_fail.ensure(16, true), "pgba workspace");
  const int32_t* b = ctx->pg_i32.as<int32_t>();
  PgbaArgs a{};
  a.poses = poses; a.disps = disps; a.intr = intrinsics; a.targets = targets; a.weights = weights; a.eta = eta;
  a.kx = b; b += r4(P.kx.size()); a.koff = b; b += r4(P.koff.size()); a.cedge = b; b += r4(P.cedge.size());
  a.ci = b; b += r4(P.ci.size()); a.cj = b; b += r4(P.cj.size()); a.pairbase = b; b += r4(P.pairbase.size());
  a.blkbase = b; b += r4(P.blkbase.size()); a.sblk_off = b; b += r4(P.sblk_off.size()); a.rhs_off = b;
  a.sblk_items = ctx->pg_items.as<Item>(); a.rhs_items = a.sblk_items + P.sblk_items.size();
  a.rel_ii = rel_ii; a.rel_jj = rel_jj; a.rel_poses = rel_poses; a.rel_infos = rel_infos; a.pw = pw; a.lm = lm; a.ep = ep;
  a.E = int(E); a.N = int(N); a.M = P.M; a.R = int(R); a.HW = int(HW); a.wd = wd; a.t0 = t0; a.Pp = Pp; a.nSblk = P.nSblk;
  a.npairs_max = P.npairs_max; a.ldA = ld;
  a.nblkmax = P.nblkmax;
  a.Hblk = ctx->pg_H.as<float>(); a.vblk = ctx->pg_v.as<float>(); a.Ep = ctx->pg_Ep.as<float>(); a.CW = ctx->pg_CW.as<float>();
  a.Q = ctx->pg_Q.as<float>(); a.Wv = ctx->pg_W.as<float>(); a.Eis = ctx->pg_Eis.as<float>(); a.Spart = ctx->pg_S.as<float>();
  a.vEpart = ctx->pg_vE.as<float>(); a.Hsp = ctx->pg_Hsp.as<float>(); a.vsp = ctx->pg_vsp.as<float>(); a.A = ctx->pg_A.as<float>();
  a.dx = ctx->pg_dx.as<float>(); a.dz = ctx->pg_dz.as<float>();
  Workspace ws{};
  ws.A = a.A; ws.ldA = ld; ws.Lm = ctx->pg_Lm.as<float>(); ws.Ldiag = ctx->pg_Ldiag.as<float>(); ws.dx = a.dx;
  ws.fail = ctx->pg_fail.as<int>();
  if (trisolve_grid_msg_floats(n7) > 0) {            // n > 768: message slots of the grid-chain triangular solves
    CK(ctx->pg_smsg.ensure(trisolve_grid_msg_floats(n7) * f), "pgba workspace");
    CK(ctx->pg_sflag.ensure(trisolve_grid_flag_words(n7) * sizeof(unsigned), true), "pgba workspace");
    ws.solve_msg = ctx->pg_smsg.as<float>(); ws.solve_flag = ctx->pg_sflag.as<unsigned>();
  }
  const int coop = cholesky_coop_max_ctas(n7);
  g_launch_count = 0;
  for (int it = 0; it < iterations; ++it) {
    CK(launch_pgba_linearize(a, st), "pgba linearize");
    if (n7 <= cholesky_small_max_n()) {
      CK(launch_cholesky_small(ws, n7, st), "cholesky (single CTA)");
    } else if (coop > 0) {
      const bool refine = ctx->opt_refine != 0 && (ctx->opt_refine >= 2 || it >= iterations - 2);   // as in run_ba
      if (refine) CK(cudaMemcpyAsync(ctx->pg_keep.p, a.A, size_t(n7 + 1) * ld * f, cudaMemcpyDeviceToDevice, st), "pgba keep");
      CK(launch_cholesky_coop(ws, n7, coop, st, !refine), "cholesky (cooperative)");
      if (refine) CK(launch_refine(ws, ctx->pg_keep.as<float>(), n7, st), "refinement");
    } else {
      return fail_arg(ctx, "pgba: reduced system too large for the cooperative factorisation");
    }
    CK(launch_pgba_update(a, (it == iterations - 1) ? dx_out : nullptr, st), "pgba update");
  }
  if (iterations > 0) {
    CK(launch_pgba_guard(disps, int64_t(t1) * int64_t(HW), st), "guard");
    if (dz_out) CK(cudaMemcpyAsync(dz_out, a.dz, size_t(P.M) * HW * f, cudaMemcpyDeviceToDevice, st), "dz_out");
  }
  ctx->last_launches = g_launch_count;
  return 0;
}

int dba_depth_cov(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics, const float* disps_sens,
                  const float* targets, const float* weights, const float* eta, const int64_t* ii, const int64_t* jj,
                  int64_t E, int64_t N, int ht, int wd, int64_t M, int t0, int t1, float lm, float ep, float* dzcov_out,
                  void* stream) {
  int rc = check_common(ctx, poses, disps, intrinsics, targets, weights, eta, ii, jj, E, N, ht, wd, M, 1, 0);
  if (rc) return rc;
  if (E <= 0 || !dzcov_out) return fail_arg(ctx, "depth covariance needs edges and an output buffer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  rc = fetch_edge_lists(ctx, ii, jj, E, st);
  if (rc) return rc;
  // The covariance linearises on the plain fp32 kernels.  The TMEM accumulation of the tcgen05 Schur products
  // truncates: eight adds per flush shrink every entry of E Q E^T by ~2e-7, COHERENTLY -- harmless for dx (within the
  // reference's own fp32 noise), but u^T S^-1 u amplifies a coherent relative error of E Q E^T ~700x on ill-conditioned
  // graphs where random rounding of the same size is amplified ~100x (tools/cov_diag.py, profiles/r2_parity_margins.md:
  // 1.4e-4 .. 2.3e-4 per pixel against 3e-5 .. 4e-5 for the reference's fp32 run; 7e-5 on the fp32 kernels).
  // Its plan has its own slot, so the ba plan of the same graph survives.
  // the state is read-only on this path (run_ba's cov branch leaves before any retraction)
  const bool tc = ctx->opt_tc;
  swap_plan_slot(ctx);
  ctx->opt_tc = false;
  rc = run_ba(ctx, const_cast<float*>(poses), const_cast<float*>(disps), intrinsics, disps_sens, targets, weights, eta,
              ctx->h_idx, ctx->h_idx + E, E, N, ht, wd, M, t0, t1, 1, lm, ep, false, nullptr, nullptr, st, dzcov_out);
  ctx->opt_tc = tc;
  swap_plan_slot(ctx);
  return rc;
}

int dba_ba_host(dba_ctx* ctx, float* poses, float* disps, const float* intrinsics, const float* disps_sens,
                const float* targets, const float* weights, const float* eta, const int64_t* ii, const int64_t* jj,
                int64_t E, int64_t N, int ht, int wd, int64_t M, int t0, int t1, int iterations, float lm, float ep,
                int motion_only, float* dx_out, float* dz_out, void* stream) {
  int rc = check_common(ctx, poses, disps, intrinsics, targets, weights, eta, ii, jj, E, N, ht, wd, M, iterations,
                        motion_only);
  if (rc) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  if (t1 - t0 <= 0 || t0 < 0 || t1 > N) return fail_arg(ctx, "need 0 <= t0 < t1 <= N");
  const size_t HW = size_t(ht) * wd, f = sizeof(float);
  const size_t Mrows = motion_only ? 0 : size_t(M);
  CK(ctx->s_poses.ensure(N * 7 * f), "staging"); CK(ctx->s_disps.ensure(N * HW * f), "staging");
  CK(ctx->s_intr.ensure(4 * f), "staging");
  CK(ctx->s_targets.ensure(size_t(E) * 2 * HW * f + 16), "staging");
  CK(ctx->s_weights.ensure(size_t(E) * 2 * HW * f + 16), "staging");
  CK(ctx->s_eta.ensure(Mrows * HW * f + 16), "staging");
  CK(ctx->s_dx.ensure(size_t(t1 - t0) * 6 * f), "staging"); CK(ctx->s_dz.ensure(Mrows * HW * f + 16), "staging");
  if (disps_sens) CK(ctx->s_sens.ensure(N * HW * f), "staging");
  auto h2d = [&](DevBuf& b, const void* src, size_t bytes) {
    return bytes ? cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, st) : cudaSuccess;
  };
  CK(h2d(ctx->s_poses, poses, N * 7 * f), "h2d"); CK(h2d(ctx->s_disps, disps, N * HW * f), "h2d");
  CK(h2d(ctx->s_intr, intrinsics, 4 * f), "h2d");
  CK(h2d(ctx->s_targets, targets, size_t(E) * 2 * HW * f), "h2d");
  CK(h2d(ctx->s_weights, weights, size_t(E) * 2 * HW * f), "h2d");
  if (Mrows) CK(h2d(ctx->s_eta, eta, Mrows * HW * f), "h2d");
  if (disps_sens) CK(h2d(ctx->s_sens, disps_sens, N * HW * f), "h2d");
  rc = run_ba(ctx, ctx->s_poses.as<float>(), ctx->s_disps.as<float>(), ctx->s_intr.as<float>(),
              disps_sens ? ctx->s_sens.as<float>() : nullptr, ctx->s_targets.as<float>(), ctx->s_weights.as<float>(),
              ctx->s_eta.as<float>(), ii, jj, E, N, ht, wd, M, t0, t1, iterations, lm, ep, motion_only != 0,
              dx_out ? ctx->s_dx.as<float>() : nullptr, (dz_out && Mrows) ? ctx->s_dz.as<float>() : nullptr, st);
  if (rc) return rc;
  CK(cudaMemcpyAsync(poses, ctx->s_poses.p, N * 7 * f, cudaMemcpyDeviceToHost, st), "d2h");
  CK(cudaMemcpyAsync(disps, ctx->s_disps.p, N * HW * f, cudaMemcpyDeviceToHost, st), "d2h");
  if (dx_out) CK(cudaMemcpyAsync(dx_out, ctx->s_dx.p, size_t(t1 - t0) * 6 * f, cudaMemcpyDeviceToHost, st), "d2h");
  if (dz_out && Mrows) CK(cudaMemcpyAsync(dz_out, ctx->s_dz.p, Mrows * HW * f, cudaMemcpyDeviceToHost, st), "d2h");
  CK(cudaStreamSynchronize(st), "d2h");
  return 0;
}

int dba_frame_distance(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics,
                       const int64_t* ii, const int64_t* jj, int64_t K, int64_t N, int ht, int wd, float beta,
                       float* out, void* stream) {
  if (!ctx) return -1;
  if (K < 0 || N <= 0 || ht <= 0 || wd <= 0) return fail_arg(ctx, "negative or zero size");
  if (K == 0) return 0;
  if (!poses || !disps || !intrinsics || !ii || !jj || !out) return fail_arg(ctx, "NULL argument");
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  g_launch_count = 0;
  CK(launch_frame_distance(poses, disps, intrinsics, ii, jj, K, int(N), ht, wd, beta, out,
                           reinterpret_cast<cudaStream_t>(stream)), "frame_distance");
  ctx->last_launches = g_launch_count;
  return 0;
}

int dba_covis_distance(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics,
                       const int64_t* ii, int64_t K, int64_t N, int ht, int wd, float* out, void* stream) {
  if (!ctx) return -1;
  if (K < 0 || N <= 1 || ht <= 0 || wd <= 0) return fail_arg(ctx, "negative or zero size (needs at least two frames)");
  if (K == 0) return 0;
  if (!poses || !disps || !intrinsics || !ii || !out) return fail_arg(ctx, "NULL argument");
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  g_launch_count = 0;
  CK(launch_covis_distance(poses, disps, intrinsics, ii, K, int(N), ht, wd, out, reinterpret_cast<cudaStream_t>(stream)),
     "covis_distance");
  ctx->last_launches = g_launch_count;
  return 0;
}

int dba_iproj(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics, int64_t K, int ht,
              int wd, float* out, void* stream) {
  if (!ctx) return -1;
  if (K < 0 || ht <= 0 || wd <= 0) return fail_arg(ctx, "negative or zero size");
  if (K == 0) return 0;
  if (!poses || !disps || !intrinsics || !out) return fail_arg(ctx, "NULL argument");
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  g_launch_count = 0;
  CK(launch_iproj(poses, disps, intrinsics, K, ht, wd, out, reinterpret_cast<cudaStream_t>(stream)), "iproj");
  ctx->last_launches = g_launch_count;
  return 0;
}

int dba_depth_filter(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics,
                     const int64_t* ix, const float* thresh, int64_t K, int64_t N, int ht, int wd, float* out,
                     void* stream) {
  if (!ctx) return -1;
  if (K < 0 || N <= 0 || ht <= 0 || wd <= 0) return fail_arg(ctx, "negative or zero size");
  if (K == 0) return 0;
  if (!poses || !disps || !intrinsics || !ix || !thresh || !out) return fail_arg(ctx, "NULL argument");
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  g_launch_count = 0;
  CK(launch_depth_filter(poses, disps, intrinsics, ix, thresh, K, int(N), ht, wd, out,
                         reinterpret_cast<cudaStream_t>(stream)), "depth_filter");
  ctx->last_launches = g_launch_count;
  return 0;
}

// the per-source-keyframe edge lists of proj_trans (cached by (ii, jj)) + its launch; *kx_out = device kx [M]
static int proj_trans_run(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics,
                          const float* targets, const float* weights, const int64_t* ii, const int64_t* jj, int64_t E,
                          int64_t N, int ht, int wd, int64_t M, float* C_out, float* w_out, cudaStream_t st,
                          const int32_t** kx_out) {
  { const int rcf = fetch_edge_lists(ctx, ii, jj, E, st); if (rcf) return rcf; }
  const int64_t* hi = ctx->h_idx; const int64_t* hj = ctx->h_idx + E;
  const bool same = int64_t(ctx->pt_ii.size()) == E && ctx->pt_N == int(N) &&
                    std::memcmp(ctx->pt_ii.data(), hi, E * sizeof(int64_t)) == 0 &&
                    std::memcmp(ctx->pt_jj.data(), hj, E * sizeof(int64_t)) == 0;
  if (!same) {   // kx = sorted unique(ii) (ba.py:203), edges of a keyframe in ascending edge id
    std::vector<int32_t> deg(N, 0);
    for (int64_t e = 0; e < E; ++e) {
      if (hi[e] < 0 || hi[e] >= N || hj[e] < 0 || hj[e] >= N) return fail_arg(ctx, "edge index outside the frame buffer");
      deg[hi[e]]++;
    }
    std::vector<int32_t> slot(N, -1), kx, koff(1, 0), cedge(E), cj(E);
    for (int f = 0; f < int(N); ++f)
      if (deg[f] > 0) { slot[f] = int32_t(kx.size()); kx.push_back(f); koff.push_back(koff.back() + deg[f]); }
    std::vector<int32_t> cur(koff.begin(), koff.end() - 1);
    for (int64_t e = 0; e < E; ++e) { const int c = cur[slot[hi[e]]]++; cedge[c] = int32_t(e); cj[c] = int32_t(hj[e]); }
    std::vector<int32_t> pack;
    auto put = [&](const std::vector<int32_t>& v) { pack.insert(pack.end(), v.begin(), v.end()); while (pack.size() % 4) pack.push_back(0); };
    put(kx); put(koff); put(cedge); put(cj);
    CK(ctx->d_pt.ensure(pack.size() * 4 + 16), "proj_trans plan");
    CK(cudaMemcpyAsync(ctx->d_pt.p, pack.data(), pack.size() * 4, cudaMemcpyHostToDevice, st), "proj_trans plan");
    CK(cudaStreamSynchronize(st), "proj_trans plan");
    ctx->pt_ii.assign(hi, hi + E); ctx->pt_jj.assign(hj, hj + E);
    ctx->pt_M = int(kx.size()); ctx->pt_N = int(N);
    ctx->pt_contig = kx.back() - kx.front() + 1 == int32_t(kx.size());
  }
  if (M != ctx->pt_M) return fail_arg(ctx, "C / w must hold one map per unique source keyframe (M != |unique(ii)|)");
  auto r4 = [](size_t n) { return (n + 3) & ~size_t(3); };
  const int32_t* base = ctx->d_pt.as<int32_t>();
  const int32_t* kx = base; const int32_t* koff = kx + r4(size_t(M)); const int32_t* cedge = koff + r4(size_t(M) + 1);
  const int32_t* cj = cedge + r4(size_t(E));
  CK(launch_proj_trans(poses, disps, intrinsics, targets, weights, kx, koff, cedge, cj, int(M), ht * wd, wd, C_out, w_out, st),
     "proj_trans");
  if (kx_out) *kx_out = kx;
  return 0;
}

int dba_proj_trans(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics, const float* targets,
                   const float* weights, const int64_t* ii, const int64_t* jj, int64_t E, int64_t N, int ht, int wd,
                   int64_t M, float* C_out, float* w_out, void* stream) {
  if (!ctx) return -1;
  if (E < 0 || N <= 0 || ht <= 0 || wd <= 0 || M < 0) return fail_arg(ctx, "negative or zero size");
  if (E == 0) return M == 0 ? 0 : fail_arg(ctx, "M must be |unique(ii)|");
  if (!poses || !disps || !intrinsics || !targets || !weights || !ii || !jj || !C_out || !w_out)
    return fail_arg(ctx, "NULL argument");
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  g_launch_count = 0;
  const int rc = proj_trans_run(ctx, poses, disps, intrinsics, targets, weights, ii, jj, E, N, ht, wd, M, C_out, w_out,
                                reinterpret_cast<cudaStream_t>(stream), nullptr);
  ctx->last_launches = g_launch_count;
  return rc;
}

int dba_jdsa(dba_ctx* ctx, const float* poses, float* disps, const float* intrinsics, const float* disps_prior,
             float* dscales, const float* targets, const float* weights, const float* eta, const int64_t* ii,
             const int64_t* jj, int64_t E, int64_t N, int ht, int wd, int hs, int ws, int64_t M, float alpha, float ep,
             float lm, float* dzcov_out, void* stream) {
  if (!ctx) return -1;
  if (E < 0 || N <= 0 || ht <= 0 || wd <= 0 || hs <= 0 || ws <= 0 || M < 0) return fail_arg(ctx, "negative or zero size");
  if (hs * ws > jdsa_max_nodes()) return fail_arg(ctx, "scale grid larger than 64 nodes per keyframe");
  if (jdsa_dynamic_smem(ht, wd) > ctx->smem_cap) return fail_arg(ctx, "ht*wd too large for the JDSA kernel's shared-memory maps");
  if (E == 0) return M == 0 ? 0 : fail_arg(ctx, "M must be |unique(ii)|");
  if (!poses || !disps || !intrinsics || !disps_prior || !dscales || !targets || !weights || !eta || !ii || !jj || !dzcov_out)
    return fail_arg(ctx, "NULL argument");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  const size_t HW = size_t(ht) * wd;
  const size_t nD = size_t(N) * size_t(hs * ws), nZ = size_t(N) * HW;
  CK(ctx->d_jd.ensure((2 * size_t(M) * HW + nZ + nD) * sizeof(float) + 16), "JDSA scratch");
  float* Cm = ctx->d_jd.as<float>();
  float* wm = Cm + size_t(M) * HW;
  g_launch_count = 0;
  const int32_t* kx = nullptr;
  const int rc = proj_trans_run(ctx, poses, disps, intrinsics, targets, weights, ii, jj, E, N, ht, wd, M, Cm, wm, st, &kx);
  if (rc != 0) return rc;
  const float* disps_in = disps;
  const float* scales_in = dscales;
  if (!ctx->pt_contig) {
    // a hole in unique(ii): slots read the prior terms of OTHER keyframes (see k_jdsa), which other CTAs update --
    // give every CTA the values from before the step, as the reference's out-of-place tensors do
    float* snap = wm + size_t(M) * HW;
    CK(cudaMemcpyAsync(snap, disps, nZ * sizeof(float), cudaMemcpyDeviceToDevice, st), "JDSA snapshot");
    CK(cudaMemcpyAsync(snap + nZ, dscales, nD * sizeof(float), cudaMemcpyDeviceToDevice, st), "JDSA snapshot");
    disps_in = snap; scales_in = snap + nZ;
  }
  CK(launch_jdsa(disps, Cm, wm, eta, disps_prior, dscales, disps_in, scales_in, kx, int(M), int(N), ht, wd, hs, ws, alpha,
                 ep, lm, dzcov_out, nullptr, nullptr, st), "jdsa");
  ctx->last_launches = g_launch_count;
  return 0;
}

int dba_bi_inter(dba_ctx* ctx, const float* scales, const float* grid, int64_t M, int ht, int wd, int hs, int ws,
                 float* vals_out, float* J_out, void* stream) {
  if (!ctx) return -1;
  if (M < 0 || ht <= 0 || wd <= 0 || hs <= 0 || ws <= 0) return fail_arg(ctx, "negative or zero size");
  if (M == 0) return 0;
  if (!scales || !grid || !vals_out || !J_out) return fail_arg(ctx, "NULL argument");
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  g_launch_count = 0;
  CK(launch_bi_inter(scales, grid, int(M), ht * wd, hs, ws, vals_out, J_out, reinterpret_cast<cudaStream_t>(stream)), "bi_inter");
  ctx->last_launches = g_launch_count;
  return 0;
}

int dba_rel_pose_cov(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics, const float* targets,
                     const float* weights, const int64_t* ii, const int64_t* jj, int64_t E, int64_t N, int ht, int wd,
                     float* rel_out, float* cov_out, float* valid_out, void* stream) {
  if (!ctx) return -1;
  if (E < 0 || N <= 0 || ht <= 0 || wd <= 0) return fail_arg(ctx, "negative or zero size");
  if (E == 0) return 0;
  if (!poses || !disps || !intrinsics || !targets || !weights || !ii || !jj || !rel_out || !cov_out || !valid_out)
    return fail_arg(ctx, "NULL argument");
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  g_launch_count = 0;
  CK(launch_rel_pose(poses, disps, intrinsics, targets, weights, ii, jj, int(E), int(N), ht * wd, wd, rel_out, cov_out, valid_out,
                     reinterpret_cast<cudaStream_t>(stream)), "rel_pose_cov");
  ctx->last_launches = g_launch_count;
  return 0;
}

int dba_reproject(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics, const int64_t* ii,
                  const int64_t* jj, int64_t E, int64_t N, int ht, int wd, float* coords_out, float* valid_out,
                  void* stream) {
  if (!ctx) return -1;
  if (E < 0 || N <= 0 || ht <= 0 || wd <= 0) return fail_arg(ctx, "negative or zero size");
  if (E > 65535) return fail_arg(ctx, "more than 65535 edges in one reproject call");
  if (E == 0) return 0;
  if (!poses || !disps || !intrinsics || !ii || !jj || !coords_out || !valid_out) return fail_arg(ctx, "NULL argument");
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  g_launch_count = 0;
  CK(launch_reproject(poses, disps, intrinsics, ii, jj, int(E), int(N), ht * wd, wd, coords_out, valid_out,
                      reinterpret_cast<cudaStream_t>(stream)), "reproject");
  ctx->last_launches = g_launch_count;
  return 0;
}

int dba_motion_features(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics, const int64_t* ii,
                        const int64_t* jj, const float* target, int64_t E, int64_t N, int ht, int wd, float* coords_out,
                        float* valid_out, float* motn_out, void* stream) {
  if (!ctx) return -1;
  if (E < 0 || N <= 0 || ht <= 0 || wd <= 0) return fail_arg(ctx, "negative or zero size");
  if (E > 65535) return fail_arg(ctx, "more than 65535 edges in one motion_features call");
  if (E == 0) return 0;
  if (!poses || !disps || !intrinsics || !ii || !jj || !target || !coords_out || !valid_out || !motn_out)
    return fail_arg(ctx, "NULL argument");
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  g_launch_count = 0;
  CK(launch_motion_features(poses, disps, intrinsics, ii, jj, target, int(E), int(N), ht * wd, wd, coords_out, valid_out,
                            motn_out, reinterpret_cast<cudaStream_t>(stream)), "motion_features");
  ctx->last_launches = g_launch_count;
  return 0;
}

int dba_damping_gather(dba_ctx* ctx, const float* damping, const int64_t* ii, int64_t E, int64_t N, int ht, int wd,
                       int64_t M, float* eta_out, void* stream) {
  if (!ctx) return -1;
  if (E < 0 || N <= 0 || ht <= 0 || wd <= 0 || M < 0) return fail_arg(ctx, "negative or zero size");
  if (N > 12000) return fail_arg(ctx, "frame buffer larger than 12000 frames");
  if (E == 0 || M == 0) return 0;
  if (!damping || !ii || !eta_out) return fail_arg(ctx, "NULL argument");
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  CK(ctx->d_slot.ensure(size_t(N) * sizeof(int32_t) + 16), "slot map");
  g_launch_count = 0;
  CK(launch_damping_gather(damping, ii, int(E), int(N), ht * wd, int(M), ctx->d_slot.as<int32_t>(), eta_out,
                           reinterpret_cast<cudaStream_t>(stream)), "damping_gather");
  ctx->last_launches = g_launch_count;
  return 0;
}

int dba_cvx_upsample(dba_ctx* ctx, const float* data, const float* mask, int64_t B, int ht, int wd, int dim,
                     float* up_out, void* stream) {
  if (!ctx) return -1;
  if (B < 0 || ht <= 0 || wd <= 0 || dim <= 0) return fail_arg(ctx, "negative or zero size");
  if (B > 65535 || ht > 65535) return fail_arg(ctx, "batch or height above 65535");
  if (B == 0) return 0;
  if (!data || !mask || !up_out) return fail_arg(ctx, "NULL argument");
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  g_launch_count = 0;
  CK(launch_cvx_upsample(data, mask, int(B), ht, wd, dim, up_out, reinterpret_cast<cudaStream_t>(stream)), "cvx_upsample");
  ctx->last_launches = g_launch_count;
  return 0;
}

int dba_nccl_unique_id(void* out128) {
  std::string err;
  if (!out128 || !load_nccl(err)) return -1;
  int r = g_nccl.GetUniqueId(out128);
  return r == 0 ? 0 : 1000 + r;
}

int dba_comm_init(dba_ctx* ctx, int nranks, int rank, const void* id128) {
  if (!ctx) return -1;
  if (nranks < 1 || rank < 0 || rank >= nranks || !id128) return fail_arg(ctx, "bad nranks / rank / id");
  if (!load_nccl(ctx->err)) return -1;
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  if (ctx->comm) { g_nccl.CommDestroy(ctx->comm); ctx->comm = nullptr; }
  Id128 id;
  std::memcpy(id.b, id128, 128);
  int r = g_nccl.CommInitRank(&ctx->comm, nranks, id, rank);
  if (r != 0) {
    ctx->err = std::string("ncclCommInitRank: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
    return 1000 + r;
  }
  ctx->nranks = nranks; ctx->rank = rank;
  return 0;
}

int dba_comm_p2p_export(dba_ctx* ctx, void* handle64_out, int64_t slab_bytes) {
  if (!ctx || !handle64_out || slab_bytes <= 0) return -1;
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  if (ctx->p2p_buf) { cudaFree(ctx->p2p_buf); ctx->p2p_buf = nullptr; ctx->p2p_ok = false; }
  const size_t slab_floats = (size_t(slab_bytes) / 4 + 3) & ~size_t(3);
  void* p = nullptr;
  // two slabs (alternating by epoch), the reduced rows of the two-shot path, two flag arrays
  CK(cudaMalloc(&p, (3 * slab_floats + 64) * sizeof(float)), "p2p buffer");
  CK(cudaMemset(p, 0, (3 * slab_floats + 64) * sizeof(float)), "p2p buffer");
  CK(cudaDeviceSynchronize(), "p2p buffer");
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) { cudaFree(p); return fail_cuda(ctx, e, "cudaIpcGetMemHandle"); }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  std::memcpy(handle64_out, &h, 64);
  ctx->p2p_buf = static_cast<float*>(p);
  ctx->p2p_slab_floats = slab_floats;
  ctx->p2p_epoch = 0;
  return 0;
}

int dba_comm_p2p_import(dba_ctx* ctx, const void* handles, int nranks) {
  if (!ctx || !handles) return -1;
  if (nranks != ctx->nranks || nranks > 8 || !ctx->p2p_buf) return fail_arg(ctx, "p2p import: call dba_comm_init and dba_comm_p2p_export first (<= 8 ranks)");
  CK(cudaSetDevice(ctx->device), "cudaSetDevice");
  for (int r = 0; r < nranks; ++r) {
    if (r == ctx->rank) { ctx->p2p_peers.base[r] = ctx->p2p_buf; continue; }
    cudaIpcMemHandle_t h;
    std::memcpy(&h, static_cast<const char*>(handles) + 64 * r, 64);
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { ctx->p2p_ok = false; return fail_cuda(ctx, e, "cudaIpcOpenMemHandle"); }
    ctx->p2p_peers.base[r] = static_cast<float*>(p);
  }
  ctx->p2p_ok = true;
  return 0;
}

int dba_comm_destroy(dba_ctx* ctx) {
  if (!ctx) return -1;
  if (ctx->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(ctx->comm);
  if (ctx->p2p_ok) {
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    for (int r = 0; r < ctx->nranks; ++r)
      if (r != ctx->rank && ctx->p2p_peers.base[r]) cudaIpcCloseMemHandle(ctx->p2p_peers.base[r]);
    ctx->p2p_ok = false;
  }
  ctx->comm = nullptr; ctx->nranks = 1; ctx->rank = 0;
  return 0;
}

int64_t dba_debug_read(dba_ctx* ctx, const char* name, void* host_out, int64_t cap) {
  if (!ctx || !name || !host_out) return -1;
  std::string n(name);
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  if (n == "dims") {
    if (cap < int64_t(sizeof(ctx->last_dims))) return -1;
    std::memcpy(host_out, ctx->last_dims, sizeof(ctx->last_dims));
    return sizeof(ctx->last_dims);
  }
  if (n == "chi2") {               // double[iterations + 1] of the last call run with option "chi2": the objective before
                                  // every iteration and after the last one (per-edge partial sums added here, edge order)
    if (ctx->chi2_rows <= 0 || cap < int64_t(ctx->chi2_rows * sizeof(double))) return -1;
    std::vector<double> h(size_t(ctx->chi2_rows) * ctx->chi2_E);
    if (cudaMemcpy(h.data(), ctx->d_chi2.p, h.size() * sizeof(double), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    double* out = static_cast<double*>(host_out);
    for (int r = 0; r < ctx->chi2_rows; ++r) {
      double t = 0.0;
      for (int64_t e = 0; e < ctx->chi2_E; ++e) t += h[size_t(r) * ctx->chi2_E + e];
      out[r] = t;
    }
    return int64_t(ctx->chi2_rows * sizeof(double));
  }
  if (n == "graph_replays") {
    if (cap < int64_t(sizeof(int64_t))) return -1;
    std::memcpy(host_out, &ctx->graph_replays, sizeof(int64_t));
    return sizeof(int64_t);
  }
  if (n == "stage_ms") {
    if (cap < int64_t(sizeof(ctx->stage_ms))) return -1;
    std::memcpy(host_out, ctx->stage_ms, sizeof(ctx->stage_ms));
    return sizeof(ctx->stage_ms);
  }
  if (n == "system") {
    if (!ctx->keep_system || !ctx->Akeep.p) return fail_arg(ctx, "keep_system option is off");
    const int ld = ctx->last_dims[1];
    int64_t bytes = int64_t(ctx->last_dims[0] + 1) * ld * sizeof(float);
    if (cap < bytes) return -1;
    if (cudaMemcpy(host_out, ctx->Akeep.p, bytes, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    return bytes;
  }
  return fail_arg(ctx, "unknown debug buffer " + n);
}

int64_t dba_last_launch_count(dba_ctx* ctx) { return ctx ? ctx->last_launches : -1; }

}  // extern "C"

// ---- host-only plan inspection (tests: the gather lists are checked on the CPU against the oracle's
// scatter-adds without a GPU) ---------------------------------------------------------------------
extern "C" {

void* dba_plan_build(const int64_t* ii, const int64_t* jj, int64_t E, int N, int t0, int t1, int motion_only,
                     int64_t M_expected, int splits, char* err, int errcap) {
  Plan* P = new Plan();
  if (!build_plan(*P, ii, jj, E, N, t0, t1, motion_only != 0, M_expected, splits)) {
    if (err && errcap > 0) { std::strncpy(err, P->error.c_str(), errcap - 1); err[errcap - 1] = 0; }
    delete P;
    return nullptr;
  }
  return P;
}

void dba_plan_free(void* plan) { delete reinterpret_cast<Plan*>(plan); }

// copies the named int32 array into out (cap entries); returns its length, or -1
int64_t dba_plan_get(void* plan, const char* name, int32_t* out, int64_t cap) {
  if (!plan || !name) return -1;
  const Plan& P = *reinterpret_cast<Plan*>(plan);
  std::string n(name);
  std::vector<int32_t> tmp;
  const std::vector<int32_t>* v = nullptr;
  if (n == "kx") v = &P.kx; else if (n == "koff") v = &P.koff; else if (n == "cedge") v = &P.cedge;
  else if (n == "ci") v = &P.ci; else if (n == "cj") v = &P.cj; else if (n == "pairbase") v = &P.pairbase;
  else if (n == "blkbase") v = &P.blkbase; else if (n == "sblk_off") v = &P.sblk_off; else if (n == "rhs_off") v = &P.rhs_off;
  else if (n == "sblk_items" || n == "rhs_items") {
    const std::vector<Item>& it = (n == "sblk_items") ? P.sblk_items : P.rhs_items;
    for (const Item& i : it) { tmp.push_back(i.flags); tmp.push_back(i.base); tmp.push_back(i.stride); tmp.push_back(i.pad); }
    v = &tmp;
  } else if (n == "dims") {
    tmp = {P.E, P.N, P.t0, P.t1, P.Pp, P.M, P.n6, P.S, P.degmax, P.nblkmax, P.npairs_max, P.nSblk,
           int32_t(P.slab_blocks), int32_t(P.vE_blocks)};
    v = &tmp;
  } else return -1;
  if (out && cap >= int64_t(v->size())) std::memcpy(out, v->data(), v->size() * sizeof(int32_t));
  return int64_t(v->size());
}

}  // extern "C"
