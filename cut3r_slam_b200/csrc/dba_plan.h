// Host-side plan for one factor graph: edges grouped by source keyframe (CSR), Schur panel / slab
// offsets and the gather lists that assemble the reduced pose system in a fixed order.
//
// Restates the index bookkeeping of the reference's BA (hislam2/geom/ba.py:70-92: kx,kk =
// unique(ii, return_inverse); ii/jj shifted by fixedp; safe_scatter_add_mat / _vec :13-19 drop
// blocks whose pose index falls outside [0, P')).  Instead of scatter-adds the plan records, for
// every lower-triangular 6x6 block of the reduced system, WHICH per-edge / per-keyframe products
// land there, so the device gathers them in a fixed order (no atomics, reproducible).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace dba {

struct Item {       // one contribution to a 6x6 block of S or a 6-vector of the rhs
  int32_t flags;    // bit0: kind (0 = per-edge H/v block, 1 = Schur slab), bit1: transpose
  int32_t base;     // kind 0: block index into Hblk (c*4+which) / vblk (c*2+which); kind 1: first slab block
  int32_t stride;   // kind 1: slab blocks between consecutive splits
  int32_t pad;
};

struct Plan {
  // problem shape
  int E = 0, N = 0, t0 = 0, t1 = 0, Pp = 0, M = 0, n6 = 0, S = 1;
  bool motion_only = false;
  int degmax = 0, nblkmax = 0, npairs_max = 0;
  int64_t slab_blocks = 0;   // total 6x6 blocks in the Schur slab (all keyframes, all splits)
  int64_t vE_blocks = 0;     // total 6-vectors in the E*Q*w slab
  int nSblk = 0;             // lower-triangular blocks of the reduced system = Pp(Pp+1)/2
  // CSR over source keyframes (slot m <-> pose index kx[m])
  std::vector<int32_t> kx, koff, cedge, ci, cj, pairbase, blkbase, slot_of_frame;
  // gather lists
  std::vector<int32_t> sblk_off, rhs_off;
  std::vector<Item> sblk_items, rhs_items;
  // identity of the graph this plan was built for (cache key)
  std::vector<int64_t> key_ii, key_jj;
  std::string error;
  // build_plan scratch (cursors / degree counts), kept between builds of the same Plan object: a SLAM loop rebuilds
  // the plan on every new edge list
  std::vector<int32_t> scratch_cur, scratch_deg;
};

// Builds the plan; returns false and sets plan.error on invalid input.
// `ii`, `jj`: host copies of the edge lists. `M_expected`: rows of eta (ignored when < 0).
bool build_plan(Plan& plan, const int64_t* ii, const int64_t* jj, int64_t E, int N, int t0, int t1,
                bool motion_only, int64_t M_expected, int splits);
// The same in two halves: what the linearisation needs (shape, CSR, slab offsets), then the gather lists that only the
// assembly reads -- libdba builds the second half while the first linearisation of a new graph already runs.
bool build_plan_csr(Plan& plan, const int64_t* ii, const int64_t* jj, int64_t E, int N, int t0, int t1,
                    bool motion_only, int64_t M_expected, int splits);
void build_plan_gather(Plan& plan);

}  // namespace dba
