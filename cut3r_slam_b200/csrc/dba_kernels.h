// Launchers of the sm_100a DBA kernels (device pointers, one stream, no host synchronisation).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

struct dba_ctx;

namespace dba {

struct Item;

// kernels launched since the last reset (per host thread); bench.py reports it as gpu_launches
extern thread_local int64_t g_launch_count;

// Device-resident view of a Plan (all int32) + workspace pointers for one `ba` call.
struct DeviceGraph {
  int E, N, t0, t1, Pp, M, n6, S;
  int degmax, nblkmax, npairs_max;
  int HW, ht, wd;
  const int32_t *kx, *koff, *cedge, *ci, *cj, *pairbase, *blkbase, *nonsrc;
  int n_nonsrc;
  // source-keyframe slots by linearisation kernel: tensor-core Schur products (out-degree <= 20) / SIMT
  const int32_t *tc_list, *simt_list, *wide_list, *wide2_list;   // wide lists: out-degree 21..25 / 26..41 (WIDE variants)
  int n_tc, n_simt, degmax_tc, n_wide, degmax_wide, n_wide2, degmax_wide2;
  const int32_t *sblk_off, *rhs_off;
  const Item *sblk_items, *rhs_items;
  int nSblk;
  // blocks of S by gather-list length: [0, n_light) have <= 8 items (8 blocks per CTA), the rest one CTA each
  const int32_t* sblk_list;
  int n_light;
};

struct BAInputs {
  float* poses;              // [N,7]   in place
  float* disps;              // [N,HW]  in place
  const float* intrinsics;   // [4] device
  const float* disps_sens;   // [N,HW] or null
  const float* targets;      // [E,2,HW]
  const float* weights;      // [E,2,HW]
  const float* eta;          // [M,HW]
};

struct Workspace {
  float* Hpart;    // [E,S,32]
  float* Spart;    // [slab_blocks,36]
  float* vEpart;   // [vE_blocks,6]
  float* Q;        // [M,HW]
  float* w;        // [M,HW]
  float* Hblk;     // [E,4,36]
  float* vblk;     // [E,2,6]
  float* A;        // [(n6+1), ldA]  reduced system (lower) + rhs row
  int ldA;
  float* Ldiag;    // per 32x32 diagonal block of the factor: L^T | 1/diag | L^-1 (cholesky_ldiag_floats)
  float* Lm;       // [(n6+1), ldA]  the factor's panels + forward-solved rhs row (cooperative path)
  float* dx;       // [Pp,6]
  float* dz;       // [M,HW]
  int* fail;       // [4]: [0] failed-factorisation flag, [1] grid-barrier counter (both self-resetting)
  float* solve_msg = nullptr;      // n > 768: [NC][NC][64] message slots of the grid-chain triangular solves
  unsigned* solve_flag = nullptr;  // ... [NC][NC] flags + the epoch counter (zeroed once)
};

// linearisation + Schur products; picks the pixel-tile width from degmax. Returns cudaError_t.
cudaError_t launch_linearize(const DeviceGraph& g, const BAInputs& in, const Workspace& ws, bool motion_only,
                             int tiles_per_split, int ppl, const int32_t* kf_list, int nkf, cudaStream_t st);
// tcgen05 variant (dba_linearize_tc.cu): keyframes of kf_list, all with out-degree <= linearize_tc_max_degree()
cudaError_t launch_linearize_tc(const DeviceGraph& g, const BAInputs& in, const Workspace& ws, int tiles_per_split,
                                const int32_t* kf_list, int nkf, int degmax, int wide, cudaStream_t st);
size_t linearize_tc_smem_bytes(int degmax, int wide = 0);
// second generation (dba_linearize_v2.cu): one warp per edge, out-degree <= linearize_v2_max_degree()
cudaError_t launch_linearize_v2(const DeviceGraph& g, const BAInputs& in, const Workspace& ws, int tiles_per_split,
                                const int32_t* kf_list, int nkf, int degmax, cudaStream_t st);
size_t linearize_v2_smem_bytes(int degmax);
int linearize_v2_max_degree();
int linearize_tc_wide_max_degree(int level);
extern int g_tc_stream_mode;
int linearize_tc_max_degree();
int linearize_tc_ppl();
size_t linearize_smem_bytes(int degmax, int ppl, bool motion_only);
int linearize_pick_ppl(int degmax, int nblkmax, bool motion_only, int HW, size_t smem_cap);

cudaError_t launch_assemble(const DeviceGraph& g, const BAInputs& in, const Workspace& ws, float lm, float ep,
                            float hscale_lm, bool motion_only, bool block_diag_quirk, cudaStream_t st);
// n6 <= cholesky_small_max_n(): factorisation + both solves in ONE single-CTA launch (frontend windows)
cudaError_t launch_cholesky_small(const Workspace& ws, int n6, cudaStream_t st);
int cholesky_small_max_n();
cudaError_t launch_cholesky_solve(const Workspace& ws, int n6, cudaStream_t st);
// single cooperative launch (one grid barrier per block step); max_ctas from cholesky_coop_max_ctas (0 = unavailable)
cudaError_t launch_cholesky_coop(const Workspace& ws, int n6, int max_ctas, cudaStream_t st, bool finalize = true);
// large systems (n >= cholesky_tc_min_n()): 128-column panels, trailing update on tcgen05 (dba_chol_tc.cuh)
cudaError_t launch_cholesky_tc(const Workspace& ws, int n6, cudaStream_t st, bool finalize = true);
int cholesky_tc_min_n();
cudaError_t launch_backsolve(const Workspace& ws, int n6, int mode, cudaStream_t st);
// one step of iterative refinement after launch_cholesky_coop(..., finalize = false): r = rhs - S dx accumulated in fp64
// from the saved system Sorig [(n6+1), ldA] (lower + rhs row), L y = r, L^T d = y, dx += d  (dba_refine.cu)
cudaError_t launch_refine(const Workspace& ws, const float* Sorig, int n6, cudaStream_t st);
int cholesky_coop_max_ctas(int n6);
size_t cholesky_ldiag_floats(int n6);   // floats in Workspace::Ldiag
cudaError_t launch_backsub(const DeviceGraph& g, const BAInputs& in, const Workspace& ws, cudaStream_t st);
cudaError_t launch_guard_nonsrc(const DeviceGraph& g, const BAInputs& in, cudaStream_t st);
cudaError_t launch_retract(const DeviceGraph& g, const BAInputs& in, const Workspace& ws, float* dx_out,
                           cudaStream_t st);

// companions
cudaError_t launch_frame_distance(const float* poses, const float* disps, const float* intr, const int64_t* ii,
                                  const int64_t* jj, int64_t K, int N, int ht, int wd, float beta, float* out,
                                  cudaStream_t st);
cudaError_t launch_chi2(const float* poses, const float* disps, const float* intr, const float* targets,
                        const float* weights, const int64_t* ii, const int64_t* jj, int64_t E, int N, int HW, int wd,
                        double* out, cudaStream_t st);
cudaError_t launch_covis_distance(const float* poses, const float* disps, const float* intr, const int64_t* ii,
                                  int64_t K, int N, int ht, int wd, float* out, cudaStream_t st);
cudaError_t launch_iproj(const float* poses, const float* disps, const float* intr, int64_t K, int ht, int wd,
                         float* out, cudaStream_t st);
cudaError_t launch_depth_filter(const float* poses, const float* disps, const float* intr, const int64_t* ix,
                                const float* thresh, int64_t K, int N, int ht, int wd, float* out,
                                cudaStream_t st);

// sharded global BA: one-shot all-reduce of the reduced system over NVLink peer memory (dba_p2p.cu)
struct P2PPeers { float* base[8]; };     // every rank's exchange buffer as mapped in THIS process (own rank: local)
cudaError_t launch_p2p_allreduce(const P2PPeers& peers, int nranks, int rank, size_t slab_floats, size_t flags_off_floats,
                                 int parity, uint32_t epoch, int ld, int n, float* out, int* err, int debug, cudaStream_t st);

cudaError_t launch_p2p_allreduce_two_shot(const P2PPeers& peers, int nranks, int rank, size_t slab_floats,
                                          size_t flags_off_floats, int parity, uint32_t epoch, int ld, int n, float* out,
                                          int* err, cudaStream_t st);

// pgba: Sim3 pose-graph BA (dba_pgba.cu); all device pointers, the plan arrays as uploaded for the 6-dof path
struct PgbaArgs {
  float* poses;              // [N,8] Sim3, in place
  float* disps;              // [N,HW] in place
  const float *intr, *targets, *weights, *eta;
  const int32_t *kx, *koff, *cedge, *ci, *cj, *pairbase, *blkbase, *sblk_off, *rhs_off;
  const Item *sblk_items, *rhs_items;
  const int64_t *rel_ii, *rel_jj;      // [R] device
  const float *rel_poses, *rel_infos;  // [R,7], [R,7,7]
  float pw, lm, ep;
  int E, N, M, R, HW, wd, t0, Pp, nSblk, npairs_max, ldA;
  int nblkmax = 0;           // largest number of panel blocks of a keyframe (out-degree + 1)
  float *Hblk, *vblk, *Ep, *CW, *Q, *Wv, *Eis, *Spart, *vEpart, *Hsp, *vsp, *A, *dx, *dz;
};
cudaError_t launch_pgba_linearize(const PgbaArgs& a, cudaStream_t st);
cudaError_t launch_pgba_update(const PgbaArgs& a, float* dx_out, cudaStream_t st);
cudaError_t launch_pgba_guard(float* disps, int64_t n, cudaStream_t st);

size_t tri_pack_floats(int n);
cudaError_t launch_tri_pack(float* A, float* buf, int ld, int n, bool pack, cudaStream_t st);

// JDSA path (dba_jdsa.cu)
cudaError_t launch_proj_trans(const float* poses, const float* disps, const float* intr, const float* targets,
                              const float* weights, const int32_t* kx, const int32_t* koff, const int32_t* cedge,
                              const int32_t* cj, int M, int HW, int wd, float* Cout, float* wout, cudaStream_t st);
cudaError_t launch_bi_inter(const float* scales, const float* grid, int M, int HW, int hs, int ws, float* vals,
                            float* J, cudaStream_t st);
cudaError_t launch_reproject(const float* poses, const float* disps, const float* intr, const int64_t* ii,
                             const int64_t* jj, int E, int N, int HW, int wd, float* coords, float* valid, cudaStream_t st);
cudaError_t launch_motion_features(const float* poses, const float* disps, const float* intr, const int64_t* ii,
                                   const int64_t* jj, const float* target, int E, int N, int HW, int wd, float* coords,
                                   float* valid, float* motn, cudaStream_t st);
cudaError_t launch_damping_gather(const float* damping, const int64_t* ii, int E, int N, int HW, int M, int32_t* slot,
                                  float* eta, cudaStream_t st);
cudaError_t launch_cvx_upsample(const float* data, const float* mask, int B, int ht, int wd, int dim, float* up,
                                cudaStream_t st);
void cholesky_ldiag_layout(int* stride, int* linv_offset);   // per 32-block: where L_kk^-1 (row-major 32x32) sits
size_t cov_scratch_floats(int n);
cudaError_t launch_depth_cov(const DeviceGraph& g, const BAInputs& in, const Workspace& ws, float* scratch,
                             size_t smem_cap, float* dzcov, cudaStream_t st);
cudaError_t launch_rel_pose(const float* poses, const float* disps, const float* intr, const float* targets,
                            const float* weights, const int64_t* ii, const int64_t* jj, int E, int N, int HW, int wd,
                            float* rel_out, float* cov_out, float* valid_out, cudaStream_t st);
int jdsa_max_nodes();
size_t jdsa_dynamic_smem(int ht, int wd);
cudaError_t launch_jdsa(float* disps, const float* Cin, const float* win, const float* eta, const float* prior,
                        float* scales, const float* disps_in, const float* scales_in, const int32_t* kx, int M, int N, int ht, int wd, int hs, int ws, float alpha,
                        float ep, float lm, float* dzcov, float* dz_out, float* dso_out, cudaStream_t st);

// triangular solves on one thread-block cluster with the factor resident in distributed shared memory
// (dba_trisolve_cluster.cu, n <= 768); cudaErrorNotSupported -> the caller uses the single-CTA kernels
extern int g_cluster_solve_mode;
cudaError_t launch_trisolve_cluster(const Workspace& ws, int n, bool forward, int mode, cudaStream_t st);
size_t trisolve_grid_msg_floats(int n);
size_t trisolve_grid_flag_words(int n);

// dba_ctx accessors (dba_api.cu) for entry points that live beside their kernels
int ctx_device(::dba_ctx* c);
int ctx_fail_arg(::dba_ctx* c, const char* m);
int ctx_fail_cuda(::dba_ctx* c, cudaError_t e, const char* where);
void ctx_set_launches(::dba_ctx* c, int64_t n);

}  // namespace dba
