/* libdba -- B200-native dense bundle adjustment engine: C ABI.
 *
 * Drop-in boundary for the `droid_backends` torch extension that the reference builds from
 * src/droid.cpp + src/droid_kernels.cu (setup.py:10-17; sources absent from the reference tree) and
 * calls from hislam2 (recovered call sites: depth_video.pyc DepthVideo.cuda_ba@232,
 * DepthVideo.distance@195-205; live: hislam2/util/droid_visualization.py:97,100-101).
 * Plain pointers and sizes only; all *device* pointers are CUDA device memory of the context's
 * device, fp32 unless stated, contiguous.  Every call is asynchronous on `stream` (a cudaStream_t
 * passed as void*) unless stated; none of them uses the CPU for arithmetic -- there is no fallback.
 *
 * Return value: 0 ok; <0 invalid argument (message via dba_last_error); >0 a cudaError_t /
 * ncclResult_t (+1000) value.
 */
#ifndef DBA_H_
#define DBA_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dba_ctx dba_ctx;

/* Context: owns the cached plan (edge CSR + gather lists), the workspace and, for sharded global
 * BA, the NCCL communicator.  One context per device per caller thread. */
dba_ctx* dba_create(int device);
void dba_destroy(dba_ctx* ctx);
const char* dba_last_error(dba_ctx* ctx);
const char* dba_version(void);

/* Options (name, value): "block_diag_quirk" 0/1 -- motion_only damping as hislam2/geom/chol.py:35-37
 * (ep + lm*H on the diagonal of EVERY 6x6 block) instead of the true diagonal (default 0);
 * "splits" n -- CTAs per source keyframe (0 = auto); "keep_system" 0/1 -- keep a copy of the reduced
 * system of the last iteration for dba_debug_read; "ppl" n -- pixels per lane override (0 = auto);
 * "profile" 0/1 -- bracket every stage with CUDA events on `stream` and accumulate per-stage times
 * (setting it resets the accumulators; the call then waits for its last event before returning);
 * "tensor_cores" 0/1 -- Schur products on tcgen05 (default 1); "linearize_v2" 0/1 -- the one-warp-per-edge kernel for
 * out-degree <= 20 (default 1); "refine" 0/1/2 -- iterative refinement of the pose solve: off / last two iterations of a
 * call (motion-only: the last one) / every iteration (default 1); "cholesky_tc" n -- 128-column-panel factorisation with tcgen05 trailing update from system
 * size n on (default 0 = off); "p2p_allreduce" 0/1, "p2p_max_n"; "stream_inputs" -1/0/1; "wide_tensor_cores" 0/1/2 -- keyframes with
 * out-degree 21..25 (1) and 26..41 (2) on the WIDE tcgen05 variants instead of the SIMT kernel (default 2);
 * "cuda_graphs" 0/1 -- dba_ba_cached calls that repeat an earlier call exactly (plan hit by graph_key, same buffer
 * addresses, sizes, iterations, lm, ep) are captured into a CUDA graph on their second occurrence and replayed with one
 * cudaGraphLaunch on `stream` afterwards (default 1; not under "profile" / "keep_system", not for sharded contexts);
 * "chi2" 0/1 -- record the reprojection objective before every LM iteration and after the last one (one extra kernel per
 * iteration; read with dba_debug_read "chi2"; default 0);
 * "nvtx" 0/1 -- NVTX ranges around every call and its stages (host side, for Nsight Systems timelines; default 0);
 * "cluster_solve" 0/1 -- triangular solves as a panel chain through one thread-block cluster (n <= 768) or a
 * co-resident grid (larger n) instead of the single-CTA slab kernels (default 1; process-wide). */
int dba_set_option(dba_ctx* ctx, const char* name, double value);

/* ---- ba: replaces droid_backends.ba(poses, disps, intrinsics, disps_sens, targets, weights, eta,
 *      ii, jj, t0, t1, iterations, lm, ep, motion_only)  [depth_video.pyc cuda_ba@232; BASELINE.json
 *      north_star argument list].  Maths: hislam2/geom/ba.py:32-158, chol.py:33-78.
 *   poses   [N,7]   in/out  [tx,ty,tz,qx,qy,qz,qw] world->camera; frames [t0,t1) are updated
 *   disps   [N,ht,wd] in/out; rows unique(ii) are updated, guards (>10 -> 0, max(.,1e-3)) on [0,t1)
 *   intrinsics [4]  device  fx,fy,cx,cy
 *   disps_sens [N,ht,wd] or NULL (sensor-depth prior; NULL or all <= 0 disables it)
 *   targets, weights [E,2,ht,wd]
 *   eta     [M,ht,wd], M = |unique(ii)|, rows in sorted(unique(ii)) order (ignored if motion_only)
 *   ii, jj  [E] int64 device
 *   dx_out  [t1-t0,6] or NULL, dz_out [M,ht,wd] or NULL: update of the LAST iteration
 * Host synchronisation: the plan (edge CSR + gather lists) is built on the host, so dba_ba copies ii / jj to the
 * host and synchronises `stream` ONCE AT ENTRY ON EVERY CALL (it then compares them with the cached plan's lists and
 * rebuilds only when they differ).  The iteration loop itself never touches the host.  Callers that re-use one edge
 * list for several calls (FactorGraph.update: depth_video.pyc cuda_ba@232 is called once per update with the same
 * ii / jj until the graph changes) use dba_ba_cached below, which does not synchronise on a hit. */
int dba_ba(dba_ctx* ctx, float* poses, float* disps, const float* intrinsics, const float* disps_sens,
           const float* targets, const float* weights, const float* eta, const int64_t* ii, const int64_t* jj,
           int64_t E, int64_t N, int ht, int wd, int64_t M, int t0, int t1, int iterations, float lm, float ep,
           int motion_only, float* dx_out, float* dz_out, void* stream);

/* dba_ba with a caller-supplied identity of the edge lists.  graph_key != 0 and equal to the key the cached plan
 * was built under (same E, N, ht, wd, t0, t1, motion_only): the plan is re-used WITHOUT reading ii / jj -- no copy,
 * no host synchronisation, the call is fully asynchronous on `stream`.  Otherwise the plan is rebuilt from
 * ii_host / jj_host (HOST copies of the lists, may be NULL) or, when those are NULL, from the device lists after one
 * synchronising copy, and remembered under graph_key.  The caller guarantees that equal keys mean equal lists
 * (cut3r_slam_b200/droid_backends.py derives the key from the tensors' storage address + version counter and keeps
 * the tensors alive while the key is cached).  graph_key = 0 behaves like dba_ba. */
int dba_ba_cached(dba_ctx* ctx, float* poses, float* disps, const float* intrinsics, const float* disps_sens,
                  const float* targets, const float* weights, const float* eta, const int64_t* ii, const int64_t* jj,
                  int64_t E, int64_t N, int ht, int wd, int64_t M, int t0, int t1, int iterations, float lm, float ep,
                  int motion_only, float* dx_out, float* dz_out, void* stream, uint64_t graph_key,
                  const int64_t* ii_host, const int64_t* jj_host);

/* Same, but every buffer is HOST memory (pinned for full speed): inputs are copied to the device,
 * the iterations run, poses/disps (and dx_out/dz_out if non-NULL) are copied back; returns after
 * the results are on the host.  This is the end-to-end entry bench.py times. */
int dba_ba_host(dba_ctx* ctx, float* poses, float* disps, const float* intrinsics, const float* disps_sens,
                const float* targets, const float* weights, const float* eta, const int64_t* ii, const int64_t* jj,
                int64_t E, int64_t N, int ht, int wd, int64_t M, int t0, int t1, int iterations, float lm, float ep,
                int motion_only, float* dx_out, float* dz_out, void* stream);

/* ---- frame_distance(poses, disps, intrinsics, ii, jj, beta) -> [K]  (depth_video.pyc distance@195-205) */
int dba_frame_distance(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics,
                       const int64_t* ii, const int64_t* jj, int64_t K, int64_t N, int ht, int wd, float beta,
                       float* out, void* stream);

/* ---- covis_distance(poses, disps, intrinsics, ii) -> [K]  (depth_video.pyc distance_covis@217; HI-SLAM2 keyframe
 * culling).  Reference CUDA absent -> PARITY UNPINNED; definition (oracle/dba.py covis_distance): mean induced-flow
 * magnitude in pixels between keyframe ii[k] and its temporal successor over the pixels visible in both (transformed
 * depth > 0.2 and projection inside the image), 1000 when none is.  The caller divides by median(disps[ii]). */
int dba_covis_distance(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics,
                       const int64_t* ii, int64_t K, int64_t N, int ht, int wd, float* out, void* stream);

/* ---- iproj(poses_inv, disps, intrinsics) -> [K,ht,wd,3]  (hislam2/util/droid_visualization.py:97) */
int dba_iproj(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics, int64_t K, int ht,
              int wd, float* out, void* stream);

/* ---- depth_filter(poses, disps, intrinsics, ix, thresh) -> [K,ht,wd]  (droid_visualization.py:100-101) */
int dba_depth_filter(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics,
                     const int64_t* ix, const float* thresh, int64_t K, int64_t N, int ht, int wd, float* out,
                     void* stream);

/* ---- JDSA ops (hislam2/geom/ba.py:161-241, the use_mono branch of cuda_ba)
 * proj_trans(poses, disps, intrinsics, target, weight, ii, jj) -> (C, w)  [hislam2/geom/ba.py:200]:
 *   C, w [M,ht,wd], rows in sorted(unique(ii)) order: C = sum over the keyframe's edges and the two residual rows
 *   of w Jz^2, w = sum w r Jz (ba.py:62-68,83-90, without eta); targets / weights [E,2,ht,wd]. */
int dba_proj_trans(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics, const float* targets,
                   const float* weights, const int64_t* ii, const int64_t* jj, int64_t E, int64_t N, int ht, int wd,
                   int64_t M, float* C_out, float* w_out, void* stream);
/* bi_inter(scales, grid) -> (vals, J)  [hislam2/geom/ba.py:167]: scales [M,hs,ws], grid [M,ht,wd,2] = (x along ws,
 *   y along hs) in node units; vals [M,ht,wd] bilinear samples, J [M,ht,wd,hs*ws] their Jacobian w.r.t. scales[m]. */
int dba_bi_inter(dba_ctx* ctx, const float* scales, const float* grid, int64_t M, int ht, int wd, int hs, int ws,
                 float* vals_out, float* J_out, void* stream);
/* JDSA(target, weight, eta, poses, disps, intrinsics, disps_sens, dscales, ii, jj, alpha)  [hislam2/geom/ba.py:173-241],
 *   one joint depth / scale-grid Gauss-Newton step with the poses fixed, in place on disps [N,ht,wd] and dscales
 *   [N,hs,ws] (hs*ws <= 64): proj_trans, the mono-prior terms (ba.py:205-229), schur_solve_mono_prior
 *   (hislam2/geom/chol.py:80-108; ep / lm are its keyword defaults 0.1 / 1e-4 upstream), retraction and the two
 *   disparity guards (ba.py:231-239).  disps_prior [N,ht,wd] (<= 0: no prior at that pixel), eta [M,ht,wd];
 *   dzcov_out [M,ht,wd] = the depth covariance the reference returns.  M = |unique(ii)|. */
int dba_jdsa(dba_ctx* ctx, const float* poses, float* disps, const float* intrinsics, const float* disps_prior,
             float* dscales, const float* targets, const float* weights, const float* eta, const int64_t* ii,
             const int64_t* jj, int64_t E, int64_t N, int ht, int wd, int hs, int ws, int64_t M, float alpha, float ep,
             float lm, float* dzcov_out, void* stream);

/* ---- pgba: Sim3 pose-graph bundle adjustment.  Replaces droid_backends.pgba (19 arguments at
 * depth_video.pyc cuda_pgba@307-309; CUDA source absent).  `iterations` steps of: the reference's BA (hislam2/geom/ba.py:
 * 32-107) with Sim3 poses (7-dof; geom/projective_ops.py:32-38, thirdparty/lietorch/lietorch/include/sim3.h) plus the
 * relative-pose factors of hislam2/pgo_buffer.py:39-62 (residual log(Gij G_i G_j^-1), numeric Jacobians in fp64, blocks
 * scaled by pw and the 7x7 information matrices) added to H, v before the LM damping (that composition is bytecode-only
 * in the reference: parity unpinned; the two parts are pinned, tests/golden/pgba.npz).
 *   poses [N,8] in/out [tx,ty,tz,qx,qy,qz,qw,s]; disps, targets, weights, eta, ii, jj as dba_ba (indices < t1);
 *   rel_ii, rel_jj [R] int64 device, rel_poses [R,7] (SE3, scale 1), rel_infos [R,7,7]; R may be 0;
 *   dx_out [t1-t0,7] or NULL, dz_out [M,ht,wd] or NULL.  Synchronises `stream` once (plan on the host). */
int dba_pgba(dba_ctx* ctx, float* poses, float* disps, const float* intrinsics, const float* targets, const float* weights,
             const float* eta, const int64_t* ii, const int64_t* jj, const int64_t* rel_ii, const int64_t* rel_jj,
             const float* rel_poses, const float* rel_infos, int64_t E, int64_t R, int64_t N, int ht, int wd, int64_t M,
             int t0, int t1, int iterations, float lm, float ep, float pw, float* dx_out, float* dz_out, void* stream);

/* ---- depth covariance of the Schur solve: the third output of BA (hislam2/geom/ba.py:98-107, chol.py:74-76;
 * SURVEY.md 8f rank 4).  Linearises at the GIVEN state exactly as one `ba` iteration would (same arguments, same
 * damping lm / ep), factors the reduced pose system and writes dzcov [M,ht,wd] = diag(E^T S^-1 E) Q^2 + Q, rows in
 * sorted(unique(ii)) order.  poses / disps are not modified.  This call always linearises on the fp32 kernels (the
 * "tensor_cores" option does not apply: S^-1 amplifies the coherent truncation of the tcgen05 accumulation) and keeps
 * a plan of its own, so alternating with dba_ba on one graph rebuilds neither plan. */
int dba_depth_cov(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics, const float* disps_sens,
                  const float* targets, const float* weights, const float* eta, const int64_t* ii, const int64_t* jj,
                  int64_t E, int64_t N, int ht, int wd, int64_t M, int t0, int t1, float lm, float ep, float* dzcov_out,
                  void* stream);

/* ---- relative pose + diagonal covariance per loop-closure edge: PGOBuffer.add_rel_poses
 * (hislam2/pgo_buffer.py:85-123).  For every edge e (no masking here: the caller drops edges
 * with valid_out[e] <= 1e-2 as pgo_buffer.py:87-89 does): rel_out [E,7] = Exp(dx)^4 (G_jj G_ii^-1) with dx the
 * Gauss-Newton step of the 6-dof reprojection problem at the current poses, cov_out [E,6] = diag(sigma^2 H^-1),
 * valid_out [E] = mean(weight > 0.1).  targets / weights [E,2,ht,wd] (the `ba` layout); ii / jj DEVICE int64. */
int dba_rel_pose_cov(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics, const float* targets,
                     const float* weights, const int64_t* ii, const int64_t* jj, int64_t E, int64_t N, int ht, int wd,
                     float* rel_out, float* cov_out, float* valid_out, void* stream);

/* ---- the elementwise steps either side of `ba` in FactorGraph.update
 * reproject(ii, jj) -> (coords, valid)  [hislam2/depth_video.pyc reproject@170-178 = projective_transform,
 *   hislam2/geom/projective_ops.py:44-74 with jacobian=False]: coords [E,ht,wd,2] = pixel of frame jj[e] that pixel
 *   (y,x) of frame ii[e] lands on, valid [E,ht,wd,1] = 1 where the transformed depth exceeds MIN_DEPTH = 0.2.
 *   ii / jj: DEVICE int64; an edge with an index outside [0, N) gets coords = 0, valid = 0 (checked on the device:
 *   no host round trip on this path). */
int dba_reproject(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics, const int64_t* ii,
                  const int64_t* jj, int64_t E, int64_t N, int ht, int wd, float* coords_out, float* valid_out,
                  void* stream);
/* motion features of FactorGraph.update (factor_graph.pyc update@195-197), fused with the reprojection:
 *   coords1, valid = reproject(ii, jj);  motn = cat([coords1 - coords0, target - coords1], -1).permute(0,1,4,2,3)
 *   .clamp(-64, 64), coords0 = the pixel grid.  target [E,ht,wd,2] (the previous update's target); outputs coords
 *   [E,ht,wd,2], valid [E,ht,wd,1], motn [E,4,ht,wd].  Bytecode-only in the reference tree -> parity unpinned
 *   (restated in oracle/glue.py from the disassembly and upstream DROID-SLAM's two torch lines). */
int dba_motion_features(dba_ctx* ctx, const float* poses, const float* disps, const float* intrinsics, const int64_t* ii,
                        const int64_t* jj, const float* target, int64_t E, int64_t N, int ht, int wd, float* coords_out,
                        float* valid_out, float* motn_out, void* stream);
/* damping gather of FactorGraph.update (factor_graph.pyc update@227): eta [M,ht,wd] = 0.2 * damping[unique(ii)] + 1e-7,
 *   rows in sorted(unique(ii)) order -- the `eta` argument of dba_ba.  damping [N,ht,wd]; M = |unique(ii)| (the
 *   caller knows it: it sizes eta); unique / rank are computed on the device, no host round trip. */
int dba_damping_gather(dba_ctx* ctx, const float* damping, const int64_t* ii, int64_t E, int64_t N, int ht, int wd,
                       int64_t M, float* eta_out, void* stream);
/* dba_damping_gather without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 *   for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 *   that forbid it).  Same arguments, same kernels and bit-identical output to dba_damping_gather; the frame -> row
 *   map, the only scratch of this op, lives in a file-scope `static __device__` array of 12000 int32 (48 000 bytes
 *   held by every GPU that loads the library until the process exits) whose device address is looked up once per
 *   device with cudaGetSymbolAddress.  Calls no cudaMalloc / cudaFree / allocating library routine.
 *   N > 12000 is rejected (<0, message via dba_last_error).  All calls on one device share the array (entries [0, N)
 *   are rewritten in full by every call, nothing needs clearing): calls that OVERLAP on the GPU -- different streams
 *   without an ordering between them -- race; order them on one stream or with events. */
int dba_damping_gather_noalloc(dba_ctx* ctx, const float* damping, const int64_t* ii, int64_t E, int64_t N, int ht,
                               int wd, int64_t M, float* eta_out, void* stream);
/* cvx_upsample(data, mask) -> up  [hislam2/modules/droid_net.py:21-38; upsample_disp :40-44 is dim = 1]:
 *   data [B,ht,wd,dim], mask [B,576,ht,wd] (= [B,9,8,8,ht,wd] logits), up [B,8ht,8wd,dim].  `mask` is read-only
 *   here (the reference writes -inf into the border entries of the caller's tensor as a side effect). */
int dba_cvx_upsample(dba_ctx* ctx, const float* data, const float* mask, int64_t B, int ht, int wd, int dim,
                     float* up_out, void* stream);

/* ---- sharded global BA (no reference counterpart: SURVEY.md section 8e).  Edges are partitioned by
 * source keyframe ii across ranks; each rank calls dba_ba with ITS edges/targets/weights/eta and a
 * full replica of poses; the reduced system is summed with one ncclAllReduce per iteration. */
int dba_nccl_unique_id(void* out128);                                   /* rank 0, then broadcast */
int dba_comm_init(dba_ctx* ctx, int nranks, int rank, const void* id128);
int dba_comm_destroy(dba_ctx* ctx);
/* Optional, after dba_comm_init: replace the per-iteration ncclAllReduce by kernels over NVLink peer memory: ONE kernel
 * for reduced systems up to option "p2p_max_n" (default 1536), a reduce-scatter + all-gather pair above it (option
 * "p2p_two_shot": 1 = for two ranks, where it measured faster than NCCL (default), 2 = always, 0 = never), as long as (n + 1) * ld floats fit slab_bytes.  Every rank exports a buffer
 * (three regions of slab_bytes + flags; 64-byte cudaIpcMemHandle_t out), the handles are exchanged by the caller
 * (rank order) and imported.  One process per GPU, all GPUs of one box with peer access. */
int dba_comm_p2p_export(dba_ctx* ctx, void* handle64_out, int64_t slab_bytes);
int dba_comm_p2p_import(dba_ctx* ctx, const void* handles /* [nranks][64] */, int nranks);

/* ---- test / profiling hooks */
/* name: "system" -> [(n6+1)*ld] floats of the last assembled reduced system (needs keep_system=1),
 * "dims" -> int32[8] {n6, ld, Pp, M, S, ppl, degmax, npairs_max}; "chi2" -> double[iterations + 1] objective of the last call (option "chi2"); "graph_replays" -> int64 calls that ran as one
 * cudaGraphLaunch; "stage_ms" -> double[8] accumulated
 * {linearize, assemble, allreduce, cholesky+solve, backsub, retract, 0, LM iterations}; returns the number of BYTES
 * written (<= cap) or <0. Synchronous. */
int64_t dba_debug_read(dba_ctx* ctx, const char* name, void* host_out, int64_t cap);
/* number of kernels launched by the last dba_ba call (for bench.py's gpu_launches) */
int64_t dba_last_launch_count(dba_ctx* ctx);

/* ---- host-only plan inspection (no GPU needed): the edge CSR and the fixed-order gather lists that
 * replace the reference's scatter_sum calls (hislam2/geom/ba.py:13-19,73-90).  name: kx, koff, cedge,
 * ci, cj, pairbase, blkbase, sblk_off, rhs_off, sblk_items, rhs_items (4 int32 per item: flags, base,
 * stride, 0), dims. */
void* dba_plan_build(const int64_t* ii, const int64_t* jj, int64_t E, int N, int t0, int t1, int motion_only,
                     int64_t M_expected, int splits, char* err, int errcap);
void dba_plan_free(void* plan);
int64_t dba_plan_get(void* plan, const char* name, int32_t* out, int64_t cap);

/* ---- edge bookkeeping on HOST int64 arrays (bit-exact restatement of hislam2/geom/graph_utils.py:20-29 and
 * hislam2/factor_graph.py:29-39,59-117; proximity selection from factor_graph.pyc add_proximity_factors@354-424).
 * Order of appended edges is the reference's.  No GPU involved. */
/* (i,j) in [t0,t1)^2, row-major, 1 <= |i-j| <= r (neighbourhood_graph / add_neighborhood_factors); returns count */
int64_t dba_graph_neighbourhood(int64_t t0, int64_t t1, int64_t r, int64_t* ii, int64_t* jj, int64_t cap);
/* keep[k] = new edge k is not in the current graph (__filter_repeated_edges); returns number kept */
int64_t dba_graph_filter_repeated(const int64_t* cur_ii, const int64_t* cur_jj, int64_t n_cur, const int64_t* new_ii,
                                  const int64_t* new_jj, int64_t n_new, uint8_t* keep);
/* add_factors' eviction mask when max_factors would be exceeded (factor_graph.py:75-77); returns number removed */
int64_t dba_graph_evict_mask(const int64_t* age, int64_t n, int64_t max_factors, int64_t n_new, uint8_t* remove);
/* rm_keyframe on the edge lists (factor_graph.py:100-106), in place; age may be NULL; returns the new count */
int64_t dba_graph_rm_keyframe(int64_t* ii, int64_t* jj, int64_t* age, int64_t n, int64_t ix);
/* add_proximity_factors' edge selection; d [(t-t0)*(t-t1)] from frame_distance is modified in place */
int64_t dba_graph_proximity(float* d, int64_t t0, int64_t t1, int64_t t, int64_t rad, int64_t nms, float thresh,
                            int64_t max_factors, const int64_t* ex_ii, const int64_t* ex_jj, int64_t n_ex,
                            int64_t* out_ii, int64_t* out_jj, int64_t cap);

#ifdef __cplusplus
}
#endif
#endif /* DBA_H_ */
