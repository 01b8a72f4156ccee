"""Host-side mirror of the reference's `droid_backends` torch extension for the DBA path.

Same function names, argument order and error behaviour as the module the reference builds in
setup.py:10-30 and calls at depth_video.pyc DepthVideo.cuda_ba@232 (`ba`), distance@195-205
(`frame_distance`) and hislam2/util/droid_visualization.py:97,100-101 (`iproj`, `depth_filter`):
CUDA + contiguous tensors or RuntimeError (the CHECK_INPUT convention of
thirdparty/lietorch/lietorch/extras/extras.cpp:6-8), poses/disps updated in place, work queued on
the current CUDA stream.  Everything below is argument checking and pointer passing into libdba.so
(include/dba.h); there is no PyTorch or CPU implementation of any op here.

Use as a drop-in:  `from cut3r_slam_b200 import droid_backends`  (or put it in sys.modules under
the name `droid_backends`).
"""
import ctypes as C

import torch

from . import _lib

_ctxs = {}


class _Ctx:
    def __init__(self, device_index):
        self.lib = _lib.load()
        self.handle = self.lib.dba_create(device_index)
        if not self.handle:
            raise RuntimeError(f"dba_create({device_index}) failed: no such CUDA device")
        self.device_index = device_index

    def check(self, rc):
        if rc != 0:
            msg = self.lib.dba_last_error(self.handle)
            raise RuntimeError(f"libdba error {rc}: {msg.decode() if msg else ''}")

    def set_option(self, name, value):
        self.check(self.lib.dba_set_option(self.handle, name.encode(), float(value)))

    def __del__(self):
        try:
            if self.handle:
                self.lib.dba_destroy(self.handle)
                self.handle = None
        except Exception:  # interpreter shutdown
            pass


def context(device=None):
    """The per-device libdba context used by the module-level ops."""
    if device is None:
        idx = torch.cuda.current_device()
    else:
        dev = torch.device(device)
        idx = dev.index if dev.index is not None else torch.cuda.current_device()
    if idx not in _ctxs:
        _ctxs[idx] = _Ctx(idx)
    return _ctxs[idx]


def set_option(name, value, device=None):
    """block_diag_quirk / splits / ppl / keep_system (include/dba.h)."""
    context(device).set_option(name, value)


def _check(t, name, dtype, ndim=None, cuda=True):
    if not isinstance(t, torch.Tensor):
        raise RuntimeError(f"{name} must be a tensor")
    if cuda and not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if not cuda and t.is_cuda:
        raise RuntimeError(f"{name} must be a CPU tensor")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    if t.dtype != dtype:
        raise RuntimeError(f"{name} must be {dtype}, got {t.dtype}")
    if ndim is not None and t.dim() != ndim:
        raise RuntimeError(f"{name} must have {ndim} dims, got {tuple(t.shape)}")


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None and t.numel() > 0 else C.c_void_p(0)


def _ba_checks(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, motion_only, cuda):
    _check(poses, "poses", torch.float32, 2, cuda)
    _check(disps, "disps", torch.float32, 3, cuda)
    _check(intrinsics, "intrinsics", torch.float32, 1, cuda)
    _check(targets, "targets", torch.float32, 4, cuda)
    _check(weights, "weights", torch.float32, 4, cuda)
    _check(ii, "ii", torch.int64, 1, cuda)
    _check(jj, "jj", torch.int64, 1, cuda)
    N, ht, wd = disps.shape
    E = ii.numel()
    if poses.shape[0] != N or poses.shape[1] != 7:
        raise RuntimeError(f"poses must be [{N},7], got {tuple(poses.shape)}")
    if intrinsics.numel() != 4:
        raise RuntimeError("intrinsics must be [4] = fx,fy,cx,cy")
    if jj.numel() != E:
        raise RuntimeError("ii and jj must have the same length")
    for t, nm in ((targets, "targets"), (weights, "weights")):
        if tuple(t.shape) != (E, 2, ht, wd):
            raise RuntimeError(f"{nm} must be [{E},2,{ht},{wd}], got {tuple(t.shape)}")
    M = 0
    if not motion_only:
        _check(eta, "eta", torch.float32, 3, cuda)
        if tuple(eta.shape[1:]) != (ht, wd):
            raise RuntimeError(f"eta must be [M,{ht},{wd}], got {tuple(eta.shape)}")
        M = eta.shape[0]
    if disps_sens is not None and disps_sens.numel() > 0:
        _check(disps_sens, "disps_sens", torch.float32, 3, cuda)
        if tuple(disps_sens.shape) != (N, ht, wd):
            raise RuntimeError("disps_sens must have the shape of disps")
    return N, ht, wd, E, M


_graph_keys = {}      # (ii ptr, ii version, jj ptr, jj version, E, device) -> (key, ii, jj): the tensors are kept alive
_next_graph_key = [1]


def _graph_key(ii, jj):
    """Identity of the edge lists for libdba's plan cache (dba_ba_cached): FactorGraph keeps `ii` / `jj` as tensors
    that are REPLACED (torch.cat, boolean indexing: hislam2/factor_graph.py:79-90) when the graph changes and passed
    unchanged to every `ba` call in between, so (storage address, version counter) identifies the content as long as
    the tensors are alive -- the cache holds a reference to them, their storage cannot be recycled for another list.
    A hit means libdba neither copies the lists to the host nor synchronises the stream."""
    k = (ii.data_ptr(), ii._version, jj.data_ptr(), jj._version, ii.numel(), ii.device.index)
    hit = _graph_keys.get(k)
    if hit is not None and hit[1] is ii and hit[2] is jj:
        return hit[0]
    if len(_graph_keys) >= 16:                       # small LRU: drop the oldest entry
        _graph_keys.pop(next(iter(_graph_keys)))
    key = _next_graph_key[0]
    _next_graph_key[0] += 1
    _graph_keys[k] = (key, ii, jj)
    return key


def ba(*args, _ctx=None):
    """droid_backends.ba -- dense bundle adjustment, in place on poses[t0:t1] and disps[unique(ii)].

    ba(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1, iterations, lm,
       ep, motion_only)                                   (15 args, BASELINE.json north_star)
    ba(poses, disps, intrinsics, targets, weights, eta, ii, jj, t0, t1, iterations, lm, ep,
       motion_only)                                       (14 args, depth_video.pyc cuda_ba@232)
    Returns [dx, dz] of the last iteration (dx [t1-t0,6]; dz [M,ht,wd], empty when motion_only).
    Asynchronous on the current stream when the same `ii` / `jj` tensor objects were seen by the previous call
    (see _graph_key); a new edge list costs one stream synchronisation (the plan is built on the host).
    """
    if len(args) == 14:
        args = args[:3] + (None,) + args[3:]
    if len(args) != 15:
        raise TypeError(f"ba() takes 14 or 15 positional arguments ({len(args)} given)")
    (poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1, iterations, lm, ep,
     motion_only) = args
    motion_only = bool(motion_only)
    N, ht, wd, E, M = _ba_checks(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, motion_only, True)
    t0, t1, iterations = int(t0), int(t1), int(iterations)
    ctx = _ctx if _ctx is not None else context(poses.device)
    alloc = torch.empty if (iterations > 0 and E > 0) else torch.zeros          # both are fully written by the call
    dx = alloc((max(t1 - t0, 0), 6), dtype=torch.float32, device=poses.device)
    dz = alloc((0 if motion_only else M, ht, wd), dtype=torch.float32, device=poses.device)
    sens = disps_sens if (disps_sens is not None and disps_sens.numel() > 0) else None
    key = _graph_key(ii, jj) if E > 0 else 0
    with torch.cuda.device(poses.device):
        stream = torch.cuda.current_stream(poses.device).cuda_stream
        rc = ctx.lib.dba_ba_cached(ctx.handle, _ptr(poses), _ptr(disps), _ptr(intrinsics), _ptr(sens), _ptr(targets),
                                   _ptr(weights), _ptr(eta if not motion_only else None), _ptr(ii), _ptr(jj), E, N, ht,
                                   wd, M, t0, t1, iterations, float(lm), float(ep), int(motion_only), _ptr(dx), _ptr(dz),
                                   C.c_void_p(stream), key, C.c_void_p(0), C.c_void_p(0))
    ctx.check(rc)
    return [dx, dz]


def pgba(poses, disps, intrinsics, targets, weights, eta, ii, jj, rel_ii, rel_jj, rel_poses, rel_infos, t0, t1,
         iterations, lm, ep, pw=1e-5, motion_only=False):
    """droid_backends.pgba -- Sim3 pose-graph bundle adjustment (19 arguments, like the absent export called at
    depth_video.pyc cuda_pgba@307-309; its exact argument order is not recoverable, this one is documented in
    include/dba.h).  poses [N,8] Sim3 ([t, q, s]: `video.poses_sim3`) and disps are updated in place;
    rel_ii / rel_jj [R] int64, rel_poses [R,7], rel_infos [R,7,7] are the relative-pose factors of PGOBuffer
    (hislam2/pgo_buffer.py:39-62; pass empty tensors for none).  Returns [dx [t1-t0,7], dz [M,ht,wd]]."""
    if motion_only:
        raise RuntimeError("pgba: motion_only is not supported")
    _check(poses, "poses", torch.float32, 2)
    if poses.shape[1] != 8:
        raise RuntimeError(f"poses must be [N,8] Sim3 data, got {tuple(poses.shape)}")
    _check(disps, "disps", torch.float32, 3)
    _check(intrinsics, "intrinsics", torch.float32, 1)
    _check(targets, "targets", torch.float32, 4)
    _check(weights, "weights", torch.float32, 4)
    _check(eta, "eta", torch.float32, 3)
    _check(ii, "ii", torch.int64, 1)
    _check(jj, "jj", torch.int64, 1)
    N, ht, wd = disps.shape
    E, M = ii.numel(), eta.shape[0]
    if poses.shape[0] != N or jj.numel() != E:
        raise RuntimeError("poses / disps / ii / jj sizes disagree")
    for t, nm in ((targets, "targets"), (weights, "weights")):
        if tuple(t.shape) != (E, 2, ht, wd):
            raise RuntimeError(f"{nm} must be [{E},2,{ht},{wd}], got {tuple(t.shape)}")
    R = 0 if rel_ii is None else rel_ii.numel()
    if R:
        _check(rel_ii, "rel_ii", torch.int64, 1)
        _check(rel_jj, "rel_jj", torch.int64, 1)
        _check(rel_poses, "rel_poses", torch.float32, 2)
        _check(rel_infos, "rel_infos", torch.float32, 3)
        if rel_jj.numel() != R or tuple(rel_poses.shape) != (R, 7) or tuple(rel_infos.shape) != (R, 7, 7):
            raise RuntimeError("relative-pose factors: rel_jj [R], rel_poses [R,7], rel_infos [R,7,7]")
    t0, t1, iterations = int(t0), int(t1), int(iterations)
    dx = torch.zeros((max(t1 - t0, 0), 7), dtype=torch.float32, device=poses.device)
    dz = torch.zeros((M, ht, wd), dtype=torch.float32, device=poses.device)
    ctx = context(poses.device)
    with torch.cuda.device(poses.device):
        stream = torch.cuda.current_stream(poses.device).cuda_stream
        rc = ctx.lib.dba_pgba(ctx.handle, _ptr(poses), _ptr(disps), _ptr(intrinsics), _ptr(targets), _ptr(weights), _ptr(eta),
                              _ptr(ii), _ptr(jj), _ptr(rel_ii if R else None), _ptr(rel_jj if R else None),
                              _ptr(rel_poses if R else None), _ptr(rel_infos if R else None), E, R, N, ht, wd, M, t0, t1,
                              iterations, float(lm), float(ep), float(pw), _ptr(dx), _ptr(dz), C.c_void_p(stream))
    ctx.check(rc)
    return [dx, dz]


def depth_cov(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1, lm, ep):
    """The depth covariance the reference's torch BA returns as its third output (hislam2/geom/ba.py:98-107,
    chol.py:74-76), at the given state and with the argument list of `ba` (no iterations / motion_only):
    dzcov [M,ht,wd], rows in sorted(unique(ii)) order.  poses and disps are not modified."""
    N, ht, wd, E, M = _ba_checks(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, False, True)
    ctx = context(poses.device)
    cov = torch.empty((M, ht, wd), dtype=torch.float32, device=poses.device)
    sens = disps_sens if (disps_sens is not None and disps_sens.numel() > 0) else None
    with torch.cuda.device(poses.device):
        stream = torch.cuda.current_stream(poses.device).cuda_stream
        rc = ctx.lib.dba_depth_cov(ctx.handle, _ptr(poses), _ptr(disps), _ptr(intrinsics), _ptr(sens), _ptr(targets),
                                   _ptr(weights), _ptr(eta), _ptr(ii), _ptr(jj), E, N, ht, wd, M, int(t0), int(t1),
                                   float(lm), float(ep), _ptr(cov), C.c_void_p(stream))
    ctx.check(rc)
    return cov


def ba_host(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1, iterations, lm, ep,
            motion_only, device=None, _ctx=None):
    """Same op on HOST tensors (pinned for full copy speed): copies in, runs, copies poses/disps back.
    The end-to-end entry point bench.py times (dba_ba_host)."""
    motion_only = bool(motion_only)
    N, ht, wd, E, M = _ba_checks(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, motion_only, False)
    ctx = _ctx if _ctx is not None else context(device)
    dx = torch.zeros((max(int(t1) - int(t0), 0), 6), dtype=torch.float32)
    sens = disps_sens if (disps_sens is not None and disps_sens.numel() > 0) else None
    with torch.cuda.device(ctx.device_index):
        stream = torch.cuda.current_stream().cuda_stream
        rc = ctx.lib.dba_ba_host(ctx.handle, _ptr(poses), _ptr(disps), _ptr(intrinsics), _ptr(sens), _ptr(targets),
                                 _ptr(weights), _ptr(eta if not motion_only else None), _ptr(ii), _ptr(jj), E, N, ht,
                                 wd, M, int(t0), int(t1), int(iterations), float(lm), float(ep), int(motion_only),
                                 _ptr(dx), C.c_void_p(0), C.c_void_p(stream))
    ctx.check(rc)
    return [dx]


def frame_distance(poses, disps, intrinsics, ii, jj, beta):
    """droid_backends.frame_distance(poses, disps, intrinsics, ii, jj, beta) -> [K] fp32."""
    _check(poses, "poses", torch.float32, 2)
    _check(disps, "disps", torch.float32, 3)
    _check(intrinsics, "intrinsics", torch.float32, 1)
    _check(ii, "ii", torch.int64, 1)
    _check(jj, "jj", torch.int64, 1)
    if ii.numel() != jj.numel():
        raise RuntimeError("ii and jj must have the same length")
    N, ht, wd = disps.shape
    K = ii.numel()
    out = torch.empty((K,), dtype=torch.float32, device=poses.device)
    ctx = context(poses.device)
    with torch.cuda.device(poses.device):
        stream = torch.cuda.current_stream(poses.device).cuda_stream
        ctx.check(ctx.lib.dba_frame_distance(ctx.handle, _ptr(poses), _ptr(disps), _ptr(intrinsics), _ptr(ii), _ptr(jj),
                                             K, N, ht, wd, float(beta), _ptr(out), C.c_void_p(stream)))
    return out


def covis_distance(poses, disps, intrinsics, ii):
    """droid_backends.covis_distance(poses, disps, intrinsics, ii) -> [K] fp32 (depth_video.pyc distance_covis@217).
    Parity unpinned (reference CUDA absent); definition in include/dba.h and oracle/dba.py."""
    _check(poses, "poses", torch.float32, 2)
    _check(disps, "disps", torch.float32, 3)
    _check(intrinsics, "intrinsics", torch.float32, 1)
    _check(ii, "ii", torch.int64, 1)
    N, ht, wd = disps.shape
    K = ii.numel()
    out = torch.empty((K,), dtype=torch.float32, device=poses.device)
    ctx = context(poses.device)
    with torch.cuda.device(poses.device):
        stream = torch.cuda.current_stream(poses.device).cuda_stream
        ctx.check(ctx.lib.dba_covis_distance(ctx.handle, _ptr(poses), _ptr(disps), _ptr(intrinsics), _ptr(ii), K, N, ht,
                                             wd, _ptr(out), C.c_void_p(stream)))
    return out


def iproj(poses, disps, intrinsics):
    """droid_backends.iproj(poses_inv, disps, intrinsics) -> [K,ht,wd,3] points G * (X0 / disp)."""
    _check(poses, "poses", torch.float32, 2)
    _check(disps, "disps", torch.float32, 3)
    _check(intrinsics, "intrinsics", torch.float32, 1)
    K, ht, wd = disps.shape
    if poses.shape[0] != K or poses.shape[1] != 7:
        raise RuntimeError(f"poses must be [{K},7]")
    out = torch.empty((K, ht, wd, 3), dtype=torch.float32, device=poses.device)
    ctx = context(poses.device)
    with torch.cuda.device(poses.device):
        stream = torch.cuda.current_stream(poses.device).cuda_stream
        ctx.check(ctx.lib.dba_iproj(ctx.handle, _ptr(poses), _ptr(disps), _ptr(intrinsics), K, ht, wd, _ptr(out),
                                    C.c_void_p(stream)))
    return out


def depth_filter(poses, disps, intrinsics, ix, thresh):
    """droid_backends.depth_filter(poses, disps, intrinsics, ix, thresh) -> [K,ht,wd] neighbour counts."""
    _check(poses, "poses", torch.float32, 2)
    _check(disps, "disps", torch.float32, 3)
    _check(intrinsics, "intrinsics", torch.float32, 1)
    _check(ix, "ix", torch.int64, 1)
    _check(thresh, "thresh", torch.float32, 1)
    if thresh.numel() != ix.numel():
        raise RuntimeError("thresh must hold one threshold per entry of ix")
    N, ht, wd = disps.shape
    K = ix.numel()
    out = torch.empty((K, ht, wd), dtype=torch.float32, device=poses.device)
    ctx = context(poses.device)
    with torch.cuda.device(poses.device):
        stream = torch.cuda.current_stream(poses.device).cuda_stream
        ctx.check(ctx.lib.dba_depth_filter(ctx.handle, _ptr(poses), _ptr(disps), _ptr(intrinsics), _ptr(ix),
                                           _ptr(thresh), K, N, ht, wd, _ptr(out), C.c_void_p(stream)))
    return out


_sources_cache = {}


def _num_sources(ii):
    """|unique(ii)| without a sort + synchronisation on every call: remembered per (storage, version, length) of the
    edge list.  A stale hit (storage reused by another tensor) cannot pass silently: the library counts the source
    keyframes itself and rejects a wrong M, and the callers below then recount."""
    key = (ii.data_ptr(), ii._version, ii.numel(), ii.device.index)
    M = _sources_cache.get(key)
    if M is None:
        if len(_sources_cache) > 64:
            _sources_cache.clear()
        M = _sources_cache[key] = int(torch.unique(ii).numel())
    return M


def _forget_sources(ii):
    _sources_cache.pop((ii.data_ptr(), ii._version, ii.numel(), ii.device.index), None)


def proj_trans(poses, disps, intrinsics, targets, weights, ii, jj):
    """droid_backends.proj_trans(poses, disps, intrinsics, target, weight, ii, jj) -> (C, w)
    (hislam2/geom/ba.py:200).  C, w [M,ht,wd] with M = |unique(ii)|, rows in sorted(unique(ii)) order."""
    _check(poses, "poses", torch.float32, 2)
    _check(disps, "disps", torch.float32, 3)
    _check(intrinsics, "intrinsics", torch.float32, 1)
    _check(targets, "targets", torch.float32, 4)
    _check(weights, "weights", torch.float32, 4)
    _check(ii, "ii", torch.int64, 1)
    _check(jj, "jj", torch.int64, 1)
    N, ht, wd = disps.shape
    E = ii.numel()
    if jj.numel() != E:
        raise RuntimeError("ii and jj must have the same length")
    for t, nm in ((targets, "targets"), (weights, "weights")):
        if tuple(t.shape) != (E, 2, ht, wd):
            raise RuntimeError(f"{nm} must be [{E},2,{ht},{wd}], got {tuple(t.shape)}")
    ctx = context(poses.device)
    for attempt in (0, 1):
        M = _num_sources(ii)
        Cm = torch.empty((M, ht, wd), dtype=torch.float32, device=poses.device)
        wm = torch.empty((M, ht, wd), dtype=torch.float32, device=poses.device)
        with torch.cuda.device(poses.device):
            stream = torch.cuda.current_stream(poses.device).cuda_stream
            rc = ctx.lib.dba_proj_trans(ctx.handle, _ptr(poses), _ptr(disps), _ptr(intrinsics), _ptr(targets),
                                        _ptr(weights), _ptr(ii), _ptr(jj), E, N, ht, wd, M, _ptr(Cm), _ptr(wm),
                                        C.c_void_p(stream))
        if rc == 0 or attempt == 1:
            break
        _forget_sources(ii)                                # possibly a stale count: recount once
    ctx.check(rc)
    return Cm, wm


def bi_inter(scales, grid):
    """droid_backends.bi_inter(scales, grid) -> (vals, J)  (hislam2/geom/ba.py:167): scales [M,hs,ws],
    grid [M,ht,wd,2] (x along ws, y along hs, node units); vals [M,ht,wd], J [M,ht,wd,hs*ws]."""
    _check(scales, "scales", torch.float32, 3)
    _check(grid, "grid", torch.float32, 4)
    M, hs, ws = scales.shape
    if grid.shape[0] != M or grid.shape[3] != 2:
        raise RuntimeError(f"grid must be [{M},ht,wd,2], got {tuple(grid.shape)}")
    ht, wd = grid.shape[1], grid.shape[2]
    vals = torch.empty((M, ht, wd), dtype=torch.float32, device=scales.device)
    J = torch.empty((M, ht, wd, hs * ws), dtype=torch.float32, device=scales.device)
    ctx = context(scales.device)
    with torch.cuda.device(scales.device):
        stream = torch.cuda.current_stream(scales.device).cuda_stream
        ctx.check(ctx.lib.dba_bi_inter(ctx.handle, _ptr(scales), _ptr(grid), M, ht, wd, hs, ws, _ptr(vals), _ptr(J),
                                       C.c_void_p(stream)))
    return vals, J


def JDSA(target, weight, eta, poses, disps, intrinsics, disps_prior, dscales, ii, jj, alpha=0.01, ep=0.1, lm=1e-4):
    """hislam2/geom/ba.py:173-241 (`JDSA`, the use_mono branch the factor graph calls between BA iterations) as ONE
    device step: proj_trans, the mono-prior normal equations, schur_solve_mono_prior (chol.py:80-108, whose keyword
    defaults are ep / lm here), retraction and the disparity guards.  Same argument order as upstream, batch
    dimension squeezed like `droid_backends.ba`: target / weight [E,2,ht,wd], eta [M,ht,wd], poses [N,7],
    disps / disps_prior [N,ht,wd], dscales [N,hs,ws] with hs*ws <= 64.  disps and dscales are updated IN PLACE
    (upstream returns new tensors: `disps, dscales, cov = JDSA(...)` keeps working); returns (disps, dscales, dzcov)."""
    _check(poses, "poses", torch.float32, 2)
    _check(disps, "disps", torch.float32, 3)
    _check(intrinsics, "intrinsics", torch.float32, 1)
    _check(disps_prior, "disps_prior", torch.float32, 3)
    _check(dscales, "dscales", torch.float32, 3)
    _check(target, "target", torch.float32, 4)
    _check(weight, "weight", torch.float32, 4)
    _check(eta, "eta", torch.float32, 3)
    _check(ii, "ii", torch.int64, 1)
    _check(jj, "jj", torch.int64, 1)
    N, ht, wd = disps.shape
    E = ii.numel()
    hs, ws = dscales.shape[1:]
    if jj.numel() != E:
        raise RuntimeError("ii and jj must have the same length")
    if tuple(disps_prior.shape) != (N, ht, wd) or dscales.shape[0] != N:
        raise RuntimeError("disps_prior / dscales must cover the same frame buffer as disps")
    for t, nm in ((target, "target"), (weight, "weight")):
        if tuple(t.shape) != (E, 2, ht, wd):
            raise RuntimeError(f"{nm} must be [{E},2,{ht},{wd}], got {tuple(t.shape)}")
    M = eta.shape[0]                     # one map per unique source keyframe: the library counts them and rejects a mismatch
    if tuple(eta.shape[1:]) != (ht, wd):
        raise RuntimeError(f"eta must be [M,{ht},{wd}] (one map per unique source keyframe), got {tuple(eta.shape)}")
    dzcov = torch.empty((M, ht, wd), dtype=torch.float32, device=poses.device)
    ctx = context(poses.device)
    with torch.cuda.device(poses.device):
        stream = torch.cuda.current_stream(poses.device).cuda_stream
        ctx.check(ctx.lib.dba_jdsa(ctx.handle, _ptr(poses), _ptr(disps), _ptr(intrinsics), _ptr(disps_prior),
                                   _ptr(dscales), _ptr(target), _ptr(weight), _ptr(eta), _ptr(ii), _ptr(jj), E, N, ht,
                                   wd, hs, ws, M, float(alpha), float(ep), float(lm), _ptr(dzcov), C.c_void_p(stream)))
    return disps, dscales, dzcov


def debug_system(device=None):
    """(S_lower [n6,n6], rhs [n6]) of the last iteration's reduced system (needs keep_system=1)."""
    ctx = context(device)
    dims = (C.c_int32 * 8)()
    n = ctx.lib.dba_debug_read(ctx.handle, b"dims", dims, C.sizeof(dims))
    if n < 0:
        ctx.check(-1)
    n6, ld = dims[0], dims[1]
    buf = torch.empty(((n6 + 1), ld), dtype=torch.float32)
    n = ctx.lib.dba_debug_read(ctx.handle, b"system", C.c_void_p(buf.data_ptr()), buf.numel() * 4)
    if n < 0:
        ctx.check(-1)
    return buf[:n6, :n6].clone(), buf[n6, :n6].clone(), list(dims)


def last_launch_count(device=None, _ctx=None):
    ctx = _ctx if _ctx is not None else context(device)
    return int(ctx.lib.dba_last_launch_count(ctx.handle))


def graph_replays(device=None, _ctx=None):
    """How many `ba` calls of this context ran as ONE cudaGraphLaunch (a call that repeats an earlier one exactly --
    same edge-list tensors, same buffers, same scalars -- is captured on its second occurrence and replayed from then
    on; option "cuda_graphs" 0 turns that off)."""
    ctx = _ctx if _ctx is not None else context(device)
    buf = (C.c_int64 * 1)()
    if ctx.lib.dba_debug_read(ctx.handle, b"graph_replays", buf, C.sizeof(buf)) < 0:
        ctx.check(-1)
    return int(buf[0])


def chi2_history(device=None, _ctx=None):
    """With option "chi2" on: the reprojection objective sum(0.001 valid w r^2) (the weighting of hislam2/geom/ba.py:44-45)
    of the last `ba` call, before every LM iteration and after the last one -- a list of iterations + 1 floats.
    A debug / convergence monitor (one extra kernel per iteration; the call graphs are off while it is on)."""
    ctx = _ctx if _ctx is not None else context(device)
    buf = (C.c_double * 256)()
    n = ctx.lib.dba_debug_read(ctx.handle, b"chi2", buf, C.sizeof(buf))
    if n < 0:
        raise RuntimeError("no chi2 record: set_option('chi2', 1) before the ba call")
    return [buf[i] for i in range(n // 8)]


STAGES = ("linearize", "assemble", "allreduce", "cholesky", "backsub", "retract")


def stage_times(device=None, _ctx=None):
    """Per-stage device time accumulated while the "profile" option was on: ({stage: ms}, LM iterations)."""
    ctx = _ctx if _ctx is not None else context(device)
    buf = (C.c_double * 8)()
    n = ctx.lib.dba_debug_read(ctx.handle, b"stage_ms", buf, C.sizeof(buf))
    if n < 0:
        ctx.check(-1)
    return {k: buf[i] for i, k in enumerate(STAGES)}, int(buf[7])
