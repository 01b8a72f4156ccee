"""Host-side mirror of the two elementwise steps FactorGraph.update runs either side of `ba`
(SURVEY.md section 8f rank 2), over `dba_reproject` / `dba_cvx_upsample` (include/dba.h):

  reproject(poses, disps, intrinsics, ii, jj)   DepthVideo.reproject (hislam2/depth_video.pyc @170-178), i.e.
                                                pops.projective_transform(Gs, disps[None], intrinsics[None], ii, jj)
                                                (hislam2/geom/projective_ops.py:44-74)
  cvx_upsample(data, mask), upsample_disp(disp, mask)   hislam2/modules/droid_net.py:21-44

There is no CPU or PyTorch fallback: without libdba.so and a CUDA device these raise.
"""
import ctypes as C

import torch

from .droid_backends import _check, _ptr, context


def reproject(poses, disps, intrinsics, ii, jj):
    """-> coords [1,E,ht,wd,2], valid [1,E,ht,wd,1] (the reference's shapes, batch dimension of 1).
    poses [N,7] (SE3.data), disps [N,ht,wd], intrinsics [4], ii / jj int64 CUDA tensors with values in [0, N)."""
    _check(poses, "poses", torch.float32, 2)
    _check(disps, "disps", torch.float32, 3)
    _check(intrinsics, "intrinsics", torch.float32, 1)
    _check(ii, "ii", torch.int64, 1)
    _check(jj, "jj", torch.int64, 1)
    N, ht, wd = disps.shape
    E = ii.numel()
    if jj.numel() != E:
        raise RuntimeError("ii and jj must have the same length")
    if poses.shape[0] < N or poses.shape[1] != 7:
        raise RuntimeError(f"poses must be [>={N},7], got {tuple(poses.shape)}")
    coords = torch.empty((1, E, ht, wd, 2), dtype=torch.float32, device=poses.device)
    valid = torch.empty((1, E, ht, wd, 1), dtype=torch.float32, device=poses.device)
    ctx = context(poses.device)
    with torch.cuda.device(poses.device):
        stream = torch.cuda.current_stream(poses.device).cuda_stream
        for lo in range(0, E, 32768):                      # grid.y limit of one launch
            n = min(32768, E - lo)
            ctx.check(ctx.lib.dba_reproject(ctx.handle, _ptr(poses), _ptr(disps), _ptr(intrinsics), _ptr(ii[lo:lo + n]),
                                            _ptr(jj[lo:lo + n]), n, N, ht, wd, _ptr(coords[0, lo:lo + n]),
                                            _ptr(valid[0, lo:lo + n]), C.c_void_p(stream)))
    return coords, valid


def cvx_upsample(data, mask):
    """data [B,ht,wd,dim], mask [B,576,ht,wd] (any shape with B*576*ht*wd elements in that order) ->
    [B,8ht,8wd,dim].  `mask` is not modified (upstream writes -inf into its border entries)."""
    _check(data, "data", torch.float32, 4)
    _check(mask, "mask", torch.float32)
    B, ht, wd, dim = data.shape
    if mask.numel() != B * 576 * ht * wd:
        raise RuntimeError(f"mask must hold {B}x576x{ht}x{wd} logits, got {tuple(mask.shape)}")
    up = torch.empty((B, 8 * ht, 8 * wd, dim), dtype=torch.float32, device=data.device)
    ctx = context(data.device)
    with torch.cuda.device(data.device):
        stream = torch.cuda.current_stream(data.device).cuda_stream
        ctx.check(ctx.lib.dba_cvx_upsample(ctx.handle, _ptr(data), _ptr(mask), B, ht, wd, dim, _ptr(up),
                                           C.c_void_p(stream)))
    return up


def upsample_disp(disp, mask):
    """droid_net.py:40-44: disp [B,num,ht,wd], mask [B,num,576,ht,wd] -> [B,num,8ht,8wd]."""
    _check(disp, "disp", torch.float32, 4)
    B, num, ht, wd = disp.shape
    return cvx_upsample(disp.view(B * num, ht, wd, 1), mask).view(B, num, 8 * ht, 8 * wd)


def add_rel_poses(poses, disps, intrinsics, ii, jj, target, weight):
    """PGOBuffer.add_rel_poses (hislam2/pgo_buffer.py:85-123) without the buffer bookkeeping: returns the rows the
    reference appends -- (ii, jj, rel_poses [n,7], rel_covs [n,6], valid_percent [n]) for the edges whose weight mask
    passes (mean(weight > 0.1) > 1e-2).  target / weight in the FactorGraph layout [1,E,ht,wd,2] (or [E,ht,wd,2])."""
    _check(poses, "poses", torch.float32, 2)
    _check(disps, "disps", torch.float32, 3)
    _check(intrinsics, "intrinsics", torch.float32, 1)
    _check(ii, "ii", torch.int64, 1)
    _check(jj, "jj", torch.int64, 1)
    N, ht, wd = disps.shape
    E = ii.numel()
    if target.dim() == 5:
        target, weight = target[0], weight[0]
    for t, nm in ((target, "target"), (weight, "weight")):
        if not t.is_cuda or t.dtype != torch.float32 or tuple(t.shape) != (E, ht, wd, 2):
            raise RuntimeError(f"{nm} must be a float32 CUDA tensor [{E},{ht},{wd},2], got {tuple(t.shape)}")
    tg = target.permute(0, 3, 1, 2).contiguous()           # the `ba` op layout [E,2,ht,wd] (depth_video.pyc cuda_ba@229-230)
    wg = weight.permute(0, 3, 1, 2).contiguous()
    rel = torch.empty((E, 7), dtype=torch.float32, device=poses.device)
    cov = torch.empty((E, 6), dtype=torch.float32, device=poses.device)
    val = torch.empty((E,), dtype=torch.float32, device=poses.device)
    ctx = context(poses.device)
    with torch.cuda.device(poses.device):
        stream = torch.cuda.current_stream(poses.device).cuda_stream
        ctx.check(ctx.lib.dba_rel_pose_cov(ctx.handle, _ptr(poses), _ptr(disps), _ptr(intrinsics), _ptr(tg), _ptr(wg),
                                           _ptr(ii), _ptr(jj), E, N, ht, wd, _ptr(rel), _ptr(cov), _ptr(val),
                                           C.c_void_p(stream)))
    keep = val > 1e-2                                       # pgo_buffer.py:87-89
    return ii[keep], jj[keep], rel[keep], cov[keep], val[keep]


def motion_features(poses, disps, intrinsics, ii, jj, target):
    """The first step of FactorGraph.update (factor_graph.pyc update@195-197) in one launch:
        coords1, valid = reproject(ii, jj)
        motn = cat([coords1 - coords0, target - coords1], -1).permute(0, 1, 4, 2, 3).clamp(-64, 64)
    target [1,E,ht,wd,2] (or [E,ht,wd,2]) -> coords1 [1,E,ht,wd,2], valid [1,E,ht,wd,1], motn [1,E,4,ht,wd]."""
    _check(poses, "poses", torch.float32, 2)
    _check(disps, "disps", torch.float32, 3)
    _check(intrinsics, "intrinsics", torch.float32, 1)
    _check(ii, "ii", torch.int64, 1)
    _check(jj, "jj", torch.int64, 1)
    _check(target, "target", torch.float32)
    N, ht, wd = disps.shape
    E = ii.numel()
    if jj.numel() != E or target.numel() != E * ht * wd * 2:
        raise RuntimeError(f"target must hold [{E},{ht},{wd},2] values and jj {E} indices")
    coords = torch.empty((1, E, ht, wd, 2), dtype=torch.float32, device=poses.device)
    valid = torch.empty((1, E, ht, wd, 1), dtype=torch.float32, device=poses.device)
    motn = torch.empty((1, E, 4, ht, wd), dtype=torch.float32, device=poses.device)
    tg = target.reshape(E, ht, wd, 2)
    ctx = context(poses.device)
    with torch.cuda.device(poses.device):
        stream = torch.cuda.current_stream(poses.device).cuda_stream
        for lo in range(0, E, 32768):
            n = min(32768, E - lo)
            ctx.check(ctx.lib.dba_motion_features(ctx.handle, _ptr(poses), _ptr(disps), _ptr(intrinsics),
                                                  _ptr(ii[lo:lo + n]), _ptr(jj[lo:lo + n]), _ptr(tg[lo:lo + n]), n, N,
                                                  ht, wd, _ptr(coords[0, lo:lo + n]), _ptr(valid[0, lo:lo + n]),
                                                  _ptr(motn[0, lo:lo + n]), C.c_void_p(stream)))
    return coords, valid, motn


def damping_eta(damping, ii):
    """The `eta` argument of `ba` as FactorGraph.update builds it (factor_graph.pyc update@227):
    0.2 * damping[unique(ii)] + 1e-7, [M,ht,wd] with rows in sorted(unique(ii)) order.  damping [N,ht,wd]."""
    from .droid_backends import _num_sources
    _check(damping, "damping", torch.float32, 3)
    _check(ii, "ii", torch.int64, 1)
    N, ht, wd = damping.shape
    M = _num_sources(ii) if ii.numel() else 0
    eta = torch.empty((M, ht, wd), dtype=torch.float32, device=damping.device)
    ctx = context(damping.device)
    with torch.cuda.device(damping.device):
        stream = torch.cuda.current_stream(damping.device).cuda_stream
        ctx.check(ctx.lib.dba_damping_gather(ctx.handle, _ptr(damping), _ptr(ii), ii.numel(), N, ht, wd, M, _ptr(eta),
                                             C.c_void_p(stream)))
    return eta


def damping_eta_noalloc(damping, ii, eta_out):
    """damping_eta through dba_damping_gather_noalloc (include/dba.h): no allocation anywhere on the path -- the caller
    passes the output eta_out [M,ht,wd] (M = |unique(ii)|, which the caller knows: it sized eta_out), the frame -> row
    map lives in the library's static device array (N <= 12000 frames).  No counterpart in the reference.  Calls that
    overlap on the GPU share that array and race: keep them on one stream."""
    _check(damping, "damping", torch.float32, 3)
    _check(ii, "ii", torch.int64, 1)
    _check(eta_out, "eta_out", torch.float32, 3)
    N, ht, wd = damping.shape
    M = eta_out.shape[0]
    if tuple(eta_out.shape[1:]) != (ht, wd) or eta_out.device != damping.device:
        raise ValueError("eta_out must be [M,ht,wd] on damping's device")
    ctx = context(damping.device)
    with torch.cuda.device(damping.device):
        stream = torch.cuda.current_stream(damping.device).cuda_stream
        ctx.check(ctx.lib.dba_damping_gather_noalloc(ctx.handle, _ptr(damping), _ptr(ii), ii.numel(), N, ht, wd, M,
                                                     _ptr(eta_out), C.c_void_p(stream)))
    return eta_out
