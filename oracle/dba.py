"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's dense bundle adjustment.

Follows hislam2/geom/ba.py (BA :32-107, MoBA :110-158), hislam2/geom/chol.py (block_solve :33-44,
schur_solve :47-78), hislam2/geom/projective_ops.py (:44-74) and hislam2/geom/pinhole.py (:6-53) of
the reference, operation for operation, on torch CPU tensors (fp32 by default, fp64 on request).
Batch dim B=1 of the reference is dropped.  `oracle_ba` is the adapter with the 15-argument
`droid_backends.ba` signature (argument order of BASELINE.json north_star; layout conversion of
pyc:DepthVideo.cuda_ba@229-230 undone).
"""
import torch

from . import se3

MIN_DEPTH = 0.2          # projective_ops.py:6
WEIGHT_SCALE = 0.001     # ba.py:44
C_EPS = 1e-7             # ba.py:92
SENS_ALPHA = 0.05        # upstream DROID-SLAM droid_kernels.cu (recollection; parity unpinned)


# --------------------------------------------------------------------------------------------
# pinhole.py
# --------------------------------------------------------------------------------------------
def iproj_pinhole(disps, intr):
    """pinhole.py:6-25. disps [N,ht,wd], intr [4] -> X0 [N,ht,wd,4], Jz0 [N,ht,wd,4]."""
    ht, wd = disps.shape[-2:]
    fx, fy, cx, cy = intr.unbind(-1)
    y, x = torch.meshgrid(torch.arange(ht, dtype=disps.dtype),
                          torch.arange(wd, dtype=disps.dtype), indexing="ij")
    i = torch.ones_like(disps)
    X = ((x - cx) / fx).expand_as(disps)
    Y = ((y - cy) / fy).expand_as(disps)
    pts = torch.stack([X, Y, i, disps], dim=-1)
    J = torch.zeros_like(pts)
    J[..., -1] = 1.0
    return pts, J


def proj_pinhole(Xs, intr, return_depth=False):
    """pinhole.py:27-53. Returns coords [...,2] and the 2x4 projection Jacobian."""
    fx, fy, cx, cy = intr.unbind(-1)
    X, Y, Z, D = Xs.unbind(dim=-1)
    Z = torch.where(Z < 0.1, torch.ones_like(Z), Z)      # :32
    d = 1.0 / Z
    x = fx * (X * d) + cx
    y = fy * (Y * d) + cy
    coords = torch.stack([x, y, D * d], dim=-1) if return_depth else torch.stack([x, y], dim=-1)
    o = torch.zeros_like(d)
    Jp = torch.stack([
        fx * d, o, -fx * X * d * d, o,
        o, fy * d, -fy * Y * d * d, o,
    ], dim=-1).reshape(d.shape + (2, 4))
    return coords, Jp


# --------------------------------------------------------------------------------------------
# projective_ops.py
# --------------------------------------------------------------------------------------------
def actp(Gij, X0):
    """projective_ops.py:15-42 (SE3 branch). Gij [E,7], X0 [E,ht,wd,4]."""
    X1 = se3.se3_act4(Gij[:, None, None, :], X0)
    X, Y, Z, d = X1.unbind(dim=-1)
    o = torch.zeros_like(d)
    Ja = torch.stack([
        d, o, o, o, Z, -Y,
        o, d, o, -Z, o, X,
        o, o, d, Y, -X, o,
        o, o, o, o, o, o,
    ], dim=-1).reshape(d.shape + (4, 6))
    return X1, Ja


def projective_transform(poses, disps, intr, ii, jj, jacobian=False, return_depth=False):
    """projective_ops.py:44-74. poses [P,7], disps [P,ht,wd], intr [4], ii/jj [E] (long).

    Returns coords [E,ht,wd,2], valid [E,ht,wd,1] and (Ji, Jj [E,ht,wd,2,6], Jz [E,ht,wd,2,1]).
    """
    X0, Jz0 = iproj_pinhole(disps[ii], intr)
    Gij = se3.se3_mul(poses[jj], se3.se3_inv(poses[ii]))                    # :51
    X1, Ja = actp(Gij, X0)
    x1, Jp = proj_pinhole(X1, intr, return_depth=return_depth)
    valid = ((X1[..., 2] > MIN_DEPTH) & (X0[..., 2] > MIN_DEPTH)).to(disps.dtype).unsqueeze(-1)  # :60-62
    if not jacobian:
        return x1, valid
    Jj = Jp @ Ja                                                           # :65
    Ad = se3.se3_adj_matrix(Gij)                                           # adjT, se3.h:84-86
    Ji = -(Jj @ Ad[:, None, None])                                         # row-wise Ad^T J^T  (:66-67)
    Jzp = se3.se3_act4(Gij[:, None, None, :], Jz0)                         # :69
    Jz = Jp @ Jzp.unsqueeze(-1)                                            # :70
    return x1, valid, (Ji, Jj, Jz)


# --------------------------------------------------------------------------------------------
# chol.py
# --------------------------------------------------------------------------------------------
def cholesky_solve(H, b):
    """chol.py:5-20: on a failed factorisation print and return zeros."""
    try:
        U = torch.linalg.cholesky(H)
        return torch.cholesky_solve(b, U), U, False
    except Exception as e:  # noqa: BLE001 -- mirrors the reference's blanket catch
        print(e)
        return torch.zeros_like(b), torch.zeros_like(H), True


def block_solve(H, b, ep=0.1, lm=0.0001, block_diag_quirk=True):
    """chol.py:33-44. H [N,N,D,D], b [N,D].

    block_diag_quirk=True is the reference as written: I = eye(D) broadcasts over [N,N,D,D], so
    ep + lm*H lands on the diagonal of EVERY DxD block, off-diagonal blocks included (:35-37).
    block_diag_quirk=False damps only the true diagonal of the (N D)x(N D) matrix (as schur_solve
    :56-57 does) -- the behaviour the product's `motion_only` path uses by default.
    """
    N, _, D, _ = H.shape
    if block_diag_quirk:
        I = torch.eye(D, dtype=H.dtype)
        H = H + (ep + lm * H) * I
        H = H.permute(0, 2, 1, 3).reshape(N * D, N * D)
    else:
        H = H.permute(0, 2, 1, 3).reshape(N * D, N * D)
        I = torch.eye(N * D, dtype=H.dtype)
        H = H + (ep + lm * H) * I
    x, _, failed = cholesky_solve(H, b.reshape(N * D, 1))
    return x.reshape(N, D), failed


def schur_solve(H, E, C, v, w, ep=0.1, lm=0.0001, with_dzcov=False):
    """chol.py:47-78. H [P,P,D,D], E [P,M,D,HW], C,w [M,HW], v [P,D] -> dx [P,D], dz [M,HW]."""
    P, M, D, HW = E.shape
    H = H.permute(0, 2, 1, 3).reshape(P * D, P * D)
    E = E.permute(0, 2, 1, 3).reshape(P * D, M * HW)
    Q = (1.0 / C).reshape(M * HW, 1)
    I = torch.eye(P * D, dtype=H.dtype)
    H = H + (ep + lm * H) * I                                              # :56-57
    v = v.reshape(P * D, 1)
    w = w.reshape(M * HW, 1)
    Et = E.transpose(0, 1)
    S = H - E @ (Q * Et)                                                   # :63
    vs = v - E @ (Q * w)                                                   # :64
    dx, L, failed = cholesky_solve(S, vs)                                  # :66
    dz = Q * (w - Et @ dx)                                                 # :70
    out = dict(dx=dx.reshape(P, D), dz=dz.reshape(M, HW), S=S, vs=vs.reshape(-1), failed=failed)
    if with_dzcov:                                                         # :74-76 (not a `ba` output)
        F = torch.linalg.inv(L) @ (E * Q[..., 0])
        out["dzcov"] = (torch.sum(torch.square(F), dim=0) + Q[..., 0]).reshape(M, HW)
    return out


# --------------------------------------------------------------------------------------------
# ba.py
# --------------------------------------------------------------------------------------------
def _scatter_sum(src, index, dim_size):
    """torch_scatter.scatter_sum over dim 0 (ba.py:15,19,24,29)."""
    out = torch.zeros((dim_size,) + src.shape[1:], dtype=src.dtype)
    return out.index_add_(0, index, src)


def safe_scatter_add_mat(A, ii, jj, n, m):
    """ba.py:13-15."""
    v = (ii >= 0) & (jj >= 0) & (ii < n) & (jj < m)
    return _scatter_sum(A[v], ii[v] * m + jj[v], n * m)


def safe_scatter_add_vec(b, ii, n):
    """ba.py:17-19."""
    v = (ii >= 0) & (ii < n)
    return _scatter_sum(b[v], ii[v], n)


def edge_blocks(target, weight, poses, disps, intr, ii, jj):
    """ba.py:39-68: per-edge normal-equation blocks.

    target, weight [E,ht,wd,2] (the reference's internal layout). Returns a dict with
    Hii,Hij,Hji,Hjj [E,6,6], vi,vj [E,6], Ei,Ej [E,6,HW], Ck,wk [E,HW].
    """
    E_, ht, wd = target.shape[:3]
    D = 6
    coords, valid, (Ji, Jj, Jz) = projective_transform(poses, disps, intr, ii, jj, jacobian=True)
    r = (target - coords).reshape(E_, -1, 1)
    w = WEIGHT_SCALE * (valid * weight).reshape(E_, -1, 1)
    Ji = Ji.reshape(E_, -1, D)
    Jj = Jj.reshape(E_, -1, D)
    wJiT = (w * Ji).transpose(1, 2)
    wJjT = (w * Jj).transpose(1, 2)
    Jz = Jz.reshape(E_, ht * wd, -1)
    out = dict(
        Hii=wJiT @ Ji, Hij=wJiT @ Jj, Hji=wJjT @ Ji, Hjj=wJjT @ Jj,
        vi=(wJiT @ r).squeeze(-1), vj=(wJjT @ r).squeeze(-1),
        Ei=(wJiT.reshape(E_, D, ht * wd, -1) * Jz[:, None]).sum(dim=-1),
        Ej=(wJjT.reshape(E_, D, ht * wd, -1) * Jz[:, None]).sum(dim=-1),
    )
    w2 = w.reshape(E_, ht * wd, -1)
    r2 = r.reshape(E_, ht * wd, -1)
    out["wk"] = torch.sum(w2 * r2 * Jz, dim=-1)
    out["Ck"] = torch.sum(w2 * Jz * Jz, dim=-1)
    return out


def linearize(target, weight, eta, poses, disps, intr, ii, jj, fixedp, disps_sens=None):
    """ba.py:39-95: assemble H [P',P',6,6], E [P',M,6,HW], v [P',6], C,w [M,HW]; kx = unique(ii)."""
    P, ht, wd = disps.shape
    blk = edge_blocks(target, weight, poses, disps, intr, ii, jj)
    kx, kk = torch.unique(ii, return_inverse=True)
    M = kx.shape[0]
    Pp = P - fixedp
    i0, j0 = ii - fixedp, jj - fixedp
    H = (safe_scatter_add_mat(blk["Hii"], i0, i0, Pp, Pp) + safe_scatter_add_mat(blk["Hij"], i0, j0, Pp, Pp) +
         safe_scatter_add_mat(blk["Hji"], j0, i0, Pp, Pp) + safe_scatter_add_mat(blk["Hjj"], j0, j0, Pp, Pp))
    Em = safe_scatter_add_mat(blk["Ei"], i0, kk, Pp, M) + safe_scatter_add_mat(blk["Ej"], j0, kk, Pp, M)
    v = safe_scatter_add_vec(blk["vi"], i0, Pp) + safe_scatter_add_vec(blk["vj"], j0, Pp)
    C = safe_scatter_add_vec(blk["Ck"], kk, M)
    w = safe_scatter_add_vec(blk["wk"], kk, M)
    eta = eta.reshape(C.shape)
    if disps_sens is not None:
        # Sensor-depth prior of upstream DROID-SLAM (recollection; PARITY UNPINNED -- the in-tree
        # torch BA has no such term). m = disps_sens > 0 ; C += m*alpha + (1-m)*eta ;
        # w -= m*alpha*(disps - disps_sens).  With disps_sens == 0 this is exactly ba.py:92.
        m = (disps_sens[kx] > 0).to(C.dtype).reshape(C.shape)
        C = C + m * SENS_ALPHA + (1 - m) * eta + C_EPS
        w = w - m * SENS_ALPHA * (disps[kx] - disps_sens[kx]).reshape(C.shape)
    else:
        C = C + eta + C_EPS                                                # :92
    return dict(H=H.reshape(Pp, Pp, 6, 6), E=Em.reshape(Pp, M, 6, ht * wd), v=v, C=C, w=w,
                kx=kx, kk=kk, blocks=blk)


def BA(target, weight, eta, poses, disps, intr, ii, jj, fixedp=1, ep=0.1, lm=0.0001,
       disps_sens=None, return_info=False):
    """ba.py:32-107, one Gauss-Newton/LM step. Returns (poses, disps[, info])."""
    P, ht, wd = disps.shape
    lin = linearize(target, weight, eta, poses, disps, intr, ii, jj, fixedp, disps_sens)
    sol = schur_solve(lin["H"], lin["E"], lin["C"], lin["v"], lin["w"], ep=ep, lm=lm)
    dx, dz = sol["dx"], sol["dz"]
    new_poses = poses.clone()
    new_poses[fixedp:] = se3.se3_retr(poses[fixedp:], dx)                  # :101
    new_disps = disps.clone()
    new_disps[lin["kx"]] += dz.reshape(-1, ht, wd)                         # :102
    new_disps = torch.where(new_disps > 10, torch.zeros_like(new_disps), new_disps)   # :104
    new_disps = new_disps.clamp(min=0.001)                                 # :105
    if return_info:
        return new_poses, new_disps, dict(lin=lin, sol=sol)
    return new_poses, new_disps


def MoBA(target, weight, poses, disps, intr, ii, jj, fixedp=1, ep=0.1, lm=0.0001,
         block_diag_quirk=True, return_info=False):
    """ba.py:110-158 (rig=1), motion-only step."""
    P = disps.shape[0]
    blk = edge_blocks(target, weight, poses, disps, intr, ii, jj)
    Pp = P - fixedp
    i0, j0 = ii - fixedp, jj - fixedp
    H = (safe_scatter_add_mat(blk["Hii"], i0, i0, Pp, Pp) + safe_scatter_add_mat(blk["Hij"], i0, j0, Pp, Pp) +
         safe_scatter_add_mat(blk["Hji"], j0, i0, Pp, Pp) + safe_scatter_add_mat(blk["Hjj"], j0, j0, Pp, Pp))
    v = safe_scatter_add_vec(blk["vi"], i0, Pp) + safe_scatter_add_vec(blk["vj"], j0, Pp)
    dx, failed = block_solve(H.reshape(Pp, Pp, 6, 6), v, ep=ep, lm=lm, block_diag_quirk=block_diag_quirk)
    new_poses = poses.clone()
    new_poses[fixedp:] = se3.se3_retr(poses[fixedp:], dx)
    if return_info:
        return new_poses, dict(H=H.reshape(Pp, Pp, 6, 6), v=v, dx=dx, failed=failed)
    return new_poses


def oracle_ba(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1,
              iterations, lm, ep, motion_only, dtype=torch.float32, block_diag_quirk=False,
              return_info=False):
    """`droid_backends.ba` with the reference's torch BA underneath.

    poses [N,7], disps [N,ht,wd], intrinsics [4], disps_sens [N,ht,wd] or None,
    targets/weights [E,2,ht,wd] (op layout, pyc:DepthVideo.cuda_ba@229-230), eta [M,ht,wd],
    ii/jj [E] int64. Returns NEW (poses, disps) tensors of the input dtype (inputs untouched) and,
    with return_info, the list of per-iteration (dx [t1-t0,6], dz [M,HW]).
    """
    out_dtype = poses.dtype
    p = poses[:t1].to(dtype).clone()
    d = disps[:t1].to(dtype).clone()
    intr = intrinsics.to(dtype)
    tgt = targets.permute(0, 2, 3, 1).to(dtype).contiguous()
    wgt = weights.permute(0, 2, 3, 1).to(dtype).contiguous()
    et = eta.to(dtype)
    ds = None
    if disps_sens is not None and bool((disps_sens[:t1] > 0).any()):
        ds = disps_sens[:t1].to(dtype)
    ii = ii.long().cpu()
    jj = jj.long().cpu()
    infos = []
    for _ in range(iterations):
        if motion_only:
            p, info = MoBA(tgt, wgt, p, d, intr, ii, jj, fixedp=t0, ep=ep, lm=lm,
                           block_diag_quirk=block_diag_quirk, return_info=True)
            infos.append((info["dx"], None))
        else:
            p, d, info = BA(tgt, wgt, et, p, d, intr, ii, jj, fixedp=t0, ep=ep, lm=lm,
                            disps_sens=ds, return_info=True)
            infos.append((info["sol"]["dx"], info["sol"]["dz"]))
    new_poses = poses.clone()
    new_disps = disps.clone()
    new_poses[:t1] = p.to(out_dtype)
    new_disps[:t1] = d.to(out_dtype)
    if return_info:
        return new_poses, new_disps, infos
    return new_poses, new_disps


def depth_cov(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1, lm, ep,
              dtype=torch.float64):
    """The third output of the reference's BA (ba.py:98,107 <- chol.py:74-76) at the given state, through the
    `droid_backends.ba` argument list: dzcov [M,ht*wd], rows in sorted(unique(ii)) order."""
    p = poses[:t1].to(dtype)
    d = disps[:t1].to(dtype)
    tgt = targets.permute(0, 2, 3, 1).to(dtype).contiguous()
    wgt = weights.permute(0, 2, 3, 1).to(dtype).contiguous()
    ds = None
    if disps_sens is not None and bool((disps_sens[:t1] > 0).any()):
        ds = disps_sens[:t1].to(dtype)
    lin = linearize(tgt, wgt, eta.to(dtype), p, d, intrinsics.to(dtype), ii.long().cpu(), jj.long().cpu(), t0, ds)
    sol = schur_solve(lin["H"], lin["E"], lin["C"], lin["v"], lin["w"], ep=ep, lm=lm, with_dzcov=True)
    return sol["dzcov"]


# --------------------------------------------------------------------------------------------
# companion ops -- PARITY UNPINNED (CUDA source absent; definitions documented in DESIGN.md)
# --------------------------------------------------------------------------------------------
FRAME_DISTANCE_SENTINEL = 1000.0


def frame_distance(poses, disps, intrinsics, ii, jj, beta=0.3):
    """Mean induced-flow magnitude from frame ii to frame jj (call shape pyc:DepthVideo.distance@195-205).

    For every pixel of frame ii: flow under the full relative pose Gij = Gj Gi^-1 and under its
    translation only; dist = sum(beta*v1*|f1| + (1-beta)*v2*|f2|) / sum(beta*v1 + (1-beta)*v2) with
    v = (Z1 > MIN_DEPTH); sentinel 1000 when the denominator is < 1 (callers treat d > 100 as
    infinity, pyc:add_proximity_factors@367). Built from projective_ops.py:44-62.
    """
    dt = poses.dtype
    ht, wd = disps.shape[-2:]
    intr = intrinsics.to(dt)
    fx, fy, cx, cy = intr.unbind(-1)
    X0, _ = iproj_pinhole(disps[ii], intr)
    y, x = torch.meshgrid(torch.arange(ht, dtype=dt), torch.arange(wd, dtype=dt), indexing="ij")
    Gij = se3.se3_mul(poses[jj], se3.se3_inv(poses[ii]))

    def flow(X1):
        Z = X1[..., 2]
        v = (Z > MIN_DEPTH).to(dt)
        Zs = torch.where(Z > MIN_DEPTH, Z, torch.ones_like(Z))
        du = fx * (X1[..., 0] / Zs) + cx - x
        dv = fy * (X1[..., 1] / Zs) + cy - y
        return torch.sqrt(du * du + dv * dv) * v, v

    X1 = se3.se3_act4(Gij[:, None, None, :], X0)
    f1, v1 = flow(X1)
    t = Gij[:, None, None, :3]
    X2 = torch.cat([X0[..., :3] + t * X0[..., 3:], X0[..., 3:]], dim=-1)
    f2, v2 = flow(X2)
    num = (beta * f1 + (1 - beta) * f2).reshape(len(ii), -1).sum(-1)
    den = (beta * v1 + (1 - beta) * v2).reshape(len(ii), -1).sum(-1)
    return torch.where(den < 1.0, torch.full_like(num, FRAME_DISTANCE_SENTINEL), num / den.clamp(min=1e-12))


def covis_distance(poses, disps, intrinsics, ii):
    """PARITY UNPINNED (call shape pyc:DepthVideo.distance_covis@217, which rescales the result by
    1 / median(disps[ii]) @218).  Definition chosen: for keyframe i and its temporal successor j = i + 1 (i - 1 for
    the last frame): mean induced-flow magnitude in pixels over the COVISIBLE pixels of frame i -- transformed depth
    > MIN_DEPTH and projection inside the image of frame j; sentinel 1000 when no pixel is covisible."""
    dt = poses.dtype
    N, ht, wd = disps.shape
    intr = intrinsics.to(dt)
    fx, fy, cx, cy = intr.unbind(-1)
    ii = ii.long()
    jj = torch.where(ii + 1 < N, ii + 1, ii - 1)
    X0, _ = iproj_pinhole(disps[ii], intr)
    y, x = torch.meshgrid(torch.arange(ht, dtype=dt), torch.arange(wd, dtype=dt), indexing="ij")
    Gij = se3.se3_mul(poses[jj], se3.se3_inv(poses[ii]))
    X1 = se3.se3_act4(Gij[:, None, None, :], X0)
    Z = X1[..., 2]
    Zs = torch.where(Z > MIN_DEPTH, Z, torch.ones_like(Z))
    u = fx * (X1[..., 0] / Zs) + cx
    v = fy * (X1[..., 1] / Zs) + cy
    vis = ((Z > MIN_DEPTH) & (u >= 0) & (u <= wd - 1) & (v >= 0) & (v <= ht - 1)).to(dt)
    f = torch.sqrt((u - x) ** 2 + (v - y) ** 2) * vis
    num = f.reshape(len(ii), -1).sum(-1)
    den = vis.reshape(len(ii), -1).sum(-1)
    return torch.where(den < 1.0, torch.full_like(num, FRAME_DISTANCE_SENTINEL), num / den.clamp(min=1.0))


def iproj(poses, disps, intrinsics):
    """Back-projection to 3-D points (call shape hislam2/util/droid_visualization.py:97).

    poses [K,7] are the transforms applied to the camera-frame points (the caller passes the
    INVERTED world->camera poses); points = G * ([(u-cx)/fx, (v-cy)/fy, 1] / disp) -> [K,ht,wd,3].
    In-tree CPU analogues: geometry_eval_utils.py:113-121,
    thirdparty/lietorch/examples/core/geom/projective_ops.py:11-23.
    """
    X0, _ = iproj_pinhole(disps, intrinsics.to(disps.dtype))
    pts = X0[..., :3] / X0[..., 3:]
    t, q = se3.se3_split(poses)
    return se3.quat_rotate(q[:, None, None, :], pts) + t[:, None, None, :]


DEPTH_FILTER_RADIUS = 3


def depth_filter(poses, disps, intrinsics, ix, thresh):
    """Multi-view depth consistency count (call shape hislam2/util/droid_visualization.py:100-101).

    For every selected frame i = ix[k] and each temporal neighbour j in {i-3..i-1, i+1..i+3} inside
    the buffer: reproject every pixel of i into j, bilinearly sample j's inverse depth at the
    landing point (only when all four taps are inside), and count the neighbour when
    |1/d_proj - 1/d_sampled| < thresh[k]. Returns float counts [K,ht,wd] (consumer keeps >= 2).
    """
    dt = disps.dtype
    N, ht, wd = disps.shape
    intr = intrinsics.to(dt)
    fx, fy, cx, cy = intr.unbind(-1)
    out = torch.zeros((len(ix), ht, wd), dtype=dt)
    for k, i in enumerate(ix.tolist()):
        for off in list(range(-DEPTH_FILTER_RADIUS, 0)) + list(range(1, DEPTH_FILTER_RADIUS + 1)):
            j = i + off
            if j < 0 or j >= N:
                continue
            X0, _ = iproj_pinhole(disps[i:i + 1], intr)
            Gij = se3.se3_mul(poses[j:j + 1], se3.se3_inv(poses[i:i + 1]))
            X1 = se3.se3_act4(Gij[:, None, None, :], X0)[0]
            Z = X1[..., 2]
            ok = Z > MIN_DEPTH
            Zs = torch.where(ok, Z, torch.ones_like(Z))
            uj = fx * (X1[..., 0] / Zs) + cx
            vj = fy * (X1[..., 1] / Zs) + cy
            dj = X1[..., 3] / Zs
            u0 = torch.floor(uj)
            v0 = torch.floor(vj)
            inb = ok & (u0 >= 0) & (v0 >= 0) & (u0 < wd - 1) & (v0 < ht - 1)
            u0c = u0.clamp(0, wd - 2).long()
            v0c = v0.clamp(0, ht - 2).long()
            wx = uj - u0
            wy = vj - v0
            D = disps[j]
            d00, d01 = D[v0c, u0c], D[v0c, u0c + 1]
            d10, d11 = D[v0c + 1, u0c], D[v0c + 1, u0c + 1]
            dhat = (1 - wy) * ((1 - wx) * d00 + wx * d01) + wy * ((1 - wx) * d10 + wx * d11)
            err = (1.0 / dj.clamp(min=1e-12) - 1.0 / dhat.clamp(min=1e-12)).abs()
            out[k] += (inb & (dj > 0) & (err < thresh[k])).to(dt)
    return out
