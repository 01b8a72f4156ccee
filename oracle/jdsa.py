"""CPU restatement of the JDSA path (HI-SLAM2's use_mono branch): hislam2/geom/ba.py:161-241 and
chol.py:80-108.  TEST INFRASTRUCTURE ONLY.

Pinning: `schur_solve_mono_prior` and `JDSA` below are checked against the reference's own functions run
unmodified on CPU (oracle/make_golden_jdsa.py -> tests/golden/jdsa.npz; the two absent CUDA ops the reference
calls there, proj_trans and bi_inter, are injected from THIS file).  `proj_trans` and `bi_inter` themselves are
PARITY UNPINNED at the droid_backends boundary (CUDA source absent): proj_trans is defined by the BA statements
it abbreviates (ba.py:62-68,83-90, already pinned through oracle/dba.py), bi_inter by its use at ba.py:161-170.
"""
import torch

from . import dba as od


def proj_trans(poses, disps, intrinsics, targets, weights, ii, jj):
    """(C, w) [M,HW]: depth block and depth gradient per source keyframe, no eta (ba.py:200 with :62-68,83-90).
    targets / weights in the op layout [E,2,ht,wd]."""
    tgt = targets.permute(0, 2, 3, 1).contiguous()
    wgt = weights.permute(0, 2, 3, 1).contiguous()
    blk = od.edge_blocks(tgt, wgt, poses, disps, intrinsics, ii, jj)
    kx, kk = torch.unique(ii, return_inverse=True)
    M = kx.shape[0]
    C = od.safe_scatter_add_vec(blk["Ck"], kk, M)
    w = od.safe_scatter_add_vec(blk["wk"], kk, M)
    return C, w


def bi_inter(scales, grid):
    """scales [M,hs,ws], grid [M,ht,wd,2] (x along ws, y along hs) -> vals [M,ht,wd], J [M,ht,wd,hs*ws]."""
    M, hs, ws = scales.shape
    ht, wd = grid.shape[1:3]
    gx, gy = grid[..., 0], grid[..., 1]
    x0 = torch.floor(gx).long().clamp(0, max(ws - 2, 0))
    y0 = torch.floor(gy).long().clamp(0, max(hs - 2, 0))
    x1 = (x0 + 1).clamp(max=ws - 1)
    y1 = (y0 + 1).clamp(max=hs - 1)
    wx, wy = gx - x0.to(gx.dtype), gy - y0.to(gy.dtype)
    w = [(1 - wy) * (1 - wx), (1 - wy) * wx, wy * (1 - wx), wy * wx]
    idx = [y0 * ws + x0, y0 * ws + x1, y1 * ws + x0, y1 * ws + x1]
    flat = scales.reshape(M, 1, 1, hs * ws).expand(M, ht, wd, hs * ws)
    vals = sum(wi * torch.gather(flat, 3, ix[..., None])[..., 0] for wi, ix in zip(w, idx))
    J = torch.zeros(M, ht, wd, hs * ws, dtype=scales.dtype)
    for wi, ix in zip(w, idx):
        J.scatter_add_(3, ix[..., None], wi[..., None])
    return vals, J


def get_prior_depth_aligned(depth_prior, scales):
    """ba.py:161-170."""
    M, ht, wd = depth_prior.shape
    hs, ws = scales.shape[-2:]
    meshx, meshy = torch.meshgrid(torch.linspace(0, hs - 1 - 1e-6, ht), torch.linspace(0, ws - 1 - 1e-6, wd), indexing="ij")
    grid = torch.stack((meshy, meshx), -1).to(scales.dtype)
    grid = grid.unsqueeze(0).expand(M, -1, -1, -1).contiguous()
    mscales_bi, Jbi = bi_inter(scales, grid)
    return depth_prior * mscales_bi, Jbi


def schur_solve_mono_prior(C, w, Hs, Es, vs, ep=0.1, lm=0.0001):
    """chol.py:80-108. C,w [1,M,HW]; Hs [1,M,M,D,D]; Es [1,M,M,D,HW]; vs [1,M,D,1] -> dso [1,M,D], dz [1,M,HW],
    dzcov [M,HW]."""
    D = Hs.shape[-1]
    B, M, HW = C.shape
    Q = (1.0 / C).view(B, M * HW, 1)
    w = w.reshape(B, M * HW, 1)
    H = Hs.permute(0, 1, 3, 2, 4).reshape(B, M * D, M * D)
    E = Es.permute(0, 1, 3, 2, 4).reshape(B, M * D, M * HW)
    v = vs.reshape(B, M * D, 1)
    I = torch.eye(M * D, dtype=H.dtype)
    H = H + (ep + lm * H) * I
    Et = E.transpose(1, 2)
    S = H - torch.matmul(E, Q * Et)
    v = v - torch.matmul(E, Q * w)
    L = torch.linalg.cholesky(S)
    dso = torch.cholesky_solve(v, L)
    dz = (Q * (w - Et @ dso)).reshape(B, M, HW)
    F = torch.linalg.inv(L) @ (E * Q[..., 0])
    dzcov = (torch.sum(torch.square(F), dim=1) + Q[..., 0]).reshape(M, HW)
    return dso.reshape(B, M, D), dz, dzcov


def JDSA(targets, weights, eta, poses, disps, intrinsics, disps_prior, dscales, ii, jj, alpha):
    """ba.py:173-241 with B = 1 squeezed out: poses [P,7], disps [P,ht,wd], targets / weights [E,2,ht,wd],
    eta [M,ht,wd], disps_prior [P,ht,wd], dscales [P,hs,ws].  Returns new (disps, dscales, dzcov [M,HW])."""
    P, ht, wd = disps.shape
    HW = ht * wd
    C, w = proj_trans(poses, disps, intrinsics, targets, weights, ii, jj)
    kx = torch.unique(ii)
    M = kx.shape[0]
    prior = disps_prior[kx]
    m = (prior > 0).to(disps.dtype).view(-1, HW)
    hs, ws = dscales.shape[-2:]
    D = hs * ws
    disps_bi, Jbi = get_prior_depth_aligned(prior, dscales[kx])
    rd = (disps[kx] - disps_bi).view(-1, HW)
    Jso = -m.unsqueeze(-1) * prior.view(-1, HW).unsqueeze(-1) * Jbi.view(M, HW, -1)            # [M,HW,D]
    al = torch.full((M, HW, 1), float(alpha), dtype=disps.dtype)
    wJsoT = (al * Jso).transpose(1, 2)                                                       # [M,D,HW]
    Hs = torch.zeros(1, M, M, D, D, dtype=disps.dtype)
    Es = torch.zeros(1, M, M, D, HW, dtype=disps.dtype)
    # ba.py:218-225: blocks are scattered at (kx - kx[0], kx - kx[0]) through safe_scatter_add_*: a keyframe whose
    # shifted index is >= M (unique(ii) not a contiguous range) silently loses its prior terms -- kept as is
    sh = kx - kx[0]
    ok = sh < M
    Hs[0, sh[ok], sh[ok]] = (wJsoT @ Jso)[ok]
    Es[0, sh[ok], sh[ok]] = wJsoT[ok]                                                        # Jd = 1
    vs = torch.zeros(1, M, D, 1, dtype=disps.dtype)
    vs[0, sh[ok]] = (-wJsoT @ rd.unsqueeze(-1))[ok]
    al = al.squeeze(-1)
    Cs = C[None] + m * al + (1 - m) * eta.view(*C.shape)
    ws_ = w[None] - m * al * rd
    dso, dz, dzcov = schur_solve_mono_prior(Cs, ws_, Hs, Es, vs)
    new_disps = disps.clone()
    new_disps[kx] += dz.view(-1, ht, wd)
    new_scales = dscales.clone()
    new_scales[kx] += dso.view(-1, hs, ws)
    new_disps = torch.where(new_disps > 10, torch.zeros_like(new_disps), new_disps).clamp(min=0.001)
    return new_disps, new_scales, dzcov
