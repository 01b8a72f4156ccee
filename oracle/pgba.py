"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the Sim3 pose-graph bundle adjustment (`pgba`).

Three parts, with their pin status:
  1. the 7-dof reprojection step = the reference's own `BA` (hislam2/geom/ba.py:32-107) called with `lietorch.Sim3`
     poses (`D = poses.manifold_dim`, ba.py:37; Sim3 column of Ja: geom/projective_ops.py:32-38; Sim3 adjoint /
     exp: thirdparty/lietorch/lietorch/include/sim3.h) -- PINNED: tests/golden/pgba.npz holds the outputs of the
     unmodified reference run through oracle/ref_harness.py (oracle/make_golden_pgba.py);
  2. the relative-pose factor blocks Hsp, vsp = `global_relative_posesim3_constraints`
     (hislam2/pgo_buffer.py:39-62, numeric Jacobians :19-37) -- PINNED the same way (the function's source is executed
     with its hard-coded device='cuda' replaced by 'cpu', nothing else);
  3. how `pgba` composes the two (adds Hsp / vsp into H / v before the Schur solve, then retracts) lives in bytecode
     only (depth_video.pyc cuda_pgba@244-314, factor_graph.pyc update_pgba@284-341) -- PARITY UNPINNED, restated here.
"""
import torch

from . import dba as od
from . import se3, sim3

EPS_NUM = 1e-8          # pgo_buffer.py:14


def edge_blocks_sim3(target, weight, poses, disps, intr, ii, jj):
    """ba.py:39-68 with D = 7.  poses [P,8]; returns the dict of od.edge_blocks with 7-wide blocks."""
    E_, ht, wd = target.shape[:3]
    D = 7
    X0, Jz0 = od.iproj_pinhole(disps[ii], intr)
    Gij = sim3.mul(poses[jj], sim3.inv(poses[ii]))                       # projective_ops.py:51
    X1 = sim3.act4(Gij[:, None, None, :], X0)
    X, Y, Z, d = X1.unbind(dim=-1)
    o = torch.zeros_like(d)
    Ja = torch.stack([                                                   # projective_ops.py:32-38
        d, o, o, o, Z, -Y, X,
        o, d, o, -Z, o, X, Y,
        o, o, d, Y, -X, o, Z,
        o, o, o, o, o, o, o,
    ], dim=-1).reshape(d.shape + (4, 7))
    coords, Jp = od.proj_pinhole(X1, intr)
    valid = ((X1[..., 2] > od.MIN_DEPTH) & (X0[..., 2] > od.MIN_DEPTH)).to(disps.dtype).unsqueeze(-1)
    Jj = Jp @ Ja
    Ad = sim3.adj_matrix(Gij)
    Ji = -(Jj @ Ad[:, None, None])                                       # projective_ops.py:66-67 (adjT per row)
    Jzp = sim3.act4(Gij[:, None, None, :], Jz0)
    Jz = Jp @ Jzp.unsqueeze(-1)
    r = (target - coords).reshape(E_, -1, 1)
    w = od.WEIGHT_SCALE * (valid * weight).reshape(E_, -1, 1)
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


def rel_pose_blocks(ii, jj, poses, rel_poses, infos, pw=1e-5):
    """pgo_buffer.py:39-62.  poses [P,8] Sim3, rel_poses [R,7] SE3 (scale 1), infos [R,7,7].
    Residual e = log(Gij * Gi * Gj^-1); Jacobians by central differences (eps = 1e-8) of left perturbations in
    fp64, cast to fp32; Hsp [4,R,7,7] = (ii,ii),(ii,jj),(jj,ii),(jj,jj) blocks, vsp [2,R,7] = -(pw J)^T info r."""
    f64 = torch.float64
    Gi = poses[ii].to(f64)
    Gj = poses[jj].to(f64)
    Gij = torch.cat([rel_poses.to(f64), torch.ones(rel_poses.shape[0], 1, dtype=f64)], dim=-1)

    def func(a, b):
        return sim3.log(sim3.mul(sim3.mul(Gij, a), sim3.inv(b)))

    def num_jac(first):
        cols = []
        for k in range(7):
            delta = torch.zeros(7, dtype=f64)
            delta[k] = EPS_NUM
            ep, em = sim3.exp(delta), sim3.exp(-delta)
            if first:
                cols.append((func(sim3.mul(ep, Gi), Gj) - func(sim3.mul(em, Gi), Gj)) / 2.0 / EPS_NUM)
            else:
                cols.append((func(Gi, sim3.mul(ep, Gj)) - func(Gi, sim3.mul(em, Gj))) / 2.0 / EPS_NUM)
        return torch.stack(cols, dim=-1)                                 # [R,7,7]

    Ji = num_jac(True).float()
    Jj = num_jac(False).float()
    r = func(Gi, Gj).unsqueeze(-1).float()
    wJiT = ((pw * Ji.double()).transpose(1, 2) @ infos.double()).float()
    wJjT = ((pw * Jj.double()).transpose(1, 2) @ infos.double()).float()
    Hsp = torch.stack([wJiT @ Ji, wJiT @ Jj, wJjT @ Ji, wJjT @ Jj])
    vsp = -torch.stack([wJiT @ r, wJjT @ r]).squeeze(-1)
    return Hsp, vsp


def pgba_step(target, weight, eta, poses, disps, intr, ii, jj, fixedp, ep, lm, rel=None, return_info=False):
    """One step: ba.py:70-107 with D = 7, plus (UNPINNED composition) the relative-pose blocks added to H, v.
    rel = (rel_ii, rel_jj, rel_poses [R,7], infos [R,7,7], pw) or None."""
    P, ht, wd = disps.shape
    D = 7
    blk = edge_blocks_sim3(target, weight, poses, disps, intr, ii, jj)
    kx, kk = torch.unique(ii, return_inverse=True)
    M = kx.shape[0]
    Pp = P - fixedp
    i0, j0 = ii - fixedp, jj - fixedp
    H = (od.safe_scatter_add_mat(blk["Hii"], i0, i0, Pp, Pp) + od.safe_scatter_add_mat(blk["Hij"], i0, j0, Pp, Pp) +
         od.safe_scatter_add_mat(blk["Hji"], j0, i0, Pp, Pp) + od.safe_scatter_add_mat(blk["Hjj"], j0, j0, Pp, Pp))
    Em = od.safe_scatter_add_mat(blk["Ei"], i0, kk, Pp, M) + od.safe_scatter_add_mat(blk["Ej"], j0, kk, Pp, M)
    v = od.safe_scatter_add_vec(blk["vi"], i0, Pp) + od.safe_scatter_add_vec(blk["vj"], j0, Pp)
    C = od.safe_scatter_add_vec(blk["Ck"], kk, M) + eta.reshape(M, -1) + od.C_EPS
    w = od.safe_scatter_add_vec(blk["wk"], kk, M)
    if rel is not None:
        rii, rjj, rel_poses, infos, pw = rel
        Hsp, vsp = rel_pose_blocks(rii, rjj, poses, rel_poses, infos, pw)
        Hsp, vsp = Hsp.to(H.dtype), vsp.to(H.dtype)
        a0, b0 = rii - fixedp, rjj - fixedp
        H = H + (od.safe_scatter_add_mat(Hsp[0], a0, a0, Pp, Pp) + od.safe_scatter_add_mat(Hsp[1], a0, b0, Pp, Pp) +
                 od.safe_scatter_add_mat(Hsp[2], b0, a0, Pp, Pp) + od.safe_scatter_add_mat(Hsp[3], b0, b0, Pp, Pp))
        v = v + od.safe_scatter_add_vec(vsp[0], a0, Pp) + od.safe_scatter_add_vec(vsp[1], b0, Pp)
    sol = od.schur_solve(H.reshape(Pp, Pp, D, D), Em.reshape(Pp, M, D, ht * wd), C, v, w, ep=ep, lm=lm)
    dx, dz = sol["dx"], sol["dz"]
    new_poses = poses.clone()
    new_poses[fixedp:] = sim3.retr(poses[fixedp:], dx)
    new_disps = disps.clone()
    new_disps[kx] += dz.reshape(-1, ht, wd)
    new_disps = torch.where(new_disps > 10, torch.zeros_like(new_disps), new_disps)
    new_disps = new_disps.clamp(min=0.001)
    if return_info:
        return new_poses, new_disps, dict(dx=dx, dz=dz, S=sol["S"], vs=sol["vs"], blk=blk)
    return new_poses, new_disps


def oracle_pgba(poses, disps, intrinsics, targets, weights, eta, ii, jj, rel_ii, rel_jj, rel_poses, rel_infos, t0, t1,
                iterations, lm, ep, pw, dtype=torch.float32, return_info=False):
    """The 19-argument `pgba` op of cut3r_slam_b200.droid_backends (device arguments minus motion_only): poses [N,8],
    targets / weights [E,2,ht,wd] (op layout), rel_* may be None / empty.  Returns new (poses, disps)."""
    out_dtype = poses.dtype
    p = poses[:t1].to(dtype).clone()
    d = disps[:t1].to(dtype).clone()
    intr = intrinsics.to(dtype)
    tgt = targets.permute(0, 2, 3, 1).to(dtype).contiguous()
    wgt = weights.permute(0, 2, 3, 1).to(dtype).contiguous()
    ii, jj = ii.long().cpu(), jj.long().cpu()
    rel = None
    if rel_ii is not None and rel_ii.numel() > 0:
        rel = (rel_ii.long().cpu(), rel_jj.long().cpu(), rel_poses, rel_infos, pw)
    infos = []
    for _ in range(iterations):
        p, d, info = pgba_step(tgt, wgt, eta.to(dtype), p, d, intr, ii, jj, t0, ep, lm, rel, return_info=True)
        infos.append((info["dx"], info["dz"]))
    new_poses, new_disps = poses.clone(), disps.clone()
    new_poses[:t1] = p.to(out_dtype)
    new_disps[:t1] = d.to(out_dtype)
    if return_info:
        return new_poses, new_disps, infos
    return new_poses, new_disps
