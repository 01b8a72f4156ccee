"""TEST INFRASTRUCTURE ONLY -- the same dense BA step as oracle/dba.py `BA` (hislam2/geom/ba.py:32-107 +
chol.py:47-78 of the reference), evaluated one SOURCE KEYFRAME at a time so that the dense
E [6P', M*HW] of chol.py:52 (19 GB at BASELINE config 4 in fp32, twice that in fp64) is never formed.

The algebra is identical: E has non-zero 6-row blocks only for the source pose and its neighbours
(ba.py:83-84: blocks are scattered at (i, kk) and (j, kk)), so
    S  = H - E Q E^T  = H - sum_k  P_k diag(Q_k) P_k^T      (chol.py:63)
    vs = v - E Q w    = v - sum_k  P_k (Q_k w_k)            (chol.py:64)
    dz_k = Q_k (w_k - P_k^T dx)                             (chol.py:70)
with P_k the [6 (deg_k + 1), HW] panel of keyframe k.  tests/test_oracle_lowmem_cpu.py pins this
restatement to oracle/dba.py `BA` (itself pinned to the unmodified reference through tests/golden) to 1e-12 in
fp64.  In fp32 the summation order differs from the reference's one big matmul; it is then another draw of the
same rounding noise (used only for the "reference's own fp32-vs-fp64 gap" at sizes the dense form cannot reach).
"""
import torch

from . import dba, se3


def BA_lowmem(target, weight, eta, poses, disps, intr, ii, jj, fixedp=1, ep=0.1, lm=0.0001, chunk=32,
              return_info=False):
    """One LM step; arguments and return values of oracle.dba.BA (reference layout: target/weight [E,ht,wd,2])."""
    P, ht, wd = disps.shape
    HW = ht * wd
    Pp = P - fixedp
    dt = disps.dtype
    kx, kk = torch.unique(ii, return_inverse=True)
    M = kx.shape[0]
    H = torch.zeros(Pp * Pp, 6, 6, dtype=dt)
    v = torch.zeros(Pp, 6, dtype=dt)
    S_sub = torch.zeros(Pp * 6, Pp * 6, dtype=dt)         # E Q E^T
    vs_sub = torch.zeros(Pp * 6, dtype=dt)                # E Q w
    panels = []                                           # per keyframe: (pose slots [nb], panel [nb,6,HW], Q, w)
    order = torch.argsort(kk, stable=True)                # edges grouped by source keyframe, ascending edge id
    counts = torch.bincount(kk, minlength=M)
    starts = torch.cumsum(counts, 0) - counts
    eta = eta.reshape(M, HW)
    for m0 in range(0, M, chunk):
        m1 = min(m0 + chunk, M)
        sel = order[starts[m0]:starts[m1 - 1] + counts[m1 - 1]]
        blk = dba.edge_blocks(target[sel], weight[sel], poses, disps, intr, ii[sel], jj[sel])
        i0, j0 = ii[sel] - fixedp, jj[sel] - fixedp
        H += (dba.safe_scatter_add_mat(blk["Hii"], i0, i0, Pp, Pp) + dba.safe_scatter_add_mat(blk["Hij"], i0, j0, Pp, Pp) +
              dba.safe_scatter_add_mat(blk["Hji"], j0, i0, Pp, Pp) + dba.safe_scatter_add_mat(blk["Hjj"], j0, j0, Pp, Pp))
        v += dba.safe_scatter_add_vec(blk["vi"], i0, Pp) + dba.safe_scatter_add_vec(blk["vj"], j0, Pp)
        kks = kk[sel]
        for m in range(m0, m1):
            loc = torch.nonzero(kks == m).reshape(-1)
            C = blk["Ck"][loc].sum(0) + eta[m] + dba.C_EPS                       # ba.py:89,92
            w = blk["wk"][loc].sum(0)                                            # ba.py:90
            Q = 1.0 / C
            # panel rows: the source pose (sum of Ei over the keyframe's edges) + one block per target pose
            slots = {}
            for e in loc.tolist():
                for s, blkE in ((int(i0[e]), blk["Ei"][e]), (int(j0[e]), blk["Ej"][e])):
                    if s < 0 or s >= Pp:
                        continue
                    slots[s] = slots[s] + blkE if s in slots else blkE.clone()
            if slots:
                ps = torch.tensor(sorted(slots), dtype=torch.long)
                Pk = torch.stack([slots[int(s)] for s in ps])                    # [nb,6,HW]
                nb = ps.numel()
                flat = Pk.reshape(nb * 6, HW)
                G = (flat * Q) @ flat.T                                          # [6nb,6nb]
                rows = (ps[:, None] * 6 + torch.arange(6)[None]).reshape(-1)
                S_sub[rows[:, None], rows[None, :]] += G
                vs_sub[rows] += flat @ (Q * w)
            else:
                ps, Pk = torch.zeros(0, dtype=torch.long), torch.zeros(0, 6, HW, dtype=dt)
            panels.append((ps, Pk, Q, w))
    Hm = H.reshape(Pp, Pp, 6, 6).permute(0, 2, 1, 3).reshape(Pp * 6, Pp * 6)
    Hm = Hm + (ep + lm * Hm) * torch.eye(Pp * 6, dtype=dt)                       # chol.py:56-57
    S = Hm - S_sub
    vs = v.reshape(-1) - vs_sub
    dx, _, failed = dba.cholesky_solve(S, vs.reshape(-1, 1))
    dx = dx.reshape(Pp, 6)
    dz = torch.zeros(M, HW, dtype=dt)
    for m, (ps, Pk, Q, w) in enumerate(panels):
        t = w.clone()
        if ps.numel():
            t -= torch.einsum("bdp,bd->p", Pk, dx[ps])
        dz[m] = Q * t                                                            # chol.py:70
    new_poses = poses.clone()
    new_poses[fixedp:] = se3.se3_retr(poses[fixedp:], dx)
    new_disps = disps.clone()
    new_disps[kx] += dz.reshape(-1, ht, wd)
    new_disps = torch.where(new_disps > 10, torch.zeros_like(new_disps), new_disps)
    new_disps = new_disps.clamp(min=0.001)
    if return_info:
        return new_poses, new_disps, dict(dx=dx, dz=dz, S=S, vs=vs, failed=failed)
    return new_poses, new_disps


def oracle_ba_lowmem(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1, iterations, lm, ep,
                     motion_only, dtype=torch.float64, return_info=False):
    """oracle.dba.oracle_ba for the dense (non motion-only, no sensor depth) case on top of BA_lowmem."""
    assert not motion_only, "motion-only needs no E: use oracle.dba.oracle_ba"
    assert disps_sens is None or not bool((disps_sens[:t1] > 0).any())
    out_dtype = poses.dtype
    p = poses[:t1].to(dtype).clone()
    d = disps[:t1].to(dtype).clone()
    intr = intrinsics.to(dtype)
    tgt = targets.permute(0, 2, 3, 1).to(dtype)
    wgt = weights.permute(0, 2, 3, 1).to(dtype)
    et = eta.to(dtype)
    ii, jj = ii.long().cpu(), jj.long().cpu()
    infos = []
    for _ in range(iterations):
        p, d, info = BA_lowmem(tgt, wgt, et, p, d, intr, ii, jj, fixedp=t0, ep=ep, lm=lm, return_info=True)
        infos.append((info["dx"], info["dz"], info["S"], info["vs"]))
    new_poses, new_disps = poses.clone(), disps.clone()
    new_poses[:t1] = p.to(out_dtype)
    new_disps[:t1] = d.to(out_dtype)
    if return_info:
        return new_poses, new_disps, infos
    return new_poses, new_disps
