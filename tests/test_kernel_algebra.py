"""The algebraic identities the CUDA kernels rely on, checked in fp64 against the oracle (CPU).

The engine never forms Ji, the four per-edge 6x6 blocks or the Ei panel per pixel.  It uses
(DESIGN.md "Kernel maths"), with Gij = (R,t), X0 = [x,y,1], d = disparity, X1 = R X0 + t d,
g_r the first three entries of row r of the projection Jacobian (pinhole.py:45-49):

  (1) Jj_r = [d g_r, X1 x g_r]                          (projective_ops.py:25-30,65)
  (2) Ji_r = -[d R^T g_r, X0 x R^T g_r]                 (== -Ad^T Jj_r, projective_ops.py:66-67)
  (3) Jz_r = g_r . t                                    (projective_ops.py:69-70)
  (4) with a_r = w_r Jz_r and h = sum_r a_r g_r:  Ej = [d h, X1 x h],  Ei = -[d R^T h, X0 x R^T h]
  (5) Hii = Ad^T Hjj Ad, Hij = -Ad^T Hjj, Hji = -Hjj Ad, vi = -Ad^T vj      (ba.py:54-59)
"""
import torch

import oracle.dba as od
import oracle.se3 as se3
from cut3r_slam_b200.synthetic import make_problem


def _setup():
    p = make_problem(P=5, E=12, ht=12, wd=16, t0=1, r=2, seed=3)
    f = lambda t: t.double()
    poses, disps, intr = f(p.poses), f(p.disps), f(p.intrinsics)
    tgt = f(p.targets).permute(0, 2, 3, 1).contiguous()
    wgt = f(p.weights).permute(0, 2, 3, 1).contiguous()
    return p, poses, disps, intr, tgt, wgt


def test_jacobian_identities():
    p, poses, disps, intr, tgt, wgt = _setup()
    ii, jj = p.ii, p.jj
    coords, valid, (Ji, Jj, Jz) = od.projective_transform(poses, disps, intr, ii, jj, jacobian=True)
    Gij = se3.se3_mul(poses[jj], se3.se3_inv(poses[ii]))
    t, q = se3.se3_split(Gij)
    R = se3.quat_to_matrix(q)
    X0, _ = od.iproj_pinhole(disps[ii], intr)
    d = X0[..., 3]
    x0 = X0[..., :3]
    X1 = torch.einsum("eab,ehwb->ehwa", R, x0) + t[:, None, None, :] * d[..., None]
    fx, fy = intr[0], intr[1]
    Z = X1[..., 2]
    Zc = torch.where(Z < 0.1, torch.ones_like(Z), Z)
    dz = 1.0 / Zc
    o = torch.zeros_like(dz)
    g = torch.stack([torch.stack([fx * dz, o, -fx * X1[..., 0] * dz * dz], -1),
                     torch.stack([o, fy * dz, -fy * X1[..., 1] * dz * dz], -1)], -2)     # [E,h,w,2,3]
    X1b = X1[..., None, :].expand_as(g)
    Jj2 = torch.cat([d[..., None, None] * g, torch.cross(X1b, g, dim=-1)], -1)
    assert torch.allclose(Jj2, Jj, atol=1e-12)
    Rtg = torch.einsum("eba,ehwrb->ehwra", R, g)
    x0b = x0[..., None, :].expand_as(g)
    Ji2 = -torch.cat([d[..., None, None] * Rtg, torch.cross(x0b, Rtg, dim=-1)], -1)
    assert torch.allclose(Ji2, Ji, atol=1e-11)
    Jz2 = (g * t[:, None, None, None, :]).sum(-1, keepdim=True)
    assert torch.allclose(Jz2, Jz, atol=1e-12)

    # (4) panel entries
    blk = od.edge_blocks(tgt, wgt, poses, disps, intr, ii, jj)
    E_, ht, wd = tgt.shape[:3]
    w = od.WEIGHT_SCALE * valid * wgt                                   # [E,h,w,2]
    a = w * Jz2[..., 0]
    h = (a[..., None] * g).sum(-2)                                      # [E,h,w,3]
    Ej = torch.cat([d[..., None] * h, torch.cross(X1, h, dim=-1)], -1)
    assert torch.allclose(Ej.reshape(E_, -1, 6).transpose(1, 2), blk["Ej"], atol=1e-12)
    Rth = torch.einsum("eba,ehwb->ehwa", R, h)
    Ei = -torch.cat([d[..., None] * Rth, torch.cross(x0, Rth, dim=-1)], -1)
    assert torch.allclose(Ei.reshape(E_, -1, 6).transpose(1, 2), blk["Ei"], atol=1e-11)

    # (5) per-edge blocks from Hjj, vj alone
    Ad = se3.se3_adj_matrix(Gij)
    AdT = Ad.transpose(1, 2)
    assert torch.allclose(AdT @ blk["Hjj"] @ Ad, blk["Hii"], atol=1e-9)
    assert torch.allclose(-AdT @ blk["Hjj"], blk["Hij"], atol=1e-9)
    assert torch.allclose(-blk["Hjj"] @ Ad, blk["Hji"], atol=1e-9)
    assert torch.allclose(-(AdT @ blk["vj"][..., None])[..., 0], blk["vi"], atol=1e-9)


def test_butterfly_index_map():
    """The warp reduce-scatter used for the 26 per-edge sums leaves value i on lane i."""
    vals = torch.arange(32 * 32, dtype=torch.float64).reshape(32, 32)   # [lane, value]
    v = vals.clone()
    cnt, mask = 32, 16
    while mask >= 1:
        half = cnt // 2
        new = torch.zeros(32, half, dtype=v.dtype)
        for lane in range(32):
            peer = lane ^ mask
            bit = (lane & mask) != 0
            for i in range(half):
                keep = v[lane, i + half] if bit else v[lane, i]
                recv = v[peer, i + half] if bit else v[peer, i]      # peer sends the half WE keep
                new[lane, i] = keep + recv
        v, cnt, mask = new, half, mask // 2
    assert torch.equal(v[:, 0], vals.sum(0))


def test_sim3_translation_coefficient_needs_fp64_for_small_scale_updates():
    """Why k_pg_retract evaluates Exp(dx) in fp64: W = A Phi + B Phi^2 + C I with C = (e^sigma - 1) / sigma
    (rxso3.h:190-233) multiplies tau directly.  The formula's series branch starts below |sigma| = 1e-6; the scale updates
    of a converging solve are 1e-6 .. 1e-5, where fp32 keeps about two digits of C."""
    import numpy as np
    worst32, worst64 = 0.0, 0.0
    for sigma in (2.5e-6, 4.7e-6, 6.5e-6, 1.2e-5, -7.5e-6, -1.3e-5):
        exact = float(np.expm1(np.float64(sigma)) / np.float64(sigma))
        s32 = np.float32(sigma)
        c32 = float((np.exp(s32, dtype=np.float32) - np.float32(1)) / s32)
        c64 = float((np.exp(np.float64(sigma)) - 1.0) / np.float64(sigma))
        worst32 = max(worst32, abs(c32 - exact) / exact)
        worst64 = max(worst64, abs(c64 - exact) / exact)
    assert worst32 > 1e-3            # fp32: 0.1 % .. several % of tau
    assert worst64 < 1e-9            # fp64: far below the fp32 rounding of the stored pose
