"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the two elementwise steps either side of `ba` in
FactorGraph.update:

  reproject(ii, jj)        hislam2/depth_video.pyc reproject@170-178 -> geom/projective_ops.py:44-74 without Jacobians
  cvx_upsample(data, mask) hislam2/modules/droid_net.py:21-38 (convex 8x upsampling; upsample_disp :40-44)

Pinned to the reference: tests/golden/glue.npz holds the outputs of the reference's own cvx_upsample / upsample_disp
(oracle/make_golden_glue.py), and the tiny_* golden files hold its projective_transform coords / valid.
Only tests/, __graft_entry__.smoke() and bench.py's CPU baseline may import this module.
"""
import torch

from . import dba as od


def reproject(poses, disps, intrinsics, ii, jj):
    """coords [E,ht,wd,2], valid [E,ht,wd,1] (projective_ops.py:44-74 with jacobian=False)."""
    coords, valid = od.projective_transform(poses, disps, intrinsics, ii, jj)
    return coords, valid


# neighbour k of the 3x3 window (row-major, F.unfold order droid_net.py:32): offset (k // 3 - 1, k % 3 - 1)
def cvx_upsample(data, mask):
    """data [B,ht,wd,dim], mask [B,576,ht,wd] -> [B,8ht,8wd,dim].  Written as explicit gathers over the nine
    neighbours (the reference uses F.unfold + a 7-d softmax).  Border rule (droid_net.py:26-28): a neighbour outside
    the image gets logit -inf, so the convex combination runs over the valid neighbours only.  The reference writes
    those -inf into the caller's `mask` storage as a side effect; this restatement (and the device op) does not."""
    B, ht, wd, dim = data.shape
    logits = mask.reshape(B, 9, 8, 8, ht, wd).clone()
    ys = torch.arange(ht).view(ht, 1).expand(ht, wd)
    xs = torch.arange(wd).view(1, wd).expand(ht, wd)
    out = torch.zeros(B, 8, 8, ht, wd, dim, dtype=data.dtype)
    nb = []
    for k in range(9):
        dy, dx = k // 3 - 1, k % 3 - 1
        yy, xx = ys + dy, xs + dx
        inside = (yy >= 0) & (yy < ht) & (xx >= 0) & (xx < wd)
        # the reference masks by position (first/last row, first/last column), which is the same set
        logits[:, k][..., ~inside] = float("-inf")
        g = data[:, yy.clamp(0, ht - 1), xx.clamp(0, wd - 1)] * inside[None, :, :, None].to(data.dtype)
        nb.append(g)                                                       # [B,ht,wd,dim], zero padded like unfold
    wgt = torch.softmax(logits, dim=1)                                     # [B,9,8,8,ht,wd]
    for k in range(9):
        out += wgt[:, k][..., None] * nb[k][:, None, None]
    # out[b,i,j,y,x,c] -> up[b, 8y+i, 8x+j, c]
    return out.permute(0, 3, 1, 4, 2, 5).reshape(B, 8 * ht, 8 * wd, dim)


def upsample_disp(disp, mask):
    """droid_net.py:40-44: disp [B,num,ht,wd], mask [B,num,576,ht,wd] -> [B,num,8ht,8wd]."""
    B, num, ht, wd = disp.shape
    up = cvx_upsample(disp.reshape(B * num, ht, wd, 1), mask.reshape(B * num, -1, ht, wd))
    return up.reshape(B, num, 8 * ht, 8 * wd)


def motion_features(poses, disps, intrinsics, ii, jj, target):
    """factor_graph.pyc update@195-197 (bytecode only in the reference tree; upstream DROID-SLAM factor_graph.py:
    `motn = torch.cat([coords1 - self.coords0, self.target - coords1], dim=-1); motn = motn.permute(0,1,4,2,3)
    .clamp(-64.0, 64.0)`): target [E,ht,wd,2] -> coords1 [E,ht,wd,2], valid [E,ht,wd,1], motn [E,4,ht,wd]."""
    coords1, valid = od.projective_transform(poses, disps, intrinsics, ii, jj)
    ht, wd = disps.shape[-2:]
    y, x = torch.meshgrid(torch.arange(ht, dtype=poses.dtype), torch.arange(wd, dtype=poses.dtype), indexing="ij")
    coords0 = torch.stack([x, y], dim=-1)
    motn = torch.cat([coords1 - coords0, target - coords1], dim=-1).permute(0, 3, 1, 2).clamp(-64.0, 64.0)
    return coords1, valid, motn


def damping_eta(damping, ii):
    """factor_graph.pyc update@227: 0.2 * self.damping[torch.unique(ii)].contiguous() + EP with EP = 1e-7."""
    return 0.2 * damping[torch.unique(ii)] + 1e-7
