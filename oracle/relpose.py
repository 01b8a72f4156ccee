"""TEST INFRASTRUCTURE ONLY -- CPU restatement of PGOBuffer.add_rel_poses (hislam2/pgo_buffer.py:85-123, SURVEY.md
section 8f rank 4): per loop-closure edge, a 6-dof Gauss-Newton estimate of the relative pose from the flow targets
with its diagonal covariance.  Pinned to the reference's own method run unmodified (oracle/make_golden_relpose.py ->
tests/golden/relpose.npz).  Only tests/, smoke() and bench.py's CPU baseline may import this module."""
import torch

from . import dba as od
from . import se3


def add_rel_poses(poses, disps, intrinsics, ii, jj, target, weight):
    """poses [P,7], disps [P,ht,wd], intrinsics [4], target / weight [E,ht,wd,2] (the FactorGraph layout without its
    leading batch dimension).  Returns the rows the reference appends to its buffers:
    (ii, jj, rel_poses [N,7], rel_covs [N,6], valid_percent [N]) for the N edges that pass the weight mask."""
    valid_pct = (weight > 0.1).to(weight.dtype).mean(dim=(-3, -2, -1))              # :87
    keep = valid_pct > 1e-2                                                          # :88
    ii, jj, target, weight = ii[keep], jj[keep], target[keep], weight[keep]
    N = ii.shape[0]
    if N < 1:                                                                        # :92-93
        z = poses.new_zeros
        return ii, jj, z((0, 7)), z((0, 6)), z((0,))
    rel = se3.se3_mul(poses[jj], se3.se3_inv(poses[ii]))                             # :97
    for _ in range(4):                                                               # :99-110: `poses` never changes,
        coords, valid, (Ji, Jj, Jz) = od.projective_transform(poses, disps, intrinsics, ii, jj, jacobian=True)
        r = (target - coords).reshape(N, -1, 1)                                      # so every pass gives the same dx
        w = 0.001 * (valid * weight).reshape(N, -1, 1)
        J = Jj.reshape(N, -1, 6)
        wJT = (w * J).transpose(1, 2)
        H = wJT @ J + 1e-4 * torch.eye(6, dtype=poses.dtype)
        v = wJT @ r
        Hinv = torch.linalg.inv(H)
        dx = Hinv @ v
        rel = se3.se3_retr(rel, dx[..., 0])
    V = J @ dx - r                                                                   # :112
    sig2 = (w * V).transpose(1, 2) @ V                                               # :113
    cov = torch.diagonal(sig2 * Hinv, dim1=1, dim2=2)                                # :114-115
    return ii, jj, rel, cov, valid_pct[keep]
