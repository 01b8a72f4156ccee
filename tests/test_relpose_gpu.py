"""GPU parity of `dba_rel_pose_cov` / glue.add_rel_poses against the reference's own
PGOBuffer.add_rel_poses outputs (tests/golden/relpose.npz) and the fp64 oracle.  Tolerances: the reference's own fp32
run sits 2.3e-5 (poses) / 7.5e-5 relative (covariances) from its fp64 run on these cases; the device accumulates in
fp32 and solves the 6x6 system in fp64 -> 1e-4 absolute on poses, 5e-4 relative on covariances."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import oracle.relpose as orp  # noqa: E402
from cut3r_slam_b200.synthetic import make_problem  # noqa: E402
from oracle.make_golden_relpose import CASES, make_case  # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(ROOT, "tests", "golden", "relpose.npz"))


def _glue():
    from cut3r_slam_b200 import glue
    return glue


@pytest.mark.parametrize("name", sorted(CASES))
def test_add_rel_poses_matches_reference_golden(name):
    gl = _glue()
    p, target, weight = make_case(CASES[name])
    ii, jj, rel, cov, vp = gl.add_rel_poses(p.poses.cuda(), p.disps.cuda(), p.intrinsics.cuda(), p.ii.cuda(), p.jj.cuda(),
                                            target[None].cuda(), weight[None].cuda())
    assert np.array_equal(ii.cpu().numpy(), GOLD[f"{name}_ii"]) and np.array_equal(jj.cpu().numpy(), GOLD[f"{name}_jj"])
    ref_c = torch.from_numpy(GOLD[f"{name}_covs_f64"])
    assert float((rel.cpu().double() - torch.from_numpy(GOLD[f"{name}_poses_f64"])).abs().max()) < 1e-4
    assert float(((cov.cpu().double() - ref_c).abs() / ref_c.abs()).max()) < 5e-4
    assert float((vp.cpu().double() - torch.from_numpy(GOLD[f"{name}_valid_f64"])).abs().max()) < 1e-6


def test_add_rel_poses_full_size_matches_oracle():
    gl = _glue()
    p = make_problem("frontend")                                        # 100 edges, 43x77
    target = p.targets.permute(0, 2, 3, 1).contiguous()
    weight = p.weights.permute(0, 2, 3, 1).contiguous()
    out = gl.add_rel_poses(p.poses.cuda(), p.disps.cuda(), p.intrinsics.cuda(), p.ii.cuda(), p.jj.cuda(), target.cuda(),
                           weight.cuda())
    d = torch.float64
    ref = orp.add_rel_poses(p.poses.to(d), p.disps.to(d), p.intrinsics.to(d), p.ii, p.jj, target.to(d), weight.to(d))
    assert torch.equal(out[0].cpu(), ref[0]) and torch.equal(out[1].cpu(), ref[1])
    assert float((out[2].cpu().double() - ref[2]).abs().max()) < 1e-4
    assert float(((out[3].cpu().double() - ref[3]).abs() / ref[3].abs()).max()) < 5e-4
    assert float((out[2][:, 3:].norm(dim=1) - 1).abs().max()) < 1e-5   # unit quaternions
    e = torch.zeros(0, dtype=torch.long, device="cuda")
    z = torch.zeros(0, *target.shape[1:], device="cuda")
    assert gl.add_rel_poses(p.poses.cuda(), p.disps.cuda(), p.intrinsics.cuda(), e, e, z, z)[2].shape == (0, 7)
    with pytest.raises(RuntimeError):
        gl.add_rel_poses(p.poses.cuda(), p.disps.cuda(), p.intrinsics.cuda(), p.ii.cuda(), p.jj.cuda(), target, weight)
