"""The compiled `droid_backends` extension against the ctypes binding (same libdba underneath: bit-identical), called
with the reference's argument lists (depth_video.pyc cuda_ba@232, distance@195-205, droid_visualization.py:97-101)."""
import pytest
import torch

from test_ext_cpu import _ext
from cut3r_slam_b200.synthetic import make_problem

pytestmark = pytest.mark.gpu


def test_ba_through_the_compiled_extension_matches_the_ctypes_binding():
    from cut3r_slam_b200 import droid_backends as db
    m = _ext()
    p = make_problem("correctness")
    a = p.to("cuda")
    dx_a, dz_a = db.ba(*a.ba_args())
    b = p.to("cuda")
    dx_b, dz_b = m.ba(*b.ba_args())
    torch.cuda.synchronize()
    assert torch.equal(b.poses, a.poses) and torch.equal(b.disps, a.disps)
    assert torch.equal(dx_b, dx_a) and torch.equal(dz_b, dz_a)
    # motion-only: eta is ignored, dz is empty
    q = make_problem(P=8, E=24, ht=16, wd=24, iters=2, t0=1, r=2, seed=3, motion_only=True)
    c, d = q.to("cuda"), q.to("cuda")
    dx_c, _ = db.ba(*c.ba_args())
    dx_d, dz_d = m.ba(*d.ba_args())
    assert torch.equal(dx_c, dx_d) and dz_d.numel() == 0 and torch.equal(c.poses, d.poses)


def test_companion_ops_through_the_compiled_extension():
    from cut3r_slam_b200 import droid_backends as db
    m = _ext()
    g = make_problem("correctness").to("cuda")
    ii = torch.arange(0, 7, device="cuda")
    jj = ii + 1
    assert torch.equal(m.frame_distance(g.poses, g.disps, g.intrinsics, ii, jj, 0.3),
                       db.frame_distance(g.poses, g.disps, g.intrinsics, ii, jj, 0.3))
    assert torch.equal(m.covis_distance(g.poses, g.disps, g.intrinsics, ii), db.covis_distance(g.poses, g.disps, g.intrinsics, ii))
    assert torch.equal(m.iproj(g.poses, g.disps, g.intrinsics), db.iproj(g.poses, g.disps, g.intrinsics))
    ix = torch.arange(1, 6, device="cuda")
    th = torch.full((5,), 0.01, device="cuda")
    assert torch.equal(m.depth_filter(g.poses, g.disps, g.intrinsics, ix, th), db.depth_filter(g.poses, g.disps, g.intrinsics, ix, th))
