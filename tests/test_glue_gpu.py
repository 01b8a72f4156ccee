"""GPU parity of reproject and cvx_upsample against oracle/glue.py and the reference's own
outputs in tests/golden/.  Tolerances: both ops are elementwise fp32 -- coords 2e-5 relative to the image extent
(the reference's fp32 run sits at the same distance from fp64), upsampled values 2e-6 relative (a 9-term softmax);
`valid` must agree except for points whose transformed depth sits within 1e-5 of MIN_DEPTH."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import oracle.glue as og  # noqa: E402
from cut3r_slam_b200.synthetic import make_problem  # noqa: E402
from oracle.make_golden_glue import CASES, make_case  # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(ROOT, "tests", "golden", "glue.npz"))


def _glue():
    from cut3r_slam_b200 import glue
    return glue


@pytest.mark.parametrize("name", sorted(CASES))
def test_cvx_upsample_matches_reference_golden(name):
    gl = _glue()
    data, mask = make_case(CASES[name])
    m = mask.float().cuda()
    keep = m.clone()
    up = gl.cvx_upsample(data.float().cuda(), m).cpu().double()
    ref = torch.from_numpy(GOLD[f"{name}_up"])
    assert up.shape == ref.shape
    assert float((up - ref).abs().max() / ref.abs().max()) < 2e-6
    assert torch.equal(m, keep)


def test_upsample_disp_full_size_matches_oracle():
    gl = _glue()
    g = torch.Generator().manual_seed(11)
    B, num, ht, wd = 1, 3, 43, 77                                    # wd not a multiple of the 32-pixel tile
    disp = torch.rand(B, num, ht, wd, generator=g) + 0.05
    mask = 4.0 * torch.randn(B, num, 576, ht, wd, generator=g)
    up = gl.upsample_disp(disp.cuda(), mask.cuda()).cpu()
    ref = og.upsample_disp(disp.double(), mask.double())
    assert up.shape == (B, num, 8 * ht, 8 * wd)
    assert float((up.double() - ref).abs().max() / ref.abs().max()) < 2e-6
    assert float(up.min()) >= float(disp.min()) - 1e-6 and float(up.max()) <= float(disp.max()) + 1e-6
    # more channels than one pass of the kernel holds
    data = torch.rand(2, 6, 9, 7, generator=g)
    mk = torch.randn(2, 576, 6, 9, generator=g)
    up = gl.cvx_upsample(data.cuda(), mk.cuda()).cpu()
    ref = og.cvx_upsample(data.double(), mk.double())
    assert float((up.double() - ref).abs().max()) < 2e-6
    with pytest.raises(RuntimeError):
        gl.cvx_upsample(data.cuda(), mk[:, :500].contiguous().cuda())
    with pytest.raises(RuntimeError):
        gl.cvx_upsample(data, mk)                                    # host tensors: no CPU fallback


@pytest.mark.parametrize("name", ["tiny_ba", "tiny_moba", "tiny_ba_backend"])
def test_reproject_matches_reference_golden(name):
    gl = _glue()
    G = np.load(os.path.join(ROOT, "tests", "golden", f"{name}.npz"))
    t = lambda k: torch.from_numpy(G[k])
    coords, valid = gl.reproject(t("poses").float().cuda(), t("disps").float().cuda(), t("intrinsics").float().cuda(),
                                 t("ii").long().cuda(), t("jj").long().cuda())
    assert coords.shape[0] == 1 and valid.shape[0] == 1 and valid.shape[-1] == 1
    sl = (slice(None), slice(1, None, 5), slice(2, None, 7))
    c, v = coords[0].cpu()[sl].double(), valid[0].cpu()[sl].double()
    ref_c, ref_v = t("jac_coords").double(), t("jac_valid").double()
    assert float((c - ref_c).abs().max()) < 2e-5 * float(ref_c.abs().max())
    assert torch.equal(v, ref_v)


def test_reproject_backend_size_matches_oracle():
    gl = _glue()
    p = make_problem("frontend")
    g = p.to("cuda")
    coords, valid = gl.reproject(g.poses, g.disps, g.intrinsics, g.ii, g.jj)
    d = torch.float64
    rc, rv = og.reproject(p.poses.to(d), p.disps.to(d), p.intrinsics.to(d), p.ii, p.jj)
    E, (ht, wd) = p.ii.numel(), p.disps.shape[1:]
    assert coords.shape == (1, E, ht, wd, 2) and valid.shape == (1, E, ht, wd, 1)
    assert float((coords[0].cpu().double() - rc).abs().max()) < 2e-5 * float(rc.abs().max())
    diff = valid[0].cpu().double() != rv
    assert float(diff.double().mean()) < 1e-4                        # only points sitting on the MIN_DEPTH cut
    # a frame reprojected onto itself is the pixel grid
    ii = torch.arange(3, device="cuda")
    c, v = gl.reproject(g.poses, g.disps, g.intrinsics, ii, ii)
    ys, xs = torch.meshgrid(torch.arange(ht, dtype=torch.float32), torch.arange(wd, dtype=torch.float32), indexing="ij")
    assert float((c[0, :, :, :, 0].cpu() - xs).abs().max()) < 1e-3 and float((c[0, :, :, :, 1].cpu() - ys).abs().max()) < 1e-3
    assert float(v.min()) == 1.0
    e = torch.zeros(0, dtype=torch.long, device="cuda")
    assert gl.reproject(g.poses, g.disps, g.intrinsics, e, e)[0].shape == (1, 0, ht, wd, 2)
    with pytest.raises(RuntimeError):
        gl.reproject(g.poses, g.disps, g.intrinsics, ii.int(), ii.int())
    # an index outside the frame buffer is answered with valid = 0 instead of an out-of-bounds read
    bad = torch.tensor([0, 10 ** 6, -1], device="cuda")
    c, v = gl.reproject(g.poses, g.disps, g.intrinsics, torch.zeros(3, dtype=torch.long, device="cuda"), bad)
    assert float(v[0, 0].min()) == 1.0 and float(v[0, 1:].abs().max()) == 0.0 and float(c[0, 1:].abs().max()) == 0.0


def test_motion_features_and_damping_gather():
    """The two remaining elementwise steps of FactorGraph.update (factor_graph.pyc update@195-197, @227) against
    their torch restatement (bytecode only in the reference tree: parity unpinned)."""
    import oracle.glue as og
    from cut3r_slam_b200 import glue
    from cut3r_slam_b200.synthetic import make_problem
    p = make_problem("frontend")
    g = p.to("cuda")
    target = p.targets.permute(0, 2, 3, 1).contiguous()                   # the reference's [E,ht,wd,2] layout
    c_ref, v_ref, m_ref = og.motion_features(p.poses.double(), p.disps.double(), p.intrinsics.double(), p.ii, p.jj,
                                             target.double())
    coords, valid, motn = glue.motion_features(g.poses, g.disps, g.intrinsics, g.ii, g.jj, target.cuda()[None])
    assert coords.shape == (1,) + tuple(c_ref.shape) and motn.shape == (1,) + tuple(m_ref.shape)
    assert float((coords[0].cpu().double() - c_ref).abs().max()) < 2e-3        # pixels, fp32 vs fp64
    assert torch.equal(valid[0].cpu().double(), v_ref)
    assert float((motn[0].cpu().double() - m_ref).abs().max()) < 2e-3
    assert float(motn.abs().max()) <= 64.0
    # damping: frames 3 and 7 are not sources -> rows follow sorted(unique(ii))
    keep = (p.ii != 3) & (p.ii != 7)
    ii = p.ii[keep].cuda()
    damping = torch.rand(p.disps.shape, generator=torch.Generator().manual_seed(3)) * 1e-3
    eta = glue.damping_eta(damping.cuda(), ii)
    ref = og.damping_eta(damping, p.ii[keep])
    assert eta.shape == ref.shape and float((eta.cpu() - ref).abs().max()) < 1e-9        # 0.2 d + 1e-7: FMA vs mul, add
