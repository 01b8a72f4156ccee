"""GPU parity of the JDSA ops (proj_trans, bi_inter, and -- once built -- the fused JDSA step) against
oracle/jdsa.py, which tests/test_jdsa_cpu.py pins to the reference's own JDSA / schur_solve_mono_prior."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import oracle.jdsa as oj  # noqa: E402
from cut3r_slam_b200.synthetic import make_problem  # noqa: E402
from oracle.make_golden_jdsa import CASES, make_case  # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(ROOT, "tests", "golden", "jdsa.npz"))


def _db():
    from cut3r_slam_b200 import droid_backends
    return droid_backends


def _rel(a, b):
    a, b = torch.as_tensor(a, dtype=torch.float64), torch.as_tensor(b, dtype=torch.float64)
    return float((a - b).abs().max() / b.abs().max())


@pytest.mark.parametrize("kw", [
    dict(P=6, E=20, ht=12, wd=16, iters=1, t0=1, r=2, seed=41),
    dict(P=9, E=40, ht=43, wd=77, iters=1, t0=1, r=3, seed=22),            # HW not a multiple of the tile
    dict(P=30, E=600, ht=20, wd=24, iters=1, t0=2, r=4, seed=83),          # out-degree > 64: chunked edge loop
])
def test_proj_trans_matches_oracle(kw):
    db = _db()
    p = make_problem(**kw)
    g = p.to("cuda")
    C, w = db.proj_trans(g.poses, g.disps, g.intrinsics, g.targets, g.weights, g.ii, g.jj)
    d = torch.float64
    Cr, wr = oj.proj_trans(p.poses.to(d), p.disps.to(d), p.intrinsics.to(d), p.targets.to(d), p.weights.to(d), p.ii, p.jj)
    M = int(torch.unique(p.ii).numel())
    assert C.shape == w.shape == (M,) + tuple(p.disps.shape[1:])
    assert _rel(C.cpu().reshape(M, -1), Cr) < 2e-5 and _rel(w.cpu().reshape(M, -1), wr) < 2e-5
    # it is the depth block of `ba` without eta: consistent with the linearisation's Q = 1 / (C + eta + 1e-7)
    with pytest.raises(RuntimeError):
        db.proj_trans(g.poses, g.disps, g.intrinsics, g.targets[:-1].contiguous(), g.weights, g.ii, g.jj)


def test_bi_inter_matches_oracle():
    db = _db()
    gen = torch.Generator().manual_seed(5)
    M, hs, ws, ht, wd = 5, 3, 4, 43, 77
    scales = torch.randn(M, hs, ws, generator=gen)
    ys, xs = torch.meshgrid(torch.linspace(0, hs - 1 - 1e-6, ht), torch.linspace(0, ws - 1 - 1e-6, wd), indexing="ij")
    grid = torch.stack((xs, ys), -1)[None].expand(M, -1, -1, -1).contiguous()     # ba.py:164-166
    vals, J = db.bi_inter(scales.cuda(), grid.cuda())
    vr, Jr = oj.bi_inter(scales.double(), grid.double())
    assert vals.shape == (M, ht, wd) and J.shape == (M, ht, wd, hs * ws)
    assert _rel(vals.cpu(), vr) < 1e-6 and _rel(J.cpu(), Jr) < 1e-6
    assert float((J.sum(-1) - 1).abs().max()) < 1e-6
