"""The memory-light oracle (oracle/lowmem.py: one source keyframe at a time, no dense E) against the dense
restatement oracle/dba.py `BA`, which tests/test_oracle_golden.py pins to the unmodified reference."""
import pytest
import torch

import oracle.dba as od
import oracle.lowmem as ol
from cut3r_slam_b200.synthetic import make_problem
from helpers import rel_err


@pytest.mark.parametrize("kw", [
    dict(P=8, E=24, ht=24, wd=32, iters=2, t0=2, r=2, seed=11),
    dict(P=12, E=70, ht=17, wd=19, iters=2, t0=1, r=3, seed=12, lm=1e-5, ep=1e-2),
    dict(P=10, E=30, ht=20, wd=28, iters=1, t0=3, r=2, seed=13),      # fixed frames with outgoing edges
])
def test_lowmem_equals_dense_fp64(kw):
    p = make_problem(**kw)
    ref_p, ref_d, infos = od.oracle_ba(*p.ba_args(), dtype=torch.float64, return_info=True)
    low_p, low_d, linfo = ol.oracle_ba_lowmem(*p.ba_args(), dtype=torch.float64, return_info=True)
    for (dx, dz), (ldx, ldz, S, vs) in zip(infos, linfo):
        assert rel_err(ldx, dx) < 1e-9
        assert rel_err(ldz, dz) < 1e-9
    assert rel_err(low_d, ref_d) < 1e-9 and rel_err(low_p, ref_p) < 1e-9


def test_lowmem_reduced_system_equals_dense():
    p = make_problem(P=9, E=40, ht=16, wd=24, iters=1, t0=2, r=3, seed=14)
    f64 = torch.float64
    tg, wg = p.targets.permute(0, 2, 3, 1).to(f64), p.weights.permute(0, 2, 3, 1).to(f64)
    lin = od.linearize(tg, wg, p.eta.to(f64), p.poses.to(f64), p.disps.to(f64), p.intrinsics.to(f64), p.ii, p.jj, p.t0)
    sol = od.schur_solve(lin["H"], lin["E"], lin["C"], lin["v"], lin["w"], ep=p.ep, lm=p.lm)
    _, _, info = ol.BA_lowmem(tg, wg, p.eta.to(f64), p.poses.to(f64), p.disps.to(f64), p.intrinsics.to(f64), p.ii,
                              p.jj, fixedp=p.t0, ep=p.ep, lm=p.lm, chunk=3, return_info=True)
    assert rel_err(info["S"], sol["S"]) < 1e-12
    assert rel_err(info["vs"], sol["vs"]) < 1e-12


def test_lowmem_fp32_is_another_draw_of_the_same_noise():
    p = make_problem(P=8, E=24, ht=24, wd=32, iters=1, t0=2, r=2, seed=15)
    _, _, i64 = od.oracle_ba(*p.ba_args(), dtype=torch.float64, return_info=True)
    _, _, i32 = od.oracle_ba(*p.ba_args(), dtype=torch.float32, return_info=True)
    _, _, l32 = ol.oracle_ba_lowmem(*p.ba_args(), dtype=torch.float32, return_info=True)
    gap_dense = rel_err(i32[0][0], i64[0][0])
    gap_low = rel_err(l32[0][0], i64[0][0])
    assert gap_low < 10 * gap_dense + 1e-6 and gap_dense < 10 * gap_low + 1e-6
