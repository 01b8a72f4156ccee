"""Ragged and degenerate factor graphs through the CUDA `ba` against the fp64 restatement of hislam2/geom/ba.py:32-107.

The generated graphs of the other parity files are regular (every keyframe has neighbours in both directions).  The
reference's scatter (`safe_scatter_add_mat` / `safe_scatter_add_vec`, geom/chol.py:22-31) and its `unique(ii)` row map
(geom/ba.py:70-75) accept anything: keyframes without outgoing or incoming edges, keyframes without edges at all,
duplicate edges, edges between two gauge frames, a window that stops before the end of the buffers, out-degrees on both
sides of every kernel-selection threshold in one graph, grids smaller than one pixel tile.  Each case: the reduced system
block by block (well conditioned: 2e-6), then dx / dz by the rule of test_ba_gpu.py (a small multiple of the reference
algorithm's own fp32-vs-fp64 gap), then a two-iteration call.
"""
import copy

import pytest
import torch

import oracle.dba as od
import oracle.se3 as ose3
from cut3r_slam_b200.synthetic import make_problem
from helpers import rel_err

pytestmark = pytest.mark.gpu

MULT = 3.0


def _db():
    from cut3r_slam_b200 import droid_backends
    return droid_backends


from ragged_cases import CASES, chain as _chain, degree_mix as _degree_mix   # noqa: E402,F401


def _oracle_step(p, dtype):
    t = lambda x: x.to(dtype)
    lin = od.linearize(t(p.targets.permute(0, 2, 3, 1)), t(p.weights.permute(0, 2, 3, 1)), t(p.eta), t(p.poses),
                       t(p.disps), t(p.intrinsics), p.ii, p.jj, p.t0)
    return od.schur_solve(lin["H"], lin["E"], lin["C"], lin["v"], lin["w"], ep=p.ep, lm=p.lm)


def _reference_fp32_gap(p, poses64, disps64, draws=3):
    """The reference algorithm's own fp32-vs-fp64 error after the whole call, as the largest of `draws` runs of the fp32
    restatement that differ only in the order of the edge list (= summation order of the scatter).  One run is one
    draw of rounding noise that the second iteration amplifies at weakly observed frames.  Measured on the radius-1
    variant of `degree_thresholds` (chain ends with ONE edge, 12 x 20 grid; profiles/r2_parity_margins.md, last section):
    eight edge orders give 1.3e-4 .. 4.0e-4 on disps, every CUDA path -- the plain fp32 SIMT kernels included -- lands at
    4.7e-4 .. 9.9e-4 after being BELOW the reference's gap after the first iteration (1.0e-5 against 1.5e-5).  That
    variant sits on the edge of the 3x rule and is kept out of the suite for that reason alone."""
    E = int(p.ii.numel())
    gd = gp = 0.0
    for s in range(draws):
        perm = torch.arange(E) if s == 0 else torch.randperm(E, generator=torch.Generator().manual_seed(s))
        q = copy.copy(p)
        q.ii, q.jj, q.targets, q.weights = p.ii[perm], p.jj[perm], p.targets[perm].contiguous(), p.weights[perm].contiguous()
        poses32, disps32 = od.oracle_ba(*q.ba_args(), dtype=torch.float32)
        gd = max(gd, rel_err(disps32, disps64))
        gp = max(gp, float(ose3.tangent_update(poses32.double(), poses64.double()).abs().max()))
    return gd, gp


@pytest.mark.parametrize("name", list(CASES))
def test_ragged_graph_matches_the_restatement(name):
    db = _db()
    p = make_problem(iters=2, **CASES[name])
    sol = _oracle_step(p, torch.float64)
    sol32 = _oracle_step(p, torch.float32)
    db.set_option("keep_system", 1)
    try:
        g = p.to("cuda")
        dx, dz = db.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj,
                       g.t0, g.t1, 1, g.lm, g.ep, False)
        S, rhs, dims = db.debug_system()
    finally:
        db.set_option("keep_system", 0)
    assert rel_err(torch.tril(S.double()), torch.tril(sol["S"])) < 2e-6
    assert rel_err(rhs, sol["vs"]) < 2e-6
    gap_dx = rel_err(sol32["dx"], sol["dx"])
    gap_dz = rel_err(sol32["dz"], sol["dz"])
    assert rel_err(dx.cpu(), sol["dx"]) < MULT * gap_dx + 1e-5, gap_dx
    assert rel_err(dz.cpu().reshape(sol["dz"].shape), sol["dz"]) < MULT * gap_dz + 1e-5, gap_dz
    # the whole call, two iterations, from the start point again
    poses64, disps64 = od.oracle_ba(*p.ba_args(), dtype=torch.float64)
    gap_d, gap_p = _reference_fp32_gap(p, poses64, disps64)
    h = p.to("cuda")
    db.ba(*h.ba_args())
    assert rel_err(h.disps.cpu(), disps64) < max(1e-4, MULT * gap_d), gap_d
    assert float(ose3.tangent_update(h.poses.cpu().double(), poses64.double()).abs().max()) < max(1e-4, MULT * gap_p), gap_p
    # frames that are no edge's source keep their disparities bit for bit (geom/ba.py:70-75: only unique(ii) rows move)
    src = torch.zeros(p.poses.shape[0], dtype=torch.bool)
    src[p.ii] = True
    assert torch.equal(h.disps.cpu()[~src], p.disps[~src])
    assert torch.equal(h.poses.cpu()[: p.t0], p.poses[: p.t0])


def test_window_that_stops_before_the_end_of_the_buffers():
    """t1 < N: frames behind the window keep poses and disparities, the system has t1 - t0 blocks (geom/ba.py:36-38)."""
    db = _db()
    full = make_problem(P=10, ht=20, wd=28, t0=1, seed=81, iters=2, edges=_chain(7, 2))     # edges among frames 0..6 only
    assert int(full.jj.max()) == 6
    full.t1 = 7
    poses64, disps64 = od.oracle_ba(*full.ba_args(), dtype=torch.float64)
    g = full.to("cuda")
    dx, dz = db.ba(*g.ba_args())
    assert dx.shape[0] == 6
    assert rel_err(g.disps.cpu(), disps64) < 1e-4
    assert float(ose3.tangent_update(g.poses.cpu().double(), poses64.double()).abs().max()) < 1e-4
    assert torch.equal(g.poses.cpu()[7:], full.poses[7:]) and torch.equal(g.disps.cpu()[7:], full.disps[7:])


def test_edges_without_weight():
    """weights = 0 on whole edges and on every edge of one keyframe: C = eta there, dz = Q w stays finite (chol.py:58-70)."""
    db = _db()
    p = make_problem(P=7, ht=20, wd=28, t0=1, seed=82, iters=2, edges=_chain(7, 2))
    p.weights[3] = 0
    p.weights[p.ii == 4] = 0
    poses64, disps64 = od.oracle_ba(*p.ba_args(), dtype=torch.float64)
    g = p.to("cuda")
    db.ba(*g.ba_args())
    assert bool(torch.isfinite(g.disps).all()) and bool(torch.isfinite(g.poses).all())
    assert rel_err(g.disps.cpu(), disps64) < 1e-4
    assert float(ose3.tangent_update(g.poses.cpu().double(), poses64.double()).abs().max()) < 1e-4


RAGGED_SUBSET = ["sink_and_source", "isolated_keyframe", "duplicate_edges", "edges_inside_the_gauge", "degree_thresholds",
                 "grid_below_one_tile", "self_edges", "random_101", "random_107"]


@pytest.mark.parametrize("name", RAGGED_SUBSET)
def test_depth_cov_on_ragged_graphs(name):
    """dzcov of the same graphs (geom/ba.py:98-107, chol.py:74-76): rows follow unique(ii), not the frame index."""
    db = _db()
    p = make_problem(iters=1, **CASES[name])
    g = p.to("cuda")
    cov = db.depth_cov(g.poses, g.disps, g.intrinsics, None, g.targets, g.weights, g.eta, g.ii, g.jj, p.t0, p.t1, p.lm, p.ep)
    ref = od.depth_cov(p.poses, p.disps, p.intrinsics, None, p.targets, p.weights, p.eta, p.ii, p.jj, p.t0, p.t1, p.lm, p.ep)
    got = cov.cpu().double().reshape(ref.shape[0], -1)
    assert got.shape == ref.reshape(ref.shape[0], -1).shape
    assert float(((got - ref.reshape(ref.shape[0], -1)).abs() / ref.reshape(ref.shape[0], -1).abs()).max()) < 1e-4


@pytest.mark.parametrize("name", RAGGED_SUBSET)
def test_pgba_on_ragged_graphs(name):
    """The Sim3 path (7-dof blocks, its own plan) on the same graphs, scales != 1, against oracle/pgba.py in fp64."""
    import oracle.pgba as op
    db = _db()
    p = make_problem(iters=2, **CASES[name])
    P = p.poses.shape[0]
    s = 1.0 + 0.05 * (2 * torch.rand(P, 1, generator=torch.Generator().manual_seed(3)) - 1)
    s[: p.t0] = 1.0
    poses8 = torch.cat([p.poses, s], -1)
    a = (poses8, p.disps, p.intrinsics, p.targets, p.weights, p.eta, p.ii, p.jj, None, None, None, None, p.t0, p.t1, 1,
         p.lm, p.ep, 0.0)
    p64, d64, i64 = op.oracle_pgba(*a, dtype=torch.float64, return_info=True)
    p32, d32, i32 = op.oracle_pgba(*a, dtype=torch.float32, return_info=True)
    e = torch.zeros(0, dtype=torch.int64, device="cuda")
    Pg, Dg = poses8.cuda(), p.disps.cuda()
    dx, dz = db.pgba(Pg, Dg, p.intrinsics.cuda(), p.targets.cuda(), p.weights.cuda(), p.eta.cuda(), p.ii.cuda(), p.jj.cuda(),
                     e, e, torch.zeros(0, 7, device="cuda"), torch.zeros(0, 7, 7, device="cuda"), p.t0, p.t1, 1, p.lm, p.ep)
    gap = rel_err(i32[0][0], i64[0][0])
    assert rel_err(dx.cpu(), i64[0][0]) < MULT * gap + 1e-5, gap
    assert rel_err(Dg.cpu(), d64) < max(1e-4, MULT * rel_err(d32, d64))
    src = torch.zeros(P, dtype=torch.bool)
    src[p.ii] = True
    assert torch.equal(Dg.cpu()[~src], p.disps[~src])
    assert torch.equal(Pg.cpu()[: p.t0], poses8[: p.t0])


import os  # noqa: E402

import numpy as np  # noqa: E402

from helpers import GOLDEN  # noqa: E402
from ragged_cases import REFERENCE_PINNED  # noqa: E402

RAGGED_REF = np.load(os.path.join(GOLDEN, "ragged.npz"))


@pytest.mark.parametrize("name", REFERENCE_PINNED)
def test_ragged_graph_matches_the_reference_itself(name):
    """tests/golden/ragged.npz = the UNMODIFIED reference BA run on these graphs (oracle/make_golden_ragged.py): the
    two-iteration CUDA call against its fp64 outputs, within a small multiple of the reference's OWN fp32-vs-fp64 gap."""
    db = _db()
    R = lambda k: torch.from_numpy(RAGGED_REF[f"{name}/{k}"])
    p = make_problem(iters=2, **CASES[name])
    g = p.to("cuda")
    dxs = []
    for _ in range(2):
        dx, dz = db.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.t0, g.t1, 1,
                       g.lm, g.ep, False)
        dxs.append(dx.cpu())
    gap0 = rel_err(R("ref_dx_f32")[0], R("ref_dx_f64")[0])
    assert rel_err(dxs[0], R("ref_dx_f64")[0]) < MULT * gap0 + 1e-5, gap0
    gap_d = rel_err(R("ref_disps_f32"), R("ref_disps_f64"))
    assert rel_err(g.disps.cpu(), R("ref_disps_f64")) < max(1e-4, MULT * gap_d), gap_d
    assert float((g.poses.cpu().double() - R("ref_poses_f64").double()).abs().max()) < 1e-4


@pytest.mark.parametrize("name", ["sink_and_source", "isolated_keyframe", "self_edges", "random_101"])
def test_depth_cov_matches_the_reference_itself_on_ragged_graphs(name):
    """tests/golden/cov.npz `ragged_*` = the third output of the UNMODIFIED reference BA on these graphs
    (oracle/make_golden_cov.py); bound: 1e-4 per pixel or 3x the reference's own fp32-vs-fp64 gap."""
    db = _db()
    cov_gold = np.load(os.path.join(GOLDEN, "cov.npz"))
    ref = torch.from_numpy(cov_gold[f"ragged_{name}_dzcov_f64"])
    ref32 = torch.from_numpy(cov_gold[f"ragged_{name}_dzcov_f32"]).double()
    p = make_problem(iters=1, **CASES[name])
    g = p.to("cuda")
    cov = db.depth_cov(g.poses, g.disps, g.intrinsics, None, g.targets, g.weights, g.eta, g.ii, g.jj, p.t0, p.t1, p.lm, p.ep)
    got = cov.cpu().double().reshape(ref.shape)
    gap = float(((ref32 - ref).abs() / ref.abs()).max())
    assert float(((got - ref).abs() / ref.abs()).max()) < max(1e-4, MULT * gap), gap


from ragged_cases import GPU_SIM3, REFERENCE_PINNED_MOBA, sim3_scales  # noqa: E402


@pytest.mark.parametrize("name", REFERENCE_PINNED_MOBA)
def test_ragged_moba_matches_the_reference_itself(name):
    """motion_only on the same graphs against the reference's MoBA (geom/ba.py:110-158; block_solve as written)."""
    db = _db()
    R = lambda k: torch.from_numpy(RAGGED_REF[f"{name}/{k}"])
    p = make_problem(iters=2, **CASES[name])
    g = p.to("cuda")
    db.set_option("block_diag_quirk", 1)
    try:
        for k in range(2):
            dx, _ = db.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.t0, g.t1, 1,
                          g.lm, g.ep, True)
            gap = rel_err(R("moba_dx_f32")[k], R("moba_dx_f64")[k])
            assert rel_err(dx.cpu(), R("moba_dx_f64")[k]) < MULT * gap + 1e-5, (k, gap)
    finally:
        db.set_option("block_diag_quirk", 0)
    assert float((g.poses.cpu().double() - R("moba_poses_f64").double()).abs().max()) < 1e-4
    assert torch.equal(g.disps.cpu(), p.disps)


@pytest.mark.parametrize("name", GPU_SIM3)
def test_ragged_sim3_ba_matches_the_reference_itself(name):
    """pgba without relative-pose factors on the same graphs against the reference's BA run with lietorch.Sim3 poses.
    These cases found the fp32 Sim3 retraction: scale updates of 1e-6 .. 1e-5 leave (e^sigma - 1) / sigma two digits in
    fp32, the state after one iteration was 2.8e-4 off and the SECOND update 4.7x the reference's own fp32 gap on
    `isolated_keyframe` (tools/sim3_state_diag.py).  With Exp(dx) in fp64 (k_pg_retract) both iterations sit at
    0.02 - 0.12x that gap (tools/sim3_ragged_diag.py)."""
    db = _db()
    R = lambda k: torch.from_numpy(RAGGED_REF[f"{name}/{k}"])
    p = make_problem(iters=2, **CASES[name])
    P8 = torch.cat([p.poses, sim3_scales(p.poses.shape[0], p.t0)], -1).cuda()
    D = p.disps.cuda()
    e = torch.zeros(0, dtype=torch.int64, device="cuda")
    a = [x.cuda() for x in (p.intrinsics, p.targets, p.weights, p.eta, p.ii, p.jj)]
    for k in range(2):
        dx, dz = db.pgba(P8, D, *a, e, e, torch.zeros(0, 7, device="cuda"), torch.zeros(0, 7, 7, device="cuda"), p.t0, p.t1, 1,
                         p.lm, p.ep)
        gap = rel_err(R("sim3_dx_f32")[k], R("sim3_dx_f64")[k])
        assert rel_err(dx.cpu(), R("sim3_dx_f64")[k]) < MULT * gap + 1e-5, (k, gap)
    gap_d = rel_err(R("sim3_disps_f32"), R("sim3_disps_f64"))
    assert rel_err(D.cpu(), R("sim3_disps_f64")) < max(1e-4, MULT * gap_d), gap_d
    # translations reach ~8 units on the longer trajectories: a rotation / scale error at the level of the reference's own
    # fp32 noise moves them by more than an absolute 1e-4, so the poses follow the same rule as dx and disps
    gap_p = float((R("sim3_poses_f32").double() - R("sim3_poses_f64").double()).abs().max())
    err_p = float((P8.cpu().double() - R("sim3_poses_f64").double()).abs().max())
    assert err_p < max(1e-4, MULT * gap_p), (err_p, gap_p)
