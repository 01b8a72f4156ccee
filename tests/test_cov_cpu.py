"""oracle.dba.depth_cov against the depth covariance the reference's own BA returns (tests/golden/cov.npz, made by
oracle/make_golden_cov.py from hislam2/geom/ba.py:98-107 / chol.py:74-76 run unmodified)."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import oracle.dba as od  # noqa: E402

GOLD = np.load(os.path.join(ROOT, "tests", "golden", "cov.npz"))


@pytest.mark.parametrize("name", ["tiny_ba", "tiny_ba_backend"])
def test_depth_cov_matches_reference(name):
    G = np.load(os.path.join(ROOT, "tests", "golden", f"{name}.npz"))
    t = lambda k: torch.from_numpy(G[k])
    cov = od.depth_cov(t("poses"), t("disps"), t("intrinsics"), None, t("targets"), t("weights"), t("eta"), t("ii"),
                       t("jj"), int(G["t0"]), int(G["t1"]), float(G["lm"]), float(G["ep"]))
    ref = torch.from_numpy(GOLD[f"{name}_dzcov_f64"])
    assert cov.shape == ref.shape
    assert float(((cov - ref).abs() / ref.abs()).max()) < 1e-10
    assert float(cov.min()) > 0                                       # Q plus a positive semi-definite term


RAGGED = ["sink_and_source", "isolated_keyframe", "self_edges", "random_101"]     # oracle/make_golden_cov.py RAGGED


@pytest.mark.parametrize("name", RAGGED)
def test_depth_cov_matches_reference_on_ragged_graphs(name):
    """Rows follow sorted(unique(ii)) (frame 4 of `sink_and_source` has no row), an isolated keyframe contributes only
    its damping, a self edge adds all four of its blocks to one diagonal block."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ragged_cases
    from cut3r_slam_b200.synthetic import make_problem
    p = make_problem(iters=1, **ragged_cases.CASES[name])
    cov = od.depth_cov(p.poses, p.disps, p.intrinsics, None, p.targets, p.weights, p.eta, p.ii, p.jj, p.t0, p.t1, p.lm, p.ep)
    ref = torch.from_numpy(GOLD[f"ragged_{name}_dzcov_f64"])
    assert cov.shape == ref.shape
    assert float(((cov - ref).abs() / ref.abs()).max()) < 1e-9
