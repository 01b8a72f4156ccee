"""Pin oracle/dba.py (the CPU restatement) to the reference's own outputs.

tests/golden/*.npz were produced by oracle/make_golden.py from the UNMODIFIED reference
hislam2/geom/ba.py (BA :32-107, MoBA :110-158) over the reference's lietorch CPU backend.
fp64: the restatement must agree to ~1e-10 (same algorithm, same precision).
fp32: both are fp32 evaluations of the same maths with different rounding order; the bound is
a small multiple of the reference's own fp32-vs-fp64 distance stored next to it.
"""
import numpy as np
import pytest
import torch

import oracle.dba as od
from helpers import input_digest, load_golden, rel_err

CASES = ["tiny_ba", "tiny_ba_backend", "tiny_moba", "correctness", "frontend"]


@pytest.mark.parametrize("name", CASES)
def test_inputs_reproducible(name):
    p, g = load_golden(name)
    assert input_digest(p) == str(g["digest"])


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_fp64(name):
    p, g = load_golden(name)
    poses, disps, infos = od.oracle_ba(*p.ba_args(), dtype=torch.float64, block_diag_quirk=True, return_info=True)
    for k, (dx, _) in enumerate(infos):
        assert rel_err(dx, g["ref_dx_f64"][k]) < 1e-9
    assert rel_err(disps, g["ref_disps_f64"]) < 1e-6          # outputs were stored as fp32
    assert float((poses - torch.from_numpy(g["ref_poses_f64"])).abs().max()) < 1e-6


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_fp32(name):
    p, g = load_golden(name)
    poses, disps, infos = od.oracle_ba(*p.ba_args(), dtype=torch.float32, block_diag_quirk=True, return_info=True)
    for k, (dx, _) in enumerate(infos):
        ref_gap = rel_err(g["ref_dx_f32"][k], g["ref_dx_f64"][k])
        assert rel_err(dx, g["ref_dx_f64"][k]) < 3 * ref_gap + 1e-5
    ref_gap = rel_err(g["ref_disps_f32"], g["ref_disps_f64"])
    assert rel_err(disps, g["ref_disps_f64"]) < 3 * ref_gap + 1e-5


@pytest.mark.parametrize("name", ["tiny_ba", "tiny_ba_backend", "tiny_moba"])
def test_jacobians_match_reference(name):
    """projective_transform (projective_ops.py:44-74) pinned on a stored pixel subsample."""
    p, g = load_golden(name)
    coords, valid, (Ji, Jj, Jz) = od.projective_transform(p.poses, p.disps, p.intrinsics, p.ii, p.jj, jacobian=True)
    sl = (slice(None), slice(1, None, 5), slice(2, None, 7))
    assert torch.equal(valid[sl], torch.from_numpy(g["jac_valid"]))
    for mine, key in ((coords, "jac_coords"), (Ji, "jac_Ji"), (Jj, "jac_Jj"), (Jz, "jac_Jz")):
        assert rel_err(mine[sl], g[key]) < 2e-6, key


# ---- ragged / degenerate graphs: the reference itself run on them (oracle/make_golden_ragged.py) -------------------
import os  # noqa: E402

from cut3r_slam_b200.synthetic import make_problem  # noqa: E402
from helpers import GOLDEN  # noqa: E402
import ragged_cases  # noqa: E402

RAGGED = np.load(os.path.join(GOLDEN, "ragged.npz"))


@pytest.mark.parametrize("name", ragged_cases.REFERENCE_PINNED)
def test_oracle_matches_reference_on_ragged_graphs(name):
    """Keyframes without outgoing / incoming / any edges, duplicate and self edges, edges inside the gauge ...: the
    restatement follows the reference's scatter and its unique(ii) row map on all of them (geom/ba.py:70-92)."""
    p = make_problem(iters=2, **ragged_cases.CASES[name])
    assert input_digest(p) == str(RAGGED[f"{name}/digest"])
    poses, disps, infos = od.oracle_ba(*p.ba_args(), dtype=torch.float64, return_info=True)
    for k, (dx, _) in enumerate(infos):
        assert rel_err(dx, RAGGED[f"{name}/ref_dx_f64"][k]) < 1e-9, k
    assert rel_err(disps, RAGGED[f"{name}/ref_disps_f64"]) < 1e-6
    assert float((poses - torch.from_numpy(RAGGED[f"{name}/ref_poses_f64"])).abs().max()) < 1e-6


@pytest.mark.parametrize("name", ragged_cases.REFERENCE_PINNED_MOBA)
def test_oracle_moba_matches_reference_on_ragged_graphs(name):
    """The reference's MoBA (geom/ba.py:110-158, block_solve as written) on the same graphs."""
    p = make_problem(iters=2, **ragged_cases.CASES[name])
    p.motion_only = True
    poses, disps, infos = od.oracle_ba(*p.ba_args(), dtype=torch.float64, block_diag_quirk=True, return_info=True)
    for k, (dx, _) in enumerate(infos):
        assert rel_err(dx, RAGGED[f"{name}/moba_dx_f64"][k]) < 1e-9, k
    assert float((poses - torch.from_numpy(RAGGED[f"{name}/moba_poses_f64"])).abs().max()) < 1e-6
    assert torch.equal(disps, p.disps)


@pytest.mark.parametrize("name", ragged_cases.REFERENCE_PINNED_SIM3)
def test_oracle_sim3_ba_matches_reference_on_ragged_graphs(name):
    """The reference's BA with lietorch.Sim3 poses (the 7-dof linearisation of pgba) on the same graphs."""
    import oracle.pgba as op
    p = make_problem(iters=2, **ragged_cases.CASES[name])
    poses8 = torch.cat([p.poses, ragged_cases.sim3_scales(p.poses.shape[0], p.t0)], -1)
    poses, disps, infos = op.oracle_pgba(poses8, p.disps, p.intrinsics, p.targets, p.weights, p.eta, p.ii, p.jj, None, None,
                                         None, None, p.t0, p.t1, 2, p.lm, p.ep, 0.0, dtype=torch.float64, return_info=True)
    for k in range(2):
        assert rel_err(infos[k][0], RAGGED[f"{name}/sim3_dx_f64"][k]) < 1e-9, k
    assert rel_err(disps.double(), RAGGED[f"{name}/sim3_disps_f64"]) < 1e-6
    assert rel_err(poses.double(), RAGGED[f"{name}/sim3_poses_f64"]) < 1e-6
