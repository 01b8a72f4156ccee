"""oracle/relpose.py against the reference's own PGOBuffer.add_rel_poses (tests/golden/relpose.npz, made by
oracle/make_golden_relpose.py from hislam2/pgo_buffer.py:85-123 run unmodified)."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import oracle.relpose as orp  # noqa: E402
from oracle.make_golden_relpose import CASES, make_case  # noqa: E402

GOLD = np.load(os.path.join(ROOT, "tests", "golden", "relpose.npz"))


@pytest.mark.parametrize("name", sorted(CASES))
def test_add_rel_poses_matches_reference(name):
    p, target, weight = make_case(CASES[name])
    d = torch.float64
    ii, jj, rel, cov, vp = orp.add_rel_poses(p.poses.to(d), p.disps.to(d), p.intrinsics.to(d), p.ii, p.jj, target.to(d),
                                             weight.to(d))
    assert ii.numel() == int(GOLD[f"{name}_n"]) < p.ii.numel()            # the masked edge was dropped
    assert np.array_equal(ii.numpy(), GOLD[f"{name}_ii"]) and np.array_equal(jj.numpy(), GOLD[f"{name}_jj"])
    ref_c = torch.from_numpy(GOLD[f"{name}_covs_f64"])
    assert float((rel - torch.from_numpy(GOLD[f"{name}_poses_f64"])).abs().max()) < 1e-12
    assert float(((cov - ref_c).abs() / ref_c.abs()).max()) < 1e-10
    assert float((vp - torch.from_numpy(GOLD[f"{name}_valid_f64"])).abs().max()) < 1e-7   # stored through fp32 upstream
