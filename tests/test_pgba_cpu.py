"""oracle/pgba.py + oracle/sim3.py against tests/golden/pgba.npz = outputs of the UNMODIFIED reference: hislam2/geom/ba.py
`BA` run with lietorch.Sim3 poses, and hislam2/pgo_buffer.py `global_relative_posesim3_constraints`."""
import os

import numpy as np
import torch

import oracle.pgba as op
from helpers import GOLDEN, rel_err


def _load():
    g = np.load(os.path.join(GOLDEN, "pgba.npz"))
    t = lambda k: torch.from_numpy(g[k])
    return g, t


def test_sim3_ba_step_matches_reference():
    g, t = _load()
    kw = dict(t0=int(g["t0"]), t1=int(g["t1"]), iterations=int(g["iterations"]), lm=float(g["lm"]), ep=float(g["ep"]))
    for dt, tag, tol in ((torch.float64, "f64", 1e-9), (torch.float32, "f32", 2e-3)):
        poses, disps, infos = op.oracle_pgba(t("poses8"), t("disps"), t("intrinsics"), t("targets"), t("weights"), t("eta"),
                                             t("ii"), t("jj"), None, None, None, None, kw["t0"], kw["t1"], kw["iterations"],
                                             kw["lm"], kw["ep"], 0.0, dtype=dt, return_info=True)
        for k in range(kw["iterations"]):
            assert rel_err(infos[k][0], g[f"ref_dx_{tag}"][k]) < tol, (tag, k)
        assert rel_err(disps.double(), g[f"ref_disps_{tag}"]) < max(tol, 1e-6)
        assert rel_err(poses.double(), g[f"ref_poses_{tag}"]) < max(tol, 1e-6)


def test_relative_pose_blocks_match_reference():
    g, t = _load()
    Hsp, vsp = op.rel_pose_blocks(t("rel_ii"), t("rel_jj"), t("poses8"), t("rel_poses"), t("rel_infos"), float(g["pw"]))
    # numeric Jacobians with eps = 1e-8 in fp64 carry ~1e-8 relative noise of their own; blocks are fp32
    assert rel_err(Hsp, g["ref_Hsp"]) < 1e-5
    assert rel_err(vsp, g["ref_vsp"]) < 1e-5
