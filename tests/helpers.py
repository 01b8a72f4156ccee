"""Shared helpers for the parity tests (CPU side)."""
import hashlib
import os

import numpy as np
import torch

from cut3r_slam_b200.synthetic import Problem, make_problem

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

GOLDEN_KW = {  # must match oracle/make_golden.py CASES
    "tiny_ba":         dict(P=6, E=14, ht=24, wd=32, iters=2, lm=1e-4, ep=0.1, t0=1, r=2, seed=11),
    "tiny_ba_backend": dict(P=7, E=24, ht=24, wd=32, iters=3, lm=1e-5, ep=1e-2, t0=2, r=2, seed=12),
    "tiny_moba":       dict(P=6, E=18, ht=24, wd=32, iters=2, lm=1e-4, ep=0.1, t0=1, r=2, motion_only=True, seed=13),
}


def input_digest(p):
    h = hashlib.sha256()
    for t in (p.poses, p.disps, p.intrinsics, p.targets, p.weights, p.eta, p.ii, p.jj):
        h.update(t.contiguous().cpu().numpy().tobytes())
    return h.hexdigest()


def load_golden(name):
    """Returns (Problem built from the stored inputs or regenerated from the seed, npz record)."""
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    if "poses" in g.files:
        t = lambda k: torch.from_numpy(g[k])
        P = g["poses"].shape[0]
        p = Problem(poses=t("poses"), disps=t("disps"), intrinsics=t("intrinsics"),
                    disps_sens=torch.zeros_like(t("disps")), targets=t("targets"), weights=t("weights"),
                    eta=t("eta"), ii=t("ii"), jj=t("jj"), t0=int(g["t0"]), t1=int(g["t1"]),
                    iterations=int(g["iterations"]), lm=float(g["lm"]), ep=float(g["ep"]),
                    motion_only=bool(g["motion_only"]), gt_poses=t("poses"), gt_disps=t("disps"))
    else:
        p = make_problem(name)
    return p, g


def rel_err(a, b):
    """max |a-b| / max |b| -- the 'max rel. err' of BASELINE.json north_star."""
    a = torch.as_tensor(a, dtype=torch.float64)
    b = torch.as_tensor(b, dtype=torch.float64)
    return float((a - b).abs().max() / b.abs().max().clamp(min=1e-30))
