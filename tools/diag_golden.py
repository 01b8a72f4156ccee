"""Developer diagnostic: golden-vector errors per kernel option (tensor cores on/off, Cholesky path)."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from cut3r_slam_b200 import droid_backends as db
from helpers import load_golden, rel_err

for name in ["tiny_ba", "tiny_ba_backend", "tiny_moba", "correctness", "frontend"]:
    p, g = load_golden(name)
    for tc in (1, 0):
        for ml in (0, 1):
            db.set_option("tensor_cores", tc); db.set_option("multilaunch_cholesky", ml)
            db.set_option("block_diag_quirk", 1 if p.motion_only else 0)
            q = p.to("cuda")
            errs = []
            for k in range(p.iterations):
                dx, dz = db.ba(q.poses, q.disps, q.intrinsics, q.disps_sens, q.targets, q.weights, q.eta, q.ii, q.jj,
                               q.t0, q.t1, 1, q.lm, q.ep, q.motion_only)
                errs.append(rel_err(dx.cpu(), g["ref_dx_f64"][k]))
            gaps = [rel_err(g["ref_dx_f32"][k], g["ref_dx_f64"][k]) for k in range(p.iterations)]
            print(f"{name:16s} tc={tc} multilaunch={ml} disps {rel_err(q.disps.cpu(), g['ref_disps_f64']):.2e} "
                  f"(ref f32 {rel_err(g['ref_disps_f32'], g['ref_disps_f64']):.2e}) dx " +
                  " ".join(f"{e:.1e}/{gp:.1e}" for e, gp in zip(errs, gaps)))
db.set_option("tensor_cores", 1); db.set_option("multilaunch_cholesky", 0); db.set_option("block_diag_quirk", 0)
