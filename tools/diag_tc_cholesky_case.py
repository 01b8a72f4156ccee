"""Developer diagnostic: disps error vs the fp64 oracle of test_tensor_core_cholesky_agrees' problems for several split
counts, with the cooperative and the tcgen05-panel Cholesky, refinement off / on."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import oracle.dba as od
from cut3r_slam_b200 import droid_backends as db
from cut3r_slam_b200.synthetic import make_problem
from helpers import rel_err

for P, E in [(40, 300), (60, 500)]:
    p = make_problem(P=P, E=E, ht=20, wd=24, iters=2, t0=2, r=4, seed=81)
    _, d64 = od.oracle_ba(*p.ba_args(), dtype=torch.float64)
    _, d32 = od.oracle_ba(*p.ba_args(), dtype=torch.float32)
    print(f"P={P} E={E} reference fp32 gap {rel_err(d32, d64):.2e}")
    for S in (0, 1, 2, 3, 4, 8):
        row = []
        for refine in (0, 1):
            for tc in (0, 64):
                db.set_option("splits", S); db.set_option("refine", refine); db.set_option("cholesky_tc", tc)
                g = p.to("cuda")
                db.ba(*g.ba_args())
                row.append(f"refine={refine} tc={tc}: {rel_err(g.disps.cpu(), d64):.2e}")
        print(f"  splits={S}: " + " | ".join(row))
db.set_option("splits", 0); db.set_option("refine", 1); db.set_option("cholesky_tc", 0)
