"""Developer diagnostic: split the dx error of one BA iteration into (a) assembly error of the reduced
system (solve the GPU's S, rhs in fp64 on the CPU) and (b) the device solver's own error."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import oracle.dba as od
from cut3r_slam_b200 import droid_backends as db
from cut3r_slam_b200.synthetic import make_problem
from helpers import rel_err

cases = [dict(P=7, E=24, ht=24, wd=32, iters=1, t0=2, r=2, seed=21),
         dict(P=9, E=40, ht=43, wd=77, iters=1, t0=1, r=3, seed=22),
         dict(P=10, E=30, ht=17, wd=19, iters=1, t0=3, r=2, seed=23),
         "correctness", "frontend"]
for kw in cases:
    p = make_problem(kw) if isinstance(kw, str) else make_problem(**kw)
    a = list(p.ba_args()); a[11] = 1
    lin = od.linearize(p.targets.permute(0, 2, 3, 1).double(), p.weights.permute(0, 2, 3, 1).double(),
                       p.eta.double(), p.poses.double(), p.disps.double(), p.intrinsics.double(), p.ii, p.jj, p.t0)
    sol = od.schur_solve(lin["H"], lin["E"], lin["C"], lin["v"], lin["w"], ep=p.ep, lm=p.lm)
    db.set_option("keep_system", 1)
    g = p.to("cuda")
    ga = list(g.ba_args()); ga[11] = 1
    dx, dz = db.ba(*ga)
    S, rhs, dims = db.debug_system()
    db.set_option("keep_system", 0)
    S64 = torch.tril(S.double()); S64 = S64 + torch.tril(S64, -1).T
    dx_fromS = torch.linalg.solve(S64, rhs.double())
    ref = sol["dx"].reshape(-1)
    cond = torch.linalg.cond(S64).item()
    print(f"{str(kw)[:40]:40s} n={S.shape[0]:4d} cond={cond:.2e}  S err {rel_err(torch.tril(S.double()), torch.tril(sol['S'])):.2e}  "
          f"dx total {rel_err(dx.cpu().reshape(-1), ref):.2e}  assembly-only {rel_err(dx_fromS, ref):.2e}  solver-only {rel_err(dx.cpu().reshape(-1), dx_fromS):.2e}")
