"""err / bound of every assertion of tests/test_ba_ragged_gpu.py::test_ragged_graph_matches_the_restatement (how far each
ragged case sits from its bound; run on the GPU box)."""
import sys, torch
sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import importlib.util
spec = importlib.util.spec_from_file_location('t', 'tests/test_ba_ragged_gpu.py'); t = importlib.util.module_from_spec(spec); spec.loader.exec_module(t)
import oracle.dba as od
import oracle.se3 as ose3
from helpers import rel_err
from cut3r_slam_b200.synthetic import make_problem
from cut3r_slam_b200 import droid_backends as db

worst = 0.0
for name, kw in t.CASES.items():
    p = make_problem(iters=2, **kw)
    sol = t._oracle_step(p, torch.float64); sol32 = t._oracle_step(p, torch.float32)
    db.set_option("keep_system", 1)
    g = p.to("cuda")
    dx, dz = db.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.t0, g.t1, 1, g.lm, g.ep, False)
    S, rhs, dims = db.debug_system()
    db.set_option("keep_system", 0)
    r = {}
    r['S'] = rel_err(torch.tril(S.double()), torch.tril(sol["S"])) / 2e-6
    r['rhs'] = rel_err(rhs, sol["vs"]) / 2e-6
    r['dx'] = rel_err(dx.cpu(), sol["dx"]) / (t.MULT * rel_err(sol32["dx"], sol["dx"]) + 1e-5)
    r['dz'] = rel_err(dz.cpu().reshape(sol["dz"].shape), sol["dz"]) / (t.MULT * rel_err(sol32["dz"], sol["dz"]) + 1e-5)
    poses64, disps64 = od.oracle_ba(*p.ba_args(), dtype=torch.float64)
    gap_d, gap_p = t._reference_fp32_gap(p, poses64, disps64)
    h = p.to("cuda"); db.ba(*h.ba_args())
    r['disps2'] = rel_err(h.disps.cpu(), disps64) / max(1e-4, t.MULT * gap_d)
    r['poses2'] = float(ose3.tangent_update(h.poses.cpu().double(), poses64.double()).abs().max()) / max(1e-4, t.MULT * gap_p)
    worst = max(worst, max(r.values()))
    print(f"{name:24s} " + "  ".join(f"{k} {v:.2f}" for k, v in r.items()))
print("worst err / bound: %.2f" % worst)
