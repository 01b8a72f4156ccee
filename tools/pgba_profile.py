"""Developer tool: one pgba call (2 iterations) on the backend graph, for an ncu launch list."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from cut3r_slam_b200 import droid_backends as db
from cut3r_slam_b200.synthetic import make_problem
p = make_problem(sys.argv[1] if len(sys.argv) > 1 else "backend")
g = p.to("cuda")
N = g.poses.shape[0]
poses8 = torch.cat([g.poses, torch.ones(N, 1, device="cuda")], 1).contiguous()
R = 16
rel_ii = torch.arange(2, 2 + R, device="cuda"); rel_jj = rel_ii + 5
rel_poses = torch.zeros(R, 7, device="cuda"); rel_poses[:, 6] = 1.0
rel_infos = torch.eye(7, device="cuda").expand(R, 7, 7).contiguous()
for _ in range(2):
    d1 = g.disps.clone(); P8 = poses8.clone()
    db.pgba(P8, d1, g.intrinsics, g.targets, g.weights, g.eta, g.ii, g.jj, rel_ii, rel_jj, rel_poses, rel_infos, p.t0, p.t1, 2, p.lm, p.ep, 1e-5)
torch.cuda.synchronize()
