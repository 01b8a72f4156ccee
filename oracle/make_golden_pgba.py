"""TEST INFRASTRUCTURE ONLY (build container) -- golden vectors for the Sim3 path from the UNMODIFIED reference:
  * hislam2/geom/ba.py `BA` run with lietorch.Sim3 poses (7-dof step: dx, new poses, new disps), fp64 and fp32;
  * hislam2/pgo_buffer.py `global_relative_posesim3_constraints` (+ `num_jacobi`, `diff`): the function sources are
    executed as they are except that the hard-coded device='cuda' of num_jacobi (:24) becomes device='cpu'.
-> tests/golden/pgba.npz.   python -m oracle.make_golden_pgba
"""
import functools
import os
import re

import numpy as np
import torch

from . import ref_harness, sim3
from cut3r_slam_b200.synthetic import make_problem

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "pgba.npz")
KW = dict(P=7, E=24, ht=24, wd=32, iters=2, lm=1e-5, ep=1e-2, t0=2, r=2, seed=17)


def sim3_problem():
    p = make_problem(**KW)
    g = torch.Generator().manual_seed(5)
    scale = 1.0 + 0.05 * (2 * torch.rand(p.poses.shape[0], 1, generator=g) - 1)
    scale[:p.t0] = 1.0
    poses8 = torch.cat([p.poses, scale], dim=-1)
    # relative-pose factors: four pairs, measured relative pose = ground truth + noise, SPD information matrices
    rii = torch.tensor([0, 1, 2, 5]); rjj = torch.tensor([4, 6, 5, 3])
    import oracle.se3 as se3
    Gij = se3.se3_mul(p.gt_poses[rii].double(), se3.se3_inv(p.gt_poses[rjj].double()))    # e = log(Gij Gi Gj^-1) ~ 0
    Gij = se3.se3_retr(Gij, 0.01 * torch.randn(4, 6, generator=g, dtype=torch.float64)).float()
    A = torch.randn(4, 7, 7, generator=g)
    infos = (A @ A.transpose(1, 2) + 7 * torch.eye(7)) * 1e4
    return p, poses8, rii, rjj, Gij, infos


def ref_rel_blocks(rii, rjj, poses8, rel_poses, infos, pw):
    rba, rchol, lietorch = ref_harness.load()
    src = open(os.path.join(ref_harness.REF, "hislam2", "pgo_buffer.py")).read()
    a = src.index("eps = 1e-8")
    b = src.index("class PGOBuffer")
    code = src[a:b].replace("device='cuda'", "device='cpu'")
    ns = dict(torch=torch, SE3=lietorch.SE3, Sim3=lietorch.Sim3)
    exec(compile(code, "pgo_buffer.py[14:62]", "exec"), ns)
    Hsp, vsp, chi2, chi2s = ns["global_relative_posesim3_constraints"](rii, rjj, poses8[None], rel_poses[None],
                                                                      infos[None], pw=pw)
    return Hsp[:, 0], vsp[:, 0]


def ref_ba_sim3(p, poses8, dtype):
    rba, rchol, lietorch = ref_harness.load()
    t1 = p.t1
    P = poses8[:t1].to(dtype).clone()[None]
    d = p.disps[:t1].to(dtype).clone()[None]
    intr = p.intrinsics.to(dtype).view(1, 1, 4).expand(1, t1, 4).contiguous()
    tgt = p.targets.permute(0, 2, 3, 1).to(dtype).contiguous()[None]
    wgt = p.weights.permute(0, 2, 3, 1).to(dtype).contiguous()[None]
    et = p.eta.to(dtype)[None]
    rba.schur_solve = functools.partial(rchol.schur_solve, ep=p.ep, lm=p.lm)
    dxs = []
    try:
        for _ in range(p.iterations):
            G = lietorch.Sim3(P)
            G2, d, _ = rba.BA(tgt, wgt, et, G, d, intr, p.ii, p.jj, fixedp=p.t0)
            dxs.append((G2 * G.inv()).log()[0, p.t0:])
            P = G2.data
    finally:
        rba.schur_solve = rchol.schur_solve
    return P[0], d[0], torch.stack(dxs)


def main():
    p, poses8, rii, rjj, rel, infos = sim3_problem()
    out = dict(poses8=poses8.numpy(), disps=p.disps.numpy(), intrinsics=p.intrinsics.numpy(), targets=p.targets.numpy(),
               weights=p.weights.numpy(), eta=p.eta.numpy(), ii=p.ii.numpy(), jj=p.jj.numpy(), t0=p.t0, t1=p.t1,
               iterations=p.iterations, lm=p.lm, ep=p.ep, rel_ii=rii.numpy(), rel_jj=rjj.numpy(), rel_poses=rel.numpy(),
               rel_infos=infos.numpy(), pw=1e-5)
    for dt, tag in ((torch.float64, "f64"), (torch.float32, "f32")):
        P, d, dxs = ref_ba_sim3(p, poses8, dt)
        out[f"ref_poses_{tag}"] = P.double().numpy()
        out[f"ref_disps_{tag}"] = d.double().numpy()
        out[f"ref_dx_{tag}"] = dxs.double().numpy()
    Hsp, vsp = ref_rel_blocks(rii, rjj, poses8, rel, infos, 1e-5)
    out["ref_Hsp"] = Hsp.numpy()
    out["ref_vsp"] = vsp.numpy()
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, {k: getattr(v, "shape", v) for k, v in out.items() if k.startswith("ref_")})


if __name__ == "__main__":
    main()
