"""pgba (Sim3, no relative-pose factors) one iteration at a time on ragged graphs against the reference's own BA run with
lietorch.Sim3 poses (tests/golden/ragged.npz): error / reference fp32 gap per iteration, and run-to-run bit identity."""
import sys, os, numpy as np, torch
sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import ragged_cases as rc
from helpers import rel_err, GOLDEN
from cut3r_slam_b200.synthetic import make_problem
from cut3r_slam_b200 import droid_backends as db
G = np.load(os.path.join(GOLDEN, 'ragged.npz'))
for name in ('isolated_keyframe', 'sink_and_source'):
    R = lambda k: torch.from_numpy(G[f'{name}/{k}'])
    outs = []
    for rep in range(2):
        p = make_problem(iters=2, **rc.CASES[name])
        P8 = torch.cat([p.poses, rc.sim3_scales(p.poses.shape[0], p.t0)], -1).cuda(); D = p.disps.cuda()
        e = torch.zeros(0, dtype=torch.int64, device='cuda')
        a = [x.cuda() for x in (p.intrinsics, p.targets, p.weights, p.eta, p.ii, p.jj)]
        dxs = []
        for k in range(2):
            dx, dz = db.pgba(P8, D, *a, e, e, torch.zeros(0, 7, device='cuda'), torch.zeros(0, 7, 7, device='cuda'), p.t0, p.t1, 1, p.lm, p.ep)
            dxs.append(dx.cpu())
        outs.append(dxs)
        if rep == 0:
            for k in range(2):
                gap = rel_err(R('sim3_dx_f32')[k], R('sim3_dx_f64')[k])
                err = rel_err(dxs[k], R('sim3_dx_f64')[k])
                d = (dxs[k].double() - R('sim3_dx_f64')[k]).abs()
                print(f'{name} it {k}: err {err:.3e} reference fp32 gap {gap:.3e} ratio {err / gap:.2f}; worst row {int(d.max(1).values.argmax())} (frame {int(d.max(1).values.argmax()) + p.t0}) component {int(d.max(0).values.argmax())}; |dx| max {float(R("sim3_dx_f64")[k].abs().max()):.3e}')
            print('   poses err %.3e (reference fp32 %.3e)' % (float((P8.cpu().double() - R('sim3_poses_f64').double()).abs().max()), float((R('sim3_poses_f32').double() - R('sim3_poses_f64').double()).abs().max())))
    print('   bit-identical between two runs:', all(torch.equal(x, y) for x, y in zip(outs[0], outs[1])))
