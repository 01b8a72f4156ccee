"""Which half owns pgba's second-iteration error on a weakly observed graph: the state its first iteration leaves, or
the second linearisation / solve itself?  (A) device iteration 1 against the fp64 restatement; (B) the fp64 restatement's
second update FROM THE DEVICE'S STATE; (C) the device's second update FROM THE fp64 STATE (rounded to fp32)."""
import sys, os, numpy as np, torch
sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import ragged_cases as rc
import oracle.pgba as op
from helpers import rel_err, GOLDEN
from cut3r_slam_b200.synthetic import make_problem
from cut3r_slam_b200 import droid_backends as db
G = np.load(os.path.join(GOLDEN, 'ragged.npz'))
e = torch.zeros(0, dtype=torch.int64, device='cuda')
for name in (sys.argv[1:] or ['isolated_keyframe', 'sink_and_source']):
    R = lambda k: torch.from_numpy(G[f'{name}/{k}'])
    p = make_problem(iters=2, **rc.CASES[name])
    P0 = torch.cat([p.poses, rc.sim3_scales(p.poses.shape[0], p.t0)], -1)
    a = (p.intrinsics, p.targets, p.weights, p.eta, p.ii, p.jj)
    ag = [x.cuda() for x in a]
    def oracle(P, D, dt):
        return op.oracle_pgba(P, D, *a, None, None, None, None, p.t0, p.t1, 1, p.lm, p.ep, 0.0, dtype=dt, return_info=True)
    def device(P, D):
        Pg, Dg = P.float().cuda().contiguous(), D.float().cuda().contiguous()
        dx, dz = db.pgba(Pg, Dg, *ag, e, e, torch.zeros(0, 7, device='cuda'), torch.zeros(0, 7, 7, device='cuda'), p.t0, p.t1, 1, p.lm, p.ep)
        return Pg.cpu(), Dg.cpu(), dx.cpu()
    P1, D1, i1 = oracle(P0, p.disps, torch.float64)
    P1s, D1s, i1s = oracle(P0, p.disps, torch.float32)
    Pg, Dg, dxg = device(P0, p.disps)
    src = torch.zeros(P0.shape[0], dtype=torch.bool); src[p.ii] = True
    print(f'{name}: (A) state after iteration 1 against fp64: disps {rel_err(Dg, D1):.2e} (fp32 restatement {rel_err(D1s, D1):.2e}), poses {float((Pg.double() - P1.double()).abs().max()):.2e} ({float((P1s.double() - P1.double()).abs().max()):.2e}), dx {rel_err(dxg, i1[0][0]):.2e} ({rel_err(i1s[0][0], i1[0][0]):.2e})')
    ed = (Dg.double() - D1.double()).abs().flatten(1).max(1).values
    print('     worst frames (disps):', sorted([(round(float(ed[f]), 7), f) for f in range(len(ed))])[-3:])
    ref2 = R('sim3_dx_f64')[1]
    _, _, iB = oracle(Pg, Dg, torch.float64)
    _, _, iBs = oracle(P1s, D1s, torch.float64)
    print(f'   (B) fp64 second update from the device state: {rel_err(iB[0][0], ref2):.2e}   (from the fp32 restatement\'s state: {rel_err(iBs[0][0], ref2):.2e})')
    _, _, dxC = device(P1, D1)
    print(f'   (C) device second update from the fp64 state: {rel_err(dxC, ref2):.2e}   (reference fp32 run, own state: {rel_err(R("sim3_dx_f32")[1], ref2):.2e})')
