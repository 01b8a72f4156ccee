"""Who owns the error of `depth_cov` on an ill-conditioned graph: the assembled system, the device factorisation, or
L^-1 / S^-1 / the quadratic form (run on the GPU box: python tools/cov_diag.py [case ...])."""
import sys, torch
sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import importlib.util
spec = importlib.util.spec_from_file_location('t', 'tests/test_ba_ragged_gpu.py'); t = importlib.util.module_from_spec(spec); spec.loader.exec_module(t)
import oracle.dba as od
from cut3r_slam_b200.synthetic import make_problem
from cut3r_slam_b200 import droid_backends as db


def relmax(a, b):
    return float(((a - b).abs() / b.abs()).max())


for name in (sys.argv[1:] or ['random_101', 'degree_thresholds']):
    p = make_problem(iters=1, **t.CASES[name])

    def lin(dt):
        c = lambda x: x.to(dt)
        l = od.linearize(c(p.targets.permute(0, 2, 3, 1)), c(p.weights.permute(0, 2, 3, 1)), c(p.eta), c(p.poses), c(p.disps),
                         c(p.intrinsics), p.ii, p.jj, p.t0)
        return od.schur_solve(l['H'], l['E'], l['C'], l['v'], l['w'], ep=p.ep, lm=p.lm, with_dzcov=True), l
    s64, l64 = lin(torch.float64)
    s32, _ = lin(torch.float32)
    ref = s64['dzcov']
    P, M, D, HW = l64['E'].shape
    E = l64['E'].permute(0, 2, 1, 3).reshape(P * D, M * HW)
    Q = (1 / l64['C']).reshape(-1)
    U = E * Q
    dsc = s64['S'].diagonal().rsqrt()
    scaled = lambda A: float(((A - s64['S']) * dsc[:, None] * dsc[None, :]).abs().max())
    print(name, 'n', s64['S'].shape[0], 'cond %.2e' % float(torch.linalg.cond(s64['S'])))
    print('  reference fp32 run                                   %.2e   scaled S error %.2e' % (relmax(s32['dzcov'].double(), ref), scaled(s32['S'].double())))
    for opts in ({}, {'tensor_cores': 0}, {'splits': 1}, {'linearize_v2': 0}):
        for k, v in opts.items(): db.set_option(k, v)
        print(' options', opts)
        g = p.to('cuda')
        db.set_option('keep_system', 1)
        cov = db.depth_cov(g.poses, g.disps, g.intrinsics, None, g.targets, g.weights, g.eta, g.ii, g.jj, p.t0, p.t1, p.lm, p.ep)
        S, rhs, dims = db.debug_system()
        db.set_option('keep_system', 0)
        for k in opts: db.set_option(k, {'tensor_cores': 1, 'splits': 0, 'linearize_v2': 1}[k])
        S = torch.tril(S.double()); S = S + torch.tril(S, -1).T
        print('  device depth_cov                                     %.2e' % relmax(cov.cpu().double().reshape(M, HW), ref))
        print('  device S vs fp64 S (max rel)                         %.2e   scaled S error %.2e' % (float((S - s64['S']).abs().max() / s64['S'].abs().max()), scaled(S)))
        f = lambda L: ((torch.linalg.inv(L) @ U).square().sum(0) + Q).reshape(M, HW)
        print('  device S, everything else fp64 on the host           %.2e' % relmax(f(torch.linalg.cholesky(S)), ref))
        print('  device S, LAPACK fp32 factor, rest fp64              %.2e' % relmax(f(torch.linalg.cholesky(S.float()).double()), ref))
