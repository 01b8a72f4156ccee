"""Where the two-iteration disparity error of a weakly observed graph comes from: every linearize / split option of
libdba against the fp64 restatement, next to the reference algorithm's own fp32 run (profiles/r2_parity_margins.md)."""
import sys, torch
sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import importlib.util
spec = importlib.util.spec_from_file_location('t', 'tests/test_ba_ragged_gpu.py'); t = importlib.util.module_from_spec(spec); spec.loader.exec_module(t)
import oracle.dba as od
import oracle.se3 as ose3
from helpers import rel_err
from cut3r_slam_b200.synthetic import make_problem
from cut3r_slam_b200 import droid_backends as db
# the radius-RADIUS variant of tests/ragged_cases.py's `degree_thresholds` (radius 1: chain ends that hang on ONE edge)
import ragged_cases as rc
RADIUS = int(sys.argv[1]) if len(sys.argv) > 1 else 1
_orig = rc.chain
rc.chain = lambda P, r=1, **k: _orig(P, RADIUS, **k)
kw = dict(rc.CASES['degree_thresholds'], edges=rc.degree_mix())
rc.chain = _orig
p = make_problem(iters=2, **kw)
deg = torch.bincount(p.ii, minlength=48)
for it in (1, 2):
    p.iterations = it
    poses64, disps64 = od.oracle_ba(*p.ba_args(), dtype=torch.float64)
    poses32, disps32 = od.oracle_ba(*p.ba_args(), dtype=torch.float32)
    e32 = (disps32.double() - disps64).abs()
    k32 = int(e32.flatten().argmax()) // (12 * 20)
    print(f'it={it} reference fp32 gap disps {rel_err(disps32, disps64):.3e} (frame {k32}, deg {int(deg[k32])}) poses {float(ose3.tangent_update(poses32.double(), poses64.double()).abs().max()):.3e}')
    for name, opts in (('default', {}), ('simt only', {'tensor_cores': 0}), ('no wide tc', {'wide_tensor_cores': 0}), ('wide 1', {'wide_tensor_cores': 1}),
                       ('splits=1', {'splits': 1}), ('v2 off', {'linearize_v2': 0}),
                       ('refine=0', {'refine': 0}), ('refine=2', {'refine': 2}), ('refine=2 simt', {'refine': 2, 'tensor_cores': 0})):
        try:
            for k, v in opts.items(): db.set_option(k, v)
            g = p.to('cuda'); db.ba(*g.ba_args())
            e = (g.disps.cpu().double() - disps64).abs()
            kf = int(e.flatten().argmax()) // (12 * 20)
            print(f'   {name:12s} disps {rel_err(g.disps.cpu(), disps64):.3e} (frame {kf}, deg {int(deg[kf])}, disp there {float(disps64.flatten()[int(e.flatten().argmax())]):.3f})  poses {float(ose3.tangent_update(g.poses.cpu().double(), poses64.double()).abs().max()):.3e}  per-frame max err top3 {sorted([(round(float(e[f].max()),6), f) for f in range(48)])[-3:]}')
        except Exception as ex:
            print('   ', name, 'raised', ex)
        finally:
            for k in opts: db.set_option(k, {'tensor_cores': 1, 'wide_tensor_cores': 2, 'splits': 0, 'linearize_v2': 1, 'refine': 1}[k])
