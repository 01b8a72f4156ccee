"""Small `ba` cases for compute-sanitizer (memcheck / racecheck / synccheck / initcheck) on the GPU box:
every hot kernel variant is launched at least once -- narrow tcgen05 (resident and evict-first loads),
WIDE 1, WIDE 2, SIMT, motion-only, the single-CTA solve, the cooperative Cholesky, the multi-launch
Cholesky, the companions.  Sizes are small: the sanitizer serialises and instruments every access.

    compute-sanitizer --tool racecheck python tools/sanitize_cases.py [case ...]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from cut3r_slam_b200 import droid_backends as db  # noqa: E402
from cut3r_slam_b200.synthetic import make_problem  # noqa: E402

CASES = {
    # name: (make_problem kwargs or config name, libdba options)
    "config1": ("correctness", {}),
    "config1_stream": ("correctness", {"stream_inputs": 1}),
    "config1_simt": ("correctness", {"tensor_cores": 0}),
    "wide1": (dict(P=40, E=760, ht=16, wd=24, iters=1, t0=1, r=11, seed=52), {"wide_tensor_cores": 1}),
    "wide2": (dict(P=44, E=1150, ht=16, wd=24, iters=1, t0=1, r=16, seed=52), {"wide_tensor_cores": 2}),
    "coop_chol": (dict(P=40, E=300, ht=20, wd=24, iters=2, t0=2, r=4, seed=81), {}),
    "multilaunch_chol": (dict(P=40, E=300, ht=20, wd=24, iters=1, t0=2, r=4, seed=81), {"multilaunch_cholesky": 1}),
    "moba": (dict(P=8, E=26, ht=30, wd=40, iters=2, t0=2, r=2, seed=32, motion_only=True), {}),
}
DEFAULTS = {"stream_inputs": -1, "tensor_cores": 1, "wide_tensor_cores": 2, "multilaunch_cholesky": 0}


def main():
    names = sys.argv[1:] or list(CASES)
    for name in names:
        spec, opts = CASES[name]
        p = make_problem(spec) if isinstance(spec, str) else make_problem(**spec)
        for k, v in {**DEFAULTS, **opts}.items():
            db.set_option(k, v)
        g = p.to("cuda")
        db.ba(*g.ba_args())
        torch.cuda.synchronize()
        ok = bool(torch.isfinite(g.poses).all() and torch.isfinite(g.disps).all())
        print(f"sanitize case {name}: launches={db.last_launch_count()} finite={ok}", flush=True)
    for k, v in DEFAULTS.items():
        db.set_option(k, v)
    g = make_problem("correctness").to("cuda")
    ii = torch.arange(0, 7, device="cuda")
    db.frame_distance(g.poses, g.disps, g.intrinsics, ii, ii + 1, 0.3)
    db.iproj(g.poses, g.disps, g.intrinsics)
    db.depth_filter(g.poses, g.disps, g.intrinsics, ii, torch.full((7,), 0.01, device="cuda"))
    torch.cuda.synchronize()
    print("sanitize companions: ok", flush=True)


if __name__ == "__main__":
    main()
