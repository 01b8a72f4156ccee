"""TEST INFRASTRUCTURE ONLY -- tests/golden/ragged.npz: the UNMODIFIED reference BA (hislam2/geom/ba.py:32-107 over the
reference's lietorch CPU backend, driven by oracle/ref_harness.py) on ragged / degenerate factor graphs.

Run in the build container (needs /root/reference): `python -m oracle.make_golden_ragged`.
The graphs are tests/ragged_cases.py's REFERENCE_PINNED subset: keyframes without outgoing / incoming / any edges,
duplicate and self edges, edges inside the gauge, a single directed edge, a grid below one pixel tile, two random graphs.
Inputs are regenerated from the case's keywords (their SHA-256 is stored); outputs: poses, disps and the per-iteration
pose updates of a two-iteration call in fp32 and fp64.  tests/test_oracle_golden.py pins oracle/dba.py to them,
tests/test_ba_ragged_gpu.py checks the CUDA path against the same vectors.
"""
import os
import sys

import numpy as np
import torch

from cut3r_slam_b200.synthetic import make_problem
from . import ref_harness
from .make_golden import input_digest

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ragged_cases  # noqa: E402


def main():
    torch.set_num_threads(8)
    rec = {}
    for name in ragged_cases.REFERENCE_PINNED:
        p = make_problem(iters=2, **ragged_cases.CASES[name])
        rec[f"{name}/digest"] = np.array(input_digest(p))
        for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
            np_, nd, dxs = ref_harness.ref_ba(p.poses, p.disps, p.intrinsics, p.targets, p.weights, p.eta, p.ii, p.jj, p.t0,
                                              p.t1, p.iterations, p.lm, p.ep, False, dtype=dt)
            rec[f"{name}/ref_poses_{tag}"] = np_.numpy()
            rec[f"{name}/ref_disps_{tag}"] = nd.numpy()
            rec[f"{name}/ref_dx_{tag}"] = torch.stack(dxs).to(torch.float64).numpy()
        print(name, "E", int(p.ii.numel()), "dx max per iteration", [float(np.abs(d).max()) for d in rec[f"{name}/ref_dx_f32"]])
    # the same graphs through the reference's MoBA (geom/ba.py:110-158) ...
    for name in ragged_cases.REFERENCE_PINNED_MOBA:
        p = make_problem(iters=2, **ragged_cases.CASES[name])
        for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
            np_, nd, dxs = ref_harness.ref_ba(p.poses, p.disps, p.intrinsics, p.targets, p.weights, p.eta, p.ii, p.jj, p.t0,
                                              p.t1, p.iterations, p.lm, p.ep, True, dtype=dt)
            rec[f"{name}/moba_poses_{tag}"] = np_.numpy()
            rec[f"{name}/moba_dx_{tag}"] = torch.stack(dxs).to(torch.float64).numpy()
        print(name, "MoBA dx max per iteration", [float(np.abs(d).max()) for d in rec[f"{name}/moba_dx_f32"]])
    # ... and through its BA with lietorch.Sim3 poses (the 7-dof linearisation of pgba), scales as in the tests
    from .make_golden_pgba import ref_ba_sim3
    for name in ragged_cases.REFERENCE_PINNED_SIM3:
        p = make_problem(iters=2, **ragged_cases.CASES[name])
        poses8 = torch.cat([p.poses, ragged_cases.sim3_scales(p.poses.shape[0], p.t0)], -1)
        for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
            P8, d, dxs = ref_ba_sim3(p, poses8, dt)
            rec[f"{name}/sim3_poses_{tag}"] = P8.numpy()
            rec[f"{name}/sim3_disps_{tag}"] = d.numpy()
            rec[f"{name}/sim3_dx_{tag}"] = dxs.to(torch.float64).numpy()
        print(name, "Sim3 dx max per iteration", [float(np.abs(d).max()) for d in rec[f"{name}/sim3_dx_f32"]])
    path = os.path.join(ROOT, "tests", "golden", "ragged.npz")
    np.savez_compressed(path, **rec)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
