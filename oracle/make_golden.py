"""TEST INFRASTRUCTURE ONLY -- generate tests/golden/*.npz from the reference's own Python BA.

Run in the build container (needs /root/reference): `python -m oracle.make_golden`.
Each file holds the inputs of a small seeded problem (or their SHA-256 when they are regenerated
from the seed) and the outputs of the UNMODIFIED reference hislam2/geom/ba.py BA / MoBA driven by
oracle/ref_harness.py (fp32 and fp64).  tests/test_oracle_golden.py pins oracle/dba.py to them;
tests/test_ba_gpu.py checks the CUDA path against the same vectors on the GPU box.
"""
import hashlib
import os

import numpy as np
import torch

from cut3r_slam_b200.synthetic import make_problem
from . import ref_harness

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")

CASES = {
    # name: make_problem kwargs + overrides, store_inputs
    "tiny_ba":         dict(kw=dict(P=6, E=14, ht=24, wd=32, iters=2, lm=1e-4, ep=0.1, t0=1, r=2, seed=11), store=True),
    "tiny_ba_backend": dict(kw=dict(P=7, E=24, ht=24, wd=32, iters=3, lm=1e-5, ep=1e-2, t0=2, r=2, seed=12), store=True),
    "tiny_moba":       dict(kw=dict(P=6, E=18, ht=24, wd=32, iters=2, lm=1e-4, ep=0.1, t0=1, r=2, motion_only=True, seed=13), store=True),
    "correctness":     dict(name="correctness", store=False),
    "frontend":        dict(name="frontend", store=False),
}


def input_digest(p):
    h = hashlib.sha256()
    for t in (p.poses, p.disps, p.intrinsics, p.targets, p.weights, p.eta, p.ii, p.jj):
        h.update(t.contiguous().numpy().tobytes())
    return h.hexdigest()


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(8)
    for name, c in CASES.items():
        p = make_problem(c["name"]) if "name" in c else make_problem(**c["kw"])
        rec = dict(digest=np.array(input_digest(p)), t0=p.t0, t1=p.t1, iterations=p.iterations,
                   lm=p.lm, ep=p.ep, motion_only=p.motion_only)
        if c["store"]:
            rec.update(poses=p.poses.numpy(), disps=p.disps.numpy(), intrinsics=p.intrinsics.numpy(),
                       targets=p.targets.numpy(), weights=p.weights.numpy(), eta=p.eta.numpy(),
                       ii=p.ii.numpy(), jj=p.jj.numpy())
        for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
            np_, nd, dxs = ref_harness.ref_ba(p.poses, p.disps, p.intrinsics, p.targets, p.weights, p.eta,
                                              p.ii, p.jj, p.t0, p.t1, p.iterations, p.lm, p.ep,
                                              p.motion_only, dtype=dt)
            rec[f"ref_poses_{tag}"] = np_.numpy()
            rec[f"ref_disps_{tag}"] = nd.numpy()
            rec[f"ref_dx_{tag}"] = torch.stack(dxs).to(torch.float64).numpy()
        if c["store"]:
            # pin the Jacobians too: a deterministic pixel subsample of projective_transform's outputs
            coords, valid, Ji, Jj, Jz = ref_harness.ref_blocks(p.poses, p.disps, p.intrinsics, p.targets,
                                                              p.weights, p.ii, p.jj)
            sl = (slice(None), slice(1, None, 5), slice(2, None, 7))
            rec.update(jac_coords=coords[sl].numpy(), jac_valid=valid[sl].numpy(), jac_Ji=Ji[sl].numpy(),
                       jac_Jj=Jj[sl].numpy(), jac_Jz=Jz[sl].numpy())
        path = os.path.join(OUT, f"{name}.npz")
        np.savez_compressed(path, **rec)
        print(name, os.path.getsize(path), "bytes; dx max per iter",
              [float(np.abs(d).max()) for d in rec["ref_dx_f32"]])


if __name__ == "__main__":
    main()
