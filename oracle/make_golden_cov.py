"""TEST INFRASTRUCTURE ONLY (build container): the depth covariance the reference's own BA returns as its third
output (hislam2/geom/ba.py:98-107, chol.py:74-76, run unmodified through oracle/ref_harness.py) on the stored inputs
of the tiny golden problems -> tests/golden/cov.npz."""
import functools
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_harness  # noqa: E402

CASES = ("tiny_ba", "tiny_ba_backend")
RAGGED = ("sink_and_source", "isolated_keyframe", "self_edges", "random_101")


def main():
    rba, rchol, lietorch = ref_harness.load()
    out = {}
    for name in CASES:
        G = np.load(os.path.join(ROOT, "tests", "golden", f"{name}.npz"))
        t = lambda k: torch.from_numpy(G[k])
        t0, t1, lm, ep = int(G["t0"]), int(G["t1"]), float(G["lm"]), float(G["ep"])
        for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
            p = t("poses")[:t1].to(dt)[None]
            d = t("disps")[:t1].to(dt)[None]
            intr = t("intrinsics").to(dt).view(1, 1, 4).expand(1, t1, 4).contiguous()
            tgt = t("targets").permute(0, 2, 3, 1).to(dt).contiguous()[None]
            wgt = t("weights").permute(0, 2, 3, 1).to(dt).contiguous()[None]
            rba.schur_solve = functools.partial(rchol.schur_solve, ep=ep, lm=lm)
            try:
                _, _, dzcov = rba.BA(tgt, wgt, t("eta").to(dt)[None], lietorch.SE3(p), d, intr, t("ii").long(), t("jj").long(),
                                     fixedp=t0)
            finally:
                rba.schur_solve = rchol.schur_solve
            out[f"{name}_dzcov_{tag}"] = dzcov.numpy()
    # ragged graphs (tests/ragged_cases.py): dzcov rows follow sorted(unique(ii)), not the frame index
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ragged_cases
    from cut3r_slam_b200.synthetic import make_problem
    for name in RAGGED:
        q = make_problem(iters=1, **ragged_cases.CASES[name])
        for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
            p = q.poses[: q.t1].to(dt)[None]
            d = q.disps[: q.t1].to(dt)[None]
            intr = q.intrinsics.to(dt).view(1, 1, 4).expand(1, q.t1, 4).contiguous()
            tgt = q.targets.permute(0, 2, 3, 1).to(dt).contiguous()[None]
            wgt = q.weights.permute(0, 2, 3, 1).to(dt).contiguous()[None]
            rba.schur_solve = functools.partial(rchol.schur_solve, ep=q.ep, lm=q.lm)
            try:
                _, _, dzcov = rba.BA(tgt, wgt, q.eta.to(dt)[None], lietorch.SE3(p), d, intr, q.ii.long(), q.jj.long(), fixedp=q.t0)
            finally:
                rba.schur_solve = rchol.schur_solve
            out[f"ragged_{name}_dzcov_{tag}"] = dzcov.numpy()
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "cov.npz"), **out)
    print({k: (v.shape, float(v.min()), float(v.max())) for k, v in out.items()})


if __name__ == "__main__":
    main()
