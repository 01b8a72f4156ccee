"""Generates tests/golden/jdsa.npz: the reference's OWN JDSA (hislam2/geom/ba.py:173-241) and
schur_solve_mono_prior (chol.py:80-108) run unmodified on CPU (build container only).  The two absent CUDA
ops JDSA calls (droid_backends.proj_trans / bi_inter) are injected from oracle/jdsa.py; `.cuda()` is a no-op
while the reference code runs."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import jdsa as oj  # noqa: E402
from oracle import ref_harness  # noqa: E402
from cut3r_slam_b200.synthetic import make_problem  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "jdsa.npz")
CASES = {"a": dict(P=6, E=20, ht=12, wd=16, iters=1, t0=1, r=2, seed=41, hs=2, ws=3, alpha=0.01),
         "b": dict(P=7, E=30, ht=10, wd=14, iters=1, t0=1, r=3, seed=42, hs=3, ws=2, alpha=0.05)}


def make_case(kw):
    kw = dict(kw)
    hs, ws, alpha = kw.pop("hs"), kw.pop("ws"), kw.pop("alpha")
    p = make_problem(**kw)
    g = torch.Generator().manual_seed(kw["seed"] + 100)
    scales = 1 + 0.1 * torch.randn(p.poses.shape[0], hs, ws, generator=g)
    prior = p.gt_disps * (1 + 0.05 * torch.randn(p.gt_disps.shape, generator=g))
    prior[:, :2, :3] = 0.0                       # pixels without a prior (m = 0)
    return p, scales.float(), prior.float(), alpha


def main():
    rba, rchol, lietorch = ref_harness.load()
    import droid_backends as stub
    out = {}
    for name, kw in CASES.items():
        p, scales, prior, alpha = make_case(kw)
        for dt, tag in ((torch.float64, "f64"), (torch.float32, "f32")):
            stub.proj_trans = lambda poses, disps, intr, tg, wg, ii, jj: oj.proj_trans(poses, disps, intr, tg, wg, ii, jj)
            stub.bi_inter = oj.bi_inter
            orig_cuda = torch.Tensor.cuda
            torch.Tensor.cuda = lambda self, *a, **k: self.to(dt) if self.is_floating_point() else self
            try:
                G = lietorch.SE3(p.poses.to(dt)[None])
                disps, dscales, dzcov = rba.JDSA(p.targets.to(dt), p.weights.to(dt), p.eta.to(dt), G, p.disps.to(dt)[None],
                                                 p.intrinsics.to(dt)[None], prior.to(dt), scales.to(dt).clone(), p.ii, p.jj, alpha)
            finally:
                torch.Tensor.cuda = orig_cuda
            out[f"{name}_disps_{tag}"] = disps[0].numpy()
            out[f"{name}_scales_{tag}"] = dscales.numpy()
            out[f"{name}_dzcov_{tag}"] = dzcov.numpy()
        out[f"{name}_scales_in"], out[f"{name}_prior"] = scales.numpy(), prior.numpy()
    # schur_solve_mono_prior alone on random block-diagonal inputs
    g = torch.Generator().manual_seed(7)
    M, D, HW = 4, 6, 40
    C = (0.5 + torch.rand(1, M, HW, generator=g, dtype=torch.float64))
    w = torch.randn(1, M, HW, generator=g, dtype=torch.float64)
    Hs = torch.zeros(1, M, M, D, D, dtype=torch.float64)
    Es = torch.zeros(1, M, M, D, HW, dtype=torch.float64)
    for m in range(M):
        A = torch.randn(D, D + 3, generator=g, dtype=torch.float64)
        Hs[0, m, m] = A @ A.T + 5 * torch.eye(D, dtype=torch.float64)
        Es[0, m, m] = 0.3 * torch.randn(D, HW, generator=g, dtype=torch.float64)
    vs = torch.randn(1, M, D, 1, generator=g, dtype=torch.float64)
    dso, dz, dzcov = rchol.schur_solve_mono_prior(C, w, Hs, Es, vs, dzcov=True)
    out.update(ss_C=C.numpy(), ss_w=w.numpy(), ss_Hs=Hs.numpy(), ss_Es=Es.numpy(), ss_vs=vs.numpy(),
               ss_dso=dso.numpy(), ss_dz=dz.numpy(), ss_dzcov=dzcov.numpy())
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, sorted(out)[:6], "...")


if __name__ == "__main__":
    main()
