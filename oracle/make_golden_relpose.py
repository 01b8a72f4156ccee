"""TEST INFRASTRUCTURE ONLY (build container): runs the reference's own PGOBuffer.add_rel_poses
(hislam2/pgo_buffer.py:85-123) unmodified on CPU -> tests/golden/relpose.npz.  The method is called on a plain
namespace standing in for `self` (video.poses / disps / intrinsics and the rel_* buffers it appends to); the only shim
is `torch.eye(..., device='cuda')` -> CPU inside the call, because this container has no GPU."""
import importlib
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from cut3r_slam_b200.synthetic import make_problem  # noqa: E402
from oracle import ref_harness  # noqa: E402
from oracle.make_golden_graph import _Stub  # noqa: E402

CASES = {"a": dict(P=6, E=14, ht=12, wd=16, iters=1, t0=1, r=2, seed=61),
         "b": dict(P=8, E=24, ht=10, wd=14, iters=1, t0=1, r=3, seed=62)}


def make_case(kw):
    p = make_problem(**kw)
    target = p.targets.permute(0, 2, 3, 1).contiguous()            # FactorGraph layout [E,ht,wd,2]
    weight = p.weights.permute(0, 2, 3, 1).contiguous().clone()
    weight[1] *= 0.05                                              # an edge that fails the weight mask
    weight[2, : p.disps.shape[1] // 2] = 0.0                       # and one that keeps half its pixels
    return p, target, weight


def load_pgo():
    ref_harness.load()
    for name in ["networkx", "matplotlib", "matplotlib.pyplot", "matplotlib.collections", "sklearn",
                 "sklearn.decomposition", "tqdm", "util", "util.utils", "factor_graph"]:
        try:
            importlib.import_module(name)
        except Exception:
            sys.modules[name] = _Stub(name)
    import pgo_buffer
    return pgo_buffer


def main():
    pgo = load_pgo()
    out = {}
    for name, kw in CASES.items():
        p, target, weight = make_case(kw)
        for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
            E = p.ii.numel()
            me = types.SimpleNamespace(
                video=types.SimpleNamespace(poses=p.poses.to(dt), disps=p.disps.to(dt),
                                            intrinsics=p.intrinsics.to(dt).view(1, 4).expand(p.poses.shape[0], 4).contiguous()),
                rel_N=types.SimpleNamespace(value=0), rel_ii=torch.zeros(E, dtype=torch.long),
                rel_jj=torch.zeros(E, dtype=torch.long), rel_poses=torch.zeros(E, 7, dtype=dt),
                rel_covs=torch.zeros(E, 6, dtype=dt), rel_valid_percent=torch.zeros(E, dtype=dt))
            eye = torch.eye
            torch.eye = lambda *a, **k: eye(*a, **{kk: vv for kk, vv in k.items() if kk != "device"}, dtype=dt)
            try:
                pgo.PGOBuffer.add_rel_poses(me, p.ii, p.jj, target.to(dt)[None], weight.to(dt)[None])
            finally:
                torch.eye = eye
            n = me.rel_N.value
            out[f"{name}_n"] = np.array(n)
            out[f"{name}_ii"], out[f"{name}_jj"] = me.rel_ii[:n].numpy(), me.rel_jj[:n].numpy()
            out[f"{name}_poses_{tag}"], out[f"{name}_covs_{tag}"] = me.rel_poses[:n].numpy(), me.rel_covs[:n].numpy()
            out[f"{name}_valid_{tag}"] = me.rel_valid_percent[:n].numpy()
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "relpose.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
