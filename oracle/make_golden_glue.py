"""TEST INFRASTRUCTURE ONLY (build container): runs the reference's own cvx_upsample / upsample_disp
(hislam2/modules/droid_net.py:21-44, imported unmodified through oracle/ref_harness.py) on seeded inputs and stores
the outputs in tests/golden/glue.npz.  Inputs are regenerated from the seed by make_case()."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_harness  # noqa: E402

CASES = {"u1": dict(B=2, ht=3, wd=4, dim=1, seed=7),
         "u3": dict(B=1, ht=5, wd=2, dim=3, seed=8),
         "row": dict(B=1, ht=1, wd=6, dim=2, seed=9),        # first row == last row
         "col": dict(B=2, ht=4, wd=1, dim=1, seed=10)}


def make_case(kw):
    g = torch.Generator().manual_seed(kw["seed"])
    data = torch.rand(kw["B"], kw["ht"], kw["wd"], kw["dim"], generator=g, dtype=torch.float64) + 0.1
    mask = 3.0 * torch.randn(kw["B"], 576, kw["ht"], kw["wd"], generator=g, dtype=torch.float64)
    return data, mask


def main():
    ref_harness.load()
    import modules.droid_net as dn                       # the reference's, unmodified
    out = {}
    for name, kw in CASES.items():
        data, mask = make_case(kw)
        out[f"{name}_up"] = dn.cvx_upsample(data.clone(), mask.clone()).numpy()
        out[f"{name}_up_f32"] = dn.cvx_upsample(data.float(), mask.float().clone()).numpy()
    data, mask = make_case(CASES["u1"])
    out["disp_up"] = dn.upsample_disp(data[None, ..., 0].clone(), mask[None].clone()).numpy()
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "glue.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
