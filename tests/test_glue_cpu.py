"""oracle/glue.py against the reference's own outputs: cvx_upsample / upsample_disp (tests/golden/glue.npz, made by
oracle/make_golden_glue.py from hislam2/modules/droid_net.py:21-44 run unmodified) and projective_transform
coords / valid (the jac_coords / jac_valid entries of the tiny_* golden files, projective_ops.py:44-74)."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import oracle.glue as og  # noqa: E402
from oracle.make_golden_glue import CASES, make_case  # noqa: E402

GOLD = np.load(os.path.join(ROOT, "tests", "golden", "glue.npz"))


@pytest.mark.parametrize("name", sorted(CASES))
def test_cvx_upsample_matches_reference(name):
    data, mask = make_case(CASES[name])
    keep = mask.clone()
    up = og.cvx_upsample(data, mask)
    assert torch.equal(mask, keep)                                    # no side effect on the caller's logits
    ref = torch.from_numpy(GOLD[f"{name}_up"])
    assert up.shape == ref.shape
    assert float((up - ref).abs().max()) < 1e-13


def test_upsample_disp_matches_reference():
    data, mask = make_case(CASES["u1"])
    up = og.upsample_disp(data[None, ..., 0], mask[None])
    assert float((up - torch.from_numpy(GOLD["disp_up"])).abs().max()) < 1e-13


def test_cvx_upsample_properties():
    """Convex combination: a constant field stays constant, outputs stay inside the neighbourhood's range, and a
    one-hot logit copies that neighbour."""
    g = torch.Generator().manual_seed(1)
    mask = torch.randn(1, 576, 4, 5, generator=g, dtype=torch.float64)
    const = torch.full((1, 4, 5, 2), 0.7, dtype=torch.float64)
    assert torch.allclose(og.cvx_upsample(const, mask), torch.full((1, 32, 40, 2), 0.7, dtype=torch.float64))
    data = torch.rand(1, 4, 5, 1, generator=g, dtype=torch.float64)
    up = og.cvx_upsample(data, mask)
    assert float(up.min()) >= float(data.min()) - 1e-12 and float(up.max()) <= float(data.max()) + 1e-12
    hot = torch.full((1, 9, 8, 8, 4, 5), -50.0, dtype=torch.float64)
    hot[:, 5] = 50.0                                                  # neighbour (0, +1): the pixel to the right
    up = og.cvx_upsample(data, hot.reshape(1, 576, 4, 5))
    assert torch.allclose(up[0, 8:16, 0:8, 0], data[0, 1, 1, 0].expand(8, 8), atol=1e-12)
    # at the last column that neighbour is outside: the remaining (equal) logits average the valid neighbours
    assert float((up[0, 8:16, 32:40, 0] - data[0, 0:3, 3:5, 0].mean()).abs().max()) < 1e-12


@pytest.mark.parametrize("name", ["tiny_ba", "tiny_moba", "tiny_ba_backend"])
def test_reproject_matches_reference(name):
    G = np.load(os.path.join(ROOT, "tests", "golden", f"{name}.npz"))
    t = lambda k: torch.from_numpy(G[k])
    coords, valid = og.reproject(t("poses").double(), t("disps").double(), t("intrinsics").double(), t("ii").long(),
                                 t("jj").long())
    sl = (slice(None), slice(1, None, 5), slice(2, None, 7))         # the golden files hold this pixel subsample
    coords, valid = coords[sl], valid[sl]
    ref_c, ref_v = t("jac_coords").double(), t("jac_valid").double()  # the reference ran in fp32
    assert coords.shape == ref_c.shape and valid.shape == ref_v.shape
    assert float((coords - ref_c).abs().max()) < 2e-5 * float(ref_c.abs().max())
    assert torch.equal(valid, ref_v)
