"""The compiled `droid_backends` torch extension (csrc/droid_backends_ext.cpp; the reference builds one from
src/droid.cpp, setup.py:7-31): it builds in-tree, imports under the reference's module name and exports the ops; its
argument checks (the reference's CHECK_INPUT convention) reject host tensors before anything touches a GPU."""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ext():
    from cut3r_slam_b200 import build as b
    path = b.ext_path()
    if not os.path.exists(path):
        b.build_ext()
    d = os.path.dirname(path)
    if d not in sys.path:
        sys.path.insert(0, d)
    import droid_backends as m
    assert os.path.samefile(m.__file__, path)
    return m


def test_extension_exports_the_reference_ops():
    m = _ext()
    for name in ("ba", "pgba", "frame_distance", "covis_distance", "iproj", "depth_filter"):
        assert callable(getattr(m, name))
    assert "libdba" in m.version()


def test_extension_rejects_host_tensors():
    m = _ext()
    z = torch.zeros(1)
    i = torch.zeros(1, dtype=torch.int64)
    with pytest.raises(RuntimeError, match="contiguous CUDA tensor"):
        m.ba(z, z, z, z, z, z, z, i, i, 0, 1, 1, 1e-4, 0.1, False)
    with pytest.raises(RuntimeError, match="contiguous CUDA tensor"):
        m.frame_distance(z, z, z, i, i, 0.3)
