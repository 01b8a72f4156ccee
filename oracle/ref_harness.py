"""TEST INFRASTRUCTURE ONLY -- run the reference's OWN Python BA (build container only).

Imports hislam2/geom/{ba,chol,projective_ops,pinhole}.py UNMODIFIED from /root/reference, over the
reference's lietorch CPU backend built by oracle/build_ref.py into oracle/_ref/.  Two shims are
needed to import the files as shipped:
  * `droid_backends` is imported at module level (geom/ba.py:4) but never called by BA/MoBA: a
    blank module object is registered under that name;
  * `scatter_sum`'s import is commented out (geom/ba.py:9): torch_scatter's published semantics
    (zeros(dim_size).index_add_(dim, index, src)) are injected into geom.ba's namespace.
/root/reference does not exist on the GPU box: nothing under tests -m gpu, smoke() or bench.py
uses this module; it only validates oracle/dba.py and generates tests/golden/*.npz here.
"""
import functools
import os
import sys
import types

import torch

REF = os.environ.get("DBA_REFERENCE_ROOT", "/root/reference")
_loaded = {}


def available():
    return os.path.isdir(os.path.join(REF, "hislam2", "geom"))


def _scatter_sum(src, index, dim=-1, dim_size=None):
    shape = list(src.shape)
    shape[dim] = dim_size
    return torch.zeros(shape, dtype=src.dtype, device=src.device).index_add_(dim, index, src)


def load():
    """Returns (geom.ba module, geom.chol module, lietorch module) of the reference."""
    if _loaded:
        return _loaded["ba"], _loaded["chol"], _loaded["lietorch"]
    from . import build_ref
    build_ref.build()
    for p in (build_ref.OUT, os.path.join(REF, "thirdparty", "lietorch"), os.path.join(REF, "hislam2")):
        if p not in sys.path:
            sys.path.insert(0, p)
    sys.modules.setdefault("droid_backends", types.ModuleType("droid_backends"))
    import lietorch  # noqa: E402  (reference's)
    import geom.ba as rba  # noqa: E402
    import geom.chol as rchol  # noqa: E402
    rba.scatter_sum = _scatter_sum
    _loaded.update(ba=rba, chol=rchol, lietorch=lietorch)
    return rba, rchol, lietorch


def ref_ba(poses, disps, intrinsics, targets, weights, eta, ii, jj, t0, t1, iterations, lm, ep,
           motion_only, dtype=torch.float32):
    """The reference's BA/MoBA driven through the `droid_backends.ba` argument list.

    lm/ep: BA() calls schur_solve with its defaults (ep=0.1, lm=1e-4, geom/ba.py:98, chol.py:47);
    to drive other values the name `schur_solve`/`block_solve` inside geom.ba is rebound to a
    functools.partial of the reference's own function -- no reference code is altered.
    Returns new poses [N,7], disps [N,ht,wd] and the per-iteration dx list.
    """
    rba, rchol, lietorch = load()
    p = poses[:t1].to(dtype).clone()[None]
    d = disps[:t1].to(dtype).clone()[None]
    intr = intrinsics.to(dtype).view(1, 1, 4).expand(1, t1, 4).contiguous()
    tgt = targets.permute(0, 2, 3, 1).to(dtype).contiguous()[None]
    wgt = weights.permute(0, 2, 3, 1).to(dtype).contiguous()[None]
    et = eta.to(dtype)[None]
    ii = ii.long()
    jj = jj.long()
    rba.schur_solve = functools.partial(rchol.schur_solve, ep=ep, lm=lm)
    rba.block_solve = functools.partial(rchol.block_solve, ep=ep, lm=lm)
    dxs = []
    try:
        for _ in range(iterations):
            G = lietorch.SE3(p)
            if motion_only:
                G2 = rba.MoBA(tgt, wgt, et, G, d, intr, ii, jj, fixedp=t0)
            else:
                G2, d, _ = rba.BA(tgt, wgt, et, G, d, intr, ii, jj, fixedp=t0)
            dxs.append((G2 * G.inv()).log()[0, t0:])
            p = G2.data
    finally:
        rba.schur_solve = rchol.schur_solve
        rba.block_solve = rchol.block_solve
    new_poses = poses.clone()
    new_disps = disps.clone()
    new_poses[:t1] = p[0].to(poses.dtype)
    new_disps[:t1] = d[0].to(disps.dtype)
    return new_poses, new_disps, dxs


def ref_blocks(poses, disps, intrinsics, targets, weights, ii, jj, dtype=torch.float32):
    """Per-edge Jacobians from the reference's projective_transform (projective_ops.py:44-74)."""
    rba, _, lietorch = load()
    import geom.projective_ops as pops
    t1 = disps.shape[0]
    G = lietorch.SE3(poses.to(dtype)[None])
    intr = intrinsics.to(dtype).view(1, 1, 4).expand(1, t1, 4).contiguous()
    coords, valid, (Ji, Jj, Jz) = pops.projective_transform(G, disps.to(dtype)[None], intr, ii.long(), jj.long(),
                                                            jacobian=True)
    return coords[0], valid[0], Ji[0], Jj[0], Jz[0]
