"""Host logic without a GPU: libdba loads and exports every symbol include/dba.h declares, and the
plan's fixed-order gather lists reproduce the reference's scatter-adds (hislam2/geom/ba.py:70-92 +
chol.py:63-64) when replayed on the CPU over the oracle's per-edge blocks."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

import oracle.dba as od
from cut3r_slam_b200 import _lib
from cut3r_slam_b200.synthetic import make_problem

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "dba.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = set(re.findall(r"\b(dba_[a-z0-9_]+)\s*\(", hdr))
    assert {"dba_ba", "dba_ba_host", "dba_frame_distance", "dba_iproj", "dba_depth_filter", "dba_create"} <= names
    lib = _lib.load()
    for n in names:
        assert hasattr(lib, n), n
    assert set(_lib.SIGNATURES) == names
    assert b"sm_100a" in lib.dba_version()


def test_no_cpu_fallback():
    """Without a CUDA device the context cannot be created and the ops raise."""
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = _lib.load()
    assert not lib.dba_create(0)
    from cut3r_slam_b200 import droid_backends as db
    p = make_problem("correctness")
    with pytest.raises(RuntimeError):
        db.ba(*p.ba_args())


def _plan(ii, jj, N, t0, t1, motion_only, M, splits=1):
    lib = _lib.load()
    ii = np.ascontiguousarray(ii.numpy().astype(np.int64))
    jj = np.ascontiguousarray(jj.numpy().astype(np.int64))
    err = C.create_string_buffer(256)
    h = lib.dba_plan_build(ii.ctypes.data, jj.ctypes.data, len(ii), N, t0, t1, int(motion_only), M, splits, err, 256)
    if not h:
        raise RuntimeError(err.value.decode())
    out = {}
    for name in ("kx", "koff", "cedge", "ci", "cj", "pairbase", "blkbase", "sblk_off", "rhs_off", "sblk_items",
                 "rhs_items", "dims"):
        n = lib.dba_plan_get(h, name.encode(), None, 0)
        a = np.zeros(max(n, 1), dtype=np.int32)
        lib.dba_plan_get(h, name.encode(), a.ctypes.data, n)
        out[name] = a[:n]
    lib.dba_plan_free(h)
    return out


@pytest.mark.parametrize("kw", [
    dict(P=6, E=14, ht=8, wd=10, t0=1, r=2, seed=3),
    dict(P=8, E=30, ht=6, wd=8, t0=3, r=3, seed=4),       # edges touching fixed frames
    dict(P=6, E=14, ht=8, wd=10, t0=1, r=2, seed=5, dup=True),   # duplicate and i == j edges
])
def test_gather_lists_reproduce_reference_scatter(kw):
    kw = dict(kw)
    dup = kw.pop("dup", False)
    p = make_problem(**kw)
    if dup:
        p.jj[3] = p.jj[2] if p.ii[3] == p.ii[2] else p.jj[3]
        p.jj[5] = p.ii[5]
        p.jj[9] = p.jj[8] if p.ii[9] == p.ii[8] else p.jj[9]
    _check_gather_lists(p)


def _ragged_names():
    import ragged_cases
    return list(ragged_cases.CASES)


@pytest.mark.parametrize("name", _ragged_names())
def test_gather_lists_on_ragged_graphs(name):
    """The plan of every ragged / degenerate graph of tests/ragged_cases.py (keyframes without outgoing / incoming / any
    edges, duplicate and self edges, hubs of out-degree 42, random graphs) replays to the reference's scatter."""
    import ragged_cases
    kw = dict(ragged_cases.CASES[name], ht=5, wd=6)         # the plan does not depend on the grid
    _check_gather_lists(make_problem(iters=1, **kw))


def _check_gather_lists(p):
    f = lambda t: t.double()
    tgt, wgt = f(p.targets).permute(0, 2, 3, 1), f(p.weights).permute(0, 2, 3, 1)
    lin = od.linearize(tgt, wgt, f(p.eta), f(p.poses), f(p.disps), f(p.intrinsics), p.ii, p.jj, p.t0)
    sol = od.schur_solve(lin["H"], lin["E"], lin["C"], lin["v"], lin["w"], ep=p.ep, lm=p.lm)
    blk = lin["blocks"]
    M = len(lin["kx"])
    pl = _plan(p.ii, p.jj, p.disps.shape[0], p.t0, p.t1, False, M)
    dims = pl["dims"]
    Pp, S = dims[4], dims[7]
    assert dims[5] == M and list(pl["kx"]) == lin["kx"].tolist() and S == 1
    # per-edge blocks in CSR order (what k_edge_finalize writes): [c,4,6,6], [c,2,6]
    ce = torch.from_numpy(pl["cedge"].astype(np.int64))
    Hblk = torch.stack([blk["Hii"][ce], blk["Hij"][ce], blk["Hji"][ce], blk["Hjj"][ce]], 1).reshape(-1, 6, 6)
    vblk = torch.stack([blk["vi"][ce], blk["vj"][ce]], 1).reshape(-1, 6)
    # Schur slab (what k_linearize writes): X_AB = E_A Q E_B^T per keyframe slot, pair order A>=B
    Q = 1.0 / lin["C"]
    slab, vE = [], []
    for m in range(M):
        e0, e1 = pl["koff"][m], pl["koff"][m + 1]
        edges = ce[e0:e1]
        panel = [blk["Ei"][edges].sum(0)] + [blk["Ej"][e] for e in edges]           # [6,HW] each
        for A in range(len(panel)):
            for B in range(A + 1):
                slab.append((panel[A] * Q[m]) @ panel[B].T)
            vE.append(panel[A] @ (Q[m] * lin["w"][m]))
    slab = torch.stack(slab)
    vE = torch.stack(vE)
    # replay the gather lists
    Sm = torch.zeros(6 * Pp, 6 * Pp, dtype=torch.float64)
    items = pl["sblk_items"].reshape(-1, 4)
    off = pl["sblk_off"]
    ob = 0
    for a in range(Pp):
        for b in range(a + 1):
            hsum = torch.zeros(6, 6, dtype=torch.float64)
            ssum = torch.zeros(6, 6, dtype=torch.float64)
            seen_schur = False
            for it in items[off[ob]:off[ob + 1]]:
                if it[0] & 1:
                    seen_schur = True
                    X = slab[it[1]]
                    ssum += X.T if it[0] & 2 else X
                else:
                    assert not seen_schur            # H items come first (damping applies to H only)
                    hsum += Hblk[it[1]]
            if a == b:
                hsum = hsum + torch.diag(p.ep + p.lm * torch.diag(hsum))
            Sm[6 * a:6 * a + 6, 6 * b:6 * b + 6] = hsum - ssum
            ob += 1
    assert torch.allclose(torch.tril(Sm), torch.tril(sol["S"]), rtol=1e-9, atol=1e-12)
    rhs = torch.zeros(Pp, 6, dtype=torch.float64)
    ritems = pl["rhs_items"].reshape(-1, 4)
    roff = pl["rhs_off"]
    for a in range(Pp):
        for it in ritems[roff[a]:roff[a + 1]]:
            rhs[a] += -vE[it[1]] if it[0] & 1 else vblk[it[1]]
    assert torch.allclose(rhs.reshape(-1), sol["vs"], rtol=1e-9, atol=1e-12)


def test_plan_rejects_bad_graphs():
    ii = torch.tensor([0, 1, 2]); jj = torch.tensor([1, 2, 5])
    with pytest.raises(RuntimeError, match="outside"):
        _plan(ii, jj, 6, 1, 4, False, 3)                    # jj >= t1
    with pytest.raises(RuntimeError, match="eta"):
        _plan(ii, torch.tensor([1, 2, 3]), 6, 1, 4, False, 2)
    with pytest.raises(RuntimeError, match="t0"):
        _plan(ii, torch.tensor([1, 2, 3]), 6, 4, 4, False, 3)


def test_plan_duplicate_and_self_edges():
    """Duplicate (i,j) edges and i == j edges are legal index patterns: every block lands somewhere."""
    ii = torch.tensor([1, 1, 2, 2, 3]); jj = torch.tensor([2, 2, 1, 2, 1])
    pl = _plan(ii, jj, 4, 1, 4, False, 3, splits=2)
    dims = pl["dims"]
    assert dims[5] == 3 and dims[8] == 2 and dims[7] == 2
    items = pl["sblk_items"].reshape(-1, 4)
    nH = int((items[:, 0] & 1 == 0).sum())
    assert nH == 3 * 4 + 4                                  # lower triangle: 3 blocks per i != j edge, 4 for i == j
    # slab stride = pairs of the keyframe; bases are scaled by the number of splits
    assert pl["pairbase"].tolist() == [0, 6, 12, 15]
