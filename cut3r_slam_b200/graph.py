"""Host-side mirror of the reference's factor-graph edge bookkeeping (int64, bit-exact, order-preserving).

Functions mirror hislam2/geom/graph_utils.py:5-29; `EdgeList` mirrors the edge part of `FactorGraph`
(hislam2/factor_graph.py:18-117 and, for `add_proximity_factors`, the stale bytecode
factor_graph.pyc@354-424 restated in SURVEY.md section 8a).  The loops live in libdba.so
(csrc/dba_graph.cpp, C ABI in include/dba.h); this file only marshals arrays.  Tensors are CPU int64
(the reference keeps them on the GPU but walks them with `.item()` per edge, factor_graph.py:33-37).
"""
import ctypes as C

import torch

from . import _lib


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None and t.numel() > 0 else C.c_void_p(0)


def _i64(x):
    return torch.as_tensor(x, dtype=torch.long).reshape(-1).cpu().contiguous()


def graph_to_edge_list(graph):
    """graph_utils.py:5-17: dict u -> iterable of v, in insertion order; kk = position of u."""
    ii, jj, kk = [], [], []
    for s, u in enumerate(graph):
        for v in graph[u]:
            ii.append(u)
            jj.append(v)
            kk.append(s)
    return torch.as_tensor(ii), torch.as_tensor(jj), torch.as_tensor(kk)


def keyframe_indicies(graph):
    """graph_utils.py:18-19."""
    return torch.as_tensor([u for u in graph])


def meshgrid(m, n, device="cpu"):
    """graph_utils.py:20-22 ('ij' order, flattened)."""
    ii = torch.arange(m).repeat_interleave(n)
    jj = torch.arange(n).repeat(m)
    return ii.to(device), jj.to(device)


def neighbourhood_graph(n, r, device="cpu"):
    """graph_utils.py:25-29."""
    lib = _lib.load()
    cnt = lib.dba_graph_neighbourhood(0, n, r, None, None, 0)
    ii, jj = torch.empty(cnt, dtype=torch.long), torch.empty(cnt, dtype=torch.long)
    lib.dba_graph_neighbourhood(0, n, r, _p(ii), _p(jj), cnt)
    return ii.to(device), jj.to(device)


class EdgeList:
    """ii / jj / age of a factor graph (factor_graph.py:23-25) with the reference's update rules."""

    def __init__(self, max_factors=-1):
        self.lib = _lib.load()
        self.max_factors = max_factors
        self.ii = torch.empty(0, dtype=torch.long)
        self.jj = torch.empty(0, dtype=torch.long)
        self.age = torch.empty(0, dtype=torch.long)
        # edges dropped with store=True (the original FactorGraph keeps them as "inactive" factors)
        self.ii_inac = torch.empty(0, dtype=torch.long)
        self.jj_inac = torch.empty(0, dtype=torch.long)

    def __len__(self):
        return int(self.ii.numel())

    def filter_repeated_edges(self, ii, jj):
        """factor_graph.py:29-39."""
        ii, jj = _i64(ii), _i64(jj)
        keep = torch.zeros(ii.numel(), dtype=torch.uint8)
        self.lib.dba_graph_filter_repeated(_p(self.ii), _p(self.jj), len(self), _p(ii), _p(jj), ii.numel(), _p(keep))
        keep = keep.bool()
        return ii[keep], jj[keep]

    def add_factors(self, ii, jj, remove=False):
        """factor_graph.py:59-82."""
        ii, jj = self.filter_repeated_edges(ii, jj)
        if ii.numel() == 0:
            return
        if self.max_factors > 0 and len(self) + ii.numel() > self.max_factors and remove:
            mask = torch.zeros(len(self), dtype=torch.uint8)
            self.lib.dba_graph_evict_mask(_p(self.age), len(self), self.max_factors, ii.numel(), _p(mask))
            self.rm_factors(mask.bool(), store=True)
        self.ii = torch.cat([self.ii, ii])
        self.jj = torch.cat([self.jj, jj])
        self.age = torch.cat([self.age, torch.zeros_like(ii)])

    def rm_factors(self, mask, store=False):
        """factor_graph.py:86-90 (store=True additionally remembers the edges as inactive, as the
        original FactorGraph.rm_factors does for the proximity NMS)."""
        mask = torch.as_tensor(mask, dtype=torch.bool).cpu()
        if store:
            self.ii_inac = torch.cat([self.ii_inac, self.ii[mask]])
            self.jj_inac = torch.cat([self.jj_inac, self.jj[mask]])
        self.ii, self.jj, self.age = self.ii[~mask], self.jj[~mask], self.age[~mask]

    def clear_edges(self):
        """factor_graph.py:55-56."""
        self.rm_factors(self.ii >= 0)

    def rm_keyframe(self, ix):
        """factor_graph.py:100-106 (the keyframe-buffer shuffling :96-99 belongs to the caller)."""
        ii, jj, age = self.ii.clone(), self.jj.clone(), self.age.clone()
        n = self.lib.dba_graph_rm_keyframe(_p(ii), _p(jj), _p(age), len(self), int(ix))
        self.ii, self.jj, self.age = ii[:n], jj[:n], age[:n]

    def add_neighborhood_factors(self, t0, t1, r=3):
        """factor_graph.py:109-117."""
        cnt = self.lib.dba_graph_neighbourhood(t0, t1, r, None, None, 0)
        ii, jj = torch.empty(cnt, dtype=torch.long), torch.empty(cnt, dtype=torch.long)
        self.lib.dba_graph_neighbourhood(t0, t1, r, _p(ii), _p(jj), cnt)
        self.add_factors(ii, jj)

    def add_proximity_factors(self, distance, t, t0=0, t1=0, rad=2, nms=2, thresh=16.0, remove=False):
        """factor_graph.pyc add_proximity_factors@354-424.  `distance(ii, jj)` returns the frame distances
        of the given pairs as a float32 tensor (droid_backends.frame_distance through
        DepthVideo.distance@195-205); `t` is the keyframe counter."""
        ii, jj = meshgrid(t - t0, t - t1)
        ii, jj = ii + t0, jj + t1
        d = distance(ii, jj).to(torch.float32).reshape(-1).cpu().contiguous().clone()
        ex_ii = torch.cat([self.ii, self.ii_inac]).contiguous()
        ex_jj = torch.cat([self.jj, self.jj_inac]).contiguous()
        cap = 2 * (int(ii.numel()) + (t - t0) * (rad + 2)) + 8
        oi, oj = torch.empty(cap, dtype=torch.long), torch.empty(cap, dtype=torch.long)
        n = self.lib.dba_graph_proximity(_p(d), t0, t1, t, rad, nms, float(thresh), int(self.max_factors), _p(ex_ii),
                                         _p(ex_jj), ex_ii.numel(), _p(oi), _p(oj), cap)
        if n > 0:
            self.add_factors(oi[:n], oj[:n], remove)
