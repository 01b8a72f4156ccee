"""CPU restatement (plain Python loops) of the reference's factor-graph index helpers.  TEST
INFRASTRUCTURE ONLY -- nothing under cut3r_slam_b200/ imports this.

Pinned: every function below except `proximity_edges` is checked in tests/test_graph_cpu.py against
tests/golden/graph_utils.npz, produced by oracle/make_golden_graph.py from the reference's own
hislam2/geom/graph_utils.py and hislam2/factor_graph.py run unmodified on CPU.
PARITY UNPINNED: `proximity_edges` restates factor_graph.pyc add_proximity_factors@354-424 (stale
bytecode, no source, cannot run) from SURVEY.md section 8a.
"""
import math


def neighbourhood(t0, t1, r):
    """graph_utils.py:25-29 / factor_graph.py:109-117."""
    ii, jj = [], []
    for i in range(t0, t1):
        for j in range(t0, t1):
            if 1 <= abs(i - j) <= r:
                ii.append(i)
                jj.append(j)
    return ii, jj


def meshgrid(m, n):
    """graph_utils.py:20-22."""
    return [i for i in range(m) for _ in range(n)], [j for _ in range(m) for j in range(n)]


class Edges:
    """FactorGraph's ii / jj / age (factor_graph.py:23-25) as Python lists."""

    def __init__(self, max_factors=-1):
        self.max_factors = max_factors
        self.ii, self.jj, self.age = [], [], []
        self.ii_inac, self.jj_inac = [], []

    def filter_repeated(self, ii, jj):
        """factor_graph.py:29-39."""
        eset = set(zip(self.ii, self.jj))
        keep = [(i, j) not in eset for i, j in zip(ii, jj)]
        return [i for i, k in zip(ii, keep) if k], [j for j, k in zip(jj, keep) if k]

    def add_factors(self, ii, jj, remove=False):
        """factor_graph.py:59-82."""
        ii, jj = self.filter_repeated(list(ii), list(jj))
        if len(ii) == 0:
            return
        if self.max_factors > 0 and len(self.ii) + len(ii) > self.max_factors and remove:
            ix = sorted(range(len(self.age)), key=lambda k: self.age[k])        # stable argsort(age)  :76
            self.rm_factors([ix[k] >= self.max_factors - len(ii) for k in range(len(ix))], store=True)   # :77
        self.ii += ii
        self.jj += jj
        self.age += [0] * len(ii)

    def rm_factors(self, mask, store=False):
        """factor_graph.py:86-90."""
        if store:
            self.ii_inac += [i for i, m in zip(self.ii, mask) if m]
            self.jj_inac += [j for j, m in zip(self.jj, mask) if m]
        self.ii = [i for i, m in zip(self.ii, mask) if not m]
        self.jj = [j for j, m in zip(self.jj, mask) if not m]
        self.age = [a for a, m in zip(self.age, mask) if not m]

    def rm_keyframe(self, ix):
        """factor_graph.py:100-106."""
        m = [(i == ix) or (j == ix) for i, j in zip(self.ii, self.jj)]
        self.ii = [i - 1 if i >= ix else i for i in self.ii]
        self.jj = [j - 1 if j >= ix else j for j in self.jj]
        self.rm_factors(m)

    def add_neighborhood_factors(self, t0, t1, r=3):
        """factor_graph.py:109-117."""
        self.add_factors(*neighbourhood(t0, t1, r))


def proximity_edges(d, t0, t1, t, rad, nms, thresh, max_factors, ex_ii, ex_jj):
    """factor_graph.pyc add_proximity_factors@354-424.  d: list of floats over the
    grid arange(t0,t) x arange(t1,t), modified in place.  Returns the list of selected (i, j)."""
    nj = t - t1
    nd = len(d)

    def kill(i, j):
        idx = (i - t0) * nj + (j - t1)
        if idx < 0:
            idx += nd
        if 0 <= idx < nd:
            d[idx] = math.inf

    def suppress(i, j):
        lim = max(min(abs(i - j) - 2, nms), 0)
        for di in range(-nms, nms + 1):
            for dj in range(-nms, nms + 1):
                if abs(di) + abs(dj) <= lim:
                    i1, j1 = i + di, j + dj
                    if t0 <= i1 < t and t1 <= j1 < t:
                        kill(i1, j1)

    for a in range(t - t0):
        for b in range(nj):
            if (t0 + a) - rad < (t1 + b):                                   # @366
                d[a * nj + b] = math.inf
            if d[a * nj + b] > 100 or d[a * nj + b] != d[a * nj + b]:        # @367
                d[a * nj + b] = math.inf
    for i, j in zip(ex_ii, ex_jj):                                          # @371-379
        suppress(i, j)
    es = []
    for i in range(t0, t):                                                  # @382-387
        for j in range(max(i - rad - 1, 0), i):
            es.append((i, j))
            es.append((j, i))
            kill(i, j)
    for k in sorted(range(nd), key=lambda q: d[q]):                         # @399 (stable)
        if d[k] > thresh:                                                   # @401
            continue
        if len(es) > max_factors:                                           # @404
            break
        i, j = t0 + k // nj, t1 + k % nj
        es.append((i, j))                                                   # @411-412
        es.append((j, i))
        suppress(i, j)                                                      # @414-421
    return es
