"""Edge bookkeeping: the C++ helpers behind cut3r_slam_b200/graph.py must be
bit-exact with the reference's own graph_utils.py / FactorGraph outputs (tests/golden/graph_utils.npz,
made by oracle/make_golden_graph.py from the unmodified reference) and with the oracle restatement."""
import ast
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import oracle.graph as og  # noqa: E402
from cut3r_slam_b200 import graph as G  # noqa: E402

GOLD = np.load(os.path.join(ROOT, "tests", "golden", "graph_utils.npz"))


def _eq(t, a):
    return np.array_equal(torch.as_tensor(t).numpy().astype(np.int64), np.asarray(a).astype(np.int64))


@pytest.mark.parametrize("n,r", [(8, 2), (12, 3), (5, 7), (1, 1)])
def test_neighbourhood_graph_matches_reference(n, r):
    ii, jj = G.neighbourhood_graph(n, r)
    assert _eq(ii, GOLD[f"nbr_{n}_{r}_ii"]) and _eq(jj, GOLD[f"nbr_{n}_{r}_jj"])
    oi, oj = og.neighbourhood(0, n, r)
    assert _eq(oi, GOLD[f"nbr_{n}_{r}_ii"]) and _eq(oj, GOLD[f"nbr_{n}_{r}_jj"])


def test_small_helpers_match_reference():
    ii, jj = G.meshgrid(3, 5)
    assert _eq(ii, GOLD["mesh_3_5_ii"]) and _eq(jj, GOLD["mesh_3_5_jj"])
    oi, oj = og.meshgrid(3, 5)
    assert _eq(oi, GOLD["mesh_3_5_ii"]) and _eq(oj, GOLD["mesh_3_5_jj"])
    graph = {4: [1, 2], 0: [3], 7: [], 2: [0, 4, 7]}
    a, b, c = G.graph_to_edge_list(graph)
    assert _eq(a, GOLD["g2e_ii"]) and _eq(b, GOLD["g2e_jj"]) and _eq(c, GOLD["g2e_kk"])
    assert _eq(G.keyframe_indicies(graph), GOLD["kfi"])


def _replay(make):
    """Replays the scripted FactorGraph session of the golden file on `make()` and checks every state."""
    script = ast.literal_eval(str(GOLD["script"]))
    g = make()
    for step, op in enumerate(script):
        if op[0] == "nbr":
            g.add_neighborhood_factors(op[1], op[2], r=op[3])
        elif op[0] == "add":
            g.add_factors(op[1], op[2], remove=bool(op[3]))
        elif op[0] == "age":                       # the golden session overwrote the ages: take its values
            age = GOLD[f"fg_age_bump{op[1]}"]
            g.age = torch.from_numpy(age.copy()) if isinstance(g, G.EdgeList) else list(age.tolist())
        elif op[0] == "rmkf":
            g.rm_keyframe(op[1])
        elif op[0] == "rm_ii_eq":
            g.rm_factors([i == op[1] for i in (g.ii.tolist() if isinstance(g, G.EdgeList) else g.ii)])
        elif op[0] == "clear":
            g.rm_factors([True] * len(g.ii))
        assert _eq(g.ii, GOLD[f"fg_{step}_ii"]), (step, op)
        assert _eq(g.jj, GOLD[f"fg_{step}_jj"]), (step, op)
        assert _eq(g.age, GOLD[f"fg_{step}_age"]), (step, op)
    return g


def test_edge_list_matches_reference_factor_graph():
    g = _replay(lambda: G.EdgeList(max_factors=48))
    assert len(g) == 0


def test_oracle_matches_reference_factor_graph():
    _replay(lambda: og.Edges(max_factors=48))


@pytest.mark.parametrize("seed", range(6))
def test_random_sessions_match_oracle(seed):
    """Random add / evict / remove sequences + proximity selection: C++ == oracle, bit for bit."""
    rng = np.random.default_rng(100 + seed)
    a, b = G.EdgeList(max_factors=40), og.Edges(max_factors=40)
    t = 6
    for it in range(25):
        op = rng.integers(0, 5)
        if op == 0:
            n = int(rng.integers(1, 12))
            ii, jj = rng.integers(0, t, n).tolist(), rng.integers(0, t, n).tolist()
            rem = bool(rng.integers(0, 2))
            a.add_factors(ii, jj, remove=rem)
            b.add_factors(ii, jj, remove=rem)
        elif op == 1 and len(a) > 0:
            age = rng.permutation(len(a)).astype(np.int64) * 3 + 1        # distinct ages: argsort ties are
            a.age, b.age = torch.from_numpy(age.copy()), age.tolist()     # implementation-defined upstream
        elif op == 2 and t > 4:
            ix = int(rng.integers(0, t))
            a.rm_keyframe(ix)
            b.rm_keyframe(ix)
            t -= 1
        elif op == 3:
            t += 1
            lo = max(t - 4, 0)
            a.add_neighborhood_factors(lo, t, r=2)
            b.add_neighborhood_factors(lo, t, r=2)
        else:
            t0, t1 = max(t - 5, 0), max(t - 12, 0)
            nd = (t - t0) * (t - t1)
            base = rng.permutation(nd).astype(np.float32) * 0.37 + 0.05   # distinct distances (no sort ties)
            base[rng.integers(0, nd, 3)] = 1000.0                         # frame_distance sentinel
            a.add_proximity_factors(lambda ii, jj: torch.from_numpy(base.copy()), t, t0=t0, t1=t1, rad=2, nms=2,
                                    thresh=12.0, remove=True)
            es = og.proximity_edges(base.astype(np.float64).tolist(), t0, t1, t, 2, 2, 12.0, b.max_factors,
                                    b.ii + b.ii_inac, b.jj + b.jj_inac)
            if es:
                b.add_factors([e[0] for e in es], [e[1] for e in es], remove=True)
        assert a.ii.tolist() == b.ii and a.jj.tolist() == b.jj and a.age.tolist() == b.age, (it, op)
        assert a.ii_inac.tolist() == b.ii_inac and a.jj_inac.tolist() == b.jj_inac
