"""The ragged / degenerate factor graphs of tests/test_ba_ragged_gpu.py (and of oracle/make_golden_ragged.py, which runs
the UNMODIFIED reference BA on a subset of them): name -> cut3r_slam_b200.synthetic.make_problem keywords."""
import torch


def chain(P, r=1, skip_src=(), skip_dst=()):
    e = [(i, j) for i in range(P) for j in range(P) if i != j and abs(i - j) <= r
         and i not in skip_src and j not in skip_dst]
    return [a for a, _ in e], [b for _, b in e]


def star(centre, leaves):
    return [centre] * len(leaves), list(leaves)


def degree_mix():
    """out-degree 20 / 21 (narrow | WIDE 1), 25 / 26 (WIDE 1 | WIDE 2), 41 / 42 (WIDE 2 | SIMT) next to a radius-2
    chain (with radius 1 the chain ends hang on ONE edge: see _reference_fp32_gap in the test file)."""
    P = 48
    ii, jj = chain(P, 2)
    for centre, deg in ((3, 20), (4, 21), (5, 25), (6, 26), (7, 41), (8, 42)):
        have = {j for i, j in zip(ii, jj) if i == centre}
        leaves = [j for j in range(P - 1, -1, -1) if j != centre and j not in have][: deg - len(have)]
        a, b = star(centre, leaves)
        ii += a
        jj += b
    return ii, jj


def random_graph(seed):
    g = torch.Generator().manual_seed(seed)
    P = int(torch.randint(3, 25, (1,), generator=g))
    E = int(torch.randint(2, 6 * P, (1,), generator=g))
    ii = torch.randint(0, P, (E,), generator=g)
    jj = (ii + torch.randint(1, P, (E,), generator=g)) % P          # never a self edge; duplicates allowed
    ht = int(torch.randint(3, 40, (1,), generator=g))
    wd = int(torch.randint(3, 50, (1,), generator=g))
    t0 = int(torch.randint(1, min(P - 1, 3) + 1, (1,), generator=g))
    return dict(P=P, ht=ht, wd=wd, t0=t0, seed=seed, edges=(ii, jj))


CASES = {
    # frame 4 is never a source (no disparity row of its own: M = P - 1), frame 6 is never a target
    "sink_and_source": dict(P=9, ht=20, wd=28, t0=1, seed=71, edges=chain(9, 2, skip_src=(4,), skip_dst=(6,))),
    # frame 5 sits inside the window with no edge at all: its pose block is damping only
    "isolated_keyframe": dict(P=9, ht=20, wd=28, t0=2, seed=72, edges=chain(9, 2, skip_src=(5,), skip_dst=(5,))),
    # every edge twice, one of them three times (factor_graph.py never produces this; the scatter sums it)
    "duplicate_edges": dict(P=6, ht=18, wd=26, t0=1, seed=73,
                            edges=tuple(x * 2 + x[:1] for x in map(list, chain(6, 2)))),
    # t0 = 3: edges gauge -> gauge, gauge -> free, free -> gauge all present
    "edges_inside_the_gauge": dict(P=8, ht=20, wd=28, t0=3, seed=74, edges=chain(8, 3)),
    "degree_thresholds": dict(P=48, ht=12, wd=20, t0=1, seed=75, edges=degree_mix()),
    "grid_below_one_tile": dict(P=6, ht=5, wd=7, t0=1, seed=76, edges=chain(6, 2)),
    "grid_exactly_one_tile": dict(P=6, ht=8, wd=16, t0=1, seed=77, edges=chain(6, 2)),
    "grid_one_over_a_tile": dict(P=6, ht=3, wd=43, t0=1, seed=78, edges=chain(6, 2)),
    "single_pair": dict(P=2, ht=16, wd=20, t0=1, seed=79, edges=([0, 1], [1, 0])),
    "one_directed_edge": dict(P=3, ht=16, wd=20, t0=1, seed=80, edges=([1], [2])),
    # ii == jj (projective_ops.py:53 has the stereo special case commented out: a self edge is a plain edge with Gij = 1);
    # all four blocks of the edge land on the same diagonal block of H
    "self_edges": dict(P=4, ht=16, wd=20, t0=1, seed=90, edges=([0, 1, 1, 2, 2, 3, 2, 3], [1, 0, 2, 1, 3, 2, 2, 3])),
}
for _s in range(100, 108):
    CASES[f"random_{_s}"] = random_graph(_s)

# the cases oracle/make_golden_ragged.py runs through the reference itself (tests/golden/ragged.npz)
REFERENCE_PINNED = ["sink_and_source", "isolated_keyframe", "duplicate_edges", "edges_inside_the_gauge", "self_edges",
                    "one_directed_edge", "grid_below_one_tile", "random_101", "random_107"]
REFERENCE_PINNED_MOBA = ["sink_and_source", "isolated_keyframe", "self_edges", "random_101"]
REFERENCE_PINNED_SIM3 = ["sink_and_source", "isolated_keyframe", "duplicate_edges", "self_edges", "random_107"]
# the CUDA pgba is held to the reference on these
GPU_SIM3 = ["sink_and_source", "isolated_keyframe", "duplicate_edges", "self_edges", "random_107"]


def sim3_scales(P, t0):
    """Per-frame scales of the Sim3 cases: 1 +- 5 %, gauge frames at 1."""
    s = 1.0 + 0.05 * (2 * torch.rand(P, 1, generator=torch.Generator().manual_seed(3)) - 1)
    s[:t0] = 1.0
    return s
