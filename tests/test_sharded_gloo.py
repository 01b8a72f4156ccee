"""World-size-2 checks of the sharded global BA host logic on CPU (gloo): the partition by source
keyframe, the additivity of the reduced pose system over shards (H, v, E Q E^T and
E Q w are sums over source keyframes) and the rank rule for the LM damping (every rank damps its share of
H with lm, rank 0 adds ep once).  The per-shard systems come from the CPU oracle (test infrastructure);
the CUDA path runs the same algebra with one ncclAllReduce (tests/test_ba_gpu.py, tools/sharded_check.py).
"""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import oracle.dba as od  # noqa: E402
from cut3r_slam_b200.sharded import partition_by_source  # noqa: E402
from cut3r_slam_b200.synthetic import make_problem  # noqa: E402

KW = dict(P=10, E=44, ht=12, wd=16, iters=1, t0=2, r=3, seed=91)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _shard_system(p, sel, lo, hi, rank):
    """Reduced system of one shard, fp64: S_r = damp_r(H_r) - E_r Q_r E_r^T, rhs_r = v_r - E_r Q_r w_r."""
    tgt = p.targets[sel].permute(0, 2, 3, 1).double()
    wgt = p.weights[sel].permute(0, 2, 3, 1).double()
    lin = od.linearize(tgt, wgt, p.eta[lo:hi].double(), p.poses.double(), p.disps.double(), p.intrinsics.double(),
                       p.ii[sel], p.jj[sel], p.t0)
    sol = od.schur_solve(lin["H"], lin["E"], lin["C"], lin["v"], lin["w"], ep=(p.ep if rank == 0 else 0.0), lm=p.lm)
    return sol["S"], sol["vs"], lin


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        p = make_problem(**KW)
        shards = partition_by_source(p.ii, world)
        mine = shards[rank]
        lo, hi = mine["slots"]
        S, vs, lin = _shard_system(p, mine["edges"], lo, hi, rank)
        buf = torch.cat([S.reshape(-1), vs.reshape(-1)])
        dist.all_reduce(buf)                                   # the one exchange step of an LM iteration
        n = S.shape[0]
        S_all, vs_all = buf[:n * n].reshape(n, n), buf[n * n:]
        dx = torch.linalg.solve(S_all, vs_all)                 # replicated solve: identical on every rank
        # rank-local back-substitution of the shard's disparity rows (chol.py:70)
        Pp, M, D, HW = lin["E"].shape
        Em = lin["E"].permute(0, 2, 1, 3).reshape(Pp * D, M * HW)
        Q = (1.0 / lin["C"]).reshape(-1)
        dz = (Q * (lin["w"].reshape(-1) - Em.T @ dx)).reshape(M, HW)
        # every rank must hold the same dx bit for bit
        gathered = [torch.zeros_like(dx) for _ in range(world)]
        dist.all_gather(gathered, dx)
        same = all(torch.equal(g, gathered[0]) for g in gathered)
        torch.save(dict(dx=dx, dz=dz, slots=(lo, hi), same=same, keyframes=mine["keyframes"]), f"{out}.{rank}")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_reduced_system_sums_to_the_global_one(tmp_path, world):
    out = str(tmp_path / "shard")
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    p = make_problem(**KW)
    lin = od.linearize(p.targets.permute(0, 2, 3, 1).double(), p.weights.permute(0, 2, 3, 1).double(), p.eta.double(),
                       p.poses.double(), p.disps.double(), p.intrinsics.double(), p.ii, p.jj, p.t0)
    sol = od.schur_solve(lin["H"], lin["E"], lin["C"], lin["v"], lin["w"], ep=p.ep, lm=p.lm)
    covered = []
    for r in range(world):
        rec = torch.load(f"{out}.{r}")
        assert rec["same"]
        assert float((rec["dx"] - sol["dx"].reshape(-1)).abs().max()) < 1e-9 * float(sol["dx"].abs().max()) + 1e-12
        lo, hi = rec["slots"]
        assert float((rec["dz"] - sol["dz"][lo:hi]).abs().max()) < 1e-9 * float(sol["dz"].abs().max()) + 1e-12
        covered += list(range(lo, hi))
        assert torch.equal(rec["keyframes"], lin["kx"][lo:hi])
    assert covered == list(range(lin["kx"].numel()))           # every source keyframe owned exactly once


def test_partition_properties():
    p = make_problem(P=40, E=300, ht=4, wd=4, iters=1, t0=2, r=4, seed=92)
    for world in (1, 2, 4, 8):
        shards = partition_by_source(p.ii, world)
        assert len(shards) == world
        edges = torch.cat([s["edges"] for s in shards])
        assert torch.equal(torch.sort(edges).values, torch.arange(p.ii.numel()))     # a partition of the edges
        kx = torch.unique(p.ii)
        assert torch.equal(torch.cat([s["keyframes"] for s in shards]), kx)          # contiguous, in order
        for s in shards:
            assert torch.equal(torch.unique(p.ii[s["edges"]]), s["keyframes"])      # disps rows stay rank-local
            assert bool((s["edges"][1:] > s["edges"][:-1]).all())                    # original edge order kept
        cost = [float(((torch.bincount(p.ii[s["edges"]]).double() + 1) ** 2).sum()) for s in shards if s["edges"].numel()]
        assert max(cost) < 2.5 * (sum(cost) / len(cost))                             # balanced on Schur cost


def test_partition_of_ragged_graphs():
    """More ranks than source keyframes, a single edge, source keyframes with gaps: every edge lands on exactly one rank,
    shards stay contiguous in sorted(unique(ii)), ranks without work get an empty (not a missing) shard -- such a rank
    still calls `ba` with E = 0 and takes part in the all-reduce of the reduced system."""
    import ragged_cases
    for name in ("one_directed_edge", "single_pair", "sink_and_source", "isolated_keyframe", "self_edges", "random_105"):
        ii = torch.as_tensor(ragged_cases.CASES[name]["edges"][0], dtype=torch.long)
        kx = torch.unique(ii)
        for world in (1, 2, 3, 8):
            shards = partition_by_source(ii, world)
            assert len(shards) == world
            edges = torch.cat([s["edges"] for s in shards])
            assert torch.equal(torch.sort(edges).values, torch.arange(ii.numel()))
            assert torch.equal(torch.cat([s["keyframes"] for s in shards]), kx)
            lo = 0
            for s in shards:
                assert s["slots"][0] == lo and s["slots"][1] == lo + s["keyframes"].numel()
                lo = s["slots"][1]
                assert torch.equal(torch.unique(ii[s["edges"]]), s["keyframes"])
            assert lo == kx.numel()
