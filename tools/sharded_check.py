"""Multi-GPU check (run under torchrun, one rank per GPU): sharded global BA over N GPUs == the same
`ba` call on one GPU, poses bit-identical across ranks.  Prints one line per workload.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
      tools/sharded_check.py [workload ...]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch
import torch.distributed as dist

from cut3r_slam_b200 import droid_backends as db
from cut3r_slam_b200.sharded import ShardedBA, partition_by_source
from cut3r_slam_b200.synthetic import make_problem
from helpers import rel_err


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    names = sys.argv[1:] or ["frontend", "backend"]
    sh = ShardedBA(dev)
    sh.init_comm()
    # single-GPU results of the whole graphs on rank 0, computed AFTER the communicator exists: ShardedBA owns a
    # private libdba context, plain droid_backends.ba calls must neither all-reduce nor hang (sharded.py:51)
    single = {}
    if rank == 0:
        for name in names:
            s = make_problem(name).to(dev)
            db.ba(*s.ba_args())
            torch.cuda.synchronize()
            single[name] = (s.poses.clone(), s.disps.cpu())
    dist.barrier()
    ok = True
    for name in names:
        p = make_problem(name)
        g = p.to(dev)
        shards = partition_by_source(p.ii, world)
        mine = shards[rank]
        sel = mine["edges"].to(dev)
        lo, hi = mine["slots"]
        sh.ba(g.poses, g.disps, g.intrinsics, None, g.targets[sel].contiguous(), g.weights[sel].contiguous(),
              g.eta[lo:hi].contiguous(), g.ii[sel].contiguous(), g.jj[sel].contiguous(), p.t0, p.t1, p.iterations,
              p.lm, p.ep)
        sh.gather_disps(g.disps, shards)
        torch.cuda.synchronize()
        ref = g.poses.clone()                                   # poses must be bit-identical on every rank
        dist.broadcast(ref, src=0)
        same = torch.tensor([int(torch.equal(ref, g.poses))], device=dev)
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        if rank == 0:
            sp, sd = single[name]
            ep = float((sp - g.poses).abs().max())
            ed = rel_err(g.disps.cpu(), sd)
            good = bool(same.item()) and ep < 1e-4 and ed < 2e-4
            ok = ok and good
            print(f"sharded_check {name}: world={world} poses identical across ranks={bool(same.item())} "
                  f"|poses - 1gpu|max={ep:.2e} disps rel={ed:.2e} -> {'OK' if good else 'FAIL'}", flush=True)
        dist.barrier()
    sh.close()
    dist.destroy_process_group()
    if rank == 0 and not ok:
        sys.exit(1)


if __name__ == "__main__":
    main()
