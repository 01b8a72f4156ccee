"""Steady-state cost of a NEW edge list per call (what a SLAM loop does): two graphs of one workload alternate, so every
call rebuilds the plan.  Run with DBA_PLAN_TRACE=1 for the host sections.  python tools/cold_plan.py [workload ...]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from cut3r_slam_b200 import droid_backends as db
from cut3r_slam_b200.synthetic import make_problem, CONFIGS

for name in (sys.argv[1:] or ["backend", "loopclosure"]):
    a = make_problem(name).to("cuda")
    b = make_problem(name, seed=CONFIGS[name]["seed"] + 1).to("cuda") if False else None
    # same inputs, second edge order: the reversed list is a different graph key and a different plan
    perm = torch.arange(a.ii.numel() - 1, -1, -1, device="cuda")
    ii2, jj2 = a.ii[perm].contiguous(), a.jj[perm].contiguous()
    tg2, wg2 = a.targets[perm].contiguous(), a.weights[perm].contiguous()
    p0, d0 = a.poses.clone(), a.disps.clone()

    def call(k):
        a.poses.copy_(p0); a.disps.copy_(d0)
        if k & 1:
            db.ba(a.poses, a.disps, a.intrinsics, a.disps_sens, tg2, wg2, a.eta, ii2, jj2, a.t0, a.t1, a.iterations, a.lm, a.ep, False)
        else:
            db.ba(a.poses, a.disps, a.intrinsics, a.disps_sens, a.targets, a.weights, a.eta, a.ii, a.jj, a.t0, a.t1, a.iterations, a.lm, a.ep, False)
    for k in range(4):
        call(k)
    torch.cuda.synchronize()
    n = 8
    t = time.perf_counter()
    for k in range(n):
        call(k)
    torch.cuda.synchronize()
    cold = (time.perf_counter() - t) / n
    for k in range(4):
        call(0)
    torch.cuda.synchronize()
    t = time.perf_counter()
    for k in range(n):
        call(0)
    torch.cuda.synchronize()
    warm = (time.perf_counter() - t) / n
    print(f"{name}: new edge list every call {1e3 * cold:.2f} ms, same edge list {1e3 * warm:.2f} ms (wall clock, state reset included)")
