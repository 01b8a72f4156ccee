"""Developer diagnostic: host time per `ba` call (Python wrapper + ctypes + libdba host code) against the device time,
for the small workloads where one of the two bounds the call rate."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from cut3r_slam_b200 import droid_backends as db
from cut3r_slam_b200.synthetic import make_problem

for name in (sys.argv[1:] or ["frontend", "correctness", "motiononly"]):
    p = make_problem(name)
    g = p.to("cuda")
    args = g.ba_args()
    for _ in range(20):
        db.ba(*args)
    torch.cuda.synchronize()
    n = 300
    t0 = time.perf_counter()
    for _ in range(n):
        db.ba(*args)
    t_issue = time.perf_counter() - t0          # host time to ISSUE n calls (the GPU runs behind)
    torch.cuda.synchronize()
    t_all = time.perf_counter() - t0
    # one call at a time: host + device latency
    t1 = time.perf_counter()
    for _ in range(n):
        db.ba(*args)
        torch.cuda.synchronize()
    t_sync = time.perf_counter() - t1
    print(f"{name}: issue {1e6 * t_issue / n:.1f} us/call, back-to-back {1e6 * t_all / n:.1f} us/call, "
          f"call + sync {1e6 * t_sync / n:.1f} us/call, graph replays {db.graph_replays()}")
