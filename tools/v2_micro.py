"""Developer experiment: the linearize stage on a graph where EVERY source keyframe takes k_linearize_v2 (no WIDE / SIMT
launch beside it): P keyframes, radius-r neighbourhood, out-degree 2r (<= 20), 43x77 grid.  Prints us per LM iteration.
    [DBA_LIB_PATH=tools/_build/libdba_<variant>.so] python tools/v2_micro.py [P] [r] [splits]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from cut3r_slam_b200 import droid_backends as db
from cut3r_slam_b200.synthetic import make_problem

P = int(sys.argv[1]) if len(sys.argv) > 1 else 148
r = int(sys.argv[2]) if len(sys.argv) > 2 else 10
S = int(sys.argv[3]) if len(sys.argv) > 3 else 0
E = sum(min(P - 1, i + r) - max(0, i - r) for i in range(P))
p = make_problem(P=P, E=E, ht=43, wd=77, iters=8, lm=1e-5, ep=0.01, t0=2, r=r, seed=5)
g = p.to("cuda")
if S:
    db.set_option("splits", S)
for _ in range(3):
    q = p.to("cuda")
    db.ba(*q.ba_args())
torch.cuda.synchronize()
db.set_option("profile", 1)
for _ in range(5):
    q = p.to("cuda")
    db.ba(*q.ba_args())
torch.cuda.synchronize()
st, n = db.stage_times()
db.set_option("profile", 0)
print(f"P={P} r={r} E={E} splits={S}: " + " ".join(f"{k}={1e3 * v / max(n, 1):.1f}us" for k, v in st.items()))
