"""Developer diagnostic (GPU box): who owns the dx error at the benchmarked sizes?  One LM iteration of the named
workloads; the reduced system the GPU assembled (keep_system) is solved (a) in fp64 on the host, (b) by LAPACK in
fp32 on the host, and compared with (c) the device solver's dx and (d) the fp64 / fp32 oracle."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

import oracle.lowmem as ol  # noqa: E402
from cut3r_slam_b200 import droid_backends as db  # noqa: E402
from cut3r_slam_b200.synthetic import make_problem  # noqa: E402
from helpers import rel_err  # noqa: E402

for name in sys.argv[1:] or ["backend", "loopclosure"]:
    p = make_problem(name)
    a = list(p.ba_args())
    a[11] = 1
    _, _, i64 = ol.oracle_ba_lowmem(*a, dtype=torch.float64, return_info=True)
    _, _, i32 = ol.oracle_ba_lowmem(*a, dtype=torch.float32, return_info=True)
    dx64, _, S64, vs64 = i64[0]
    dx32, _, S32, vs32 = i32[0]
    db.set_option("keep_system", 1)
    g = p.to("cuda")
    ga = list(g.ba_args())
    ga[11] = 1
    dx, dz = db.ba(*ga)
    S, rhs, dims = db.debug_system()
    db.set_option("keep_system", 0)
    Sf = torch.tril(S.double())
    Sf = Sf + torch.tril(Sf, -1).T
    x64 = torch.linalg.solve(Sf, rhs.double())
    L32 = torch.linalg.cholesky(Sf.float())
    x32 = torch.cholesky_solve(rhs.float().reshape(-1, 1), L32).double().reshape(-1)
    ref = dx64.reshape(-1)
    ev = torch.linalg.eigvalsh(S64)
    d = torch.diagonal(S64).rsqrt()
    evs = torch.linalg.eigvalsh(S64 * d[:, None] * d[None, :])
    print(f"{name}: n={S.shape[0]} cond(S)={float(ev[-1] / ev[0]):.2e} cond(DSD)={float(evs[-1] / evs[0]):.2e}")
    print(f"  oracle fp32 vs fp64 (the reference's own gap)        : {rel_err(dx32.reshape(-1), ref):.2e}")
    print(f"  S: GPU vs oracle fp64 {rel_err(torch.tril(S.double()), torch.tril(S64)):.2e}; oracle fp32 vs fp64 {rel_err(torch.tril(S32.double()), torch.tril(S64)):.2e}")
    print(f"  GPU system solved in fp64 on the host (assembly-only)  : {rel_err(x64, ref):.2e}")
    print(f"  GPU system solved by LAPACK fp32 on the host vs fp64   : {rel_err(x32, x64):.2e}")
    print(f"  device solver vs fp64 solve of the same system         : {rel_err(dx.cpu().reshape(-1).double(), x64):.2e}")
    print(f"  device dx vs oracle fp64 (total)                       : {rel_err(dx.cpu().reshape(-1), ref):.2e}")
    x64r = torch.linalg.solve(S64.float().double(), vs64.float().double())
    print(f"  exact S rounded to fp32, solved in fp64                : {rel_err(x64r, ref):.2e}", flush=True)
