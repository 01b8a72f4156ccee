"""Parity margins table (GPU box): for every golden file and every linearisation variant, the error of
the CUDA fp32 result against the reference's fp64 result, divided by the reference's OWN fp32-vs-fp64
gap on the same inputs; and, per case, which stage owns the dx error: the reduced system the GPU
assembled is solved in fp64 on the host ("assembly-only") and compared with what the device solver
returned from the same system ("solver-only").

    python tools/parity_margins.py > profiles/r2_parity_margins.md
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402

import oracle.dba as od  # noqa: E402
from cut3r_slam_b200 import droid_backends as db  # noqa: E402
from cut3r_slam_b200.synthetic import make_problem  # noqa: E402
from helpers import load_golden, rel_err  # noqa: E402

VARIANTS = {
    "tcgen05": dict(tensor_cores=1, wide_tensor_cores=2),
    "SIMT": dict(tensor_cores=0),
}
WIDE_VARIANTS = {
    "WIDE": dict(tensor_cores=1),      # wide level filled per case
    "SIMT>20": dict(tensor_cores=1, wide_tensor_cores=0),
    "SIMT": dict(tensor_cores=0),
}
DEFAULTS = dict(tensor_cores=1, wide_tensor_cores=2, block_diag_quirk=0, keep_system=0)


def set_opts(o):
    for k, v in {**DEFAULTS, **o}.items():
        db.set_option(k, v)


def stepwise(p, opts, quirk=None):
    """one LM iteration per call; returns disps, [dx], [(assembly-only, solver-only)]"""
    if quirk is None:
        quirk = p.motion_only       # golden MoBA = the reference's block_solve as written (chol.py:35-37)
    set_opts({**opts, "block_diag_quirk": 1 if quirk else 0, "keep_system": 1})
    g = p.to("cuda")
    dxs, split = [], []
    for _ in range(p.iterations):
        dx, dz = db.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj,
                       g.t0, g.t1, 1, g.lm, g.ep, g.motion_only)
        S, rhs, _ = db.debug_system()
        S64 = torch.tril(S.double())
        S64 = S64 + torch.tril(S64, -1).T
        dxs.append(dx.cpu())
        split.append(torch.linalg.solve(S64, rhs.double()))
    set_opts({})
    return g.poses.cpu(), g.disps.cpu(), dxs, split


def row(case, variant, p, ref_dx64, ref_dx32, ref_d64, ref_d32, opts, quirk=None):
    poses, disps, dxs, fromS = stepwise(p, opts, quirk)
    cells = []
    for k, dx in enumerate(dxs):
        gap = rel_err(ref_dx32[k], ref_dx64[k])
        err = rel_err(dx, ref_dx64[k])
        asm = rel_err(fromS[k].reshape(-1), torch.as_tensor(ref_dx64[k]).reshape(-1))
        sol = rel_err(dx.reshape(-1), fromS[k].reshape(-1))
        cells.append(f"{err:.1e} / {gap:.1e} = **{err / max(gap, 1e-30):.2f}x** (asm {asm:.1e}, solve {sol:.1e})")
    gd = rel_err(ref_d32, ref_d64)
    ed = rel_err(disps, ref_d64)
    print(f"| {case} | {variant} | " + "<br>".join(cells) + f" | {ed:.1e} / {gd:.1e} = **{ed / max(gd, 1e-30):.2f}x** |", flush=True)


def main():
    print("# Parity margins: CUDA fp32 error vs fp64 reference, over the reference's own fp32-vs-fp64 gap\n")
    print("`err / gap = ratio` per LM iteration (dx: max-rel on the pose tangent update; disps: max-rel after the last")
    print("iteration). `asm` = dx from the GPU-assembled reduced system solved in fp64 on the host vs the fp64 reference")
    print("(error owned by linearize + Schur + assembly); `solve` = device dx vs that fp64 solve (error owned by the")
    print("device Cholesky).  Golden files: outputs of the UNMODIFIED reference `geom/ba.py` (tests/golden).\n")
    print("| case | variant | dx per iteration | disps |")
    print("|---|---|---|---|")
    for name in ["tiny_ba", "tiny_ba_backend", "tiny_moba", "correctness", "frontend"]:
        p, g = load_golden(name)
        for vname, o in VARIANTS.items():
            if p.motion_only and vname != "SIMT":
                continue        # motion-only always runs the SIMT kernel
            row(name, vname, p, g["ref_dx_f64"], g["ref_dx_f32"], g["ref_disps_f64"], g["ref_disps_f32"], o)
    # graphs with wide keyframes (not in the golden set: the reference fp64 is the oracle restatement here)
    for level, P, r, E in [(1, 40, 11, 760), (2, 44, 16, 1150)]:
        p = make_problem(P=P, E=E, ht=16, wd=24, iters=1, t0=1, r=r, seed=52)
        p64, d64, i64 = od.oracle_ba(*p.ba_args(), dtype=torch.float64, return_info=True)
        p32, d32, i32 = od.oracle_ba(*p.ba_args(), dtype=torch.float32, return_info=True)
        for vname, o in WIDE_VARIANTS.items():
            oo = dict(o)
            if vname == "WIDE":
                oo["wide_tensor_cores"] = level
            row(f"wide{level} (P={P}, E={E}, oracle)", vname if vname != "WIDE" else f"WIDE {level}", p,
                [i[0] for i in i64], [i[0] for i in i32], d64, d32, oo)
    # full-size graphs, 2 iterations (oracle restatement as the fp64 reference)
    for name in sys.argv[1:]:
        p = make_problem(name)
        p.iterations = 2
        a = list(p.ba_args())
        p64, d64, i64 = od.oracle_ba(*a, dtype=torch.float64, return_info=True)
        p32, d32, i32 = od.oracle_ba(*a, dtype=torch.float32, return_info=True)
        for vname, o in VARIANTS.items():
            if p.motion_only and vname != "SIMT":
                continue
            row(f"{name} (full size, oracle)", vname, p, [i[0] for i in i64], [i[0] for i in i32], d64, d32, o, quirk=False)


if __name__ == "__main__":
    main()
