"""Developer tool (GPU box): device time of the ops built around `ba` at the backend
configuration (128 keyframes, 2000 edges, 43x77), CUDA events on the launching stream, a 256 MB write between timed
calls (inputs < L2 otherwise), and -- for the streaming ones -- algorithmic bytes / time against the measured HBM peak
(MEASURED_PEAKS.json).  Prints one JSON line per op.  usage: python tools/ops_bench.py [--reps 20]"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from cut3r_slam_b200 import droid_backends as db, glue  # noqa: E402
from cut3r_slam_b200.synthetic import make_problem  # noqa: E402


def timeit(fn, reps, flush):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(reps):
        flush.fill_(1.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2] * 1e-3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--workload", default="backend")
    a = ap.parse_args()
    peak = 6535.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    p = make_problem(a.workload)
    g = p.to("cuda")
    N, ht, wd = g.disps.shape
    HW, E = ht * wd, g.ii.numel()
    M = int(torch.unique(g.ii).numel())
    flush = torch.empty(64 * 1024 * 1024, device="cuda")
    gen = torch.Generator().manual_seed(1)
    rows = []

    def row(name, t, nbytes=None, **kw):
        r = dict(op=name, workload=a.workload, us=round(t * 1e6, 1))
        if nbytes is not None:
            r.update(algorithmic_MB=round(nbytes / 1e6, 2), GBps=round(nbytes / t / 1e9, 1), frac_of_hbm_peak=round(nbytes / t / 1e9 / peak, 3))
        r.update(kw)
        rows.append(r)
        print(json.dumps(r), flush=True)

    # reproject: disps of the source frame read, coords (8 B) + valid (4 B) written per edge-pixel
    row("reproject", timeit(lambda: glue.reproject(g.poses, g.disps, g.intrinsics, g.ii, g.jj), a.reps, flush), E * HW * 16)
    # proj_trans: targets + weights (16 B per edge-pixel) read once, C and w written, disps read per keyframe
    row("proj_trans", timeit(lambda: db.proj_trans(g.poses, g.disps, g.intrinsics, g.targets, g.weights, g.ii, g.jj), a.reps, flush),
        E * HW * 16 + M * HW * 12)
    # convex upsampling of every keyframe's disparity: 576 logits read, 64 values written per coarse pixel
    mask = torch.randn(N, 576, ht, wd, device="cuda")
    disp4 = g.disps[None].contiguous()
    row("upsample_disp", timeit(lambda: glue.upsample_disp(disp4, mask), a.reps, flush), N * HW * (2304 + 4 + 256), frames=N)
    del mask
    # bi_inter at the JDSA grid (8x8 nodes): dense Jacobian written
    hs = ws = 8
    scales = (1 + 0.1 * torch.randn(M, hs, ws, generator=gen)).cuda()
    ys, xs = torch.meshgrid(torch.linspace(0, hs - 1 - 1e-6, ht), torch.linspace(0, ws - 1 - 1e-6, wd), indexing="ij")
    grid = torch.stack((xs, ys), -1)[None].expand(M, -1, -1, -1).contiguous().cuda()
    row("bi_inter", timeit(lambda: db.bi_inter(scales, grid), a.reps, flush), M * HW * (8 + 4 + 4 * hs * ws))
    # JDSA step (proj_trans + per-keyframe mono-prior solve), 8x8 and 2x2 scale grids
    prior = (g.disps * (1 + 0.05 * torch.randn(g.disps.shape, generator=gen).cuda())).contiguous()
    for hs, ws in ((8, 8), (2, 2)):
        dsc = (1 + 0.1 * torch.randn(N, hs, ws, generator=gen)).cuda()
        d0 = g.disps.clone()

        def step():
            db.JDSA(g.targets, g.weights, g.eta, g.poses, d0, g.intrinsics, prior, dsc, g.ii, g.jj, 0.01)
        row(f"JDSA_{hs}x{ws}", timeit(step, a.reps, flush), None, launches=3)
    # companions
    K = min(E, 2000)
    row("frame_distance", timeit(lambda: db.frame_distance(g.poses, g.disps, g.intrinsics, g.ii[:K].contiguous(), g.jj[:K].contiguous(), 0.3),
                                 a.reps, flush), None, pairs=K)
    t, q = p.poses[:, :3].double(), p.poses[:, 3:].double()          # inverse poses (host, not timed): (-R^T t, q*)
    qc = torch.cat((-q[:, :3], q[:, 3:]), 1)
    uv = torch.cross(qc[:, :3], t, dim=1)
    Rt_t = t + 2 * (qc[:, 3:] * uv + torch.cross(qc[:, :3], uv, dim=1))
    Ginv = torch.cat((-Rt_t, qc), 1).float().cuda().contiguous()
    row("iproj", timeit(lambda: db.iproj(Ginv, g.disps, g.intrinsics), a.reps, flush), N * HW * 16)
    ix = torch.arange(N, device="cuda")
    th = torch.full((N,), 0.1, device="cuda")
    row("depth_filter", timeit(lambda: db.depth_filter(g.poses, g.disps, g.intrinsics, ix, th), a.reps, flush), None, frames=N)
    # depth covariance: one linearisation + factorisation + S^-1 + per-pixel quadratic forms
    row("depth_cov", timeit(lambda: db.depth_cov(g.poses, g.disps, g.intrinsics, None, g.targets, g.weights, g.eta, g.ii, g.jj,
                                                 p.t0, p.t1, p.lm, p.ep), a.reps, flush), None, n=6 * (p.t1 - p.t0))
    # pgba: Sim3 pose-graph BA on the same graph (scales 1, 16 relative-pose factors), 2 iterations per call
    poses8 = torch.cat([g.poses, torch.ones(N, 1, device="cuda")], 1).contiguous()
    R = 16
    rel_ii = torch.arange(2, 2 + R, device="cuda")
    rel_jj = rel_ii + 5
    rel_poses = torch.zeros(R, 7, device="cuda"); rel_poses[:, 6] = 1.0
    rel_infos = torch.eye(7, device="cuda").expand(R, 7, 7).contiguous()
    d1 = g.disps.clone()

    def pg():
        db.pgba(poses8, d1, g.intrinsics, g.targets, g.weights, g.eta, g.ii, g.jj, rel_ii, rel_jj, rel_poses, rel_infos,
                p.t0, p.t1, 2, p.lm, p.ep, 1e-5)
    row("pgba_2it", timeit(pg, max(3, a.reps // 4), flush), None, n=7 * (p.t1 - p.t0), rel_factors=R)
    # the other elementwise steps of FactorGraph.update
    tgt = g.targets.permute(0, 2, 3, 1).contiguous()[None]
    row("motion_features", timeit(lambda: glue.motion_features(g.poses, g.disps, g.intrinsics, g.ii, g.jj, tgt), a.reps, flush), E * HW * (4 + 8 + 12 + 16))
    damping = torch.rand_like(g.disps) * 1e-3
    row("damping_eta", timeit(lambda: glue.damping_eta(damping, g.ii), a.reps, flush), M * HW * 8)
    row("covis_distance", timeit(lambda: db.covis_distance(g.poses, g.disps, g.intrinsics, torch.arange(N - 1, device="cuda")), a.reps, flush), None, pairs=N - 1)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", f"ops_bench_{a.workload}.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
