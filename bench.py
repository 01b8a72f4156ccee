#!/usr/bin/env python
"""bench.py -- dense bundle adjustment throughput on B200 (BASELINE.json metric).

A "step" is ONE `ba` call on the named workload: `iterations` Levenberg-Marquardt iterations over the
whole factor graph (backend global BA: 128 keyframes, 2000 edges, 43x77 grid, 8 iterations).
  value : edge-pixels/s = E*ht*wd*iterations / device time, inputs resident in HBM (poses/disps are
          restored from a pristine device copy inside the timed region: the op works in place)
  e2e   : same metric through the host-buffer entry (dba_ba_host): pinned host inputs are copied
          to the device, the iterations run, poses + disps are copied back, all inside the timed region
  N > 1 : the SAME graph sharded by source keyframe over N GPUs (strong scaling), one NCCL all-reduce
          of the reduced pose system per LM iteration, time = max over ranks.
`--impl reference` times the reference's CPU algorithm (hislam2/geom/ba.py restated in oracle/dba.py
-- the reference's own Python cannot travel to the GPU box) on the host cores.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

WORKLOAD = "backend"


def algorithmic_bytes(E, HW, M, Pp, motion_only=False):
    """bytes one LM iteration must move. 20 B per edge-pixel (targets 8 + weights 8 +
    disps[ii] 4), 12 B per source-keyframe pixel (eta 4 + disps read-modify-write 8), the reduced
    system written and read once, poses read + written."""
    b = 20 * E * HW + 8 * (6 * Pp) ** 2 + 56 * Pp
    if not motion_only:
        b += 12 * M * HW
    return b


def linearize_bytes(E, HW, M):
    """Algorithmic bytes of ONE launch of the dominant kernel (k_linearize): the per-edge-pixel
    stream (20 B) plus eta in and Q, w out per source-keyframe pixel (12 B)."""
    return 20 * E * HW + 12 * M * HW


def workload_string(name, cfg, p):
    return (f"{name}: {cfg['P']} keyframes, {cfg['E']} edges, {cfg['ht']}x{cfg['wd']} grid, "
            f"{p.iterations} LM iterations per ba call, lm={p.lm}, ep={p.ep}, t0={p.t0}")


def measured_traffic(name, world):
    """DRAM bytes of one launch of the dominant kernel from the committed `ncu --set full` capture
    (profiles/linearize_traffic.json), or None when there is none for this workload / GPU count."""
    try:
        rec = json.load(open(os.path.join(ROOT, "profiles", "linearize_traffic.json")))
        return rec.get(f"{name}@{world}")
    except Exception:
        return None


class ClockSampler(threading.Thread):
    """Polls SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._stop_ev = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop_ev.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.02)

    def stop(self):
        self._stop_ev.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def run_reference(args, rank):
    """The reference's CPU algorithm (oracle port) on the host cores; rank 0 only."""
    if rank != 0:
        return
    import oracle.dba as od
    from cut3r_slam_b200.synthetic import CONFIGS, make_problem
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = CONFIGS[WORKLOAD]
    p = make_problem(WORKLOAD)
    sample = "1 LM iteration of the full workload graph per step"
    a = list(p.ba_args())
    a[11] = 1
    t = time.perf_counter()
    od.oracle_ba(*a, dtype=torch.float32)
    probe = time.perf_counter() - t
    if probe * (args.steps + args.warmup) > 240:
        # bounded sample: the sub-graph induced by the first 48 keyframes (dense E shrinks cubically)
        keep = (p.ii < 48) & (p.jj < 48)
        M = int(torch.unique(p.ii[keep]).numel())
        a = [p.poses[:48].clone(), p.disps[:48].clone(), p.intrinsics, None, p.targets[keep].contiguous(),
             p.weights[keep].contiguous(), p.eta[:M].contiguous(), p.ii[keep], p.jj[keep], p.t0, 48, 1, p.lm, p.ep, False]
        sample = f"1 LM iteration of the sub-graph on the first 48 keyframes ({int(keep.sum())} edges) per step"
    E = a[7].numel()
    HW = cfg["ht"] * cfg["wd"]
    for _ in range(args.warmup):
        od.oracle_ba(*a, dtype=torch.float32)
    t = time.perf_counter()
    for _ in range(args.steps):
        od.oracle_ba(*a, dtype=torch.float32)
    dt = (time.perf_counter() - t) / max(args.steps, 1)
    val = E * HW * 1 / dt
    line = {
        "impl": "reference", "metric": "ba_edge_pixels_per_s", "value": val, "unit": "edge-pixels/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_string(WORKLOAD, cfg, p), "parallelism": "host CPU, all threads",
                   "sample": sample},
        "ba_iters_per_s": 1.0 / dt,
        "cpu_baseline": {"value": val, "unit": "edge-pixels/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "edge-pixels/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def sharded_parity(db, sh, p, world, rank, dev, shards):
    """Correctness signal for the N > 1 lines (outside any timed region): the sharded call against the same `ba`
    on one GPU (rank 0, whole graph), and whether the replicated poses are bit-identical on every rank."""
    import torch.distributed as dist
    single = None
    if rank == 0:
        s = p.to(dev)
        db.ba(*s.ba_args())
        torch.cuda.synchronize()
        single = (s.poses.clone(), s.disps.clone())
        del s
    g = p.to(dev)
    mine = shards[rank]
    sel = mine["edges"].to(dev)
    lo, hi = mine["slots"]
    sh.ba(g.poses, g.disps, g.intrinsics, None, g.targets[sel].contiguous(), g.weights[sel].contiguous(),
          g.eta[lo:hi].contiguous(), g.ii[sel].contiguous(), g.jj[sel].contiguous(), p.t0, p.t1, p.iterations, p.lm, p.ep)
    sh.gather_disps(g.disps, shards)
    torch.cuda.synchronize()
    ref = g.poses.clone()
    dist.broadcast(ref, src=0)
    same = torch.tensor([int(torch.equal(ref, g.poses))], device=dev)
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    out = None
    if rank == 0:
        sp, sd = single
        out = {"max_abs_poses_vs_1gpu": float((sp - g.poses).abs().max()),
               "disps_rel_vs_1gpu": float((sd - g.disps).abs().max() / sd.abs().max()),
               "poses_identical_across_ranks": bool(same.item())}
    dist.barrier()
    return out


def run_workload(name, args, db, sh, world, rank, dev, steps, warmup, flush, full):
    """Times one workload; returns the dict of measurements (rank 0) -- `full` adds e2e / cold-plan / parity."""
    import torch.distributed as dist
    from cut3r_slam_b200.sharded import partition_by_source
    from cut3r_slam_b200.synthetic import CONFIGS, make_problem
    cfg = CONFIGS[name]
    p = make_problem(name)
    E, HW, Pp = p.ii.numel(), p.disps.shape[1] * p.disps.shape[2], p.t1 - p.t0
    M = int(torch.unique(p.ii).numel())
    iters = p.iterations
    shards = None
    if world > 1:
        shards = partition_by_source(p.ii, world)
        mine = shards[rank]
        sel = mine["edges"]
        lo, hi = mine["slots"]
    else:
        sel = torch.arange(E)
        lo, hi = 0, M
    h_poses, h_disps = p.poses.pin_memory(), p.disps.pin_memory()
    h_targets, h_weights = p.targets[sel].contiguous().pin_memory(), p.weights[sel].contiguous().pin_memory()
    h_eta = p.eta[lo:hi].contiguous().pin_memory()
    h_ii, h_jj = p.ii[sel].contiguous(), p.jj[sel].contiguous()
    h_intr = p.intrinsics.pin_memory()
    poses0, disps0 = h_poses.to(dev), h_disps.to(dev)
    poses, disps = poses0.clone(), disps0.clone()
    intr, targets, weights, eta = h_intr.to(dev), h_targets.to(dev), h_weights.to(dev), h_eta.to(dev)
    ii, jj = h_ii.to(dev), h_jj.to(dev)

    def step_device():
        poses.copy_(poses0)
        disps.copy_(disps0)
        if world > 1:       # the step ends with complete disparities on every rank
            sh.ba(poses, disps, intr, None, targets, weights, eta, ii, jj, p.t0, p.t1, iters, p.lm, p.ep)
            sh.gather_disps(disps, shards)
        else:
            db.ba(poses, disps, intr, None, targets, weights, eta, ii, jj, p.t0, p.t1, iters, p.lm, p.ep, p.motion_only)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, nsteps, nwarm):
        for _ in range(nwarm):
            flush.fill_(1.0)
            fn()
        barrier()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(nsteps)]
        launches = 0
        for a, b in evs:
            flush.fill_(1.0)                   # L2 flush between timed steps (outside the event pair)
            if world > 1:
                dist.barrier()
            a.record()
            fn()
            b.record()
            launches += sh.last_launch_count() if world > 1 else db.last_launch_count()
        barrier()
        ms = sum(a.elapsed_time(b) for a, b in evs)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / nsteps, launches

    sampler = ClockSampler(dev.index)
    sampler.start()
    ms_step, launches = timed(step_device, steps, warmup)
    clocks = sampler.stop()
    out = {"name": name, "cfg": cfg, "p": p, "E": E, "HW": HW, "M": M, "Pp": Pp, "iters": iters, "ms_step": ms_step,
           "launches": launches, "clocks": clocks, "E_loc": int(sel.numel()), "M_loc": hi - lo}

    # ---- end to end: pinned host buffers in, poses + disps out, through dba_ba_host ------------------
    if full:
        w_poses, w_disps = h_poses.clone().pin_memory(), h_disps.clone().pin_memory()
        ctx = sh.ctx if world > 1 else None

        def step_host():
            w_poses.copy_(h_poses)
            w_disps.copy_(h_disps)
            db.ba_host(w_poses, w_disps, h_intr, None, h_targets, h_weights, h_eta, h_ii, h_jj, p.t0, p.t1, iters,
                       p.lm, p.ep, p.motion_only, device=dev, _ctx=ctx)

        out["ms_e2e"], _ = timed(step_host, max(3, min(steps, 10)), 3)
        out["h2d"] = sum(t.numel() * t.element_size() for t in (h_poses, h_disps, h_intr, h_targets, h_weights, h_eta))
        out["d2h"] = sum(t.numel() * t.element_size() for t in (h_poses, h_disps)) + Pp * 6 * 4

    # ---- cold plan: the edge list changes on every call (a real SLAM loop edits the graph between updates), so
    # each call copies ii / jj to the host, rebuilds the plan (CSR + gather lists) and uploads it
    if full and world == 1:
        perm = torch.randperm(E, generator=torch.Generator().manual_seed(7))
        alt = [(ii, jj, targets, weights), (ii[perm.to(dev)].contiguous(), jj[perm.to(dev)].contiguous(),
                                            targets[perm.to(dev)].contiguous(), weights[perm.to(dev)].contiguous())]
        flip = [0]

        def step_cold():
            poses.copy_(poses0)
            disps.copy_(disps0)
            a = alt[flip[0] & 1]
            flip[0] += 1
            db.ba(poses, disps, intr, None, a[2], a[3], eta, a[0].clone(), a[1].clone(), p.t0, p.t1, iters, p.lm, p.ep,
                  p.motion_only)

        out["ms_cold"], _ = timed(step_cold, max(3, min(steps, 6)), 2)
        del alt

    # ---- per-stage device times (CUDA events inside libdba, on the launching stream) ------------
    setopt = (lambda k, v: sh.ctx.set_option(k, v)) if world > 1 else (lambda k, v: db.set_option(k, v, dev))
    setopt("profile", 1)
    timed(step_device, max(3, min(steps, 10)), 1)
    stages, n_it = sh.stage_times() if world > 1 else db.stage_times(dev)
    setopt("profile", 0)
    out["stage_ms"] = {k: v / max(n_it, 1) for k, v in stages.items()}          # per LM iteration
    out["stage_all"] = None
    if world > 1:                                                         # every rank's stage times (diagnosis of skew)
        keys = sorted(out["stage_ms"])
        mine_t = torch.tensor([out["stage_ms"][k] for k in keys], dtype=torch.float64, device=dev)
        allt = [torch.zeros_like(mine_t) for _ in range(world)]
        dist.all_gather(allt, mine_t)
        out["stage_all"] = {k: [round(float(t[i]) * 1e3, 1) for t in allt] for i, k in enumerate(keys)}
        out["sharded_parity"] = sharded_parity(db, sh, p, world, rank, dev, shards)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=WORKLOAD)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-loopclosure", action="store_true", help="skip the secondary loop-closure object")
    ap.add_argument("--opt", action="append", default=[], help="libdba option name=value (developer experiments)")
    ap.add_argument("--brief", action="store_true", help="print only ms/step and the per-stage times")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch.distributed as dist
    from cut3r_slam_b200 import droid_backends as db
    from cut3r_slam_b200.sharded import ShardedBA

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    sh = ShardedBA(dev)
    if world > 1:
        sh.init_comm()
    for o in args.opt:
        k_, v_ = o.split("=")
        db.set_option(k_, float(v_), dev)
        sh.ctx.set_option(k_, float(v_))
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)    # > 126 MB L2

    r = run_workload(args.workload, args, db, sh, world, rank, dev, args.steps, args.warmup, flush, full=not args.brief)
    lc = None
    if not args.brief and not args.no_loopclosure and args.workload != "loopclosure":
        # BASELINE.json configs[3] (512 keyframes, 12k edges): the configuration where sharding pays; fewer steps
        lc = run_workload("loopclosure", args, db, sh, world, rank, dev, max(3, min(args.steps, 5)), 3, flush, full=False)

    p, cfg, E, HW, M, Pp, iters = r["p"], r["cfg"], r["E"], r["HW"], r["M"], r["Pp"], r["iters"]
    ms_step, stage_ms = r["ms_step"], r["stage_ms"]
    lin_bytes = linearize_bytes(r["E_loc"], HW, r["M_loc"])
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    lin_ms = max(stage_ms["linearize"], 1e-9)
    achieved = lin_bytes / (lin_ms * 1e-3) / 1e9
    b_alg = algorithmic_bytes(E, HW, M, Pp, p.motion_only)
    whole = b_alg * iters / (ms_step * 1e-3) / 1e9

    if rank == 0 and args.brief:
        print(json.dumps({"opts": args.opt, "ms_per_step": round(ms_step, 4),
                          "stage_us": {k: round(v * 1e3, 1) for k, v in stage_ms.items()},
                          "stage_us_per_rank": r["stage_all"]}), flush=True)
    elif rank == 0:
        ep_per_s = E * HW * iters / (ms_step * 1e-3)
        ms_e2e = r["ms_e2e"]
        line = {
            "metric": "ba_edge_pixels_per_s", "value": ep_per_s, "unit": "edge-pixels/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_string(args.workload, cfg, p),
                       "parallelism": f"edges sharded by source keyframe over {world} GPU(s); the step ends with the "
                                      f"disparities gathered on every rank" if world > 1 else "1 GPU",
                       "l2": "256 MB buffer written between timed steps (inputs 106 MB < 126 MB L2)"},
            "ba_iters_per_s": iters / (ms_step * 1e-3),
            "clocks": r["clocks"],
            "e2e": {"value": E * HW * iters / (ms_e2e * 1e-3), "unit": "edge-pixels/s", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"],
                    "entry": "dba_ba_host (C ABI, pinned host buffers)"},
            "gpu_launches": r["launches"],
            "roofline": {"bound": "hbm", "kernel": "k_linearize (residual/Jacobian/Hessian + Schur, fused)",
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "peak_source": peak_src, "algorithmic_bytes_per_launch": lin_bytes,
                         "ms_per_launch": lin_ms, "traffic": measured_traffic(args.workload, world),
                         "limiter": "instruction issue, not DRAM (ncu, profiles/): the kernel's bytes are HBM-type "
                                    "work by SURVEY 8(d), its measured bound is the SIMT front end",
                         "whole_iteration": {"algorithmic_bytes": b_alg, "achieved": whole, "frac": whole / peak,
                                             "frac_of_8TBs_nominal": whole / 8000.0}},
            "stage_ms_per_lm_iteration": stage_ms,
            "stage_us_per_rank": r["stage_all"],
        }
        if "ms_cold" in r:
            line["cold_plan"] = {"ms_per_step": r["ms_cold"],
                                 "what": "every call sees a NEW edge list: ii/jj to the host (one sync), plan rebuild, upload"}
        if r.get("sharded_parity") is not None:
            line["sharded_parity"] = r["sharded_parity"]
        if lc is not None:
            lp = lc["p"]
            line["loopclosure"] = {
                "workload": workload_string("loopclosure", lc["cfg"], lp), "ms_per_step": lc["ms_step"],
                "value": lc["E"] * lc["HW"] * lc["iters"] / (lc["ms_step"] * 1e-3), "unit": "edge-pixels/s",
                "steps": max(3, min(args.steps, 5)), "gpu_launches": lc["launches"],
                "stage_ms_per_lm_iteration": lc["stage_ms"], "stage_us_per_rank": lc["stage_all"],
                "sharded_parity": lc.get("sharded_parity"),
                "l2": "inputs 590 MB > 126 MB L2"}
        if not args.no_cpu_baseline and world == 1:
            import oracle.dba as od
            cores = os.cpu_count() or 1
            torch.set_num_threads(cores)
            a = list(p.ba_args())
            a[11] = 2
            t = time.perf_counter()
            od.oracle_ba(*a, dtype=torch.float32)
            dt = time.perf_counter() - t
            line["cpu_baseline"] = {"value": E * HW * 2 / dt, "unit": "edge-pixels/s", "cores": cores, "kind": "port",
                                    "sample": "2 LM iterations of the same graph, oracle/dba.py (restated "
                                              "hislam2/geom/ba.py BA) in fp32 on all host threads",
                                    "seconds": dt}
        print(json.dumps(line), flush=True)
    if world > 1:
        sh.close()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
