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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=WORKLOAD)
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
    from cut3r_slam_b200.sharded import ShardedBA, partition_by_source
    from cut3r_slam_b200.synthetic import CONFIGS, make_problem

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    for o in args.opt:
        k_, v_ = o.split("=")
        db.set_option(k_, float(v_), dev)
    cfg = CONFIGS[args.workload]
    p = make_problem(args.workload)
    E, HW, Pp = p.ii.numel(), p.disps.shape[1] * p.disps.shape[2], p.t1 - p.t0
    M = int(torch.unique(p.ii).numel())
    iters = p.iterations

    # ---- shard (N > 1) or take the whole graph -------------------------------------------------
    sh = ShardedBA(dev)
    if world > 1:
        sh.init_comm()
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
    sens = None
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)    # > 126 MB L2

    def step_device():
        poses.copy_(poses0)
        disps.copy_(disps0)
        db.ba(poses, disps, intr, sens, targets, weights, eta, ii, jj, p.t0, p.t1, iters, p.lm, p.ep, p.motion_only)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            flush.fill_(1.0)
            fn()
        barrier()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        launches = 0
        for a, b in evs:
            flush.fill_(1.0)                   # L2 flush between timed steps (outside the event pair)
            if world > 1:
                dist.barrier()
            a.record()
            fn()
            b.record()
            launches += db.last_launch_count()
        barrier()
        ms = sum(a.elapsed_time(b) for a, b in evs)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / steps, launches

    sampler = ClockSampler(local_rank)
    sampler.start()
    ms_step, launches = timed(step_device, args.steps, args.warmup)
    clocks = sampler.stop()

    # ---- end to end: pinned host buffers in, poses + disps out, through dba_ba_host ------------------
    w_poses, w_disps = h_poses.clone().pin_memory(), h_disps.clone().pin_memory()

    def step_host():
        w_poses.copy_(h_poses)
        w_disps.copy_(h_disps)
        db.ba_host(w_poses, w_disps, h_intr, None, h_targets, h_weights, h_eta, h_ii, h_jj, p.t0, p.t1, iters, p.lm,
                   p.ep, p.motion_only, device=dev)

    e2e_steps = max(3, min(args.steps, 10))
    ms_e2e, _ = timed(step_host, e2e_steps, 3)
    h2d = sum(t.numel() * t.element_size() for t in (h_poses, h_disps, h_intr, h_targets, h_weights, h_eta))
    d2h = sum(t.numel() * t.element_size() for t in (h_poses, h_disps)) + Pp * 6 * 4

    # ---- per-stage device times (CUDA events inside libdba, on the launching stream) ------------
    db.set_option("profile", 1, dev)
    prof_steps = max(3, min(args.steps, 10))
    timed(step_device, prof_steps, 1)
    stages, n_it = db.stage_times(dev)
    db.set_option("profile", 0, dev)
    stage_ms = {k: v / max(n_it, 1) for k, v in stages.items()}          # per LM iteration
    stage_all = None
    if world > 1:                                                         # every rank's stage times (diagnosis of skew)
        keys = sorted(stage_ms)
        mine_t = torch.tensor([stage_ms[k] for k in keys], dtype=torch.float64, device=dev)
        allt = [torch.zeros_like(mine_t) for _ in range(world)]
        dist.all_gather(allt, mine_t)
        stage_all = {k: [round(float(t[i]) * 1e3, 1) for t in allt] for i, k in enumerate(keys)}
    E_loc = int(sel.numel())
    lin_bytes = linearize_bytes(E_loc, HW, hi - lo)
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
        print(json.dumps({"opts": args.opt, "ms_per_step": round(ms_step, 4), "ms_e2e": round(ms_e2e, 4),
                          "stage_us": {k: round(v * 1e3, 1) for k, v in stage_ms.items()},
                          "stage_us_per_rank": stage_all}), flush=True)
    elif rank == 0:
        ep_per_s = E * HW * iters / (ms_step * 1e-3)
        line = {
            "metric": "ba_edge_pixels_per_s", "value": ep_per_s, "unit": "edge-pixels/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_string(args.workload, cfg, p),
                       "parallelism": f"edges sharded by source keyframe over {world} GPU(s)" if world > 1 else "1 GPU",
                       "l2": "256 MB buffer written between timed steps (inputs 106 MB < 126 MB L2)"},
            "ba_iters_per_s": iters / (ms_step * 1e-3),
            "clocks": clocks,
            "e2e": {"value": E * HW * iters / (ms_e2e * 1e-3), "unit": "edge-pixels/s", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "entry": "dba_ba_host (C ABI, pinned host buffers)"},
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "kernel": "k_linearize (residual/Jacobian/Hessian + Schur, fused)",
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "peak_source": peak_src, "algorithmic_bytes_per_launch": lin_bytes,
                         "ms_per_launch": lin_ms, "traffic": measured_traffic(args.workload, world),
                         "whole_iteration": {"algorithmic_bytes": b_alg, "achieved": whole, "frac": whole / peak,
                                             "frac_of_8TBs_nominal": whole / 8000.0}},
            "stage_ms_per_lm_iteration": stage_ms,
            "stage_us_per_rank": stage_all,
        }
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
