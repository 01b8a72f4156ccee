"""Global BA sharded by source keyframe over the GPUs of one box (no reference
counterpart -- the reference runs everything on cuda:0).

Every disparity map belongs to exactly one source keyframe k and only edges with ii == k touch it
(hislam2/geom/ba.py:70,83-90: E, C, w are indexed by kk = inverse(unique(ii))).  So a partition of
the edges by ii keeps C_k, w_k, E_.k, Q_k, dz_k and disps[k] rank-local; only the reduced pose system
(H - E Q E^T, v - E Q w) couples the shards, and it is a SUM over k: one ncclAllReduce per LM
iteration, issued by libdba on the compute stream between assembly and the (replicated) Cholesky.
The LM damping H_dd (1+lm) + ep is linear in H: every rank damps its share, rank 0 adds ep.
"""
import ctypes as C

import torch

from . import droid_backends


def partition_by_source(ii, nranks, cost="schur"):
    """Contiguous ranges of sorted(unique(ii)) per rank, balanced on the measured per-keyframe
    linearisation cost.  Returns [(keyframes [m_r], edge_index [E_r])] per rank; edge
    order inside a shard is the original order; eta rows of a shard are the contiguous slice
    [first_slot, first_slot + m_r) of the global eta."""
    ii = ii.cpu()
    kx, kk = torch.unique(ii, return_inverse=True)
    deg = torch.bincount(kk, minlength=kx.numel()).double()
    # measured kernel costs per 64-pixel tile (csrc/dba_api.cu prepare_plan, profiles/r2_linearize_v2.md): the
    # one-warp-per-edge kernel up to out-degree 20 (the tile time hardly depends on the degree), the WIDE tcgen05
    # variants beyond (one CTA per SM each, ~0.1 (deg + 1) of a narrow tile)
    w = torch.where(deg <= 20, 1.0 + 0.03 * deg, 0.1 * (deg + 1)) if cost == "schur" else deg
    cum = torch.cumsum(w, 0)
    total = float(cum[-1]) if cum.numel() else 0.0
    bounds = [0]
    for r in range(1, nranks):
        target = total * r / nranks
        b = int(torch.searchsorted(cum, torch.tensor(target, dtype=cum.dtype)))
        bounds.append(min(max(b, bounds[-1]), kx.numel()))
    bounds.append(kx.numel())
    out = []
    for r in range(nranks):
        lo, hi = bounds[r], bounds[r + 1]
        sel = torch.nonzero((kk >= lo) & (kk < hi)).reshape(-1)
        out.append(dict(slots=(lo, hi), keyframes=kx[lo:hi], edges=sel))
    return out


class ShardedBA:
    """One instance per rank (process).  `init_comm` wires libdba's NCCL communicator using the
    existing torch.distributed group only to broadcast the NCCL unique id."""

    def __init__(self, device):
        self.device = torch.device(device)
        # a PRIVATE libdba context: once the communicator is attached every `ba` through this context all-reduces
        # its reduced system, so plain droid_backends.ba / depth_cov calls (the module-level per-device context:
        # frontend windows, motion-only BA, a graph that differs per rank) must not share it
        idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.ctx = droid_backends._Ctx(idx)
        self.rank, self.nranks = 0, 1

    def init_comm(self, p2p=True, p2p_slab_bytes=40 << 20):
        import torch.distributed as dist
        self.rank, self.nranks = dist.get_rank(), dist.get_world_size()
        if self.nranks == 1:
            return
        idbuf = (C.c_char * 128)()
        if self.rank == 0:
            rc = self.ctx.lib.dba_nccl_unique_id(idbuf)
            if rc != 0:
                raise RuntimeError(f"dba_nccl_unique_id failed ({rc})")
        obj = [bytes(idbuf.raw) if self.rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        idbuf.raw = obj[0]
        with torch.cuda.device(self.device):
            self.ctx.check(self.ctx.lib.dba_comm_init(self.ctx.handle, self.nranks, self.rank, idbuf))
            # peer-memory all-reduce for the small / medium systems (csrc/dba_p2p.cu); NCCL stays the path for the
            # large ones and the fallback when the GPUs cannot map each other's memory
            self.p2p = False
            if p2p and self.nranks <= 8:
                h = (C.c_char * 64)()
                ok = self.ctx.lib.dba_comm_p2p_export(self.ctx.handle, h, int(p2p_slab_bytes)) == 0
                handles = [None] * self.nranks
                dist.all_gather_object(handles, bytes(h.raw) if ok else None)
                if all(x is not None for x in handles):
                    allh = (C.c_char * (64 * self.nranks))()
                    allh.raw = b"".join(handles)
                    self.p2p = self.ctx.lib.dba_comm_p2p_import(self.ctx.handle, allh, self.nranks) == 0
                flag = torch.tensor([int(self.p2p)], device=self.device)
                dist.all_reduce(flag, op=dist.ReduceOp.MIN)       # all ranks or none
                self.p2p = bool(flag.item())
                if not self.p2p:
                    self.ctx.set_option("p2p_allreduce", 0)

    def close(self):
        self.ctx.lib.dba_comm_destroy(self.ctx.handle)

    def ba(self, poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1, iterations, lm, ep):
        """The rank's shard: targets/weights/eta/ii/jj hold ITS edges (global frame indices); poses is
        a full replica (updated identically on every rank); disps rows of ITS source keyframes are
        updated, the others are stale until `gather_disps`."""
        return droid_backends.ba(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1,
                                 iterations, lm, ep, False, _ctx=self.ctx)

    def last_launch_count(self):
        return droid_backends.last_launch_count(_ctx=self.ctx)

    def stage_times(self):
        return droid_backends.stage_times(_ctx=self.ctx)

    def gather_disps(self, disps, shards):
        """Make every rank's disps complete again: ONE all-gather of the ranks' keyframe rows (padded to the longest
        shard) and one indexed copy; the index tensors are built once per shard layout."""
        import torch.distributed as dist
        if self.nranks == 1:
            return disps
        key = (id(shards), disps.device.index)
        lay = getattr(self, "_gather_layout", None)
        if lay is None or lay[0] != key:
            lens = [int(sh["keyframes"].numel()) for sh in shards]
            maxr = max(max(lens), 1)
            rows = torch.cat([sh["keyframes"] for sh in shards]).to(disps.device)
            src = torch.cat([r * maxr + torch.arange(n) for r, n in enumerate(lens)]).to(disps.device)
            mine = shards[self.rank]["keyframes"].to(disps.device)
            send = torch.zeros((maxr,) + tuple(disps.shape[1:]), dtype=disps.dtype, device=disps.device)
            recv = torch.empty((self.nranks * maxr,) + tuple(disps.shape[1:]), dtype=disps.dtype, device=disps.device)
            lay = (key, rows, src, mine, send, recv)
            self._gather_layout = lay
        _, rows, src, mine, send, recv = lay
        if mine.numel():
            send[:mine.numel()] = disps[mine]
        dist.all_gather_into_tensor(recv, send)
        disps[rows] = recv[src]
        return disps
