"""Seeded synthetic factor graphs of the BASELINE.json shapes.

Host-side input generation only (torch, CPU or CUDA tensors) -- not the hot path, never timed.
Conventions follow the reference: poses [tx,ty,tz,qx,qy,qz,qw] world->camera
(thirdparty/lietorch/lietorch/include/se3.h:34), disparities at 1/8 resolution, intrinsics
[fx,fy,cx,cy], targets/weights in the op layout [E,2,ht,wd] (pyc:DepthVideo.cuda_ba@229-230),
eta rows in sorted(unique(ii)) order (pyc:FactorGraph.update@227), ii/jj int64.
"""
from dataclasses import dataclass

import torch

CONFIGS = {
    # name: keyframes, edges, ht, wd, iterations, lm, ep, t0, radius, motion_only
    "correctness":  dict(P=8,   E=24,    ht=48, wd=64, iters=2, lm=1e-4, ep=0.1,  t0=2, r=2, motion_only=False, seed=1235),
    "frontend":     dict(P=12,  E=100,   ht=43, wd=77, iters=2, lm=1e-4, ep=0.1,  t0=2, r=3, motion_only=False, seed=1236),
    "backend":      dict(P=128, E=2000,  ht=43, wd=77, iters=8, lm=1e-5, ep=1e-2, t0=2, r=3, motion_only=False, seed=1237),
    "loopclosure":  dict(P=512, E=12000, ht=48, wd=64, iters=8, lm=1e-5, ep=1e-2, t0=2, r=3, motion_only=False, seed=1238),
    "motiononly":   dict(P=64,  E=372,   ht=48, wd=64, iters=2, lm=1e-4, ep=0.1,  t0=2, r=3, motion_only=True,  seed=1239),
}


@dataclass
class Problem:
    poses: torch.Tensor        # [P,7]   start point (perturbed)
    disps: torch.Tensor        # [P,ht,wd]
    intrinsics: torch.Tensor   # [4]
    disps_sens: torch.Tensor   # [P,ht,wd] zeros (no sensor depth)
    targets: torch.Tensor      # [E,2,ht,wd]
    weights: torch.Tensor      # [E,2,ht,wd]
    eta: torch.Tensor          # [M,ht,wd]
    ii: torch.Tensor           # [E] int64
    jj: torch.Tensor           # [E] int64
    t0: int
    t1: int
    iterations: int
    lm: float
    ep: float
    motion_only: bool
    gt_poses: torch.Tensor
    gt_disps: torch.Tensor

    def to(self, device):
        kw = {}
        for k, v in self.__dict__.items():
            kw[k] = v.to(device) if isinstance(v, torch.Tensor) else v
        return Problem(**kw)

    def ba_args(self):
        return (self.poses, self.disps, self.intrinsics, self.disps_sens, self.targets, self.weights,
                self.eta, self.ii, self.jj, self.t0, self.t1, self.iterations, self.lm, self.ep,
                self.motion_only)

    @property
    def edge_pixels(self):
        return self.ii.numel() * self.disps.shape[1] * self.disps.shape[2]


# ---- minimal SE3 helpers for generation (double precision, then cast) -------------------------
def _hat(p):
    x, y, z = p.unbind(-1)
    o = torch.zeros_like(x)
    return torch.stack([o, -z, y, z, o, -x, -y, x, o], -1).reshape(p.shape[:-1] + (3, 3))


def _exp(xi):
    """SE3 exponential -> (R [..,3,3], t [..,3]); Rodrigues + left Jacobian."""
    tau, phi = xi[..., :3], xi[..., 3:]
    th = phi.norm(dim=-1, keepdim=True).clamp(min=1e-12)[..., None]
    K = _hat(phi)
    I = torch.eye(3, dtype=xi.dtype).expand(K.shape)
    R = I + torch.sin(th) / th * K + (1 - torch.cos(th)) / th ** 2 * (K @ K)
    V = I + (1 - torch.cos(th)) / th ** 2 * K + (th - torch.sin(th)) / th ** 3 * (K @ K)
    return R, (V @ tau[..., None])[..., 0]


def _mat_to_quat(R):
    """Rotation matrix -> unit quaternion xyzw (w >= 0 branch is enough for the small rotations used)."""
    w = torch.sqrt((1 + R[..., 0, 0] + R[..., 1, 1] + R[..., 2, 2]).clamp(min=1e-12)) / 2
    x = (R[..., 2, 1] - R[..., 1, 2]) / (4 * w)
    y = (R[..., 0, 2] - R[..., 2, 0]) / (4 * w)
    z = (R[..., 1, 0] - R[..., 0, 1]) / (4 * w)
    q = torch.stack([x, y, z, w], -1)
    return q / q.norm(dim=-1, keepdim=True)


def _reproject(R, t, disps, intr, ii, jj):
    """coords of the pixels of frame ii seen in frame jj (pinhole; projective_ops.py:44-62 semantics)."""
    P, ht, wd = disps.shape
    fx, fy, cx, cy = intr.tolist()
    y, x = torch.meshgrid(torch.arange(ht, dtype=disps.dtype), torch.arange(wd, dtype=disps.dtype), indexing="ij")
    X0 = torch.stack([(x - cx) / fx, (y - cy) / fy, torch.ones_like(x)], -1)           # [ht,wd,3]
    Rij = R[jj] @ R[ii].transpose(-1, -2)                                               # [E,3,3]
    tij = t[jj] - (Rij @ t[ii][..., None])[..., 0]
    X1 = torch.einsum("eab,hwb->ehwa", Rij, X0) + tij[:, None, None, :] * disps[ii][..., None]
    Z = X1[..., 2]
    Z = torch.where(Z < 0.1, torch.ones_like(Z), Z)
    return torch.stack([fx * X1[..., 0] / Z + cx, fy * X1[..., 1] / Z + cy], 1)         # [E,2,ht,wd]


def make_edges(P, E, r, gen):
    """Radius-r temporal neighbours in both directions (hislam2/geom/graph_utils.py:25-29 order),
    then random loop edge pairs (i,j),(j,i) with |i-j| > r until E edges; no duplicates
    (hislam2/factor_graph.py:29-39 semantics). Generation order is kept (NOT sorted by ii)."""
    a = torch.arange(P)
    gi, gj = torch.meshgrid(a, a, indexing="ij")
    gi, gj = gi.reshape(-1), gj.reshape(-1)
    d = (gi - gj).abs()
    keep = (d >= 1) & (d <= r)
    edges = list(zip(gi[keep].tolist(), gj[keep].tolist()))[:E]
    seen = set(edges)
    while len(edges) < E:
        i, j = torch.randint(0, P, (2,), generator=gen).tolist()
        if abs(i - j) <= r or (i, j) in seen:
            continue
        for e in ((i, j), (j, i)):
            if len(edges) < E and e not in seen:
                edges.append(e)
                seen.add(e)
    ii = torch.tensor([e[0] for e in edges], dtype=torch.long)
    jj = torch.tensor([e[1] for e in edges], dtype=torch.long)
    return ii, jj


def make_problem(name=None, *, P=None, E=None, ht=None, wd=None, iters=2, lm=1e-4, ep=0.1, t0=2, r=3,
                 motion_only=False, seed=1234, pose_noise=0.02, disp_noise=0.1, flow_noise=0.1,
                 device="cpu", edges=None):
    """Build a seeded problem. `name` picks one of CONFIGS; keyword overrides apply on top.
    `edges=(ii, jj)` replaces the generated edge list (ragged / degenerate graphs of the parity tests)."""
    if name is not None:
        c = dict(CONFIGS[name])
        P, E, ht, wd = P or c["P"], E or c["E"], ht or c["ht"], wd or c["wd"]
        iters, lm, ep, t0, r, motion_only, seed = c["iters"], c["lm"], c["ep"], c["t0"], c["r"], c["motion_only"], c["seed"]
    gen = torch.Generator().manual_seed(seed)
    f64 = torch.float64
    # trajectory: smooth back-and-forth arc so that loop edges between distant keyframes still overlap
    k = torch.arange(P, dtype=f64)
    s = 8.0 * torch.sin(2 * torch.pi * k / 48.0)
    base = torch.tensor([0.05, 0, 0, 0, 0.02, 0], dtype=f64)
    xi = s[:, None] * base[None] + 0.01 * torch.randn(P, 6, generator=gen, dtype=f64)
    xi[0] = 0
    Rg, tg = _exp(xi)
    # disparities: 1/depth, depth ~ U(1,4) smoothed by a 5x5 box
    depth = 1 + 3 * torch.rand(P, 1, ht, wd, generator=gen, dtype=f64)
    depth = torch.nn.functional.avg_pool2d(torch.nn.functional.pad(depth, (2, 2, 2, 2), mode="replicate"), 5, 1)[:, 0]
    gt_disps = 1.0 / depth
    intr = torch.tensor([0.8 * wd, 0.8 * wd, wd / 2.0, ht / 2.0], dtype=f64)
    if edges is None:
        ii, jj = make_edges(P, E, r, gen)
    else:
        ii, jj = (torch.as_tensor(e, dtype=torch.long).reshape(-1) for e in edges)
        E = int(ii.numel())
    targets = _reproject(Rg, tg, gt_disps, intr, ii, jj) + flow_noise * torch.randn(E, 2, ht, wd, generator=gen, dtype=f64)
    weights = torch.rand(E, 2, ht, wd, generator=gen, dtype=f64)
    M = int(torch.unique(ii).numel())
    eta = torch.full((M, ht, wd), 0.2 * 1e-6 + 1e-7, dtype=f64)
    # start point
    Rn, tn = _exp(pose_noise * torch.randn(P, 6, generator=gen, dtype=f64))
    R0 = Rn @ Rg
    t0v = (Rn @ tg[..., None])[..., 0] + tn
    R0[:t0], t0v[:t0] = Rg[:t0], tg[:t0]      # gauge frames start (and stay) at ground truth
    disps0 = gt_disps * (1 + disp_noise * (2 * torch.rand(P, ht, wd, generator=gen, dtype=f64) - 1))
    f32 = torch.float32
    prob = Problem(
        poses=torch.cat([t0v, _mat_to_quat(R0)], -1).to(f32).contiguous(),
        disps=disps0.to(f32).contiguous(),
        intrinsics=intr.to(f32),
        disps_sens=torch.zeros(P, ht, wd, dtype=f32),
        targets=targets.to(f32).contiguous(), weights=weights.to(f32).contiguous(),
        eta=eta.to(f32), ii=ii, jj=jj, t0=t0, t1=P, iterations=iters, lm=lm, ep=ep,
        motion_only=motion_only,
        gt_poses=torch.cat([tg, _mat_to_quat(Rg)], -1).to(f32), gt_disps=gt_disps.to(f32),
    )
    return prob.to(device) if device != "cpu" else prob
