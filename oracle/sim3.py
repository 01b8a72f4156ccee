"""TEST INFRASTRUCTURE ONLY -- Sim3 maths of the reference's lietorch, restated on torch CPU tensors.

Storage [tx,ty,tz, qx,qy,qz,qw, s] (thirdparty/lietorch/lietorch/include/sim3.h:39-40: translation(data),
rxso3(data+3); rxso3.h: quaternion then scale); tangent order [tau(3), phi(3), sigma] (sim3.h:122-127).
All functions broadcast over leading dims and keep the input dtype.  Pinned to the reference's own lietorch CPU
build in tests/test_pgba_cpu.py (oracle/_ref, build container only) through tests/golden/pgba.npz.
"""
import torch

from . import se3

EPS = 1e-6  # common.h:7


def split(G):
    return G[..., :3], se3.quat_normalize(G[..., 3:7]), G[..., 7:8]


def mul(A, B):
    """sim3.h:51-53, rxso3.h:26-28: (s1 R1, t1)(s2 R2, t2) = (s1 s2 R1 R2, t1 + s1 R1 t2)."""
    ta, qa, sa = split(A)
    tb, qb, sb = split(B)
    return torch.cat([ta + sa * se3.quat_rotate(qa, tb), se3.quat_normalize(se3.quat_mul(qa, qb)), sa * sb], dim=-1)


def inv(G):
    """sim3.h:42-44: (rxso3^-1, -(rxso3^-1 t))."""
    t, q, s = split(G)
    qi = se3.quat_conj(q)
    return torch.cat([-(se3.quat_rotate(qi, t) / s), qi, 1.0 / s], dim=-1)


def act4(G, X):
    """sim3.h:59-62: [s R X.xyz + t X.w, X.w]."""
    t, q, s = split(G)
    return torch.cat([s * se3.quat_rotate(q, X[..., :3]) + t * X[..., 3:], X[..., 3:]], dim=-1)


def adj_matrix(G):
    """sim3.h:91-104: Ad = [[sR, [t]x R, -t], [0, R, 0], [0, 0, 1]]  (7 x 7)."""
    t, q, s = split(G)
    R = se3.quat_to_matrix(q)
    Ad = torch.zeros(G.shape[:-1] + (7, 7), dtype=G.dtype)
    Ad[..., :3, :3] = s[..., None] * R
    Ad[..., :3, 3:6] = se3.hat(t) @ R
    Ad[..., :3, 6] = -t
    Ad[..., 3:6, 3:6] = R
    Ad[..., 6, 6] = 1.0
    return Ad


def calc_w(phi, sigma):
    """rxso3.h:190-233 (calcW): W = A Phi + B Phi^2 + C I."""
    dt = phi.dtype
    theta = phi.norm(dim=-1)
    scale = torch.exp(sigma)
    one = torch.ones_like(theta)
    small_s = sigma.abs() < EPS
    small_t = theta.abs() < EPS
    ts = torch.where(small_t, one, theta)
    ss = torch.where(small_s, one, sigma)
    tsq = ts * ts
    # sigma ~ 0
    A0 = torch.where(small_t, 0.5 * one, (1 - torch.cos(ts)) / tsq)
    B0 = torch.where(small_t, one / 6.0, (ts - torch.sin(ts)) / (tsq * ts))
    # general sigma
    C1 = (scale - 1) / ss
    ssq = ss * ss
    A1t0 = ((ss - 1) * scale + 1) / ssq
    B1t0 = (scale * 0.5 * ssq + scale - 1 - ss * scale) / (ssq * ss)
    a = scale * torch.sin(ts)
    b = scale * torch.cos(ts)
    c = tsq + ssq
    A1 = (a * ss + (1 - b) * ts) / (ts * c)
    B1 = (C1 - ((b - 1) * ss + a * ts) / c) / tsq
    A = torch.where(small_s, A0, torch.where(small_t, A1t0, A1))
    B = torch.where(small_s, B0, torch.where(small_t, B1t0, B1))
    C = torch.where(small_s, one, C1)
    Phi = se3.hat(phi)
    I = torch.eye(3, dtype=dt).expand(Phi.shape)
    return A[..., None, None] * Phi + B[..., None, None] * (Phi @ Phi) + C[..., None, None] * I


def exp(xi):
    """sim3.h:156-166: Exp(tau, phi, sigma) = (RxSO3::Exp(phi, sigma), W tau)."""
    tau, phi, sigma = xi[..., :3], xi[..., 3:6], xi[..., 6]
    q = se3.so3_exp(phi)
    W = calc_w(phi, sigma)
    t = (W @ tau[..., None])[..., 0]
    return torch.cat([t, q, torch.exp(sigma)[..., None]], dim=-1)


def log(G):
    """sim3.h:145-154: [W^-1 t, rxso3.Log()] with rxso3.Log() = [SO3 log, log(scale)] (rxso3.h:131-166)."""
    t, q, s = split(G)
    phi = se3.so3_log(q)
    sigma = torch.log(s[..., 0])
    W = calc_w(phi, sigma)
    tau = torch.linalg.solve(W, t[..., None])[..., 0]
    return torch.cat([tau, phi, sigma[..., None]], dim=-1)


def retr(G, xi):
    """groups.py:153-156: left-multiplicative retraction Exp(xi) * G."""
    return mul(exp(xi), G)
