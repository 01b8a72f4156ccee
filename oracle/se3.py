"""TEST INFRASTRUCTURE ONLY -- SE3 maths of the reference's lietorch, restated on torch CPU tensors.

Pose storage [tx,ty,tz, qx,qy,qz,qw] (thirdparty/lietorch/lietorch/include/se3.h:34,
groups.py:271-272); tangent order [tau(3), phi(3)] (se3.h:90-91,135-136).
All functions broadcast over leading dims and keep the input dtype (fp32 or fp64).
"""
import torch

EPS = 1e-6  # thirdparty/lietorch/lietorch/include/common.h:7


def quat_normalize(q):
    """so3.h:31-37 -- every SO3 constructor normalises its quaternion."""
    return q / torch.sqrt((q * q).sum(dim=-1, keepdim=True))


def quat_mul(a, b):
    """Hamilton product, xyzw storage (Eigen::Quaternion operator*, used by so3.h:51-53)."""
    ax, ay, az, aw = a.unbind(-1)
    bx, by, bz, bw = b.unbind(-1)
    return torch.stack([
        aw * bx + ax * bw + ay * bz - az * by,
        aw * by + ay * bw + az * bx - ax * bz,
        aw * bz + az * bw + ax * by - ay * bx,
        aw * bw - ax * bx - ay * by - az * bz,
    ], dim=-1)


def quat_conj(q):
    return torch.cat([-q[..., :3], q[..., 3:]], dim=-1)


def quat_rotate(q, p):
    """so3.h:55-60: uv = 2 (q.vec x p); p' = p + q.w uv + q.vec x uv."""
    qv, qw = q[..., :3], q[..., 3:]
    uv = torch.cross(qv.expand(torch.broadcast_shapes(qv.shape, p.shape)),
                     p.expand(torch.broadcast_shapes(qv.shape, p.shape)), dim=-1)
    uv = uv + uv
    return p + qw * uv + torch.cross(qv.expand(uv.shape), uv, dim=-1)


def quat_to_matrix(q):
    """Eigen::Quaternion::toRotationMatrix (used by so3.h:68-74)."""
    x, y, z, w = q.unbind(-1)
    tx, ty, tz = 2 * x, 2 * y, 2 * z
    twx, twy, twz = tx * w, ty * w, tz * w
    txx, txy, txz = tx * x, ty * x, tz * x
    tyy, tyz, tzz = ty * y, tz * y, tz * z
    R = torch.stack([
        1 - (tyy + tzz), txy - twz, txz + twy,
        txy + twz, 1 - (txx + tzz), tyz - twx,
        txz - twy, tyz + twx, 1 - (txx + tyy),
    ], dim=-1)
    return R.reshape(q.shape[:-1] + (3, 3))


def hat(phi):
    """so3.h:101-109."""
    x, y, z = phi.unbind(-1)
    o = torch.zeros_like(x)
    return torch.stack([o, -z, y, z, o, -x, -y, x, o], dim=-1).reshape(phi.shape[:-1] + (3, 3))


def se3_split(G):
    return G[..., :3], quat_normalize(G[..., 3:7])


def se3_inv(G):
    """se3.h:36-38."""
    t, q = se3_split(G)
    qi = quat_normalize(quat_conj(q))
    return torch.cat([-quat_rotate(qi, t), qi], dim=-1)


def se3_mul(A, B):
    """se3.h:45-47: (R1,t1)(R2,t2) = (R1 R2, t1 + R1 t2)."""
    ta, qa = se3_split(A)
    tb, qb = se3_split(B)
    return torch.cat([ta + quat_rotate(qa, tb), quat_normalize(quat_mul(qa, qb))], dim=-1)


def se3_act4(G, X):
    """se3.h:53-56: [R X.xyz + t X.w, X.w]."""
    t, q = se3_split(G)
    return torch.cat([quat_rotate(q, X[..., :3]) + t * X[..., 3:], X[..., 3:]], dim=-1)


def se3_adj_matrix(G):
    """se3.h:58-67: Ad = [[R, [t]x R], [0, R]]."""
    t, q = se3_split(G)
    R = quat_to_matrix(q)
    tR = hat(t) @ R
    Z = torch.zeros_like(R)
    top = torch.cat([R, tR], dim=-1)
    bot = torch.cat([Z, R], dim=-1)
    return torch.cat([top, bot], dim=-2)


def se3_adjT(G, a):
    """se3.h:84-86: Ad^T a  (groups.py:162-164 `adjT`)."""
    Ad = se3_adj_matrix(G)
    return (Ad.transpose(-1, -2) @ a.unsqueeze(-1)).squeeze(-1)


def so3_exp(phi):
    """so3.h:153-170."""
    theta2 = (phi * phi).sum(-1, keepdim=True)
    theta = torch.sqrt(theta2)
    theta4 = theta2 * theta2
    small = theta < EPS
    safe = torch.where(small, torch.ones_like(theta), theta)
    imag = torch.where(small, 0.5 - theta2 / 48.0 + theta4 / 3840.0, torch.sin(0.5 * safe) / safe)
    real = torch.where(small, 1.0 - theta2 / 8.0 + theta4 / 384.0, torch.cos(0.5 * safe))
    return quat_normalize(torch.cat([imag * phi, real], dim=-1))


def so3_left_jacobian(phi):
    """so3.h:172-190."""
    theta2 = (phi * phi).sum(-1, keepdim=True)
    theta = torch.sqrt(theta2)
    small = theta < EPS
    safe2 = torch.where(small, torch.ones_like(theta2), theta2)
    safe = torch.where(small, torch.ones_like(theta), theta)
    c1 = torch.where(small, 0.5 - theta2 / 24.0, (1.0 - torch.cos(safe)) / safe2)
    c2 = torch.where(small, 1.0 / 6.0 - theta2 / 120.0, (safe - torch.sin(safe)) / (safe2 * safe))
    Phi = hat(phi)
    I = torch.eye(3, dtype=phi.dtype).expand(Phi.shape)
    return I + c1.unsqueeze(-1) * Phi + c2.unsqueeze(-1) * (Phi @ Phi)


def se3_exp(xi):
    """se3.h:134-142: Exp(tau,phi) = (Exp_SO3(phi), J_l(phi) tau)."""
    tau, phi = xi[..., :3], xi[..., 3:]
    q = so3_exp(phi)
    t = (so3_left_jacobian(phi) @ tau.unsqueeze(-1)).squeeze(-1)
    return torch.cat([t, q], dim=-1)


def se3_retr(G, xi):
    """groups.py:153-156: Exp(xi) * G."""
    return se3_mul(se3_exp(xi), G)


def so3_log(q):
    """so3.h:119-151 (atan-based)."""
    q = quat_normalize(q)
    v, w = q[..., :3], q[..., 3:]
    n2 = (v * v).sum(-1, keepdim=True)
    n = torch.sqrt(n2)
    small = n2 < EPS * EPS
    safe_n = torch.where(small, torch.ones_like(n), n)
    wz = w.abs() < EPS
    safe_w = torch.where(wz, torch.ones_like(w), w)
    f_small = 2.0 / safe_w - (2.0 / 3.0) * n2 / (safe_w * safe_w * safe_w)
    f_wz = torch.where(w > 0, torch.pi / safe_n, -torch.pi / safe_n)
    f_reg = 2.0 * torch.atan(safe_n / safe_w) / safe_n
    f = torch.where(small, f_small, torch.where(wz, f_wz, f_reg))
    return f * v


def so3_left_jacobian_inverse(phi):
    """so3.h:192-209."""
    theta2 = (phi * phi).sum(-1, keepdim=True)
    theta = torch.sqrt(theta2)
    small = theta < EPS
    safe = torch.where(small, torch.ones_like(theta), theta)
    half = 0.5 * safe
    c2 = torch.where(small, torch.full_like(theta, 1.0 / 12.0),
                     (1.0 - safe * torch.cos(half) / (2.0 * torch.sin(half))) / (safe * safe))
    Phi = hat(phi)
    I = torch.eye(3, dtype=phi.dtype).expand(Phi.shape)
    return I - 0.5 * Phi + c2.unsqueeze(-1) * (Phi @ Phi)


def se3_log(G):
    """se3.h:124-132."""
    t, q = se3_split(G)
    phi = so3_log(q)
    tau = (so3_left_jacobian_inverse(phi) @ t.unsqueeze(-1)).squeeze(-1)
    return torch.cat([tau, phi], dim=-1)


def tangent_update(G_new, G_old):
    """dx such that G_new = Exp(dx) * G_old (the quantity north_star's pose tolerance is stated on)."""
    return se3_log(se3_mul(G_new, se3_inv(G_old)))
