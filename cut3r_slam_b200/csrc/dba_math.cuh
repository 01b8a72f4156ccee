// SE3 / pinhole maths shared by all DBA kernels (fp32, device only).
//
// Conventions restated from the reference (nothing here is a translation of its Eigen code):
//   pose storage [tx,ty,tz, qx,qy,qz,qw], world->camera   thirdparty/lietorch/lietorch/include/se3.h:34
//   quaternion normalised on load                          so3.h:31-37
//   rotate: uv = 2 (q.vec x p); p' = p + q.w uv + q.vec x uv   so3.h:55-60
//   compose / invert                                       se3.h:36-47
//   Ad = [[R, [t]x R],[0, R]]                              se3.h:58-67
//   Exp(tau,phi) = (Exp_SO3(phi), J_l(phi) tau)            se3.h:134-142, so3.h:153-190
//   relative pose Gij = Gj * Gi^-1                         hislam2/geom/projective_ops.py:51
#pragma once
#include <cuda_runtime.h>

namespace dba {

constexpr float kMinDepth = 0.2f;      // projective_ops.py:6
constexpr float kWeightScale = 0.001f; // ba.py:44
constexpr float kCEps = 1e-7f;         // ba.py:92
constexpr float kSensAlpha = 0.05f;    // upstream DROID-SLAM sensor-depth prior (parity unpinned)
constexpr float kLieEps = 1e-6f;       // lietorch common.h:7

struct V3 { float x, y, z; };

__device__ __forceinline__ V3 cross(const V3& a, const V3& b) {
  return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
__device__ __forceinline__ float dot(const V3& a, const V3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

struct Quat { float x, y, z, w; };

__device__ __forceinline__ Quat qnormalize(Quat q) {
  float n = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
  float s = 1.0f / n;
  return Quat{q.x * s, q.y * s, q.z * s, q.w * s};
}
__device__ __forceinline__ Quat qmul(const Quat& a, const Quat& b) {
  return Quat{a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
              a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z,
              a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x,
              a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z};
}
__device__ __forceinline__ V3 qrot(const Quat& q, const V3& p) {
  V3 qv{q.x, q.y, q.z};
  V3 uv = cross(qv, p);
  uv.x += uv.x; uv.y += uv.y; uv.z += uv.z;
  V3 c = cross(qv, uv);
  return V3{p.x + q.w * uv.x + c.x, p.y + q.w * uv.y + c.y, p.z + q.w * uv.z + c.z};
}

struct SE3 { V3 t; Quat q; };

__device__ __forceinline__ SE3 se3_load(const float* __restrict__ p) {
  SE3 g;
  g.t = V3{p[0], p[1], p[2]};
  g.q = qnormalize(Quat{p[3], p[4], p[5], p[6]});
  return g;
}
__device__ __forceinline__ SE3 se3_inv(const SE3& g) {
  SE3 r;
  r.q = qnormalize(Quat{-g.q.x, -g.q.y, -g.q.z, g.q.w});
  V3 rt = qrot(r.q, g.t);
  r.t = V3{-rt.x, -rt.y, -rt.z};
  return r;
}
__device__ __forceinline__ SE3 se3_mul(const SE3& a, const SE3& b) {
  SE3 r;
  V3 rt = qrot(a.q, b.t);
  r.t = V3{a.t.x + rt.x, a.t.y + rt.y, a.t.z + rt.z};
  r.q = qnormalize(qmul(a.q, b.q));
  return r;
}
// rotation matrix, row-major R[9] (Eigen::Quaternion::toRotationMatrix form)
__device__ __forceinline__ void quat_to_R(const Quat& q, float* R) {
  float tx = 2.f * q.x, ty = 2.f * q.y, tz = 2.f * q.z;
  float twx = tx * q.w, twy = ty * q.w, twz = tz * q.w;
  float txx = tx * q.x, txy = ty * q.x, txz = tz * q.x;
  float tyy = ty * q.y, tyz = tz * q.y, tzz = tz * q.z;
  R[0] = 1.f - (tyy + tzz); R[1] = txy - twz;         R[2] = txz + twy;
  R[3] = txy + twz;         R[4] = 1.f - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy;         R[7] = tyz + twx;         R[8] = 1.f - (txx + tyy);
}

// Relative pose of an edge as 12 floats: R (row-major 9) then t (3).
__device__ __forceinline__ void relative_pose_Rt(const float* __restrict__ poses, int i, int j, float* Rt) {
  SE3 gi = se3_load(poses + 7 * i);
  SE3 gj = se3_load(poses + 7 * j);
  SE3 gij = se3_mul(gj, se3_inv(gi));
  quat_to_R(gij.q, Rt);
  Rt[9] = gij.t.x; Rt[10] = gij.t.y; Rt[11] = gij.t.z;
}

// 6x6 adjoint, row-major, from (R,t): [[R, [t]x R],[0, R]]
__device__ __forceinline__ void adjoint_from_Rt(const float* Rt, float* Ad) {
  const float* R = Rt;
  float tx = Rt[9], ty = Rt[10], tz = Rt[11];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float r0 = R[0 + c], r1 = R[3 + c], r2 = R[6 + c];
    Ad[0 * 6 + c] = r0; Ad[1 * 6 + c] = r1; Ad[2 * 6 + c] = r2;
    Ad[3 * 6 + c] = 0.f; Ad[4 * 6 + c] = 0.f; Ad[5 * 6 + c] = 0.f;
    Ad[3 * 6 + 3 + c] = r0; Ad[4 * 6 + 3 + c] = r1; Ad[5 * 6 + 3 + c] = r2;
    // [t]x R column c = t x R[:,c]
    Ad[0 * 6 + 3 + c] = ty * r2 - tz * r1;
    Ad[1 * 6 + 3 + c] = tz * r0 - tx * r2;
    Ad[2 * 6 + 3 + c] = tx * r1 - ty * r0;
  }
}

// pose <- Exp(xi) * pose  (groups.py:153-156), xi = [tau, phi]
__device__ __forceinline__ void se3_retract(float* __restrict__ pose, const float* xi) {
  V3 tau{xi[0], xi[1], xi[2]};
  V3 phi{xi[3], xi[4], xi[5]};
  float th2 = dot(phi, phi);
  float th = sqrtf(th2);
  float th4 = th2 * th2;
  float imag, real, c1, c2;
  if (th < kLieEps) {
    imag = 0.5f - th2 / 48.0f + th4 / 3840.0f;
    real = 1.0f - th2 / 8.0f + th4 / 384.0f;
    c1 = 0.5f - th2 / 24.0f;
    c2 = 1.0f / 6.0f - th2 / 120.0f;
  } else {
    float s, c;
    sincosf(0.5f * th, &s, &c);
    imag = s / th;
    real = c;
    float st, ct;
    sincosf(th, &st, &ct);
    c1 = (1.0f - ct) / th2;
    c2 = (th - st) / (th2 * th);
  }
  SE3 e;
  e.q = qnormalize(Quat{imag * phi.x, imag * phi.y, imag * phi.z, real});
  // J_l tau = tau + c1 (phi x tau) + c2 (phi x (phi x tau))
  V3 pt = cross(phi, tau);
  V3 ppt = cross(phi, pt);
  e.t = V3{tau.x + c1 * pt.x + c2 * ppt.x, tau.y + c1 * pt.y + c2 * ppt.y, tau.z + c1 * pt.z + c2 * ppt.z};
  SE3 g = se3_load(pose);
  SE3 r = se3_mul(e, g);
  pose[0] = r.t.x; pose[1] = r.t.y; pose[2] = r.t.z;
  pose[3] = r.q.x; pose[4] = r.q.y; pose[5] = r.q.z; pose[6] = r.q.w;
}

// ---------------------------------------------------------------------------------------------
// Per edge-pixel projection (pinhole.py:6-53, projective_ops.py:15-74), in the reduced form
// proven in tests/test_kernel_algebra.py:
//   X1 = R [x0,y0,1] + t d ;  Zc = (Z < 0.1 ? 1 : Z) ; iz = 1/Zc
//   g0 = [fx iz, 0, -fx X iz^2], g1 = [0, fy iz, -fy Y iz^2]   (rows of the projection Jacobian)
//   Jj_r = [d g_r, X1 x g_r] ; Jz_r = g_r . t ; valid = Z > 0.2
// ---------------------------------------------------------------------------------------------
struct Proj {
  V3 X1;
  float u, v;        // reprojected coordinates
  float a0, c0;      // g0 = [a0, 0, c0]
  float b1, c1;      // g1 = [0, b1, c1]
  float jz0, jz1;
  float valid;
};

__device__ __forceinline__ Proj project(const float* Rt, float x0, float y0, float d,
                                        float fx, float fy, float cx, float cy) {
  Proj P;
  P.X1.x = Rt[0] * x0 + Rt[1] * y0 + Rt[2] + Rt[9] * d;
  P.X1.y = Rt[3] * x0 + Rt[4] * y0 + Rt[5] + Rt[10] * d;
  P.X1.z = Rt[6] * x0 + Rt[7] * y0 + Rt[8] + Rt[11] * d;
  float Z = P.X1.z;
  P.valid = (Z > kMinDepth) ? 1.0f : 0.0f;
  float Zc = (Z < 0.1f) ? 1.0f : Z;
  float iz;                                  // 1/Zc: MUFU.RCP + one Newton step (~0.5 ulp) instead of the IEEE division
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(iz) : "f"(Zc));
  iz = iz * fmaf(-Zc, iz, 2.0f);
  P.u = fx * (P.X1.x * iz) + cx;
  P.v = fy * (P.X1.y * iz) + cy;
  P.a0 = fx * iz;
  P.b1 = fy * iz;
  P.c0 = -P.a0 * P.X1.x * iz;
  P.c1 = -P.b1 * P.X1.y * iz;
  P.jz0 = P.a0 * Rt[9] + P.c0 * Rt[11];
  P.jz1 = P.b1 * Rt[10] + P.c1 * Rt[11];
  return P;
}

}  // namespace dba
