// Per-sample geometry shared by the forward and backward kernels: the two Rodrigues variants of the
// reference, the perspective projection and their adjoints.  Accurate sinf/cosf/division (the library
// is built WITHOUT --use_fast_math) so results stay within 1e-5 of the fp64 oracle.
#pragma once
#include "common.cuh"

namespace smplb {

// batch_rodrigues, deps/tf_smpl/batch_lbs.py:38-57.  eps is added to every component inside the norm
// only (:45); the axis divides the un-shifted theta (:46).  R row-major.
__device__ __forceinline__ void rodrigues_smpl(float tx, float ty, float tz, float R[9]) {
  const float ex = tx + 1e-8f, ey = ty + 1e-8f, ez = tz + 1e-8f;
  const float angle = sqrtf(ex * ex + ey * ey + ez * ez);
  const float rx = tx / angle, ry = ty / angle, rz = tz / angle;
  float s, c;
  sincosf(angle, &s, &c);
  const float k = 1.0f - c;
  R[0] = c + k * rx * rx;
  R[1] = k * rx * ry - s * rz;
  R[2] = k * rx * rz + s * ry;
  R[3] = k * ry * rx + s * rz;
  R[4] = c + k * ry * ry;
  R[5] = k * ry * rz - s * rx;
  R[6] = k * rz * rx - s * ry;
  R[7] = k * rz * ry + s * rx;
  R[8] = c + k * rz * rz;
}

// rodrigues_batch, deps/tf_perspective_projection/project.py:51-81: R = I + sin K + (1-cos) K^2,
// identity when the norm is exactly zero (:80-81).
__device__ __forceinline__ void rodrigues_cam(float ax, float ay, float az, float R[9]) {
  const float th = sqrtf(ax * ax + ay * ay + az * az);
  if (th == 0.0f) {
    R[0] = 1.f; R[1] = 0.f; R[2] = 0.f;
    R[3] = 0.f; R[4] = 1.f; R[5] = 0.f;
    R[6] = 0.f; R[7] = 0.f; R[8] = 1.f;
    return;
  }
  const float ux = ax / th, uy = ay / th, uz = az / th;
  float s, c;
  sincosf(th, &s, &c);
  const float k = 1.0f - c;
  // K = [[0,-uz,uy],[uz,0,-ux],[-uy,ux,0]] ; K^2 = u u^T - (u.u) I (computed as the matrix product)
  const float k00 = -uz * uz - uy * uy, k01 = uy * ux, k02 = uz * ux;
  const float k10 = ux * uy, k11 = -uz * uz - ux * ux, k12 = uz * uy;
  const float k20 = ux * uz, k21 = uy * uz, k22 = -uy * uy - ux * ux;
  R[0] = 1.0f + k * k00;
  R[1] = -s * uz + k * k01;
  R[2] = s * uy + k * k02;
  R[3] = s * uz + k * k10;
  R[4] = 1.0f + k * k11;
  R[5] = -s * ux + k * k12;
  R[6] = -s * uy + k * k20;
  R[7] = s * ux + k * k21;
  R[8] = 1.0f + k * k22;
}

// project(), deps/tf_perspective_projection/project.py:19-48: q = R^T (p - c); uv = f q_xy / q_z.
__device__ __forceinline__ void project_point(const float Rc[9], float px, float py, float pz, float cx,
                                              float cy, float cz, float f, float& u, float& v,
                                              float q[3]) {
  const float dx = px - cx, dy = py - cy, dz = pz - cz;
  q[0] = Rc[0] * dx + Rc[3] * dy + Rc[6] * dz;
  q[1] = Rc[1] * dx + Rc[4] * dy + Rc[7] * dz;
  q[2] = Rc[2] * dx + Rc[5] * dy + Rc[8] * dz;
  u = (f * q[0]) / q[2];
  v = (f * q[1]) / q[2];
}

// pose_model_3d.py:229-231: uv / ss * img_dim[1] + img_dim / 2 with img_dim = [W, H].
__device__ __forceinline__ void to_pixels(float u, float v, const smplb_camcfg& cfg, float& pxx,
                                          float& pxy) {
  pxx = u / cfg.ss * cfg.img_h + cfg.img_w * 0.5f;
  pxy = v / cfg.ss * cfg.img_h + cfg.img_h * 0.5f;
}

__device__ __forceinline__ float tf32_rna(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

}  // namespace smplb

namespace smplb {

// Adjoint of rodrigues_smpl (batch_lbs.py:38-57): gR = dL/dR (row-major) -> dL/dtheta.  SURVEY.md appendix A.8.
__device__ __forceinline__ void rodrigues_smpl_bwd(float tx, float ty, float tz, const float gR[9], float gth[3]) {
  const float ex = tx + 1e-8f, ey = ty + 1e-8f, ez = tz + 1e-8f;
  const float a = sqrtf(ex * ex + ey * ey + ez * ez);
  const float rx = tx / a, ry = ty / a, rz = tz / a;
  float s, c;
  sincosf(a, &s, &c);
  const float k = 1.0f - c;
  // Gr = gR r ; r^T gR r ; trace
  const float g0 = gR[0] * rx + gR[1] * ry + gR[2] * rz;
  const float g1 = gR[3] * rx + gR[4] * ry + gR[5] * rz;
  const float g2 = gR[6] * rx + gR[7] * ry + gR[8] * rz;
  const float h0 = gR[0] * rx + gR[3] * ry + gR[6] * rz;   // gR^T r
  const float h1 = gR[1] * rx + gR[4] * ry + gR[7] * rz;
  const float h2 = gR[2] * rx + gR[5] * ry + gR[8] * rz;
  const float rGr = rx * g0 + ry * g1 + rz * g2;
  const float gc = (gR[0] + gR[4] + gR[8]) - rGr;          // d/dcos : I - r r^T
  const float ax = gR[7] - gR[5], ay = gR[2] - gR[6], az = gR[3] - gR[1];   // skew part of gR
  const float gs = rx * ax + ry * ay + rz * az;            // d/dsin : [r]x
  const float ga = -s * gc + c * gs;
  const float grx = k * (g0 + h0) + s * ax;
  const float gry = k * (g1 + h1) + s * ay;
  const float grz = k * (g2 + h2) + s * az;
  // r = theta / a (theta without eps), a = |theta + eps|
  const float t = (ga - (grx * tx + gry * ty + grz * tz) / (a * a)) / a;
  gth[0] = grx / a + t * ex;
  gth[1] = gry / a + t * ey;
  gth[2] = grz / a + t * ez;
}

// Adjoint of rodrigues_cam (project.py:51-81): R = I + sin K + (1-cos) K^2.  Zero gradient at exactly zero
// rotation (TF propagates NaN there through the dead tf.where branch).
__device__ __forceinline__ void rodrigues_cam_bwd(float ax, float ay, float az, const float gR[9], float gr[3]) {
  const float th = sqrtf(ax * ax + ay * ay + az * az);
  if (th == 0.0f) {
    gr[0] = gr[1] = gr[2] = 0.f;
    return;
  }
  const float ux = ax / th, uy = ay / th, uz = az / th;
  float s, c;
  sincosf(th, &s, &c);
  const float k = 1.0f - c;
  const float K[9] = {0.f, -uz, uy, uz, 0.f, -ux, -uy, ux, 0.f};
  float K2[9], gK[9];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) K2[i * 3 + j] = K[i * 3] * K[j] + K[i * 3 + 1] * K[3 + j] + K[i * 3 + 2] * K[6 + j];
  float gs = 0.f, gk = 0.f;
#pragma unroll
  for (int e = 0; e < 9; ++e) {
    gs += gR[e] * K[e];
    gk += gR[e] * K2[e];
  }
  const float gth = c * gs + s * gk;
  // d(K K)/dK : gR K^T + K^T gR
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      float v = s * gR[i * 3 + j];
#pragma unroll
      for (int m = 0; m < 3; ++m) v += k * (gR[i * 3 + m] * K[j * 3 + m] + K[m * 3 + i] * gR[m * 3 + j]);
      gK[i * 3 + j] = v;
    }
  const float gux = gK[7] - gK[5], guy = gK[2] - gK[6], guz = gK[3] - gK[1];
  const float t = (gth - (gux * ax + guy * ay + guz * az) / (th * th)) / th;
  gr[0] = gux / th + t * ax;
  gr[1] = guy / th + t * ay;
  gr[2] = guz / th + t * az;
}

// Adjoint of project_point wrt the camera-space point: g_uv -> g_q.
__device__ __forceinline__ void project_bwd_q(const float q[3], float f, float gu, float gv, float gq[3]) {
  gq[0] = f * gu / q[2];
  gq[1] = f * gv / q[2];
  gq[2] = -f * (gu * q[0] + gv * q[1]) / (q[2] * q[2]);
}

}  // namespace smplb
