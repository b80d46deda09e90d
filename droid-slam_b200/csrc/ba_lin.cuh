// Per pixel-edge linearisation shared by the BA kernels (ba.cu, schur_tc.cu).
#pragma once
#include "common.cuh"

namespace droid {

// ------------------------------------------------------------------------------------------------
// per pixel-edge linearisation (src/droid_kernels.cu:283-375), target-pose Jacobian only.
// The Jacobian w.r.t. the SOURCE pose is not formed per pixel: Ji = -Ad(g_ij)^T Jj is a LINEAR image of Jj with a matrix that
// depends on the edge only, not on the pixel, so the kernels keep just Jj per pixel and apply the 6x6 matrix
// N = -Ad(g_ij)^T (edge_adjoint below, stored next to the relative pose) once per edge / once per 6-vector:
// the reference evaluates the adjoint twice per pixel-edge (src/droid_kernels.cu:336-338, 368-371), which is
// ~220 of the ~250 flops of its linearisation.
struct PixLinJ {
  float Jj_u[6], Jj_v[6];
  float Jz_u, Jz_v;
  float wu_d, wv_d;
  float wu, wv;
  float ru, rv;
  float xd, yd, hd;   // x/z, y/z, h/z of the transformed point (what the Jacobian rows are made of)
};

// The Schur kernels need, per pixel-edge, only the 6-vector E_j = wu Jz_u Jj_u + wv Jz_v Jj_v.  The accumulation
// kernel has the linearisation in registers anyway and leaves E_j behind in 5 floats,
//   a = wu Jz_u fx,  b = wv Jz_v fy,  xd,  yd,  hd,
// from which the six components follow with a dozen flops (Jj_u = fx [hd, 0, -xd hd, -xd yd, 1 + xd^2, -yd],
// Jj_v = fy [0, hd, -yd hd, -1 - yd^2, xd yd, xd]); the Schur kernels then stream 20 bytes per pixel-edge instead of
// redoing the projection.  Layout: ejc[(edge * 5 + component) * HW + pixel].
static constexpr int kEjComp = 5;

__device__ __forceinline__ void ej_pack(const PixLinJ& L, const Intr& K, float (&c)[5]) {
  c[0] = (L.wu * L.Jz_u) * K.fx;
  c[1] = (L.wv * L.Jz_v) * K.fy;
  c[2] = L.xd; c[3] = L.yd; c[4] = L.hd;
}

__device__ __forceinline__ void ej_unpack(const float (&c)[5], float (&ej)[6]) {
  const float a = c[0], b = c[1], xd = c[2], yd = c[3], hd = c[4];
  const float xy = xd * yd;
  ej[0] = a * hd;
  ej[1] = b * hd;
  ej[2] = -(a * xd + b * yd) * hd;
  ej[3] = -a * xy - b * (1.f + yd * yd);
  ej[4] = a * (1.f + xd * xd) + b * xy;
  ej[5] = b * xd - a * yd;
}

// x0, y0 = normalised image coordinates (u - cx) / fx, (v - cy) / fy of the pixel
__device__ __forceinline__ void linearise_j_xy(const SE3& g, const Intr& K, float x0, float y0, float disp,
                                               float tu, float tv, float wru, float wrv, bool stereo,
                                               PixLinJ& L) {
  float Xi[4], Xj[4];
  Xi[0] = x0;
  Xi[1] = y0;
  Xi[2] = 1;
  Xi[3] = disp;
  act_se3(g.t, g.q, Xi, Xj);
  const float x = Xj[0], y = Xj[1], h = Xj[3];
  const bool behind = Xj[2] < DROID_MIN_DEPTH;
  // the reference divides in double and rounds to float (:305); the IEEE-rounded float reciprocal is the same value
  // except for double-rounding ties (~2^-29 of the inputs) and keeps the slow fp64 pipe off the critical chain
  const float d = behind ? 0.0f : rcp_depth(Xj[2]);
  const float d2 = d * d;
  L.wu_d = behind ? 0.0f : (float)(.001 * (double)wru);
  L.wv_d = behind ? 0.0f : (float)(.001 * (double)wrv);
  L.ru = tu - (K.fx * d * x + K.cx);
  L.rv = tv - (K.fy * d * y + K.cy);
  L.Jj_u[0] = K.fx * (h * d);
  L.Jj_u[1] = K.fx * 0;
  L.Jj_u[2] = K.fx * (-x * h * d2);
  L.Jj_u[3] = K.fx * (-x * y * d2);
  L.Jj_u[4] = K.fx * (1 + x * x * d2);
  L.Jj_u[5] = K.fx * (-y * d);
  L.Jz_u = K.fx * (g.t[0] * d - g.t[2] * (x * d2));
  L.Jj_v[0] = K.fy * 0;
  L.Jj_v[1] = K.fy * (h * d);
  L.Jj_v[2] = K.fy * (-y * h * d2);
  L.Jj_v[3] = K.fy * (-1 - y * y * d2);
  L.Jj_v[4] = K.fy * (x * y * d2);
  L.Jj_v[5] = K.fy * (x * d);
  L.Jz_v = K.fy * (g.t[1] * d - g.t[2] * (y * d2));
  L.wu = stereo ? 0.f : L.wu_d;
  L.wv = stereo ? 0.f : L.wv_d;
  L.xd = x * d; L.yd = y * d; L.hd = h * d;
}

__device__ __forceinline__ void linearise_j(const SE3& g, const Intr& K, int wd, int k, float disp,
                                            float tu, float tv, float wru, float wrv, bool stereo,
                                            PixLinJ& L) {
  const int i = k / wd, j = k - i * wd;
  linearise_j_xy(g, K, ((float)j - K.cx) / K.fx, ((float)i - K.cy) / K.fy, disp, tu, tv, wru, wrv, stereo, L);
}

// N[n][a] with Ji[n] = sum_a N[n][a] Jj[a]  (row-major 6x6): column a is -adj_se3(g, e_a)
__device__ __forceinline__ void edge_adjoint(const SE3& g, float* N /*[36]*/) {
#pragma unroll
  for (int a = 0; a < 6; a++) {
    float e[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, y[6];
    e[a] = 1.f;
    adj_se3(g.t, g.q, e, y);
#pragma unroll
    for (int n = 0; n < 6; n++) N[n * 6 + a] = -y[n];
  }
}

}  // namespace droid
