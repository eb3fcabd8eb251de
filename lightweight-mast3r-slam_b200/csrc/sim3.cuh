// Sim(3) group helpers shared by the graph Gauss-Newton (gn.cu) and the tracking pose solve
// (tracking.cu).  Pose layout (t xyz, q xyzw, s) as lietorch.Sim3.data / the reference's Twc rows.
#pragma once

namespace m3r {

struct Pose {
  float t[3], q[4], s;
};

__device__ __forceinline__ Pose load_pose(const float *__restrict__ Twc, int k) {
  Pose P;
  const float4 a = *reinterpret_cast<const float4 *>(Twc + k * 8);
  const float4 b = *reinterpret_cast<const float4 *>(Twc + k * 8 + 4);
  P.t[0] = a.x, P.t[1] = a.y, P.t[2] = a.z;
  P.q[0] = a.w, P.q[1] = b.x, P.q[2] = b.y, P.q[3] = b.z;
  P.s = b.w;
  return P;
}

template <typename T>
__device__ __forceinline__ void act_so3(const T q[4], const T X[3], T Y[3]) {
  // gn_kernels.cu:195-205
  const T uv0 = T(2) * (q[1] * X[2] - q[2] * X[1]);
  const T uv1 = T(2) * (q[2] * X[0] - q[0] * X[2]);
  const T uv2 = T(2) * (q[0] * X[1] - q[1] * X[0]);
  const T y0 = X[0] + q[3] * uv0 + (q[1] * uv2 - q[2] * uv1);
  const T y1 = X[1] + q[3] * uv1 + (q[2] * uv0 - q[0] * uv2);
  const T y2 = X[2] + q[3] * uv2 + (q[0] * uv1 - q[1] * uv0);
  Y[0] = y0, Y[1] = y1, Y[2] = y2;
}

__device__ __forceinline__ void quat_mul(const float a[4], const float b[4], float o[4]) {
  // gn_kernels.cu:178-184
  o[0] = a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1];
  o[1] = a[3] * b[1] - a[0] * b[2] + a[1] * b[3] + a[2] * b[0];
  o[2] = a[3] * b[2] + a[0] * b[1] - a[1] * b[0] + a[2] * b[3];
  o[3] = a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2];
}

// T_ij = T_i^-1 T_j  (relSim3, gn_kernels.cu:252-272)
__device__ __forceinline__ Pose rel_sim3(const Pose &Ti, const Pose &Tj) {
  Pose R;
  const float si_inv = __frcp_rn(Ti.s);
  R.s = si_inv * Tj.s;
  const float qi_inv[4] = {-Ti.q[0], -Ti.q[1], -Ti.q[2], Ti.q[3]};
  quat_mul(qi_inv, Tj.q, R.q);
  float d[3] = {Tj.t[0] - Ti.t[0], Tj.t[1] - Ti.t[1], Tj.t[2] - Ti.t[2]};
  act_so3<float>(qi_inv, d, d);
  R.t[0] = d[0] * si_inv, R.t[1] = d[1] * si_inv, R.t[2] = d[2] * si_inv;
  return R;
}

__device__ __forceinline__ void act_sim3(const Pose &T, const float X[3], float Y[3]) {
  // gn_kernels.cu:207-219
  act_so3<float>(T.q, X, Y);
  Y[0] = Y[0] * T.s + T.t[0];
  Y[1] = Y[1] * T.s + T.t[1];
  Y[2] = Y[2] * T.s + T.t[2];
}

__device__ inline void exp_so3(const float phi[3], float q[4]) {
  // gn_kernels.cu:299-321
  const float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  float imag, real;
  if (theta_sq < 1e-6f) {
    const float theta_p4 = theta_sq * theta_sq;
    imag = 0.5f - (1.0f / 48.0f) * theta_sq + (1.0f / 3840.0f) * theta_p4;
    real = 1.0f - (1.0f / 8.0f) * theta_sq + (1.0f / 384.0f) * theta_p4;
  } else {
    const float theta = sqrtf(theta_sq);
    imag = sinf(0.5f * theta) / theta;
    real = cosf(0.5f * theta);
  }
  q[0] = imag * phi[0], q[1] = imag * phi[1], q[2] = imag * phi[2], q[3] = real;
}

__device__ __forceinline__ void cross_inplace(const float a[3], float b[3]) {
  const float x0 = a[1] * b[2] - a[2] * b[1];
  const float x1 = a[2] * b[0] - a[0] * b[2];
  const float x2 = a[0] * b[1] - a[1] * b[0];
  b[0] = x0, b[1] = x1, b[2] = x2;
}

__device__ inline void exp_sim3(const float xi[7], float t[3], float q[4], float &s) {
  // gn_kernels.cu:323-390 (lietorch rxso3 W-matrix coefficients, small-angle/-scale branches)
  float tau[3] = {xi[0], xi[1], xi[2]};
  const float phi[3] = {xi[3], xi[4], xi[5]};
  const float sigma = xi[6];
  const float scale = expf(sigma);
  exp_so3(phi, q);
  s = scale;
  const float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  const float theta = sqrtf(theta_sq);
  float A, B, C;
  const float EPSF = 1e-6f;
  if (fabsf(sigma) < EPSF) {
    C = 1.0f;
    if (fabsf(theta) < EPSF) {
      A = 0.5f;
      B = 1.0f / 6.0f;
    } else {
      A = (1.0f - cosf(theta)) / theta_sq;
      B = (theta - sinf(theta)) / (theta_sq * theta);
    }
  } else {
    C = (scale - 1.0f) / sigma;
    if (fabsf(theta) < EPSF) {
      const float sigma_sq = sigma * sigma;
      A = ((sigma - 1.0f) * scale + 1.0f) / sigma_sq;
      B = (scale * 0.5f * sigma_sq + scale - 1.0f - sigma * scale) / (sigma_sq * sigma);
    } else {
      const float a = scale * sinf(theta);
      const float b = scale * cosf(theta);
      const float c = theta_sq + sigma * sigma;
      A = (a * sigma + (1.0f - b) * theta) / (theta * c);
      B = (C - ((b - 1.0f) * sigma + a * theta) / c) / theta_sq;
    }
  }
  t[0] = C * tau[0], t[1] = C * tau[1], t[2] = C * tau[2];
  cross_inplace(phi, tau);
  t[0] += A * tau[0], t[1] += A * tau[1], t[2] += A * tau[2];
  cross_inplace(phi, tau);
  t[0] += B * tau[0], t[1] += B * tau[1], t[2] += B * tau[2];
}


// X <- Exp(xi) * X  (retrSim3, gn_kernels.cu:392-413; lietorch's retr is this left update)
__device__ inline void retr_sim3(const float xi[7], Pose &X) {
  float dt[3], dq[4], ds;
  exp_sim3(xi, dt, dq, ds);
  float q1[4], t1[3];
  quat_mul(dq, X.q, q1);
  act_so3<float>(dq, X.t, t1);
  X.t[0] = t1[0] * ds + dt[0];
  X.t[1] = t1[1] * ds + dt[1];
  X.t[2] = t1[2] * ds + dt[2];
  X.q[0] = q1[0], X.q[1] = q1[1], X.q[2] = q1[2], X.q[3] = q1[3];
  X.s = ds * X.s;
}

// A * B  (composition: t = sA RA tB + tA, q = qA qB, s = sA sB)
__device__ inline Pose mul_sim3(const Pose &A, const Pose &B) {
  Pose R;
  float t[3];
  act_so3<float>(A.q, B.t, t);
  R.t[0] = A.s * t[0] + A.t[0], R.t[1] = A.s * t[1] + A.t[1], R.t[2] = A.s * t[2] + A.t[2];
  quat_mul(A.q, B.q, R.q);
  R.s = A.s * B.s;
  return R;
}

}  // namespace m3r
