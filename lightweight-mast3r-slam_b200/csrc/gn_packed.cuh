// Two points per thread in packed fp32 (sm_100a FFMA2 / FMUL2 / FADD2 on register pairs).
// (Included by gn.cu inside namespace m3r, after the scalar Residual<> definitions it restates.)
//
// gn_accumulate_kernel is bound by instruction issue, not by memory: 247 warp instructions per round of 32
// list entries, 164 of them scalar fp32 arithmetic (95 FFMA, 61 FMUL, 8 FADD) and ~55 ring
// management.  With lane .x = entry e and lane .y = entry e + 32 of the same warp round every one of those
// arithmetic instructions serves two points, and so does every instruction of the ring.
// The formulas are those of Residual<KIND>::point (same reference lines); sums run over the .x and .y
// streams separately and meet in the warp reduction, so the per-item partial sums are, as before, a
// fixed function of the item: results do not depend on how the edges are sharded.
#pragma once

struct F2 {
  float2 v;
  __device__ __forceinline__ F2() {}
  __device__ __forceinline__ F2(float a, float b) : v(make_float2(a, b)) {}
  __device__ __forceinline__ explicit F2(float s) : v(make_float2(s, s)) {}
};
__device__ __forceinline__ F2 operator+(F2 a, F2 b) { F2 r; r.v = __fadd2_rn(a.v, b.v); return r; }
__device__ __forceinline__ F2 operator*(F2 a, F2 b) { F2 r; r.v = __fmul2_rn(a.v, b.v); return r; }
__device__ __forceinline__ F2 fma2(F2 a, F2 b, F2 c) { F2 r; r.v = __ffma2_rn(a.v, b.v, c.v); return r; }
__device__ __forceinline__ F2 neg(F2 a) { return F2(-a.v.x, -a.v.y); }
__device__ __forceinline__ F2 operator-(F2 a, F2 b) { return fma2(b, F2(-1.f), a); }  // one FFMA2
// c - a * b
__device__ __forceinline__ F2 fnma2(F2 a, F2 b, F2 c) { return fma2(neg(a), b, c); }
__device__ __forceinline__ F2 rsqrt2(F2 a) { return F2(rsqrt_fast(a.v.x), rsqrt_fast(a.v.y)); }
__device__ __forceinline__ F2 sqrt2(F2 a) { return F2(sqrt_fast(a.v.x), sqrt_fast(a.v.y)); }
__device__ __forceinline__ F2 rcp2(F2 a) { return F2(rcp_fast(a.v.x), rcp_fast(a.v.y)); }
__device__ __forceinline__ F2 huber_w2(F2 r) {  // min(1, k / |r|), gn_kernels.cu:172-175
  const F2 m = F2(rcp_fast(fabsf(r.v.x)), rcp_fast(fabsf(r.v.y))) * F2(1.345f);
  return F2(fminf(1.0f, m.v.x), fminf(1.0f, m.v.y));
}

template <unsigned MASK>
__device__ __forceinline__ void add_row2(F2 w, F2 e, const F2 J[7], F2 acc[kNAcc]) {
  const F2 we = w * e;
#pragma unroll
  for (int a = 0; a < 7; a++) {
    if (!((MASK >> a) & 1u)) continue;
    const F2 wa = w * J[a];
#pragma unroll
    for (int c = 0; c <= a; c++) {
      if (!((MASK >> c) & 1u)) continue;
      acc[a * (a + 1) / 2 + c] = fma2(wa, J[c], acc[a * (a + 1) / 2 + c]);
    }
    acc[kNS + a] = fma2(we, J[a], acc[kNS + a]);
  }
}

// the relative pose with every component in both halves of a register pair
struct Pose2 {
  F2 q[4], t[3], s;
  __device__ __forceinline__ explicit Pose2(const Pose &T) {
#pragma unroll
    for (int k = 0; k < 4; k++) q[k] = F2(T.q[k]);
#pragma unroll
    for (int k = 0; k < 3; k++) t[k] = F2(T.t[k]);
    s = F2(T.s);
  }
};

__device__ __forceinline__ void act_sim3_2(const Pose2 &T, const F2 X[3], F2 Y[3]) {
  // act_so3 (gn_kernels.cu:195-205) and scale / translation (:207-219)
  const F2 two(2.f);
  const F2 uv0 = two * fnma2(T.q[2], X[1], T.q[1] * X[2]);
  const F2 uv1 = two * fnma2(T.q[0], X[2], T.q[2] * X[0]);
  const F2 uv2 = two * fnma2(T.q[1], X[0], T.q[0] * X[1]);
  const F2 y0 = fma2(T.q[3], uv0, X[0]) + fnma2(T.q[2], uv1, T.q[1] * uv2);
  const F2 y1 = fma2(T.q[3], uv1, X[1]) + fnma2(T.q[0], uv2, T.q[2] * uv0);
  const F2 y2 = fma2(T.q[3], uv2, X[2]) + fnma2(T.q[1], uv0, T.q[0] * uv1);
  Y[0] = fma2(y0, T.s, T.t[0]);
  Y[1] = fma2(y1, T.s, T.t[1]);
  Y[2] = fma2(y2, T.s, T.t[2]);
}

template <int KIND>
struct Residual2;

// ray_align_kernel :910-1090
template <>
struct Residual2<M3R_GN_RAYS> {
  static __device__ __forceinline__ void point(const GnParams &P, const Pose2 &Tij, const F2 Xi[3],
                                               const F2 Xj[3], F2 q, int, int, F2 acc[kNAcc]) {
    const F2 n2i = fma2(Xi[0], Xi[0], fma2(Xi[1], Xi[1], Xi[2] * Xi[2]));
    const F2 n1i_inv = rsqrt2(n2i);
    const F2 n1i = n2i * n1i_inv;
    F2 Pj[3];
    act_sim3_2(Tij, Xj, Pj);
    const F2 n2j = fma2(Pj[0], Pj[0], fma2(Pj[1], Pj[1], Pj[2] * Pj[2]));
    const F2 n1j_inv = rsqrt2(n2j);
    const F2 n1j = n2j * n1j_inv;
    const F2 r[3] = {n1j_inv * Pj[0], n1j_inv * Pj[1], n1j_inv * Pj[2]};
    const F2 m1i = neg(n1i_inv);
    const F2 e0 = fma2(m1i, Xi[0], r[0]);
    const F2 e1 = fma2(m1i, Xi[1], r[1]);
    const F2 e2 = fma2(m1i, Xi[2], r[2]);
    const F2 e3 = n1j - n1i;
    const F2 sq = sqrt2(q);
    const F2 swr = F2(P.sigma_a_inv) * sq;
    const F2 swd = F2(P.sigma_b_inv) * sq;
    const F2 wr = swr * swr, wd = swd * swd;
    const F2 w0 = huber_w2(swr * e0) * wr;
    const F2 w1 = huber_w2(swr * e1) * wr;
    const F2 w2 = huber_w2(swr * e2) * wr;
    const F2 w3 = huber_w2(swd * e3) * wd;
    const F2 m3 = neg(n1j_inv * n1j_inv * n1j_inv);  // -1/|P|^3 (:982)
    const F2 axx = fma2(Pj[0] * Pj[0], m3, n1j_inv);
    const F2 ayy = fma2(Pj[1] * Pj[1], m3, n1j_inv);
    const F2 azz = fma2(Pj[2] * Pj[2], m3, n1j_inv);
    const F2 axy = Pj[0] * Pj[1] * m3;
    const F2 axz = Pj[0] * Pj[2] * m3;
    const F2 ayz = Pj[1] * Pj[2] * m3;
    const F2 z(0.f);
    const F2 nr0 = neg(r[0]), nr1 = neg(r[1]), nr2 = neg(r[2]);
    {
      const F2 J[7] = {axx, axy, axz, z, r[2], nr1, z};  // :991-997
      add_row2<0b0110111>(w0, e0, J, acc);
    }
    {
      const F2 J[7] = {axy, ayy, ayz, nr2, z, r[0], z};  // :1016-1022
      add_row2<0b0101111>(w1, e1, J, acc);
    }
    {
      const F2 J[7] = {axz, ayz, azz, r[1], nr0, z, z};  // :1041-1047
      add_row2<0b0011111>(w2, e2, J, acc);
    }
    {
      const F2 J[7] = {r[0], r[1], r[2], z, z, z, n1j};  // :1067-1073
      add_row2<0b1000111>(w3, e3, J, acc);
    }
  }
};

// point_align_kernel :540-700
template <>
struct Residual2<M3R_GN_POINTS> {
  static __device__ __forceinline__ void point(const GnParams &P, const Pose2 &Tij, const F2 Xi[3],
                                               const F2 Xj[3], F2 q, int, int, F2 acc[kNAcc]) {
    F2 Pj[3];
    act_sim3_2(Tij, Xj, Pj);
    const F2 e0 = Pj[0] - Xi[0], e1 = Pj[1] - Xi[1], e2 = Pj[2] - Xi[2];
    const F2 sw = F2(P.sigma_a_inv) * sqrt2(q);
    const F2 wc = sw * sw;
    const F2 w0 = huber_w2(sw * e0) * wc;
    const F2 w1 = huber_w2(sw * e1) * wc;
    const F2 w2 = huber_w2(sw * e2) * wc;
    const F2 z(0.f), one(1.f);
    const F2 n0 = neg(Pj[0]), n1 = neg(Pj[1]), n2 = neg(Pj[2]);
    {
      const F2 J[7] = {one, z, z, z, Pj[2], n1, Pj[0]};  // :602-608
      add_row2<0b1110001>(w0, e0, J, acc);
    }
    {
      const F2 J[7] = {z, one, z, n2, z, Pj[0], Pj[1]};  // :627-633
      add_row2<0b1101010>(w1, e1, J, acc);
    }
    {
      const F2 J[7] = {z, z, one, Pj[1], n0, z, Pj[2]};  // :652-658
      add_row2<0b1011100>(w2, e2, J, acc);
    }
  }
};

// calib_proj_kernel :1346-1497.  The scalar kernel leaves a point whose projection is not usable
// (:1368, :1382-1383); here such a half gets harmless inputs and a zero confidence, which zeroes all
// its weights (sqrt(0) = 0), and exact zero products change no sum.
template <>
struct Residual2<M3R_GN_CALIB> {
  static __device__ __forceinline__ void point(const GnParams &P, const Pose2 &Tij, const F2 Xi[3],
                                               const F2 Xj[3], F2 q, int ind0, int ind1, F2 acc[kNAcc]) {
    F2 Pj[3];
    act_sim3_2(Tij, Xj, Pj);
    const bool vz0 = (Pj[2].v.x > P.z_eps) && (Xi[2].v.x > P.z_eps);  // :1368
    const bool vz1 = (Pj[2].v.y > P.z_eps) && (Xi[2].v.y > P.z_eps);
    const F2 zj(vz0 ? Pj[2].v.x : 1.f, vz1 ? Pj[2].v.y : 1.f);
    const F2 zi(vz0 ? Xi[2].v.x : 1.f, vz1 ? Xi[2].v.y : 1.f);
    const F2 iz = rcp2(zj);
    const F2 x = Pj[0] * iz, y = Pj[1] * iz;
    const F2 fx(P.fx), fy(P.fy);
    const F2 u = fma2(fx, x, F2(P.cx));
    const F2 v = fma2(fy, y, F2(P.cy));
    const bool ok0 = vz0 && (u.v.x > P.u_lo) && (u.v.x < P.u_hi) && (v.v.x > P.v_lo) && (v.v.x < P.v_hi);  // :1382-1383
    const bool ok1 = vz1 && (u.v.y > P.u_lo) && (u.v.y < P.u_hi) && (v.v.y > P.v_lo) && (v.v.y < P.v_hi);
    if (!ok0 && !ok1) return;
    const int vt0 = ind0 / P.width, vt1 = ind1 / P.width;  // :1361-1362
    const F2 ut(static_cast<float>(ind0 - vt0 * P.width), static_cast<float>(ind1 - vt1 * P.width));
    const F2 vt(static_cast<float>(vt0), static_cast<float>(vt1));
    const F2 e0 = u - ut;
    const F2 e1 = v - vt;
    const F2 zr = zj * rcp2(zi);
    const F2 e2(__logf(zr.v.x), __logf(zr.v.y));  // log z_j->i - log z_i
    const F2 sq = sqrt2(F2(ok0 ? q.v.x : 0.f, ok1 ? q.v.y : 0.f));
    const F2 swp = F2(P.sigma_a_inv) * sq;
    const F2 swd = F2(P.sigma_b_inv) * sq;
    const F2 wp = swp * swp, wd = swd * swd;
    const F2 w0 = huber_w2(swp * e0) * wp;
    const F2 w1 = huber_w2(swp * e1) * wp;
    const F2 w2 = huber_w2(swd * e2) * wd;
    const F2 z(0.f), one(1.f);
    const F2 fxiz = fx * iz, fyiz = fy * iz;
    const F2 fxx = fx * x, fyy = fy * y;
    {
      const F2 J[7] = {fxiz, z, neg(fxx * iz), neg(fxx * y), fma2(fxx, x, fx), neg(fx * y), z};  // :1422-1428
      add_row2<0b0111101>(w0, e0, J, acc);
    }
    {
      const F2 J[7] = {z, fyiz, neg(fyy * iz), neg(fma2(fyy, y, fy)), fyy * x, fy * x, z};  // :1448-1454
      add_row2<0b0111110>(w1, e1, J, acc);
    }
    {
      const F2 J[7] = {z, z, iz, y, neg(x), z, one};  // :1473-1479
      add_row2<0b1011100>(w2, e2, J, acc);
    }
  }
};
