// viekf_math.cuh -- device-side manifold math of the batch engine (sm_100a).
//
// Replaces, for the batch engine, the parts of lib/geometry/include/geometry/quat.h and include/math_helper.h the
// filter core touches (reference quat.h:128-162 exp/log, :306-322 R, :362-391 rotp/rota, :419-445 otimes/boxplus/
// boxminus, :233-255 from_two_unit_vectors; math_helper.h:14-48 zeta/T_zeta/q_feat_boxplus/boxminus).
// Conventions: Hamilton quaternion [w,x,y,z]; rotp(v) = R v (passive), rota(v) = R^T v (active).
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace vk {

struct V3 {
  double x, y, z;
};
struct Q4 {
  double w, x, y, z;
};

__device__ __forceinline__ V3 mk3(double x, double y, double z) { return V3{x, y, z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator*(double s, V3 a) { return V3{s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) { return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
__device__ __forceinline__ double norm(V3 a) { return sqrt(dot(a, a)); }
__device__ __forceinline__ double comp(V3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

__device__ __forceinline__ Q4 ldq(const double* p) { return Q4{p[0], p[1], p[2], p[3]}; }
__device__ __forceinline__ void stq(double* p, Q4 q) {
  p[0] = q.w;
  p[1] = q.x;
  p[2] = q.y;
  p[3] = q.z;
}
__device__ __forceinline__ V3 ld3(const double* p) { return V3{p[0], p[1], p[2]}; }

// t = 2 v x qbar;  rota = v - w t + t x qbar;  rotp = v + w t + t x qbar   (kept algebraically consistent with
// Rmat() below for non-unit q)
__device__ __forceinline__ V3 rota(Q4 q, V3 v) {
  V3 b = mk3(q.x, q.y, q.z);
  V3 t = 2.0 * cross(v, b);
  return v - q.w * t + cross(t, b);
}
__device__ __forceinline__ V3 rotp(Q4 q, V3 v) {
  V3 b = mk3(q.x, q.y, q.z);
  V3 t = 2.0 * cross(v, b);
  return v + q.w * t + cross(t, b);
}
// passive rotation matrix, row-major m[3*r+c]
__device__ __forceinline__ void Rmat(Q4 q, double* m) {
  double wx = q.w * q.x, wy = q.w * q.y, wz = q.w * q.z;
  double xx = q.x * q.x, xy = q.x * q.y, xz = q.x * q.z;
  double yy = q.y * q.y, yz = q.y * q.z, zz = q.z * q.z;
  m[0] = 1. - 2. * yy - 2. * zz;
  m[1] = 2. * xy + 2. * wz;
  m[2] = 2. * xz - 2. * wy;
  m[3] = 2. * xy - 2. * wz;
  m[4] = 1. - 2. * xx - 2. * zz;
  m[5] = 2. * yz + 2. * wx;
  m[6] = 2. * xz + 2. * wy;
  m[7] = 2. * yz - 2. * wx;
  m[8] = 1. - 2. * xx - 2. * yy;
}
__device__ __forceinline__ Q4 otimes(Q4 a, Q4 b) {
  return Q4{a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z, a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
            a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x, a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w};
}
__device__ __forceinline__ Q4 qinv(Q4 a) { return Q4{a.w, -a.x, -a.y, -a.z}; }
// Quat::exp (quat.h:128-146): [cos(|v|/2), sin(|v|/2) v/|v|] for |v| > 1e-4, the normalised first-order form below it.
// For 1e-4 < |v| < 0.25 -- every state correction and every propagate step of a healthy filter -- the two functions of |v|^2
// are evaluated as truncated Maclaurin series in h^2 = |v|^2/4 (first neglected terms h^14/14! < 3e-24 and h^14/15!): no square
// root, no division and no sincos argument reduction on what is the longest dependent chain of the state update (boxplus sits
// on the x warp's critical path once per event).  The result is within an ulp of the sincos form.
__device__ __forceinline__ Q4 qexp(V3 v) {
  const double n2 = dot(v, v);
  if (n2 < 0.0625 && n2 > 1e-8) {
    const double h2 = 0.25 * n2;
    double cs = fma(h2, 1.0 / 479001600.0, -1.0 / 3628800.0);
    double sc = fma(h2, 1.0 / 6227020800.0, -1.0 / 39916800.0);
    cs = fma(cs, h2, 1.0 / 40320.0);
    sc = fma(sc, h2, 1.0 / 362880.0);
    cs = fma(cs, h2, -1.0 / 720.0);
    sc = fma(sc, h2, -1.0 / 5040.0);
    cs = fma(cs, h2, 1.0 / 24.0);
    sc = fma(sc, h2, 1.0 / 120.0);
    cs = fma(cs, h2, -0.5);
    sc = fma(sc, h2, -1.0 / 6.0);
    cs = fma(cs, h2, 1.0);
    sc = fma(sc, h2, 1.0);
    const double s = 0.5 * sc;  // sin(|v|/2) / |v|
    return Q4{cs, s * v.x, s * v.y, s * v.z};
  }
  const double n = sqrt(n2);
  if (n > 1e-4) {
    double sn, cs;
    sincos(n / 2.0, &sn, &cs);  // one argument reduction for both
    const double s = sn / n;
    return Q4{cs, s * v.x, s * v.y, s * v.z};
  }
  Q4 o{1.0, v.x / 2.0, v.y / 2.0, v.z / 2.0};
  const double inn = 1.0 / sqrt(o.w * o.w + o.x * o.x + o.y * o.y + o.z * o.z);  // one reciprocal for the four components
  return Q4{o.w * inn, o.x * inn, o.y * inn, o.z * inn};
}
__device__ __forceinline__ V3 qlog(Q4 q) {
  V3 v = mk3(q.x, q.y, q.z);
  double n = norm(v);
  if (n < 1e-8) return mk3(0, 0, 0);
  double s = (2.0 * atan2(n, q.w)) / n;
  return mk3(s * v.x, s * v.y, s * v.z);
}
// a [-] b
__device__ __forceinline__ V3 qboxminus(Q4 a, Q4 b) {
  Q4 d = otimes(qinv(b), a);
  if (d.w < 0.0) d = Q4{-d.w, -d.x, -d.y, -d.z};
  return qlog(d);
}
__device__ __forceinline__ Q4 from_two_unit_vectors(V3 u, V3 v) {
  double d = dot(u, v);
  if (d < 0.99999999 && d > -0.99999999) {
    double invs = 1.0 / sqrt(2.0 * (1.0 + d));
    V3 c = cross(u, invs * v);
    Q4 o{0.5 / invs, c.x, c.y, c.z};
    double nn = sqrt(o.w * o.w + o.x * o.x + o.y * o.y + o.z * o.z);
    return Q4{o.w / nn, o.x / nn, o.y / nn, o.z / nn};
  } else if (d < -0.99999999) {
    return Q4{0, 1, 0, 0};
  }
  return Q4{1, 0, 0, 0};
}
__device__ __forceinline__ Q4 from_euler(double roll, double pitch, double yaw) {
  double cp = cos(roll / 2.0), ct = cos(pitch / 2.0), cs = cos(yaw / 2.0);
  double sp = sin(roll / 2.0), st = sin(pitch / 2.0), ss = sin(yaw / 2.0);
  return Q4{cp * ct * cs + sp * st * ss, sp * ct * cs - cp * st * ss, cp * st * cs + sp * ct * ss, cp * ct * ss - sp * st * cs};
}
__device__ __forceinline__ double qroll(Q4 q) { return atan2(2.0 * (q.w * q.x + q.y * q.z), 1.0 - 2.0 * (q.x * q.x + q.y * q.y)); }
__device__ __forceinline__ double qpitch(Q4 q) {
  double v = 2.0 * (q.w * q.y - q.x * q.z);
  if (fabs(v) > 1.0) return copysign(1.0, v) * M_PI / 2.0;
  return asin(v);
}
__device__ __forceinline__ double qyaw(Q4 q) { return atan2(2.0 * (q.w * q.z + q.x * q.y), 1.0 - 2.0 * (q.y * q.y + q.z * q.z)); }

// q.rota(e_x), q.rota(e_y), q.rota(e_z) = the rows of the passive rotation matrix (Rmat above), written out: the generic rota()
// spends two thirds of its multiplications on the zeros of the unit vector (x * 0 cannot be folded in IEEE arithmetic)
__device__ __forceinline__ V3 rota_ex(Q4 q) {
  return V3{1.0 - 2.0 * (q.y * q.y + q.z * q.z), 2.0 * (q.x * q.y + q.w * q.z), 2.0 * (q.x * q.z - q.w * q.y)};
}
__device__ __forceinline__ V3 rota_ey(Q4 q) {
  return V3{2.0 * (q.x * q.y - q.w * q.z), 1.0 - 2.0 * (q.x * q.x + q.z * q.z), 2.0 * (q.y * q.z + q.w * q.x)};
}
__device__ __forceinline__ V3 rota_ez(Q4 q) {
  return V3{2.0 * (q.x * q.z + q.w * q.y), 2.0 * (q.y * q.z - q.w * q.x), 1.0 - 2.0 * (q.x * q.x + q.y * q.y)};
}
// bearing-vector manifold: zeta = q.rota(e_z), T_zeta = [q.rota(e_x) q.rota(e_y)]
__device__ __forceinline__ void zeta_T(Q4 q, V3& zeta, V3& T0, V3& T1) {
  zeta = rota_ez(q);
  T0 = rota_ex(q);
  T1 = rota_ey(q);
}
// exp(T_zeta d) (x) q      (left multiplication)
__device__ __forceinline__ Q4 qfeat_boxplus(Q4 q, double d0, double d1) {
  V3 T0 = rota_ex(q), T1 = rota_ey(q);
  V3 d = mk3(T0.x * d0 + T1.x * d1, T0.y * d0 + T1.y * d1, T0.z * d0 + T1.z * d1);
  return otimes(qexp(d), q);
}
// q_j [-] q_i
__device__ __forceinline__ void qfeat_boxminus(Q4 qj, Q4 qi, double out[2]) {
  V3 zi = rota_ez(qi), zj = rota_ez(qj);
  if (norm(zi - zj) > 1e-8) {
    V3 s = cross(zi, zj);
    double ns = norm(s);
    s = mk3(s.x / ns, s.y / ns, s.z / ns);
    double th = acos(dot(zi, zj));
    V3 T0 = rota_ex(qi), T1 = rota_ey(qi);
    V3 ts = th * s;
    out[0] = dot(T0, ts);
    out[1] = dot(T1, ts);
  } else {
    out[0] = out[1] = 0.0;
  }
}

// packed lower-triangular (row-major) index; P(i,j) with i >= j lives at i(i+1)/2 + j
__device__ __forceinline__ int plo(int i, int j) { return ((i * (i + 1)) >> 1) + j; }
__device__ __forceinline__ int psym(int i, int j) { return i >= j ? plo(i, j) : plo(j, i); }

}  // namespace vk
