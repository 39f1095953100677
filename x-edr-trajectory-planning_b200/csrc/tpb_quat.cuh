// Quaternion arithmetic of the reference's quaternion B-spline (splines/bsplineq.cc:99-160)
// with Eigen's scalar formulas for product, inverse, stableNorm and stableNormalized
// (DESIGN.md lists them as assumptions about the un-vendored Eigen).  FP64, no FMA
// contraction.
#pragma once

#include "tpb_device.cuh"

namespace tpb {

constexpr double kDummyPrecision = 1e-12;  // Eigen::NumTraits<double>::dummy_precision()

struct Quat {
  double x, y, z, w;
};

__device__ __forceinline__ double sq_norm(const Quat& q) {
  return q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w;
}
// Eigen stableNorm() of the vector part: scale * sqrt(sum((v/scale)^2))
__device__ __forceinline__ double stable_norm3(const Quat& q) {
  double scale = fabs(q.x);
  if (fabs(q.y) > scale) scale = fabs(q.y);
  if (fabs(q.z) > scale) scale = fabs(q.z);
  if (!(scale > 0)) return 0.0;
  const double inv = 1.0 / scale;
  const double a = q.x * inv, b = q.y * inv, c = q.z * inv;
  return scale * sqrt(a * a + b * b + c * c);
}
// Eigen stableNormalized() of the vector part
__device__ __forceinline__ void stable_normalized3(const Quat& q, double* out) {
  double w = fabs(q.x);
  if (fabs(q.y) > w) w = fabs(q.y);
  if (fabs(q.z) > w) w = fabs(q.z);
  double z = 0.0;
  if (w > 0) {
    const double a = q.x / w, b = q.y / w, c = q.z / w;
    z = a * a + b * b + c * c;
  }
  if (z > 0) {
    const double d = sqrt(z) * w;
    out[0] = q.x / d;
    out[1] = q.y / d;
    out[2] = q.z / d;
  } else {
    out[0] = q.x;
    out[1] = q.y;
    out[2] = q.z;
  }
}
// Eigen quat_product, scalar formula
__device__ __forceinline__ Quat qmul(const Quat& a, const Quat& b) {
  Quat r;
  r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
  r.y = a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z;
  r.z = a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x;
  r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
  return r;
}
// QuaternionBase::inverse(): conjugate / squaredNorm
__device__ __forceinline__ Quat qinv(const Quat& q) {
  const double n2 = sq_norm(q);
  Quat r;
  if (n2 > 0) {
    r.x = -q.x / n2;
    r.y = -q.y / n2;
    r.z = -q.z / n2;
    r.w = q.w / n2;
  } else {
    r.x = r.y = r.z = r.w = 0.0;
  }
  return r;
}
// QuaternionBase::_transformVector
__device__ __forceinline__ void qrot(const Quat& q, const double* v, double* out) {
  double ux = q.y * v[2] - q.z * v[1];
  double uy = q.z * v[0] - q.x * v[2];
  double uz = q.x * v[1] - q.y * v[0];
  ux += ux;
  uy += uy;
  uz += uz;
  const double cx = q.y * uz - q.z * uy;
  const double cy = q.z * ux - q.x * uz;
  const double cz = q.x * uy - q.y * ux;
  out[0] = v[0] + q.w * ux + cx;
  out[1] = v[1] + q.w * uy + cy;
  out[2] = v[2] + q.w * uz + cz;
}
// NormalizeIfNecessaryAndEnsurePositiveReal, splines/bsplineq.cc:99-109
__device__ __forceinline__ void normalize_positive_real(Quat& q) {
  if (q.w < 0) {
    q.x *= -1.0;
    q.y *= -1.0;
    q.z *= -1.0;
    q.w *= -1.0;
  }
  if (fabs(sq_norm(q) - 1.0) > kDummyPrecision) {
    const double n = sqrt(sq_norm(q));
    q.x /= n;
    q.y /= n;
    q.z /= n;
    q.w /= n;
  }
}
// QuatLog, splines/bsplineq.cc:113-126
__device__ __forceinline__ Quat qlog(const Quat& q) {
  const double squared_norm_q = sq_norm(q);
  const double norm_v = stable_norm3(q);
  Quat r;
  r.w = 0.5 * log(squared_norm_q);
  if (norm_v > kDummyPrecision) {
    double n[3];
    stable_normalized3(q, n);
    const double ang = atan2(norm_v, q.w);
    r.x = n[0] * ang;
    r.y = n[1] * ang;
    r.z = n[2] * ang;
  } else {
    r.x = q.x;
    r.y = q.y;
    r.z = q.z;
  }
  return r;
}
// QuatExp, splines/bsplineq.cc:128-135
__device__ __forceinline__ Quat qexp(const Quat& q) {
  const double norm_v = stable_norm3(q);
  double n[3];
  stable_normalized3(q, n);
  const double s = sin(norm_v);
  const double e = exp(q.w);
  Quat r;
  r.w = cos(norm_v);
  r.x = n[0] * s;
  r.y = n[1] * s;
  r.z = n[2] * s;
  r.x *= e;
  r.y *= e;
  r.z *= e;
  r.w *= e;
  return r;
}
// QuatPower, splines/bsplineq.cc:137-146
__device__ __forceinline__ Quat qpow(Quat q, double power) {
  normalize_positive_real(q);
  Quat l = qlog(q);
  l.x *= power;
  l.y *= power;
  l.z *= power;
  l.w *= power;
  return qexp(l);
}

__device__ __forceinline__ Quat load_quat(const double* p) { return Quat{p[0], p[1], p[2], p[3]}; }

// Eigen QuaternionBase::slerp(t, other)
__device__ __forceinline__ Quat qslerp(const Quat& a, double t, const Quat& b) {
  const double one = 1.0 - 2.220446049250313e-16;
  const double d = a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w;
  const double abs_d = fabs(d);
  double scale0, scale1;
  if (abs_d >= one) {
    scale0 = 1.0 - t;
    scale1 = t;
  } else {
    const double theta = acos(abs_d);
    const double sin_theta = sin(theta);
    scale0 = sin((1.0 - t) * theta) / sin_theta;
    scale1 = sin(t * theta) / sin_theta;
  }
  if (d < 0.0) scale1 = -scale1;
  return Quat{scale0 * a.x + scale1 * b.x, scale0 * a.y + scale1 * b.y,
              scale0 * a.z + scale1 * b.z, scale0 * a.w + scale1 * b.w};
}

// QuatPowerDerivative, splines/bsplineq.cc:148-160: exp(p log q) * log q
__device__ __forceinline__ Quat qpow_derivative(Quat q, double power) {
  normalize_positive_real(q);
  const Quat l = qlog(q);
  Quat s = l;
  s.x *= power;
  s.y *= power;
  s.z *= power;
  s.w *= power;
  return qmul(qexp(s), l);
}

}  // namespace tpb
