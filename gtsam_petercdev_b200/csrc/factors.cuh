// factors.cuh -- per-bucket factor kernels: linearize (Jacobians + whitening), nonlinear error, linear error,
// Hessian diagonal and retract.  One thread per factor / variable, one launch per bucket; replaces the
// per-factor virtual dispatch of NonlinearFactorGraph::linearize (nonlinear/NonlinearFactorGraph.cpp:239-278).
//
// All arithmetic is FP64 and follows the reference's default-build conventions:
//   Pose2 first-order chart            geometry/Pose2.cpp:100-122
//   Pose3 full SE(3) Expmap / Logmap   geometry/Pose3.cpp:184-245, geometry/SO3.cpp:61-112, 299-375
//   BetweenFactor: H1 = -Ad(hx^-1), H2 = I, Local Jacobian NOT chained     slam/BetweenFactor.h:111-124
//   PriorFactor:   H = I                                                    nonlinear/PriorFactor.h:98-102
//   GeneralSFMFactor<PinholeCamera<Cal3Bundler>>                            slam/GeneralSFMFactor.h:141-177
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace gtb {

enum { NK_UNIT = 0, NK_ISO = 1, NK_DIAG = 2, NK_FULL = 3 };

// device-side state sizes (doubles per row)
__host__ __device__ constexpr int state_size(int vtype) { return vtype == 0 ? 4 : vtype == 1 ? 12 : vtype == 2 ? 17 : 3; }
__host__ __device__ constexpr int tangent_dim(int vtype) { return vtype == 0 ? 3 : vtype == 1 ? 6 : vtype == 2 ? 9 : 3; }

struct BucketDev {
  int32_t n;
  int32_t any_noise;         // 0: every factor is Unit
  const int32_t* row0;       // row of key 0 in its type array
  const int32_t* row1;       // row of key 1 (binary buckets)
  const int32_t* goff0;      // global scalar offset of key 0's delta block
  const int32_t* goff1;
  const double* z;           // n x ZD
  const int8_t* nk;          // resolved noise kind per factor
  const double* noise;       // n x d*d
  double* Ab;                // n x m x W, row-major per factor
  int32_t loss;              // robust m-estimator of the whole bucket (0 = none), Block reweighting
  int32_t pad;
  double loss_k;
};

template <int B> struct BucketTraits;
//                                              m   n0  n1  W            ZD  vtype0 vtype1
template <> struct BucketTraits<0> { enum { M = 3, N0 = 3, N1 = 3, W = 7,  ZD = 4,  T0 = 0, T1 = 0 }; };
template <> struct BucketTraits<1> { enum { M = 6, N0 = 6, N1 = 6, W = 13, ZD = 12, T0 = 1, T1 = 1 }; };
template <> struct BucketTraits<2> { enum { M = 2, N0 = 9, N1 = 3, W = 13, ZD = 2,  T0 = 2, T1 = 3 }; };
template <> struct BucketTraits<3> { enum { M = 3, N0 = 3, N1 = 0, W = 4,  ZD = 4,  T0 = 0, T1 = -1 }; };
template <> struct BucketTraits<4> { enum { M = 6, N0 = 6, N1 = 0, W = 7,  ZD = 12, T0 = 1, T1 = -1 }; };
template <> struct BucketTraits<5> { enum { M = 9, N0 = 9, N1 = 0, W = 10, ZD = 17, T0 = 2, T1 = -1 }; };
template <> struct BucketTraits<6> { enum { M = 3, N0 = 3, N1 = 0, W = 4,  ZD = 3,  T0 = 3, T1 = -1 }; };

// ----------------------------------------------------------------------------------------------------
// geometry
// ----------------------------------------------------------------------------------------------------
struct P2 { double c, s, x, y; };

// Rot2::fromCosSin -> normalize (geometry/Rot2.cpp:27-30, 56-64)
__device__ __forceinline__ void rot2_normalize(double& c, double& s) {
  double scale = c * c + s * s;
  if (fabs(scale - 1.0) > 1e-10) {
    scale = 1.0 / sqrt(scale);
    c *= scale;
    s *= scale;
  }
}
// Pose2::inverse (geometry/Pose2.cpp:202-204)
__device__ __forceinline__ P2 pose2_inverse(const P2& p) {
  P2 r;
  r.c = p.c;
  r.s = -p.s;
  const double px = -p.x, py = -p.y;
  r.x = p.c * px + p.s * py;      // Rot2::unrotate (geometry/Rot2.cpp:110-116)
  r.y = -p.s * px + p.c * py;
  return r;
}
// Pose2::operator* (geometry/Pose2.h:138-140), Rot2::operator* (geometry/Rot2.h:120-122)
__device__ __forceinline__ P2 pose2_compose(const P2& a, const P2& b) {
  P2 r;
  r.c = a.c * b.c - a.s * b.s;
  r.s = a.s * b.c + a.c * b.s;
  rot2_normalize(r.c, r.s);
  r.x = a.x + (a.c * b.x + -a.s * b.y);
  r.y = a.y + (a.s * b.x + a.c * b.y);
  return r;
}
__device__ __forceinline__ P2 pose2_load(const double* p) { return P2{p[0], p[1], p[2], p[3]}; }

// Pose3 as R (row-major) + t
struct P3 { double R[9]; double t[3]; };
__device__ __forceinline__ P3 pose3_load(const double* p) {
  P3 r;
#pragma unroll
  for (int i = 0; i < 9; ++i) r.R[i] = p[i];
#pragma unroll
  for (int i = 0; i < 3; ++i) r.t[i] = p[9 + i];
  return r;
}
// Pose3::inverse (geometry/Pose3.cpp:61-64): Rt, Rt * (-t)
__device__ __forceinline__ P3 pose3_inverse(const P3& p) {
  P3 r;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) r.R[3 * i + j] = p.R[3 * j + i];
  const double nx = -p.t[0], ny = -p.t[1], nz = -p.t[2];
#pragma unroll
  for (int i = 0; i < 3; ++i) r.t[i] = r.R[3 * i] * nx + r.R[3 * i + 1] * ny + r.R[3 * i + 2] * nz;
  return r;
}
// Pose3::operator* (geometry/Pose3.h:117-119)
__device__ __forceinline__ P3 pose3_compose(const P3& a, const P3& b) {
  P3 r;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
      r.R[3 * i + j] = a.R[3 * i] * b.R[j] + a.R[3 * i + 1] * b.R[3 + j] + a.R[3 * i + 2] * b.R[6 + j];
#pragma unroll
  for (int i = 0; i < 3; ++i)
    r.t[i] = a.t[i] + (a.R[3 * i] * b.t[0] + a.R[3 * i + 1] * b.t[1] + a.R[3 * i + 2] * b.t[2]);
  return r;
}

// SO3::Logmap (geometry/SO3.cpp:299-375)
__device__ inline void so3_logmap(const double* R, double* omega) {
  const double R11 = R[0], R12 = R[1], R13 = R[2];
  const double R21 = R[3], R22 = R[4], R23 = R[5];
  const double R31 = R[6], R32 = R[7], R33 = R[8];
  const double tr = R11 + R22 + R33;
  const double kPi = 3.14159265358979323846;
  if (tr + 1.0 < 1e-3) {
    double W, Q1, Q2, Q3;
    int perm;
    if (R33 > R22 && R33 > R11) {
      W = R21 - R12; Q1 = 2.0 + 2.0 * R33; Q2 = R31 + R13; Q3 = R23 + R32; perm = 0;
    } else if (R22 > R11) {
      W = R13 - R31; Q1 = 2.0 + 2.0 * R22; Q2 = R23 + R32; Q3 = R12 + R21; perm = 1;
    } else {
      W = R32 - R23; Q1 = 2.0 + 2.0 * R11; Q2 = R12 + R21; Q3 = R31 + R13; perm = 2;
    }
    const double r = sqrt(Q1);
    const double one_over_r = 1 / r;
    const double norm = sqrt(Q1 * Q1 + Q2 * Q2 + Q3 * Q3 + W * W);
    const double sgn_w = W < 0 ? -1.0 : 1.0;
    const double mag = kPi - (2 * sgn_w * W) / norm;
    const double scale = 0.5 * one_over_r * mag;
    const double k = sgn_w * scale;
    if (perm == 0) { omega[0] = k * Q2; omega[1] = k * Q3; omega[2] = k * Q1; }
    else if (perm == 1) { omega[0] = k * Q3; omega[1] = k * Q1; omega[2] = k * Q2; }
    else { omega[0] = k * Q1; omega[1] = k * Q2; omega[2] = k * Q3; }
  } else {
    double magnitude;
    const double tr_3 = tr - 3.0;
    if (tr_3 < -1e-6) {
      const double theta = acos((tr - 1.0) / 2.0);
      magnitude = theta / (2.0 * sin(theta));
    } else {
      magnitude = 0.5 - tr_3 / 12.0 + tr_3 * tr_3 / 60.0;
    }
    omega[0] = magnitude * (R32 - R23);
    omega[1] = magnitude * (R13 - R31);
    omega[2] = magnitude * (R21 - R12);
  }
}

// Pose3::Logmap (geometry/Pose3.cpp:225-245); tangent order (omega, v)
__device__ inline void pose3_logmap(const P3& p, double* xi) {
  double w[3];
  so3_logmap(p.R, w);
  const double* T = p.t;
  const double t = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
  xi[0] = w[0]; xi[1] = w[1]; xi[2] = w[2];
  if (t < 1e-10) {
    xi[3] = T[0]; xi[4] = T[1]; xi[5] = T[2];
  } else {
    const double ax = w[0] / t, ay = w[1] / t, az = w[2] / t;  // W = skew(w / t)
    const double Tan = tan(0.5 * t);
    // WT = W * T
    const double WT0 = -az * T[1] + ay * T[2];
    const double WT1 = az * T[0] - ax * T[2];
    const double WT2 = -ay * T[0] + ax * T[1];
    const double WWT0 = -az * WT1 + ay * WT2;
    const double WWT1 = az * WT0 - ax * WT2;
    const double WWT2 = -ay * WT0 + ax * WT1;
    const double a = 0.5 * t, b = 1 - t / (2. * Tan);
    xi[3] = T[0] - a * WT0 + b * WWT0;
    xi[4] = T[1] - a * WT1 + b * WWT1;
    xi[5] = T[2] - a * WT2 + b * WWT2;
  }
}

// Pose3::Expmap (geometry/Pose3.cpp:184-222) with so3::DexpFunctor coefficients (geometry/SO3.cpp:61-112)
__device__ inline P3 pose3_expmap(const double* xi) {
  const double wx = xi[0], wy = xi[1], wz = xi[2];
  const double vx = xi[3], vy = xi[4], vz = xi[5];
  const double theta2 = wx * wx + wy * wy + wz * wz;
  const bool nearZero = (theta2 <= 1e-5) || (theta2 <= 2.220446049250313e-16);
  double A, B, C;
  if (!nearZero) {
    const double theta = sqrt(theta2);
    A = sin(theta) / theta;
    const double s2 = sin(theta / 2.0);
    B = 2.0 * s2 * s2 / theta2;
    C = (1 - A) / theta2;
  } else {
    A = 1.0 - theta2 * (1.0 / 6.0);
    B = 0.5 - theta2 * (1.0 / 24.0);
    C = (1.0 / 6.0) - theta2 * (1.0 / 120.0);
  }
  // W = skew(w), WW = W * W
  const double W[9] = {0.0, -wz, wy, wz, 0.0, -wx, -wy, wx, 0.0};
  double WW[9];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) WW[3 * i + j] = W[3 * i] * W[j] + W[3 * i + 1] * W[3 + j] + W[3 * i + 2] * W[6 + j];
  P3 r;
#pragma unroll
  for (int i = 0; i < 9; ++i) r.R[i] = ((i % 4 == 0) ? 1.0 : 0.0) + A * W[i] + B * WW[i];
  // t = v + B (w x v) + C (w x (w x v))   (applyLeftJacobian, SO3.cpp:165-174)
  const double c0 = wy * vz - wz * vy, c1 = wz * vx - wx * vz, c2 = wx * vy - wy * vx;
  const double d0 = wy * c2 - wz * c1, d1 = wz * c0 - wx * c2, d2 = wx * c1 - wy * c0;
  r.t[0] = vx + B * c0 + C * d0;
  r.t[1] = vy + B * c1 + C * d1;
  r.t[2] = vz + B * c2 + C * d2;
  return r;
}

// ----------------------------------------------------------------------------------------------------
// whitening of an m x W row-major block (A | b) in registers
// Gaussian: R * A (linear/NoiseModel.cpp:217-220); Diagonal: rows scaled by 1/sigma (:335-341);
// Isotropic: everything scaled by 1/sigma (:662-674)
// ----------------------------------------------------------------------------------------------------
// Robust m-estimators of noiseModel::Robust with the default Block reweighting (linear/NoiseModel.cpp:709-737,
// linear/LossFunctions.cpp:51-90): after whitening, A and b are scaled by sqrt(w(|b|)); the factor's error is rho(|b|).
// w and rho as in linear/LossFunctions.cpp (Fair :146-155, Huber :179-192, Cauchy :217-224, Tukey :250-268,
// Welsch :289-297, GemanMcClure :320-331, DCS :354-373, L2WithDeadZone :400-412).
enum { LOSS_NONE = 0, LOSS_FAIR, LOSS_HUBER, LOSS_CAUCHY, LOSS_TUKEY, LOSS_WELSCH, LOSS_GEMANMCCLURE, LOSS_DCS, LOSS_L2DEADZONE };
__device__ __forceinline__ double loss_weight(int kind, double k, double d) {
  const double a = fabs(d);
  switch (kind) {
    case LOSS_FAIR: return 1.0 / (1.0 + a / k);
    case LOSS_HUBER: return a <= k ? 1.0 : k / a;
    case LOSS_CAUCHY: return (k * k) / (k * k + d * d);
    case LOSS_TUKEY: {
      if (a > k) return 0.0;
      const double u = 1.0 - d * d / (k * k);
      return u * u;
    }
    case LOSS_WELSCH: return exp(-(d * d) / (k * k));
    case LOSS_GEMANMCCLURE: {
      const double c2 = k * k, c4 = c2 * c2, q = c2 + d * d;
      return c4 / (q * q);
    }
    case LOSS_DCS: {
      const double e2 = d * d;
      if (e2 > k) {
        const double w = 2.0 * k / (k + e2);
        return w * w;
      }
      return 1.0;
    }
    case LOSS_L2DEADZONE: return a <= k ? 0.0 : (d > k ? (-k + d) / d : (k + d) / d);
    default: return 1.0;
  }
}
__device__ __forceinline__ double loss_rho(int kind, double k, double d) {
  const double a = fabs(d);
  switch (kind) {
    case LOSS_FAIR: {
      const double n = a / k;
      return k * k * (n - log1p(n));
    }
    case LOSS_HUBER: return a <= k ? d * d / 2 : k * (a - k / 2);
    case LOSS_CAUCHY: return k * k * log1p(d * d / (k * k)) * 0.5;
    case LOSS_TUKEY: {
      if (a > k) return k * k / 6.0;
      const double u = 1.0 - d * d / (k * k);
      return k * k * (1 - u * u * u) / 6.0;
    }
    case LOSS_WELSCH: return k * k * 0.5 * -expm1(-(d * d) / (k * k));
    case LOSS_GEMANMCCLURE: return 0.5 * (k * k * d * d) / (k * k + d * d);
    case LOSS_DCS: {
      const double e2 = d * d, e4 = e2 * e2, c2 = k * k;
      return (c2 * e2 + k * e4) / ((e2 + k) * (e2 + k));
    }
    case LOSS_L2DEADZONE: return a < k ? 0.0 : 0.5 * (k - a) * (k - a);
    default: return 0.5 * d * d;
  }
}

template <int M, int W>
__device__ __forceinline__ void whiten_rows(double* Ab, int kind, const double* nz) {
  if (kind == NK_UNIT) return;
  if (kind == NK_ISO) {
    const double s = nz[0];
#pragma unroll
    for (int i = 0; i < M * W; ++i) Ab[i] *= s;
  } else if (kind == NK_DIAG) {
#pragma unroll
    for (int r = 0; r < M; ++r) {
      const double s = nz[r];
#pragma unroll
      for (int c = 0; c < W; ++c) Ab[r * W + c] *= s;
    }
  } else {
    // full upper-triangular sqrt information; column by column so only M temporaries are live
#pragma unroll
    for (int c = 0; c < W; ++c) {
      double col[M];
#pragma unroll
      for (int r = 0; r < M; ++r) col[r] = Ab[r * W + c];
#pragma unroll
      for (int r = 0; r < M; ++r) {
        double acc = 0.0;
#pragma unroll
        for (int k = 0; k < M; ++k) acc += nz[r * M + k] * col[k];
        Ab[r * W + c] = acc;
      }
    }
  }
}

// ----------------------------------------------------------------------------------------------------
// per-bucket evaluation: fills Ab (M x W row-major, UNWHITENED, b = -(h(x) - z)) when JAC, else only the
// last column.  Returns nothing; residual sign convention as NoiseModelFactor::linearize
// (nonlinear/NonlinearFactor.cpp:161).
// ----------------------------------------------------------------------------------------------------
template <int B, bool JAC> struct Eval;

// BetweenFactor<Pose2>
template <bool JAC> struct Eval<0, JAC> {
  __device__ static void run(const double* s0, const double* s1, const double* z, double* Ab) {
    const P2 p1 = pose2_load(s0), p2 = pose2_load(s1), zm = pose2_load(z);
    const P2 hx = pose2_compose(pose2_inverse(p1), p2);
    const P2 d = pose2_compose(pose2_inverse(zm), hx);
    const int W = 7;
    Ab[0 * W + 6] = -d.x;
    Ab[1 * W + 6] = -d.y;
    Ab[2 * W + 6] = -atan2(d.s, d.c);
    if (JAC) {
      const P2 hi = pose2_inverse(hx);  // H1 = -AdjointMap(hx^-1)   (geometry/Pose2.cpp:127-135)
      Ab[0 * W + 0] = -hi.c; Ab[0 * W + 1] = hi.s;  Ab[0 * W + 2] = -hi.y;
      Ab[1 * W + 0] = -hi.s; Ab[1 * W + 1] = -hi.c; Ab[1 * W + 2] = hi.x;
      Ab[2 * W + 0] = -0.0;  Ab[2 * W + 1] = -0.0;  Ab[2 * W + 2] = -1.0;
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) Ab[r * W + 3 + c] = (r == c) ? 1.0 : 0.0;
    }
  }
};

// BetweenFactor<Pose3>
template <bool JAC> struct Eval<1, JAC> {
  __device__ static void run(const double* s0, const double* s1, const double* z, double* Ab) {
    const P3 p1 = pose3_load(s0), p2 = pose3_load(s1), zm = pose3_load(z);
    const P3 hx = pose3_compose(pose3_inverse(p1), p2);
    const P3 d = pose3_compose(pose3_inverse(zm), hx);
    double xi[6];
    pose3_logmap(d, xi);
    const int W = 13;
#pragma unroll
    for (int r = 0; r < 6; ++r) Ab[r * W + 12] = -xi[r];
    if (JAC) {
      const P3 hi = pose3_inverse(hx);  // H1 = -[R 0; skew(t) R, R]   (geometry/Pose3.cpp:69-75)
      const double tx = hi.t[0], ty = hi.t[1], tz = hi.t[2];
      const double Sk[9] = {0.0, -tz, ty, tz, 0.0, -tx, -ty, tx, 0.0};
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          const double a = Sk[3 * i] * hi.R[j] + Sk[3 * i + 1] * hi.R[3 + j] + Sk[3 * i + 2] * hi.R[6 + j];
          Ab[i * W + j] = -hi.R[3 * i + j];
          Ab[i * W + 3 + j] = -0.0;
          Ab[(3 + i) * W + j] = -a;
          Ab[(3 + i) * W + 3 + j] = -hi.R[3 * i + j];
        }
#pragma unroll
      for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int c = 0; c < 6; ++c) Ab[r * W + 6 + c] = (r == c) ? 1.0 : 0.0;
    }
  }
};

// GeneralSFMFactor<PinholeCamera<Cal3Bundler>, Point3>
template <bool JAC> struct Eval<2, JAC> {
  __device__ static void run(const double* cam, const double* pt, const double* z, double* Ab) {
    const int W = 13;
    const double* R = cam;
    const double dx = pt[0] - cam[9], dy = pt[1] - cam[10], dz = pt[2] - cam[11];
    // q = R^T (p - t)   (geometry/Pose3.cpp:380-397)
    const double qx = R[0] * dx + R[3] * dy + R[6] * dz;
    const double qy = R[1] * dx + R[4] * dy + R[7] * dz;
    const double qz = R[2] * dx + R[5] * dy + R[8] * dz;
    if (qz <= 0) {  // CheiralityException -> zero rows (slam/GeneralSFMFactor.h:153-158)
      if (JAC) {
#pragma unroll
        for (int i = 0; i < 2 * W; ++i) Ab[i] = 0.0;
      } else {
        Ab[12] = 0.0;
        Ab[W + 12] = 0.0;
      }
      return;
    }
    const double d = 1.0 / qz;
    const double x = qx * d, y = qy * d;  // pn  (geometry/CalibratedCamera.cpp:88-94)
    const double f = cam[12], k1 = cam[13], k2 = cam[14], u0 = cam[15], v0 = cam[16];
    // Cal3Bundler::uncalibrate (geometry/Cal3Bundler.cpp:64-90)
    const double r = x * x + y * y;
    const double g = 1. + (k1 + k2 * r) * r;
    const double u = g * x, v = g * y;
    Ab[12] = z[0] - (u0 + f * u);
    Ab[W + 12] = z[1] - (v0 + f * v);
    if (JAC) {
      const double rx = r * x, ry = r * y;
      const double a = 2. * (k1 + 2. * k2 * r);
      const double axx = a * x * x, axy = a * x * y, ayy = a * y * y;
      const double D00 = (g + axx) * f, D01 = axy * f, D10 = axy * f, D11 = (g + ayy) * f;
      // Dpose (geometry/CalibratedCamera.cpp:27-34)
      const double uv = x * y, uu = x * x, vv = y * y;
      const double P0[6] = {uv, -1 - uu, y, -d, 0, d * x};
      const double P1[6] = {1 + vv, -uv, -x, 0, -d, d * y};
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        Ab[c] = D00 * P0[c] + D01 * P1[c];
        Ab[W + c] = D10 * P0[c] + D11 * P1[c];
      }
      // Dcal
      Ab[6] = u; Ab[7] = f * rx; Ab[8] = f * r * rx;
      Ab[W + 6] = v; Ab[W + 7] = f * ry; Ab[W + 8] = f * r * ry;
      // Dpoint = d * (Rt rows 0,1 - (x,y) * Rt row 2)   (geometry/CalibratedCamera.cpp:37-46); Rt(i,j) = R(j,i)
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const double q0 = (R[3 * c + 0] - x * R[3 * c + 2]) * d;
        const double q1 = (R[3 * c + 1] - y * R[3 * c + 2]) * d;
        Ab[9 + c] = D00 * q0 + D01 * q1;
        Ab[W + 9 + c] = D10 * q0 + D11 * q1;
      }
    }
  }
};

// PriorFactor<Pose2>: error = -Local(x, prior), H = I, b = Local(x^-1 * prior)
template <bool JAC> struct Eval<3, JAC> {
  __device__ static void run(const double* s0, const double*, const double* z, double* Ab) {
    const P2 d = pose2_compose(pose2_inverse(pose2_load(s0)), pose2_load(z));
    const int W = 4;
    Ab[0 * W + 3] = d.x;
    Ab[1 * W + 3] = d.y;
    Ab[2 * W + 3] = atan2(d.s, d.c);
    if (JAC) {
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) Ab[r * W + c] = (r == c) ? 1.0 : 0.0;
    }
  }
};

// PriorFactor<Pose3>
template <bool JAC> struct Eval<4, JAC> {
  __device__ static void run(const double* s0, const double*, const double* z, double* Ab) {
    const P3 d = pose3_compose(pose3_inverse(pose3_load(s0)), pose3_load(z));
    double xi[6];
    pose3_logmap(d, xi);
    const int W = 7;
#pragma unroll
    for (int r = 0; r < 6; ++r) Ab[r * W + 6] = xi[r];
    if (JAC) {
#pragma unroll
      for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int c = 0; c < 6; ++c) Ab[r * W + c] = (r == c) ? 1.0 : 0.0;
    }
  }
};

// PriorFactor<PinholeCamera<Cal3Bundler>>: localCoordinates (geometry/PinholeCamera.h:206-211)
template <bool JAC> struct Eval<5, JAC> {
  __device__ static void run(const double* s0, const double*, const double* z, double* Ab) {
    const P3 d = pose3_compose(pose3_inverse(pose3_load(s0)), pose3_load(z));
    double xi[6];
    pose3_logmap(d, xi);
    const int W = 10;
#pragma unroll
    for (int r = 0; r < 6; ++r) Ab[r * W + 9] = xi[r];
#pragma unroll
    for (int r = 0; r < 3; ++r) Ab[(6 + r) * W + 9] = z[12 + r] - s0[12 + r];  // Cal3Bundler.h:139-141
    if (JAC) {
#pragma unroll
      for (int r = 0; r < 9; ++r)
#pragma unroll
        for (int c = 0; c < 9; ++c) Ab[r * W + c] = (r == c) ? 1.0 : 0.0;
    }
  }
};

// PriorFactor<Point3>
template <bool JAC> struct Eval<6, JAC> {
  __device__ static void run(const double* s0, const double*, const double* z, double* Ab) {
    const int W = 4;
#pragma unroll
    for (int r = 0; r < 3; ++r) Ab[r * W + 3] = z[r] - s0[r];
    if (JAC) {
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) Ab[r * W + c] = (r == c) ? 1.0 : 0.0;
    }
  }
};

// ----------------------------------------------------------------------------------------------------
// kernels
// ----------------------------------------------------------------------------------------------------
constexpr int kFactorThreads = 128;

// Linearize one bucket.  Each thread evaluates one factor in registers, whitens, and the CTA stages its
// kFactorThreads x (M*W) block through shared memory so the [A|b] rows leave as fully coalesced stores.
template <int B>
__global__ void __launch_bounds__(kFactorThreads)
linearize_kernel(BucketDev bk, const double* __restrict__ st0, const double* __restrict__ st1) {
  typedef BucketTraits<B> T;
  constexpr int MW = T::M * T::W;
  constexpr int LDS = (MW % 2 == 0) ? MW + 1 : MW;  // odd stride: conflict-free 64-bit shared stores
  extern __shared__ double stage[];
  const int base = blockIdx.x * kFactorThreads;
  const int i = base + threadIdx.x;
  if (i < bk.n) {
    double Ab[MW];
    const double* s0 = st0 + (int64_t)bk.row0[i] * state_size(T::T0);
    const double* s1 = (T::T1 >= 0) ? st1 + (int64_t)bk.row1[i] * state_size(T::T1 >= 0 ? T::T1 : 0) : nullptr;
    Eval<B, true>::run(s0, s1, bk.z + (int64_t)i * T::ZD, Ab);
    if (bk.any_noise) whiten_rows<T::M, T::W>(Ab, bk.nk[i], bk.noise + (int64_t)i * T::M * T::M);
    if (bk.loss) {
      double n2 = 0.0;
#pragma unroll
      for (int r = 0; r < T::M; ++r) n2 += Ab[r * T::W + T::W - 1] * Ab[r * T::W + T::W - 1];
      // GeneralSFMFactor::linearize whitens H1, H2 and b one by one through Whiten(Matrix) (slam/GeneralSFMFactor.h:161-167),
      // and Robust::Whiten(Matrix) reweights with an EMPTY error vector (linear/NoiseModel.h:713-714), i.e. with
      // sqrt(w(0)) whatever the residual is; only the factor's error() sees rho.  Reproduced as is: parity, not a judgement.
      const double w = sqrt(loss_weight(bk.loss, bk.loss_k, B == 2 ? 0.0 : sqrt(n2)));
#pragma unroll
      for (int k = 0; k < MW; ++k) Ab[k] *= w;
    }
#pragma unroll
    for (int k = 0; k < MW; ++k) stage[threadIdx.x * LDS + k] = Ab[k];
  }
  __syncthreads();
  const int nvalid = min(kFactorThreads, bk.n - base);
  double* out = bk.Ab + (int64_t)base * MW;
  for (int k = threadIdx.x; k < nvalid * MW; k += kFactorThreads) out[k] = stage[(k / MW) * LDS + (k % MW)];
}

// Deterministic block reduction of one double per thread (fixed tree); result valid in thread 0.
__device__ __forceinline__ double block_reduce_sum(double v, double* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) red[w] = v;
  __syncthreads();
  if (w == 0) {
    v = (l < (int)(blockDim.x >> 5)) ? red[l] : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  }
  __syncthreads();
  return v;
}

// NonlinearFactorGraph::error: 0.5 * || whiten(h(x) - z) ||^2 per factor, one partial per CTA.
template <int B>
__global__ void __launch_bounds__(kFactorThreads)
error_kernel(BucketDev bk, const double* __restrict__ st0, const double* __restrict__ st1,
             double* __restrict__ partials) {
  typedef BucketTraits<B> T;
  __shared__ double red[kFactorThreads / 32];
  const int i = blockIdx.x * kFactorThreads + threadIdx.x;
  double e = 0.0;
  if (i < bk.n) {
    double Ab[T::M * T::W];
    const double* s0 = st0 + (int64_t)bk.row0[i] * state_size(T::T0);
    const double* s1 = (T::T1 >= 0) ? st1 + (int64_t)bk.row1[i] * state_size(T::T1 >= 0 ? T::T1 : 0) : nullptr;
    Eval<B, false>::run(s0, s1, bk.z + (int64_t)i * T::ZD, Ab);
    double r[T::M];
#pragma unroll
    for (int k = 0; k < T::M; ++k) r[k] = Ab[k * T::W + T::W - 1];
    const int kind = bk.any_noise ? bk.nk[i] : NK_UNIT;
    if (kind != NK_UNIT) whiten_rows<T::M, 1>(r, kind, bk.noise + (int64_t)i * T::M * T::M);
#pragma unroll
    for (int k = 0; k < T::M; ++k) e += r[k] * r[k];
    e = bk.loss ? loss_rho(bk.loss, bk.loss_k, sqrt(e)) : 0.5 * e;
  }
  e = block_reduce_sum(e, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = e;
}

// GaussianFactorGraph::error at zero and at delta: 0.5 ||b||^2 and 0.5 ||A delta - b||^2
// (linear/JacobianFactor.cpp:509-514), two partials per CTA.
template <int B>
__global__ void __launch_bounds__(kFactorThreads)
linear_error_kernel(BucketDev bk, const double* __restrict__ delta, double* __restrict__ partials0,
                    double* __restrict__ partials1) {
  typedef BucketTraits<B> T;
  __shared__ double red[kFactorThreads / 32];
  const int i = blockIdx.x * kFactorThreads + threadIdx.x;
  double e0 = 0.0, e1 = 0.0;
  if (i < bk.n) {
    const double* Ab = bk.Ab + (int64_t)i * T::M * T::W;
    double d0[T::N0], d1[T::N1 > 0 ? T::N1 : 1];
    const double* p0 = delta + bk.goff0[i];
#pragma unroll
    for (int c = 0; c < T::N0; ++c) d0[c] = p0[c];
    if (T::N1 > 0) {
      const double* p1 = delta + bk.goff1[i];
#pragma unroll
      for (int c = 0; c < T::N1; ++c) d1[c] = p1[c];
    }
#pragma unroll
    for (int r = 0; r < T::M; ++r) {
      const double* row = Ab + r * T::W;
      double acc = 0.0;
#pragma unroll
      for (int c = 0; c < T::N0; ++c) acc += row[c] * d0[c];
#pragma unroll
      for (int c = 0; c < T::N1; ++c) acc += row[T::N0 + c] * d1[c];
      const double b = row[T::W - 1];
      e0 += b * b;
      e1 += (acc - b) * (acc - b);
    }
    e0 *= 0.5;
    e1 *= 0.5;
  }
  e0 = block_reduce_sum(e0, red);
  e1 = block_reduce_sum(e1, red);
  if (threadIdx.x == 0) { partials0[blockIdx.x] = e0; partials1[blockIdx.x] = e1; }
}

// Final fixed-order sum of the per-CTA partials (single CTA).
__global__ void __launch_bounds__(256) sum_partials_kernel(const double* __restrict__ p, int n, double* out) {
  __shared__ double red[8];
  double v = 0.0;
  for (int i = threadIdx.x; i < n; i += 256) v += p[i];
  v = block_reduce_sum(v, red);
  if (threadIdx.x == 0) *out = v;
}

// failure flag of a factorisation (min failing clique id, INT_MAX = none) as a double next to the error scalars
__global__ void fail_flag_kernel(const int32_t* __restrict__ fail, double* __restrict__ out) { *out = (double)*fail; }

// Hessian diagonal (linear/JacobianFactor.cpp:539-564): one warp per variable gathers the squared column norms
// of every incident factor block in graph order (deterministic, no atomics).
struct IncidenceDev {
  const int64_t* off;      // [n_vars + 1]
  const int32_t* bucket;   // per incidence
  const int32_t* index;    // factor index in its bucket
  const int32_t* slot;     // 0 / 1
  const int32_t* var_goff; // global scalar offset per variable
  const int32_t* var_dim;
  int32_t n_vars;
};
struct AbTable { const double* Ab[7]; };

__global__ void __launch_bounds__(128)
hessian_diag_kernel(IncidenceDev inc, AbTable tab, double* __restrict__ hdiag) {
  const int v = blockIdx.x * 4 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (v >= inc.n_vars) return;
  const int dim = inc.var_dim[v];
  double acc[9];
#pragma unroll
  for (int c = 0; c < 9; ++c) acc[c] = 0.0;
  const int Ms[7] = {3, 6, 2, 3, 6, 9, 3};
  const int Ws[7] = {7, 13, 13, 4, 7, 10, 4};
  const int N0s[7] = {3, 6, 9, 3, 6, 9, 3};
  for (int64_t q = inc.off[v] + lane; q < inc.off[v + 1]; q += 32) {
    const int b = inc.bucket[q];
    const int M = Ms[b], W = Ws[b];
    const double* Ab = tab.Ab[b] + (int64_t)inc.index[q] * M * W + (inc.slot[q] ? N0s[b] : 0);
    for (int r = 0; r < M; ++r)
#pragma unroll
      for (int c = 0; c < 9; ++c)
        if (c < dim) { const double a = Ab[r * W + c]; acc[c] += a * a; }
  }
#pragma unroll
  for (int c = 0; c < 9; ++c)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[c] += __shfl_down_sync(0xffffffffu, acc[c], o);
  if (lane == 0)
    for (int c = 0; c < dim; ++c) hdiag[inc.var_goff[v] + c] = acc[c];
}

// damping vector: identity mode d = s*s ; diagonal mode d = (s * sqrt(clamp(h)))^2 with s = 1/(1/sqrt(lambda))
// (nonlinear/internal/LevenbergMarquardtState.h:98-156, LevenbergMarquardtOptimizer.cpp:293-299)
__global__ void damping_kernel(int n, double s, int diag_mode, double lo, double hi, const double* __restrict__ hdiag,
                               double* __restrict__ damp) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (!diag_mode) {
    damp[i] = s * s;
  } else {
    const double a = s * sqrt(fmin(fmax(hdiag[i], lo), hi));
    damp[i] = a * a;
  }
}

// Several ranks: keep the entries of delta this rank solved for and clear the rest, so that a sum over ranks rebuilds
// the whole vector exactly (every entry has one owner).
__global__ void __launch_bounds__(256)
own_delta_kernel(int n, const double* __restrict__ delta, const uint8_t* __restrict__ own, double* __restrict__ out,
                 const double* __restrict__ fail_flag) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = own[i] ? delta[i] : 0.0;
  else if (i == n) out[n] = (*fail_flag < 2147483647.0) ? 1.0 : 0.0;   // this rank's factorisation failed (INT_MAX = no failure)
}

// Values::retract (nonlinear/Values.cpp:53-63): one thread per variable of one type.
template <int VT>
__global__ void __launch_bounds__(128)
retract_kernel(int n, const double* __restrict__ cur, const int32_t* __restrict__ goff,
               const double* __restrict__ delta, double* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double* s = cur + (int64_t)i * state_size(VT);
  double* o = out + (int64_t)i * state_size(VT);
  const double* d = delta + goff[i];
  if (VT == 0) {
    // p * Pose2(v0, v1, v2)   (geometry/Pose2.cpp:100-110, base/Lie.h:131-133)
    P2 v;
    v.c = cos(d[2]); v.s = sin(d[2]); v.x = d[0]; v.y = d[1];
    const P2 r = pose2_compose(pose2_load(s), v);
    o[0] = r.c; o[1] = r.s; o[2] = r.x; o[3] = r.y;
  } else if (VT == 1 || VT == 2) {
    double xi[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) xi[k] = d[k];
    const P3 r = pose3_compose(pose3_load(s), pose3_expmap(xi));
#pragma unroll
    for (int k = 0; k < 9; ++k) o[k] = r.R[k];
#pragma unroll
    for (int k = 0; k < 3; ++k) o[9 + k] = r.t[k];
    if (VT == 2) {  // Cal3Bundler::retract (geometry/Cal3Bundler.h:134-136)
      o[12] = s[12] + d[6]; o[13] = s[13] + d[7]; o[14] = s[14] + d[8];
      o[15] = s[15]; o[16] = s[16];
    }
  } else {
#pragma unroll
    for (int k = 0; k < 3; ++k) o[k] = s[k] + d[k];
  }
}

// ISAM2's fluid relinearisation (nonlinear/ISAM2.cpp:456-468, ISAM2-impl.h:341-351 CheckRelinearizationFull): a variable
// whose |delta|_inf reaches the threshold moves its linearisation point, theta_j := theta_j (+) delta_j
// (Values::retractMasked).  `retracted` already holds theta (+) delta of every variable (retract_kernel), so the marked rows
// are copied.  One thread per variable of one type.
template <int VT>
__global__ void __launch_bounds__(128)
relinearize_marked_kernel(int n, const int32_t* __restrict__ goff, const double* __restrict__ delta, double threshold,
                          const double* __restrict__ retracted, double* __restrict__ cur, uint8_t* __restrict__ marks,
                          int* __restrict__ count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  constexpr int dim = VT == 0 ? 3 : VT == 1 ? 6 : VT == 2 ? 9 : 3;
  const double* d = delta + goff[i];
  double m = 0.0;
#pragma unroll
  for (int k = 0; k < dim; ++k) m = fmax(m, fabs(d[k]));
  const bool mark = m >= threshold;
  if (marks) marks[i] = mark ? 1 : 0;
  if (!mark) return;
#pragma unroll
  for (int k = 0; k < state_size(VT); ++k) cur[(int64_t)i * state_size(VT) + k] = retracted[(int64_t)i * state_size(VT) + k];
  atomicAdd(count, 1);
}

}  // namespace gtb
