// ba_math.cuh — per-observation reprojection residual + analytic Jacobian, and the small dense
// block helpers.  Everything is __host__ __device__ so tests can exercise the exact same
// arithmetic on the CPU (tests/test_math_host.py builds a tiny g++ harness around this header);
// the product only ever calls it from kernels.
//
// Follows the reference functor statement by statement (CTracker.cpp:585-604: rotate, translate,
// divide by z, apply k0,k1,k2,k4,k5, subtract the observation) and differentiates it by the
// chain rule over the SAME intermediates and the SAME theta^2 > epsilon branch that
// ceres::AngleAxisRotatePoint takes under autodiff, so the result
// matches AutoDiffCostFunction<…,2,3,3,3> (CTracker.h:108-110) to rounding.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define BA_HD __host__ __device__ __forceinline__
#else
#define BA_HD inline
#endif

namespace hsfm {

// Per-camera row prepared once per evaluation point by cam_prep_kernel (so that no observation
// pays for sqrt/sincos).  24 doubles = 192 B, 16-byte aligned for double2 loads.
struct CamRow {
  double u0, u1, u2;  // branch A: w/theta ; Taylor branch (theta^2 <= eps): w itself
  double ti;          // 1/theta, or 0 to select the Taylor branch
  double c, s;        // cos(theta), sin(theta)
  double t0, t1, t2;  // translation
  double fx, sk, cx, fy, cy;  // K[0],K[1],K[2],K[4],K[5]  (CTracker.cpp:597-598)
  double k1, k2;      // radial distortion extension (0,0 = reference model)
  double sc[6];       // Jacobi column scaling of this camera's 6 columns (1 = unscaled)
  double pad0, pad1;
};
static_assert(sizeof(CamRow) == 192, "CamRow must be 24 doubles");

// Build a CamRow from (w,t), intrinsics K9 (row-major 3x3), distortion and scaling.
BA_HD void cam_prepare(const double* wt, const double* K9, const double* dist, const double* scale6, CamRow& C) {
  const double w0 = wt[0], w1 = wt[1], w2 = wt[2];
  const double theta2 = w0 * w0 + w1 * w1 + w2 * w2;
  if (theta2 > 2.220446049250313e-16 /* std::numeric_limits<double>::epsilon() */) {
    const double theta = sqrt(theta2);
    C.c = cos(theta);
    C.s = sin(theta);
    const double ti = 1.0 / theta;
    C.ti = ti;
    C.u0 = w0 * ti; C.u1 = w1 * ti; C.u2 = w2 * ti;
  } else {
    C.c = 1.0; C.s = 0.0; C.ti = 0.0;
    C.u0 = w0; C.u1 = w1; C.u2 = w2;
  }
  C.t0 = wt[3]; C.t1 = wt[4]; C.t2 = wt[5];
  C.fx = K9[0]; C.sk = K9[1]; C.cx = K9[2]; C.fy = K9[4]; C.cy = K9[5];
  C.k1 = dist ? dist[0] : 0.0;
  C.k2 = dist ? dist[1] : 0.0;
  for (int k = 0; k < 6; ++k) C.sc[k] = scale6 ? scale6[k] : 1.0;
  C.pad0 = 0.0; C.pad1 = 0.0;
}

// Residual only (candidate-cost evaluation).
BA_HD void reproject_r(const CamRow& C, double X0, double X1, double X2, double ox, double oy, double& r0, double& r1) {
  double p0, p1, p2;
  if (C.ti != 0.0) {
    const double x0 = C.u1 * X2 - C.u2 * X1, x1 = C.u2 * X0 - C.u0 * X2, x2 = C.u0 * X1 - C.u1 * X0;
    const double tmp = (C.u0 * X0 + C.u1 * X1 + C.u2 * X2) * (1.0 - C.c);
    p0 = X0 * C.c + x0 * C.s + C.u0 * tmp;
    p1 = X1 * C.c + x1 * C.s + C.u1 * tmp;
    p2 = X2 * C.c + x2 * C.s + C.u2 * tmp;
  } else {
    p0 = X0 + (C.u1 * X2 - C.u2 * X1);
    p1 = X1 + (C.u2 * X0 - C.u0 * X2);
    p2 = X2 + (C.u0 * X1 - C.u1 * X0);
  }
  p0 += C.t0; p1 += C.t1; p2 += C.t2;
  const double iz = 1.0 / p2;
  double xp = p0 * iz, yp = p1 * iz;
  const double r2 = xp * xp + yp * yp;
  const double f = 1.0 + C.k1 * r2 + C.k2 * (r2 * r2);
  xp *= f; yp *= f;
  r0 = (C.fx * xp + C.sk * yp + C.cx) - ox;
  r1 = (C.fy * yp + C.cy) - oy;
}

// Residual + Jacobian, column by column.  emit(col, du, dv) receives the nine Jacobian columns
// (col 0..5 = w0,w1,w2,t0,t1,t2 scaled by C.sc; col 6..8 = X0,X1,X2 scaled by sp).
//
// The chain rule is grouped so that one observation costs ~230 FP64 instructions with short
// dependency chains:
//   * everything after the rigid transform is linear in dp = (a,b,cz): (du,dv) = M dp with one 2x3
//     matrix M per observation (perspective divide, radial factor, intrinsics), so a column is six
//     multiply-adds;
//   * d p / d w_k = alpha_k X + beta_k (u x X) + gamma_k u + (s/theta) e_k x X + (tmp/theta) e_k — the
//     products of the statement-by-statement derivative of ceres::AngleAxisRotatePoint regrouped by
//     the vector they multiply.  The same intermediates (1-cos, u.X, ...) enter with the same values,
//     so the ill-conditioned small-angle terms round the way the autodiff jets do.
template <class Emit>
BA_HD void reproject_lin(const CamRow& C, double X0, double X1, double X2, double ox, double oy, const double* sp, double& r0,
                         double& r1, Emit&& emit) {
  const double u0 = C.u0, u1 = C.u1, u2 = C.u2;
  const bool full = C.ti != 0.0;
  const double c = C.c, s = C.s, ti = C.ti;
  const double omc = 1.0 - c;
  // x = u x X (w x X on the Taylor branch), d = u.X
  const double x0 = u1 * X2 - u2 * X1, x1 = u2 * X0 - u0 * X2, x2 = u0 * X1 - u1 * X0;
  const double d = u0 * X0 + u1 * X1 + u2 * X2;
  const double tmp = d * omc;
  double p0, p1, p2;
  if (full) {
    p0 = X0 * c + x0 * s + u0 * tmp;
    p1 = X1 * c + x1 * s + u1 * tmp;
    p2 = X2 * c + x2 * s + u2 * tmp;
  } else {
    p0 = X0 + x0; p1 = X1 + x1; p2 = X2 + x2;
  }
  p0 += C.t0; p1 += C.t1; p2 += C.t2;
  const double iz = 1.0 / p2;
  const double xp = p0 * iz, yp = p1 * iz;
  const double r2 = xp * xp + yp * yp;
  const double f = 1.0 + C.k1 * r2 + C.k2 * (r2 * r2);
  const double xd = xp * f, yd = yp * f;
  r0 = (C.fx * xd + C.sk * yd + C.cx) - ox;
  r1 = (C.fy * yd + C.cy) - oy;
  // M: d(xd,yd)/d(a,b,cz), then the intrinsics
  const double g = 2.0 * (C.k1 + 2.0 * C.k2 * r2) * iz;  // d f / d(a,b,cz) = g * (xp, yp, -r2)
  const double h = iz * f;
  const double xg = xp * g, yg = yp * g;
  const double q = -(g * r2 + h);
  const double A = h + xp * xg, B = xp * yg, Cc = xp * q;
  const double E = h + yp * yg, F = yp * q;
  const double m00 = C.fx * A + C.sk * B, m01 = C.fx * B + C.sk * E, m02 = C.fx * Cc + C.sk * F;
  const double m10 = C.fy * B, m11 = C.fy * E, m12 = C.fy * F;
  // translation columns: dp = e_k
  emit(3, m00 * C.sc[3], m10 * C.sc[3]);
  emit(4, m01 * C.sc[4], m11 * C.sc[4]);
  emit(5, m02 * C.sc[5], m12 * C.sc[5]);
  if (full) {
    // point columns: dp = R e_m,  R = c I + s [u]x + (1-c) u u^T
    const double o0 = omc * u0, o1 = omc * u1, o2 = omc * u2;
    const double su0 = s * u0, su1 = s * u1, su2 = s * u2;
    {
      const double a = c + o0 * u0, b = su2 + o1 * u0, cz = o2 * u0 - su1;
      emit(6, (m00 * a + m01 * b + m02 * cz) * sp[0], (m10 * a + m11 * b + m12 * cz) * sp[0]);
    }
    {
      const double a = o0 * u1 - su2, b = c + o1 * u1, cz = su0 + o2 * u1;
      emit(7, (m00 * a + m01 * b + m02 * cz) * sp[1], (m10 * a + m11 * b + m12 * cz) * sp[1]);
    }
    {
      const double a = su1 + o0 * u2, b = o1 * u2 - su0, cz = c + o2 * u2;
      emit(8, (m00 * a + m01 * b + m02 * cz) * sp[2], (m10 * a + m11 * b + m12 * cz) * sp[2]);
    }
    // rotation columns
    const double sti = s * ti, cms = c - sti, tt = ti * tmp;
    const double X[3] = {X0, X1, X2};
    const double uu[3] = {u0, u1, u2};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double uk = uu[k];
      const double alpha = -s * uk, beta = uk * cms;
      const double dd = ti * (X[k] - uk * d);
      const double dtmp = dd * omc + d * (s * uk);
      const double gamma = dtmp - tt * uk;
      double a = alpha * X0 + beta * x0 + gamma * u0;
      double b = alpha * X1 + beta * x1 + gamma * u1;
      double cz = alpha * X2 + beta * x2 + gamma * u2;
      // + sti * (e_k x X) + tt * e_k
      if (k == 0) { a += tt;  b -= sti * X2; cz += sti * X1; }
      if (k == 1) { a += sti * X2; b += tt;  cz -= sti * X0; }
      if (k == 2) { a -= sti * X1; b += sti * X0; cz += tt; }
      emit(k, (m00 * a + m01 * b + m02 * cz) * C.sc[k], (m10 * a + m11 * b + m12 * cz) * C.sc[k]);
    }
  } else {
    // p = X + w x X:  d/dX = I + [w]x,  d/dw_k = e_k x X
    emit(6, (m00 + m01 * u2 - m02 * u1) * sp[0], (m10 + m11 * u2 - m12 * u1) * sp[0]);
    emit(7, (m01 - m00 * u2 + m02 * u0) * sp[1], (m11 - m10 * u2 + m12 * u0) * sp[1]);
    emit(8, (m02 + m00 * u1 - m01 * u0) * sp[2], (m12 + m10 * u1 - m11 * u0) * sp[2]);
    emit(0, (m02 * X1 - m01 * X2) * C.sc[0], (m12 * X1 - m11 * X2) * C.sc[0]);
    emit(1, (m00 * X2 - m02 * X0) * C.sc[1], (m10 * X2 - m12 * X0) * C.sc[1]);
    emit(2, (m01 * X0 - m00 * X1) * C.sc[2], (m11 * X0 - m10 * X1) * C.sc[2]);
  }
}

// Residual + Jacobian into arrays.  jc[12] = {du/dc_0..5, dv/dc_0..5} (c = w0,w1,w2,t0,t1,t2), jp[6] =
// {du/dX_0..2, dv/dX_0..2}; camera columns are multiplied by C.sc, point columns by sp[3].
BA_HD void reproject_rj(const CamRow& C, double X0, double X1, double X2, double ox, double oy, const double* sp,
                        double& r0, double& r1, double* jc, double* jp) {
  reproject_lin(C, X0, X1, X2, ox, oy, sp, r0, r1, [&](int col, double du, double dv) {
    if (col < 6) { jc[col] = du; jc[6 + col] = dv; }
    else { jp[col - 6] = du; jp[col - 3] = dv; }
  });
}

// Symmetric 3x3 stored as 6 (00,01,02,11,12,22).  SPD inverse through Cholesky (as Ceres'
// InvertPSDMatrix does with Eigen LLT).  Returns false when not positive definite.
BA_HD bool inv3_spd(const double* H6, double* Hi6) {
  const double a00 = H6[0];
  if (!(a00 > 0.0)) return false;
  const double l00 = sqrt(a00);
  const double l10 = H6[1] / l00, l20 = H6[2] / l00;
  const double a11 = H6[3] - l10 * l10;
  if (!(a11 > 0.0)) return false;
  const double l11 = sqrt(a11);
  const double l21 = (H6[4] - l20 * l10) / l11;
  const double a22 = H6[5] - l20 * l20 - l21 * l21;
  if (!(a22 > 0.0)) return false;
  const double l22 = sqrt(a22);
  const double i00 = 1.0 / l00, i11 = 1.0 / l11, i22 = 1.0 / l22;
  const double i10 = -l10 * i00 * i11;
  const double i21 = -l21 * i11 * i22;
  const double i20 = -(l20 * i00 + l21 * i10) * i22;
  Hi6[0] = i00 * i00 + i10 * i10 + i20 * i20;
  Hi6[1] = i10 * i11 + i20 * i21;
  Hi6[2] = i20 * i22;
  Hi6[3] = i11 * i11 + i21 * i21;
  Hi6[4] = i21 * i22;
  Hi6[5] = i22 * i22;
  return true;
}

// y = H6 (symmetric 3x3) * v
BA_HD void sym3_mul(const double* H6, double v0, double v1, double v2, double& y0, double& y1, double& y2) {
  y0 = H6[0] * v0 + H6[1] * v1 + H6[2] * v2;
  y1 = H6[1] * v0 + H6[3] * v1 + H6[4] * v2;
  y2 = H6[2] * v0 + H6[4] * v1 + H6[5] * v2;
}

// In-place SPD inverse of a full row-major 6x6 (Cholesky, forward/back solves on identity).
BA_HD bool inv6_spd(double* A /*36 in, 36 out*/) {
  double L[36];
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    double d = A[j * 6 + j];
    for (int k = 0; k < j; ++k) d -= L[j * 6 + k] * L[j * 6 + k];
    if (!(d > 0.0)) return false;
    d = sqrt(d);
    L[j * 6 + j] = d;
    const double inv = 1.0 / d;
    for (int i = j + 1; i < 6; ++i) {
      double s = A[i * 6 + j];
      for (int k = 0; k < j; ++k) s -= L[i * 6 + k] * L[j * 6 + k];
      L[i * 6 + j] = s * inv;
    }
  }
  // Li = L^-1 (lower)
  double Li[36];
#pragma unroll
  for (int i = 0; i < 36; ++i) Li[i] = 0.0;
#pragma unroll
  for (int c = 0; c < 6; ++c) {
    Li[c * 6 + c] = 1.0 / L[c * 6 + c];
    for (int i = c + 1; i < 6; ++i) {
      double s = 0.0;
      for (int k = c; k < i; ++k) s -= L[i * 6 + k] * Li[k * 6 + c];
      Li[i * 6 + c] = s / L[i * 6 + i];
    }
  }
  // A^-1 = Li^T Li
#pragma unroll
  for (int a = 0; a < 6; ++a)
    for (int b = 0; b <= a; ++b) {
      double s = 0.0;
      for (int k = a; k < 6; ++k) s += Li[k * 6 + a] * Li[k * 6 + b];
      A[a * 6 + b] = s;
      A[b * 6 + a] = s;
    }
  return true;
}

}  // namespace hsfm
