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

// Residual + Jacobian.  jc[12] = {du/dc_0..5, dv/dc_0..5} (c = w0,w1,w2,t0,t1,t2), jp[6] =
// {du/dX_0..2, dv/dX_0..2}; camera columns are multiplied by C.sc, point columns by sp[3].
BA_HD void reproject_rj(const CamRow& C, double X0, double X1, double X2, double ox, double oy, const double* sp,
                        double& r0, double& r1, double* jc, double* jp) {
  // dP[k][j] = d p_j / d param_k for the 9 parameters (w0,w1,w2,t0,t1,t2,X0,X1,X2); t columns are e_k.
  double p0, p1, p2;
  double dw[3][3];  // d p / d w_k
  double R[3][3];   // d p / d X  (rotation matrix, R[j][m] = d p_j / d X_m)
  const double u0 = C.u0, u1 = C.u1, u2 = C.u2;
  if (C.ti != 0.0) {
    const double c = C.c, s = C.s, ti = C.ti;
    const double omc = 1.0 - c;
    const double x0 = u1 * X2 - u2 * X1, x1 = u2 * X0 - u0 * X2, x2 = u0 * X1 - u1 * X0;
    const double d = u0 * X0 + u1 * X1 + u2 * X2;
    const double tmp = d * omc;
    p0 = X0 * c + x0 * s + u0 * tmp;
    p1 = X1 * c + x1 * s + u1 * tmp;
    p2 = X2 * c + x2 * s + u2 * tmp;
    const double uu[3] = {u0, u1, u2};
    const double X[3] = {X0, X1, X2};
    const double x[3] = {x0, x1, x2};
    // e_k x X
    const double eX[3][3] = {{0.0, -X2, X1}, {X2, 0.0, -X0}, {-X1, X0, 0.0}};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double uk = uu[k];
      // du_j/dw_k = ti (delta_jk - u_j u_k);  d(u x X)/dw_k = ti (e_k x X - u_k x);  d(u.X)/dw_k = ti (X_k - u_k d)
      const double dc = -s * uk, ds = c * uk;
      const double dd = ti * (X[k] - uk * d);
      const double dtmp = dd * omc + d * (s * uk);
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const double du_j = ti * ((j == k ? 1.0 : 0.0) - uu[j] * uk);
        const double dx_j = ti * (eX[k][j] - uk * x[j]);
        dw[k][j] = X[j] * dc + dx_j * s + x[j] * ds + du_j * tmp + uu[j] * dtmp;
      }
    }
    // R = c I + s [u]x + (1-c) u u^T
    R[0][0] = c + omc * u0 * u0;      R[0][1] = -s * u2 + omc * u0 * u1; R[0][2] = s * u1 + omc * u0 * u2;
    R[1][0] = s * u2 + omc * u1 * u0; R[1][1] = c + omc * u1 * u1;       R[1][2] = -s * u0 + omc * u1 * u2;
    R[2][0] = -s * u1 + omc * u2 * u0; R[2][1] = s * u0 + omc * u2 * u1; R[2][2] = c + omc * u2 * u2;
  } else {
    // p = X + w x X   (w stored in u)
    p0 = X0 + (u1 * X2 - u2 * X1);
    p1 = X1 + (u2 * X0 - u0 * X2);
    p2 = X2 + (u0 * X1 - u1 * X0);
    // d/dw_k = e_k x X
    dw[0][0] = 0.0; dw[0][1] = -X2; dw[0][2] = X1;
    dw[1][0] = X2;  dw[1][1] = 0.0; dw[1][2] = -X0;
    dw[2][0] = -X1; dw[2][1] = X0;  dw[2][2] = 0.0;
    // d/dX = I + [w]x
    R[0][0] = 1.0; R[0][1] = -u2; R[0][2] = u1;
    R[1][0] = u2;  R[1][1] = 1.0; R[1][2] = -u0;
    R[2][0] = -u1; R[2][1] = u0;  R[2][2] = 1.0;
  }
  p0 += C.t0; p1 += C.t1; p2 += C.t2;
  const double iz = 1.0 / p2;
  const double xp = p0 * iz, yp = p1 * iz;
  const double r2 = xp * xp + yp * yp;
  const double f = 1.0 + C.k1 * r2 + C.k2 * (r2 * r2);
  const double xd = xp * f, yd = yp * f;
  r0 = (C.fx * xd + C.sk * yd + C.cx) - ox;
  r1 = (C.fy * yd + C.cy) - oy;
  const double dfdr2 = C.k1 + 2.0 * C.k2 * r2;
  // one column: given dP = (a,b,cz) produce (du, dv)
  auto column = [&](double a, double b, double cz, double scale, double& du, double& dv) {
    const double dxp = (a - xp * cz) * iz;
    const double dyp = (b - yp * cz) * iz;
    const double dr2 = 2.0 * (xp * dxp + yp * dyp);
    const double df = dfdr2 * dr2;
    const double dxd = dxp * f + xp * df;
    const double dyd = dyp * f + yp * df;
    du = (C.fx * dxd + C.sk * dyd) * scale;
    dv = (C.fy * dyd) * scale;
  };
#pragma unroll
  for (int k = 0; k < 3; ++k) column(dw[k][0], dw[k][1], dw[k][2], C.sc[k], jc[k], jc[6 + k]);
  column(1.0, 0.0, 0.0, C.sc[3], jc[3], jc[9]);
  column(0.0, 1.0, 0.0, C.sc[4], jc[4], jc[10]);
  column(0.0, 0.0, 1.0, C.sc[5], jc[5], jc[11]);
#pragma unroll
  for (int m = 0; m < 3; ++m) column(R[0][m], R[1][m], R[2][m], sp[m], jp[m], jp[3 + m]);
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
