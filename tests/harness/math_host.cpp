// CPU harness around sfm_b200/csrc/ba_math.cuh (the product's __host__ __device__ arithmetic),
// built by tests/test_math_host.py with g++.  Lets the "not gpu" suite compare the analytic
// Jacobian against the dual-number oracle without a GPU.  Test infrastructure only.
#include <cstdint>
#include "../../sfm_b200/csrc/ba_math.cuh"
using namespace hsfm;
extern "C" int math_host_eval(int64_t n_obs, const double* obs, const int32_t* cam_idx, const int32_t* pt_idx,
                              int32_t n_cams, const double* K9, const double* dist, const double* cam, const double* pts,
                              const double* scale_c, const double* scale_p, double* r, double* J, double* r_only) {
  for (int64_t i = 0; i < n_obs; ++i) {
    const int c = cam_idx[i];
    const int64_t p = pt_idx[i];
    CamRow C;
    cam_prepare(cam + 6 * c, K9 + 9 * c, dist ? dist + 2 * c : nullptr, scale_c ? scale_c + 6 * c : nullptr, C);
    double one[3] = {1.0, 1.0, 1.0};
    double jc[12], jp[6], r0, r1;
    reproject_rj(C, pts[3 * p], pts[3 * p + 1], pts[3 * p + 2], obs[2 * i], obs[2 * i + 1], scale_p ? scale_p + 3 * p : one, r0, r1, jc, jp);
    r[2 * i] = r0; r[2 * i + 1] = r1;
    for (int row = 0; row < 2; ++row) {
      for (int j = 0; j < 6; ++j) J[18 * i + 9 * row + j] = jc[6 * row + j];
      for (int j = 0; j < 3; ++j) J[18 * i + 9 * row + 6 + j] = jp[3 * row + j];
    }
    reproject_r(C, pts[3 * p], pts[3 * p + 1], pts[3 * p + 2], obs[2 * i], obs[2 * i + 1], r_only[2 * i], r_only[2 * i + 1]);
  }
  return 0;
}
extern "C" int math_host_inv(const double* H6, double* Hi6, double* A36) {
  bool a = inv3_spd(H6, Hi6);
  bool b = inv6_spd(A36);
  return (a ? 1 : 0) | (b ? 2 : 0);
}
