// CTrackerBA.cpp — see CTrackerBA.h.  Replaces CTracker.cpp:564-702 of the reference.
#include "CTrackerBA.h"

#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>

static_assert(sizeof(int) == sizeof(int32_t), "camIdx is passed through as int32");

CTracker::CTracker(int device_id) : _ctx(nullptr), _status(0) {
  std::memset(&_summ, 0, sizeof(_summ));
  const int rc = hsfm_ba_create(&_ctx, device_id);
  if (rc) throw std::runtime_error(std::string("hsfm_ba_create: ") + hsfm_ba_error_string(rc));
  initialiseBAOptions();
}

CTracker::~CTracker() { hsfm_ba_destroy(_ctx); }

void CTracker::initialiseBAOptions() {
  // reference: _opts.linear_solver_type = ceres::DENSE_SCHUR; progress off; logging SILENT — all else default
  hsfm_ba_default_options(&_opts);
  _opts.linear_solver = HSFM_BA_LINEAR_DENSE_SCHUR;
  _opts.verbose = 0;
}

void CTracker::bundleAdjustmentStructAndPose(const std::vector<cv::Point2d>& observations, const std::vector<int>& camIdx,
                                             const std::vector<cv::Matx33d>& K, std::vector<double*>& R, std::vector<double*>& t,
                                             std::vector<double*>& pts3D, int isStructOrPose) {
  // The reference loops i < observations.size() and indexes camIdx[i], pts3D[i], R[camNo], t[camNo],
  // K[camNo] unchecked (CTracker.cpp:676-690); the C ABI validates instead of reading out of bounds.
  const int64_t n = (int64_t)observations.size();
  std::memset(&_summ, 0, sizeof(_summ));
  if ((int64_t)camIdx.size() < n || (int64_t)pts3D.size() < n || R.size() != t.size() || K.size() < R.size()) {
    _status = HSFM_BA_ERR_ARG;
    std::fprintf(stderr, "[CTracker] bundleAdjustmentStructAndPose: inconsistent argument sizes, nothing done\n");
    return;
  }
  _status = hsfm_ba_solve_ptr(_ctx, isStructOrPose, n, n ? &observations[0].x : nullptr, reinterpret_cast<const int32_t*>(camIdx.data()),
                              (int32_t)R.size(), K.empty() ? nullptr : K[0].val, R.data(), t.data(), pts3D.data(), &_opts, &_summ);
  if (_status) std::fprintf(stderr, "[CTracker] bundle adjustment failed (%s); poses and points left untouched\n", hsfm_ba_error_string(_status));
}
