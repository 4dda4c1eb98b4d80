// CTrackerBA.h — host-side (C++) mirror of the bundle-adjustment facade of the reference's CTracker
// (CTracker.h:64-67,99; CTracker.cpp:571-577,670-702).  Same method name, same argument list, same
// void return, parameters optimised in place through the double* the caller hands in.  The body that
// used to build a ceres::Problem and call ceres::Solve now packs nothing but pointers and calls the
// C ABI (include/hsfm_ba.h); all arithmetic runs on the GPU.  There is no CPU fallback: without a
// usable sm_100 device the constructor throws std::runtime_error.
#pragma once
#include <vector>

#include "../../include/hsfm_ba.h"
#include "cv_standins.h"

class CTracker {
 public:
  explicit CTracker(int device_id = 0);
  ~CTracker();
  CTracker(const CTracker&) = delete;
  CTracker& operator=(const CTracker&) = delete;

  // bundle adjustment — CTracker.h:65
  void bundleAdjustmentStructAndPose(const std::vector<cv::Point2d>& observations, const std::vector<int>& camIdx,
                                     const std::vector<cv::Matx33d>& K, std::vector<double*>& R, std::vector<double*>& t,
                                     std::vector<double*>& pts3D, int isStructOrPose);

  enum BA_TYPE { STRUCT_ONLY = 0, POSE_ONLY = 1, STRUCT_AND_POSE = 2 };  // CTracker.h:67

  // What ceres::Solver::Summary would have carried (the reference drops it, CTracker.cpp:700-701);
  // status of the last call: 0 ok, otherwise an hsfm_ba error code (parameters were left untouched).
  const hsfm_ba_summary& lastSummary() const { return _summ; }
  int lastStatus() const { return _status; }

 private:
  void initialiseBAOptions();  // CTracker.cpp:571-577
  hsfm_ba_ctx* _ctx;
  hsfm_ba_options _opts;  // stands where ceres::Solver::Options _opts stood (CTracker.h:99)
  hsfm_ba_summary _summ;
  int _status;
};
