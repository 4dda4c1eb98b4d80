// Stand-in for ceres/rotation.h (Ceres Solver 1.12, absent from /root/reference and from this image) —
// CPU ORACLE / test infrastructure, see oracle/ba_oracle.h.  Only what CTracker.cpp:588,620,652 calls.
#ifndef REF_STANDIN_CERES_ROTATION_H_
#define REF_STANDIN_CERES_ROTATION_H_
#include "../../ba_oracle_jet.h"

namespace ceres {
// ceres::AngleAxisRotatePoint — the oracle's restatement of the published routine (ba_oracle_jet.h).
template <typename T>
inline void AngleAxisRotatePoint(const T angle_axis[3], const T pt[3], T result[3]) {
  ba_oracle::AngleAxisRotatePoint(angle_axis, pt, result);
}
}  // namespace ceres
#endif
