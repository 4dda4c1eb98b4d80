// C++ driver for tests/test_gpu_host_shim.py: builds the argument vectors exactly as
// CSfM::bundleAdjustment does (CSfM.cpp:311-341) from a flat binary dump and calls the drop-in
// CTracker::bundleAdjustmentStructAndPose.  Test infrastructure.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../sfm_b200/host/CTrackerBA.h"

template <typename T>
static std::vector<T> rd(FILE* f, size_t n) {
  std::vector<T> v(n);
  if (n && fread(v.data(), sizeof(T), n, f) != n) { std::fprintf(stderr, "short read\n"); std::exit(2); }
  return v;
}

int main(int argc, char** argv) {
  if (argc < 4) return 2;
  FILE* f = fopen(argv[1], "rb");
  if (!f) return 2;
  const int mode = std::atoi(argv[3]);
  auto hdr = rd<int64_t>(f, 3);
  const size_t n = hdr[0], nc = hdr[1], np = hdr[2];
  auto obs = rd<double>(f, 2 * n);
  auto cam_idx = rd<int32_t>(f, n);
  auto pt_idx = rd<int32_t>(f, n);
  auto K9 = rd<double>(f, 9 * nc);
  auto cam = rd<double>(f, 6 * nc);
  auto pts = rd<double>(f, 3 * np);
  fclose(f);
  // keyframe / map storage as the reference owns it: per-frame _rot, _t (CFrame.h:70,72) and vector<Matx31d> _pts3D (CMap.h:83)
  struct Frame { double rot[3], t[3]; };
  std::vector<Frame> frames(nc);
  for (size_t c = 0; c < nc; ++c) for (int k = 0; k < 3; ++k) { frames[c].rot[k] = cam[6 * c + k]; frames[c].t[k] = cam[6 * c + 3 + k]; }
  std::vector<cv::Point2d> observations;
  std::vector<int> camIdx;
  std::vector<cv::Matx33d> K(nc);
  std::vector<double*> R, t, pts3D;
  for (size_t c = 0; c < nc; ++c) {
    for (int k = 0; k < 9; ++k) K[c].val[k] = K9[9 * c + k];
    t.push_back(frames[c].t);
    R.push_back(frames[c].rot);
  }
  for (size_t i = 0; i < n; ++i) {
    observations.emplace_back(obs[2 * i], obs[2 * i + 1]);
    camIdx.push_back(cam_idx[i]);
    pts3D.push_back(&pts[3 * (size_t)pt_idx[i]]);  // same pointer for every observation of a point
  }
  CTracker tracker(0);
  tracker.bundleAdjustmentStructAndPose(observations, camIdx, K, R, t, pts3D, mode);
  if (tracker.lastStatus()) return 3;
  FILE* o = fopen(argv[2], "wb");
  for (size_t c = 0; c < nc; ++c) { fwrite(frames[c].rot, sizeof(double), 3, o); fwrite(frames[c].t, sizeof(double), 3, o); }
  fwrite(pts.data(), sizeof(double), pts.size(), o);
  const hsfm_ba_summary& s = tracker.lastSummary();
  double tail[3] = {s.final_cost, (double)s.num_iterations, (double)s.termination_type};
  fwrite(tail, sizeof(double), 3, o);
  fclose(o);
  return 0;
}
