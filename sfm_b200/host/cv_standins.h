// cv_standins.h — layout-compatible stand-ins for the two OpenCV types in the reference signature
// (CTracker.h:65): cv::Point2d = {double x, y}; cv::Matx33d = {double val[9]} row-major.  OpenCV C++
// headers are not installed in this image; a build inside hulop/SfM includes <opencv2/core.hpp>
// instead and defines HSFM_HAVE_OPENCV so the real types are used unchanged.
#pragma once
#ifdef HSFM_HAVE_OPENCV
#include <opencv2/core.hpp>
#else
namespace cv {
struct Point2d {
  double x, y;
  Point2d() : x(0), y(0) {}
  Point2d(double x_, double y_) : x(x_), y(y_) {}
};
struct Matx33d {
  double val[9];
  Matx33d() : val{0, 0, 0, 0, 0, 0, 0, 0, 0} {}
  Matx33d(double a, double b, double c, double d, double e, double f, double g, double h, double i) : val{a, b, c, d, e, f, g, h, i} {}
};
}  // namespace cv
#endif
static_assert(sizeof(cv::Point2d) == 2 * sizeof(double), "Point2d must be two packed doubles");
static_assert(sizeof(cv::Matx33d) == 9 * sizeof(double), "Matx33d must be nine packed doubles");
