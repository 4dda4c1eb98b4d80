// Stand-ins for the two OpenCV types in the signature of CTracker::bundleAdjustmentStructAndPose (CTracker.h:65):
// cv::Point2d (two doubles x, y) and cv::Matx33d (nine doubles, row-major, member `val`) — CPU ORACLE / test
// infrastructure.  OpenCV's C++ headers are not in this image.  The reference has
// `using namespace cv; using namespace std;` in scope (CTracker.h), reproduced by the generated shell.
#ifndef REF_STANDIN_CV_TYPES_H_
#define REF_STANDIN_CV_TYPES_H_
namespace cv {
struct Point2d { double x, y; };
struct Matx33d { double val[9]; };
static_assert(sizeof(Point2d) == 16 && sizeof(Matx33d) == 72, "layout of cv::Point2d / cv::Matx33d");
}  // namespace cv
#endif
