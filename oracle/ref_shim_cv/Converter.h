// ORACLE / test infrastructure: shadows the reference's include/Converter.h (which needs Eigen + g2o) with the one
// declaration the compiled sources use; defined in oracle/ref_orbmatcher.cc.
#pragma once
#include <vector>
#include <opencv2/core/core.hpp>
namespace ORB_SLAM {
class Converter {
 public:
  static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& Descriptors);
};
}  // namespace ORB_SLAM
