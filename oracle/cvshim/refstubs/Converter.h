// ORACLE test infrastructure.  Shadows the reference's include/Converter.h (which needs Eigen and g2o, neither on
// this box) when src/Frame.cc, src/KeyFrame.cc ... are compiled in place for oracle/_ref.  Only the member those
// files call on the paths oracle/_ref drives is provided: Converter::toDescriptorVector (src/Converter.cc:27-35),
// one cv::Mat row per descriptor.
#ifndef CONVERTER_H
#define CONVERTER_H
#include <opencv2/core/core.hpp>
#include <vector>
namespace ORB_SLAM2 {
class Converter {
 public:
  static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& Descriptors) {
    std::vector<cv::Mat> v;
    v.reserve(Descriptors.rows);
    for (int j = 0; j < Descriptors.rows; j++) v.push_back(Descriptors.row(j));
    return v;
  }
};
}  // namespace ORB_SLAM2
#endif
