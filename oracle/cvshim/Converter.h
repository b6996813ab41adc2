// TEST INFRASTRUCTURE — stands in for the reference's include/Converter.h (which needs Eigen and
// g2o, neither installed) in the verbatim build of src/Frame.cc and src/KeyFrame.cc.  Those two
// files use one function of it: toDescriptorVector (src/Converter.cc:27-35: one cv::Mat row
// header per descriptor).
#pragma once
#include <vector>
#include <opencv2/core/core.hpp>
namespace ORB_SLAM2 {
class Converter {
public:
    static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& Descriptors) {
        std::vector<cv::Mat> vDesc;
        vDesc.reserve(Descriptors.rows);
        for (int j = 0; j < Descriptors.rows; j++) vDesc.push_back(Descriptors.row(j));
        return vDesc;
    }
};
}
