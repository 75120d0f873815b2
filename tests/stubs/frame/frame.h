// Test stub of the data carrier at the boundary: the members of visionx::Frame / visionx::Feature
// that Extract and Match read and write (image, features, descriptors).  Pose, camera, depth and
// Clone are not touched by the feature path and are left out.
#pragma once
#include <Eigen/Core>
#include <cstdint>
#include <memory>
#include <opencv2/core.hpp>
#include <vector>

namespace visionx {

struct Feature {
    Eigen::Vector2d position;
    float response = 0.0f;
    uint64_t landmark_id_ = 0;
    bool has_landmark = false;
    bool is_outlier = false;
};

class Frame {
public:
    using Ptr = std::shared_ptr<Frame>;
    Frame(uint64_t id, double timestamp, const cv::Mat& image) : id_(id), timestamp_(timestamp), image_(image.clone()) {}
    uint64_t Id() const { return id_; }
    double Timestamp() const { return timestamp_; }
    const cv::Mat& Image() const { return image_; }
    std::vector<Feature>& Features() { return features_; }
    const std::vector<Feature>& Features() const { return features_; }
    cv::Mat& Descriptors() { return descriptors_; }
    const cv::Mat& Descriptors() const { return descriptors_; }

private:
    uint64_t id_;
    double timestamp_;
    cv::Mat image_;
    std::vector<Feature> features_;
    cv::Mat descriptors_;
};

}  // namespace visionx
