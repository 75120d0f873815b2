// Test stub of the matcher plug-in point (abstract Match(last, curr, matches)).
#pragma once
#include <memory>
#include <vector>
#include "frame/frame.h"
namespace visionx {
class FeatureMatcher {
public:
    using Ptr = std::shared_ptr<FeatureMatcher>;
    virtual ~FeatureMatcher() = default;
    virtual int Match(const Frame::Ptr& last, const Frame::Ptr& curr, std::vector<cv::DMatch>& matches) = 0;
};
}  // namespace visionx
