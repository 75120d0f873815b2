// Test stub of the extractor plug-in point (abstract Extract(Frame&)).
#pragma once
#include <memory>
#include "frame/frame.h"
namespace visionx {
class FeatureExtractor {
public:
    using Ptr = std::shared_ptr<FeatureExtractor>;
    virtual ~FeatureExtractor() = default;
    virtual void Extract(Frame& frame) = 0;
};
}  // namespace visionx
