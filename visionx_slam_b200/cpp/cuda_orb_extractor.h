// cuda_orb_extractor.h — drop-in for visionx::ORBExtractor (core/feature/orb_extractor.h:9-20).
//
// Same constructor arguments and the same Extract(Frame&) contract as the reference class
// (core/feature/orb_extractor.cpp:9-27): frame.Features() is replaced by one Feature per keypoint
// (position = level-0 pixel coordinates as double, response = Harris response) and
// frame.Descriptors() is assigned a freshly owned N x 32 CV_8UC1 matrix.  The arithmetic
// cv::ORB::detectAndCompute did on the CPU (orb_extractor.cpp:13) runs in libvxorb.so's sm_100a
// kernels through the C ABI of include/vxorb.h.  There is no CPU fallback: construction throws
// std::runtime_error when no B200-class device / library is usable, and a failing device call
// leaves the frame with zero features (the reference's "no features" path, tracking.cpp:179-184)
// after logging the vxorb error text.
//
// Keypoint order is canonical (octave, y, x) instead of OpenCV's nth_element artefact order; the
// keypoint SET, responses and descriptors are identical.
#pragma once

#include <memory>
#include <string>
#include <vector>

#include "feature/feature_extractor.h"
#include "vxorb.h"

namespace visionx {

class CudaORBExtractor : public FeatureExtractor {
public:
    explicit CudaORBExtractor(int n_features = 1000, float scale_factor = 1.2f, int n_levels = 8,
                              int device_id = 0, int max_width = 1920, int max_height = 1088);
    ~CudaORBExtractor() override;
    CudaORBExtractor(const CudaORBExtractor&) = delete;
    CudaORBExtractor& operator=(const CudaORBExtractor&) = delete;

    void Extract(Frame& frame) override;

    // Per-frame quantities of the last Extract, computed on the device beside the extraction
    // (vxorb_extract_info): what Tracking::CheckImageQuality (tracking.cpp:120-139) and
    // Tracking::CheckFeatureDistribution (tracking.cpp:93-118) recompute on the host.
    const vxorb_frame_info& last_frame_info() const { return info_; }
    // mean in [30, 225] and stddev >= 20 (tracking.cpp:130-137)
    bool ImageQualityOk() const { return !(info_.mean < 30 || info_.mean > 225) && !(info_.stddev < 20); }
    // at least half of the 5 x 5 grid cells hold a feature (tracking.cpp:116)
    bool FeatureDistributionOk() const { return info_.grid_occupied >= (5 * 5 * 0.5); }

    // Last vxorb status / message (VXORB_OK after a successful Extract).
    int last_status() const { return last_status_; }
    const std::string& last_error() const { return last_error_; }
    // A CudaORBMatcher constructed with this extractor shares the handle (one device context, one set of
    // streams) and matches the descriptors Extract left on the device instead of uploading them again.
    // The pointer changes when a frame larger than the current limits makes Extract rebuild the handle: read it per
    // call, do not keep it.
    vxorb_handle* handle() const { return handle_; }
    // Frames larger than max_width x max_height are accepted like cv::ORB accepts them: the handle is rebuilt for
    // the larger size on the first such frame (a one-off cost of a few milliseconds).  Number of rebuilds so far:
    int rebuilds() const { return rebuilds_; }
    // Device-resident id of the descriptors Extract produced for `frame` (vxorb_extract_id), or 0 when the
    // frame was not extracted here, has been overwritten (ring of 8 frames) or its descriptor matrix no longer
    // is what Extract wrote (row count and a fingerprint of the first and last rows are compared).
    long long ResidentId(const Frame& frame) const;
    // The per-feature depth Tracking::CreateLandmarksFromDepth (tracking.cpp:586-650) reads from frame.Depth():
    // depth_m[i] in metres for feature i, 0 where the reference skips it (outside the image, 16-bit depth 0, outside
    // [0.1, 10] m).  Uses the positions Extract left on the device (only the depth image goes up); returns false
    // when the frame is no longer resident or the depth type is not CV_16U / CV_32F / CV_64F — the caller then runs
    // its own loop.
    bool FeatureDepths(const Frame& frame, const cv::Mat& depth, std::vector<double>& depth_m);

private:
    struct Resident {
        uint64_t frame_id = 0;
        long long vx_id = 0;
        int rows = 0;
        uint64_t fingerprint = 0;
    };
    static uint64_t Fingerprint(const cv::Mat& desc);
    bool Reserve(int width, int height);
    vxorb_params params_{};
    int rebuilds_ = 0;
    Resident resident_[8];
    int resident_next_ = 0;
    vxorb_handle* handle_ = nullptr;
    int kp_cap_ = 0;
    std::vector<vxorb_keypoint> kp_;
    vxorb_frame_info info_{};
    int last_status_ = VXORB_OK;
    std::string last_error_;
};

}  // namespace visionx
