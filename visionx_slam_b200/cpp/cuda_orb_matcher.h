// cuda_orb_matcher.h — drop-in for visionx::ORBMatcher (core/feature/orb_matcher.h:11-29).
//
// Same Options{nn_ratio = 0.8f, min_matches = 50} and the same Match(last, curr, matches)
// contract as core/feature/orb_matcher.cpp:11-43: matches is cleared; 0 is returned when either
// descriptor matrix is empty; otherwise every query row of `last` whose best neighbour in `curr`
// passes  d1 < nn_ratio * d2  is appended in ascending queryIdx order as
// cv::DMatch{queryIdx, trainIdx, imgIdx = 0, distance}; a warning is logged below min_matches and
// the count is returned.  The 2-NN search that cv::BFMatcher(NORM_HAMMING).knnMatch did
// (orb_matcher.cpp:22-25) runs in libvxorb.so's sm_100a kernels; ties go to the lowest train index
// exactly as OpenCV's batchDistance.
#pragma once

#include <memory>
#include <string>
#include <vector>

#include "cuda_orb_extractor.h"
#include "feature/feature_matcher.h"
#include "vxorb.h"

namespace visionx {

class CudaORBMatcher : public FeatureMatcher {
public:
    struct Options {
        float nn_ratio = 0.8f;
        int min_matches = 50;
        int device_id = 0;
        int max_descriptors = 16384;  // sizing hint; the device scratch grows with the descriptor counts it meets
    };

    CudaORBMatcher() : CudaORBMatcher(Options()) {}
    explicit CudaORBMatcher(const Options& options);
    // Shares the extractor's handle: frames that extractor processed are matched from the descriptors it left on
    // the device (vxorb_match_ids, only the matches cross the link); any other descriptor matrix is uploaded as
    // before.  This is the constructor system.cpp:15-16 should use.
    explicit CudaORBMatcher(std::shared_ptr<CudaORBExtractor> extractor) : CudaORBMatcher(std::move(extractor), Options()) {}
    CudaORBMatcher(std::shared_ptr<CudaORBExtractor> extractor, const Options& options);
    ~CudaORBMatcher() override;
    CudaORBMatcher(const CudaORBMatcher&) = delete;
    CudaORBMatcher& operator=(const CudaORBMatcher&) = delete;

    int Match(const Frame::Ptr& last, const Frame::Ptr& curr, std::vector<cv::DMatch>& matches) override;

    // Match() followed by the post-filter every caller applies (tracking.cpp:211-222, 292-304,
    // 343-355): keep distance <= max(2 * min_dist, 30) with min_dist starting at 100.  Fused into
    // the match kernel's epilogue; not part of the reference interface.
    int MatchFiltered(const Frame::Ptr& last, const Frame::Ptr& curr, std::vector<cv::DMatch>& matches);

    int last_status() const { return last_status_; }
    // Match calls served from device-resident descriptors / through the upload path since construction.
    int resident_calls() const { return resident_calls_; }
    int upload_calls() const { return upload_calls_; }

private:
    int Run(bool filtered, const Frame::Ptr& last, const Frame::Ptr& curr, std::vector<cv::DMatch>& matches);
    Options options_;
    std::shared_ptr<CudaORBExtractor> extractor_;  // keeps the shared handle alive
    vxorb_handle* handle_ = nullptr;
    bool owns_handle_ = true;
    int resident_calls_ = 0, upload_calls_ = 0;
    std::vector<vxorb_match_t> buf_;
    int last_status_ = VXORB_OK;
};

}  // namespace visionx
