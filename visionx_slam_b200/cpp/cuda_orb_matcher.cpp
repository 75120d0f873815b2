// cuda_orb_matcher.cpp — see cuda_orb_matcher.h.
#include "cuda_orb_matcher.h"

#include <glog/logging.h>

#include <stdexcept>

namespace visionx {

CudaORBMatcher::CudaORBMatcher(const Options& options) : options_(options) {
    vxorb_params p;
    vxorb_default_params(&p);
    p.nn_ratio = options.nn_ratio;
    p.min_matches = options.min_matches;
    p.device_id = options.device_id;
    p.max_train = options.max_descriptors;
    p.max_width = p.max_height = 64;  // the matcher handle never extracts
    const int rc = vxorb_create(&p, &handle_);
    if (rc != VXORB_OK)
        throw std::runtime_error(std::string("CudaORBMatcher: vxorb_create failed: ") + vxorb_last_error(nullptr));
}

CudaORBMatcher::CudaORBMatcher(std::shared_ptr<CudaORBExtractor> extractor, const Options& options)
    : options_(options), extractor_(std::move(extractor)) {
    if (!extractor_) throw std::runtime_error("CudaORBMatcher: null extractor");
    handle_ = extractor_->handle();
    owns_handle_ = false;
}

CudaORBMatcher::~CudaORBMatcher() {
    if (owns_handle_) vxorb_destroy(handle_);
}

int CudaORBMatcher::Run(bool filtered, const Frame::Ptr& last, const Frame::Ptr& curr,
                        std::vector<cv::DMatch>& matches) {
    matches.clear();  // orb_matcher.cpp:13
    const cv::Mat& desc1 = last->Descriptors();
    const cv::Mat& desc2 = curr->Descriptors();
    if (desc1.empty() || desc2.empty()) return 0;  // orb_matcher.cpp:18-20
    // ORB descriptors come out of Extract() continuous; clone only if a caller handed us a view
    const cv::Mat q = desc1.isContinuous() ? desc1 : desc1.clone();
    const cv::Mat t = desc2.isContinuous() ? desc2 : desc2.clone();
    buf_.resize(static_cast<size_t>(q.rows));
    int n = 0;
    if (extractor_) handle_ = extractor_->handle();  // the extractor rebuilds its handle for larger frames
    const long long id_q = extractor_ ? extractor_->ResidentId(*last) : 0;
    const long long id_t = extractor_ ? extractor_->ResidentId(*curr) : 0;
    if (id_q != 0 && id_t != 0) {  // both descriptor sets are still on the device
        last_status_ = vxorb_match_ids(handle_, id_q, id_t, 0.f, options_.nn_ratio, filtered ? 1 : 0, q.rows, buf_.data(), &n);
        ++resident_calls_;
    } else {
        auto fn = filtered ? vxorb_match_filtered : vxorb_match;
        last_status_ = fn(handle_, q.data, q.rows, t.data, t.rows, options_.nn_ratio, q.rows, buf_.data(), &n);
        ++upload_calls_;
    }
    if (last_status_ != VXORB_OK) {
        // an error, not "no matches": the reference's matcher has no failure path, so say it loudly
        LOG(ERROR) << "[CudaORBMatcher] vxorb status " << last_status_ << ": " << vxorb_last_error(handle_);
        return 0;
    }
    matches.reserve(static_cast<size_t>(n));
    for (int i = 0; i < n; ++i)  // orb_matcher.cpp:34: push_back(m1)
        matches.emplace_back(buf_[i].query_idx, buf_[i].train_idx, 0, static_cast<float>(buf_[i].distance));
    if (matches.size() < static_cast<size_t>(options_.min_matches))  // orb_matcher.cpp:38-40
        LOG(WARNING) << "[ORBMatcher] Too few matches: " << matches.size();
    return static_cast<int>(matches.size());
}

int CudaORBMatcher::Match(const Frame::Ptr& last, const Frame::Ptr& curr, std::vector<cv::DMatch>& matches) {
    return Run(false, last, curr, matches);
}

int CudaORBMatcher::MatchFiltered(const Frame::Ptr& last, const Frame::Ptr& curr,
                                  std::vector<cv::DMatch>& matches) {
    return Run(true, last, curr, matches);
}

}  // namespace visionx
