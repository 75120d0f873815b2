// cuda_orb_extractor.cpp — see cuda_orb_extractor.h.  Marshals between Frame (core/frame/frame.h:
// 25-64) and the C ABI; no image arithmetic happens here.
#include "cuda_orb_extractor.h"

#include <glog/logging.h>

#include <algorithm>
#include <cstring>
#include <stdexcept>

namespace visionx {

CudaORBExtractor::CudaORBExtractor(int n_features, float scale_factor, int n_levels, int device_id,
                                   int max_width, int max_height) {
    vxorb_default_params(&params_);
    params_.n_features = n_features;      // orb_extractor.cpp:6 cv::ORB::create(n_features, scale, levels)
    params_.scale_factor = scale_factor;
    params_.n_levels = n_levels;
    params_.device_id = device_id;
    params_.max_width = max_width;
    params_.max_height = max_height;
    const int rc = vxorb_create(&params_, &handle_);
    if (rc != VXORB_OK)
        throw std::runtime_error(std::string("CudaORBExtractor: vxorb_create failed: ") + vxorb_last_error(nullptr));
    kp_cap_ = vxorb_keypoint_capacity(handle_);
    kp_.resize(static_cast<size_t>(kp_cap_));
}

// cv::ORB takes frames of any size; a vxorb handle is sized at creation.  A frame beyond the current limits gets a
// new, larger handle (limits rounded up to 64 px); the old one is destroyed only after the new one exists, and the
// frames it kept resident are forgotten (a later Match of those frames uploads their descriptors instead).
bool CudaORBExtractor::Reserve(int width, int height) {
    if (width <= params_.max_width && height <= params_.max_height) return true;
    vxorb_params p = params_;
    p.max_width = std::max(p.max_width, (width + 63) / 64 * 64);
    p.max_height = std::max(p.max_height, (height + 63) / 64 * 64);
    vxorb_handle* bigger = nullptr;
    if (vxorb_create(&p, &bigger) != VXORB_OK) {
        last_error_ = std::string("cannot grow to ") + std::to_string(width) + "x" + std::to_string(height) + ": " +
                      vxorb_last_error(nullptr);
        return false;
    }
    LOG(INFO) << "[CudaORBExtractor] frame " << width << "x" << height << " exceeds " << params_.max_width << "x"
              << params_.max_height << ": handle rebuilt for " << p.max_width << "x" << p.max_height;
    vxorb_destroy(handle_);
    handle_ = bigger;
    params_ = p;
    for (Resident& r : resident_) r = Resident{};
    resident_next_ = 0;
    kp_cap_ = vxorb_keypoint_capacity(handle_);
    kp_.resize(static_cast<size_t>(kp_cap_));
    ++rebuilds_;
    return true;
}

CudaORBExtractor::~CudaORBExtractor() { vxorb_destroy(handle_); }

uint64_t CudaORBExtractor::Fingerprint(const cv::Mat& desc) {
    if (desc.empty() || desc.cols != VXORB_DESC_BYTES) return 0;
    uint64_t a[4], b[4], f = 0x9e3779b97f4a7c15ull ^ static_cast<uint64_t>(desc.rows);
    std::memcpy(a, desc.data, 32);
    std::memcpy(b, desc.data + static_cast<size_t>(desc.rows - 1) * desc.step, 32);
    for (int i = 0; i < 4; ++i) f = (f ^ a[i]) * 0x100000001b3ull ^ (b[i] + (f << 7));
    return f;
}

long long CudaORBExtractor::ResidentId(const Frame& frame) const {
    const cv::Mat& d = frame.Descriptors();
    for (const Resident& r : resident_)
        if (r.vx_id != 0 && r.frame_id == frame.Id() && r.rows == d.rows && r.fingerprint == Fingerprint(d) &&
            vxorb_set_size(handle_, r.vx_id) == d.rows)
            return r.vx_id;
    return 0;
}

bool CudaORBExtractor::FeatureDepths(const Frame& frame, const cv::Mat& depth, std::vector<double>& depth_m) {
    depth_m.clear();
    const long long id = ResidentId(frame);
    const int type = depth.depth() == CV_16U ? 0 : depth.depth() == CV_32F ? 1 : depth.depth() == CV_64F ? 2 : -1;
    if (id == 0 || type < 0 || depth.empty() || depth.channels() != 1) return false;
    depth_m.resize(frame.Features().size());
    int n = 0;
    const int rc = vxorb_depth_lookup(handle_, id, depth.data, type, depth.cols, depth.rows, static_cast<size_t>(depth.step),
                                      static_cast<int>(depth_m.size()), depth_m.data(), &n);
    if (rc != VXORB_OK || n != static_cast<int>(depth_m.size())) {
        LOG(ERROR) << "[CudaORBExtractor] depth look-up: " << vxorb_last_error(handle_);
        depth_m.clear();
        return false;
    }
    return true;
}

void CudaORBExtractor::Extract(Frame& frame) {
    auto& features = frame.Features();
    features.clear();  // orb_extractor.cpp:16
    const cv::Mat& image = frame.Image();
    const bool usable = !image.empty() && image.depth() == CV_8U && (image.channels() == 1 || image.channels() == 3);
    const bool fits = usable && Reserve(image.cols, image.rows);  // may replace handle_ and kp_cap_
    cv::Mat desc(kp_cap_, VXORB_DESC_BYTES, CV_8UC1);
    int n = 0;
    int rc = VXORB_E_INVALID;
    long long vx_id = 0;
    if (fits) {
        // cv::ORB converts 3-channel input to gray itself; system.cpp:46 feeds IMREAD_COLOR frames
        rc = vxorb_extract_id(handle_, image.data, image.channels(), image.cols, image.rows,
                              static_cast<size_t>(image.step), kp_cap_, kp_.data(), desc.data, &n, &info_, &vx_id);
    }
    if (rc != VXORB_OK && rc != VXORB_E_CAPACITY) info_ = vxorb_frame_info{};
    last_status_ = rc;
    if (rc != VXORB_OK && rc != VXORB_E_CAPACITY) {
        if (!usable) last_error_ = "empty or unsupported image";
        else if (fits) last_error_ = vxorb_last_error(handle_);  // else: Reserve left its message
        LOG(ERROR) << "[CudaORBExtractor] " << last_error_;
        frame.Descriptors() = cv::Mat();
        return;
    }
    if (rc == VXORB_E_CAPACITY) {
        last_error_ = vxorb_last_error(handle_);
        LOG(WARNING) << "[CudaORBExtractor] " << last_error_;
    }
    if (n > kp_cap_) n = kp_cap_;
    features.reserve(static_cast<size_t>(n));
    for (int i = 0; i < n; ++i) {  // orb_extractor.cpp:19-24
        Feature f;
        f.position = Eigen::Vector2d(kp_[i].x, kp_[i].y);
        f.response = kp_[i].response;
        features.emplace_back(f);
    }
    frame.Descriptors() = n > 0 ? desc.rowRange(0, n).clone() : cv::Mat();  // orb_extractor.cpp:26
    if (rc == VXORB_OK && vx_id != 0 && n > 0) {  // remember where this frame's descriptors live on the device
        Resident& r = resident_[resident_next_];
        resident_next_ = (resident_next_ + 1) % 8;
        r.frame_id = frame.Id();
        r.vx_id = vx_id;
        r.rows = n;
        r.fingerprint = Fingerprint(frame.Descriptors());
    }
}

}  // namespace visionx
