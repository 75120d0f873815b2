// Test stub: the slice of cv::Mat / cv::DMatch the adapters use (see tests/stubs/README.md).
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#define CV_8U 0
#define CV_16U 2
#define CV_32F 5
#define CV_64F 6
#define CV_MAKETYPE(depth, cn) ((depth) + (((cn) - 1) << 3))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)

namespace cv {

class Mat {
public:
    int rows = 0, cols = 0;
    unsigned char* data = nullptr;
    size_t step = 0;

    Mat() = default;
    Mat(int r, int c, int type) : rows(r), cols(c), type_(type) {
        step = static_cast<size_t>(c) * channels();
        buf_ = std::shared_ptr<unsigned char>(new unsigned char[step * (r > 0 ? r : 1)], std::default_delete<unsigned char[]>());
        data = buf_.get();
    }
    // view on caller memory (no ownership), like cv::Mat(rows, cols, type, data, step)
    Mat(int r, int c, int type, void* p, size_t s) : rows(r), cols(c), data(static_cast<unsigned char*>(p)), step(s), type_(type) {}

    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int depth() const { return type_ & 7; }
    int channels() const { return (type_ >> 3) + 1; }
    int type() const { return type_; }
    bool isContinuous() const { return step == static_cast<size_t>(cols) * channels(); }
    Mat clone() const {
        if (empty()) return Mat();
        Mat m(rows, cols, type_);
        for (int r = 0; r < rows; ++r) std::memcpy(m.data + r * m.step, data + r * step, m.step);
        return m;
    }
    Mat rowRange(int a, int b) const {
        Mat m = *this;
        m.data = data + static_cast<size_t>(a) * step;
        m.rows = b - a;
        return m;
    }

private:
    int type_ = CV_8UC1;
    std::shared_ptr<unsigned char> buf_;
};

struct DMatch {
    DMatch() = default;
    DMatch(int q, int t, int img, float d) : queryIdx(q), trainIdx(t), imgIdx(img), distance(d) {}
    int queryIdx = -1, trainIdx = -1, imgIdx = -1;
    float distance = 0.f;
};

}  // namespace cv
