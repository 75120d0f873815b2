// adapter_main.cpp — drives the C++ adapters the way the reference's front end does
// (tracking.cpp:42 Extract on every frame; tracking.cpp:208 Match(init, current)).
//
//   adapter_main <in.bin> <out.bin> [initial_max_side]
// in.bin : int32 n_frames, w, h, channels; then n_frames * h * w * channels bytes
// out.bin: per frame: int32 n; n * (double x, double y, float response); n * 32 descriptor bytes
//          then per consecutive pair (i-1 -> i): int32 m; m * (int32 q, int32 t, float dist)
//          then the same for MatchFiltered.
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <vector>

#include "cuda_orb_extractor.h"
#include "cuda_orb_matcher.h"

using namespace visionx;

int main(int argc, char** argv) {
    if (argc != 3 && argc != 4) return 2;
    const int initial_max = argc == 4 ? atoi(argv[3]) : 0;  // > 0: start with a handle too small for the frames
    FILE* fi = fopen(argv[1], "rb");
    if (!fi) return 3;
    int hdr[4];
    if (fread(hdr, 4, 4, fi) != 4) return 3;
    const int n = hdr[0], w = hdr[1], h = hdr[2], ch = hdr[3];
    std::vector<unsigned char> pix((size_t)w * h * ch);
    auto cuda_extractor = initial_max > 0 ? std::make_shared<CudaORBExtractor>(1000, 1.2f, 8, 0, initial_max, initial_max)
                                          : std::make_shared<CudaORBExtractor>();
    std::shared_ptr<FeatureExtractor> extractor = cuda_extractor;                                       // system.cpp:15
    std::shared_ptr<FeatureMatcher> matcher = std::make_shared<CudaORBMatcher>(cuda_extractor);         // system.cpp:16
    std::shared_ptr<FeatureMatcher> standalone = std::make_shared<CudaORBMatcher>();  // own handle: always uploads
    std::vector<Frame::Ptr> frames;
    for (int i = 0; i < n; ++i) {
        if (fread(pix.data(), 1, pix.size(), fi) != pix.size()) return 3;
        cv::Mat img(h, w, ch == 3 ? CV_8UC3 : CV_8UC1, pix.data(), (size_t)w * ch);
        auto f = std::make_shared<Frame>((uint64_t)i, 0.1 * i, img);
        extractor->Extract(*f);
        // the caller-side checks of tracking.cpp:93-139, answered from the device-computed frame info
        auto* ce = dynamic_cast<CudaORBExtractor*>(extractor.get());
        const vxorb_frame_info& fi_ = ce->last_frame_info();
        fprintf(stderr, "FRAMEINFO %d %.17g %.17g %d %d %d\n", i, fi_.mean, fi_.stddev, fi_.grid_occupied,
                ce->ImageQualityOk() ? 1 : 0, ce->FeatureDistributionOk() ? 1 : 0);
        // the depth look-up of CreateLandmarksFromDepth (tracking.cpp:586-650) on a synthetic 16-bit depth image
        if (ch == 1) {
            std::vector<uint16_t> dep((size_t)w * h);
            for (int y = 0; y < h; ++y)
                for (int x = 0; x < w; ++x) dep[(size_t)y * w + x] = (uint16_t)((x * 37 + y * 101 + i * 977) % 60000);
            cv::Mat depth(h, w, CV_MAKETYPE(CV_16U, 1), dep.data(), (size_t)w * 2);
            std::vector<double> dm;
            const bool ok = ce->FeatureDepths(*f, depth, dm);
            double sum = 0.0;
            int valid = 0;
            for (double d : dm) {
                sum += d;
                valid += d > 0.0 ? 1 : 0;
            }
            fprintf(stderr, "DEPTH %d %d %d %.17g\n", i, ok ? 1 : 0, valid, sum);
        }
        frames.push_back(f);
    }
    fclose(fi);
    FILE* fo = fopen(argv[2], "wb");
    if (!fo) return 4;
    for (auto& f : frames) {
        const int k = (int)f->Features().size();
        fwrite(&k, 4, 1, fo);
        for (auto& ft : f->Features()) {
            const double xy[2] = {ft.position.x(), ft.position.y()};
            fwrite(xy, 8, 2, fo);
            fwrite(&ft.response, 4, 1, fo);
        }
        if (k != f->Descriptors().rows || (k && f->Descriptors().cols != 32)) return 5;
        for (int r = 0; r < k; ++r) fwrite(f->Descriptors().data + r * f->Descriptors().step, 1, 32, fo);
    }
    auto* cm = dynamic_cast<CudaORBMatcher*>(matcher.get());
    for (int pass = 0; pass < 2; ++pass)
        for (int i = 1; i < n; ++i) {
            std::vector<cv::DMatch> matches(3);  // must be cleared by Match
            const int m = pass == 0 ? matcher->Match(frames[i - 1], frames[i], matches)
                                    : cm->MatchFiltered(frames[i - 1], frames[i], matches);
            if (m != (int)matches.size()) return 6;
            if (pass == 0) {  // the upload path of a matcher with its own handle must give the same rows
                std::vector<cv::DMatch> again;
                if (standalone->Match(frames[i - 1], frames[i], again) != m) return 8;
                for (int k = 0; k < m; ++k)
                    if (again[k].queryIdx != matches[k].queryIdx || again[k].trainIdx != matches[k].trainIdx ||
                        again[k].distance != matches[k].distance)
                        return 8;
            }
            fwrite(&m, 4, 1, fo);
            for (auto& d : matches) {
                if (d.imgIdx != 0) return 7;
                fwrite(&d.queryIdx, 4, 1, fo);
                fwrite(&d.trainIdx, 4, 1, fo);
                fwrite(&d.distance, 4, 1, fo);
            }
        }
    fclose(fo);
    fprintf(stderr, "REBUILDS %d\n", cuda_extractor->rebuilds());
    fprintf(stderr, "MATCHPATH resident=%d upload=%d standalone_upload=%d\n", cm->resident_calls(), cm->upload_calls(),
            dynamic_cast<CudaORBMatcher*>(standalone.get())->upload_calls());
    return 0;
}
