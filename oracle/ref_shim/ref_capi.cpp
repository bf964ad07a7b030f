// oracle/_ref/libmslam_ref.so: the REFERENCE's own sources for the hot path, compiled unchanged from where they lie
// under /root/reference (oracle/Makefile target `_ref`), behind a small C ABI for the tests.  TEST INFRASTRUCTURE.
//
// What is the reference's own code in this library:
//   frame.cpp                      mslam::toGrayScale                                  (separate TU, unchanged)
//   orb_feature.cpp                OrbOpenCvDetector::detect / OrbOpenCvMatcher::match  (separate TU, unchanged)
//   distributed_cv_feature.cpp     DistributedOrbOpenCvDetector and everything under it (included below, unchanged;
//                                  `private` is widened so that OrbExtractorPimpl::extract's full cv::KeyPoint output --
//                                  angle, octave, size, response -- can be compared, not only the (x, y, descriptor) that
//                                  detect() keeps)
//   rgbd_feature_frontend.cpp      reconstructPoint + pointsFromRgbdKeypoints, cut out by line pattern at build time into
//                                  oracle/_ref/gen_backproject.inc (the rest of that file needs the map / Boost.Range)
// What is NOT the reference's: Eigen / OpenCV / spdlog, replaced by oracle/ref_shim/ (see mslam_cv_shim.hpp for which
// OpenCV primitives forward to the cv2-pinned C oracle).
#include <algorithm>
#include <array>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <iterator>
#include <limits>
#include <list>
#include <map>
#include <memory>
#include <optional>
#include <string>
#include <utility>
#include <vector>
#include <x86intrin.h>

#include "mslam_cv_shim.hpp"
#include "mslam_eigen_shim.hpp"

#define private public
#include REF_DIST_CPP // distributed_cv_feature.cpp
#undef private

#include "modular_slam/orb_feature.hpp"
#include "modular_slam/types/depth_frame.hpp"

namespace mslam
{
#include "gen_backproject.inc" // rgbd_feature_frontend.cpp: reconstructPoint, pointsFromRgbdKeypoints
}

static_assert(sizeof(mslam::Keypoint) == 32 && sizeof(mslam::OrbKeypoint) == 64 && offsetof(mslam::OrbKeypoint, descriptor) == 32 &&
                  sizeof(mslam::DescriptorMatch) == 16,
              "SURVEY 8(b) layouts");

#define REF_API extern "C" __attribute__((visibility("default")))

namespace
{
mslam::RgbFrame make_frame(const uint8_t* rgb, int w, int h)
{
    mslam::RgbFrame f;
    f.data.assign(rgb, rgb + (size_t)w * h * 3);
    f.size = mslam::Size{w, h};
    return f;
}
struct kp_out
{
    float x, y, size, angle, response;
    int octave;
};
int export_keypoints(const std::vector<mslam::OrbKeypoint>& v, double* xy, uint64_t* ids, uint8_t* desc, int cap)
{
    const int n = (int)v.size();
    for(int i = 0; i < n && i < cap; ++i)
    {
        xy[2 * i] = v[(size_t)i].keypoint.coordinates.x();
        xy[2 * i + 1] = v[(size_t)i].keypoint.coordinates.y();
        if(ids)
            ids[i] = v[(size_t)i].keypoint.id;
        std::memcpy(desc + (size_t)i * 32, v[(size_t)i].descriptor.data(), 32);
    }
    return n;
}
} // namespace

REF_API void ref_layout(int* out4)
{
    out4[0] = (int)sizeof(mslam::Keypoint);
    out4[1] = (int)sizeof(mslam::OrbKeypoint);
    out4[2] = (int)offsetof(mslam::OrbKeypoint, descriptor);
    out4[3] = (int)sizeof(mslam::DescriptorMatch);
}

// frame.cpp:6-27
REF_API void ref_gray(const uint8_t* rgb, int w, int h, uint8_t* out)
{
    const auto g = mslam::toGrayScale(make_frame(rgb, w, h));
    std::memcpy(out, g.data.data(), g.data.size());
}

// distributed_cv_feature.cpp:1190-1222 -- the public detect(): (id, x, y as double, descriptor) per keypoint
REF_API int ref_dist_detect(const uint8_t* rgb, int w, int h, double* xy, uint64_t* ids, uint8_t* desc, int cap)
{
    mslam::DistributedOrbOpenCvDetector det;
    return export_keypoints(det.detect(make_frame(rgb, w, h)), xy, ids, desc, cap);
}

// distributed_cv_feature.cpp:719-809 -- OrbExtractorPimpl::extract on a gray image: every cv::KeyPoint field
REF_API int ref_dist_extract(const uint8_t* gray, int w, int h, kp_out* kps, uint8_t* desc, int cap)
{
    mslam::DistributedOrbOpenCvDetector det;
    cv::Mat g(h, w, CV_8UC1, const_cast<uint8_t*>(gray));
    std::vector<cv::KeyPoint> k;
    cv::Mat d;
    det.pimpl->extract(g, cv::Mat(), k, d);
    const int n = (int)k.size();
    for(int i = 0; i < n && i < cap; ++i)
    {
        kps[i] = kp_out{k[(size_t)i].pt.x, k[(size_t)i].pt.y, k[(size_t)i].size, k[(size_t)i].angle, k[(size_t)i].response, k[(size_t)i].octave};
        std::memcpy(desc + (size_t)i * 32, d.ptr(i), 32);
    }
    return n;
}

// distributed_cv_feature.cpp:981-1022 -- the quad-tree alone: (x, y, response) in, surviving keypoints out in list order
REF_API int ref_dist_tree(const int* xys, int n, int min_x, int max_x, int min_y, int max_y, float scale_factor, int* out_xys, int cap)
{
    mslam::DistributedOrbOpenCvDetector det;
    std::vector<cv::KeyPoint> in;
    for(int i = 0; i < n; ++i)
        in.emplace_back((float)xys[3 * i], (float)xys[3 * i + 1], 7.f, -1.f, (float)xys[3 * i + 2]);
    const auto out = det.pimpl->distribute_keypoints_via_tree(in, min_x, max_x, min_y, max_y, scale_factor);
    for(size_t i = 0; i < out.size() && (int)i < cap; ++i)
    {
        out_xys[3 * i] = (int)out[i].pt.x;
        out_xys[3 * i + 1] = (int)out[i].pt.y;
        out_xys[3 * i + 2] = (int)out[i].response;
    }
    return (int)out.size();
}

// distributed_cv_feature.cpp:456-503, 519-541, 413-421
REF_API float ref_dist_cos(float v) { return mslam::util::cos(v); }
REF_API float ref_dist_sin(float v) { return mslam::util::sin(v); }
REF_API void ref_dist_umax(int* out16)
{
    mslam::orb_impl impl;
    for(int i = 0; i < 16; ++i)
        out16[i] = impl.u_max_.at((size_t)i);
}
REF_API void ref_dist_scale_factors(float* out8)
{
    const auto s = mslam::orb_params::calc_scale_factors(8, 1.2f);
    std::copy(s.begin(), s.end(), out8);
}
REF_API void ref_dist_descriptor(const uint8_t* blurred, int w, int h, int cx, int cy, float angle_deg, uint8_t* desc32)
{
    mslam::orb_impl impl;
    cv::Mat img(h, w, CV_8UC1, const_cast<uint8_t*>(blurred));
    impl.compute_orb_descriptor(cv::KeyPoint((float)cx, (float)cy, 31.f, angle_deg), img, desc32);
}
REF_API float ref_dist_ic_angle(const uint8_t* img_, int w, int h, int cx, int cy)
{
    mslam::orb_impl impl;
    cv::Mat img(h, w, CV_8UC1, const_cast<uint8_t*>(img_));
    return impl.ic_angle(img, cv::Point2f((float)cx, (float)cy));
}

// orb_feature.cpp:33-65
REF_API int ref_orb_detect(const uint8_t* rgb, int w, int h, double* xy, uint8_t* desc, int cap)
{
    mslam::OrbOpenCvDetector det;
    return export_keypoints(det.detect(make_frame(rgb, w, h)), xy, nullptr, desc, cap);
}

// orb_feature.cpp:84-117 -- records are 64-byte OrbKeypoint images (descriptor at offset 32), exactly what match() views
REF_API int ref_orb_match(const uint8_t* from_recs, int n_from, const uint8_t* to_recs, int n_to, uint64_t* from_idx, uint64_t* to_idx, int cap)
{
    std::vector<mslam::OrbKeypoint> from((size_t)n_from), to((size_t)n_to);
    if(n_from)
        std::memcpy((void*)from.data(), from_recs, (size_t)n_from * 64);
    if(n_to)
        std::memcpy((void*)to.data(), to_recs, (size_t)n_to * 64);
    mslam::OrbOpenCvMatcher m;
    const auto r = m.match(from, to);
    for(size_t i = 0; i < r.size() && (int)i < cap; ++i)
    {
        from_idx[i] = r[i].fromIndex;
        to_idx[i] = r[i].toIndex;
    }
    return (int)r.size();
}

// rgbd_feature_frontend.cpp:101-138 + types/depth_frame.hpp:20-30
REF_API void ref_backproject(const double* xy, int n, const uint16_t* depth, int w, int h, double cx, double cy, double fx, double fy,
                             float factor, double* xyz, uint8_t* valid)
{
    mslam::DepthFrame df;
    df.data.assign(depth, depth + (size_t)w * h);
    df.size = mslam::Size{w, h};
    df.cameraParameters.principalPoint = mslam::Vector2{cx, cy};
    df.cameraParameters.focal = mslam::Vector2{fx, fy};
    df.cameraParameters.factor = factor;
    std::vector<mslam::OrbKeypoint> kps((size_t)n);
    for(int i = 0; i < n; ++i)
        kps[(size_t)i].keypoint.coordinates = mslam::Vector2{xy[2 * i], xy[2 * i + 1]};
    const auto pts = mslam::pointsFromRgbdKeypoints(df, kps);
    for(int i = 0; i < n; ++i)
    {
        valid[i] = pts[(size_t)i].has_value();
        for(int c = 0; c < 3; ++c)
            xyz[3 * i + c] = valid[i] ? (*pts[(size_t)i])[(size_t)c] : 0.0;
    }
}
