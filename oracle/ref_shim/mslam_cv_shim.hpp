// Stand-in for the slice of OpenCV's C++ API that the reference's hot-path sources use
// (distributed_cv_feature.cpp, orb_feature.cpp, rgbd_feature_frontend.cpp:101-138), so that those files compile
// UNCHANGED from /root/reference into oracle/_ref/ (OpenCV's C++ headers are not in this image; only cv2 is).
//
// TEST INFRASTRUCTURE.  Nothing under modular-slam_b200/ may include this.
//
// Containers (cv::Mat, cv::KeyPoint, Point_, Size, InputArray / OutputArray) are written out here with OpenCV's
// semantics where the reference depends on them -- in particular Mat::create() keeping an already matching buffer,
// which is what makes `descriptors = cv::Mat::zeros(...)` (distributed_cv_feature.cpp:1233) write through the
// rowRange view it was handed.  The ARITHMETIC primitives forward to oracle/liborb_oracle.so, each of which is pinned
// bit-exact to cv2 4.13.0 by the committed goldens (tests/test_oracle_golden.py, tests/test_dist_oracle.py):
//   cv::resize(INTER_LINEAR) -> orc_resize_linear      cv::FAST(img, kps, thr, true) -> orc_fast_subimage
//   cv::GaussianBlur(7x7, 2) -> orc_gauss7_u8          cv::fastAtan2 -> orc_fast_atan2
//   cv::ORB::detectAndCompute -> orc_orb_detect_gray   BFMatcher(HAMMING)::knnMatch -> orc_knn2
// GUI calls (imshow / waitKey / pollKey / drawKeypoints) do nothing: unavailable on this box.
#pragma once
#include <algorithm>
#include <array>
#include <cassert>
#include <limits>
#include <list>
#include <map>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>
#include <xmmintrin.h>
#include <emmintrin.h>

typedef unsigned char uchar;

#define CV_8U 0
#define CV_8UC1 0

extern "C"
{
    struct orc_kp_shim
    {
        float x, y, size, angle, response;
        int octave;
    };
    struct orc_params_shim
    {
        int nfeatures, nlevels, fast_threshold, edge;
        float scale_factor;
    };
    void orc_resize_linear(const uint8_t* src, int sw, int sh, uint8_t* dst, int dw, int dh);
    int orc_fast_subimage(const uint8_t* img, int stride, int w, int h, int thr, int* xys, int cap);
    void orc_gauss7_u8(const uint8_t* src, int w, int h, uint8_t* dst);
    float orc_fast_atan2(float y, float x);
    int orc_orb_detect_gray(const uint8_t* gray, int W, int H, const orc_params_shim* prm, orc_kp_shim* kps, uint8_t* desc, int cap,
                            const void* dump);
    void orc_knn2(const uint8_t* query, int nq, int qstride, const uint8_t* train, int nt, int tstride, int* idx0, int* d0,
                  int* idx1, int* d1);
}

// OpenCV's x86-64 definitions (core/fast_math.hpp): cvRound = cvtsd2si / cvtss2si (round-half-even), cvFloor = truncate and fix up
inline int cvRound(double v) { return _mm_cvtsd_si32(_mm_set_sd(v)); }
inline int cvRound(float v) { return _mm_cvtss_si32(_mm_set_ss(v)); }
inline int cvRound(int v) { return v; }
inline int cvFloor(double v)
{
    const int i = (int)v;
    return i - (i > v);
}
inline int cvFloor(float v)
{
    const int i = (int)v;
    return i - (i > v);
}

inline int cvCeil(double v)
{
    const int i = (int)v;
    return i + (i < v);
}
inline int cvCeil(float v)
{
    const int i = (int)v;
    return i + (i < v);
}

namespace cv
{
template <typename T>
using Ptr = std::shared_ptr<T>;

template <typename T>
struct Point_
{
    T x{}, y{};
    Point_() = default;
    Point_(T x_, T y_) : x(x_), y(y_) {}
    template <typename S>
    Point_& operator*=(S s)
    {
        // OpenCV: a.x = saturate_cast<T>(a.x * b)
        x = static_cast<T>(x * s);
        y = static_cast<T>(y * s);
        return *this;
    }
};
using Point2i = Point_<int>;
using Point2f = Point_<float>;
using Point = Point2i;

struct Size
{
    int width = 0, height = 0;
    Size() = default;
    Size(int w, int h) : width(w), height(h) {}
};

struct Scalar
{
    double val[4] = {0, 0, 0, 0};
    Scalar() = default;
    Scalar(double v0) { val[0] = v0; }
};

struct KeyPoint
{
    Point2f pt;
    float size = 0.f, angle = -1.f, response = 0.f;
    int octave = 0, class_id = -1;
    KeyPoint() = default;
    KeyPoint(float x, float y, float size_, float angle_ = -1.f, float response_ = 0.f, int octave_ = 0, int class_id_ = -1)
        : pt(x, y), size(size_), angle(angle_), response(response_), octave(octave_), class_id(class_id_)
    {
    }
};

struct DMatch
{
    int queryIdx = -1, trainIdx = -1, imgIdx = -1;
    float distance = 0.f;
};

enum
{
    INTER_LINEAR = 1,
    BORDER_REFLECT_101 = 4,
    LINE_AA = 16
};

struct MatExpr
{
    int rows, cols, type;
};

struct MatStep
{
    size_t v = 0;
    operator size_t() const { return v; }
};

// 8-bit single-channel matrices only (all the hot path uses)
class Mat
{
  public:
    int rows = 0, cols = 0;
    uchar* data = nullptr;
    MatStep step;
    static constexpr size_t AUTO_STEP = 0;

    Mat() = default;
    Mat(int r, int c, int type, void* user, size_t step_ = AUTO_STEP) : rows(r), cols(c), data((uchar*)user)
    {
        assert(type == CV_8UC1);
        step.v = step_ == AUTO_STEP ? (size_t)c : step_;
    }
    Mat(int r, int c, int type, const Scalar& s)
    {
        create(r, c, type);
        for(int y = 0; y < rows; ++y)
            std::memset(data + y * step.v, (int)s.val[0], (size_t)cols);
    }
    // cv::Mat::create: keeps the current buffer when shape and type already match
    void create(int r, int c, int type)
    {
        assert(type == CV_8UC1);
        if(data && r == rows && c == cols)
            return;
        rows = r;
        cols = c;
        step.v = (size_t)c;
        owner = std::shared_ptr<uchar>(new uchar[(size_t)r * c + 1], std::default_delete<uchar[]>());
        data = owner.get();
    }
    void release()
    {
        owner.reset();
        data = nullptr;
        rows = cols = 0;
        step.v = 0;
    }
    Mat clone() const
    {
        Mat m;
        copyTo(m);
        return m;
    }
    void copyTo(Mat& dst) const
    {
        if(empty())
        {
            dst.release();
            return;
        }
        dst.create(rows, cols, CV_8UC1);
        for(int y = 0; y < rows; ++y)
            std::memcpy(dst.data + y * dst.step.v, data + y * step.v, (size_t)cols);
    }
    Mat rowRange(int a, int b) const
    {
        assert(0 <= a && a <= b && b <= rows);
        Mat m(*this);
        m.data = data + (size_t)a * step.v;
        m.rows = b - a;
        return m;
    }
    Mat colRange(int a, int b) const
    {
        assert(0 <= a && a <= b && b <= cols);
        Mat m(*this);
        m.data = data + a;
        m.cols = b - a;
        return m;
    }
    template <typename T>
    T& at(int y, int x)
    {
        static_assert(sizeof(T) == 1, "8-bit only");
        return *(T*)(data + (size_t)y * step.v + x);
    }
    template <typename T>
    const T& at(int y, int x) const
    {
        static_assert(sizeof(T) == 1, "8-bit only");
        return *(const T*)(data + (size_t)y * step.v + x);
    }
    template <typename T>
    T* ptr(int y = 0)
    {
        return (T*)(data + (size_t)y * step.v);
    }
    template <typename T>
    const T* ptr(int y = 0) const
    {
        return (const T*)(data + (size_t)y * step.v);
    }
    uchar* ptr(int y = 0) { return data + (size_t)y * step.v; }
    const uchar* ptr(int y = 0) const { return data + (size_t)y * step.v; }
    size_t step1() const { return step.v; }
    int type() const { return CV_8UC1; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    bool isContinuous() const { return step.v == (size_t)cols || rows <= 1; }
    static MatExpr zeros(int r, int c, int type) { return MatExpr{r, c, type}; }
    // MatOp_Initializer::assign: m.create(size, type) -- a no-op on a matching view -- then m = Scalar(0)
    Mat& operator=(const MatExpr& e)
    {
        create(e.rows, e.cols, e.type);
        for(int y = 0; y < rows; ++y)
            std::memset(data + y * step.v, 0, (size_t)cols);
        return *this;
    }
    Mat(const MatExpr& e) { *this = e; }
    // dense copy of the pixels (for the primitives that want contiguous input)
    std::vector<uchar> dense() const
    {
        std::vector<uchar> v((size_t)rows * cols);
        for(int y = 0; y < rows; ++y)
            std::memcpy(v.data() + (size_t)y * cols, data + y * step.v, (size_t)cols);
        return v;
    }

  private:
    std::shared_ptr<uchar> owner;
};

class _InputArray
{
  public:
    _InputArray() = default;
    _InputArray(const Mat& m) : m_(&m) {}
    bool empty() const { return !m_ || m_->empty(); }
    Mat getMat() const { return m_ ? *m_ : Mat(); }

  protected:
    const Mat* m_ = nullptr;
};

class _OutputArray : public _InputArray
{
  public:
    _OutputArray() = default;
    _OutputArray(Mat& m) : _InputArray(m), out_(&m) {}
    void create(int r, int c, int type) const
    {
        if(out_)
            out_->create(r, c, type);
    }
    void release() const
    {
        if(out_)
            out_->release();
    }
    Mat getMat() const { return out_ ? *out_ : Mat(); }
    Mat& getMatRef() const { return *out_; }
    bool needed() const { return out_ != nullptr; }

  private:
    Mat* out_ = nullptr;
};
using InputArray = const _InputArray&;
using OutputArray = const _OutputArray&;
using InputOutputArray = const _OutputArray&;
inline InputOutputArray noArray()
{
    static const _OutputArray none;
    return none;
}

// ---- arithmetic primitives: forwarded to the cv2-pinned C oracle -----------------------------------------------------
inline float fastAtan2(float y, float x) { return orc_fast_atan2(y, x); }

inline void resize(InputArray src_, OutputArray dst_, Size dsize, double fx = 0, double fy = 0, int interpolation = INTER_LINEAR)
{
    if(fx != 0 || fy != 0 || interpolation != INTER_LINEAR)
        throw std::runtime_error("cv shim: only resize(dsize, INTER_LINEAR) is provided");
    const Mat src = src_.getMat();
    const std::vector<uchar> s = src.dense();
    Mat out; // never aliases src (OpenCV allocates a new buffer when the size differs)
    out.create(dsize.height, dsize.width, CV_8UC1);
    orc_resize_linear(s.data(), src.cols, src.rows, out.data, dsize.width, dsize.height);
    dst_.getMatRef() = out;
}

inline void FAST(InputArray image_, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression = true)
{
    if(!nonmaxSuppression)
        throw std::runtime_error("cv shim: FAST without non-max suppression is not provided");
    const Mat img = image_.getMat();
    keypoints.clear();
    if(img.empty())
        return;
    std::vector<int> xys((size_t)img.rows * img.cols * 3 + 3);
    const int n = orc_fast_subimage(img.data, (int)img.step.v, img.cols, img.rows, threshold, xys.data(), img.rows * img.cols);
    keypoints.reserve((size_t)n);
    for(int i = 0; i < n; ++i) // OpenCV: KeyPoint((float)x, (float)y, 7.f, -1, (float)score)
        keypoints.emplace_back((float)xys[3 * i], (float)xys[3 * i + 1], 7.f, -1.f, (float)xys[3 * i + 2]);
}

inline void GaussianBlur(InputArray src_, OutputArray dst_, Size ksize, double sx, double sy = 0, int border = BORDER_REFLECT_101)
{
    if(ksize.width != 7 || ksize.height != 7 || sx != 2 || sy != 2 || border != BORDER_REFLECT_101)
        throw std::runtime_error("cv shim: only GaussianBlur(7x7, 2, 2, REFLECT_101) is provided");
    const Mat src = src_.getMat();
    const std::vector<uchar> s = src.dense();
    std::vector<uchar> d(s.size());
    orc_gauss7_u8(s.data(), src.cols, src.rows, d.data());
    Mat& dst = dst_.getMatRef();
    dst.create(src.rows, src.cols, CV_8UC1);
    for(int y = 0; y < dst.rows; ++y)
        std::memcpy(dst.ptr(y), d.data() + (size_t)y * dst.cols, (size_t)dst.cols);
}

inline void rectangle(Mat&, Point2i, Point2i, const Scalar&, int = 1, int = 8)
{
    // only reached with mask rectangles (distributed_cv_feature.cpp:811-828); DistributedOrbOpenCvDetector passes none
    throw std::runtime_error("cv shim: cv::rectangle is not provided (mask rectangles are never set by the detector)");
}

// ---- GUI: unavailable on this box ---------------------------------------------------------------------------------------
inline void drawKeypoints(InputArray, const std::vector<KeyPoint>&, InputOutputArray) {}
inline void imshow(const std::string&, InputArray) {}
inline int waitKey(int = 0) { return -1; }
inline int pollKey() { return -1; }

// ---- features2d classes used by orb_feature.cpp ---------------------------------------------------------------------------
class Feature2D
{
  public:
    virtual ~Feature2D() = default;
    virtual void detectAndCompute(InputArray image, InputArray mask, std::vector<KeyPoint>& keypoints, OutputArray descriptors,
                                  bool useProvidedKeypoints = false) = 0;
};

// cv::ORB with OpenCV's defaults (scaleFactor 1.2f, nlevels 8, edgeThreshold 31, HARRIS_SCORE, patchSize 31, fastThreshold 20).
// Keypoint order: the oracle's canonical (octave, y, x) -- OpenCV's own order inside a level is a std::nth_element artefact.
class ORB : public Feature2D
{
  public:
    static Ptr<ORB> create(int nfeatures = 500) { return Ptr<ORB>(new ORB(nfeatures)); }
    void detectAndCompute(InputArray image_, InputArray mask, std::vector<KeyPoint>& keypoints, OutputArray descriptors,
                          bool useProvidedKeypoints = false) override
    {
        if(!mask.empty() || useProvidedKeypoints)
            throw std::runtime_error("cv shim: ORB masks / provided keypoints are not provided");
        const Mat img = image_.getMat();
        const std::vector<uchar> g = img.dense();
        const orc_params_shim prm{nfeatures_, 8, 20, 31, 1.2f};
        const int cap = nfeatures_ * 4 + 64;
        std::vector<orc_kp_shim> kps((size_t)cap);
        std::vector<uchar> desc((size_t)cap * 32);
        const int n = orc_orb_detect_gray(g.data(), img.cols, img.rows, &prm, kps.data(), desc.data(), cap, nullptr);
        if(n > cap)
            throw std::runtime_error("cv shim: ORB keypoint capacity exceeded");
        keypoints.clear();
        for(int i = 0; i < n; ++i)
            keypoints.emplace_back(kps[i].x, kps[i].y, kps[i].size, kps[i].angle, kps[i].response, kps[i].octave, -1);
        if(n == 0)
        {
            descriptors.release();
            return;
        }
        descriptors.create(n, 32, CV_8U);
        Mat d = descriptors.getMat();
        for(int i = 0; i < n; ++i)
            std::memcpy(d.ptr(i), desc.data() + (size_t)i * 32, 32);
    }

  private:
    explicit ORB(int n) : nfeatures_(n) {}
    int nfeatures_;
};

class DescriptorMatcher
{
  public:
    enum MatcherType
    {
        FLANNBASED = 1,
        BRUTEFORCE = 2,
        BRUTEFORCE_L1 = 3,
        BRUTEFORCE_HAMMING = 4
    };
    static Ptr<DescriptorMatcher> create(const MatcherType& type)
    {
        if(type != BRUTEFORCE_HAMMING)
            throw std::runtime_error("cv shim: only BRUTEFORCE_HAMMING is provided");
        return Ptr<DescriptorMatcher>(new DescriptorMatcher());
    }
    // BFMatcher::knnMatchImpl, k <= 2: per query the k smallest distances ascending, ties -> lower train index;
    // fewer than k entries when the train set is smaller
    void knnMatch(InputArray query_, InputArray train_, std::vector<std::vector<DMatch>>& matches, int k) const
    {
        if(k < 1 || k > 2)
            throw std::runtime_error("cv shim: knnMatch k must be 1 or 2");
        const Mat q = query_.getMat(), t = train_.getMat();
        matches.clear();
        const int nq = q.empty() ? 0 : q.rows, nt = t.empty() ? 0 : t.rows;
        if(nq == 0)
            return;
        std::vector<int> r((size_t)nq * 4);
        orc_knn2(q.data, nq, (int)q.step.v, t.data, nt, nt ? (int)t.step.v : 32, r.data(), r.data() + nq, r.data() + 2 * nq,
                 r.data() + 3 * nq);
        matches.resize((size_t)nq);
        for(int i = 0; i < nq; ++i)
        {
            if(r[i] >= 0)
                matches[i].push_back(DMatch{i, r[i], 0, (float)r[nq + i]});
            if(k == 2 && r[2 * nq + i] >= 0)
                matches[i].push_back(DMatch{i, r[2 * nq + i], 0, (float)r[3 * nq + i]});
        }
    }
};
} // namespace cv
