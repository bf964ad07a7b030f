// orb_b200_plugin.cpp -- see orb_b200_plugin.hpp.  Host side only: every number is computed by the CUDA library behind
// include/mslam_b200.h; this file converts between the reference's containers and the C ABI's flat buffers.
#include "orb_b200_plugin.hpp"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <mutex>
#include <sstream>
#include <stdexcept>

namespace mslam
{

// the layouts the C ABI is fed with (the matcher passes &v[0].descriptor[0] with stride sizeof(OrbKeypoint), exactly as
// orb_feature.cpp:88-91 hands the same memory to cv::Mat)
static_assert(sizeof(Keypoint) == 32, "Keypoint layout (feature_interface.hpp:18-22)");
static_assert(sizeof(OrbKeypoint) == 64, "KeypointDescriptor<uint8_t,32> layout (feature_interface.hpp:24-29)");
static_assert(offsetof(OrbKeypoint, descriptor) == 32, "descriptor offset inside OrbKeypoint");
static_assert(sizeof(DescriptorMatch) == 16, "DescriptorMatch layout (feature_interface.hpp:37-41)");

// ---- shared context ------------------------------------------------------------------------------------------------
class OrbB200Context
{
  public:
    explicit OrbB200Context(const OrbB200Options& o) : options(o)
    {
        mslam_b200_config cfg;
        mslam_b200_default_config(&cfg);
        cfg.device = o.device;
        cfg.max_width = o.maxWidth;
        cfg.max_height = o.maxHeight;
        cfg.max_batch = maxBatch = o.maxBatch > 0 ? o.maxBatch : 8;
        cfg.nfeatures = o.nfeatures;
        cfg.nlevels = o.nlevels;
        cfg.scale_factor = o.scaleFactor;
        cfg.fast_threshold = o.fastThreshold;
        cfg.ratio = o.ratio;
        cfg.max_descriptors = o.maxDescriptors > 0 ? o.maxDescriptors : 4 * o.nfeatures;
        const int rc = mslam_b200_create(&cfg, &handle);
        if(rc != MSLAM_B200_OK)
            throw std::runtime_error(std::string("mslam_b200_create failed: ") + mslam_b200_status_string(rc) +
                                     " (the B200 plugin has no CPU fallback)");
        kpCapacity = o.nfeatures + o.nfeatures / 4 + 64; // the library's default per-frame capacity (ties of retainBest)
        descCapacity = cfg.max_descriptors;
    }
    ~OrbB200Context() { mslam_b200_destroy(handle); }
    OrbB200Context(const OrbB200Context&) = delete;
    OrbB200Context& operator=(const OrbB200Context&) = delete;

    void report(const char* what, int rc)
    {
        if(!reported)
            std::fprintf(stderr, "[mslam_b200 plugin] %s: %s %s\n", what, mslam_b200_status_string(rc),
                         mslam_b200_last_cuda_error(handle));
        reported = true;
    }

    int maxBatch = 8; // frames per batched call: OrbB200Options::maxBatch (plugin_config.json "maxBatch")
    OrbB200Options options;
    mslam_b200_ctx* handle = nullptr;
    int kpCapacity = 0, descCapacity = 0;
    std::mutex mutex; // detector and matcher may share one context
    bool reported = false;
};

static std::mutex g_defaultMutex;
static OrbB200Options g_defaultOptions;
static bool g_defaultLoaded = false;
static std::weak_ptr<OrbB200Context> g_sharedContext;

void setDefaultOrbB200Options(const OrbB200Options& options)
{
    std::lock_guard<std::mutex> lock(g_defaultMutex);
    g_defaultOptions = options;
    g_defaultLoaded = true;
}

OrbB200Options defaultOrbB200Options()
{
    std::lock_guard<std::mutex> lock(g_defaultMutex);
    if(!g_defaultLoaded)
    {
        g_defaultLoaded = true;
        if(const char* path = std::getenv("MSLAM_B200_PLUGIN_CONFIG"))
        {
            std::ifstream in(path);
            std::stringstream ss;
            ss << in.rdbuf();
            std::vector<PluginEntry> entries;
            std::string err;
            if(in && parsePluginConfig(ss.str(), &entries, &err) && !entries.empty())
                g_defaultOptions = entries[0].options;
            else
                std::fprintf(stderr, "[mslam_b200 plugin] cannot use %s: %s\n", path, err.c_str());
        }
    }
    return g_defaultOptions;
}

static bool sameOptions(const OrbB200Options& a, const OrbB200Options& b)
{
    return a.device == b.device && a.maxWidth == b.maxWidth && a.maxHeight == b.maxHeight && a.nfeatures == b.nfeatures &&
           a.nlevels == b.nlevels && a.scaleFactor == b.scaleFactor && a.fastThreshold == b.fastThreshold &&
           a.ratio == b.ratio && a.maxDescriptors == b.maxDescriptors && a.maxBatch == b.maxBatch;
}

// detector and matcher built with the same options share one context (one stream, one set of HBM buffers)
static std::shared_ptr<OrbB200Context> acquireContext(const OrbB200Options& o)
{
    std::lock_guard<std::mutex> lock(g_defaultMutex);
    if(auto sp = g_sharedContext.lock())
        if(sameOptions(sp->options, o))
            return sp;
    auto sp = std::make_shared<OrbB200Context>(o);
    g_sharedContext = sp;
    return sp;
}

// ---- detector ------------------------------------------------------------------------------------------------------
OrbB200Detector::OrbB200Detector() : OrbB200Detector(defaultOrbB200Options()) {}

OrbB200Detector::OrbB200Detector(const OrbB200Options& options) : ctx(acquireContext(options))
{
    kpBuf.resize((size_t)ctx->kpCapacity * ctx->maxBatch);
    descBuf.resize((size_t)ctx->kpCapacity * ctx->maxBatch * MSLAM_B200_DESC_BYTES);
}

OrbB200Detector::~OrbB200Detector() = default;

static void fillKeypoints(const mslam_b200_keypoint* kp, const std::uint8_t* desc, int n, std::vector<OrbKeypoint>* out)
{
    out->resize((size_t)n);
    for(int i = 0; i < n; ++i)
    {
        OrbKeypoint& k = (*out)[(size_t)i];
        k.keypoint.id = (Id)i; // the reference leaves id uninitialised (orb_feature.cpp:52-58); the index is harmless
        k.keypoint.coordinates[0] = (double)kp[i].x; // float -> double, orb_feature.cpp:57
        k.keypoint.coordinates[1] = (double)kp[i].y;
        std::memcpy(k.descriptor.data(), desc + (size_t)i * MSLAM_B200_DESC_BYTES, MSLAM_B200_DESC_BYTES);
    }
}

std::vector<OrbKeypoint> OrbB200Detector::detect(const RgbFrame& sensorData)
{
    std::vector<OrbKeypoint> result;
    const int w = sensorData.size.width, h = sensorData.size.height;
    if(w <= 0 || h <= 0 || sensorData.data.size() < (size_t)w * h * 3)
        return result;
    std::lock_guard<std::mutex> lock(ctx->mutex);
    int n = 0;
    const int rc = mslam_b200_detect(ctx->handle, sensorData.data.data(), w, h, 3 * w, kpBuf.data(), descBuf.data(),
                                     ctx->kpCapacity, &n);
    if(rc != MSLAM_B200_OK && rc != MSLAM_B200_E_CAPACITY)
    {
        ctx->report("detect", rc);
        return result;
    }
    fillKeypoints(kpBuf.data(), descBuf.data(), n < ctx->kpCapacity ? n : ctx->kpCapacity, &result);
    return result;
}

std::vector<std::vector<OrbKeypoint>> OrbB200Detector::detectBatch(const std::vector<const RgbFrame*>& frames)
{
    std::vector<std::vector<OrbKeypoint>> result(frames.size());
    std::lock_guard<std::mutex> lock(ctx->mutex);
    std::vector<const std::uint8_t*> ptrs((size_t)ctx->maxBatch);
    std::vector<int> counts((size_t)ctx->maxBatch);
    for(size_t base = 0; base < frames.size(); base += (size_t)ctx->maxBatch)
    {
        const int nb = (int)std::min<size_t>((size_t)ctx->maxBatch, frames.size() - base);
        const int w = frames[base]->size.width, h = frames[base]->size.height;
        bool uniform = w > 0 && h > 0;
        for(int i = 0; i < nb; ++i)
        {
            const RgbFrame* f = frames[base + (size_t)i];
            uniform = uniform && f->size.width == w && f->size.height == h && f->data.size() >= (size_t)w * h * 3;
            ptrs[i] = f->data.data();
        }
        if(!uniform)
            continue; // frames of one batch must share a size; callers mix sizes through detect()
        std::fill(counts.begin(), counts.end(), 0);
        const int rc = mslam_b200_detect_batch(ctx->handle, ptrs.data(), nb, w, h, 3 * w, kpBuf.data(), descBuf.data(),
                                               ctx->kpCapacity, counts.data());
        if(rc != MSLAM_B200_OK && rc != MSLAM_B200_E_CAPACITY)
        {
            ctx->report("detectBatch", rc);
            continue;
        }
        for(int i = 0; i < nb; ++i)
            fillKeypoints(kpBuf.data() + (size_t)i * ctx->kpCapacity,
                          descBuf.data() + (size_t)i * ctx->kpCapacity * MSLAM_B200_DESC_BYTES,
                          counts[i] < ctx->kpCapacity ? counts[i] : ctx->kpCapacity, &result[base + (size_t)i]);
    }
    return result;
}

// ---- the throughput path: flat pinned results, two batches in flight --------------------------------------------------
OrbB200BatchResult::OrbB200BatchResult(int maxFrames_, int capacity_) : maxFrames(maxFrames_), capacity(capacity_)
{
    void *a = nullptr, *b = nullptr, *c = nullptr;
    const size_t n = (size_t)maxFrames * capacity;
    if(maxFrames < 1 || capacity < 1 || mslam_b200_host_alloc(&a, n * sizeof(mslam_b200_keypoint)) != MSLAM_B200_OK ||
       mslam_b200_host_alloc(&b, n * MSLAM_B200_DESC_BYTES) != MSLAM_B200_OK ||
       mslam_b200_host_alloc(&c, (size_t)maxFrames * sizeof(int)) != MSLAM_B200_OK)
    {
        mslam_b200_host_free(a), mslam_b200_host_free(b), mslam_b200_host_free(c);
        throw std::runtime_error("OrbB200BatchResult: pinned host allocation failed (no CUDA runtime?)");
    }
    kp = static_cast<mslam_b200_keypoint*>(a), desc = static_cast<std::uint8_t*>(b), counts = static_cast<int*>(c);
    std::memset(counts, 0, (size_t)maxFrames * sizeof(int));
}

OrbB200BatchResult::~OrbB200BatchResult()
{
    mslam_b200_host_free(kp), mslam_b200_host_free(desc), mslam_b200_host_free(counts);
}

std::vector<OrbKeypoint> OrbB200BatchResult::keypoints(int frame) const
{
    std::vector<OrbKeypoint> out;
    if(frame >= 0 && frame < nFrames)
        fillKeypoints(raw(frame), descriptors(frame), count(frame), &out);
    return out;
}

int OrbB200Detector::maxBatch() const { return ctx->maxBatch; }
int OrbB200Detector::keypointCapacity() const { return ctx->kpCapacity; }

bool OrbB200Detector::detectBatchBegin(const std::uint8_t* const* frames, int nFrames, int width, int height, OrbB200BatchResult* out)
{
    if(!frames || !out || nFrames < 1 || nFrames > ctx->maxBatch || nFrames > out->maxFrames || out->capacity != ctx->kpCapacity)
        return false;
    std::lock_guard<std::mutex> lock(ctx->mutex);
    const int rc = mslam_b200_detect_batch_begin(ctx->handle, frames, nFrames, width, height, 3 * width, out->kp, out->desc,
                                                 out->capacity, out->counts);
    if(rc != MSLAM_B200_OK)
    {
        ctx->report("detectBatchBegin", rc);
        return false;
    }
    out->nFrames = nFrames;
    return true;
}

bool OrbB200Detector::detectBatchEnd()
{
    std::lock_guard<std::mutex> lock(ctx->mutex);
    const int rc = mslam_b200_detect_batch_end(ctx->handle);
    if(rc != MSLAM_B200_OK && rc != MSLAM_B200_E_CAPACITY)
    {
        ctx->report("detectBatchEnd", rc);
        return false;
    }
    return true;
}

// ---- second detector ---------------------------------------------------------------------------------------------
DistributedOrbB200Detector::DistributedOrbB200Detector() : DistributedOrbB200Detector(defaultOrbB200Options()) {}

DistributedOrbB200Detector::DistributedOrbB200Detector(const OrbB200Options& o)
{
    mslam_b200_dist_config cfg;
    mslam_b200_dist_default_config(&cfg);
    cfg.device = o.device;
    cfg.max_width = o.maxWidth;
    cfg.max_height = o.maxHeight;
    cfg.max_batch = 1;
    cfg.nlevels = o.nlevels;
    cfg.scale_factor = o.scaleFactor;
    cfg.ini_fast_threshold = o.iniFastThreshold;
    cfg.min_fast_threshold = o.minFastThreshold;
    cfg.min_size = o.minSize;
    const int rc = mslam_b200_dist_create(&cfg, &handle);
    if(rc != MSLAM_B200_OK)
        throw std::runtime_error(std::string("mslam_b200_dist_create failed: ") + mslam_b200_status_string(rc) +
                                 " (the B200 plugin has no CPU fallback)");
    capacity = mslam_b200_dist_keypoint_capacity(handle);
    kpBuf.resize((size_t)capacity);
    descBuf.resize((size_t)capacity * MSLAM_B200_DESC_BYTES);
}

DistributedOrbB200Detector::~DistributedOrbB200Detector() { mslam_b200_dist_destroy(handle); }

std::vector<OrbKeypoint> DistributedOrbB200Detector::detect(const RgbFrame& sensorData)
{
    std::vector<OrbKeypoint> result;
    const int w = sensorData.size.width, h = sensorData.size.height;
    if(w <= 0 || h <= 0 || sensorData.data.size() < (size_t)w * h * 3)
        return result;
    int n = 0;
    const int rc = mslam_b200_dist_detect(handle, sensorData.data.data(), w, h, 3 * w, kpBuf.data(), descBuf.data(), capacity, &n);
    if(rc != MSLAM_B200_OK && rc != MSLAM_B200_E_CAPACITY)
    {
        if(!reported)
            std::fprintf(stderr, "[mslam_b200 plugin] distributed detect: %s %s\n", mslam_b200_status_string(rc),
                         mslam_b200_dist_last_cuda_error(handle));
        reported = true;
        return result;
    }
    fillKeypoints(kpBuf.data(), descBuf.data(), n < capacity ? n : capacity, &result); // id = index (:1201-1205)
    return result;
}

// ---- matcher -------------------------------------------------------------------------------------------------------
OrbB200Matcher::OrbB200Matcher() : OrbB200Matcher(defaultOrbB200Options()) {}

OrbB200Matcher::OrbB200Matcher(const OrbB200Options& options) : ctx(acquireContext(options)) {}

OrbB200Matcher::~OrbB200Matcher()
{
    mslam_b200_host_free(mA), mslam_b200_host_free(mB), mslam_b200_host_free(mD), mslam_b200_host_free(mN);
}

bool OrbB200Matcher::ensureBatchBuffers()
{
    if(mA)
        return true;
    void* p[4] = {nullptr, nullptr, nullptr, nullptr};
    const size_t n = (size_t)ctx->maxBatch * ctx->kpCapacity * sizeof(int);
    bool ok = true;
    for(int i = 0; i < 4; ++i)
        ok = ok && mslam_b200_host_alloc(&p[i], i < 3 ? n : (size_t)ctx->maxBatch * sizeof(int)) == MSLAM_B200_OK;
    if(!ok)
    {
        for(void* q : p)
            mslam_b200_host_free(q);
        return false;
    }
    mA = static_cast<int*>(p[0]), mB = static_cast<int*>(p[1]), mD = static_cast<int*>(p[2]), mN = static_cast<int*>(p[3]);
    return true;
}

bool OrbB200Matcher::matchBatchBegin(const OrbB200BatchResult& first, const OrbB200BatchResult& second)
{
    const int P = first.frames(), cap = first.capacity;
    if(P < 1 || P != second.frames() || cap != second.capacity || P > ctx->maxBatch || cap > ctx->descCapacity || cap != ctx->kpCapacity)
        return false;
    std::lock_guard<std::mutex> lock(ctx->mutex);
    if(pendingPairs || !ensureBatchBuffers())
        return false;
    // the C ABI's batch form: mode 0 = knn + ratio, out_a = fromIndex (index into `first`), out_b = toIndex (into `second`)
    const int rc = mslam_b200_match_batch_begin(ctx->handle, 0, first.desc, first.counts, second.desc, second.counts, P, cap, mA, mB, mD, mN);
    if(rc != MSLAM_B200_OK)
    {
        ctx->report("matchBatchBegin", rc);
        return false;
    }
    pendingPairs = P, pendingCap = cap;
    return true;
}

bool OrbB200Matcher::matchBatchEnd(std::vector<std::vector<DescriptorMatch>>* out)
{
    std::lock_guard<std::mutex> lock(ctx->mutex);
    if(!pendingPairs || !out)
        return false;
    const int P = pendingPairs, cap = pendingCap;
    pendingPairs = 0;
    const int rc = mslam_b200_match_batch_end(ctx->handle);
    if(rc != MSLAM_B200_OK)
    {
        ctx->report("matchBatchEnd", rc);
        return false;
    }
    out->resize((size_t)P);
    for(int i = 0; i < P; ++i)
    {
        std::vector<DescriptorMatch>& m = (*out)[(size_t)i];
        m.resize((size_t)mN[i]);
        for(int j = 0; j < mN[i]; ++j)
        {
            m[(size_t)j].fromIndex = (std::size_t)mA[(size_t)i * cap + j];
            m[(size_t)j].toIndex = (std::size_t)mB[(size_t)i * cap + j];
        }
    }
    return true;
}

bool OrbB200Matcher::matchBatch(const OrbB200BatchResult& first, const OrbB200BatchResult& second,
                                std::vector<std::vector<DescriptorMatch>>* out)
{
    return matchBatchBegin(first, second) && matchBatchEnd(out);
}

std::vector<DescriptorMatch> OrbB200Matcher::match(const std::vector<OrbKeypoint>& from, const std::vector<OrbKeypoint>& to)
{
    // OrbOpenCvMatcher::Pimpl::match(from, to), orb_feature.cpp:84-117.  The reference dereferences &v[0] of empty
    // vectors and match[1] of a 1-row train set (undefined behaviour); the replacement returns no matches there.
    std::vector<DescriptorMatch> result;
    if(from.size() < 2 || to.empty() || from.size() > (size_t)ctx->descCapacity || to.size() > (size_t)ctx->descCapacity)
    {
        if(from.size() > (size_t)ctx->descCapacity || to.size() > (size_t)ctx->descCapacity)
            ctx->report("match: descriptor set larger than maxDescriptors", MSLAM_B200_E_CAPACITY);
        return result;
    }
    std::lock_guard<std::mutex> lock(ctx->mutex);
    fromBuf.resize(to.size());
    toBuf.resize(to.size());
    int n = 0;
    const int rc = mslam_b200_match_knn_ratio(ctx->handle, from[0].descriptor.data(), (int)from.size(), (int)sizeof(OrbKeypoint),
                                              to[0].descriptor.data(), (int)to.size(), (int)sizeof(OrbKeypoint), fromBuf.data(),
                                              toBuf.data(), nullptr, (int)to.size(), &n);
    if(rc != MSLAM_B200_OK)
    {
        ctx->report("match", rc);
        return result;
    }
    result.resize((size_t)n);
    for(int i = 0; i < n; ++i)
    {
        result[(size_t)i].fromIndex = (std::size_t)fromBuf[(size_t)i]; // trainIdx  (orb_feature.cpp:112)
        result[(size_t)i].toIndex = (std::size_t)toBuf[(size_t)i];     // queryIdx  (orb_feature.cpp:113)
    }
    return result;
}

} // namespace mslam
