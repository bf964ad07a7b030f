// orb_b200_plugin.hpp -- the drop-in plugin classes: modular-slam's own detector / matcher interfaces implemented on
// the B200 C ABI (include/mslam_b200.h).
//
//   mslam::OrbB200Detector  replaces  mslam::OrbOpenCvDetector   (orb_feature.hpp:19-31, orb_feature.cpp:19-72)
//   mslam::OrbB200Matcher   replaces  mslam::OrbOpenCvMatcher    (orb_feature.hpp:33-47, orb_feature.cpp:74-117)
//   mslam::DistributedOrbB200Detector  replaces  mslam::DistributedOrbOpenCvDetector
//                                      (distributed_cv_feature.hpp:10-21, distributed_cv_feature.cpp:1181-1222; viewer.cpp:183)
//
// Both are constructed the way the reference apps construct theirs (std::make_shared<...>(), rgbd_slam.cpp:74-76) or
// through the boost::dll factory aliases exported by the plugin library (plugin_loader.hpp:19-25):
//   orbDetectorFactory : std::unique_ptr<mslam::IOrbFeatureDetector>()
//   orbMatcherFactory  : std::unique_ptr<mslam::IOrbMatcher>()
//   distributedOrbDetectorFactory : std::unique_ptr<mslam::IOrbFeatureDetector>()
// Error behaviour mirrors the reference: no exceptions out of detect()/match(); an empty vector means "nothing"
// (a CUDA failure is logged to stderr once per object and also yields an empty vector).  Constructors throw
// std::runtime_error when no CUDA device / context can be created -- there is no CPU fallback.
#pragma once
#ifdef MSLAM_B200_WITH_REFERENCE_HEADERS
#include "modular_slam/orb_feature.hpp"
#else
#include "mslam_shim.hpp"
#endif
#include <memory>
#include <string>
#include <vector>

#include "../../include/mslam_b200.h"

namespace mslam
{

// parameters the reference hard-codes (cv::ORB::create(1000), ratio 0.7); overridable through plugin_config.json
struct OrbB200Options
{
    int device = 0;
    int maxWidth = 1920, maxHeight = 1080;
    int nfeatures = 1000;
    int nlevels = 8;
    float scaleFactor = 1.2f;
    int fastThreshold = 20;
    double ratio = 0.7;
    int maxDescriptors = 0; // matcher capacity per side; 0 = derived from nfeatures
    int maxBatch = 8;       // frames per batched call (detectBatch*, matchBatch); plugin_config.json "maxBatch"
    // DistributedOrbB200Detector only: orb_params("orb", 1.2f, 8, 20, 7), OrbExtractorPimpl(params, 1000)
    // (distributed_cv_feature.cpp:1184-1186); it also reads device, maxWidth, maxHeight, nlevels, scaleFactor
    int iniFastThreshold = 20, minFastThreshold = 7, minSize = 1000;
};

// one context shared by the detector and the matcher of a process (one CUDA stream, one set of HBM buffers)
class OrbB200Context;

// What a batched detect leaves behind: the C ABI's flat outputs for up to maxBatch frames, in pinned host memory so that the
// device->host copies of a pipelined call run at full speed.  The matcher's batched form reads these directly (no
// per-keypoint conversion); keypoints(i) converts one frame to the reference's container on demand.
class OrbB200BatchResult
{
  public:
    OrbB200BatchResult(int maxFrames, int capacity);
    ~OrbB200BatchResult();
    OrbB200BatchResult(const OrbB200BatchResult&) = delete;
    OrbB200BatchResult& operator=(const OrbB200BatchResult&) = delete;
    int frames() const { return nFrames; }
    int count(int frame) const { return counts[frame] < capacity ? counts[frame] : capacity; }
    const mslam_b200_keypoint* raw(int frame) const { return kp + (std::size_t)frame * capacity; }
    const std::uint8_t* descriptors(int frame) const { return desc + (std::size_t)frame * capacity * MSLAM_B200_DESC_BYTES; }
    std::vector<OrbKeypoint> keypoints(int frame) const; // id = index, coordinates widened to double (orb_feature.cpp:52-62)

  private:
    friend class OrbB200Detector;
    friend class OrbB200Matcher;
    int maxFrames, capacity, nFrames = 0;
    mslam_b200_keypoint* kp = nullptr;
    std::uint8_t* desc = nullptr;
    int* counts = nullptr;
};

class OrbB200Detector : public IOrbFeatureDetector
{
  public:
    OrbB200Detector();
    explicit OrbB200Detector(const OrbB200Options& options);
    ~OrbB200Detector() override;
    std::vector<OrbKeypoint> detect(const RgbFrame& sensorData) override;
    // batched variant for callers that hold several frames (independent sequences); same results frame by frame
    std::vector<std::vector<OrbKeypoint>> detectBatch(const std::vector<const RgbFrame*>& frames);
    // The throughput path: up to maxBatch frames of one size per call, results left in `out` (flat, pinned).  Begin returns as
    // soon as the batch is enqueued; up to TWO batches may be in flight, End waits for the oldest -- so the copy-in of the next
    // batches overlaps the kernels and the copy-out of the current one.  Frame pointers (RgbFrame::data layout, 3 bytes per
    // pixel) and `out` must stay valid until the matching End; pinned frame memory (mslam_b200_host_alloc,
    // RgbdFileB200Provider::fetchBatch) is what makes the copies asynchronous.  false = rejected / failed (logged once).
    bool detectBatchBegin(const std::uint8_t* const* frames, int nFrames, int width, int height, OrbB200BatchResult* out);
    bool detectBatchEnd();
    int maxBatch() const;
    int keypointCapacity() const;

  private:
    std::shared_ptr<OrbB200Context> ctx;
    std::vector<mslam_b200_keypoint> kpBuf;
    std::vector<std::uint8_t> descBuf;
};

// Keypoints come back in the reference's order and Keypoint::id = index (distributed_cv_feature.cpp:1201-1210).
class DistributedOrbB200Detector : public IOrbFeatureDetector
{
  public:
    DistributedOrbB200Detector();
    explicit DistributedOrbB200Detector(const OrbB200Options& options);
    ~DistributedOrbB200Detector() override;
    DistributedOrbB200Detector(const DistributedOrbB200Detector&) = delete;
    DistributedOrbB200Detector& operator=(const DistributedOrbB200Detector&) = delete;
    std::vector<OrbKeypoint> detect(const RgbFrame& sensorData) override;

  private:
    mslam_b200_dist* handle = nullptr;
    int capacity = 0;
    bool reported = false;
    std::vector<mslam_b200_keypoint> kpBuf;
    std::vector<std::uint8_t> descBuf;
};

class OrbB200Matcher : public IOrbMatcher
{
  public:
    OrbB200Matcher();
    explicit OrbB200Matcher(const OrbB200Options& options);
    ~OrbB200Matcher() override;
    std::vector<DescriptorMatch> match(const std::vector<OrbKeypoint>& firstDescriptors,
                                       const std::vector<OrbKeypoint>& secondDescriptors) override;
    // match(first = frame i of `first`, second = frame i of `second`) for every frame of two batched detects, in one call:
    // the descriptor blocks go to the device as they are.  (*out)[i] is what match() would return for frame i.
    bool matchBatch(const OrbB200BatchResult& first, const OrbB200BatchResult& second, std::vector<std::vector<DescriptorMatch>>* out);
    // the same in two halves (one match batch in flight): Begin enqueues behind whatever the context already runs and returns,
    // End waits and converts -- a pipelined caller puts the next detectBatchBegin between them
    bool matchBatchBegin(const OrbB200BatchResult& first, const OrbB200BatchResult& second);
    bool matchBatchEnd(std::vector<std::vector<DescriptorMatch>>* out);

  private:
    std::shared_ptr<OrbB200Context> ctx;
    std::vector<std::uint64_t> fromBuf, toBuf;
    int *mA = nullptr, *mB = nullptr, *mD = nullptr, *mN = nullptr; // pinned outputs of matchBatch
    int pendingPairs = 0, pendingCap = 0;
    bool ensureBatchBuffers();
};

// ---- plugin_config.json (test/plugin_config.json:1-11 is the only schema hint in the reference) ----------------------
//   {"plugins":[{"library":"libmslam_b200_plugin.so",
//                "factoryMethods":{"featureDetector":"orbDetectorFactory","featureMatcher":"orbMatcherFactory"},
//                "parameters":{"nfeatures":1000,"nlevels":8,"scaleFactor":1.2,"fastThreshold":20,"ratio":0.7,
//                              "device":0,"maxWidth":1920,"maxHeight":1080}}]}
struct PluginEntry
{
    std::string library;
    std::string detectorFactory, matcherFactory; // factoryMethods.featureDetector / .featureMatcher
    OrbB200Options options;
};
// returns false (and fills *error) on malformed input
bool parsePluginConfig(const std::string& jsonText, std::vector<PluginEntry>* out, std::string* error);
// options used by the default constructors / factories; set from the environment variable MSLAM_B200_PLUGIN_CONFIG
// (path to a plugin_config.json) on first use, or explicitly here
void setDefaultOrbB200Options(const OrbB200Options& options);
OrbB200Options defaultOrbB200Options();

} // namespace mslam
