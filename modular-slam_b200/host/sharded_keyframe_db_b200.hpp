// sharded_keyframe_db_b200.hpp -- the keyframe-descriptor database of loop-candidate verification, one shard per GPU.
//
// The reference only sketches this step (rgbd_feature_frontend.cpp:495-534, commented out: for every candidate keyframe run
// matchLandmarks, score what survives, keep the best; orb_relocalizer.cpp:32-50 loads a DBoW3 vocabulary and stops).  The
// build definition:
//     score(Q, K) = | OrbOpenCvMatcher::match(from = Q, to = K) |          (orb_feature.cpp:84-117 semantics, ratio 0.7)
//     per query frame: the k keyframes with the best (score desc, keyframe id asc)
//
// One process per GPU.  Every process constructs the class with its (rank, world) and the same 128-byte NCCL id (rank 0 makes
// it with createUniqueId() and passes it on by whatever channel the application has), loads the keyframes -- each rank keeps
// the contiguous id range shardRange() gives it, resident in its GPU's memory -- and calls query() with the same query
// frames; all ranks return the same global answer, bit-identical to a single-GPU database.  Underneath:
// mslam_b200_db_attach / _db_load_shard / _db_query_loaded of include/mslam_b200.h (score on the tensor cores -> local top-k
// -> ncclAllGather -> merge, one stream).  Constructors throw std::runtime_error without a CUDA device, or without NCCL when
// world > 1: there is no fallback.
#pragma once
#ifdef MSLAM_B200_WITH_REFERENCE_HEADERS
#include "modular_slam/orb_feature.hpp"
#else
#include "mslam_shim.hpp"
#endif
#include <array>
#include <cstddef>
#include <cstdint>
#include <utility>
#include <vector>

#include "../../include/mslam_b200.h"

namespace mslam
{

struct ShardedKeyframeDbOptions
{
    int device = 0;
    int rank = 0, world = 1;
    std::array<char, MSLAM_B200_NCCL_ID_BYTES> ncclUniqueId{}; // world > 1: the id every rank shares ...
    void* ncclComm = nullptr;                                  // ... or a communicator (ncclComm_t) the application already has
    int maxDescriptors = 1000;                                 // descriptors per frame (cv::ORB::create(1000), orb_feature.cpp:25)
    double ratio = 0.7;                                        // orb_feature.cpp:101
};

struct LoopCandidate
{
    Id keyframeId;     // global keyframe id (index in the list given to load())
    std::int32_t score; // number of ratio-test matches
};

class ShardedKeyframeDbB200
{
  public:
    explicit ShardedKeyframeDbB200(const ShardedKeyframeDbOptions& options);
    ~ShardedKeyframeDbB200();
    ShardedKeyframeDbB200(const ShardedKeyframeDbB200&) = delete;
    ShardedKeyframeDbB200& operator=(const ShardedKeyframeDbB200&) = delete;

    static std::array<char, MSLAM_B200_NCCL_ID_BYTES> createUniqueId();
    // [first, last) of the n keyframe ids that `rank` holds: the first n % world ranks hold one more
    static std::pair<std::size_t, std::size_t> shardRange(std::size_t n, int rank, int world);

    // All keyframes of the map, in id order (id = index); this rank uploads its own range only.  Replaces the previous content.
    void load(const std::vector<std::vector<OrbKeypoint>>& keyframes);
    // Same, from this rank's range alone: `mine` = keyframes [firstId, firstId + mine.size()) of a database of `total`.
    void loadRange(const std::vector<std::vector<OrbKeypoint>>& mine, std::size_t firstId, std::size_t total);
    std::size_t localSize() const { return nLocal; }
    std::size_t firstLocalId() const { return idBase; }

    // top-k candidates per query frame (fewer when the whole database is smaller).  Collective: every rank calls it with the
    // same frames.  An empty vector means the call failed (logged once), as the reference's interfaces report "nothing".
    std::vector<std::vector<LoopCandidate>> query(const std::vector<std::vector<OrbKeypoint>>& queryFrames, int k);

  private:
    ShardedKeyframeDbOptions options;
    mslam_b200_ctx* handle = nullptr;
    std::size_t nLocal = 0, idBase = 0;
    std::vector<std::uint8_t> packed; // dense [frame][maxDescriptors][32] staging
    std::vector<int> counts, outScore, outId;
    bool reported = false;

    void pack(const std::vector<std::vector<OrbKeypoint>>& frames);
};

} // namespace mslam
