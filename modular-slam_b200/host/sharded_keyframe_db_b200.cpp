// sharded_keyframe_db_b200.cpp -- see the header.  Host side only: container conversion round the C ABI.
#include "sharded_keyframe_db_b200.hpp"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>

namespace mslam
{

ShardedKeyframeDbB200::ShardedKeyframeDbB200(const ShardedKeyframeDbOptions& o) : options(o)
{
    if(o.world < 1 || o.rank < 0 || o.rank >= o.world || o.maxDescriptors < 2)
        throw std::invalid_argument("ShardedKeyframeDbB200: bad rank / world / maxDescriptors");
    mslam_b200_config cfg;
    mslam_b200_default_config(&cfg);
    cfg.device = o.device;
    cfg.max_width = cfg.max_height = 64; // the detector side of the context is not used here
    cfg.max_batch = 1;
    cfg.nfeatures = 64;
    cfg.max_descriptors = o.maxDescriptors;
    cfg.ratio = o.ratio;
    int rc = mslam_b200_create(&cfg, &handle);
    if(rc != MSLAM_B200_OK)
        throw std::runtime_error(std::string("mslam_b200_create failed: ") + mslam_b200_status_string(rc) + " (no CPU fallback)");
    if(o.world > 1)
    {
        rc = o.ncclComm ? mslam_b200_db_attach_comm(handle, o.ncclComm, o.rank, o.world)
                        : mslam_b200_db_attach(handle, o.rank, o.world, o.ncclUniqueId.data());
        if(rc != MSLAM_B200_OK)
        {
            const std::string why = mslam_b200_last_cuda_error(handle);
            mslam_b200_destroy(handle);
            handle = nullptr;
            throw std::runtime_error("ShardedKeyframeDbB200: cannot join the NCCL communicator: " + why);
        }
    }
}

ShardedKeyframeDbB200::~ShardedKeyframeDbB200()
{
    mslam_b200_destroy(handle); // detaches (and destroys a communicator the context created itself)
}

std::array<char, MSLAM_B200_NCCL_ID_BYTES> ShardedKeyframeDbB200::createUniqueId()
{
    std::array<char, MSLAM_B200_NCCL_ID_BYTES> id{};
    if(mslam_b200_nccl_unique_id(id.data()) != MSLAM_B200_OK)
        throw std::runtime_error("ShardedKeyframeDbB200: ncclGetUniqueId failed (libnccl.so.2 not loadable?)");
    return id;
}

std::pair<std::size_t, std::size_t> ShardedKeyframeDbB200::shardRange(std::size_t n, int rank, int world)
{
    long long a = 0, b = 0;
    mslam_b200_db_shard_range((long long)n, rank, world, &a, &b);
    return {(std::size_t)a, (std::size_t)b};
}

// frames -> dense [frame][maxDescriptors][32] + counts (a frame with more descriptors than the capacity is truncated, and
// said so once: the reference's detector never returns more than it was created for)
void ShardedKeyframeDbB200::pack(const std::vector<std::vector<OrbKeypoint>>& frames)
{
    const std::size_t cap = (std::size_t)options.maxDescriptors;
    packed.assign(frames.size() * cap * MSLAM_B200_DESC_BYTES, 0);
    counts.assign(frames.size(), 0);
    for(std::size_t f = 0; f < frames.size(); ++f)
    {
        std::size_t n = frames[f].size();
        if(n > cap)
        {
            if(!reported)
                std::fprintf(stderr, "[mslam_b200 db] frame with %zu descriptors truncated to maxDescriptors = %zu\n", n, cap);
            reported = true;
            n = cap;
        }
        counts[f] = (int)n;
        std::uint8_t* dst = packed.data() + f * cap * MSLAM_B200_DESC_BYTES;
        for(std::size_t i = 0; i < n; ++i)
            std::memcpy(dst + i * MSLAM_B200_DESC_BYTES, frames[f][i].descriptor.data(), MSLAM_B200_DESC_BYTES);
    }
}

void ShardedKeyframeDbB200::loadRange(const std::vector<std::vector<OrbKeypoint>>& mine, std::size_t firstId, std::size_t total)
{
    const auto range = shardRange(total, options.rank, options.world);
    if(firstId != range.first || mine.size() != range.second - range.first)
        throw std::invalid_argument("ShardedKeyframeDbB200::loadRange: not this rank's id range");
    pack(mine);
    const int rc = mslam_b200_db_load_shard(handle, packed.data(), counts.data(), (int)mine.size(), options.maxDescriptors, (int)firstId);
    if(rc != MSLAM_B200_OK)
        throw std::runtime_error(std::string("mslam_b200_db_load_shard failed: ") + mslam_b200_status_string(rc) + " " +
                                 mslam_b200_last_cuda_error(handle));
    nLocal = mine.size();
    idBase = firstId;
}

void ShardedKeyframeDbB200::load(const std::vector<std::vector<OrbKeypoint>>& keyframes)
{
    const auto range = shardRange(keyframes.size(), options.rank, options.world);
    const std::vector<std::vector<OrbKeypoint>> mine(keyframes.begin() + (std::ptrdiff_t)range.first,
                                                     keyframes.begin() + (std::ptrdiff_t)range.second);
    loadRange(mine, range.first, keyframes.size());
}

std::vector<std::vector<LoopCandidate>> ShardedKeyframeDbB200::query(const std::vector<std::vector<OrbKeypoint>>& queryFrames, int k)
{
    std::vector<std::vector<LoopCandidate>> result;
    if(queryFrames.empty() || k < 1)
        return result;
    pack(queryFrames);
    outScore.assign(queryFrames.size() * (std::size_t)k, -1);
    outId.assign(queryFrames.size() * (std::size_t)k, -1);
    const int rc = mslam_b200_db_query_loaded(handle, packed.data(), counts.data(), (int)queryFrames.size(), options.maxDescriptors, k,
                                              outScore.data(), outId.data());
    if(rc != MSLAM_B200_OK)
    {
        if(!reported)
            std::fprintf(stderr, "[mslam_b200 db] query: %s %s\n", mslam_b200_status_string(rc), mslam_b200_last_cuda_error(handle));
        reported = true;
        return result;
    }
    result.resize(queryFrames.size());
    for(std::size_t q = 0; q < queryFrames.size(); ++q)
        for(int r = 0; r < k; ++r)
        {
            const std::size_t at = q * (std::size_t)k + (std::size_t)r;
            if(outId[at] < 0)
                break; // (-1, -1) pads: the database holds fewer than k keyframes
            result[q].push_back(LoopCandidate{(Id)outId[at], outScore[at]});
        }
    return result;
}

} // namespace mslam
