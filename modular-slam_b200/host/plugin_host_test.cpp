// plugin_host_test.cpp -- a stand-in for the reference's side of the plugin boundary, used by tests/:
// loads the plugin the way mslam::loadFactoryMethod does (plugin_loader.hpp:19-25: boost::dll::import_alias reads the
// exported alias variable and calls the function it points to), then drives IFeatureDetector::detect /
// IFeatureMatcher::match exactly as RgbdFeatureFrontend does (rgbd_feature_frontend.cpp:187,237).
//
//   plugin_host_test symbols <plugin.so>
//   plugin_host_test config  <plugin_config.json>
//   plugin_host_test provider <associations.txt> <out.bin> [batch size]   (RgbdFileB200Provider; batch > 0 => fetchBatch, pinned)
//   plugin_host_test pnp     <plugin.so> <in.bin> <out.bin>   (RansacPnpB200 through the ransacPnpFactory alias; in.bin = i64 n,
//                            fx fy cx cy f64, n x 3 f64 landmarks, n x 2 f64 pixels; called twice: solvePnp, then solvePnpBatch)
//   plugin_host_test db      <in.bin> <out.bin> [rank world idfile]   (ShardedKeyframeDbB200; in.bin = i32 K, n, Q, k, then
//                            K x n x 32 keyframe descriptor bytes, Q x n x 32 query bytes; out.bin = Q x k x (i64 id, i32 score,
//                            i32 pad); with world > 1 each process is one rank on device `rank`, rank 0 writes the NCCL id to idfile)
//   plugin_host_test bench   <frames.raw> <W> <H> <B> <T> <iters>   (C++ timing loops through the plugin classes; frames.raw =
//                            (T + 1) batches x B frames x H x W x 3 bytes, batch t = frame t of B sequences.  Prints one JSON
//                            object: per-frame detect + match of one sequence -- what RgbdFeatureFrontend does,
//                            rgbd_feature_frontend.cpp:187,237 -- and B-frame pipelined batches from pinned memory)
//   plugin_host_test run     <plugin.so> <frame0.raw> <frame1.raw> <W> <H> <out.bin> [detector alias]
#include <dlfcn.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>

#include "orb_b200_plugin.hpp"
#include "ransac_pnp_b200.hpp"
#include "rgbd_file_b200_provider.hpp"
#include "sharded_keyframe_db_b200.hpp"

#include <chrono>
#include <thread>

template <typename BlockType>
using BlockFactoryCreator = std::unique_ptr<BlockType>(); // plugin_loader.hpp:13-14

template <typename BlockType>
static BlockFactoryCreator<BlockType>* loadFactoryMethod(void* lib, const char* name)
{
    // import_alias<T>(lib, name): the symbol is a variable holding the function's address
    void* sym = dlsym(lib, name);
    if(!sym)
        return nullptr;
    return *reinterpret_cast<BlockFactoryCreator<BlockType>**>(sym);
}

static bool readFile(const char* path, std::vector<std::uint8_t>* out)
{
    std::ifstream in(path, std::ios::binary);
    if(!in)
        return false;
    out->assign(std::istreambuf_iterator<char>(in), std::istreambuf_iterator<char>());
    return true;
}

int main(int argc, char** argv)
{
    if(argc >= 3 && !std::strcmp(argv[1], "config"))
    {
        std::ifstream in(argv[2]);
        std::stringstream ss;
        ss << in.rdbuf();
        std::vector<mslam::PluginEntry> entries;
        std::string err;
        if(!in || !mslam::parsePluginConfig(ss.str(), &entries, &err))
        {
            std::printf("error: %s\n", err.c_str());
            return 2;
        }
        for(const auto& e : entries)
            std::printf("library=%s detector=%s matcher=%s nfeatures=%d nlevels=%d scaleFactor=%.3f fastThreshold=%d ratio=%.3f\n",
                        e.library.c_str(), e.detectorFactory.c_str(), e.matcherFactory.c_str(), e.options.nfeatures,
                        e.options.nlevels, e.options.scaleFactor, e.options.fastThreshold, e.options.ratio);
        return 0;
    }
    if(argc >= 8 && !std::strcmp(argv[1], "bench"))
    {
        const int W = std::atoi(argv[3]), H = std::atoi(argv[4]), B = std::atoi(argv[5]), T = std::atoi(argv[6]), iters = std::atoi(argv[7]);
        const size_t frameBytes = (size_t)W * H * 3, batchBytes = frameBytes * (size_t)B;
        std::vector<std::uint8_t> raw;
        if(W < 1 || H < 1 || B < 1 || T < 1 || iters < 1 || !readFile(argv[2], &raw) || raw.size() < batchBytes * (size_t)(T + 1))
            return 5;
        try
        {
            using clock = std::chrono::steady_clock;
            mslam::OrbB200Options o;
            o.maxWidth = W, o.maxHeight = H, o.maxBatch = B;
            mslam::OrbB200Detector detector(o);
            mslam::OrbB200Matcher matcher(o);
            // ---- (1) one frame at a time, as the reference's front-end calls the interfaces ----
            std::vector<mslam::RgbFrame> seq((size_t)T + 1);
            for(int t = 0; t <= T; ++t)
            {
                seq[(size_t)t].size = mslam::Size{W, H};
                seq[(size_t)t].data.assign(raw.data() + batchBytes * (size_t)t, raw.data() + batchBytes * (size_t)t + frameBytes);
            }
            std::vector<mslam::OrbKeypoint> prev = detector.detect(seq[0]), cur;
            size_t matches = 0, keypoints = 0;
            const int singleIters = iters * 8;
            double singleMs = 0;
            for(int pass = 0; pass < 2; ++pass) // pass 0 warms up (kernel graph capture, first-touch)
            {
                const auto t0 = clock::now();
                for(int i = 0; i < (pass ? singleIters : 16); ++i)
                {
                    cur = detector.detect(seq[(size_t)(1 + i % T)]);
                    const std::vector<mslam::DescriptorMatch> m = matcher.match(cur, prev); // rgbd_feature_frontend.cpp:237
                    matches = m.size(), keypoints = cur.size();
                    prev.swap(cur);
                }
                singleMs = std::chrono::duration<double, std::milli>(clock::now() - t0).count();
            }
            // ---- (2) B-frame batches from pinned memory, two detect batches in flight, match between them ----
            void* pinned = nullptr;
            if(mslam_b200_host_alloc(&pinned, batchBytes * (size_t)(T + 1)) != MSLAM_B200_OK)
                return 7;
            std::memcpy(pinned, raw.data(), batchBytes * (size_t)(T + 1));
            std::vector<std::vector<const std::uint8_t*>> ptrs((size_t)T + 1, std::vector<const std::uint8_t*>((size_t)B));
            for(int t = 0; t <= T; ++t)
                for(int b = 0; b < B; ++b)
                    ptrs[(size_t)t][(size_t)b] = static_cast<const std::uint8_t*>(pinned) + batchBytes * (size_t)t + frameBytes * (size_t)b;
            std::vector<std::unique_ptr<mslam::OrbB200BatchResult>> sets;
            for(int i = 0; i < 4; ++i)
                sets.emplace_back(new mslam::OrbB200BatchResult(B, detector.keypointCapacity()));
            std::vector<std::vector<mslam::DescriptorMatch>> bm;
            size_t batchMatches = 0, batchKeypoints = 0;
            bool ok = detector.detectBatchBegin(ptrs[0].data(), B, W, H, sets[3].get()) && detector.detectBatchEnd(); // "previous"
            long long step = 0;
            auto runPipelined = [&](int n) {
                const long long base = step;
                for(long long j = base; j < base + std::min(2, n); ++j)
                    ok = ok && detector.detectBatchBegin(ptrs[(size_t)(1 + j % T)].data(), B, W, H, sets[(size_t)(j % 4)].get());
                for(long long i = base; i < base + n; ++i)
                {
                    ok = ok && detector.detectBatchEnd();
                    const mslam::OrbB200BatchResult &c = *sets[(size_t)(i % 4)], &p = *sets[(size_t)((i + 3) % 4)];
                    ok = ok && matcher.matchBatchBegin(c, p);
                    if(i + 2 < base + n)
                        ok = ok && detector.detectBatchBegin(ptrs[(size_t)(1 + (i + 2) % T)].data(), B, W, H, sets[(size_t)((i + 2) % 4)].get());
                    ok = ok && matcher.matchBatchEnd(&bm);
                    batchMatches = 0, batchKeypoints = 0;
                    for(int b = 0; ok && b < B; ++b)
                        batchMatches += bm[(size_t)b].size(), batchKeypoints += (size_t)c.count(b);
                }
                step = base + n;
            };
            runPipelined(3);
            const auto t1 = clock::now();
            runPipelined(iters);
            const double batchMs = std::chrono::duration<double, std::milli>(clock::now() - t1).count();
            // a converted frame of the batched path equals the per-frame call on the same image
            const int lastT = (int)(1 + (step - 1) % T);
            const std::vector<mslam::OrbKeypoint> viaBatch = sets[(size_t)((step - 1) % 4)]->keypoints(0), viaSingle = detector.detect(seq[(size_t)lastT]);
            bool same = viaBatch.size() == viaSingle.size();
            for(size_t i = 0; same && i < viaBatch.size(); ++i)
                same = viaBatch[i].keypoint.coordinates[0] == viaSingle[i].keypoint.coordinates[0] &&
                       viaBatch[i].keypoint.coordinates[1] == viaSingle[i].keypoint.coordinates[1] && viaBatch[i].descriptor == viaSingle[i].descriptor;
            mslam_b200_host_free(pinned);
            if(!ok)
                return 8;
            std::printf("{\"plugin_single_frame\": {\"value\": %.2f, \"unit\": \"frames/s\", \"ms_per_frame\": %.4f, \"frames\": %d, "
                        "\"keypoints\": %zu, \"matches\": %zu, \"call\": \"OrbB200Detector::detect + OrbB200Matcher::match per frame, pageable RgbFrame\"}, "
                        "\"plugin_batched\": {\"value\": %.2f, \"unit\": \"frames/s\", \"ms_per_step\": %.4f, \"batch\": %d, \"steps\": %d, "
                        "\"keypoints_per_batch\": %zu, \"matches_per_batch\": %zu, \"batched_equals_single\": %s, "
                        "\"call\": \"detectBatchBegin/End (two batches in flight, pinned frames) + matchBatchBegin/End\"}}\n",
                        singleIters / (singleMs * 1e-3), singleMs / singleIters, singleIters, keypoints, matches,
                        (double)B * iters / (batchMs * 1e-3), batchMs / iters, B, iters, batchKeypoints, batchMatches, same ? "true" : "false");
        }
        catch(const std::exception& e)
        {
            std::fprintf(stderr, "bench failed: %s\n", e.what());
            return 6;
        }
        return 0;
    }
    if(argc >= 4 && !std::strcmp(argv[1], "db"))
    {
        std::vector<std::uint8_t> raw;
        if(!readFile(argv[2], &raw) || raw.size() < 16)
            return 5;
        int hdr[4];
        std::memcpy(hdr, raw.data(), 16);
        const int K = hdr[0], n = hdr[1], Q = hdr[2], k = hdr[3];
        if(raw.size() < 16 + ((size_t)K + Q) * n * 32)
            return 5;
        auto frames = [&](const std::uint8_t* p, int count) {
            std::vector<std::vector<mslam::OrbKeypoint>> v((size_t)count);
            for(int f = 0; f < count; ++f)
            {
                v[(size_t)f].resize((size_t)n);
                for(int i = 0; i < n; ++i)
                    std::memcpy(v[(size_t)f][(size_t)i].descriptor.data(), p + ((size_t)f * n + i) * 32, 32);
            }
            return v;
        };
        try
        {
            mslam::ShardedKeyframeDbOptions o;
            o.maxDescriptors = n;
            if(argc >= 7)
            {
                o.rank = std::atoi(argv[4]), o.world = std::atoi(argv[5]), o.device = o.rank;
                if(o.world > 1 && o.rank == 0)
                {
                    o.ncclUniqueId = mslam::ShardedKeyframeDbB200::createUniqueId();
                    std::ofstream tmp(std::string(argv[6]) + ".tmp", std::ios::binary);
                    tmp.write(o.ncclUniqueId.data(), (std::streamsize)o.ncclUniqueId.size());
                    tmp.close();
                    std::rename((std::string(argv[6]) + ".tmp").c_str(), argv[6]);
                }
                else if(o.world > 1)
                {
                    std::vector<std::uint8_t> id;
                    for(int tries = 0; tries < 600 && (!readFile(argv[6], &id) || id.size() != o.ncclUniqueId.size()); ++tries)
                        std::this_thread::sleep_for(std::chrono::milliseconds(100));
                    if(id.size() != o.ncclUniqueId.size())
                        return 10;
                    std::memcpy(o.ncclUniqueId.data(), id.data(), id.size());
                }
            }
            mslam::ShardedKeyframeDbB200 db(o);
            db.load(frames(raw.data() + 16, K));
            const auto queries = frames(raw.data() + 16 + (size_t)K * n * 32, Q);
            auto result = db.query(queries, k);
            result = db.query(queries, k); // a second query reuses the resident shard and the communicator
            if(result.size() != (size_t)Q)
                return 11;
            std::ofstream out(argv[3], std::ios::binary);
            for(int q = 0; q < Q; ++q)
                for(int r = 0; r < k; ++r)
                {
                    const bool have = (size_t)r < result[(size_t)q].size();
                    const long long id = have ? (long long)result[(size_t)q][(size_t)r].keyframeId : -1;
                    const int rec[2] = {have ? result[(size_t)q][(size_t)r].score : -1, 0};
                    out.write(reinterpret_cast<const char*>(&id), 8);
                    out.write(reinterpret_cast<const char*>(rec), 8);
                }
            std::printf("db rank %d/%d: %zu local keyframes from id %zu, %d query frames\n", o.rank, o.world, db.localSize(),
                        db.firstLocalId(), Q);
        }
        catch(const std::exception& e)
        {
            std::fprintf(stderr, "db failed: %s\n", e.what());
            return 6;
        }
        return 0;
    }
    if(argc >= 4 && !std::strcmp(argv[1], "provider"))
    {
        // the way the reference apps build their provider (rgbd_slam.cpp:66-70): readTumRgbdDataset + tumRgbdCameraParams
        mslam::RgbdFileB200Provider provider(mslam::readTumRgbdDataset(argv[2]), mslam::tumRgbdCameraParams());
        if(!provider.init())
        {
            std::printf("init failed\n");
            return 7;
        }
        const int batch = argc >= 5 ? std::atoi(argv[4]) : 0;
        std::ofstream out(argv[3], std::ios::binary);
        int frames = 0;
        auto dump = [&out](int w, int h, double ts, const std::uint8_t* rgb, const std::uint16_t* depth, float factor) {
            const double hdr[4] = {(double)w, (double)h, ts, (double)factor};
            out.write(reinterpret_cast<const char*>(hdr), sizeof(hdr));
            out.write(reinterpret_cast<const char*>(rgb), (std::streamsize)((size_t)w * h * 3));
            out.write(reinterpret_cast<const char*>(depth), (std::streamsize)((size_t)w * h * 2));
        };
        if(batch > 0)
        {
            for(;;)
            {
                const mslam::RgbdFileB200Provider::Batch b = provider.fetchBatch(batch);
                if(b.count == 0)
                    break;
                for(int i = 0; i < b.count; ++i, ++frames)
                    dump(b.width, b.height, b.timestamps[(size_t)i], b.rgb[(size_t)i], b.depth[(size_t)i], mslam::tumRgbdCameraParams().factor);
            }
        }
        else
        {
            while(provider.fetch())
            {
                const std::shared_ptr<mslam::RgbdFrame> f = provider.recentData();
                if(f->depth.size.width != f->rgb.size.width || f->depth.size.height != f->rgb.size.height)
                    return 8;
                dump(f->rgb.size.width, f->rgb.size.height, f->timestamp.timePoint, f->rgb.data.data(), f->depth.data.data(),
                     f->depth.cameraParameters.factor);
                ++frames;
            }
            if(provider.recentData() != nullptr)
                return 9; // the reference resets recentFrame when the sequence ends (rgbd_file_provider.cpp:57-61)
        }
        std::printf("frames=%d position=%zu\n", frames, provider.position());
        return 0;
    }
    if(argc < 3)
    {
        std::fprintf(stderr, "usage: see the header of plugin_host_test.cpp\n");
        return 64;
    }
    void* lib = dlopen(argv[2], RTLD_NOW | RTLD_LOCAL);
    if(!lib)
    {
        std::fprintf(stderr, "dlopen: %s\n", dlerror());
        return 3;
    }
    if(!std::strcmp(argv[1], "pnp") && argc >= 5)
    {
        auto* makePnp = loadFactoryMethod<mslam::IRansacPnp>(lib, "ransacPnpFactory");
        if(!makePnp)
            return 4;
        std::vector<std::uint8_t> raw;
        if(!readFile(argv[3], &raw) || raw.size() < 40)
            return 5;
        long long n = 0;
        std::memcpy(&n, raw.data(), 8);
        const double* d = reinterpret_cast<const double*>(raw.data() + 8);
        if(raw.size() < 40 + (size_t)n * 40)
            return 5;
        try
        {
            std::shared_ptr<mslam::IRansacPnp> pnp = makePnp();
            mslam::CameraParameters cam{};
            cam.focal.x() = d[0], cam.focal.y() = d[1], cam.principalPoint.x() = d[2], cam.principalPoint.y() = d[3];
            cam.factor = 1.f;
            pnp->setCameraParameters(cam); // rgbd_feature_frontend.cpp:345
            std::vector<std::shared_ptr<mslam::Landmark<mslam::Vector3>>> landmarks;
            std::vector<mslam::Vector2> points((size_t)n);
            for(long long i = 0; i < n; ++i)
            {
                auto lm = std::make_shared<mslam::Landmark<mslam::Vector3>>();
                lm->id = (mslam::Id)i;
                lm->state = mslam::Vector3{{d[4 + 3 * i], d[4 + 3 * i + 1], d[4 + 3 * i + 2]}};
                landmarks.push_back(lm);
                points[(size_t)i].x() = d[4 + 3 * n + 2 * i];
                points[(size_t)i].y() = d[4 + 3 * n + 2 * i + 1];
            }
            const auto result = pnp->solvePnp(landmarks, points); // :347
            const auto none = pnp->solvePnp({}, {});
            std::ofstream out(argv[4], std::ios::binary);
            const std::uint64_t hdr[2] = {result ? 1u : 0u, none ? 1u : 0u};
            out.write(reinterpret_cast<const char*>(hdr), sizeof(hdr));
            if(result)
            {
                const double pose[7] = {result->pose.position[0], result->pose.position[1], result->pose.position[2],
                                        result->pose.orientation.x(), result->pose.orientation.y(), result->pose.orientation.z(),
                                        result->pose.orientation.w()};
                out.write(reinterpret_cast<const char*>(pose), sizeof(pose));
                for(size_t i = 0; i < result->inliers.size(); ++i)
                    out.put(result->inliers[i] ? 1 : 0);
                std::printf("pnp inliers %zu / %zu\n", result->inliers.count(), landmarks.size()); // :75
            }
            else
                std::printf("Didnt find pnp solution\n");
        }
        catch(const std::exception& e)
        {
            std::fprintf(stderr, "plugin failed: %s\n", e.what());
            return 6;
        }
        return 0;
    }
    const char* detectorAlias = (argc >= 9 && !std::strcmp(argv[1], "run")) ? argv[8] : "orbDetectorFactory";
    auto* makeDetector = loadFactoryMethod<mslam::IOrbFeatureDetector>(lib, detectorAlias);
    auto* makeMatcher = loadFactoryMethod<mslam::IOrbMatcher>(lib, "orbMatcherFactory");
    if(!makeDetector || !makeMatcher || !loadFactoryMethod<mslam::IOrbFeatureDetector>(lib, "distributedOrbDetectorFactory") ||
       !loadFactoryMethod<mslam::IRansacPnp>(lib, "ransacPnpFactory"))
    {
        std::fprintf(stderr, "factory aliases missing\n");
        return 4;
    }
    if(!std::strcmp(argv[1], "symbols"))
    {
        std::printf("aliases ok: orbDetectorFactory orbMatcherFactory distributedOrbDetectorFactory ransacPnpFactory\n");
        return 0;
    }
    if(std::strcmp(argv[1], "run") || argc < 8)
        return 64;
    mslam::RgbFrame f0, f1;
    f0.size = f1.size = mslam::Size{std::atoi(argv[5]), std::atoi(argv[6])};
    if(!readFile(argv[3], &f0.data) || !readFile(argv[4], &f1.data))
        return 5;
    try
    {
        std::shared_ptr<mslam::IOrbFeatureDetector> detector = makeDetector(); // unique_ptr -> shared_ptr, as the apps hold them
        std::shared_ptr<mslam::IOrbMatcher> matcher = makeMatcher();
        const std::vector<mslam::OrbKeypoint> k0 = detector->detect(f0);
        const std::vector<mslam::OrbKeypoint> k1 = detector->detect(f1);
        const std::vector<mslam::DescriptorMatch> m = matcher->match(k0, k1);
        const std::vector<mslam::DescriptorMatch> none = matcher->match({}, k1);
        std::ofstream out(argv[7], std::ios::binary);
        const std::uint64_t hdr[4] = {k0.size(), k1.size(), m.size(), none.size()};
        out.write(reinterpret_cast<const char*>(hdr), sizeof(hdr));
        out.write(reinterpret_cast<const char*>(k0.data()), (std::streamsize)(k0.size() * sizeof(mslam::OrbKeypoint)));
        out.write(reinterpret_cast<const char*>(k1.data()), (std::streamsize)(k1.size() * sizeof(mslam::OrbKeypoint)));
        out.write(reinterpret_cast<const char*>(m.data()), (std::streamsize)(m.size() * sizeof(mslam::DescriptorMatch)));
        std::printf("detect %zu + %zu keypoints, %zu matches\n", k0.size(), k1.size(), m.size());
    }
    catch(const std::exception& e)
    {
        std::fprintf(stderr, "plugin failed: %s\n", e.what());
        return 6;
    }
    return 0;
}
