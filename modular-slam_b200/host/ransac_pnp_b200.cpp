// ransac_pnp_b200.cpp -- see ransac_pnp_b200.hpp
#include "ransac_pnp_b200.hpp"

#include <algorithm>
#include <cstdio>
#include <stdexcept>
#include <string>

namespace mslam
{
RansacPnpB200::RansacPnpB200(int device, unsigned seedValue) : seed(seedValue)
{
    mslam_b200_config cfg;
    mslam_b200_default_config(&cfg);
    cfg.device = device;
    cfg.max_width = cfg.max_height = 64; // the PnP entry points use none of the image buffers
    cfg.max_batch = 1;
    cfg.nfeatures = 64;
    const int rc = mslam_b200_create(&cfg, &handle);
    if(rc != MSLAM_B200_OK)
        throw std::runtime_error(std::string("mslam_b200_create failed: ") + mslam_b200_status_string(rc) +
                                 " (the B200 plugin has no CPU fallback)");
}

RansacPnpB200::~RansacPnpB200()
{
    mslam_b200_destroy(handle);
}

std::vector<std::optional<IRansacPnp::PnpResult>>
RansacPnpB200::solvePnpBatch(const std::vector<const std::vector<std::shared_ptr<Landmark<Vector3>>>*>& landmarks,
                             const std::vector<const std::vector<Vector2>*>& sensorPoints)
{
    const std::size_t F = landmarks.size();
    std::vector<std::optional<PnpResult>> out(F);
    if(F == 0 || sensorPoints.size() != F)
        return out;
    std::size_t cap = 1;
    std::vector<int> counts(F);
    for(std::size_t f = 0; f < F; ++f)
    {
        const std::size_t n = std::min(landmarks[f]->size(), sensorPoints[f]->size()); // the reference asserts equality (:19)
        counts[f] = (int)n;
        cap = std::max(cap, n);
    }
    objBuf.assign(F * cap * 3, 0.f);
    imgBuf.assign(F * cap * 2, 0.f);
    inlBuf.assign(F * cap, 0);
    for(std::size_t f = 0; f < F; ++f)
        for(int i = 0; i < counts[f]; ++i)
        {
            const auto& lm = (*landmarks[f])[(std::size_t)i];
            float* o = &objBuf[(f * cap + (std::size_t)i) * 3];
            o[0] = static_cast<float>(lm->state.x()); // cv_ransac_pnp.cpp:27-29
            o[1] = static_cast<float>(lm->state.y());
            o[2] = static_cast<float>(lm->state.z());
            const Vector2& p = (*sensorPoints[f])[(std::size_t)i];
            float* q = &imgBuf[(f * cap + (std::size_t)i) * 2];
            q[0] = static_cast<float>(p.x()); // :38-39
            q[1] = static_cast<float>(p.y());
        }
    mslam_b200_pnp_params params;
    mslam_b200_pnp_default_params(&params); // 100 hypotheses, 5 px, 0.99 (:56-57)
    params.fx = cameraParams.focal.x();     // :52-53
    params.fy = cameraParams.focal.y();
    params.cx = cameraParams.principalPoint.x();
    params.cy = cameraParams.principalPoint.y();
    params.seed = seed;
    params.frame_base = calls;
    calls += (int)F;
    std::vector<double> rt(F * 6), pose(F * 7);
    std::vector<int> info(F * 4);
    const int rc = mslam_b200_pnp_ransac_batch(handle, objBuf.data(), imgBuf.data(), counts.data(), (int)F, (int)cap, &params,
                                               rt.data(), pose.data(), inlBuf.data(), info.data());
    if(rc != MSLAM_B200_OK)
    {
        if(!reported)
            std::fprintf(stderr, "[mslam_b200 plugin] pnp_ransac_batch: %s %s\n", mslam_b200_status_string(rc),
                         mslam_b200_last_cuda_error(handle));
        reported = true;
        return out;
    }
    for(std::size_t f = 0; f < F; ++f)
    {
        if(!info[4 * f])
            continue; // "Didnt find pnp solution" (:59-63)
        PnpResult r;
        const double* p = &pose[7 * f];
#ifdef MSLAM_B200_WITH_REFERENCE_HEADERS
        r.pose.orientation = Quaternion(p[0], p[1], p[2], p[3]);
        r.pose.position = Vector3(p[4], p[5], p[6]);
#else
        r.pose.orientation.c[0] = p[1], r.pose.orientation.c[1] = p[2], r.pose.orientation.c[2] = p[3], r.pose.orientation.c[3] = p[0];
        r.pose.position = Vector3{{p[4], p[5], p[6]}};
#endif
        r.inliers.resize(landmarks[f]->size(), false); // :79-82
        for(int i = 0; i < counts[f]; ++i)
            if(inlBuf[f * cap + (std::size_t)i])
                r.inliers[(std::size_t)i] = true;
        out[f] = std::move(r);
    }
    return out;
}

std::optional<IRansacPnp::PnpResult> RansacPnpB200::solvePnp(const std::vector<std::shared_ptr<Landmark<Vector3>>>& landmarks,
                                                             const std::vector<Vector2>& sensorPoints,
                                                             const slam3d::SensorState& /*initial*/)
{
    return std::move(solvePnpBatch({&landmarks}, {&sensorPoints})[0]);
}

std::unique_ptr<IRansacPnp> createRansacPnpB200()
{
    return std::unique_ptr<IRansacPnp>(new RansacPnpB200());
}
} // namespace mslam
