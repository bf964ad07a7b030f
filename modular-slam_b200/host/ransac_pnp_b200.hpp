// ransac_pnp_b200.hpp -- SURVEY 8(f) row f4: the reference's PnP interface on the B200 C ABI.
//
//   mslam::RansacPnpB200  replaces  mslam::OpenCvRansacPnp   (cv_ransac_pnp.hpp:10-17, cv_ransac_pnp.cpp:15-85)
//
// Constructed like the reference's (std::make_shared<mslam::OpenCvRansacPnp>(), rgbd_slam.cpp / viewer.cpp) or through the
// boost::dll alias `ransacPnpFactory : std::unique_ptr<mslam::IPnpAlgorithm<slam3d::SensorState, Vector3>>()`.
// solvePnp() narrows landmarks / pixels / camera matrix to float exactly as the reference does (:20-41,:52-53), runs
// mslam_b200_pnp_ransac_batch with the reference's constants (100 hypotheses, 5 px, :56-57) and returns the pose the
// reference returns (orientation = R^-1, position = -(R^-1 t), :65-77) plus one inlier bit per landmark (:79-82);
// std::nullopt where the reference returns it (no model) and for fewer than 4 points (where cv::solvePnPRansac throws).
// `initial` only seeds OpenCV's final fit in the reference; here the fit starts from the winning hypothesis.
#pragma once
#ifdef MSLAM_B200_WITH_REFERENCE_HEADERS
#include "modular_slam/pnp.hpp"
#include "modular_slam/types/slam3d_types.hpp"
#else
#include "mslam_shim.hpp"
#endif
#include <memory>
#include <optional>
#include <vector>

#include "../../include/mslam_b200.h"

namespace mslam
{
using IRansacPnp = IPnpAlgorithm<slam3d::SensorState, Vector3>;

class RansacPnpB200 : public IRansacPnp
{
  public:
    explicit RansacPnpB200(int device = 0, unsigned seed = 0);
    ~RansacPnpB200() override;
    RansacPnpB200(const RansacPnpB200&) = delete;
    RansacPnpB200& operator=(const RansacPnpB200&) = delete;

    std::optional<PnpResult> solvePnp(const std::vector<std::shared_ptr<Landmark<Vector3>>>& landmarks,
                                      const std::vector<Vector2>& sensorPoints,
                                      const slam3d::SensorState& initial = slam3d::SensorState()) override;
    // several frames in one launch (independent sequences); element i == solvePnp(landmarks[i], sensorPoints[i])
    // for the same call counter
    std::vector<std::optional<PnpResult>>
    solvePnpBatch(const std::vector<const std::vector<std::shared_ptr<Landmark<Vector3>>>*>& landmarks,
                  const std::vector<const std::vector<Vector2>*>& sensorPoints);

  private:
    mslam_b200_ctx* handle = nullptr;
    unsigned seed = 0;
    int calls = 0; // frame counter: hypothesis draws differ from call to call, repeatably
    bool reported = false;
    std::vector<float> objBuf, imgBuf;
    std::vector<std::uint8_t> inlBuf;
};
} // namespace mslam
