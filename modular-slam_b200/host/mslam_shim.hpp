// mslam_shim.hpp -- layout-compatible stand-ins for the reference types that cross the plugin boundary.
//
// The reference headers (modular_slam/frontend/feature/feature_interface.hpp, orb_feature.hpp, types/rgb_frame.hpp,
// types/basic_types.hpp) pull in Eigen, Boost and OpenCV, none of which exist in this build image.  When they ARE
// available, compile the adapter with -DMSLAM_B200_WITH_REFERENCE_HEADERS -I<modular_slam include dir> and this shim
// is skipped; the static_asserts in orb_b200_plugin.cpp hold for both.
//
// Layouts follow (x86-64, Eigen 3.4):
//   Id = std::uint64_t                                              types/basic_types.hpp:12
//   Vector2 = Eigen::Vector2d  -> 16 bytes, 16-byte aligned         types/basic_types.hpp:14
//   Keypoint { Id id; Vector2 coordinates; }          = 32 bytes   feature_interface.hpp:18-22
//   KeypointDescriptor<u8, 32> { Keypoint; array<u8,32> } = 64 bytes, descriptor at +32   feature_interface.hpp:24-29
//   DescriptorMatch { size_t fromIndex, toIndex; }    = 16 bytes   feature_interface.hpp:37-41
//   Size { int width, height; }                                    types/basic_types.hpp:23-27
//   RgbFrame { std::vector<uint8_t> data; Size size; }              types/rgb_frame.hpp:12-16
#pragma once
#include <array>
#include <cstddef>
#include <cstdint>
#include <memory>
#include <optional>
#include <vector>

namespace mslam
{
using Id = std::uint64_t;

// Eigen::Vector2d stand-in: two doubles, 16-byte aligned, Eigen-style accessors
struct alignas(16) Vector2
{
    double v[2];
    double& x() { return v[0]; }
    double& y() { return v[1]; }
    double x() const { return v[0]; }
    double y() const { return v[1]; }
    double& operator[](int i) { return v[i]; }
    double operator[](int i) const { return v[i]; }
};

struct Size
{
    int width;
    int height;
};

struct RgbFrame
{
    std::vector<std::uint8_t> data;
    Size size;
};

struct Keypoint
{
    Id id;
    Vector2 coordinates;
};

template <typename DescriptorType, int Length = 32>
struct KeypointDescriptor
{
    Keypoint keypoint;
    std::array<DescriptorType, Length> descriptor;
};

struct DescriptorMatch
{
    std::size_t fromIndex;
    std::size_t toIndex;
};

template <typename SensorData, typename DescriptorType, int Length>
class IFeatureDetector
{
  public:
    virtual std::vector<KeypointDescriptor<DescriptorType, Length>> detect(const SensorData& sensorData) = 0;
    virtual ~IFeatureDetector() {}
};

template <typename DescriptorType, int Length>
class IFeatureMatcher
{
  public:
    virtual std::vector<DescriptorMatch>
    match(const std::vector<KeypointDescriptor<DescriptorType, Length>>& firstDescriptors,
          const std::vector<KeypointDescriptor<DescriptorType, Length>>& secondDescriptors) = 0;
    virtual ~IFeatureMatcher() {}
};

// ---- sensor side (rgbd_file_b200_provider.hpp): types/timestamp.hpp:7-10, sensors/camera_parameters.hpp:7-12,
// types/depth_frame.hpp:12-17, types/rgbd_frame.hpp:13-19, slam_component.hpp:9-23, sensors/data_provider.hpp:10-17
struct Timestamp
{
    double timePoint;
};

struct CameraParameters
{
    Vector2 principalPoint;
    Vector2 focal;
    float factor;
};

struct DepthFrame
{
    std::vector<std::uint16_t> data;
    Size size;
    CameraParameters cameraParameters;
};

struct RgbdFrame
{
    Timestamp timestamp;
    RgbFrame rgb;
    DepthFrame depth;
};

class ParametersHandlerInterface; // parameters/parameters_handler.hpp:79-95 (not used by the provider)

class SlamComponent
{
  public:
    SlamComponent(std::shared_ptr<ParametersHandlerInterface> handler = nullptr) : parametersHandler{handler} {}
    void setParameterHandler(std::shared_ptr<ParametersHandlerInterface> newParameterHendler)
    {
        parametersHandler = std::move(newParameterHendler);
    }
    virtual bool init() { return true; }
    virtual ~SlamComponent() {}

  protected:
    std::shared_ptr<ParametersHandlerInterface> parametersHandler;
};

template <typename DataType>
class DataProviderInterface : public SlamComponent
{
  public:
    virtual bool init() = 0;
    virtual bool fetch() = 0;
    virtual std::shared_ptr<DataType> recentData() const = 0;
};

// ---- PnP side (ransac_pnp_b200.hpp): types/basic_types.hpp:16-17, types/state.hpp:7-12, types/slam3d_types.hpp:14,
// types/landmark.hpp:11-16, pnp.hpp:14-35
struct Vector3 // Eigen::Vector3d: three doubles, not over-aligned
{
    double v[3];
    double x() const { return v[0]; }
    double y() const { return v[1]; }
    double z() const { return v[2]; }
    double& operator[](int i) { return v[i]; }
    double operator[](int i) const { return v[i]; }
};

struct alignas(16) Quaternion // Eigen::Quaterniond: coefficients stored x, y, z, w; identity by default here
{
    double c[4] = {0.0, 0.0, 0.0, 1.0};
    double x() const { return c[0]; }
    double y() const { return c[1]; }
    double z() const { return c[2]; }
    double w() const { return c[3]; }
};

template <typename PositionType, typename OrientationType>
struct State
{
    PositionType position;
    OrientationType orientation;
};

namespace slam3d
{
using SensorState = mslam::State<Vector3, Quaternion>;
}

template <typename StateType>
struct Landmark
{
    Id id;
    StateType state;
};

namespace shim
{
// the part of boost::dynamic_bitset<> the reference touches (resize(n, false), operator[], count, size)
class dynamic_bitset
{
  public:
    void resize(std::size_t n, bool value = false) { bits.resize(n, value); }
    std::vector<bool>::reference operator[](std::size_t i) { return bits[i]; }
    bool operator[](std::size_t i) const { return bits[i]; }
    std::size_t size() const { return bits.size(); }
    std::size_t count() const
    {
        std::size_t n = 0;
        for(bool b : bits)
            n += b;
        return n;
    }

  private:
    std::vector<bool> bits;
};
} // namespace shim

template <typename SensorStateType, typename LandmarkStateType>
class IPnpAlgorithm
{
  public:
    struct PnpResult
    {
        SensorStateType pose;
        shim::dynamic_bitset inliers;
    };
    virtual std::optional<PnpResult> solvePnp(const std::vector<std::shared_ptr<Landmark<LandmarkStateType>>>& landmarks,
                                              const std::vector<mslam::Vector2>& imgPoints,
                                              const SensorStateType& initial = SensorStateType()) = 0;
    void setCameraParameters(const CameraParameters& newParameters) { cameraParams = newParameters; }
    [[nodiscard]] const CameraParameters& cameraParameters() const { return cameraParams; }
    virtual ~IPnpAlgorithm() = default;

  protected:
    CameraParameters cameraParams;
};

using IOrbFeatureDetector = IFeatureDetector<RgbFrame, std::uint8_t, 32>;
using IOrbMatcher = IFeatureMatcher<std::uint8_t, 32>;
using OrbKeypoint = KeypointDescriptor<std::uint8_t, 32>;
} // namespace mslam
