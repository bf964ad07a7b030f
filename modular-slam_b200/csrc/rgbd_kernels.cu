// rgbd_kernels.cu -- SURVEY.md section 8(f) row f2: depth back-projection of matched keypoints, the data-parallel
// step that follows the matcher in RgbdFeatureFrontend::track (src/lib/modular_slam/rgbd_feature_frontend.cpp:292-293).
//
// Restated arithmetic (rgbd_feature_frontend.cpp:101-138, types/depth_frame.hpp:20-30):
//   imgPoint = coordinates.cast<int>()                     -- double -> int, truncation toward zero
//   depth    = (float)data[width * y + x] * factor         -- uint16 -> int -> float, one rounded f32 product
//   valid    = depth > FLT_EPSILON
//   z = depth;  x = (u - cx) * (double)z * (1.0 / fx);  y = (v - cy) * (double)z * (1.0 / fy)      -- IEEE double, left to
//   right, no FMA (the reference is built for baseline x86-64), with u, v the UN-truncated double coordinates.
#include "common.cuh"
#include "kernels.h"

namespace mslam_b200
{

struct Point3
{
    double x, y, z;
    bool ok;
};
__device__ __forceinline__ Point3 backproject_one(double u, double v, const uint16_t* __restrict__ depth, int width, int height,
                                                  const mslam_b200_camera& cam)
{
    Point3 r = {0.0, 0.0, 0.0, false};
    if(!(u > -2147483648.0 && u < 2147483647.0 && v > -2147483648.0 && v < 2147483647.0))
        return r;
    const int xi = (int)u, yi = (int)v; // cast<int>(): truncation toward zero
    if(xi < 0 || yi < 0 || xi >= width || yi >= height)
        return r;
    const float d = __fmul_rn((float)(int)depth[(long long)yi * width + xi], cam.factor);
    if(!(d > 1.1920928955078125e-07f)) // std::numeric_limits<float>::epsilon()
        return r;
    const double inv_fx = __ddiv_rn(1.0, cam.fx), inv_fy = __ddiv_rn(1.0, cam.fy);
    r.z = (double)d;
    r.x = __dmul_rn(__dmul_rn(__dsub_rn(u, cam.cx), r.z), inv_fx);
    r.y = __dmul_rn(__dmul_rn(__dsub_rn(v, cam.cy), r.z), inv_fy);
    r.ok = true;
    return r;
}

// one thread per (pair, slot).  `index` (optional) selects keypoints, e.g. the matcher's fromIndex list, so that the
// association "match -> 3D point of its keypoint" never leaves the device.
__global__ void __launch_bounds__(256) k_backproject(const mslam_b200_keypoint* __restrict__ kp, int kp_cap,
                                                     const int* __restrict__ index, const int* __restrict__ counts, int cap,
                                                     const uint16_t* __restrict__ depth, int width, int height,
                                                     mslam_b200_camera cam, double* __restrict__ xyz,
                                                     uint8_t* __restrict__ valid)
{
    const int p = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = min(counts[p], cap);
    if(i >= n)
        return;
    const int src = index ? index[(long long)p * cap + i] : i;
    Point3 r = {0.0, 0.0, 0.0, false};
    if(src >= 0 && src < kp_cap)
    {
        const mslam_b200_keypoint k = kp[(long long)p * kp_cap + src];
        // Keypoint::coordinates = (double)pt (orb_feature.cpp:57)
        r = backproject_one((double)k.x, (double)k.y, depth + (long long)p * height * width, width, height, cam);
    }
    double* o = xyz + ((long long)p * cap + i) * 3;
    o[0] = r.x;
    o[1] = r.y;
    o[2] = r.z;
    valid[(long long)p * cap + i] = r.ok ? 1 : 0;
}

// host entry point: coordinates as the reference stores them (two doubles per record)
__global__ void __launch_bounds__(256) k_backproject_xy(const double* __restrict__ xy, int n, const uint16_t* __restrict__ depth,
                                                        int width, int height, mslam_b200_camera cam,
                                                        double* __restrict__ xyz, uint8_t* __restrict__ valid)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if(i >= n)
        return;
    const Point3 r = backproject_one(xy[2 * i], xy[2 * i + 1], depth, width, height, cam);
    xyz[3 * i] = r.x;
    xyz[3 * i + 1] = r.y;
    xyz[3 * i + 2] = r.z;
    valid[i] = r.ok ? 1 : 0;
}

void launch_backproject_xy(cudaStream_t st, const double* xy, int n, const uint16_t* depth, int width, int height,
                           const mslam_b200_camera& cam, double* xyz, uint8_t* valid, long long* launches)
{
    k_backproject_xy<<<(n + 255) / 256, 256, 0, st>>>(xy, n, depth, width, height, cam, xyz, valid);
    ++*launches;
}

void launch_backproject(cudaStream_t st, const mslam_b200_keypoint* kp, int kp_cap, const int* index, const int* counts,
                        int n_pairs, int cap, const uint16_t* depth, int width, int height, const mslam_b200_camera& cam,
                        double* xyz, uint8_t* valid, long long* launches)
{
    k_backproject<<<dim3((cap + 255) / 256, n_pairs), 256, 0, st>>>(kp, kp_cap, index, counts, cap, depth, width, height, cam,
                                                                    xyz, valid);
    ++*launches;
}

} // namespace mslam_b200
