// common.cuh -- shared device/host structures of the B200 ORB + Hamming hot path.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/mslam_b200.h"

namespace mslam_b200
{

constexpr int kMaxLevels = MSLAM_B200_MAX_LEVELS;
constexpr int kPatchRadius = 15; // ORB patchSize 31 (cv::ORB default; orb_feature.cpp:25)
constexpr int kHarrisBlock = 7;

// One pyramid level of the batch.  Level planes are level-major in HBM:
//   pixel (b, y, x) of level l lives at  pyr + img_off + b*img_frame_stride + y*pitch + x
struct LevelDesc
{
    int w, h, pitch; // pitch: bytes per row, multiple of 16
    int n_keep;      // cv::ORB nfeaturesPerLevel[l]
    int cand_cap;    // capacity of this level's candidate list (upper bound of strict 3x3 maxima)
    int tiles_x, tiles_y; // FAST tiling of the interior
    int tile_sh;          // scored rows per FAST tile (32 or 64)
    int tile_base;   // first FAST tile index of this level inside the fused launch
    int blur_xblocks, blur_warp_base; // blur: 128-px column blocks per row band; first warp index of the level
    long long img_off;          // bytes
    long long img_frame_stride; // bytes
    long long cand_off;         // elements, within one frame's candidate block
    int xtab_off, ytab_off;     // offsets into the resize tap tables (short2 {ofs, alpha1})
    int qtab_off, rs_quads;     // resize quad table (uint4 per 4 output columns); rs_quads = 0 => generic kernel
    float scale;                // (float)pow((double)scaleFactor, l)
    float inv_scale;            // 1.f / scale
};

struct PipeParams
{
    int nlevels, n_frames, W, H;
    int edge, fast_thr, kp_cap;
    int frame0; // first frame (absolute index in the context's batch buffers) this launch covers; n_frames from there
    long long cand_frame_stride; // elements per frame in the candidate arrays
    int total_fast_tiles, pad1;
    int blur_band;      // rows one warp of k_blur walks
    int one, minus_one; // +1 / -1 as run-time values (IMAD pipe steering in k_fast)
    LevelDesc lv[kMaxLevels];
};

// per-(frame, level) bookkeeping in HBM, zeroed at the start of every batch
struct LevelCounters
{
    unsigned int n_fast;  // NMS + border survivors
    unsigned int n_cut1;  // survivors of retainBest(2 n_l) on the FAST score
    unsigned int n_final; // survivors of retainBest(n_l) on the Harris response
    unsigned int pad;
};

// selected keypoint of one level, before orientation / description
struct SelRec
{
    unsigned int yx_score; // y << 20 | x << 8 | fast score
    float response;
};

#ifdef __CUDACC__
// cv::fastAtan2 (degrees), every operation rounded separately (OpenCV core mathfuncs_core atan_f32; shared by the ORB and the
// distributed-ORB orientation kernels)
__device__ __forceinline__ float fast_atan2_deg(float y, float x)
{
    const float p1 = __int_as_float(0x4265226f), p3 = __int_as_float(0xc19556ee), p5 = __int_as_float(0x410e9fbf),
                p7 = __int_as_float(0xc0228ad9);
    const float eps = 2.2204460492503131e-16f; // (float)DBL_EPSILON
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if(ax >= ay)
    {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(
            __fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    }
    else
    {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(
            90.f,
            __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if(x < 0)
        a = __fsub_rn(180.f, a);
    if(y < 0)
        a = __fsub_rn(360.f, a);
    return a;
}
#endif

#define MSLAM_CUDA_OK(expr)                                                                                            \
    do                                                                                                                 \
    {                                                                                                                  \
        cudaError_t e__ = (expr);                                                                                      \
        if(e__ != cudaSuccess)                                                                                         \
            return ctx_fail(ctx, e__, #expr);                                                                          \
    } while(0)

} // namespace mslam_b200
