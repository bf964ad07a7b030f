// kernels.h -- host-side launchers of the sm_100a kernels (orb_kernels.cu, match_kernels.cu).
#pragma once
#include "common.cuh"

namespace mslam_b200
{
// ORB path
void launch_gray(cudaStream_t st, const uint8_t* d_bgr, uint8_t* pyr, const PipeParams& P, long long* launches);
void launch_pyramid(cudaStream_t st, uint8_t* pyr, const PipeParams& P, const short2* xtab, const short2* ytab,
                    const uint4* qtab, int sm_count, int l_begin, int l_end, long long* launches);
void launch_blur(cudaStream_t st, const uint8_t* pyr, uint8_t* blur, const PipeParams& P, int l_begin, int l_end,
                 long long* launches);
void fast_tiles(int w, int h, int edge, bool short_tiles, int* tx, int* ty, int* tile_sh);
void blur_warps(int w, int h, int* xblocks, int* warps);
int blur_band_rows();
// one CUtensorMap (128 opaque bytes) per pyramid level: the FAST kernel's tile loads (k_fast, orb_kernels.cu)
struct alignas(64) FastMaps
{
    unsigned char m[kMaxLevels][128];
};
cudaError_t build_fast_maps(const PipeParams& P, const uint8_t* pyr, int max_batch, FastMaps* out);
void launch_fast(cudaStream_t st, const uint8_t* pyr, unsigned* cand, LevelCounters* ctr, unsigned* hist,
                 const PipeParams& P, const FastMaps& maps, int l_begin, int l_end, long long* launches);
void launch_select(cudaStream_t st, const uint8_t* pyr, unsigned* cand, unsigned* cand2, unsigned* keys, unsigned* keys2,
                   LevelCounters* ctr, const unsigned* hist, SelRec* sel, int* err, const PipeParams& P, int sm_count,
                   long long* launches);
void launch_orient_describe(cudaStream_t st, const uint8_t* pyr, const uint8_t* blur, const LevelCounters* ctr,
                            const SelRec* sel, const float4* pat_t, const uint2* ic_w, mslam_b200_keypoint* out_kp,
                            uint8_t* out_desc, int* out_counts, int out_cap, int n_slots, const PipeParams& P, int sm_count,
                            long long* launches);
void ic_weight_table(unsigned* out /* 4 * 31 * 9 * 2 */);
void brief_pattern_transposed(float* out1024);
// matching path
constexpr int kKnnMaxSlices = 8; // train-set slices of the small-batch kNN (scratch: kKnnMaxSlices x [pair][cap] keys)
void launch_knn2(cudaStream_t st, const uint8_t* q, const int* qn, const uint8_t* t, const int* tn, int n_pairs, int cap,
                 unsigned* key0, unsigned* key1, unsigned* part0, unsigned* part1, int sm_count, long long* launches);
// tensor-core kNN (knn_mma.cu): descriptors expanded to +1/-1 int8 rows, then D = Q T^T on tcgen05 with a top-2 epilogue
constexpr int kDbDpt = 4;                // keyframe descriptors per thread of the scalar DB kernel
constexpr int kDbMaxSmem = 200 * 1024;   // largest resident query frame of the scalar DB kernel (6400 descriptors)
cudaError_t configure_match_kernels();   // per-device kernel attributes (dynamic shared memory opt-in)
cudaError_t configure_mma_kernels();
int knn_mma_pad(int cap);
size_t knn_mma_bytes(int n_pairs, int cap);
void launch_expand_pm1(cudaStream_t st, const uint8_t* desc, const int* counts, int n_pairs, int cap, uint8_t* out,
                       long long* launches);
void launch_expand_pm1_x2(cudaStream_t st, const uint8_t* desc0, const int* counts0, uint8_t* out0, const uint8_t* desc1,
                          const int* counts1, uint8_t* out1, int n_pairs, int cap, long long* launches);
cudaError_t launch_knn2_mma(cudaStream_t st, const uint8_t* xq, const int* qn, const uint8_t* xt, const int* tn, int n_pairs,
                            int cap, unsigned* key0, unsigned* key1, long long* launches);
cudaError_t launch_db_scores_mma(cudaStream_t st, const uint8_t* xqueries, const int* q_counts, int nq_frames, int q_cap,
                                 const uint8_t* xdb, const int* k_counts, int n_keyframes, int k_cap, double ratio, int* scores,
                                 int score_stride, long long* launches);
void launch_ratio_compact(cudaStream_t st, const unsigned* key0, const unsigned* key1, const int* qn, const int* tn,
                          int n_pairs, int cap, double ratio, int* out_a, int* out_b, int* out_d, int* n_out,
                          long long* launches);
void launch_mutual_compact(cudaStream_t st, const unsigned* key0_q, const unsigned* key0_t, const int* qn, const int* tn,
                           int n_pairs, int cap, int* out_a, int* out_b, int* out_d, int* n_out, long long* launches);
void launch_unpack_knn(cudaStream_t st, const unsigned* key0, const unsigned* key1, int n, int* idx0, int* d0, int* idx1,
                       int* d1, long long* launches);
cudaError_t launch_db_scores(cudaStream_t st, const uint8_t* queries, const int* q_counts, int nq_frames, int q_cap,
                             const uint8_t* db, const int* k_counts, int n_keyframes, int k_cap, double ratio, int* scores,
                             int sm_count, long long* launches);
void launch_topk(cudaStream_t st, const int* scores, const int* ids, int n_lists, int per_list, long long list_stride,
                 int n_queries, int id_base, int k, int* out_score, int* out_id, long long* launches);
// RGB-D association (rgbd_kernels.cu)
void launch_backproject(cudaStream_t st, const mslam_b200_keypoint* kp, int kp_cap, const int* index, const int* counts,
                        int n_pairs, int cap, const uint16_t* depth, int width, int height, const mslam_b200_camera& cam,
                        double* xyz, uint8_t* valid, long long* launches);
void launch_backproject_xy(cudaStream_t st, const double* xy, int n, const uint16_t* depth, int width, int height,
                           const mslam_b200_camera& cam, double* xyz, uint8_t* valid, long long* launches);
// RANSAC PnP (pnp_kernels.cu)
void launch_pnp_ransac(cudaStream_t st, const float* obj, const float* img, const int* counts, int n_frames, int cap, int frame0,
                       double fx, double fy, double cx, double cy, int iterations, double thr2, unsigned seed, int refine_iters,
                       double* out_rt, double* out_pose, uint8_t* out_mask, int* out_info, long long* launches);
int pnp_max_hypotheses();
cudaError_t measure_tensor_peak(cudaStream_t st, int sm_count, double* out4, long long* launches);
cudaError_t measure_int_peaks(cudaStream_t st, int sm_count, double* out4, long long* launches);
} // namespace mslam_b200
