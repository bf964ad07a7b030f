/*
 * mslam_b200.h -- C ABI of the B200-native ORB + Hamming-matching hot path.
 *
 * This is the drop-in boundary for modular-slam's feature front-end.  Every entry point cites the
 * reference interface it replaces (paths relative to the reference repository root):
 *
 *   IFeatureDetector<RgbFrame,uint8_t,32>::detect   src/lib/modular_slam/include/modular_slam/frontend/feature/feature_interface.hpp:50-56
 *     implemented by OrbOpenCvDetector::Pimpl::detect  src/lib/modular_slam/orb_feature.cpp:33-65
 *     (toGrayScale src/lib/modular_slam/frame.cpp:6-27, then cv::ORB::create(1000)->detectAndCompute, orb_feature.cpp:25,40)
 *   IFeatureMatcher<uint8_t,32>::match              feature_interface.hpp:62-70
 *     implemented by OrbOpenCvMatcher::Pimpl::match    src/lib/modular_slam/orb_feature.cpp:84-117
 *
 * Conventions
 *   - plain pointers and sizes only; no allocation crosses the boundary (caller provides capacity, gets a count);
 *   - every function returns MSLAM_B200_OK (0) or a negative mslam_b200_status;
 *   - "host" entry points take host pointers (pageable or pinned), run synchronously and include the
 *     host<->device copies; "_dev" entry points take device pointers, enqueue on the context's stream and
 *     return immediately (call mslam_b200_synchronize, or order further work on the same stream);
 *   - there is no CPU fallback: without a CUDA device mslam_b200_create fails with MSLAM_B200_E_CUDA.
 *   - a context is single-threaded (the reference caller is: slam.hpp:74-99), one context per thread/GPU.
 */
#ifndef MSLAM_B200_H_
#define MSLAM_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define MSLAM_B200_API __attribute__((visibility("default")))
#else
#define MSLAM_B200_API
#endif

#define MSLAM_B200_VERSION 100 /* major*100 + minor */
#define MSLAM_B200_MAX_LEVELS 16
#define MSLAM_B200_DESC_BYTES 32

typedef enum mslam_b200_status
{
    MSLAM_B200_OK = 0,
    MSLAM_B200_E_INVALID = -1,  /* bad argument (null pointer, size outside the context's limits, ...) */
    MSLAM_B200_E_CUDA = -2,     /* CUDA runtime error or no device; see mslam_b200_last_cuda_error */
    MSLAM_B200_E_CAPACITY = -3, /* an output (or internal candidate list) did not fit; counts report the need */
    MSLAM_B200_E_NOMEM = -4
} mslam_b200_status;

typedef struct mslam_b200_ctx mslam_b200_ctx;

/* Defaults (mslam_b200_default_config) are the reference's hard-coded values:
 * cv::ORB::create(1000) => nfeatures 1000, scaleFactor 1.2f, nlevels 8, edgeThreshold 31, fastThreshold 20,
 * HARRIS_SCORE, patchSize 31 (orb_feature.cpp:25); ratio 0.7 (orb_feature.cpp:101). */
typedef struct mslam_b200_config
{
    int struct_size;    /* = sizeof(mslam_b200_config), for forward compatibility */
    int device;         /* CUDA device ordinal */
    int max_width;      /* largest frame the context accepts */
    int max_height;
    int max_batch;      /* frames per detect_batch call / frame pairs per match batch */
    int nfeatures;      /* ORB nfeatures */
    int nlevels;        /* ORB nlevels (<= MSLAM_B200_MAX_LEVELS) */
    float scale_factor; /* ORB scaleFactor */
    int fast_threshold; /* FAST threshold */
    int edge_threshold; /* ORB edgeThreshold (border filter); patch size stays 31, so must be >= 31 */
    int max_keypoints;  /* per-frame output capacity inside the context (retainBest keeps ties, so the count can
                           exceed nfeatures); 0 => nfeatures + nfeatures/4 + 64 */
    int max_descriptors;/* largest descriptor set per side in the matcher; 0 => max_keypoints */
    double ratio;       /* Lowe ratio of the kNN matcher, compared exactly as the reference does:
                           (double)(float)d0 < ratio * (double)(float)d1 */
    void* stream;       /* optional cudaStream_t to run on (e.g. the caller's); NULL => context-owned stream */
    int flags;          /* MSLAM_B200_FLAG_* */
    int reserved;
} mslam_b200_config;
/* run every kernel of detect on the one context stream (default: the pyramid chain, FAST + selection and the blur run on
 * three forked streams that join before the describe kernel, so that small levels and latency-bound stages overlap) */
#define MSLAM_B200_FLAG_SERIAL 1
/* kNN / keyframe-DB scoring kernel choice.  Default: by problem size -- the tensor-core kernel (Hamming distances as an
 * int8 GEMM on tcgen05, knn_mma.cu) for large problems, the scalar XOR + POPC kernel for small ones.  Both give the same
 * bits; the flags force one of them (tests, A/B measurements).  Environment MSLAM_KNN_MMA=0|1|2 does the same at create. */
#define MSLAM_B200_FLAG_KNN_SCALAR 2
#define MSLAM_B200_FLAG_KNN_TENSOR 4

/* what cv::KeyPoint carries out of detectAndCompute (the reference interface keeps only pt, orb_feature.cpp:52-58) */
typedef struct mslam_b200_keypoint
{
    float x, y;     /* level-0 pixel coordinates: (float)x_l * scale_l */
    float size;     /* 31 * scale_l */
    float angle;    /* degrees, cv::fastAtan2 of the intensity centroid */
    float response; /* Harris response */
    int octave;     /* pyramid level */
} mslam_b200_keypoint;

/* ---- library / context -------------------------------------------------------------------------------------- */
MSLAM_B200_API int mslam_b200_version(void);
MSLAM_B200_API const char* mslam_b200_status_string(int status);
MSLAM_B200_API void mslam_b200_default_config(mslam_b200_config* cfg);
MSLAM_B200_API int mslam_b200_create(const mslam_b200_config* cfg, mslam_b200_ctx** out_ctx);
MSLAM_B200_API void mslam_b200_destroy(mslam_b200_ctx* ctx);
MSLAM_B200_API int mslam_b200_set_stream(mslam_b200_ctx* ctx, void* cuda_stream);
MSLAM_B200_API int mslam_b200_synchronize(mslam_b200_ctx* ctx);
/* change mslam_b200_config::flags of a live context (synchronises first): the kernel choice of the matcher follows the new
 * flags from the next call on (expansion buffers are allocated if the tensor-core kernel becomes selectable), captured
 * single-frame kernel graphs are dropped.  KNN_SCALAR | KNN_TENSOR together is MSLAM_B200_E_INVALID. */
MSLAM_B200_API int mslam_b200_set_flags(mslam_b200_ctx* ctx, int flags);
/* text of the last CUDA error seen by this context ("" if none) */
MSLAM_B200_API const char* mslam_b200_last_cuda_error(mslam_b200_ctx* ctx);
/* number of kernels this library has launched on the context since creation (bench.py "gpu_launches") */
MSLAM_B200_API long long mslam_b200_launch_count(mslam_b200_ctx* ctx);
/* pinned host memory for callers that want full-speed H2D/D2H through the host entry points */
MSLAM_B200_API int mslam_b200_host_alloc(void** ptr, size_t bytes);
MSLAM_B200_API int mslam_b200_host_free(void* ptr);
/* per-level geometry the context derived (cv::ORB layer sizes / scales / feature budget); arrays of nlevels */
MSLAM_B200_API int mslam_b200_level_geometry(mslam_b200_ctx* ctx, int width, int height, int* w, int* h, float* scale,
                                             int* n_per_level);

/* ---- detection: replaces IFeatureDetector::detect / OrbOpenCvDetector::detect (orb_feature.cpp:33-65) ---------- */
/* One frame.  `bgr` is RgbFrame::data (3 interleaved bytes per pixel, weights applied in memory order exactly as
 * toGrayScale does, frame.cpp:15-19); row stride in bytes (>= 3*width).  Keypoints come out in canonical order
 * (octave, y, x); OpenCV's own order inside a level is a std::nth_element artefact.  n_out receives the count
 * found; MSLAM_B200_E_CAPACITY if it exceeds `capacity` (the first `capacity` are still written). */
MSLAM_B200_API int mslam_b200_detect(mslam_b200_ctx* ctx, const uint8_t* bgr, int width, int height, int stride_bytes,
                                     mslam_b200_keypoint* out_kp, uint8_t* out_desc /* capacity x 32 */, int capacity,
                                     int* n_out);
/* A batch of independent frames of one size (frames[i] = pointer to frame i).  Outputs are [n_frames][capacity]. */
MSLAM_B200_API int mslam_b200_detect_batch(mslam_b200_ctx* ctx, const uint8_t* const* frames, int n_frames, int width,
                                           int height, int stride_bytes, mslam_b200_keypoint* out_kp, uint8_t* out_desc,
                                           int capacity, int* n_out /* [n_frames] */);
/* Pipelined form of mslam_b200_detect_batch for callers that stream batches: _begin enqueues the copy-in, the kernels
 * and the copy-out of one batch and returns without waiting; _end waits for the OLDEST batch in flight and fills its
 * n_out.  Up to two batches may be in flight, so the copy-in of batch t+1 (and t+2 once t has ended) overlaps the
 * kernels and the copy-out of batch t -- the PCIe link never idles.  Frames and outputs must stay valid (and, for real
 * overlap, be pinned: mslam_b200_host_alloc) until the matching _end.  mslam_b200_detect_batch == _begin + _end. */
MSLAM_B200_API int mslam_b200_detect_batch_begin(mslam_b200_ctx* ctx, const uint8_t* const* frames, int n_frames, int width,
                                                 int height, int stride_bytes, mslam_b200_keypoint* out_kp,
                                                 uint8_t* out_desc, int capacity, int* n_out);
MSLAM_B200_API int mslam_b200_detect_batch_end(mslam_b200_ctx* ctx);
/* Device-resident variant: d_bgr is [n_frames][height][width][3] dense in device memory; outputs are device arrays
 * [n_frames][capacity] (keypoints, descriptors) and [n_frames] (counts, un-clamped).  Asynchronous. */
MSLAM_B200_API int mslam_b200_detect_batch_dev(mslam_b200_ctx* ctx, const uint8_t* d_bgr, int n_frames, int width,
                                               int height, mslam_b200_keypoint* d_kp, uint8_t* d_desc, int capacity,
                                               int* d_counts);
/* Describe GIVEN keypoints (parity aid: "descriptors bit-exact given identical keypoints").  kp[].x/.y/.octave are
 * read (level-0 coordinates, as detect returns them); angle is recomputed and written back together with desc. */
MSLAM_B200_API int mslam_b200_describe(mslam_b200_ctx* ctx, const uint8_t* bgr, int width, int height, int stride_bytes,
                                       mslam_b200_keypoint* kp, int n_kp, uint8_t* out_desc);

/* ---- matching: replaces IFeatureMatcher::match / OrbOpenCvMatcher::match (orb_feature.cpp:84-117) -------------- */
/* match(from, to): kNN (k=2) with query = `to`, train = `from`, ties -> lower train index; keep the best iff
 * d0 < ratio*d1; emit {fromIndex = trainIdx, toIndex = queryIdx} in ascending toIndex (orb_feature.cpp:96-114).
 * Strides are in bytes so the adapter can pass &v[0].descriptor[0] with stride sizeof(OrbKeypoint) = 64, as the
 * reference itself does (orb_feature.cpp:88-91).  n_from < 2 or n_to < 1 => zero matches (reference: undefined). */
MSLAM_B200_API int mslam_b200_match_knn_ratio(mslam_b200_ctx* ctx, const uint8_t* from, int n_from, int from_stride,
                                              const uint8_t* to, int n_to, int to_stride, uint64_t* out_from_idx,
                                              uint64_t* out_to_idx, int* out_dist /* may be NULL */, int capacity,
                                              int* n_out);
/* Raw k=2 result per query row (cv::BFMatcher::knnMatch(query, train, 2)); -1 where train has fewer rows. */
MSLAM_B200_API int mslam_b200_knn2(mslam_b200_ctx* ctx, const uint8_t* query, int n_query, int query_stride,
                                   const uint8_t* train, int n_train, int train_stride, int* idx0, int* dist0, int* idx1,
                                   int* dist1);
/* BFMatcher(NORM_HAMMING, crossCheck=true).match(query, train): mutual nearest neighbours, lowest-index argmin on
 * both sides, ascending query index (BASELINE.json configs[2]; not used by the reference itself). */
MSLAM_B200_API int mslam_b200_match_crosscheck(mslam_b200_ctx* ctx, const uint8_t* query, int n_query, int query_stride,
                                               const uint8_t* train, int n_train, int train_stride, int* out_query_idx,
                                               int* out_train_idx, int* out_dist, int capacity, int* n_out);
/* Batches of independent (from, to) pairs, dense [n_pairs][capacity][32] descriptor blocks and [n_pairs] counts.
 * Host variant copies in/out; outputs [n_pairs][capacity] (+ [n_pairs] match counts).  mode 0 = knn+ratio
 * (out_a = fromIndex, out_b = toIndex), mode 1 = cross-check with query = `to`, train = `from`
 * (out_a = trainIdx, out_b = queryIdx). */
MSLAM_B200_API int mslam_b200_match_batch(mslam_b200_ctx* ctx, int mode, const uint8_t* from, const int* n_from,
                                          const uint8_t* to, const int* n_to, int n_pairs, int capacity, int* out_a,
                                          int* out_b, int* out_dist, int* n_out);
/* pipelined form (one match batch in flight): _begin enqueues on the context stream, _end waits for the results */
MSLAM_B200_API int mslam_b200_match_batch_begin(mslam_b200_ctx* ctx, int mode, const uint8_t* from, const int* n_from,
                                                const uint8_t* to, const int* n_to, int n_pairs, int capacity, int* out_a,
                                                int* out_b, int* out_dist, int* n_out);
MSLAM_B200_API int mslam_b200_match_batch_end(mslam_b200_ctx* ctx);
MSLAM_B200_API int mslam_b200_match_batch_dev(mslam_b200_ctx* ctx, int mode, const uint8_t* d_from, const int* d_n_from,
                                              const uint8_t* d_to, const int* d_n_to, int n_pairs, int capacity,
                                              int* d_out_a, int* d_out_b, int* d_out_dist, int* d_n_out);

/* ---- matching without any allocation ------------------------------------------------------------------------------- */
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  One (from, to) pair, the knn + ratio
 * semantics of mslam_b200_match_knn_ratio (query = `to`, train = `from`, k = 2, ties to the lower train index, keep iff
 * (double)(float)d0 < ratio * (double)(float)d1, ascending toIndex), device in / device out, NO context:
 *   d_from [n_from][32], d_to [n_to][32]       dense descriptor rows in device memory
 *   d_out_from / d_out_to / d_out_dist [n_to]  fromIndex, toIndex, Hamming distance of the kept matches (d_out_dist may be NULL)
 *   d_n_out [1]                                number of matches (0 when n_to < 1 or n_from < 2, as the other entry points)
 * The kernels' scratch (the packed top-2 keys of every query, the partial lists of the train-set slices, the two counts:
 * 164 KB) lives in a file-scope __device__ array of the library, which every GPU that loads the library holds until the
 * process exits; nothing is allocated or freed by this call, ever (no cudaMalloc / cudaFree / allocating library routine).
 * Limit: n_from, n_to <= MSLAM_B200_NOALLOC_MAX_DESC; larger sets are rejected with MSLAM_B200_E_INVALID and a message in
 * err_msg (optional buffer of err_len bytes, always NUL-terminated when given; also filled for CUDA errors).
 * `device` is made current; work is enqueued on `cuda_stream` (NULL = the legacy default stream) and the call returns
 * without waiting.  The first call per device looks up the scratch array's device address (cudaGetSymbolAddress) and the
 * SM count -- make that first call outside any graph capture.
 * ALL calls in the process that target one device share the one scratch array: calls whose kernels may overlap on the GPU
 * (different streams without an ordering between them) race on it; calls on one stream, or ordered by events, are safe.
 * The array keeps the previous call's contents; every call rewrites the counts and every key it later reads. */
#define MSLAM_B200_NOALLOC_MAX_DESC 4096
MSLAM_B200_API int mslam_b200_match_knn_ratio_noalloc_dev(int device, void* cuda_stream, const uint8_t* d_from, int n_from,
                                                          const uint8_t* d_to, int n_to, double ratio, int* d_out_from,
                                                          int* d_out_to, int* d_out_dist, int* d_n_out, char* err_msg,
                                                          size_t err_len);

/* ---- keyframe-database scoring (loop-candidate verification; rgbd_feature_frontend.cpp:499-533 intent) --------- */
/* score[q][k] = |match(from = query frame q, to = keyframe k)| with the knn+ratio semantics above.
 * Descriptor blocks are device arrays [n][cap][32] with per-frame counts; scores is a device array
 * [n_query_frames][n_keyframes] (int32).  Asynchronous. */
MSLAM_B200_API int mslam_b200_db_scores_dev(mslam_b200_ctx* ctx, const uint8_t* d_queries, const int* d_q_counts,
                                            int n_query_frames, int q_cap, const uint8_t* d_db, const int* d_k_counts,
                                            int n_keyframes, int k_cap, int* d_scores);
/* per query frame: top-k keyframes by (score desc, id asc); ids are local index + id_base (shard start).
 * d_out_score / d_out_id are device arrays [n_query_frames][k]; (-1,-1) pads when n_keyframes < k. */
MSLAM_B200_API int mslam_b200_db_topk_dev(mslam_b200_ctx* ctx, const int* d_scores, int n_query_frames, int n_keyframes,
                                          int id_base, int k, int* d_out_score, int* d_out_id);
/* merge n_lists candidate lists per query (e.g. the all-gathered per-shard top-k, laid out
 * [n_lists][n_query_frames][k]) into one top-k with the same (score desc, id asc) order. */
MSLAM_B200_API int mslam_b200_db_merge_topk_dev(mslam_b200_ctx* ctx, const int* d_scores, const int* d_ids, int n_lists,
                                                int n_query_frames, int k, int* d_out_score, int* d_out_id);
/* host convenience wrapper: scores + top-k of one shard, host in / host out */
MSLAM_B200_API int mslam_b200_db_query(mslam_b200_ctx* ctx, const uint8_t* queries, const int* q_counts,
                                       int n_query_frames, int q_cap, const uint8_t* db, const int* k_counts,
                                       int n_keyframes, int k_cap, int id_base, int k, int* out_score, int* out_id,
                                       int* out_all_scores /* may be NULL; [n_query_frames][n_keyframes] */);

/* ---- sharded keyframe database (SURVEY 8(b) "_db_build / _db_query: sharded, takes ncclComm_t or rank/size"; 8(e)) ---------
 * Keyframes are sharded by contiguous id range, one shard per rank (one process per GPU), each shard resident in its rank's
 * HBM as a caller-owned device array [n_local][k_cap][32] + counts; every rank sees all query frames.  A query runs
 *   score the shard -> local top-k -> ONE ncclAllGather of the [2][n_query_frames][k] (score, id) block -> k-way merge
 * on the context's stream with no host synchronisation in between, and leaves the same global top-k -- bit-identical to the
 * single-GPU answer, (score desc, keyframe id asc) -- on every rank.  NCCL is bound at run time (dlopen libnccl.so.2, so a
 * process that already holds PyTorch's NCCL shares it); without it the attach calls fail with MSLAM_B200_E_CUDA for
 * world > 1 -- there is no fallback.  world == 1 needs no NCCL and no attach. */
#define MSLAM_B200_NCCL_ID_BYTES 128
/* contiguous id range [k0, k1) of `rank`: the first n_keyframes % world ranks hold one keyframe more */
MSLAM_B200_API void mslam_b200_db_shard_range(long long n_keyframes, int rank, int world, long long* k0, long long* k1);
/* ncclGetUniqueId: rank 0 calls this and hands the 128 bytes to every rank by any means (MPI, a file, torch.distributed) */
MSLAM_B200_API int mslam_b200_nccl_unique_id(void* out_id128);
/* join the database's communicator: ncclCommInitRank(world, id, rank) on the context's device (collective over all ranks) */
MSLAM_B200_API int mslam_b200_db_attach(mslam_b200_ctx* ctx, int rank, int world, const void* nccl_unique_id);
/* ... or use a communicator the caller already has (an ncclComm_t; checked against rank / world; not destroyed by detach) */
MSLAM_B200_API int mslam_b200_db_attach_comm(mslam_b200_ctx* ctx, void* nccl_comm, int rank, int world);
MSLAM_B200_API int mslam_b200_db_detach(mslam_b200_ctx* ctx);
MSLAM_B200_API int mslam_b200_db_shard_info(mslam_b200_ctx* ctx, int* rank, int* world);
/* Device pointers, asynchronous.  n_local may be 0 (an empty shard still takes part in the collective).  Every rank must
 * call with the same n_query_frames and k.  d_out_score / d_out_id: [n_query_frames][k], (-1, -1) pads when the whole
 * database holds fewer than k keyframes. */
MSLAM_B200_API int mslam_b200_db_query_sharded_dev(mslam_b200_ctx* ctx, const uint8_t* d_queries, const int* d_q_counts,
                                                   int n_query_frames, int q_cap, const uint8_t* d_db_shard,
                                                   const int* d_k_counts, int n_local, int k_cap, int id_base, int k,
                                                   int* d_out_score, int* d_out_id);
/* Host query frames in, host top-k out (synchronous); the shard arguments are still device pointers (resident). */
MSLAM_B200_API int mslam_b200_db_query_sharded(mslam_b200_ctx* ctx, const uint8_t* queries, const int* q_counts,
                                               int n_query_frames, int q_cap, const uint8_t* d_db_shard,
                                               const int* d_k_counts, int n_local, int k_cap, int id_base, int k,
                                               int* out_score, int* out_id);

/* A shard owned by the context, for callers without device allocations of their own (the C++ class
 * mslam::ShardedKeyframeDbB200): _load_shard uploads host descriptors [n_local][k_cap][32] + counts once (replacing any
 * previous shard; n_local may be 0), _query_loaded runs mslam_b200_db_query_sharded against it. */
MSLAM_B200_API int mslam_b200_db_load_shard(mslam_b200_ctx* ctx, const uint8_t* db, const int* k_counts, int n_local, int k_cap,
                                            int id_base);
MSLAM_B200_API int mslam_b200_db_query_loaded(mslam_b200_ctx* ctx, const uint8_t* queries, const int* q_counts,
                                              int n_query_frames, int q_cap, int k, int* out_score, int* out_id);

/* ---- RGB-D association: SURVEY 8(f) row f2 ------------------------------------------------------------------------ */
/* Replaces pointsFromRgbdKeypoints / reconstructPoint (src/lib/modular_slam/rgbd_feature_frontend.cpp:101-138) with
 * getDepth / isDepthValid (include/modular_slam/types/depth_frame.hpp:20-30), the step RgbdFeatureFrontend::track runs
 * on the matcher's output (rgbd_feature_frontend.cpp:292-293).  CameraParameters (sensors/camera_parameters.hpp:7-12). */
typedef struct mslam_b200_camera
{
    double cx, cy;  /* principalPoint */
    double fx, fy;  /* focal */
    float factor;   /* depth unit -> metres */
    int reserved;
} mslam_b200_camera;
/* Device variant, batched: kp [n_pairs][kp_cap] as detect wrote them; d_index [n_pairs][cap] picks keypoints (e.g. the
 * fromIndex list of match_batch_dev; NULL = identity), d_counts [n_pairs]; depth [n_pairs][height][width] u16.
 * Outputs d_xyz [n_pairs][cap][3] double (camera frame) and d_valid [n_pairs][cap] (0 where the reference returns
 * std::nullopt: depth <= FLT_EPSILON; also 0 for a pixel outside the frame, where the reference reads out of bounds). */
MSLAM_B200_API int mslam_b200_backproject_dev(mslam_b200_ctx* ctx, const mslam_b200_keypoint* d_kp, int kp_cap,
                                              const int* d_index, const int* d_counts, int n_pairs, int cap,
                                              const uint16_t* d_depth, int width, int height, const mslam_b200_camera* cam,
                                              double* d_xyz, uint8_t* d_valid);
/* Host variant, one frame: xy = &keypoints[0].keypoint.coordinates (two doubles, stride_bytes = sizeof(OrbKeypoint) = 64
 * for the reference's records); DepthFrame::data in `depth`.  out_xyz [n][3], out_valid [n]. */
MSLAM_B200_API int mslam_b200_backproject(mslam_b200_ctx* ctx, const double* xy, int n, int stride_bytes,
                                          const uint16_t* depth, int width, int height, const mslam_b200_camera* cam,
                                          double* out_xyz, uint8_t* out_valid);

/* ---- RANSAC PnP: SURVEY 8(f) row f4 -------------------------------------------------------------------------------- */
/* Replaces OpenCvRansacPnp::solvePnp (src/lib/modular_slam/cv_ransac_pnp.cpp:15-85), i.e. the call
 *   cv::solvePnPRansac(objectPoints f32, imagePoints f32, cameraMatrix f32, noArray(), rvec, tvec, true, 100, 5.0, 0.99,
 *                      inliers)                                                                        (:56-57)
 * made by RgbdFeatureFrontend::track on the matcher's output (rgbd_feature_frontend.cpp:345-361).  All hypotheses of a
 * frame are evaluated in parallel and a call covers a batch of frames (csrc/pnp_kernels.cu).  The minimal solver (P3P +
 * a 4th point) and the random stream (a counter-based hash of seed / frame / hypothesis) differ from OpenCV's EPnP-on-5 +
 * cv::RNG, so agreement with cv::solvePnPRansac is by tolerance (same inlier set on separable data, same
 * least-squares pose over it), not by bits; `confidence` is accepted for interface parity and unused (no early exit:
 * the hypotheses do not run one after another).  The extrinsic guess the reference passes only seeds OpenCV's final
 * Levenberg-Marquardt fit; here the fit starts from the winning hypothesis, so no guess is taken. */
typedef struct mslam_b200_pnp_params
{
    int struct_size;          /* = sizeof(mslam_b200_pnp_params) */
    int iterations;           /* hypotheses per frame; reference: 100 (1..256) */
    double reproj_error;      /* inlier threshold in pixels; reference: 5.0 (squared error <= reproj_error^2) */
    double confidence;        /* reference: 0.99; unused */
    unsigned int seed;        /* random stream; frame f of a batch call uses (seed, frame_base + f) */
    int frame_base;
    int refine_iterations;    /* Levenberg-Marquardt cap; default 30 */
    int reserved;
    double fx, fy, cx, cy;    /* narrowed to float like the reference's cameraMatrix (:52-53) */
} mslam_b200_pnp_params;
MSLAM_B200_API void mslam_b200_pnp_default_params(mslam_b200_pnp_params* p);
/* Device variant: d_obj [n_frames][cap][3] f32 landmark positions (world), d_img [n_frames][cap][2] f32 pixels,
 * d_counts [n_frames].  Outputs per frame: d_rvec_tvec [6] f64 (world->camera Rodrigues vector, translation),
 * d_pose [7] f64 (the reference's result.pose: orientation quaternion w,x,y,z = R^-1 and position = -(R^-1 t), :65-77),
 * d_inliers [cap] u8 (result.inliers, :79-82), d_info [4] i32 = {found, n_inliers, winning hypothesis, LM iterations}.
 * found = 0 where the reference returns std::nullopt (and for n < 4, where cv::solvePnPRansac throws). */
MSLAM_B200_API int mslam_b200_pnp_ransac_batch_dev(mslam_b200_ctx* ctx, const float* d_obj, const float* d_img,
                                                   const int* d_counts, int n_frames, int cap,
                                                   const mslam_b200_pnp_params* params, double* d_rvec_tvec, double* d_pose,
                                                   uint8_t* d_inliers, int* d_info);
/* Host variant (host in / host out, same layouts). */
MSLAM_B200_API int mslam_b200_pnp_ransac_batch(mslam_b200_ctx* ctx, const float* obj, const float* img, const int* counts,
                                               int n_frames, int cap, const mslam_b200_pnp_params* params,
                                               double* rvec_tvec, double* pose, uint8_t* inliers, int* info);

/* ---- second detector: DistributedOrbOpenCvDetector, SURVEY 8(f) row f1 ---------------------------------------- */
/* Replaces DistributedOrbOpenCvDetector::detect (src/lib/modular_slam/distributed_cv_feature.cpp:1189-1222: toGrayScale,
 * then OrbExtractorPimpl::extract :719-809), the detector the Viewer app wires in (src/app/viewer/viewer.cpp:183):
 * chained cv::resize INTER_LINEAR pyramid (:830-841), cv::FAST per 64-px cell with threshold 20 then 7 (:843-939), quad-tree
 * keypoint distribution (:941-1106), intensity-centroid angle (:543-570), fixed-point 7x7 Gaussian blur (:797-798) and
 * rotated BRIEF with the polynomial cos/sin (:456-503, :572-629).  Keypoints come out in the reference's order (level,
 * then the tree's final list order), so index i is the reference's Keypoint::id (:1201-1210).
 * It is a separate handle: its pyramid geometry differs from cv::ORB's (round(W / scale_l), float product-chain scales). */
typedef struct mslam_b200_dist mslam_b200_dist;
/* Defaults = orb_params("orb", 1.2f, 8, 20, 7) and OrbExtractorPimpl(params, 1000)  (distributed_cv_feature.cpp:1184-1186) */
typedef struct mslam_b200_dist_config
{
    int struct_size;        /* = sizeof(mslam_b200_dist_config) */
    int device;
    int max_width;          /* largest frame the handle accepts */
    int max_height;
    int max_batch;
    int nlevels;            /* orb_params::num_levels_ */
    float scale_factor;     /* orb_params::scale_factor_ */
    int ini_fast_threshold; /* orb_params::ini_fast_thr_ */
    int min_fast_threshold; /* orb_params::min_fast_thr_ */
    int min_size;           /* OrbExtractorPimpl::min_size_: a tree node stops splitting at area * scale^2 <= min_size */
    int max_keypoints;      /* per-frame capacity inside the handle; 0 => the tree's own upper bound for max_width x max_height */
    void* stream;           /* optional cudaStream_t; NULL => handle-owned stream */
    int flags;
    int reserved;
} mslam_b200_dist_config;
MSLAM_B200_API void mslam_b200_dist_default_config(mslam_b200_dist_config* cfg);
MSLAM_B200_API int mslam_b200_dist_create(const mslam_b200_dist_config* cfg, mslam_b200_dist** out);
MSLAM_B200_API void mslam_b200_dist_destroy(mslam_b200_dist* d);
MSLAM_B200_API int mslam_b200_dist_synchronize(mslam_b200_dist* d);
MSLAM_B200_API const char* mslam_b200_dist_last_cuda_error(mslam_b200_dist* d);
MSLAM_B200_API long long mslam_b200_dist_launch_count(mslam_b200_dist* d);
/* per-frame keypoint capacity the handle was created with (size outputs with it) */
MSLAM_B200_API int mslam_b200_dist_keypoint_capacity(mslam_b200_dist* d);
/* per-level sizes / scales for a frame size (compute_image_pyramid, :830-841); arrays of nlevels */
MSLAM_B200_API int mslam_b200_dist_level_geometry(mslam_b200_dist* d, int width, int height, int* w, int* h, float* scale);
/* One frame (host pointers, synchronous).  `bgr` is RgbFrame::data; keypoint.response is the FAST score, .size is
 * (float)(unsigned)(31 * scale_l), x / y are level coordinates times scale_l (:1117-1130).  MSLAM_B200_E_CAPACITY if the
 * count exceeds `capacity` (n_out still reports it).  Frames whose smallest level is under 38 px on a side are
 * MSLAM_B200_E_INVALID (the reference's unsigned border arithmetic underflows there). */
MSLAM_B200_API int mslam_b200_dist_detect(mslam_b200_dist* d, const uint8_t* bgr, int width, int height, int stride_bytes,
                                          mslam_b200_keypoint* out_kp, uint8_t* out_desc, int capacity, int* n_out);
/* A batch of independent frames of one size; outputs [n_frames][capacity], counts [n_frames] */
MSLAM_B200_API int mslam_b200_dist_detect_batch(mslam_b200_dist* d, const uint8_t* const* frames, int n_frames, int width,
                                                int height, int stride_bytes, mslam_b200_keypoint* out_kp,
                                                uint8_t* out_desc, int capacity, int* n_out);
/* pipelined form, as mslam_b200_detect_batch_begin / _end: up to two batches in flight, _end waits for the oldest */
MSLAM_B200_API int mslam_b200_dist_detect_batch_begin(mslam_b200_dist* d, const uint8_t* const* frames, int n_frames,
                                                      int width, int height, int stride_bytes, mslam_b200_keypoint* out_kp,
                                                      uint8_t* out_desc, int capacity, int* n_out);
MSLAM_B200_API int mslam_b200_dist_detect_batch_end(mslam_b200_dist* d);
/* Device-resident variant (asynchronous on the handle's stream): d_bgr [n_frames][height][width][3] dense */
MSLAM_B200_API int mslam_b200_dist_detect_batch_dev(mslam_b200_dist* d, const uint8_t* d_bgr, int n_frames, int width,
                                                    int height, mslam_b200_keypoint* d_kp, uint8_t* d_desc, int capacity,
                                                    int* d_counts);
typedef enum mslam_b200_dist_stage
{
    MSLAM_B200_DIST_PYRAMID = 0, /* u8 level image, dense w_l*h_l */
    MSLAM_B200_DIST_BLURRED = 1, /* u8 blurred level image */
    MSLAM_B200_DIST_N_FAST = 2,  /* int: keypts_to_distribute of the level */
    MSLAM_B200_DIST_N_TREE = 3,  /* int: keypoints the tree kept on the level */
    MSLAM_B200_DIST_PASSES = 4   /* int: passes of the tree's while(true) loop */
} mslam_b200_dist_stage;
/* copy an intermediate of the LAST detect call (frame index in that batch, level) to host memory */
MSLAM_B200_API int mslam_b200_dist_debug_read(mslam_b200_dist* d, int stage, int frame, int level, void* out, size_t bytes);
/* per-stage device time of the calls since the last read: ms[6] = gray, resize, fast cells, tree, blur, describe
 * (CUDA events on the handle's stream; enable = 0/1 switches the bracketing, ms may be NULL) */
MSLAM_B200_API int mslam_b200_dist_stage_times(mslam_b200_dist* d, int enable, double* ms6);

/* ---- measurement / parity aids ------------------------------------------------------------------------------- */
typedef enum mslam_b200_stage
{
    MSLAM_B200_STAGE_PYRAMID = 0,  /* u8 level image, dense w_l*h_l */
    MSLAM_B200_STAGE_BLURRED = 1,  /* u8 blurred level image */
    MSLAM_B200_STAGE_N_FAST = 2,   /* int: NMS+border survivors of the level */
    MSLAM_B200_STAGE_N_CUT1 = 3,   /* int: survivors of retainBest(2*n_l) on the FAST score */
    MSLAM_B200_STAGE_N_FINAL = 4   /* int: keypoints kept on the level */
} mslam_b200_stage;
/* copy an intermediate of the LAST detect call (frame index in that batch, level) to host memory */
MSLAM_B200_API int mslam_b200_debug_read(mslam_b200_ctx* ctx, int stage, int frame, int level, void* out, size_t bytes);
/* Per-stage device timing (tracing aid for bench.py's roofline): when enabled, every stage of the detect / match
 * entry points is bracketed by CUDA events on the context's stream.  mslam_b200_stage_times synchronises the
 * stream, folds the recorded intervals into per-stage totals since the last call and resets them. */
typedef enum mslam_b200_timed_stage
{
    MSLAM_B200_T_GRAY = 0,     /* toGrayScale -> level 0 */
    MSLAM_B200_T_RESIZE = 1,   /* chained INTER_LINEAR_EXACT levels 1..n-1 */
    MSLAM_B200_T_FAST = 2,     /* FAST-9 score + NMS + border filter, all levels */
    MSLAM_B200_T_SELECT = 3,   /* retainBest x2 + Harris */
    MSLAM_B200_T_BLUR = 4,     /* 7x7 float blur, all levels */
    MSLAM_B200_T_DESCRIBE = 5, /* IC angle + rBRIEF */
    MSLAM_B200_T_KNN = 6,      /* brute-force Hamming kNN (k = 2) */
    MSLAM_B200_T_COMPACT = 7,  /* ratio test / mutual check + ordered compaction */
    MSLAM_B200_T_DB = 8,       /* keyframe-database scoring */
    MSLAM_B200_T_COUNT = 9
} mslam_b200_timed_stage;
MSLAM_B200_API int mslam_b200_enable_stage_timing(mslam_b200_ctx* ctx, int enable);
/* total_ms[MSLAM_B200_T_COUNT], intervals[MSLAM_B200_T_COUNT] (number of bracketed intervals), launches[..] (kernel
 * launches inside them) */
MSLAM_B200_API int mslam_b200_stage_times(mslam_b200_ctx* ctx, double* total_ms, long long* intervals,
                                          long long* launches);
/* integer-pipe roofline denominators, measured on the context's device with register-only loops over all SMs:
 * out[0] = POPC.32 results/s, out[1] = LOP3 results/s, out[2] = 256-bit pair evaluations/s of the plain
 * 8x(xor,popc,add) + top-2 sequence, out[3] = SM clock (MHz) the loops ran at (from clock64/globaltimer),
 * out[4] = pair evaluations/s of the carry-save sequence the matcher uses (8 xor, 8 lop3, 4 popc) + top-2,
 * out[5] = LOP3 results/s when all three operands are live registers (12 independent chains per thread). */
MSLAM_B200_API int mslam_b200_measure_int_peaks(mslam_b200_ctx* ctx, double* out6);
/* int8 tensor-core roofline denominator of the tensor-core matcher (csrc/knn_mma.cu), measured on the context's device:
 * tcgen05.mma.cta_group::1.kind::i8 issued back to back on every SM from operands resident in shared memory, accumulators
 * in tensor memory, no loads and no epilogue.  out[0] / out[1] = int8 operations per second (2 M N K per instruction) of
 * the M128 N128 K32 shape the matcher issues / of M128 N256 K32, SUSTAINED (the last quarter of ~0.4 s of back-to-back
 * launches: the power-limited clock a long scoring step runs at); out[2] / out[3] = the same for one short launch (burst). */
MSLAM_B200_API int mslam_b200_measure_tensor_peak(mslam_b200_ctx* ctx, double* out4);

#ifdef __cplusplus
}
#endif
#endif /* MSLAM_B200_H_ */
