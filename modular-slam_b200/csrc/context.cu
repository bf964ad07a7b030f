// context.cu -- host side of the C ABI declared in include/mslam_b200.h: context, HBM layout, launch
// sequencing.  The reference interfaces each entry point replaces are cited in the header.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <algorithm>
#include <new>
#include <vector>

#include "common.cuh"
#include "kernels.h"
#include "nccl_dyn.h"

using namespace mslam_b200;

constexpr int kDefaultChunks = 4;
constexpr int kMaxChunks = 8; // host detect_batch: most chunks in flight through copy-in / kernels / copy-out

struct mslam_b200_ctx
{
    mslam_b200_config cfg;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    long long launches = 0;
    char err[512] = {0};

    // geometry of the frame size currently configured (cv::ORB layer table)
    int geo_w = 0, geo_h = 0;
    PipeParams P;
    FastMaps fast_maps; // tensor maps of k_fast's tile loads for the current geometry
    PipeParams P_few;   // the same geometry cut into short FAST tiles (calls of one or two frames), and its tensor maps
    FastMaps fast_maps_few;
    int kp_cap = 0, desc_cap = 0;

    // ---- HBM: detector ----
    uint8_t* d_bgr = nullptr;  // [max_batch][H][W][3]
    uint8_t* d_pyr = nullptr;  // level-major planes, see LevelDesc
    uint8_t* d_blur = nullptr; // same layout
    size_t pyr_bytes = 0, bgr_bytes = 0;
    short2 *d_xtab = nullptr, *d_ytab = nullptr;
    uint4* d_qtab = nullptr; // resize quad table (k_resize_quads)
    int tab_cap = 0;
    unsigned *d_cand = nullptr, *d_cand2 = nullptr, *d_keys = nullptr, *d_keys2 = nullptr;
    long long cand_elems = 0; // per frame
    uint8_t* d_zero = nullptr; // one block, zeroed per batch: counters | hist | err
    size_t zero_bytes = 0;
    LevelCounters* d_ctr = nullptr;
    unsigned* d_hist = nullptr;
    int* d_err = nullptr;
    SelRec* d_sel = nullptr;
    mslam_b200_keypoint* d_kp = nullptr;
    uint8_t* d_desc = nullptr;
    int* d_counts = nullptr;
    float4* d_pat = nullptr; // rBRIEF pattern as floats, transposed (256 x float4)
    uint2* d_icw = nullptr;  // intensity-centroid byte weights (ic_weight_table)
    // ---- pinned host staging ----
    int* h_counts = nullptr;
    int* h_err = nullptr;
    int last_frames = 0;
    // ---- host pipeline: up to two batches in flight (mslam_b200_detect_batch_begin / _end) ----
    struct PendingBatch
    {
        bool active = false;
        int n_frames = 0, capacity = 0;
        int* n_out = nullptr;
    } pend[2];
    int pend_next = 0;
    cudaEvent_t ev_slot_done[2] = {nullptr, nullptr}, ev_out_done = nullptr, ev_match_done = nullptr,
                ev_match_in = nullptr, ev_match_k = nullptr;
    bool match_pending = false;
    bool out_done_valid = false;
    // ---- single-frame kernel graph ----
    // one frame is ~16 small launches on three streams: enqueueing them costs more host time than they run for, so the
    // single-frame calls replay a captured graph of the same launches (one per frame slot; re-captured on a new geometry)
    struct FrameGraph
    {
        cudaGraphExec_t exec = nullptr;
        int width = 0, height = 0, uses = 0;
        long long launches = 0;
    } graph1[2];
    int graphs_enabled = -1; // -1: read MSLAM_B200_GRAPH on first use
    // ---- single-frame latency path (mslam_b200_detect with nothing else in flight) ----
    // one result block per frame: [count, err, pad, pad | kp_cap keypoints | kp_cap descriptors], written by the kernels of a
    // captured graph that also clears the frame's counters first and copies the block to its pinned mirror last, so a call is
    // copy-in + graph launch + one wait on ONE stream (the general path hops through two copy streams and four D2H copies).
    // (Measured and not kept: the describe kernel writing straight into mapped pinned memory instead of the copy node --
    // the kernel grows from 10 to 14 us of posted PCIe writes, which is what the 75 KB copy node costs.)
    uint8_t *d_one = nullptr, *h_one = nullptr;
    size_t one_bytes = 0, one_kp_off = 0, one_desc_off = 0;
    FrameGraph graph_lat;
    // ---- single-pair matcher call (mslam_b200_match_knn_ratio / _crosscheck): one staged block in, one block out ----
    // in:  [n_from, n_to, pad.. (32 B) | from descriptors, packed | to descriptors, packed]   (pinned mirror h_m1)
    // out: [n_out, pad.. (16 B) | a[desc_cap] | b[desc_cap] | dist[desc_cap]]               (pinned mirror h_mo)
    uint8_t *d_m1 = nullptr, *h_m1 = nullptr, *d_mo = nullptr, *h_mo = nullptr;
    size_t m1_bytes = 0, mo_bytes = 0;
    const int *lat_a = nullptr, *lat_b = nullptr, *lat_d = nullptr; // where the last single-pair call left its results
    // ---- HBM: matcher ----
    uint8_t *d_mfrom = nullptr, *d_mto = nullptr, *d_raw = nullptr;
    int *d_nfrom = nullptr, *d_nto = nullptr;
    unsigned *d_key0 = nullptr, *d_key1 = nullptr, *d_key0b = nullptr, *d_key1b = nullptr;
    unsigned *d_part0 = nullptr, *d_part1 = nullptr; // per-slice partial top-2 keys of the small-batch kNN
    uint8_t *d_xa = nullptr, *d_xb = nullptr;        // +1/-1 int8 expansions of the two descriptor sets (knn_mma.cu)
    uint8_t* d_pnp = nullptr;                        // scratch of the host back-projection / RANSAC-PnP entry points, grow-only
    size_t pnp_bytes = 0;
    uint8_t* h_pnp = nullptr;                        // pinned staging of the RANSAC-PnP outputs, grow-only
    size_t h_pnp_bytes = 0;
    uint8_t *d_xdb = nullptr, *d_xqf = nullptr;      // ... of a chunk of keyframes / of the query frames (DB scoring), lazily sized
    size_t xdb_bytes = 0, xqf_bytes = 0;
    int knn_mma = 1;                                 // MSLAM_KNN_MMA: 0 scalar kernel only, 1 by problem size, 2 always
    int knn_mma_env = -1;                            // what MSLAM_KNN_MMA said at create (-1: unset)
    // ---- sharded keyframe database (mslam_b200_db_attach* / _db_query_sharded_dev) ----
    ncclComm_t db_comm = nullptr;
    bool db_own_comm = false;
    int db_rank = 0, db_world = 1;
    int* d_dbs = nullptr;                            // grow-only: [Q][n_local] scores | local top-k [2][Q][k] | gathered [world][2][Q][k]
    size_t dbs_bytes = 0;
    uint8_t* d_dbq = nullptr;                        // device copy of the host entry point's query frames / outputs, grow-only
    size_t dbq_bytes = 0;
    uint8_t* d_shard = nullptr;                      // context-owned resident shard (mslam_b200_db_load_shard): descriptors | counts
    size_t shard_bytes = 0;
    int shard_n = 0, shard_cap = 0, shard_id_base = 0;
    int *d_out_a = nullptr, *d_out_b = nullptr, *d_out_d = nullptr, *d_nout = nullptr;
    int *h_out_a = nullptr, *h_out_b = nullptr, *h_out_d = nullptr, *h_nout = nullptr, *h_n2 = nullptr;
    // ---- copy engines of the host entry points (chunked H2D / compute / D2H pipeline) ----
    cudaStream_t copy[2] = {nullptr, nullptr};
    cudaEvent_t ev_h2d[kMaxChunks] = {nullptr}, ev_done[kMaxChunks + 1] = {nullptr};
    // ---- forked streams of the detect pipeline ----
    cudaStream_t aux[2] = {nullptr, nullptr};
    cudaEvent_t ev_level[kMaxLevels] = {nullptr};
    cudaEvent_t ev_join[2] = {nullptr, nullptr};
    cudaEvent_t ev_fast0 = nullptr; // few-frames structure: FAST of level 0 done (it runs on the blur fork)
    // ---- per-stage event timing (mslam_b200_enable_stage_timing) ----
    bool timing = false;
    bool timing_in_graph = false; // the stage events become real record nodes of a graph being captured (latency trace)
    std::vector<cudaEvent_t> ev_pool; // pairs: [2i] start, [2i+1] stop
    std::vector<int> ev_stage;        // stage of pair i
    std::vector<long long> ev_launches;
    size_t ev_used = 0;               // pairs in use
    double t_ms[MSLAM_B200_T_COUNT] = {0};
    long long t_n[MSLAM_B200_T_COUNT] = {0}, t_l[MSLAM_B200_T_COUNT] = {0};
};

// RAII bracket of one timed stage; no-op unless timing is enabled
struct StageTimer
{
    mslam_b200_ctx* c;
    long long l0;
    int pair = -1;
    cudaStream_t st;
    StageTimer(mslam_b200_ctx* ctx, int stage, cudaStream_t on = nullptr) : c(ctx), l0(ctx->launches), st(on ? on : ctx->stream)
    {
        if(!c->timing || c->ev_used >= c->ev_stage.size())
            return;
        pair = (int)c->ev_used++;
        c->ev_stage[pair] = stage;
        if(c->timing_in_graph)
            cudaEventRecordWithFlags(c->ev_pool[2 * pair], st, cudaEventRecordExternal);
        else
            cudaEventRecord(c->ev_pool[2 * pair], st);
    }
    ~StageTimer()
    {
        if(pair < 0)
            return;
        if(c->timing_in_graph)
            cudaEventRecordWithFlags(c->ev_pool[2 * pair + 1], st, cudaEventRecordExternal);
        else
            cudaEventRecord(c->ev_pool[2 * pair + 1], st);
        c->ev_launches[pair] = c->launches - l0;
    }
};

static int ctx_fail(mslam_b200_ctx* ctx, cudaError_t e, const char* what)
{
    if(ctx)
        snprintf(ctx->err, sizeof(ctx->err), "%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
    return MSLAM_B200_E_CUDA;
}

static inline size_t align_up(size_t v, size_t a)
{
    return (v + a - 1) / a * a;
}

// ---- cv::ORB layer geometry (OpenCV orb.cpp getScale / detectAndCompute / computeKeyPoints) ----
static void level_geometry(int W, int H, int nlevels, float scale_factor, int* w, int* h, float* scale)
{
    for(int l = 0; l < nlevels; ++l)
    {
        const float s = (float)pow((double)scale_factor, (double)l);
        scale[l] = s;
        const float inv = 1.f / s; // cv::ORB multiplies by the f32 reciprocal (237 rows -> 198 at level 1, not 197)
        w[l] = (int)lrintf((float)W * inv);
        h[l] = (int)lrintf((float)H * inv);
    }
}

static void features_per_level(int nfeatures, int nlevels, float scale_factor, int* n_l)
{
    const float factor = (float)(1.0 / (double)scale_factor);
    float nd = (float)nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
    int sum = 0;
    for(int l = 0; l < nlevels - 1; ++l)
    {
        n_l[l] = (int)lrintf(nd);
        sum += n_l[l];
        nd *= factor;
    }
    n_l[nlevels - 1] = std::max(nfeatures - sum, 0);
}

// INTER_LINEAR_EXACT taps: IEEE double, no FMA (file compiled with -ffp-contract=off)
static void resize_taps(int src, int dst, short2* tab)
{
    const double inv = (double)dst / (double)src;
    const double scale = 1.0 / inv;
    for(int v = 0; v < dst; ++v)
    {
        const double f = scale * ((double)v + 0.5) - 0.5;
        int i = (int)floor(f);
        int a1 = 0;
        if(i < 0)
            i = 0;
        else if(i >= src - 1)
            i = src - 1;
        else
            a1 = (int)lrint((f - (double)i) * 256.0);
        tab[v].x = (short)i;
        tab[v].y = (short)a1;
    }
}

static int reflect101_host(int p, int n)
{
    if(n == 1)
        return 0;
    while(p < 0 || p >= n)
        p = p < 0 ? -p : 2 * n - 2 - p;
    return p;
}

// Quad table of k_resize_quads for one level: entry q covers output columns 4q .. 4q+3 (columns >= dw are the
// reflect-101 padding the blur reads).  {x: first source byte of the 8-byte window, y: PRMT selectors (pixels 0,1 |
// pixels 2,3 << 16), z: alpha1 of pixels 0,1, w: alpha1 of pixels 2,3}.  Returns false when a quad's taps do not fit
// the window (scale factors > 2), in which case the level uses the generic kernel.
static bool resize_quads(const short2* xtab, int dw, std::vector<uint4>* out)
{
    const int nq = (dw + 4 + 3) / 4;
    for(int q = 0; q < nq; ++q)
    {
        int i0[4], a1[4], base = 1 << 30;
        for(int j = 0; j < 4; ++j)
        {
            const int x = 4 * q + j;
            const short2 t = xtab[x < dw ? x : reflect101_host(x, dw)];
            i0[j] = t.x;
            a1[j] = t.y;
            base = std::min(base, i0[j]);
        }
        unsigned sel[4];
        for(int j = 0; j < 4; ++j)
        {
            const int o = i0[j] - base;
            if(o + 1 > 7)
                return false;
            sel[j] = (unsigned)o | ((unsigned)(o + 1) << 4);
        }
        uint4 e;
        e.x = (unsigned)base;
        e.y = (sel[0] | (sel[1] << 8)) | ((sel[2] | (sel[3] << 8)) << 16);
        e.z = (unsigned)a1[0] | ((unsigned)a1[1] << 16);
        e.w = (unsigned)a1[2] | ((unsigned)a1[3] << 16);
        out->push_back(e);
    }
    return true;
}

struct Geometry
{
    PipeParams P;
    size_t pyr_bytes;
    long long cand_elems;
    int tab_elems;
    std::vector<short2> xtab, ytab;
    std::vector<uint4> qtab;
};

static void build_geometry(const mslam_b200_config& cfg, int kp_cap, int W, int H, Geometry* g, bool short_tiles = false)
{
    PipeParams& P = g->P;
    memset(&P, 0, sizeof(P));
    P.nlevels = cfg.nlevels;
    P.W = W;
    P.H = H;
    P.edge = cfg.edge_threshold;
    P.fast_thr = cfg.fast_threshold;
    P.kp_cap = kp_cap;
    P.one = 1;
    P.blur_band = blur_band_rows();
    P.minus_one = -1;
    int w[kMaxLevels], h[kMaxLevels], nl[kMaxLevels];
    float sc[kMaxLevels];
    level_geometry(W, H, cfg.nlevels, cfg.scale_factor, w, h, sc);
    features_per_level(cfg.nfeatures, cfg.nlevels, cfg.scale_factor, nl);
    size_t off = 0;
    long long coff = 0;
    int toff_x = 0, toff_y = 0, tiles = 0, bwarps = 0;
    g->xtab.clear();
    g->qtab.clear();
    g->ytab.clear();
    for(int l = 0; l < cfg.nlevels; ++l)
    {
        LevelDesc& L = P.lv[l];
        L.w = std::max(w[l], 1);
        L.h = std::max(h[l], 1);
        L.pitch = (int)align_up((size_t)L.w + 12, 64); // the 4 reflect-101 padding columns + aligned over-reads (resize window: 12 B)
        L.n_keep = nl[l];
        L.scale = sc[l];
        L.inv_scale = 1.f / sc[l];
        L.img_off = (long long)off;
        L.img_frame_stride = (long long)align_up((size_t)L.pitch * L.h, 256);
        off += (size_t)L.img_frame_stride * cfg.max_batch;
        const int wi = L.w - 2 * cfg.edge_threshold, hi = L.h - 2 * cfg.edge_threshold;
        L.cand_cap = (wi > 0 && hi > 0) ? ((wi + 1) / 2) * ((hi + 1) / 2) : 0;
        L.cand_off = coff;
        coff += (long long)align_up((size_t)L.cand_cap, 4);
        fast_tiles(L.w, L.h, cfg.edge_threshold, short_tiles, &L.tiles_x, &L.tiles_y, &L.tile_sh);
        L.tile_base = tiles;
        tiles += L.tiles_x * L.tiles_y;
        int bw = 0;
        blur_warps(L.w, L.h, &L.blur_xblocks, &bw);
        L.blur_warp_base = bwarps;
        bwarps += bw;
        L.xtab_off = toff_x;
        L.ytab_off = toff_y;
        if(l > 0)
        {
            g->xtab.resize(toff_x + L.w);
            g->ytab.resize(toff_y + L.h);
            resize_taps(P.lv[l - 1].w, L.w, g->xtab.data() + toff_x);
            resize_taps(P.lv[l - 1].h, L.h, g->ytab.data() + toff_y);
            L.qtab_off = (int)g->qtab.size();
            const size_t before = g->qtab.size();
            if(resize_quads(g->xtab.data() + toff_x, L.w, &g->qtab))
                L.rs_quads = (int)(g->qtab.size() - before);
            else
            {
                g->qtab.resize(before);
                L.rs_quads = 0;
            }
            toff_x += L.w;
            toff_y += L.h;
        }
    }
    P.total_fast_tiles = tiles;
    P.cand_frame_stride = std::max<long long>(coff, 4);
    g->pyr_bytes = off;
    g->cand_elems = P.cand_frame_stride;
    g->tab_elems = std::max(toff_x, toff_y) + 1;
}

static int set_geometry(mslam_b200_ctx* ctx, int W, int H)
{
    if(W == ctx->geo_w && H == ctx->geo_h)
        return MSLAM_B200_OK;
    if(W < 1 || H < 1 || W > ctx->cfg.max_width || H > ctx->cfg.max_height || W > 4095 || H > 4095)
        return MSLAM_B200_E_INVALID;
    Geometry g;
    build_geometry(ctx->cfg, ctx->kp_cap, W, H, &g);
    if(g.pyr_bytes > ctx->pyr_bytes || g.cand_elems > ctx->cand_elems || g.tab_elems > ctx->tab_cap)
        return MSLAM_B200_E_INVALID; // cannot happen for W,H <= max (geometry is monotone); guard anyway
    if(!g.xtab.empty())
    {
        MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->d_xtab, g.xtab.data(), g.xtab.size() * sizeof(short2), cudaMemcpyHostToDevice,
                                      ctx->stream));
        MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->d_ytab, g.ytab.data(), g.ytab.size() * sizeof(short2), cudaMemcpyHostToDevice,
                                      ctx->stream));
        if(!g.qtab.empty())
            MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->d_qtab, g.qtab.data(), g.qtab.size() * sizeof(uint4), cudaMemcpyHostToDevice,
                                          ctx->stream));
        MSLAM_CUDA_OK(cudaStreamSynchronize(ctx->stream)); // the host vectors die with this scope
    }
    MSLAM_CUDA_OK(build_fast_maps(g.P, ctx->d_pyr, ctx->cfg.max_batch, &ctx->fast_maps));
    ctx->P = g.P;
    {
        Geometry gs;
        build_geometry(ctx->cfg, ctx->kp_cap, W, H, &gs, true);
        MSLAM_CUDA_OK(build_fast_maps(gs.P, ctx->d_pyr, ctx->cfg.max_batch, &ctx->fast_maps_few));
        ctx->P_few = gs.P;
    }
    ctx->geo_w = W;
    ctx->geo_h = H;
    return MSLAM_B200_OK;
}

// ------------------------------------------------------------------------------------------------------------------
extern "C" {

int mslam_b200_version(void)
{
    return MSLAM_B200_VERSION;
}

const char* mslam_b200_status_string(int s)
{
    switch(s)
    {
    case MSLAM_B200_OK: return "ok";
    case MSLAM_B200_E_INVALID: return "invalid argument";
    case MSLAM_B200_E_CUDA: return "CUDA error (no device, or a runtime failure; see mslam_b200_last_cuda_error)";
    case MSLAM_B200_E_CAPACITY: return "output or candidate capacity exceeded";
    case MSLAM_B200_E_NOMEM: return "out of memory";
    default: return "unknown status";
    }
}

void mslam_b200_default_config(mslam_b200_config* cfg)
{
    if(!cfg)
        return;
    memset(cfg, 0, sizeof(*cfg));
    cfg->struct_size = (int)sizeof(*cfg);
    cfg->device = 0;
    cfg->max_width = 640;
    cfg->max_height = 480;
    cfg->max_batch = 1;
    cfg->nfeatures = 1000; // cv::ORB::create(1000), orb_feature.cpp:25
    cfg->nlevels = 8;
    cfg->scale_factor = 1.2f;
    cfg->fast_threshold = 20;
    cfg->edge_threshold = 31;
    cfg->max_keypoints = 0;
    cfg->max_descriptors = 0;
    cfg->ratio = 0.7; // orb_feature.cpp:101
    cfg->stream = nullptr;
}

const char* mslam_b200_last_cuda_error(mslam_b200_ctx* ctx)
{
    return ctx ? ctx->err : "";
}

long long mslam_b200_launch_count(mslam_b200_ctx* ctx)
{
    return ctx ? ctx->launches : 0;
}

int mslam_b200_host_alloc(void** ptr, size_t bytes)
{
    if(!ptr)
        return MSLAM_B200_E_INVALID;
    return cudaHostAlloc(ptr, bytes, cudaHostAllocDefault) == cudaSuccess ? MSLAM_B200_OK : MSLAM_B200_E_NOMEM;
}

int mslam_b200_host_free(void* ptr)
{
    return cudaFreeHost(ptr) == cudaSuccess ? MSLAM_B200_OK : MSLAM_B200_E_CUDA;
}

void mslam_b200_destroy(mslam_b200_ctx* ctx)
{
    if(!ctx)
        return;
    cudaSetDevice(ctx->cfg.device);
    if(ctx->stream)
        cudaStreamSynchronize(ctx->stream);
    mslam_b200_db_detach(ctx);
    if(ctx->d_dbs)
        cudaFree(ctx->d_dbs);
    if(ctx->d_dbq)
        cudaFree(ctx->d_dbq);
    if(ctx->d_shard)
        cudaFree(ctx->d_shard);
    void* dev[] = {ctx->d_qtab, ctx->d_bgr,  ctx->d_pyr,   ctx->d_blur,  ctx->d_xtab,  ctx->d_ytab,  ctx->d_cand,  ctx->d_cand2,
                   ctx->d_keys, ctx->d_keys2, ctx->d_zero,  ctx->d_sel,   ctx->d_kp,    ctx->d_desc,  ctx->d_counts,
                   ctx->d_mfrom, ctx->d_mto,  ctx->d_raw,   ctx->d_nfrom, ctx->d_nto,   ctx->d_key0,  ctx->d_key1,
                   ctx->d_part0, ctx->d_part1, ctx->d_key0b, ctx->d_key1b, ctx->d_out_a, ctx->d_out_b, ctx->d_out_d, ctx->d_nout, ctx->d_pat, ctx->d_icw, ctx->d_xa, ctx->d_xb, ctx->d_xdb, ctx->d_xqf, ctx->d_pnp, ctx->d_one, ctx->d_m1, ctx->d_mo};
    for(void* p : dev)
        if(p)
            cudaFree(p);
    void* host[] = {ctx->h_counts, ctx->h_err, ctx->h_out_a, ctx->h_out_b, ctx->h_out_d, ctx->h_nout, ctx->h_n2, ctx->h_pnp, ctx->h_one, ctx->h_m1, ctx->h_mo};
    for(void* p : host)
        if(p)
            cudaFreeHost(p);
    for(cudaEvent_t e : ctx->ev_pool)
        cudaEventDestroy(e);
    for(int i = 0; i < 2; ++i)
    {
        if(ctx->aux[i])
        {
            cudaStreamSynchronize(ctx->aux[i]);
            cudaStreamDestroy(ctx->aux[i]);
        }
        if(ctx->ev_join[i])
            cudaEventDestroy(ctx->ev_join[i]);
    }
    for(int i = 0; i < kMaxLevels; ++i)
        if(ctx->ev_level[i])
            cudaEventDestroy(ctx->ev_level[i]);
    for(int i = 0; i < 2; ++i)
        if(ctx->copy[i])
        {
            cudaStreamSynchronize(ctx->copy[i]);
            cudaStreamDestroy(ctx->copy[i]);
        }
    for(int i = 0; i < kMaxChunks; ++i)
        if(ctx->ev_h2d[i])
            cudaEventDestroy(ctx->ev_h2d[i]);
    for(int i = 0; i <= kMaxChunks; ++i)
        if(ctx->ev_done[i])
            cudaEventDestroy(ctx->ev_done[i]);
    for(cudaEvent_t e : {ctx->ev_slot_done[0], ctx->ev_slot_done[1], ctx->ev_out_done, ctx->ev_match_done, ctx->ev_match_in, ctx->ev_match_k, ctx->ev_fast0})
        if(e)
            cudaEventDestroy(e);
    for(auto& g : ctx->graph1)
        if(g.exec)
            cudaGraphExecDestroy(g.exec);
    if(ctx->graph_lat.exec)
        cudaGraphExecDestroy(ctx->graph_lat.exec);
    if(ctx->own_stream && ctx->stream)
        cudaStreamDestroy(ctx->stream);
    delete ctx;
}

// kernel choice of the matcher from cfg.flags (and MSLAM_KNN_MMA as read at create); allocates the +1/-1 expansion buffers the
// tensor-core kernel needs the first time it becomes selectable
static int apply_knn_flags(mslam_b200_ctx* ctx)
{
    ctx->knn_mma = ctx->knn_mma_env >= 0 ? ctx->knn_mma_env : 1;
    if(ctx->cfg.flags & MSLAM_B200_FLAG_KNN_SCALAR)
        ctx->knn_mma = 0;
    if(ctx->cfg.flags & MSLAM_B200_FLAG_KNN_TENSOR)
        ctx->knn_mma = 2;
    if(ctx->knn_mma && !ctx->d_xa)
    {
        const size_t bytes = knn_mma_bytes(ctx->cfg.max_batch, ctx->desc_cap);
        MSLAM_CUDA_OK(cudaMalloc(&ctx->d_xa, bytes));
        MSLAM_CUDA_OK(cudaMalloc(&ctx->d_xb, bytes));
    }
    return MSLAM_B200_OK;
}

int mslam_b200_create(const mslam_b200_config* cfg_in, mslam_b200_ctx** out_ctx)
{
    if(!cfg_in || !out_ctx || cfg_in->struct_size != (int)sizeof(mslam_b200_config))
        return MSLAM_B200_E_INVALID;
    *out_ctx = nullptr;
    const mslam_b200_config& c = *cfg_in;
    if(c.max_width < 1 || c.max_height < 1 || c.max_width > 4095 || c.max_height > 4095 || c.max_batch < 1 ||
       c.max_batch > 65535 || c.nfeatures < 1 || c.nlevels < 1 || c.nlevels > kMaxLevels || !(c.scale_factor > 1.f) ||
       c.fast_threshold < 1 || c.fast_threshold > 254 || c.edge_threshold < 31 || !(c.ratio > 0.0))
        return MSLAM_B200_E_INVALID;
    int ndev = 0;
    if(cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1 || c.device < 0 || c.device >= ndev)
        return MSLAM_B200_E_CUDA; // no CPU fallback by design
    mslam_b200_ctx* ctx = new(std::nothrow) mslam_b200_ctx();
    if(!ctx)
        return MSLAM_B200_E_NOMEM;
    ctx->cfg = c;
    ctx->kp_cap = c.max_keypoints > 0 ? c.max_keypoints : c.nfeatures + c.nfeatures / 4 + 64;
    ctx->desc_cap = c.max_descriptors > 0 ? c.max_descriptors : ctx->kp_cap;
    if(ctx->desc_cap > 65535)
    {
        delete ctx;
        return MSLAM_B200_E_INVALID; // train index is packed into 16 bits of the kNN key
    }
#define CREATE_OK(expr)                                                                                                \
    do                                                                                                                 \
    {                                                                                                                  \
        cudaError_t e__ = (expr);                                                                                      \
        if(e__ != cudaSuccess)                                                                                         \
        {                                                                                                              \
            fprintf(stderr, "mslam_b200_create: %s failed: %s\n", #expr, cudaGetErrorString(e__));                     \
            const int rc__ = e__ == cudaErrorMemoryAllocation ? MSLAM_B200_E_NOMEM : MSLAM_B200_E_CUDA;                \
            mslam_b200_destroy(ctx);                                                                                   \
            return rc__;                                                                                               \
        }                                                                                                              \
    } while(0)
    CREATE_OK(cudaSetDevice(c.device));
    cudaDeviceProp prop;
    CREATE_OK(cudaGetDeviceProperties(&prop, c.device));
    ctx->sm_count = prop.multiProcessorCount;
    // dynamic shared-memory opt-ins are per device: set them for THIS context's device (a process may hold one context per GPU)
    CREATE_OK(configure_match_kernels());
    CREATE_OK(configure_mma_kernels());
    if(c.stream)
        ctx->stream = (cudaStream_t)c.stream;
    else
    {
        CREATE_OK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
        ctx->own_stream = true;
    }
    {
        // the FAST + selection fork is the long pole of a batch (its kernels are ALU-bound and 40 % of the step): its CTAs go
        // first, the blur fork fills what they leave.  MSLAM_B200_STREAM_PRIO=0 creates both forks at the default priority.
        int prio_lo = 0, prio_hi = 0;
        CREATE_OK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
        const char* e = getenv("MSLAM_B200_STREAM_PRIO");
        const int mode = e ? atoi(e) : 1; // A/B aid: bit 0 = the FAST fork high, bit 1 = the blur fork high
        for(int i = 0; i < 2; ++i)
        {
            CREATE_OK(cudaStreamCreateWithPriority(&ctx->aux[i], cudaStreamNonBlocking, ((mode >> i) & 1) ? prio_hi : prio_lo));
            CREATE_OK(cudaEventCreateWithFlags(&ctx->ev_join[i], cudaEventDisableTiming));
        }
    }
    for(int i = 0; i < kMaxLevels; ++i)
        CREATE_OK(cudaEventCreateWithFlags(&ctx->ev_level[i], cudaEventDisableTiming));
    CREATE_OK(cudaEventCreateWithFlags(&ctx->ev_fast0, cudaEventDisableTiming));
    for(int i = 0; i < 2; ++i)
        CREATE_OK(cudaStreamCreateWithFlags(&ctx->copy[i], cudaStreamNonBlocking));
    for(int i = 0; i < kMaxChunks; ++i)
        CREATE_OK(cudaEventCreateWithFlags(&ctx->ev_h2d[i], cudaEventDisableTiming));
    for(int i = 0; i <= kMaxChunks; ++i)
        CREATE_OK(cudaEventCreateWithFlags(&ctx->ev_done[i], cudaEventDisableTiming));
    Geometry g;
    build_geometry(c, ctx->kp_cap, c.max_width, c.max_height, &g);
    const size_t B = (size_t)c.max_batch;
    ctx->pyr_bytes = g.pyr_bytes;
    ctx->cand_elems = g.cand_elems;
    ctx->tab_cap = g.tab_elems + 64;
    ctx->bgr_bytes = B * c.max_width * c.max_height * 3;
    CREATE_OK(cudaMalloc(&ctx->d_bgr, 2 * ctx->bgr_bytes)); // two pipeline slots
    CREATE_OK(cudaMalloc(&ctx->d_pyr, ctx->pyr_bytes));
    CREATE_OK(cudaMalloc(&ctx->d_blur, ctx->pyr_bytes));
    CREATE_OK(cudaMalloc(&ctx->d_xtab, ctx->tab_cap * sizeof(short2)));
    CREATE_OK(cudaMalloc(&ctx->d_ytab, ctx->tab_cap * sizeof(short2)));
    CREATE_OK(cudaMalloc(&ctx->d_qtab, ctx->tab_cap * sizeof(uint4))); // <= (w + 7) / 4 quads per level, far below tab_cap
    const size_t cand_bytes = B * (size_t)ctx->cand_elems * sizeof(unsigned);
    CREATE_OK(cudaMalloc(&ctx->d_cand, cand_bytes));
    CREATE_OK(cudaMalloc(&ctx->d_cand2, cand_bytes));
    CREATE_OK(cudaMalloc(&ctx->d_keys, cand_bytes));
    CREATE_OK(cudaMalloc(&ctx->d_keys2, cand_bytes));
    const size_t ctr_bytes = align_up(B * c.nlevels * sizeof(LevelCounters), 256);
    const size_t hist_bytes = align_up(B * c.nlevels * 256 * sizeof(unsigned), 256);
    const size_t err_bytes = align_up(B * sizeof(int), 256);
    ctx->zero_bytes = ctr_bytes + hist_bytes + err_bytes;
    CREATE_OK(cudaMalloc(&ctx->d_zero, ctx->zero_bytes));
    ctx->d_ctr = reinterpret_cast<LevelCounters*>(ctx->d_zero);
    ctx->d_hist = reinterpret_cast<unsigned*>(ctx->d_zero + ctr_bytes);
    ctx->d_err = reinterpret_cast<int*>(ctx->d_zero + ctr_bytes + hist_bytes);
    CREATE_OK(cudaMalloc(&ctx->d_sel, B * c.nlevels * (size_t)ctx->kp_cap * sizeof(SelRec)));
    CREATE_OK(cudaMalloc(&ctx->d_kp, B * (size_t)ctx->kp_cap * sizeof(mslam_b200_keypoint)));
    CREATE_OK(cudaMalloc(&ctx->d_desc, B * (size_t)ctx->kp_cap * 32));
    CREATE_OK(cudaMalloc(&ctx->d_counts, B * sizeof(int)));
    {
        float pat[1024];
        brief_pattern_transposed(pat);
        CREATE_OK(cudaMalloc(&ctx->d_pat, sizeof(pat)));
        CREATE_OK(cudaMemcpy(ctx->d_pat, pat, sizeof(pat), cudaMemcpyHostToDevice));
        std::vector<unsigned> icw(4 * 31 * 9 * 2);
        ic_weight_table(icw.data());
        CREATE_OK(cudaMalloc(&ctx->d_icw, icw.size() * sizeof(unsigned)));
        CREATE_OK(cudaMemcpy(ctx->d_icw, icw.data(), icw.size() * sizeof(unsigned), cudaMemcpyHostToDevice));
    }
    ctx->one_kp_off = 16;
    ctx->one_desc_off = align_up(ctx->one_kp_off + (size_t)ctx->kp_cap * sizeof(mslam_b200_keypoint), 32);
    ctx->one_bytes = ctx->one_desc_off + (size_t)ctx->kp_cap * 32;
    CREATE_OK(cudaMalloc(&ctx->d_one, ctx->one_bytes));
    CREATE_OK(cudaHostAlloc(&ctx->h_one, ctx->one_bytes, cudaHostAllocDefault));
    CREATE_OK(cudaHostAlloc(&ctx->h_counts, 2 * B * sizeof(int), cudaHostAllocDefault));
    CREATE_OK(cudaHostAlloc(&ctx->h_err, 2 * B * sizeof(int), cudaHostAllocDefault));
    for(int i = 0; i < 2; ++i)
        CREATE_OK(cudaEventCreateWithFlags(&ctx->ev_slot_done[i], cudaEventDisableTiming));
    CREATE_OK(cudaEventCreateWithFlags(&ctx->ev_out_done, cudaEventDisableTiming));
    CREATE_OK(cudaEventCreateWithFlags(&ctx->ev_match_done, cudaEventDisableTiming));
    CREATE_OK(cudaEventCreateWithFlags(&ctx->ev_match_in, cudaEventDisableTiming));
    CREATE_OK(cudaEventCreateWithFlags(&ctx->ev_match_k, cudaEventDisableTiming));
    // matcher
    const size_t dc = (size_t)ctx->desc_cap;
    CREATE_OK(cudaMalloc(&ctx->d_mfrom, B * dc * 32));
    CREATE_OK(cudaMalloc(&ctx->d_mto, B * dc * 32));
    CREATE_OK(cudaMalloc(&ctx->d_raw, dc * 128));
    CREATE_OK(cudaMalloc(&ctx->d_nfrom, B * sizeof(int)));
    CREATE_OK(cudaMalloc(&ctx->d_nto, B * sizeof(int)));
    CREATE_OK(cudaMalloc(&ctx->d_key0, B * dc * sizeof(unsigned)));
    CREATE_OK(cudaMalloc(&ctx->d_key1, B * dc * sizeof(unsigned)));
    CREATE_OK(cudaMalloc(&ctx->d_key0b, B * dc * sizeof(unsigned)));
    CREATE_OK(cudaMalloc(&ctx->d_key1b, B * dc * sizeof(unsigned)));
    CREATE_OK(cudaMalloc(&ctx->d_part0, kKnnMaxSlices * B * dc * sizeof(unsigned)));
    CREATE_OK(cudaMalloc(&ctx->d_part1, kKnnMaxSlices * B * dc * sizeof(unsigned)));
    if(const char* e = getenv("MSLAM_KNN_MMA"))
        ctx->knn_mma_env = atoi(e);
    {
        const int rc__ = apply_knn_flags(ctx);
        if(rc__ != MSLAM_B200_OK)
        {
            mslam_b200_destroy(ctx);
            return rc__;
        }
    }
    CREATE_OK(cudaMalloc(&ctx->d_out_a, B * dc * sizeof(int)));
    CREATE_OK(cudaMalloc(&ctx->d_out_b, B * dc * sizeof(int)));
    CREATE_OK(cudaMalloc(&ctx->d_out_d, B * dc * sizeof(int)));
    CREATE_OK(cudaMalloc(&ctx->d_nout, B * sizeof(int)));
    CREATE_OK(cudaHostAlloc(&ctx->h_out_a, B * dc * sizeof(int), cudaHostAllocDefault));
    CREATE_OK(cudaHostAlloc(&ctx->h_out_b, B * dc * sizeof(int), cudaHostAllocDefault));
    CREATE_OK(cudaHostAlloc(&ctx->h_out_d, B * dc * sizeof(int), cudaHostAllocDefault));
    CREATE_OK(cudaHostAlloc(&ctx->h_nout, B * sizeof(int), cudaHostAllocDefault));
    CREATE_OK(cudaHostAlloc(&ctx->h_n2, 2 * B * sizeof(int), cudaHostAllocDefault));
    ctx->m1_bytes = 32 + 2 * dc * 32;
    ctx->mo_bytes = 16 + 3 * dc * sizeof(int);
    CREATE_OK(cudaMalloc(&ctx->d_m1, ctx->m1_bytes));
    CREATE_OK(cudaMalloc(&ctx->d_mo, ctx->mo_bytes));
    CREATE_OK(cudaHostAlloc(&ctx->h_m1, ctx->m1_bytes, cudaHostAllocDefault));
    CREATE_OK(cudaHostAlloc(&ctx->h_mo, ctx->mo_bytes, cudaHostAllocDefault));
#undef CREATE_OK
    *out_ctx = ctx;
    return MSLAM_B200_OK;
}

int mslam_b200_set_stream(mslam_b200_ctx* ctx, void* cuda_stream)
{
    if(!ctx)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    if(ctx->own_stream)
    {
        cudaStreamDestroy(ctx->stream);
        ctx->own_stream = false;
    }
    if(cuda_stream)
        ctx->stream = (cudaStream_t)cuda_stream;
    else
    {
        MSLAM_CUDA_OK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
        ctx->own_stream = true;
    }
    return MSLAM_B200_OK;
}

int mslam_b200_synchronize(mslam_b200_ctx* ctx)
{
    if(!ctx)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    return MSLAM_B200_OK;
}

int mslam_b200_set_flags(mslam_b200_ctx* ctx, int flags)
{
    if(!ctx)
        return MSLAM_B200_E_INVALID;
    if((flags & MSLAM_B200_FLAG_KNN_SCALAR) && (flags & MSLAM_B200_FLAG_KNN_TENSOR))
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    MSLAM_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    ctx->cfg.flags = flags;
    // the single-frame kernel graphs were captured under the old flags (MSLAM_B200_FLAG_SERIAL changes the stream structure)
    for(auto& g : ctx->graph1)
    {
        if(g.exec)
            cudaGraphExecDestroy(g.exec);
        g = mslam_b200_ctx::FrameGraph{};
    }
    if(ctx->graph_lat.exec)
        cudaGraphExecDestroy(ctx->graph_lat.exec);
    ctx->graph_lat = mslam_b200_ctx::FrameGraph{};
    return apply_knn_flags(ctx);
}

int mslam_b200_level_geometry(mslam_b200_ctx* ctx, int width, int height, int* w, int* h, float* scale, int* n_per_level)
{
    if(!ctx || width < 1 || height < 1)
        return MSLAM_B200_E_INVALID;
    int lw[kMaxLevels], lh[kMaxLevels], nl[kMaxLevels];
    float sc[kMaxLevels];
    level_geometry(width, height, ctx->cfg.nlevels, ctx->cfg.scale_factor, lw, lh, sc);
    features_per_level(ctx->cfg.nfeatures, ctx->cfg.nlevels, ctx->cfg.scale_factor, nl);
    for(int l = 0; l < ctx->cfg.nlevels; ++l)
    {
        if(w) w[l] = lw[l];
        if(h) h[l] = lh[l];
        if(scale) scale[l] = sc[l];
        if(n_per_level) n_per_level[l] = nl[l];
    }
    return MSLAM_B200_OK;
}

// ---- detection ---------------------------------------------------------------------------------------------------
static int run_pyramid(mslam_b200_ctx* ctx, const uint8_t* d_bgr, int n_frames)
{
    ctx->P.n_frames = n_frames;
    {
        StageTimer t(ctx, MSLAM_B200_T_GRAY);
        launch_gray(ctx->stream, d_bgr, ctx->d_pyr, ctx->P, &ctx->launches);
    }
    {
        StageTimer t(ctx, MSLAM_B200_T_RESIZE);
        launch_pyramid(ctx->stream, ctx->d_pyr, ctx->P, ctx->d_xtab, ctx->d_ytab, ctx->d_qtab, ctx->sm_count, 1, ctx->P.nlevels, &ctx->launches);
    }
    return MSLAM_B200_OK;
}

// enqueue the detect pipeline for frames [frame0, frame0 + n_frames) of the batch buffers (geometry already set, the
// zero block already cleared); outputs are indexed by absolute frame
static int detect_range(mslam_b200_ctx* ctx, const uint8_t* d_bgr, int frame0, int n_frames, mslam_b200_keypoint* d_kp,
                        uint8_t* d_desc, int capacity, int* d_counts, int* d_err = nullptr)
{
    cudaStream_t st = ctx->stream;
    if(!d_err)
        d_err = ctx->d_err; // per-frame "selection overflowed" flags, part of the zeroed block
    ctx->P.frame0 = frame0;
    const int n_slots = std::min(capacity, ctx->kp_cap * ctx->P.nlevels);
    if(ctx->cfg.flags & MSLAM_B200_FLAG_SERIAL)
    {
        run_pyramid(ctx, d_bgr, n_frames);
        const PipeParams& P = ctx->P;
        {
            StageTimer t(ctx, MSLAM_B200_T_FAST);
            launch_fast(st, ctx->d_pyr, ctx->d_cand, ctx->d_ctr, ctx->d_hist, P, ctx->fast_maps, 0, P.nlevels, &ctx->launches);
        }
        {
            StageTimer t(ctx, MSLAM_B200_T_SELECT);
            launch_select(st, ctx->d_pyr, ctx->d_cand, ctx->d_cand2, ctx->d_keys, ctx->d_keys2, ctx->d_ctr, ctx->d_hist,
                          ctx->d_sel, d_err, P, ctx->sm_count, &ctx->launches);
        }
        {
            StageTimer t(ctx, MSLAM_B200_T_BLUR);
            launch_blur(st, ctx->d_pyr, ctx->d_blur, P, 0, P.nlevels, &ctx->launches);
        }
        StageTimer t(ctx, MSLAM_B200_T_DESCRIBE);
        launch_orient_describe(st, ctx->d_pyr, ctx->d_blur, ctx->d_ctr, ctx->d_sel, ctx->d_pat, ctx->d_icw, d_kp, d_desc, d_counts, capacity,
                               n_slots, P, ctx->sm_count, &ctx->launches);
    }
    else
    {
        // Dependency graph of one batch:  gray -> L1 -> L2 -> ... (chained resize, main stream)
        //   FAST(l) and blur(l) need only level l; select needs every FAST; describe needs select and every blur.
        // The chain, FAST + select (aux 0) and the blur (aux 1) run on three streams that fork at the level events and
        // join before describe, so the latency-bound pieces (small levels, the one-CTA-per-level selection) hide
        // under the throughput-bound ones.
        ctx->P.n_frames = n_frames;
        const PipeParams& P = ctx->P;
        cudaStream_t s_fast = ctx->aux[0], s_blur = ctx->aux[1];
        if(n_frames <= 2)
        {
            // One or two frames: every launch is a step of a latency chain (a tile's walk, ~10 us), not throughput, and launches
            // on one stream run one after the other.  The MSLAM_B200_LAT_TRACE=2 timeline of the batch structure below showed the
            // chain done at 40 us while FAST(0), FAST(1), FAST(2..) in a row on the fork kept select waiting until 72.  Here
            // FAST(0) runs under the chain on the blur stream, FAST(1..) is ONE launch behind the chain on the other fork (all
            // its tiles fit the machine at once), the blur is one launch behind FAST(0), select waits for both FAST launches.
            const int nl = P.nlevels;
            // FAST in short tiles: these launches never fill the machine, so each lasts as long as ONE tile
            PipeParams Pf = ctx->P_few;
            Pf.frame0 = P.frame0, Pf.n_frames = P.n_frames;
            {
                StageTimer t(ctx, MSLAM_B200_T_GRAY);
                launch_gray(st, d_bgr, ctx->d_pyr, P, &ctx->launches);
            }
            MSLAM_CUDA_OK(cudaEventRecord(ctx->ev_level[0], st));
            if(nl > 1)
            {
                StageTimer t(ctx, MSLAM_B200_T_RESIZE);
                launch_pyramid(st, ctx->d_pyr, P, ctx->d_xtab, ctx->d_ytab, ctx->d_qtab, ctx->sm_count, 1, nl, &ctx->launches);
            }
            MSLAM_CUDA_OK(cudaEventRecord(ctx->ev_level[nl - 1], st));
            MSLAM_CUDA_OK(cudaStreamWaitEvent(s_blur, ctx->ev_level[0], 0));
            {
                StageTimer t(ctx, MSLAM_B200_T_FAST, s_blur);
                launch_fast(s_blur, ctx->d_pyr, ctx->d_cand, ctx->d_ctr, ctx->d_hist, Pf, ctx->fast_maps_few, 0, 1, &ctx->launches);
            }
            MSLAM_CUDA_OK(cudaEventRecord(ctx->ev_fast0, s_blur));
            MSLAM_CUDA_OK(cudaStreamWaitEvent(s_blur, ctx->ev_level[nl - 1], 0));
            {
                StageTimer t(ctx, MSLAM_B200_T_BLUR, s_blur);
                launch_blur(s_blur, ctx->d_pyr, ctx->d_blur, P, 0, nl, &ctx->launches);
            }
            MSLAM_CUDA_OK(cudaEventRecord(ctx->ev_join[1], s_blur));
            MSLAM_CUDA_OK(cudaStreamWaitEvent(s_fast, ctx->ev_level[nl - 1], 0));
            if(nl > 1)
            {
                StageTimer t(ctx, MSLAM_B200_T_FAST, s_fast);
                launch_fast(s_fast, ctx->d_pyr, ctx->d_cand, ctx->d_ctr, ctx->d_hist, Pf, ctx->fast_maps_few, 1, nl, &ctx->launches);
            }
            MSLAM_CUDA_OK(cudaStreamWaitEvent(s_fast, ctx->ev_fast0, 0));
            {
                StageTimer t(ctx, MSLAM_B200_T_SELECT, s_fast);
                launch_select(s_fast, ctx->d_pyr, ctx->d_cand, ctx->d_cand2, ctx->d_keys, ctx->d_keys2, ctx->d_ctr, ctx->d_hist,
                              ctx->d_sel, d_err, P, ctx->sm_count, &ctx->launches);
            }
            MSLAM_CUDA_OK(cudaEventRecord(ctx->ev_join[0], s_fast));
            MSLAM_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_join[0], 0));
            MSLAM_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_join[1], 0));
            StageTimer t(ctx, MSLAM_B200_T_DESCRIBE);
            launch_orient_describe(st, ctx->d_pyr, ctx->d_blur, ctx->d_ctr, ctx->d_sel, ctx->d_pat, ctx->d_icw, d_kp, d_desc, d_counts, capacity,
                                   n_slots, P, ctx->sm_count, &ctx->launches);
            MSLAM_CUDA_OK(cudaGetLastError());
            return MSLAM_B200_OK;
        }
        static int split_env = -1; // A/B aid: MSLAM_FAST_SPLIT = number of leading levels with their own FAST / blur launches
        if(split_env < 0)
        {
            const char* e = getenv("MSLAM_FAST_SPLIT");
            split_env = e ? atoi(e) : 2;
            if(split_env < 1 || split_env > 4)
                split_env = 2;
        }
        const int nl = P.nlevels, l_split = std::min(split_env, nl); // levels < l_split get their own FAST launch
        {
            StageTimer t(ctx, MSLAM_B200_T_GRAY);
            launch_gray(st, d_bgr, ctx->d_pyr, P, &ctx->launches);
        }
        MSLAM_CUDA_OK(cudaEventRecord(ctx->ev_level[0], st));
        // levels below l_split are waited for one by one (their own FAST / blur launches); the rest of the chain is one call
        // (which fuses the trailing small levels into a single launch) with one event at its end
        for(int l = 1; l < std::min(l_split, nl); ++l)
        {
            {
                StageTimer t(ctx, MSLAM_B200_T_RESIZE);
                launch_pyramid(st, ctx->d_pyr, P, ctx->d_xtab, ctx->d_ytab, ctx->d_qtab, ctx->sm_count, l, l + 1, &ctx->launches);
            }
            MSLAM_CUDA_OK(cudaEventRecord(ctx->ev_level[l], st));
        }
        if(std::max(l_split, 1) < nl)
        {
            {
                StageTimer t(ctx, MSLAM_B200_T_RESIZE);
                launch_pyramid(st, ctx->d_pyr, P, ctx->d_xtab, ctx->d_ytab, ctx->d_qtab, ctx->sm_count, std::max(l_split, 1), nl, &ctx->launches);
            }
            MSLAM_CUDA_OK(cudaEventRecord(ctx->ev_level[nl - 1], st));
        }
        for(int l = 0; l < l_split; ++l)
        {
            MSLAM_CUDA_OK(cudaStreamWaitEvent(s_fast, ctx->ev_level[l], 0));
            StageTimer t(ctx, MSLAM_B200_T_FAST, s_fast);
            launch_fast(s_fast, ctx->d_pyr, ctx->d_cand, ctx->d_ctr, ctx->d_hist, P, ctx->fast_maps, l, l + 1, &ctx->launches);
        }
        if(l_split < nl)
        {
            MSLAM_CUDA_OK(cudaStreamWaitEvent(s_fast, ctx->ev_level[nl - 1], 0));
            StageTimer t(ctx, MSLAM_B200_T_FAST, s_fast);
            launch_fast(s_fast, ctx->d_pyr, ctx->d_cand, ctx->d_ctr, ctx->d_hist, P, ctx->fast_maps, l_split, nl, &ctx->launches);
        }
        {
            StageTimer t(ctx, MSLAM_B200_T_SELECT, s_fast);
            launch_select(s_fast, ctx->d_pyr, ctx->d_cand, ctx->d_cand2, ctx->d_keys, ctx->d_keys2, ctx->d_ctr, ctx->d_hist,
                          ctx->d_sel, d_err, P, ctx->sm_count, &ctx->launches);
        }
        MSLAM_CUDA_OK(cudaEventRecord(ctx->ev_join[0], s_fast));
        for(int l = 0; l < l_split; ++l)
        {
            MSLAM_CUDA_OK(cudaStreamWaitEvent(s_blur, ctx->ev_level[l], 0));
            StageTimer t(ctx, MSLAM_B200_T_BLUR, s_blur);
            launch_blur(s_blur, ctx->d_pyr, ctx->d_blur, P, l, l + 1, &ctx->launches);
        }
        if(l_split < nl)
        {
            MSLAM_CUDA_OK(cudaStreamWaitEvent(s_blur, ctx->ev_level[nl - 1], 0));
            StageTimer t(ctx, MSLAM_B200_T_BLUR, s_blur);
            launch_blur(s_blur, ctx->d_pyr, ctx->d_blur, P, l_split, nl, &ctx->launches);
        }
        MSLAM_CUDA_OK(cudaEventRecord(ctx->ev_join[1], s_blur));
        MSLAM_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_join[0], 0));
        MSLAM_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_join[1], 0));
        StageTimer t(ctx, MSLAM_B200_T_DESCRIBE);
        launch_orient_describe(st, ctx->d_pyr, ctx->d_blur, ctx->d_ctr, ctx->d_sel, ctx->d_pat, ctx->d_icw, d_kp, d_desc, d_counts, capacity,
                               n_slots, P, ctx->sm_count, &ctx->launches);
    }
    MSLAM_CUDA_OK(cudaGetLastError());
    return MSLAM_B200_OK;
}

int mslam_b200_detect_batch_dev(mslam_b200_ctx* ctx, const uint8_t* d_bgr, int n_frames, int width, int height,
                                mslam_b200_keypoint* d_kp, uint8_t* d_desc, int capacity, int* d_counts)
{
    if(!ctx || !d_bgr || !d_kp || !d_desc || !d_counts || n_frames < 1 || n_frames > ctx->cfg.max_batch || capacity < 1)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    int rc = set_geometry(ctx, width, height);
    if(rc)
        return rc;
    MSLAM_CUDA_OK(cudaMemsetAsync(ctx->d_zero, 0, ctx->zero_bytes, ctx->stream));
    ctx->last_frames = n_frames;
    return detect_range(ctx, d_bgr, 0, n_frames, d_kp, d_desc, capacity, d_counts);
}

// enqueue one host batch into pipeline slot `slot` (copy-in, kernels, copy-out); nothing here waits for the GPU
// One frame through the kernels of detect_range, replayed from a captured graph once the geometry has been seen (the
// first call of a geometry launches directly, which also performs any one-time kernel attribute set-up outside capture).
// A capture that fails switches the graphs off for this context and the launches go out directly, as for batches.
static int detect_one_graphed(mslam_b200_ctx* ctx, int slot, const uint8_t* d_bgr, int width, int height)
{
    cudaStream_t st = ctx->stream;
    if(ctx->graphs_enabled < 0)
    {
        const char* e = getenv("MSLAM_B200_GRAPH");
        ctx->graphs_enabled = e ? atoi(e) != 0 : 1;
    }
    auto& g = ctx->graph1[slot];
    const bool usable = ctx->graphs_enabled && !ctx->timing && !(ctx->cfg.flags & MSLAM_B200_FLAG_SERIAL);
    if(!usable)
        return detect_range(ctx, d_bgr, 0, 1, ctx->d_kp, ctx->d_desc, ctx->kp_cap, ctx->d_counts);
    if(g.width != width || g.height != height)
    {
        if(g.exec)
            cudaGraphExecDestroy(g.exec);
        g.exec = nullptr;
        g.width = width, g.height = height, g.uses = 0;
    }
    if(!g.exec && g.uses++ >= 1)
    {
        const long long l0 = ctx->launches;
        cudaGraph_t graph = nullptr;
        bool ok = cudaStreamBeginCapture(st, cudaStreamCaptureModeRelaxed) == cudaSuccess;
        if(ok)
        {
            ok = detect_range(ctx, d_bgr, 0, 1, ctx->d_kp, ctx->d_desc, ctx->kp_cap, ctx->d_counts) == MSLAM_B200_OK;
            ok = (cudaStreamEndCapture(st, &graph) == cudaSuccess) && ok && graph;
        }
        if(ok)
            ok = cudaGraphInstantiate(&g.exec, graph, 0) == cudaSuccess;
        if(graph)
            cudaGraphDestroy(graph);
        g.launches = ctx->launches - l0;
        ctx->launches = l0; // nothing ran yet
        if(!ok)
        {
            cudaGetLastError();
            g.exec = nullptr;
            ctx->graphs_enabled = 0;
        }
    }
    if(!g.exec)
        return detect_range(ctx, d_bgr, 0, 1, ctx->d_kp, ctx->d_desc, ctx->kp_cap, ctx->d_counts);
    ctx->P.frame0 = 0, ctx->P.n_frames = 1;
    MSLAM_CUDA_OK(cudaGraphLaunch(g.exec, st));
    ctx->launches += g.launches;
    return MSLAM_B200_OK;
}

static int detect_batch_enqueue(mslam_b200_ctx* ctx, int slot, const uint8_t* const* frames, int n_frames, int width, int height,
                                int stride_bytes, mslam_b200_keypoint* out_kp, uint8_t* out_desc, int capacity)
{
    cudaStream_t st = ctx->stream;
    const size_t row = (size_t)width * 3, frame_bytes = row * height;
    int rc = set_geometry(ctx, width, height);
    if(rc)
        return rc;
    // The batch is cut into chunks that flow through three engines at once: chunk c+1 is copied in (copy stream) while
    // chunk c runs through the kernels (context stream + its forks) and chunk c-1's results are copied out (second
    // copy stream).  Per-frame scratch in HBM is indexed by absolute frame, so chunks never touch each other's data.
    const int icap = ctx->kp_cap, ncopy = std::min(capacity, icap);
    static int want_chunks = -1;
    if(want_chunks < 0)
    {
        const char* e = getenv("MSLAM_B200_CHUNKS");
        want_chunks = e ? std::max(1, std::min(kMaxChunks, atoi(e))) : kDefaultChunks;
    }
    const int nchunks = n_frames >= 8 ? std::min(want_chunks, n_frames / 2) : 1;
    // chunk boundaries: the first chunk's copy-in and the last chunk's kernels are the only parts nothing overlaps, so
    // with four chunks and a large batch those two are made small (1 : 3 : 3 : 1)
    static int tapered = -1;
    if(tapered < 0)
    {
        const char* e = getenv("MSLAM_B200_TAPER");
        tapered = e ? atoi(e) : 1;
    }
    int bound[kMaxChunks + 1];
    for(int c = 0; c <= nchunks; ++c)
        bound[c] = (int)((long long)n_frames * c / nchunks);
    if(tapered && nchunks == 4 && n_frames >= 32)
        bound[1] = n_frames / 8, bound[2] = n_frames / 2, bound[3] = n_frames - n_frames / 8;
    cudaStream_t s_in = ctx->copy[0], s_out = ctx->copy[1];
    uint8_t* d_bgr = ctx->d_bgr + (size_t)slot * ctx->bgr_bytes;
    const bool pipelined = ctx->pend[slot ^ 1].active;
    // the previous batch's results must have left d_kp / d_desc / d_counts / d_err before this batch's kernels reuse them
    if(ctx->out_done_valid)
        MSLAM_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_out_done, 0));
    MSLAM_CUDA_OK(cudaMemsetAsync(ctx->d_zero, 0, ctx->zero_bytes, st));
    if(!pipelined)
    {
        // earlier single-frame work on the context stream may still read this frame buffer (mslam_b200_describe); with a
        // batch already in flight the only work queued since is that batch's, which lives in the other slot
        MSLAM_CUDA_OK(cudaEventRecord(ctx->ev_done[kMaxChunks], st));
        MSLAM_CUDA_OK(cudaStreamWaitEvent(s_in, ctx->ev_done[kMaxChunks], 0));
    }
    for(int c = 0; c < nchunks; ++c)
    {
        const int b0 = bound[c], b1 = bound[c + 1];
        bool contiguous = (size_t)stride_bytes == row;
        for(int i = b0 + 1; i < b1 && contiguous; ++i)
            contiguous = frames[i] == frames[i - 1] + frame_bytes;
        if(contiguous && b1 > b0)
            MSLAM_CUDA_OK(cudaMemcpyAsync(d_bgr + b0 * frame_bytes, frames[b0], (size_t)(b1 - b0) * frame_bytes,
                                          cudaMemcpyHostToDevice, s_in));
        else
            for(int i = b0; i < b1; ++i)
                MSLAM_CUDA_OK(cudaMemcpy2DAsync(d_bgr + i * frame_bytes, row, frames[i], (size_t)stride_bytes, row,
                                                (size_t)height, cudaMemcpyHostToDevice, s_in));
        MSLAM_CUDA_OK(cudaEventRecord(ctx->ev_h2d[c], s_in));
    }
    for(int c = 0; c < nchunks; ++c)
    {
        const int b0 = bound[c], b1 = bound[c + 1];
        if(b1 <= b0)
            continue;
        MSLAM_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_h2d[c], 0));
        if(n_frames == 1)
            rc = detect_one_graphed(ctx, slot, d_bgr, width, height);
        else
            rc = detect_range(ctx, d_bgr, b0, b1 - b0, ctx->d_kp, ctx->d_desc, icap, ctx->d_counts);
        if(rc)
            return rc;
        MSLAM_CUDA_OK(cudaEventRecord(ctx->ev_done[c], st));
        MSLAM_CUDA_OK(cudaStreamWaitEvent(s_out, ctx->ev_done[c], 0));
        MSLAM_CUDA_OK(cudaMemcpy2DAsync(out_kp + (size_t)b0 * capacity, (size_t)capacity * sizeof(mslam_b200_keypoint),
                                        ctx->d_kp + (size_t)b0 * icap, (size_t)icap * sizeof(mslam_b200_keypoint),
                                        (size_t)ncopy * sizeof(mslam_b200_keypoint), (size_t)(b1 - b0), cudaMemcpyDeviceToHost,
                                        s_out));
        MSLAM_CUDA_OK(cudaMemcpy2DAsync(out_desc + (size_t)b0 * capacity * 32, (size_t)capacity * 32,
                                        ctx->d_desc + (size_t)b0 * icap * 32, (size_t)icap * 32, (size_t)ncopy * 32,
                                        (size_t)(b1 - b0), cudaMemcpyDeviceToHost, s_out));
    }
    ctx->last_frames = n_frames;
    const size_t B = (size_t)ctx->cfg.max_batch;
    MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->h_counts + slot * B, ctx->d_counts, n_frames * sizeof(int), cudaMemcpyDeviceToHost, s_out));
    MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->h_err + slot * B, ctx->d_err, n_frames * sizeof(int), cudaMemcpyDeviceToHost, s_out));
    MSLAM_CUDA_OK(cudaEventRecord(ctx->ev_slot_done[slot], s_out));
    MSLAM_CUDA_OK(cudaEventRecord(ctx->ev_out_done, s_out));
    ctx->out_done_valid = true;
    return MSLAM_B200_OK;
}

// wait for the batch in `slot`, hand its counts to the caller
static int detect_batch_finish(mslam_b200_ctx* ctx, int slot)
{
    mslam_b200_ctx::PendingBatch& pb = ctx->pend[slot];
    pb.active = false;
    MSLAM_CUDA_OK(cudaEventSynchronize(ctx->ev_slot_done[slot]));
    if(!ctx->pend[slot ^ 1].active)
        ctx->out_done_valid = false; // the copy-out stream has drained: nothing for the next call to wait for
    const size_t B = (size_t)ctx->cfg.max_batch;
    int rc = MSLAM_B200_OK;
    for(int i = 0; i < pb.n_frames; ++i)
    {
        const int n = ctx->h_counts[slot * B + i];
        pb.n_out[i] = n;
        if(n > pb.capacity || ctx->h_err[slot * B + i])
            rc = MSLAM_B200_E_CAPACITY;
    }
    return rc;
}

static int detect_batch_args_ok(mslam_b200_ctx* ctx, const uint8_t* const* frames, int n_frames, int width, int height,
                                int stride_bytes, const void* out_kp, const void* out_desc, int capacity, const int* n_out)
{
    if(!ctx || !frames || !out_kp || !out_desc || !n_out || n_frames < 1 || n_frames > ctx->cfg.max_batch ||
       capacity < 1 || width < 1 || height < 1 || stride_bytes < 3 * width || width > ctx->cfg.max_width ||
       height > ctx->cfg.max_height)
        return 0;
    for(int i = 0; i < n_frames; ++i)
        if(!frames[i])
            return 0;
    return 1;
}

int mslam_b200_detect_batch_begin(mslam_b200_ctx* ctx, const uint8_t* const* frames, int n_frames, int width, int height,
                                  int stride_bytes, mslam_b200_keypoint* out_kp, uint8_t* out_desc, int capacity, int* n_out)
{
    if(!detect_batch_args_ok(ctx, frames, n_frames, width, height, stride_bytes, out_kp, out_desc, capacity, n_out))
        return MSLAM_B200_E_INVALID;
    const int slot = ctx->pend_next;
    if(ctx->pend[slot].active)
        return MSLAM_B200_E_INVALID; // two batches are already in flight: call mslam_b200_detect_batch_end first
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    const int rc = detect_batch_enqueue(ctx, slot, frames, n_frames, width, height, stride_bytes, out_kp, out_desc, capacity);
    if(rc)
        return rc;
    ctx->pend[slot].active = true;
    ctx->pend[slot].n_frames = n_frames;
    ctx->pend[slot].capacity = capacity;
    ctx->pend[slot].n_out = n_out;
    ctx->pend_next = slot ^ 1;
    return MSLAM_B200_OK;
}

int mslam_b200_detect_batch_end(mslam_b200_ctx* ctx)
{
    if(!ctx)
        return MSLAM_B200_E_INVALID;
    // the oldest batch in flight: the slot the next begin would take if it is busy, else the other one
    int slot = ctx->pend_next;
    if(!ctx->pend[slot].active)
        slot ^= 1;
    if(!ctx->pend[slot].active)
        return MSLAM_B200_E_INVALID; // nothing in flight
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    return detect_batch_finish(ctx, slot);
}

// The call the reference's frontend makes once per frame (IFeatureDetector::detect, rgbd_feature_frontend.cpp:187) with nothing
// else in flight on the context.  Returns kTakeGeneralPath when the call has to take the general path (first call of a geometry, graphs off,
// stage timing on, a pipelined batch in flight), else the status of the finished call.
// Measured (tools/latency_breakdown.py, 640x480 / 1000 keypoints): the kernels of one frame span ~57 us on the GPU and the
// 0.9 MB frame copy 18 us, but the general path took 194 us wall -- three streams, six event hops, a memset, four device->host
// copies into caller memory.  Here: copy-in, ONE graph launch (clear counters -> kernels -> result block to its pinned
// mirror), one stream wait, two host memcpy into the caller's arrays.
constexpr int kTakeGeneralPath = 1; // not a status of the C ABI (those are <= 0)
static int detect_one_latency(mslam_b200_ctx* ctx, const uint8_t* bgr, int width, int height, int stride_bytes,
                              mslam_b200_keypoint* out_kp, uint8_t* out_desc, int capacity, int* n_out)
{
    if(ctx->graphs_enabled < 0)
    {
        const char* e = getenv("MSLAM_B200_GRAPH");
        ctx->graphs_enabled = e ? atoi(e) != 0 : 1;
    }
    static int lat_env = -1; // A/B aid: MSLAM_B200_LATENCY_PATH=0 sends single frames through the general path
    if(lat_env < 0)
    {
        const char* e = getenv("MSLAM_B200_LATENCY_PATH");
        lat_env = e ? atoi(e) != 0 : 1;
    }
    if(!lat_env || !ctx->graphs_enabled || ctx->timing || (ctx->cfg.flags & MSLAM_B200_FLAG_SERIAL) || ctx->pend[0].active ||
       ctx->pend[1].active)
        return kTakeGeneralPath;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    int rc = set_geometry(ctx, width, height);
    if(rc)
        return rc;
    cudaStream_t st = ctx->stream;
    auto& g = ctx->graph_lat;
    if(g.width != width || g.height != height)
    {
        if(g.exec)
            cudaGraphExecDestroy(g.exec);
        g = mslam_b200_ctx::FrameGraph{};
        g.width = width, g.height = height;
    }
    // MSLAM_B200_LAT_TRACE=1: where a call's time goes (host clock at each step, GPU time of the copy-in and of the graph), averaged
    // over 100 calls to stderr.  The events sit between the operations, so a traced call is a few microseconds slower.
    // =2: the graph is also captured with the stage timers' event pairs in it, and every 100th call prints when each stage began
    // and ended on the GPU (microseconds after the copy-in).
    static int trace = -1;
    static cudaEvent_t tev[3];
    static double tacc[6];
    static int tn = 0, tpairs = 0;
    if(trace < 0)
    {
        const char* e = getenv("MSLAM_B200_LAT_TRACE");
        trace = e ? atoi(e) : 0;
        if(trace)
            for(auto& ev : tev)
                cudaEventCreate(&ev);
    }
    int* d_hdr = reinterpret_cast<int*>(ctx->d_one);
    mslam_b200_keypoint* d_kp1 = reinterpret_cast<mslam_b200_keypoint*>(ctx->d_one + ctx->one_kp_off);
    uint8_t* d_desc1 = ctx->d_one + ctx->one_desc_off;
    if(!g.exec)
    {
        if(g.uses++ < 1)
            return kTakeGeneralPath; // first call of a geometry: direct launches (one-time kernel set-up happens outside any capture)
        const long long l0 = ctx->launches;
        cudaGraph_t graph = nullptr;
        bool ok = cudaStreamBeginCapture(st, cudaStreamCaptureModeRelaxed) == cudaSuccess;
        if(ok)
        {
            // frame 0's slice of the zeroed block (counters and histograms are [frame][level]) and the result header
            ok = cudaMemsetAsync(ctx->d_ctr, 0, ctx->P.nlevels * sizeof(LevelCounters), st) == cudaSuccess && ok;
            ok = cudaMemsetAsync(ctx->d_hist, 0, (size_t)ctx->P.nlevels * 256 * sizeof(unsigned), st) == cudaSuccess && ok;
            ok = cudaMemsetAsync(ctx->d_one, 0, 16, st) == cudaSuccess && ok;
            if(trace >= 2)
            {
                if(ctx->ev_pool.empty())
                {
                    ctx->ev_pool.resize(128);
                    for(auto& ev : ctx->ev_pool)
                        cudaEventCreate(&ev);
                    ctx->ev_stage.assign(64, 0);
                    ctx->ev_launches.assign(64, 0);
                }
                ctx->timing = ctx->timing_in_graph = true, ctx->ev_used = 0;
            }
            ok = detect_range(ctx, ctx->d_bgr, 0, 1, d_kp1, d_desc1, ctx->kp_cap, d_hdr, d_hdr + 1) == MSLAM_B200_OK && ok;
            if(trace >= 2)
            {
                tpairs = (int)ctx->ev_used;
                ctx->timing = ctx->timing_in_graph = false, ctx->ev_used = 0;
            }
            ok = cudaMemcpyAsync(ctx->h_one, ctx->d_one, ctx->one_bytes, cudaMemcpyDeviceToHost, st) == cudaSuccess && ok;
            ok = (cudaStreamEndCapture(st, &graph) == cudaSuccess) && ok && graph;
        }
        if(ok)
            ok = cudaGraphInstantiate(&g.exec, graph, 0) == cudaSuccess;
        if(graph)
            cudaGraphDestroy(graph);
        g.launches = ctx->launches - l0;
        ctx->launches = l0; // nothing ran yet
        if(!ok)
        {
            cudaGetLastError();
            g.exec = nullptr;
            ctx->graphs_enabled = 0;
            return kTakeGeneralPath;
        }
    }
    // results of an earlier batch may still be leaving through the copy-out stream, which reads the zeroed block's err flags
    if(ctx->out_done_valid)
    {
        MSLAM_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_out_done, 0));
        ctx->out_done_valid = false;
    }
    auto now_us = []() {
        timespec ts;
        clock_gettime(CLOCK_MONOTONIC, &ts);
        return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3;
    };
    const double t0 = trace ? now_us() : 0.0;
    if(trace)
        cudaEventRecord(tev[0], st);
    const size_t row = (size_t)width * 3;
    if((size_t)stride_bytes == row)
        MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->d_bgr, bgr, row * height, cudaMemcpyHostToDevice, st));
    else
        MSLAM_CUDA_OK(cudaMemcpy2DAsync(ctx->d_bgr, row, bgr, (size_t)stride_bytes, row, (size_t)height, cudaMemcpyHostToDevice, st));
    const double t1 = trace ? now_us() : 0.0;
    if(trace)
        cudaEventRecord(tev[1], st);
    ctx->P.frame0 = 0, ctx->P.n_frames = 1;
    ctx->last_frames = 1;
    MSLAM_CUDA_OK(cudaGraphLaunch(g.exec, st));
    ctx->launches += g.launches;
    const double t2 = trace ? now_us() : 0.0;
    if(trace)
        cudaEventRecord(tev[2], st);
    MSLAM_CUDA_OK(cudaStreamSynchronize(st));
    const double t3 = trace ? now_us() : 0.0;
    const int* hdr = reinterpret_cast<const int*>(ctx->h_one);
    const int n = hdr[0], m = std::max(0, std::min(std::min(n, capacity), ctx->kp_cap));
    memcpy(out_kp, ctx->h_one + ctx->one_kp_off, (size_t)m * sizeof(mslam_b200_keypoint));
    memcpy(out_desc, ctx->h_one + ctx->one_desc_off, (size_t)m * 32);
    n_out[0] = n;
    if(trace)
    {
        float g01 = 0.f, g12 = 0.f;
        cudaEventElapsedTime(&g01, tev[0], tev[1]);
        cudaEventElapsedTime(&g12, tev[1], tev[2]);
        const double t4 = now_us();
        const double v[6] = {t1 - t0, t2 - t1, t3 - t2, t4 - t3, g01 * 1e3, g12 * 1e3};
        for(int i = 0; i < 6; ++i)
            tacc[i] += v[i];
        if(++tn == 100)
        {
            fprintf(stderr, "[mslam_b200 latency path, us] host: copy-in enqueue %.1f, graph launch %.1f, wait %.1f, unpack %.1f | gpu: copy-in %.1f, graph %.1f\n",
                    tacc[0] / tn, tacc[1] / tn, tacc[2] / tn, tacc[3] / tn, tacc[4] / tn, tacc[5] / tn);
            tn = 0;
            for(double& a : tacc)
                a = 0.0;
            if(trace >= 2)
            {
                static const char* names[] = {"gray", "resize", "fast", "select", "blur", "describe", "knn", "compact"};
                for(int i = 0; i < tpairs; ++i)
                {
                    float a = 0.f, b = 0.f;
                    // relative to the first event inside the graph (gray's start)
                    const cudaError_t ea = cudaEventElapsedTime(&a, ctx->ev_pool[0], ctx->ev_pool[2 * i]);
                    const cudaError_t eb = cudaEventElapsedTime(&b, ctx->ev_pool[0], ctx->ev_pool[2 * i + 1]);
                    if(ea != cudaSuccess || eb != cudaSuccess)
                    {
                        fprintf(stderr, "    (event query: %s / %s)\n", cudaGetErrorName(ea), cudaGetErrorName(eb));
                        cudaGetLastError();
                    }
                    const int sg = ctx->ev_stage[i];
                    fprintf(stderr, "    %-8s %6.1f .. %6.1f\n", sg >= 0 && sg < 8 ? names[sg] : "?", a * 1e3, b * 1e3);
                }
                fprintf(stderr, "    graph end %6.1f\n", g12 * 1e3);
            }
        }
    }
    return (n > capacity || hdr[1]) ? MSLAM_B200_E_CAPACITY : MSLAM_B200_OK;
}

int mslam_b200_detect_batch(mslam_b200_ctx* ctx, const uint8_t* const* frames, int n_frames, int width, int height,
                            int stride_bytes, mslam_b200_keypoint* out_kp, uint8_t* out_desc, int capacity, int* n_out)
{
    if(!detect_batch_args_ok(ctx, frames, n_frames, width, height, stride_bytes, out_kp, out_desc, capacity, n_out) ||
       ctx->pend[0].active || ctx->pend[1].active)
        return MSLAM_B200_E_INVALID;
    if(n_frames == 1)
    {
        const int rc1 = detect_one_latency(ctx, frames[0], width, height, stride_bytes, out_kp, out_desc, capacity, n_out);
        if(rc1 != kTakeGeneralPath)
            return rc1;
    }
    int rc = mslam_b200_detect_batch_begin(ctx, frames, n_frames, width, height, stride_bytes, out_kp, out_desc, capacity, n_out);
    if(rc)
        return rc;
    rc = mslam_b200_detect_batch_end(ctx);
    MSLAM_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    return rc;
}

int mslam_b200_detect(mslam_b200_ctx* ctx, const uint8_t* bgr, int width, int height, int stride_bytes,
                      mslam_b200_keypoint* out_kp, uint8_t* out_desc, int capacity, int* n_out)
{
    const uint8_t* frames[1] = {bgr};
    return mslam_b200_detect_batch(ctx, frames, 1, width, height, stride_bytes, out_kp, out_desc, capacity, n_out);
}

int mslam_b200_describe(mslam_b200_ctx* ctx, const uint8_t* bgr, int width, int height, int stride_bytes,
                        mslam_b200_keypoint* kp, int n_kp, uint8_t* out_desc)
{
    if(!ctx || !bgr || !kp || !out_desc || n_kp < 0 || width < 1 || height < 1 || stride_bytes < 3 * width ||
       width > ctx->cfg.max_width || height > ctx->cfg.max_height)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    int rc = set_geometry(ctx, width, height);
    if(rc)
        return rc;
    cudaStream_t st = ctx->stream;
    const PipeParams& P0 = ctx->P;
    const int nl = P0.nlevels, icap = ctx->kp_cap;
    // bucket the given keypoints by level, keeping the caller's order inside a level
    std::vector<std::vector<int>> per(nl);
    for(int i = 0; i < n_kp; ++i)
    {
        const int l = kp[i].octave;
        if(l < 0 || l >= nl)
            return MSLAM_B200_E_INVALID;
        per[l].push_back(i);
    }
    std::vector<LevelCounters> ctr(nl);
    std::vector<SelRec> sel((size_t)nl * icap);
    std::vector<int> slot_to_src;
    for(int l = 0; l < nl; ++l)
    {
        if((int)per[l].size() > icap)
            return MSLAM_B200_E_CAPACITY;
        memset(&ctr[l], 0, sizeof(LevelCounters));
        ctr[l].n_final = (unsigned)per[l].size();
        const LevelDesc& L = P0.lv[l];
        for(size_t r = 0; r < per[l].size(); ++r)
        {
            const mslam_b200_keypoint& k = kp[per[l][r]];
            // same mapping OpenCV's computeOrbDescriptors uses: cvRound(pt * (1.f / scale_l))
            const int x = (int)lrintf(k.x * L.inv_scale), y = (int)lrintf(k.y * L.inv_scale);
            if(x < P0.edge || y < P0.edge || x >= L.w - P0.edge || y >= L.h - P0.edge)
                return MSLAM_B200_E_INVALID;
            sel[(size_t)l * icap + r].yx_score = ((unsigned)y << 20) | ((unsigned)x << 8);
            sel[(size_t)l * icap + r].response = k.response;
            slot_to_src.push_back(per[l][r]);
        }
    }
    if(n_kp > icap)
        return MSLAM_B200_E_CAPACITY;
    const size_t row = (size_t)width * 3;
    MSLAM_CUDA_OK(cudaMemcpy2DAsync(ctx->d_bgr, row, bgr, (size_t)stride_bytes, row, (size_t)height, cudaMemcpyHostToDevice, st));
    MSLAM_CUDA_OK(cudaMemsetAsync(ctx->d_zero, 0, ctx->zero_bytes, st));
    MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->d_ctr, ctr.data(), nl * sizeof(LevelCounters), cudaMemcpyHostToDevice, st));
    MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->d_sel, sel.data(), sel.size() * sizeof(SelRec), cudaMemcpyHostToDevice, st));
    ctx->P.frame0 = 0;
    run_pyramid(ctx, ctx->d_bgr, 1);
    launch_blur(st, ctx->d_pyr, ctx->d_blur, ctx->P, 0, ctx->P.nlevels, &ctx->launches);
    launch_orient_describe(st, ctx->d_pyr, ctx->d_blur, ctx->d_ctr, ctx->d_sel, ctx->d_pat, ctx->d_icw, ctx->d_kp, ctx->d_desc, ctx->d_counts, icap,
                           std::max(n_kp, 1), ctx->P, ctx->sm_count, &ctx->launches);
    std::vector<mslam_b200_keypoint> hk((size_t)std::max(n_kp, 1));
    std::vector<uint8_t> hd((size_t)std::max(n_kp, 1) * 32);
    if(n_kp > 0)
    {
        MSLAM_CUDA_OK(cudaMemcpyAsync(hk.data(), ctx->d_kp, n_kp * sizeof(mslam_b200_keypoint), cudaMemcpyDeviceToHost, st));
        MSLAM_CUDA_OK(cudaMemcpyAsync(hd.data(), ctx->d_desc, (size_t)n_kp * 32, cudaMemcpyDeviceToHost, st));
    }
    MSLAM_CUDA_OK(cudaStreamSynchronize(st));
    ctx->last_frames = 1;
    for(int s = 0; s < n_kp; ++s)
    {
        const int i = slot_to_src[s];
        kp[i].angle = hk[s].angle;
        kp[i].size = hk[s].size;
        memcpy(out_desc + (size_t)i * 32, hd.data() + (size_t)s * 32, 32);
    }
    return MSLAM_B200_OK;
}

int mslam_b200_debug_read(mslam_b200_ctx* ctx, int stage, int frame, int level, void* out, size_t bytes)
{
    if(!ctx || !out || frame < 0 || frame >= ctx->cfg.max_batch || level < 0 || level >= ctx->P.nlevels || ctx->geo_w == 0)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    const LevelDesc& L = ctx->P.lv[level];
    MSLAM_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    if(stage == MSLAM_B200_STAGE_PYRAMID || stage == MSLAM_B200_STAGE_BLURRED)
    {
        if(bytes < (size_t)L.w * L.h)
            return MSLAM_B200_E_INVALID;
        const uint8_t* base = (stage == MSLAM_B200_STAGE_PYRAMID ? ctx->d_pyr : ctx->d_blur) + L.img_off +
                              (long long)frame * L.img_frame_stride;
        MSLAM_CUDA_OK(cudaMemcpy2D(out, (size_t)L.w, base, (size_t)L.pitch, (size_t)L.w, (size_t)L.h, cudaMemcpyDeviceToHost));
        return MSLAM_B200_OK;
    }
    if(stage >= MSLAM_B200_STAGE_N_FAST && stage <= MSLAM_B200_STAGE_N_FINAL)
    {
        if(bytes < sizeof(int))
            return MSLAM_B200_E_INVALID;
        LevelCounters c;
        MSLAM_CUDA_OK(cudaMemcpy(&c, ctx->d_ctr + (size_t)frame * ctx->P.nlevels + level, sizeof(c), cudaMemcpyDeviceToHost));
        const unsigned v = stage == MSLAM_B200_STAGE_N_FAST ? c.n_fast : stage == MSLAM_B200_STAGE_N_CUT1 ? c.n_cut1 : c.n_final;
        *(int*)out = (int)v;
        return MSLAM_B200_OK;
    }
    return MSLAM_B200_E_INVALID;
}

// ---- matching ----------------------------------------------------------------------------------------------------
// gather 32-byte descriptors out of a strided record array (e.g. OrbKeypoint, 64 B, descriptor at +32)
__global__ void k_pack_desc(const uint8_t* __restrict__ raw, int stride, int n, uint8_t* __restrict__ dense)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x; // one 4-byte word per thread
    if(i >= n * 8)
        return;
    const int row = i >> 3, wd = i & 7;
    const uint8_t* s = raw + (size_t)row * stride + wd * 4;
    uchar4 v = make_uchar4(s[0], s[1], s[2], s[3]);
    reinterpret_cast<uchar4*>(dense)[i] = v;
}

static int upload_desc(mslam_b200_ctx* ctx, const uint8_t* src, int n, int stride, uint8_t* d_dst)
{
    if(n <= 0)
        return MSLAM_B200_OK;
    cudaStream_t st = ctx->stream;
    if(stride == 32)
        MSLAM_CUDA_OK(cudaMemcpyAsync(d_dst, src, (size_t)n * 32, cudaMemcpyHostToDevice, st));
    else if(stride > 32 && stride <= 128)
    {
        MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->d_raw, src, (size_t)(n - 1) * stride + 32, cudaMemcpyHostToDevice, st));
        k_pack_desc<<<(n * 8 + 255) / 256, 256, 0, st>>>(ctx->d_raw, stride, n, d_dst);
        ++ctx->launches;
    }
    else
        MSLAM_CUDA_OK(cudaMemcpy2DAsync(d_dst, 32, src, (size_t)stride, 32, (size_t)n, cudaMemcpyHostToDevice, st));
    return MSLAM_B200_OK;
}

// kNN (k = 2) of every query set against its train set -> packed keys; `both` also runs train against query (cross
// check).  Large problems go to the tensor-core kernel (one int8 expansion of each set, shared by both directions), small
// ones to the scalar kernel, whose one-warp-per-64-queries granularity fills the GPU better when there is little work.
static int run_knn(mslam_b200_ctx* ctx, const uint8_t* d_q, const int* d_nq, const uint8_t* d_t, const int* d_nt, int n_pairs,
                   int cap, bool both)
{
    cudaStream_t st = ctx->stream;
    const long long ctas = (long long)((cap + 255) / 256) * n_pairs;
    const bool mma = ctx->knn_mma == 2 || (ctx->knn_mma == 1 && cap >= 256 && ctas >= ctx->sm_count / 2);
    if(mma)
    {
        launch_expand_pm1_x2(st, d_q, d_nq, ctx->d_xa, d_t, d_nt, ctx->d_xb, n_pairs, cap, &ctx->launches);
        MSLAM_CUDA_OK(launch_knn2_mma(st, ctx->d_xa, d_nq, ctx->d_xb, d_nt, n_pairs, cap, ctx->d_key0, ctx->d_key1, &ctx->launches));
        if(both)
            MSLAM_CUDA_OK(launch_knn2_mma(st, ctx->d_xb, d_nt, ctx->d_xa, d_nq, n_pairs, cap, ctx->d_key0b, ctx->d_key1b, &ctx->launches));
        return MSLAM_B200_OK;
    }
    launch_knn2(st, d_q, d_nq, d_t, d_nt, n_pairs, cap, ctx->d_key0, ctx->d_key1, ctx->d_part0, ctx->d_part1, ctx->sm_count, &ctx->launches);
    if(both)
        launch_knn2(st, d_t, d_nt, d_q, d_nq, n_pairs, cap, ctx->d_key0b, ctx->d_key1b, ctx->d_part0, ctx->d_part1, ctx->sm_count, &ctx->launches);
    return MSLAM_B200_OK;
}

static int upload_counts(mslam_b200_ctx* ctx, int n_from, int n_to)
{
    ctx->h_n2[0] = n_from;
    ctx->h_n2[1] = n_to;
    MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->d_nfrom, &ctx->h_n2[0], sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->d_nto, &ctx->h_n2[1], sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    return MSLAM_B200_OK;
}

int mslam_b200_match_batch_dev(mslam_b200_ctx* ctx, int mode, const uint8_t* d_from, const int* d_n_from, const uint8_t* d_to,
                               const int* d_n_to, int n_pairs, int capacity, int* d_out_a, int* d_out_b, int* d_out_dist,
                               int* d_n_out)
{
    if(!ctx || !d_from || !d_to || !d_n_from || !d_n_to || !d_out_a || !d_out_b || !d_n_out || n_pairs < 1 ||
       n_pairs > ctx->cfg.max_batch || capacity < 1 || capacity > ctx->desc_cap || (mode != 0 && mode != 1))
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    cudaStream_t st = ctx->stream;
    // query = `to`, train = `from` (orb_feature.cpp:96)
    {
        StageTimer t(ctx, MSLAM_B200_T_KNN);
        const int rc = run_knn(ctx, d_to, d_n_to, d_from, d_n_from, n_pairs, capacity, mode == 1);
        if(rc)
            return rc;
    }
    {
        StageTimer t(ctx, MSLAM_B200_T_COMPACT);
        if(mode == 0)
            launch_ratio_compact(st, ctx->d_key0, ctx->d_key1, d_n_to, d_n_from, n_pairs, capacity, ctx->cfg.ratio, d_out_a,
                                 d_out_b, d_out_dist, d_n_out, &ctx->launches);
        else
            launch_mutual_compact(st, ctx->d_key0, ctx->d_key0b, d_n_to, d_n_from, n_pairs, capacity, d_out_a, d_out_b,
                                  d_out_dist, d_n_out, &ctx->launches);
    }
    MSLAM_CUDA_OK(cudaGetLastError());
    return MSLAM_B200_OK;
}

int mslam_b200_match_batch_begin(mslam_b200_ctx* ctx, int mode, const uint8_t* from, const int* n_from, const uint8_t* to,
                                 const int* n_to, int n_pairs, int capacity, int* out_a, int* out_b, int* out_dist, int* n_out)
{
    if(!ctx || !from || !to || !n_from || !n_to || !out_a || !out_b || !n_out || n_pairs < 1 ||
       n_pairs > ctx->cfg.max_batch || capacity < 1 || capacity > ctx->desc_cap || ctx->match_pending)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    cudaStream_t st = ctx->stream, s_in = ctx->copy[0], s_out = ctx->copy[1];
    const size_t blk = (size_t)n_pairs * capacity;
    // descriptors go in on the copy-in engine and results come out on the copy-out engine, so that neither transfer
    // stalls kernels of a detect batch queued on the context stream around this match
    // only the rows that hold descriptors cross the bus: one strided copy per set, as wide as the pair with the most of them
    // (the counts are host arrays; capacity is typically 30 % above the feature count)
    int max_from = 0, max_to = 0;
    for(int i = 0; i < n_pairs; ++i)
    {
        max_from = std::max(max_from, std::min(std::max(n_from[i], 0), capacity));
        max_to = std::max(max_to, std::min(std::max(n_to[i], 0), capacity));
    }
    const size_t pitch = (size_t)capacity * 32;
    if(max_from > 0)
        MSLAM_CUDA_OK(cudaMemcpy2DAsync(ctx->d_mfrom, pitch, from, pitch, (size_t)max_from * 32, (size_t)n_pairs, cudaMemcpyHostToDevice, s_in));
    if(max_to > 0)
        MSLAM_CUDA_OK(cudaMemcpy2DAsync(ctx->d_mto, pitch, to, pitch, (size_t)max_to * 32, (size_t)n_pairs, cudaMemcpyHostToDevice, s_in));
    MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->d_nfrom, n_from, n_pairs * sizeof(int), cudaMemcpyHostToDevice, s_in));
    MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->d_nto, n_to, n_pairs * sizeof(int), cudaMemcpyHostToDevice, s_in));
    MSLAM_CUDA_OK(cudaEventRecord(ctx->ev_match_in, s_in));
    MSLAM_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_match_in, 0));
    int rc = mslam_b200_match_batch_dev(ctx, mode, ctx->d_mfrom, ctx->d_nfrom, ctx->d_mto, ctx->d_nto, n_pairs, capacity,
                                        ctx->d_out_a, ctx->d_out_b, ctx->d_out_d, ctx->d_nout);
    if(rc)
        return rc;
    MSLAM_CUDA_OK(cudaEventRecord(ctx->ev_match_k, st));
    MSLAM_CUDA_OK(cudaStreamWaitEvent(s_out, ctx->ev_match_k, 0));
    MSLAM_CUDA_OK(cudaMemcpyAsync(out_a, ctx->d_out_a, blk * sizeof(int), cudaMemcpyDeviceToHost, s_out));
    MSLAM_CUDA_OK(cudaMemcpyAsync(out_b, ctx->d_out_b, blk * sizeof(int), cudaMemcpyDeviceToHost, s_out));
    if(out_dist)
        MSLAM_CUDA_OK(cudaMemcpyAsync(out_dist, ctx->d_out_d, blk * sizeof(int), cudaMemcpyDeviceToHost, s_out));
    MSLAM_CUDA_OK(cudaMemcpyAsync(n_out, ctx->d_nout, n_pairs * sizeof(int), cudaMemcpyDeviceToHost, s_out));
    MSLAM_CUDA_OK(cudaEventRecord(ctx->ev_match_done, s_out));
    ctx->match_pending = true;
    return MSLAM_B200_OK;
}

int mslam_b200_match_batch_end(mslam_b200_ctx* ctx)
{
    if(!ctx || !ctx->match_pending)
        return MSLAM_B200_E_INVALID;
    ctx->match_pending = false;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    MSLAM_CUDA_OK(cudaEventSynchronize(ctx->ev_match_done));
    return MSLAM_B200_OK;
}

int mslam_b200_match_batch(mslam_b200_ctx* ctx, int mode, const uint8_t* from, const int* n_from, const uint8_t* to,
                           const int* n_to, int n_pairs, int capacity, int* out_a, int* out_b, int* out_dist, int* n_out)
{
    const int rc = mslam_b200_match_batch_begin(ctx, mode, from, n_from, to, n_to, n_pairs, capacity, out_a, out_b, out_dist, n_out);
    return rc ? rc : mslam_b200_match_batch_end(ctx);
}

static int match_single(mslam_b200_ctx* ctx, int mode, const uint8_t* from, int n_from, int from_stride, const uint8_t* to,
                        int n_to, int to_stride, int* n_found)
{
    // results land in ctx->lat_{a,b,d}; *n_found = count
    *n_found = 0;
    if(n_from < 0 || n_to < 0 || n_from > ctx->desc_cap || n_to > ctx->desc_cap || from_stride < 32 || to_stride < 32)
        return MSLAM_B200_E_INVALID;
    if(n_to < 1 || n_from < (mode == 0 ? 2 : 1))
        return MSLAM_B200_OK; // the reference is undefined here (orb_feature.cpp:88-91,101); contract: no match
    if(!from || !to)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    cudaStream_t st = ctx->stream;
    // The call the frontend makes once per frame (rgbd_feature_frontend.cpp:237).  Both descriptor sets are packed by the host
    // into ONE pinned block with their counts in front (the caller's arrays are pageable, 64-byte OrbKeypoint records), one
    // copy takes it in, the kernels run on the packed sets in place, one copy brings count + indices + distances back:
    // two copies, the kernels and a wait instead of two pageable copies, two pack kernels, two count copies and four
    // result copies (0.072 -> see tools/latency.py).
    int* hh = reinterpret_cast<int*>(ctx->h_m1);
    hh[0] = n_from, hh[1] = n_to;
    auto pack = [](uint8_t* dst, const uint8_t* src, int n, int stride) {
        if(stride == 32)
            memcpy(dst, src, (size_t)n * 32);
        else
            for(int i = 0; i < n; ++i)
                memcpy(dst + (size_t)i * 32, src + (size_t)i * stride, 32);
    };
    pack(ctx->h_m1 + 32, from, n_from, from_stride);
    pack(ctx->h_m1 + 32 + (size_t)n_from * 32, to, n_to, to_stride);
    MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->d_m1, ctx->h_m1, 32 + (size_t)(n_from + n_to) * 32, cudaMemcpyHostToDevice, st));
    const int* d_nf = reinterpret_cast<const int*>(ctx->d_m1);
    const uint8_t* d_from = ctx->d_m1 + 32;
    const uint8_t* d_to = d_from + (size_t)n_from * 32;
    const int cap = ctx->desc_cap;
    int* o_n = reinterpret_cast<int*>(ctx->d_mo);
    int* o_a = reinterpret_cast<int*>(ctx->d_mo + 16);
    int rc = mslam_b200_match_batch_dev(ctx, mode, d_from, d_nf, d_to, d_nf + 1, 1, cap, o_a, o_a + cap, o_a + 2 * cap, o_n);
    if(rc)
        return rc;
    const int* h_n = reinterpret_cast<const int*>(ctx->h_mo);
    const int* h_a = reinterpret_cast<const int*>(ctx->h_mo + 16);
    if(ctx->mo_bytes <= (64u << 10))
        MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->h_mo, ctx->d_mo, ctx->mo_bytes, cudaMemcpyDeviceToHost, st));
    else
    {
        // a context sized for very large sets: only the n_to entries that can be filled
        MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->h_mo, ctx->d_mo, 16, cudaMemcpyDeviceToHost, st));
        for(int k = 0; k < 3; ++k)
            MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->h_mo + 16 + (size_t)k * cap * sizeof(int), ctx->d_mo + 16 + (size_t)k * cap * sizeof(int),
                                          (size_t)n_to * sizeof(int), cudaMemcpyDeviceToHost, st));
    }
    MSLAM_CUDA_OK(cudaStreamSynchronize(st));
    ctx->lat_a = h_a, ctx->lat_b = h_a + cap, ctx->lat_d = h_a + 2 * cap;
    *n_found = h_n[0];
    return MSLAM_B200_OK;
}

int mslam_b200_match_knn_ratio(mslam_b200_ctx* ctx, const uint8_t* from, int n_from, int from_stride, const uint8_t* to,
                               int n_to, int to_stride, uint64_t* out_from_idx, uint64_t* out_to_idx, int* out_dist,
                               int capacity, int* n_out)
{
    if(!ctx || !n_out || capacity < 0 || (capacity > 0 && (!out_from_idx || !out_to_idx)))
        return MSLAM_B200_E_INVALID;
    int n = 0;
    int rc = match_single(ctx, 0, from, n_from, from_stride, to, n_to, to_stride, &n);
    if(rc)
        return rc;
    *n_out = n;
    const int m = std::min(n, capacity);
    for(int i = 0; i < m; ++i)
    {
        out_from_idx[i] = (uint64_t)ctx->lat_a[i];
        out_to_idx[i] = (uint64_t)ctx->lat_b[i];
        if(out_dist)
            out_dist[i] = ctx->lat_d[i];
    }
    return n > capacity ? MSLAM_B200_E_CAPACITY : MSLAM_B200_OK;
}

int mslam_b200_match_crosscheck(mslam_b200_ctx* ctx, const uint8_t* query, int n_query, int query_stride,
                                const uint8_t* train, int n_train, int train_stride, int* out_query_idx, int* out_train_idx,
                                int* out_dist, int capacity, int* n_out)
{
    if(!ctx || !n_out || capacity < 0 || (capacity > 0 && (!out_query_idx || !out_train_idx)))
        return MSLAM_B200_E_INVALID;
    int n = 0;
    // match_single: "to" is the kNN query, "from" the train set
    int rc = match_single(ctx, 1, train, n_train, train_stride, query, n_query, query_stride, &n);
    if(rc)
        return rc;
    *n_out = n;
    const int m = std::min(n, capacity);
    for(int i = 0; i < m; ++i)
    {
        out_train_idx[i] = ctx->lat_a[i];
        out_query_idx[i] = ctx->lat_b[i];
        if(out_dist)
            out_dist[i] = ctx->lat_d[i];
    }
    return n > capacity ? MSLAM_B200_E_CAPACITY : MSLAM_B200_OK;
}

int mslam_b200_knn2(mslam_b200_ctx* ctx, const uint8_t* query, int n_query, int query_stride, const uint8_t* train,
                    int n_train, int train_stride, int* idx0, int* dist0, int* idx1, int* dist1)
{
    if(!ctx || !idx0 || !dist0 || !idx1 || !dist1 || n_query < 0 || n_train < 0 || n_query > ctx->desc_cap ||
       n_train > ctx->desc_cap || query_stride < 32 || train_stride < 32)
        return MSLAM_B200_E_INVALID;
    if(n_query == 0)
        return MSLAM_B200_OK;
    if(!query || (n_train > 0 && !train))
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    cudaStream_t st = ctx->stream;
    int rc = upload_desc(ctx, train, n_train, train_stride, ctx->d_mfrom);
    if(rc)
        return rc;
    rc = upload_desc(ctx, query, n_query, query_stride, ctx->d_mto);
    if(rc)
        return rc;
    rc = upload_counts(ctx, n_train, n_query);
    if(rc)
        return rc;
    const int cap = ctx->desc_cap;
    rc = run_knn(ctx, ctx->d_mto, ctx->d_nto, ctx->d_mfrom, ctx->d_nfrom, 1, cap, false);
    if(rc)
        return rc;
    // unpack into the four int scratch arrays (out_a, out_b, out_d, key0b reinterpreted)
    int* d_i1 = reinterpret_cast<int*>(ctx->d_key0b);
    launch_unpack_knn(st, ctx->d_key0, ctx->d_key1, n_query, ctx->d_out_a, ctx->d_out_d, d_i1, ctx->d_out_b, &ctx->launches);
    MSLAM_CUDA_OK(cudaMemcpyAsync(idx0, ctx->d_out_a, n_query * sizeof(int), cudaMemcpyDeviceToHost, st));
    MSLAM_CUDA_OK(cudaMemcpyAsync(dist0, ctx->d_out_d, n_query * sizeof(int), cudaMemcpyDeviceToHost, st));
    MSLAM_CUDA_OK(cudaMemcpyAsync(idx1, d_i1, n_query * sizeof(int), cudaMemcpyDeviceToHost, st));
    MSLAM_CUDA_OK(cudaMemcpyAsync(dist1, ctx->d_out_b, n_query * sizeof(int), cudaMemcpyDeviceToHost, st));
    MSLAM_CUDA_OK(cudaStreamSynchronize(st));
    return MSLAM_B200_OK;
}

// ---- keyframe database -------------------------------------------------------------------------------------------
int mslam_b200_db_scores_dev(mslam_b200_ctx* ctx, const uint8_t* d_queries, const int* d_q_counts, int n_query_frames,
                             int q_cap, const uint8_t* d_db, const int* d_k_counts, int n_keyframes, int k_cap, int* d_scores)
{
    if(!ctx || !d_queries || !d_q_counts || !d_db || !d_k_counts || !d_scores || n_query_frames < 1 ||
       n_query_frames > 65535 || n_keyframes < 1 || q_cap < 1 || k_cap < 1 || (size_t)q_cap * 32 > 200 * 1024)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    StageTimer t(ctx, MSLAM_B200_T_DB);
    cudaStream_t st = ctx->stream;
    // Large scorings go to the tensor-core kernel: the query frames and (a chunk at a time, bounded scratch) the keyframes
    // are expanded to +1/-1 int8 rows, which costs ~0.2 % of the scoring itself.
    const long long work = (long long)n_query_frames * n_keyframes * ((k_cap + 255) / 256);
    if(ctx->knn_mma == 2 || (ctx->knn_mma == 1 && q_cap >= 256 && work >= 4LL * ctx->sm_count))
    {
        const size_t per_kf = (size_t)knn_mma_pad(k_cap) * 256, per_qf = (size_t)knn_mma_pad(q_cap) * 256;
        const int kf_chunk = (int)std::min<size_t>(std::min<size_t>((size_t)n_keyframes, 65535), std::max<size_t>(1, ((size_t)4 << 30) / per_kf));
        if(ctx->xdb_bytes < kf_chunk * per_kf || ctx->xqf_bytes < n_query_frames * per_qf)
        {
            MSLAM_CUDA_OK(cudaStreamSynchronize(st));
            if(ctx->xdb_bytes < kf_chunk * per_kf)
            {
                cudaFree(ctx->d_xdb);
                ctx->d_xdb = nullptr, ctx->xdb_bytes = 0;
                MSLAM_CUDA_OK(cudaMalloc(&ctx->d_xdb, kf_chunk * per_kf));
                ctx->xdb_bytes = kf_chunk * per_kf;
            }
            if(ctx->xqf_bytes < n_query_frames * per_qf)
            {
                cudaFree(ctx->d_xqf);
                ctx->d_xqf = nullptr, ctx->xqf_bytes = 0;
                MSLAM_CUDA_OK(cudaMalloc(&ctx->d_xqf, n_query_frames * per_qf));
                ctx->xqf_bytes = n_query_frames * per_qf;
            }
        }
        MSLAM_CUDA_OK(cudaMemsetAsync(d_scores, 0, (size_t)n_query_frames * n_keyframes * sizeof(int), st));
        launch_expand_pm1(st, d_queries, d_q_counts, n_query_frames, q_cap, ctx->d_xqf, &ctx->launches);
        for(int kf0 = 0; kf0 < n_keyframes; kf0 += kf_chunk)
        {
            const int nk = std::min(kf_chunk, n_keyframes - kf0);
            launch_expand_pm1(st, d_db + (size_t)kf0 * k_cap * 32, d_k_counts + kf0, nk, k_cap, ctx->d_xdb, &ctx->launches);
            MSLAM_CUDA_OK(launch_db_scores_mma(st, ctx->d_xqf, d_q_counts, n_query_frames, q_cap, ctx->d_xdb, d_k_counts + kf0, nk,
                                               k_cap, ctx->cfg.ratio, d_scores + kf0, n_keyframes, &ctx->launches));
        }
        return MSLAM_B200_OK;
    }
    MSLAM_CUDA_OK(launch_db_scores(ctx->stream, d_queries, d_q_counts, n_query_frames, q_cap, d_db, d_k_counts, n_keyframes,
                                   k_cap, ctx->cfg.ratio, d_scores, ctx->sm_count, &ctx->launches));
    return MSLAM_B200_OK;
}

int mslam_b200_db_topk_dev(mslam_b200_ctx* ctx, const int* d_scores, int n_query_frames, int n_keyframes, int id_base, int k,
                           int* d_out_score, int* d_out_id)
{
    if(!ctx || !d_scores || !d_out_score || !d_out_id || n_query_frames < 1 || n_keyframes < 1 || k < 1 || id_base < 0)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    launch_topk(ctx->stream, d_scores, nullptr, 1, n_keyframes, 0, n_query_frames, id_base, k, d_out_score, d_out_id,
                &ctx->launches);
    MSLAM_CUDA_OK(cudaGetLastError());
    return MSLAM_B200_OK;
}

int mslam_b200_db_merge_topk_dev(mslam_b200_ctx* ctx, const int* d_scores, const int* d_ids, int n_lists, int n_query_frames,
                                 int k, int* d_out_score, int* d_out_id)
{
    if(!ctx || !d_scores || !d_ids || !d_out_score || !d_out_id || n_lists < 1 || n_query_frames < 1 || k < 1)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    launch_topk(ctx->stream, d_scores, d_ids, n_lists, k, (long long)n_query_frames * k, n_query_frames, 0, k, d_out_score, d_out_id,
                &ctx->launches);
    MSLAM_CUDA_OK(cudaGetLastError());
    return MSLAM_B200_OK;
}

int mslam_b200_db_query(mslam_b200_ctx* ctx, const uint8_t* queries, const int* q_counts, int n_query_frames, int q_cap,
                        const uint8_t* db, const int* k_counts, int n_keyframes, int k_cap, int id_base, int k,
                        int* out_score, int* out_id, int* out_all_scores)
{
    if(!ctx || !queries || !q_counts || !db || !k_counts || !out_score || !out_id || n_query_frames < 1 ||
       n_keyframes < 1 || k < 1)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    cudaStream_t st = ctx->stream;
    uint8_t *d_q = nullptr, *d_db = nullptr;
    int *d_qc = nullptr, *d_kc = nullptr, *d_sc = nullptr, *d_os = nullptr, *d_oi = nullptr;
    const size_t qb = (size_t)n_query_frames * q_cap * 32, kb = (size_t)n_keyframes * k_cap * 32;
    const size_t sb = (size_t)n_query_frames * n_keyframes * sizeof(int), ob = (size_t)n_query_frames * k * sizeof(int);
    int rc = MSLAM_B200_OK;
    cudaError_t e;
#define DBQ(expr)                                                                                                      \
    if(rc == MSLAM_B200_OK && (e = (expr)) != cudaSuccess)                                                             \
    rc = ctx_fail(ctx, e, #expr)
    DBQ(cudaMalloc(&d_q, qb));
    DBQ(cudaMalloc(&d_db, kb));
    DBQ(cudaMalloc(&d_qc, n_query_frames * sizeof(int)));
    DBQ(cudaMalloc(&d_kc, n_keyframes * sizeof(int)));
    DBQ(cudaMalloc(&d_sc, sb));
    DBQ(cudaMalloc(&d_os, ob));
    DBQ(cudaMalloc(&d_oi, ob));
    DBQ(cudaMemcpyAsync(d_q, queries, qb, cudaMemcpyHostToDevice, st));
    DBQ(cudaMemcpyAsync(d_db, db, kb, cudaMemcpyHostToDevice, st));
    DBQ(cudaMemcpyAsync(d_qc, q_counts, n_query_frames * sizeof(int), cudaMemcpyHostToDevice, st));
    DBQ(cudaMemcpyAsync(d_kc, k_counts, n_keyframes * sizeof(int), cudaMemcpyHostToDevice, st));
    if(rc == MSLAM_B200_OK)
        rc = mslam_b200_db_scores_dev(ctx, d_q, d_qc, n_query_frames, q_cap, d_db, d_kc, n_keyframes, k_cap, d_sc);
    if(rc == MSLAM_B200_OK)
        rc = mslam_b200_db_topk_dev(ctx, d_sc, n_query_frames, n_keyframes, id_base, k, d_os, d_oi);
    DBQ(cudaMemcpyAsync(out_score, d_os, ob, cudaMemcpyDeviceToHost, st));
    DBQ(cudaMemcpyAsync(out_id, d_oi, ob, cudaMemcpyDeviceToHost, st));
    if(out_all_scores)
        DBQ(cudaMemcpyAsync(out_all_scores, d_sc, sb, cudaMemcpyDeviceToHost, st));
    DBQ(cudaStreamSynchronize(st));
#undef DBQ
    void* frees[] = {d_q, d_db, d_qc, d_kc, d_sc, d_os, d_oi};
    for(void* p : frees)
        if(p)
            cudaFree(p);
    return rc;
}

// ---- sharded keyframe database: one shard per rank, per-query top-k all-gathered over NCCL and merged on the device ------------
void mslam_b200_db_shard_range(long long n_keyframes, int rank, int world, long long* k0, long long* k1)
{
    long long a = 0, b = 0;
    if(world >= 1 && rank >= 0 && rank < world && n_keyframes >= 0)
    {
        const long long base = n_keyframes / world, extra = n_keyframes % world;
        a = rank * base + std::min<long long>(rank, extra);
        b = a + base + (rank < extra ? 1 : 0);
    }
    if(k0)
        *k0 = a;
    if(k1)
        *k1 = b;
}

static int nccl_fail(mslam_b200_ctx* ctx, int r, const char* what)
{
    NcclApi& api = nccl_api();
    snprintf(ctx->err, sizeof(ctx->err), "%s: NCCL error %d (%s)", what, r, api.GetErrorString ? api.GetErrorString(r) : "?");
    return MSLAM_B200_E_CUDA;
}

int mslam_b200_nccl_unique_id(void* out_id128)
{
    if(!out_id128)
        return MSLAM_B200_E_INVALID;
    NcclApi& api = nccl_api();
    if(!api.ok())
        return MSLAM_B200_E_CUDA;
    ncclUniqueId id;
    if(api.GetUniqueId(&id) != 0)
        return MSLAM_B200_E_CUDA;
    memcpy(out_id128, &id, sizeof(id));
    return MSLAM_B200_OK;
}

int mslam_b200_db_detach(mslam_b200_ctx* ctx)
{
    if(!ctx)
        return MSLAM_B200_E_INVALID;
    if(ctx->db_comm)
    {
        cudaSetDevice(ctx->cfg.device);
        cudaStreamSynchronize(ctx->stream);
        // every collective this context enqueued has completed (stream synchronised), so nothing of ours is outstanding;
        // ncclCommAbort releases the communicator without waiting for the peers -- ncclCommDestroy can block for ever when a
        // peer process has already exited (measured: a rank that outlived its peer hung in teardown)
        if(ctx->db_own_comm)
            nccl_api().CommAbort(ctx->db_comm);
    }
    ctx->db_comm = nullptr, ctx->db_own_comm = false, ctx->db_rank = 0, ctx->db_world = 1;
    return MSLAM_B200_OK;
}

int mslam_b200_db_attach(mslam_b200_ctx* ctx, int rank, int world, const void* nccl_unique_id)
{
    if(!ctx || world < 1 || rank < 0 || rank >= world || (world > 1 && !nccl_unique_id))
        return MSLAM_B200_E_INVALID;
    mslam_b200_db_detach(ctx);
    if(world > 1)
    {
        NcclApi& api = nccl_api();
        if(!api.ok())
        {
            snprintf(ctx->err, sizeof(ctx->err), "%s", api.why);
            return MSLAM_B200_E_CUDA; // no fallback: a sharded database without its collective is not a database
        }
        MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
        ncclUniqueId id;
        memcpy(&id, nccl_unique_id, sizeof(id));
        const int r = api.CommInitRank(&ctx->db_comm, world, id, rank);
        if(r != 0)
        {
            ctx->db_comm = nullptr;
            return nccl_fail(ctx, r, "ncclCommInitRank");
        }
        ctx->db_own_comm = true;
    }
    ctx->db_rank = rank, ctx->db_world = world;
    return MSLAM_B200_OK;
}

int mslam_b200_db_attach_comm(mslam_b200_ctx* ctx, void* nccl_comm, int rank, int world)
{
    if(!ctx || world < 1 || rank < 0 || rank >= world || (world > 1 && !nccl_comm))
        return MSLAM_B200_E_INVALID;
    mslam_b200_db_detach(ctx);
    if(world > 1)
    {
        NcclApi& api = nccl_api();
        if(!api.ok())
        {
            snprintf(ctx->err, sizeof(ctx->err), "%s", api.why);
            return MSLAM_B200_E_CUDA;
        }
        int n = 0, me = -1;
        if(api.CommCount((ncclComm_t)nccl_comm, &n) != 0 || api.CommUserRank((ncclComm_t)nccl_comm, &me) != 0 || n != world || me != rank)
            return MSLAM_B200_E_INVALID; // the communicator is not the (rank, world) the caller says
        ctx->db_comm = (ncclComm_t)nccl_comm;
    }
    ctx->db_rank = rank, ctx->db_world = world;
    return MSLAM_B200_OK;
}

int mslam_b200_db_shard_info(mslam_b200_ctx* ctx, int* rank, int* world)
{
    if(!ctx)
        return MSLAM_B200_E_INVALID;
    if(rank)
        *rank = ctx->db_rank;
    if(world)
        *world = ctx->db_world;
    return MSLAM_B200_OK;
}

// scores -> local top-k -> ncclAllGather -> k-way merge, all enqueued on the context stream: no host synchronisation
// between them (scratch growth, on the first call of a shape, is the one exception)
int mslam_b200_db_query_sharded_dev(mslam_b200_ctx* ctx, const uint8_t* d_queries, const int* d_q_counts, int n_query_frames,
                                    int q_cap, const uint8_t* d_db_shard, const int* d_k_counts, int n_local, int k_cap,
                                    int id_base, int k, int* d_out_score, int* d_out_id)
{
    if(!ctx || !d_queries || !d_q_counts || !d_out_score || !d_out_id || n_query_frames < 1 || n_query_frames > 65535 || q_cap < 1 ||
       n_local < 0 || (n_local > 0 && (!d_db_shard || !d_k_counts || k_cap < 1)) || k < 1 || id_base < 0)
        return MSLAM_B200_E_INVALID;
    const int world = ctx->db_world;
    if(world > 1 && !ctx->db_comm)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    cudaStream_t st = ctx->stream;
    const size_t qk = (size_t)n_query_frames * k;
    const size_t n_sc = (size_t)n_query_frames * std::max(n_local, 1);
    const size_t need = (n_sc + 2 * qk + (size_t)world * 2 * qk) * sizeof(int);
    if(ctx->dbs_bytes < need)
    {
        MSLAM_CUDA_OK(cudaStreamSynchronize(st));
        cudaFree(ctx->d_dbs);
        ctx->d_dbs = nullptr, ctx->dbs_bytes = 0;
        MSLAM_CUDA_OK(cudaMalloc(&ctx->d_dbs, need));
        ctx->dbs_bytes = need;
    }
    int* d_sc = ctx->d_dbs;
    int* d_local = d_sc + n_sc;        // [2][Q][k]: scores, then ids
    int* d_gather = d_local + 2 * qk;  // [world][2][Q][k]
    if(n_local > 0)
    {
        int rc = mslam_b200_db_scores_dev(ctx, d_queries, d_q_counts, n_query_frames, q_cap, d_db_shard, d_k_counts, n_local, k_cap, d_sc);
        if(rc != MSLAM_B200_OK)
            return rc;
        rc = mslam_b200_db_topk_dev(ctx, d_sc, n_query_frames, n_local, id_base, k, d_local, d_local + qk);
        if(rc != MSLAM_B200_OK)
            return rc;
    }
    else // an empty shard still takes part in the collective: (-1, -1) pads
        MSLAM_CUDA_OK(cudaMemsetAsync(d_local, 0xff, 2 * qk * sizeof(int), st));
    if(world == 1)
    {
        MSLAM_CUDA_OK(cudaMemcpyAsync(d_out_score, d_local, qk * sizeof(int), cudaMemcpyDeviceToDevice, st));
        MSLAM_CUDA_OK(cudaMemcpyAsync(d_out_id, d_local + qk, qk * sizeof(int), cudaMemcpyDeviceToDevice, st));
        return MSLAM_B200_OK;
    }
    const int r = nccl_api().AllGather(d_local, d_gather, 2 * qk, kNcclInt32, ctx->db_comm, st);
    if(r != 0)
        return nccl_fail(ctx, r, "ncclAllGather");
    launch_topk(st, d_gather, d_gather + qk, world, k, (long long)(2 * qk), n_query_frames, 0, k, d_out_score, d_out_id, &ctx->launches);
    MSLAM_CUDA_OK(cudaGetLastError());
    return MSLAM_B200_OK;
}

// host queries in, host top-k out; the shard stays resident on the device (loaded once by the caller)
int mslam_b200_db_query_sharded(mslam_b200_ctx* ctx, const uint8_t* queries, const int* q_counts, int n_query_frames, int q_cap,
                                const uint8_t* d_db_shard, const int* d_k_counts, int n_local, int k_cap, int id_base, int k,
                                int* out_score, int* out_id)
{
    if(!ctx || !queries || !q_counts || !out_score || !out_id || n_query_frames < 1 || q_cap < 1 || k < 1)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    cudaStream_t st = ctx->stream;
    const size_t qb = (size_t)n_query_frames * q_cap * 32, cb = ((size_t)n_query_frames * sizeof(int) + 15) / 16 * 16;
    const size_t ob = (size_t)n_query_frames * k * sizeof(int);
    const size_t need = qb + cb + 2 * ob;
    if(ctx->dbq_bytes < need)
    {
        MSLAM_CUDA_OK(cudaStreamSynchronize(st));
        cudaFree(ctx->d_dbq);
        ctx->d_dbq = nullptr, ctx->dbq_bytes = 0;
        MSLAM_CUDA_OK(cudaMalloc(&ctx->d_dbq, need));
        ctx->dbq_bytes = need;
    }
    uint8_t* d_q = ctx->d_dbq;
    int* d_qc = reinterpret_cast<int*>(d_q + qb);
    int* d_os = reinterpret_cast<int*>(d_q + qb + cb);
    int* d_oi = d_os + (size_t)n_query_frames * k;
    MSLAM_CUDA_OK(cudaMemcpyAsync(d_q, queries, qb, cudaMemcpyHostToDevice, st));
    MSLAM_CUDA_OK(cudaMemcpyAsync(d_qc, q_counts, n_query_frames * sizeof(int), cudaMemcpyHostToDevice, st));
    const int rc = mslam_b200_db_query_sharded_dev(ctx, d_q, d_qc, n_query_frames, q_cap, d_db_shard, d_k_counts, n_local, k_cap,
                                                   id_base, k, d_os, d_oi);
    if(rc != MSLAM_B200_OK)
        return rc;
    MSLAM_CUDA_OK(cudaMemcpyAsync(out_score, d_os, ob, cudaMemcpyDeviceToHost, st));
    MSLAM_CUDA_OK(cudaMemcpyAsync(out_id, d_oi, ob, cudaMemcpyDeviceToHost, st));
    MSLAM_CUDA_OK(cudaStreamSynchronize(st));
    return MSLAM_B200_OK;
}

// a shard the context owns: uploaded once, resident in HBM, queried by mslam_b200_db_query_loaded (the form a C++ caller
// without its own device allocations uses -- host/sharded_keyframe_db_b200.cpp)
int mslam_b200_db_load_shard(mslam_b200_ctx* ctx, const uint8_t* db, const int* k_counts, int n_local, int k_cap, int id_base)
{
    if(!ctx || n_local < 0 || (n_local > 0 && (!db || !k_counts || k_cap < 1)) || id_base < 0)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    cudaStream_t st = ctx->stream;
    MSLAM_CUDA_OK(cudaStreamSynchronize(st));
    const size_t db_b = ((size_t)n_local * k_cap * 32 + 255) / 256 * 256, need = db_b + (size_t)n_local * sizeof(int);
    if(ctx->shard_bytes < need)
    {
        cudaFree(ctx->d_shard);
        ctx->d_shard = nullptr, ctx->shard_bytes = 0;
        MSLAM_CUDA_OK(cudaMalloc(&ctx->d_shard, std::max<size_t>(need, 256)));
        ctx->shard_bytes = std::max<size_t>(need, 256);
    }
    if(n_local > 0)
    {
        MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->d_shard, db, (size_t)n_local * k_cap * 32, cudaMemcpyHostToDevice, st));
        MSLAM_CUDA_OK(cudaMemcpyAsync(ctx->d_shard + db_b, k_counts, (size_t)n_local * sizeof(int), cudaMemcpyHostToDevice, st));
        MSLAM_CUDA_OK(cudaStreamSynchronize(st));
    }
    ctx->shard_n = n_local, ctx->shard_cap = n_local > 0 ? k_cap : 1, ctx->shard_id_base = id_base;
    return MSLAM_B200_OK;
}

int mslam_b200_db_query_loaded(mslam_b200_ctx* ctx, const uint8_t* queries, const int* q_counts, int n_query_frames, int q_cap, int k,
                               int* out_score, int* out_id)
{
    if(!ctx)
        return MSLAM_B200_E_INVALID;
    const size_t db_b = ((size_t)ctx->shard_n * ctx->shard_cap * 32 + 255) / 256 * 256;
    return mslam_b200_db_query_sharded(ctx, queries, q_counts, n_query_frames, q_cap, ctx->shard_n ? ctx->d_shard : nullptr,
                                       ctx->shard_n ? reinterpret_cast<const int*>(ctx->d_shard + db_b) : nullptr, ctx->shard_n,
                                       ctx->shard_cap, ctx->shard_id_base, k, out_score, out_id);
}

// grow-only device scratch of the host convenience entry points (back-projection, RANSAC PnP): a per-call cudaMalloc /
// cudaFree pair costs more than the kernels they feed
static int host_scratch(mslam_b200_ctx* ctx, size_t bytes, uint8_t** out)
{
    if(ctx->pnp_bytes < bytes)
    {
        MSLAM_CUDA_OK(cudaStreamSynchronize(ctx->stream));
        cudaFree(ctx->d_pnp);
        ctx->d_pnp = nullptr, ctx->pnp_bytes = 0;
        MSLAM_CUDA_OK(cudaMalloc(&ctx->d_pnp, bytes));
        ctx->pnp_bytes = bytes;
    }
    *out = ctx->d_pnp;
    return MSLAM_B200_OK;
}

// ---- RGB-D association ------------------------------------------------------------------------------
static bool camera_ok(const mslam_b200_camera* cam)
{
    return cam && cam->fx != 0.0 && cam->fy != 0.0;
}

int mslam_b200_backproject_dev(mslam_b200_ctx* ctx, const mslam_b200_keypoint* d_kp, int kp_cap, const int* d_index,
                               const int* d_counts, int n_pairs, int cap, const uint16_t* d_depth, int width, int height,
                               const mslam_b200_camera* cam, double* d_xyz, uint8_t* d_valid)
{
    if(!ctx || !d_kp || !d_counts || !d_depth || !d_xyz || !d_valid || !camera_ok(cam) || kp_cap < 1 || n_pairs < 1 ||
       n_pairs > 65535 || cap < 1 || width < 1 || height < 1)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    launch_backproject(ctx->stream, d_kp, kp_cap, d_index, d_counts, n_pairs, cap, d_depth, width, height, *cam, d_xyz, d_valid,
                       &ctx->launches);
    MSLAM_CUDA_OK(cudaGetLastError());
    return MSLAM_B200_OK;
}

int mslam_b200_backproject(mslam_b200_ctx* ctx, const double* xy, int n, int stride_bytes, const uint16_t* depth, int width,
                           int height, const mslam_b200_camera* cam, double* out_xyz, uint8_t* out_valid)
{
    if(!ctx || n < 0 || !camera_ok(cam) || width < 1 || height < 1 || stride_bytes < 16 || (stride_bytes & 7))
        return MSLAM_B200_E_INVALID;
    if(n == 0)
        return MSLAM_B200_OK;
    if(!xy || !depth || !out_xyz || !out_valid)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    cudaStream_t st = ctx->stream;
    const size_t depth_bytes = (size_t)width * height * sizeof(uint16_t);
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const size_t o_xyz = up((size_t)n * 16), o_depth = o_xyz + up((size_t)n * 24), o_valid = o_depth + up(depth_bytes);
    uint8_t* d = nullptr;
    int rc = host_scratch(ctx, o_valid + up((size_t)n), &d);
    if(rc != MSLAM_B200_OK)
        return rc;
    double *d_xy = (double*)d, *d_xyz = (double*)(d + o_xyz);
    uint16_t* d_depth = (uint16_t*)(d + o_depth);
    uint8_t* d_valid = d + o_valid;
    cudaError_t e;
#define BPQ(expr)                                                                                                      \
    if(rc == MSLAM_B200_OK && (e = (expr)) != cudaSuccess)                                                             \
    rc = ctx_fail(ctx, e, #expr)
    BPQ(cudaMemcpy2DAsync(d_xy, 16, xy, (size_t)stride_bytes, 16, (size_t)n, cudaMemcpyHostToDevice, st));
    BPQ(cudaMemcpyAsync(d_depth, depth, depth_bytes, cudaMemcpyHostToDevice, st));
    if(rc == MSLAM_B200_OK)
        launch_backproject_xy(st, d_xy, n, d_depth, width, height, *cam, d_xyz, d_valid, &ctx->launches);
    BPQ(cudaGetLastError());
    BPQ(cudaMemcpyAsync(out_xyz, d_xyz, (size_t)n * 24, cudaMemcpyDeviceToHost, st));
    BPQ(cudaMemcpyAsync(out_valid, d_valid, (size_t)n, cudaMemcpyDeviceToHost, st));
    BPQ(cudaStreamSynchronize(st));
#undef BPQ
    return rc;
}

// ---- RANSAC PnP ----------------------------------------------------------------------------------------
void mslam_b200_pnp_default_params(mslam_b200_pnp_params* p)
{
    if(!p)
        return;
    memset(p, 0, sizeof(*p));
    p->struct_size = (int)sizeof(*p);
    p->iterations = 100;    // cv_ransac_pnp.cpp:57
    p->reproj_error = 5.0;
    p->confidence = 0.99;
    p->refine_iterations = 30;
    p->fx = p->fy = 1.0;
}

static bool pnp_params_ok(const mslam_b200_pnp_params* p)
{
    return p && p->struct_size == (int)sizeof(*p) && p->iterations >= 1 && p->iterations <= pnp_max_hypotheses() &&
           p->reproj_error > 0.0 && p->refine_iterations >= 0 && (float)p->fx != 0.f && (float)p->fy != 0.f;
}

int mslam_b200_pnp_ransac_batch_dev(mslam_b200_ctx* ctx, const float* d_obj, const float* d_img, const int* d_counts,
                                    int n_frames, int cap, const mslam_b200_pnp_params* params, double* d_rvec_tvec,
                                    double* d_pose, uint8_t* d_inliers, int* d_info)
{
    if(!ctx || !d_obj || !d_img || !d_counts || !d_rvec_tvec || !d_pose || !d_inliers || !d_info || n_frames < 1 || cap < 1 ||
       !pnp_params_ok(params))
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    launch_pnp_ransac(ctx->stream, d_obj, d_img, d_counts, n_frames, cap, params->frame_base, (double)(float)params->fx,
                      (double)(float)params->fy, (double)(float)params->cx, (double)(float)params->cy, params->iterations,
                      params->reproj_error * params->reproj_error, params->seed, params->refine_iterations, d_rvec_tvec,
                      d_pose, d_inliers, d_info, &ctx->launches);
    MSLAM_CUDA_OK(cudaGetLastError());
    return MSLAM_B200_OK;
}

int mslam_b200_pnp_ransac_batch(mslam_b200_ctx* ctx, const float* obj, const float* img, const int* counts, int n_frames,
                                int cap, const mslam_b200_pnp_params* params, double* rvec_tvec, double* pose,
                                uint8_t* inliers, int* info)
{
    if(!ctx || !obj || !img || !counts || !rvec_tvec || !pose || !inliers || !info || n_frames < 1 || cap < 1 ||
       !pnp_params_ok(params))
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    cudaStream_t st = ctx->stream;
    const size_t F = (size_t)n_frames;
    const size_t b_obj = F * cap * 3 * sizeof(float), b_img = F * cap * 2 * sizeof(float), b_cnt = F * sizeof(int);
    const size_t b_rt = F * 6 * sizeof(double), b_pose = F * 7 * sizeof(double), b_inl = F * cap, b_info = F * 4 * sizeof(int);
    // one allocation, 256-byte aligned sections
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const size_t o_img = up(b_obj), o_cnt = o_img + up(b_img), o_rt = o_cnt + up(b_cnt), o_pose = o_rt + up(b_rt),
                 o_inl = o_pose + up(b_pose), o_info = o_inl + up(b_inl), total = o_info + up(b_info);
    uint8_t* d = nullptr;
    int rc = host_scratch(ctx, total, &d);
    if(rc != MSLAM_B200_OK)
        return rc;
    cudaError_t e;
#define PNQ(expr)                                                                                                      \
    if(rc == MSLAM_B200_OK && (e = (expr)) != cudaSuccess)                                                             \
    rc = ctx_fail(ctx, e, #expr)
    PNQ(cudaMemcpyAsync(d, obj, b_obj, cudaMemcpyHostToDevice, st));
    PNQ(cudaMemcpyAsync(d + o_img, img, b_img, cudaMemcpyHostToDevice, st));
    PNQ(cudaMemcpyAsync(d + o_cnt, counts, b_cnt, cudaMemcpyHostToDevice, st));
    if(rc == MSLAM_B200_OK)
        rc = mslam_b200_pnp_ransac_batch_dev(ctx, (const float*)d, (const float*)(d + o_img), (const int*)(d + o_cnt), n_frames,
                                             cap, params, (double*)(d + o_rt), (double*)(d + o_pose), d + o_inl,
                                             (int*)(d + o_info));
    // the four outputs are adjacent in the scratch block: one device->host copy into a pinned staging block, then plain
    // host copies (four small DMA transfers cost more than the 72 KB they move)
    const size_t b_out = total - o_rt;
    if(rc == MSLAM_B200_OK && ctx->h_pnp_bytes < b_out)
    {
        if(ctx->h_pnp)
            cudaFreeHost(ctx->h_pnp);
        ctx->h_pnp = nullptr, ctx->h_pnp_bytes = 0;
        PNQ(cudaHostAlloc((void**)&ctx->h_pnp, b_out, cudaHostAllocDefault));
        if(rc == MSLAM_B200_OK)
            ctx->h_pnp_bytes = b_out;
    }
    PNQ(cudaMemcpyAsync(ctx->h_pnp, d + o_rt, b_out, cudaMemcpyDeviceToHost, st));
    PNQ(cudaStreamSynchronize(st));
#undef PNQ
    if(rc == MSLAM_B200_OK)
    {
        const uint8_t* h = ctx->h_pnp;
        memcpy(rvec_tvec, h, b_rt);
        memcpy(pose, h + (o_pose - o_rt), b_pose);
        memcpy(inliers, h + (o_inl - o_rt), b_inl);
        memcpy(info, h + (o_info - o_rt), b_info);
    }
    return rc;
}

int mslam_b200_enable_stage_timing(mslam_b200_ctx* ctx, int enable)
{
    if(!ctx)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    MSLAM_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    if(enable && ctx->ev_pool.empty())
    {
        const size_t pairs = 8192;
        ctx->ev_pool.resize(2 * pairs);
        for(size_t i = 0; i < 2 * pairs; ++i)
            MSLAM_CUDA_OK(cudaEventCreate(&ctx->ev_pool[i]));
        ctx->ev_stage.assign(pairs, 0);
        ctx->ev_launches.assign(pairs, 0);
    }
    ctx->timing = enable != 0;
    ctx->ev_used = 0;
    for(int i = 0; i < MSLAM_B200_T_COUNT; ++i)
        ctx->t_ms[i] = 0, ctx->t_n[i] = 0, ctx->t_l[i] = 0;
    return MSLAM_B200_OK;
}

int mslam_b200_stage_times(mslam_b200_ctx* ctx, double* total_ms, long long* intervals, long long* launches)
{
    if(!ctx || !total_ms)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    MSLAM_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    for(size_t i = 0; i < ctx->ev_used; ++i)
    {
        float ms = 0.f;
        MSLAM_CUDA_OK(cudaEventElapsedTime(&ms, ctx->ev_pool[2 * i], ctx->ev_pool[2 * i + 1]));
        const int s = ctx->ev_stage[i];
        ctx->t_ms[s] += ms;
        ctx->t_n[s] += 1;
        ctx->t_l[s] += ctx->ev_launches[i];
    }
    ctx->ev_used = 0;
    for(int i = 0; i < MSLAM_B200_T_COUNT; ++i)
    {
        total_ms[i] = ctx->t_ms[i];
        if(intervals)
            intervals[i] = ctx->t_n[i];
        if(launches)
            launches[i] = ctx->t_l[i];
        ctx->t_ms[i] = 0, ctx->t_n[i] = 0, ctx->t_l[i] = 0;
    }
    return MSLAM_B200_OK;
}

int mslam_b200_measure_int_peaks(mslam_b200_ctx* ctx, double* out4 /* 5 values */)
{
    if(!ctx || !out4)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    MSLAM_CUDA_OK(measure_int_peaks(ctx->stream, ctx->sm_count, out4, &ctx->launches));
    return MSLAM_B200_OK;
}

int mslam_b200_measure_tensor_peak(mslam_b200_ctx* ctx, double* out4)
{
    if(!ctx || !out4)
        return MSLAM_B200_E_INVALID;
    MSLAM_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    MSLAM_CUDA_OK(measure_tensor_peak(ctx->stream, ctx->sm_count, out4, &ctx->launches));
    return MSLAM_B200_OK;
}

} // extern "C"
