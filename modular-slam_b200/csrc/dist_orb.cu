// dist_orb.cu -- the second detector of the reference, DistributedOrbOpenCvDetector, built B200-first.
//
// Reference: src/lib/modular_slam/distributed_cv_feature.cpp (a stella_vslam-derived extractor).  What it computes:
//   toGrayScale (frame.cpp:6-27) -> chained cv::resize INTER_LINEAR pyramid (:830-841) -> cv::FAST with non-max suppression
//   on every 64-px cell (+6 px overlap) inside a 19-px border, threshold 20, then 7 if the cell came back empty (:843-939)
//   -> quad-tree distribution of the level's keypoints, one keypoint per final node, maximum response (:941-1106)
//   -> intensity-centroid angle on the unblurred level (:543-570) -> cv::GaussianBlur 7x7 sigma 2 on u8 (fixed point,
//   :797-798) -> 256 rotated BRIEF tests with polynomial cos/sin (:456-503, :572-629) -> coordinates * scale_l (:1117-1130).
//
// How it is mapped to the GPU (one launch per stage covers every frame of the batch and every level):
//   k_dist_resize   level l-1 -> l, 11-bit fixed-point bilinear taps from per-level tables (tables built on the host in
//                   IEEE double / float exactly as OpenCV builds them); 4 output pixels per thread
//   k_dist_fast     one CTA per (frame, level, cell): the cell's 70x70 tile in shared memory, the exact FAST-9 classification
//                   of the 64x64 cell in byte-SIMD (fast_simd.cuh, shared with k_fast), a queue of the corners, their exact
//                   cornerScore two at a time in one packed u16x2 min/max network, strict 3x3 NMS inside the cell, threshold
//                   fallback, one atomic per cell to append the survivors to the (frame, level) list
//   k_dist_tree     one CTA per (frame, level).  The reference's std::list surgery is replaced by an order-key
//                   formulation: every keypoint's path through the tree is a pure function of the node geometry, a
//                   pass of the reference divides exactly the divisible nodes of one depth, and the final list order
//                   is (depth descending, key ascending) with key_{d+1} = (nodes_d - 1 - key_d) * 4 + (3 - child)
//                   (children are pushed to the FRONT of the list in order 0..3, which reverses the order every depth).
//                   Per-depth occupancy counters in shared memory, the loop's "list size unchanged" exit reproduced
//                   (created == divided), per-leaf argmax by atomicMax on (score, arrival order), block scan for the
//                   output rank.
//   k_dist_gauss    64x32 tiles, 8.8 horizontal / 16.16 vertical fixed point, REFLECT_101
//   k_dist_describe one warp per keypoint: moments as IDP.4A of aligned patch words against packed disc weights, cv::fastAtan2,
//                   polynomial cos/sin in separately rounded f32, lane = descriptor byte (float pattern table, rounding
//                   by the 1.5 * 2^23 magic add: nothing on the XU pipe)
// All integer stages are exact by construction; every float operation is an explicit round-to-nearest intrinsic
// (the library is also compiled with -fmad=false).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <new>
#include <set>
#include <utility>
#include <vector>

#include "kernels.h"
#include "fast_simd.cuh"

using namespace mslam_b200;

namespace
{
constexpr int kBorder = 19;   // orb_patch_radius_ (:695)
constexpr int kCell = 64;     // cell_size (:854)
constexpr int kFastOrg = 22;  // first pixel cv::FAST can report in a cell that starts at the border: 19 + 3
constexpr int kMaxInit = 64;  // initial tree nodes supported (round(aspect ratio))
constexpr int kTreeThreads = 1024;
constexpr int kTreeSmemBudget = 200 * 1024;
constexpr unsigned kFlag = 0x80000000u;
constexpr int kDistChunks = 4; // host detect_batch: chunks in flight through copy-in / kernels / copy-out

struct DistLevel
{
    int w, h, pitch;
    int ncx, ncy, cell_base;  // FAST cells
    int gtx, gty, gtile_base; // blur tiles
    int cand_cap, sel_cap;
    int sel_off;              // elements, within one frame's selection block
    int ninit_x, ninit_y;
    int depth;                // deepest depth a node can have (all nodes there stop by area)
    int tree_slots;           // sum over depths 0..depth of ninit * 4^d
    int tree_in_smem;
    int xtab_off, ytab_off;
    int qtab_off, nquads;     // k_dist_resize_quads' table (uint4 pairs); nquads = 0: the level uses k_dist_resize
    long long img_off, img_frame_stride;
    long long cand_off;       // elements, within one frame's candidate block
    long long tree_off;       // words, within one frame's global tree scratch (levels that do not fit in smem)
    double delta_x, delta_y;
    float scale, size;
};

struct DistParams
{
    int nlevels, n_frames, W, H;
    int ini_thr, min_thr;
    int one, minus_one;       // +1 / -1 the compiler cannot see: byte-lane additions of k_dist_fast as IMADs on the FMA pipe
    unsigned min_size;
    int kp_cap;               // per-frame keypoint capacity of the handle
    int total_cells, total_gtiles;
    int sel_frame_stride;
    int frame0;               // first frame (absolute index in the handle's per-frame buffers) a launch covers
    long long cand_frame_stride, tree_frame_stride;
    DistLevel lv[kMaxLevels];
};

struct DistCounters
{
    unsigned n_fast, n_tree, passes, err;
};


// =================================================================================================
// resize: cv::resize INTER_LINEAR, CV_8UC1 (OpenCV resize.cpp HResizeLinear / VResizeLinear, 11-bit coefficients)
// =================================================================================================
struct XTap
{
    int ofs;
    unsigned a; // a0 | a1 << 16
};
struct YTap
{
    int sy0, sy1; // clamped source rows
    unsigned b;   // b0 | b1 << 16
    int pad;
};

__global__ void __launch_bounds__(256) k_dist_resize(const uint8_t* __restrict__ src, int spitch, long long sfs, int sw,
                                                     uint8_t* __restrict__ dst, int dpitch, long long dfs, int dw, int dh,
                                                     const XTap* __restrict__ xt, const YTap* __restrict__ yt)
{
    const int x4 = (blockIdx.x * 64 + (threadIdx.x & 63)) * 4;
    const int y = blockIdx.y * 4 + (threadIdx.x >> 6);
    if(x4 >= dw || y >= dh)
        return;
    const YTap Y = yt[y];
    const uint8_t* s0 = src + blockIdx.z * sfs + (long long)Y.sy0 * spitch;
    const uint8_t* s1 = src + blockIdx.z * sfs + (long long)Y.sy1 * spitch;
    const int b0 = Y.b & 0xffff, b1 = Y.b >> 16;
    unsigned out = 0;
#pragma unroll
    for(int i = 0; i < 4; ++i)
    {
        const int x = x4 + i < dw ? x4 + i : dw - 1;
        const XTap X = xt[x];
        const int a0 = X.a & 0xffff, a1 = X.a >> 16;
        const int sx1 = X.ofs + 1 < sw ? X.ofs + 1 : X.ofs;
        const int r0 = s0[X.ofs] * a0 + s0[sx1] * a1;
        const int r1 = s1[X.ofs] * a0 + s1[sx1] * a1;
        const unsigned o = (unsigned)((((b0 * (r0 >> 4)) >> 16) + ((b1 * (r1 >> 4)) >> 16) + 2) >> 2);
        out |= (o & 0xff) << (8 * i);
    }
    *reinterpret_cast<unsigned*>(dst + blockIdx.z * dfs + (long long)y * dpitch + x4) = out; // pitch covers the tail
}

// The same arithmetic, four output pixels per lane from one table entry: entry q of a level covers columns 4q .. 4q + 3 --
// {first source byte of an 8-byte window, PRMT selectors that pull (left, right) taps of pixels 0, 1 and of pixels 2, 3 out of
// the window | the four (a0 | a1 << 16) weight pairs}.  Per source row 3 aligned word loads, 2 funnel shifts, 2 PRMT and
// 4 IDP.2A (both taps of a pixel in one instruction) instead of four table loads, eight byte loads and eight IMADs; used
// for every level whose quads fit the window (scale factors up to 2), k_dist_resize above stays as the fallback.
__global__ void __launch_bounds__(256) k_dist_resize_quads(const uint8_t* __restrict__ src, int spitch, long long sfs,
                                                           uint8_t* __restrict__ dst, int dpitch, long long dfs, int nquads, int dh,
                                                           const uint4* __restrict__ qt, const YTap* __restrict__ yt)
{
    const int q = blockIdx.x * 64 + (threadIdx.x & 63);
    const int y = blockIdx.y * 4 + (threadIdx.x >> 6);
    if(q >= nquads || y >= dh)
        return;
    const uint4 e0 = __ldg(qt + 2 * q), wts = __ldg(qt + 2 * q + 1);
    const YTap Y = yt[y];
    const int wi = (int)(e0.x >> 2), sh = (int)(e0.x & 3u) * 8;
    const unsigned selA = e0.y & 0xffffu, selB = e0.y >> 16;
    const unsigned* s0 = reinterpret_cast<const unsigned*>(src + blockIdx.z * sfs + (long long)Y.sy0 * spitch) + wi;
    const unsigned* s1 = reinterpret_cast<const unsigned*>(src + blockIdx.z * sfs + (long long)Y.sy1 * spitch) + wi;
    const unsigned a0 = __ldg(s0), a1 = __ldg(s0 + 1), a2 = __ldg(s0 + 2), c0 = __ldg(s1), c1 = __ldg(s1 + 1), c2 = __ldg(s1 + 2);
    unsigned r0[4], r1[4];
    {
        const unsigned S0 = __funnelshift_r(a0, a1, sh), S1 = __funnelshift_r(a1, a2, sh);
        const unsigned pA = __byte_perm(S0, S1, selA), pB = __byte_perm(S0, S1, selB);
        r0[0] = __dp2a_lo(wts.x, pA, 0u), r0[1] = __dp2a_hi(wts.y, pA, 0u), r0[2] = __dp2a_lo(wts.z, pB, 0u), r0[3] = __dp2a_hi(wts.w, pB, 0u);
    }
    {
        const unsigned S0 = __funnelshift_r(c0, c1, sh), S1 = __funnelshift_r(c1, c2, sh);
        const unsigned pA = __byte_perm(S0, S1, selA), pB = __byte_perm(S0, S1, selB);
        r1[0] = __dp2a_lo(wts.x, pA, 0u), r1[1] = __dp2a_hi(wts.y, pA, 0u), r1[2] = __dp2a_lo(wts.z, pB, 0u), r1[3] = __dp2a_hi(wts.w, pB, 0u);
    }
    const unsigned b0 = Y.b & 0xffffu, b1 = Y.b >> 16;
    unsigned out = 0;
#pragma unroll
    for(int i = 0; i < 4; ++i)
    {
        const unsigned o = (((b0 * (r0[i] >> 4)) >> 16) + ((b1 * (r1[i] >> 4)) >> 16) + 2u) >> 2;
        out |= (o & 0xffu) << (8 * i);
    }
    *reinterpret_cast<unsigned*>(dst + blockIdx.z * dfs + (long long)y * dpitch + 4 * q) = out; // pitch covers the tail
}

// =================================================================================================
// FAST per cell (cv::FAST(cell, thr, true), :915-922)
// =================================================================================================
constexpr int FT_PITCH = 80; // tile row pitch: 19 aligned words cover the 70 columns of a cell at any alignment
constexpr int FS_PITCH = 68; // score row pitch (66 used: 64 + zero ring)

__device__ __forceinline__ int level_of_cell(const DistParams& P, int idx)
{
    int l = 0;
    while(l + 1 < P.nlevels && idx >= P.lv[l + 1].cell_base)
        ++l;
    return l;
}
__device__ __forceinline__ int level_of_gtile(const DistParams& P, int idx)
{
    int l = 0;
    while(l + 1 < P.nlevels && idx >= P.lv[l + 1].gtile_base)
        ++l;
    return l;
}

__global__ void __launch_bounds__(256) k_dist_fast(const __grid_constant__ DistParams P, const uint8_t* __restrict__ pyr,
                                                   unsigned* __restrict__ cand, DistCounters* __restrict__ ctr)
{
    __shared__ __align__(16) uint8_t tile[70 * FT_PITCH + 16];
    __shared__ __align__(16) uint8_t sc[66 * FS_PITCH];
    __shared__ unsigned short q[4096];
    __shared__ unsigned surv[1024];
    __shared__ int s_nq, s_ns, s_base;

    const int tid = threadIdx.x, lane = tid & 31;
    const int l = level_of_cell(P, blockIdx.x);
    const DistLevel& L = P.lv[l];
    const int frame = P.frame0 + blockIdx.y;
    const int c = blockIdx.x - L.cell_base;
    const int ci = c / L.ncx, cj = c - ci * L.ncx;
    const int x0 = kFastOrg + kCell * cj, y0 = kFastOrg + kCell * ci;
    const int cw = min(kCell, L.w - kFastOrg - x0), ch = min(kCell, L.h - kFastOrg - y0);
    // the tile starts at column x0 - 3 = 19 + 64 cj; rows are fetched as aligned words from column 16 + 64 cj, so the
    // tile's column 0 sits at byte 3 of a shared-memory row (T below)
    const uint8_t* img = pyr + L.img_off + frame * L.img_frame_stride + (long long)(y0 - 3) * L.pitch + (x0 - 6);
    // ... and is stored shifted down by two bytes, which puts cell pixel lx at byte lx + 4 of a tile row: the four centres of
    // a byte-SIMD group are one aligned word (T below points at pixel (-3, -3) of the cell, byte 1 of the tile)
    const int th = ch + 6, nwords = (cw + 12) >> 2;
    for(int r = tid >> 5; r < th; r += 8)
        if(lane < nwords)
        {
            const unsigned* g = reinterpret_cast<const unsigned*>(img + (long long)r * L.pitch) + lane;
            reinterpret_cast<unsigned*>(tile + r * FT_PITCH)[lane] = __funnelshift_r(__ldg(g), __ldg(g + 1), 16);
        }
    const uint8_t* T = tile + 1;
    int thr = P.ini_thr;
    int ns = 0;
    for(int attempt = 0; attempt < 2; ++attempt)
    {
        for(int i = tid; i < 66 * FS_PITCH / 4; i += 256)
            reinterpret_cast<unsigned*>(sc)[i] = 0;
        if(tid == 0)
            s_nq = 0, s_ns = 0;
        __syncthreads();
        // phase A: exact FAST-9 classification, byte-SIMD (fast_simd.cuh, as k_fast): a thread owns 4 adjacent centres x 4 rows
        // -- 16 word columns x 16 row groups = the 256 threads of the CTA cover the 64 x 64 cell in one pass.  Only the corners
        // go to the queue (the earlier compass pre-test let 55 % of the bench's pixels through to the 94-operation exact
        // score; a fifth of them are corners).
        {
            const int wcol = tid & 15, rb = (tid >> 4) * 4; // cell columns 4 * wcol .. + 3, rows rb .. rb + 3
            unsigned flags16 = 0, bright16 = 0;
            const unsigned t4 = (unsigned)thr * 0x01010101u;
            const unsigned one = (unsigned)P.one, mone = (unsigned)P.minus_one;
            const unsigned* wbase = reinterpret_cast<const unsigned*>(tile + rb * FT_PITCH) + wcol;
            unsigned wm[10], w0[10], wp[10]; // tile rows rb .. rb + 9, words wcol, wcol + 1 (the centres), wcol + 2
#pragma unroll
            for(int r = 0; r < 10; ++r)
            {
                wm[r] = wbase[r * (FT_PITCH / 4)];
                w0[r] = wbase[r * (FT_PITCH / 4) + 1];
                wp[r] = wbase[r * (FT_PITCH / 4) + 2];
            }
#pragma unroll
            for(int row = 0; row < 4; ++row)
            {
                const unsigned C = w0[row + 3];
                const unsigned H = __vaddus4(C, t4), Lo = __vsubus4(C, t4);
                const unsigned nH7 = ~H & 0x7f7f7f7fu;
                const unsigned K = (Lo & 0x7f7f7f7fu) + 0x7f7f7f7fu;
                unsigned mb[16], md[16];
                constexpr int DX[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
                constexpr int DY[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
#pragma unroll
                for(int k = 0; k < 16; ++k)
                {
                    const int r = row + 3 + DY[k];
                    const unsigned R = ring4(wm[r], w0[r], wp[r], DX[k]);
                    const unsigned r7 = R & 0x7f7f7f7fu;
                    mb[k] = lop3_maj_a_nb_c(R, H, r7 * one + nH7); // bit 7: R > centre + thr
                    md[k] = lop3_maj_a_nb_c(Lo, R, r7 * mone + K); // bit 7: R < centre - thr
                }
                const unsigned fb = any_run9(mb) & 0x80808080u, f = fb | (any_run9(md) & 0x80808080u);
                flags16 |= (((f >> 7) * 0x00204081u) >> 21 & 0xfu) << (4 * row);
                bright16 |= (((fb >> 7) * 0x00204081u) >> 21 & 0xfu) << (4 * row);
            }
            // positions inside the cell only (cells at the right / bottom edge are narrower)
            {
                const int nx = min(4, max(0, cw - 4 * wcol)), ny = min(4, max(0, ch - rb));
                flags16 &= (((1u << nx) - 1u) * 0x1111u) & ((1u << (4 * ny)) - 1u);
            }
            const int cnt = __popc(flags16);
            int incl = cnt;
#pragma unroll
            for(int o = 1; o < 32; o <<= 1)
            {
                const int t = __shfl_up_sync(0xffffffffu, incl, o);
                if(lane >= o)
                    incl += t;
            }
            int wq = 0;
            if(lane == 31)
                wq = atomicAdd(&s_nq, incl);
            wq = __shfl_sync(0xffffffffu, wq, 31);
            unsigned short* qp = q + (wq + incl - cnt);
            const unsigned ebase = (unsigned)(rb << 8 | 4 * wcol);
            const unsigned b2 = bright16 << 15;
            unsigned f = flags16;
            while(f)
            {
                unsigned bit;
                asm("bfind.u32 %0, %1;" : "=r"(bit) : "r"(f));
                f ^= 1u << bit;
                // entry: row (rb + bit / 4) << 8 | column 4 * wcol + bit % 4 | bright << 15
                *qp++ = (unsigned short)(ebase + ((bit * 65u) & 0x303u) + ((b2 >> bit) & 0x8000u));
            }
        }
        __syncthreads();
        // phase B: exact score of the queued corners, two per thread in the lanes of one packed min / max network
        const int nq = s_nq;
        for(int i = 2 * tid; i < nq; i += 512)
        {
            const unsigned ea = q[i], eb = q[min(i + 1, nq - 1)];
            const int ya = (ea >> 8) & 127, xa = ea & 255, yb = (eb >> 8) & 127, xb = eb & 255;
            const unsigned scs = corner_score2<FT_PITCH>(T + (ya + 3) * FT_PITCH + xa + 3, T + (yb + 3) * FT_PITCH + xb + 3, ea & 0x8000u,
                                                        eb & 0x8000u);
            sc[(ya + 1) * FS_PITCH + xa + 1] = (uint8_t)(scs & 0xffu);
            sc[(yb + 1) * FS_PITCH + xb + 1] = (uint8_t)(scs >> 16);
        }
        __syncthreads();
        // phase C: strict 3x3 non-max suppression inside the cell
        for(int i = tid; i < nq; i += 256)
        {
            const int ly = (q[i] >> 8) & 127, lx = q[i] & 255;
            const uint8_t* s = sc + (ly + 1) * FS_PITCH + lx + 1;
            const int v = s[0];
            if(v && v > s[-1] && v > s[1] && v > s[-FS_PITCH - 1] && v > s[-FS_PITCH] && v > s[-FS_PITCH + 1] &&
               v > s[FS_PITCH - 1] && v > s[FS_PITCH] && v > s[FS_PITCH + 1])
            {
                const int k = atomicAdd(&s_ns, 1);
                // coordinates relative to the border origin, as keypts_to_distribute holds them (:928-931)
                surv[k] = (unsigned)(y0 + ly - kBorder) << 20 | (unsigned)(x0 + lx - kBorder) << 8 | (unsigned)v;
            }
        }
        __syncthreads();
        ns = s_ns;
        if(ns > 0 || thr == P.min_thr)
            break;
        thr = P.min_thr; // keypts_in_cell.empty() -> second cv::FAST with min_fast_thr_ (:919-923)
        __syncthreads();
    }
    if(ns == 0)
        return;
    if(tid == 0)
        s_base = (int)atomicAdd(&ctr[frame * P.nlevels + l].n_fast, (unsigned)ns);
    __syncthreads();
    unsigned* out = cand + frame * P.cand_frame_stride + L.cand_off;
    for(int i = tid; i < ns; i += 256)
        if(s_base + i < L.cand_cap)
            out[s_base + i] = surv[i];
}

// =================================================================================================
// quad-tree distribution (:941-1106), order-key formulation -- see the file header
// =================================================================================================
struct Walk
{
    int bx, by, ex, ey;
    int depth, key, off, nslots;
};

__device__ __forceinline__ bool area_stops(int bx, int by, int ex, int ey, float s, unsigned min_size)
{
    const unsigned area = (unsigned)((ex - bx) * (ey - by)); // orb_extractor_node::size()
    return __fmul_rn(__fmul_rn((float)area, s), s) <= (float)min_size;
}

// place a keypoint in its initial node (initialize_nodes, :1032-1041): ix = (unsigned)(pt.x / delta_x).  Only one axis
// ever has more than one initial node; its quotients come from a per-CTA table filled once with the same double
// division (init_tab[c] = (unsigned)((double)(float)c / delta)), and the initial nodes' bounds ((int)(delta * i), double
// products in the reference) from a second per-CTA table, so the walk itself is integer-only -- it runs once per keypoint and
// pass, and four FP64 multiplies + conversions per call were a tenth of the kernel's instructions on a pipe B200 barely has.
__device__ __forceinline__ void walk_init(const DistLevel& L, const uint8_t* __restrict__ init_tab, const int4* __restrict__ init_box,
                                          int x, int y, Walk& w)
{
    const int ninit = L.ninit_x * L.ninit_y;
    unsigned ix = L.ninit_x > 1 ? init_tab[x] : 0u, iy = L.ninit_y > 1 ? init_tab[y] : 0u;
    unsigned i0 = ix + iy * (unsigned)L.ninit_x;
    if(i0 >= (unsigned)ninit)
        i0 = ninit - 1;
    const int4 bx = init_box[i0];
    w.bx = bx.x, w.by = bx.y, w.ex = bx.z, w.ey = bx.w;
    w.depth = 0, w.key = (int)i0, w.off = 0, w.nslots = ninit;
}

// one divide_node step (:306-355): returns the child index and moves w into the child
__device__ __forceinline__ void walk_down(int x, int y, Walk& w)
{
    const int cx = w.bx + ((w.ex - w.bx + 1) >> 1), cy = w.by + ((w.ey - w.by + 1) >> 1); // cvCeil(len / 2.0)
    int ci = 0;
    if(cx <= x)
        ci += 1, w.bx = cx;
    else
        w.ex = cx;
    if(cy <= y)
        ci += 2, w.by = cy;
    else
        w.ey = cy;
    w.key = (w.nslots - 1 - w.key) * 4 + (3 - ci);
    w.off += w.nslots;
    w.nslots *= 4;
    w.depth += 1;
}

// walk from the root towards depth `target`; stops early at a node that is final (one keypoint, or small area).
// Returns true when the walk arrived at `target` through divided nodes only.
__device__ __forceinline__ bool walk_to(const DistLevel& L, const uint8_t* __restrict__ init_tab, const int4* __restrict__ init_box,
                                        const unsigned* cnt, int x, int y, int target, Walk& w)
{
    walk_init(L, init_tab, init_box, x, y, w);
    while(w.depth < target)
    {
        const unsigned c = cnt[w.off + w.key];
        if((c & ~kFlag) == 1u || (c & kFlag))
            return false;
        walk_down(x, y, w);
    }
    return true;
}

__global__ void __launch_bounds__(kTreeThreads) k_dist_tree(const __grid_constant__ DistParams P, const unsigned* __restrict__ cand,
                                                           DistCounters* __restrict__ ctr, unsigned* __restrict__ sel,
                                                           unsigned* __restrict__ gtree)
{
    extern __shared__ unsigned s_dyn[];
    __shared__ unsigned s_created, s_divided;
    __shared__ unsigned s_warp[32];
    __shared__ uint8_t s_init[4096]; // initial-node index of a coordinate along the axis that has several initial nodes
    __shared__ int4 s_box[kMaxInit];  // bounds of the initial nodes: (int)(delta_x * ix), (int)(delta_y * iy), ... (ix + 1), ... (iy + 1)
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int l = blockIdx.y, frame = P.frame0 + blockIdx.x; // level-major launch order: the big levels start first
    const DistLevel& L = P.lv[l];
    DistCounters* C = ctr + frame * P.nlevels + l;
    const int n = (int)min(C->n_fast, (unsigned)L.cand_cap);
    if(n == 0)
        return; // n_tree and passes stay 0
    if(C->n_fast > (unsigned)L.cand_cap && tid == 0)
        C->err = 1;
    unsigned* cnt = L.tree_in_smem ? s_dyn : gtree + frame * P.tree_frame_stride + L.tree_off;
    unsigned* best = cnt + L.tree_slots;
    for(int i = tid; i < 2 * L.tree_slots; i += kTreeThreads)
        cnt[i] = 0;
    if(tid < L.ninit_x * L.ninit_y)
    {
        const unsigned ix = (unsigned)tid % (unsigned)L.ninit_x, iy = (unsigned)tid / (unsigned)L.ninit_x;
        s_box[tid] = make_int4((int)(L.delta_x * ix), (int)(L.delta_y * iy), (int)(L.delta_x * (ix + 1)), (int)(L.delta_y * (iy + 1)));
    }
    if(L.ninit_x * L.ninit_y > 1)
    {
        const double delta = L.ninit_x > 1 ? L.delta_x : L.delta_y;
        for(int i = tid; i < 4096; i += kTreeThreads)
            s_init[i] = (uint8_t)min(255u, (unsigned)((double)(float)i / delta));
    }
    __syncthreads();
    const unsigned* list = cand + frame * P.cand_frame_stride + L.cand_off;
    const int ninit = L.ninit_x * L.ninit_y;
    const float s = L.scale;

    // depth 0: occupancy of the initial nodes
    for(int k = tid; k < n; k += kTreeThreads)
    {
        const unsigned r = list[k];
        const int x = (r >> 8) & 0xfff, y = r >> 20;
        Walk w;
        walk_init(L, s_init, s_box, x, y, w);
        const unsigned old = atomicAdd(&cnt[w.key], 1u);
        if(old == 0 && area_stops(w.bx, w.by, w.ex, w.ey, s, P.min_size))
            atomicOr(&cnt[w.key], kFlag);
    }
    __syncthreads();

    // pass p divides every divisible node of depth p (the reference's while(true) body, :954-976)
    int t_end = L.depth, passes = 0;
    bool broke = false;
    int off_p = 0, nslots_p = ninit;
    for(int p = 0; p < L.depth; ++p)
    {
        if(tid == 0)
            s_created = 0, s_divided = 0;
        __syncthreads();
        unsigned created = 0, divided = 0;
        for(int k = tid; k < n; k += kTreeThreads)
        {
            const unsigned r = list[k];
            const int x = (r >> 8) & 0xfff, y = r >> 20;
            Walk w;
            if(!walk_to(L, s_init, s_box, cnt, x, y, p, w))
                continue;
            const unsigned c = cnt[w.off + w.key];
            if((c & ~kFlag) == 1u || (c & kFlag))
                continue;
            walk_down(x, y, w);
            const unsigned old = atomicAdd(&cnt[w.off + w.key], 1u);
            if((old & ~kFlag) == 0)
            {
                ++created;
                if(area_stops(w.bx, w.by, w.ex, w.ey, s, P.min_size))
                    atomicOr(&cnt[w.off + w.key], kFlag);
            }
        }
        for(int i = tid; i < nslots_p; i += kTreeThreads)
        {
            const unsigned c = cnt[off_p + i];
            divided += (c & ~kFlag) >= 2u && !(c & kFlag);
        }
        created = __reduce_add_sync(0xffffffffu, created);
        divided = __reduce_add_sync(0xffffffffu, divided);
        if(lane == 0 && (created | divided))
        {
            atomicAdd(&s_created, created);
            atomicAdd(&s_divided, divided);
        }
        __syncthreads();
        const unsigned tc = s_created, td = s_divided;
        ++passes;
        off_p += nslots_p;
        nslots_p *= 4;
        if(tc == td) // nodes.size() == prev_size (:972): also when every divided node had exactly one non-empty child
        {
            t_end = p + 1;
            broke = true;
            break;
        }
        __syncthreads();
    }
    if(!broke)
        ++passes; // every pass changed the list: the reference runs one more, which finds every node stopped by area
    __syncthreads();

    // final nodes: per node the keypoint of maximum response, first in arrival order on ties (:1079-1106).
    // arrival order = cell-major raster (cells rows x cols, raster inside the cell), 24 bits.
    for(int k = tid; k < n; k += kTreeThreads)
    {
        const unsigned r = list[k];
        const int x = (r >> 8) & 0xfff, y = r >> 20;
        Walk w;
        walk_to(L, s_init, s_box, cnt, x, y, t_end, w);
        const int cj = (x - 3) >> 6, ci = (y - 3) >> 6;
        const unsigned ord = (unsigned)(ci * L.ncx + cj) << 12 | (unsigned)((y - 3) & 63) << 6 | (unsigned)((x - 3) & 63);
        atomicMax(&best[w.off + w.key], (r & 0xffu) << 24 | (0xffffffu - ord));
    }
    __syncthreads();

    // final list order: depth t_end first, then the final nodes of depth t_end - 1, ... , depth 0; key ascending
    int total_seq = 0;
    {
        int ns = ninit;
        for(int d = 0; d <= t_end; ++d, ns *= 4)
            total_seq += ns;
    }
    const int per = (total_seq + kTreeThreads - 1) / kTreeThreads;
    const int q0 = min(tid * per, total_seq), q1 = min(q0 + per, total_seq);
    // depth / offset of sequence position q0
    int d = t_end, ns_d = ninit, off_d = 0, seq_d = 0;
    for(int e = 0; e < t_end; ++e)
        off_d += ns_d, ns_d *= 4;
    while(d > 0 && q0 >= seq_d + ns_d)
    {
        seq_d += ns_d;
        ns_d >>= 2;
        off_d -= ns_d;
        --d;
    }
    int mine = 0;
    {
        int dd = d, nsd = ns_d, offd = off_d, seqd = seq_d;
        for(int qq = q0; qq < q1; ++qq)
        {
            while(qq >= seqd + nsd)
                seqd += nsd, nsd >>= 2, offd -= nsd, --dd;
            const unsigned c = cnt[offd + (qq - seqd)];
            const unsigned m = c & ~kFlag;
            mine += m > 0 && (dd == t_end || m == 1u || (c & kFlag));
        }
    }
    // block exclusive scan of `mine`
    int incl = mine;
#pragma unroll
    for(int o = 1; o < 32; o <<= 1)
    {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if(lane >= o)
            incl += v;
    }
    if(lane == 31)
        s_warp[wid] = (unsigned)incl;
    __syncthreads();
    if(wid == 0)
    {
        int v = (int)s_warp[lane], iv = v;
#pragma unroll
        for(int o = 1; o < 32; o <<= 1)
        {
            const int u = __shfl_up_sync(0xffffffffu, iv, o);
            if(lane >= o)
                iv += u;
        }
        s_warp[lane] = (unsigned)(iv - v);
        if(lane == 31)
            s_created = (unsigned)iv; // total
    }
    __syncthreads();
    int pos = (int)s_warp[wid] + incl - mine;
    const int total = (int)s_created;
    unsigned* out = sel + (long long)frame * P.sel_frame_stride + L.sel_off;
    {
        int dd = d, nsd = ns_d, offd = off_d, seqd = seq_d;
        for(int qq = q0; qq < q1; ++qq)
        {
            while(qq >= seqd + nsd)
                seqd += nsd, nsd >>= 2, offd -= nsd, --dd;
            const unsigned c = cnt[offd + (qq - seqd)];
            const unsigned m = c & ~kFlag;
            if(m > 0 && (dd == t_end || m == 1u || (c & kFlag)))
            {
                const unsigned b = best[offd + (qq - seqd)];
                const unsigned ord = 0xffffffu - (b & 0xffffffu);
                const int cell = ord >> 12, ci = cell / L.ncx, cj = cell - ci * L.ncx;
                const int y = ci * 64 + ((ord >> 6) & 63) + 3 + kBorder, x = cj * 64 + (ord & 63) + 3 + kBorder;
                if(pos < L.sel_cap)
                    out[pos] = (unsigned)y << 20 | (unsigned)x << 8 | (b >> 24); // absolute level coordinates
                ++pos;
            }
        }
    }
    if(tid == 0)
    {
        C->n_tree = (unsigned)min(total, L.sel_cap);
        C->passes = (unsigned)passes;
        if(total > L.sel_cap)
            C->err = 2;
    }
}

// =================================================================================================
// cv::GaussianBlur(7x7, sigma 2) on u8: OpenCV's fixed-point path, kernel [18 34 48 56 48 34 18] / 256
// =================================================================================================
__device__ __forceinline__ int reflect101_d(int p, int n)
{
    if(n == 1)
        return 0;
    while(p < 0 || p >= n)
        p = p < 0 ? -p : 2 * n - 2 - p;
    return p;
}

// tile: 64 columns x GT_H rows.  GT_H + 6 = 64 source rows = 32 row pairs x 16 column groups = two full rounds of the CTA in the
// horizontal pass (a 32-row tile had 19 pairs: 304 items for 256 threads), and 6 halo rows per 58 instead of per 32
constexpr int GT_H = 58, GT_IN = GT_H + 6, GT_RP = GT_IN / 2;
__global__ void __launch_bounds__(256) k_dist_gauss(const __grid_constant__ DistParams P, const uint8_t* __restrict__ pyr,
                                                    uint8_t* __restrict__ blur)
{
    __shared__ __align__(16) uint8_t in[GT_IN][76];
    __shared__ __align__(16) unsigned hp[GT_RP][64]; // horizontally filtered rows, stored as row pairs
    const int tid = threadIdx.x;
    const int l = level_of_gtile(P, blockIdx.x);
    const DistLevel& L = P.lv[l];
    const int t = blockIdx.x - L.gtile_base;
    const int ty = t / L.gtx, tx = t - ty * L.gtx;
    const int x0 = tx * 64, y0 = ty * GT_H;
    const uint8_t* img = pyr + L.img_off + (P.frame0 + blockIdx.y) * L.img_frame_stride;
    // shared row r holds columns x0 - 4 .. x0 + 71 of source row reflect(y0 + r - 3): 19 aligned words when the tile
    // is clear of the left / right image border, byte-wise with REFLECT_101 otherwise
    if(x0 >= 4 && x0 + 72 <= L.w)
    {
        const int lane = tid & 31;
        for(int r = tid >> 5; r < GT_IN; r += 8)
            if(lane < 19)
                reinterpret_cast<unsigned*>(&in[r][0])[lane] = __ldg(
                    reinterpret_cast<const unsigned*>(img + (long long)reflect101_d(y0 + r - 3, L.h) * L.pitch + x0 - 4) + lane);
    }
    else
        for(int i = tid; i < GT_IN * 70; i += 256)
        {
            const int r = i / 70, c = i - r * 70;
            const int sy = reflect101_d(y0 + r - 3, L.h), sx = reflect101_d(x0 + c - 3, L.w);
            in[r][c + 1] = __ldg(img + (long long)sy * L.pitch + sx);
        }
    __syncthreads();
    // horizontal pass, exact in 8.8: an item is (row pair, 4 columns); the 10 bytes a row needs come from three aligned
    // words, every output is two IDP.4A over byte windows cut with PRMT.  Results are stored as row PAIRS -- word
    // (rp, x) = h[2 rp][x] | h[2 rp + 1][x] << 16 -- so that the vertical pass can run on IDP.2A.
    for(int i = tid; i < GT_RP * 16; i += 256)
    {
        const int rp = i >> 4, c = (i & 15) * 4;
        unsigned h[2][4];
#pragma unroll
        for(int k = 0; k < 2; ++k)
        {
            const unsigned* w = reinterpret_cast<const unsigned*>(&in[2 * rp + k][c]);
            const unsigned w0 = w[0], w1 = w[1], w2 = w[2];
            const unsigned lo[4] = {__byte_perm(w0, w1, 0x4321), __byte_perm(w0, w1, 0x5432), __byte_perm(w0, w1, 0x6543), w1};
            const unsigned hi[4] = {__byte_perm(w1, w2, 0x4321), __byte_perm(w1, w2, 0x5432), __byte_perm(w1, w2, 0x6543), w2};
#pragma unroll
            for(int j = 0; j < 4; ++j) // taps 18 34 48 56 | 48 34 18 0
                h[k][j] = __dp4a(hi[j], 0x00122230u, __dp4a(lo[j], 0x38302212u, 0u));
        }
        *reinterpret_cast<uint4*>(&hp[rp][c]) =
            make_uint4(h[0][0] | h[1][0] << 16, h[0][1] | h[1][1] << 16, h[0][2] | h[1][2] << 16, h[0][3] | h[1][3] << 16);
    }
    __syncthreads();
    // vertical pass, exact in 16.16: output row y uses h rows y .. y+6 = row pairs (y >> 1) .. (y >> 1) + 3, with the tap
    // pairs shifted by one row when y is odd
    uint8_t* dst = blur + L.img_off + (P.frame0 + blockIdx.y) * L.img_frame_stride;
    for(int i = tid; i < GT_H * 16; i += 256)
    {
        const int y = i >> 4, xq = (i & 15) * 4;
        if(y0 + y >= L.h || x0 + xq >= L.w)
            continue;
        const bool odd = y & 1;
        const unsigned c0 = odd ? 0x30221200u : 0x38302212u; // (0 18 | 34 48)   or (18 34 | 48 56)
        const unsigned c1 = odd ? 0x12223038u : 0x00122230u; // (56 48 | 34 18)  or (48 34 | 18 0)
        const int rp = y >> 1;
        const uint4 a = *reinterpret_cast<const uint4*>(&hp[rp][xq]), b = *reinterpret_cast<const uint4*>(&hp[rp + 1][xq]);
        const uint4 c = *reinterpret_cast<const uint4*>(&hp[rp + 2][xq]), d = *reinterpret_cast<const uint4*>(&hp[rp + 3][xq]);
        const unsigned pa[4] = {a.x, a.y, a.z, a.w}, pb[4] = {b.x, b.y, b.z, b.w}, pc[4] = {c.x, c.y, c.z, c.w},
                       pd[4] = {d.x, d.y, d.z, d.w};
        unsigned o = 0;
#pragma unroll
        for(int j = 0; j < 4; ++j)
        {
            unsigned v = __dp2a_lo(pa[j], c0, 32768u);
            v = __dp2a_hi(pb[j], c0, v);
            v = __dp2a_lo(pc[j], c1, v);
            v = __dp2a_hi(pd[j], c1, v);
            o |= (v >> 16) << (8 * j);
        }
        *reinterpret_cast<unsigned*>(dst + (long long)(y0 + y) * L.pitch + x0 + xq) = o;
    }
}

// =================================================================================================
// orientation + descriptor (:543-570, :572-629): one warp per keypoint
// =================================================================================================
__device__ __forceinline__ float poly_cos(float v)
{
    const float c1 = 0.99940307f, c2 = -0.49558072f, c3 = 0.03679168f;
    const float v2 = __fmul_rn(v, v);
    return __fadd_rn(c1, __fmul_rn(v2, __fadd_rn(c2, __fmul_rn(c3, v2))));
}
__device__ __forceinline__ float util_cos(float v)
{
    const float PI = 3.14159265358979f;
    const float PI_2 = __fdiv_rn(PI, 2.0f), TWO_PI = __fmul_rn(2.0f, PI), INV_TWO_PI = __fdiv_rn(1.0f, TWO_PI),
                THREE_PI_2 = __fmul_rn(3.0f, PI_2);
    v = __fsub_rn(v, __fmul_rn((float)(int)floorf(__fmul_rn(v, INV_TWO_PI)), TWO_PI));
    v = (0.0f < v) ? v : -v;
    if(v < PI_2)
        return poly_cos(v);
    else if(v < PI)
        return -poly_cos(__fsub_rn(PI, v));
    else if(v < THREE_PI_2)
        return -poly_cos(__fsub_rn(v, PI));
    else
        return poly_cos(__fsub_rn(TWO_PI, v));
}

__global__ void __launch_bounds__(256) k_dist_describe(const __grid_constant__ DistParams P, const uint8_t* __restrict__ pyr,
                                                       const uint8_t* __restrict__ blur, const DistCounters* __restrict__ ctr,
                                                       const unsigned* __restrict__ sel, const float4* __restrict__ pat_t,
                                                       const uint2* __restrict__ ic_w, mslam_b200_keypoint* __restrict__ out_kp,
                                                       uint8_t* __restrict__ out_desc, int* __restrict__ out_counts, int out_cap)
{
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int frame = P.frame0 + blockIdx.y;
    const DistCounters* C = ctr + frame * P.nlevels;
    int l = 0, base = 0, total = 0;
    {
        int acc = 0;
        bool found = false;
        for(int i = 0; i < P.nlevels; ++i)
        {
            const int n = (int)C[i].n_tree;
            if(!found && g < acc + n)
                l = i, base = acc, found = true;
            acc += n;
        }
        total = acc;
    }
    if(g == 0 && lane == 0)
        out_counts[frame] = total;
    if(g >= total || g >= out_cap)
        return;
    const DistLevel& L = P.lv[l];
    const unsigned r = sel[(long long)frame * P.sel_frame_stride + L.sel_off + (g - base)];
    const int x = (r >> 8) & 0xfff, y = r >> 20;
    const uint8_t* img = pyr + L.img_off + frame * L.img_frame_stride;
    // intensity-centroid moments over the radius-15 disc (u_max of orb_impl::orb_impl, :521-541 -- OpenCV's table): the 31 x 31
    // patch as 31 rows x 9 ALIGNED words, every word against two packed signed-byte weight words (u resp. v inside the disc,
    // 0 outside; ic_weight_table of the ORB path, per alignment of the patch) in two IDP.4A -- 9 word loads per lane where
    // the first version issued 31 predicated byte loads
    int m10 = 0, m01 = 0;
    {
        constexpr int kIcWords = 31 * 9;
        const unsigned pitch = (unsigned)L.pitch;
        const int al = (x - 15) & 3;
        const uint8_t* base = img + (long long)(y - 15) * pitch + ((x - 15) & ~3); // keypoints keep 19 px from the border
        const uint2* wtab = ic_w + al * kIcWords;
#pragma unroll
        for(int it = 0; it < (kIcWords + 31) / 32; ++it)
        {
            const unsigned i = (unsigned)(lane + 32 * it);
            if(i < (unsigned)kIcWords)
            {
                const unsigned rr = i / 9u, kk = i - 9u * rr;
                const unsigned word = __ldg(reinterpret_cast<const unsigned*>(base + (rr * pitch + 4u * kk)));
                const uint2 wgt = __ldg(wtab + i);
                asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(m10) : "r"(word), "r"(wgt.x));
                asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(m01) : "r"(word), "r"(wgt.y));
            }
        }
        m10 = __reduce_add_sync(0xffffffffu, m10);
        m01 = __reduce_add_sync(0xffffffffu, m01);
    }
    const float angle = fast_atan2_deg((float)m01, (float)m10);
    // keypt.angle * M_PI / 180.0 in double, then float (:574)
    const float rad = (float)((double)angle * 3.14159265358979323846 / 180.0);
    const float ca = util_cos(rad), sa = util_cos(__fsub_rn(__fdiv_rn(3.14159265358979f, 2.0f), rad));
    // rotated tests: cvRound(v) = low bits of (v + 1.5 * 2^23) - 0x4B400000 (round-to-nearest-even, what __float2int_rn gave);
    // the two subtractions are folded into the base pointer: iy * pitch + ix = by * pitch + bx - 0x4B400000 * (pitch + 1)
    // (mod 2^32).  Pattern as floats (brief_pattern_transposed of the ORB path): no I2F / F2I, which run on the 16-lane XU
    // pipe and were 64 of them per lane and keypoint.
    const int pitch = L.pitch;
    const unsigned kfold = 0x4B400000u * (unsigned)(pitch + 1);
    const uint8_t* bc = blur + L.img_off + frame * L.img_frame_stride + (long long)y * pitch + x;
    unsigned byte = 0;
#pragma unroll
    for(int b = 0; b < 8; ++b)
    {
        const float4 t = __ldg(pat_t + b * 32 + lane); // test (lane * 8 + b): x0, y0, x1, y1
        const unsigned by0 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(t.x, sa), __fmul_rn(t.y, ca)), 12582912.f));
        const unsigned bx0 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(t.x, ca), __fmul_rn(t.y, sa)), 12582912.f));
        const unsigned by1 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(t.z, sa), __fmul_rn(t.w, ca)), 12582912.f));
        const unsigned bx1 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(t.z, ca), __fmul_rn(t.w, sa)), 12582912.f));
        const int o0 = (int)(by0 * (unsigned)pitch + bx0 - kfold), o1 = (int)(by1 * (unsigned)pitch + bx1 - kfold);
        const int v0 = __ldg(bc + o0), v1 = __ldg(bc + o1);
        byte |= (unsigned)(v0 < v1) << b;
    }
    out_desc[((long long)frame * out_cap + g) * 32 + lane] = (uint8_t)byte;
    if(lane == 0)
    {
        mslam_b200_keypoint k;
        const float px = (float)x, py = (float)y;
        k.x = l == 0 ? px : __fmul_rn(px, L.scale);
        k.y = l == 0 ? py : __fmul_rn(py, L.scale);
        k.size = L.size;
        k.angle = angle;
        k.response = (float)(r & 0xff);
        k.octave = l;
        out_kp[(long long)frame * out_cap + g] = k;
    }
}

// =================================================================================================
// host side
// =================================================================================================
inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// cv::resize INTER_LINEAR coefficient tables (OpenCV resize.cpp), built exactly as OpenCV builds them
void linear_taps_x(int src, int dst, XTap* t)
{
    const double inv_scale = (double)dst / (double)src, scale = 1.0 / inv_scale;
    for(int d = 0; d < dst; ++d)
    {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)std::floor(f);
        f -= (float)s;
        if(s < 0)
            f = 0.f, s = 0;
        if(s >= src - 1)
            f = 0.f, s = src - 1;
        t[d].ofs = s;
        t[d].a = (unsigned)lrintf((1.f - f) * 2048.f) | (unsigned)lrintf(f * 2048.f) << 16;
    }
}
void linear_taps_y(int src, int dst, YTap* t)
{
    const double inv_scale = (double)dst / (double)src, scale = 1.0 / inv_scale;
    for(int d = 0; d < dst; ++d)
    {
        float f = (float)((d + 0.5) * scale - 0.5);
        const int s = (int)std::floor(f);
        f -= (float)s;
        t[d].sy0 = std::min(std::max(s, 0), src - 1);
        t[d].sy1 = std::min(std::max(s + 1, 0), src - 1);
        t[d].b = (unsigned)lrintf((1.f - f) * 2048.f) | (unsigned)lrintf(f * 2048.f) << 16;
        t[d].pad = 0;
    }
}

// quad table of k_dist_resize_quads for one level (see the kernel); false when a quad's taps do not fit the 8-byte window
bool linear_quads(const XTap* xt, int dw, std::vector<uint4>* out)
{
    const int nq = (dw + 3) / 4;
    const size_t before = out->size();
    for(int q = 0; q < nq; ++q)
    {
        int ofs[4], base = 1 << 30;
        unsigned a[4];
        for(int j = 0; j < 4; ++j)
        {
            const XTap& t = xt[std::min(4 * q + j, dw - 1)];
            ofs[j] = t.ofs, a[j] = t.a;
            base = std::min(base, ofs[j]);
        }
        unsigned sel[4];
        for(int j = 0; j < 4; ++j)
        {
            if(ofs[j] - base + 1 > 7) // the right tap must lie inside the 8-byte window that starts at `base`
            {
                out->resize(before);
                return false;
            }
            sel[j] = (unsigned)(ofs[j] - base) | ((unsigned)(ofs[j] - base + 1) << 4);
        }
        out->push_back(make_uint4((unsigned)base, (sel[0] | (sel[1] << 8)) | ((sel[2] | (sel[3] << 8)) << 16), 0u, 0u));
        out->push_back(make_uint4(a[0], a[1], a[2], a[3]));
    }
    return true;
}

struct DistGeometry
{
    DistParams P;
    size_t pyr_bytes;
    long long cand_elems, tree_words;
    int sel_elems, kp_bound;
    size_t tree_smem;
    std::vector<XTap> xtab;
    std::vector<YTap> ytab;
    std::vector<uint4> qtab;
    bool ok;
};

void build_dist_geometry(const mslam_b200_dist_config& cfg, int W, int H, DistGeometry* g)
{
    DistParams& P = g->P;
    memset(&P, 0, sizeof(P));
    g->ok = false;
    g->xtab.clear();
    g->ytab.clear();
    g->qtab.clear();
    P.nlevels = cfg.nlevels;
    P.W = W, P.H = H;
    P.ini_thr = cfg.ini_fast_threshold;
    P.one = 1, P.minus_one = -1;
    P.min_thr = cfg.min_fast_threshold;
    P.min_size = (unsigned)cfg.min_size;
    size_t off = 0, tree_smem = 0;
    long long coff = 0, toff = 0;
    int cells = 0, gtiles = 0, soff = 0, xo = 0, yo = 0;
    float s = 1.0f;
    for(int l = 0; l < cfg.nlevels; ++l)
    {
        DistLevel& L = P.lv[l];
        if(l > 0)
            s = cfg.scale_factor * s; // orb_params::calc_scale_factors (:413-421)
        L.scale = s;
        L.size = (float)(unsigned)(31 * s); // fast_patch_size_ * scale_factor, truncated (:931)
        const double sc = (double)s;
        L.w = l == 0 ? W : (int)std::round(W * 1.0 / sc); // compute_image_pyramid (:835-836)
        L.h = l == 0 ? H : (int)std::round(H * 1.0 / sc);
        if(L.w < 2 * kBorder || L.h < 2 * kBorder)
            return; // the reference's unsigned border arithmetic underflows
        L.pitch = (int)align_up((size_t)L.w + 4, 64);
        L.img_off = (long long)off;
        L.img_frame_stride = (long long)align_up((size_t)L.pitch * L.h, 256);
        off += (size_t)L.img_frame_stride * cfg.max_batch;
        // cells: the region cv::FAST can report in, [22, w-22) x [22, h-22), in 64-px steps
        const int rw = L.w - 2 * kFastOrg, rh = L.h - 2 * kFastOrg;
        L.ncx = rw > 0 ? (rw + kCell - 1) / kCell : 0;
        L.ncy = rh > 0 ? (rh + kCell - 1) / kCell : 0;
        if(L.ncx == 0 || L.ncy == 0)
            L.ncx = L.ncy = 0;
        L.cell_base = cells;
        cells += L.ncx * L.ncy;
        int cap = 0;
        for(int i = 0; i < L.ncy; ++i)
            for(int j = 0; j < L.ncx; ++j)
            {
                const int cw = std::min(kCell, rw - j * kCell), ch = std::min(kCell, rh - i * kCell);
                cap += ((cw + 1) / 2) * ((ch + 1) / 2); // strict 3x3 maxima
            }
        L.cand_cap = cap;
        L.cand_off = coff;
        coff += (long long)align_up((size_t)cap, 4);
        L.gtx = (L.w + 63) / 64, L.gty = (L.h + GT_H - 1) / GT_H;
        L.gtile_base = gtiles;
        gtiles += L.gtx * L.gty;
        // tree geometry (initialize_nodes, :989-1011)
        const int tw = L.w - 2 * kBorder, th = L.h - 2 * kBorder;
        L.ninit_x = L.ninit_y = 1;
        L.delta_x = tw, L.delta_y = th;
        L.depth = 0;
        if(tw > 0 && th > 0)
        {
            const double ratio = (double)tw / th;
            if(ratio > 1)
            {
                L.ninit_x = (int)(unsigned)std::round(ratio);
                L.delta_x = (double)tw / (unsigned)L.ninit_x;
                L.delta_y = th;
            }
            else
            {
                L.ninit_y = (int)(unsigned)std::round(1 / ratio);
                L.delta_x = tw; // max_x - min_y in the reference (:1018); min_x == min_y == 19
                L.delta_y = (double)th / (unsigned)L.ninit_y;
            }
            if(L.ninit_x * L.ninit_y > kMaxInit)
                return;
            // deepest depth: descend the distinct node shapes until every one stops by area
            std::set<std::pair<int, int>> shapes;
            for(int iy = 0; iy < L.ninit_y; ++iy)
                for(int ix = 0; ix < L.ninit_x; ++ix)
                    shapes.insert({(int)(L.delta_x * (ix + 1)) - (int)(L.delta_x * ix), (int)(L.delta_y * (iy + 1)) - (int)(L.delta_y * iy)});
            int depth = 0;
            for(;; ++depth)
            {
                std::set<std::pair<int, int>> next;
                for(const auto& sh : shapes)
                {
                    const unsigned area = (unsigned)(sh.first * sh.second);
                    if((float)area * s * s <= (float)P.min_size)
                        continue;
                    const int hx = (sh.first + 1) / 2, hy = (sh.second + 1) / 2;
                    const int xs[2] = {hx, sh.first - hx}, ys[2] = {hy, sh.second - hy};
                    for(int a = 0; a < 2; ++a)
                        for(int b = 0; b < 2; ++b)
                            next.insert({xs[a], ys[b]});
                }
                if(next.empty() || depth >= 11)
                    break;
                shapes.swap(next);
            }
            L.depth = depth;
        }
        long long slots = 0, ns = (long long)L.ninit_x * L.ninit_y;
        for(int d = 0; d <= L.depth; ++d, ns *= 4)
            slots += ns;
        if(slots > (1 << 24))
            return;
        L.tree_slots = (int)slots;
        L.tree_in_smem = (size_t)slots * 8 <= (size_t)kTreeSmemBudget;
        L.tree_off = toff;
        if(L.tree_in_smem)
            tree_smem = std::max(tree_smem, (size_t)slots * 8);
        else
            toff += 2 * slots;
        long long deepest = (long long)L.ninit_x * L.ninit_y;
        for(int d = 0; d < L.depth; ++d)
            deepest *= 4;
        L.sel_cap = (int)std::min<long long>(deepest, L.cand_cap);
        L.sel_off = soff;
        soff += (int)align_up((size_t)L.sel_cap, 4);
        L.xtab_off = xo, L.ytab_off = yo;
        if(l > 0)
        {
            g->xtab.resize(xo + L.w);
            g->ytab.resize(yo + L.h);
            linear_taps_x(P.lv[l - 1].w, L.w, g->xtab.data() + xo);
            linear_taps_y(P.lv[l - 1].h, L.h, g->ytab.data() + yo);
            L.qtab_off = (int)g->qtab.size();
            L.nquads = linear_quads(g->xtab.data() + xo, L.w, &g->qtab) ? (L.w + 3) / 4 : 0;
            xo += L.w, yo += L.h;
        }
    }
    P.total_cells = cells;
    P.total_gtiles = gtiles;
    P.cand_frame_stride = std::max<long long>(coff, 4);
    P.tree_frame_stride = toff;
    P.sel_frame_stride = std::max(soff, 4);
    g->pyr_bytes = off;
    g->cand_elems = P.cand_frame_stride;
    g->tree_words = toff;
    g->sel_elems = P.sel_frame_stride;
    g->kp_bound = soff;
    g->tree_smem = tree_smem;
    g->ok = true;
}
} // namespace

struct mslam_b200_dist
{
    mslam_b200_dist_config cfg;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    long long launches = 0;
    char err[512] = {0};
    int geo_w = 0, geo_h = 0;
    DistParams P;
    size_t tree_smem = 0, tree_smem_max = 0;
    int kp_cap = 0;
    uint8_t *d_bgr = nullptr, *d_pyr = nullptr, *d_blur = nullptr;
    size_t pyr_bytes = 0;
    XTap* d_xtab = nullptr;
    YTap* d_ytab = nullptr;
    uint4* d_qtab = nullptr;
    float4* d_pat = nullptr; // rBRIEF pattern as floats, transposed (brief_pattern_transposed)
    uint2* d_icw = nullptr;  // intensity-centroid byte weights (ic_weight_table)
    int xtab_cap = 0, ytab_cap = 0, qtab_cap = 0;
    unsigned *d_cand = nullptr, *d_sel = nullptr, *d_tree = nullptr;
    long long cand_elems = 0, tree_words = 0;
    int sel_elems = 0;
    DistCounters* d_ctr = nullptr;
    mslam_b200_keypoint* d_kp = nullptr;
    uint8_t* d_desc = nullptr;
    int* d_counts = nullptr;
    int* h_counts = nullptr;      // [2 slots][max_batch]
    DistCounters* h_ctr = nullptr; // [2 slots][max_batch][kMaxLevels]
    size_t bgr_bytes = 0;
    struct PendingBatch
    {
        bool active = false;
        int n_frames = 0, capacity = 0;
        int* n_out = nullptr;
    } pend[2];
    int pend_next = 0;
    cudaEvent_t ev_slot_done[2] = {nullptr, nullptr}, ev_out_done = nullptr;
    bool out_done_valid = false;
    // single-frame kernel graph (one per frame slot; the same launches as dist_run, replayed)
    struct FrameGraph
    {
        cudaGraphExec_t exec = nullptr;
        int width = 0, height = 0, uses = 0;
        long long launches = 0;
    } graph1[2];
    int graphs_enabled = -1;
    // single-frame latency path (as mslam_b200_detect's, context.cu): result block [count, pad.. (16 B) | counters of the levels
    // (kMaxLevels) | kp_cap keypoints | kp_cap descriptors] + pinned mirror, filled by one captured graph
    uint8_t *d_one = nullptr, *h_one = nullptr;
    size_t one_bytes = 0, one_ctr_off = 0, one_kp_off = 0, one_desc_off = 0;
    FrameGraph graph_lat;
    int last_frames = 0;
    bool timing = false;
    cudaEvent_t ev[7] = {nullptr};
    cudaStream_t aux = nullptr; // forked stream of the blur
    cudaStream_t copy[2] = {nullptr, nullptr}; // copy-in / copy-out engines of the host entry point
    cudaEvent_t ev_h2d[kDistChunks] = {nullptr}, ev_done[kDistChunks] = {nullptr}, ev_cnt[kDistChunks] = {nullptr};
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    double t_ms[6] = {0};
};

static int dist_fail(mslam_b200_dist* ctx, cudaError_t e, const char* what)
{
    if(ctx)
        snprintf(ctx->err, sizeof(ctx->err), "%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
    return MSLAM_B200_E_CUDA;
}
#define DIST_CUDA_OK(expr)                                                                                             \
    do                                                                                                                 \
    {                                                                                                                  \
        cudaError_t e__ = (expr);                                                                                      \
        if(e__ != cudaSuccess)                                                                                         \
            return dist_fail(ctx, e__, #expr);                                                                         \
    } while(0)

static int dist_set_geometry(mslam_b200_dist* ctx, int W, int H)
{
    if(W == ctx->geo_w && H == ctx->geo_h)
        return MSLAM_B200_OK;
    if(W < 1 || H < 1 || W > ctx->cfg.max_width || H > ctx->cfg.max_height || W > 4095 || H > 4095)
        return MSLAM_B200_E_INVALID;
    DistGeometry g;
    build_dist_geometry(ctx->cfg, W, H, &g);
    if(!g.ok)
        return MSLAM_B200_E_INVALID;
    if(g.pyr_bytes > ctx->pyr_bytes || g.cand_elems > ctx->cand_elems || g.tree_words > ctx->tree_words ||
       g.sel_elems > ctx->sel_elems || (int)g.xtab.size() > ctx->xtab_cap || (int)g.ytab.size() > ctx->ytab_cap || (int)g.qtab.size() > ctx->qtab_cap ||
       g.tree_smem > ctx->tree_smem_max)
        return MSLAM_B200_E_INVALID;
    if(!g.xtab.empty())
    {
        DIST_CUDA_OK(cudaMemcpyAsync(ctx->d_xtab, g.xtab.data(), g.xtab.size() * sizeof(XTap), cudaMemcpyHostToDevice, ctx->stream));
        DIST_CUDA_OK(cudaMemcpyAsync(ctx->d_ytab, g.ytab.data(), g.ytab.size() * sizeof(YTap), cudaMemcpyHostToDevice, ctx->stream));
        if(!g.qtab.empty())
            DIST_CUDA_OK(cudaMemcpyAsync(ctx->d_qtab, g.qtab.data(), g.qtab.size() * sizeof(uint4), cudaMemcpyHostToDevice, ctx->stream));
        DIST_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    }
    ctx->P = g.P;
    ctx->P.kp_cap = ctx->kp_cap;
    ctx->tree_smem = g.tree_smem;
    ctx->geo_w = W, ctx->geo_h = H;
    return MSLAM_B200_OK;
}

extern "C" {

void mslam_b200_dist_default_config(mslam_b200_dist_config* cfg)
{
    if(!cfg)
        return;
    memset(cfg, 0, sizeof(*cfg));
    cfg->struct_size = (int)sizeof(*cfg);
    cfg->max_width = 640;
    cfg->max_height = 480;
    cfg->max_batch = 1;
    cfg->nlevels = 8;
    cfg->scale_factor = 1.2f;
    cfg->ini_fast_threshold = 20;
    cfg->min_fast_threshold = 7;
    cfg->min_size = 1000;
}

void mslam_b200_dist_destroy(mslam_b200_dist* ctx)
{
    if(!ctx)
        return;
    cudaSetDevice(ctx->cfg.device);
    if(ctx->stream)
        cudaStreamSynchronize(ctx->stream);
    void* dev[] = {ctx->d_pat, ctx->d_icw, ctx->d_qtab, ctx->d_bgr, ctx->d_pyr, ctx->d_blur, ctx->d_xtab, ctx->d_ytab, ctx->d_cand, ctx->d_sel, ctx->d_tree,
                   ctx->d_ctr, ctx->d_kp,  ctx->d_desc, ctx->d_counts};
    for(void* p : dev)
        if(p)
            cudaFree(p);
    if(ctx->h_counts)
        cudaFreeHost(ctx->h_counts);
    if(ctx->h_ctr)
        cudaFreeHost(ctx->h_ctr);
    if(ctx->h_one)
        cudaFreeHost(ctx->h_one);
    if(ctx->d_one)
        cudaFree(ctx->d_one);
    if(ctx->graph_lat.exec)
        cudaGraphExecDestroy(ctx->graph_lat.exec);
    for(cudaEvent_t e : ctx->ev)
        if(e)
            cudaEventDestroy(e);
    for(int i = 0; i < kDistChunks; ++i)
        for(cudaEvent_t e : {ctx->ev_h2d[i], ctx->ev_done[i], ctx->ev_cnt[i]})
            if(e)
                cudaEventDestroy(e);
    for(int i = 0; i < 2; ++i)
        if(ctx->copy[i])
        {
            cudaStreamSynchronize(ctx->copy[i]);
            cudaStreamDestroy(ctx->copy[i]);
        }
    for(auto& g : ctx->graph1)
        if(g.exec)
            cudaGraphExecDestroy(g.exec);
    for(cudaEvent_t e : {ctx->ev_slot_done[0], ctx->ev_slot_done[1], ctx->ev_out_done})
        if(e)
            cudaEventDestroy(e);
    if(ctx->ev_fork)
        cudaEventDestroy(ctx->ev_fork);
    if(ctx->ev_join)
        cudaEventDestroy(ctx->ev_join);
    if(ctx->aux)
    {
        cudaStreamSynchronize(ctx->aux);
        cudaStreamDestroy(ctx->aux);
    }
    if(ctx->own_stream && ctx->stream)
        cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int mslam_b200_dist_create(const mslam_b200_dist_config* cfg_in, mslam_b200_dist** out)
{
    if(!cfg_in || !out || cfg_in->struct_size != (int)sizeof(mslam_b200_dist_config))
        return MSLAM_B200_E_INVALID;
    *out = nullptr;
    const mslam_b200_dist_config& c = *cfg_in;
    if(c.max_width < 2 * kBorder || c.max_height < 2 * kBorder || c.max_width > 4095 || c.max_height > 4095 || c.max_batch < 1 ||
       c.nlevels < 1 || c.nlevels > kMaxLevels || !(c.scale_factor > 1.f) || c.ini_fast_threshold < 1 ||
       c.ini_fast_threshold > 254 || c.min_fast_threshold < 1 || c.min_fast_threshold > c.ini_fast_threshold || c.min_size < 1)
        return MSLAM_B200_E_INVALID;
    int ndev = 0;
    if(cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1 || c.device < 0 || c.device >= ndev)
        return MSLAM_B200_E_CUDA; // no CPU fallback
    mslam_b200_dist* ctx = new(std::nothrow) mslam_b200_dist;
    if(!ctx)
        return MSLAM_B200_E_NOMEM;
    ctx->cfg = c;
    DistGeometry g;
    build_dist_geometry(c, c.max_width, c.max_height, &g);
    if(!g.ok)
    {
        delete ctx;
        return MSLAM_B200_E_INVALID;
    }
    ctx->kp_cap = c.max_keypoints > 0 ? c.max_keypoints : g.kp_bound;
    auto fail = [&](int rc) {
        mslam_b200_dist_destroy(ctx);
        return rc;
    };
#define CREATE_OK(expr)                                                                                                \
    do                                                                                                                 \
    {                                                                                                                  \
        if((expr) != cudaSuccess)                                                                                      \
            return fail(MSLAM_B200_E_CUDA);                                                                            \
    } while(0)
    CREATE_OK(cudaSetDevice(c.device));
    if(c.stream)
        ctx->stream = (cudaStream_t)c.stream;
    else
    {
        CREATE_OK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
        ctx->own_stream = true;
    }
    for(auto& e : ctx->ev)
        CREATE_OK(cudaEventCreate(&e));
    CREATE_OK(cudaStreamCreateWithFlags(&ctx->aux, cudaStreamNonBlocking));
    for(int i = 0; i < 2; ++i)
        CREATE_OK(cudaStreamCreateWithFlags(&ctx->copy[i], cudaStreamNonBlocking));
    for(int i = 0; i < kDistChunks; ++i)
    {
        CREATE_OK(cudaEventCreateWithFlags(&ctx->ev_h2d[i], cudaEventDisableTiming));
        CREATE_OK(cudaEventCreateWithFlags(&ctx->ev_done[i], cudaEventDisableTiming));
        CREATE_OK(cudaEventCreateWithFlags(&ctx->ev_cnt[i], cudaEventDisableTiming));
    }
    CREATE_OK(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
    CREATE_OK(cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
    const size_t B = (size_t)c.max_batch;
    // a little slack on the geometry-dependent sizes: level sizes are monotone in (W, H) but row pitches round up
    ctx->pyr_bytes = g.pyr_bytes + 4096;
    ctx->cand_elems = g.cand_elems + 64;
    ctx->tree_words = g.tree_words;
    ctx->sel_elems = g.sel_elems + 64;
    ctx->xtab_cap = (int)g.xtab.size() + 16;
    ctx->ytab_cap = (int)g.ytab.size() + 16;
    ctx->qtab_cap = ctx->xtab_cap / 2 + 4 * kMaxLevels + 16; // two entries per quad, (w + 3) / 4 quads per level: monotone in the frame size
    ctx->tree_smem_max = g.tree_smem;
    ctx->bgr_bytes = B * c.max_width * c.max_height * 3;
    CREATE_OK(cudaMalloc(&ctx->d_bgr, 2 * ctx->bgr_bytes)); // two pipeline slots
    // + 256: k_dist_resize_quads reads whole 12-byte windows, up to 10 bytes past a row's last pixel (weights there are 0)
    CREATE_OK(cudaMalloc(&ctx->d_pyr, ctx->pyr_bytes + 256));
    CREATE_OK(cudaMalloc(&ctx->d_blur, ctx->pyr_bytes));
    CREATE_OK(cudaMemset(ctx->d_pyr, 0, ctx->pyr_bytes));
    CREATE_OK(cudaMemset(ctx->d_blur, 0, ctx->pyr_bytes));
    CREATE_OK(cudaMalloc(&ctx->d_xtab, ctx->xtab_cap * sizeof(XTap)));
    CREATE_OK(cudaMalloc(&ctx->d_ytab, ctx->ytab_cap * sizeof(YTap)));
    CREATE_OK(cudaMalloc(&ctx->d_qtab, ctx->qtab_cap * sizeof(uint4)));
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
    CREATE_OK(cudaMalloc(&ctx->d_cand, B * ctx->cand_elems * sizeof(unsigned)));
    CREATE_OK(cudaMalloc(&ctx->d_sel, B * ctx->sel_elems * sizeof(unsigned)));
    if(ctx->tree_words > 0)
        CREATE_OK(cudaMalloc(&ctx->d_tree, B * ctx->tree_words * sizeof(unsigned)));
    CREATE_OK(cudaMalloc(&ctx->d_ctr, B * kMaxLevels * sizeof(DistCounters)));
    CREATE_OK(cudaMalloc(&ctx->d_kp, B * ctx->kp_cap * sizeof(mslam_b200_keypoint)));
    CREATE_OK(cudaMalloc(&ctx->d_desc, B * ctx->kp_cap * 32));
    CREATE_OK(cudaMalloc(&ctx->d_counts, B * sizeof(int)));
    CREATE_OK(cudaMallocHost(&ctx->h_counts, 2 * B * sizeof(int)));
    CREATE_OK(cudaMallocHost(&ctx->h_ctr, 2 * B * kMaxLevels * sizeof(DistCounters)));
    ctx->one_ctr_off = 16;
    ctx->one_kp_off = ctx->one_ctr_off + kMaxLevels * sizeof(DistCounters);
    ctx->one_desc_off = (ctx->one_kp_off + (size_t)ctx->kp_cap * sizeof(mslam_b200_keypoint) + 31) / 32 * 32;
    ctx->one_bytes = ctx->one_desc_off + (size_t)ctx->kp_cap * 32;
    CREATE_OK(cudaMalloc(&ctx->d_one, ctx->one_bytes));
    CREATE_OK(cudaMallocHost(&ctx->h_one, ctx->one_bytes));
    for(int i = 0; i < 2; ++i)
        CREATE_OK(cudaEventCreateWithFlags(&ctx->ev_slot_done[i], cudaEventDisableTiming));
    CREATE_OK(cudaEventCreateWithFlags(&ctx->ev_out_done, cudaEventDisableTiming));
    ctx->tree_smem_max = kTreeSmemBudget; // any level that fits the budget runs from shared memory
    CREATE_OK(cudaFuncSetAttribute(k_dist_tree, cudaFuncAttributeMaxDynamicSharedMemorySize, kTreeSmemBudget));
#undef CREATE_OK
    *out = ctx;
    return MSLAM_B200_OK;
}

int mslam_b200_dist_synchronize(mslam_b200_dist* ctx)
{
    if(!ctx)
        return MSLAM_B200_E_INVALID;
    DIST_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    return MSLAM_B200_OK;
}

const char* mslam_b200_dist_last_cuda_error(mslam_b200_dist* ctx) { return ctx ? ctx->err : ""; }
long long mslam_b200_dist_launch_count(mslam_b200_dist* ctx) { return ctx ? ctx->launches : 0; }
int mslam_b200_dist_keypoint_capacity(mslam_b200_dist* ctx) { return ctx ? ctx->kp_cap : 0; }

int mslam_b200_dist_level_geometry(mslam_b200_dist* ctx, int width, int height, int* w, int* h, float* scale)
{
    if(!ctx || width < 1 || height < 1)
        return MSLAM_B200_E_INVALID;
    DistGeometry g;
    build_dist_geometry(ctx->cfg, width, height, &g);
    if(!g.ok)
        return MSLAM_B200_E_INVALID;
    for(int l = 0; l < ctx->cfg.nlevels; ++l)
    {
        if(w) w[l] = g.P.lv[l].w;
        if(h) h[l] = g.P.lv[l].h;
        if(scale) scale[l] = g.P.lv[l].scale;
    }
    return MSLAM_B200_OK;
}

// enqueue the whole detector for n_frames frames already in d_bgr
static int dist_run(mslam_b200_dist* ctx, const uint8_t* d_bgr, int frame0, int n_frames, mslam_b200_keypoint* d_kp,
                    uint8_t* d_desc, int capacity, int* d_counts)
{
    DistParams& P = ctx->P;
    P.n_frames = n_frames;
    P.frame0 = frame0;
    cudaStream_t st = ctx->stream;
    const bool tm = ctx->timing;
    DIST_CUDA_OK(cudaMemsetAsync(ctx->d_ctr + (size_t)frame0 * P.nlevels, 0, (size_t)n_frames * P.nlevels * sizeof(DistCounters), st));
    if(tm)
        cudaEventRecord(ctx->ev[0], st);
    {
        // toGrayScale into level 0 (k_gray16 / k_gray1 of the ORB path; same RgbFrame layout, frame.cpp:6-27)
        PipeParams G;
        memset(&G, 0, sizeof(G));
        G.nlevels = 1, G.n_frames = n_frames, G.W = P.W, G.H = P.H, G.frame0 = frame0;
        G.lv[0].w = P.W, G.lv[0].h = P.H, G.lv[0].pitch = P.lv[0].pitch;
        G.lv[0].img_off = P.lv[0].img_off, G.lv[0].img_frame_stride = P.lv[0].img_frame_stride;
        launch_gray(st, d_bgr, ctx->d_pyr, G, &ctx->launches);
    }
    if(tm)
        cudaEventRecord(ctx->ev[1], st);
    for(int l = 1; l < P.nlevels; ++l)
    {
        const DistLevel &S = P.lv[l - 1], &D = P.lv[l];
        if(D.nquads > 0)
            k_dist_resize_quads<<<dim3((D.nquads + 63) / 64, (D.h + 3) / 4, n_frames), 256, 0, st>>>(
                ctx->d_pyr + S.img_off + frame0 * S.img_frame_stride, S.pitch, S.img_frame_stride,
                ctx->d_pyr + D.img_off + frame0 * D.img_frame_stride, D.pitch, D.img_frame_stride, D.nquads, D.h,
                ctx->d_qtab + D.qtab_off, ctx->d_ytab + D.ytab_off);
        else
        k_dist_resize<<<dim3((D.w + 255) / 256, (D.h + 3) / 4, n_frames), 256, 0, st>>>(
            ctx->d_pyr + S.img_off + frame0 * S.img_frame_stride, S.pitch, S.img_frame_stride, S.w,
            ctx->d_pyr + D.img_off + frame0 * D.img_frame_stride, D.pitch, D.img_frame_stride, D.w,
            D.h, ctx->d_xtab + D.xtab_off, ctx->d_ytab + D.ytab_off);
        ++ctx->launches;
    }
    if(tm)
        cudaEventRecord(ctx->ev[2], st);
    // the blur depends on the pyramid only: unless stages are being timed it runs on a forked stream beside FAST + tree
    cudaStream_t bs = tm ? st : ctx->aux;
    if(!tm)
    {
        DIST_CUDA_OK(cudaEventRecord(ctx->ev_fork, st));
        DIST_CUDA_OK(cudaStreamWaitEvent(ctx->aux, ctx->ev_fork, 0));
        k_dist_gauss<<<dim3(P.total_gtiles, n_frames), 256, 0, bs>>>(P, ctx->d_pyr, ctx->d_blur);
        ++ctx->launches;
        DIST_CUDA_OK(cudaEventRecord(ctx->ev_join, ctx->aux));
    }
    if(P.total_cells > 0)
    {
        k_dist_fast<<<dim3(P.total_cells, n_frames), 256, 0, st>>>(P, ctx->d_pyr, ctx->d_cand, ctx->d_ctr);
        ++ctx->launches;
    }
    if(tm)
        cudaEventRecord(ctx->ev[3], st);
    k_dist_tree<<<dim3(n_frames, P.nlevels), kTreeThreads, ctx->tree_smem, st>>>(P, ctx->d_cand, ctx->d_ctr, ctx->d_sel, ctx->d_tree);
    ++ctx->launches;
    if(tm)
    {
        cudaEventRecord(ctx->ev[4], st);
        k_dist_gauss<<<dim3(P.total_gtiles, n_frames), 256, 0, st>>>(P, ctx->d_pyr, ctx->d_blur);
        ++ctx->launches;
        cudaEventRecord(ctx->ev[5], st);
    }
    else
        DIST_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_join, 0));
    const int slots = std::min(capacity, ctx->kp_cap);
    k_dist_describe<<<dim3((std::max(slots, 1) + 7) / 8, n_frames), 256, 0, st>>>(P, ctx->d_pyr, ctx->d_blur, ctx->d_ctr, ctx->d_sel,
                                                                                  ctx->d_pat, ctx->d_icw, d_kp, d_desc, d_counts, capacity);
    ++ctx->launches;
    if(tm)
        cudaEventRecord(ctx->ev[6], st);
    DIST_CUDA_OK(cudaGetLastError());
    if(tm)
    {
        DIST_CUDA_OK(cudaEventSynchronize(ctx->ev[6]));
        for(int i = 0; i < 6; ++i)
        {
            float ms = 0;
            cudaEventElapsedTime(&ms, ctx->ev[i], ctx->ev[i + 1]);
            ctx->t_ms[i] += ms;
        }
    }
    ctx->last_frames = n_frames;
    return MSLAM_B200_OK;
}

// One frame through dist_run, replayed from a captured graph once its geometry has been seen (a frame is ~13 small
// launches: enqueueing them costs more host time than they run for).  A failed capture switches the graphs off.
static int dist_one_graphed(mslam_b200_dist* ctx, int slot, const uint8_t* d_bgr, int width, int height)
{
    if(ctx->graphs_enabled < 0)
    {
        const char* e = getenv("MSLAM_B200_GRAPH");
        ctx->graphs_enabled = e ? atoi(e) != 0 : 1;
    }
    auto& g = ctx->graph1[slot];
    if(!ctx->graphs_enabled || ctx->timing)
        return dist_run(ctx, d_bgr, 0, 1, ctx->d_kp, ctx->d_desc, ctx->kp_cap, ctx->d_counts);
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
        bool ok = cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeRelaxed) == cudaSuccess;
        if(ok)
        {
            ok = dist_run(ctx, d_bgr, 0, 1, ctx->d_kp, ctx->d_desc, ctx->kp_cap, ctx->d_counts) == MSLAM_B200_OK;
            ok = (cudaStreamEndCapture(ctx->stream, &graph) == cudaSuccess) && ok && graph;
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
        return dist_run(ctx, d_bgr, 0, 1, ctx->d_kp, ctx->d_desc, ctx->kp_cap, ctx->d_counts);
    ctx->P.n_frames = 1, ctx->P.frame0 = 0;
    ctx->last_frames = 1;
    DIST_CUDA_OK(cudaGraphLaunch(g.exec, ctx->stream));
    ctx->launches += g.launches;
    return MSLAM_B200_OK;
}

int mslam_b200_dist_detect_batch_dev(mslam_b200_dist* ctx, const uint8_t* d_bgr, int n_frames, int width, int height,
                                     mslam_b200_keypoint* d_kp, uint8_t* d_desc, int capacity, int* d_counts)
{
    if(!ctx || !d_bgr || !d_kp || !d_desc || !d_counts || n_frames < 1 || n_frames > ctx->cfg.max_batch || capacity < 1)
        return MSLAM_B200_E_INVALID;
    DIST_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    const int rc = dist_set_geometry(ctx, width, height);
    if(rc)
        return rc;
    return dist_run(ctx, d_bgr, 0, n_frames, d_kp, d_desc, capacity, d_counts);
}

// Host batches flow through a chunked pipeline: copy-in of chunk c+1 (copy engine 1) overlaps the kernels of chunk c and
// the copy-out of chunk c-1 (copy engine 2); with mslam_b200_dist_detect_batch_begin / _end two whole batches may be in
// flight (two frame-buffer slots), so the copy-in of the next batch overlaps everything of the current one.
static int dist_batch_args_ok(mslam_b200_dist* ctx, const uint8_t* const* frames, int n_frames, int width, int stride_bytes,
                              const void* out_kp, const void* out_desc, int capacity, const int* n_out)
{
    if(!ctx || !frames || !n_out || n_frames < 1 || n_frames > ctx->cfg.max_batch || capacity < 0 ||
       (capacity > 0 && (!out_kp || !out_desc)) || stride_bytes < 3 * width)
        return 0;
    for(int i = 0; i < n_frames; ++i)
        if(!frames[i])
            return 0;
    return 1;
}

int mslam_b200_dist_detect_batch_begin(mslam_b200_dist* ctx, const uint8_t* const* frames, int n_frames, int width, int height,
                                       int stride_bytes, mslam_b200_keypoint* out_kp, uint8_t* out_desc, int capacity, int* n_out)
{
    if(!dist_batch_args_ok(ctx, frames, n_frames, width, stride_bytes, out_kp, out_desc, capacity, n_out))
        return MSLAM_B200_E_INVALID;
    const int slot = ctx->pend_next;
    if(ctx->pend[slot].active)
        return MSLAM_B200_E_INVALID; // two batches are already in flight
    DIST_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    int rc = dist_set_geometry(ctx, width, height);
    if(rc)
        return rc;
    const size_t fbytes = (size_t)width * height * 3, B = (size_t)ctx->cfg.max_batch;
    uint8_t* d_bgr = ctx->d_bgr + (size_t)slot * ctx->bgr_bytes;
    // with another batch in flight the overlap comes from the batch pipeline itself; whole-batch launches then keep the
    // kernels at their best occupancy (MSLAM_B200_DIST_CHUNKS overrides)
    static int forced_chunks = -1;
    if(forced_chunks < 0)
    {
        const char* e = getenv("MSLAM_B200_DIST_CHUNKS");
        forced_chunks = e ? std::max(1, std::min(kDistChunks, atoi(e))) : 0;
    }
    const bool pipelined = ctx->pend[slot ^ 1].active;
    int nchunks = forced_chunks ? forced_chunks : (pipelined ? 1 : kDistChunks);
    nchunks = ctx->timing ? 1 : std::min(nchunks, std::max(1, n_frames / 4));
    const int nl = ctx->P.nlevels, icap = ctx->kp_cap, ncopy = std::min(capacity, icap);
    int bounds[kDistChunks + 1];
    for(int c = 0; c <= nchunks; ++c)
        bounds[c] = (int)((long long)n_frames * c / nchunks);
    // the previous batch's results must have left d_kp / d_desc / d_counts / d_ctr before this batch's kernels reuse them
    if(ctx->out_done_valid)
        DIST_CUDA_OK(cudaStreamWaitEvent(ctx->stream, ctx->ev_out_done, 0));
    for(int c = 0; c < nchunks; ++c)
    {
        for(int i = bounds[c]; i < bounds[c + 1]; ++i)
            DIST_CUDA_OK(cudaMemcpy2DAsync(d_bgr + i * fbytes, (size_t)width * 3, frames[i], (size_t)stride_bytes,
                                           (size_t)width * 3, height, cudaMemcpyHostToDevice, ctx->copy[0]));
        DIST_CUDA_OK(cudaEventRecord(ctx->ev_h2d[c], ctx->copy[0]));
    }
    for(int c = 0; c < nchunks; ++c)
    {
        const int b0 = bounds[c], nb = bounds[c + 1] - bounds[c];
        DIST_CUDA_OK(cudaStreamWaitEvent(ctx->stream, ctx->ev_h2d[c], 0));
        rc = n_frames == 1 ? dist_one_graphed(ctx, slot, d_bgr, width, height)
                           : dist_run(ctx, d_bgr, b0, nb, ctx->d_kp, ctx->d_desc, icap, ctx->d_counts);
        if(rc)
            return rc;
        DIST_CUDA_OK(cudaEventRecord(ctx->ev_done[c], ctx->stream));
        DIST_CUDA_OK(cudaStreamWaitEvent(ctx->copy[1], ctx->ev_done[c], 0));
        DIST_CUDA_OK(cudaMemcpyAsync(ctx->h_counts + slot * B + b0, ctx->d_counts + b0, nb * sizeof(int), cudaMemcpyDeviceToHost,
                                     ctx->copy[1]));
        DIST_CUDA_OK(cudaMemcpyAsync(ctx->h_ctr + (slot * B + b0) * kMaxLevels, ctx->d_ctr + (size_t)b0 * nl,
                                     (size_t)nb * nl * sizeof(DistCounters), cudaMemcpyDeviceToHost, ctx->copy[1]));
        if(ncopy > 0)
        {
            // whole rows of the per-frame output blocks (the count is not known on the host yet)
            DIST_CUDA_OK(cudaMemcpy2DAsync(out_kp + (size_t)b0 * capacity, (size_t)capacity * sizeof(mslam_b200_keypoint),
                                           ctx->d_kp + (size_t)b0 * icap, (size_t)icap * sizeof(mslam_b200_keypoint),
                                           (size_t)ncopy * sizeof(mslam_b200_keypoint), (size_t)nb, cudaMemcpyDeviceToHost, ctx->copy[1]));
            DIST_CUDA_OK(cudaMemcpy2DAsync(out_desc + (size_t)b0 * capacity * 32, (size_t)capacity * 32,
                                           ctx->d_desc + (size_t)b0 * icap * 32, (size_t)icap * 32, (size_t)ncopy * 32, (size_t)nb,
                                           cudaMemcpyDeviceToHost, ctx->copy[1]));
        }
    }
    DIST_CUDA_OK(cudaEventRecord(ctx->ev_slot_done[slot], ctx->copy[1]));
    DIST_CUDA_OK(cudaEventRecord(ctx->ev_out_done, ctx->copy[1]));
    ctx->out_done_valid = true;
    ctx->pend[slot].active = true;
    ctx->pend[slot].n_frames = n_frames;
    ctx->pend[slot].capacity = capacity;
    ctx->pend[slot].n_out = n_out;
    ctx->pend_next = slot ^ 1;
    return MSLAM_B200_OK;
}

int mslam_b200_dist_detect_batch_end(mslam_b200_dist* ctx)
{
    if(!ctx)
        return MSLAM_B200_E_INVALID;
    int slot = ctx->pend_next; // the oldest batch in flight
    if(!ctx->pend[slot].active)
        slot ^= 1;
    if(!ctx->pend[slot].active)
        return MSLAM_B200_E_INVALID;
    DIST_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    mslam_b200_dist::PendingBatch& pb = ctx->pend[slot];
    pb.active = false;
    DIST_CUDA_OK(cudaEventSynchronize(ctx->ev_slot_done[slot]));
    if(!ctx->pend[slot ^ 1].active)
        ctx->out_done_valid = false; // the copy-out stream has drained
    const size_t B = (size_t)ctx->cfg.max_batch;
    const int nl = ctx->P.nlevels;
    int rc = MSLAM_B200_OK;
    for(int i = 0; i < pb.n_frames; ++i)
    {
        for(int l = 0; l < nl; ++l)
            if(ctx->h_ctr[(slot * B + i) * kMaxLevels + l].err)
                rc = MSLAM_B200_E_CAPACITY; // an internal list overflowed (cannot happen: capacities are upper bounds)
        const int n = ctx->h_counts[slot * B + i];
        pb.n_out[i] = n;
        if(n > pb.capacity || n > ctx->kp_cap)
            rc = MSLAM_B200_E_CAPACITY;
    }
    return rc;
}

// One frame with nothing else in flight (DistributedOrbOpenCvDetector::detect as the Viewer calls it, viewer.cpp:183): copy-in,
// ONE captured graph (dist_run + the frame's counters and result block to their pinned mirror) and one wait, all on the context
// stream; the caller's arrays are filled by the host.  Returns kDistGeneralPath when the call has to take the batch machinery
// (first call of a geometry, graphs off, stage timing on, a pipelined batch in flight).
constexpr int kDistGeneralPath = 1; // not a status of the C ABI (those are <= 0)
static int dist_one_latency(mslam_b200_dist* ctx, const uint8_t* bgr, int width, int height, int stride_bytes,
                            mslam_b200_keypoint* out_kp, uint8_t* out_desc, int capacity, int* n_out)
{
    if(ctx->graphs_enabled < 0)
    {
        const char* e = getenv("MSLAM_B200_GRAPH");
        ctx->graphs_enabled = e ? atoi(e) != 0 : 1;
    }
    static int lat_env = -1; // A/B aid: MSLAM_B200_LATENCY_PATH=0
    if(lat_env < 0)
    {
        const char* e = getenv("MSLAM_B200_LATENCY_PATH");
        lat_env = e ? atoi(e) != 0 : 1;
    }
    if(!lat_env || !ctx->graphs_enabled || ctx->timing || ctx->pend[0].active || ctx->pend[1].active)
        return kDistGeneralPath;
    DIST_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    const int rc = dist_set_geometry(ctx, width, height);
    if(rc)
        return rc;
    cudaStream_t st = ctx->stream;
    auto& g = ctx->graph_lat;
    if(g.width != width || g.height != height)
    {
        if(g.exec)
            cudaGraphExecDestroy(g.exec);
        g = mslam_b200_dist::FrameGraph{};
        g.width = width, g.height = height;
    }
    const int nl = ctx->P.nlevels;
    if(!g.exec)
    {
        if(g.uses++ < 1)
            return kDistGeneralPath; // first call of a geometry: direct launches
        const long long l0 = ctx->launches;
        cudaGraph_t graph = nullptr;
        bool ok = cudaStreamBeginCapture(st, cudaStreamCaptureModeRelaxed) == cudaSuccess;
        if(ok)
        {
            ok = dist_run(ctx, ctx->d_bgr, 0, 1, reinterpret_cast<mslam_b200_keypoint*>(ctx->d_one + ctx->one_kp_off),
                          ctx->d_one + ctx->one_desc_off, ctx->kp_cap, reinterpret_cast<int*>(ctx->d_one)) == MSLAM_B200_OK;
            // the level counters stay where mslam_b200_dist_debug_read looks for them; a copy joins the result block
            ok = cudaMemcpyAsync(ctx->d_one + ctx->one_ctr_off, ctx->d_ctr, nl * sizeof(DistCounters), cudaMemcpyDeviceToDevice, st) == cudaSuccess && ok;
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
            return kDistGeneralPath;
        }
    }
    if(ctx->out_done_valid)
    {
        DIST_CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_out_done, 0));
        ctx->out_done_valid = false;
    }
    const size_t row = (size_t)width * 3;
    if((size_t)stride_bytes == row)
        DIST_CUDA_OK(cudaMemcpyAsync(ctx->d_bgr, bgr, row * height, cudaMemcpyHostToDevice, st));
    else
        DIST_CUDA_OK(cudaMemcpy2DAsync(ctx->d_bgr, row, bgr, (size_t)stride_bytes, row, (size_t)height, cudaMemcpyHostToDevice, st));
    ctx->P.n_frames = 1, ctx->P.frame0 = 0;
    ctx->last_frames = 1;
    DIST_CUDA_OK(cudaGraphLaunch(g.exec, st));
    ctx->launches += g.launches;
    DIST_CUDA_OK(cudaStreamSynchronize(st));
    const int n = *reinterpret_cast<const int*>(ctx->h_one);
    const DistCounters* hc = reinterpret_cast<const DistCounters*>(ctx->h_one + ctx->one_ctr_off);
    int status = MSLAM_B200_OK;
    for(int l = 0; l < nl; ++l)
        if(hc[l].err)
            status = MSLAM_B200_E_CAPACITY;
    if(n > capacity || n > ctx->kp_cap)
        status = MSLAM_B200_E_CAPACITY;
    const int m = std::max(0, std::min(std::min(n, capacity), ctx->kp_cap));
    if(m > 0)
    {
        memcpy(out_kp, ctx->h_one + ctx->one_kp_off, (size_t)m * sizeof(mslam_b200_keypoint));
        memcpy(out_desc, ctx->h_one + ctx->one_desc_off, (size_t)m * 32);
    }
    n_out[0] = n;
    return status;
}

int mslam_b200_dist_detect_batch(mslam_b200_dist* ctx, const uint8_t* const* frames, int n_frames, int width, int height,
                                 int stride_bytes, mslam_b200_keypoint* out_kp, uint8_t* out_desc, int capacity, int* n_out)
{
    if(!dist_batch_args_ok(ctx, frames, n_frames, width, stride_bytes, out_kp, out_desc, capacity, n_out) || ctx->pend[0].active ||
       ctx->pend[1].active)
        return MSLAM_B200_E_INVALID;
    if(n_frames == 1 && width >= 1 && height >= 1)
    {
        const int rc1 = dist_one_latency(ctx, frames[0], width, height, stride_bytes, out_kp, out_desc, capacity, n_out);
        if(rc1 != kDistGeneralPath)
            return rc1;
    }
    int rc = mslam_b200_dist_detect_batch_begin(ctx, frames, n_frames, width, height, stride_bytes, out_kp, out_desc, capacity, n_out);
    if(rc)
        return rc;
    rc = mslam_b200_dist_detect_batch_end(ctx);
    DIST_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    return rc;
}

int mslam_b200_dist_detect(mslam_b200_dist* ctx, const uint8_t* bgr, int width, int height, int stride_bytes,
                           mslam_b200_keypoint* out_kp, uint8_t* out_desc, int capacity, int* n_out)
{
    const uint8_t* frames[1] = {bgr};
    return mslam_b200_dist_detect_batch(ctx, frames, 1, width, height, stride_bytes, out_kp, out_desc, capacity, n_out);
}

int mslam_b200_dist_debug_read(mslam_b200_dist* ctx, int stage, int frame, int level, void* out, size_t bytes)
{
    if(!ctx || !out || frame < 0 || frame >= ctx->cfg.max_batch || level < 0 || ctx->geo_w == 0 || level >= ctx->P.nlevels)
        return MSLAM_B200_E_INVALID;
    DIST_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    DIST_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    const DistLevel& L = ctx->P.lv[level];
    if(stage == MSLAM_B200_DIST_PYRAMID || stage == MSLAM_B200_DIST_BLURRED)
    {
        if(bytes < (size_t)L.w * L.h)
            return MSLAM_B200_E_INVALID;
        const uint8_t* src = (stage == MSLAM_B200_DIST_PYRAMID ? ctx->d_pyr : ctx->d_blur) + L.img_off + frame * L.img_frame_stride;
        DIST_CUDA_OK(cudaMemcpy2D(out, L.w, src, L.pitch, L.w, L.h, cudaMemcpyDeviceToHost));
        return MSLAM_B200_OK;
    }
    if(stage >= MSLAM_B200_DIST_N_FAST && stage <= MSLAM_B200_DIST_PASSES)
    {
        if(bytes < sizeof(int))
            return MSLAM_B200_E_INVALID;
        DistCounters c;
        DIST_CUDA_OK(cudaMemcpy(&c, ctx->d_ctr + frame * ctx->P.nlevels + level, sizeof(c), cudaMemcpyDeviceToHost));
        *(int*)out = (int)(stage == MSLAM_B200_DIST_N_FAST ? c.n_fast : stage == MSLAM_B200_DIST_N_TREE ? c.n_tree : c.passes);
        return MSLAM_B200_OK;
    }
    return MSLAM_B200_E_INVALID;
}

int mslam_b200_dist_stage_times(mslam_b200_dist* ctx, int enable, double* ms6)
{
    if(!ctx)
        return MSLAM_B200_E_INVALID;
    DIST_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    if(ms6)
        for(int i = 0; i < 6; ++i)
            ms6[i] = ctx->t_ms[i];
    for(double& v : ctx->t_ms)
        v = 0;
    ctx->timing = enable != 0;
    return MSLAM_B200_OK;
}

} // extern "C"
