// orb_kernels.cu -- sm_100a kernels of the ORB detect + describe path.
//
// What each kernel must compute is OpenCV's cv::ORB::detectAndCompute as called by the reference
// (src/lib/modular_slam/orb_feature.cpp:25,40) after toGrayScale (src/lib/modular_slam/frame.cpp:6-27);
// the arithmetic contract is restated next to every kernel.
// All float stages use explicit round-to-nearest intrinsics so that nvcc can never contract them into
// FMAs (the reference path is non-FMA float except the 7x7 blur, which is an explicit FMA chain).
#include <stdlib.h>
#include <string.h>

#include <cuda.h> // CUtensorMap types only; the encoder is fetched through cudaGetDriverEntryPoint

#include "common.cuh"
#include "kernels.h"
#include "fast_simd.cuh"

namespace mslam_b200
{

// =================================================================================================
// K1  gray : interleaved 3-channel u8 -> level 0      (frame.cpp:15-19)
//   g = (u8) min(255.f, .299f*c0 + .587f*c1 + .114f*c2)  three rounded products, two rounded adds,
//   truncating cast; weights in MEMORY order.  HBM-bound: 3 B read + 1 B written per pixel.
// =================================================================================================
__device__ __forceinline__ int reflect101(int p, int n)
{
    if(n == 1)
        return 0;
    while(p < 0 || p >= n)
        p = p < 0 ? -p : 2 * n - 2 - p;
    return p;
}

__device__ __forceinline__ unsigned gray_px(unsigned c0, unsigned c1, unsigned c2)
{
    float v = __fadd_rn(__fadd_rn(__fmul_rn(0.299f, (float)c0), __fmul_rn(0.587f, (float)c1)),
                        __fmul_rn(0.114f, (float)c2));
    v = v < 255.f ? v : 255.f;
    return __float2uint_rz(v);
}

// exact u8 -> f32 without the XU pipe's I2F: 0x4B000000 | b is the float 8388608 + b
__device__ __forceinline__ float byte_to_float(unsigned word, int i)
{
    return __fsub_rn(__uint_as_float(__byte_perm(word, 0x4B000000u, 0x7440 + i)), 8388608.f);
}

// gray_px with the conversions kept off the XU pipe (I2F / F2I run at a quarter of the FMA rate and made the kernel
// XU-bound): bytes enter through PRMT + FSUB, and the truncating cast is the low byte of (v + 2^23) rounded toward zero.
__device__ __forceinline__ unsigned gray_px_fast(float c0, float c1, float c2)
{
    float v = __fadd_rn(__fadd_rn(__fmul_rn(0.299f, c0), __fmul_rn(0.587f, c1)), __fmul_rn(0.114f, c2));
    v = v < 255.f ? v : 255.f;
    return __float_as_uint(__fadd_rz(v, 8388608.f)); // bits 0..7 = (int)v for 0 <= v <= 255
}

// 16 pixels (48 B in, 16 B out) per thread with 128-bit accesses.  Requires W % 16 == 0.
__global__ void __launch_bounds__(256) k_gray16(const uint8_t* __restrict__ bgr, long long in_fs, uint8_t* __restrict__ out,
                                                long long out_fs, int pitch, int W, int H)
{
    const int cpr = W >> 4;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if(idx >= cpr * H)
        return;
    const int y = idx / cpr, c = idx - y * cpr;
    const uint4* src = reinterpret_cast<const uint4*>(bgr + (long long)blockIdx.y * in_fs + ((long long)y * W + c * 16) * 3);
    const uint4 a = __ldg(src), b = __ldg(src + 1), d = __ldg(src + 2);
    const unsigned w[12] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, d.x, d.y, d.z, d.w};
    unsigned o[4];
#pragma unroll
    for(int q = 0; q < 4; ++q)
    {
        unsigned g[4];
#pragma unroll
        for(int i = 0; i < 4; ++i)
        {
            const int k = 3 * (q * 4 + i);
            g[i] = gray_px_fast(byte_to_float(w[k >> 2], k & 3), byte_to_float(w[(k + 1) >> 2], (k + 1) & 3),
                                byte_to_float(w[(k + 2) >> 2], (k + 2) & 3));
        }
        o[q] = __byte_perm(__byte_perm(g[0], g[1], 0x0040), __byte_perm(g[2], g[3], 0x0040), 0x5410);
    }
    uint8_t* dst = out + (long long)blockIdx.y * out_fs + (long long)y * pitch + c * 16;
    *reinterpret_cast<uint4*>(dst) = make_uint4(o[0], o[1], o[2], o[3]);
    if(c == cpr - 1) // reflect-101 padding columns W .. W+3 = pixels W-2, W-3, W-4, W-5 (read by the blur's row pass)
        *reinterpret_cast<unsigned*>(dst + 16) = __byte_perm(o[3], o[2], 0x7012);
}

// generic width: one pixel per thread
__global__ void __launch_bounds__(256) k_gray1(const uint8_t* __restrict__ bgr, long long in_fs, uint8_t* __restrict__ out,
                                               long long out_fs, int pitch, int W, int H)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    if(x >= W)
        return;
    const uint8_t* p = bgr + (long long)blockIdx.z * in_fs + ((long long)y * W + x) * 3;
    const uint8_t g = (uint8_t)gray_px(p[0], p[1], p[2]);
    uint8_t* row = out + (long long)blockIdx.z * out_fs + (long long)y * pitch;
    row[x] = g;
    for(int px = W; px < W + 4; ++px) // reflect-101 padding columns (read by the blur's row pass)
        if(reflect101(px, W) == x)
            row[px] = g;
}

// =================================================================================================
// K2  resize : level l-1 -> level l, cv::resize INTER_LINEAR_EXACT (item 2)
//   per axis the host precomputes (ofs, alpha1) in IEEE double; horizontal pass in 8.8 fixed point,
//   vertical pass (a0*h0 + a1*h1 + 32768) >> 16.
//
//   k_resize_quads (the fast path): a lane owns 4 adjacent output columns and walks down a band of output rows.
//   Everything that depends on x only is decoded once per lane from one uint4 of the per-level quad table
//   (ResizeQuad): the source byte the quad's window starts at, two PRMT selectors that pull the (left, right) taps of
//   two pixels each out of the 8-byte window, and the four alpha1.  Per SOURCE row a lane then issues 3 aligned 32-bit
//   loads, 2 funnel shifts, 2 PRMT and 4 IDP.2A (h = a0*p[i] + a1*p[i+1], both taps in one instruction); consecutive
//   output rows share their source rows (1.2 new source rows per output row), so the horizontal pass is never
//   repeated.  Per OUTPUT row: 8 IMAD, 3 PRMT, one 32-bit store.  Against ~60 instructions per pixel for the
//   per-pixel table walk of k_resize_generic, which stays as the fallback for scale factors whose 4-pixel footprint
//   does not fit the 8-byte window.
// =================================================================================================
constexpr int RS_BAND_MAX = 16; // output rows per warp (chosen per launch: short bands when the level alone cannot fill the machine)

struct Row3
{
    unsigned a, b, c;
};
__device__ __forceinline__ Row3 resize_load(const uint8_t* __restrict__ row, int wi)
{
    const unsigned* p = reinterpret_cast<const unsigned*>(row) + wi;
    Row3 r;
    r.a = __ldg(p), r.b = __ldg(p + 1), r.c = __ldg(p + 2);
    return r;
}
__device__ __forceinline__ void resize_hrow(const Row3& W, int sh, unsigned selA, unsigned selB, unsigned w0, unsigned w1,
                                            unsigned w2, unsigned w3, unsigned (&h)[4])
{
    const unsigned S0 = __funnelshift_r(W.a, W.b, sh), S1 = __funnelshift_r(W.b, W.c, sh);
    const unsigned pA = __byte_perm(S0, S1, selA), pB = __byte_perm(S0, S1, selB);
    h[0] = __dp2a_lo(w0, pA, 0u);
    h[1] = __dp2a_hi(w1, pA, 0u);
    h[2] = __dp2a_lo(w2, pB, 0u);
    h[3] = __dp2a_hi(w3, pB, 0u);
}

// one warp, one band of output rows [yb, yb + rs_band) x 128 columns (quads xb * 32 .. + 31) of one frame.  LDG: the source may
// go through the read-only path (it was written by an earlier kernel); the fused tail kernel reads levels it wrote itself.
// Walks the band OUTPUT row by output row.  Output row y needs the horizontal pass of source rows n0(y) and n0(y) + 1 (the
// last one clamped); from one output row to the next n0 advances by 1 (the lower tap row of y becomes the upper one of y + 1,
// one new source row) or by 2 (two new rows) for every scale factor the quad table accepts, so the lane keeps the horizontal
// pass of the current tap pair (hP, hC), the raw words of the next two source rows (already in flight while this row is
// emitted), and does per output row: one or two horizontal passes, one or two row loads, the vertical blend, one store.  The
// per-row (ofs, alpha1) of the band sit in lanes 0..31 of one coalesced load and are broadcast with shuffles.  An advance
// outside {1, 2} (not produced by the host's tables today) reloads both tap rows directly.
// (Round 1 walked SOURCE rows with a four-row prefetch ring and an inner `while` that emitted whatever output rows had
// become ready: 30 instructions per pixel, 40 % of them loop bookkeeping -- profiles/r2_v2_*.)
template <bool LDG>
__device__ __forceinline__ void resize_band(const uint8_t* __restrict__ simg, int spitch, int sh_rows, uint8_t* __restrict__ dbase,
                                            int dpitch, int dh, int nquads, int xb, int yb, int rs_band,
                                            const uint4* __restrict__ qtab, const short2* __restrict__ ytab, int lane)
{
    const int quad_raw = xb * 32 + lane;
    const bool live = quad_raw < nquads;
    const int quad = live ? quad_raw : nquads - 1; // dead lanes shadow the last quad (they must stay for the shuffles)
    const uint4 q = __ldg(qtab + quad);
    const int wi = (int)(q.x >> 2), sh = (int)(q.x & 3u) * 8;
    const unsigned selA = q.y & 0xffffu, selB = q.y >> 16;
    const unsigned a10 = q.z & 0xffffu, a11 = q.z >> 16, a12 = q.w & 0xffffu, a13 = q.w >> 16;
    const unsigned w0 = (256u - a10) | (a10 << 16), w1 = (256u - a11) | (a11 << 16), w2 = (256u - a12) | (a12 << 16),
                   w3 = (256u - a13) | (a13 << 16);
    const int ye = min(yb + rs_band, dh);
    const short2 ty_lane = __ldg(ytab + min(yb + lane, dh - 1));
    const int last = sh_rows - 1;
    const uint8_t* srow = simg + 4 * (long long)wi; // byte address of the lane's first word in source row 0
    auto load = [&](int r) {
        const unsigned* p = reinterpret_cast<const unsigned*>(srow + (long long)min(r, last) * spitch);
        Row3 w;
        if(LDG)
            w.a = __ldg(p), w.b = __ldg(p + 1), w.c = __ldg(p + 2);
        else
            w.a = p[0], w.b = p[1], w.c = p[2];
        return w;
    };
    int n0 = __shfl_sync(0xffffffffu, (int)ty_lane.x, 0);
    unsigned hP[4], hC[4];
    {
        const Row3 t0 = load(n0), t1 = load(n0 + 1);
        resize_hrow(t0, sh, selA, selB, w0, w1, w2, w3, hP);
        resize_hrow(t1, sh, selA, selB, w0, w1, w2, w3, hC);
    }
    Row3 R0 = load(n0 + 2), R1 = load(n0 + 3); // the next two source rows, in flight
    uint8_t* drow = dbase + 4 * quad + (long long)yb * dpitch;
    for(int y = yb; y < ye; ++y, drow += dpitch)
    {
        const unsigned b1 = (unsigned)__shfl_sync(0xffffffffu, (int)ty_lane.y, y - yb);
        const unsigned b0 = 256u - b1;
        unsigned v[4];
#pragma unroll
        for(int j = 0; j < 4; ++j)
            v[j] = b0 * hP[j] + (b1 * hC[j] + 32768u); // result byte = bits 16..23
        const unsigned lo = __byte_perm(v[0], v[1], 0x0062), hi = __byte_perm(v[2], v[3], 0x0062);
        if(live)
            *reinterpret_cast<unsigned*>(drow) = __byte_perm(lo, hi, 0x5410);
        if(y + 1 < ye) // warp-uniform: set up the tap pair of the next output row
        {
            const int n1 = __shfl_sync(0xffffffffu, (int)ty_lane.x, y + 1 - yb);
            const int adv = n1 - n0;
            n0 = n1;
            if(adv == 1)
            {
#pragma unroll
                for(int j = 0; j < 4; ++j)
                    hP[j] = hC[j];
                resize_hrow(R0, sh, selA, selB, w0, w1, w2, w3, hC);
                R0 = R1;
                R1 = load(n0 + 3);
            }
            else if(adv == 2)
            {
                resize_hrow(R0, sh, selA, selB, w0, w1, w2, w3, hP);
                resize_hrow(R1, sh, selA, selB, w0, w1, w2, w3, hC);
                R0 = load(n0 + 2);
                R1 = load(n0 + 3);
            }
            else if(adv != 0)
            {
                const Row3 t0 = load(n0), t1 = load(n0 + 1);
                resize_hrow(t0, sh, selA, selB, w0, w1, w2, w3, hP);
                resize_hrow(t1, sh, selA, selB, w0, w1, w2, w3, hC);
                R0 = load(n0 + 2);
                R1 = load(n0 + 3);
            }
        }
    }
}

__global__ void __launch_bounds__(128) k_resize_quads(const uint8_t* __restrict__ src, int spitch, long long sfs, int sh_rows,
                                                      uint8_t* __restrict__ dst, int dpitch, long long dfs, int dh, int nquads,
                                                      int xblocks, int rs_band, const uint4* __restrict__ qtab,
                                                      const short2* __restrict__ ytab)
{
    static_assert(RS_BAND_MAX <= 32, "one lane per output row of the band");
    const int lane = threadIdx.x & 31;
    const int wg = blockIdx.x * 4 + (threadIdx.x >> 5);
    const int band = wg / xblocks, xb = wg - band * xblocks;
    const int yb = band * rs_band;
    if(yb >= dh) // warp-uniform
        return;
    resize_band<true>(src + (long long)blockIdx.y * sfs, spitch, sh_rows, dst + (long long)blockIdx.y * dfs, dpitch, dh, nquads, xb, yb,
                      rs_band, qtab, ytab, lane);
}

// The small levels of the chain in ONE launch: a CLUSTER of RT_CLUSTER CTAs owns a frame and walks levels [l_begin, nlevels),
// every warp of the cluster taking bands of the current level, a cluster barrier (release / acquire at cluster scope: the level
// just written is visible to the other CTAs' plain loads) between levels.  A level of ~70k pixels and less is a
// 10-microsecond launch that cannot fill the machine, and the chain pays one such launch per level; here the frames of the
// batch run side by side.  (One CTA per frame -- 64 busy SMs of 148 -- was measured slower than the four launches it replaced.)
constexpr int RT_THREADS = 512, RT_BAND = 4, RT_CLUSTER = 4;
__global__ void __cluster_dims__(RT_CLUSTER, 1, 1) __launch_bounds__(RT_THREADS, 2)
    k_resize_tail(const __grid_constant__ PipeParams P, int l_begin, int rt_band, int rt_adapt, uint8_t* __restrict__ pyr,
                  const uint4* __restrict__ qtab, const short2* __restrict__ ytab)
{
    const int lane = threadIdx.x & 31;
    const int crank = blockIdx.x % RT_CLUSTER;
    const int worker = crank * (RT_THREADS / 32) + (threadIdx.x >> 5), n_workers = RT_CLUSTER * (RT_THREADS / 32);
    const int b = blockIdx.x / RT_CLUSTER + P.frame0;
    for(int l = l_begin; l < P.nlevels; ++l)
    {
        const LevelDesc &S = P.lv[l - 1], &D = P.lv[l];
        const uint8_t* simg = pyr + S.img_off + (long long)b * S.img_frame_stride;
        uint8_t* dimg = pyr + D.img_off + (long long)b * D.img_frame_stride;
        // Band height per level: a warp's walk is a chain (row after row), so a level costs (rounds of tasks) x (rows per band).
        // The shortest band that gives every warp of the cluster at most ONE task wastes the least of that product (VGA levels
        // 2..7 with 64 warps: 21, 14, 11, 10, 6, 5 rows = 67 per warp against 80 with four-row bands in 6, 4, 3, 3, 2, 2 rounds);
        // rt_band remains the fallback for a level too large for one round of bands of at most 32 rows.
        const int xblocks = (D.rs_quads + 31) / 32;
        int bh = rt_band;
        if(rt_adapt && xblocks <= n_workers)
        {
            const int one_round = (D.h + n_workers / xblocks - 1) / (n_workers / xblocks);
            if(one_round <= 32)
                bh = max(one_round, 2);
        }
        const int bands = (D.h + bh - 1) / bh;
        for(int task = worker; task < xblocks * bands; task += n_workers)
        {
            const int band = task / xblocks, xb = task - band * xblocks;
            if(l == l_begin)
                resize_band<true>(simg, S.pitch, S.h, dimg, D.pitch, D.h, D.rs_quads, xb, band * bh, bh, qtab + D.qtab_off,
                                  ytab + D.ytab_off, lane);
            else
                resize_band<false>(simg, S.pitch, S.h, dimg, D.pitch, D.h, D.rs_quads, xb, band * bh, bh, qtab + D.qtab_off,
                                   ytab + D.ytab_off, lane);
        }
        if(l + 1 < P.nlevels)
        {
            asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
            asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
        }
    }
}

// fallback: 4 output pixels per thread, per-pixel table walk
__global__ void __launch_bounds__(256) k_resize_generic(const uint8_t* __restrict__ src, int spitch, long long sfs, int sw, int sh,
                                                        uint8_t* __restrict__ dst, int dpitch, long long dfs, int dw, int dh,
                                                        const short2* __restrict__ xtab, const short2* __restrict__ ytab)
{
    const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if(x4 >= dw + 4 || y >= dh) // columns dw .. dw+3 are the reflect-101 padding the blur's row pass reads
        return;
    const short2 ty = __ldg(ytab + y);
    const int y0 = ty.x, y1 = min(y0 + 1, sh - 1);
    const unsigned b1 = (unsigned)ty.y, b0 = 256u - b1;
    const uint8_t* r0 = src + (long long)blockIdx.z * sfs + (long long)y0 * spitch;
    const uint8_t* r1 = src + (long long)blockIdx.z * sfs + (long long)y1 * spitch;
    unsigned outw = 0;
#pragma unroll
    for(int i = 0; i < 4; ++i)
    {
        const int x = x4 + i;
        if(x < dw + 4)
        {
            const short2 tx = __ldg(xtab + (x < dw ? x : reflect101(x, dw)));
            const int i0 = tx.x, i1 = min(i0 + 1, sw - 1);
            const unsigned a1 = (unsigned)tx.y, a0 = 256u - a1;
            const unsigned h0 = a0 * __ldg(r0 + i0) + a1 * __ldg(r0 + i1);
            const unsigned h1 = a0 * __ldg(r1 + i0) + a1 * __ldg(r1 + i1);
            outw |= ((b0 * h0 + b1 * h1 + 32768u) >> 16) << (8 * i);
        }
    }
    *reinterpret_cast<unsigned*>(dst + (long long)blockIdx.z * dfs + (long long)y * dpitch + x4) = outw;
}

// =================================================================================================
// K7' blur : per level 7x7 sigma=2 separable float blur, BORDER_REFLECT_101 (item 9)
//   row pass  acc = k0*p[x-3]; acc = fma(k_j, p[x-3+j], acc)  j = 1..6
//   col pass  acc = k3*r[y];   acc = fma(k_{3+i}, r[y+i] + r[y-i], acc)  i = 1..3 ; out = sat_u8(rint(acc))
// =================================================================================================
constexpr int BT_W = 128; // one warp: 128 columns (4 per lane) x PipeParams::blur_band rows

// the three aligned words covering source pixels x-4 .. x+7 of one row.  The producers pad every pyramid row with its
// reflect-101 columns w .. w+3, so only the left border needs care: for x = 0 pixels -3, -2, -1 are pixels 3, 2, 1.
struct Row12
{
    unsigned a, b, c;
};
__device__ __forceinline__ Row12 blur_load(const uint8_t* __restrict__ row, int x)
{
    Row12 r;
    const unsigned* wp = reinterpret_cast<const unsigned*>(row + x);
    r.b = __ldg(wp), r.c = __ldg(wp + 1);
    r.a = __ldg(wp - (x > 0)); // x = 0: placeholder, blur_row4 reflects pixels 3, 2, 1 out of r.b when it consumes the row
    return r;                  // (nothing may depend on a loaded value here: the loads are prefetches)
}

// row pass for the 4 pixels x .. x+3:  acc = k0 * p[x-3]; acc = fma(k_j, p[x-3+j], acc)
__device__ __forceinline__ void blur_row4(const Row12& wd_in, bool left_edge, float (&o)[4])
{
    Row12 wd = wd_in;
    if(left_edge)
        wd.a = __byte_perm(wd.b, 0u, 0x1230);
    const float kk[7] = {__int_as_float(0x3d8fafb1), __int_as_float(0x3e06387e), __int_as_float(0x3e434a39),
                         __int_as_float(0x3e5d4ae0), __int_as_float(0x3e434a39), __int_as_float(0x3e06387e),
                         __int_as_float(0x3d8fafb1)};
    const float p[10] = {byte_to_float(wd.a, 1), byte_to_float(wd.a, 2), byte_to_float(wd.a, 3), byte_to_float(wd.b, 0),
                         byte_to_float(wd.b, 1), byte_to_float(wd.b, 2), byte_to_float(wd.b, 3), byte_to_float(wd.c, 0),
                         byte_to_float(wd.c, 1), byte_to_float(wd.c, 2)};
#pragma unroll
    for(int q = 0; q < 4; ++q)
    {
        float acc = __fmul_rn(kk[0], p[q]);
#pragma unroll
        for(int j = 1; j < 7; ++j)
            acc = __fmaf_rn(kk[j], p[q + j], acc);
        o[q] = acc;
    }
}

// column pass + store of one output row: ring slot (k + j) % 7 holds the row pass of source row y - 3 + j
__device__ __forceinline__ void blur_col_store(const float (&r)[7][4], const int k, uint8_t* __restrict__ dst_px)
{
    const float k1 = __int_as_float(0x3e434a39), k2 = __int_as_float(0x3e06387e), k3 = __int_as_float(0x3d8fafb1),
                kc = __int_as_float(0x3e5d4ae0); // kernel[3 + 1], [3 + 2], [3 + 3], [3]
    unsigned bits[4];
#pragma unroll
    for(int q = 0; q < 4; ++q)
    {
        float acc = __fmul_rn(kc, r[(k + 3) % 7][q]);
        acc = __fmaf_rn(k1, __fadd_rn(r[(k + 4) % 7][q], r[(k + 2) % 7][q]), acc);
        acc = __fmaf_rn(k2, __fadd_rn(r[(k + 5) % 7][q], r[(k + 1) % 7][q]), acc);
        acc = __fmaf_rn(k3, __fadd_rn(r[(k + 6) % 7][q], r[k % 7][q]), acc);
        // 0 <= acc < 256 (weights sum to 1 within rounding): the low byte of acc + 1.5 * 2^23 is rint(acc)
        bits[q] = __float_as_uint(__fadd_rn(acc, 12582912.f));
    }
    const unsigned lo = __byte_perm(bits[0], bits[1], 0x0040), hi = __byte_perm(bits[2], bits[3], 0x0040);
    *reinterpret_cast<unsigned*>(dst_px) = __byte_perm(lo, hi, 0x5410);
}

// A lane owns a 4-pixel-wide column strip and walks down a band of BT_H rows: the row pass of each new source row goes
// into a 7-deep register ring, the column pass reads the ring -- no shared memory, no barriers, every source byte is
// read once per band (+6 halo rows).  Source rows sit in a 7-slot register buffer that is refilled right after a slot
// is consumed, i.e. they are fetched 7 rows ahead of their use, so the walk is not bound by memory latency.
// One launch covers every level of every frame: warp w of the grid is (level, column block, band) through the
// per-level warp offsets in LevelDesc, so that small levels do not leave the machine idle.
// Rounding is the 1.5 * 2^23 magic add (round-to-nearest-even, like cvRound / saturate_cast<uchar>(float)).
__global__ void __launch_bounds__(128, 6) k_blur(const __grid_constant__ PipeParams P, int l_begin, int l_end,
                                                 const uint8_t* __restrict__ src, uint8_t* __restrict__ dst)
{
    const int lane = threadIdx.x & 31;
    int wg = blockIdx.x * 4 + (threadIdx.x >> 5) + P.lv[l_begin].blur_warp_base;
    int l = l_begin;
    while(l + 1 < l_end && wg >= P.lv[l + 1].blur_warp_base)
        ++l;
    const LevelDesc& L = P.lv[l];
    wg -= L.blur_warp_base;
    const int band = wg / L.blur_xblocks, xb = wg - band * L.blur_xblocks;
    const int w = L.w, h = L.h, pitch = L.pitch;
    const int x = (xb * 32 + lane) * 4;
    const int y0 = band * P.blur_band;
    if(x >= w || y0 >= h)
        return;
    const long long plane = L.img_off + (long long)(blockIdx.y + P.frame0) * L.img_frame_stride;
    const uint8_t* img = src + plane;
    uint8_t* out = dst + plane;
    const int rows = min(P.blur_band, h - y0);
    float r[7][4];
    Row12 buf[7];
    {
        Row12 pre[6];
#pragma unroll
        for(int s = 0; s < 6; ++s)
            pre[s] = blur_load(img + (long long)reflect101(y0 - 3 + s, h) * pitch, x);
#pragma unroll
        for(int k = 0; k < 7; ++k)
            buf[k] = blur_load(img + (long long)reflect101(y0 + 3 + k, h) * pitch, x);
#pragma unroll
        for(int s = 0; s < 6; ++s)
            blur_row4(pre[s], x == 0, r[s]);
    }
    if(rows % 7 == 0 && y0 + rows + 2 < h) // warp-uniform
    {
        // A whole band whose look-ahead rows (up to y0 + rows + 2) stay inside the image -- every band but the last one or two
        // of a level: no reflection and no row guards in the walk, the source / destination addresses advance by one pitch
        // per row, blocks of seven rows without a branch (the last block fetches nothing).  The general walk below spends a
        // fifth of its instructions on those guards and on 64-bit row addresses.
        const unsigned* ldw = reinterpret_cast<const unsigned*>(img + (long long)(y0 + 10) * pitch + x);
        const unsigned* lda = ldw - (x > 0);
        uint8_t* stp = out + (long long)y0 * pitch + x;
        const int pw = pitch >> 2;
        for(int i0 = 0; i0 < rows; i0 += 7)
        {
            const bool fetch = i0 + 7 < rows;
#pragma unroll
            for(int k = 0; k < 7; ++k)
            {
                blur_row4(buf[k], x == 0, r[(k + 6) % 7]);
                if(fetch)
                {
                    buf[k].b = __ldg(ldw), buf[k].c = __ldg(ldw + 1), buf[k].a = __ldg(lda);
                    ldw += pw, lda += pw;
                }
                blur_col_store(r, k, stp);
                stp += pitch;
            }
        }
        return;
    }
    for(int i0 = 0; i0 < rows; i0 += 7)
    {
#pragma unroll
        for(int k = 0; k < 7; ++k)
        {
            const int i = i0 + k;
            if(i < rows)
            {
                blur_row4(buf[k], x == 0, r[(k + 6) % 7]);
                if(i + 7 < rows)
                    buf[k] = blur_load(img + (long long)reflect101(y0 + 3 + i + 7, h) * pitch, x);
                blur_col_store(r, k, out + (long long)(y0 + i) * pitch + x);
            }
        }
    }
}

// =================================================================================================
// K3  FAST-9/16 + 3x3 strict NMS + border filter (items 3, 4), any range of levels in one launch.
//
//   One CTA scores a 128 x 32 block of positions (assigned interior 124 x 30, the rest is the 1-px halo the NMS
//   needs plus 4-byte alignment), pixels staged once in shared memory (136 x 38).
//   Phase A, corner classification, is branch-free byte-SIMD: a thread owns 4 horizontally adjacent centres x 4
//   rows.  For each of the 16 ring offsets the 4 ring pixels are one PRMT of two staged words; "ring > centre + t"
//   and "ring < centre - t" are the carry-outs of byte-wise additions, evaluated for 4 pixels at once with one
//   7-bit add and one majority LOP3 each (bit 7 of every byte carries the answer).  The "9 contiguous" test is
//   bit-sliced over the 16 masks: a3_k = m_k & m_k+1 & m_k+2, a9_k = a3_k & a3_k+3 & a3_k+6, any = OR a9_k -- 40 LOP3
//   per side per 4 pixels, no divergence however dense the corners are.
//   Phase B scores only the corners, compacted CTA-wide into a queue: cornerScore = max over the 16 arcs of the arc
//   minimum of |centre - ring| on the bright or the dark side, minus 1.  Both sides run through ONE packed u16x2
//   max/min network (lane 0 = ring value, lane 1 = 255 - ring value; VIMNMX.U16x2), with the arc maxima built from
//   block prefix/suffix maxima (28 + 16 ops) instead of 16 independent 9-element reductions.
//   Phase C walks the same queue for the strict 3x3 NMS and the border filter; survivors go to the level's
//   candidate list (one global atomic per CTA) and its 256-bin score histogram (first retainBest cut).
// =================================================================================================
constexpr int FS_SW = 128, FS_SH_MAX = 64;      // scored positions per CTA: 128 x (32 | 64), LevelDesc::tile_sh rows
constexpr int FS_IW = FS_SW - 4;                // assigned interior: 124 x (tile_sh - 2)
// staged pixels: 136 columns are used.  The innermost coordinate of a tiled TMA load must be a multiple of 16 BYTES (measured:
// tools/tma_probe.cu -- x0 = 24 raises cudaErrorIllegalInstruction, 32 is fine), and a tile starts at column 24 + 124 tx, so
// the box starts at that rounded down to 16 (up to 12 columns early) and is 160 wide
constexpr int FS_PW = FS_SW + 32, FS_PH_MAX = FS_SH_MAX + 6;
constexpr int FS_PWW = FS_PW / 4;               // staged words per row (40)

__device__ __forceinline__ unsigned fast_smem_u32(const void* p)
{
    return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void fast_mbar_init(unsigned bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
// bounded wait: a protocol error becomes a trap (a CUDA error on the host), never a hung GPU
__device__ __forceinline__ void fast_mbar_wait(unsigned bar, unsigned parity, int dbg = 0)
{
    for(unsigned spins = 0;; ++spins)
    {
        unsigned ok;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok)
                     : "r"(bar), "r"(parity)
                     : "memory");
        if(ok)
            return;
        if(spins > (1u << 24))
        {
            if(dbg & 1)
                return; // debugging aid (MSLAM_FAST_DBG=1): carry on with whatever the tile holds instead of trapping
            __trap();
        }
    }
}

#ifndef MSLAM_FAST_CTAS
#define MSLAM_FAST_CTAS 4 // resident CTAs per SM the register allocation aims for (A/B aid: -DMSLAM_FAST_CTAS=5 caps at 48 registers)
#endif
__global__ void __launch_bounds__(256, MSLAM_FAST_CTAS) k_fast(const __grid_constant__ PipeParams P, const __grid_constant__ FastMaps M, const FastMaps* __restrict__ gmaps, int dbg, int l_begin, int l_end,
                                              const uint8_t* __restrict__ pyr, unsigned* __restrict__ cand,
                                              LevelCounters* __restrict__ ctr, unsigned* __restrict__ hist)
{
    __shared__ __align__(128) uint8_t s_px[FS_PH_MAX][FS_PW];
    __shared__ __align__(8) unsigned long long s_bar;
    __shared__ __align__(16) uint8_t s_sc[FS_SH_MAX][FS_SW];
    __shared__ unsigned short s_q[FS_SH_MAX * FS_SW];
    __shared__ unsigned s_out[FS_IW * (FS_SH_MAX - 2) / 4 + 64];
    __shared__ int s_qn, s_on;
    __shared__ unsigned s_base;

    // which level does this tile belong to?
    int l = l_begin;
    int tile = (int)blockIdx.x + P.lv[l_begin].tile_base;
    while(l + 1 < l_end && tile >= P.lv[l + 1].tile_base)
        ++l;
    const LevelDesc& L = P.lv[l];
    tile -= L.tile_base;
    const int ty = tile / L.tiles_x, tx = tile - ty * L.tiles_x;
    const int b = blockIdx.y + P.frame0;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int edge = P.edge, thr = P.fast_thr;
    const int sx0 = ((edge - 1) & ~3) + FS_IW * tx; // first scored column (multiple of 4)
    const int SH = L.tile_sh, PH = SH + 6;         // scored / staged rows of this level's tiles
    const int sy0 = edge - 1 + (SH - 2) * ty;       // first scored row
    const int px0 = sx0 - 4, py0 = sy0 - 3;         // first pixel the tile needs
    const int px0a = px0 & ~15, xo = px0 - px0a;    // first staged column (16-byte aligned for the TMA) and where px0 sits in it

    // ---- tile staging: ONE 3-D bulk-tensor (TMA) load per CTA -- box = 160 x PH x 1 of the level's (pitch, h, frame) plane
    // stack, zero-filled past the image / the row pitch (what the per-word guards of the earlier version did by hand) ----
    if(tid == 0)
    {
        fast_mbar_init(fast_smem_u32(&s_bar), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fast_smem_u32(&s_bar)), "r"(PH * FS_PW) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(fast_smem_u32(&s_px[0][0])), "l"(gmaps ? reinterpret_cast<const void*>(&gmaps->m[l][0]) : reinterpret_cast<const void*>(&M.m[l][0])), "r"(px0a), "r"(py0), "r"(b),
                     "r"(fast_smem_u32(&s_bar))
                     : "memory");
    }
    if(tid == 0)
    {
        s_qn = 0;
        s_on = 0;
    }
    __syncthreads(); // also publishes the barrier's initialisation to the waiting threads
    fast_mbar_wait(fast_smem_u32(&s_bar), 0, dbg);

    // ---- phase A: classify 4 rows x 4 pixels per thread and pass (32 rows per pass) ----
    // A tile of the level's last column may hold only a few scored columns (tiles are 128 wide whatever the level's width).  When
    // the last column that matters lies in the tile's left half, a warp covers TWO row groups of 16 column groups each instead of
    // one of 32, and the tile takes half the passes (VGA: the last tile columns of levels 2, 3 and 6).
    const bool narrow = L.w - edge - sx0 < 62; // CTA-uniform: every scored or NMS-halo column has sx <= 61, its ring sx <= 64
    const int clane = narrow ? (lane & 15) : lane;            // column group of this thread
    const int rows_per_pass = narrow ? 64 : 32;
#pragma unroll 1
    for(int pass = 0; pass * rows_per_pass < SH; ++pass)
    {
    const int rb = rows_per_pass * pass + 4 * (narrow ? 2 * warp + (lane >> 4) : warp); // first scored row of this thread's group
    unsigned flags16 = 0; // bit (4 * row + j): pixel j of the group in row rb + row is a corner
    // Rows below the level's last scored row (y > h - edge) exist only because tiles have a fixed height: a warp whose rows
    // all lie there (or past the tile) has nothing to classify, nothing to queue, and nobody reads its part of the score map (the
    // NMS of an owned position looks at most one row past the last owned row, which is y = h - edge).  Over the eight levels of
    // a VGA frame that is 9 % of the tile rows.
    const bool rows_live = rb < SH && L.h - edge - (sy0 + rb) + 1 > 0;
    if(!__any_sync(0xffffffffu, rows_live))
        continue;
    if(rows_live)
    {
        const unsigned t4 = (unsigned)thr * 0x01010101u;
        const unsigned one = (unsigned)P.one, mone = (unsigned)P.minus_one;
        const unsigned* wbase = reinterpret_cast<const unsigned*>(&s_px[rb][0]) + clane + (xo >> 2);
        unsigned wm[10], w0[10], wp[10]; // staged rows 4 * warp .. 4 * warp + 9, words lane, lane + 1, lane + 2
#pragma unroll
        for(int r = 0; r < 10; ++r)
        {
            wm[r] = wbase[r * FS_PWW];
            w0[r] = wbase[r * FS_PWW + 1];
            wp[r] = wbase[r * FS_PWW + 2];
        }
#pragma unroll
        for(int row = 0; row < 4; ++row)
        {
            const unsigned C = w0[row + 3];
            const unsigned H = __vaddus4(C, t4), Lo = __vsubus4(C, t4);
            const unsigned nH7 = ~H & 0x7f7f7f7fu;
            const unsigned K = (Lo & 0x7f7f7f7fu) + 0x7f7f7f7fu;
            unsigned mb[16], md[16];
            // ring offsets (dx, dy), k = 0..15, as in cv::FAST's 16-pixel circle
            constexpr int DX[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
            constexpr int DY[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
#pragma unroll
            for(int k = 0; k < 16; ++k)
            {
                const int r = row + 3 + DY[k];
                const unsigned R = ring4(wm[r], w0[r], wp[r], DX[k]);
                const unsigned r7 = R & 0x7f7f7f7fu;
                // the two byte-lane additions go through IMAD with multipliers the compiler cannot see (+1 / -1 from the
                // parameter block): they issue on the FMA pipe, which idles while the ALU pipe is the one that binds here
                mb[k] = lop3_maj_a_nb_c(R, H, r7 * one + nH7);  // bit 7: R > centre + t
                md[k] = lop3_maj_a_nb_c(Lo, R, r7 * mone + K);  // bit 7: R < centre - t
            }
            const unsigned fb = any_run9(mb), f = (fb | any_run9(md)) & 0x80808080u;
            // gather the four bit-7 flags into a nibble
            flags16 |= (((f >> 7) * 0x00204081u) >> 21 & 0xfu) << (4 * row);
            // "the arc is brighter than the centre" goes into the score map as 0 / 1 per position -- one word store per row,
            // which also is the map's initialisation (every position of the tile is classified by exactly one thread);
            // phase B reads the flag of a corner before it overwrites it with the score (>= fast_thr - 1 > 1)
            reinterpret_cast<unsigned*>(&s_sc[rb + row][0])[clane] = (fb >> 7) & 0x01010101u;
        }
    }
    // only positions that can matter: inside the image interior +- 1 (the NMS halo):
    //   edge - 1 <= x <= w - edge (a nibble, the same for the 4 rows) and y <= h - edge (y >= edge - 1 by construction)
    {
        const int x = sx0 + 4 * clane;
        const int jlo = max(0, edge - 1 - x), jhi = min(3, L.w - edge - x);
        const unsigned nib = jlo <= jhi ? ((0xfu >> (3 - jhi)) & (0xfu << jlo) & 0xfu) : 0u;
        const int nrows = rows_live ? min(4, max(0, L.h - edge - (sy0 + rb) + 1)) : 0;
        flags16 &= (nib * 0x1111u) & ((1u << (4 * nrows)) - 1u);
    }
    // warp-aggregated append to the CTA's corner queue
    {
        const int cnt = __popc(flags16);
        int incl = cnt;
#pragma unroll
        for(int o = 1; o < 32; o <<= 1)
        {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if(lane >= o)
                incl += t;
        }
        int wbase_q = 0;
        if(lane == 31)
            wbase_q = atomicAdd(&s_qn, incl);
        wbase_q = __shfl_sync(0xffffffffu, wbase_q, 31);
        // The loop runs as often as the busiest lane of the warp has corners, so its body is kept to the bone: highest set bit
        // (one FLO), entry = position of the group | (bit & 3) | (bit >> 2) << 7 -- the group's position has zeros in those four
        // bits (4 * lane and a row that is a multiple of 4), and bit * 33 puts both fields in place with one IMAD.
        unsigned short* q = s_q + (wbase_q + incl - cnt);
        const unsigned ebase = (unsigned)(rb * FS_SW + 4 * clane);
        unsigned f = flags16;
        while(f)
        {
            unsigned bit; // index of the highest set bit: one FLO (31 - __clz(f) compiles to three instructions)
            asm("bfind.u32 %0, %1;" : "=r"(bit) : "r"(f));
            f ^= 1u << bit;
            *q++ = (unsigned short)(ebase | ((bit * 33u) & 0x183u));
        }
    }
    } // passes
    __syncthreads();

    // ---- phase B: exact cornerScore of the queued corners ----
    const int nq = s_qn;
    for(int j = 2 * tid; j < nq; j += 512)
    {
        const int ia = s_q[j], ib = s_q[min(j + 1, nq - 1)];
        const int ya = ia / FS_SW, xa = ia - ya * FS_SW, yb = ib / FS_SW, xb = ib - yb * FS_SW;
        const unsigned sc = corner_score2<FS_PW>(&s_px[ya + 3][xa + 4 + xo], &s_px[yb + 3][xb + 4 + xo], s_sc[ya][xa], s_sc[yb][xb]);
        s_sc[ya][xa] = (uint8_t)(sc & 0xffu);
        s_sc[yb][xb] = (uint8_t)(sc >> 16);
    }
    __syncthreads();

    // ---- phase C: 3x3 strict NMS + border filter over the queue ----
    const int own_x0 = max(2, edge - sx0), own_x1 = min(FS_SW - 3, L.w - edge - 1 - sx0), own_y1 = min(SH - 2, L.h - edge - 1 - sy0);
    for(int j0 = 0; j0 < nq; j0 += 256)
    {
        const int j = j0 + tid;
        bool keep = false;
        unsigned rec = 0;
        if(j < nq)
        {
            const int i = s_q[j];
            const int sy = i / FS_SW, sx = i - sy * FS_SW;
            const int x = sx0 + sx, y = sy0 + sy;
            // assigned interior of the tile and of the level's border (sy >= 1 already means y >= edge), as CTA-wide bounds
            if(sx >= own_x0 && sx <= own_x1 && sy >= 1 && sy <= own_y1)
            {
                const uint8_t* s = &s_sc[sy][sx];
                // all eight neighbours are requested before the first comparison (no short-circuit chain of dependent
                // shared-memory loads); strict maximum <=> v > max of the neighbours
                const int v = s[0];
                const int n0 = s[-FS_SW - 1], n1 = s[-FS_SW], n2 = s[-FS_SW + 1], n3 = s[-1], n4 = s[1], n5 = s[FS_SW - 1],
                          n6 = s[FS_SW], n7 = s[FS_SW + 1];
                keep = v > max(max(max(n0, n1), max(n2, n3)), max(max(n4, n5), max(n6, n7)));
                rec = ((unsigned)y << 20) | ((unsigned)x << 8) | (unsigned)v;
            }
        }
        const unsigned bal = __ballot_sync(0xffffffffu, keep);
        if(bal)
        {
            int wb = 0;
            if(lane == 0)
                wb = atomicAdd(&s_on, __popc(bal));
            wb = __shfl_sync(0xffffffffu, wb, 0);
            if(keep)
                s_out[wb + __popc(bal & ((1u << lane) - 1u))] = rec;
        }
    }
    __syncthreads();
    const int n_out = s_on;
    if(n_out == 0)
        return;
    LevelCounters* C = ctr + (long long)b * P.nlevels + l;
    if(tid == 0)
        s_base = atomicAdd(&C->n_fast, (unsigned)n_out);
    __syncthreads();
    unsigned* list = cand + (long long)b * P.cand_frame_stride + L.cand_off;
    unsigned* hh = hist + ((long long)b * P.nlevels + l) * 256;
    for(int j = tid; j < n_out; j += 256)
    {
        const unsigned rec = s_out[j];
        const unsigned idx = s_base + j;
        if(idx < (unsigned)L.cand_cap)
            list[idx] = rec;
        atomicAdd(hh + (rec & 255u), 1u);
    }
}

// =================================================================================================
// K4  select : one CTA per (frame, level)  (items 5, 6)
//   retainBest(2 n_l) on the integer FAST score (exact k-th largest from the histogram, ties kept)
//   -> Harris response of the survivors (7x7 block of 3x3 Sobel sums in int32; non-FMA f32 formula)
//   -> retainBest(n_l) on the response (exact k-th largest by 4-pass radix select on the ordered
//      float bits, ties kept) -> survivors written in canonical (y, x) order.
// =================================================================================================
__device__ __forceinline__ unsigned fkey(float f)
{
    const unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float funkey(unsigned k)
{
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

__device__ float harris_response(const uint8_t* __restrict__ img, int pitch, int x, int y)
{
    // rows y-4 .. y+4, columns x-4 .. x+4.  All 27 aligned words of the window are requested before the first one is
    // used (one exposed memory latency per keypoint instead of one per row), then a three-row sliding window.
    const int off8 = ((x - 4) & 3) * 8;
    const unsigned* p = reinterpret_cast<const unsigned*>(img + (long long)(y - 4) * pitch + ((x - 4) & ~3));
    const int pw = pitch >> 2; // pitch is a multiple of 64
    unsigned w[9][3];
#pragma unroll
    for(int r = 0; r < 9; ++r)
    {
        w[r][0] = __ldg(p + r * pw);
        w[r][1] = __ldg(p + r * pw + 1);
        w[r][2] = __ldg(p + r * pw + 2);
    }
    int a = 0, b = 0, c = 0;
    int r0[9], r1[9], r2[9];
    auto unpack = [&](int r, int (&o)[9]) {
        const unsigned v0 = __funnelshift_r(w[r][0], w[r][1], off8), v1 = __funnelshift_r(w[r][1], w[r][2], off8);
#pragma unroll
        for(int j = 0; j < 4; ++j)
        {
            o[j] = (int)((v0 >> (8 * j)) & 255u);
            o[4 + j] = (int)((v1 >> (8 * j)) & 255u);
        }
        o[8] = (int)((w[r][2] >> off8) & 255u);
    };
    unpack(0, r0);
    unpack(1, r1);
#pragma unroll
    for(int row = 0; row < 7; ++row)
    {
        unpack(row + 2, r2);
#pragma unroll
        for(int j = 1; j < 8; ++j)
        {
            const int Ix = (r1[j + 1] - r1[j - 1]) * 2 + (r0[j + 1] - r0[j - 1]) + (r2[j + 1] - r2[j - 1]);
            const int Iy = (r2[j] - r0[j]) * 2 + (r2[j - 1] - r0[j - 1]) + (r2[j + 1] - r0[j + 1]);
            a += Ix * Ix;
            b += Iy * Iy;
            c += Ix * Iy;
        }
#pragma unroll
        for(int j = 0; j < 9; ++j)
        {
            r0[j] = r1[j];
            r1[j] = r2[j];
        }
    }
    const float s = __fdiv_rn(1.f, __fmul_rn(28.f, 255.f)); // 1.f/((1 << 2) * 7 * 255.f)
    const float s4 = __fmul_rn(__fmul_rn(__fmul_rn(s, s), s), s);
    const float fa = (float)a, fb = (float)b, fc = (float)c;
    const float sum = __fadd_rn(fa, fb);
    const float t = __fsub_rn(__fsub_rn(__fmul_rn(fa, fb), __fmul_rn(fc, fc)), __fmul_rn(__fmul_rn(0.04f, sum), sum));
    return __fmul_rn(t, s4);
}

constexpr int SEL_SMEM = 2048; // the survivor lists of both cuts live in shared memory up to this many entries (global beyond)

// Block-wide "k-th largest bin" over a 256-bin histogram, one bin per thread (bin = tid):
//   S[v] = sum of bins >= v.  The unique v with S[v] >= k > S[v + 1] is written to *s_thr and k - S[v + 1] to *s_k.
// Requires 1 <= k <= S[0].  Two barriers; replaces a 256-step serial walk by one thread.  The CTA may have more than
// 256 threads: the extra warps pass hval = 0 and only meet the barriers.
__device__ __forceinline__ void suffix_select(unsigned hval, unsigned k, unsigned* s_w /* 8 */, unsigned* s_thr, unsigned* s_k)
{
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    unsigned sfx = hval;
#pragma unroll
    for(int o = 1; o < 32; o <<= 1)
    {
        const unsigned t = __shfl_down_sync(0xffffffffu, sfx, o);
        if(lane + o < 32)
            sfx += t;
    }
    if(lane == 0 && wid < 8)
        s_w[wid] = sfx;
    __syncthreads();
#pragma unroll
    for(int w = 1; w < 8; ++w)
        sfx += (w > wid) ? s_w[w] : 0u;
    const unsigned next = sfx - hval;
    if(tid < 256 && sfx >= k && next < k)
    {
        *s_thr = (unsigned)tid;
        *s_k = k - next;
    }
    __syncthreads();
}

// The CTA is a chain of dependent steps (scan -> Harris -> four radix passes -> compaction -> ranking); with the lists in global
// memory every step began with an exposed L2 round trip (~14 of them, most of the kernel's 17 us on one frame).  Now the
// first-cut survivors (record + response key) and the final survivors sit in shared memory whenever they fit -- 2 n_l + ties
// and n_l + ties entries: always, for the reference's feature counts -- and the global lists are written only as the
// overflow fallback reads them; the candidate scan keeps four loads in flight per thread.
template <int SEL_THREADS>
__global__ void __launch_bounds__(SEL_THREADS) k_select(const __grid_constant__ PipeParams P, const uint8_t* __restrict__ pyr,
                                                        unsigned* __restrict__ cand, unsigned* __restrict__ cand2,
                                                        unsigned* __restrict__ keys, unsigned* __restrict__ keys2,
                                                        LevelCounters* __restrict__ ctr, const unsigned* __restrict__ hist,
                                                        SelRec* __restrict__ sel, int* __restrict__ err)
{
    __shared__ unsigned s_hist[256];
    __shared__ unsigned s_w[8];
    __shared__ unsigned s_rec[SEL_SMEM], s_key[SEL_SMEM];   // first-cut survivors: candidate record, ordered response key
    __shared__ unsigned s_frec[SEL_SMEM], s_fkey[SEL_SMEM]; // final survivors
    __shared__ unsigned s_cnt, s_thr, s_k;
    const int l = blockIdx.x, b = blockIdx.y + P.frame0, tid = threadIdx.x;
    const LevelDesc& L = P.lv[l];
    LevelCounters* C = ctr + (long long)b * P.nlevels + l;
    unsigned* c1 = cand + (long long)b * P.cand_frame_stride + L.cand_off;
    unsigned* c2 = cand2 + (long long)b * P.cand_frame_stride + L.cand_off;
    unsigned* k1v = keys + (long long)b * P.cand_frame_stride + L.cand_off;
    unsigned* k2v = keys2 + (long long)b * P.cand_frame_stride + L.cand_off;
    const uint8_t* img = pyr + L.img_off + (long long)b * L.img_frame_stride;

    // the counter and the histogram bin travel together (the bin is only needed when the cut is)
    const unsigned hval = tid < 256 ? hist[((long long)b * P.nlevels + l) * 256 + tid] : 0u;
    unsigned n = C->n_fast;
    if(n > (unsigned)L.cand_cap)
    {
        if(tid == 0)
            atomicOr(err + b, 1);
        n = (unsigned)L.cand_cap;
    }
    // ---- cut 1: k-th largest integer score ----
    const unsigned kcut1 = 2u * (unsigned)L.n_keep;
    unsigned thr1 = 0;
    if(n > kcut1)
    {
        if(kcut1 == 0)
            thr1 = 256; // keep nothing
        else
        {
            // the histogram counts every NMS survivor (>= n > kcut1 entries), so the k-th largest exists
            suffix_select(hval, kcut1, s_w, &s_thr, &s_k);
            thr1 = s_thr;
        }
    }
    if(tid == 0)
        s_cnt = 0;
    __syncthreads();
    for(unsigned i0 = 0; i0 < n; i0 += 4 * SEL_THREADS)
    {
        unsigned rec[4];
#pragma unroll
        for(int j = 0; j < 4; ++j)
        {
            const unsigned i = i0 + j * SEL_THREADS + tid;
            rec[j] = i < n ? c1[i] : 0u;
        }
#pragma unroll
        for(int j = 0; j < 4; ++j)
        {
            const unsigned i = i0 + j * SEL_THREADS + tid;
            if(i < n && (rec[j] & 255u) >= thr1)
            {
                const unsigned pos = atomicAdd(&s_cnt, 1u);
                if(pos < (unsigned)SEL_SMEM)
                    s_rec[pos] = rec[j];
                c2[pos] = rec[j];
            }
        }
    }
    __syncthreads();
    const unsigned m = s_cnt;
    const bool m_smem = m <= (unsigned)SEL_SMEM;
    if(tid == 0)
        C->n_cut1 = m;
    // ---- Harris on the survivors ----
    for(unsigned i = tid; i < m; i += SEL_THREADS)
    {
        const unsigned rec = m_smem ? s_rec[i] : c2[i];
        const unsigned k = fkey(harris_response(img, L.pitch, (rec >> 8) & 0xfffu, rec >> 20));
        if(m_smem)
            s_key[i] = k;
        else
            k1v[i] = k;
    }
    __syncthreads();
    // ---- cut 2: k-th largest response (radix select, MSB first) ----
    const unsigned kcut2 = (unsigned)L.n_keep;
    unsigned thrkey = 0;
    if(m > kcut2)
    {
        if(kcut2 == 0)
            thrkey = 0xffffffffu; // keep nothing (handled below)
        else
        {
            unsigned prefix = 0, mask = 0, krem = kcut2;
            for(int shift = 24; shift >= 0; shift -= 8)
            {
                if(tid < 256)
                    s_hist[tid] = 0;
                __syncthreads();
                for(unsigned i = tid; i < m; i += SEL_THREADS)
                {
                    const unsigned k = m_smem ? s_key[i] : k1v[i];
                    if((k & mask) == prefix)
                        atomicAdd(&s_hist[(k >> shift) & 255u], 1u);
                }
                __syncthreads();
                // by induction at least krem keys carry the current prefix, so the digit exists
                suffix_select(tid < 256 ? s_hist[tid] : 0u, krem, s_w, &s_thr, &s_k);
                prefix |= s_thr << shift;
                mask |= 255u << shift;
                krem = s_k;
                __syncthreads();
            }
            thrkey = prefix;
        }
    }
    // ---- final compaction (the global list reuses the dead first-cut input) ----
    if(tid == 0)
        s_cnt = 0;
    __syncthreads();
    const bool keep_none = (m > kcut2 && kcut2 == 0);
    for(unsigned i = tid; i < m; i += SEL_THREADS)
    {
        const unsigned k = m_smem ? s_key[i] : k1v[i];
        if(!keep_none && k >= thrkey)
        {
            const unsigned pos = atomicAdd(&s_cnt, 1u);
            const unsigned rec = m_smem ? s_rec[i] : c2[i];
            if(pos < (unsigned)SEL_SMEM)
            {
                s_frec[pos] = rec;
                s_fkey[pos] = k;
            }
            c1[pos] = rec;
            k2v[pos] = k;
        }
    }
    __syncthreads();
    const unsigned nf = s_cnt;
    if(tid == 0)
    {
        C->n_final = nf;
        if(nf > (unsigned)P.kp_cap)
            atomicOr(err + b, 2);
    }
    // ---- canonical order: rank by (y, x) ----
    SelRec* out = sel + ((long long)b * P.nlevels + l) * P.kp_cap;
    const bool in_smem = nf <= (unsigned)SEL_SMEM;
    for(unsigned i = tid; i < nf; i += SEL_THREADS)
    {
        const unsigned rec = in_smem ? s_frec[i] : c1[i], me = rec >> 8;
        unsigned rank = 0;
        if(in_smem)
        {
#pragma unroll 4
            for(unsigned j = 0; j < nf; ++j)
                rank += (s_frec[j] >> 8) < me;
        }
        else
            for(unsigned j = 0; j < nf; ++j)
                rank += (c1[j] >> 8) < me;
        if(rank < (unsigned)P.kp_cap)
        {
            SelRec r;
            r.yx_score = rec;
            r.response = funkey(in_smem ? s_fkey[i] : k2v[i]);
            out[rank] = r;
        }
    }
}

// =================================================================================================
// K5  orient + describe : one warp per keypoint (items 7, 8, 10)
//   intensity-centroid moments over the radius-15 disc of the UNBLURRED level (int), cv::fastAtan2
//   (non-FMA f32 polynomial), then the 256 rotated BRIEF tests on the BLURRED level; lane i produces
//   descriptor byte i.
// =================================================================================================

constexpr int OD_SLOTS = 32; // keypoints per CTA on a full machine: 8 warps x 4 rounds
constexpr int OD_SLOTS_FEW = 8; // ... when 32 per CTA would leave SMs idle (single frames): one round per warp
constexpr int IC_WORDS = 31 * 9; // aligned words covering the 31 x 31 orientation patch (any alignment)

// One CTA takes 32 consecutive output slots of one frame.  Phase 0: thread i finds (level, rank) of slot i.  Phase 1: a
// warp per keypoint computes the intensity-centroid angle.  Phase 2: the 32 double-precision sincos of the CTA run in
// the 32 lanes of ONE warp (as one warp-per-keypoint step they would cost every warp ~240 issue slots for a single
// active lane).  Phase 3: a warp per keypoint evaluates the 256 rotated tests, lane i -> descriptor byte i.
template <int SLOTS>
__global__ void __launch_bounds__(256) k_orient_describe(const __grid_constant__ PipeParams P, const uint8_t* __restrict__ pyr,
                                                         const uint8_t* __restrict__ blur,
                                                         const LevelCounters* __restrict__ ctr, const SelRec* __restrict__ sel,
                                                         const float4* __restrict__ pat_t, const uint2* __restrict__ ic_w,
                                                         mslam_b200_keypoint* __restrict__ out_kp, uint8_t* __restrict__ out_desc,
                                                         int* __restrict__ out_counts, int out_cap)
{
    __shared__ int s_l[SLOTS];
    __shared__ unsigned s_yx[SLOTS];
    __shared__ float s_resp[SLOTS], s_ang[SLOTS], s_cos[SLOTS], s_sin[SLOTS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int b = blockIdx.y + P.frame0;
    const int slot0 = blockIdx.x * SLOTS;
    // ---- phase 0: slot -> (level, rank) ----
    if(threadIdx.x < SLOTS)
    {
        const int slot = slot0 + threadIdx.x;
        int base = 0, l = -1, r = 0;
        for(int i = 0; i < P.nlevels; ++i)
        {
            const int c = min((int)ctr[(long long)b * P.nlevels + i].n_final, P.kp_cap);
            if(l < 0 && slot < base + c)
            {
                l = i;
                r = slot - base;
            }
            base += c;
        }
        if(slot == 0)
            out_counts[b] = base;
        if(slot >= out_cap)
            l = -1;
        s_l[threadIdx.x] = l;
        if(l >= 0)
        {
            const SelRec rec = sel[((long long)b * P.nlevels + l) * P.kp_cap + r];
            s_yx[threadIdx.x] = rec.yx_score;
            s_resp[threadIdx.x] = rec.response;
        }
    }
    __syncthreads();
    // ---- phase 1: intensity centroid.  The 31 x 31 disc is read as 31 rows x 9 ALIGNED words (279 words, lane i takes
    // words i, i + 32, ...); every word meets two packed signed-byte weight words -- u * inside(u, v) and
    // v * inside(u, v) for its four columns, tabulated per alignment of the patch -- in two IDP.4A instructions:
    // m10 += sum u * I, m01 += sum v * I.  9 word loads per lane instead of 31 predicated byte loads.
    // (row, byte offset of the word) of this lane's nine patch words do not depend on the keypoint: once per CTA
    unsigned ic_row[(IC_WORDS + 31) / 32], ic_col[(IC_WORDS + 31) / 32];
#pragma unroll
    for(int it = 0; it < (IC_WORDS + 31) / 32; ++it)
    {
        const unsigned i = min(lane + 32 * it, IC_WORDS - 1);
        ic_row[it] = i / 9u;
        ic_col[it] = 4u * (i - 9u * ic_row[it]);
    }
#pragma unroll 1
    for(int rnd = 0; rnd < SLOTS / 8; ++rnd)
    {
        const int s = rnd * 8 + warp;
        const int l = s_l[s];
        if(l < 0) // warp-uniform
            continue;
        const LevelDesc& L = P.lv[l];
        const int x = (s_yx[s] >> 8) & 0xfff, y = s_yx[s] >> 20;
        const unsigned pitch = (unsigned)L.pitch;
        const int al = (x - 15) & 3;
        // keypoints keep 31 pixels from the level's border: the patch origin is inside the plane, every offset is unsigned
        const uint8_t* base = pyr + L.img_off + (long long)b * L.img_frame_stride + (long long)(y - 15) * pitch + ((x - 15) & ~3);
        const uint2* wtab = ic_w + al * IC_WORDS;
        int m10 = 0, m01 = 0;
#pragma unroll
        for(int it = 0; it < (IC_WORDS + 31) / 32; ++it)
        {
            const int i = lane + 32 * it;
            if(i < IC_WORDS)
            {
                const unsigned word = __ldg(reinterpret_cast<const unsigned*>(base + (ic_row[it] * pitch + ic_col[it])));
                const uint2 wgt = __ldg(wtab + i);
                asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(m10) : "r"(word), "r"(wgt.x));
                asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(m01) : "r"(word), "r"(wgt.y));
            }
        }
#pragma unroll
        for(int o = 16; o > 0; o >>= 1)
        {
            m10 += __shfl_xor_sync(0xffffffffu, m10, o);
            m01 += __shfl_xor_sync(0xffffffffu, m01, o);
        }
        if(lane == 0)
            s_ang[s] = fast_atan2_deg((float)m01, (float)m10);
    }
    __syncthreads();
    // ---- phase 2: (float)cos((double)a), (float)sin((double)a) of OpenCV, one keypoint per lane ----
    if(warp == 0 && lane < SLOTS && s_l[lane] >= 0)
    {
        const float ang = __fmul_rn(s_ang[lane], (float)(3.1415926535897932384626433832795 / 180.f));
        double sd, cd;
        sincos((double)ang, &sd, &cd);
        s_cos[lane] = (float)cd;
        s_sin[lane] = (float)sd;
    }
    __syncthreads();
    // ---- phase 3: rotated BRIEF, lane i produces descriptor byte i ----
#pragma unroll 1
    for(int rnd = 0; rnd < SLOTS / 8; ++rnd)
    {
        const int s = rnd * 8 + warp;
        const int l = s_l[s];
        if(l < 0)
            continue;
        const int slot = slot0 + s;
        const LevelDesc& L = P.lv[l];
        const int x = (s_yx[s] >> 8) & 0xfff, y = s_yx[s] >> 20;
        const int pitch = L.pitch;
        const float ca = s_cos[s], sa = s_sin[s];
        const float kx = __fmul_rn((float)x, L.scale), ky = __fmul_rn((float)y, L.scale);
        const int cx = __float2int_rn(__fmul_rn(kx, L.inv_scale)), cy = __float2int_rn(__fmul_rn(ky, L.inv_scale));
        // cvRound(v) = low bits of (v + 1.5 * 2^23) - 0x4B400000 (round-to-nearest-even, |v| < 2^22); the two subtractions
        // are folded into the base pointer:  iy * pitch + ix = by * pitch + bx - 0x4B400000 * (pitch + 1)  (mod 2^32)
        const unsigned kfold = 0x4B400000u * (unsigned)(pitch + 1);
        const uint8_t* centre = blur + L.img_off + (long long)b * L.img_frame_stride + (long long)cy * pitch + cx;
        unsigned byte = 0;
#pragma unroll
        for(int j = 0; j < 8; ++j)
        {
            const float4 t = __ldg(pat_t + j * 32 + lane); // test (lane * 8 + j): x0, y0, x1, y1
            const unsigned bx0 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(t.x, ca), __fmul_rn(t.y, sa)), 12582912.f));
            const unsigned by0 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(t.x, sa), __fmul_rn(t.y, ca)), 12582912.f));
            const unsigned bx1 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(t.z, ca), __fmul_rn(t.w, sa)), 12582912.f));
            const unsigned by1 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(t.z, sa), __fmul_rn(t.w, ca)), 12582912.f));
            const int o0 = (int)(by0 * (unsigned)pitch + bx0 - kfold), o1 = (int)(by1 * (unsigned)pitch + bx1 - kfold);
            const int v0 = __ldg(centre + o0);
            const int v1 = __ldg(centre + o1);
            byte |= (unsigned)(v0 < v1) << j;
        }
        out_desc[((long long)b * out_cap + slot) * 32 + lane] = (uint8_t)byte;
        if(lane == 0)
        {
            mslam_b200_keypoint kp;
            kp.x = kx;
            kp.y = ky;
            kp.size = __fmul_rn(31.f, L.scale);
            kp.angle = s_ang[s];
            kp.response = s_resp[s];
            kp.octave = l;
            out_kp[(long long)b * out_cap + slot] = kp;
        }
    }
}

// intensity-centroid weights: for alignment a = (x - 15) & 3, word (row r, k) of the aligned window holds columns
// u = 4k + j - a - 15 (j = byte) of row v = r - 15; weight bytes are u resp. v inside the radius-15 disc
// (|u| <= u_max[|v|], OpenCV's umax table), 0 outside.  out: 4 * IC_WORDS * {wu, wv}.
void ic_weight_table(unsigned* out)
{
    static const int umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};
    for(int a = 0; a < 4; ++a)
        for(int r = 0; r < 31; ++r)
            for(int k = 0; k < 9; ++k)
            {
                unsigned wu = 0, wv = 0;
                const int v = r - 15, av = v < 0 ? -v : v;
                for(int j = 0; j < 4; ++j)
                {
                    const int u = 4 * k + j - a - 15, au = u < 0 ? -u : u;
                    if(au <= umax[av])
                    {
                        wu |= (unsigned)(u & 0xff) << (8 * j);
                        wv |= (unsigned)(v & 0xff) << (8 * j);
                    }
                }
                out[((a * 31 + r) * 9 + k) * 2] = wu;
                out[((a * 31 + r) * 9 + k) * 2 + 1] = wv;
            }
}

// rBRIEF pattern as floats, transposed for the warp-per-keypoint kernel: entry j * 32 + lane = test lane * 8 + j
void brief_pattern_transposed(float* out1024)
{
    static const int8_t pat[256 * 4] = {
#include "brief_pattern.inc"
    };
    for(int lane = 0; lane < 32; ++lane)
        for(int j = 0; j < 8; ++j)
            for(int c = 0; c < 4; ++c)
                out1024[(j * 32 + lane) * 4 + c] = (float)pat[(lane * 8 + j) * 4 + c];
}

// =================================================================================================
// launchers
// =================================================================================================
void launch_gray(cudaStream_t st, const uint8_t* d_bgr, uint8_t* pyr, const PipeParams& P, long long* launches)
{
    const LevelDesc& L0 = P.lv[0];
    const long long in_fs = (long long)P.W * P.H * 3;
    if((P.W & 15) == 0 && (reinterpret_cast<uintptr_t>(d_bgr) & 15) == 0)
    {
        const int n = (P.W >> 4) * P.H;
        k_gray16<<<dim3((n + 255) / 256, P.n_frames), 256, 0, st>>>(d_bgr + P.frame0 * in_fs, in_fs,
                                                                    pyr + L0.img_off + P.frame0 * L0.img_frame_stride, L0.img_frame_stride,
                                                                    L0.pitch, P.W, P.H);
    }
    else
        k_gray1<<<dim3((P.W + 255) / 256, P.H, P.n_frames), 256, 0, st>>>(d_bgr + P.frame0 * in_fs, in_fs,
                                                                          pyr + L0.img_off + P.frame0 * L0.img_frame_stride,
                                                                          L0.img_frame_stride, L0.pitch, P.W, P.H);
    ++*launches;
}

static long long resize_tail_max_pixels()
{
    static long long v = -1;
    if(v < 0)
    {
        const char* e = getenv("MSLAM_RESIZE_TAIL_MAX_PX"); // A/B aid
        v = e ? atoll(e) : 150000; // VGA: levels 2..7 (measured: c2 107.1k -> 110.8k frames/s; 80000 = levels 4..7 was neutral)
    }
    return v;
}
static int resize_tail_min_frames()
{
    static int v = -1;
    if(v < 0)
    {
        const char* e = getenv("MSLAM_RESIZE_TAIL_MIN_FRAMES"); // A/B aid: a huge value switches the fused tail off
        v = e ? atoi(e) : 16;
    }
    return v;
}

void launch_pyramid(cudaStream_t st, uint8_t* pyr, const PipeParams& P, const short2* xtab, const short2* ytab,
                    const uint4* qtab, int sm_count, int l_begin, int l_end, long long* launches)
{
    // trailing small levels (all on the quad fast path) of a batch go to the fused per-frame kernel
    int l_tail = l_end;
    if(l_end == P.nlevels && P.n_frames >= resize_tail_min_frames())
    {
        while(l_tail - 1 >= (l_begin < 1 ? 1 : l_begin) && P.lv[l_tail - 1].rs_quads > 0 &&
              (long long)P.lv[l_tail - 1].w * P.lv[l_tail - 1].h <= resize_tail_max_pixels())
            --l_tail;
        if(l_end - l_tail < 2)
            l_tail = l_end; // one level: nothing to fuse
    }
    for(int l = l_begin < 1 ? 1 : l_begin; l < l_tail; ++l)
    {
        const LevelDesc &S = P.lv[l - 1], &D = P.lv[l];
        if(D.rs_quads > 0)
        {
            // the shortest band that still keeps the launch within one wave of resident warps: a warp's walk is a serial
            // chain (~90 instructions per row), so small levels want many short walks rather than few long ones
            const int xblocks = (D.rs_quads + 31) / 32;
            int rs_band = 2;
            while(rs_band < RS_BAND_MAX &&
                  (long long)xblocks * ((D.h + rs_band - 1) / rs_band) * P.n_frames > (long long)sm_count * 64)
                rs_band *= 2;
            const int bands = (D.h + rs_band - 1) / rs_band;
            k_resize_quads<<<dim3((xblocks * bands + 3) / 4, P.n_frames), 128, 0, st>>>(
                pyr + S.img_off + P.frame0 * S.img_frame_stride, S.pitch, S.img_frame_stride, S.h,
                pyr + D.img_off + P.frame0 * D.img_frame_stride, D.pitch, D.img_frame_stride, D.h, D.rs_quads, xblocks, rs_band,
                qtab + D.qtab_off, ytab + D.ytab_off);
        }
        else
        {
            dim3 blk(32, 8);
            dim3 grd((D.w + 4 + 127) / 128, (D.h + 7) / 8, P.n_frames);
            k_resize_generic<<<grd, blk, 0, st>>>(pyr + S.img_off + P.frame0 * S.img_frame_stride, S.pitch, S.img_frame_stride,
                                                  S.w, S.h, pyr + D.img_off + P.frame0 * D.img_frame_stride, D.pitch,
                                                  D.img_frame_stride, D.w, D.h, xtab + D.xtab_off, ytab + D.ytab_off);
        }
        ++*launches;
    }
    if(l_tail < l_end)
    {
        static int rt_band = 0;
        if(rt_band == 0)
        {
            const char* e = getenv("MSLAM_RESIZE_TAIL_BAND"); // A/B aid
            rt_band = e ? atoi(e) : RT_BAND;
            if(rt_band < 1 || rt_band > 32)
                rt_band = RT_BAND;
        }
        static int rt_adapt = -1;
        if(rt_adapt < 0)
        {
            const char* e = getenv("MSLAM_RESIZE_TAIL_ADAPT"); // A/B aid: 0 = the fixed band height everywhere
            rt_adapt = e ? atoi(e) : 1;
        }
        k_resize_tail<<<P.n_frames * RT_CLUSTER, RT_THREADS, 0, st>>>(P, l_tail, rt_band, rt_adapt, pyr, qtab, ytab);
        ++*launches;
    }
}

int blur_band_rows()
{
    static int band = 0;
    if(band == 0)
    {
        const char* e = getenv("MSLAM_BLUR_BAND");
        band = e ? atoi(e) : 35; // a multiple of 7: whole bands take k_blur's guard-free walk
        if(band < 7)
            band = 35;
    }
    return band;
}

void blur_warps(int w, int h, int* xblocks, int* warps)
{
    const int BT_H = blur_band_rows();
    *xblocks = (w + BT_W - 1) / BT_W;
    *warps = *xblocks * ((h + BT_H - 1) / BT_H);
}

// levels [l_begin, l_end) of every frame in one launch
void launch_blur(cudaStream_t st, const uint8_t* pyr, uint8_t* blur, const PipeParams& P, int l_begin, int l_end,
                 long long* launches)
{
    int warps = 0;
    for(int l = l_begin; l < l_end; ++l)
        warps += P.lv[l].blur_xblocks * ((P.lv[l].h + P.blur_band - 1) / P.blur_band);
    if(warps <= 0)
        return;
    k_blur<<<dim3((warps + 3) / 4, P.n_frames), 128, 0, st>>>(P, l_begin, l_end, pyr, blur);
    ++*launches;
}

void fast_tiles(int w, int h, int edge, bool short_tiles, int* tx, int* ty, int* tile_sh)
{
    const int ox0 = ((edge - 1) & ~3) + 2; // first assigned interior column of tile 0
    const int wi = w - edge - ox0, hi = h - 2 * edge;
    // tall tiles (64 scored rows) halve the per-tile fixed cost and the halo share; short ones waste fewer rows on levels
    // whose interior is only a few tiles high -- and are what a call of one or two frames uses everywhere (short_tiles): its
    // launches never fill the machine, so they last as long as ONE tile
    const int sh = (hi > 150 && !short_tiles) ? 64 : 32;
    *tile_sh = sh;
    *tx = wi > 0 ? (wi + FS_IW - 1) / FS_IW : 0;
    *ty = hi > 0 ? (hi + (sh - 2) - 1) / (sh - 2) : 0;
    if(*tx == 0 || *ty == 0)
        *tx = *ty = 0;
}

// levels [l_begin, l_end) in one launch
// tensor maps of the FAST tile loads: level l = a 3-D u8 tensor (pitch, h, max_batch) with strides (pitch, frame stride),
// box (FS_PW, tile_sh + 6, 1).  cuTensorMapEncodeTiled is fetched through the runtime (no link against libcuda).
cudaError_t build_fast_maps(const PipeParams& P, const uint8_t* pyr, int max_batch, FastMaps* out)
{
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn encode = nullptr;
    if(!encode)
    {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
        if(e != cudaSuccess)
            return e;
        if(!fn || q != cudaDriverEntryPointSuccess)
            return cudaErrorNotSupported;
        encode = reinterpret_cast<EncodeFn>(fn);
    }
    static_assert(sizeof(CUtensorMap) == 128, "FastMaps holds raw CUtensorMap bytes");
    memset(out, 0, sizeof(*out));
    for(int l = 0; l < P.nlevels; ++l)
    {
        const LevelDesc& L = P.lv[l];
        if(L.tiles_x * L.tiles_y == 0)
            continue;
        const cuuint64_t dims[3] = {(cuuint64_t)L.pitch, (cuuint64_t)L.h, (cuuint64_t)max_batch};
        const cuuint64_t strides[2] = {(cuuint64_t)L.pitch, (cuuint64_t)L.img_frame_stride};
        const cuuint32_t box[3] = {(cuuint32_t)FS_PW, (cuuint32_t)(L.tile_sh + 6), 1u};
        const cuuint32_t estr[3] = {1u, 1u, 1u};
        const CUresult r = encode(reinterpret_cast<CUtensorMap*>(&out->m[l][0]), CU_TENSOR_MAP_DATA_TYPE_UINT8, 3,
                                  const_cast<uint8_t*>(pyr) + L.img_off, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                  CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if(r != CUDA_SUCCESS)
            return cudaErrorInvalidValue;
    }
    return cudaSuccess;
}

void launch_fast(cudaStream_t st, const uint8_t* pyr, unsigned* cand, LevelCounters* ctr, unsigned* hist,
                 const PipeParams& P, const FastMaps& maps, int l_begin, int l_end, long long* launches)
{
    int tiles = 0;
    for(int l = l_begin; l < l_end; ++l)
        tiles += P.lv[l].tiles_x * P.lv[l].tiles_y;
    if(tiles <= 0)
        return;
    static int dbg = -1;
    static FastMaps* d_maps = nullptr;
    if(dbg < 0)
    {
        const char* e = getenv("MSLAM_FAST_DBG"); // 1: no trap on a lost tile load, 2: tensor maps read from global memory
        dbg = e ? atoi(e) : 0;
    }
    if(dbg & 2)
    {
        if(!d_maps)
            cudaMalloc(&d_maps, sizeof(FastMaps));
        cudaMemcpyAsync(d_maps, &maps, sizeof(FastMaps), cudaMemcpyHostToDevice, st);
    }
    k_fast<<<dim3(tiles, P.n_frames), 256, 0, st>>>(P, maps, (dbg & 2) ? d_maps : nullptr, dbg, l_begin, l_end, pyr, cand, ctr, hist);
    ++*launches;
}

void launch_select(cudaStream_t st, const uint8_t* pyr, unsigned* cand, unsigned* cand2, unsigned* keys, unsigned* keys2,
                   LevelCounters* ctr, const unsigned* hist, SelRec* sel, int* err, const PipeParams& P, int sm_count,
                   long long* launches)
{
    // one CTA per (frame, level): with few of them the longest one (level 0) is the critical path, so it gets 1024 threads
    if((long long)P.nlevels * P.n_frames <= 2LL * sm_count)
        k_select<1024><<<dim3(P.nlevels, P.n_frames), 1024, 0, st>>>(P, pyr, cand, cand2, keys, keys2, ctr, hist, sel, err);
    else
        k_select<256><<<dim3(P.nlevels, P.n_frames), 256, 0, st>>>(P, pyr, cand, cand2, keys, keys2, ctr, hist, sel, err);
    ++*launches;
}

void launch_orient_describe(cudaStream_t st, const uint8_t* pyr, const uint8_t* blur, const LevelCounters* ctr,
                            const SelRec* sel, const float4* pat_t, const uint2* ic_w, mslam_b200_keypoint* out_kp,
                            uint8_t* out_desc, int* out_counts, int out_cap, int n_slots, const PipeParams& P, int sm_count, long long* launches)
{
    const int slots = n_slots < 1 ? 1 : n_slots;
    // 32 keypoints per CTA amortise the per-CTA phases; with few frames that grid is smaller than the machine and the four
    // rounds of a warp run one after the other on mostly idle SMs, so small batches take 8 per CTA (one round per warp)
    if((long long)((slots + OD_SLOTS - 1) / OD_SLOTS) * P.n_frames < 2LL * sm_count)
        k_orient_describe<OD_SLOTS_FEW><<<dim3((slots + OD_SLOTS_FEW - 1) / OD_SLOTS_FEW, P.n_frames), 256, 0, st>>>(
            P, pyr, blur, ctr, sel, pat_t, ic_w, out_kp, out_desc, out_counts, out_cap);
    else
        k_orient_describe<OD_SLOTS><<<dim3((slots + OD_SLOTS - 1) / OD_SLOTS, P.n_frames), 256, 0, st>>>(P, pyr, blur, ctr, sel, pat_t, ic_w,
                                                                                              out_kp, out_desc, out_counts, out_cap);
    ++*launches;
}

} // namespace mslam_b200
