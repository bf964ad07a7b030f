// orb_kernels.cu -- sm_100a kernels of the ORB detect + describe path.
//
// What each kernel must compute is OpenCV's cv::ORB::detectAndCompute as called by the reference
// (src/lib/modular_slam/orb_feature.cpp:25,40) after toGrayScale (src/lib/modular_slam/frame.cpp:6-27);
// the arithmetic contract is restated next to every kernel.
// All float stages use explicit round-to-nearest intrinsics so that nvcc can never contract them into
// FMAs (the reference path is non-FMA float except the 7x7 blur, which is an explicit FMA chain).
#include "common.cuh"
#include "kernels.h"

namespace mslam_b200
{

// =================================================================================================
// K1  gray : interleaved 3-channel u8 -> level 0      (frame.cpp:15-19)
//   g = (u8) min(255.f, .299f*c0 + .587f*c1 + .114f*c2)  three rounded products, two rounded adds,
//   truncating cast; weights in MEMORY order.  HBM-bound: 3 B read + 1 B written per pixel.
// =================================================================================================
__device__ __forceinline__ unsigned gray_px(unsigned c0, unsigned c1, unsigned c2)
{
    float v = __fadd_rn(__fadd_rn(__fmul_rn(0.299f, (float)c0), __fmul_rn(0.587f, (float)c1)),
                        __fmul_rn(0.114f, (float)c2));
    v = v < 255.f ? v : 255.f;
    return __float2uint_rz(v);
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
        unsigned r = 0;
#pragma unroll
        for(int i = 0; i < 4; ++i)
        {
            const int k = 3 * (q * 4 + i);
            const unsigned c0 = (w[k >> 2] >> ((k & 3) * 8)) & 255u;
            const unsigned c1 = (w[(k + 1) >> 2] >> (((k + 1) & 3) * 8)) & 255u;
            const unsigned c2 = (w[(k + 2) >> 2] >> (((k + 2) & 3) * 8)) & 255u;
            r |= gray_px(c0, c1, c2) << (8 * i);
        }
        o[q] = r;
    }
    *reinterpret_cast<uint4*>(out + (long long)blockIdx.y * out_fs + (long long)y * pitch + c * 16) =
        make_uint4(o[0], o[1], o[2], o[3]);
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
    out[(long long)blockIdx.z * out_fs + (long long)y * pitch + x] = (uint8_t)gray_px(p[0], p[1], p[2]);
}

// =================================================================================================
// K2  resize : level l-1 -> level l, cv::resize INTER_LINEAR_EXACT (item 2)
//   per axis the host precomputes (ofs, alpha1) in IEEE double; horizontal pass in 8.8 fixed point,
//   vertical pass (a0*h0 + a1*h1 + 32768) >> 16.  4 output pixels per thread, 32-bit stores.
// =================================================================================================
__global__ void __launch_bounds__(256) k_resize(const uint8_t* __restrict__ src, int spitch, long long sfs, int sw, int sh,
                                                uint8_t* __restrict__ dst, int dpitch, long long dfs, int dw, int dh,
                                                const short2* __restrict__ xtab, const short2* __restrict__ ytab)
{
    const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if(x4 >= dw || y >= dh)
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
        if(x < dw)
        {
            const short2 tx = __ldg(xtab + x);
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
constexpr int BT_W = 64, BT_H = 16;

__device__ __forceinline__ int reflect101(int p, int n)
{
    if(n == 1)
        return 0;
    while(p < 0 || p >= n)
        p = p < 0 ? -p : 2 * n - 2 - p;
    return p;
}

__global__ void __launch_bounds__(256) k_blur(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int pitch,
                                              long long fs, int w, int h)
{
    __shared__ uint8_t s_in[BT_H + 6][BT_W + 8];
    __shared__ float s_mid[BT_H + 6][BT_W];
    const float k0 = __int_as_float(0x3d8fafb1), k1 = __int_as_float(0x3e06387e), k2 = __int_as_float(0x3e434a39),
                k3 = __int_as_float(0x3e5d4ae0);
    const float kk[7] = {k0, k1, k2, k3, k2, k1, k0};
    const int tx0 = blockIdx.x * BT_W, ty0 = blockIdx.y * BT_H;
    const uint8_t* img = src + (long long)blockIdx.z * fs;
    for(int i = threadIdx.x; i < (BT_H + 6) * (BT_W + 6); i += 256)
    {
        const int ry = i / (BT_W + 6), rx = i - ry * (BT_W + 6);
        const int gy = reflect101(ty0 - 3 + ry, h), gx = reflect101(tx0 - 3 + rx, w);
        s_in[ry][rx] = __ldg(img + (long long)gy * pitch + gx);
    }
    __syncthreads();
    for(int i = threadIdx.x; i < (BT_H + 6) * BT_W; i += 256)
    {
        const int ry = i / BT_W, x = i - ry * BT_W;
        float acc = __fmul_rn(kk[0], (float)s_in[ry][x]);
#pragma unroll
        for(int j = 1; j < 7; ++j)
            acc = __fmaf_rn(kk[j], (float)s_in[ry][x + j], acc);
        s_mid[ry][x] = acc;
    }
    __syncthreads();
    for(int i = threadIdx.x; i < BT_H * BT_W; i += 256)
    {
        const int y = i / BT_W, x = i - y * BT_W;
        if(tx0 + x >= w || ty0 + y >= h)
            continue;
        float acc = __fmul_rn(k3, s_mid[y + 3][x]);
#pragma unroll
        for(int j = 1; j <= 3; ++j)
            acc = __fmaf_rn(kk[3 + j], __fadd_rn(s_mid[y + 3 + j][x], s_mid[y + 3 - j][x]), acc);
        int v = __float2int_rn(acc);
        v = min(max(v, 0), 255);
        dst[(long long)blockIdx.z * fs + (long long)(ty0 + y) * pitch + tx0 + x] = (uint8_t)v;
    }
}

// =================================================================================================
// K3  FAST-9/16 + 3x3 strict NMS + border filter (items 3, 4) on one level of the whole batch.
//   One CTA = 64x16 interior pixels.  Pixels are staged once in shared memory (4-px halo, 32-bit
//   aligned loads); every position first runs the two-pixel high-speed rejects, then the exact
//   16-bit bright/dark ring masks (9 contiguous set bits); corner positions are queued and only
//   those get the exact cornerScore (max over the 16 arcs of the arc minimum, minus 1).
//   Survivors of the NMS are appended to the level's candidate list (one global atomic per CTA) and
//   counted into the level's 256-bin score histogram that retainBest's first cut needs.
// =================================================================================================
constexpr int FT_W = 64, FT_H = 16;
constexpr int FS_W = FT_W + 8, FS_H = FT_H + 8; // staged pixels
constexpr int SC_W = FT_W + 2, SC_H = FT_H + 2; // scored positions
constexpr int SC_P = 68;                        // score row pitch

__device__ __forceinline__ bool has_run9(unsigned m)
{
    unsigned x = m | (m << 16);
    x &= x >> 1;
    x &= x >> 2;
    x &= x >> 4;
    x &= x >> 1;
    return x != 0;
}

__global__ void __launch_bounds__(256) k_fast(const __grid_constant__ PipeParams P, int l, const uint8_t* __restrict__ pyr,
                                              unsigned* __restrict__ cand, LevelCounters* __restrict__ ctr,
                                              unsigned* __restrict__ hist)
{
    __shared__ __align__(16) uint8_t s_px[FS_H][FS_W];
    __shared__ __align__(4) uint8_t s_sc[SC_H][SC_P];
    __shared__ unsigned short s_q[SC_H * SC_W];
    __shared__ unsigned s_out[FT_W * FT_H / 4];
    __shared__ int s_qn, s_on;
    __shared__ unsigned s_base;

    const LevelDesc& L = P.lv[l];
    const int b = blockIdx.z;
    const int tid = threadIdx.x;
    const int edge = P.edge, thr = P.fast_thr;
    const int ox0 = (edge & ~3) + FT_W * blockIdx.x, oy0 = edge + FT_H * blockIdx.y; // first interior pixel of the tile
    const int hx0 = ox0 - 4, hy0 = oy0 - 4;                                           // first staged pixel
    const uint8_t* img = pyr + L.img_off + (long long)b * L.img_frame_stride;

    for(int i = tid; i < FS_H * (FS_W / 4); i += 256)
    {
        const int r = i / (FS_W / 4), c = i - r * (FS_W / 4);
        const int gy = hy0 + r, gx = hx0 + 4 * c;
        unsigned v = 0;
        if(gy < L.h && gx < L.pitch)
            v = __ldg(reinterpret_cast<const unsigned*>(img + (long long)gy * L.pitch + gx));
        reinterpret_cast<unsigned*>(&s_px[r][0])[c] = v;
    }
    for(int i = tid; i < SC_H * SC_P / 4; i += 256)
        reinterpret_cast<unsigned*>(&s_sc[0][0])[i] = 0;
    if(tid == 0)
    {
        s_qn = 0;
        s_on = 0;
    }
    __syncthreads();

    // ---- corner test on every scored position ----
    for(int i = tid; i < SC_H * SC_W; i += 256)
    {
        const int sy = i / SC_W, sx = i - sy * SC_W;
        const int x = ox0 - 1 + sx, y = oy0 - 1 + sy;
        if(x < edge - 1 || x > L.w - edge || y > L.h - edge)
            continue;
        const uint8_t* c = &s_px[sy + 3][sx + 3];
        const int v = c[0], hi = v + thr, lo = v - thr;
        const int p0 = c[3 * FS_W], p8 = c[-3 * FS_W];
        bool br = (p0 > hi) | (p8 > hi), dk = (p0 < lo) | (p8 < lo);
        if(!(br | dk))
            continue;
        const int p4 = c[3], p12 = c[-3];
        br &= (p4 > hi) | (p12 > hi);
        dk &= (p4 < lo) | (p12 < lo);
        if(!(br | dk))
            continue;
        const int r1 = c[3 * FS_W + 1], r2 = c[2 * FS_W + 2], r3 = c[FS_W + 3], r5 = c[-FS_W + 3], r6 = c[-2 * FS_W + 2],
                  r7 = c[-3 * FS_W + 1], r9 = c[-3 * FS_W - 1], r10 = c[-2 * FS_W - 2], r11 = c[-FS_W - 3],
                  r13 = c[FS_W - 3], r14 = c[2 * FS_W - 2], r15 = c[3 * FS_W - 1];
        const int ring[16] = {p0, r1, r2, r3, p4, r5, r6, r7, p8, r9, r10, r11, p12, r13, r14, r15};
        unsigned mb = 0, md = 0;
#pragma unroll
        for(int k = 0; k < 16; ++k)
        {
            mb |= (unsigned)(ring[k] > hi) << k;
            md |= (unsigned)(ring[k] < lo) << k;
        }
        if(has_run9(mb) || has_run9(md))
            s_q[atomicAdd(&s_qn, 1)] = (unsigned short)i;
    }
    __syncthreads();

    // ---- exact cornerScore for the queued corners ----
    const int nq = s_qn;
    for(int j = tid; j < nq; j += 256)
    {
        const int i = s_q[j];
        const int sy = i / SC_W, sx = i - sy * SC_W;
        const uint8_t* c = &s_px[sy + 3][sx + 3];
        const int v = c[0];
        int mn[16], mx[16];
        {
            const int ring[16] = {c[3 * FS_W],      c[3 * FS_W + 1],  c[2 * FS_W + 2],  c[FS_W + 3],
                                  c[3],             c[-FS_W + 3],     c[-2 * FS_W + 2], c[-3 * FS_W + 1],
                                  c[-3 * FS_W],     c[-3 * FS_W - 1], c[-2 * FS_W - 2], c[-FS_W - 3],
                                  c[-3],            c[FS_W - 3],      c[2 * FS_W - 2],  c[3 * FS_W - 1]};
            // sliding window of 9 over the cyclic ring: 2 -> 4 -> 8 -> 9
            int a2[16], b2[16], a4[16], b4[16];
#pragma unroll
            for(int k = 0; k < 16; ++k)
            {
                a2[k] = min(ring[k], ring[(k + 1) & 15]);
                b2[k] = max(ring[k], ring[(k + 1) & 15]);
            }
#pragma unroll
            for(int k = 0; k < 16; ++k)
            {
                a4[k] = min(a2[k], a2[(k + 2) & 15]);
                b4[k] = max(b2[k], b2[(k + 2) & 15]);
            }
#pragma unroll
            for(int k = 0; k < 16; ++k)
            {
                mn[k] = min(min(a4[k], a4[(k + 4) & 15]), ring[(k + 8) & 15]);
                mx[k] = max(max(b4[k], b4[(k + 4) & 15]), ring[(k + 8) & 15]);
            }
        }
        int best_min = mn[0], best_max = mx[0]; // max over arcs of arc-min ; min over arcs of arc-max
#pragma unroll
        for(int k = 1; k < 16; ++k)
        {
            best_min = max(best_min, mn[k]);
            best_max = min(best_max, mx[k]);
        }
        const int score = max(best_min - v, v - best_max) - 1; // >= thr because the position is a corner
        s_sc[sy][sx] = (uint8_t)score;
    }
    __syncthreads();

    // ---- 3x3 strict NMS + border filter ----
    for(int i = tid; i < FT_W * FT_H; i += 256)
    {
        const int oy = i / FT_W, ox = i - oy * FT_W;
        const uint8_t* s = &s_sc[oy + 1][ox + 1];
        const int v = s[0];
        if(v == 0)
            continue;
        const int x = ox0 + ox, y = oy0 + oy;
        if(x < edge || x >= L.w - edge || y >= L.h - edge)
            continue;
        if(v > s[-1] && v > s[1] && v > s[-SC_P - 1] && v > s[-SC_P] && v > s[-SC_P + 1] && v > s[SC_P - 1] &&
           v > s[SC_P] && v > s[SC_P + 1])
            s_out[atomicAdd(&s_on, 1)] = ((unsigned)y << 20) | ((unsigned)x << 8) | (unsigned)v;
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
    int a = 0, b = 0, c = 0;
    // rows y-4 .. y+4, columns x-4 .. x+4; three-row sliding window
    int r0[9], r1[9], r2[9];
    const uint8_t* p = img + (long long)(y - 4) * pitch + (x - 4);
#pragma unroll
    for(int j = 0; j < 9; ++j)
    {
        r0[j] = __ldg(p + j);
        r1[j] = __ldg(p + pitch + j);
    }
#pragma unroll
    for(int row = 0; row < 7; ++row)
    {
        const uint8_t* q = p + (long long)(row + 2) * pitch;
#pragma unroll
        for(int j = 0; j < 9; ++j)
            r2[j] = __ldg(q + j);
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

constexpr int SEL_THREADS = 256;

__global__ void __launch_bounds__(SEL_THREADS) k_select(const __grid_constant__ PipeParams P, const uint8_t* __restrict__ pyr,
                                                        unsigned* __restrict__ cand, unsigned* __restrict__ cand2,
                                                        unsigned* __restrict__ keys, unsigned* __restrict__ keys2,
                                                        LevelCounters* __restrict__ ctr, const unsigned* __restrict__ hist,
                                                        SelRec* __restrict__ sel, int* __restrict__ err)
{
    __shared__ unsigned s_hist[256];
    __shared__ unsigned s_cnt, s_thr, s_k;
    const int l = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
    const LevelDesc& L = P.lv[l];
    LevelCounters* C = ctr + (long long)b * P.nlevels + l;
    unsigned* c1 = cand + (long long)b * P.cand_frame_stride + L.cand_off;
    unsigned* c2 = cand2 + (long long)b * P.cand_frame_stride + L.cand_off;
    unsigned* k1v = keys + (long long)b * P.cand_frame_stride + L.cand_off;
    unsigned* k2v = keys2 + (long long)b * P.cand_frame_stride + L.cand_off;
    const uint8_t* img = pyr + L.img_off + (long long)b * L.img_frame_stride;

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
        s_hist[tid] = hist[((long long)b * P.nlevels + l) * 256 + tid];
        __syncthreads();
        if(tid == 0)
        {
            unsigned cum = 0, t = 256; // t = 256 => keep nothing (kcut1 == 0)
            if(kcut1 > 0)
                for(int v = 255; v >= 0; --v)
                {
                    cum += s_hist[v];
                    if(cum >= kcut1)
                    {
                        t = (unsigned)v;
                        break;
                    }
                }
            s_thr = t;
        }
        __syncthreads();
        thr1 = s_thr;
    }
    if(tid == 0)
        s_cnt = 0;
    __syncthreads();
    for(unsigned i = tid; i < n; i += SEL_THREADS)
    {
        const unsigned rec = c1[i];
        if((rec & 255u) >= thr1)
            c2[atomicAdd(&s_cnt, 1u)] = rec;
    }
    __syncthreads();
    const unsigned m = s_cnt;
    if(tid == 0)
        C->n_cut1 = m;
    // ---- Harris on the survivors ----
    for(unsigned i = tid; i < m; i += SEL_THREADS)
    {
        const unsigned rec = c2[i];
        k1v[i] = fkey(harris_response(img, L.pitch, (rec >> 8) & 0xfffu, rec >> 20));
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
            unsigned prefix = 0, mask = 0;
            if(tid == 0)
                s_k = kcut2;
            for(int shift = 24; shift >= 0; shift -= 8)
            {
                s_hist[tid] = 0;
                __syncthreads();
                for(unsigned i = tid; i < m; i += SEL_THREADS)
                {
                    const unsigned k = k1v[i];
                    if((k & mask) == prefix)
                        atomicAdd(&s_hist[(k >> shift) & 255u], 1u);
                }
                __syncthreads();
                if(tid == 0)
                {
                    unsigned cum = 0, k = s_k;
                    int v = 255;
                    for(; v > 0; --v)
                    {
                        if(cum + s_hist[v] >= k)
                            break;
                        cum += s_hist[v];
                    }
                    s_k = k - cum;
                    s_thr = (unsigned)v;
                }
                __syncthreads();
                prefix |= s_thr << shift;
                mask |= 255u << shift;
                __syncthreads();
            }
            thrkey = prefix;
        }
    }
    // ---- final compaction (list reuses the dead first-cut input) ----
    if(tid == 0)
        s_cnt = 0;
    __syncthreads();
    const bool keep_none = (m > kcut2 && kcut2 == 0);
    for(unsigned i = tid; i < m; i += SEL_THREADS)
    {
        const unsigned k = k1v[i];
        if(!keep_none && k >= thrkey)
        {
            const unsigned pos = atomicAdd(&s_cnt, 1u);
            c1[pos] = c2[i];
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
    for(unsigned i = tid; i < nf; i += SEL_THREADS)
    {
        const unsigned rec = c1[i], me = rec >> 8;
        unsigned rank = 0;
        for(unsigned j = 0; j < nf; ++j)
            rank += (c1[j] >> 8) < me;
        if(rank < (unsigned)P.kp_cap)
        {
            SelRec r;
            r.yx_score = rec;
            r.response = funkey(k2v[i]);
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

__global__ void __launch_bounds__(256) k_orient_describe(const __grid_constant__ PipeParams P, const uint8_t* __restrict__ pyr,
                                                         const uint8_t* __restrict__ blur,
                                                         const LevelCounters* __restrict__ ctr, const SelRec* __restrict__ sel,
                                                         const unsigned* __restrict__ pat_t,
                                                         mslam_b200_keypoint* __restrict__ out_kp, uint8_t* __restrict__ out_desc,
                                                         int* __restrict__ out_counts, int out_cap)
{
    const int lane = threadIdx.x & 31;
    const int slot = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int b = blockIdx.y;
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
    if(slot == 0 && lane == 0)
        out_counts[b] = base;
    if(l < 0 || slot >= out_cap)
        return;
    const LevelDesc& L = P.lv[l];
    const SelRec rec = sel[((long long)b * P.nlevels + l) * P.kp_cap + r];
    const int x = (rec.yx_score >> 8) & 0xfff, y = rec.yx_score >> 20;
    const long long plane = L.img_off + (long long)b * L.img_frame_stride;
    const int pitch = L.pitch;

    // ---- intensity centroid: the warp walks the 31 rows of the disc, lane = column (coalesced 31-byte row reads) ----
    int m10 = 0, m01 = 0;
    {
        const int u = lane - 15, au = u < 0 ? -u : u;
        const uint8_t* c = pyr + plane + (long long)y * pitch + x + u;
#pragma unroll
        for(int v = -15; v <= 15; ++v)
        {
            const int av = v < 0 ? -v : v;
            const int d = (int)((0x3689ABCDDEEEFFFFull >> (4 * av)) & 15ull); // u_max[|v|]
            const int val = (lane < 31 && au <= d) ? (int)__ldg(c + v * pitch) : 0;
            m10 += u * val;
            m01 += v * val;
        }
    }
#pragma unroll
    for(int o = 16; o > 0; o >>= 1)
    {
        m10 += __shfl_xor_sync(0xffffffffu, m10, o);
        m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    }
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    // ---- rotated BRIEF: lane i produces descriptor byte i ----
    const float kx = __fmul_rn((float)x, L.scale), ky = __fmul_rn((float)y, L.scale);
    const int cx = __float2int_rn(__fmul_rn(kx, L.inv_scale)), cy = __float2int_rn(__fmul_rn(ky, L.inv_scale));
    float ca = 0.f, sa = 0.f;
    if(lane == 0)
    {
        const float ang = __fmul_rn(angle, (float)(3.1415926535897932384626433832795 / 180.f));
        ca = (float)cos((double)ang);
        sa = (float)sin((double)ang);
    }
    ca = __shfl_sync(0xffffffffu, ca, 0);
    sa = __shfl_sync(0xffffffffu, sa, 0);
    const uint8_t* centre = blur + plane + (long long)cy * pitch + cx;
    unsigned byte = 0;
#pragma unroll
    for(int j = 0; j < 8; ++j)
    {
        const unsigned t = __ldg(pat_t + j * 32 + lane); // test (lane * 8 + j): x0 | y0 << 8 | x1 << 16 | y1 << 24 (int8 each)
        const float px0 = (float)(int)(signed char)(t & 255u), py0 = (float)(int)(signed char)((t >> 8) & 255u),
                    px1 = (float)(int)(signed char)((t >> 16) & 255u), py1 = (float)(int)(signed char)(t >> 24);
        const int x0 = __float2int_rn(__fsub_rn(__fmul_rn(px0, ca), __fmul_rn(py0, sa)));
        const int y0 = __float2int_rn(__fadd_rn(__fmul_rn(px0, sa), __fmul_rn(py0, ca)));
        const int x1 = __float2int_rn(__fsub_rn(__fmul_rn(px1, ca), __fmul_rn(py1, sa)));
        const int y1 = __float2int_rn(__fadd_rn(__fmul_rn(px1, sa), __fmul_rn(py1, ca)));
        const int v0 = __ldg(centre + y0 * pitch + x0);
        const int v1 = __ldg(centre + y1 * pitch + x1);
        byte |= (unsigned)(v0 < v1) << j;
    }
    out_desc[((long long)b * out_cap + slot) * 32 + lane] = (uint8_t)byte;
    if(lane == 0)
    {
        mslam_b200_keypoint kp;
        kp.x = kx;
        kp.y = ky;
        kp.size = __fmul_rn(31.f, L.scale);
        kp.angle = angle;
        kp.response = rec.response;
        kp.octave = l;
        out_kp[(long long)b * out_cap + slot] = kp;
    }
}

// rBRIEF pattern, transposed for the warp-per-keypoint kernel: word j * 32 + lane = test lane * 8 + j
void brief_pattern_transposed(unsigned* out256)
{
    static const int8_t pat[256 * 4] = {
#include "brief_pattern.inc"
    };
    for(int lane = 0; lane < 32; ++lane)
        for(int j = 0; j < 8; ++j)
        {
            const int8_t* t = pat + (lane * 8 + j) * 4;
            out256[j * 32 + lane] = (unsigned)(uint8_t)t[0] | ((unsigned)(uint8_t)t[1] << 8) | ((unsigned)(uint8_t)t[2] << 16) |
                                    ((unsigned)(uint8_t)t[3] << 24);
        }
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
        k_gray16<<<dim3((n + 255) / 256, P.n_frames), 256, 0, st>>>(d_bgr, in_fs, pyr + L0.img_off, L0.img_frame_stride,
                                                                    L0.pitch, P.W, P.H);
    }
    else
        k_gray1<<<dim3((P.W + 255) / 256, P.H, P.n_frames), 256, 0, st>>>(d_bgr, in_fs, pyr + L0.img_off,
                                                                          L0.img_frame_stride, L0.pitch, P.W, P.H);
    ++*launches;
}

void launch_pyramid(cudaStream_t st, uint8_t* pyr, const PipeParams& P, const short2* xtab, const short2* ytab,
                    long long* launches)
{
    for(int l = 1; l < P.nlevels; ++l)
    {
        const LevelDesc &S = P.lv[l - 1], &D = P.lv[l];
        dim3 blk(32, 8);
        dim3 grd((D.w + 127) / 128, (D.h + 7) / 8, P.n_frames);
        k_resize<<<grd, blk, 0, st>>>(pyr + S.img_off, S.pitch, S.img_frame_stride, S.w, S.h, pyr + D.img_off, D.pitch,
                                      D.img_frame_stride, D.w, D.h, xtab + D.xtab_off, ytab + D.ytab_off);
        ++*launches;
    }
}

void launch_blur(cudaStream_t st, const uint8_t* pyr, uint8_t* blur, const PipeParams& P, long long* launches)
{
    for(int l = 0; l < P.nlevels; ++l)
    {
        const LevelDesc& L = P.lv[l];
        dim3 grd((L.w + BT_W - 1) / BT_W, (L.h + BT_H - 1) / BT_H, P.n_frames);
        k_blur<<<grd, 256, 0, st>>>(pyr + L.img_off, blur + L.img_off, L.pitch, L.img_frame_stride, L.w, L.h);
        ++*launches;
    }
}

void fast_tiles(int w, int h, int edge, int* tx, int* ty)
{
    const int ox0 = edge & ~3;
    const int wi = w - edge - ox0, hi = h - 2 * edge;
    *tx = wi > 0 ? (wi + FT_W - 1) / FT_W : 0;
    *ty = hi > 0 ? (hi + FT_H - 1) / FT_H : 0;
}

void launch_fast(cudaStream_t st, const uint8_t* pyr, unsigned* cand, LevelCounters* ctr, unsigned* hist,
                 const PipeParams& P, long long* launches)
{
    for(int l = 0; l < P.nlevels; ++l)
    {
        const LevelDesc& L = P.lv[l];
        if(L.tiles_x <= 0 || L.tiles_y <= 0)
            continue;
        k_fast<<<dim3(L.tiles_x, L.tiles_y, P.n_frames), 256, 0, st>>>(P, l, pyr, cand, ctr, hist);
        ++*launches;
    }
}

void launch_select(cudaStream_t st, const uint8_t* pyr, unsigned* cand, unsigned* cand2, unsigned* keys, unsigned* keys2,
                   LevelCounters* ctr, const unsigned* hist, SelRec* sel, int* err, const PipeParams& P,
                   long long* launches)
{
    k_select<<<dim3(P.nlevels, P.n_frames), SEL_THREADS, 0, st>>>(P, pyr, cand, cand2, keys, keys2, ctr, hist, sel, err);
    ++*launches;
}

void launch_orient_describe(cudaStream_t st, const uint8_t* pyr, const uint8_t* blur, const LevelCounters* ctr,
                            const SelRec* sel, const unsigned* pat_t, mslam_b200_keypoint* out_kp, uint8_t* out_desc, int* out_counts,
                            int out_cap, int n_slots, const PipeParams& P, long long* launches)
{
    const int slots = n_slots < 1 ? 1 : n_slots;
    k_orient_describe<<<dim3((slots + 7) / 8, P.n_frames), 256, 0, st>>>(P, pyr, blur, ctr, sel, pat_t, out_kp,
                                                                         out_desc, out_counts, out_cap);
    ++*launches;
}

} // namespace mslam_b200
