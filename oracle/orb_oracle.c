/*
 * orb_oracle.c -- TEST INFRASTRUCTURE ONLY (parity oracle + timed CPU baseline).
 *
 * A plain-C CPU restatement of what modular-slam's OrbOpenCvDetector / OrbOpenCvMatcher compute
 * (reference: src/lib/modular_slam/orb_feature.cpp:33-65 detect, :84-117 match;
 *  src/lib/modular_slam/frame.cpp:6-27 toGrayScale).  The arithmetic of those two classes lives in
 * a third-party dependency that is NOT vendored in the reference tree: OpenCV 4.8.1
 * (conanfile.py:28), calls cv::ORB::create(1000)->detectAndCompute (orb_feature.cpp:25,40) and
 * cv::DescriptorMatcher::create(BRUTEFORCE_HAMMING)->knnMatch(..., 2) (orb_feature.cpp:81,96).
 * This file restates OpenCV's published algorithm for those calls (features2d/src/orb.cpp,
 * fast.cpp, fast_score.cpp, imgproc resize INTER_LINEAR_EXACT, sepFilter2D float path,
 * core fastAtan2, batch_distance) following SURVEY.md section 8(c).
 *
 * PARITY PIN: the reference's own tests hold no golden vector for this path, so the
 * oracle is pinned against outputs of OpenCV itself run in the build container
 * (opencv-python-headless 4.13.0; tests/golden/make_golden.py generated tests/golden/*.npz from the
 * reference's bundled data/rgb/000{0,1}.png) and, wherever cv2 is importable, live against cv2
 * (tests/test_oracle_vs_cv2.py).  Version caveat: oracle == cv2 4.13.0, reference pins 4.8.1.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (libmslam_b200.so) never links or calls it.
 *
 * Build: see oracle/Makefile  (gcc -O2 -ffp-contract=off: no implicit FMA contraction anywhere;
 * the one place OpenCV's float path uses FMA, the 7x7 blur, calls fmaf() explicitly).
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))

static const int8_t kPattern[256 * 4] = {
#include "../modular-slam_b200/csrc/brief_pattern.inc"
};

/* round-half-to-even, as cvRound (SSE cvtsd2si / lrint in default rounding mode) */
static inline int cv_round_d(double v) { return (int)lrint(v); }
static inline int cv_round_f(float v) { return (int)lrintf(v); }

/* ------------------------------------------------------------------------------------------------
 * Level geometry and per-level feature budget   (OpenCV orb.cpp: getScale, detectAndCompute,
 * computeKeyPoints; call site orb_feature.cpp:25 -- ORB::create(1000) => scaleFactor 1.2f, 8 levels)
 * ---------------------------------------------------------------------------------------------- */
ORC_API void orc_level_geometry(int W, int H, int nlevels, float scale_factor, int* w, int* h, float* scale)
{
    for(int l = 0; l < nlevels; ++l)
    {
        float s = (float)pow((double)scale_factor, (double)l);
        scale[l] = s;
        /* orb.cpp: float inv_scale = 1.f / scale; Size sz(cvRound(cols * inv_scale), cvRound(rows * inv_scale)).
         * Pinned by tests/golden/synthetic.npz (237 rows -> 198 at level 1; a division would give 197). */
        float inv = 1.f / s;
        w[l] = cv_round_f((float)W * inv);
        h[l] = cv_round_f((float)H * inv);
    }
}

ORC_API void orc_features_per_level(int nfeatures, int nlevels, float scale_factor, int* n_l)
{
    float factor = (float)(1.0 / (double)scale_factor);
    float nd = (float)nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
    int sum = 0;
    for(int l = 0; l < nlevels - 1; ++l)
    {
        n_l[l] = cv_round_f(nd);
        sum += n_l[l];
        nd *= factor;
    }
    n_l[nlevels - 1] = nfeatures - sum > 0 ? nfeatures - sum : 0;
}

/* ------------------------------------------------------------------------------------------------
 * toGrayScale  (frame.cpp:6-27; weights :18-19).  Three rounded f32 products, two rounded f32 adds
 * left to right, min(255.f, .), truncating cast.  Channel weights are applied in MEMORY order.
 * ---------------------------------------------------------------------------------------------- */
ORC_API void orc_gray(const uint8_t* rgb, long npx, uint8_t* gray)
{
    for(long i = 0; i < npx; ++i)
    {
        float r = (float)rgb[3 * i], g = (float)rgb[3 * i + 1], b = (float)rgb[3 * i + 2];
        float v = 0.299f * r + 0.587f * g + 0.114f * b;
        if(v > 255.f)
            v = 255.f;
        gray[i] = (uint8_t)v;
    }
}

/* ------------------------------------------------------------------------------------------------
 * INTER_LINEAR_EXACT resize (OpenCV imgproc resize.cpp resize_bitExact, 8.8 fixed point taps,
 * horizontal first).  SURVEY 8(c) item 2.
 * ---------------------------------------------------------------------------------------------- */
ORC_API void orc_resize_coeffs(int src, int dst, int* ofs, int* a1)
{
    /* all in IEEE double, no FMA (file is compiled with -ffp-contract=off) */
    double inv = (double)dst / (double)src;
    double scale = 1.0 / inv;
    for(int v = 0; v < dst; ++v)
    {
        double f = scale * ((double)v + 0.5) - 0.5;
        int i = (int)floor(f);
        if(i < 0)
        {
            ofs[v] = 0;
            a1[v] = 0;
        }
        else if(i >= src - 1)
        {
            ofs[v] = src - 1;
            a1[v] = 0;
        }
        else
        {
            ofs[v] = i;
            a1[v] = cv_round_d((f - (double)i) * 256.0);
        }
    }
}

ORC_API void orc_resize_exact(const uint8_t* src, int sw, int sh, uint8_t* dst, int dw, int dh)
{
    int* xo = (int*)malloc(sizeof(int) * (size_t)dw * 2);
    int* xa = xo + dw;
    int* yo = (int*)malloc(sizeof(int) * (size_t)dh * 2);
    int* ya = yo + dh;
    orc_resize_coeffs(sw, dw, xo, xa);
    orc_resize_coeffs(sh, dh, yo, ya);
    for(int y = 0; y < dh; ++y)
    {
        const uint8_t* r0 = src + (size_t)yo[y] * sw;
        const uint8_t* r1 = src + (size_t)(yo[y] + 1 < sh ? yo[y] + 1 : sh - 1) * sw;
        uint32_t b1 = (uint32_t)ya[y], b0 = 256u - b1;
        for(int x = 0; x < dw; ++x)
        {
            int i0 = xo[x], i1 = i0 + 1 < sw ? i0 + 1 : sw - 1;
            uint32_t a1 = (uint32_t)xa[x], a0 = 256u - a1;
            uint32_t h0 = a0 * r0[i0] + a1 * r0[i1];
            uint32_t h1 = a0 * r1[i0] + a1 * r1[i1];
            dst[(size_t)y * dw + x] = (uint8_t)((b0 * h0 + b1 * h1 + 32768u) >> 16);
        }
    }
    free(xo);
    free(yo);
}

/* ------------------------------------------------------------------------------------------------
 * FAST-9/16 score (OpenCV fast.cpp FAST_t<16> + fast_score.cpp cornerScore<16>), SURVEY 8(c) item 3.
 * score map: 0 where not a corner, else cornerScore (>= threshold).  Rows/cols outside [3, dim-3) are 0.
 * ---------------------------------------------------------------------------------------------- */
static const int kRingDx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int kRingDy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

static inline int fast_score_px(const uint8_t* p, int stride, int thr)
{
    int v = p[0];
    int d[25];
    /* quick reject: any 9-arc contains ring pixel 0 or 8, and 4 or 12 */
    int d0 = v - p[3 * stride], d8 = v - p[-3 * stride];
    int hi = (d0 > thr) | (d8 > thr), lo = (d0 < -thr) | (d8 < -thr);
    if(!(hi | lo))
        return 0;
    for(int k = 0; k < 16; ++k)
        d[k] = v - p[kRingDy[k] * stride + kRingDx[k]];
    for(int k = 16; k < 25; ++k)
        d[k] = d[k - 16];
    int best = 0;
    for(int k = 0; k < 16; ++k)
    {
        int mn = d[k], mx = d[k];
        for(int j = 1; j < 9; ++j)
        {
            if(d[k + j] < mn) mn = d[k + j];
            if(d[k + j] > mx) mx = d[k + j];
        }
        if(mn > best) best = mn;   /* centre brighter-than-arc by at least mn ("dark arc") */
        if(-mx > best) best = -mx; /* centre darker-than-arc */
    }
    /* corner at threshold thr  <=>  best > thr ; cornerScore = best - 1 */
    return best > thr ? best - 1 : 0;
}

/* the per-pixel score, for dist_orb_oracle.c (cv::FAST on sub-images) */
ORC_API int orc_fast_score_px(const uint8_t* p, int stride, int thr) { return fast_score_px(p, stride, thr); }

ORC_API void orc_fast_score_map(const uint8_t* img, int w, int h, int thr, uint8_t* score)
{
    memset(score, 0, (size_t)w * h);
    for(int y = 3; y < h - 3; ++y)
        for(int x = 3; x < w - 3; ++x)
            score[(size_t)y * w + x] = (uint8_t)fast_score_px(img + (size_t)y * w + x, w, thr);
}

/* 3x3 strict non-max suppression + border filter (KeyPointsFilter::runByImageBorder, edge=31).
 * Emits (x, y, score) triples in raster order; returns the count (may exceed cap: caller checks). */
ORC_API int orc_fast_nms_border(const uint8_t* score, int w, int h, int edge, int* xys, int cap)
{
    int n = 0;
    for(int y = edge; y < h - edge; ++y)
        for(int x = edge; x < w - edge; ++x)
        {
            int s = score[(size_t)y * w + x];
            if(!s)
                continue;
            const uint8_t* c = score + (size_t)y * w + x;
            if(s > c[-1] && s > c[1] && s > c[-w - 1] && s > c[-w] && s > c[-w + 1] && s > c[w - 1] && s > c[w] &&
               s > c[w + 1])
            {
                if(n < cap)
                {
                    xys[3 * n] = x;
                    xys[3 * n + 1] = y;
                    xys[3 * n + 2] = s;
                }
                ++n;
            }
        }
    return n;
}

/* ------------------------------------------------------------------------------------------------
 * Harris response at integer (x,y) (orb.cpp HarrisResponses, blockSize 7, k 0.04), item 6.
 * ---------------------------------------------------------------------------------------------- */
ORC_API float orc_harris(const uint8_t* img, int stride, int x, int y)
{
    int a = 0, b = 0, c = 0;
    for(int dy = -3; dy <= 3; ++dy)
        for(int dx = -3; dx <= 3; ++dx)
        {
            const uint8_t* p = img + (size_t)(y + dy) * stride + (x + dx);
            int Ix = (p[1] - p[-1]) * 2 + (p[-stride + 1] - p[-stride - 1]) + (p[stride + 1] - p[stride - 1]);
            int Iy = (p[stride] - p[-stride]) * 2 + (p[stride - 1] - p[-stride - 1]) + (p[stride + 1] - p[-stride + 1]);
            a += Ix * Ix;
            b += Iy * Iy;
            c += Ix * Iy;
        }
    float scale = 1.f / ((1 << 2) * 7 * 255.f);
    float s4 = scale * scale * scale * scale;
    float fa = (float)a, fb = (float)b, fc = (float)c;
    return (fa * fb - fc * fc - 0.04f * (fa + fb) * (fa + fb)) * s4;
}

/* ------------------------------------------------------------------------------------------------
 * cv::fastAtan2 (core mathfuncs_core atan_f32), degrees, item 7.
 * ---------------------------------------------------------------------------------------------- */
ORC_API float orc_fast_atan2(float y, float x)
{
    const float p1 = 0.9997878412794807f * (float)(180 / 3.1415926535897932384626433832795);
    const float p3 = -0.3258083974640975f * (float)(180 / 3.1415926535897932384626433832795);
    const float p5 = 0.1555786518463281f * (float)(180 / 3.1415926535897932384626433832795);
    const float p7 = -0.04432655554792128f * (float)(180 / 3.1415926535897932384626433832795);
    float ax = fabsf(x), ay = fabsf(y), a, c, c2;
    if(ax >= ay)
    {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    else
    {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if(x < 0) a = 180.f - a;
    if(y < 0) a = 360.f - a;
    return a;
}

static const int kUmax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

/* intensity-centroid angle (orb.cpp ICAngles), on the UNBLURRED level */
ORC_API float orc_ic_angle(const uint8_t* img, int stride, int x, int y)
{
    const uint8_t* c = img + (size_t)y * stride + x;
    int m01 = 0, m10 = 0;
    for(int u = -15; u <= 15; ++u)
        m10 += u * c[u];
    for(int v = 1; v <= 15; ++v)
    {
        int vs = 0, d = kUmax[v];
        for(int u = -d; u <= d; ++u)
        {
            int vp = c[u + v * stride], vm = c[u - v * stride];
            vs += vp - vm;
            m10 += u * (vp + vm);
        }
        m01 += v * vs;
    }
    return orc_fast_atan2((float)m01, (float)m10);
}

/* ------------------------------------------------------------------------------------------------
 * 7x7 sigma=2 separable float blur, BORDER_REFLECT_101 (sepFilter2D float path), item 9.
 * ---------------------------------------------------------------------------------------------- */
static inline float f32_from_bits(uint32_t u)
{
    float f;
    memcpy(&f, &u, 4);
    return f;
}
static inline int reflect101(int p, int n)
{
    if(n == 1) return 0;
    while(p < 0 || p >= n)
        p = p < 0 ? -p : 2 * n - 2 - p;
    return p;
}

ORC_API void orc_blur7(const uint8_t* src, int w, int h, uint8_t* dst)
{
    const float k[7] = {f32_from_bits(0x3d8fafb1u), f32_from_bits(0x3e06387eu), f32_from_bits(0x3e434a39u),
                        f32_from_bits(0x3e5d4ae0u), f32_from_bits(0x3e434a39u), f32_from_bits(0x3e06387eu),
                        f32_from_bits(0x3d8fafb1u)};
    float* rows = (float*)malloc(sizeof(float) * (size_t)w * h);
    for(int y = 0; y < h; ++y)
    {
        const uint8_t* s = src + (size_t)y * w;
        float* r = rows + (size_t)y * w;
        for(int x = 0; x < w; ++x)
        {
            float acc = k[0] * (float)s[reflect101(x - 3, w)];
            for(int j = 1; j < 7; ++j)
                acc = fmaf(k[j], (float)s[reflect101(x - 3 + j, w)], acc);
            r[x] = acc;
        }
    }
    for(int y = 0; y < h; ++y)
    {
        const float* r0 = rows + (size_t)y * w;
        const float *rp[4], *rm[4];
        for(int i = 1; i <= 3; ++i)
        {
            rp[i] = rows + (size_t)reflect101(y + i, h) * w;
            rm[i] = rows + (size_t)reflect101(y - i, h) * w;
        }
        for(int x = 0; x < w; ++x)
        {
            float acc = k[3] * r0[x];
            for(int i = 1; i <= 3; ++i)
                acc = fmaf(k[3 + i], rp[i][x] + rm[i][x], acc);
            int v = (int)lrintf(acc);
            dst[(size_t)y * w + x] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
        }
    }
    free(rows);
}

/* ------------------------------------------------------------------------------------------------
 * rotated BRIEF-256 (orb.cpp computeOrbDescriptors, WTA_K=2), item 10.
 * (cx,cy) integer centre on the BLURRED level, angle in degrees.
 * ---------------------------------------------------------------------------------------------- */
ORC_API void orc_describe(const uint8_t* blurred, int stride, int cx, int cy, float angle_deg, uint8_t* desc)
{
    float ang = angle_deg * (float)(3.1415926535897932384626433832795 / 180.f);
    float a = (float)cos((double)ang), b = (float)sin((double)ang);
    const uint8_t* c = blurred + (size_t)cy * stride + cx;
    for(int i = 0; i < 32; ++i)
    {
        int byte = 0;
        for(int j = 0; j < 8; ++j)
        {
            const int8_t* t = kPattern + (i * 8 + j) * 4;
            float x0 = (float)t[0] * a - (float)t[1] * b, y0 = (float)t[0] * b + (float)t[1] * a;
            float x1 = (float)t[2] * a - (float)t[3] * b, y1 = (float)t[2] * b + (float)t[3] * a;
            int v0 = c[cv_round_f(y0) * stride + cv_round_f(x0)];
            int v1 = c[cv_round_f(y1) * stride + cv_round_f(x1)];
            byte |= (v0 < v1) << j;
        }
        desc[i] = (uint8_t)byte;
    }
}

/* ------------------------------------------------------------------------------------------------
 * KeyPointsFilter::retainBest threshold: k-th largest value (ties kept by the caller).
 * ---------------------------------------------------------------------------------------------- */
static int cmp_float_desc(const void* a, const void* b)
{
    float x = *(const float*)a, y = *(const float*)b;
    return x > y ? -1 : x < y ? 1 : 0;
}
static float kth_largest(const float* v, int n, int k)
{
    float* t = (float*)malloc(sizeof(float) * (size_t)n);
    memcpy(t, v, sizeof(float) * (size_t)n);
    qsort(t, (size_t)n, sizeof(float), cmp_float_desc);
    float r = t[k - 1];
    free(t);
    return r;
}

typedef struct
{
    float x, y, size, angle, response;
    int octave;
} orc_kp;

typedef struct
{
    int nfeatures, nlevels, fast_threshold, edge;
    float scale_factor;
} orc_params;

/* Optional stage dumps (any pointer may be NULL):
 *   pyr     : concatenated unblurred levels (dense, w_l*h_l each)
 *   blurred : concatenated blurred levels
 *   n_fast  : per level, number of NMS+border survivors before retainBest
 *   n_cut1  : per level, survivors of retainBest(2*n_l) on the FAST score */
typedef struct
{
    uint8_t* pyr;
    uint8_t* blurred;
    int* n_fast;
    int* n_cut1;
} orc_dumps;

/* cv::ORB::detectAndCompute restated; keypoints are emitted in canonical order (octave, y, x)
 * (OpenCV's own order inside a level is a std::nth_element artefact -- compare as sets).
 * Returns the keypoint count (all are written if <= cap, else nothing beyond cap). */
ORC_API int orc_orb_detect_gray(const uint8_t* gray, int W, int H, const orc_params* prm, orc_kp* kps, uint8_t* desc,
                                int cap, const orc_dumps* dump)
{
    int L = prm->nlevels;
    int w[32], h[32], nl[32];
    float sc[32];
    orc_level_geometry(W, H, L, prm->scale_factor, w, h, sc);
    orc_features_per_level(prm->nfeatures, L, prm->scale_factor, nl);
    size_t off[33];
    off[0] = 0;
    for(int l = 0; l < L; ++l)
        off[l + 1] = off[l] + (size_t)w[l] * h[l];
    uint8_t* pyr = (uint8_t*)malloc(off[L]);
    uint8_t* blr = (uint8_t*)malloc(off[L]);
    memcpy(pyr, gray, (size_t)W * H);
    for(int l = 1; l < L; ++l)
        orc_resize_exact(pyr + off[l - 1], w[l - 1], h[l - 1], pyr + off[l], w[l], h[l]);

    int total = 0;
    for(int l = 0; l < L; ++l)
    {
        const uint8_t* img = pyr + off[l];
        int wl = w[l], hl = h[l];
        int n_fast = 0, n_cut1 = 0, m = 0;
        int* xys = NULL;
        float* resp = NULL;
        if(wl > 2 * prm->edge && hl > 2 * prm->edge)
        {
            uint8_t* score = (uint8_t*)malloc((size_t)wl * hl);
            orc_fast_score_map(img, wl, hl, prm->fast_threshold, score);
            int cap_l = (wl * hl) / 4 + 16;
            xys = (int*)malloc(sizeof(int) * 3 * (size_t)cap_l);
            n_fast = orc_fast_nms_border(score, wl, hl, prm->edge, xys, cap_l);
            free(score);
            /* retainBest(2*n_l) on FAST score */
            int k1 = 2 * nl[l];
            m = n_fast;
            if(m > k1)
            {
                if(k1 == 0)
                    m = 0;
                else
                {
                    float* s = (float*)malloc(sizeof(float) * (size_t)m);
                    for(int i = 0; i < m; ++i)
                        s[i] = (float)xys[3 * i + 2];
                    float thr = kth_largest(s, m, k1);
                    free(s);
                    int j = 0;
                    for(int i = 0; i < m; ++i)
                        if((float)xys[3 * i + 2] >= thr)
                        {
                            xys[3 * j] = xys[3 * i];
                            xys[3 * j + 1] = xys[3 * i + 1];
                            xys[3 * j + 2] = xys[3 * i + 2];
                            ++j;
                        }
                    m = j;
                }
            }
            n_cut1 = m;
            /* Harris, retainBest(n_l) */
            resp = (float*)malloc(sizeof(float) * (size_t)(m + 1));
            for(int i = 0; i < m; ++i)
                resp[i] = orc_harris(img, wl, xys[3 * i], xys[3 * i + 1]);
            int k2 = nl[l];
            if(m > k2)
            {
                if(k2 == 0)
                    m = 0;
                else
                {
                    float thr = kth_largest(resp, m, k2);
                    int j = 0;
                    for(int i = 0; i < m; ++i)
                        if(resp[i] >= thr)
                        {
                            xys[3 * j] = xys[3 * i];
                            xys[3 * j + 1] = xys[3 * i + 1];
                            resp[j] = resp[i];
                            ++j;
                        }
                    m = j;
                }
            }
        }
        if(dump && dump->n_fast) dump->n_fast[l] = n_fast;
        if(dump && dump->n_cut1) dump->n_cut1[l] = n_cut1;
        /* blur this level (detectAndCompute blurs every level before describing) */
        orc_blur7(img, wl, hl, blr + off[l]);
        /* raster order of orc_fast_nms_border is preserved by both cuts => already (y, x) sorted */
        for(int i = 0; i < m; ++i)
        {
            int x = xys[3 * i], y = xys[3 * i + 1];
            if(total < cap)
            {
                orc_kp* kp = kps + total;
                kp->angle = orc_ic_angle(img, wl, x, y);
                kp->x = (float)x * sc[l];
                kp->y = (float)y * sc[l];
                kp->size = 31.f * sc[l];
                kp->response = resp[i];
                kp->octave = l;
                float inv = 1.f / sc[l];
                int cx = cv_round_f(kp->x * inv), cy = cv_round_f(kp->y * inv);
                orc_describe(blr + off[l], wl, cx, cy, kp->angle, desc + (size_t)total * 32);
            }
            ++total;
        }
        free(xys);
        free(resp);
    }
    if(dump && dump->pyr) memcpy(dump->pyr, pyr, off[L]);
    if(dump && dump->blurred) memcpy(dump->blurred, blr, off[L]);
    free(pyr);
    free(blr);
    return total;
}

/* OrbOpenCvDetector::detect restated: toGrayScale then ORB (orb_feature.cpp:35-40). */
ORC_API int orc_orb_detect_rgb(const uint8_t* rgb, int W, int H, const orc_params* prm, orc_kp* kps, uint8_t* desc,
                               int cap)
{
    uint8_t* gray = (uint8_t*)malloc((size_t)W * H);
    orc_gray(rgb, (long)W * H, gray);
    int n = orc_orb_detect_gray(gray, W, H, prm, kps, desc, cap, NULL);
    free(gray);
    return n;
}

/* batch of independent frames, one OpenMP thread per frame (timed CPU baseline "port") */
ORC_API void orc_orb_detect_rgb_batch(const uint8_t* rgb, int nframes, int W, int H, const orc_params* prm, orc_kp* kps,
                                      uint8_t* desc, int cap, int* counts, int nthreads)
{
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads > 0 ? nthreads : 1)
    for(int f = 0; f < nframes; ++f)
        counts[f] = orc_orb_detect_rgb(rgb + (size_t)f * W * H * 3, W, H, prm, kps + (size_t)f * cap,
                                       desc + (size_t)f * cap * 32, cap);
}

/* ------------------------------------------------------------------------------------------------
 * Matching.  knn k=2: per query the two smallest Hamming distances, ties -> lower train index
 * (cv::batchDistance).
 * ---------------------------------------------------------------------------------------------- */
static inline int hamming256(const uint8_t* a, const uint8_t* b)
{
    uint64_t x[4], y[4];
    memcpy(x, a, 32);
    memcpy(y, b, 32);
    return __builtin_popcountll(x[0] ^ y[0]) + __builtin_popcountll(x[1] ^ y[1]) + __builtin_popcountll(x[2] ^ y[2]) +
           __builtin_popcountll(x[3] ^ y[3]);
}

/* out arrays sized n_query: idx0,d0,idx1,d1 ; idx = -1 / d = -1 where train has fewer rows */
ORC_API void orc_knn2(const uint8_t* query, int nq, int qstride, const uint8_t* train, int nt, int tstride, int* idx0,
                      int* d0, int* idx1, int* d1)
{
    for(int q = 0; q < nq; ++q)
    {
        int b0 = 1 << 30, b1 = 1 << 30, i0 = -1, i1 = -1;
        const uint8_t* qd = query + (size_t)q * qstride;
        for(int t = 0; t < nt; ++t)
        {
            int d = hamming256(qd, train + (size_t)t * tstride);
            if(d < b0)
            {
                b1 = b0;
                i1 = i0;
                b0 = d;
                i0 = t;
            }
            else if(d < b1)
            {
                b1 = d;
                i1 = t;
            }
        }
        idx0[q] = i0;
        d0[q] = i0 < 0 ? -1 : b0;
        idx1[q] = i1;
        d1[q] = i1 < 0 ? -1 : b1;
    }
}

/* OrbOpenCvMatcher::match(from, to) restated (orb_feature.cpp:84-117): query = to, train = from,
 * keep m[0] iff d0 < 0.7*d1 (double compare, as the reference writes it), output
 * {fromIndex = trainIdx, toIndex = queryIdx} ascending toIndex.  The reference is undefined for empty
 * inputs / from.size() < 2; the replacement contract is "no match", restated here. */
ORC_API int orc_match_ratio(const uint8_t* from, int nfrom, int fstride, const uint8_t* to, int nto, int tstride,
                            double ratio, int64_t* from_idx, int64_t* to_idx, int* dist)
{
    if(nfrom < 2 || nto < 1)
        return 0;
    int n = 0;
    int* buf = (int*)malloc(sizeof(int) * 4 * (size_t)nto);
    orc_knn2(to, nto, tstride, from, nfrom, fstride, buf, buf + nto, buf + 2 * nto, buf + 3 * nto);
    for(int q = 0; q < nto; ++q)
    {
        float fd0 = (float)buf[nto + q], fd1 = (float)buf[3 * nto + q];
        if((double)fd0 < ratio * (double)fd1)
        {
            from_idx[n] = buf[q];
            to_idx[n] = q;
            dist[n] = buf[nto + q];
            ++n;
        }
    }
    free(buf);
    return n;
}

/* BFMatcher(NORM_HAMMING, crossCheck=true).match(query, train) as cv2 4.13 computes it:
 * mutual nearest neighbours, lowest-index argmin on both sides, ascending query. */
ORC_API int orc_match_crosscheck(const uint8_t* query, int nq, int qstride, const uint8_t* train, int nt, int tstride,
                                 int* q_idx, int* t_idx, int* dist)
{
    if(nq < 1 || nt < 1)
        return 0;
    int* bq = (int*)malloc(sizeof(int) * (size_t)nq);
    int* bt = (int*)malloc(sizeof(int) * (size_t)nt);
    int* dq = (int*)malloc(sizeof(int) * (size_t)nq);
    int* dt = (int*)malloc(sizeof(int) * (size_t)nt);
    for(int q = 0; q < nq; ++q) dq[q] = 1 << 30, bq[q] = -1;
    for(int t = 0; t < nt; ++t) dt[t] = 1 << 30, bt[t] = -1;
    for(int q = 0; q < nq; ++q)
        for(int t = 0; t < nt; ++t)
        {
            int d = hamming256(query + (size_t)q * qstride, train + (size_t)t * tstride);
            if(d < dq[q]) dq[q] = d, bq[q] = t;
            if(d < dt[t]) dt[t] = d, bt[t] = q;
        }
    int n = 0;
    for(int q = 0; q < nq; ++q)
        if(bt[bq[q]] == q)
        {
            q_idx[n] = q;
            t_idx[n] = bq[q];
            dist[n] = dq[q];
            ++n;
        }
    free(bq);
    free(bt);
    free(dq);
    free(dt);
    return n;
}

/* Loop-candidate verification score:
 * score[q][k] = OrbOpenCvMatcher::match(from = query frame q, to = keyframe k).size(), integer ratio test
 * 10*d0 < 7*d1 (== the reference's float test for all 0 <= d <= 256, tests/test_oracle.py proves it). */
ORC_API void orc_db_scores(const uint8_t* queries, int nq_frames, const int* q_counts, int q_cap, const uint8_t* db,
                           int n_keyframes, const int* k_counts, int k_cap, int* scores, int nthreads)
{
#pragma omp parallel for schedule(dynamic, 1) collapse(2) num_threads(nthreads > 0 ? nthreads : 1)
    for(int q = 0; q < nq_frames; ++q)
        for(int k = 0; k < n_keyframes; ++k)
        {
            const uint8_t* qd = queries + (size_t)q * q_cap * 32;
            const uint8_t* kd = db + (size_t)k * k_cap * 32;
            int nfrom = q_counts[q], nto = k_counts[k], cnt = 0;
            if(nfrom >= 2)
                for(int i = 0; i < nto; ++i)
                {
                    int b0 = 1 << 30, b1 = 1 << 30;
                    for(int j = 0; j < nfrom; ++j)
                    {
                        int d = hamming256(kd + (size_t)i * 32, qd + (size_t)j * 32);
                        if(d < b0) b1 = b0, b0 = d;
                        else if(d < b1) b1 = d;
                    }
                    cnt += 10 * b0 < 7 * b1;
                }
            scores[(size_t)q * n_keyframes + k] = cnt;
        }
}

/* per query frame: top-k keyframes by (score desc, id asc); ids offset by id_base (shard start) */
ORC_API void orc_db_topk(const int* scores, int nq_frames, int n_keyframes, int id_base, int k, int* out_score,
                         int* out_id)
{
    for(int q = 0; q < nq_frames; ++q)
    {
        const int* s = scores + (size_t)q * n_keyframes;
        uint8_t* used = (uint8_t*)calloc((size_t)n_keyframes + 1, 1);
        for(int r = 0; r < k; ++r)
        {
            int best = -1;
            for(int i = 0; i < n_keyframes; ++i)
                if(!used[i] && (best < 0 || s[i] > s[best]))
                    best = i;
            if(best >= 0)
            {
                used[best] = 1;
                out_score[(size_t)q * k + r] = s[best];
                out_id[(size_t)q * k + r] = id_base + best;
            }
            else
            {
                out_score[(size_t)q * k + r] = -1;
                out_id[(size_t)q * k + r] = -1;
            }
        }
        free(used);
    }
}

/* match a batch of independent frame pairs (prev -> cur), OpenMP over pairs; CPU baseline helper */
ORC_API void orc_match_ratio_batch(const uint8_t* from, const int* nfrom, const uint8_t* to, const int* nto, int npairs,
                                   int cap, int64_t* from_idx, int64_t* to_idx, int* dist, int* counts, int nthreads)
{
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads > 0 ? nthreads : 1)
    for(int p = 0; p < npairs; ++p)
        counts[p] = orc_match_ratio(from + (size_t)p * cap * 32, nfrom[p], 32, to + (size_t)p * cap * 32, nto[p], 32, 0.7,
                                    from_idx + (size_t)p * cap, to_idx + (size_t)p * cap, dist + (size_t)p * cap);
}

/* ---- depth back-projection ---------------------------------------------------------------------
 * pointsFromRgbdKeypoints / reconstructPoint (src/lib/modular_slam/rgbd_feature_frontend.cpp:101-138) with getDepth /
 * isDepthValid (include/modular_slam/types/depth_frame.hpp:20-30):
 *   imgPoint = coordinates.cast<int>()  (truncation);  depth = (float)data[w * y + x] * factor  (one f32 product);
 *   nullopt unless depth > FLT_EPSILON;  z = depth;  x = (u - cx) * z * (1.0 / fx),  y likewise -- IEEE double, left to
 *   right, no FMA (this file is compiled with -ffp-contract=off).  xy: n records of two doubles, stride in bytes.
 * Parity unpinned: the reference holds no golden vector for this step and cannot be built here; the arithmetic is
 * restated from the lines cited. */
ORC_API void orc_backproject(const uint8_t* xy, int n, int stride_bytes, const uint16_t* depth, int width, int height,
                             double cx, double cy, double fx, double fy, float factor, double* out_xyz, uint8_t* out_valid)
{
    const double inv_fx = 1.0 / fx, inv_fy = 1.0 / fy;
    for(int i = 0; i < n; ++i)
    {
        const double* c = (const double*)(xy + (size_t)i * stride_bytes);
        const double u = c[0], v = c[1];
        out_xyz[3 * i] = out_xyz[3 * i + 1] = out_xyz[3 * i + 2] = 0.0;
        out_valid[i] = 0;
        if(!(u > -2147483648.0 && u < 2147483647.0 && v > -2147483648.0 && v < 2147483647.0))
            continue;
        const int xi = (int)u, yi = (int)v;
        if(xi < 0 || yi < 0 || xi >= width || yi >= height)
            continue; /* the reference reads out of bounds here */
        const float d = (float)(int)depth[(size_t)width * yi + xi] * factor;
        if(!(d > 1.1920928955078125e-07f))
            continue;
        const double z = (double)d;
        out_xyz[3 * i] = (u - cx) * z * inv_fx;
        out_xyz[3 * i + 1] = (v - cy) * z * inv_fy;
        out_xyz[3 * i + 2] = z;
        out_valid[i] = 1;
    }
}
