// pnp_kernels.cu -- SURVEY.md section 8(f) row f4: RANSAC PnP, batched over frames.
//
// Replaces OpenCvRansacPnp::solvePnp (src/lib/modular_slam/cv_ransac_pnp.cpp:15-85), i.e.
//   cv::solvePnPRansac(obj f32, img f32, K f32, noArray(), rvec, tvec, true, 100, 5.0, 0.99, inliers)      (:56-57)
// the consumer of the matcher's output in RgbdFeatureFrontend::track (rgbd_feature_frontend.cpp:345-361).
//
// cv::solvePnPRansac is a sequential loop (draw 5 points from cv::RNG, EPnP, count inliers, adapt the iteration count)
// followed by a Levenberg-Marquardt fit over the inliers.  Here all hypotheses of a frame are evaluated at once, and a
// launch covers a whole batch of frames: a cluster of two 512-thread CTAs per frame, of which rank 1 only helps with
// the inlier counting (phase 2) and leaves:
//   phase 1  thread h < iterations: 4 distinct indices from a counter-based hash, Grunert's P3P on the first three
//            (quartic in v = s3 / s1 from polynomial products, Ferrari roots + 2 Newton steps), triangle alignment
//            -> (R, t) per real root, the 4th point picks the root;  float64 throughout; both CTAs build all of them
//   phase 2  inlier counts: squared reprojection error <= thr^2 (the test of PnPRansacCallback::computeError +
//            findInliers, evaluated without the division); a warp takes two hypotheses x 256 points at a time, the
//            items alternate between the two CTAs, rank 1 adds its counts into rank 0's through distributed shared memory;
//   phase 3  best = most inliers, lowest h on ties, >= 5 inliers (cv: goodCount > modelPoints - 1); inlier mask written;
//   phase 4  Levenberg-Marquardt over the inliers from the best hypothesis: block-reduced 6x6 normal equations (halving
//            butterfly over the 27 sums), solved by one warp (row per lane, partial pivoting); the first damping trial of
//            an iteration runs alone, trials 2..8 together (warp per damping, one pass for the seven costs);
//   phase 5  rvec = log(R), tvec, and the reference's sensor pose (orientation = R^-1 as a quaternion,
//            position = -(R^-1 t), cv_ransac_pnp.cpp:65-77).
// Every step mirrors oracle/pnp_oracle.py (same draws, same formulas); parity with OpenCV itself is by tolerance
// (different minimal solver and random stream) -- DESIGN.md section 4e.
#include "common.cuh"
#include "kernels.h"

#include <cooperative_groups.h>
#include <type_traits>

namespace mslam_b200
{
namespace
{
constexpr int PNP_THREADS = 512;
constexpr int PNP_MAX_HYP = 256;
constexpr int PNP_CLUSTER = 2; // CTAs per frame: they share the inlier counting (phase 2), rank 0 carries on alone

__device__ __forceinline__ unsigned lowbias32(unsigned x)
{
    x ^= x >> 16;
    x *= 0x7FEB352Du;
    x ^= x >> 15;
    x *= 0x846CA68Bu;
    x ^= x >> 16;
    return x;
}
__device__ __forceinline__ int pnp_draw(unsigned seed, unsigned frame, unsigned hyp, unsigned j, int n)
{
    const unsigned r = lowbias32(seed ^ lowbias32(frame * 0x9E3779B9u + hyp * 0x85EBCA6Bu + j * 0xC2B2AE35u + 0x27D4EB2Fu));
    return (int)(((unsigned long long)r * (unsigned long long)n) >> 32);
}

struct V3
{
    double x, y, z;
};
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator*(V3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }

struct Pose
{
    double R[9]; // row major
    double t[3];
};

__device__ double cubic_largest_real_root(double p, double q, double r)
{
    const double Q = (p * p - 3.0 * q) / 9.0;
    const double R = (2.0 * p * p * p - 9.0 * p * q + 27.0 * r) / 54.0;
    const double Q3 = Q * Q * Q;
    if(R * R < Q3)
    {
        const double th = acos(fmax(-1.0, fmin(1.0, R / sqrt(Q3))));
        const double sq = -2.0 * sqrt(Q);
        const double pi = 3.14159265358979323846;
        const double y0 = sq * cos(th / 3.0), y1 = sq * cos((th + 2.0 * pi) / 3.0), y2 = sq * cos((th - 2.0 * pi) / 3.0);
        return fmax(y0, fmax(y1, y2)) - p / 3.0;
    }
    const double A = -copysign(1.0, R) * cbrt(fabs(R) + sqrt(fmax(R * R - Q3, 0.0)));
    const double B = A != 0.0 ? Q / A : 0.0;
    return A + B - p / 3.0;
}

// real roots of c4 x^4 + ... + c0; returns how many were written to out[4]
__device__ int quartic_real_roots(double c4, double c3, double c2, double c1, double c0, double* out)
{
    if(c4 == 0.0 || !isfinite(c4))
        return 0;
    const double a = c3 / c4, b = c2 / c4, c = c1 / c4, d = c0 / c4;
    if(!(isfinite(a) && isfinite(b) && isfinite(c) && isfinite(d)))
        return 0;
    const double a2 = a * a;
    const double y = cubic_largest_real_root(-b, a * c - 4.0 * d, 4.0 * b * d - a2 * d - c * c);
    const double R2 = 0.25 * a2 - b + y;
    const double scale = fabs(0.25 * a2) + fabs(b) + fabs(y) + 1e-300;
    if(!(R2 >= -1e-12 * scale))
        return 0;
    double R = sqrt(fmax(R2, 0.0));
    double D2, E2;
    if(R > 1e-8 * sqrt(scale))
    {
        const double t = 0.25 * (4.0 * a * b - 8.0 * c - a2 * a) / R;
        D2 = 0.75 * a2 - R * R - 2.0 * b + t;
        E2 = 0.75 * a2 - R * R - 2.0 * b - t;
    }
    else
    {
        R = 0.0;
        double w = y * y - 4.0 * d;
        if(w < 0.0)
            w = 0.0;
        D2 = 0.75 * a2 - 2.0 * b + 2.0 * sqrt(w);
        E2 = 0.75 * a2 - 2.0 * b - 2.0 * sqrt(w);
    }
    const double tol = 1e-10 * (fabs(0.75 * a2) + fabs(2.0 * b) + 1e-300);
    int n = 0;
    if(D2 >= -tol)
    {
        const double D = sqrt(fmax(D2, 0.0));
        out[n++] = -0.25 * a + 0.5 * R + 0.5 * D;
        out[n++] = -0.25 * a + 0.5 * R - 0.5 * D;
    }
    if(E2 >= -tol)
    {
        const double E = sqrt(fmax(E2, 0.0));
        out[n++] = -0.25 * a - 0.5 * R + 0.5 * E;
        out[n++] = -0.25 * a - 0.5 * R - 0.5 * E;
    }
    for(int i = 0; i < n; ++i)
    {
        double x = out[i];
        for(int k = 0; k < 2; ++k)
        {
            const double f = (((x + a) * x + b) * x + c) * x + d;
            const double fp = ((4.0 * x + 3.0 * a) * x + 2.0 * b) * x + c;
            if(fp != 0.0)
                x = x - f / fp;
        }
        out[i] = x;
    }
    return n;
}

// orthonormal frame of a triangle: columns e1, e2, e3 (B[r][c], row major)
__device__ bool triad(V3 p0, V3 p1, V3 p2, double* B)
{
    V3 e1 = p1 - p0;
    const double n1 = sqrt(dot(e1, e1));
    if(n1 == 0.0)
        return false;
    e1 = e1 * (1.0 / n1);
    V3 e3 = cross(e1, p2 - p0);
    const double n3 = sqrt(dot(e3, e3));
    if(n3 == 0.0)
        return false;
    e3 = e3 * (1.0 / n3);
    const V3 e2 = cross(e3, e1);
    B[0] = e1.x, B[1] = e2.x, B[2] = e3.x;
    B[3] = e1.y, B[4] = e2.y, B[5] = e3.y;
    B[6] = e1.z, B[7] = e2.z, B[8] = e3.z;
    return true;
}

struct Cam
{
    double fx, fy, cx, cy;
};

__device__ __forceinline__ V3 load_obj(const float* __restrict__ obj, int i)
{
    return {(double)__ldg(obj + 3 * i), (double)__ldg(obj + 3 * i + 1), (double)__ldg(obj + 3 * i + 2)};
}
__device__ __forceinline__ V3 bearing(const float* __restrict__ img, int i, const Cam& K)
{
    const double bx = ((double)__ldg(img + 2 * i) - K.cx) / K.fx, by = ((double)__ldg(img + 2 * i + 1) - K.cy) / K.fy;
    const double nn = sqrt(bx * bx + by * by + 1.0);
    return {bx / nn, by / nn, 1.0 / nn};
}
__device__ __forceinline__ V3 transform(const double* R, const double* t, V3 X)
{
    return {R[0] * X.x + R[1] * X.y + R[2] * X.z + t[0], R[3] * X.x + R[4] * X.y + R[5] * X.z + t[1],
            R[6] * X.x + R[7] * X.y + R[8] * X.z + t[2]};
}
__device__ __forceinline__ double err2(const double* R, const double* t, V3 X, double u0, double v0, const Cam& K)
{
    const V3 c = transform(R, t, X);
    const double u = K.fx * (c.x / c.z) + K.cx, v = K.fy * (c.y / c.z) + K.cy;
    const double du = u - u0, dv = v - v0;
    return du * du + dv * dv;
}

// the inlier test |proj(X) - (u0, v0)|^2 <= thr2 without the division: with c = R X + t,
//   (fx c.x + (cx - u0) c.z)^2 + (fy c.y + (cy - v0) c.z)^2 <= thr2 c.z^2          (both sides multiplied by c.z^2 > 0)
// c.z == 0 (where the projection is infinite) is never an inlier.  oracle/pnp_oracle.py::inlier_mask is the same sequence.
__device__ __forceinline__ bool is_inlier(const double* R, const double* t, V3 X, double u0, double v0, const Cam& K, double thr2)
{
    const V3 c = transform(R, t, X);
    const double a = K.fx * c.x + (K.cx - u0) * c.z, b = K.fy * c.y + (K.cy - v0) * c.z;
    return c.z != 0.0 && a * a + b * b <= thr2 * (c.z * c.z);
}

// one hypothesis from points idx[0..3]; false if the sample yields no pose
__device__ bool p3p_hypothesis(const float* __restrict__ obj, const float* __restrict__ img, const int* idx, const Cam& K,
                               Pose& best)
{
    const V3 P0 = load_obj(obj, idx[0]), P1 = load_obj(obj, idx[1]), P2 = load_obj(obj, idx[2]), P3 = load_obj(obj, idx[3]);
    const V3 f0 = bearing(img, idx[0], K), f1 = bearing(img, idx[1], K), f2 = bearing(img, idx[2], K);
    const double a2 = dot(P1 - P2, P1 - P2), b2 = dot(P0 - P2, P0 - P2), c2 = dot(P0 - P1, P0 - P1);
    if(a2 == 0.0 || b2 == 0.0 || c2 == 0.0)
        return false;
    const double ca = dot(f1, f2), cb = dot(f0, f2), cg = dot(f0, f1);
    const double k1 = (a2 - c2) / b2, k2 = c2 / b2;
    const double n2 = k1 - 1.0, n1 = -2.0 * k1 * cb, n0 = 1.0 + k1;
    const double d1 = -2.0 * ca, d0 = 2.0 * cg;
    const double NN[5] = {n0 * n0, 2.0 * n0 * n1, n1 * n1 + 2.0 * n0 * n2, 2.0 * n1 * n2, n2 * n2};
    const double ND[5] = {n0 * d0, n0 * d1 + n1 * d0, n1 * d1 + n2 * d0, n2 * d1, 0.0};
    const double DD[3] = {d0 * d0, 2.0 * d0 * d1, d1 * d1};
    const double q0 = 1.0 - k2, q1 = 2.0 * k2 * cb, q2 = -k2;
    const double DQ[5] = {DD[0] * q0, DD[0] * q1 + DD[1] * q0, DD[0] * q2 + DD[1] * q1 + DD[2] * q0, DD[1] * q2 + DD[2] * q1,
                          DD[2] * q2};
    double co[5];
#pragma unroll
    for(int i = 0; i < 5; ++i)
        co[i] = NN[i] - 2.0 * cg * ND[i] + DQ[i];
    double Bw[9];
    if(!triad(P0, P1, P2, Bw))
        return false;
    double roots[4];
    const int nr = quartic_real_roots(co[4], co[3], co[2], co[1], co[0], roots);
    const double u3 = (double)__ldg(img + 2 * idx[3]), v3 = (double)__ldg(img + 2 * idx[3] + 1);
    bool found = false;
    double best_e = 0.0;
    for(int k = 0; k < nr; ++k)
    {
        const double v = roots[k];
        if(!(v > 0.0) || !isfinite(v))
            continue;
        const double den = d1 * v + d0;
        if(fabs(den) < 1e-12)
            continue;
        const double u = ((n2 * v + n1) * v + n0) / den;
        if(!(u > 0.0))
            continue;
        const double w = 1.0 + v * v - 2.0 * v * cb;
        if(!(w > 0.0))
            continue;
        const double s1 = sqrt(b2 / w);
        const V3 C0 = f0 * s1, C1 = f1 * (u * s1), C2 = f2 * (v * s1);
        double Bc[9];
        if(!triad(C0, C1, C2, Bc))
            continue;
        Pose p;
#pragma unroll
        for(int r = 0; r < 3; ++r)
#pragma unroll
            for(int c = 0; c < 3; ++c) // R = Bc Bw^T
                p.R[3 * r + c] = Bc[3 * r] * Bw[3 * c] + Bc[3 * r + 1] * Bw[3 * c + 1] + Bc[3 * r + 2] * Bw[3 * c + 2];
        const double zero[3] = {0.0, 0.0, 0.0};
        const V3 RP0 = transform(p.R, zero, P0);
        p.t[0] = C0.x - RP0.x, p.t[1] = C0.y - RP0.y, p.t[2] = C0.z - RP0.z;
        const double e = err2(p.R, p.t, P3, u3, v3, K);
        const double zc = transform(p.R, p.t, P3).z;
        if(zc > 0.0 && (!found || e < best_e) && e < INFINITY)
        {
            found = true;
            best_e = e;
            best = p;
        }
    }
    return found;
}

__device__ void rodrigues_exp(const double* w, double* R)
{
    const double th = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
    const double K[9] = {0.0, -w[2], w[1], w[2], 0.0, -w[0], -w[1], w[0], 0.0};
    double A = 1.0, B = 0.0;
    if(th >= 1e-12)
    {
        A = sin(th) / th;
        B = (1.0 - cos(th)) / (th * th);
    }
#pragma unroll
    for(int r = 0; r < 3; ++r)
#pragma unroll
        for(int c = 0; c < 3; ++c)
        {
            const double kk = K[3 * r] * K[c] + K[3 * r + 1] * K[3 + c] + K[3 * r + 2] * K[6 + c];
            R[3 * r + c] = (r == c ? 1.0 : 0.0) + A * K[3 * r + c] + B * kk;
        }
}

__device__ void rodrigues_log(const double* R, double* out)
{
    double c = 0.5 * (R[0] + R[4] + R[8] - 1.0);
    c = fmax(-1.0, fmin(1.0, c));
    const double th = acos(c);
    const double r[3] = {R[7] - R[5], R[2] - R[6], R[3] - R[1]};
    const double s = 0.5 * sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
    if(s < 1e-10)
    {
        if(c > 0.0)
        {
            out[0] = 0.5 * r[0], out[1] = 0.5 * r[1], out[2] = 0.5 * r[2];
            return;
        }
        double d[3] = {sqrt(fmax((R[0] + 1.0) * 0.5, 0.0)), sqrt(fmax((R[4] + 1.0) * 0.5, 0.0)), sqrt(fmax((R[8] + 1.0) * 0.5, 0.0))};
        if(R[1] < 0.0)
            d[1] = -d[1];
        if(R[2] < 0.0)
            d[2] = -d[2];
        const double n = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        const double k = n > 0.0 ? th / n : 0.0;
        out[0] = d[0] * k, out[1] = d[1] * k, out[2] = d[2] * k;
        return;
    }
    const double k = 0.5 * th / s;
    out[0] = r[0] * k, out[1] = r[1] * k, out[2] = r[2] * k;
}

// rotation matrix (row major) -> unit quaternion (w, x, y, z), w >= 0
__device__ void quat_from_R(const double* R, double* q)
{
    const double tr = R[0] + R[4] + R[8];
    if(tr > 0.0)
    {
        const double s = sqrt(tr + 1.0) * 2.0;
        q[0] = 0.25 * s, q[1] = (R[7] - R[5]) / s, q[2] = (R[2] - R[6]) / s, q[3] = (R[3] - R[1]) / s;
    }
    else if(R[0] > R[4] && R[0] > R[8])
    {
        const double s = sqrt(1.0 + R[0] - R[4] - R[8]) * 2.0;
        q[0] = (R[7] - R[5]) / s, q[1] = 0.25 * s, q[2] = (R[1] + R[3]) / s, q[3] = (R[2] + R[6]) / s;
    }
    else if(R[4] > R[8])
    {
        const double s = sqrt(1.0 + R[4] - R[0] - R[8]) * 2.0;
        q[0] = (R[2] - R[6]) / s, q[1] = (R[1] + R[3]) / s, q[2] = 0.25 * s, q[3] = (R[5] + R[7]) / s;
    }
    else
    {
        const double s = sqrt(1.0 + R[8] - R[0] - R[4]) * 2.0;
        q[0] = (R[3] - R[1]) / s, q[1] = (R[2] + R[6]) / s, q[2] = (R[5] + R[7]) / s, q[3] = 0.25 * s;
    }
    if(q[0] < 0.0)
        q[0] = -q[0], q[1] = -q[1], q[2] = -q[2], q[3] = -q[3];
}

// (H + lam diag(H) + 1e-12 I) dx = -g for the 6x6 normal equations of one Levenberg-Marquardt trial: Gaussian elimination
// with partial pivoting, run by one whole warp.  Lane r < 6 keeps row r of [A | b] in 7 registers; rows are never
// exchanged -- the pivot of column c is the largest |A[r][c]| among the rows not used yet (lowest row on ties), its
// row is broadcast with shuffles, and the back substitution visits the pivot lanes in reverse order.  The same
// elimination on one thread needs 42 doubles: with run-time indices it sat in local memory and was 40 % of the kernel's
// stall samples, fully unrolled it still spilled under the 128-register cap of a 512-thread CTA (27 % of the samples;
// profiles/r1_v12_pnp_source_lines.txt).
// hg = 21 upper-triangle entries of H row by row, then the 6 of g.  All 32 lanes must call; the result is uniform.
// false: a zero / infinite pivot (a NaN elsewhere propagates into dx and the trial's cost, which rejects the trial too).
__device__ __forceinline__ bool solve6_damped_warp(const double* hg, double lam, double (&x)[6])
{
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, r = lane < 6 ? lane : 5;
    double a[7];
#pragma unroll
    for(int c = 0; c < 6; ++c)
    {
        const int lo = min(r, c), hi = max(r, c);
        a[c] = hg[6 * lo - lo * (lo - 1) / 2 + hi - lo];
        if(c == r)
            a[c] += lam * a[c] + 1e-12;
    }
    a[6] = -hg[21 + r];
    bool ok = true, used = lane >= 6;
    int order[6];
#pragma unroll
    for(int col = 0; col < 6; ++col)
    {
        double v = used ? -1.0 : fabs(a[col]);
        int piv = lane;
#pragma unroll
        for(int o = 4; o > 0; o >>= 1)
        {
            const double ov = __shfl_xor_sync(full, v, o);
            const int op = __shfl_xor_sync(full, piv, o);
            if(ov > v || (ov == v && op < piv))
                v = ov, piv = op;
        }
        v = __shfl_sync(full, v, 0);
        piv = __shfl_sync(full, piv, 0);
        order[col] = piv;
        ok = ok && v > 0.0 && v < INFINITY;
        double prow[7];
#pragma unroll
        for(int k = col; k < 7; ++k)
            prow[k] = __shfl_sync(full, a[k], piv);
        if(lane == piv)
            used = true;
        if(!used)
        {
            const double m = a[col] / prow[col];
#pragma unroll
            for(int k = col + 1; k < 7; ++k)
                a[k] -= m * prow[k];
        }
    }
#pragma unroll
    for(int col = 5; col >= 0; --col)
    {
        double v = a[6];
#pragma unroll
        for(int k = col + 1; k < 6; ++k)
            v -= a[k] * x[k];
        x[col] = __shfl_sync(full, v / a[col], order[col]);
    }
    return ok;
}

// block-wide sum of NV doubles per thread; result in s_red[0..NV) (valid for all threads after the call)
template <int NV>
__device__ void block_sum(double* v, double* s_part /* [PNP_THREADS / 32][NV] */, double* s_red /* [NV] */)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if constexpr(NV > 8)
    {
        // many values: halve the value count at every butterfly stage (a lane keeps the half of the values its bit selects
        // and sends the other half), 31 shuffles for up to 32 values instead of 5 per value; lane k ends with total k
        static_assert(NV <= 32, "block_sum: at most 32 values");
        double w[32];
#pragma unroll
        for(int k = 0; k < 32; ++k)
            w[k] = k < NV ? v[k] : 0.0;
#pragma unroll
        for(int o = 16; o > 0; o >>= 1)
        {
            const bool hi = (lane & o) != 0;
#pragma unroll
            for(int k = 0; k < o; ++k)
            {
                const double send = hi ? w[k] : w[k + o], keep = hi ? w[k + o] : w[k];
                w[k] = keep + __shfl_xor_sync(0xffffffffu, send, o);
            }
        }
        if(lane < NV)
            s_part[warp * NV + lane] = w[0];
    }
    else
    {
#pragma unroll
        for(int k = 0; k < NV; ++k)
        {
            double x = v[k];
#pragma unroll
            for(int o = 16; o > 0; o >>= 1)
                x += __shfl_xor_sync(0xffffffffu, x, o);
            if(lane == 0)
                s_part[warp * NV + k] = x;
        }
    }
    __syncthreads();
    if(threadIdx.x < NV)
    {
        double x = 0.0;
#pragma unroll
        for(int w = 0; w < PNP_THREADS / 32; ++w)
            x += s_part[w * NV + threadIdx.x];
        s_red[threadIdx.x] = x;
    }
    __syncthreads();
}

__global__ void __cluster_dims__(PNP_CLUSTER, 1, 1) __launch_bounds__(PNP_THREADS) k_pnp_ransac(const float* __restrict__ obj_all, const float* __restrict__ img_all,
                                                           const int* __restrict__ counts, int cap, int frame0, Cam K,
                                                           int iterations, double thr2, unsigned seed, int refine_iters,
                                                           double* __restrict__ out_rt, double* __restrict__ out_pose,
                                                           uint8_t* __restrict__ out_mask, int* __restrict__ out_info)
{
    __shared__ Pose s_hyp[PNP_MAX_HYP];
    __shared__ int s_cnt[PNP_MAX_HYP];
    __shared__ double s_part[(PNP_THREADS / 32) * 28], s_red[28];
    __shared__ Pose s_cur;
    __shared__ int s_best;
    __shared__ double s_H[28], s_tdx[7][6], s_tlam[7];
    __shared__ Pose s_try[7];
    __shared__ int s_tok[7];

    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    const int crank = (int)cluster.block_rank();
    const int f = blockIdx.x / PNP_CLUSTER, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = min(max(counts[f], 0), cap);
    const float* obj = obj_all + (long long)f * cap * 3;
    const float* img = img_all + (long long)f * cap * 2;
    uint8_t* mask = out_mask + (long long)f * cap;
    int* info = out_info + 4 * f;
    double* rt = out_rt + 6 * f;
    double* pose = out_pose + 7 * f;

    if(crank == 0)
        for(int i = tid; i < cap; i += PNP_THREADS)
            mask[i] = 0;
    if(n < 4)
    {
        if(crank != 0)
            return;
        if(tid == 0)
            info[0] = 0, info[1] = 0, info[2] = -1, info[3] = 0;
        if(tid < 6)
            rt[tid] = 0.0;
        if(tid < 7)
            pose[tid] = tid == 0 ? 1.0 : 0.0;
        return;
    }
    const int need = n >= 5 ? 5 : 4;

    // ---- phase 1: one hypothesis per thread ----
    for(int h = tid; h < iterations; h += PNP_THREADS)
    {
        int idx[4], got = 0;
        for(unsigned j = 0; j < 64 && got < 4; ++j)
        {
            const int i = pnp_draw(seed, (unsigned)(frame0 + f), (unsigned)h, j, n);
            bool dup = false;
            for(int k = 0; k < got; ++k)
                dup |= idx[k] == i;
            if(!dup)
                idx[got++] = i;
        }
        Pose p;
        const bool ok = got == 4 && p3p_hypothesis(obj, img, idx, K, p);
        if(ok)
            s_hyp[h] = p;
        s_cnt[h] = ok ? 0 : -1;
    }
    // every CTA of the frame's cluster has built the same hypotheses (one thread each: splitting them would not shorten
    // the phase); nobody may add into rank 0's counts before rank 0 has initialised them
    cluster.sync();

    // ---- phase 2: inlier counts ----
    // work item = (pair of hypotheses, chunk of 256 points): a warp loads each point of the chunk once (32 lanes x 8 points)
    // and tests it against both poses held in registers; counts are accumulated with one shared-memory atomic per item
    {
        constexpr int CHUNK = 256;
        const int n_pairs = (iterations + 1) >> 1, n_chunks = (n + CHUNK - 1) / CHUNK;
        for(int item = warp + (PNP_THREADS / 32) * crank; item < n_pairs * n_chunks; item += PNP_CLUSTER * (PNP_THREADS / 32))
        {
            const int pr = item / n_chunks, ch = item - pr * n_chunks;
            const int h0 = 2 * pr, h1 = h0 + 1;
            const bool ok0 = s_cnt[h0] >= 0, ok1 = h1 < iterations && s_cnt[h1] >= 0;
            if(!ok0 && !ok1)
                continue;
            const Pose p0 = s_hyp[ok0 ? h0 : h1], p1 = s_hyp[ok1 ? h1 : h0];
            int c0 = 0, c1 = 0;
            const int i_end = min(n, (ch + 1) * CHUNK);
            for(int i = ch * CHUNK + lane; i < i_end; i += 32)
            {
                const V3 X = load_obj(obj, i);
                const double u0 = (double)__ldg(img + 2 * i), v0 = (double)__ldg(img + 2 * i + 1);
                c0 += is_inlier(p0.R, p0.t, X, u0, v0, K, thr2);
                c1 += is_inlier(p1.R, p1.t, X, u0, v0, K, thr2);
            }
            c0 = __reduce_add_sync(0xffffffffu, c0);
            c1 = __reduce_add_sync(0xffffffffu, c1);
            if(lane == 0)
            {
                if(ok0 && c0)
                    atomicAdd(&s_cnt[h0], c0);
                if(ok1 && c1)
                    atomicAdd(&s_cnt[h1], c1);
            }
        }
    }
    __syncthreads();
    if(crank != 0)
    {
        // hand this CTA's share of the counts to rank 0 through distributed shared memory, then leave
        int* cnt0 = cluster.map_shared_rank(s_cnt, 0);
        for(int h = tid; h < iterations; h += PNP_THREADS)
            if(s_cnt[h] > 0)
                atomicAdd(cnt0 + h, s_cnt[h]);
    }
    cluster.sync();
    if(crank != 0)
        return;

    // ---- phase 3: best hypothesis (most inliers, lowest index) ----
    if(warp == 0)
    {
        int bc = -1, bh = -1;
        for(int h = lane; h < iterations; h += 32)
            if(s_cnt[h] > bc)
                bc = s_cnt[h], bh = h;
#pragma unroll
        for(int o = 16; o > 0; o >>= 1)
        {
            const int oc = __shfl_xor_sync(0xffffffffu, bc, o), oh = __shfl_xor_sync(0xffffffffu, bh, o);
            if(oc > bc || (oc == bc && oh >= 0 && (bh < 0 || oh < bh)))
                bc = oc, bh = oh;
        }
        if(lane == 0)
        {
            s_best = bc >= need ? bh : -1;
            if(bc >= need)
                s_cur = s_hyp[bh];
            info[0] = bc >= need;
            info[1] = bc >= need ? bc : 0;
            info[2] = bc >= need ? bh : -1;
        }
    }
    __syncthreads();
    if(s_best < 0)
    {
        if(tid == 0)
            info[3] = 0;
        if(tid < 6)
            rt[tid] = 0.0;
        if(tid < 7)
            pose[tid] = tid == 0 ? 1.0 : 0.0;
        return;
    }
    // inlier mask of the best hypothesis (kept in registers as well: point i = tid + 256 k)
    for(int i = tid; i < n; i += PNP_THREADS)
        mask[i] = is_inlier(s_cur.R, s_cur.t, load_obj(obj, i), (double)__ldg(img + 2 * i), (double)__ldg(img + 2 * i + 1), K, thr2);
    __syncthreads(); // mask[] is re-read below by the threads that wrote it (same thread, same index): no fence needed

    // ---- phase 4: Levenberg-Marquardt over the inliers ----
    auto cost_of = [&](const Pose& p) -> double {
        double c[1] = {0.0};
        for(int i = tid; i < n; i += PNP_THREADS)
            if(mask[i])
                c[0] += err2(p.R, p.t, load_obj(obj, i), (double)__ldg(img + 2 * i), (double)__ldg(img + 2 * i + 1), K);
        block_sum<1>(c, s_part, s_red);
        const double r = s_red[0];
        __syncthreads();
        return r;
    };
    double lam = 1e-3;
    double cost = cost_of(s_cur);
    int it = 0;
    for(; it < refine_iters; ++it)
    {
        double acc[27]; // 21 upper-triangle entries of J^T J, then 6 of J^T r
#pragma unroll
        for(int k = 0; k < 27; ++k)
            acc[k] = 0.0;
        for(int i = tid; i < n; i += PNP_THREADS)
        {
            if(!mask[i])
                continue;
            const V3 c = transform(s_cur.R, s_cur.t, load_obj(obj, i));
            const double iz = 1.0 / c.z, px = c.x * iz, py = c.y * iz;
            const double ru = K.fx * px + K.cx - (double)__ldg(img + 2 * i), rv = K.fy * py + K.cy - (double)__ldg(img + 2 * i + 1);
            const V3 ju = {K.fx * iz, 0.0, -K.fx * px * iz}, jv = {0.0, K.fy * iz, -K.fy * py * iz};
            const V3 wu = cross(c, ju), wv = cross(c, jv);
            const double Ju[6] = {wu.x, wu.y, wu.z, ju.x, ju.y, ju.z}, Jv[6] = {wv.x, wv.y, wv.z, jv.x, jv.y, jv.z};
            int k = 0;
#pragma unroll
            for(int r = 0; r < 6; ++r)
#pragma unroll
                for(int cc = r; cc < 6; ++cc)
                    acc[k++] += Ju[r] * Ju[cc] + Jv[r] * Jv[cc];
#pragma unroll
            for(int r = 0; r < 6; ++r)
                acc[21 + r] += Ju[r] * ru + Jv[r] * rv;
        }
        block_sum<27>(acc, s_part, s_H); // s_H stays valid through the trials below (cost_of reduces into s_red)
        bool improved = false;
        double step = 0.0, drop = 0.0;
        // The up-to-8 trials of an iteration differ only in the damping (lam, 10 lam, 100 lam, ...) and the first one that
        // does not raise the cost is taken.  Trial 0 is accepted in every iteration but the last, where all eight are
        // rejected (the fit has converged): it runs alone; trials 1..7 then run TOGETHER -- warp w solves trial w + 1,
        // one pass over the inliers yields the seven costs -- and the first acceptable one in trial order is taken, so the
        // outcome is that of the sequential loop (oracle/pnp_oracle.py::refine) in two rounds instead of eight.
        auto run_trials = [&](auto nt_c) {
            constexpr int NT = decltype(nt_c)::value;
            if(warp < NT)
            {
                double lk = lam; // damping of this warp's trial: every earlier trial multiplied it by 10
                for(int k = 0; k < warp; ++k)
                    lk *= 10.0;
                double dx[6];
                const bool ok = solve6_damped_warp(s_H, lk, dx);
                if(lane == 0)
                {
                    s_tok[warp] = ok;
                    s_tlam[warp] = lk;
#pragma unroll
                    for(int k = 0; k < 6; ++k)
                        s_tdx[warp][k] = dx[k];
                    Pose pn = s_cur;
                    if(ok)
                    {
                        double dR[9];
                        rodrigues_exp(dx, dR);
#pragma unroll
                        for(int r = 0; r < 3; ++r)
                        {
#pragma unroll
                            for(int c = 0; c < 3; ++c)
                                pn.R[3 * r + c] = dR[3 * r] * s_cur.R[c] + dR[3 * r + 1] * s_cur.R[3 + c] + dR[3 * r + 2] * s_cur.R[6 + c];
                            pn.t[r] = dR[3 * r] * s_cur.t[0] + dR[3 * r + 1] * s_cur.t[1] + dR[3 * r + 2] * s_cur.t[2] + dx[3 + r];
                        }
                    }
                    s_try[warp] = pn;
                }
            }
            __syncthreads();
            double c[NT];
#pragma unroll
            for(int k = 0; k < NT; ++k)
                c[k] = 0.0;
            for(int i = tid; i < n; i += PNP_THREADS)
                if(mask[i])
                {
                    const V3 X = load_obj(obj, i);
                    const double u0 = (double)__ldg(img + 2 * i), v0 = (double)__ldg(img + 2 * i + 1);
#pragma unroll
                    for(int k = 0; k < NT; ++k)
                        c[k] += err2(s_try[k].R, s_try[k].t, X, u0, v0, K);
                }
            block_sum<NT>(c, s_part, s_red);
            int pick = -1;
            double cn = 0.0;
#pragma unroll
            for(int k = NT - 1; k >= 0; --k)
                if(s_tok[k] && isfinite(s_red[k]) && s_red[k] <= cost)
                    pick = k, cn = s_red[k];
            if(pick >= 0)
            {
                step = 0.0;
                for(int k = 0; k < 6; ++k)
                    step = fmax(step, fabs(s_tdx[pick][k]));
                drop = cost - cn;
                cost = cn;
                lam = fmax(s_tlam[pick] * 0.1, 1e-12);
                improved = true;
            }
            else
                for(int k = 0; k < NT; ++k)
                    lam *= 10.0;
            __syncthreads();
            if(pick >= 0)
            {
                if(tid == 0)
                    s_cur = s_try[pick];
                __syncthreads();
            }
        };
        run_trials(std::integral_constant<int, 1>());
        if(!improved)
            run_trials(std::integral_constant<int, 7>());
        // the fit converges quadratically (steps ~ 4e-3, 3e-5, 2e-8, 5e-12 on the bench scenes): a step below 1e-10 is the
        // last one that can matter in float64 (OpenCV's own refinement stops at FLT_EPSILON); going on only buys an
        // iteration whose eight trials are all rejected
        if(!improved || step < 1e-10 || drop <= 1e-15 * fmax(cost, 1e-300))
        {
            ++it;
            break;
        }
    }

    // ---- phase 5: outputs ----
    if(tid == 0)
    {
        info[3] = it;
        rodrigues_log(s_cur.R, rt);
        rt[3] = s_cur.t[0], rt[4] = s_cur.t[1], rt[5] = s_cur.t[2];
        // sensor pose (cv_ransac_pnp.cpp:65-77): orientation = R^-1, position = -(R^-1 t)
        const double Rt[9] = {s_cur.R[0], s_cur.R[3], s_cur.R[6], s_cur.R[1], s_cur.R[4], s_cur.R[7], s_cur.R[2], s_cur.R[5], s_cur.R[8]};
        quat_from_R(Rt, pose);
        for(int r = 0; r < 3; ++r)
            pose[4 + r] = -(Rt[3 * r] * s_cur.t[0] + Rt[3 * r + 1] * s_cur.t[1] + Rt[3 * r + 2] * s_cur.t[2]);
    }
}
} // namespace

void launch_pnp_ransac(cudaStream_t st, const float* obj, const float* img, const int* counts, int n_frames, int cap, int frame0,
                       double fx, double fy, double cx, double cy, int iterations, double thr2, unsigned seed, int refine_iters,
                       double* out_rt, double* out_pose, uint8_t* out_mask, int* out_info, long long* launches)
{
    const Cam K{fx, fy, cx, cy};
    k_pnp_ransac<<<PNP_CLUSTER * n_frames, PNP_THREADS, 0, st>>>(obj, img, counts, cap, frame0, K, iterations, thr2, seed, refine_iters, out_rt,
                                                   out_pose, out_mask, out_info);
    if(launches)
        ++*launches;
}
int pnp_max_hypotheses() { return PNP_MAX_HYP; }
} // namespace mslam_b200
