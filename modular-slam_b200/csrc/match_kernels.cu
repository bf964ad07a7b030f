// match_kernels.cu -- sm_100a kernels of the 256-bit Hamming matching path.
//
// Semantics are those of the reference matcher, OrbOpenCvMatcher::Pimpl::match
// (src/lib/modular_slam/orb_feature.cpp:84-117): cv::BFMatcher(NORM_HAMMING).knnMatch(query = to, train = from, k = 2)
// -- per query the two smallest distances, ties to the lower train index (cv::batchDistance) -- then the
// 0.7 ratio test (orb_feature.cpp:101) and {fromIndex = trainIdx, toIndex = queryIdx} in ascending toIndex
// (orb_feature.cpp:110-114).  Everything here is XOR + POPC on the integer pipes; no tensor cores.
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"

namespace mslam_b200
{

constexpr unsigned kNoKey = 0xffffffffu;

// three-input logic on the ALU pipe (one LOP3 each)
__device__ __forceinline__ unsigned xor3(unsigned a, unsigned b, unsigned c)
{
    unsigned r;
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ unsigned maj3(unsigned a, unsigned b, unsigned c)
{
    unsigned r;
    asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}

// 256-bit Hamming distance of two descriptors held as 2 x uint4.
// POPC runs on the XU pipe at 16 results/clk/SM -- a quarter of the ALU pipe's LOP3 rate (measured:
// mslam_b200_measure_int_peaks) -- so the eight XOR words are first compressed with four carry-save adders
// (x0+x1+x2 = s + 2c per bit position): sum(popc(x0..x6)) = popc(s2) + 2 popc(s3) + 4 popc(c3), and only 4 POPC
// (instead of 8) are issued per pair.  Exact integer arithmetic; nothing is approximated.
__device__ __forceinline__ int ham256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1)
{
    const unsigned x0 = a0.x ^ b0.x, x1 = a0.y ^ b0.y, x2 = a0.z ^ b0.z, x3 = a0.w ^ b0.w;
    const unsigned x4 = a1.x ^ b1.x, x5 = a1.y ^ b1.y, x6 = a1.z ^ b1.z, x7 = a1.w ^ b1.w;
    const unsigned s0 = xor3(x0, x1, x2), c0 = maj3(x0, x1, x2);
    const unsigned s1 = xor3(x3, x4, x5), c1 = maj3(x3, x4, x5);
    const unsigned s2 = xor3(s0, s1, x6), c2 = maj3(s0, s1, x6);
    const unsigned s3 = xor3(c0, c1, c2), c3 = maj3(c0, c1, c2);
    return (__popc(s2) + __popc(x7)) + 2 * (__popc(s3) + 2 * __popc(c3));
}

// =================================================================================================
// K8  brute-force kNN (k = 2) for a batch of independent (query set, train set) pairs.
//   Each thread owns QPT query descriptors in registers (8 x u32 each); the train set streams through
//   shared memory in chunks and is read with broadcast 128-bit loads.  The running top-2 is kept as
//   packed keys (distance << 16 | train index): two min/max per candidate give exactly "smallest
//   distance, ties to the lower index" with no branches.
//   Output: key0/key1 per query ([pair][cap]); kNoKey where the train set is too small.
// =================================================================================================
constexpr int KNN_CHUNK = 512; // train descriptors per shared-memory chunk (16 KB)

template <int QPT, int KNN_THREADS>
__global__ void __launch_bounds__(KNN_THREADS) k_knn2(const uint8_t* __restrict__ q_base, const int* __restrict__ q_counts,
                                                      const uint8_t* __restrict__ t_base, const int* __restrict__ t_counts,
                                                      int cap, unsigned* __restrict__ key0, unsigned* __restrict__ key1)
{
    // blockIdx.z = slice of the train set (gridDim.z slices; partial top-2 lists are merged by k_merge_keys)
    __shared__ uint4 s_t[KNN_CHUNK * 2];
    const int p = blockIdx.y;
    const int nq = min(q_counts[p], cap), nt_all = min(t_counts[p], cap);
    const int per = (nt_all + (int)gridDim.z - 1) / (int)gridDim.z;
    const int t_begin = min((int)blockIdx.z * per, nt_all), nt = min(t_begin + per, nt_all);
    key0 += (long long)blockIdx.z * gridDim.y * cap;
    key1 += (long long)blockIdx.z * gridDim.y * cap;
    const int q0 = blockIdx.x * (KNN_THREADS * QPT);
    if(q0 >= nq)
        return;
    const uint4* qv = reinterpret_cast<const uint4*>(q_base + (long long)p * cap * 32);
    const uint4* tv = reinterpret_cast<const uint4*>(t_base + (long long)p * cap * 32);
    uint4 qa[QPT], qb[QPT];
    unsigned k0[QPT], k1[QPT];
#pragma unroll
    for(int i = 0; i < QPT; ++i)
    {
        const int q = q0 + i * KNN_THREADS + threadIdx.x;
        qa[i] = q < nq ? __ldg(qv + 2 * q) : make_uint4(0, 0, 0, 0);
        qb[i] = q < nq ? __ldg(qv + 2 * q + 1) : make_uint4(0, 0, 0, 0);
        k0[i] = kNoKey;
        k1[i] = kNoKey;
    }
    for(int t0 = t_begin; t0 < nt; t0 += KNN_CHUNK)
    {
        const int nc = min(KNN_CHUNK, nt - t0);
        __syncthreads();
        for(int i = threadIdx.x; i < nc * 2; i += KNN_THREADS)
            s_t[i] = __ldg(tv + 2 * t0 + i);
        __syncthreads();
#pragma unroll(QPT == 1 ? 4 : 2)
        for(int j = 0; j < nc; ++j)
        {
            const uint4 ta = s_t[2 * j], tb = s_t[2 * j + 1];
#pragma unroll
            for(int i = 0; i < QPT; ++i)
            {
                const unsigned key = ((unsigned)ham256(qa[i], qb[i], ta, tb) << 16) | (unsigned)(t0 + j);
                k1[i] = min(k1[i], max(k0[i], key));
                k0[i] = min(k0[i], key);
            }
        }
    }
#pragma unroll
    for(int i = 0; i < QPT; ++i)
    {
        const int q = q0 + i * KNN_THREADS + threadIdx.x;
        if(q < nq)
        {
            key0[(long long)p * cap + q] = k0[i];
            key1[(long long)p * cap + q] = k1[i];
        }
    }
}

// merge the per-slice top-2 lists ([slice][pair][cap]) into the final one: the two smallest of all 2 * slices keys
__global__ void __launch_bounds__(256) k_merge_keys(const unsigned* __restrict__ part0, const unsigned* __restrict__ part1,
                                                    int slices, long long n, unsigned* __restrict__ key0,
                                                    unsigned* __restrict__ key1)
{
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if(i >= n)
        return;
    unsigned a = kNoKey, b = kNoKey;
    for(int s = 0; s < slices; ++s)
    {
        const unsigned x = part0[s * n + i], y = part1[s * n + i];
        b = min(b, max(a, x));
        a = min(a, x);
        b = min(b, max(a, y));
        a = min(a, y);
    }
    key0[i] = a;
    key1[i] = b;
}

// block-wide exclusive scan of a small count per thread (256 threads); returns the offset, total in *total
__device__ __forceinline__ int block_scan_count(int cnt, int* total, int* s_warp /* >= 8 ints */)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int incl = cnt;
#pragma unroll
    for(int o = 1; o < 32; o <<= 1)
    {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if(lane >= o)
            incl += t;
    }
    if(lane == 31)
        s_warp[wid] = incl;
    __syncthreads();
    int off = 0, tot = 0;
#pragma unroll
    for(int w = 0; w < 8; ++w)
    {
        const int c = s_warp[w];
        off += w < wid ? c : 0;
        tot += c;
    }
    __syncthreads();
    *total = tot;
    return off + incl - cnt;
}

// ratio test + ordered compaction: one CTA per pair, ascending query (= toIndex) order.
//   keep iff (double)(float)d0 < ratio * (double)(float)d1          (orb_feature.cpp:101)
__global__ void __launch_bounds__(256) k_ratio_compact(const unsigned* __restrict__ key0, const unsigned* __restrict__ key1,
                                                       const int* __restrict__ q_counts, const int* __restrict__ t_counts,
                                                       int cap, double ratio, int* __restrict__ out_a, int* __restrict__ out_b,
                                                       int* __restrict__ out_d, int* __restrict__ n_out)
{
    __shared__ int s_warp[8];
    const int p = blockIdx.x;
    const int nq = min(q_counts[p], cap), nt = min(t_counts[p], cap);
    int run = 0;
    // four consecutive queries per thread: a frame's ~1000 queries are one round of the CTA (one scan, two barriers) instead of four
    if(nt >= 2)
        for(int base = 0; base < nq; base += 1024)
        {
            const int q0 = base + 4 * threadIdx.x;
            unsigned a[4];
            bool keep[4];
            int cnt = 0;
#pragma unroll
            for(int j = 0; j < 4; ++j)
            {
                const int q = q0 + j;
                keep[j] = false;
                a[j] = 0;
                if(q < nq)
                {
                    a[j] = key0[(long long)p * cap + q];
                    const unsigned b = key1[(long long)p * cap + q];
                    keep[j] = (double)(float)(a[j] >> 16) < ratio * (double)(float)(b >> 16);
                }
                cnt += keep[j] ? 1 : 0;
            }
            int tot;
            int pos = run + block_scan_count(cnt, &tot, s_warp);
#pragma unroll
            for(int j = 0; j < 4; ++j)
                if(keep[j])
                {
                    out_a[(long long)p * cap + pos] = (int)(a[j] & 0xffffu); // fromIndex = trainIdx
                    out_b[(long long)p * cap + pos] = q0 + j;                // toIndex = queryIdx
                    if(out_d)
                        out_d[(long long)p * cap + pos] = (int)(a[j] >> 16);
                    ++pos;
                }
            run += tot;
        }
    if(threadIdx.x == 0)
        n_out[p] = run;
}

// mutual-nearest-neighbour compaction (cv::BFMatcher crossCheck=true): key0_q from knn(query, train),
// key0_t from knn(train, query); ascending query order.
__global__ void __launch_bounds__(256) k_mutual_compact(const unsigned* __restrict__ key0_q, const unsigned* __restrict__ key0_t,
                                                        const int* __restrict__ q_counts, const int* __restrict__ t_counts,
                                                        int cap, int* __restrict__ out_a, int* __restrict__ out_b,
                                                        int* __restrict__ out_d, int* __restrict__ n_out)
{
    __shared__ int s_warp[8];
    const int p = blockIdx.x;
    const int nq = min(q_counts[p], cap), nt = min(t_counts[p], cap);
    int run = 0;
    if(nt >= 1)
        for(int base = 0; base < nq; base += 1024) // four consecutive queries per thread, as k_ratio_compact
        {
            const int q0 = base + 4 * threadIdx.x;
            unsigned a[4];
            bool keep[4];
            int cnt = 0;
#pragma unroll
            for(int j = 0; j < 4; ++j)
            {
                const int q = q0 + j;
                keep[j] = false;
                a[j] = 0;
                if(q < nq)
                {
                    a[j] = key0_q[(long long)p * cap + q];
                    const unsigned back = key0_t[(long long)p * cap + (a[j] & 0xffffu)];
                    keep[j] = (int)(back & 0xffffu) == q;
                }
                cnt += keep[j] ? 1 : 0;
            }
            int tot;
            int pos = run + block_scan_count(cnt, &tot, s_warp);
#pragma unroll
            for(int j = 0; j < 4; ++j)
                if(keep[j])
                {
                    out_a[(long long)p * cap + pos] = (int)(a[j] & 0xffffu); // trainIdx
                    out_b[(long long)p * cap + pos] = q0 + j;                // queryIdx
                    if(out_d)
                        out_d[(long long)p * cap + pos] = (int)(a[j] >> 16);
                    ++pos;
                }
            run += tot;
        }
    if(threadIdx.x == 0)
        n_out[p] = run;
}

// unpack keys to (idx, dist) int arrays for the raw knn2 entry point
__global__ void k_unpack_knn(const unsigned* __restrict__ key0, const unsigned* __restrict__ key1, int n, int* idx0, int* d0,
                             int* idx1, int* d1)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if(i >= n)
        return;
    const unsigned a = key0[i], b = key1[i];
    idx0[i] = a == kNoKey ? -1 : (int)(a & 0xffffu);
    d0[i] = a == kNoKey ? -1 : (int)(a >> 16);
    idx1[i] = b == kNoKey ? -1 : (int)(b & 0xffffu);
    d1[i] = b == kNoKey ? -1 : (int)(b >> 16);
}

// =================================================================================================
// K10  keyframe-database scoring:
//   score[q][k] = | match(from = query frame q, to = keyframe k) |
//   knn query role = keyframe descriptors (registers, DPT per thread), train role = the query frame's
//   descriptors (whole frame resident in shared memory).  Only distances are needed, so the top-2 is
//   two integers per descriptor; the ratio test is the integer form 10*d0 < 7*d1 when ratio == 0.7,
//   else the double comparison.  One CTA walks KF_PER_CTA keyframes against one query frame.
// =================================================================================================
constexpr int DB_THREADS = 256;

template <int DPT>
__global__ void __launch_bounds__(DB_THREADS) k_db_scores(const uint8_t* __restrict__ queries, const int* __restrict__ q_counts,
                                                          int q_cap, const uint8_t* __restrict__ db,
                                                          const int* __restrict__ k_counts, int k_cap, int n_keyframes,
                                                          int kf_per_cta, double ratio, int* __restrict__ scores)
{
    extern __shared__ uint4 s_q[]; // q_cap * 2
    __shared__ int s_red[DB_THREADS / 32];
    const int qf = blockIdx.y;
    const int nfrom = min(q_counts[qf], q_cap);
    const uint4* qv = reinterpret_cast<const uint4*>(queries + (long long)qf * q_cap * 32);
    for(int i = threadIdx.x; i < nfrom * 2; i += DB_THREADS)
        s_q[i] = __ldg(qv + i);
    __syncthreads();
    const bool int_ratio = ratio == 0.7;
    const int kf0 = blockIdx.x * kf_per_cta, kf1 = min(kf0 + kf_per_cta, n_keyframes);
    for(int kf = kf0; kf < kf1; ++kf)
    {
        const int nto = min(k_counts[kf], k_cap);
        const uint4* kv = reinterpret_cast<const uint4*>(db + (long long)kf * k_cap * 32);
        int cnt = 0;
        if(nfrom >= 2)
            for(int base = 0; base < nto; base += DB_THREADS * DPT)
            {
                uint4 da[DPT], dbb[DPT];
                int d0[DPT], d1[DPT];
#pragma unroll
                for(int i = 0; i < DPT; ++i)
                {
                    const int d = base + i * DB_THREADS + threadIdx.x;
                    da[i] = d < nto ? __ldg(kv + 2 * d) : make_uint4(0, 0, 0, 0);
                    dbb[i] = d < nto ? __ldg(kv + 2 * d + 1) : make_uint4(0, 0, 0, 0);
                    d0[i] = 1 << 20;
                    d1[i] = 1 << 20;
                }
#pragma unroll 2
                for(int j = 0; j < nfrom; ++j)
                {
                    const uint4 ta = s_q[2 * j], tb = s_q[2 * j + 1];
#pragma unroll
                    for(int i = 0; i < DPT; ++i)
                    {
                        const int d = ham256(da[i], dbb[i], ta, tb);
                        d1[i] = min(d1[i], max(d0[i], d));
                        d0[i] = min(d0[i], d);
                    }
                }
#pragma unroll
                for(int i = 0; i < DPT; ++i)
                {
                    const int d = base + i * DB_THREADS + threadIdx.x;
                    const bool pass = int_ratio ? (10 * d0[i] < 7 * d1[i])
                                                : ((double)(float)d0[i] < ratio * (double)(float)d1[i]);
                    cnt += (d < nto) && pass;
                }
            }
#pragma unroll
        for(int o = 16; o > 0; o >>= 1)
            cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        if((threadIdx.x & 31) == 0)
            s_red[threadIdx.x >> 5] = cnt;
        __syncthreads();
        if(threadIdx.x == 0)
        {
            int t = 0;
#pragma unroll
            for(int w = 0; w < DB_THREADS / 32; ++w)
                t += s_red[w];
            scores[(long long)qf * n_keyframes + kf] = t;
        }
        __syncthreads();
    }
}

// per query frame: top-k of n entries by (score desc, id asc).  Entry e of query q is read through
// (list = e / per_list, j = e % per_list) at  list * list_stride + q * per_list + j  so the same kernel serves a plain
// score row (n_lists = 1, ids implicit = id_base + j), dense [n_lists][n_queries][per_list] candidate lists
// (list_stride = n_queries * per_list) and the all-gathered shard lists, where every rank's block is its scores followed
// by its ids (list_stride = 2 * n_queries * per_list, ids = scores + n_queries * per_list).
__global__ void __launch_bounds__(256) k_topk(const int* __restrict__ scores, const int* __restrict__ ids, int n_lists,
                                              int per_list, long long list_stride, int n_queries, int id_base, int k,
                                              int* __restrict__ out_score, int* __restrict__ out_id)
{
    __shared__ unsigned long long s_best[8];
    __shared__ unsigned long long s_last;
    const int q = blockIdx.x;
    const int n = n_lists * per_list;
    unsigned long long last = ~0ull; // strictly-below bound for the next pick
    for(int r = 0; r < k; ++r)
    {
        unsigned long long best = 0; // 0 = nothing (valid keys are > 0 because score+1 >= 1 sits in the high word)
        for(int e = threadIdx.x; e < n; e += 256)
        {
            const int list = e / per_list, j = e - list * per_list;
            const long long at = (long long)list * list_stride + (long long)q * per_list + j;
            const int sc = scores[at];
            const int id = ids ? ids[at] : id_base + j;
            if(sc < 0 || id < 0)
                continue;
            const unsigned long long key = ((unsigned long long)(unsigned)(sc + 1) << 32) | (0xffffffffu - (unsigned)id);
            if(key < last && key > best)
                best = key;
        }
#pragma unroll
        for(int o = 16; o > 0; o >>= 1)
        {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
            best = other > best ? other : best;
        }
        if((threadIdx.x & 31) == 0)
            s_best[threadIdx.x >> 5] = best;
        __syncthreads();
        if(threadIdx.x == 0)
        {
            unsigned long long m = 0;
#pragma unroll
            for(int w = 0; w < 8; ++w)
                m = s_best[w] > m ? s_best[w] : m;
            s_last = m;
            out_score[(long long)q * k + r] = m ? (int)(m >> 32) - 1 : -1;
            out_id[(long long)q * k + r] = m ? (int)(0xffffffffu - (unsigned)(m & 0xffffffffu)) : -1;
        }
        __syncthreads();
        last = s_last; // 0 once exhausted: no key is < 0, so every further pick is (-1, -1)
        __syncthreads();
    }
}

// =================================================================================================
// integer-pipe roofline probes (register-only, all SMs)
// =================================================================================================
__global__ void __launch_bounds__(256) k_peak_popc(unsigned* out, int iters)
{
    unsigned v[8];
#pragma unroll
    for(int i = 0; i < 8; ++i)
        v[i] = threadIdx.x * 2654435761u + i * 40503u + blockIdx.x;
#pragma unroll 8
    for(int it = 0; it < iters; ++it)
    {
#pragma unroll
        for(int i = 0; i < 8; ++i)
            v[i] = __popc(v[i]) * 0x01010101u + (unsigned)i; // POPC feeds the next POPC; the IMAD (fma pipe) keeps chains apart
    }
    unsigned s = 0;
#pragma unroll
    for(int i = 0; i < 8; ++i)
        s ^= v[i];
    if(s == 0x12345u)
        out[0] = s;
}

__global__ void __launch_bounds__(256) k_peak_lop3(unsigned* out, int iters, unsigned a, unsigned b)
{
    unsigned v[8];
#pragma unroll
    for(int i = 0; i < 8; ++i)
        v[i] = threadIdx.x * 2654435761u + i * 40503u + blockIdx.x;
#pragma unroll 8
    for(int it = 0; it < iters; ++it)
    {
#pragma unroll
        for(int i = 0; i < 8; ++i)
            v[i] = v[i] ^ (a & v[(i + 1) & 7]); // one LOP3 per result (3-input function)
    }
    unsigned s = 0;
#pragma unroll
    for(int i = 0; i < 8; ++i)
        s ^= v[i];
    if(s == 0x12345u)
        out[0] = s;
}

// 3-register-operand LOP3 (what the bit-sliced FAST test and the carry-save adders issue): 12 independent chains per
// thread, every instruction reads three different live registers
__global__ void __launch_bounds__(256) k_peak_lop3r(unsigned* out, int iters)
{
    unsigned v[12];
#pragma unroll
    for(int i = 0; i < 12; ++i)
        v[i] = threadIdx.x * 2654435761u + i * 40503u + blockIdx.x;
#pragma unroll 4
    for(int it = 0; it < iters; ++it)
    {
#pragma unroll
        for(int i = 0; i < 12; ++i)
            v[i] = maj3(v[i], v[(i + 4) % 12], v[(i + 8) % 12]) ^ 0u; // majority of three registers, one LOP3
    }
    unsigned s = 0;
#pragma unroll
    for(int i = 0; i < 12; ++i)
        s ^= v[i];
    if(s == 0x12345u)
        out[0] = s;
}

// the pair sequence of the matcher with operands in registers: 8 x xor, then either 8 POPC + adds (CSA = false) or the
// carry-save form ham256 uses (CSA = true).  All eight train words change every iteration through an IMAD (fma pipe,
// otherwise idle) so that nothing can be hoisted out of the loop.
template <bool CSA>
__global__ void __launch_bounds__(256) k_peak_pairs(unsigned* out, int iters)
{
    uint4 qa, qb, ta, tb;
    qa = make_uint4(threadIdx.x * 2654435761u, threadIdx.x * 40503u + 1u, threadIdx.x + 7u, threadIdx.x * 3u);
    qb = make_uint4(qa.x ^ 0x9e3779b9u, qa.y + 11u, qa.z * 5u, qa.w ^ 0x7f4a7c15u);
    ta = make_uint4(blockIdx.x * 2246822519u, blockIdx.x + 3u, blockIdx.x * 7u + 1u, blockIdx.x ^ 0x55555555u);
    tb = make_uint4(ta.x + 1u, ta.y * 3u, ta.z ^ 0x33333333u, ta.w + 5u);
    unsigned best = 0xffffffffu, second = 0xffffffffu;
#pragma unroll 4
    for(int it = 0; it < iters; ++it)
    {
        unsigned d;
        if(CSA)
            d = (unsigned)ham256(qa, qb, ta, tb);
        else
            d = __popc(qa.x ^ ta.x) + __popc(qa.y ^ ta.y) + __popc(qa.z ^ ta.z) + __popc(qa.w ^ ta.w) + __popc(qb.x ^ tb.x) +
                __popc(qb.y ^ tb.y) + __popc(qb.z ^ tb.z) + __popc(qb.w ^ tb.w);
        const unsigned key = (d << 16) | (unsigned)(it & 0xffff);
        second = min(second, max(best, key));
        best = min(best, key);
        ta.x = ta.x * 3u + d; ta.y = ta.y * 5u + d; ta.z = ta.z * 7u + d; ta.w = ta.w * 9u + d;
        tb.x = tb.x * 11u + d; tb.y = tb.y * 13u + d; tb.z = tb.z * 17u + d; tb.w = tb.w * 19u + d;
    }
    if((best ^ second) == 0x12345u)
        out[0] = best;
}

__global__ void k_clock_probe(long long* out)
{
    // SM cycles and nanoseconds across a fixed spin, for the MHz estimate
    const long long c0 = clock64();
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
    unsigned v = threadIdx.x;
    for(int i = 0; i < 200000; ++i)
        v = v * 1664525u + 1013904223u;
    const long long c1 = clock64();
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
    if(threadIdx.x == 0 && blockIdx.x == 0)
    {
        out[0] = c1 - c0;
        out[1] = (long long)(t1 - t0);
        out[2] = v;
    }
}

// =================================================================================================
// launchers
// =================================================================================================
void launch_knn2(cudaStream_t st, const uint8_t* q, const int* qn, const uint8_t* t, const int* tn, int n_pairs, int cap,
                 unsigned* key0, unsigned* key1, unsigned* part0, unsigned* part1, int sm_count, long long* launches)
{
    // Work granularity: the kernel is compute-bound on the integer pipes, so what matters is enough warps on every SM and
    // an even split over the 148 SMs.  Few frame pairs -> 64 queries per CTA and the train set cut into slices (grid z)
    // until there are ~50 warps per SM, partial top-2 lists merged by a second tiny kernel; lots of work -> 256 queries
    // per CTA (fewer shared-memory refills of the train set), no slicing.
    const long long small_ctas = (long long)((cap + 63) / 64) * n_pairs;
    if(small_ctas <= 24LL * sm_count)
    {
        static int forced = -1;
        if(forced < 0)
        {
            const char* e = getenv("MSLAM_KNN_SLICES");
            forced = e ? atoi(e) : 0;
        }
        int slices = 1;
        while(slices < 4 && small_ctas * slices < 28LL * sm_count && cap / (slices * 2) >= 128)
            slices *= 2;
        if(forced > 0)
            slices = forced < kKnnMaxSlices ? forced : kKnnMaxSlices;
        dim3 grd((cap + 63) / 64, n_pairs, slices);
        if(slices == 1)
            k_knn2<1, 64><<<grd, 64, 0, st>>>(q, qn, t, tn, cap, key0, key1);
        else
        {
            k_knn2<1, 64><<<grd, 64, 0, st>>>(q, qn, t, tn, cap, part0, part1);
            const long long n = (long long)n_pairs * cap;
            k_merge_keys<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(part0, part1, slices, n, key0, key1);
            ++*launches;
        }
    }
    else
    {
        // 256 queries per CTA (two per thread) halve the shared-memory traffic per pair, but leave only
        // n_pairs * cap / 64 warps; below ~48 warps per SM one query per thread hides latency better (measured on the
        // distributed-ORB workload, 64 pairs x ~3260 descriptors: 534 -> 704 Gpair/s)
        static int mode = -1;
        if(mode < 0)
        {
            const char* e = getenv("MSLAM_KNN_MODE");
            mode = e ? atoi(e) : 0;
        }
        const long long warps2 = (long long)((cap + 255) / 256) * n_pairs * 4;
        const bool one_per_thread = mode == 2 || (mode == 0 && warps2 < 48LL * sm_count);
        if(one_per_thread)
        {
            dim3 grd((cap + 127) / 128, n_pairs);
            k_knn2<1, 128><<<grd, 128, 0, st>>>(q, qn, t, tn, cap, key0, key1);
        }
        else
        {
            dim3 grd((cap + 255) / 256, n_pairs);
            k_knn2<2, 128><<<grd, 128, 0, st>>>(q, qn, t, tn, cap, key0, key1);
        }
    }
    ++*launches;
}

void launch_ratio_compact(cudaStream_t st, const unsigned* key0, const unsigned* key1, const int* qn, const int* tn,
                          int n_pairs, int cap, double ratio, int* out_a, int* out_b, int* out_d, int* n_out,
                          long long* launches)
{
    k_ratio_compact<<<n_pairs, 256, 0, st>>>(key0, key1, qn, tn, cap, ratio, out_a, out_b, out_d, n_out);
    ++*launches;
}

void launch_mutual_compact(cudaStream_t st, const unsigned* key0_q, const unsigned* key0_t, const int* qn, const int* tn,
                           int n_pairs, int cap, int* out_a, int* out_b, int* out_d, int* n_out, long long* launches)
{
    k_mutual_compact<<<n_pairs, 256, 0, st>>>(key0_q, key0_t, qn, tn, cap, out_a, out_b, out_d, n_out);
    ++*launches;
}

void launch_unpack_knn(cudaStream_t st, const unsigned* key0, const unsigned* key1, int n, int* idx0, int* d0, int* idx1,
                       int* d1, long long* launches)
{
    k_unpack_knn<<<(n + 255) / 256, 256, 0, st>>>(key0, key1, n, idx0, d0, idx1, d1);
    ++*launches;
}

// per DEVICE: called from mslam_b200_create after cudaSetDevice (the attribute applies to the current device only)
cudaError_t configure_match_kernels()
{
    return cudaFuncSetAttribute(k_db_scores<kDbDpt>, cudaFuncAttributeMaxDynamicSharedMemorySize, kDbMaxSmem);
}

cudaError_t launch_db_scores(cudaStream_t st, const uint8_t* queries, const int* q_counts, int nq_frames, int q_cap,
                             const uint8_t* db, const int* k_counts, int n_keyframes, int k_cap, double ratio, int* scores,
                             int sm_count, long long* launches)
{
    constexpr int DPT = kDbDpt;
    const size_t smem = (size_t)q_cap * 32; // <= kDbMaxSmem (checked by the entry point), opted in by configure_match_kernels
    // enough CTAs to fill the machine a few times over, but long enough walks to amortise the query-frame load
    int kf_per_cta = 8;
    while(kf_per_cta > 1 && (long long)((n_keyframes + kf_per_cta - 1) / kf_per_cta) * nq_frames < 4LL * sm_count)
        kf_per_cta >>= 1;
    dim3 grd((n_keyframes + kf_per_cta - 1) / kf_per_cta, nq_frames);
    k_db_scores<DPT><<<grd, DB_THREADS, smem, st>>>(queries, q_counts, q_cap, db, k_counts, k_cap, n_keyframes, kf_per_cta,
                                                    ratio, scores);
    ++*launches;
    return cudaGetLastError();
}

void launch_topk(cudaStream_t st, const int* scores, const int* ids, int n_lists, int per_list, long long list_stride,
                 int n_queries, int id_base, int k, int* out_score, int* out_id, long long* launches)
{
    k_topk<<<n_queries, 256, 0, st>>>(scores, ids, n_lists, per_list, list_stride, n_queries, id_base, k, out_score, out_id);
    ++*launches;
}

cudaError_t measure_int_peaks(cudaStream_t st, int sm_count, double* out4 /* 5 values */, long long* launches)
{
    unsigned* d_out = nullptr;
    long long* d_clk = nullptr;
    cudaError_t e = cudaMalloc(&d_out, 64);
    if(e != cudaSuccess)
        return e;
    e = cudaMalloc(&d_clk, 64);
    if(e != cudaSuccess)
        return e;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int grid = sm_count * 8, iters = 4096;
    float ms[5] = {0, 0, 0, 0, 0};
    for(int which = 0; which < 5; ++which)
    {
        for(int rep = 0; rep < 3; ++rep) // last repetition is the one kept (warm clocks)
        {
            cudaEventRecord(e0, st);
            if(which == 0)
                k_peak_popc<<<grid, 256, 0, st>>>(d_out, iters);
            else if(which == 1)
                k_peak_lop3<<<grid, 256, 0, st>>>(d_out, iters, 0x9e3779b9u, 0x7f4a7c15u);
            else if(which == 2)
                k_peak_pairs<false><<<grid, 256, 0, st>>>(d_out, iters);
            else if(which == 3)
                k_peak_pairs<true><<<grid, 256, 0, st>>>(d_out, iters);
            else
                k_peak_lop3r<<<grid, 256, 0, st>>>(d_out, iters);
            cudaEventRecord(e1, st);
            cudaEventSynchronize(e1);
            cudaEventElapsedTime(&ms[which], e0, e1);
            ++*launches;
        }
    }
    k_clock_probe<<<1, 32, 0, st>>>(d_clk);
    ++*launches;
    long long h_clk[3] = {0, 1, 0};
    e = cudaMemcpyAsync(h_clk, d_clk, sizeof(h_clk), cudaMemcpyDeviceToHost, st);
    cudaStreamSynchronize(st);
    const double threads = (double)grid * 256.0;
    out4[0] = threads * iters * 8.0 / (ms[0] * 1e-3);
    out4[1] = threads * iters * 8.0 / (ms[1] * 1e-3);
    out4[2] = threads * iters / (ms[2] * 1e-3);
    out4[3] = h_clk[1] > 0 ? (double)h_clk[0] / (double)h_clk[1] * 1e3 : 0.0;
    out4[4] = threads * iters / (ms[3] * 1e-3);
    out4[5] = threads * iters * 12.0 / (ms[4] * 1e-3);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d_out);
    cudaFree(d_clk);
    return e == cudaSuccess ? cudaGetLastError() : e;
}


// =================================================================================================
// matching without any allocation (mslam_b200_match_knn_ratio_noalloc_dev, include/mslam_b200.h)
//   The kNN kernel's scratch -- packed top-2 keys per query, the per-slice partial lists, the two counts -- lives in a
//   file-scope __device__ array instead of a context's cudaMalloc'ed buffers.  In HOST code the array's name is only a
//   stand-in for the symbol: its device address comes from cudaGetSymbolAddress, once per device with that device current.
// =================================================================================================
constexpr int kNaMax = MSLAM_B200_NOALLOC_MAX_DESC, kNaSlices = 4;
constexpr int kNaCounts = 64; // words in front: [0] n_from, [1] n_to (256-byte block)
constexpr int kNaWords = kNaCounts + 2 * kNaMax + 2 * kNaSlices * kNaMax;
static __device__ __align__(256) unsigned g_noalloc_scratch[kNaWords];

__global__ void k_noalloc_set(unsigned* scratch, int n_from, int n_to, int* n_out)
{
    if(threadIdx.x == 0)
    {
        scratch[0] = (unsigned)n_from;
        scratch[1] = (unsigned)n_to;
        *n_out = 0; // what the reference's matcher leaves for an empty / one-descriptor set; k_ratio_compact overwrites it
    }
}

struct NoAllocDevice
{
    int state = 0; // 0 not set up, 1 ready, -1 failed
    unsigned* scratch = nullptr;
    int sm_count = 0;
    cudaError_t err = cudaSuccess;
};
constexpr int kNaMaxDevices = 64;
static NoAllocDevice g_na_dev[kNaMaxDevices];
static pthread_mutex_t g_na_mutex = PTHREAD_MUTEX_INITIALIZER;

// once per device, with that device current
static cudaError_t noalloc_setup(int device, NoAllocDevice** out)
{
    NoAllocDevice& d = g_na_dev[device];
    pthread_mutex_lock(&g_na_mutex);
    if(d.state == 0)
    {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if(e == cudaSuccess)
            e = cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, device);
        d.scratch = static_cast<unsigned*>(p);
        d.err = e;
        d.state = (e == cudaSuccess && p) ? 1 : -1;
        if(e == cudaSuccess && !p)
            d.err = cudaErrorInvalidSymbol;
    }
    pthread_mutex_unlock(&g_na_mutex);
    *out = &d;
    return d.state == 1 ? cudaSuccess : d.err;
}

} // namespace mslam_b200

using namespace mslam_b200;

extern "C" int mslam_b200_match_knn_ratio_noalloc_dev(int device, void* cuda_stream, const uint8_t* d_from, int n_from,
                                                      const uint8_t* d_to, int n_to, double ratio, int* d_out_from, int* d_out_to,
                                                      int* d_out_dist, int* d_n_out, char* err_msg, size_t err_len)
{
    auto fail = [&](int status, const char* fmt, auto... a) {
        if(err_msg && err_len > 0)
            snprintf(err_msg, err_len, fmt, a...);
        return status;
    };
    if(err_msg && err_len > 0)
        err_msg[0] = 0;
    if(n_from > kNaMax || n_to > kNaMax)
        return fail(MSLAM_B200_E_INVALID,
                    "mslam_b200_match_knn_ratio_noalloc_dev: n_from = %d, n_to = %d exceed MSLAM_B200_NOALLOC_MAX_DESC = %d, the "
                    "size of the library's static scratch (use a context and mslam_b200_match_batch_dev for larger sets)",
                    n_from, n_to, kNaMax);
    if(n_from < 0 || n_to < 0 || !d_n_out || !(ratio > 0.0) || device < 0 || device >= kNaMaxDevices ||
       (n_to > 0 && n_from > 0 && (!d_from || !d_to || !d_out_from || !d_out_to)))
        return fail(MSLAM_B200_E_INVALID, "mslam_b200_match_knn_ratio_noalloc_dev: bad argument (null pointer, negative count, "
                                          "ratio <= 0 or device outside 0..%d)", kNaMaxDevices - 1);
    cudaError_t e = cudaSetDevice(device);
    NoAllocDevice* dv = nullptr;
    if(e == cudaSuccess)
        e = noalloc_setup(device, &dv);
    if(e != cudaSuccess)
        return fail(MSLAM_B200_E_CUDA, "mslam_b200_match_knn_ratio_noalloc_dev: set-up on device %d: %s (%s)", device,
                    cudaGetErrorName(e), cudaGetErrorString(e));
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    unsigned* sc = dv->scratch;
    const int* d_nfrom = reinterpret_cast<const int*>(sc);
    const int* d_nto = d_nfrom + 1;
    unsigned *key0 = sc + kNaCounts, *key1 = key0 + kNaMax, *part0 = key1 + kNaMax, *part1 = part0 + kNaSlices * kNaMax;
    k_noalloc_set<<<1, 32, 0, st>>>(sc, n_from, n_to, d_n_out);
    if(n_to >= 1 && n_from >= 2) // otherwise: no match (the reference is undefined there, orb_feature.cpp:88-91,101)
    {
        // one pair: `cap` is only the bound of the counts and the row stride between pairs, so the larger set's size does;
        // train set cut into slices until the 64-query CTAs fill the machine (the rule of launch_knn2 for small problems)
        const int cap = n_to > n_from ? n_to : n_from;
        const int ctas = (n_to + 63) / 64;
        int slices = 1;
        while(slices < kNaSlices && (long long)ctas * slices < 28LL * dv->sm_count && n_from / (slices * 2) >= 128)
            slices *= 2;
        dim3 grd(ctas, 1, slices);
        if(slices == 1)
            k_knn2<1, 64><<<grd, 64, 0, st>>>(d_to, d_nto, d_from, d_nfrom, cap, key0, key1);
        else
        {
            // slice z writes its partial lists at part + z * cap (gridDim.y = 1): kNaSlices * kNaMax words hold them
            k_knn2<1, 64><<<grd, 64, 0, st>>>(d_to, d_nto, d_from, d_nfrom, cap, part0, part1);
            k_merge_keys<<<(unsigned)((cap + 255) / 256), 256, 0, st>>>(part0, part1, slices, (long long)cap, key0, key1);
        }
        k_ratio_compact<<<1, 256, 0, st>>>(key0, key1, d_nto, d_nfrom, cap, ratio, d_out_from, d_out_to, d_out_dist, d_n_out);
    }
    e = cudaGetLastError();
    if(e != cudaSuccess)
        return fail(MSLAM_B200_E_CUDA, "mslam_b200_match_knn_ratio_noalloc_dev: launch: %s (%s)", cudaGetErrorName(e),
                    cudaGetErrorString(e));
    return MSLAM_B200_OK;
}
