// knn_mma.cu -- brute-force Hamming kNN (k = 2) on the 5th-generation tensor cores (tcgen05, sm_100a).
//
// Replaces, for large problems, the scalar XOR + POPC kernel of match_kernels.cu behind the same launcher contract
// (reference: cv::BFMatcher(NORM_HAMMING).knnMatch(query, train, k = 2), orb_feature.cpp:92-104 -- smallest distance,
// ties to the lower train index).
//
// A 256-bit descriptor is expanded to 256 int8 values +1 / -1 (k_expand_pm1).  For two descriptors a, b
//      dot(a, b) = (#equal bits) - (#different bits) = 256 - 2 * hamming(a, b)
// exactly, in int32.  So the all-pairs distance matrix of a (query set, train set) pair is one int8 GEMM with K = 256:
//      D[128 x 256] = Q[128 x 256] * T[256 x 256]^T        -> 8 tcgen05.mma.kind::i8 (M128 N256 K32) per tile
// and the packed result key of the scalar kernel, (hamming << 16) | train index, is one IMAD of the accumulator:
//      key = D * (-32768) + (0x800000 + index).
//
// One CTA = one (frame pair, 128 queries): the query tile stays in shared memory, the train set streams through a
// two-stage ring of 64 KB tiles (cp.async.bulk from the expanded copy in HBM/L2, completion on mbarriers), accumulators
// are double-buffered in tensor memory (2 x 256 columns), and eight epilogue warps read them back with tcgen05.ld
// (one query row per lane) and keep the running top-2 in registers.  Roles: warp 0 lane 0 = bulk-copy producer,
// warp 1 lane 0 = MMA issuer, warps 2..9 = epilogue (warp w reads TMEM lanes 32 * (w % 4) .. + 31; the two warps of a
// lane quarter split the 256 columns).
//
// Shared-memory operand layout (and, so that a tile is one contiguous range, the layout of the expanded copy in
// HBM): the no-swizzle K-major canonical layout -- core matrices of 8 rows x 16 bytes stored as 128 contiguous
// bytes;  [row / 8][k_byte / 16][row % 8][k_byte % 16]  =>  K-direction core-matrix stride (LBO) 128 B, 8-row group
// stride (SBO) 2048 B, and a K step of 32 bytes advances the descriptor start address by 256 B.
#include "common.cuh"
#include "kernels.h"

namespace mslam_b200
{
namespace
{
constexpr unsigned kNoKey = 0xffffffffu;
constexpr int MMA_M = 128, MMA_N = 256;
constexpr int ROW_BYTES = 256;               // expanded descriptor: 256 x int8
constexpr int A_BYTES = MMA_M * ROW_BYTES;   // 32 KB
constexpr int B_BYTES = MMA_N * ROW_BYTES;   // 64 KB
constexpr int STAGES = 2;
constexpr int BULK = 16384;                  // bytes per cp.async.bulk
constexpr int MMA_THREADS = 64 + 256;        // producer warp + MMA warp + 8 epilogue warps
constexpr int TMEM_COLS = 512;
constexpr int SMEM_BYTES = A_BYTES + STAGES * B_BYTES + 256;

// instruction descriptor of tcgen05.mma.kind::i8: D = S32 (bits 4-5 = 2), A = B = signed 8 bit (bits 7-9, 10-12 = 1),
// both operands K-major (bits 15, 16 = 0), N >> 3 at bits 17-22, M >> 4 at bits 24-28
constexpr uint32_t kIdesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(MMA_N >> 3) << 17) | ((uint32_t)(MMA_M >> 4) << 24);

__device__ __forceinline__ uint32_t smem_u32(const void* p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok)
                 : "r"(bar), "r"(parity)
                 : "memory");
    return ok != 0;
}
// bounded wait: a protocol error becomes a trap (CUDA error on the host), never a hung GPU
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    for(unsigned spins = 0; !mbar_try_wait(bar, parity); ++spins)
        if(spins > (1u << 22))
            __trap();
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                 "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void tc_fence_before()
{
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after()
{
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// shared-memory matrix descriptor: start >> 4 | LBO >> 4 at bit 16 | SBO >> 4 at bit 32 | version 1 at bit 46 | no swizzle
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr)
{
    return (uint64_t)((addr >> 4) & 0x3fffu) | ((uint64_t)(128 >> 4) << 16) | ((uint64_t)(2048 >> 4) << 32) | (1ull << 46);
}
__device__ __forceinline__ void umma_i8(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
                 "l"(adesc), "l"(bdesc), "r"(kIdesc), "r"(accumulate)
                 : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// 32 consecutive accumulator columns of this thread's TMEM lane (asynchronous: the registers are valid after tmem_wait)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                   "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
                   "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
                   "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(taddr)
                 : "memory");
}
// tcgen05.wait::ld, tied to the registers of the load it completes so that no use of them can be scheduled above it
__device__ __forceinline__ void tmem_wait(uint32_t (&v)[32])
{
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]),
                   "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]), "+r"(v[16]),
                   "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]), "+r"(v[23]), "+r"(v[24]),
                   "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]), "+r"(v[30]), "+r"(v[31])
                 :
                 : "memory");
}

// running top-2 (smallest keys) as two independent chains, merged at the end
struct Top2
{
    unsigned a0 = kNoKey, a1 = kNoKey, b0 = kNoKey, b1 = kNoKey;
    __device__ __forceinline__ void push_a(unsigned key)
    {
        a1 = min(a1, max(a0, key));
        a0 = min(a0, key);
    }
    __device__ __forceinline__ void push_b(unsigned key)
    {
        b1 = min(b1, max(b0, key));
        b0 = min(b0, key);
    }
};
__device__ __forceinline__ void merge2(unsigned& k0, unsigned& k1, unsigned o0, unsigned o1)
{
    // two sorted pairs -> the two smallest of the four
    const unsigned lo = min(k0, o0), hi = max(k0, o0);
    k1 = min(hi, min(k1, o1));
    k0 = lo;
}

// 32 accumulators (columns col0 .. col0 + 31 of the tile) -> keys -> top-2 chains.  MASKED: columns >= n_valid are
// train descriptors that do not exist (last tile of a ragged set).
template <bool MASKED>
__device__ __forceinline__ void push32(Top2& t, const uint32_t (&v)[32], int col0, int n_valid)
{
#pragma unroll
    for(int i = 0; i < 32; ++i)
    {
        // (256 - D) << 15 == hamming << 16 ; + local train index
        unsigned key = (unsigned)((int)v[i] * -32768) + (0x800000u + (unsigned)(col0 + i));
        if(MASKED)
            key = (col0 + i) < n_valid ? key : kNoKey;
        if(i & 1)
            t.push_b(key);
        else
            t.push_a(key);
    }
}

// One accumulator stage, this warp's half of the columns (128 columns = 4 loads of 32), double-buffered in registers.
template <bool MASKED>
__device__ __forceinline__ void drain_half(Top2& t, uint32_t taddr, int col_begin, int n_valid)
{
    uint32_t va[32], vb[32];
    tmem_ld32(taddr + col_begin, va);
    tmem_wait(va);
    tmem_ld32(taddr + col_begin + 32, vb);
    push32<MASKED>(t, va, col_begin, n_valid);
    tmem_wait(vb);
    tmem_ld32(taddr + col_begin + 64, va);
    push32<MASKED>(t, vb, col_begin + 32, n_valid);
    tmem_wait(va);
    tmem_ld32(taddr + col_begin + 96, vb);
    push32<MASKED>(t, va, col_begin + 64, n_valid);
    tmem_wait(vb);
    push32<MASKED>(t, vb, col_begin + 96, n_valid);
}

// ---------------------------------------------------------------------------------------------------------------
// descriptors (32 bytes) -> +1 / -1 int8 rows in the canonical core-matrix order described at the top.
// One thread = one 16-byte K chunk of one descriptor (16 descriptor bits).  Rows >= the set's count are left as they
// are: they only ever reach accumulator rows that are not stored / columns that the epilogue masks.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_expand_pm1(const uint8_t* __restrict__ desc, const int* __restrict__ counts, int cap,
                                                    int cap_pad, uint8_t* __restrict__ out)
{
    const int p = blockIdx.y;
    const int n = min(counts[p], cap);
    const int gid = blockIdx.x * 256 + threadIdx.x; // (row, chunk): chunk fastest within a row's 16 chunks
    const int row = gid >> 4, chunk = gid & 15;
    if(row >= n)
        return;
    const unsigned bits = reinterpret_cast<const unsigned short*>(desc + ((size_t)p * cap + row) * 32)[chunk];
    uint4 o;
    unsigned w[4];
#pragma unroll
    for(int i = 0; i < 4; ++i)
    {
        const unsigned nib = (bits >> (4 * i)) & 15u;
        const unsigned spread = (nib * 0x00204081u) & 0x01010101u; // bit b of the nibble -> byte b, value 0 / 1
        w[i] = ~(spread * 0xfeu);                                  // 1 -> 0x01 (+1), 0 -> 0xff (-1)
    }
    o = make_uint4(w[0], w[1], w[2], w[3]);
    uint8_t* dst = out + (size_t)p * cap_pad * ROW_BYTES + (size_t)(row >> 3) * 2048 + chunk * 128 + (row & 7) * 16;
    *reinterpret_cast<uint4*>(dst) = o;
}

// ---------------------------------------------------------------------------------------------------------------
// the kNN kernel.  grid = (ceil(cap / 128), n_pairs); 320 threads; one CTA per SM (all 512 TMEM columns).
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(MMA_THREADS, 1)
    k_knn2_mma(const uint8_t* __restrict__ xq, const int* __restrict__ q_counts, const uint8_t* __restrict__ xt,
               const int* __restrict__ t_counts, int cap, int cap_pad, unsigned* __restrict__ key0, unsigned* __restrict__ key1)
{
    extern __shared__ __align__(128) uint8_t smem[];
    const int p = blockIdx.y;
    const int nq = min(q_counts[p], cap), nt = min(t_counts[p], cap);
    const int q0 = blockIdx.x * MMA_M;
    if(q0 >= nq)
        return; // whole CTA, before any barrier or allocation
    const int n_tiles = (nt + MMA_N - 1) / MMA_N;
    uint8_t* sA = smem;
    uint8_t* sB = smem + A_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + A_BYTES + STAGES * B_BYTES);
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 16);
    __shared__ unsigned s_part[2][MMA_M];
    const uint32_t bar_a = smem_u32(bars + 0);
    const uint32_t bar_bfull = smem_u32(bars + 1), bar_bempty = smem_u32(bars + 3);   // [STAGES] each
    const uint32_t bar_cfull = smem_u32(bars + 5), bar_cempty = smem_u32(bars + 7);   // accumulator stages
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if(threadIdx.x == 0)
    {
        mbar_init(bar_a, 1);
        for(int s = 0; s < STAGES; ++s)
        {
            mbar_init(bar_bfull + 8 * s, 1);
            mbar_init(bar_bempty + 8 * s, 1);
            mbar_init(bar_cfull + 8 * s, 1);
            mbar_init(bar_cempty + 8 * s, 8); // one arrival per epilogue warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if(warp == 0)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s_tmem;

    if(warp == 0)
    {
        if(lane == 0)
        {
            // ---- producer: query tile once, then the train tiles through the ring ----
            const uint8_t* a_src = xq + ((size_t)p * cap_pad + q0) * ROW_BYTES;
            mbar_expect_tx(bar_a, A_BYTES);
            for(int o = 0; o < A_BYTES; o += BULK)
                bulk_g2s(smem_u32(sA + o), a_src + o, BULK, bar_a);
            for(int it = 0; it < n_tiles; ++it)
            {
                const int s = it % STAGES;
                mbar_wait(bar_bempty + 8 * s, ((it / STAGES) & 1) ^ 1);
                mbar_expect_tx(bar_bfull + 8 * s, B_BYTES);
                const uint8_t* b_src = xt + ((size_t)p * cap_pad + (size_t)it * MMA_N) * ROW_BYTES;
                for(int o = 0; o < B_BYTES; o += BULK)
                    bulk_g2s(smem_u32(sB + s * B_BYTES + o), b_src + o, BULK, bar_bfull + 8 * s);
            }
        }
    }
    else if(warp == 1)
    {
        if(lane == 0)
        {
            // ---- MMA issuer ----
            mbar_wait(bar_a, 0);
            const uint32_t a_addr = smem_u32(sA);
            for(int it = 0; it < n_tiles; ++it)
            {
                const int s = it % STAGES;
                const uint32_t ph = (it / STAGES) & 1;
                mbar_wait(bar_bfull + 8 * s, ph);
                mbar_wait(bar_cempty + 8 * s, ph ^ 1);
                tc_fence_after();
                const uint32_t b_addr = smem_u32(sB + s * B_BYTES);
#pragma unroll
                for(int k = 0; k < ROW_BYTES / 32; ++k)
                    umma_i8(tmem + s * MMA_N, smem_desc(a_addr + k * 256), smem_desc(b_addr + k * 256), k > 0);
                umma_commit(bar_bempty + 8 * s); // the train tile may be overwritten once these MMAs have read it
                umma_commit(bar_cfull + 8 * s);  // ... and the accumulator stage is complete
            }
        }
    }
    else
    {
        // ---- epilogue: warp w owns TMEM lanes 32 * (w % 4) ..; the two warps of a quarter split the columns ----
        const int quarter = warp & 3, half = (warp - 2) >> 2;
        const int row = quarter * 32 + lane;
        unsigned k0 = kNoKey, k1 = kNoKey;
        for(int it = 0; it < n_tiles; ++it)
        {
            const int s = it % STAGES;
            mbar_wait(bar_cfull + 8 * s, (it / STAGES) & 1);
            tc_fence_after();
            const uint32_t taddr = tmem + ((uint32_t)(quarter * 32) << 16) + s * MMA_N;
            const int n_valid = nt - it * MMA_N; // >= 1
            Top2 t;
            if(n_valid >= MMA_N)
                drain_half<false>(t, taddr, half * 128, n_valid);
            else
                drain_half<true>(t, taddr, half * 128, n_valid);
            tc_fence_before();
            __syncwarp();
            if(lane == 0)
                mbar_arrive(bar_cempty + 8 * s);
            unsigned t0 = t.a0, t1 = t.a1;
            merge2(t0, t1, t.b0, t.b1);
            const unsigned base = (unsigned)(it * MMA_N);
            t0 = t0 == kNoKey ? t0 : t0 + base;
            t1 = t1 == kNoKey ? t1 : t1 + base;
            merge2(k0, k1, t0, t1);
        }
        // the upper-half warp hands its pair to the lower-half warp of the same lane quarter
        if(half == 1)
        {
            s_part[0][row] = k0;
            s_part[1][row] = k1;
        }
        asm volatile("bar.sync 1, 256;" ::: "memory"); // the eight epilogue warps only
        if(half == 0)
        {
            merge2(k0, k1, s_part[0][row], s_part[1][row]);
            const int q = q0 + row;
            if(q < nq)
            {
                key0[(long long)p * cap + q] = k0;
                key1[(long long)p * cap + q] = k1;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if(warp == 0)
    {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS) : "memory");
    }
}
} // namespace

int knn_mma_pad(int cap)
{
    return (cap + MMA_N - 1) / MMA_N * MMA_N;
}
size_t knn_mma_bytes(int n_pairs, int cap)
{
    return (size_t)n_pairs * knn_mma_pad(cap) * ROW_BYTES;
}
void launch_expand_pm1(cudaStream_t st, const uint8_t* desc, const int* counts, int n_pairs, int cap, uint8_t* out,
                       long long* launches)
{
    dim3 grd((cap * 16 + 255) / 256, n_pairs);
    k_expand_pm1<<<grd, 256, 0, st>>>(desc, counts, cap, knn_mma_pad(cap), out);
    ++*launches;
}
cudaError_t launch_knn2_mma(cudaStream_t st, const uint8_t* xq, const int* qn, const uint8_t* xt, const int* tn, int n_pairs,
                            int cap, unsigned* key0, unsigned* key1, long long* launches)
{
    static bool configured = false;
    if(!configured)
    {
        cudaError_t e = cudaFuncSetAttribute(k_knn2_mma, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
        if(e != cudaSuccess)
            return e;
        configured = true;
    }
    dim3 grd((cap + MMA_M - 1) / MMA_M, n_pairs);
    k_knn2_mma<<<grd, MMA_THREADS, SMEM_BYTES, st>>>(xq, qn, xt, tn, cap, knn_mma_pad(cap), key0, key1);
    ++*launches;
    return cudaSuccess;
}
} // namespace mslam_b200
