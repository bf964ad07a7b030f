// knn_mma.cu -- brute-force Hamming kNN (k = 2) on the 5th-generation tensor cores (tcgen05, sm_100a).
//
// Replaces, for large problems, the scalar XOR + POPC kernel of match_kernels.cu behind the same launcher contract
// (reference: cv::BFMatcher(NORM_HAMMING).knnMatch(query, train, k = 2), orb_feature.cpp:92-104 -- smallest distance,
// ties to the lower train index).
//
// A 256-bit descriptor is expanded to 256 int8 values +1 / -1 (k_expand_pm1).  For two descriptors a, b
//      dot(a, b) = (#equal bits) - (#different bits) = 256 - 2 * hamming(a, b)
// exactly, in int32.  So the all-pairs distance matrix of a (query set, train set) pair is one int8 GEMM with K = 256:
//      D[128 x 128] = Q[128 x 256] * T[128 x 256]^T        -> 8 tcgen05.mma.kind::i8 (M128 N128 K32) per tile
// and the packed result key of the scalar kernel, (hamming << 16) | train index, is one IMAD of the accumulator:
//      key = D * (-32768) + (0x800000 + index).
//
// One CTA = one (frame pair, 256 queries): the two 128-row query tiles stay in shared memory (64 KB), the train set
// streams through a four-stage ring of 128-descriptor tiles (32 KB each, cp.async.bulk from the expanded copy in HBM/L2,
// completion on mbarriers -- three tiles in flight per SM cover the L2 latency), each train tile feeds 2 x 8 MMAs
// (M128 N128 K32), accumulators are double-buffered in tensor memory (2 stages x 2 query tiles x 128 columns = all 512),
// and eight epilogue warps read them back with tcgen05.ld (one query row per lane) and keep the running top-2 in
// registers.  Roles: warps 0..7 = epilogue (warp w reads TMEM lanes 32 * (w % 4) .. + 31 of query tile w / 4), warp 8
// lane 0 = bulk-copy producer, warp 9 lane 0 = MMA issuer.
//
// Shared-memory operand layout (and, so that a tile is one contiguous range, the layout of the expanded copy in
// HBM): the no-swizzle K-major canonical layout -- core matrices of 8 rows x 16 bytes stored as 128 contiguous
// bytes;  [row / 8][k_byte / 16][row % 8][k_byte % 16]  =>  K-direction core-matrix stride (LBO) 128 B, 8-row group
// stride (SBO) 2048 B, and a K step of 32 bytes advances the descriptor start address by 256 B.
#include <cstdio>
#include "common.cuh"
#include "kernels.h"

namespace mslam_b200
{
namespace
{
constexpr unsigned kNoKey = 0xffffffffu;
constexpr int MMA_M = 128, MMA_N = 128;
constexpr int CTA_ROWS = 2 * MMA_M;          // queries per CTA (two MMA row tiles)
constexpr int ROW_BYTES = 256;               // expanded descriptor: 256 x int8
constexpr int A_BYTES = CTA_ROWS * ROW_BYTES; // 64 KB
constexpr int B_BYTES = MMA_N * ROW_BYTES;   // 32 KB
constexpr int STAGES = 4;                    // train-tile ring
constexpr int ACC_STAGES = 2;                // accumulator stages in tensor memory
constexpr int BULK = 16384;                  // bytes per cp.async.bulk
constexpr int MMA_THREADS = 256 + 64;        // 8 epilogue warps + producer warp + MMA warp
// The SMSP arbiter prefers the highest warp id among eligible warps: the two single-thread roles sit above the epilogue
// warps so that an MMA or a bulk copy is issued the moment its barrier opens, not when the ALU-bound epilogue warps of
// that SMSP happen to stall
constexpr int kProducerWarp = 8, kMmaWarp = 9;
constexpr int TMEM_COLS = 512;
constexpr int kDbGroup = 8;                   // query frames one DB-scoring CTA walks
constexpr int PAD_ROWS = 256;                // rows of an expanded set are padded to a multiple of this
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

// ---- CTA-pair (cta_group::2) forms: two CTAs of a cluster on the two SMs of a TPC execute one M = 256 MMA -- each computes
// its own 128 rows, reads its own query rows (A) and HALF of the train tile (B) from its shared memory and gets the other
// half from its peer, which halves the shared-memory bytes an MMA needs per SM ----
__device__ __forceinline__ void cluster_sync_all()
{
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// address of the same shared-memory location in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank)
{
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr)
{
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void umma_i8_pair(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
                 "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
                 : "memory");
}
// arrives (once the MMAs issued so far have completed) on the barrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit_pair(uint32_t bar)
{
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
                 "h"((unsigned short)3)
                 : "memory");
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

__device__ __forceinline__ void merge2(unsigned& k0, unsigned& k1, unsigned o0, unsigned o1)
{
    // two sorted pairs -> the two smallest of the four
    const unsigned lo = min(k0, o0), hi = max(k0, o0);
    k1 = min(hi, min(k1, o1));
    k0 = lo;
}

// Running top-2 of one row over the 128 columns of a tile, two columns per register.
// Inside a tile a key fits 16 bits: hamming (0..256) * 128 + column (0..127) <= 32895, so the even and the odd
// columns run as the two halves of VIMNMX.U16x2 (1.5 ALU-pipe operations per accumulator instead of 3), and both halves of a register are produced by two IMADs on the FMA pipe, no packing:
//      hamming * 128 + col = 16384 - 64 * D + col;   pk = D_even * -64 + C_even;   pk = D_odd * (-64 << 16) + pk
// 0xffff marks "no candidate" (columns that do not exist in a ragged last tile).
struct Top2x2
{
    unsigned k0 = 0xffffffffu, k1 = 0xffffffffu;
    unsigned a0 = 0xffffffffu, a1 = 0xffffffffu; // the second, independent chain
    __device__ __forceinline__ void push2_alt(unsigned a, unsigned b)
    {
        const unsigned m = __vminu2(a, b), M = __vmaxu2(a, b);
        a1 = __vimin3_u16x2(a1, __vmaxu2(a0, m), M);
        a0 = __vminu2(a0, m);
    }
    // merge the second chain into (k0, k1): two sorted pairs -> the two smallest of the four, per 16-bit lane
    __device__ __forceinline__ void fold()
    {
        const unsigned lo = __vminu2(k0, a0), hi = __vmaxu2(k0, a0);
        k1 = __vimin3_u16x2(hi, k1, a1);
        k0 = lo;
        a0 = a1 = 0xffffffffu;
    }
    // two new values at once: 5 packed min / max operations instead of 6 (second smallest of {k0, k1, a, b} =
    // min3(k1, max(k0, min(a, b)), max(a, b)); VIMNMX3.U16x2 is a native instruction)
    __device__ __forceinline__ void push2(unsigned a, unsigned b)
    {
        const unsigned m = __vminu2(a, b), M = __vmaxu2(a, b);
        k1 = __vimin3_u16x2(k1, __vmaxu2(k0, m), M);
        k0 = __vminu2(k0, m);
    }
};
__device__ __forceinline__ unsigned widen_key(unsigned k16, unsigned col_base)
{
    // 16-bit tile key -> (hamming << 16) | train index
    return k16 == 0xffffu ? kNoKey : ((k16 >> 7) << 16) | (col_base + (k16 & 127u));
}

// run-time copies of two constants (-64 and 65536): multiplying by a value the compiler cannot see keeps the key arithmetic
// in IMAD form on the FMA pipe; with literals it becomes LEA / shift sequences on the ALU pipe, which the min / max work binds
struct KeyMul
{
    unsigned neg64, two16;
};

// 32 accumulators (columns col0 .. col0 + 31 of the tile) -> 16 packed key pairs -> top-2 chains.
//      pk = (16384 - 64 D_even + col) | (16384 - 64 D_odd + col + 1) << 16  =  (D_odd * 65536 + D_even) * -64 + C
// (every lane value lies in 0 .. 32895, so the sum is exact and the lanes do not disturb each other): two IMADs per register
// (the pair updates alternate between two independent top-2 chains, t and its shadow t.alt, merged by fold(): the serial
// dependency through k0 / k1 is the epilogue's critical path -- ncu: "wait" is its first stall reason)
template <bool MASKED>
__device__ __forceinline__ void push32(Top2x2& t, const uint32_t (&v)[32], int col0, int n_valid, KeyMul km)
{
#pragma unroll
    for(int i = 0; i < 32; i += 4)
    {
        unsigned pk[2];
#pragma unroll
        for(int h = 0; h < 2; ++h)
        {
            const int c0 = col0 + i + 2 * h;
            const unsigned c = (16384u + (unsigned)c0) | ((16384u + (unsigned)(c0 + 1)) << 16);
            pk[h] = (v[i + 2 * h + 1] * km.two16 + v[i + 2 * h]) * km.neg64 + c;
            if(MASKED)
            {
                pk[h] |= c0 < n_valid ? 0u : 0x0000ffffu;
                pk[h] |= (c0 + 1) < n_valid ? 0u : 0xffff0000u;
            }
        }
        if(i & 4)
            t.push2_alt(pk[0], pk[1]);
        else
            t.push2(pk[0], pk[1]);
    }
}

// One 128-column accumulator (4 loads of 32 columns), double-buffered in registers.  RAGGED: only the first n_valid columns
// exist (the last tile of a set); the one 32-column group that straddles n_valid is masked, the groups past it are skipped.
template <bool RAGGED>
__device__ __forceinline__ void drain_tile(Top2x2& t, uint32_t taddr, int n_valid, KeyMul km)
{
    uint32_t va[32], vb[32];
    auto group = [&](const uint32_t (&v)[32], int col0) {
        if(!RAGGED || col0 + 32 <= n_valid) // warp-uniform
            push32<false>(t, v, col0, n_valid, km);
        else if(col0 < n_valid)
            push32<true>(t, v, col0, n_valid, km);
    };
    tmem_ld32(taddr, va);
    tmem_wait(va);
    if(!RAGGED || n_valid > 32)
        tmem_ld32(taddr + 32, vb);
    group(va, 0);
    if(RAGGED && n_valid <= 32)
        return;
    tmem_wait(vb);
    if(!RAGGED || n_valid > 64)
        tmem_ld32(taddr + 64, va);
    group(vb, 32);
    if(RAGGED && n_valid <= 64)
        return;
    tmem_wait(va);
    if(!RAGGED || n_valid > 96)
        tmem_ld32(taddr + 96, vb);
    group(va, 64);
    if(RAGGED && n_valid <= 96)
        return;
    tmem_wait(vb);
    group(vb, 96);
}

// ---- keyframe-database scoring needs distances only, not indices, and tcgen05.ld can PACK: with .pack::16b one register receives
// the low 16 bits of two adjacent 32-bit accumulator columns, (column 2i + 1) << 16 | column 2i (measured: tools/tmem_pack_probe.cu)
// -- exactly the pair the epilogue used to build with an IMAD, for dots in [-256, 256] as two's-complement s16 lanes.  So the
// database epilogue loads 64 columns into 32 registers, keeps the two LARGEST dots (= the two nearest neighbours) per signed
// 16-bit lane with VIMNMX.S16x2, and needs neither packing arithmetic nor biased accumulators (the first version of this path
// wrote 256 into every accumulator behind each drain to make the values non-negative for an IMAD pack). ----
__device__ __forceinline__ void tmem_ld64_pack(uint32_t taddr, uint32_t (&v)[32])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.pack::16b.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                   "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
                   "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
                   "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(taddr)
                 : "memory");
}
constexpr unsigned kNothingS16x2 = 0x80008000u; // -32768 in both lanes: below every dot
constexpr int kDotBias = 256;                   // the scalar results leave Top2MaxS as dot + 256 = 512 - 2 hamming, 0 = nothing
struct Top2MaxS
{
    unsigned k0 = kNothingS16x2, k1 = kNothingS16x2, a0 = kNothingS16x2, a1 = kNothingS16x2; // two independent chains
    __device__ __forceinline__ void push2(unsigned a, unsigned b)
    {
        const unsigned M = __vmaxs2(a, b), m = __vmins2(a, b);
        k1 = __vimax3_s16x2(k1, __vmins2(k0, M), m);
        k0 = __vmaxs2(k0, M);
    }
    __device__ __forceinline__ void push2_alt(unsigned a, unsigned b)
    {
        const unsigned M = __vmaxs2(a, b), m = __vmins2(a, b);
        a1 = __vimax3_s16x2(a1, __vmins2(a0, M), m);
        a0 = __vmaxs2(a0, M);
    }
    // the two largest of all four lanes' (largest, second largest), as dot + 256 (0 = nothing)
    __device__ __forceinline__ void top2(unsigned& m0, unsigned& m1) const
    {
        const unsigned hi = __vmaxs2(k0, a0), lo = __vmins2(k0, a0);
        const unsigned se = __vimax3_s16x2(lo, k1, a1); // per lane: (hi, se) = its two largest
        const int h0 = (short)(hi & 0xffffu), h1 = (short)(hi >> 16), s0 = (short)(se & 0xffffu), s1 = (short)(se >> 16);
        m0 = (unsigned)max(max(h0, h1) + kDotBias, 0);
        m1 = (unsigned)max(max(min(h0, h1), max(s0, s1)) + kDotBias, 0);
    }
};
__device__ __forceinline__ void merge2max(unsigned& m0, unsigned& m1, unsigned o0, unsigned o1)
{
    const unsigned hi = max(m0, o0), lo = min(m0, o0);
    m1 = max(lo, max(m1, o1));
    m0 = hi;
}
// 32 packed registers = 64 columns starting at col0 -> the max chains; MASKED: columns >= n_valid count as "nothing"
template <bool MASKED>
__device__ __forceinline__ void push64_pack(Top2MaxS& t, const uint32_t (&v)[32], int col0, int n_valid)
{
#pragma unroll
    for(int i = 0; i < 32; i += 2)
    {
        unsigned pk[2] = {v[i], v[i + 1]};
        if(MASKED)
        {
#pragma unroll
            for(int h = 0; h < 2; ++h)
            {
                const int c0 = col0 + 2 * (i + h);
                const unsigned keep = (c0 < n_valid ? 0x0000ffffu : 0u) | ((c0 + 1) < n_valid ? 0xffff0000u : 0u);
                pk[h] = (pk[h] & keep) | (kNothingS16x2 & ~keep);
            }
        }
        if(i & 2)
            t.push2_alt(pk[0], pk[1]);
        else
            t.push2(pk[0], pk[1]);
    }
}
// one 128-column key window: two packed loads of 64 columns, the second in flight while the first is processed
__device__ __forceinline__ void drain_tile_pack(Top2MaxS& t, uint32_t taddr, int n_valid)
{
    uint32_t va[32], vb[32];
    auto group = [&](const uint32_t (&v)[32], int col0) {
        if(col0 + 64 <= n_valid) // warp-uniform
            push64_pack<false>(t, v, col0, n_valid);
        else if(col0 < n_valid)
            push64_pack<true>(t, v, col0, n_valid);
    };
    tmem_ld64_pack(taddr, va);
    tmem_wait(va);
    tmem_ld64_pack(taddr + 64, vb);
    group(va, 0);
    tmem_wait(vb);
    group(vb, 64);
}

// NG groups of 32 accumulator columns starting at column col0 of a 128-column key window of which the first n_valid exist
// (n_valid may be <= col0: nothing to do).  One register buffer: this form runs with four epilogue warps per scheduler, which
// cover each other's tensor-memory load latency, and has to fit 112 registers.
template <int NG>
__device__ __forceinline__ void drain_cols(Top2x2& t, uint32_t taddr, int col0, int n_valid, KeyMul km)
{
    uint32_t v[32];
#pragma unroll
    for(int g = 0; g < NG; ++g)
    {
        const int c = col0 + 32 * g;
        if(c >= n_valid) // warp-uniform
            break;
        tmem_ld32(taddr + 32 * g, v);
        tmem_wait(v);
        if(c + 32 <= n_valid)
            push32<false>(t, v, c, n_valid, km);
        else
            push32<true>(t, v, c, n_valid, km);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// descriptors (32 bytes) -> +1 / -1 int8 rows in the canonical core-matrix order described at the top.
// One thread = four 16-byte K chunks of one descriptor (16 descriptor bits each).  Rows >= the set's count are left as they
// are: they only ever reach accumulator rows that are not stored / columns that the epilogue masks.
// ---------------------------------------------------------------------------------------------------------------
struct ExpandSet
{
    const uint8_t* desc;
    const int* counts;
    uint8_t* out;
};
__global__ void __launch_bounds__(256) k_expand_pm1(ExpandSet s0, ExpandSet s1, int cap, int cap_pad)
{
    const ExpandSet s = blockIdx.z ? s1 : s0;
    const int p = blockIdx.y;
    const int n = min(s.counts[p], cap);
    // (row % 8) fastest, then the K chunk, then the 8-row group: 8 consecutive lanes write one 128-byte core matrix.  A thread
    // takes chunks c, c + 4, c + 8, c + 12 of its row (four independent 16-byte stores; a quarter of the threads of the
    // one-chunk-per-thread version, which spent its time launching and retiring them)
    const int gid = blockIdx.x * 256 + threadIdx.x;
    const int row = ((gid >> 5) << 3) | (gid & 7), chunk0 = (gid >> 3) & 3;
    if(row >= n)
        return;
    const unsigned short* bits16 = reinterpret_cast<const unsigned short*>(s.desc + ((size_t)p * cap + row) * 32);
    uint8_t* dst = s.out + (size_t)p * cap_pad * ROW_BYTES + (size_t)(row >> 3) * 2048 + (size_t)(row & 7) * 16;
#pragma unroll
    for(int k = 0; k < 4; ++k)
    {
        const int chunk = chunk0 + 4 * k;
        const unsigned bits = bits16[chunk];
        unsigned w[4];
#pragma unroll
        for(int i = 0; i < 4; ++i)
        {
            const unsigned nib = (bits >> (4 * i)) & 15u;
            const unsigned spread = (nib * 0x00204081u) & 0x01010101u; // bit b of the nibble -> byte b, value 0 / 1
            w[i] = ~(spread * 0xfeu);                                  // 1 -> 0x01 (+1), 0 -> 0xff (-1)
        }
        *reinterpret_cast<uint4*>(dst + chunk * 128) = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// the kNN kernel.  grid = (ceil(cap / 256), n_pairs); 320 threads; one CTA per SM (all 512 TMEM columns).
// ---------------------------------------------------------------------------------------------------------------
// DB = false: kNN of pair p = blockIdx.y (rows: set q of the pair, columns: set t of the pair) -> key0 / key1.
// DB = true : keyframe-database scoring (k_db_scores of match_kernels.cu on the tensor cores): rows = descriptors of
//             keyframe blockIdx.z (set q, capacity cap), columns = descriptors of query frame blockIdx.y (set t, capacity
//             t_cap); a row counts when its two smallest distances pass the ratio test; the CTA adds its count to
//             scores[query frame][keyframe] (zeroed by the launcher; integer adds commute, so the result is exact).
__device__ unsigned long long g_mma_clk[2]; // MSLAM_MMA_DEBUG & 32: sum of SM cycles / of nanoseconds over sampled CTAs

// Timing experiments (switching parts of the kernel off; results are WRONG) exist only in builds with -DMSLAM_MMA_DEBUG
// (tools/gpu_mma_dbg.sh); the product library compiles the switches out.
#ifdef MSLAM_MMA_DEBUG
#define MMA_DBG(mask) (db.debug & (mask))
#else
#define MMA_DBG(mask) 0
#endif

struct DbArgs
{
    int t_cap, t_cap_pad, n_keyframes, int_ratio;
    int group, n_sets; // query frames per CTA, query frames in all
    int debug;         // timing experiments only (MSLAM_MMA_DEBUG; results are wrong): 1 = one K step per tile, 2 = epilogue
                       // loads tensor memory but does not process, 4 = epilogue does not load at all
    double ratio;
    int* scores;
    KeyMul km; // {-64, 65536}, set by the launchers
};
// CG = 1: one CTA per SM works alone (tcgen05 cta_group::1).  CG = 2: the CTAs 2k, 2k + 1 of grid.x form a cluster on one TPC
// and run as a pair (cta_group::2): CTA r keeps its own 256 query rows and loads train rows [64 r, 64 r + 64) of every
// 128-row tile (8 ring stages of 16 KB instead of 4 of 32 KB); the LEADER (rank 0) issues M = 256 MMAs for both; its
// tcgen05.commit arrives on both CTAs' "slot free" / "accumulators full" barriers (multicast); the peer forwards "my half of
// the tile has landed" to the leader's full barrier and both CTAs' epilogue warps report "accumulators drained" to the leader.
template <bool DB, int CG>
__global__ void __launch_bounds__(MMA_THREADS, 1)
    k_knn2_mma(const uint8_t* __restrict__ xq, const int* __restrict__ q_counts, const uint8_t* __restrict__ xt,
               const int* __restrict__ t_counts, int cap, int cap_pad, unsigned* __restrict__ key0, unsigned* __restrict__ key1,
               DbArgs db)
{
    extern __shared__ __align__(128) uint8_t smem[];
    // DB: the query frame is the faster grid dimension, so the CTAs that run together share one keyframe's rows (read from
    // HBM once) while the query frames (a few MB) stay in L2
    // ... and one CTA walks a group of db.group query frames against its resident keyframe rows: the pipeline (ring, tensor
    // memory, barriers) is set up once per group instead of once per (keyframe, query frame)
    const int p = DB ? blockIdx.z : blockIdx.y; // row set
    const int nq = min(q_counts[p], cap);
    const int q0 = blockIdx.x * CTA_ROWS;
    const uint32_t crank = CG == 2 ? (blockIdx.x & 1u) : 0u; // == %cluster_ctarank for cluster dims (2, 1, 1)
    const int q0p = CG == 2 ? (int)(blockIdx.x & ~1u) * CTA_ROWS : q0; // first row of the pair
    if(q0p >= nq)
        return; // whole CTA (CG = 2: the whole pair), before any barrier or allocation
    constexpr int NSTAGES = CG == 2 ? 2 * STAGES : STAGES;   // ring depth
    constexpr int STAGE_BYTES = B_BYTES / CG;                 // this CTA's share of a 128-row train tile
    long long clk0 = 0;
    unsigned long long ns0 = 0;
    if(MMA_DBG(32) && threadIdx.x == 0)
    {
        clk0 = clock64();
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(ns0));
    }
    const int t_pad = DB ? db.t_cap_pad : cap_pad, t_cap = DB ? db.t_cap : cap;
    const int pt_begin = DB ? blockIdx.y * db.group : p, pt_end = DB ? min(pt_begin + db.group, db.n_sets) : p + 1;
    // the column sets of this CTA, in order; a set with fewer than 2 descriptors scores 0 in DB mode and is skipped.
    // Every role walks the same sequence; `it` numbers the tiles of all sets consecutively (ring / accumulator phases).
    auto for_each_tile = [&](auto&& set_begin, auto&& tile, auto&& set_end) {
        for(int pt = pt_begin, it = 0; pt < pt_end; ++pt)
        {
            const int nt = min(t_counts[pt], t_cap);
            if(DB && nt < 2)
                continue;
            const int n_tiles = (nt + MMA_N - 1) / MMA_N;
            set_begin();
            for(int j = 0; j < n_tiles; ++j, ++it)
                tile(pt, nt, j, it);
            set_end(pt);
        }
    };
    // query tiles that exist: in this CTA (epilogue, A load) / in the pair (the MMAs cover row tile a of both CTAs)
    const int n_row_tiles = nq - q0 > MMA_M ? 2 : (nq - q0 > 0 ? 1 : 0);
    const int n_row_tiles_mma = nq - q0p > MMA_M ? 2 : 1;
    uint8_t* sA = smem;
    uint8_t* sB = smem + A_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + A_BYTES + STAGES * B_BYTES);
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 24);
    const uint32_t bar_a = smem_u32(bars + 0);
    const uint32_t bar_bfull = smem_u32(bars + 1), bar_bempty = smem_u32(bars + 1 + NSTAGES);          // [NSTAGES] each
    const uint32_t bar_cfull = smem_u32(bars + 1 + 2 * NSTAGES), bar_cempty = bar_cfull + 8 * ACC_STAGES; // [ACC_STAGES] each
    static_assert(1 + 2 * 2 * STAGES + 2 * ACC_STAGES <= 24, "barrier block");
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if(threadIdx.x == 0)
    {
        mbar_init(bar_a, 1);
        for(int s = 0; s < NSTAGES; ++s)
        {
            // full: the local producer's expect_tx arrival (+ its bytes); the leader of a pair also waits for the peer's relay
            mbar_init(bar_bfull + 8 * s, (CG == 2 && crank == 0) ? 2 : 1);
            mbar_init(bar_bempty + 8 * s, 1);
        }
        for(int s = 0; s < ACC_STAGES; ++s)
        {
            mbar_init(bar_cfull + 8 * s, 1);
            mbar_init(bar_cempty + 8 * s, 8 * CG); // one arrival per epilogue warp (of both CTAs, at the leader)
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if(warp == 0)
    {
        if(CG == 2)
        {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(TMEM_COLS)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        }
        else
        {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(TMEM_COLS)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    tc_fence_before();
    __syncthreads();
    if(CG == 2)
        cluster_sync_all(); // both CTAs' barriers are initialised before anything can arrive on them from the other side
    tc_fence_after();
    const uint32_t tmem = *s_tmem;

    if(warp == kProducerWarp)
    {
        if(lane == 0)
        {
            // ---- producer: query tiles once, then (this CTA's share of) the train tiles through the ring ----
            const uint8_t* a_src = xq + ((size_t)p * cap_pad + q0) * ROW_BYTES;
            const int a_bytes = n_row_tiles * MMA_M * ROW_BYTES;
            if(a_bytes > 0)
            {
                mbar_expect_tx(bar_a, a_bytes);
                for(int o = 0; o < a_bytes; o += BULK)
                    bulk_g2s(smem_u32(sA + o), a_src + o, BULK, bar_a);
            }
            else
                mbar_arrive(bar_a); // a peer without rows of its own still takes part (its half of every train tile)
            for_each_tile([] {}, [&](int pt, int, int j, int it) {
                    const int s = it % NSTAGES;
                    mbar_wait(bar_bempty + 8 * s, ((it / NSTAGES) & 1) ^ 1);
                    mbar_expect_tx(bar_bfull + 8 * s, STAGE_BYTES);
                    // rows [j * 128 + crank * (128 / CG), + 128 / CG) of the set: contiguous in the expanded layout
                    const uint8_t* b_src = xt + ((size_t)pt * t_pad + (size_t)j * MMA_N + (size_t)crank * (MMA_N / CG)) * ROW_BYTES;
                    for(int o = 0; o < STAGE_BYTES; o += BULK)
                        bulk_g2s(smem_u32(sB + s * STAGE_BYTES + o), b_src + o, BULK, bar_bfull + 8 * s);
                }, [](int) {});
        }
    }
    else if(warp == kMmaWarp)
    {
        if(lane == 0)
        {
            if(CG == 2 && crank != 0)
            {
                // ---- peer of a pair: no MMAs to issue; forward "my query rows / my half of tile `it` have landed" to the leader ----
                mbar_wait(bar_a, 0);
                for_each_tile([] {}, [&](int, int, int, int it) {
                        const int s = it % NSTAGES;
                        mbar_wait(bar_bfull + 8 * s, (it / NSTAGES) & 1);
                        mbar_arrive_cluster(mapa_u32(bar_bfull + 8 * s, 0));
                    }, [](int) {});
            }
            else
            {
                // ---- MMA issuer (CG = 2: for both CTAs of the pair) ----
                mbar_wait(bar_a, 0);
                const uint32_t a_addr = smem_u32(sA);
                constexpr uint32_t kIdescPair = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(MMA_N >> 3) << 17) | ((uint32_t)((2 * MMA_M) >> 4) << 24);
                for_each_tile([] {}, [&](int, int, int, int it) {
                        const int s = it % NSTAGES, as = it % ACC_STAGES;
                        mbar_wait(bar_bfull + 8 * s, (it / NSTAGES) & 1);
                        mbar_wait(bar_cempty + 8 * as, ((it / ACC_STAGES) & 1) ^ 1);
                        tc_fence_after();
                        const uint32_t b_addr = smem_u32(sB + s * STAGE_BYTES);
                        for(int a = 0; a < (MMA_DBG(16) ? 1 : n_row_tiles_mma); ++a)
                        {
#pragma unroll
                            for(int k = 0; k < ROW_BYTES / 32; ++k)
                                if(k == 0 || !MMA_DBG(1))
                                {
                                    if(CG == 2)
                                        umma_i8_pair(tmem + (as * 2 + a) * MMA_N, smem_desc(a_addr + a * (MMA_M * ROW_BYTES) + k * 256),
                                                     smem_desc(b_addr + k * 256), kIdescPair, k > 0);
                                    else
                                        umma_i8(tmem + (as * 2 + a) * MMA_N, smem_desc(a_addr + a * (MMA_M * ROW_BYTES) + k * 256),
                                                smem_desc(b_addr + k * 256), k > 0);
                                }
                        }
                        if(CG == 2)
                        {
                            umma_commit_pair(bar_bempty + 8 * s); // both CTAs' slot s may be overwritten once these MMAs have read it
                            umma_commit_pair(bar_cfull + 8 * as); // ... and the accumulator stage is complete in both CTAs
                        }
                        else
                        {
                            umma_commit(bar_bempty + 8 * s); // the train tile may be overwritten once these MMAs have read it
                            umma_commit(bar_cfull + 8 * as); // ... and the accumulator stage is complete
                        }
                    }, [](int) {});
            }
        }
    }
    else
    {
        // ---- epilogue: warp w (0..7) owns TMEM lanes 32 * (w % 4) .. of query tile a = w / 4 ----
        const int quarter = warp & 3, a = warp >> 2;
        const int row = a * MMA_M + quarter * 32 + lane;
        const bool active = a < n_row_tiles;
        const int q = q0 + row;
        unsigned k0 = kNoKey, k1 = kNoKey;
        for_each_tile(
            [&] { k0 = kNoKey, k1 = kNoKey; },
            [&](int, int nt, int j, int it) {
                const int as = it % ACC_STAGES;
                mbar_wait(bar_cfull + 8 * as, (it / ACC_STAGES) & 1);
                tc_fence_after();
                const uint32_t taddr = tmem + ((uint32_t)(quarter * 32) << 16) + (as * 2 + a) * MMA_N;
                const int col_base = j * MMA_N;
                const int n_valid = nt - col_base; // >= 1
                Top2x2 t;
                if(MMA_DBG(8))
                {
                    // ALU work of the epilogue on whatever the registers hold, no tensor-memory load
                    uint32_t va[32];
#pragma unroll
                    for(int i = 0; i < 32; ++i)
                        va[i] = (uint32_t)(lane * 7 + i + it);
                    if(active)
                        for(int c = 0; c < MMA_N; c += 32)
                        {
                            push32<false>(t, va, 0, MMA_N, db.km);
                            asm volatile("" : "+r"(va[0]), "+r"(va[1]));
                        }
                }
                else if(MMA_DBG(6))
                {
                    if(active && MMA_DBG(2))
                    {
                        uint32_t va[32];
                        for(int c = 0; c < MMA_N; c += 32)
                        {
                            tmem_ld32(taddr + c, va);
                            tmem_wait(va);
                            t.k0 ^= va[0] ^ va[31];
                        }
                    }
                }
                else if(active)
                {
                    if(n_valid >= MMA_N)
                        drain_tile<false>(t, taddr, n_valid, db.km);
                    else
                        drain_tile<true>(t, taddr, n_valid, db.km);
                }
                tc_fence_before();
                __syncwarp();
                if(lane == 0)
                {
                    if(CG == 2 && crank != 0)
                        mbar_arrive_cluster(mapa_u32(bar_cempty + 8 * as, 0)); // the leader issues the MMAs: it collects both CTAs' arrivals
                    else
                        mbar_arrive(bar_cempty + 8 * as);
                }
                t.fold();
                unsigned t0 = widen_key(t.k0 & 0xffffu, col_base), t1 = widen_key(t.k1 & 0xffffu, col_base);
                merge2(t0, t1, widen_key(t.k0 >> 16, col_base), widen_key(t.k1 >> 16, col_base));
                merge2(k0, k1, t0, t1);
            },
            [&](int pt) {
                if(!DB)
                {
                    if(active && q < nq)
                    {
                        key0[(long long)p * cap + q] = k0;
                        key1[(long long)p * cap + q] = k1;
                    }
                }
                else
                {
                    const int d0 = (int)(k0 >> 16), d1 = (int)(k1 >> 16); // nt >= 2: both exist
                    const bool pass = db.int_ratio ? (10 * d0 < 7 * d1) : ((double)(float)d0 < db.ratio * (double)(float)d1);
                    const int cnt = __reduce_add_sync(0xffffffffu, (active && q < nq && pass) ? 1 : 0);
                    if(lane == 0 && cnt)
                        atomicAdd(db.scores + (long long)pt * db.n_keyframes + p, cnt);
                }
            });
    }
    tc_fence_before();
    __syncthreads();
    if(MMA_DBG(32) && threadIdx.x == 0)
    {
        unsigned long long ns1;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(ns1));
        atomicAdd(&g_mma_clk[0], (unsigned long long)(clock64() - clk0));
        atomicAdd(&g_mma_clk[1], ns1 - ns0);
    }
    if(CG == 2)
        cluster_sync_all(); // neither CTA may leave (or free its tensor memory) while the other can still reach into it
    if(warp == 0)
    {
        __syncwarp();
        if(CG == 2)
            asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS) : "memory");
        else
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS) : "memory");
    }
}
// ---------------------------------------------------------------------------------------------------------------
// The same kernel with N = 256 instructions ("wide"; the default, MSLAM_MMA_WIDE=0 selects the N = 128 form above).  Why: an M128 N128 K32 MMA reads 4 KB of A and 4 KB of B
// from shared memory in its 64 cycles -- the whole 128 B/clk port -- so the bulk copies of the next train tile only land while
// the tensor pipe idles; an M128 N256 K32 MMA reads 4 + 8 KB in 128 cycles = 96 B/clk and leaves room for them.  A 256-column
// accumulator per query tile fills tensor memory (2 x 256 = 512 columns), so the two accumulators cannot be double-buffered
// per tile; instead the two QUERY TILES alternate: while the MMAs of query tile 1 run, all eight epilogue warps drain query
// tile 0 (warp w: lane quarter w % 4, column half w / 4), and vice versa.  The two column halves of a row meet in shared
// memory at the end of a set.  Train tiles are 256 descriptors (two adjacent 32 KB ring slots = one 64 KB stage, two stages).
// ---------------------------------------------------------------------------------------------------------------
constexpr int WIDE_N = 256, WIDE_STAGES = 2, WIDE_B_BYTES = WIDE_N * ROW_BYTES;
constexpr int WIDE_X_BYTES = 2 * 2 * MMA_M * 3 * 2 * 4; // [set parity][query tile][row][column part 1..3][2 keys] handed to column part 0
constexpr int SMEM_BYTES_WIDE = A_BYTES + WIDE_STAGES * WIDE_B_BYTES + 256 + WIDE_X_BYTES;
// EW epilogue warps (8 or 16): warp w drains lane quarter w % 4, columns (256 / (EW / 4)) * (w / 4) .. of the accumulator
template <bool DB, int EW>
__global__ void __launch_bounds__(EW * 32 + 64, 1)
    k_knn2_mma_wide(const uint8_t* __restrict__ xq, const int* __restrict__ q_counts, const uint8_t* __restrict__ xt,
                    const int* __restrict__ t_counts, int cap, int cap_pad, unsigned* __restrict__ key0, unsigned* __restrict__ key1,
                    DbArgs db)
{
    extern __shared__ __align__(128) uint8_t smem[];
    const int p = DB ? blockIdx.z : blockIdx.y; // row set
    const int nq = min(q_counts[p], cap);
    const int q0 = blockIdx.x * CTA_ROWS;
    if(q0 >= nq)
        return;
    const int t_pad = DB ? db.t_cap_pad : cap_pad, t_cap = DB ? db.t_cap : cap;
    const int pt_begin = DB ? blockIdx.y * db.group : p, pt_end = DB ? min(pt_begin + db.group, db.n_sets) : p + 1;
    auto for_each_tile = [&](auto&& set_begin, auto&& tile, auto&& set_end) {
        for(int pt = pt_begin, it = 0, set = 0; pt < pt_end; ++pt)
        {
            const int nt = min(t_counts[pt], t_cap);
            if(DB && nt < 2)
                continue;
            const int n_tiles = (nt + WIDE_N - 1) / WIDE_N;
            set_begin();
            for(int j = 0; j < n_tiles; ++j, ++it)
                tile(pt, nt, j, it);
            set_end(pt, set++);
        }
    };
    const int n_row_tiles = nq - q0 > MMA_M ? 2 : 1;
    // database scoring with the eight-warp epilogue reads its accumulators as packed s16 pairs (see drain_tile_pack)
    constexpr bool PACK = DB && EW == 8;
    uint8_t* sA = smem;
    uint8_t* sB = smem + A_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + A_BYTES + WIDE_STAGES * WIDE_B_BYTES);
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 24);
    unsigned* s_x = reinterpret_cast<unsigned*>(smem + A_BYTES + WIDE_STAGES * WIDE_B_BYTES + 256);
    const uint32_t bar_a = smem_u32(bars + 0);
    const uint32_t bar_bfull = smem_u32(bars + 1), bar_bempty = smem_u32(bars + 3); // [2] each
    const uint32_t bar_cfull = smem_u32(bars + 5), bar_cempty = smem_u32(bars + 7); // [query tile] each
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if(threadIdx.x == 0)
    {
        mbar_init(bar_a, 1);
        for(int s = 0; s < 2; ++s)
        {
            mbar_init(bar_bfull + 8 * s, 1);
            mbar_init(bar_bempty + 8 * s, 1);
            mbar_init(bar_cfull + 8 * s, 1);
            mbar_init(bar_cempty + 8 * s, EW); // one arrival per epilogue warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if(warp == 0)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s_tmem;

    constexpr int NPARTS = EW / 4, CPW = WIDE_N / NPARTS; // column parts of an accumulator, columns per warp (128 or 64)
    if(warp == EW) // producer (the two single-thread roles sit above the epilogue warps: arbiter priority)
    {
        if(lane == 0)
        {
            const uint8_t* a_src = xq + ((size_t)p * cap_pad + q0) * ROW_BYTES;
            const int a_bytes = n_row_tiles * MMA_M * ROW_BYTES;
            mbar_expect_tx(bar_a, a_bytes);
            for(int o = 0; o < a_bytes; o += BULK)
                bulk_g2s(smem_u32(sA + o), a_src + o, BULK, bar_a);
            for_each_tile([] {}, [&](int pt, int, int j, int it) {
                    const int s = it & 1;
                    mbar_wait(bar_bempty + 8 * s, ((it >> 1) & 1) ^ 1);
                    mbar_expect_tx(bar_bfull + 8 * s, WIDE_B_BYTES);
                    const uint8_t* b_src = xt + ((size_t)pt * t_pad + (size_t)j * WIDE_N) * ROW_BYTES;
                    for(int o = 0; o < WIDE_B_BYTES; o += BULK)
                        bulk_g2s(smem_u32(sB + s * WIDE_B_BYTES + o), b_src + o, BULK, bar_bfull + 8 * s);
                }, [](int, int) {});
        }
    }
    else if(warp == EW + 1)
    {
        if(lane == 0)
        {
            mbar_wait(bar_a, 0);
            const uint32_t a_addr = smem_u32(sA);
            constexpr uint32_t kIdescWide = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(WIDE_N >> 3) << 17) | ((uint32_t)(MMA_M >> 4) << 24);
            for_each_tile([] {}, [&](int, int, int, int it) {
                    const int s = it & 1;
                    mbar_wait(bar_bfull + 8 * s, (it >> 1) & 1);
                    const uint32_t b_addr = smem_u32(sB + s * WIDE_B_BYTES);
                    for(int a = 0; a < n_row_tiles; ++a)
                    {
                        // query tile a's accumulator of the previous train tile is drained
                        mbar_wait(bar_cempty + 8 * a, (it & 1) ^ 1);
                        tc_fence_after();
#pragma unroll
                        for(int k = 0; k < ROW_BYTES / 32; ++k)
                            if(k == 0 || !MMA_DBG(1))
                            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                         "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem + a * WIDE_N),
                                         "l"(smem_desc(a_addr + a * (MMA_M * ROW_BYTES) + k * 256)), "l"(smem_desc(b_addr + k * 256)),
                                         "r"(kIdescWide), "r"((uint32_t)(k > 0))
                                         : "memory");
                        umma_commit(bar_cfull + 8 * a);
                    }
                    umma_commit(bar_bempty + 8 * s);
                }, [](int, int) {});
        }
    }
    else
    {
        // ---- epilogue: warp w owns TMEM lanes 32 * (w % 4) .. and columns CPW * (w / 4) .. of BOTH query tiles, in turn ----
        const int quarter = warp & 3, part = warp >> 2;
        const int win = (part * CPW) / MMA_N, col0 = (part * CPW) % MMA_N; // 128-column key window of the part, first column in it
        unsigned k0[2], k1[2];
        for_each_tile(
            [&] { k0[0] = k0[1] = k1[0] = k1[1] = PACK ? 0u : kNoKey; },
            [&](int, int nt, int j, int it) {
                const int col_base = j * WIDE_N + win * MMA_N;
                const int n_valid = nt - col_base; // columns of this key window that exist (may be <= col0: nothing for this warp)
#pragma unroll
                for(int a = 0; a < 2; ++a)
                {
                    if(a >= n_row_tiles) // CTA-uniform
                        continue;
                    mbar_wait(bar_cfull + 8 * a, it & 1);
                    tc_fence_after();
                    const uint32_t taddr = tmem + ((uint32_t)(quarter * 32) << 16) + a * WIDE_N + part * CPW;
                    if(PACK)
                    {
                        Top2MaxS tm;
                        if(!MMA_DBG(4)) // debug 4: the epilogue only hands the accumulator back
                            drain_tile_pack(tm, taddr, max(n_valid, 0));
                        tc_fence_before();
                        __syncwarp();
                        if(lane == 0)
                            mbar_arrive(bar_cempty + 8 * a);
                        unsigned m0, m1;
                        tm.top2(m0, m1);
                        merge2max(k0[a], k1[a], m0, m1);
                        continue;
                    }
                    Top2x2 t;
                    if(CPW == MMA_N) // a whole key window per warp: the hand-scheduled four-group form
                    {
                        if(n_valid >= MMA_N)
                            drain_tile<false>(t, taddr, n_valid, db.km);
                        else if(n_valid > 0)
                            drain_tile<true>(t, taddr, n_valid, db.km);
                    }
                    else
                        drain_cols<CPW / 32>(t, taddr, col0, n_valid, db.km);
                    tc_fence_before();
                    __syncwarp();
                    if(lane == 0)
                        mbar_arrive(bar_cempty + 8 * a);
                    if(n_valid > col0)
                    {
                        t.fold();
                unsigned t0 = widen_key(t.k0 & 0xffffu, col_base), t1 = widen_key(t.k1 & 0xffffu, col_base);
                        merge2(t0, t1, widen_key(t.k0 >> 16, col_base), widen_key(t.k1 >> 16, col_base));
                        merge2(k0[a], k1[a], t0, t1);
                    }
                }
            },
            [&](int pt, int set) {
                // column parts 1.. hand their top-2 to part 0 (same rows), which finishes the set
                unsigned* xs = s_x + (set & 1) * (2 * MMA_M * 3 * 2);
                if(part > 0)
                {
#pragma unroll
                    for(int a = 0; a < 2; ++a)
                    {
                        unsigned* e = xs + (((a * MMA_M + quarter * 32 + lane) * 3) + (part - 1)) * 2;
                        e[0] = k0[a];
                        e[1] = k1[a];
                    }
                }
                asm volatile("bar.sync 1, %0;" ::"n"(EW * 32) : "memory"); // the epilogue warps
                if(part == 0)
                {
                    int cnt = 0;
#pragma unroll
                    for(int a = 0; a < 2; ++a)
                    {
                        const int q = q0 + a * MMA_M + quarter * 32 + lane;
#pragma unroll
                        for(int o = 0; o < NPARTS - 1; ++o)
                        {
                            const unsigned* e = xs + (((a * MMA_M + quarter * 32 + lane) * 3) + o) * 2;
                            if(PACK)
                                merge2max(k0[a], k1[a], e[0], e[1]);
                            else
                                merge2(k0[a], k1[a], e[0], e[1]);
                        }
                        const bool live = a < n_row_tiles && q < nq;
                        if(!DB)
                        {
                            if(live)
                            {
                                key0[(long long)p * cap + q] = k0[a];
                                key1[(long long)p * cap + q] = k1[a];
                            }
                        }
                        else
                        {
                            // nt >= 2: both exist.  PACK: the values are dot + 256 = 512 - 2 hamming
                            const int d0 = PACK ? (int)(512u - k0[a]) >> 1 : (int)(k0[a] >> 16), d1 = PACK ? (int)(512u - k1[a]) >> 1 : (int)(k1[a] >> 16);
                            const bool pass = db.int_ratio ? (10 * d0 < 7 * d1) : ((double)(float)d0 < db.ratio * (double)(float)d1);
                            cnt += (live && pass) ? 1 : 0;
                        }
                    }
                    if(DB)
                    {
                        cnt = __reduce_add_sync(0xffffffffu, cnt);
                        if(lane == 0 && cnt)
                            atomicAdd(db.scores + (long long)pt * db.n_keyframes + p, cnt);
                    }
                }
            });
    }
    tc_fence_before();
    __syncthreads();
    if(warp == 0)
    {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS) : "memory");
    }
}

// ---------------------------------------------------------------------------------------------------------------
// roofline denominator of the tensor-core matcher, MEASURED: the rate at which this GPU retires tcgen05.mma.kind::i8
// when nothing else runs -- operands resident in shared memory (+1/-1 bytes, like the matcher's), accumulators in tensor
// memory, one issuing thread per SM, no loads, no epilogue.  N = 128 is the instruction shape k_knn2_mma issues, N = 256
// the widest cta_group::1 shape.
// ---------------------------------------------------------------------------------------------------------------
template <int N>
__global__ void __launch_bounds__(128, 1) k_peak_i8(int iters, unsigned* sink)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* sA = smem;
    uint8_t* sB = smem + MMA_M * ROW_BYTES;
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + (MMA_M + N) * ROW_BYTES);
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar + 1);
    for(int i = threadIdx.x; i < (MMA_M + N) * ROW_BYTES / 4; i += blockDim.x)
    {
        unsigned h = (i + blockIdx.x * 7919u) * 2654435761u;
        h ^= h >> 15;
        // four bytes, each +1 or -1
        reinterpret_cast<unsigned*>(smem)[i] = 0x01010101u + ((h & 1u) * 0xfeu) + (((h >> 1) & 1u) * 0xfe00u) +
                                               (((h >> 2) & 1u) * 0xfe0000u) + (((h >> 3) & 1u) * 0xfe000000u);
    }
    if(threadIdx.x == 0)
    {
        mbar_init(smem_u32(bar), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if(threadIdx.x < 32)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // generic-proxy writes of the operands -> visible to the MMA's async proxy
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s_tmem;
    if(threadIdx.x == 0)
    {
        constexpr uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(MMA_M >> 4) << 24);
        const uint32_t a_addr = smem_u32(sA), b_addr = smem_u32(sB);
        for(int it = 0; it < iters; ++it)
        {
            const uint32_t d = tmem + (uint32_t)(it % (TMEM_COLS / N)) * N;
#pragma unroll
            for(int k = 0; k < ROW_BYTES / 32; ++k)
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                             "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                             "l"(smem_desc(a_addr + k * 256)), "l"(smem_desc(b_addr + k * 256)), "r"(idesc), "r"((uint32_t)(k > 0))
                             : "memory");
        }
        umma_commit(smem_u32(bar));
        for(unsigned spins = 0; !mbar_try_wait(smem_u32(bar), 0); ++spins)
            if(spins > (1u << 28))
                __trap();
        tc_fence_after();
    }
    __syncthreads();
    if(threadIdx.x < 32)
    {
        uint32_t v[32];
        tmem_ld32(tmem + ((uint32_t)0 << 16), v);
        tmem_wait(v);
        if(sink && v[threadIdx.x & 31] == 0x7fffffffu)
            sink[0] = v[0];
        tc_fence_before();
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS) : "memory");
    }
}
} // namespace

// out[0] / out[1]: int8 tensor operations per second (2 * M * N * K per MMA) of the N = 128 / N = 256 instruction shape,
// sustained (timed over the last quarter of ~0.4 s of back-to-back launches, i.e. at the power-limited clock a long scoring
// step runs at); out[2] / out[3]: the same for the first short launch after idle (burst).
cudaError_t measure_tensor_peak(cudaStream_t st, int sm_count, double* out4, long long* launches)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaError_t err = cudaSuccess;
    for(int shape = 0; shape < 2 && err == cudaSuccess; ++shape)
    {
        const int n = shape ? 256 : 128;
        const int smem = (MMA_M + n) * ROW_BYTES + 64;
        auto launch = [&](int iters) {
            if(shape)
                k_peak_i8<256><<<sm_count, 128, smem, st>>>(iters, nullptr);
            else
                k_peak_i8<128><<<sm_count, 128, smem, st>>>(iters, nullptr);
            ++*launches;
        };
        const int iters = shape ? 8192 : 16384; // ~2^30 cycles-worth / 128: about 4 ms per launch
        const double ops = (double)sm_count * iters * (ROW_BYTES / 32) * 2.0 * MMA_M * n * 32;
        float ms = 0.f;
        launch(64); // cold start (module load, first-touch of tensor memory)
        cudaStreamSynchronize(st);
        cudaEventRecord(e0, st);
        launch(iters);
        cudaEventRecord(e1, st);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        out4[2 + shape] = ops / (ms * 1e-3);
        const int reps = 100;
        for(int r = 0; r < reps; ++r)
        {
            if(r == reps * 3 / 4)
                cudaEventRecord(e0, st);
            launch(iters);
        }
        cudaEventRecord(e1, st);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        out4[shape] = ops * (reps - reps * 3 / 4) / (ms * 1e-3);
        err = cudaGetLastError();
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return err;
}
namespace
{
} // namespace

static int mma_debug()
{
#ifdef MSLAM_MMA_DEBUG
    static int v = -1;
    if(v < 0)
    {
        const char* e = getenv("MSLAM_MMA_DEBUG");
        v = e ? atoi(e) : 0;
    }
    return v;
#else
    return 0;
#endif
}
// cta_group of the matcher kernels: 1 = single CTAs (default), 2 = CTA pairs (MSLAM_MMA_CG=2).  The pair form is correct (same
// bits, tests/test_knn_tensor.py runs under either) but measured SLOWER on this workload: 3.49 vs 3.64 T pairs/s, because its
// feed path tops out at 4.8 T however few MMAs a tile gets (profiles/r2_v3_mma_experiments.txt): the leader's MMAs of tile
// it + 2 wait for the PEER's epilogue warps to report tile `it` drained, a commit-multicast + remote-arrive round trip per
// accumulator stage, and with all 512 tensor-memory columns in use there are only two stages to hide it behind.
static int mma_cta_group()
{
    static int v = -1;
    if(v < 0)
    {
        const char* e = getenv("MSLAM_MMA_CG");
        v = (e && atoi(e) == 2) ? 2 : 1;
    }
    return v;
}
// Which form of the kernel runs.  Default: the N = 256 form (k_knn2_mma_wide) with 8 epilogue warps -- measured on c5:
// 4.82 T pairs/s against 3.87 T for the N = 128 form (MSLAM_MMA_WIDE=0) and 3.86 T for 16 epilogue warps (=16: 96 registers,
// spills, and three times the end-of-set merging).  Same bits in every form.
static int mma_wide()
{
    static int v = -1;
    if(v < 0)
    {
        const char* e = getenv("MSLAM_MMA_WIDE");
        v = e ? atoi(e) : 8;
        if(v != 8 && v != 16)
            v = 0;
    }
    return v;
}
template <typename K, typename... Args>
static cudaError_t launch_pair(K kernel, dim3 grd, cudaStream_t st, Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grd;
    cfg.blockDim = dim3(MMA_THREADS);
    cfg.dynamicSmemBytes = SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2, at[0].val.clusterDim.y = 1, at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, args...);
}
// per DEVICE (cudaFuncSetAttribute applies to the current device only): called from mslam_b200_create after cudaSetDevice
cudaError_t configure_mma_kernels()
{
    cudaError_t e = cudaFuncSetAttribute(k_knn2_mma<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if(e != cudaSuccess)
        return e;
    e = cudaFuncSetAttribute(k_knn2_mma<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if(e != cudaSuccess)
        return e;
    e = cudaFuncSetAttribute(k_knn2_mma_wide<false, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES_WIDE);
    if(e != cudaSuccess)
        return e;
    e = cudaFuncSetAttribute(k_knn2_mma_wide<true, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES_WIDE);
    if(e != cudaSuccess)
        return e;
    e = cudaFuncSetAttribute(k_knn2_mma_wide<false, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES_WIDE);
    if(e != cudaSuccess)
        return e;
    e = cudaFuncSetAttribute(k_knn2_mma_wide<true, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES_WIDE);
    if(e != cudaSuccess)
        return e;
    e = cudaFuncSetAttribute(k_knn2_mma<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if(e != cudaSuccess)
        return e;
    e = cudaFuncSetAttribute(k_knn2_mma<true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if(e != cudaSuccess)
        return e;
    e = cudaFuncSetAttribute(k_peak_i8<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (MMA_M + 128) * ROW_BYTES + 64);
    if(e != cudaSuccess)
        return e;
    return cudaFuncSetAttribute(k_peak_i8<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (MMA_M + 256) * ROW_BYTES + 64);
}
int knn_mma_pad(int cap)
{
    return (cap + PAD_ROWS - 1) / PAD_ROWS * PAD_ROWS;
}
size_t knn_mma_bytes(int n_pairs, int cap)
{
    return (size_t)n_pairs * knn_mma_pad(cap) * ROW_BYTES;
}
void launch_expand_pm1(cudaStream_t st, const uint8_t* desc, const int* counts, int n_pairs, int cap, uint8_t* out,
                       long long* launches)
{
    const ExpandSet s{desc, counts, out};
    dim3 grd(((cap + 7) / 8 * 32 + 255) / 256, n_pairs, 1);
    k_expand_pm1<<<grd, 256, 0, st>>>(s, s, cap, knn_mma_pad(cap));
    ++*launches;
}
void launch_expand_pm1_x2(cudaStream_t st, const uint8_t* desc0, const int* counts0, uint8_t* out0, const uint8_t* desc1,
                          const int* counts1, uint8_t* out1, int n_pairs, int cap, long long* launches)
{
    dim3 grd(((cap + 7) / 8 * 32 + 255) / 256, n_pairs, 2);
    k_expand_pm1<<<grd, 256, 0, st>>>(ExpandSet{desc0, counts0, out0}, ExpandSet{desc1, counts1, out1}, cap, knn_mma_pad(cap));
    ++*launches;
}
cudaError_t launch_knn2_mma(cudaStream_t st, const uint8_t* xq, const int* qn, const uint8_t* xt, const int* tn, int n_pairs,
                            int cap, unsigned* key0, unsigned* key1, long long* launches)
{
    dim3 grd((cap + CTA_ROWS - 1) / CTA_ROWS, n_pairs);
    DbArgs none{};
    none.debug = mma_debug();
    none.km = KeyMul{(unsigned)-64, 65536u};
    if(mma_wide() == 16)
        k_knn2_mma_wide<false, 16><<<grd, 16 * 32 + 64, SMEM_BYTES_WIDE, st>>>(xq, qn, xt, tn, cap, knn_mma_pad(cap), key0, key1, none);
    else if(mma_wide() == 8)
        k_knn2_mma_wide<false, 8><<<grd, 8 * 32 + 64, SMEM_BYTES_WIDE, st>>>(xq, qn, xt, tn, cap, knn_mma_pad(cap), key0, key1, none);
    else if(mma_cta_group() == 2)
    {
        grd.x = (grd.x + 1) & ~1u; // pairs: CTAs 2k, 2k + 1 are one cluster
        cudaError_t e = launch_pair(k_knn2_mma<false, 2>, grd, st, xq, qn, xt, tn, cap, knn_mma_pad(cap), key0, key1, none);
        if(e != cudaSuccess)
            return e;
    }
    else
        k_knn2_mma<false, 1><<<grd, MMA_THREADS, SMEM_BYTES, st>>>(xq, qn, xt, tn, cap, knn_mma_pad(cap), key0, key1, none);
    ++*launches;
    return cudaSuccess;
}
// xdb: expanded descriptors of keyframes [0, n_keyframes) of this call [n_keyframes][pad(k_cap)], xqueries: expanded query
// frames [nq_frames][pad(q_cap)].  Adds into scores[q * score_stride + kf] -- the caller zeroes the score matrix.
cudaError_t launch_db_scores_mma(cudaStream_t st, const uint8_t* xqueries, const int* q_counts, int nq_frames, int q_cap,
                                 const uint8_t* xdb, const int* k_counts, int n_keyframes, int k_cap, double ratio, int* scores,
                                 int score_stride, long long* launches)
{
    DbArgs db;
    db.t_cap = q_cap, db.t_cap_pad = knn_mma_pad(q_cap), db.n_keyframes = score_stride, db.int_ratio = ratio == 0.7;
    db.ratio = ratio, db.scores = scores;
    db.group = kDbGroup, db.n_sets = nq_frames;
    db.debug = mma_debug();
    db.km = KeyMul{(unsigned)-64, 65536u};
    // grid.y = query frames, grid.z = keyframes (<= 65535 per launch)
    for(int kf0 = 0; kf0 < n_keyframes; kf0 += 65535)
    {
        const int nk = n_keyframes - kf0 < 65535 ? n_keyframes - kf0 : 65535;
        DbArgs d = db;
        d.scores = scores + kf0;
        dim3 grd((k_cap + CTA_ROWS - 1) / CTA_ROWS, (nq_frames + kDbGroup - 1) / kDbGroup, nk);
        if(mma_wide() == 16)
            k_knn2_mma_wide<true, 16><<<grd, 16 * 32 + 64, SMEM_BYTES_WIDE, st>>>(xdb + (size_t)kf0 * knn_mma_pad(k_cap) * ROW_BYTES, k_counts + kf0,
                                                                                 xqueries, q_counts, k_cap, knn_mma_pad(k_cap), nullptr, nullptr, d);
        else if(mma_wide() == 8)
            k_knn2_mma_wide<true, 8><<<grd, 8 * 32 + 64, SMEM_BYTES_WIDE, st>>>(xdb + (size_t)kf0 * knn_mma_pad(k_cap) * ROW_BYTES, k_counts + kf0,
                                                                               xqueries, q_counts, k_cap, knn_mma_pad(k_cap), nullptr, nullptr, d);
        else if(mma_cta_group() == 2)
        {
            grd.x = (grd.x + 1) & ~1u;
            cudaError_t e = launch_pair(k_knn2_mma<true, 2>, grd, st, xdb + (size_t)kf0 * knn_mma_pad(k_cap) * ROW_BYTES, k_counts + kf0,
                                        xqueries, q_counts, k_cap, knn_mma_pad(k_cap), (unsigned*)nullptr, (unsigned*)nullptr, d);
            if(e != cudaSuccess)
                return e;
        }
        else
            k_knn2_mma<true, 1><<<grd, MMA_THREADS, SMEM_BYTES, st>>>(xdb + (size_t)kf0 * knn_mma_pad(k_cap) * ROW_BYTES, k_counts + kf0,
                                                                  xqueries, q_counts, k_cap, knn_mma_pad(k_cap), nullptr, nullptr, d);
        ++*launches;
    }
    if(MMA_DBG(32))
    {
        unsigned long long h[2] = {0, 0};
        cudaStreamSynchronize(st);
        cudaMemcpyFromSymbol(h, g_mma_clk, sizeof(h));
        fprintf(stderr, "k_knn2_mma: %.1f MHz SM clock inside the kernel (%llu cycles / %llu ns summed over CTAs)\n",
                h[1] ? 1e3 * (double)h[0] / (double)h[1] : 0.0, h[0], h[1]);
        unsigned long long z[2] = {0, 0};
        cudaMemcpyToSymbol(g_mma_clk, z, sizeof(z));
    }
    return cudaGetLastError();
}
} // namespace mslam_b200
