// fast_simd.cuh -- the byte-SIMD pieces of FAST-9/16 shared by the two detectors' kernels (k_fast in orb_kernels.cu, k_dist_fast
// in dist_orb.cu): carry-out comparisons of four pixels per register, the bit-sliced "9 contiguous of 16" test, the ring
// word extraction and the exact cornerScore of two corners in one packed min / max network.  See k_fast's header comment.
#pragma once
#include <stdint.h>

namespace mslam_b200
{
__device__ __forceinline__ unsigned lop3_maj_a_nb_c(unsigned a, unsigned b, unsigned c)
{
    unsigned r; // majority(a, ~b, c)
    asm("lop3.b32 %0, %1, %2, %3, 0xB2;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ unsigned and3(unsigned a, unsigned b, unsigned c)
{
    unsigned r;
    asm("lop3.b32 %0, %1, %2, %3, 0x80;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ unsigned or3(unsigned a, unsigned b, unsigned c)
{
    unsigned r;
    asm("lop3.b32 %0, %1, %2, %3, 0xFE;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}

// bit 7 of each byte of the result: does the pixel have 9 contiguous set masks among m[0..15] (cyclic)?
// Every 9-arc is an even-aligned 8-arc 2j .. 2j+7 plus the position after it (start 2j) or the one before it (start 2j - 1):
//   e_j = pair (2j, 2j+1), f_j = e_j & e_j+1 (4 from 2j), arc8_j = f_j & f_j+2, any = OR_j arc8_j & (m_2j+8 | m_2j-1)
// -- 36 LOP3 (8 + 8 + 8 for the neighbours + 8 + 4) against 40 for the "three triples" form it replaces.
__device__ __forceinline__ unsigned any_run9(const unsigned (&m)[16])
{
    unsigned e[8], f[8], c[8];
#pragma unroll
    for(int j = 0; j < 8; ++j)
        e[j] = m[2 * j] & m[2 * j + 1];
#pragma unroll
    for(int j = 0; j < 8; ++j)
        f[j] = e[j] & e[(j + 1) & 7];
#pragma unroll
    for(int j = 0; j < 8; ++j)
    {
        const unsigned nb = m[(2 * j + 8) & 15] | m[(2 * j + 15) & 15];
        c[j] = and3(f[j], f[(j + 2) & 7], nb);
    }
    return or3(or3(c[0], c[1], c[2]), or3(c[3], c[4], c[5]), c[6] | c[7]);
}

// the 4 pixels starting dx (in -3..3) bytes from the start of word w0, out of the row's three staged words
__device__ __forceinline__ unsigned ring4(unsigned wm, unsigned w0, unsigned wp, int dx)
{
    switch(dx)
    {
    case -3: return __byte_perm(wm, w0, 0x4321);
    case -2: return __byte_perm(wm, w0, 0x5432);
    case -1: return __byte_perm(wm, w0, 0x6543);
    case 1: return __byte_perm(w0, wp, 0x4321);
    case 2: return __byte_perm(w0, wp, 0x5432);
    case 3: return __byte_perm(w0, wp, 0x6543);
    default: return w0;
    }
}

// exact cornerScore of TWO pixels known to be FAST corners, one per u16 lane of a packed min/max network.
// Each corner is read in the domain in which it is "centre above the whole arc": the raw tile for a corner whose arc is
// darker than the centre, the complemented tile (255 - pixel) for a corner whose arc is brighter (phase A knows which;
// the other side cannot win: two 9-arcs of a 16-ring always overlap).  In that domain
//   score = centre - min over the 16 arcs (max over the 9 ring pixels of the arc) - 1,
// with the arc maxima built from prefix/suffix maxima of the two half rings (28 + 32 VIMNMX.U16x2 for both corners).
template <int S /* row pitch of the staged tile in bytes */>
__device__ __forceinline__ unsigned corner_score2(const uint8_t* ca, const uint8_t* cb, unsigned bright_a, unsigned bright_b)
{
    constexpr int OFF[16] = {3 * S,  3 * S + 1,  2 * S + 2,  S + 3,  3,  -S + 3, -2 * S + 2, -3 * S + 1,
                             -3 * S, -3 * S - 1, -2 * S - 2, -S - 3, -3, S - 3,  2 * S - 2,  3 * S - 1};
    // lane value = pixel (arc darker than the centre) or 255 - pixel (arc brighter): p * (+-1) + (0 | 255), both lanes
    // through two IMADs (fma pipe); exact modulo 2^32 because every lane result lies in 0..255
    const unsigned sa = bright_a ? 0xffffffffu : 1u, sb = bright_b ? 0xffff0000u : 0x00010000u;
    const unsigned kk = (bright_a ? 255u : 0u) + (bright_b ? 255u << 16 : 0u);
    unsigned x[16];
#pragma unroll
    for(int k = 0; k < 16; ++k)
        x[k] = (unsigned)cb[OFF[k]] * sb + ((unsigned)ca[OFF[k]] * sa + kk);
    unsigned pa[8], sfa[8], pb[8], sfb[8]; // prefix / suffix maxima of the two half rings
    pa[0] = x[0], pb[0] = x[8], sfa[7] = x[7], sfb[7] = x[15];
#pragma unroll
    for(int i = 1; i < 8; ++i)
    {
        pa[i] = __vmaxu2(pa[i - 1], x[i]);
        pb[i] = __vmaxu2(pb[i - 1], x[8 + i]);
        sfa[7 - i] = __vmaxu2(sfa[8 - i], x[7 - i]);
        sfb[7 - i] = __vmaxu2(sfb[8 - i], x[15 - i]);
    }
    // arc k..k+8: k < 8 -> suffix of A from k + prefix of B to k ; k = 8 + j -> suffix of B from j + prefix of A to j
    unsigned m = 0xffffffffu;
#pragma unroll
    for(int k = 0; k < 8; ++k)
    {
        m = __vminu2(m, __vmaxu2(sfa[k], pb[k]));
        m = __vminu2(m, __vmaxu2(sfb[k], pa[k]));
    }
    const unsigned v = (unsigned)cb[0] * sb + ((unsigned)ca[0] * sa + kk);
    return v - m - 0x00010001u; // both lanes are >= fast_thr - 1 >= 0: no borrow crosses the lanes
}

} // namespace mslam_b200
