// bc45_block.cuh - the BC4 / BC5 block arithmetic shared by the stand-alone encoder (bc45_encode.cu) and the
// fused mip + encode kernel (mip_kernels.cu).  Integer restatement of the reference's vendored
// /root/reference/src/libs/stb/stb_dxt.h:547-597 (stb__CompressAlphaBlock), bit exact against that code.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rosy_tex {
namespace bc45 {

// floor(2^20 / d) + 1 for d = 1..255 (0 for d = 0): exact reciprocals for quotients a / d with a * d < 2^20
__device__ const uint32_t g_RECIP20[256] = {
    0, 1048577, 524289, 349526, 262145, 209716, 174763, 149797, 131073, 116509, 104858, 95326,
    87382, 80660, 74899, 69906, 65537, 61681, 58255, 55189, 52429, 49933, 47663, 45591,
    43691, 41944, 40330, 38837, 37450, 36158, 34953, 33826, 32769, 31776, 30841, 29960,
    29128, 28340, 27595, 26887, 26215, 25576, 24967, 24386, 23832, 23302, 22796, 22311,
    21846, 21400, 20972, 20561, 20165, 19785, 19419, 19066, 18725, 18397, 18079, 17773,
    17477, 17190, 16913, 16645, 16385, 16132, 15888, 15651, 15421, 15197, 14980, 14769,
    14564, 14365, 14170, 13982, 13798, 13618, 13444, 13274, 13108, 12946, 12788, 12634,
    12484, 12337, 12193, 12053, 11916, 11782, 11651, 11523, 11398, 11276, 11156, 11038,
    10923, 10811, 10700, 10592, 10486, 10382, 10281, 10181, 10083, 9987, 9893, 9800,
    9710, 9620, 9533, 9447, 9363, 9280, 9199, 9119, 9040, 8963, 8887, 8812,
    8739, 8666, 8595, 8526, 8457, 8389, 8323, 8257, 8193, 8129, 8066, 8005,
    7944, 7885, 7826, 7768, 7711, 7654, 7599, 7544, 7490, 7437, 7385, 7333,
    7282, 7232, 7183, 7134, 7085, 7038, 6991, 6945, 6899, 6854, 6809, 6766,
    6722, 6679, 6637, 6595, 6554, 6513, 6473, 6433, 6394, 6356, 6317, 6279,
    6242, 6205, 6169, 6133, 6097, 6062, 6027, 5992, 5958, 5925, 5891, 5858,
    5826, 5794, 5762, 5730, 5699, 5668, 5638, 5608, 5578, 5549, 5519, 5490,
    5462, 5434, 5406, 5378, 5350, 5323, 5296, 5270, 5243, 5217, 5191, 5166,
    5141, 5116, 5091, 5066, 5042, 5018, 4994, 4970, 4947, 4923, 4900, 4878,
    4855, 4833, 4810, 4789, 4767, 4745, 4724, 4703, 4682, 4661, 4640, 4620,
    4600, 4579, 4560, 4540, 4520, 4501, 4482, 4463, 4444, 4425, 4406, 4388,
    4370, 4351, 4333, 4316, 4298, 4280, 4263, 4246, 4229, 4212, 4195, 4178,
    4162, 4145, 4129, 4113};

// 16 samples packed 4 per word (texel 4k+j in byte j of v[k]) -> 8-byte alpha-style block.
//
// stb__CompressAlphaBlock picks, per sample, ind = floor(a / dist) by three compare-and-subtract
// steps (a = 7 v + bias, 0 <= a < 8 dist) and maps it to the block's index order
// {1,7,6,5,4,3,2,0}[ind].  The same integers are produced here with far fewer instructions (the
// kernels have to encode at HBM speed): the quotient is one multiply by the exact 20-bit reciprocal
// recip[dist] = floor(2^20 / dist) + 1 (exact because a * dist < 2^20), the sixteen 3-bit quotients
// are gathered as nibbles of two words, and the index map and the 4-bit -> 3-bit packing are done on
// eight samples at a time with word-wide logic.

// {1,7,6,5,4,3,2,0}[q] on every nibble of x: ((8 - q) & 7) ^ (1 if the result is 0 or 1)
__device__ __forceinline__ uint32_t map_index_nibbles(uint32_t x) {
    uint32_t r = (0x88888888u - x) & 0x77777777u;
    r ^= ~((r >> 1) | (r >> 2)) & 0x11111111u;
    return r;
}
__device__ __forceinline__ uint32_t map_and_pack8(uint32_t x /* eight quotients, one per nibble */) {
    const uint32_t r = map_index_nibbles(x);
    // nibbles -> 3-bit fields: pairs, then quads, then all eight (24 bits)
    uint32_t t = (r & 0x07070707u) | ((r >> 1) & 0x38383838u);
    t = (t & 0x003F003Fu) | ((t >> 2) & 0x0FC00FC0u);
    return (t & 0x00000FFFu) | ((t >> 4) & 0x00FFF000u);
}

__device__ __forceinline__ uint32_t mad_u32(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;  // kept opaque so that the compiler does not re-factor the folded constants back into three multiplies
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// the per-block constants of the quotient (7 s + bias) / dist as one multiply-add: t = s * M7 + biasM, quotient = t >> 20
__device__ __forceinline__ void quotient_consts(int mn, int mx, const uint32_t* recip, uint32_t& M7, uint32_t& biasM) {
    const int dist = mx - mn;
    int bias = dist < 8 ? dist - 1 : dist / 2 + 2;
    bias -= mn * 7;
    const uint32_t M = recip[dist];  // 0 for dist == 0: every quotient is 0, as in stb (a = -1 fails every compare)
    M7 = 7u * M;
    biasM = (uint32_t)(bias * (int)M);
}

__device__ __forceinline__ uint2 alpha_block(const uint32_t v[4], const uint32_t* recip) {
    // the sixteen samples as integers (needed for the multiplies anyway), min / max by 3-input min / max
    // (stb's `else if` update gives the same mn/mx: a sample that lowers mn can never also raise mx past
    // the first sample)
    int sv[16];
#pragma unroll
    for (int i = 0; i < 16; i++) sv[i] = (int)__byte_perm(v[i >> 2], 0u, 0x4440 + (i & 3));
    int mn = sv[0], mx = sv[0];
#pragma unroll
    for (int i = 1; i < 15; i += 2) {
        mn = __vimin3_s32(mn, sv[i], sv[i + 1]);
        mx = __vimax3_s32(mx, sv[i], sv[i + 1]);
    }
    mn = min(mn, sv[15]);
    mx = max(mx, sv[15]);
    uint32_t M7, biasM;
    quotient_consts(mn, mx, recip, M7, biasM);
    uint32_t nib[2] = {0u, 0u};
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const uint32_t t = mad_u32((uint32_t)sv[i], M7, biasM);  // (7 s + bias) * M, below 2^24
        const int j = i & 7, sh = 20 - 4 * j;
        nib[i >> 3] |= (sh >= 0 ? t >> sh : t << -sh) & (7u << (4 * j));
    }
    const uint32_t lo = map_and_pack8(nib[0]), hi = map_and_pack8(nib[1]);  // 24 bits each
    // block = mx, mn, then 48 index bits
    return make_uint2((uint32_t)mx | ((uint32_t)mn << 8) | (lo << 16), (lo >> 16) | (hi << 8));
}

// The same block from sixteen samples given as BASE + v (v = the 8-bit value; BASE = e.g. the exponent bits of a
// float whose low mantissa byte holds v): min / max compare the raw words, and BASE * M7 is folded into the additive
// constant of the quotient's multiply-add (arithmetic modulo 2^32), so the samples are never unpacked.
template <uint32_t BASE>
__device__ __forceinline__ uint2 alpha_block_raw(const uint32_t s[16], const uint32_t* recip) {
    int mn = (int)s[0], mx = (int)s[0];
#pragma unroll
    for (int i = 1; i < 15; i += 2) {
        mn = __vimin3_s32(mn, (int)s[i], (int)s[i + 1]);
        mx = __vimax3_s32(mx, (int)s[i], (int)s[i + 1]);
    }
    mn = min(mn, (int)s[15]) - (int)BASE;
    mx = max(mx, (int)s[15]) - (int)BASE;
    uint32_t M7, biasM;
    quotient_consts(mn, mx, recip, M7, biasM);
    biasM -= BASE * M7;
    uint32_t nib[2] = {0u, 0u};
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const uint32_t t = mad_u32(s[i], M7, biasM);
        const int j = i & 7, sh = 20 - 4 * j;
        nib[i >> 3] |= (sh >= 0 ? t >> sh : t << -sh) & (7u << (4 * j));
    }
    const uint32_t lo = map_and_pack8(nib[0]), hi = map_and_pack8(nib[1]);
    return make_uint2((uint32_t)mx | ((uint32_t)mn << 8) | (lo << 16), (lo >> 16) | (hi << 8));
}

// sixteen RGBA8 texels of a block, row major -> byte-planar R (and G): texel 4k+j in byte j of word k
__device__ __forceinline__ void planes_rg(const uint32_t px[16], uint32_t r4[4], uint32_t g4[4]) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t t0 = __byte_perm(px[4 * k], px[4 * k + 1], 0x5140);
        const uint32_t t1 = __byte_perm(px[4 * k + 2], px[4 * k + 3], 0x5140);
        r4[k] = __byte_perm(t0, t1, 0x5410);
        g4[k] = __byte_perm(t0, t1, 0x7632);
    }
}

// One 4x4 block held by FOUR lanes {l, l^1, l^16, l^17} of a warp, a 2x2 quad of texels each (lane bit 0: right
// half, lane bit 4: lower half).  rg[0..3] = the lane's texels in the order (0,0) (0,1) (1,0) (1,1) as R | G << 16.
// Every lane computes the indices of its own samples; the block's min / max (both channels at once, 16x2 SIMD
// min / max) and the 48 index bits travel by xor shuffles (redux.sync would serialise the eight quads of a
// warp: its result is warp-uniform).  Returns the two words of the 8-byte block of channel C on every lane.
struct QuadMinMax {
    uint32_t mn, mx;  // R | G << 16
};
__device__ __forceinline__ QuadMinMax quad_minmax(const uint32_t rg[4]) {
    uint32_t mn = __vimin3_s16x2(__vimin3_s16x2(rg[0], rg[1], rg[2]), rg[3], rg[3]);
    uint32_t mx = __vimax3_s16x2(__vimax3_s16x2(rg[0], rg[1], rg[2]), rg[3], rg[3]);
    mn = __vimin3_s16x2(mn, mn, __shfl_xor_sync(0xFFFFFFFFu, mn, 1));
    mx = __vimax3_s16x2(mx, mx, __shfl_xor_sync(0xFFFFFFFFu, mx, 1));
    mn = __vimin3_s16x2(mn, mn, __shfl_xor_sync(0xFFFFFFFFu, mn, 16));
    mx = __vimax3_s16x2(mx, mx, __shfl_xor_sync(0xFFFFFFFFu, mx, 16));
    return {mn, mx};
}
template <int C>
__device__ __forceinline__ uint2 alpha_block_quad(const uint32_t rg[4], QuadMinMax mm, const uint32_t* recip, int dx, int dy) {
    const int mn = (int)((mm.mn >> (16 * C)) & 0xFFFFu), mx = (int)((mm.mx >> (16 * C)) & 0xFFFFu);
    uint32_t M7, biasM;
    quotient_consts(mn, mx, recip, M7, biasM);
    uint32_t nib = 0u;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const uint32_t t = mad_u32((rg[i] >> (16 * C)) & 0xFFFFu, M7, biasM);
        const int sh = 20 - 4 * i;
        nib |= (t >> sh) & (7u << (4 * i));
    }
    const uint32_t r = map_index_nibbles(nib) & 0x7777u;
    const uint32_t pairs = (r & 0x0707u) | ((r >> 1) & 0x3838u);            // row 0 of the quad in bits 0..5, row 1 in bits 8..13
    const uint32_t V = ((pairs & 0x3Fu) | ((pairs & 0x3F00u) << 4)) << (6 * dx);  // texel (y, x) of a block half sits at bit 3 (4 y + x)
    const uint32_t W = V | __shfl_xor_sync(0xFFFFFFFFu, V, 1);   // this half of the block (rows 2 dy, 2 dy + 1), 24 bits
    const uint32_t O = __shfl_xor_sync(0xFFFFFFFFu, W, 16);      // the other half
    const uint32_t lo = dy ? O : W, hi = dy ? W : O;
    return make_uint2((uint32_t)mx | ((uint32_t)mn << 8) | (lo << 16), (lo >> 16) | (hi << 8));
}

}  // namespace bc45
}  // namespace rosy_tex
