// bc7_encode.cu - hand-written sm_100a BC7 encoder (kernel K4 of SURVEY.md 2.4).
//
// Replaces nvtt::Context::compress(..., Format_BC7) at /root/reference/src/Packager/Packager.cpp:245
// (colour maps) and :296 (normal maps).  One launch encodes the blocks of every mip level.
//
// Mapping: one 4x4 block per thread; the kernel is instruction-issue bound, so the design goals
// are (a) few instructions per fit and (b) a small instruction footprint (the first version of this
// kernel was 172 KB of SASS against a 32 KB L1.5 instruction cache and stalled on fetch).
//   * The search is a warp-uniform JOB LIST held in constant memory ("fit subset k of partition
//     #i of mode m", "rank the partitions", ...) walked by one loop, so every piece of code - the
//     three `prepare<NC>` and three `refine<IB>` instances, the estimator, the packer - exists once.
//   * Per-thread state lives in shared memory columns [field][thread] (conflict free, no barrier
//     is ever needed because a thread only touches its own column): the 16 texels, the same
//     texels as byte planes (channel c of texels 4k..4k+3 in one word), the cached all-texel
//     RGB fit, the chosen partitions and the best solution so far.  No local memory is used.
//   * All subset statistics are integer dot products of byte planes against byte masks (dp4a):
//     moments for the PCA, and the least-squares sums from per-texel weight bytes that one PRMT
//     looks up from the index nibbles.
//   * index search: exhaustive, 3 integer instructions per (texel, palette entry):
//         dot = dp4a(texel, entry);  key = K[entry] - 32*dot;  best = min(best, key)
//     with K = (|entry|^2 << 4) | entry_index, so one signed min yields both error and index.
// fp32 expression order is part of the specification ("spec v2"): this file is built
// with -fmad=false and IEEE division / sqrt, so its decisions coincide with the scalar statement
// in oracle/bc7.c (which the tests use as the checker; nothing here includes it).
//
// Tensor cores are not used: there is no dense contraction anywhere in the search.
#ifdef BC7_HOST_EMU
// tools/bc7_host_emu: the device functions of this file compiled for the host (one 4x4 block at a
// time, NT = 1) so that every change to the search can be checked against tests/golden/bc7_spec_v2.npz
// on a machine without a GPU.  Test tooling only; the product is the sm_100a build below.
#include "bc7_host_emu.h"
#else
#include <cuda_runtime.h>
#endif
#include <stdint.h>

#include <mutex>
#include <vector>

#include "bc7_tables.cuh"
#include "tex_internal.h"

namespace rosy_tex {
namespace {

// unroll factors of the texel loop in evaluate<IB> (8 texels per index word): measured on B200,
// 2/2/2 3947, 8/4/2 4025, 8/8/2 4113, 8/4/4 4108, 8/8/4 4133, 8/8/8 4014 Mpix/s (less loop and address
// arithmetic until the code outgrows the instruction cache); the projection loop of prepare: 4 beats 8 and 16
#ifndef BC7_U2
#define BC7_U2 8
#endif
#ifndef BC7_U3
#define BC7_U3 8
#endif
#ifndef BC7_U4
#define BC7_U4 4
#endif
#ifndef BC7_UP
#define BC7_UP 4
#endif
#ifndef BC7_MAGIC_I2F
#define BC7_MAGIC_I2F 1
#endif
#ifndef BC7_JOB_SYNC
#define BC7_JOB_SYNC 0
#endif
constexpr int EVAL_U2 = BC7_U2, EVAL_U3 = BC7_U3, EVAL_U4 = BC7_U4, PREP_U = BC7_UP;
constexpr int TOPK = 6;   // most partitions refined per mode (high tier)
constexpr int MAXJOBS = 96;

__constant__ uint8_t c_SOLID5[256][2];       // best 7-bit end-point pair for a solid value (mode 5)
__constant__ uint32_t c_PART2_BYTES[64][4];  // 0xFF per texel of subset 1, texels 4k..4k+3 in word k
__constant__ uint32_t c_PART3_BYTES[64][2][4];
__constant__ uint16_t c_P3M1[64], c_P3M2[64];  // 3-subset partitions as 16-bit masks of subsets 1, 2
__constant__ uint8_t c_ANCHOR2[64], c_ANCHOR3A[64], c_ANCHOR3B[64];
__constant__ uint2 c_JOBS[3][MAXJOBS];       // per quality tier
__constant__ int c_NJOBS[3];
// 1/n rounded to fp32 (the oracle's 1.0f / (float)n), n = texels of a subset
__constant__ float c_RCPN[17] = {0.0f, 1.0f / 1.0f, 1.0f / 2.0f, 1.0f / 3.0f, 1.0f / 4.0f, 1.0f / 5.0f, 1.0f / 6.0f,
                                 1.0f / 7.0f, 1.0f / 8.0f, 1.0f / 9.0f, 1.0f / 10.0f, 1.0f / 11.0f, 1.0f / 12.0f,
                                 1.0f / 13.0f, 1.0f / 14.0f, 1.0f / 15.0f, 1.0f / 16.0f};

// ---- job encoding (built on the host, bc7_upload_tables) ---------------------------------------
enum { JOB_PREP_P3 = 0, JOB_FIT = 1, JOB_EST = 2 };
enum { PB_NONE = 0, PB_SHARED = 1, PB_PER_EP = 2 };
enum { PREP_COMPUTE = 0, PREP_FROM_P3 = 1, PREP_MODE6 = 2 };
enum { AREQ_ANY = 0, AREQ_OPAQUE = 1, AREQ_ALPHA = 2 };
enum { DEST_PART = 0, DEST_COLOUR = 1, DEST_SCALAR = 2 };
// views: 0 RGBA, 1 RGB, 2..4 colour of rotation 1..3, 5..8 scalar of rotation 0..3
//   sel   = __byte_perm selector producing the view's texel (unused bytes zero)
//   pcmap = physical byte plane of logical channel c in nibble c
__constant__ uint32_t c_VIEW_SEL[9] = {0x3210u, 0x4210u, 0x4213u, 0x4230u, 0x4310u,
                                       0x4443u, 0x4440u, 0x4441u, 0x4442u};
//   inv   = __byte_perm selector that moves a palette entry in view order (logical channel k in byte k) to the
//           texel's physical byte order (bytes the view does not use become zero): dot(view(p), e) == dot(p, inv(e)),
//           so the index search reads the stored texels as they are and permutes the few palette entries instead
__constant__ uint32_t c_VIEW_INV[9] = {0x3210u, 0x4210u, 0x0214u, 0x1240u, 0x2410u,
                                       0x0444u, 0x4440u, 0x4404u, 0x4044u};
__constant__ uint32_t c_VIEW_PC[9] = {0x3210u, 0x0210u, 0x0213u, 0x0230u, 0x0310u,
                                      0x0003u, 0x0000u, 0x0001u, 0x0002u};
// w0: type[0:2] mode[2:5] ns[5:7] k[7:9] prow[9:15] view[15:19] nc[19:22] ib[22:25] bits[25:29] pbmode[29:31]
// w1: prep[0:2] first[2] last[3] areq[4:6] dest[6:8] idxsel[8] rot[9:11] estkind[11:13]
struct JobFields {
    int type, mode, ns, k, prow, view, nc, ib, bits, pbmode, prep, first, last, areq, dest, idxsel, rot, estkind;
};
inline uint2 pack_job(const JobFields& j) {
    uint2 w;
    w.x = (uint32_t)j.type | (uint32_t)j.mode << 2 | (uint32_t)j.ns << 5 | (uint32_t)j.k << 7 |
          (uint32_t)j.prow << 9 | (uint32_t)j.view << 15 | (uint32_t)j.nc << 19 | (uint32_t)j.ib << 22 |
          (uint32_t)j.bits << 25 | (uint32_t)j.pbmode << 29;
    w.y = (uint32_t)j.prep | (uint32_t)j.first << 2 | (uint32_t)j.last << 3 | (uint32_t)j.areq << 4 |
          (uint32_t)j.dest << 6 | (uint32_t)j.idxsel << 8 | (uint32_t)j.rot << 9 | (uint32_t)j.estkind << 11;
    return w;
}

struct Params {
    Chain chain;
    int tier;               // 0 fast, 1 normal, 2 high: selects the job list
    int ls_rounds, pb_search;
    const uint8_t* level0;  // level 0 pixels (colour maps: the input image itself)
    const uint8_t* chainbuf;
    uint8_t* out;
    int first_level;
    int num_levels;
    uint64_t first_block;  // chain.lv[first_level].blk_offset
    uint64_t num_blocks;
};

// per-thread state columns.  NT = threads (= 4x4 blocks) per CTA: the kernel exists for several NT
// (launch_bc7 picks the CTA size per launch so that the last wave of CTAs is as full as possible) and
// for NT = 1 on the host (BC7_HOST_EMU).  Device functions take the instance as `SH`.
template <int NT>
struct Shared {
    uint32_t px[16][NT];   // texels, RGBA8
    uint32_t pl[16][NT];   // byte planes: [4*c + k] = channel c of texels 4k..4k+3
    uint32_t p3[12][NT];   // all-texel RGB prepare: lo[4], hi[4] (float bits), S1[4]
    uint32_t best[12][NT]; // q[3][2], pb, idx lo/hi, idx2 lo/hi, meta
    uint8_t parts[5 * TOPK][NT];  // chosen partitions: lists l1, l3, l0, l2, l7
};

struct Fit {
    uint32_t q0, q1;  // end-point fields, one byte per channel
    uint32_t pb;      // bit 0 / bit 1: p-bit of end point 0 / 1
    uint64_t idx;     // 4 bits per texel (only texels of the subset are non-zero)
    uint32_t err;
};

__device__ __forceinline__ int dp4a_uu(uint32_t a, uint32_t b, int c) {
    return (int)__dp4a(a, b, (uint32_t)c);
}

template <int IB>
__device__ __forceinline__ int weight_of(int j) {
    return IB == 2 ? (j * 85 + 2) >> 2 : (IB == 3 ? (j * 73 + 4) >> 3 : (j * 68 + 8) >> 4);
}

__device__ __forceinline__ int expand_bits(int v, int bits) {
    return bits >= 8 ? v : ((v << (8 - bits)) | (v >> (2 * bits - 8)));
}

__device__ __forceinline__ float clampf(float x, float lo, float hi) {
    return x < lo ? lo : (x > hi ? hi : x);
}

// 4 mask bits -> 4 byte masks
__device__ __forceinline__ uint32_t nibble_to_bytes(uint32_t nib) {
    return ((nib * 0x00204081u) & 0x01010101u) * 0xFFu;
}

// ---- prepare: integer moments -> principal axis -> extremes along the axis (spec v2 steps 1-3) ----
struct Prep {
    float lo[4], hi[4];
    int S1[4];
    int n;
    int trace;  // sum over the subset's texels of |texel|^2 (view channels): the constant term of every fit error
};

template <int NC, class SH>
__device__ __forceinline__ void prepare(const SH& S, int tid, uint32_t sel, uint32_t pcmap,
                                        uint32_t mask, Prep& P) {
    const int n = __popc(mask);
    int S1[NC], S2[NC][NC];
#pragma unroll
    for (int a = 0; a < NC; a++) {
        S1[a] = 0;
#pragma unroll
        for (int b = 0; b < NC; b++) S2[a][b] = 0;
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t mk = nibble_to_bytes((mask >> (4 * k)) & 15u);
        uint32_t pw[NC];
#pragma unroll
        for (int a = 0; a < NC; a++) pw[a] = S.pl[4 * ((pcmap >> (4 * a)) & 15u) + k][tid];
#pragma unroll
        for (int a = 0; a < NC; a++) {
            const uint32_t m = pw[a] & mk;
            S1[a] = dp4a_uu(m, 0x01010101u, S1[a]);
#pragma unroll
            for (int b = a; b < NC; b++) S2[a][b] = dp4a_uu(m, pw[b], S2[a][b]);
        }
    }
    float C[NC][NC];
#pragma unroll
    for (int a = 0; a < NC; a++)
#pragma unroll
        for (int b = a; b < NC; b++) C[a][b] = C[b][a] = (float)(n * S2[a][b] - S1[a] * S1[b]);
    // principal axis: 4 power iterations from the covariance column with the largest diagonal
    float cbest = C[0][0], v[NC];
#pragma unroll
    for (int a = 0; a < NC; a++) v[a] = C[a][0];
#pragma unroll
    for (int c = 1; c < NC; c++) {
        if (C[c][c] > cbest) {
            cbest = C[c][c];
#pragma unroll
            for (int a = 0; a < NC; a++) v[a] = C[a][c];
        }
    }
    if (!(cbest > 0.0f)) {
#pragma unroll
        for (int a = 0; a < NC; a++) v[a] = 1.0f;
    }
#pragma unroll 1
    for (int it = 0; it < 4; it++) {
        float t[NC], m = 0.0f;
#pragma unroll
        for (int a = 0; a < NC; a++) {
            float s = 0.0f;
#pragma unroll
            for (int b = 0; b < NC; b++) s = __fmaf_rn(C[a][b], v[b], s);
            t[a] = s;
            m = fmaxf(m, fabsf(s));
        }
        if (m == 0.0f) break;
        // rescale by the power of two that brings the largest component into [1, 2): exact
        const float r = __uint_as_float(0x7F000000u - (__float_as_uint(m) & 0x7F800000u));
#pragma unroll
        for (int a = 0; a < NC; a++) v[a] = t[a] * r;
    }
    float len2 = 0.0f;
#pragma unroll
    for (int c = 0; c < NC; c++) len2 = __fmaf_rn(v[c], v[c], len2);
    const float rl = 1.0f / sqrtf(len2);
#pragma unroll
    for (int c = 0; c < NC; c++) v[c] = v[c] * rl;
    float mean[NC];
    const float rn = 1.0f / (float)n;
#pragma unroll
    for (int c = 0; c < NC; c++) mean[c] = (float)S1[c] * rn;
    // projections of the raw texels; the mean's projection is taken off the two extremes afterwards
    float tmin = 3.0e38f, tmax = -3.0e38f, mv = 0.0f;
#pragma unroll
    for (int c = 0; c < NC; c++) mv = __fmaf_rn(mean[c], v[c], mv);
#if BC7_MAGIC_I2F
    // byte -> float without the conversion unit: one PRMT drops view channel c of the stored texel into the
    // mantissa of 2^23 (selector 4 reads the zero low byte of the constant: channels outside the view are 0),
    // one FADD takes 2^23 off again; exact
    uint32_t csel[NC];
#pragma unroll
    for (int c = 0; c < NC; c++) csel[c] = 0x7440u | ((sel >> (4 * c)) & 7u);
#endif
#pragma unroll(PREP_U)
    for (int i = 0; i < 16; i++) {
#if BC7_MAGIC_I2F
        const uint32_t praw = S.px[i][tid];
        float t = 0.0f;
#pragma unroll
        for (int c = 0; c < NC; c++)
            t = __fmaf_rn(__uint_as_float(__byte_perm(praw, 0x4B000000u, csel[c])) - 8388608.0f, v[c], t);
#else
        const uint32_t p = __byte_perm(S.px[i][tid], 0u, sel);
        float t = 0.0f;
#pragma unroll
        for (int c = 0; c < NC; c++) t = __fmaf_rn((float)((p >> (8 * c)) & 255u), v[c], t);
#endif
        if ((mask >> i) & 1u) {
            tmin = t < tmin ? t : tmin;
            tmax = t > tmax ? t : tmax;
        }
    }
    tmin = tmin - mv;
    tmax = tmax - mv;
#pragma unroll
    for (int c = 0; c < 4; c++) {
        if (c < NC) {
            P.lo[c] = clampf(__fmaf_rn(tmin, v[c], mean[c]), 0.0f, 255.0f);
            P.hi[c] = clampf(__fmaf_rn(tmax, v[c], mean[c]), 0.0f, 255.0f);
            P.S1[c] = S1[c];
        } else {
            P.lo[c] = P.hi[c] = 0.0f;
            P.S1[c] = 0;
        }
    }
    P.n = n;
    int tr = 0;
#pragma unroll
    for (int a = 0; a < NC; a++) tr += S2[a][a];
    P.trace = tr;
}

// ---- refine: quantise, exhaustive indices, least-squares rounds (spec v2 steps 4-6) ----------------

// Exhaustive index search of the subset's texels against the palette of (e0, e1) (end points in view order).
// Returns  sum over the subset of  min_j (|P_j|^2 - 2 p.P_j)  =  integer SSD - Prep::trace  (may be negative):
// the texels' own |p|^2 does not depend on the end points and is added once per fit by the caller.
// The palette entries are moved to the texels' byte order (`inv`), so the loop reads S.px as stored.
template <int IB, class SH>
__device__ __forceinline__ int evaluate(const SH& S, int tid, uint32_t inv, uint32_t mask,
                                        uint32_t e0, uint32_t e1, uint64_t& idx_out) {
    constexpr int L = 1 << IB;
    uint32_t P[L];
    int K[L];
    const uint32_t e0rb = e0 & 0x00FF00FFu, e0ga = (e0 >> 8) & 0x00FF00FFu;
    const uint32_t e1rb = e1 & 0x00FF00FFu, e1ga = (e1 >> 8) & 0x00FF00FFu;
#pragma unroll
    for (int j = 0; j < L; j++) {
        const uint32_t w = (uint32_t)weight_of<IB>(j);
        uint32_t rb = (((64u - w) * e0rb + w * e1rb + 0x00200020u) >> 6) & 0x00FF00FFu;
        uint32_t ga = (((64u - w) * e0ga + w * e1ga + 0x00200020u) >> 6) & 0x00FF00FFu;
        const uint32_t pv = rb | (ga << 8);
        K[j] = (dp4a_uu(pv, pv, 0) << 4) | j;
        P[j] = __byte_perm(pv, 0u, inv);
    }
    // Texels from 15 down to 0, eight per index word: the word is shifted up one nibble per texel
    // (one IMAD), the subset mask sits with the current texel's bit in the sign position.
    int total = 0;
    uint32_t word_hi = 0, word_lo = 0;
    uint32_t m = mask << 16;
#pragma unroll 1
    for (int h = 1; h >= 0; h--) {
        uint32_t acc = 0;
#pragma unroll(IB == 2 ? EVAL_U2 : (IB == 3 ? EVAL_U3 : EVAL_U4))
        for (int i = 8 * h + 7; i >= 8 * h; i--) {
            const uint32_t p = S.px[i][tid];
            int best = 0x7FFFFFFF;
#pragma unroll
            for (int j = 0; j < L; j += 2)  // two entries per 3-input min (VIMNMX3)
                best = __vimin3_s32(best, K[j] - (dp4a_uu(p, P[j], 0) << 5), K[j + 1] - (dp4a_uu(p, P[j + 1], 0) << 5));
            const int in = (int)m >> 31;  // all ones for a texel of the subset
            m <<= 1;
            // 4-bit indices only occur in mode 6, whose single subset is the whole block: no mask there
            const int b = IB == 4 ? best : (best & in);  // 16 * (|P|^2 - 2 p.P) + index; the low nibble survives the sign
            total += b >> 4;
            acc = acc * 16u + ((uint32_t)b & 15u);
        }
        word_hi = h ? acc : word_hi;
        word_lo = h ? word_lo : acc;
    }
    idx_out = ((uint64_t)word_hi << 32) | word_lo;
    return total;
}

// interpolation weights of 4 texels (index nibbles in the low 16 bits) as 4 bytes
template <int IB>
__device__ __forceinline__ uint32_t weight_bytes(uint32_t nib16) {
    if (IB == 2) return __byte_perm(0x402B1500u, 0u, nib16);                 // 0 21 43 64
    if (IB == 3) return __byte_perm(0x1B120900u, 0x40372E25u, nib16);        // 0 9 18 27 | 37 46 55 64
    const uint32_t s = nib16 & 0x7777u;
    const uint32_t wl = __byte_perm(0x0D090400u, 0x1E1A1511u, s);            // 0 4 9 13 | 17 21 26 30
    const uint32_t wh = __byte_perm(0x2F2B2622u, 0x403C3733u, s);            // 34 38 43 47 | 51 55 60 64
    uint32_t h = (nib16 >> 3) & 0x1111u;       // bit 3 of every nibble, at bits 0 4 8 12
    h = (h | (h << 8)) & 0x00110011u;          // at bits 0 4 16 20
    h = (h | (h << 4)) & 0x01010101u;          // at bits 0 8 16 24
    h *= 0xFFu;
    return (wl & ~h) | (wh & h);
}

// floor of a float in [0, 2^23) without the conversion unit: a round-down add of 2^23 leaves the integer in
// the low mantissa bits (FADD.RM + LOP instead of F2I.FLOOR on the quarter-rate pipe)
constexpr float FLOOR_MAGIC = 8388608.0f;
__device__ __forceinline__ int floor_small(float t) {
    return (int)(__float_as_uint(__fadd_rd(t, FLOOR_MAGIC)) & 0x1FFu);  // callers stay below 512
}
// end-point channel x in [0,255] -> field value; `scale` = (2^(bits+1) - 1) / 255 with a p-bit
// (LSB forced to p), (2^bits - 1) / 255 without.  The rounded argument is never negative
// ((y - 1) / 2 + 1/2 >= 0 for y >= 0), so only the upper clamp of the specification can act.
__device__ __forceinline__ int quant_pbit(float x, float scale, int maxq, int p) {
    const float y = x * scale;
    return min(floor_small(__fmaf_rn(y - (float)p, 0.5f, 0.5f)), maxq);  // halving is exact: same value as mul, add
}
__device__ __forceinline__ int quant_plain(float x, float scale, int maxq) {
    return min(floor_small(x * scale + 0.5f), maxq);
}

// One fit after `prepare`: quantise (p-bit by nearest end point, or every combination when
// pb_search), exhaustive indices, then `ls_rounds` least-squares refinements (spec v2 steps 4-6).
// Only `evaluate` and the weight lookup depend on the index width, so they are the only code that
// exists per width.
template <class SH>
__device__ __forceinline__ void refine(const SH& S, int tid, uint32_t inv, uint32_t pcmap,
                                       uint32_t mask, int ib, int nc, int bits, int pbmode,
                                       int ls_rounds, int pb_search, Prep& P, Fit& out) {
    out.err = 0xFFFFFFFFu;
    out.q0 = out.q1 = out.pb = 0;
    out.idx = 0;
    const int maxq = (1 << bits) - 1;
    const float scale = pbmode == PB_NONE ? (float)maxq / 255.0f : (float)((2 << bits) - 1) / 255.0f;
    const uint32_t chmask = nc == 4 ? 0xFFFFFFFFu : (nc == 3 ? 0x00FFFFFFu : 0x000000FFu);
#pragma unroll 1
    for (int round = 0; round <= ls_rounds; round++) {
        if (round > 0) {
            // least squares with the current indices.  Sums over the subset, u = 64 - w:
            //   sum u^2 = 4096 n - 128 sum w + sum w^2,  sum u w = 64 sum w - sum w^2,
            //   X = sum u p = 64 S1 - Y,  Y = sum w p        (all exact integers)
            int sw = 0, sww = 0, Y[4] = {0, 0, 0, 0};
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const uint32_t nib = (uint32_t)(out.idx >> (16 * k)) & 0xFFFFu;
                const uint32_t wq = ib == 2 ? weight_bytes<2>(nib) : (ib == 3 ? weight_bytes<3>(nib) : weight_bytes<4>(nib));
                sw = dp4a_uu(wq, 0x01010101u, sw);
                sww = dp4a_uu(wq, wq, sww);
#pragma unroll
                for (int c = 0; c < 4; c++)
                    if (c < nc) Y[c] = dp4a_uu(wq, S.pl[4 * ((pcmap >> (4 * c)) & 15u) + k][tid], Y[c]);
            }
            const float fA = (float)(4096 * P.n - 128 * sw + sww), fB = (float)(64 * sw - sww), fC = (float)sww;
            const float det = fA * fC - fB * fB;
            if (det == 0.0f) break;
            const float kk = 64.0f / det;
#pragma unroll
            for (int c = 0; c < 4; c++) {
                if (c >= nc) break;  // uniform: channels outside the view stay zero (and are masked out of the end points)
                const float fx = (float)(64 * P.S1[c] - Y[c]), fy = (float)Y[c];
                P.lo[c] = clampf((fC * fx - fB * fy) * kk, 0.0f, 255.0f);
                P.hi[c] = clampf((fA * fy - fB * fx) * kk, 0.0f, 255.0f);
            }
        }
        const uint32_t before = out.err;
        // ---- p-bits ----
        int ncombo = 1;
        uint32_t combo0 = 0;  // bit e = p-bit of end point e for combination 0
        bool decided = false;  // end points already quantised by the p-bit decision below
        uint32_t dq0 = 0, dq1 = 0, de0 = 0, de1 = 0;
        if (pbmode != PB_NONE) {
            if (pb_search) {
                ncombo = pbmode == PB_PER_EP ? 4 : 2;
            } else {
                // squared distance of the float end point to its quantised value, for p = 0 and p = 1:
                // per end point (sums over c), and over both end points in (e, c) order for a shared p-bit.
                // Both quantisations are kept (bytes per channel), the chosen one is used as it is.
                float e0s[2] = {0.0f, 0.0f}, e1s[2] = {0.0f, 0.0f}, s0 = 0.0f, s1 = 0.0f;
                uint32_t qq[2][2] = {{0, 0}, {0, 0}}, ee[2][2] = {{0, 0}, {0, 0}};  // [end point][p]
#pragma unroll
                for (int e = 0; e < 2; e++) {
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        if (c < nc) {
                            const float x = e ? P.hi[c] : P.lo[c];
                            // both quantisations from one product: p = 0 is floor(y/2 + 1/2) (at most maxq + 1),
                            // p = 1 is floor(fma(y - 1, 1/2, 1/2)) == floor(y/2) <= maxq for every y >= 0 (y - 1 is
                            // exact for y >= 1 and the result is 0 either way below that)
                            const float y = x * scale;
                            int q = min(floor_small(__fmaf_rn(y, 0.5f, 0.5f)), maxq);
                            int ex = expand_bits(q << 1, bits + 1);
                            float d = (float)ex - x;
                            const float d0 = d * d;
                            qq[e][0] |= (uint32_t)q << (8 * c);
                            ee[e][0] |= (uint32_t)ex << (8 * c);
                            q = floor_small(y * 0.5f);
                            ex = expand_bits((q << 1) | 1, bits + 1);
                            d = (float)ex - x;
                            const float d1 = d * d;
                            qq[e][1] |= (uint32_t)q << (8 * c);
                            ee[e][1] |= (uint32_t)ex << (8 * c);
                            e0s[e] = e0s[e] + d0;
                            e1s[e] = e1s[e] + d1;
                            s0 = s0 + d0;
                            s1 = s1 + d1;
                        }
                    }
                }
                if (pbmode == PB_PER_EP)
                    combo0 = (e1s[0] < e0s[0] ? 1u : 0u) | (e1s[1] < e0s[1] ? 2u : 0u);
                else
                    combo0 = s1 < s0 ? 3u : 0u;
                decided = true;
                dq0 = (combo0 & 1u) ? qq[0][1] : qq[0][0];
                de0 = (combo0 & 1u) ? ee[0][1] : ee[0][0];
                dq1 = (combo0 & 2u) ? qq[1][1] : qq[1][0];
                de1 = (combo0 & 2u) ? ee[1][1] : ee[1][0];
            }
        }
#pragma unroll 1
        for (int k = 0; k < ncombo; k++) {
            uint32_t pb = combo0;
            if (ncombo == 4) pb = (uint32_t)k;  // (k & 1, k >> 1)
            if (ncombo == 2) pb = k ? 3u : 0u;
            uint32_t q0 = dq0, q1 = dq1, e0 = de0, e1 = de1;
            if (!decided) {
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    if (c >= nc) break;  // uniform
                    int qa, qb, ea, eb;
                    if (pbmode == PB_NONE) {
                        qa = quant_plain(P.lo[c], scale, maxq);
                        qb = quant_plain(P.hi[c], scale, maxq);
                        ea = expand_bits(qa, bits);
                        eb = expand_bits(qb, bits);
                    } else {
                        const int pa = pb & 1, pbb = (pb >> 1) & 1;
                        qa = quant_pbit(P.lo[c], scale, maxq, pa);
                        qb = quant_pbit(P.hi[c], scale, maxq, pbb);
                        ea = expand_bits((qa << 1) | pa, bits + 1);
                        eb = expand_bits((qb << 1) | pbb, bits + 1);
                    }
                    q0 |= (uint32_t)qa << (8 * c);
                    q1 |= (uint32_t)qb << (8 * c);
                    e0 |= (uint32_t)ea << (8 * c);
                    e1 |= (uint32_t)eb << (8 * c);
                }
            }
            q0 &= chmask;
            q1 &= chmask;
            e0 &= chmask;
            e1 &= chmask;
            uint64_t idx;
            int rel;
            if (ib == 2)
                rel = evaluate<2>(S, tid, inv, mask, e0, e1, idx);
            else if (ib == 3)
                rel = evaluate<3>(S, tid, inv, mask, e0, e1, idx);
            else
                rel = evaluate<4>(S, tid, inv, mask, e0, e1, idx);
            const uint32_t err = (uint32_t)(rel + P.trace);
            if (err < out.err) {
                out.err = err;
                out.q0 = q0;
                out.q1 = q1;
                out.pb = pb;
                out.idx = idx;
            }
        }
        if (round > 0 && out.err >= before) break;
    }
}

// ---- partition ranking -----------------------------------------------------------------------

// line-fit statistics of one subset from integer moments: tr/n and lambda_max/n
template <int NC>
__device__ __forceinline__ void subset_line_stats(int n, const int S1[4], const int S2[4][4],
                                                  float& tr_n, float& lam_n) {
    float C[4][4];
#pragma unroll
    for (int a = 0; a < NC; a++)
#pragma unroll
        for (int b = a; b < NC; b++) C[a][b] = C[b][a] = (float)(n * S2[a][b] - S1[a] * S1[b]);
    float tr = 0.0f, cbest = C[0][0], v[4] = {C[0][0], C[1][0], C[2][0], NC > 3 ? C[3][0] : 0.0f};
#pragma unroll
    for (int a = 0; a < NC; a++) {
        tr = tr + C[a][a];
        if (a > 0 && C[a][a] > cbest) {
            cbest = C[a][a];
#pragma unroll
            for (int b = 0; b < NC; b++) v[b] = C[b][a];
        }
    }
    // 1/n comes rounded once from a table (n is a property of the partition: uniform over the warp)
    const float rn = c_RCPN[n];
    tr_n = tr * rn;
    lam_n = 0.0f;
    if (tr == 0.0f) return;
    // one power iteration, no rescaling: |C| < 2^22, |w| < 2^46, |C w| < 2^70, sums < 2^118
    float w[4];
#pragma unroll
    for (int a = 0; a < NC; a++) {
        float s = 0.0f;
#pragma unroll
        for (int b = 0; b < NC; b++) s = __fmaf_rn(C[a][b], v[b], s);
        w[a] = s;
    }
    float wu = 0.0f, ww = 0.0f;
#pragma unroll
    for (int a = 0; a < NC; a++) {
        float s = 0.0f;
#pragma unroll
        for (int b = 0; b < NC; b++) s = __fmaf_rn(C[a][b], w[b], s);
        wu = __fmaf_rn(w[a], s, wu);
        ww = __fmaf_rn(w[a], w[a], ww);
    }
    if (ww == 0.0f) return;
    float lam = wu / ww;
    lam = lam > tr ? tr : lam;
    lam = lam < 0.0f ? 0.0f : lam;
    lam_n = lam * rn;
}

struct TopList {
    uint32_t key[TOPK];  // (score bits & ~63) | partition, ascending
};
__device__ __forceinline__ void top_init(TopList& t) {
#pragma unroll
    for (int k = 0; k < TOPK; k++) t.key[k] = 0xFFFFFFFFu;
}
// KEEP = how many entries of the list are ever read: the first KEEP entries of the sorted insertion do not depend
// on the ones behind them, so those are not maintained (fast / normal tiers refine at most 3 + 2 partitions)
template <int KEEP>
__device__ __forceinline__ void top_insert(TopList& t, float score, int part) {
    score = score < 0.0f ? 0.0f : score;
    uint32_t key = (__float_as_uint(score) & ~63u) | (uint32_t)part;
#pragma unroll
    for (int k = 0; k < KEEP; k++) {
        const uint32_t lo = min(t.key[k], key);
        key = max(t.key[k], key);
        t.key[k] = lo;
    }
}

__device__ __forceinline__ float kfac(int ib) {
    const float l = (float)((1 << ib) - 1);
    return 1.0f - 1.0f / (l * l);
}

// integer moments of the texels selected by the byte masks mk[4]
template <int NC>
__device__ __forceinline__ void masked_moments(const uint32_t pl[4][4], const uint32_t mk[4],
                                               int S1[4], int S2[4][4]) {
#pragma unroll
    for (int a = 0; a < 4; a++) {
        S1[a] = 0;
#pragma unroll
        for (int b = 0; b < 4; b++) S2[a][b] = 0;
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
        uint32_t m[4];
#pragma unroll
        for (int a = 0; a < NC; a++) m[a] = pl[a][k] & mk[k];
#pragma unroll
        for (int a = 0; a < NC; a++) {
            S1[a] = dp4a_uu(m[a], 0x01010101u, S1[a]);
#pragma unroll
            for (int b = a; b < NC; b++) S2[a][b] = dp4a_uu(m[a], pl[b][k], S2[a][b]);
        }
    }
}

// Rank the partitions of a 2- or 3-subset table for two index widths at once; the best TOPK of
// each list go to S.parts rows rowA.. and rowB.. (rowB < 0: second list not wanted).  `small` (uniform over
// the launch): only the first 3 / 2 entries of the lists will be read.
template <int NC, int NS, class SH>
__device__ __forceinline__ void estimate_partitions(SH& S, int tid, int nparts, int ibA, int ibB,
                                                    int rowA, int rowB, bool small) {
    uint32_t pl[4][4];
#pragma unroll
    for (int c = 0; c < 4; c++)
#pragma unroll
        for (int k = 0; k < 4; k++) pl[c][k] = S.pl[4 * c + k][tid];
    const uint32_t ones[4] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
    int T1[4], T2[4][4];
    masked_moments<NC>(pl, ones, T1, T2);
    TopList listA, listB;
    top_init(listA);
    top_init(listB);
    const float kA = kfac(ibA), kB = kfac(ibB);
#pragma unroll 1
    for (int p = 0; p < nparts; p++) {
        int S1[3][4], S2[3][4][4], n[3];
        uint32_t mk[4];
        if (NS == 2) {
#pragma unroll
            for (int k = 0; k < 4; k++) mk[k] = c_PART2_BYTES[p][k];
            masked_moments<NC>(pl, mk, S1[1], S2[1]);
            n[1] = __popc(c_PART2_MASK[p]);
            n[2] = 0;
        } else {
#pragma unroll
            for (int k = 0; k < 4; k++) mk[k] = c_PART3_BYTES[p][0][k];
            masked_moments<NC>(pl, mk, S1[1], S2[1]);
#pragma unroll
            for (int k = 0; k < 4; k++) mk[k] = c_PART3_BYTES[p][1][k];
            masked_moments<NC>(pl, mk, S1[2], S2[2]);
            const uint32_t pm = c_PART3_MASK[p];
            n[1] = __popc(pm & 0x55555555u);
            n[2] = __popc(pm & 0xAAAAAAAAu);
        }
        n[0] = 16 - n[1] - n[2];
#pragma unroll
        for (int a = 0; a < 4; a++) {
            S1[0][a] = T1[a] - S1[1][a] - (NS == 3 ? S1[2][a] : 0);
#pragma unroll
            for (int b = a; b < 4; b++)
                S2[0][a][b] = T2[a][b] - S2[1][a][b] - (NS == 3 ? S2[2][a][b] : 0);
        }
        float A = 0.0f, B = 0.0f;
#pragma unroll
        for (int s = 0; s < NS; s++) {
            float tr_n, lam_n;
            subset_line_stats<NC>(n[s], S1[s], S2[s], tr_n, lam_n);
            A = A + tr_n;
            B = B + lam_n;
        }
        if (small) {
            top_insert<3>(listA, A - kA * B, p);
            top_insert<2>(listB, A - kB * B, p);
        } else {
            top_insert<TOPK>(listA, A - kA * B, p);
            top_insert<TOPK>(listB, A - kB * B, p);
        }
    }
#pragma unroll
    for (int k = 0; k < TOPK; k++) {
        S.parts[rowA + k][tid] = (uint8_t)(listA.key[k] & 63u);
        if (rowB >= 0) S.parts[rowB + k][tid] = (uint8_t)(listB.key[k] & 63u);
    }
}

// ---- bit packing -----------------------------------------------------------------------------

__device__ __forceinline__ void subset_masks(int ns, int partition, uint32_t masks[3]) {
    masks[0] = 0xFFFFu;
    masks[1] = masks[2] = 0u;
    if (ns == 2) {
        masks[1] = c_PART2_MASK[partition];
        masks[0] = 0xFFFFu & ~masks[1];
    } else if (ns == 3) {
        masks[1] = c_P3M1[partition];
        masks[2] = c_P3M2[partition];
        masks[0] = 0xFFFFu & ~(masks[1] | masks[2]);
    }
}

struct Bits128 {
    uint64_t lo, hi;
    int pos;
};
__device__ __forceinline__ void put(Bits128& b, uint32_t v, int n) {
    if (n <= 0) return;
    if (b.pos < 64) {
        b.lo |= (uint64_t)v << b.pos;
        if (b.pos + n > 64) b.hi |= (uint64_t)v >> (64 - b.pos);
    } else {
        b.hi |= (uint64_t)v << (b.pos - 64);
    }
    b.pos += n;
}

// 16-bit texel mask -> one bit per 4-bit lane
__device__ __forceinline__ uint64_t lanes_of(uint32_t mask16) {
    uint64_t x = mask16;
    x = (x | (x << 24)) & 0x000000FF000000FFull;
    x = (x | (x << 12)) & 0x000F000F000F000Full;
    x = (x | (x << 6)) & 0x0303030303030303ull;
    x = (x | (x << 3)) & 0x1111111111111111ull;
    return x;
}

__device__ __forceinline__ void swap_endpoint_bytes(uint32_t& a, uint32_t& b, uint32_t bytemask) {
    const uint32_t na = (a & ~bytemask) | (b & bytemask), nb = (b & ~bytemask) | (a & bytemask);
    a = na;
    b = nb;
}

// the thread's best solution (S.best column) -> 128-bit block
template <class SH>
__device__ __forceinline__ uint4 pack_solution(const SH& S, int tid) {
    uint32_t q[3][2];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        q[k][0] = S.best[2 * k][tid];
        q[k][1] = S.best[2 * k + 1][tid];
    }
    uint32_t pb = S.best[6][tid];
    uint64_t idx = (uint64_t)S.best[7][tid] | ((uint64_t)S.best[8][tid] << 32);
    uint64_t idx2 = (uint64_t)S.best[9][tid] | ((uint64_t)S.best[10][tid] << 32);
    const uint32_t meta = S.best[11][tid];
    const int mode = meta & 7, partition = (meta >> 3) & 63, rotation = (meta >> 9) & 3, idxsel = (meta >> 11) & 1;
    const int ns = (mode == 0 || mode == 2) ? 3 : ((mode == 1 || mode == 3 || mode == 7) ? 2 : 1);
    const int part_bits = mode == 0 ? 4 : (ns > 1 ? 6 : 0);
    const int rb = (mode == 4 || mode == 5) ? 2 : 0, isb = mode == 4 ? 1 : 0;
    // per-mode field widths, one nibble per mode (mode 0 = lowest nibble)
    const int cb = (0x57757564u >> (4 * mode)) & 15, ab = (0x57860000u >> (4 * mode)) & 15;
    const int ib = (0x24222233u >> (4 * mode)) & 15, ib2 = (0x00230000u >> (4 * mode)) & 15;
    const bool epb = (mode == 0 || mode == 3 || mode == 6 || mode == 7), spb = mode == 1;
    int anchor[3] = {0, -1, -1};
    uint32_t masks[3];
    subset_masks(ns, partition, masks);
    if (ns == 2) anchor[1] = c_ANCHOR2[partition];
    if (ns == 3) {
        anchor[1] = c_ANCHOR3A[partition];
        anchor[2] = c_ANCHOR3B[partition];
    }
    // anchor rule: the anchor texel's index must have a clear MSB; swap the end points if not
    const uint64_t maxi = (1u << ib) - 1u;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        if (k < ns && ((idx >> (4 * anchor[k] + ib - 1)) & 1ull)) {
            const uint64_t lanes = lanes_of(masks[k]);
            idx = (idx & ~(lanes * 15ull)) | (lanes * maxi - (idx & (lanes * 15ull)));
            if (ib2 == 0) {  // one index set drives every channel (and the p-bits)
                swap_endpoint_bytes(q[k][0], q[k][1], 0xFFFFFFFFu);
                const uint32_t p0 = (pb >> (2 * k)) & 1u, p1 = (pb >> (2 * k + 1)) & 1u;
                pb = (pb & ~(3u << (2 * k))) | (p1 << (2 * k)) | (p0 << (2 * k + 1));
            } else {  // modes 4/5: the primary set is colour (idxsel 0) or alpha (idxsel 1)
                swap_endpoint_bytes(q[k][0], q[k][1], idxsel ? 0xFF000000u : 0x00FFFFFFu);
            }
        }
    }
    if (ib2 && ((idx2 >> (ib2 - 1)) & 1ull)) {
        const uint64_t maxi2 = (1u << ib2) - 1u;
        idx2 = 0x1111111111111111ull * maxi2 - idx2;
        swap_endpoint_bytes(q[0][0], q[0][1], idxsel ? 0x00FFFFFFu : 0xFF000000u);
    }
    Bits128 w = {0ull, 0ull, 0};
    put(w, 1u << mode, mode + 1);
    put(w, (uint32_t)partition, part_bits);
    put(w, (uint32_t)rotation, rb);
    put(w, (uint32_t)idxsel, isb);
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
        for (int k = 0; k < 3; k++)
            if (k < ns) {
                put(w, (q[k][0] >> (8 * c)) & 255u, cb);
                put(w, (q[k][1] >> (8 * c)) & 255u, cb);
            }
    if (ab) {
#pragma unroll
        for (int k = 0; k < 3; k++)
            if (k < ns) {
                put(w, q[k][0] >> 24, ab);
                put(w, q[k][1] >> 24, ab);
            }
    }
    if (epb) put(w, pb & ((1u << (2 * ns)) - 1u), 2 * ns);
    if (spb) {
        put(w, pb & 1u, 1);
        put(w, (pb >> 2) & 1u, 1);
    }
#pragma unroll 1
    for (int i = 0; i < 16; i++) {
        const int is_anchor = (i == anchor[0]) || (i == anchor[1]) || (i == anchor[2]);
        put(w, (uint32_t)((idx >> (4 * i)) & 15ull), ib - is_anchor);
    }
    if (ib2) {
#pragma unroll 1
        for (int i = 0; i < 16; i++) put(w, (uint32_t)((idx2 >> (4 * i)) & 15ull), ib2 - (i == 0));
    }
    return make_uint4((uint32_t)w.lo, (uint32_t)(w.lo >> 32), (uint32_t)w.hi, (uint32_t)(w.hi >> 32));
}

// ---- kernel ----------------------------------------------------------------------------------

// The search for the 4x4 block whose texels sit in column `tid` of S.px (spec v2) -> the
// 128-bit block.  tier / ls / pbs are uniform over the launch, so the job decode stays uniform.
template <class SH>
__device__ __forceinline__ uint4 encode_texels(SH& S, const int tid, const bool live, const int tier, const int ls,
                                               const int pbs) {
    // every thread only ever touches its own column of S: no barrier needed anywhere
    const uint32_t p0 = S.px[0][tid];
    bool solid = true, has_alpha = false;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t a = S.px[4 * k][tid], bq = S.px[4 * k + 1][tid], c = S.px[4 * k + 2][tid],
                       d = S.px[4 * k + 3][tid];
        solid = solid && a == p0 && bq == p0 && c == p0 && d == p0;
        has_alpha = has_alpha || ((a & bq & c & d) >> 24) != 255u;
        // byte planes: channel c of texels 4k..4k+3
        const uint32_t t0 = __byte_perm(a, bq, 0x5140), t1 = __byte_perm(c, d, 0x5140);
        const uint32_t t2 = __byte_perm(a, bq, 0x7362), t3 = __byte_perm(c, d, 0x7362);
        S.pl[0 + k][tid] = __byte_perm(t0, t1, 0x5410);
        S.pl[4 + k][tid] = __byte_perm(t0, t1, 0x7632);
        S.pl[8 + k][tid] = __byte_perm(t2, t3, 0x5410);
        S.pl[12 + k][tid] = __byte_perm(t2, t3, 0x7632);
    }
    uint32_t best_err = 0xFFFFFFFFu;
    if (solid) {
        // exact (or nearest) single colour: mode 5, rotation 0, colour index 1, alpha index 0
        uint32_t q0 = 0, q1 = 0;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const uint32_t v = (p0 >> (8 * c)) & 255u;
            q0 |= (uint32_t)c_SOLID5[v][0] << (8 * c);
            q1 |= (uint32_t)c_SOLID5[v][1] << (8 * c);
        }
        S.best[0][tid] = q0 | (p0 & 0xFF000000u);
        S.best[1][tid] = q1 | (p0 & 0xFF000000u);
#pragma unroll
        for (int f = 2; f < 11; f++) S.best[f][tid] = 0u;
        S.best[7][tid] = 0x11111111u;
        S.best[8][tid] = 0x11111111u;
        S.best[11][tid] = 5u;
    }
    const bool search = live && !solid;
    // candidate under construction
    uint32_t cq[3][2] = {{0, 0}, {0, 0}, {0, 0}}, cpb = 0, cerr = 0;
    uint64_t cidx = 0, cidx2 = 0;
    const int njobs = c_NJOBS[tier];
#pragma unroll 1
    for (int j = 0; j < njobs; j++) {
#if BC7_JOB_SYNC
        // keep the CTA's warps on the same job: the code of one job fits the instruction caches,
        // the whole search does not (data hazards do not exist: a thread only touches its own column)
        if ((j % BC7_JOB_SYNC) == 0) __syncthreads();
#endif
        const uint2 jw = c_JOBS[tier][j];
        const int type = jw.x & 3, areq = (jw.y >> 4) & 3;
        const bool on = search && (areq == AREQ_ANY || (areq == AREQ_ALPHA) == has_alpha);
        if (!on) continue;
        if (type == JOB_EST) {
            const int kind = (jw.y >> 11) & 3;
            if (kind == 0)
                estimate_partitions<3, 2>(S, tid, 64, 3, 2, 0 * TOPK, 1 * TOPK, tier < 2);
            else if (kind == 3)
                estimate_partitions<4, 2>(S, tid, 64, 2, 2, 4 * TOPK, -1, tier < 2);
            else
                estimate_partitions<3, 3>(S, tid, kind == 1 ? 16 : 64, kind == 1 ? 3 : 2, 2,
                                          kind == 1 ? 2 * TOPK : 3 * TOPK, -1, false);
            continue;
        }
        Prep pr;
        if (type == JOB_PREP_P3) {
            prepare<3>(S, tid, c_VIEW_SEL[1], c_VIEW_PC[1], 0xFFFFu, pr);
#pragma unroll
            for (int c = 0; c < 4; c++) {
                S.p3[c][tid] = __float_as_uint(pr.lo[c]);
                S.p3[4 + c][tid] = __float_as_uint(pr.hi[c]);
                S.p3[8 + c][tid] = (uint32_t)pr.S1[c];
            }
            S.p3[11][tid] = (uint32_t)pr.trace;  // S1[3] of the RGB view is zero by construction: its slot carries the trace
            continue;
        }
        // ---- a fit ----
        const int mode = (jw.x >> 2) & 7, ns = (jw.x >> 5) & 3, k = (jw.x >> 7) & 3, prow = (jw.x >> 9) & 63;
        const int view = (jw.x >> 15) & 15, nc = (jw.x >> 19) & 7, ib = (jw.x >> 22) & 7;
        const int bits = (jw.x >> 25) & 15, pbmode = (jw.x >> 29) & 3;
        const int prep = jw.y & 3, first = (jw.y >> 2) & 1, last = (jw.y >> 3) & 1, dest = (jw.y >> 6) & 3;
        const int idxsel = (jw.y >> 8) & 1, rot = (jw.y >> 9) & 3;
        const uint32_t sel = c_VIEW_SEL[view], pcmap = c_VIEW_PC[view];
        const int partition = prow < 63 ? (int)S.parts[prow][tid] : 0;
        uint32_t mask = 0xFFFFu;
        if (ns == 2) {
            const uint32_t m1 = c_PART2_MASK[partition];
            mask = k ? m1 : (0xFFFFu & ~m1);
        } else if (ns == 3) {
            const uint32_t m1 = c_P3M1[partition], m2 = c_P3M2[partition];
            mask = k == 0 ? (0xFFFFu & ~(m1 | m2)) : (k == 1 ? m1 : m2);
        }
        Fit f;
        // A scalar channel that is constant over the block (alpha of an opaque texture, R of a metallic-roughness map):
        // prepare gives lo = hi = v (zero covariance, unit axis, mean v), the fit is q0 = q1 = Q(v) with every index 0
        // and error 16 (v - E)^2, and the first least-squares round stops on det = 0.  Taken in closed form: same
        // bits, none of the work (all scalar fits are without p-bits).  -1.6 % warp instructions on opaque textures.
        bool konst = false;
        if (nc == 1) {
            const uint32_t c4 = 4u * (pcmap & 15u);
            const uint32_t w0 = S.pl[c4][tid];
            const uint32_t v4 = __byte_perm(w0, 0u, 0x0000);
            konst = w0 == v4 && S.pl[c4 + 1][tid] == v4 && S.pl[c4 + 2][tid] == v4 && S.pl[c4 + 3][tid] == v4;
            if (konst) {
                const int maxq = (1 << bits) - 1;
                const int v = (int)(w0 & 255u);
                const int q = quant_plain((float)v, (float)maxq / 255.0f, maxq);
                const int d = v - expand_bits(q, bits);
                f.q0 = f.q1 = (uint32_t)q;
                f.pb = 0u;
                f.idx = 0ull;
                f.err = (uint32_t)(16 * d * d);
            }
        }
        if (konst) {
            // nothing to prepare
        } else if (prep == PREP_COMPUTE || (prep == PREP_MODE6 && has_alpha)) {
            if (nc == 4)
                prepare<4>(S, tid, sel, pcmap, mask, pr);
            else if (nc == 3)
                prepare<3>(S, tid, sel, pcmap, mask, pr);
            else
                prepare<1>(S, tid, sel, pcmap, mask, pr);
        } else {
#pragma unroll
            for (int c = 0; c < 4; c++) {
                pr.lo[c] = __uint_as_float(S.p3[c][tid]);
                pr.hi[c] = __uint_as_float(S.p3[4 + c][tid]);
                pr.S1[c] = (int)S.p3[8 + c][tid];
            }
            pr.n = 16;
            pr.trace = pr.S1[3];
            pr.S1[3] = 0;
            if (prep == PREP_MODE6) {  // opaque block: the RGBA fit is the RGB fit with alpha = 255
                pr.lo[3] = pr.hi[3] = 255.0f;
                pr.S1[3] = 16 * 255;
                pr.trace += 16 * 255 * 255;
            }
        }
        if (!konst) refine(S, tid, c_VIEW_INV[view], pcmap, mask, ib, nc, bits, pbmode, ls, pbs, pr, f);
        // ---- accumulate into the candidate ----
        if (first) {
#pragma unroll
            for (int s = 0; s < 3; s++) cq[s][0] = cq[s][1] = 0;
            cpb = 0;
            cerr = 0;
            cidx = cidx2 = 0;
        }
        if (dest == DEST_PART) {
            if (k == 0) {
                cq[0][0] = f.q0;
                cq[0][1] = f.q1;
            } else if (k == 1) {
                cq[1][0] = f.q0;
                cq[1][1] = f.q1;
            } else {
                cq[2][0] = f.q0;
                cq[2][1] = f.q1;
            }
            cpb |= f.pb << (2 * k);
            cidx |= f.idx;
            cerr += f.err;
        } else if (dest == DEST_COLOUR) {
            cq[0][0] = f.q0 & 0x00FFFFFFu;
            cq[0][1] = f.q1 & 0x00FFFFFFu;
            cidx = f.idx;
            cerr = f.err;
        } else {
            cq[0][0] |= f.q0 << 24;
            cq[0][1] |= f.q1 << 24;
            cerr += f.err;
            // the primary index set is the 2-bit one: colour when idxsel == 0, the scalar otherwise
            if (idxsel) {
                cidx2 = cidx;
                cidx = f.idx;
            } else {
                cidx2 = f.idx;
            }
        }
        if (last && cerr < best_err) {
            best_err = cerr;
#pragma unroll
            for (int s = 0; s < 3; s++) {
                S.best[2 * s][tid] = cq[s][0];
                S.best[2 * s + 1][tid] = cq[s][1];
            }
            S.best[6][tid] = cpb;
            S.best[7][tid] = (uint32_t)cidx;
            S.best[8][tid] = (uint32_t)(cidx >> 32);
            S.best[9][tid] = (uint32_t)cidx2;
            S.best[10][tid] = (uint32_t)(cidx2 >> 32);
            S.best[11][tid] = (uint32_t)mode | ((uint32_t)partition << 3) | ((uint32_t)rot << 9) | ((uint32_t)idxsel << 11);
        }
    }
    return pack_solution(S, tid);
}

#ifndef BC7_HOST_EMU
// One texture (P1, plist == nullptr) or a batch of textures of one quality tier: plist[t] describes
// texture t, whose blocks are threads base[t] .. base[t+1]-1 of the launch.  A CTA may straddle two
// textures: every thread is independent.
template <int NT>
__global__ void __launch_bounds__(NT, 1)
bc7_encode_kernel(const __grid_constant__ Params P1, const Params* __restrict__ plist,
                  const uint64_t* __restrict__ base_of, int ntex) {
    extern __shared__ __align__(16) uint8_t bc7_smem[];
    Shared<NT>& S = *reinterpret_cast<Shared<NT>*>(bc7_smem);
    const int tid = threadIdx.x;
    uint64_t b = (uint64_t)blockIdx.x * NT + tid;
    const Params* Pp = &P1;
    if (plist) {
        int lo = 0, hi = ntex;  // base_of[lo] <= b < base_of[hi]  (threads past the end use the last texture)
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (b >= base_of[mid]) lo = mid; else hi = mid;
        }
        Pp = plist + lo;
        b -= base_of[lo];
    }
    const Params& P = *Pp;
    const bool live = b < P.num_blocks;
    uint4* const outp = live ? reinterpret_cast<uint4*>(P.out) + b : nullptr;  // the only thing kept of P and b
    // which level does this block belong to?
    const uint64_t gb = P.first_block + (live ? b : 0);
    int level = P.first_level;
    for (int l = P.first_level + 1; l < P.first_level + P.num_levels; l++)
        if (gb >= P.chain.lv[l].blk_offset) level = l;
    const Level lv = P.chain.lv[level];
    const uint32_t lb = (uint32_t)(gb - lv.blk_offset);
    const uint32_t bx = lb % lv.bw, by = lb / lv.bw;
    const uint8_t* base = (level == 0 && P.level0) ? P.level0 : P.chainbuf + 4ull * lv.px_offset;
    // load the 4x4 texels (partial blocks repeat the edge texels)
    if (4 * bx + 3 < lv.w && 4 * by + 3 < lv.h && (lv.w & 3u) == 0 && ((uintptr_t)base & 15u) == 0) {
#pragma unroll
        for (int y = 0; y < 4; y++) {
            const uint4 r = __ldg(reinterpret_cast<const uint4*>(
                base + 4ull * ((uint64_t)(4 * by + y) * lv.w + 4 * bx)));
            S.px[4 * y + 0][tid] = r.x;
            S.px[4 * y + 1][tid] = r.y;
            S.px[4 * y + 2][tid] = r.z;
            S.px[4 * y + 3][tid] = r.w;
        }
    } else {
        for (int y = 0; y < 4; y++)
            for (int x = 0; x < 4; x++) {
                const uint32_t sx = min(4 * bx + x, lv.w - 1), sy = min(4 * by + y, lv.h - 1);
                S.px[4 * y + x][tid] =
                    __ldg(reinterpret_cast<const uint32_t*>(base) + (uint64_t)sy * lv.w + sx);
            }
    }
    // every thread only ever touches its own column of S: no barrier needed anywhere
    const uint4 blk = encode_texels(S, tid, live, P1.tier, P1.ls_rounds, P1.pb_search);
    if (outp) *outp = blk;
}
#endif  // !BC7_HOST_EMU

}  // namespace

// fast / normal / high  ~  nvtt::Quality_Fastest / Normal / Production
static const Bc7Tier kTiers[3] = {
    {1, 0, 0, 0, 1, 1, 0, 1, 0},
    {3, 2, 0, 0, 2, 4, 1, 2, 0},
    {6, 4, 3, 3, 4, 4, 4, 3, 1},
};
Bc7Tier bc7_tier(int quality) { return kTiers[quality < 0 ? 0 : (quality > 2 ? 2 : quality)]; }

namespace {
// the search of one tier as a job list, in the order of the specification (ties keep the earlier candidate)
int build_jobs(const Bc7Tier& t, uint2* out) {
    int n = 0;
    auto fit = [&](int mode, int ns, int k, int prow, int view, int nc, int ib, int bits, int pbmode, int prep,
                   int first, int last, int areq, int dest, int idxsel, int rot) {
        JobFields j = {JOB_FIT, mode, ns, k, prow, view, nc, ib, bits, pbmode, prep, first, last, areq, dest, idxsel, rot, 0};
        out[n++] = pack_job(j);
    };
    auto est = [&](int kind, int areq) {
        JobFields j = {JOB_EST, 0, 0, 0, 63, 0, 0, 0, 0, 0, 0, 0, 0, areq, 0, 0, 0, kind};
        out[n++] = pack_job(j);
    };
    JobFields p3 = {JOB_PREP_P3, 0, 0, 0, 63, 1, 3, 0, 0, 0, 0, 0, 0, AREQ_ANY, 0, 0, 0, 0};
    out[n++] = pack_job(p3);
    fit(6, 1, 0, 63, 0, 4, 4, 7, PB_PER_EP, PREP_MODE6, 1, 1, AREQ_ANY, DEST_PART, 0, 0);
    if (t.n1 | t.n3) {
        est(0, AREQ_OPAQUE);
        for (int i = 0; i < t.n1; i++)
            for (int k = 0; k < 2; k++)
                fit(1, 2, k, 0 * TOPK + i, 1, 3, 3, 6, PB_SHARED, PREP_COMPUTE, k == 0, k == 1, AREQ_OPAQUE, DEST_PART, 0, 0);
        for (int i = 0; i < t.n3; i++)
            for (int k = 0; k < 2; k++)
                fit(3, 2, k, 1 * TOPK + i, 1, 3, 2, 7, PB_PER_EP, PREP_COMPUTE, k == 0, k == 1, AREQ_OPAQUE, DEST_PART, 0, 0);
    }
    if (t.n0 | t.n2) {
        if (t.n0) est(1, AREQ_OPAQUE);
        if (t.n2) est(2, AREQ_OPAQUE);
        for (int i = 0; i < t.n0; i++)
            for (int k = 0; k < 3; k++)
                fit(0, 3, k, 2 * TOPK + i, 1, 3, 3, 4, PB_PER_EP, PREP_COMPUTE, k == 0, k == 2, AREQ_OPAQUE, DEST_PART, 0, 0);
        for (int i = 0; i < t.n2; i++)
            for (int k = 0; k < 3; k++)
                fit(2, 3, k, 3 * TOPK + i, 1, 3, 2, 5, PB_NONE, PREP_COMPUTE, k == 0, k == 2, AREQ_OPAQUE, DEST_PART, 0, 0);
    }
    if (t.n7) {
        est(3, AREQ_ALPHA);
        for (int i = 0; i < t.n7; i++)
            for (int k = 0; k < 2; k++)
                fit(7, 2, k, 4 * TOPK + i, 0, 4, 2, 5, PB_PER_EP, PREP_COMPUTE, k == 0, k == 1, AREQ_ALPHA, DEST_PART, 0, 0);
    }
    for (int r = 0; r < t.rot5; r++) {
        fit(5, 1, 0, 63, r == 0 ? 1 : 1 + r, 3, 2, 7, PB_NONE, r == 0 ? PREP_FROM_P3 : PREP_COMPUTE, 1, 0, AREQ_ANY, DEST_COLOUR, 0, r);
        fit(5, 1, 0, 63, 5 + r, 1, 2, 8, PB_NONE, PREP_COMPUTE, 0, 1, AREQ_ANY, DEST_SCALAR, 0, r);
    }
    for (int r = 0; r < t.rot4; r++)
        for (int s = 0; s < 2; s++) {
            fit(4, 1, 0, 63, r == 0 ? 1 : 1 + r, 3, s ? 3 : 2, 5, PB_NONE, r == 0 ? PREP_FROM_P3 : PREP_COMPUTE, 1, 0, AREQ_ANY, DEST_COLOUR, s, r);
            fit(4, 1, 0, 63, 5 + r, 1, s ? 2 : 3, 6, PB_NONE, PREP_COMPUTE, 0, 1, AREQ_ANY, DEST_SCALAR, s, r);
        }
    return n;
}
}  // namespace

// Derived lookup tables, uploaded once per device.
int bc7_upload_tables() {
    static uint8_t solid5[256][2];
    static uint32_t p2b[64][4], p3b[64][2][4];
    static uint16_t p3m1[64], p3m2[64];
    auto interp = [](int e0, int e1, int w) { return ((64 - w) * e0 + w * e1 + 32) >> 6; };
    auto expand7 = [](int v) { return (v << 1) | (v >> 6); };
    for (int c = 0; c < 256; c++) {
        int best = 1 << 30, bl = 0, bh = 0;
        for (int l = 0; l < 128; l++)
            for (int h = 0; h < 128; h++) {
                const int v = interp(expand7(l), expand7(h), 21);
                const int e = (v - c) * (v - c) * 256 + (l > h ? l - h : h - l);
                if (e < best) {
                    best = e;
                    bl = l;
                    bh = h;
                }
            }
        solid5[c][0] = (uint8_t)bl;
        solid5[c][1] = (uint8_t)bh;
    }
    for (int p = 0; p < 64; p++) {
        p3m1[p] = p3m2[p] = 0;
        for (int k = 0; k < 4; k++) p2b[p][k] = p3b[p][0][k] = p3b[p][1][k] = 0;
        for (int i = 0; i < 16; i++) {
            if ((h_PART2_MASK[p] >> i) & 1) p2b[p][i >> 2] |= 0xFFu << (8 * (i & 3));
            const int s3 = (h_PART3_MASK[p] >> (2 * i)) & 3;
            if (s3 == 1) {
                p3b[p][0][i >> 2] |= 0xFFu << (8 * (i & 3));
                p3m1[p] |= (uint16_t)(1u << i);
            }
            if (s3 == 2) {
                p3b[p][1][i >> 2] |= 0xFFu << (8 * (i & 3));
                p3m2[p] |= (uint16_t)(1u << i);
            }
        }
    }
    static uint2 jobs[3][MAXJOBS];
    int njobs[3];
    for (int t = 0; t < 3; t++) {
        njobs[t] = build_jobs(kTiers[t], jobs[t]);
        if (njobs[t] > MAXJOBS) return -1;
        // the kernel keeps only 3 + 2 list entries below the high tier (estimate_partitions, `small`)
        if (t < 2 && (kTiers[t].n1 > 3 || kTiers[t].n3 > 2 || kTiers[t].n7 > 3)) return -1;
    }
    cudaError_t e = cudaSuccess;
    auto ok = [&](cudaError_t r) {
        if (e == cudaSuccess) e = r;
    };
    ok(cudaMemcpyToSymbol(c_SOLID5, solid5, sizeof solid5));
    ok(cudaMemcpyToSymbol(c_PART2_BYTES, p2b, sizeof p2b));
    ok(cudaMemcpyToSymbol(c_PART3_BYTES, p3b, sizeof p3b));
    ok(cudaMemcpyToSymbol(c_P3M1, p3m1, sizeof p3m1));
    ok(cudaMemcpyToSymbol(c_P3M2, p3m2, sizeof p3m2));
    ok(cudaMemcpyToSymbol(c_ANCHOR2, h_ANCHOR2, sizeof h_ANCHOR2));
    ok(cudaMemcpyToSymbol(c_ANCHOR3A, h_ANCHOR3A, sizeof h_ANCHOR3A));
    ok(cudaMemcpyToSymbol(c_ANCHOR3B, h_ANCHOR3B, sizeof h_ANCHOR3B));
    ok(cudaMemcpyToSymbol(c_JOBS, jobs, sizeof jobs));
    ok(cudaMemcpyToSymbol(c_NJOBS, njobs, sizeof njobs));
    return e == cudaSuccess ? 0 : -1;
}

#ifndef BC7_HOST_EMU
// ---- CTA size per launch ---------------------------------------------------------------------
// One CTA per SM (registers), so a launch runs in waves of SMs x NT blocks and a mostly empty last wave
// costs as much as a full one: one 1024^2 chain is 87 383 blocks = 1.15 waves of 148 x 512 and ran at 52 %
// of the batched rate.  The kernel therefore exists for three CTA sizes and every launch takes the one
// with the smallest modelled time  ceil(blocks / (SMs * NT)) * wave_time(NT):
//   NT 512 (128 registers)  the steady-state size: best time per block
//   NT 640 (96 registers)   one 1024^2 chain fits ONE wave of 148 x 640 (0.424 -> 0.276 ms, measured)
//   NT 256                  launches below one 148 x 256 wave: fewer warps per scheduler, each block finishes
//                           sooner (512^2 chain 0.220 -> 0.151 ms): a short launch waits for latency, not throughput
// wave_time relative to NT 512, measured on B200 (profiles/r2b_ab.log: 4096^2 chain 3.982 ms in 19 waves of 512,
// 4.019 ms in 15 waves of 640; 512^2 chain in one wave of 256).  ROSY_TEX_BC7_NT=<256|512|640> forces a size.
template <int NT>
static cudaError_t launch_nt(uint64_t nblocks, const Params& p1, const Params* d_ps, const uint64_t* d_base, int ntex,
                             cudaStream_t s) {
    static std::once_flag once[64];  // function attributes are per device: set once on each device that launches
    int dev = 0;
    cudaGetDevice(&dev);
    std::call_once(once[dev & 63], [] {
        cudaFuncSetAttribute(bc7_encode_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Shared<NT>));
    });
    const uint64_t grid = (nblocks + NT - 1) / NT;
    bc7_encode_kernel<NT><<<(unsigned)grid, NT, sizeof(Shared<NT>), s>>>(p1, d_ps, d_base, ntex);
    return cudaGetLastError();
}

static int sm_count() {
    static int sms[64];
    int dev = 0;
    cudaGetDevice(&dev);
    int& n = sms[dev & 63];
    if (n == 0 && cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) n = 148;
    return n > 0 ? n : 148;
}

int bc7_pick_cta_size(uint64_t nblocks, int sms) {
    static const int forced = [] {
        const char* e = getenv("ROSY_TEX_BC7_NT");
        const int v = e ? atoi(e) : 0;
        return (v == 256 || v == 512 || v == 640) ? v : 0;
    }();
    if (forced) return forced;
    struct V { int nt; double wave; };
    static const V v[3] = {{512, 1.0}, {640, 1.275}, {256, 0.72}};
    int best = 512;
    double best_t = 0.0;
    for (int i = 0; i < 3; i++) {
        const uint64_t per_wave = (uint64_t)sms * (uint64_t)v[i].nt;
        const uint64_t waves = (nblocks + per_wave - 1) / per_wave;
        if (v[i].nt == 256 && waves > 1) continue;  // only for launches below one small wave
        const double t = (double)waves * v[i].wave;
        if (i == 0 || t < best_t * 0.995) {  // ties (within 0.5 %) keep the earlier entry
            best = v[i].nt;
            best_t = t;
        }
    }
    return best;
}

static cudaError_t launch_any(uint64_t nblocks, const Params& p1, const Params* d_ps, const uint64_t* d_base, int ntex,
                              cudaStream_t s) {
    switch (bc7_pick_cta_size(nblocks, sm_count())) {
        case 640: return launch_nt<640>(nblocks, p1, d_ps, d_base, ntex, s);
        case 256: return launch_nt<256>(nblocks, p1, d_ps, d_base, ntex, s);
        default: return launch_nt<512>(nblocks, p1, d_ps, d_base, ntex, s);
    }
}

static void fill_params(Params& P, const Chain& c, int first_level, int num_levels, int quality,
                        const uint8_t* d_level0, const uint8_t* d_chain, uint8_t* d_blocks);

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }
size_t bc7_batch_desc_bytes(int n) {
    return align256((size_t)n * sizeof(Params)) + align256((size_t)(n + 1) * sizeof(uint64_t));
}

// BC7 blocks of n textures (all levels each): one launch per quality tier present.  d_desc:
// bc7_batch_desc_bytes(n) bytes of device memory that must stay untouched until the launches ran.
int launch_bc7_batch(int n, const Bc7BatchItem* items, void* d_desc, cudaStream_t s) {
    int launches = 0;
    uint8_t* dd = reinterpret_cast<uint8_t*>(d_desc);
    size_t used = 0;
    std::vector<Params> ps;
    std::vector<uint64_t> base;
    for (int tier = 0; tier < 3; tier++) {
        ps.clear();
        base.clear();
        uint64_t total = 0;
        for (int i = 0; i < n; i++) {
            const int q = items[i].quality < 0 ? 0 : (items[i].quality > 2 ? 2 : items[i].quality);
            if (q != tier) continue;
            Params P;
            fill_params(P, items[i].chain, 0, items[i].chain.levels, q, items[i].d_level0, items[i].d_chain, items[i].d_blocks);
            base.push_back(total);
            total += P.num_blocks;
            ps.push_back(P);
        }
        if (ps.empty()) continue;
        base.push_back(total);
        const size_t pb = align256(ps.size() * sizeof(Params)), bb = align256(base.size() * sizeof(uint64_t));
        Params* d_ps = reinterpret_cast<Params*>(dd + used);
        uint64_t* d_base = reinterpret_cast<uint64_t*>(dd + used + pb);
        used += pb + bb;
        if (cudaMemcpyAsync(d_ps, ps.data(), ps.size() * sizeof(Params), cudaMemcpyHostToDevice, s) != cudaSuccess ||
            cudaMemcpyAsync(d_base, base.data(), base.size() * sizeof(uint64_t), cudaMemcpyHostToDevice, s) != cudaSuccess)
            return -1;
        if (launch_any(total, ps[0], d_ps, d_base, (int)ps.size(), s) != cudaSuccess) return -1;
        launches++;
    }
    return launches;
}

static void fill_params(Params& P, const Chain& c, int first_level, int num_levels, int quality,
                        const uint8_t* d_level0, const uint8_t* d_chain, uint8_t* d_blocks) {
    P.chain = c;
    P.tier = quality < 0 ? 0 : (quality > 2 ? 2 : quality);
    P.ls_rounds = kTiers[P.tier].ls_rounds;
    P.pb_search = kTiers[P.tier].pb_search;
    P.level0 = d_level0;
    P.chainbuf = d_chain;
    P.first_level = first_level;
    P.num_levels = num_levels;
    P.first_block = c.lv[first_level].blk_offset;
    const int last = first_level + num_levels - 1;
    P.num_blocks = c.lv[last].blk_offset + (uint64_t)c.lv[last].bw * c.lv[last].bh - P.first_block;
    P.out = d_blocks + 16ull * P.first_block;
}

int launch_bc7(const Chain& c, int first_level, int num_levels, int quality, const uint8_t* d_level0,
               const uint8_t* d_chain, uint8_t* d_blocks, cudaStream_t s) {
    Params P;
    fill_params(P, c, first_level, num_levels, quality, d_level0, d_chain, d_blocks);
    return launch_any(P.num_blocks, P, nullptr, nullptr, 1, s) == cudaSuccess ? 1 : -1;
}

#else   // BC7_HOST_EMU: one block at a time through the same device functions
}  // namespace rosy_tex
extern "C" int bc7_emu_encode_image(const uint8_t* rgba, int w, int h, int quality, uint8_t* blocks) {
    using namespace rosy_tex;
    static bool ready = false;
    if (!ready) {
        if (bc7_upload_tables() != 0) return -1;
        ready = true;
    }
    const Bc7Tier t = bc7_tier(quality);
    const int tier = quality < 0 ? 0 : (quality > 2 ? 2 : quality);
    const int bw = (w + 3) / 4, bh = (h + 3) / 4;
    static Shared<1> S;
    for (int by = 0; by < bh; by++)
        for (int bx = 0; bx < bw; bx++) {
            for (int y = 0; y < 4; y++)
                for (int x = 0; x < 4; x++) {
                    const int sx = 4 * bx + x < w ? 4 * bx + x : w - 1, sy = 4 * by + y < h ? 4 * by + y : h - 1;
                    memcpy(&S.px[4 * y + x][0], rgba + 4 * ((size_t)sy * w + sx), 4);
                }
            const uint4 blk = encode_texels(S, 0, true, tier, t.ls_rounds, t.pb_search);
            memcpy(blocks + 16 * ((size_t)by * bw + bx), &blk, 16);
        }
    return 0;
}
namespace rosy_tex {
#endif  // BC7_HOST_EMU
}  // namespace rosy_tex
