// bc7_encode.cu - hand-written sm_100a BC7 encoder (kernel K4 of SURVEY.md 2.4).
//
// Replaces nvtt::Context::compress(..., Format_BC7) at /root/reference/src/Packager/Packager.cpp:245
// (colour maps) and :296 (normal maps).  One launch encodes the blocks of every mip level.
//
// Mapping: one 4x4 block per thread.  The 16 texels live in shared memory, transposed
// ([texel][thread], conflict free), so that every search loop is rolled over texels while the
// palette of the candidate end-point pair lives in registers.  All decisions are taken on
// integer sums of squared differences:
//   * index search: exhaustive, 3 integer instructions per (texel, palette entry):
//         dot = dp4a(texel, entry);  key = K[entry] - 32*dot;  best = min(best, key)
//     with K = (|entry|^2 << 4) | entry_index, so one signed min yields both error and index;
//   * subset moments for the PCA: integer, exact; converted to fp32 only for the 4 power iterations;
//   * partition ranking: per-partition integer moments from byte-planar texels with dp4a against
//     0xFF byte masks, then a line-fit residual estimate in fp32.
// fp32 expression order is part of the specification ("spec v1"): this file is built
// with -fmad=false and uses IEEE division / sqrt, so its decisions coincide with the scalar
// statement in oracle/bc7.c (which the tests use as the checker; nothing here includes it).
//
// Tensor cores are not used: there is no dense contraction anywhere in the search.
#include <cuda_runtime.h>
#include <stdint.h>

#include "bc7_tables.cuh"
#include "tex_internal.h"

namespace rosy_tex {
namespace {

constexpr int NT = 128;  // threads (= blocks) per CTA
constexpr int TOPK = 6;  // most partitions refined per mode (high tier)

__constant__ uint8_t c_SOLID5[256][2];       // best 7-bit end-point pair for a solid value (mode 5)
__constant__ uint32_t c_PART2_BYTES[64][4];  // 0xFF per texel of subset 1, texels 4k..4k+3 in word k
__constant__ uint32_t c_PART3_BYTES[64][2][4];
__device__ uint16_t g_P3M1[64], g_P3M2[64];  // 3-subset partitions as 16-bit masks of subsets 1, 2
__device__ uint16_t g_P2M[64];
__device__ uint8_t g_ANCHOR2[64], g_ANCHOR3A[64], g_ANCHOR3B[64];

struct Params {
    Chain chain;
    Bc7Tier tier;
    const uint8_t* level0;  // level 0 pixels (colour maps: the input image itself)
    const uint8_t* chainbuf;
    uint8_t* out;
    int first_level;
    int num_levels;
    uint64_t first_block;  // chain.lv[first_level].blk_offset
    uint64_t num_blocks;
};

enum { PB_NONE = 0, PB_SHARED = 1, PB_PER_EP = 2 };

struct Fit {
    uint32_t q0, q1;  // end-point fields, one byte per channel
    uint32_t pb;      // bit 0 / bit 1: p-bit of end point 0 / 1
    uint64_t idx;     // 4 bits per texel (only texels of the subset are non-zero)
    uint32_t err;
};

struct Solution {
    uint32_t q[3][2];
    uint32_t pb;  // bit 2k+e
    uint64_t idx, idx2;
    uint32_t err;
    int mode, partition, rotation, idxsel;
};

__device__ __forceinline__ int dp4a_uu(uint32_t a, uint32_t b, int c) {
    return (int)__dp4a(a, b, (uint32_t)c);
}

template <int IB>
__device__ __forceinline__ int weight_of(int j) {
    return IB == 2 ? (j * 85 + 2) >> 2 : (IB == 3 ? (j * 73 + 4) >> 3 : (j * 68 + 8) >> 4);
}
__device__ __forceinline__ int weight_rt(int ib, int j) {
    return ib == 2 ? (j * 85 + 2) >> 2 : (ib == 3 ? (j * 73 + 4) >> 3 : (j * 68 + 8) >> 4);
}

__device__ __forceinline__ int expand_bits(int v, int bits) {
    return bits >= 8 ? v : ((v << (8 - bits)) | (v >> (2 * bits - 8)));
}

__device__ __forceinline__ int quant_pbit(float x, int bits, int p) {
    const int tb = bits + 1;
    float y = x * ((float)((1 << tb) - 1) / 255.0f);
    int q = (int)floorf((y - (float)p) * 0.5f + 0.5f);
    return min(max(q, 0), (1 << bits) - 1);
}

__device__ __forceinline__ int quant_plain(float x, int bits) {
    int q = (int)floorf(x * ((float)((1 << bits) - 1) / 255.0f) + 0.5f);
    return min(max(q, 0), (1 << bits) - 1);
}

__device__ __forceinline__ float clampf(float x, float lo, float hi) {
    return x < lo ? lo : (x > hi ? hi : x);
}

// The texel view of a fit: byte permutation (channel rotation / scalar extraction) + byte mask.
struct View {
    uint32_t sel, msk;
};
__device__ __forceinline__ uint32_t fetch(const uint32_t (*spx)[NT], int i, int tid, View v) {
    return __byte_perm(spx[i][tid], 0u, v.sel) & v.msk;
}

// Integer SSD of the subset's texels against the palette of (e0, e1); exhaustive index search.
template <int IB>
__device__ __forceinline__ uint32_t evaluate(const uint32_t (*spx)[NT], int tid, View view,
                                             uint32_t mask, uint32_t e0, uint32_t e1,
                                             uint64_t& idx_out) {
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
        P[j] = rb | (ga << 8);
        K[j] = (dp4a_uu(P[j], P[j], 0) << 4) | j;
    }
    uint32_t total = 0;
    uint64_t idx = 0;
#pragma unroll 2
    for (int i = 0; i < 16; i++) {
        const uint32_t p = fetch(spx, i, tid, view);
        int best = 0x7FFFFFFF;
#pragma unroll
        for (int j = 0; j < L; j++) best = min(best, K[j] - (dp4a_uu(p, P[j], 0) << 5));
        best += dp4a_uu(p, p, 0) << 4;
        const bool in = (mask >> i) & 1u;
        total += in ? (uint32_t)(best >> 4) : 0u;
        idx |= (uint64_t)(in ? (uint32_t)(best & 15) : 0u) << (4 * i);
    }
    idx_out = idx;
    return total;
}

// quantise the float end points, choose p-bits, evaluate, keep the best in `best`
template <int IB>
__device__ __forceinline__ void quantise_and_try(const uint32_t (*spx)[NT], int tid, View view,
                                                 uint32_t mask, int nc, int bits, int pbmode,
                                                 int pb_search, const float lo[4],
                                                 const float hi[4], Fit& best) {
    int ncombo = 1;
    uint32_t combo0 = 0;  // bit e = p-bit of end point e for combination 0
    if (pbmode == PB_PER_EP) {
        if (pb_search) {
            ncombo = 4;
        } else {
#pragma unroll
            for (int e = 0; e < 2; e++) {
                float err0 = 0.0f, err1 = 0.0f;
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    if (c < nc) {
                        const float x = e ? hi[c] : lo[c];
                        int q = quant_pbit(x, bits, 0);
                        float d = (float)expand_bits(q << 1, bits + 1) - x;
                        err0 = err0 + d * d;
                        q = quant_pbit(x, bits, 1);
                        d = (float)expand_bits((q << 1) | 1, bits + 1) - x;
                        err1 = err1 + d * d;
                    }
                }
                combo0 |= (err1 < err0 ? 1u : 0u) << e;
            }
        }
    } else if (pbmode == PB_SHARED) {
        if (pb_search) {
            ncombo = 2;
        } else {
            float err0 = 0.0f, err1 = 0.0f;
#pragma unroll
            for (int e = 0; e < 2; e++) {
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    if (c < nc) {
                        const float x = e ? hi[c] : lo[c];
                        int q = quant_pbit(x, bits, 0);
                        float d = (float)expand_bits(q << 1, bits + 1) - x;
                        err0 = err0 + d * d;
                    }
                }
            }
#pragma unroll
            for (int e = 0; e < 2; e++) {
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    if (c < nc) {
                        const float x = e ? hi[c] : lo[c];
                        int q = quant_pbit(x, bits, 1);
                        float d = (float)expand_bits((q << 1) | 1, bits + 1) - x;
                        err1 = err1 + d * d;
                    }
                }
            }
            combo0 = err1 < err0 ? 3u : 0u;
        }
    }
    const uint32_t chmask = nc == 4 ? 0xFFFFFFFFu : (nc == 3 ? 0x00FFFFFFu : 0x000000FFu);
    for (int k = 0; k < ncombo; k++) {
        uint32_t pb = combo0;
        if (ncombo == 4) pb = (uint32_t)k;           // (k & 1, k >> 1)
        if (ncombo == 2) pb = k ? 3u : 0u;
        uint32_t q0 = 0, q1 = 0, e0 = 0, e1 = 0;
#pragma unroll
        for (int c = 0; c < 4; c++) {
            int qa, qb, ea, eb;
            if (pbmode == PB_NONE) {
                qa = quant_plain(lo[c], bits);
                qb = quant_plain(hi[c], bits);
                ea = expand_bits(qa, bits);
                eb = expand_bits(qb, bits);
            } else {
                const int pa = pb & 1, pbb = (pb >> 1) & 1;
                qa = quant_pbit(lo[c], bits, pa);
                qb = quant_pbit(hi[c], bits, pbb);
                ea = expand_bits((qa << 1) | pa, bits + 1);
                eb = expand_bits((qb << 1) | pbb, bits + 1);
            }
            q0 |= (uint32_t)qa << (8 * c);
            q1 |= (uint32_t)qb << (8 * c);
            e0 |= (uint32_t)ea << (8 * c);
            e1 |= (uint32_t)eb << (8 * c);
        }
        q0 &= chmask;
        q1 &= chmask;
        e0 &= chmask;
        e1 &= chmask;
        uint64_t idx;
        const uint32_t err = evaluate<IB>(spx, tid, view, mask, e0, e1, idx);
        if (err < best.err) {
            best.err = err;
            best.q0 = q0;
            best.q1 = q1;
            best.pb = pb;
            best.idx = idx;
        }
    }
}

// Fit one subset (texels selected by mask) with one end-point pair: spec v1 steps 1-6.
template <int IB>
__device__ __noinline__ void fit_subset(const uint32_t (*spx)[NT], int tid, View view, uint32_t mask,
                                        int nc, int bits, int pbmode, int ls_rounds, int pb_search,
                                        Fit& out) {
    // 1. integer moments of the subset
    const int n = __popc(mask);
    int S1[4] = {0, 0, 0, 0};
    int S2[4][4];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 4; b++) S2[a][b] = 0;
#pragma unroll 2
    for (int i = 0; i < 16; i++) {
        uint32_t p = fetch(spx, i, tid, view);
        p = ((mask >> i) & 1u) ? p : 0u;
        int ch[4] = {(int)(p & 255u), (int)((p >> 8) & 255u), (int)((p >> 16) & 255u), (int)(p >> 24)};
#pragma unroll
        for (int a = 0; a < 4; a++) {
            S1[a] += ch[a];
#pragma unroll
            for (int b = a; b < 4; b++) S2[a][b] += ch[a] * ch[b];
        }
    }
    float C[4][4];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = a; b < 4; b++) C[a][b] = C[b][a] = (float)(n * S2[a][b] - S1[a] * S1[b]);
    // 2. principal axis: 4 power iterations from the covariance column with the largest diagonal
    float cbest = C[0][0], col[4] = {C[0][0], C[1][0], C[2][0], C[3][0]};
#pragma unroll
    for (int c = 1; c < 4; c++) {
        if (C[c][c] > cbest) {
            cbest = C[c][c];
#pragma unroll
            for (int a = 0; a < 4; a++) col[a] = C[a][c];
        }
    }
    float v[4];
#pragma unroll
    for (int c = 0; c < 4; c++) v[c] = c < nc ? (cbest > 0.0f ? col[c] : 1.0f) : 0.0f;
    for (int it = 0; it < 4; it++) {
        float t[4], m = 0.0f;
#pragma unroll
        for (int a = 0; a < 4; a++) {
            float s = 0.0f;
#pragma unroll
            for (int b = 0; b < 4; b++) s = s + C[a][b] * v[b];
            t[a] = s;
            m = fmaxf(m, fabsf(s));
        }
        if (m == 0.0f) break;
        const float r = 1.0f / m;
#pragma unroll
        for (int a = 0; a < 4; a++) v[a] = t[a] * r;
    }
    float len2 = 0.0f;
#pragma unroll
    for (int c = 0; c < 4; c++) len2 = len2 + v[c] * v[c];
    const float rl = 1.0f / sqrtf(len2);
#pragma unroll
    for (int c = 0; c < 4; c++) v[c] = v[c] * rl;
    // 3. project, extremes along the axis
    float mean[4];
    const float fn = (float)n;
#pragma unroll
    for (int c = 0; c < 4; c++) mean[c] = (float)S1[c] / fn;
    float tmin = 3.0e38f, tmax = -3.0e38f;
#pragma unroll 2
    for (int i = 0; i < 16; i++) {
        const uint32_t p = fetch(spx, i, tid, view);
        float t = 0.0f;
        t = t + ((float)(p & 255u) - mean[0]) * v[0];
        t = t + ((float)((p >> 8) & 255u) - mean[1]) * v[1];
        t = t + ((float)((p >> 16) & 255u) - mean[2]) * v[2];
        t = t + ((float)(p >> 24) - mean[3]) * v[3];
        if ((mask >> i) & 1u) {
            tmin = t < tmin ? t : tmin;
            tmax = t > tmax ? t : tmax;
        }
    }
    float lo[4], hi[4];
#pragma unroll
    for (int c = 0; c < 4; c++) {
        lo[c] = clampf(mean[c] + tmin * v[c], 0.0f, 255.0f);
        hi[c] = clampf(mean[c] + tmax * v[c], 0.0f, 255.0f);
    }
    // 4/5/6. quantise + evaluate; then least-squares refinement rounds with the current indices
    out.err = 0xFFFFFFFFu;
    out.q0 = out.q1 = out.pb = 0;
    out.idx = 0;
    for (int round = 0; round <= ls_rounds; round++) {
        if (round > 0) {
            int A = 0, B = 0, Cc = 0, X[4] = {0, 0, 0, 0}, Y[4] = {0, 0, 0, 0};
#pragma unroll 2
            for (int i = 0; i < 16; i++) {
                const uint32_t p = fetch(spx, i, tid, view);
                const bool in = (mask >> i) & 1u;
                const int w = in ? weight_of<IB>((int)((out.idx >> (4 * i)) & 15u)) : 0;
                const int u = in ? 64 - w : 0;
                A += u * u;
                B += u * w;
                Cc += w * w;
                const int ch[4] = {(int)(p & 255u), (int)((p >> 8) & 255u),
                                   (int)((p >> 16) & 255u), (int)(p >> 24)};
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    X[c] += u * ch[c];
                    Y[c] += w * ch[c];
                }
            }
            const float fA = (float)A, fB = (float)B, fC = (float)Cc;
            const float det = fA * fC - fB * fB;
            if (det == 0.0f) break;
            const float k = 64.0f / det;
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const float fx = (float)X[c], fy = (float)Y[c];
                lo[c] = clampf((fC * fx - fB * fy) * k, 0.0f, 255.0f);
                hi[c] = clampf((fA * fy - fB * fx) * k, 0.0f, 255.0f);
            }
        }
        const uint32_t before = out.err;
        quantise_and_try<IB>(spx, tid, view, mask, nc, bits, pbmode, pb_search, lo, hi, out);
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
    const float fn = (float)n;
    tr_n = tr / fn;
    lam_n = 0.0f;
    if (tr == 0.0f) return;
    float w[4], m = 0.0f;
#pragma unroll
    for (int a = 0; a < NC; a++) m = fmaxf(m, fabsf(v[a]));
    float r = 1.0f / m;
#pragma unroll
    for (int a = 0; a < NC; a++) v[a] = v[a] * r;
    m = 0.0f;
#pragma unroll
    for (int a = 0; a < NC; a++) {
        float s = 0.0f;
#pragma unroll
        for (int b = 0; b < NC; b++) s = s + C[a][b] * v[b];
        w[a] = s;
        m = fmaxf(m, fabsf(s));
    }
    if (m == 0.0f) return;
    r = 1.0f / m;
#pragma unroll
    for (int a = 0; a < NC; a++) w[a] = w[a] * r;
    float wu = 0.0f, ww = 0.0f;
#pragma unroll
    for (int a = 0; a < NC; a++) {
        float s = 0.0f;
#pragma unroll
        for (int b = 0; b < NC; b++) s = s + C[a][b] * w[b];
        wu = wu + w[a] * s;
        ww = ww + w[a] * w[a];
    }
    float lam = wu / ww;
    lam = lam > tr ? tr : lam;
    lam = lam < 0.0f ? 0.0f : lam;
    lam_n = lam / fn;
}

struct TopList {
    uint32_t key[TOPK];  // (score bits & ~63) | partition, ascending
};
__device__ __forceinline__ void top_init(TopList& t) {
#pragma unroll
    for (int k = 0; k < TOPK; k++) t.key[k] = 0xFFFFFFFFu;
}
__device__ __forceinline__ void top_insert(TopList& t, float score, int part) {
    score = score < 0.0f ? 0.0f : score;
    uint32_t key = (__float_as_uint(score) & ~63u) | (uint32_t)part;
#pragma unroll
    for (int k = 0; k < TOPK; k++) {
        const uint32_t lo = min(t.key[k], key);
        key = max(t.key[k], key);
        t.key[k] = lo;
    }
}

__device__ __forceinline__ float kfac(int ib) {
    const float l = (float)((1 << ib) - 1);
    return 1.0f - 1.0f / (l * l);
}

// byte-planar copy of the block: plane[c][k] holds channel c of texels 4k..4k+3
__device__ __forceinline__ void make_planes(const uint32_t (*spx)[NT], int tid, uint32_t pl[4][4]) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t p0 = spx[4 * k][tid], p1 = spx[4 * k + 1][tid], p2 = spx[4 * k + 2][tid],
                       p3 = spx[4 * k + 3][tid];
        const uint32_t t0 = __byte_perm(p0, p1, 0x5140), t1 = __byte_perm(p2, p3, 0x5140);
        const uint32_t t2 = __byte_perm(p0, p1, 0x7362), t3 = __byte_perm(p2, p3, 0x7362);
        pl[0][k] = __byte_perm(t0, t1, 0x5410);
        pl[1][k] = __byte_perm(t0, t1, 0x7632);
        pl[2][k] = __byte_perm(t2, t3, 0x5410);
        pl[3][k] = __byte_perm(t2, t3, 0x7632);
    }
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

// Rank the partitions of a 2- or 3-subset table for two index widths at once.
template <int NC, int NS>
__device__ __noinline__ void estimate_partitions(const uint32_t (*spx)[NT], int tid, int nparts,
                                                 int ibA, int ibB, TopList& listA, TopList& listB) {
    uint32_t pl[4][4];
    make_planes(spx, tid, pl);
    const uint32_t ones[4] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
    int T1[4], T2[4][4];
    masked_moments<NC>(pl, ones, T1, T2);
    top_init(listA);
    top_init(listB);
    const float kA = kfac(ibA), kB = kfac(ibB);
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
        top_insert(listA, A - kA * B, p);
        top_insert(listB, A - kB * B, p);
    }
}

// ---- candidates ------------------------------------------------------------------------------

__device__ __forceinline__ void subset_masks(int ns, int partition, uint32_t masks[3]) {
    masks[0] = 0xFFFFu;
    masks[1] = masks[2] = 0u;
    if (ns == 2) {
        masks[1] = g_P2M[partition];
        masks[0] = 0xFFFFu & ~masks[1];
    } else if (ns == 3) {
        masks[1] = g_P3M1[partition];
        masks[2] = g_P3M2[partition];
        masks[0] = 0xFFFFu & ~(masks[1] | masks[2]);
    }
}

template <int IB>
__device__ __noinline__ void try_partitioned(const uint32_t (*spx)[NT], int tid, int mode, int ns,
                                             int nc, int bits, int pbmode, int partition,
                                             int ls_rounds, int pb_search, Solution& best) {
    uint32_t masks[3];
    subset_masks(ns, partition, masks);
    const View view = {0x3210u, nc == 4 ? 0xFFFFFFFFu : 0x00FFFFFFu};
    Solution s;
    s.mode = mode;
    s.partition = partition;
    s.rotation = 0;
    s.idxsel = 0;
    s.err = 0;
    s.pb = 0;
    s.idx = 0;
    s.idx2 = 0;
#pragma unroll
    for (int k = 0; k < 3; k++) s.q[k][0] = s.q[k][1] = 0;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        if (k < ns) {
            Fit f;
            fit_subset<IB>(spx, tid, view, masks[k], nc, bits, pbmode, ls_rounds, pb_search, f);
            s.q[k][0] = f.q0;
            s.q[k][1] = f.q1;
            s.pb |= f.pb << (2 * k);
            s.idx |= f.idx;
            s.err += f.err;
        }
    }
    if (s.err < best.err) best = s;
}

__device__ __noinline__ void try_mode45(const uint32_t (*spx)[NT], int tid, int mode, int rotation,
                                        int idxsel, int ls_rounds, int pb_search, Solution& best) {
    // rotation r swaps alpha with channel r-1.  __byte_perm selectors (nibble i picks result
    // byte i; 4 = zero): colour view = rotated (c0,c1,c2,0), scalar view = rotated alpha in byte 0
    uint32_t colsel, scasel;
    switch (rotation) {
        default: colsel = 0x4210u; scasel = 0x4443u; break;  // (R,G,B) | A
        case 1:  colsel = 0x4213u; scasel = 0x4440u; break;  // (A,G,B) | R
        case 2:  colsel = 0x4230u; scasel = 0x4441u; break;  // (R,A,B) | G
        case 3:  colsel = 0x4310u; scasel = 0x4442u; break;  // (R,G,A) | B
    }
    const View cview = {colsel, 0xFFFFFFFFu}, aview = {scasel, 0xFFFFFFFFu};
    const int cb = mode == 4 ? 5 : 7, ab = mode == 4 ? 6 : 8;
    Fit fc, fa;
    if (mode == 5 || !idxsel) {  // colour 2-bit, scalar 2-bit (mode 5) or 3-bit (mode 4)
        fit_subset<2>(spx, tid, cview, 0xFFFFu, 3, cb, PB_NONE, ls_rounds, pb_search, fc);
        if (mode == 5)
            fit_subset<2>(spx, tid, aview, 0xFFFFu, 1, ab, PB_NONE, ls_rounds, pb_search, fa);
        else
            fit_subset<3>(spx, tid, aview, 0xFFFFu, 1, ab, PB_NONE, ls_rounds, pb_search, fa);
    } else {  // mode 4, index selector 1: colour 3-bit, scalar 2-bit
        fit_subset<3>(spx, tid, cview, 0xFFFFu, 3, cb, PB_NONE, ls_rounds, pb_search, fc);
        fit_subset<2>(spx, tid, aview, 0xFFFFu, 1, ab, PB_NONE, ls_rounds, pb_search, fa);
    }
    const uint32_t err = fc.err + fa.err;
    if (err < best.err) {
        best.mode = mode;
        best.partition = 0;
        best.rotation = rotation;
        best.idxsel = idxsel;
        best.err = err;
        best.pb = 0;
        best.q[0][0] = (fc.q0 & 0x00FFFFFFu) | (fa.q0 << 24);
        best.q[0][1] = (fc.q1 & 0x00FFFFFFu) | (fa.q1 << 24);
        best.q[1][0] = best.q[1][1] = best.q[2][0] = best.q[2][1] = 0;
        // the primary index set is the 2-bit one: colour when idxsel == 0, the scalar otherwise
        best.idx = idxsel ? fa.idx : fc.idx;
        best.idx2 = idxsel ? fc.idx : fa.idx;
    }
}

// ---- bit packing -----------------------------------------------------------------------------

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

__device__ __noinline__ void pack_solution(Solution& s, uint4& out) {
    const int mode = s.mode;
    const int ns = (mode == 0 || mode == 2) ? 3 : ((mode == 1 || mode == 3 || mode == 7) ? 2 : 1);
    const int part_bits = mode == 0 ? 4 : (ns > 1 ? 6 : 0);
    const int rb = (mode == 4 || mode == 5) ? 2 : 0, isb = mode == 4 ? 1 : 0;
    // per-mode field widths, one nibble per mode (mode 0 = lowest nibble)
    const int cb = (0x57757564u >> (4 * mode)) & 15, ab = (0x57860000u >> (4 * mode)) & 15;
    const int ib = (0x24222233u >> (4 * mode)) & 15, ib2 = (0x00230000u >> (4 * mode)) & 15;
    const bool epb = (mode == 0 || mode == 3 || mode == 6 || mode == 7), spb = mode == 1;
    int anchor[3] = {0, -1, -1};
    uint32_t masks[3];
    subset_masks(ns, s.partition, masks);
    if (ns == 2) anchor[1] = g_ANCHOR2[s.partition];
    if (ns == 3) {
        anchor[1] = g_ANCHOR3A[s.partition];
        anchor[2] = g_ANCHOR3B[s.partition];
    }
    // anchor rule: the anchor texel's index must have a clear MSB; swap the end points if not
    const uint64_t maxi = (1u << ib) - 1u;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        if (k < ns && ((s.idx >> (4 * anchor[k] + ib - 1)) & 1ull)) {
            const uint64_t lanes = lanes_of(masks[k]);
            s.idx = (s.idx & ~(lanes * 15ull)) | (lanes * maxi - (s.idx & (lanes * 15ull)));
            if (ib2 == 0) {  // one index set drives every channel (and the p-bits)
                swap_endpoint_bytes(s.q[k][0], s.q[k][1], 0xFFFFFFFFu);
                const uint32_t p0 = (s.pb >> (2 * k)) & 1u, p1 = (s.pb >> (2 * k + 1)) & 1u;
                s.pb = (s.pb & ~(3u << (2 * k))) | (p1 << (2 * k)) | (p0 << (2 * k + 1));
            } else {  // modes 4/5: the primary set is colour (idxsel 0) or alpha (idxsel 1)
                swap_endpoint_bytes(s.q[k][0], s.q[k][1], s.idxsel ? 0xFF000000u : 0x00FFFFFFu);
            }
        }
    }
    if (ib2 && ((s.idx2 >> (ib2 - 1)) & 1ull)) {
        const uint64_t maxi2 = (1u << ib2) - 1u;
        s.idx2 = 0x1111111111111111ull * maxi2 - s.idx2;
        swap_endpoint_bytes(s.q[0][0], s.q[0][1], s.idxsel ? 0x00FFFFFFu : 0xFF000000u);
    }
    Bits128 w = {0ull, 0ull, 0};
    put(w, 1u << mode, mode + 1);
    put(w, (uint32_t)s.partition, part_bits);
    put(w, (uint32_t)s.rotation, rb);
    put(w, (uint32_t)s.idxsel, isb);
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
        for (int k = 0; k < 3; k++)
            if (k < ns) {
                put(w, (s.q[k][0] >> (8 * c)) & 255u, cb);
                put(w, (s.q[k][1] >> (8 * c)) & 255u, cb);
            }
    if (ab) {
#pragma unroll
        for (int k = 0; k < 3; k++)
            if (k < ns) {
                put(w, s.q[k][0] >> 24, ab);
                put(w, s.q[k][1] >> 24, ab);
            }
    }
    if (epb) put(w, s.pb & ((1u << (2 * ns)) - 1u), 2 * ns);
    if (spb) {
        put(w, s.pb & 1u, 1);
        put(w, (s.pb >> 2) & 1u, 1);
    }
    for (int i = 0; i < 16; i++) {
        const int is_anchor = (i == anchor[0]) || (i == anchor[1]) || (i == anchor[2]);
        put(w, (uint32_t)((s.idx >> (4 * i)) & 15ull), ib - is_anchor);
    }
    if (ib2)
        for (int i = 0; i < 16; i++) put(w, (uint32_t)((s.idx2 >> (4 * i)) & 15ull), ib2 - (i == 0));
    out = make_uint4((uint32_t)w.lo, (uint32_t)(w.lo >> 32), (uint32_t)w.hi, (uint32_t)(w.hi >> 32));
}

__device__ __forceinline__ int top_pop(TopList& t) {
    const int part = (int)(t.key[0] & 63u);
#pragma unroll
    for (int k = 0; k + 1 < TOPK; k++) t.key[k] = t.key[k + 1];
    return part;
}

// ---- kernel ----------------------------------------------------------------------------------

__global__ void __launch_bounds__(NT) bc7_encode_kernel(const __grid_constant__ Params P) {
    __shared__ uint32_t spx[16][NT];
    const int tid = threadIdx.x;
    const uint64_t b = (uint64_t)blockIdx.x * NT + tid;
    const bool live = b < P.num_blocks;
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
    if (4 * bx + 3 < lv.w && 4 * by + 3 < lv.h && (lv.w & 3u) == 0) {
#pragma unroll
        for (int y = 0; y < 4; y++) {
            const uint4 r = __ldg(reinterpret_cast<const uint4*>(
                base + 4ull * ((uint64_t)(4 * by + y) * lv.w + 4 * bx)));
            spx[4 * y + 0][tid] = r.x;
            spx[4 * y + 1][tid] = r.y;
            spx[4 * y + 2][tid] = r.z;
            spx[4 * y + 3][tid] = r.w;
        }
    } else {
        for (int y = 0; y < 4; y++)
            for (int x = 0; x < 4; x++) {
                const uint32_t sx = min(4 * bx + x, lv.w - 1), sy = min(4 * by + y, lv.h - 1);
                spx[4 * y + x][tid] =
                    __ldg(reinterpret_cast<const uint32_t*>(base) + (uint64_t)sy * lv.w + sx);
            }
    }
    // every thread only ever reads its own column of spx: no barrier needed
    const uint32_t p0 = spx[0][tid];
    bool solid = true, has_alpha = false;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const uint32_t p = spx[i][tid];
        solid = solid && (p == p0);
        has_alpha = has_alpha || ((p >> 24) != 255u);
    }
    const int n1 = P.tier.n1, n3 = P.tier.n3, n0 = P.tier.n0, n2 = P.tier.n2, n7 = P.tier.n7;
    const int ls = P.tier.ls_rounds, pbs = P.tier.pb_search;
    Solution best;
    best.err = 0xFFFFFFFFu;
    best.mode = 6;
    best.partition = best.rotation = best.idxsel = 0;
    best.pb = 0;
    best.idx = best.idx2 = 0;
#pragma unroll
    for (int k = 0; k < 3; k++) best.q[k][0] = best.q[k][1] = 0;
    if (solid) {
        // exact (or nearest) single colour: mode 5, rotation 0, colour index 1, alpha index 0
        best.mode = 5;
        uint32_t q0 = 0, q1 = 0;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const uint32_t v = (p0 >> (8 * c)) & 255u;
            q0 |= (uint32_t)c_SOLID5[v][0] << (8 * c);
            q1 |= (uint32_t)c_SOLID5[v][1] << (8 * c);
        }
        best.q[0][0] = q0 | (p0 & 0xFF000000u);
        best.q[0][1] = q1 | (p0 & 0xFF000000u);
        best.idx = 0x1111111111111111ull;
    } else {
        // mode 6: one subset, RGBA 7.7.7.7 + p-bit per end point, 4-bit indices
        try_partitioned<4>(spx, tid, 6, 1, 4, 7, PB_PER_EP, 0, ls, pbs, best);
        if (!has_alpha) {
            if (n1 | n3) {
                TopList l1, l3;
                estimate_partitions<3, 2>(spx, tid, 64, 3, 2, l1, l3);
                for (int i = 0; i < n1; i++)
                    try_partitioned<3>(spx, tid, 1, 2, 3, 6, PB_SHARED, top_pop(l1), ls, pbs, best);
                for (int i = 0; i < n3; i++)
                    try_partitioned<2>(spx, tid, 3, 2, 3, 7, PB_PER_EP, top_pop(l3), ls, pbs, best);
            }
            if (n0 | n2) {
                TopList l0, l2, dummy;
                estimate_partitions<3, 3>(spx, tid, 16, 3, 2, l0, dummy);
                estimate_partitions<3, 3>(spx, tid, 64, 2, 2, l2, dummy);
                for (int i = 0; i < n0; i++)
                    try_partitioned<3>(spx, tid, 0, 3, 3, 4, PB_PER_EP, top_pop(l0), ls, pbs, best);
                for (int i = 0; i < n2; i++)
                    try_partitioned<2>(spx, tid, 2, 3, 3, 5, PB_NONE, top_pop(l2), ls, pbs, best);
            }
        } else if (n7) {
            TopList l7, dummy;
            estimate_partitions<4, 2>(spx, tid, 64, 2, 2, l7, dummy);
            for (int i = 0; i < n7; i++)
                try_partitioned<2>(spx, tid, 7, 2, 4, 5, PB_PER_EP, top_pop(l7), ls, pbs, best);
        }
        for (int r = 0; r < P.tier.rot5; r++) try_mode45(spx, tid, 5, r, 0, ls, pbs, best);
        for (int r = 0; r < P.tier.rot4; r++) {
            try_mode45(spx, tid, 4, r, 0, ls, pbs, best);
            try_mode45(spx, tid, 4, r, 1, ls, pbs, best);
        }
    }
    uint4 outv;
    pack_solution(best, outv);
    if (live) reinterpret_cast<uint4*>(P.out)[b] = outv;
}

}  // namespace

// fast / normal / high  ~  nvtt::Quality_Fastest / Normal / Production
Bc7Tier bc7_tier(int quality) {
    static const Bc7Tier tiers[3] = {
        {1, 0, 0, 0, 1, 1, 0, 1, 0},
        {3, 2, 0, 0, 2, 4, 1, 2, 0},
        {6, 4, 3, 3, 4, 4, 4, 3, 1},
    };
    return tiers[quality < 0 ? 0 : (quality > 2 ? 2 : quality)];
}

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
    cudaError_t e = cudaSuccess;
    auto ok = [&](cudaError_t r) {
        if (e == cudaSuccess) e = r;
    };
    ok(cudaMemcpyToSymbol(c_SOLID5, solid5, sizeof solid5));
    ok(cudaMemcpyToSymbol(c_PART2_BYTES, p2b, sizeof p2b));
    ok(cudaMemcpyToSymbol(c_PART3_BYTES, p3b, sizeof p3b));
    ok(cudaMemcpyToSymbol(g_P3M1, p3m1, sizeof p3m1));
    ok(cudaMemcpyToSymbol(g_P3M2, p3m2, sizeof p3m2));
    ok(cudaMemcpyToSymbol(g_P2M, h_PART2_MASK, sizeof h_PART2_MASK));
    ok(cudaMemcpyToSymbol(g_ANCHOR2, h_ANCHOR2, sizeof h_ANCHOR2));
    ok(cudaMemcpyToSymbol(g_ANCHOR3A, h_ANCHOR3A, sizeof h_ANCHOR3A));
    ok(cudaMemcpyToSymbol(g_ANCHOR3B, h_ANCHOR3B, sizeof h_ANCHOR3B));
    return e == cudaSuccess ? 0 : -1;
}

int launch_bc7(const Chain& c, int first_level, int num_levels, const Bc7Tier& t,
               const uint8_t* d_level0, const uint8_t* d_chain, uint8_t* d_blocks, cudaStream_t s) {
    Params P;
    P.chain = c;
    P.tier = t;
    P.level0 = d_level0;
    P.chainbuf = d_chain;
    P.first_level = first_level;
    P.num_levels = num_levels;
    P.first_block = c.lv[first_level].blk_offset;
    const int last = first_level + num_levels - 1;
    P.num_blocks = c.lv[last].blk_offset + (uint64_t)c.lv[last].bw * c.lv[last].bh - P.first_block;
    P.out = d_blocks + 16ull * P.first_block;
    const uint64_t grid = (P.num_blocks + NT - 1) / NT;
    bc7_encode_kernel<<<(unsigned)grid, NT, 0, s>>>(P);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

}  // namespace rosy_tex
