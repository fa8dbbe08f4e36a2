// ab_bc7_mapping.cu - A/B of the two thread mappings for the BC7 fit (VERDICT r1 task 3.iv; north_star:
// "one 4x4 block per thread group, warp-shuffle endpoint PCA").  Development / evidence tool, not product code.
//
// Work per 4x4 block, identical in both variants (results are compared bit for bit):
//   fit A  mode 6: all 16 texels, RGBA 7.7.7.7 + p-bit per end point, 4-bit indices
//   fit B  mode 1: the two subsets of 2-subset partition (block index & 63), RGB 6.6.6 + shared p-bit, 3-bit indices
// each fit = integer moments -> covariance -> 4 power iterations -> extremes along the axis -> quantise
// (nearest p-bit) -> exhaustive index search -> integer SSD   (spec v2 steps 1-5, no least-squares rounds).
//
//   variant T  one block per THREAD: the product kernel's own prepare<NC> / refine from bc7_encode.cu
//   variant G  one block per 16-lane GROUP (one texel per lane, two blocks per warp): moments, extremes, error
//              sum and index word by __shfl_xor_sync butterflies; palette entry j built by lane j and fetched by
//              __shfl_sync in the search loop; the scalar steps (power iteration, quantisation) replicated on
//              every lane of the group
//
// Build + run (B200):  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -fmad=false -prec-div=true \
//                           -prec-sqrt=true -Irosy_b200/csrc -o /tmp/ab_bc7_mapping tools/ab_bc7_mapping.cu && /tmp/ab_bc7_mapping
#include "../rosy_b200/csrc/bc7_encode.cu"

#include <stdio.h>
#include <stdlib.h>

#include <vector>

using namespace rosy_tex;

struct FitOut {
    uint32_t q0, q1, pb, err;
    uint64_t idx;
};

constexpr int NT_T = 512;

// ---------------------------------------------------------------- variant T: one block per thread
__global__ void __launch_bounds__(NT_T, 1) fit_thread_kernel(const uint32_t* __restrict__ texels, FitOut* __restrict__ out, int nblocks) {
    extern __shared__ __align__(16) uint8_t smem[];
    Shared<NT_T>& S = *reinterpret_cast<Shared<NT_T>*>(smem);
    const int tid = threadIdx.x;
    const int b = blockIdx.x * NT_T + tid;
    const int bb = b < nblocks ? b : nblocks - 1;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint4 r = __ldg(reinterpret_cast<const uint4*>(texels) + 4 * (size_t)bb + k);
        S.px[4 * k + 0][tid] = r.x;
        S.px[4 * k + 1][tid] = r.y;
        S.px[4 * k + 2][tid] = r.z;
        S.px[4 * k + 3][tid] = r.w;
        const uint32_t t0 = __byte_perm(r.x, r.y, 0x5140), t1 = __byte_perm(r.z, r.w, 0x5140);
        const uint32_t t2 = __byte_perm(r.x, r.y, 0x7362), t3 = __byte_perm(r.z, r.w, 0x7362);
        S.pl[0 + k][tid] = __byte_perm(t0, t1, 0x5410);
        S.pl[4 + k][tid] = __byte_perm(t0, t1, 0x7632);
        S.pl[8 + k][tid] = __byte_perm(t2, t3, 0x5410);
        S.pl[12 + k][tid] = __byte_perm(t2, t3, 0x7632);
    }
    Prep pr;
    Fit f;
    FitOut o[3];
    prepare<4>(S, tid, c_VIEW_SEL[0], c_VIEW_PC[0], 0xFFFFu, pr);
    refine(S, tid, c_VIEW_INV[0], c_VIEW_PC[0], 0xFFFFu, 4, 4, 7, PB_PER_EP, 0, 0, pr, f);
    o[0] = {f.q0, f.q1, f.pb, f.err, f.idx};
    const uint32_t m1 = c_PART2_MASK[bb & 63];
#pragma unroll 1
    for (int k = 0; k < 2; k++) {
        const uint32_t mask = k ? m1 : (0xFFFFu & ~m1);
        prepare<3>(S, tid, c_VIEW_SEL[1], c_VIEW_PC[1], mask, pr);
        refine(S, tid, c_VIEW_INV[1], c_VIEW_PC[1], mask, 3, 3, 6, PB_SHARED, 0, 0, pr, f);
        o[1 + k] = {f.q0, f.q1, f.pb, f.err, f.idx};
    }
    if (b < nblocks)
        for (int k = 0; k < 3; k++) out[3 * (size_t)b + k] = o[k];
}

// ---------------------------------------------------------------- variant G: one block per 16 lanes
__device__ __forceinline__ int gsum(int v) {  // sum over the 16 lanes of the group
#pragma unroll
    for (int m = 1; m < 16; m <<= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, m);
    return v;
}

// one fit of the group's block over the lanes with `in` set; NC view channels (bytes 0..NC-1 of p; others zero)
template <int NC, int IB>
__device__ __forceinline__ FitOut fit_group(uint32_t p, bool in, int lane16, int bits, int pbmode) {
    constexpr int L = 1 << IB;
    // 1. integer moments: masked contributions, butterfly sums (every lane ends with the totals)
    int ch[NC];
#pragma unroll
    for (int a = 0; a < NC; a++) ch[a] = in ? (int)((p >> (8 * a)) & 255u) : 0;
    const int n = gsum(in ? 1 : 0);
    int S1[NC], S2[NC][NC];
#pragma unroll
    for (int a = 0; a < NC; a++) {
        S1[a] = gsum(ch[a]);
#pragma unroll
        for (int b = a; b < NC; b++) S2[a][b] = gsum(ch[a] * ch[b]);
    }
    // 2. covariance, principal axis (replicated on the 16 lanes)
    float C[NC][NC];
#pragma unroll
    for (int a = 0; a < NC; a++)
#pragma unroll
        for (int b = a; b < NC; b++) C[a][b] = C[b][a] = (float)(n * S2[a][b] - S1[a] * S1[b]);
    float cbest = C[0][0], v[NC];
#pragma unroll
    for (int a = 0; a < NC; a++) v[a] = C[a][0];
#pragma unroll
    for (int c = 1; c < NC; c++)
        if (C[c][c] > cbest) {
            cbest = C[c][c];
#pragma unroll
            for (int a = 0; a < NC; a++) v[a] = C[a][c];
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
        const float r = __uint_as_float(0x7F000000u - (__float_as_uint(m) & 0x7F800000u));
#pragma unroll
        for (int a = 0; a < NC; a++) v[a] = t[a] * r;
    }
    float len2 = 0.0f;
#pragma unroll
    for (int c = 0; c < NC; c++) len2 = __fmaf_rn(v[c], v[c], len2);
    const float rl = 1.0f / sqrtf(len2);
    float mean[NC], mv = 0.0f;
    const float rn = 1.0f / (float)n;
#pragma unroll
    for (int c = 0; c < NC; c++) {
        v[c] = v[c] * rl;
        mean[c] = (float)S1[c] * rn;
    }
#pragma unroll
    for (int c = 0; c < NC; c++) mv = __fmaf_rn(mean[c], v[c], mv);
    // 3. this lane's projection, extremes over the group
    float t = 0.0f;
#pragma unroll
    for (int c = 0; c < NC; c++) t = __fmaf_rn((float)((p >> (8 * c)) & 255u), v[c], t);
    float tmin = in ? t : 3.0e38f, tmax = in ? t : -3.0e38f;
#pragma unroll
    for (int m = 1; m < 16; m <<= 1) {
        tmin = fminf(tmin, __shfl_xor_sync(0xFFFFFFFFu, tmin, m));
        tmax = fmaxf(tmax, __shfl_xor_sync(0xFFFFFFFFu, tmax, m));
    }
    tmin = tmin - mv;
    tmax = tmax - mv;
    // 4. quantise, p-bit by nearest end point (replicated)
    const int maxq = (1 << bits) - 1;
    const float scale = (float)((2 << bits) - 1) / 255.0f;
    float e0s[2] = {0.0f, 0.0f}, e1s[2] = {0.0f, 0.0f}, s0 = 0.0f, s1 = 0.0f;
    uint32_t qq[2][2] = {{0, 0}, {0, 0}}, ee[2][2] = {{0, 0}, {0, 0}};
#pragma unroll
    for (int e = 0; e < 2; e++)
#pragma unroll
        for (int c = 0; c < NC; c++) {
            const float x = clampf(__fmaf_rn(e ? tmax : tmin, v[c], mean[c]), 0.0f, 255.0f);
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
    const uint32_t pb = pbmode == PB_PER_EP ? ((e1s[0] < e0s[0] ? 1u : 0u) | (e1s[1] < e0s[1] ? 2u : 0u)) : (s1 < s0 ? 3u : 0u);
    const uint32_t q0 = (pb & 1u) ? qq[0][1] : qq[0][0], e0 = (pb & 1u) ? ee[0][1] : ee[0][0];
    const uint32_t q1 = (pb & 2u) ? qq[1][1] : qq[1][0], e1 = (pb & 2u) ? ee[1][1] : ee[1][0];
    // 5. palette: lane j builds entry j (L <= 16 lanes busy); search: entries fetched by shuffle
    const int j = lane16 & (L - 1);
    const uint32_t w = (uint32_t)weight_of<IB>(j);
    const uint32_t rb = (((64u - w) * (e0 & 0x00FF00FFu) + w * (e1 & 0x00FF00FFu) + 0x00200020u) >> 6) & 0x00FF00FFu;
    const uint32_t ga = (((64u - w) * ((e0 >> 8) & 0x00FF00FFu) + w * ((e1 >> 8) & 0x00FF00FFu) + 0x00200020u) >> 6) & 0x00FF00FFu;
    const uint32_t Pj = rb | (ga << 8);
    const int Kj = (dp4a_uu(Pj, Pj, 0) << 4) | j;
    int best = 0x7FFFFFFF;
#pragma unroll
    for (int k = 0; k < L; k++) {
        const uint32_t Pk = __shfl_sync(0xFFFFFFFFu, Pj, k, 16);
        const int Kk = __shfl_sync(0xFFFFFFFFu, Kj, k, 16);
        best = min(best, Kk - (dp4a_uu(p, Pk, 0) << 5));
    }
    best += dp4a_uu(p, p, 0) << 4;
    const int b = in ? best : 0;
    const uint32_t err = (uint32_t)gsum(b >> 4);
    // index word: lane i contributes its nibble at bit 4 i
    uint32_t lo = lane16 < 8 ? (uint32_t)(b & 15) << (4 * lane16) : 0u, hi = lane16 >= 8 ? (uint32_t)(b & 15) << (4 * (lane16 - 8)) : 0u;
#pragma unroll
    for (int m = 1; m < 16; m <<= 1) {
        lo |= __shfl_xor_sync(0xFFFFFFFFu, lo, m);
        hi |= __shfl_xor_sync(0xFFFFFFFFu, hi, m);
    }
    return {q0, q1, pb, err, ((uint64_t)hi << 32) | lo};
}

__global__ void __launch_bounds__(512) fit_group_kernel(const uint32_t* __restrict__ texels, FitOut* __restrict__ out, int nblocks) {
    const int gid = (blockIdx.x * 512 + threadIdx.x) >> 4, lane16 = threadIdx.x & 15;
    const int bb = gid < nblocks ? gid : nblocks - 1;
    const uint32_t p = __ldg(texels + 16 * (size_t)bb + lane16);
    const FitOut a = fit_group<4, 4>(p, true, lane16, 7, PB_PER_EP);
    const uint32_t m1 = c_PART2_MASK[bb & 63];
    const bool in1 = (m1 >> lane16) & 1u;
    const uint32_t p3 = p & 0x00FFFFFFu;
    const FitOut b0 = fit_group<3, 3>(p3, !in1, lane16, 6, PB_SHARED);
    const FitOut b1 = fit_group<3, 3>(p3, in1, lane16, 6, PB_SHARED);
    if (gid < nblocks && lane16 == 0) {
        out[3 * (size_t)gid + 0] = a;
        out[3 * (size_t)gid + 1] = b0;
        out[3 * (size_t)gid + 2] = b1;
    }
}

int main() {
    const int nblocks = 1 << 20;
    std::vector<uint32_t> h((size_t)nblocks * 16);
    uint32_t s = 0x524F5359u;
    for (int b = 0; b < nblocks; b++) {  // smooth-ish blocks: a base colour, a gradient and some noise; every 8th has alpha
        uint32_t base[4], grad[4];
        for (int c = 0; c < 4; c++) {
            s = s * 1664525u + 1013904223u;
            base[c] = (s >> 24);
            s = s * 1664525u + 1013904223u;
            grad[c] = (s >> 28);
        }
        for (int i = 0; i < 16; i++) {
            uint32_t px = 0;
            for (int c = 0; c < 4; c++) {
                s = s * 1664525u + 1013904223u;
                int v = (int)base[c] + (int)grad[c] * (i & 3) - (int)grad[(c + 1) & 3] * (i >> 2) + (int)((s >> 29) & 7) - 3;
                v = v < 0 ? 0 : (v > 255 ? 255 : v);
                if (c == 3 && (b & 7) != 7) v = 255;
                px |= (uint32_t)v << (8 * c);
            }
            h[(size_t)b * 16 + i] = px;
        }
    }
    if (bc7_upload_tables() != 0) return 1;
    uint32_t* d_tex;
    FitOut *d_a, *d_b;
    cudaMalloc(&d_tex, h.size() * 4);
    cudaMalloc(&d_a, sizeof(FitOut) * 3 * (size_t)nblocks);
    cudaMalloc(&d_b, sizeof(FitOut) * 3 * (size_t)nblocks);
    cudaMemcpy(d_tex, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(fit_thread_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Shared<NT_T>));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float ms_t = 0, ms_g = 0;
    for (int rep = 0; rep < 4; rep++) {
        cudaEventRecord(e0);
        fit_thread_kernel<<<(nblocks + NT_T - 1) / NT_T, NT_T, sizeof(Shared<NT_T>)>>>(d_tex, d_a, nblocks);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms_t, e0, e1);
        cudaEventRecord(e0);
        fit_group_kernel<<<(nblocks * 16 + 511) / 512, 512>>>(d_tex, d_b, nblocks);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms_g, e0, e1);
    }
    if (cudaGetLastError() != cudaSuccess) {
        printf("cuda error\n");
        return 1;
    }
    std::vector<FitOut> ra(3 * (size_t)nblocks), rb(3 * (size_t)nblocks);
    cudaMemcpy(ra.data(), d_a, sizeof(FitOut) * ra.size(), cudaMemcpyDeviceToHost);
    cudaMemcpy(rb.data(), d_b, sizeof(FitOut) * rb.size(), cudaMemcpyDeviceToHost);
    size_t diff = 0;
    for (size_t i = 0; i < ra.size(); i++)
        if (ra[i].q0 != rb[i].q0 || ra[i].q1 != rb[i].q1 || ra[i].pb != rb[i].pb || ra[i].err != rb[i].err || ra[i].idx != rb[i].idx) diff++;
    printf("{\"tool\": \"ab_bc7_mapping\", \"blocks\": %d, \"fits_per_block\": 3, \"thread_per_block_ms\": %.3f, \"group16_per_block_ms\": %.3f, "
           "\"thread_per_block_Mblocks_s\": %.1f, \"group16_Mblocks_s\": %.1f, \"group16_over_thread_time\": %.2f, \"fits_differing\": %zu}\n",
           nblocks, ms_t, ms_g, nblocks / ms_t / 1e3, nblocks / ms_g / 1e3, ms_g / ms_t, diff);
    return diff == 0 ? 0 : 2;
}
