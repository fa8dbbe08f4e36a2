// bc45_encode.cu - hand-written sm_100a BC5 / BC4 encoders (kernels K5 / K6 of SURVEY.md 2.4).
//
// north_star extension (the reference itself only emits BC7): BC5 for normal maps (R,G), BC4 for
// single-channel maps (R).  The block algorithm is the one of the reference's vendored
// /root/reference/src/libs/stb/stb_dxt.h:547-597 (stb__CompressAlphaBlock; entry points :617-626):
// end points = (max, min), then the optimal 3-bit index on the 8-value ramp per sample.
// Integer only, bit exact against that code.
//
// HBM bound (read 64 B, write 16 B / 8 B per block): one block per thread, four 16-byte row
// loads per thread (a warp reads 4 x 512 contiguous bytes), one 16-byte or 8-byte store per
// thread (a warp writes 512 / 256 contiguous bytes).  One launch covers every mip level.
#include <cuda_runtime.h>
#include <stdint.h>

#include "tex_internal.h"

namespace rosy_tex {
namespace {

constexpr int NT = 256;

struct Params {
    Chain chain;
    const uint8_t* level0;
    const uint8_t* chainbuf;
    uint8_t* out;
    int first_level, num_levels;
    uint64_t first_block, num_blocks;
};

// 16 samples packed 4 per word (texel 4k+j in byte j of v[k]) -> 8-byte alpha-style block
__device__ __forceinline__ uint2 alpha_block(const uint32_t v[4]) {
    // min / max with packed byte SIMD (stb's `else if` update gives the same mn/mx: a sample
    // that lowers mn can never also raise mx past the first sample)
    uint32_t mn4 = __vminu4(__vminu4(v[0], v[1]), __vminu4(v[2], v[3]));
    uint32_t mx4 = __vmaxu4(__vmaxu4(v[0], v[1]), __vmaxu4(v[2], v[3]));
    mn4 = __vminu4(mn4, mn4 >> 16);
    mx4 = __vmaxu4(mx4, mx4 >> 16);
    const int mn = (int)(__vminu4(mn4, mn4 >> 8) & 255u);
    const int mx = (int)(__vmaxu4(mx4, mx4 >> 8) & 255u);
    const int dist = mx - mn, dist4 = dist * 4, dist2 = dist * 2;
    int bias = dist < 8 ? dist - 1 : dist / 2 + 2;
    bias -= mn * 7;
    uint64_t bits = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        int a = (int)((v[i >> 2] >> (8 * (i & 3))) & 255u) * 7 + bias;
        int ind = 0;
        if (a >= dist4) {
            ind = 4;
            a -= dist4;
        }
        if (a >= dist2) {
            ind += 2;
            a -= dist2;
        }
        if (a >= dist) ind += 1;
        ind = (-ind) & 7;
        ind ^= (2 > ind);
        bits |= (uint64_t)ind << (3 * i);
    }
    bits = (bits << 16) | ((uint64_t)mn << 8) | (uint64_t)mx;
    return make_uint2((uint32_t)bits, (uint32_t)(bits >> 32));
}

template <bool BC5>
__global__ void __launch_bounds__(NT) bc45_encode_kernel(const __grid_constant__ Params P) {
    const uint64_t b = (uint64_t)blockIdx.x * NT + threadIdx.x;
    if (b >= P.num_blocks) return;
    const uint64_t gb = P.first_block + b;
    int level = P.first_level;
    for (int l = P.first_level + 1; l < P.first_level + P.num_levels; l++)
        if (gb >= P.chain.lv[l].blk_offset) level = l;
    const Level lv = P.chain.lv[level];
    const uint32_t lb = (uint32_t)(gb - lv.blk_offset);
    const uint32_t bx = lb % lv.bw, by = lb / lv.bw;
    const uint8_t* base = (level == 0 && P.level0) ? P.level0 : P.chainbuf + 4ull * lv.px_offset;
    uint32_t px[16];
    if (4 * bx + 3 < lv.w && 4 * by + 3 < lv.h && (lv.w & 3u) == 0 && ((uintptr_t)base & 15u) == 0) {
#pragma unroll
        for (int y = 0; y < 4; y++) {
            const uint4 r = __ldg(reinterpret_cast<const uint4*>(
                base + 4ull * ((uint64_t)(4 * by + y) * lv.w + 4 * bx)));
            px[4 * y + 0] = r.x;
            px[4 * y + 1] = r.y;
            px[4 * y + 2] = r.z;
            px[4 * y + 3] = r.w;
        }
    } else {
#pragma unroll
        for (int y = 0; y < 4; y++)
#pragma unroll
            for (int x = 0; x < 4; x++) {
                const uint32_t sx = min(4 * bx + x, lv.w - 1), sy = min(4 * by + y, lv.h - 1);
                px[4 * y + x] =
                    __ldg(reinterpret_cast<const uint32_t*>(base) + (uint64_t)sy * lv.w + sx);
            }
    }
    // byte-planar R (and G): texel 4k+j -> byte j of word k
    uint32_t r4[4], g4[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t t0 = __byte_perm(px[4 * k], px[4 * k + 1], 0x5140);
        const uint32_t t1 = __byte_perm(px[4 * k + 2], px[4 * k + 3], 0x5140);
        r4[k] = __byte_perm(t0, t1, 0x5410);
        g4[k] = __byte_perm(t0, t1, 0x7632);
    }
    const uint2 rb = alpha_block(r4);
    if (BC5) {
        const uint2 gbk = alpha_block(g4);
        reinterpret_cast<uint4*>(P.out)[b] = make_uint4(rb.x, rb.y, gbk.x, gbk.y);
    } else {
        reinterpret_cast<uint2*>(P.out)[b] = rb;
    }
}

}  // namespace

int launch_bc45(const Chain& c, int first_level, int num_levels, bool bc5, const uint8_t* d_level0,
                const uint8_t* d_chain, uint8_t* d_blocks, cudaStream_t s) {
    Params P;
    P.chain = c;
    P.level0 = d_level0;
    P.chainbuf = d_chain;
    P.first_level = first_level;
    P.num_levels = num_levels;
    P.first_block = c.lv[first_level].blk_offset;
    const int last = first_level + num_levels - 1;
    P.num_blocks = c.lv[last].blk_offset + (uint64_t)c.lv[last].bw * c.lv[last].bh - P.first_block;
    P.out = d_blocks + (bc5 ? 16ull : 8ull) * P.first_block;
    const unsigned grid = (unsigned)((P.num_blocks + NT - 1) / NT);
    if (bc5)
        bc45_encode_kernel<true><<<grid, NT, 0, s>>>(P);
    else
        bc45_encode_kernel<false><<<grid, NT, 0, s>>>(P);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

}  // namespace rosy_tex
