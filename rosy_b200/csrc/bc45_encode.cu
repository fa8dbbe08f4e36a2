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

#include <vector>

#include "bc45_block.cuh"
#include "tex_internal.h"

namespace rosy_tex {
namespace {

using namespace bc45;

constexpr int NT = 256;  // also the size of the reciprocal table

struct Params {
    Chain chain;
    const uint8_t* level0;
    const uint8_t* chainbuf;
    uint8_t* out;
    int first_level, num_levels;
    uint64_t first_block, num_blocks;
};

// One texture (P1, plist == nullptr) or a batch: plist[t] describes texture t, whose blocks are
// threads base_of[t] .. base_of[t+1]-1 of the launch.
template <bool BC5>
__global__ void __launch_bounds__(NT)
bc45_encode_kernel(const __grid_constant__ Params P1, const Params* __restrict__ plist,
                   const uint64_t* __restrict__ base_of, int ntex) {
    __shared__ uint32_t recip[256];
    recip[threadIdx.x] = g_RECIP20[threadIdx.x];  // NT == 256
    __syncthreads();
    uint64_t b = (uint64_t)blockIdx.x * NT + threadIdx.x;
    const Params* Pp = &P1;
    if (plist) {
        int lo = 0, hi = ntex;
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (b >= base_of[mid]) lo = mid; else hi = mid;
        }
        Pp = plist + lo;
        b -= base_of[lo];
    }
    const Params& P = *Pp;
    if (b >= P.num_blocks) return;
    const uint64_t gb = P.first_block + b;
    // three blocks in four belong to the first level: walk up only as far as needed
    int level = P.first_level;
    const int end_level = P.first_level + P.num_levels;
    while (level + 1 < end_level && gb >= P.chain.lv[level + 1].blk_offset) level++;
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
    uint32_t r4[4], g4[4];
    planes_rg(px, r4, g4);
    const uint2 rb = alpha_block(r4, recip);
    if (BC5) {
        const uint2 gbk = alpha_block(g4, recip);
        reinterpret_cast<uint4*>(P.out)[b] = make_uint4(rb.x, rb.y, gbk.x, gbk.y);
    } else {
        reinterpret_cast<uint2*>(P.out)[b] = rb;
    }
}

}  // namespace

static void fill_params(Params& P, const Chain& c, int first_level, int num_levels, bool bc5, const uint8_t* d_level0,
                        const uint8_t* d_chain, uint8_t* d_blocks) {
    P.chain = c;
    P.level0 = d_level0;
    P.chainbuf = d_chain;
    P.first_level = first_level;
    P.num_levels = num_levels;
    P.first_block = c.lv[first_level].blk_offset;
    const int last = first_level + num_levels - 1;
    P.num_blocks = c.lv[last].blk_offset + (uint64_t)c.lv[last].bw * c.lv[last].bh - P.first_block;
    P.out = d_blocks + (bc5 ? 16ull : 8ull) * P.first_block;
}

int launch_bc45(const Chain& c, int first_level, int num_levels, bool bc5, const uint8_t* d_level0,
                const uint8_t* d_chain, uint8_t* d_blocks, cudaStream_t s) {
    Params P;
    fill_params(P, c, first_level, num_levels, bc5, d_level0, d_chain, d_blocks);
    const unsigned grid = (unsigned)((P.num_blocks + NT - 1) / NT);
    if (bc5)
        bc45_encode_kernel<true><<<grid, NT, 0, s>>>(P, nullptr, nullptr, 1);
    else
        bc45_encode_kernel<false><<<grid, NT, 0, s>>>(P, nullptr, nullptr, 1);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }
size_t bc45_batch_desc_bytes(int n) {
    return align256((size_t)n * sizeof(Params)) + align256((size_t)(n + 2) * sizeof(uint64_t));
}

// BC5 / BC4 blocks of n textures (all levels each): one launch per format present.  d_desc:
// bc45_batch_desc_bytes(n) bytes of device memory that must stay untouched until the launches ran.
int launch_bc45_batch(int n, const Bc45BatchItem* items, void* d_desc, cudaStream_t s) {
    int launches = 0;
    uint8_t* dd = reinterpret_cast<uint8_t*>(d_desc);
    Params* d_ps = reinterpret_cast<Params*>(dd);
    uint64_t* d_base = reinterpret_cast<uint64_t*>(dd + align256((size_t)n * sizeof(Params)));
    std::vector<Params> ps;
    std::vector<uint64_t> base;
    size_t used_p = 0, used_b = 0;
    for (int fmt = 0; fmt < 2; fmt++) {
        const bool bc5 = fmt == 0;
        ps.clear();
        base.clear();
        uint64_t total = 0;
        for (int i = 0; i < n; i++) {
            if (items[i].bc5 != bc5) continue;
            Params P;
            const int fl = items[i].first_level;  // levels below it were encoded by the fused mip kernel
            if (fl >= items[i].chain.levels) continue;
            fill_params(P, items[i].chain, fl, items[i].chain.levels - fl, bc5, items[i].d_level0, items[i].d_chain, items[i].d_blocks);
            base.push_back(total);
            total += P.num_blocks;
            ps.push_back(P);
        }
        if (ps.empty()) continue;
        base.push_back(total);
        if (cudaMemcpyAsync(d_ps + used_p, ps.data(), ps.size() * sizeof(Params), cudaMemcpyHostToDevice, s) != cudaSuccess ||
            cudaMemcpyAsync(d_base + used_b, base.data(), base.size() * sizeof(uint64_t), cudaMemcpyHostToDevice, s) != cudaSuccess)
            return -1;
        const unsigned grid = (unsigned)((total + NT - 1) / NT);
        if (bc5)
            bc45_encode_kernel<true><<<grid, NT, 0, s>>>(ps[0], d_ps + used_p, d_base + used_b, (int)ps.size());
        else
            bc45_encode_kernel<false><<<grid, NT, 0, s>>>(ps[0], d_ps + used_p, d_base + used_b, (int)ps.size());
        if (cudaGetLastError() != cudaSuccess) return -1;
        used_p += ps.size();
        used_b += base.size();
        launches++;
    }
    return launches;
}

}  // namespace rosy_tex
