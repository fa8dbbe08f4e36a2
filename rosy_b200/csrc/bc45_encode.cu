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

#include "tex_internal.h"

namespace rosy_tex {
namespace {

constexpr int NT = 256;  // also the size of the reciprocal table

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

struct Params {
    Chain chain;
    const uint8_t* level0;
    const uint8_t* chainbuf;
    uint8_t* out;
    int first_level, num_levels;
    uint64_t first_block, num_blocks;
};

// 16 samples packed 4 per word (texel 4k+j in byte j of v[k]) -> 8-byte alpha-style block.
//
// stb__CompressAlphaBlock picks, per sample, ind = floor(a / dist) by three compare-and-subtract
// steps (a = 7 v + bias, 0 <= a < 8 dist) and maps it to the block's index order
// {1,7,6,5,4,3,2,0}[ind].  The same integers are produced here with far fewer instructions (the
// kernel has to encode at HBM speed): the quotient is one multiply by the exact 20-bit reciprocal
// recip[dist] = floor(2^20 / dist) + 1 (exact because a * dist < 2^20), the sixteen 3-bit quotients
// are gathered as nibbles of two words, and the index map and the 4-bit -> 3-bit packing are done on
// eight samples at a time with word-wide logic.
__device__ __forceinline__ uint32_t map_and_pack8(uint32_t x /* eight quotients, one per nibble */) {
    // {1,7,6,5,4,3,2,0}[q] = ((8 - q) & 7) ^ (1 if the result is 0 or 1)
    uint32_t r = (0x88888888u - x) & 0x77777777u;
    r ^= ~((r >> 1) | (r >> 2)) & 0x11111111u;
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
    const int dist = mx - mn;
    int bias = dist < 8 ? dist - 1 : dist / 2 + 2;
    bias -= mn * 7;
    const uint32_t M = recip[dist];  // 0 for dist == 0: every quotient is 0, as in stb (a = -1 fails every compare)
    const uint32_t M7 = 7u * M, biasM = (uint32_t)(bias * (int)M);
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
    // byte-planar R (and G): texel 4k+j -> byte j of word k
    uint32_t r4[4], g4[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t t0 = __byte_perm(px[4 * k], px[4 * k + 1], 0x5140);
        const uint32_t t1 = __byte_perm(px[4 * k + 2], px[4 * k + 3], 0x5140);
        r4[k] = __byte_perm(t0, t1, 0x5410);
        g4[k] = __byte_perm(t0, t1, 0x7632);
    }
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
            fill_params(P, items[i].chain, 0, items[i].chain.levels, bc5, items[i].d_level0, items[i].d_chain, items[i].d_blocks);
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
