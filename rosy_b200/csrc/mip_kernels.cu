// mip_kernels.cu - hand-written sm_100a mip-chain kernels (kernels K2 / K3 of SURVEY.md 2.4).
//
// Replaces the nvtt::Surface calls between two compress() calls of the reference:
//   colour  (/root/reference/src/Packager/Packager.cpp:254-258): toLinearFromSrgb, premultiplyAlpha,
//           buildNextMipmap(Box), demultiplyAlpha, toSrgb
//   normal  (Packager.cpp:294-295, :305): buildNextMipmap(Box) on the raw surface; every level
//           (level 0 included) is re-normalised on a copy before it is compressed.
// The reference keeps one fp32 surface and quantises only inside compress() (F11); here every
// level is likewise derived from fp32 values of the level above (never from quantised bytes), in
// linear premultiplied space for colour, so the result is within 1 LSB of the iterated chain.
//
// Two paths:
//   fused   - ONE launch for the whole chain.  Extents that are multiples of 64 (every config).
//             Persistent CTAs; each 64x64 RGBA8 tile (16 KB) is brought into shared memory by a
//             2-D TMA tensor copy (cp.async.bulk.tensor) completing on an mbarrier, 3 stages deep;
//             the 256 threads read 4x4 texel patches as uint4 rows, reduce levels 1-3 in registers /
//             shuffles, levels 4-6 in one warp, and the last CTA to finish (atomic ticket) reduces
//             the per-tile level-6 values down to 1x1.  HBM traffic = the algorithmic 5.33 B/px
//             (colour) or 9.33 B/px (normal maps, whose level 0 is rewritten).
//   generic - any extent (odd sizes use NVTT-2's polyphase box weights): one small
//             kernel per level over an fp32 ping-pong scratch.
#include <cuda.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdlib.h>
#include <stdint.h>

#include <mutex>

#include "tex_internal.h"

namespace rosy_tex {
namespace {

__constant__ float c_SRGB_TO_LINEAR[256];

__device__ __forceinline__ float linear_to_srgb(float c) {
    // Surface::toSrgb (NVTT-2 used the literal exponent 0.41666f), clamped to [0,1]
    const float s = c <= 0.0031308f ? 12.92f * c : 1.055f * __powf(c, 0.41666f) - 0.055f;
    return fminf(fmaxf(s, 0.0f), 1.0f);
}

__device__ __forceinline__ uint32_t quant8(float x) {
    return (uint32_t)__float2int_rd(fminf(fmaxf(x, 0.0f), 1.0f) * 255.0f + 0.5f);
}

// fp32 working value (colour: linear premultiplied rgb + alpha; normal: raw rgba) -> RGBA8 as
// Context::compress would see it
template <bool NORMAL>
__device__ __forceinline__ uint32_t finish_pixel(float4 v, int normalize) {
    float r = v.x, g = v.y, b = v.z;
    if (NORMAL) {
        if (normalize) {  // Surface::normalizeNormalMap on the copy (H4: unpack, normalise, repack)
            float x = 2.0f * r - 1.0f, y = 2.0f * g - 1.0f, z = 2.0f * b - 1.0f;
            const float l2 = x * x + y * y + z * z;
            const float il = l2 > 0.0f ? rsqrtf(l2) : 0.0f;
            r = x * il * 0.5f + 0.5f;
            g = y * il * 0.5f + 0.5f;
            b = z * il * 0.5f + 0.5f;
        }
    } else {
        const float ia = fabsf(v.w) > 1e-12f ? __frcp_rn(v.w) : 1.0f;  // demultiplyAlpha
        r = linear_to_srgb(r * ia);
        g = linear_to_srgb(g * ia);
        b = linear_to_srgb(b * ia);
    }
    return quant8(r) | (quant8(g) << 8) | (quant8(b) << 16) | (quant8(v.w) << 24);
}

// RGBA8 texel -> fp32 working value; lut = sRGB->linear table (colour only)
template <bool NORMAL>
__device__ __forceinline__ float4 load_pixel(uint32_t p, const float* lut) {
    const float a = (float)(p >> 24) * (1.0f / 255.0f);
    if (NORMAL) {
        return make_float4((float)(p & 255u) * (1.0f / 255.0f), (float)((p >> 8) & 255u) * (1.0f / 255.0f),
                           (float)((p >> 16) & 255u) * (1.0f / 255.0f), a);
    }
    return make_float4(lut[p & 255u] * a, lut[(p >> 8) & 255u] * a, lut[(p >> 16) & 255u] * a, a);
}

__device__ __forceinline__ float4 add4(float4 a, float4 b) {
    return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}
__device__ __forceinline__ float4 scale4(float4 a, float s) {
    return make_float4(a.x * s, a.y * s, a.z * s, a.w * s);
}
__device__ __forceinline__ float4 fma4(float s, float4 a, float4 acc) {
    return make_float4(acc.x + s * a.x, acc.y + s * a.y, acc.z + s * a.z, acc.w + s * a.w);
}

// 1-D box taps for output x of an extent-n input (row A7 / H7): extent 1 -> identity,
// even -> two taps of 1/2, odd n = 2k+1 -> three taps ((k-x), k, (x+1)) / n.
__device__ __forceinline__ int box_taps(int n, int on, int x, int& first, float wgt[3]) {
    if (n == 1) {
        first = 0;
        wgt[0] = 1.0f;
        return 1;
    }
    first = 2 * x;
    if ((n & 1) == 0) {
        wgt[0] = wgt[1] = 0.5f;
        return 2;
    }
    const float s = 1.0f / (float)n;
    wgt[0] = s * (float)(on - x);
    wgt[1] = s * (float)on;
    wgt[2] = s * (float)(1 + x);
    return 3;
}

// ------------------------------------------------------------------------------------------
// generic path: one level per launch
// ------------------------------------------------------------------------------------------
template <bool NORMAL, bool FROM_U8>
__global__ void __launch_bounds__(256)
mip_level_kernel(const void* __restrict__ src, int sw, int sh, float4* __restrict__ dst_f32,
                 uint32_t* __restrict__ dst_u8, int ow, int oh, int normalize) {
    __shared__ float lut[256];
    if (!NORMAL && FROM_U8) {
        lut[threadIdx.x] = c_SRGB_TO_LINEAR[threadIdx.x];
        __syncthreads();
    }
    const int x = blockIdx.x * 32 + (threadIdx.x & 31);
    const int y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= ow || y >= oh) return;
    int fx, fy;
    float wx[3], wy[3];
    const int nx = box_taps(sw, ow, x, fx, wx), ny = box_taps(sh, oh, y, fy, wy);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int j = 0; j < ny; j++) {
        float4 row = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int i = 0; i < nx; i++) {
            const size_t at = (size_t)(fy + j) * sw + (fx + i);
            float4 v;
            if (FROM_U8)
                v = load_pixel<NORMAL>(reinterpret_cast<const uint32_t*>(src)[at], lut);
            else
                v = reinterpret_cast<const float4*>(src)[at];
            row = fma4(wx[i], v, row);
        }
        acc = fma4(wy[j], row, acc);
    }
    const size_t o = (size_t)y * ow + x;
    dst_f32[o] = acc;
    dst_u8[o] = finish_pixel<NORMAL>(acc, normalize);
}

// level 0 of a normal map: re-normalised copy of the input (Packager.cpp:294-296 at mip 0)
__global__ void __launch_bounds__(256)
normal_level0_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n4, size_t npx,
                     int normalize) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i < n4) {
        const uint4 p = __ldg(src + i);
        uint4 o;
        o.x = finish_pixel<true>(load_pixel<true>(p.x, nullptr), normalize);
        o.y = finish_pixel<true>(load_pixel<true>(p.y, nullptr), normalize);
        o.z = finish_pixel<true>(load_pixel<true>(p.z, nullptr), normalize);
        o.w = finish_pixel<true>(load_pixel<true>(p.w, nullptr), normalize);
        dst[i] = o;
    } else if (i == n4) {  // up to 3 trailing texels when the pixel count is not a multiple of 4
        const uint32_t* s = reinterpret_cast<const uint32_t*>(src);
        uint32_t* d = reinterpret_cast<uint32_t*>(dst);
        for (size_t k = 4 * n4; k < npx; k++)
            d[k] = finish_pixel<true>(load_pixel<true>(s[k], nullptr), normalize);
    }
}

// ------------------------------------------------------------------------------------------
// fused path: whole chain in one launch, TMA-staged 64x64 tiles
// ------------------------------------------------------------------------------------------
constexpr int TILE = 64;
constexpr int STAGES = 3;
constexpr int TILE_BYTES = TILE * TILE * 4;

struct FusedParams {
    Chain chain;
    uint8_t* chainbuf;
    float4* tile_top;   // per-tile level-6 fp32 value, [tiles_y][tiles_x]
    float4* tail_a;     // ping-pong scratch for the tail levels
    float4* tail_b;
    unsigned int* ticket;
    int tiles_x, tiles_y;
    int normalize;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_tile(void* dst, const CUtensorMap* map, int x, int y,
                                              uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, "
        "%3}], [%4];" ::"r"(smem_u32(dst)),
        "l"(map), "r"(x), "r"(y), "r"(smem_u32(bar))
        : "memory");
}

template <bool NORMAL>
__global__ void __launch_bounds__(256)
mip_fused_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ FusedParams P) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* tiles = smem;                                            // STAGES x 16 KB
    float4* l3 = reinterpret_cast<float4*>(smem + STAGES * TILE_BYTES);  // 2 x 64 float4
    float* lut = reinterpret_cast<float*>(l3 + 128);                  // 256 floats
    uint64_t* full = reinterpret_cast<uint64_t*>(lut + 256);          // STAGES barriers
    __shared__ unsigned int s_ticket;

    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int ntiles = P.tiles_x * P.tiles_y;
    if (!NORMAL) lut[tid] = c_SRGB_TO_LINEAR[tid];
    if (tid == 0) {
        for (int s = 0; s < STAGES; s++) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap) : "memory");
    }
    __syncthreads();
    // prologue: fill the pipeline
    if (tid == 0) {
        for (int s = 0; s < STAGES; s++) {
            const int t = blockIdx.x + s * gridDim.x;
            if (t < ntiles) {
                mbar_expect_tx(&full[s], TILE_BYTES);
                tma_load_tile(tiles + s * TILE_BYTES, &tmap, (t % P.tiles_x) * TILE,
                              (t / P.tiles_x) * TILE, &full[s]);
            }
        }
    }
    const Level L1 = P.chain.lv[1], L2 = P.chain.lv[2], L3 = P.chain.lv[3];
    uint32_t* c1 = reinterpret_cast<uint32_t*>(P.chainbuf) + L1.px_offset;
    uint32_t* c2 = reinterpret_cast<uint32_t*>(P.chainbuf) + L2.px_offset;
    uint32_t* c3 = reinterpret_cast<uint32_t*>(P.chainbuf) + L3.px_offset;
    int it = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, it++) {
        const int s = it % STAGES;
        const int tile_x = t % P.tiles_x, tile_y = t / P.tiles_x;
        mbar_wait(&full[s], (uint32_t)((it / STAGES) & 1));
        // each thread owns a 4x4 texel patch: four 16-byte rows
        const uint4* trow = reinterpret_cast<const uint4*>(tiles + s * TILE_BYTES);
        uint4 rows[4];
#pragma unroll
        for (int r = 0; r < 4; r++) rows[r] = trow[(4 * ty + r) * (TILE / 4) + tx];
        if (NORMAL) {  // level 0 of a normal map is the re-normalised input
            uint4* c0 = reinterpret_cast<uint4*>(P.chainbuf);
#pragma unroll
            for (int r = 0; r < 4; r++) {
                uint4 o;
                o.x = finish_pixel<true>(load_pixel<true>(rows[r].x, nullptr), P.normalize);
                o.y = finish_pixel<true>(load_pixel<true>(rows[r].y, nullptr), P.normalize);
                o.z = finish_pixel<true>(load_pixel<true>(rows[r].z, nullptr), P.normalize);
                o.w = finish_pixel<true>(load_pixel<true>(rows[r].w, nullptr), P.normalize);
                const size_t y = (size_t)tile_y * TILE + 4 * ty + r;
                c0[(y * P.chain.lv[0].w + (size_t)tile_x * TILE) / 4 + tx] = o;
            }
        }
        // level 1: 2x2 per thread
        float4 q[2][2];
#pragma unroll
        for (int j = 0; j < 2; j++) {
            const uint4 a = rows[2 * j], b = rows[2 * j + 1];
            q[j][0] = scale4(add4(add4(load_pixel<NORMAL>(a.x, lut), load_pixel<NORMAL>(a.y, lut)),
                                  add4(load_pixel<NORMAL>(b.x, lut), load_pixel<NORMAL>(b.y, lut))),
                             0.25f);
            q[j][1] = scale4(add4(add4(load_pixel<NORMAL>(a.z, lut), load_pixel<NORMAL>(a.w, lut)),
                                  add4(load_pixel<NORMAL>(b.z, lut), load_pixel<NORMAL>(b.w, lut))),
                             0.25f);
        }
#pragma unroll
        for (int j = 0; j < 2; j++) {
            const size_t y = (size_t)tile_y * 32 + 2 * ty + j;
            uint2 o = make_uint2(finish_pixel<NORMAL>(q[j][0], P.normalize),
                                 finish_pixel<NORMAL>(q[j][1], P.normalize));
            *reinterpret_cast<uint2*>(c1 + y * L1.w + (size_t)tile_x * 32 + 2 * tx) = o;
        }
        // level 2: one texel per thread
        float4 v2 = scale4(add4(add4(q[0][0], q[0][1]), add4(q[1][0], q[1][1])), 0.25f);
        c2[((size_t)tile_y * 16 + ty) * L2.w + (size_t)tile_x * 16 + tx] =
            finish_pixel<NORMAL>(v2, P.normalize);
        // level 3: 2x2 threads (lane ^ 1 = tx neighbour, lane ^ 16 = ty neighbour)
        float4 v3 = v2;
        v3.x += __shfl_xor_sync(0xFFFFFFFFu, v3.x, 1);
        v3.y += __shfl_xor_sync(0xFFFFFFFFu, v3.y, 1);
        v3.z += __shfl_xor_sync(0xFFFFFFFFu, v3.z, 1);
        v3.w += __shfl_xor_sync(0xFFFFFFFFu, v3.w, 1);
        v3.x += __shfl_xor_sync(0xFFFFFFFFu, v3.x, 16);
        v3.y += __shfl_xor_sync(0xFFFFFFFFu, v3.y, 16);
        v3.z += __shfl_xor_sync(0xFFFFFFFFu, v3.z, 16);
        v3.w += __shfl_xor_sync(0xFFFFFFFFu, v3.w, 16);
        v3 = scale4(v3, 0.25f);
        float4* l3buf = l3 + (it & 1) * 64;
        if (((tx | ty) & 1) == 0) {
            const int x3 = tx >> 1, y3 = ty >> 1;
            l3buf[y3 * 8 + x3] = v3;
            c3[((size_t)tile_y * 8 + y3) * L3.w + (size_t)tile_x * 8 + x3] =
                finish_pixel<NORMAL>(v3, P.normalize);
        }
        __syncthreads();  // every thread has consumed stage s; level-3 values are visible
        if (tid == 0) {
            const int tn = t + STAGES * gridDim.x;
            if (tn < ntiles) {
                mbar_expect_tx(&full[s], TILE_BYTES);
                tma_load_tile(tiles + s * TILE_BYTES, &tmap, (tn % P.tiles_x) * TILE,
                              (tn / P.tiles_x) * TILE, &full[s]);
            }
        }
        if (tid < 32) {  // levels 4, 5, 6 of this tile in one warp
            const int lane = tid & 15;  // lanes 16..31 mirror 0..15
            const int x4 = lane & 3, y4 = lane >> 2;
            float4 v4 = scale4(add4(add4(l3buf[(2 * y4) * 8 + 2 * x4], l3buf[(2 * y4) * 8 + 2 * x4 + 1]),
                                    add4(l3buf[(2 * y4 + 1) * 8 + 2 * x4],
                                         l3buf[(2 * y4 + 1) * 8 + 2 * x4 + 1])),
                               0.25f);
            const Level L4 = P.chain.lv[4], L5 = P.chain.lv[5], L6 = P.chain.lv[6];
            uint32_t* cb = reinterpret_cast<uint32_t*>(P.chainbuf);
            if (tid < 16)
                cb[L4.px_offset + ((size_t)tile_y * 4 + y4) * L4.w + (size_t)tile_x * 4 + x4] =
                    finish_pixel<NORMAL>(v4, P.normalize);
            float4 v5 = v4;
            v5.x += __shfl_xor_sync(0xFFFFFFFFu, v5.x, 1);
            v5.y += __shfl_xor_sync(0xFFFFFFFFu, v5.y, 1);
            v5.z += __shfl_xor_sync(0xFFFFFFFFu, v5.z, 1);
            v5.w += __shfl_xor_sync(0xFFFFFFFFu, v5.w, 1);
            v5.x += __shfl_xor_sync(0xFFFFFFFFu, v5.x, 4);
            v5.y += __shfl_xor_sync(0xFFFFFFFFu, v5.y, 4);
            v5.z += __shfl_xor_sync(0xFFFFFFFFu, v5.z, 4);
            v5.w += __shfl_xor_sync(0xFFFFFFFFu, v5.w, 4);
            v5 = scale4(v5, 0.25f);
            if (tid < 16 && ((x4 | y4) & 1) == 0)
                cb[L5.px_offset + ((size_t)tile_y * 2 + (y4 >> 1)) * L5.w + (size_t)tile_x * 2 + (x4 >> 1)] =
                    finish_pixel<NORMAL>(v5, P.normalize);
            float4 v6 = v5;
            v6.x += __shfl_xor_sync(0xFFFFFFFFu, v6.x, 2);
            v6.y += __shfl_xor_sync(0xFFFFFFFFu, v6.y, 2);
            v6.z += __shfl_xor_sync(0xFFFFFFFFu, v6.z, 2);
            v6.w += __shfl_xor_sync(0xFFFFFFFFu, v6.w, 2);
            v6.x += __shfl_xor_sync(0xFFFFFFFFu, v6.x, 8);
            v6.y += __shfl_xor_sync(0xFFFFFFFFu, v6.y, 8);
            v6.z += __shfl_xor_sync(0xFFFFFFFFu, v6.z, 8);
            v6.w += __shfl_xor_sync(0xFFFFFFFFu, v6.w, 8);
            v6 = scale4(v6, 0.25f);
            if (tid == 0) {
                cb[L6.px_offset + (size_t)tile_y * L6.w + tile_x] = finish_pixel<NORMAL>(v6, P.normalize);
                P.tile_top[(size_t)tile_y * P.tiles_x + tile_x] = v6;
            }
        }
    }
    // ---- tail: the last CTA to finish reduces the per-tile values (level 6) down to 1x1 ----
    if (P.chain.levels <= 7) return;
    __threadfence();
    __syncthreads();
    if (tid == 0) s_ticket = atomicAdd(P.ticket, 1u);
    __syncthreads();
    if (s_ticket != gridDim.x - 1) return;
    __threadfence();
    const float4* src = P.tile_top;
    float4* dst = P.tail_a;
    int sw = P.tiles_x, sh = P.tiles_y;
    for (int l = 7; l < P.chain.levels; l++) {
        const Level L = P.chain.lv[l];
        uint32_t* out = reinterpret_cast<uint32_t*>(P.chainbuf) + L.px_offset;
        for (int o = tid; o < (int)(L.w * L.h); o += 256) {
            const int x = o % (int)L.w, y = o / (int)L.w;
            int fx, fy;
            float wx[3], wy[3];
            const int nx = box_taps(sw, (int)L.w, x, fx, wx), ny = box_taps(sh, (int)L.h, y, fy, wy);
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int j = 0; j < ny; j++) {
                float4 row = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int i = 0; i < nx; i++)
                    row = fma4(wx[i], __ldcg(src + (size_t)(fy + j) * sw + fx + i), row);
                acc = fma4(wy[j], row, acc);
            }
            dst[o] = acc;
            out[o] = finish_pixel<NORMAL>(acc, P.normalize);
        }
        __threadfence();
        __syncthreads();
        src = dst;
        dst = (dst == P.tail_a) ? P.tail_b : P.tail_a;
        sw = (int)L.w;
        sh = (int)L.h;
    }
    if (tid == 0) *P.ticket = 0u;  // ready for the next launch
}

// ---- host side ----------------------------------------------------------------------------
typedef CUresult (*encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                    CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                    CUtensorMapFloatOOBfill);

encode_tiled_fn get_encode_tiled() {
    static encode_tiled_fn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<encode_tiled_fn>(p);
    });
    return fn;
}

constexpr size_t FUSED_SMEM = STAGES * TILE_BYTES + 128 * sizeof(float4) + 256 * sizeof(float) +
                              STAGES * sizeof(uint64_t);

bool fused_ok(const Chain& c) {
    return c.lv[0].w % TILE == 0 && c.lv[0].h % TILE == 0 && c.levels >= 7;
}

size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

}  // namespace

static int g_force_generic = -1;

size_t mips_scratch_bytes(const Chain& c) {
    // generic: fp32 level 1 + fp32 level 2 (ping-pong);  fused: tile tops + two tail buffers + ticket
    const size_t l1 = c.levels > 1 ? (size_t)c.lv[1].w * c.lv[1].h * sizeof(float4) : 0;
    const size_t l2 = c.levels > 2 ? (size_t)c.lv[2].w * c.lv[2].h * sizeof(float4) : 0;
    const size_t generic = align256(l1) + align256(l2) + 256;
    const size_t ntiles = (size_t)((c.lv[0].w + TILE - 1) / TILE) * ((c.lv[0].h + TILE - 1) / TILE);
    const size_t fused = 256 + 3 * ntiles * sizeof(float4);
    return generic > fused ? generic : fused;
}

int mips_upload_tables() {
    float lut[256];
    for (int i = 0; i < 256; i++) {
        // Surface::toLinearFromSrgb (row A5) of i/255, computed in fp32 like the reference's surface
        const float c = (float)i / 255.0f;
        lut[i] = c <= 0.04045f ? c / 12.92f : powf((c + 0.055f) / 1.055f, 2.4f);
    }
    return cudaMemcpyToSymbol(c_SRGB_TO_LINEAR, lut, sizeof lut) == cudaSuccess ? 0 : -1;
}

int launch_mips(const Chain& c, bool normal_map, int normalize, const uint8_t* d_rgba,
                uint8_t* d_chain, void* d_scratch, cudaStream_t s) {
    if (g_force_generic < 0) {
        const char* e = getenv("ROSY_TEX_MIPS");
        g_force_generic = (e && e[0] == 'g') ? 1 : 0;
    }
    int launches = 0;
    const size_t npx0 = (size_t)c.lv[0].w * c.lv[0].h;
    if (fused_ok(c) && !g_force_generic && get_encode_tiled()) {
        CUtensorMap tmap;
        const cuuint64_t dims[2] = {c.lv[0].w, c.lv[0].h};
        const cuuint64_t strides[1] = {(cuuint64_t)c.lv[0].w * 4};
        const cuuint32_t box[2] = {TILE, TILE};
        const cuuint32_t estr[2] = {1, 1};
        CUresult r = get_encode_tiled()(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<uint8_t*>(d_rgba),
                                        dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                        CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return -1;
        FusedParams P;
        P.chain = c;
        P.chainbuf = d_chain;
        P.tiles_x = (int)(c.lv[0].w / TILE);
        P.tiles_y = (int)(c.lv[0].h / TILE);
        P.normalize = normalize;
        const size_t ntiles = (size_t)P.tiles_x * P.tiles_y;
        uint8_t* sc = reinterpret_cast<uint8_t*>(d_scratch);
        P.ticket = reinterpret_cast<unsigned int*>(sc);
        P.tile_top = reinterpret_cast<float4*>(sc + 256);
        P.tail_a = P.tile_top + ntiles;
        P.tail_b = P.tail_a + ntiles;
        int dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        const unsigned grid = (unsigned)(ntiles < (size_t)(3 * sms) ? ntiles : (size_t)(3 * sms));
        cudaFuncSetAttribute(mip_fused_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FUSED_SMEM);
        cudaFuncSetAttribute(mip_fused_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FUSED_SMEM);
        if (normal_map)
            mip_fused_kernel<true><<<grid, 256, FUSED_SMEM, s>>>(tmap, P);
        else
            mip_fused_kernel<false><<<grid, 256, FUSED_SMEM, s>>>(tmap, P);
        return cudaGetLastError() == cudaSuccess ? 1 : -1;
    }
    // ---- generic path ----
    if (normal_map) {
        const size_t n4 = npx0 / 4;
        normal_level0_kernel<<<(unsigned)((n4 + 1 + 255) / 256), 256, 0, s>>>(
            reinterpret_cast<const uint4*>(d_rgba), reinterpret_cast<uint4*>(d_chain), n4, npx0, normalize);
        launches++;
    }
    uint8_t* sc = reinterpret_cast<uint8_t*>(d_scratch) + 256;
    const size_t l1 = c.levels > 1 ? (size_t)c.lv[1].w * c.lv[1].h * sizeof(float4) : 0;
    float4* buf[2] = {reinterpret_cast<float4*>(sc), reinterpret_cast<float4*>(sc + align256(l1))};
    for (int l = 1; l < c.levels; l++) {
        const Level& S = c.lv[l - 1];
        const Level& D = c.lv[l];
        uint32_t* out = reinterpret_cast<uint32_t*>(d_chain) + D.px_offset;
        float4* dst = buf[(l - 1) & 1];
        const dim3 grid((D.w + 31) / 32, (D.h + 7) / 8);
        if (l == 1) {
            if (normal_map)
                mip_level_kernel<true, true><<<grid, 256, 0, s>>>(d_rgba, S.w, S.h, dst, out, D.w, D.h, normalize);
            else
                mip_level_kernel<false, true><<<grid, 256, 0, s>>>(d_rgba, S.w, S.h, dst, out, D.w, D.h, normalize);
        } else {
            const float4* src = buf[l & 1];
            if (normal_map)
                mip_level_kernel<true, false><<<grid, 256, 0, s>>>(src, S.w, S.h, dst, out, D.w, D.h, normalize);
            else
                mip_level_kernel<false, false><<<grid, 256, 0, s>>>(src, S.w, S.h, dst, out, D.w, D.h, normalize);
        }
        launches++;
    }
    return cudaGetLastError() == cudaSuccess ? launches : -1;
}

}  // namespace rosy_tex
