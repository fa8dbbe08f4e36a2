/* rosy_oracle.h - CPU ORACLE for the Rosy Packager texture stage.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (rosy_b200/, include/) may
 * include, link or call this.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and only as the checker.
 *
 * PARITY UNPINNED versus NVTT: the reference's arithmetic for this path lives inside
 * NVIDIA Texture Tools 3.2.5 (closed source DLL `nvtt30205`, /root/reference/src/premake5.lua:123),
 * which is absent from /root/reference and from this image, and the reference has no tests or
 * golden vectors for the path.  This file restates
 *   - the call sequence of src/Packager/Packager.cpp:216-308 literally,
 *   - NVTT's documented per-call semantics,
 *   - the public BC7 (BPTC) specification for block layout / decoding,
 *   - src/libs/stb/stb_dxt.h:547-626 for BC4/BC5 (this piece IS pinned: oracle/_ref compiles
 *     the reference's stb_dxt.h and tests/golden holds vectors generated from it),
 *   - src/libs/dds_formats.hpp:251-311 + src/libs/dds.hpp:409-574 for the container
 *     (pinned by parsing emitted files with the reference's own dds::readImage, oracle/_ref).
 */
#ifndef ROSY_ORACLE_H
#define ROSY_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* image kinds: Asset.h:109-112 image_type 0/1/2/3; only the routing matters:
 * colour / metallic-roughness / mixmap -> srgb function, normal -> normal function. */
enum { ORACLE_KIND_SRGB = 0, ORACLE_KIND_NORMAL = 1 };
enum { ORACLE_FMT_BC7 = 0, ORACLE_FMT_BC5 = 1, ORACLE_FMT_BC4 = 2 };
/* Quality tiers: nvtt::Quality_Fastest / Normal / Production (reference uses Normal only,
 * Packager.cpp:277 and NVTT default at :228-229). */
enum { ORACLE_Q_FAST = 0, ORACLE_Q_NORMAL = 1, ORACLE_Q_HIGH = 2 };
/* Oracle-only ceiling: every partition / rotation / index selector / p-bit combination, least
 * squares to convergence.  Not a product tier; it bounds what the three tiers above can reach. */
enum { ORACLE_Q_EXHAUSTIVE = 3 };
/* normalizeNormalMap reading: 1 = unpack 2v-1, safe-normalise, repack (default);
 * 0 = literal NVTT-2 no-op (surface was never flagged setNormalMap(true)). */
enum { ORACLE_NORMALIZE_OFF = 0, ORACLE_NORMALIZE_UNPACK = 1 };

/* nvtt::Surface::countMipmaps(): Packager.cpp:235,284 */
int oracle_count_mipmaps(int w, int h);
/* extent of level `level` */
void oracle_level_extent(int w, int h, int level, int* lw, int* lh);
/* total pixels over all levels, total 4x4 blocks over all levels */
size_t oracle_chain_pixels(int w, int h);
size_t oracle_chain_blocks(int w, int h);

/* Colour chain, Packager.cpp:243-259: out receives every level as RGBA8 (what
 * Context::compress quantises to), level 0 first, tightly packed. */
void oracle_mip_chain_srgb(const uint8_t* rgba, int w, int h, uint8_t* out);
/* Normal chain, Packager.cpp:292-306. */
void oracle_mip_chain_normal(const uint8_t* rgba, int w, int h, int normalize_mode, uint8_t* out);
/* The same chains with the rows of every pass shared by `threads` host threads (<= 0: all cores); results
 * are identical for every thread count (the passes are element-wise or row-local). */
void oracle_mip_chain_srgb_mt(const uint8_t* rgba, int w, int h, uint8_t* out, int threads);
void oracle_mip_chain_normal_mt(const uint8_t* rgba, int w, int h, int normalize_mode, uint8_t* out, int threads);

/* --- BC7 (public BPTC spec) --- */
void oracle_bc7_decode_block(const uint8_t blk[16], uint8_t rgba[64]);
/* returns the mode (0..7), or -1 for the reserved encoding (first byte 0) */
int oracle_bc7_block_mode(const uint8_t blk[16]);
void oracle_bc7_encode_block(const uint8_t rgba[64], int quality, uint8_t blk[16]);

/* --- BC4 / BC5, stb_dxt.h:547-626 --- */
void oracle_bc4_encode_block(const uint8_t src16[16], uint8_t blk[8]);
void oracle_bc5_encode_block(const uint8_t rg32[32], uint8_t blk[16]);
void oracle_bc4_decode_block(const uint8_t blk[8], uint8_t out16[16]);

/* Image-level helpers (row-major blocks, edge blocks padded by clamping coordinates).
 * threads <= 0: all cores (OpenMP). */
void oracle_encode_image(const uint8_t* rgba, int w, int h, int fmt, int quality, uint8_t* blocks,
                         int threads);
/* decode to RGBA8 (BC4: r,0,0,255; BC5: r,g,0,255) */
void oracle_decode_image(const uint8_t* blocks, int w, int h, int fmt, uint8_t* rgba);

/* --- DDS container, dds_formats.hpp:251-311; values per SURVEY.md row A10 --- */
#define ORACLE_DDS_HEADER_BYTES 148
void oracle_dds_header(int w, int h, int levels, int fmt, uint8_t hdr[ORACLE_DDS_HEADER_BYTES]);
size_t oracle_dds_size(int w, int h, int fmt);

/* Whole stage, memory to memory: what generate_srgb_texture / generate_normal_map_texture
 * write to <image>.dds for an already decoded RGBA8 image.  Returns bytes written. */
size_t oracle_generate_texture(const uint8_t* rgba, int w, int h, int kind, int fmt, int quality,
                               int normalize_mode, uint8_t* dds_out, int threads);

int oracle_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
