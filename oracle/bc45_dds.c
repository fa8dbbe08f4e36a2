/* bc45_dds.c - CPU ORACLE (test infrastructure, see rosy_oracle.h):
 *   BC4 / BC5 block encode  - restates /root/reference/src/libs/stb/stb_dxt.h:547-597
 *                             (stb__CompressAlphaBlock) and :617-626 (bc4 / bc5 entry points).
 *                             PINNED: tests compare against oracle/_ref (the reference's own
 *                             stb_dxt.h compiled as is) and against tests/golden vectors made from it.
 *   BC4 decode              - public BC4 definition (8-value / 6-value ramps).
 *   DDS header              - /root/reference/src/libs/dds_formats.hpp:251-311 structs, field values
 *                             per SURVEY.md row A10; PINNED by parsing with the reference's
 *                             dds::readImage (src/libs/dds.hpp:409-574) in oracle/_ref.
 */
#include <string.h>

#include "rosy_oracle.h"

/* One 8-byte alpha-style block over 16 samples `stride` bytes apart (stb_dxt.h:547-597). */
static void alpha_block(uint8_t* dest, const uint8_t* src, int stride) {
    int mn = src[0], mx = src[0];
    for (int i = 1; i < 16; i++) { /* :554-558 (note the else: one update per sample) */
        int v = src[i * stride];
        if (v < mn)
            mn = v;
        else if (v > mx)
            mx = v;
    }
    dest[0] = (uint8_t)mx; /* :561-562 */
    dest[1] = (uint8_t)mn;
    const int dist = mx - mn; /* :568-573 */
    const int dist4 = dist * 4, dist2 = dist * 2;
    int bias = dist < 8 ? dist - 1 : dist / 2 + 2;
    bias -= mn * 7;
    uint64_t bits = 0;
    for (int i = 0; i < 16; i++) { /* :576-596 */
        int a = src[i * stride] * 7 + bias;
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
        ind = (-ind) & 7; /* linear scale -> DXT index, 0/1 are the end points */
        ind ^= (2 > ind);
        bits |= (uint64_t)ind << (3 * i);
    }
    for (int i = 0; i < 6; i++) dest[2 + i] = (uint8_t)(bits >> (8 * i));
}

void oracle_bc4_encode_block(const uint8_t src16[16], uint8_t blk[8]) { alpha_block(blk, src16, 1); }

void oracle_bc5_encode_block(const uint8_t rg32[32], uint8_t blk[16]) {
    alpha_block(blk, rg32, 2);
    alpha_block(blk + 8, rg32 + 1, 2);
}

void oracle_bc4_decode_block(const uint8_t blk[8], uint8_t out16[16]) {
    int a0 = blk[0], a1 = blk[1];
    int ramp[8];
    ramp[0] = a0;
    ramp[1] = a1;
    if (a0 > a1) {
        for (int i = 1; i < 7; i++) ramp[i + 1] = ((7 - i) * a0 + i * a1) / 7;
    } else {
        for (int i = 1; i < 5; i++) ramp[i + 1] = ((5 - i) * a0 + i * a1) / 5;
        ramp[6] = 0;
        ramp[7] = 255;
    }
    uint64_t bits = 0;
    for (int i = 0; i < 6; i++) bits |= (uint64_t)blk[2 + i] << (8 * i);
    for (int i = 0; i < 16; i++) out16[i] = (uint8_t)ramp[(bits >> (3 * i)) & 7];
}

static void put32(uint8_t* p, uint32_t v) {
    p[0] = (uint8_t)v;
    p[1] = (uint8_t)(v >> 8);
    p[2] = (uint8_t)(v >> 16);
    p[3] = (uint8_t)(v >> 24);
}

static int block_bytes(int fmt) { return fmt == ORACLE_FMT_BC4 ? 8 : 16; }

/* Context::outputHeader, Packager.cpp:237 / :286 (row A10).  DX10 header, dxgiFormat
 * 98 = BC7_UNORM (sRGB flag never set, F13), 83 = BC5_UNORM, 80 = BC4_UNORM. */
void oracle_dds_header(int w, int h, int levels, int fmt, uint8_t hdr[ORACLE_DDS_HEADER_BYTES]) {
    memset(hdr, 0, ORACLE_DDS_HEADER_BYTES);
    memcpy(hdr, "DDS ", 4);
    uint8_t* p = hdr + 4;
    put32(p + 0, 124);
    put32(p + 4, 0x1 | 0x2 | 0x4 | 0x1000 | 0x20000 | 0x80000); /* 0xA1007 */
    put32(p + 8, (uint32_t)h);
    put32(p + 12, (uint32_t)w);
    put32(p + 16, (uint32_t)(((w + 3) / 4) * ((h + 3) / 4) * block_bytes(fmt))); /* linear size */
    put32(p + 20, 0);
    put32(p + 24, (uint32_t)levels);
    /* reserved[11] at p+28..p+71 left zero (NVTT stamps 'NVTT'+version into [9..10]; ignored) */
    uint8_t* pf = p + 72;
    put32(pf + 0, 32);
    put32(pf + 4, 0x4); /* FOURCC */
    memcpy(pf + 8, "DX10", 4);
    put32(p + 104, 0x401008); /* caps1: TEXTURE | MIPMAP | COMPLEX */
    uint8_t* dx = hdr + 128;
    put32(dx + 0, fmt == ORACLE_FMT_BC7 ? 98u : (fmt == ORACLE_FMT_BC5 ? 83u : 80u));
    put32(dx + 4, 3); /* TEXTURE2D */
    put32(dx + 8, 0);
    put32(dx + 12, 1);
    put32(dx + 16, 0);
}

size_t oracle_dds_size(int w, int h, int fmt) {
    return ORACLE_DDS_HEADER_BYTES + oracle_chain_blocks(w, h) * (size_t)block_bytes(fmt);
}
