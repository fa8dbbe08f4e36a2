// image_io.cpp - host-side image file decode for the path entry points.
//
// Stands in for nvtt::Surface::load (/root/reference/src/Packager/Packager.cpp:222, :269; SURVEY
// row A3): decode a file to 8-bit RGBA, images without alpha get alpha = 255.  glTF image URIs
// (src/Packager/Gltf.cpp:98-111) are PNG / JPEG; the FBX path feeds TGA (src/Packager/Fbx.cpp:421-438).
// Written from the PNG / TGA / PNM specifications (zlib does the inflate); JPEG lives in jpeg_decode.cpp.  Raw ".rgba8" files (u32 width, u32 height,
// pixels) are accepted for benchmarks.
#include <ctype.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

#include <string>
#include <vector>

#include "../../include/rosy_tex_cuda.h"

namespace rosy_tex {

// jpeg_decode.cpp
int decode_jpeg(const std::vector<uint8_t>& f, std::vector<uint8_t>& rgba, uint32_t& W, uint32_t& H, std::string& err);
bool inflate_fast(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_len, bool raw);  // inflate_fast.cpp

static bool read_file(const char* path, std::vector<uint8_t>& out) {
    FILE* f = fopen(path, "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    if (n < 0) {
        fclose(f);
        return false;
    }
    out.resize((size_t)n);
    size_t got = n ? fread(out.data(), 1, (size_t)n, f) : 0;
    fclose(f);
    return got == (size_t)n;
}

static uint32_t be32(const uint8_t* p) {
    return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3];
}

// ---------------------------------------------------------------- PNG
static int paeth(int a, int b, int c) {
    int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// undo the scanline filters of one (sub)image in place; returns false on a bad filter byte
// One row, filter type FT, BPP bytes per complete pixel (a compile-time constant for the common layouts so
// that the left-neighbour distance is fixed and the loops unroll).  The file pipeline is decode-bound
// (DESIGN.md section 5), so the row loops are kept free of per-byte dispatch.
template <int FT, size_t BPP>
static void png_unfilter_row(uint8_t* cur, const uint8_t* prev, size_t stride, size_t bpp_rt) {
    const size_t bpp = BPP ? BPP : bpp_rt;
    const size_t head = bpp < stride ? bpp : stride;
    for (size_t i = 0; i < head; i++) {  // no left neighbour yet: a = c = 0
        const int b = prev[i];
        int v = cur[i];
        if (FT == 2) v += b;
        if (FT == 3) v += b >> 1;
        if (FT == 4) v += paeth(0, b, 0);
        cur[i] = (uint8_t)v;
    }
    for (size_t i = bpp; i < stride; i++) {
        const int a = cur[i - bpp], b = prev[i], c = prev[i - bpp];
        int v = cur[i];
        if (FT == 1) v += a;
        if (FT == 2) v += b;
        if (FT == 3) v += (a + b) >> 1;
        if (FT == 4) v += paeth(a, b, c);
        cur[i] = (uint8_t)v;
    }
}
#if defined(__SSE2__)
// Sub / Average / Paeth carry a dependency from pixel to pixel, but the 3 or 4 channels of a pixel are independent:
// one pixel per step in 16-bit SSE2 lanes (the branchy scalar Paeth costs ~10 ns per byte on noisy textures, this
// ~1.5).  Same arithmetic, same tie order (a, then b, then c).  Reads 4 bytes per pixel: for BPP = 3 one byte past the
// pixel, which is the next pixel, the next row's filter byte or the buffer's slack; stores exactly BPP bytes.
template <int FT, size_t BPP>
static void png_unfilter_row_sse2(uint8_t* cur, const uint8_t* prev, size_t stride) {
    const __m128i zero = _mm_setzero_si128(), mask8 = _mm_set1_epi16(0x00FF);
    __m128i a = zero, c = zero;  // left and upper-left pixel, one channel per 16-bit lane
    for (size_t i = 0; i + BPP <= stride; i += BPP) {
        uint32_t wx, wb;
        memcpy(&wx, cur + i, 4);
        memcpy(&wb, prev + i, 4);
        __m128i x = _mm_unpacklo_epi8(_mm_cvtsi32_si128((int)wx), zero);
        const __m128i b = _mm_unpacklo_epi8(_mm_cvtsi32_si128((int)wb), zero);
        if (FT == 1) {
            x = _mm_add_epi16(x, a);
        } else if (FT == 3) {
            x = _mm_add_epi16(x, _mm_srli_epi16(_mm_add_epi16(a, b), 1));
        } else {
            __m128i pa = _mm_sub_epi16(b, c), pb = _mm_sub_epi16(a, c);
            __m128i pc = _mm_add_epi16(pa, pb);
            pa = _mm_max_epi16(pa, _mm_sub_epi16(zero, pa));
            pb = _mm_max_epi16(pb, _mm_sub_epi16(zero, pb));
            pc = _mm_max_epi16(pc, _mm_sub_epi16(zero, pc));
            const __m128i smallest = _mm_min_epi16(pc, _mm_min_epi16(pa, pb));
            const __m128i is_a = _mm_cmpeq_epi16(smallest, pa), is_b = _mm_cmpeq_epi16(smallest, pb);
            const __m128i bc = _mm_or_si128(_mm_and_si128(is_b, b), _mm_andnot_si128(is_b, c));
            x = _mm_add_epi16(x, _mm_or_si128(_mm_and_si128(is_a, a), _mm_andnot_si128(is_a, bc)));
        }
        x = _mm_and_si128(x, mask8);
        const uint32_t out = (uint32_t)_mm_cvtsi128_si32(_mm_packus_epi16(x, x));
        memcpy(cur + i, &out, BPP);
        a = x;
        c = b;
    }
}
#endif

template <size_t BPP>
static bool png_unfilter_rows(uint8_t* data, size_t stride, size_t h, size_t bpp_rt) {
    std::vector<uint8_t> zero(stride + 4, 0);
    const uint8_t* prev = zero.data();
    for (size_t y = 0; y < h; y++) {
        uint8_t* row = data + y * (stride + 1);
        uint8_t* cur = row + 1;
#if defined(__SSE2__)
        constexpr bool simd = BPP == 3 || BPP == 4;
#else
        constexpr bool simd = false;
#endif
        switch (row[0]) {
            case 0: break;
            case 2: png_unfilter_row<2, BPP>(cur, prev, stride, bpp_rt); break;  // no carried dependency: the compiler vectorises it
#if defined(__SSE2__)
            case 1: if (simd) png_unfilter_row_sse2<1, BPP ? BPP : 4>(cur, prev, stride); else png_unfilter_row<1, BPP>(cur, prev, stride, bpp_rt); break;
            case 3: if (simd) png_unfilter_row_sse2<3, BPP ? BPP : 4>(cur, prev, stride); else png_unfilter_row<3, BPP>(cur, prev, stride, bpp_rt); break;
            case 4: if (simd) png_unfilter_row_sse2<4, BPP ? BPP : 4>(cur, prev, stride); else png_unfilter_row<4, BPP>(cur, prev, stride, bpp_rt); break;
#else
            case 1: png_unfilter_row<1, BPP>(cur, prev, stride, bpp_rt); break;
            case 3: png_unfilter_row<3, BPP>(cur, prev, stride, bpp_rt); break;
            case 4: png_unfilter_row<4, BPP>(cur, prev, stride, bpp_rt); break;
#endif
            default: return false;
        }
        prev = cur;
    }
    return true;
}
static bool png_unfilter(uint8_t* data, size_t w, size_t h, int bpp_bits) {
    const size_t stride = (w * (size_t)bpp_bits + 7) / 8, bpp = (size_t)(bpp_bits + 7) / 8;
    if (bpp == 4) return png_unfilter_rows<4>(data, stride, h, bpp);
    if (bpp == 3) return png_unfilter_rows<3>(data, stride, h, bpp);
    return png_unfilter_rows<0>(data, stride, h, bpp);
}

static int decode_png(const std::vector<uint8_t>& f, std::vector<uint8_t>& rgba, uint32_t& W, uint32_t& H) {
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (f.size() < 8 + 25 || memcmp(f.data(), sig, 8)) return ROSY_TEX_ERR_DECODE;
    size_t pos = 8;
    uint32_t w = 0, h = 0;
    int depth = 0, ctype = 0, interlace = 0;
    std::vector<uint8_t> idat, plte, trns;
    bool have_ihdr = false, done = false, cgbi = false;
    while (!done && pos + 12 <= f.size()) {
        const uint32_t len = be32(&f[pos]);
        const uint8_t* type = &f[pos + 4];
        const uint8_t* body = &f[pos + 8];
        if (pos + 12 + (size_t)len > f.size()) return ROSY_TEX_ERR_DECODE;
        if (!memcmp(type, "IHDR", 4)) {
            if (len < 13) return ROSY_TEX_ERR_DECODE;
            w = be32(body);
            h = be32(body + 4);
            depth = body[8];
            ctype = body[9];
            interlace = body[12];
            if (body[10] != 0 || body[11] != 0 || interlace > 1) return ROSY_TEX_ERR_DECODE;
            have_ihdr = true;
        } else if (!memcmp(type, "CgBI", 4)) {
            cgbi = true;  // Apple's variant: raw deflate stream; channels stay as stored, like stb_image's default
        } else if (!memcmp(type, "PLTE", 4)) {
            plte.assign(body, body + len);
        } else if (!memcmp(type, "tRNS", 4)) {
            trns.assign(body, body + len);
        } else if (!memcmp(type, "IDAT", 4)) {
            idat.insert(idat.end(), body, body + len);
        } else if (!memcmp(type, "IEND", 4)) {
            done = true;
        }
        pos += 12 + (size_t)len;
    }
    if (!have_ihdr || w == 0 || h == 0 || w > (1u << 15) || h > (1u << 15)) return ROSY_TEX_ERR_DECODE;
    int channels;
    switch (ctype) {
        case 0: channels = 1; break;
        case 2: channels = 3; break;
        case 3: channels = 1; break;
        case 4: channels = 2; break;
        case 6: channels = 4; break;
        default: return ROSY_TEX_ERR_DECODE;
    }
    if (!(depth == 1 || depth == 2 || depth == 4 || depth == 8 || depth == 16)) return ROSY_TEX_ERR_DECODE;
    if ((ctype == 2 || ctype == 4 || ctype == 6) && depth < 8) return ROSY_TEX_ERR_DECODE;
    if (ctype == 3 && (depth == 16 || plte.empty())) return ROSY_TEX_ERR_DECODE;
    const int bpp_bits = channels * depth;
    // passes: {x0, y0, dx, dy}
    static const int adam7[7][4] = {{0, 0, 8, 8}, {4, 0, 8, 8}, {0, 4, 4, 8}, {2, 0, 4, 4},
                                    {0, 2, 2, 4}, {1, 0, 2, 2}, {0, 1, 1, 2}};
    const int npass = interlace ? 7 : 1;
    size_t raw_size = 0;
    for (int p = 0; p < npass; p++) {
        const size_t pw = interlace ? (w - adam7[p][0] + adam7[p][2] - 1) / adam7[p][2] : w;
        const size_t ph = interlace ? (h - adam7[p][1] + adam7[p][3] - 1) / adam7[p][3] : h;
        if (interlace && (w <= (uint32_t)adam7[p][0] || h <= (uint32_t)adam7[p][1])) continue;
        if (pw && ph) raw_size += ph * (1 + (pw * (size_t)bpp_bits + 7) / 8);
    }
    // inflate_fast.cpp: whole-buffer decoder (1.5x zlib on texture data); anything it does not accept - malformed or
    // unusual streams - goes through zlib below, which then yields the data or the error
    std::vector<uint8_t> raw(raw_size + 16);
    const size_t idat_len = idat.size();
    idat.resize(idat_len + 32, 0);  // readable padding for the decoder's 8-byte loads
    uLongf got = (uLongf)raw_size;
    if (inflate_fast(idat.data(), idat_len, raw.data(), raw_size, cgbi)) {
        // done
    } else if (cgbi) {
        z_stream zs;
        memset(&zs, 0, sizeof zs);
        if (inflateInit2(&zs, -15) != Z_OK) return ROSY_TEX_ERR_DECODE;
        zs.next_in = idat.data();
        zs.avail_in = (uInt)idat_len;
        zs.next_out = raw.data();
        zs.avail_out = (uInt)raw_size;
        const int zr = inflate(&zs, Z_FINISH);
        got = (uLongf)zs.total_out;
        inflateEnd(&zs);
        if ((zr != Z_STREAM_END && zr != Z_BUF_ERROR && zr != Z_OK) || got != raw_size) return ROSY_TEX_ERR_DECODE;
    } else {
        const int zr = uncompress(raw.data(), &got, idat.data(), (uLong)idat_len);
        if ((zr != Z_OK && zr != Z_BUF_ERROR) || got != raw_size) return ROSY_TEX_ERR_DECODE;
    }
    W = w;
    H = h;
    rgba.assign((size_t)w * h * 4, 255);
    size_t off = 0;
    const int maxv = (1 << (depth > 8 ? 8 : depth)) - 1;
    for (int p = 0; p < npass; p++) {
        size_t pw = w, ph = h;
        int x0 = 0, y0 = 0, dx = 1, dy = 1;
        if (interlace) {
            x0 = adam7[p][0];
            y0 = adam7[p][1];
            dx = adam7[p][2];
            dy = adam7[p][3];
            if (w <= (uint32_t)x0 || h <= (uint32_t)y0) continue;
            pw = (w - x0 + dx - 1) / dx;
            ph = (h - y0 + dy - 1) / dy;
        }
        if (!pw || !ph) continue;
        const size_t stride = (pw * (size_t)bpp_bits + 7) / 8;
        uint8_t* sub = raw.data() + off;
        if (!png_unfilter(sub, pw, ph, bpp_bits)) return ROSY_TEX_ERR_DECODE;
        // the two layouts almost every texture uses: 8-bit RGBA / RGB, not interlaced, no colour key
        const bool fast = !interlace && depth == 8 && (ctype == 6 || (ctype == 2 && trns.size() < 6));
        for (size_t y = 0; y < ph; y++) {
            const uint8_t* row = sub + y * (stride + 1) + 1;
            if (fast) {
                uint8_t* o = &rgba[y * (size_t)w * 4];
                if (ctype == 6) {
                    memcpy(o, row, (size_t)w * 4);
                } else {
                    for (size_t x = 0; x < pw; x++) {
                        o[4 * x] = row[3 * x];
                        o[4 * x + 1] = row[3 * x + 1];
                        o[4 * x + 2] = row[3 * x + 2];
                    }  // alpha is already 255
                }
                continue;
            }
            for (size_t x = 0; x < pw; x++) {
                int s[4] = {0, 0, 0, 0};   // samples as 8-bit (16-bit: high byte, like stb_image)
                int s16[4] = {0, 0, 0, 0};  // full samples for the tRNS colour key
                for (int c = 0; c < channels; c++) {
                    if (depth == 16) {
                        s16[c] = (row[(x * channels + c) * 2] << 8) | row[(x * channels + c) * 2 + 1];
                        s[c] = s16[c] >> 8;
                    } else if (depth == 8) {
                        s16[c] = s[c] = row[x * channels + c];
                    } else {
                        const size_t bit = x * (size_t)depth;
                        s16[c] = s[c] = (row[bit >> 3] >> (8 - depth - (bit & 7))) & maxv;
                    }
                }
                uint8_t* o = &rgba[(((size_t)y0 + y * dy) * w + x0 + x * dx) * 4];
                switch (ctype) {
                    case 0: {
                        const int g = depth < 8 ? s[0] * 255 / maxv : s[0];
                        o[0] = o[1] = o[2] = (uint8_t)g;
                        o[3] = (trns.size() >= 2 && s16[0] == ((trns[0] << 8) | trns[1])) ? 0 : 255;
                        break;
                    }
                    case 2:
                        o[0] = (uint8_t)s[0];
                        o[1] = (uint8_t)s[1];
                        o[2] = (uint8_t)s[2];
                        o[3] = (trns.size() >= 6 && s16[0] == ((trns[0] << 8) | trns[1]) &&
                                s16[1] == ((trns[2] << 8) | trns[3]) && s16[2] == ((trns[4] << 8) | trns[5]))
                                   ? 0
                                   : 255;
                        break;
                    case 3: {
                        const size_t k = (size_t)s[0];
                        if (3 * k + 2 >= plte.size()) return ROSY_TEX_ERR_DECODE;
                        o[0] = plte[3 * k];
                        o[1] = plte[3 * k + 1];
                        o[2] = plte[3 * k + 2];
                        o[3] = k < trns.size() ? trns[k] : 255;
                        break;
                    }
                    case 4:
                        o[0] = o[1] = o[2] = (uint8_t)s[0];
                        o[3] = (uint8_t)s[1];
                        break;
                    default:
                        o[0] = (uint8_t)s[0];
                        o[1] = (uint8_t)s[1];
                        o[2] = (uint8_t)s[2];
                        o[3] = (uint8_t)s[3];
                }
            }
        }
        off += ph * (stride + 1);
    }
    return ROSY_TEX_OK;
}

// ---------------------------------------------------------------- TGA
// Types 1/9 (colour-mapped), 2/10 (true colour), 3/11 (grey), raw or RLE; 8/15/16/24/32 bits per pixel
// or per palette entry, 8- or 16-bit palette indices.  Behaviour follows the reference tree's
// stb_image.h:5735-6074 so decoded bytes are identical (tests/golden/stb_image_tga.npz): 15/16-bit
// pixels are 5-5-5 expanded with (v*255)/31 and carry no alpha, 16-bit grey is grey + alpha, the
// right-to-left descriptor bit is ignored, `first entry index` of a colour map is a byte skip, and an
// out-of-range palette index reads entry 0.  Unlike stb, truncated files are an error.
static int decode_tga(const std::vector<uint8_t>& f, std::vector<uint8_t>& rgba, uint32_t& W, uint32_t& H) {
    if (f.size() < 18) return ROSY_TEX_ERR_DECODE;
    const int idlen = f[0], cmap = f[1];
    int type = f[2];
    const size_t pal_start = f[3] | (f[4] << 8), pal_len = f[5] | (f[6] << 8);
    const int pal_bits = f[7];
    const uint32_t w = f[12] | (f[13] << 8), h = f[14] | (f[15] << 8);
    const int bpp = f[16], desc = f[17];
    if (cmap > 1 || w == 0 || h == 0 || w > 32768 || h > 32768) return ROSY_TEX_ERR_DECODE;
    if (cmap ? !(type == 1 || type == 9) : !(type == 2 || type == 3 || type == 10 || type == 11))
        return ROSY_TEX_ERR_DECODE;
    const bool rle = type >= 8;
    if (rle) type -= 8;
    if (cmap && bpp != 8 && bpp != 16) return ROSY_TEX_ERR_DECODE;
    // components of a pixel (or palette entry): 1 grey, 2 grey + alpha, 3 RGB (also 5-5-5), 4 RGBA
    const int cbits = cmap ? pal_bits : bpp;
    int comp = 0;
    bool rgb16 = false;
    switch (cbits) {
        case 8: comp = 1; break;
        case 16:
            if (!cmap && type == 3) {
                comp = 2;
                break;
            }
            [[fallthrough]];
        case 15:
            comp = 3;
            rgb16 = true;
            break;
        case 24: comp = 3; break;
        case 32: comp = 4; break;
        default: return ROSY_TEX_ERR_DECODE;
    }
    const size_t n = (size_t)w * h;
    const size_t src_bytes = cmap ? (size_t)bpp / 8 : (rgb16 ? 2 : (size_t)comp);  // per pixel in the file
    size_t pos = 18 + (size_t)idlen;
    if (pos > f.size()) return ROSY_TEX_ERR_DECODE;
    // palette
    std::vector<uint8_t> pal;
    auto rgb555 = [](const uint8_t* s, uint8_t* o) {
        const uint32_t px = s[0] | (s[1] << 8);
        o[0] = (uint8_t)((((px >> 10) & 31u) * 255u) / 31u);
        o[1] = (uint8_t)((((px >> 5) & 31u) * 255u) / 31u);
        o[2] = (uint8_t)(((px & 31u) * 255u) / 31u);
    };
    if (cmap) {
        if (pal_len == 0) return ROSY_TEX_ERR_DECODE;
        pos += pal_start;
        const size_t eb = rgb16 ? 2 : (size_t)comp;
        if (pos > f.size() || pal_len * eb > f.size() - pos) return ROSY_TEX_ERR_DECODE;
        pal.resize(pal_len * comp);
        if (rgb16)
            for (size_t i = 0; i < pal_len; i++) rgb555(&f[pos + 2 * i], &pal[3 * i]);
        else
            memcpy(pal.data(), &f[pos], pal_len * comp);
        pos += pal_len * eb;
    }
    // the payload must be able to hold the image before anything of its size is allocated
    const size_t avail = f.size() - pos;
    if (rle ? n > avail * 128 : n > avail / src_bytes) return ROSY_TEX_ERR_DECODE;
    std::vector<uint8_t> px;
    auto read_pixel = [&](const uint8_t* s, uint8_t* o) {
        if (cmap) {
            size_t idx = bpp == 8 ? s[0] : (size_t)(s[0] | (s[1] << 8));
            if (idx >= pal_len) idx = 0;
            memcpy(o, &pal[idx * comp], (size_t)comp);
        } else if (rgb16) {
            rgb555(s, o);
        } else if (comp == 4) {
            memcpy(o, s, 4);  // fixed sizes: the FBX route feeds 24- / 32-bit TGA only and a run-time-sized memcpy is a call
        } else if (comp == 3) {
            memcpy(o, s, 3);
        } else {
            memcpy(o, s, (size_t)comp);
        }
    };
    const bool plain = !cmap && !rgb16 && (comp == 3 || comp == 4);
    if (!rle && plain) {
        // uncompressed true colour: straight from the file to RGBA8, one row at a time
        W = w;
        H = h;
        rgba.resize(n * 4);
        const bool top_first = desc & 0x20;
        for (uint32_t y = 0; y < h; y++) {
            const uint8_t* s = &f[pos + (size_t)(top_first ? y : h - 1 - y) * w * comp];
            uint8_t* o = &rgba[(size_t)y * w * 4];
            if (comp == 4) {
                for (uint32_t x = 0; x < w; x++) {
                    uint32_t v;
                    memcpy(&v, s + 4 * (size_t)x, 4);
                    v = (v & 0xFF00FF00u) | ((v >> 16) & 0xFFu) | ((v & 0xFFu) << 16);  // BGRA -> RGBA
                    memcpy(o + 4 * (size_t)x, &v, 4);
                }
            } else {
                for (uint32_t x = 0; x < w; x++) {
                    o[4 * (size_t)x] = s[3 * (size_t)x + 2];
                    o[4 * (size_t)x + 1] = s[3 * (size_t)x + 1];
                    o[4 * (size_t)x + 2] = s[3 * (size_t)x];
                    o[4 * (size_t)x + 3] = 255;
                }
            }
        }
        return ROSY_TEX_OK;
    }
    px.resize(n * comp);
    if (!rle) {
        for (size_t i = 0; i < n; i++) read_pixel(&f[pos + i * src_bytes], &px[i * comp]);
    } else {
        size_t o = 0;
        while (o < n) {
            if (pos >= f.size()) return ROSY_TEX_ERR_DECODE;
            const int hd = f[pos++];
            size_t cnt = (size_t)(hd & 127) + 1;
            if (cnt > n - o) cnt = n - o;  // a packet may run past the last pixel (stb stops there too)
            if (hd & 128) {
                if (src_bytes > f.size() - pos) return ROSY_TEX_ERR_DECODE;
                uint8_t one[4];
                read_pixel(&f[pos], one);
                for (size_t k = 0; k < cnt; k++) memcpy(&px[(o + k) * comp], one, (size_t)comp);
                pos += src_bytes;
            } else {
                if (cnt * src_bytes > f.size() - pos) return ROSY_TEX_ERR_DECODE;
                for (size_t k = 0; k < cnt; k++) read_pixel(&f[pos + k * src_bytes], &px[(o + k) * comp]);
                pos += cnt * src_bytes;
            }
            o += cnt;
        }
    }
    W = w;
    H = h;
    rgba.resize(n * 4);
    const bool top = desc & 0x20;
    const bool swap_rb = comp >= 3 && !rgb16;  // stored BGR(A); 5-5-5 pixels were expanded in RGB order
    for (uint32_t y = 0; y < h; y++) {
        const size_t sy = top ? y : h - 1 - y;
        const uint8_t* s = &px[sy * w * comp];
        uint8_t* o = &rgba[(size_t)y * w * 4];
        if (plain && comp == 4) {
            for (uint32_t x = 0; x < w; x++) {
                uint32_t v;
                memcpy(&v, s + 4 * (size_t)x, 4);
                v = (v & 0xFF00FF00u) | ((v >> 16) & 0xFFu) | ((v & 0xFFu) << 16);
                memcpy(o + 4 * (size_t)x, &v, 4);
            }
            continue;
        }
        for (uint32_t x = 0; x < w; x++, s += comp, o += 4) {
            if (comp <= 2) {
                o[0] = o[1] = o[2] = s[0];
                o[3] = comp == 2 ? s[1] : 255;
            } else {
                o[0] = swap_rb ? s[2] : s[0];
                o[1] = s[1];
                o[2] = swap_rb ? s[0] : s[2];
                o[3] = comp == 4 ? s[3] : 255;
            }
        }
    }
    return ROSY_TEX_OK;
}

// ---------------------------------------------------------------- binary PNM (P5 / P6, maxval 255)
static int decode_pnm(const std::vector<uint8_t>& f, std::vector<uint8_t>& rgba, uint32_t& W, uint32_t& H) {
    size_t pos = 2;
    uint32_t vals[3];
    for (int k = 0; k < 3; k++) {
        for (;;) {
            while (pos < f.size() && (f[pos] == ' ' || f[pos] == '\n' || f[pos] == '\r' || f[pos] == '\t')) pos++;
            if (pos < f.size() && f[pos] == '#')
                while (pos < f.size() && f[pos] != '\n') pos++;
            else
                break;
        }
        uint32_t v = 0;
        bool any = false;
        while (pos < f.size() && f[pos] >= '0' && f[pos] <= '9') {
            v = v * 10 + (uint32_t)(f[pos++] - '0');
            any = true;
        }
        if (!any) return ROSY_TEX_ERR_DECODE;
        vals[k] = v;
    }
    pos++;  // single whitespace after maxval
    const int ch = f[1] == '6' ? 3 : 1;
    if (vals[0] == 0 || vals[1] == 0 || vals[0] > 32768 || vals[1] > 32768 || pos > f.size()) return ROSY_TEX_ERR_DECODE;
    const size_t n = (size_t)vals[0] * vals[1];
    if (vals[2] != 255 || n > (f.size() - pos) / (size_t)ch) return ROSY_TEX_ERR_DECODE;
    W = vals[0];
    H = vals[1];
    rgba.resize(n * 4);
    for (size_t i = 0; i < n; i++) {
        const uint8_t* s = &f[pos + i * ch];
        rgba[4 * i] = s[0];
        rgba[4 * i + 1] = ch == 3 ? s[1] : s[0];
        rgba[4 * i + 2] = ch == 3 ? s[2] : s[0];
        rgba[4 * i + 3] = 255;
    }
    return ROSY_TEX_OK;
}

int decode_image_memory(const uint8_t* data, size_t len, const char* name_hint, std::vector<uint8_t>& rgba,
                        uint32_t& w, uint32_t& h, std::string& err) {
    std::vector<uint8_t> f(data, data + len);
    int rc = ROSY_TEX_ERR_DECODE;
    const std::string name = name_hint ? name_hint : "";
    auto ends_with = [&](const char* ext) {
        const size_t n = strlen(ext);
        if (name.size() < n) return false;
        for (size_t i = 0; i < n; i++)
            if (tolower((unsigned char)name[name.size() - n + i]) != ext[i]) return false;
        return true;
    };
    if (len >= 8 && f[0] == 0x89 && f[1] == 'P') {
        rc = decode_png(f, rgba, w, h);
    } else if (len >= 2 && f[0] == 'P' && (f[1] == '5' || f[1] == '6')) {
        rc = decode_pnm(f, rgba, w, h);
    } else if (len >= 3 && f[0] == 0xFF && f[1] == 0xD8) {
        rc = decode_jpeg(f, rgba, w, h, err);
        if (rc != ROSY_TEX_OK) return rc;
    } else if (ends_with(".rgba8")) {
        if (len >= 8) {
            memcpy(&w, &f[0], 4);
            memcpy(&h, &f[4], 4);
            if (w && h && (size_t)w * h * 4 + 8 == len) {
                rgba.assign(f.begin() + 8, f.end());
                rc = ROSY_TEX_OK;
            }
        }
    } else if (ends_with(".tga") || name.empty()) {
        rc = decode_tga(f, rgba, w, h);
    }
    if (rc != ROSY_TEX_OK) err = "unsupported or corrupt image '" + name + "'";
    return rc;
}

int load_image(const char* path, std::vector<uint8_t>& rgba, uint32_t& w, uint32_t& h, std::string& err) {
    std::vector<uint8_t> f;
    if (!read_file(path, f)) {
        err = std::string("cannot read '") + path + "'";
        return ROSY_TEX_ERR_IO;
    }
    return decode_image_memory(f.data(), f.size(), path, rgba, w, h, err);
}

}  // namespace rosy_tex
