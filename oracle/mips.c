/* mips.c - CPU ORACLE (test infrastructure, see rosy_oracle.h): the two mip chains of the
 * Packager texture stage, restated literally from the reference call sequence.
 *
 *   colour : /root/reference/src/Packager/Packager.cpp:243-259
 *   normal : /root/reference/src/Packager/Packager.cpp:292-306
 *
 * The nvtt::Surface calls are restated from NVTT's documented semantics;
 * NVTT itself is closed source => PARITY UNPINNED for these functions.
 * The surface is float32 RGBA and is only quantised to 8 bits when a level is handed to
 * Context::compress (F11); every level is derived from the previous float level.
 */
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#include "rosy_oracle.h"

typedef struct {
    int w, h;
    float* px; /* interleaved RGBA */
} surface;

/* ---- rows of a pass shared by host threads (the timed CPU baseline uses every core, SURVEY 8d:
 * "OpenMP over blocks and rows").  Every pass is element-wise or row-local, so the result does not
 * depend on the thread count.  t_threads is set by the public entry points. */
static __thread int t_threads = 1;

typedef void (*row_fn)(void* ctx, int y0, int y1);
typedef struct {
    row_fn fn;
    void* ctx;
    int y0, y1;
} row_task;

static void* row_worker(void* arg) {
    row_task* t = (row_task*)arg;
    t->fn(t->ctx, t->y0, t->y1);
    return NULL;
}

static void for_rows(row_fn fn, void* ctx, int rows, size_t elems_per_row) {
    int n = t_threads;
    if ((size_t)rows * elems_per_row < 32768) n = 1; /* small levels: not worth a thread */
    if (n > rows) n = rows;
    if (n > 256) n = 256;
    if (n <= 1) {
        fn(ctx, 0, rows);
        return;
    }
    pthread_t tid[256];
    row_task task[256];
    char started[256];
    for (int i = 0; i < n; i++) {
        task[i].fn = fn;
        task[i].ctx = ctx;
        task[i].y0 = (int)((long long)rows * i / n);
        task[i].y1 = (int)((long long)rows * (i + 1) / n);
        started[i] = i > 0 && pthread_create(&tid[i], NULL, row_worker, &task[i]) == 0;
    }
    row_worker(&task[0]);
    for (int i = 1; i < n; i++) {
        /* no thread to be had (EAGAIN): the range is done here instead - a row range is never skipped */
        if (started[i])
            pthread_join(tid[i], NULL);
        else
            row_worker(&task[i]);
    }
}

/* nvtt::Surface::countMipmaps, Packager.cpp:235 / :284 */
int oracle_count_mipmaps(int w, int h) {
    int n = 1;
    while (w > 1 || h > 1) {
        w = w / 2 > 1 ? w / 2 : 1;
        h = h / 2 > 1 ? h / 2 : 1;
        n++;
    }
    return n;
}

void oracle_level_extent(int w, int h, int level, int* lw, int* lh) {
    for (int i = 0; i < level; i++) {
        w = w / 2 > 1 ? w / 2 : 1;
        h = h / 2 > 1 ? h / 2 : 1;
    }
    *lw = w;
    *lh = h;
}

size_t oracle_chain_pixels(int w, int h) {
    size_t n = 0;
    int levels = oracle_count_mipmaps(w, h);
    for (int l = 0; l < levels; l++) {
        int lw, lh;
        oracle_level_extent(w, h, l, &lw, &lh);
        n += (size_t)lw * lh;
    }
    return n;
}

size_t oracle_chain_blocks(int w, int h) {
    size_t n = 0;
    int levels = oracle_count_mipmaps(w, h);
    for (int l = 0; l < levels; l++) {
        int lw, lh;
        oracle_level_extent(w, h, l, &lw, &lh);
        n += (size_t)((lw + 3) / 4) * ((lh + 3) / 4);
    }
    return n;
}

/* nvtt::Surface::load, Packager.cpp:222 / :269 (row A3): 8-bit -> float = byte / 255 */
typedef struct {
    const uint8_t* rgba;
    float* px;
    int w;
} load_ctx;
static void load_rows(void* ctx, int y0, int y1) {
    load_ctx* c = (load_ctx*)ctx;
    for (size_t i = (size_t)y0 * c->w * 4; i < (size_t)y1 * c->w * 4; i++) c->px[i] = (float)c->rgba[i] / 255.0f;
}
static surface surface_from_rgba8(const uint8_t* rgba, int w, int h) {
    surface s;
    s.w = w;
    s.h = h;
    s.px = (float*)malloc(sizeof(float) * 4 * (size_t)w * h);
    load_ctx c = {rgba, s.px, w};
    for_rows(load_rows, &c, h, (size_t)w * 4);
    return s;
}

static float clamp01(float x) { return x < 0.0f ? 0.0f : (x > 1.0f ? 1.0f : x); }

/* What Context::compress does to the float surface before the block search (row A9):
 * round(clamp(x) * 255). */
typedef struct {
    const surface* s;
    uint8_t* out;
} quant_ctx;
static void quantise_rows(void* ctx, int y0, int y1) {
    quant_ctx* c = (quant_ctx*)ctx;
    const size_t n = (size_t)c->s->w * 4;
    for (size_t i = n * y0; i < n * y1; i++) c->out[i] = (uint8_t)floorf(clamp01(c->s->px[i]) * 255.0f + 0.5f);
}
static void quantise(const surface* s, uint8_t* out) {
    quant_ctx c = {s, out};
    for_rows(quantise_rows, &c, s->h, (size_t)s->w * 4);
}

/* Surface::toLinearFromSrgb, Packager.cpp:254 (row A5); RGB only */
static void to_linear_rows(void* ctx, int y0, int y1) {
    surface* s = (surface*)ctx;
    for (size_t i = (size_t)y0 * s->w; i < (size_t)y1 * s->w; i++)
        for (int c = 0; c < 3; c++) {
            float v = s->px[4 * i + c];
            v = v <= 0.04045f ? v / 12.92f : powf((v + 0.055f) / 1.055f, 2.4f);
            s->px[4 * i + c] = clamp01(v);
        }
}
static void to_linear_from_srgb(surface* s) { for_rows(to_linear_rows, s, s->h, (size_t)s->w * 64); }

/* Surface::toSrgb, Packager.cpp:258 (row A5); NVTT-2 used the literal exponent 0.41666f */
static void to_srgb_rows(void* ctx, int y0, int y1) {
    surface* s = (surface*)ctx;
    for (size_t i = (size_t)y0 * s->w; i < (size_t)y1 * s->w; i++)
        for (int c = 0; c < 3; c++) {
            float v = s->px[4 * i + c];
            v = v <= 0.0031308f ? 12.92f * v : 1.055f * powf(v, 0.41666f) - 0.055f;
            s->px[4 * i + c] = clamp01(v);
        }
}
static void to_srgb(surface* s) { for_rows(to_srgb_rows, s, s->h, (size_t)s->w * 64); }

/* Surface::premultiplyAlpha, Packager.cpp:255 (row A6) */
static void premultiply_rows(void* ctx, int y0, int y1) {
    surface* s = (surface*)ctx;
    for (size_t i = (size_t)y0 * s->w; i < (size_t)y1 * s->w; i++) {
        float a = s->px[4 * i + 3];
        for (int c = 0; c < 3; c++) s->px[4 * i + c] *= a;
    }
}
static void premultiply_alpha(surface* s) { for_rows(premultiply_rows, s, s->h, (size_t)s->w * 4); }

/* Surface::demultiplyAlpha(epsilon = 1e-12), Packager.cpp:257 (row A6) */
static void demultiply_rows(void* ctx, int y0, int y1) {
    surface* s = (surface*)ctx;
    for (size_t i = (size_t)y0 * s->w; i < (size_t)y1 * s->w; i++) {
        float a = s->px[4 * i + 3];
        if (fabsf(a) > 1e-12f) {
            float ia = 1.0f / a;
            for (int c = 0; c < 3; c++) s->px[4 * i + c] *= ia;
        }
    }
}
static void demultiply_alpha(surface* s) { for_rows(demultiply_rows, s, s->h, (size_t)s->w * 4); }

/* One separable pass of the box filter along x.  Even extent: mean of the pair.  Odd extent
 * n = 2k+1 -> k outputs with NVTT-2's polyphase box weights ((k-i), k, (i+1)) / n over inputs
 * 2i, 2i+1, 2i+2 (row A7, H7).  Extent 1 stays 1 (identity). */
typedef struct {
    const float* src;
    float* dst;
    int w, h, o;
} box_ctx;
static void box_rows_range(void* ctx, int y0, int y1) {
    const box_ctx* b = (const box_ctx*)ctx;
    const float* src = b->src;
    float* dst = b->dst;
    const int w = b->w, ow = b->o;
    for (int y = y0; y < y1; y++) {
        const float* s = src + (size_t)4 * w * y;
        float* d = dst + (size_t)4 * ow * y;
        if (w == 1) {
            memcpy(d, s, sizeof(float) * 4);
        } else if ((w & 1) == 0) {
            for (int x = 0; x < ow; x++)
                for (int c = 0; c < 4; c++) d[4 * x + c] = (s[8 * x + c] + s[8 * x + 4 + c]) * 0.5f;
        } else {
            const float scale = 1.0f / (float)w;
            for (int x = 0; x < ow; x++) {
                float w0 = (float)(ow - x), w1 = (float)ow, w2 = (float)(1 + x);
                for (int c = 0; c < 4; c++)
                    d[4 * x + c] =
                        scale * (w0 * s[8 * x + c] + w1 * s[8 * x + 4 + c] + w2 * s[8 * x + 8 + c]);
            }
        }
    }
}
static void box_rows(const float* src, int w, int h, float* dst, int ow) {
    box_ctx b = {src, dst, w, h, ow};
    for_rows(box_rows_range, &b, h, (size_t)w * 4);
}

static void box_cols_range(void* ctx, int y0, int y1) {
    const box_ctx* b = (const box_ctx*)ctx;
    const float* src = b->src;
    float* dst = b->dst;
    const int w = b->w, h = b->h, oh = b->o;
    for (int y = y0; y < y1; y++) {
        float* d = dst + (size_t)4 * w * y;
        if ((h & 1) == 0) {
            const float* s0 = src + (size_t)4 * w * (2 * y);
            const float* s1 = s0 + (size_t)4 * w;
            for (int i = 0; i < 4 * w; i++) d[i] = (s0[i] + s1[i]) * 0.5f;
        } else {
            const float scale = 1.0f / (float)h;
            float w0 = (float)(oh - y), w1 = (float)oh, w2 = (float)(1 + y);
            const float* s0 = src + (size_t)4 * w * (2 * y);
            const float* s1 = s0 + (size_t)4 * w;
            const float* s2 = s1 + (size_t)4 * w;
            for (int i = 0; i < 4 * w; i++) d[i] = scale * (w0 * s0[i] + w1 * s1[i] + w2 * s2[i]);
        }
    }
}
static void box_cols(const float* src, int w, int h, float* dst, int oh) {
    if (h == 1) {
        memcpy(dst, src, sizeof(float) * 4 * (size_t)w);
        return;
    }
    box_ctx b = {src, dst, w, h, oh};
    for_rows(box_cols_range, &b, oh, (size_t)w * 8);
}

/* Surface::buildNextMipmap(MipmapFilter_Box), Packager.cpp:256 / :305 (row A7).
 * New extent max(1,w/2) x max(1,h/2); all four channels are filtered. */
static void build_next_mipmap_box(surface* s) {
    int ow = s->w / 2 > 1 ? s->w / 2 : 1;
    int oh = s->h / 2 > 1 ? s->h / 2 : 1;
    float* tmp = (float*)malloc(sizeof(float) * 4 * (size_t)ow * s->h);
    float* out = (float*)malloc(sizeof(float) * 4 * (size_t)ow * oh);
    box_rows(s->px, s->w, s->h, tmp, ow);
    box_cols(tmp, ow, s->h, out, oh);
    free(tmp);
    free(s->px);
    s->px = out;
    s->w = ow;
    s->h = oh;
}

/* Surface::normalizeNormalMap on a copy, Packager.cpp:294-295 (row A8, H4):
 * unpack 2v-1, normalizeSafe(v, 0, 0), repack; alpha untouched. */
static void normalize_rows(void* ctx, int y0, int y1) {
    surface* s = (surface*)ctx;
    for (size_t i = (size_t)y0 * s->w; i < (size_t)y1 * s->w; i++) {
        float x = 2.0f * s->px[4 * i + 0] - 1.0f;
        float y = 2.0f * s->px[4 * i + 1] - 1.0f;
        float z = 2.0f * s->px[4 * i + 2] - 1.0f;
        float l = sqrtf(x * x + y * y + z * z);
        if (l > 0.0f) {
            float il = 1.0f / l;
            x *= il;
            y *= il;
            z *= il;
        } else {
            x = y = z = 0.0f;
        }
        s->px[4 * i + 0] = x * 0.5f + 0.5f;
        s->px[4 * i + 1] = y * 0.5f + 0.5f;
        s->px[4 * i + 2] = z * 0.5f + 0.5f;
    }
}
static void normalize_normal_map(surface* s) { for_rows(normalize_rows, s, s->h, (size_t)s->w * 16); }

void oracle_mip_chain_srgb(const uint8_t* rgba, int w, int h, uint8_t* out) {
    oracle_mip_chain_srgb_mt(rgba, w, h, out, 1);
}

void oracle_mip_chain_srgb_mt(const uint8_t* rgba, int w, int h, uint8_t* out, int threads) {
    t_threads = threads <= 0 ? oracle_max_threads() : threads;
    surface image = surface_from_rgba8(rgba, w, h); /* Packager.cpp:221-222 */
    const int n = oracle_count_mipmaps(w, h);        /* :235 */
    for (int mip = 0; mip < n; mip++) {              /* :243 */
        quantise(&image, out);                       /* :245 compress() sees this level */
        out += (size_t)4 * image.w * image.h;
        if (mip == n - 1) break;                     /* :250-253 */
        to_linear_from_srgb(&image);                 /* :254 */
        premultiply_alpha(&image);                   /* :255 */
        build_next_mipmap_box(&image);               /* :256 */
        demultiply_alpha(&image);                    /* :257 */
        to_srgb(&image);                             /* :258 */
    }
    free(image.px);
}

void oracle_mip_chain_normal(const uint8_t* rgba, int w, int h, int normalize_mode, uint8_t* out) {
    oracle_mip_chain_normal_mt(rgba, w, h, normalize_mode, out, 1);
}

void oracle_mip_chain_normal_mt(const uint8_t* rgba, int w, int h, int normalize_mode, uint8_t* out, int threads) {
    t_threads = threads <= 0 ? oracle_max_threads() : threads;
    surface image = surface_from_rgba8(rgba, w, h); /* Packager.cpp:268-269 */
    const int n = oracle_count_mipmaps(w, h);        /* :284 */
    for (int mip = 0; mip < n; mip++) {              /* :292 */
        surface temp = image;                        /* :294 copy */
        temp.px = (float*)malloc(sizeof(float) * 4 * (size_t)image.w * image.h);
        memcpy(temp.px, image.px, sizeof(float) * 4 * (size_t)image.w * image.h);
        if (normalize_mode == ORACLE_NORMALIZE_UNPACK) normalize_normal_map(&temp); /* :295 */
        quantise(&temp, out);                        /* :296 compress() sees the copy */
        out += (size_t)4 * image.w * image.h;
        free(temp.px);
        if (mip == n - 1) break;                     /* :301-304 */
        build_next_mipmap_box(&image);               /* :305 un-normalised chain */
    }
    free(image.px);
}
