/* ref_stb.c - build recipe glue for oracle/_ref/libstb_ref.so (test infrastructure).
 * Compiles the REFERENCE's own vendored stb headers from /root/reference/src/libs/stb
 * (include path given by the Makefile; no reference source is copied into this repo):
 *   stb_dxt.h:617-626          stb_compress_bc4_block / stb_compress_bc5_block  (rows A11/A12)
 *   stb_image.h                stbi_load_from_memory  (Surface::load stand-in, row A3)
 *   stb_image_resize2.h:508    stbir_resize, sRGB-aware box filter (independent check C4)
 */
#define STB_DXT_IMPLEMENTATION
#include "stb_dxt.h"
#define STB_IMAGE_IMPLEMENTATION
#include "stb_image.h"
#define STB_IMAGE_RESIZE_IMPLEMENTATION
#include "stb_image_resize2.h"

void ref_bc4_block(unsigned char* dst, const unsigned char* src16) { stb_compress_bc4_block(dst, src16); }
void ref_bc5_block(unsigned char* dst, const unsigned char* rg32) { stb_compress_bc5_block(dst, rg32); }

/* returns malloc'd RGBA8 (caller frees with ref_free) */
unsigned char* ref_image_load(const unsigned char* data, int len, int* w, int* h) {
    int comp;
    return stbi_load_from_memory(data, len, w, h, &comp, 4);
}
void ref_free(void* p) { free(p); }

/* sRGB-correct, alpha-weighted box downsample to half size (one level) */
int ref_resize_srgb_box(const unsigned char* src, int w, int h, unsigned char* dst, int ow, int oh) {
    return stbir_resize(src, w, h, 0, dst, ow, oh, 0, STBIR_RGBA, STBIR_TYPE_UINT8_SRGB,
                        STBIR_EDGE_CLAMP, STBIR_FILTER_BOX) != 0;
}
