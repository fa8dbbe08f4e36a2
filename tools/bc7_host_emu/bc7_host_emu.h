// bc7_host_emu.h - TEST TOOLING: lets rosy_b200/csrc/bc7_encode.cu's device functions compile with g++ and run
// on the host, one 4x4 block at a time (NT = 1), so that a change to the kernel's search can be checked against
// tests/golden/bc7_spec_v2.npz without a GPU (tests/test_bc7_host_emu.py).  Nothing in the product includes this.
// Intrinsics are restated from the PTX ISA / CUDA math API documentation; fp32 behaviour matches the device
// build's flags (-fmad=false, IEEE division and square root) when this is compiled with -ffp-contract=off.
#pragma once
#include <cuda_runtime.h>  // vector types (uint2, uint4, make_uint4), cudaError_t: host-side declarations only
#include <math.h>
#include <stdint.h>
#include <string.h>

static inline unsigned int __dp4a(unsigned int a, unsigned int b, unsigned int c) {
    for (int i = 0; i < 4; i++) c += ((a >> (8 * i)) & 255u) * ((b >> (8 * i)) & 255u);
    return c;
}
static inline unsigned int __byte_perm(unsigned int x, unsigned int y, unsigned int s) {
    const uint64_t v = ((uint64_t)y << 32) | x;
    unsigned int r = 0;
    for (int i = 0; i < 4; i++) r |= (unsigned int)((v >> (8 * ((s >> (4 * i)) & 7u))) & 255u) << (8 * i);
    return r;
}
static inline int __vimin3_s32(int a, int b, int c) { return a < b ? (a < c ? a : c) : (b < c ? b : c); }
static inline int __popc(unsigned int x) { return __builtin_popcount(x); }
static inline float __fmaf_rn(float a, float b, float c) { return fmaf(a, b, c); }
static inline float __uint_as_float(unsigned int u) { float f; memcpy(&f, &u, 4); return f; }
static inline unsigned int __float_as_uint(float f) { unsigned int u; memcpy(&u, &f, 4); return u; }
// round-toward-minus-infinity add: the double sum of two floats is exact for the magnitudes used here
static inline float __fadd_rd(float a, float b) {
    const double s = (double)a + (double)b;
    float r = (float)s;
    if ((double)r > s) r = nextafterf(r, -INFINITY);
    return r;
}
static inline int min(int a, int b) { return a < b ? a : b; }
static inline int max(int a, int b) { return a > b ? a : b; }
static inline unsigned int min(unsigned int a, unsigned int b) { return a < b ? a : b; }
static inline unsigned int max(unsigned int a, unsigned int b) { return a > b ? a : b; }
// constant-memory symbols are plain globals on the host
#define cudaMemcpyToSymbol(sym, src, n) (memcpy((void*)&(sym), (src), (n)), cudaSuccess)
