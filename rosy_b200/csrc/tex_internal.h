// tex_internal.h - shared between the kernels and the C-ABI host code of librosy_tex_cuda.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/rosy_tex_cuda.h"

namespace rosy_tex {

// One mip level as the kernels see it.  The chain buffer holds RGBA8 levels back to back,
// level 0 first (offsets in pixels); blocks of all levels are emitted back to back, largest
// level first (dds.hpp:551-565 consumer contract).
struct Level {
    uint32_t w, h;        // extent in pixels
    uint32_t bw, bh;      // extent in 4x4 blocks
    uint64_t px_offset;   // first pixel of this level in the chain buffer
    uint64_t blk_offset;  // first block of this level in the block stream
};

struct Chain {
    int levels;
    uint32_t pad;
    uint64_t total_px;
    uint64_t total_blocks;
    Level lv[ROSY_TEX_MAX_LEVELS];
};

Chain make_chain(uint32_t w, uint32_t h);

// Search effort of the BC7 encoder ("spec v1").  Mirrors nvtt::Quality_*.
struct Bc7Tier {
    int n1, n3, n0, n2, n7;  // partitions refined in modes 1,3,0,2,7 (0 = mode off)
    int rot5;                // rotations tried in mode 5 (1 or 4)
    int rot4;                // rotations tried in mode 4 (0 = off), both index selectors
    int ls_rounds;           // least-squares refinement rounds per fit
    int pb_search;           // 1: evaluate every p-bit combination
};
Bc7Tier bc7_tier(int quality);

// ---- kernel launchers (each returns the number of kernels it launched, or -1 on error) ----
// mip_kernels.cu
// enc_format / d_blocks: ROSY_TEX_FORMAT_BC5 / _BC4 and the texture's block stream let the fused kernel encode levels
// 0 .. mips_fused_encode_levels() - 1 itself (they are then NOT written to d_chain); -1 / nullptr: mips only
int launch_mips(const Chain& c, bool normal_map, int normalize, const uint8_t* d_rgba,
                uint8_t* d_chain, void* d_scratch, cudaStream_t s, int enc_format = -1, uint8_t* d_blocks = nullptr);
int mips_fused_encode_levels(const Chain& c, int format);
size_t mips_scratch_bytes(const Chain& c);
// several textures, one fused launch per image kind (ragged extents fall back to per-level launches)
struct MipBatchItem {
    Chain chain;
    bool normal_map;
    int normalize;
    const uint8_t* d_rgba;
    uint8_t* d_chain;
    void* d_scratch;  // mips_scratch_bytes(chain), first 256 bytes zero
    int enc_format = -1;          // see launch_mips
    uint8_t* d_blocks = nullptr;
};
size_t mips_batch_desc_bytes(int n);
// mip chain over the library's own fixed scratch array (no device allocation anywhere on the path); returns a
// rosy_tex_status, *launches = kernels enqueued, *err = static message text
size_t mips_static_scratch_need(const Chain& c);
int mips_static_launch(const Chain& c, bool normal_map, int normalize, const uint8_t* d_rgba, uint8_t* d_chain,
                       cudaStream_t s, int* launches, const char** err);
int launch_mips_batch(int n, const MipBatchItem* items, void* d_desc, cudaStream_t s);
// bc45_encode.cu
int launch_bc45(const Chain& c, int first_level, int num_levels, bool bc5, const uint8_t* d_level0,
                const uint8_t* d_chain, uint8_t* d_blocks, cudaStream_t s);
struct Bc45BatchItem {
    Chain chain;
    bool bc5;
    int first_level;          // first level this launch encodes (levels below it came out of the fused mip kernel)
    const uint8_t* d_level0;  // level 0 pixels when they are not in the chain buffer (colour maps), else nullptr
    const uint8_t* d_chain;
    uint8_t* d_blocks;
};
size_t bc45_batch_desc_bytes(int n);
int launch_bc45_batch(int n, const Bc45BatchItem* items, void* d_desc, cudaStream_t s);
// bc7_encode.cu
int launch_bc7(const Chain& c, int first_level, int num_levels, int quality,
               const uint8_t* d_level0, const uint8_t* d_chain, uint8_t* d_blocks, cudaStream_t s);
// several textures, one launch per quality tier present
struct Bc7BatchItem {
    Chain chain;
    int quality;
    const uint8_t* d_level0;  // level 0 pixels when they are not in the chain buffer (colour maps), else nullptr
    const uint8_t* d_chain;
    uint8_t* d_blocks;
};
size_t bc7_batch_desc_bytes(int n);
int launch_bc7_batch(int n, const Bc7BatchItem* items, void* d_desc, cudaStream_t s);
int bc7_upload_tables();   // per device, once
int mips_upload_tables();  // per device, once (also probes the shared-window layout the fused mip kernel needs)
// a fused mip launch on the current device refused to run since the last call (its output is invalid): clears the
// flag and moves the device to the per-level kernels, so running the texture again succeeds
bool mips_take_layout_fault();
long long mips_check_faults();  // ROSY_TEX_CHECKED builds: out-of-bounds stores counted so far on the current device; -1 otherwise

}  // namespace rosy_tex
