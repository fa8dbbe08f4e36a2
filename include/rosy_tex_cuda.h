/* rosy_tex_cuda.h - C ABI of the B200-native replacement for the Rosy Packager texture stage.
 *
 * Drop-in boundary.  The reference links NVTT directly; its texture stage is two
 * C++ free functions declared at /root/reference/src/Packager/Packager.h:8-9 and defined at
 * /root/reference/src/Packager/Packager.cpp:216-261 (generate_srgb_texture) and :263-308
 * (generate_normal_map_texture).  Those two functions keep their signatures and become thin
 * wrappers over the first two entry points below (INTEGRATION.md shows the patch).  Everything
 * else here exists for tests, measurement and the multi-GPU work queue.
 *
 * Plain C: no C++ types, no exceptions across the boundary, no torch types.  Every function is
 * thread-safe.  There is NO CPU fallback: without a usable sm_100 device every compute entry
 * point returns ROSY_TEX_ERR_NO_DEVICE.
 */
#ifndef ROSY_TEX_CUDA_H
#define ROSY_TEX_CUDA_H

#include <stddef.h>
#include <stdint.h>

#if defined(_WIN32)
#define ROSY_TEX_API __declspec(dllexport)
#else
#define ROSY_TEX_API __attribute__((visibility("default")))
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* Replaces the `false` returns of nvtt::Surface::load / Context::outputHeader / Context::compress
 * (Packager.cpp:222-226, :237-241, :245-249); the wrapper maps != 0 to rosy::result::error. */
typedef enum rosy_tex_status {
    ROSY_TEX_OK = 0,
    ROSY_TEX_ERR_IO = 1,        /* cannot read the image / cannot write the .dds */
    ROSY_TEX_ERR_DECODE = 2,    /* unsupported or corrupt image file */
    ROSY_TEX_ERR_ARG = 3,       /* null pointer, zero extent, buffer too small, bad enum */
    ROSY_TEX_ERR_CUDA = 4,      /* a CUDA call failed; see rosy_tex_last_error() */
    ROSY_TEX_ERR_NO_DEVICE = 5, /* no sm_100 device: there is no CPU fallback */
    ROSY_TEX_ERR_NOMEM = 6
} rosy_tex_status;

/* image_type of the .rsy asset, /root/reference/src/Asset/Asset.h:109-112.  Routing follows
 * src/Packager/Gltf.cpp:225,252,280 and Fbx.cpp:444,453,462: everything except the normal map
 * goes through the sRGB function. */
typedef enum rosy_tex_kind {
    ROSY_TEX_KIND_COLOR = 0,
    ROSY_TEX_KIND_NORMAL_MAP = 1,
    ROSY_TEX_KIND_METALLIC_ROUGHNESS = 2,
    ROSY_TEX_KIND_MIXMAP = 3
} rosy_tex_kind;

/* nvtt::Format.  The reference only ever uses Format_BC7 (Packager.cpp:229, :276) and the Engine
 * binds BC7 regardless of the header (src/Engine/Graphics.cpp:3631-3647), so BC7 is the default
 * for every kind; BC5 / BC4 are opt-in. */
typedef enum rosy_tex_format {
    ROSY_TEX_FORMAT_BC7 = 0, /* DXGI_FORMAT_BC7_UNORM (98) */
    ROSY_TEX_FORMAT_BC5 = 1, /* DXGI_FORMAT_BC5_UNORM (83), channels R,G */
    ROSY_TEX_FORMAT_BC4 = 2  /* DXGI_FORMAT_BC4_UNORM (80), channel R   */
} rosy_tex_format;

/* nvtt::Quality_Fastest / Quality_Normal / Quality_Production.  The reference uses Normal
 * (Packager.cpp:277; NVTT default at :228-229). */
typedef enum rosy_tex_quality {
    ROSY_TEX_QUALITY_FAST = 0,
    ROSY_TEX_QUALITY_NORMAL = 1,
    ROSY_TEX_QUALITY_HIGH = 2
} rosy_tex_quality;

/* How nvtt::Surface::normalizeNormalMap (Packager.cpp:295) is read. */
typedef enum rosy_tex_normalize {
    ROSY_TEX_NORMALIZE_OFF = 0,   /* literal NVTT-2 reading: no-op on an unflagged surface */
    ROSY_TEX_NORMALIZE_UNPACK = 1 /* n = 2v-1, n /= |n| (0 if |n| = 0), v = (n+1)/2; default */
} rosy_tex_normalize;

#define ROSY_TEX_MAX_DEVICES 16
#define ROSY_TEX_DDS_HEADER_BYTES 148 /* 'DDS ' + DDS_HEADER (124) + DDS_HEADER_DXT10 (20) */
#define ROSY_TEX_MAX_LEVELS 16

/* Process-wide options.  Zero-initialise, set struct_size, change what you need.
 * Environment overrides, read by rosy_tex_init (so the Packager CLI stays flag-free):
 *   ROSY_TEX_DEVICES=0,1,..   ROSY_TEX_QUALITY=fast|normal|high
 *   ROSY_TEX_FORMAT_NORMAL=bc7|bc5   ROSY_TEX_FORMAT_COLOR=bc7   ROSY_TEX_FORMAT_MR=bc7|bc5|bc4
 *   ROSY_TEX_NORMALIZE=unpack|off
 *   ROSY_TEX_LINEAR_DATA=1   ROSY_TEX_SRGB_FLAG=1   (opt-in corrections; both change
 *   bytes versus the reference and are therefore off by default)
 *   ROSY_TEX_IO_THREADS=n     decode / write threads of the file queue (default: host cores, max 32)
 *   ROSY_TEX_STAGING_MB=n     pinned staging budget of the file queue (default 4096)
 *   ROSY_TEX_QUEUE_BATCH=0    work queue: one texture per pass whatever its size (default 1: a backlog of
 *                             textures of at most 512x512 pixels is taken up to 64 at a time / 4 Mpix and goes
 *                             through one set of kernel launches; the bytes are the same either way) */
typedef struct rosy_tex_config {
    uint32_t struct_size;
    int32_t num_devices;                   /* 0: every visible sm_100 device */
    int32_t devices[ROSY_TEX_MAX_DEVICES]; /* CUDA ordinals */
    int32_t quality;                       /* rosy_tex_quality, default NORMAL */
    int32_t format_for_kind[4];            /* rosy_tex_format per rosy_tex_kind, default BC7 */
    int32_t normalize;                     /* rosy_tex_normalize, default UNPACK */
    int32_t streams_per_device;            /* work-queue depth per device, default 2 */
    /* -- fields below were added later; struct_size tells which ones the caller knows -- */
    int32_t linear_data_maps;              /* 1: metallic-roughness and mixmap images are data, not colour: plain box
                                              filter on the stored values, no toLinearFromSrgb / toSrgb round trip
                                              (the reference sends them through the sRGB function, Gltf.cpp:252,
                                              Fbx.cpp:453).  Default 0 = like the reference. */
    int32_t srgb_flag;                     /* 1: OutputOptions::setSrgbFlag(true): colour maps are written as
                                              DXGI_FORMAT_BC7_UNORM_SRGB (99).  The reference never sets it
                                              (Packager.cpp:282) and writes 98.  Default 0. */
} rosy_tex_config;

/* One texture.  format / quality / normalize < 0 mean "use the configured default". */
typedef struct rosy_tex_desc {
    uint32_t width;
    uint32_t height;
    int32_t kind;      /* rosy_tex_kind */
    int32_t format;    /* rosy_tex_format or -1 */
    int32_t quality;   /* rosy_tex_quality or -1 */
    int32_t normalize; /* rosy_tex_normalize or -1 */
} rosy_tex_desc;

/* ---- lifetime.  init is optional: the first compute call initialises with defaults. ---- */
ROSY_TEX_API int rosy_tex_init(const rosy_tex_config* config /* may be NULL */);
ROSY_TEX_API void rosy_tex_shutdown(void);
ROSY_TEX_API int rosy_tex_device_count(void); /* sm_100 devices in use (0 if none) */
/* thread-local, never NULL; valid until the next call on this thread */
ROSY_TEX_API const char* rosy_tex_last_error(void);
ROSY_TEX_API const char* rosy_tex_status_string(int status);

/* ---- 1:1 with the reference --------------------------------------------------------------
 * Read the image at image_path (PNG / JPEG baseline + progressive / TGA / PNM / raw .rgba8), build the full mip chain, block
 * compress every level and write <image_path with its extension replaced by ".dds"> with a
 * DDS + DX10 header.  Return 0 on success.  The path is borrowed.  The file is written under a
 * temporary name and renamed, so a failed call leaves no partial .dds behind.
 *   rosy_tex_generate_srgb_texture        == Packager.cpp:216  generate_srgb_texture
 *   rosy_tex_generate_normal_map_texture  == Packager.cpp:263  generate_normal_map_texture */
ROSY_TEX_API int rosy_tex_generate_srgb_texture(const char* image_path);
ROSY_TEX_API int rosy_tex_generate_normal_map_texture(const char* image_path);
/* same, with the asset's image_type so that per-kind format overrides apply, and an explicit
 * output path (NULL: derive as above) */
ROSY_TEX_API int rosy_tex_generate_texture_file(const char* image_path, int kind,
                                                const char* dds_path);

/* nvtt::Surface::load stand-in (Packager.cpp:222, :269), exposed so the decoder can be tested on
 * its own: decodes PNG / JPEG / TGA / PNM / .rgba8 into a malloc'd RGBA8 buffer (free with
 * rosy_tex_free_image).  Host only; needs no device. */
ROSY_TEX_API int rosy_tex_load_image(const char* image_path, uint8_t** rgba, uint32_t* width,
                                     uint32_t* height);
ROSY_TEX_API void rosy_tex_free_image(uint8_t* rgba);

/* ---- geometry helpers (nvtt::Surface::countMipmaps, Packager.cpp:235; dds.hpp:527-565) ---- */
ROSY_TEX_API int rosy_tex_count_mipmaps(uint32_t width, uint32_t height);
ROSY_TEX_API void rosy_tex_level_extent(uint32_t width, uint32_t height, int level, uint32_t* w,
                                        uint32_t* h);
/* bytes of the complete .dds (header + every level) for this descriptor */
ROSY_TEX_API size_t rosy_tex_dds_size(const rosy_tex_desc* desc);
/* bytes of the RGBA8 mip chain, level 0 included */
ROSY_TEX_API size_t rosy_tex_chain_size(uint32_t width, uint32_t height);
/* writes the 148-byte header (Context::outputHeader, Packager.cpp:237 / :286) */
ROSY_TEX_API int rosy_tex_dds_header(const rosy_tex_desc* desc, uint8_t* out148);

/* ---- memory to memory (host buffers; H2D, kernels, D2H inside) ---------------------------
 * rgba: width*height*4 bytes, tightly packed, already decoded.  out receives the whole .dds
 * image (header + levels, largest first).  The caller owns both buffers; nothing is retained. */
ROSY_TEX_API int rosy_tex_encode_rgba8(const rosy_tex_desc* desc, const uint8_t* rgba, void* out,
                                       size_t out_capacity, size_t* out_len);
/* the RGBA8 mip chain exactly as the encoder sees it (what Context::compress quantises to),
 * level 0 first; for the <= 1 LSB parity check */
ROSY_TEX_API int rosy_tex_mips_rgba8(const rosy_tex_desc* desc, const uint8_t* rgba, uint8_t* out,
                                     size_t out_capacity, size_t* out_len);
/* block-compress ONE already built level (no mips, no header); out = raw blocks, row major */
ROSY_TEX_API int rosy_tex_compress_level_rgba8(const rosy_tex_desc* desc, const uint8_t* rgba,
                                               void* out, size_t out_capacity, size_t* out_len);

/* ---- asynchronous per-texture work queue over every configured device ------
 * No collectives: textures are independent.  submit returns at once; results are complete after
 * wait_all, which returns the first non-zero status of the batch (0 if all succeeded).  For the
 * memory variant the buffers must stay valid until wait_all returns; *out_len is written by then.
 * Pinned host buffers (rosy_tex_alloc_host) make the copies asynchronous. */
ROSY_TEX_API int rosy_tex_submit(const char* image_path, int kind);
ROSY_TEX_API int rosy_tex_submit_rgba8(const rosy_tex_desc* desc, const uint8_t* rgba, void* out,
                                       size_t out_capacity, size_t* out_len);
ROSY_TEX_API int rosy_tex_wait_all(void);
ROSY_TEX_API void* rosy_tex_alloc_host(size_t bytes); /* cudaHostAlloc, portable */
ROSY_TEX_API void rosy_tex_free_host(void* p);

/* ---- device-resident entry points (measurement, callers that already hold device memory) ---
 * All pointers are 16-byte aligned device pointers on the CURRENT CUDA device; `stream` is a cudaStream_t
 * (NULL: legacy default stream).  Nothing synchronises; kernels are only enqueued.
 *   d_chain  : rosy_tex_chain_size() bytes.  Receives levels 1.. (and level 0 for normal maps,
 *              whose level 0 is re-normalised); for colour maps level 0 is read from d_rgba.
 *              Exception: rosy_tex_encode_device / _batch_device with BOTH stages and a BC5 / BC4
 *              target on extents that are multiples of 64: levels 0..4 are block-compressed inside
 *              the mip kernel and never written, and only the channels the format stores are
 *              meaningful in levels 5.. - treat d_chain as scratch there (rosy_tex_mips_device
 *              always writes the complete chain).
 *   d_scratch: rosy_tex_device_scratch_size() bytes; its first 256 bytes must be zero before the
 *              first use (the kernels leave them zero again).
 *   d_blocks : rosy_tex_dds_size() - 148 bytes, every level's blocks, largest level first. */
ROSY_TEX_API size_t rosy_tex_device_scratch_size(const rosy_tex_desc* desc);
ROSY_TEX_API int rosy_tex_mips_device(const rosy_tex_desc* desc, const void* d_rgba, void* d_chain,
                                      void* d_scratch, void* stream);
ROSY_TEX_API int rosy_tex_blocks_device(const rosy_tex_desc* desc, const void* d_rgba,
                                        const void* d_chain, void* d_blocks, void* stream);
/* both of the above: the whole hot path with inputs resident in HBM */
ROSY_TEX_API int rosy_tex_encode_device(const rosy_tex_desc* desc, const void* d_rgba,
                                        void* d_chain, void* d_scratch, void* d_blocks,
                                        void* stream);

/* ---- the mip chain without any allocation ----------------------------------------------------
 * No counterpart in the reference (NVTT allocates whatever it likes inside Surface::buildNextMipmap).
 *
 * Same result as rosy_tex_mips_device - the complete RGBA8 chain in d_chain, d_rgba / d_chain / stream exactly as
 * described above - but the caller brings no scratch buffer and the library asks nobody for one: the fp32 hand-over
 * buffers between levels live in a fixed array compiled into the library.  Neither this call nor anything it reaches
 * uses cudaMalloc, cudaFree, cudaHostAlloc or any other allocator.
 *
 * Who needs that: code embedded in a host that has already handed the whole of device memory to its own arena (a
 * game or render engine, a caching tensor allocator) and forbids plug-ins to go to the driver for more - there
 * cudaMalloc simply fails; and code on a frame or streaming thread that must never stall, because cudaMalloc and
 * cudaFree can make the whole device wait, whatever streams other threads are running.
 *
 * What it costs and where it stops:
 *   - the array is ROSY_TEX_STATIC_SCRATCH_BYTES (5 MiB + 256) of device memory on EVERY GPU the process touches
 *     with this library loaded, held until the process exits, used or not;
 *   - a texture is accepted when levels 1 and 2 of its chain, as fp32 RGBA, fit in it: that is every texture of
 *     at most 1 048 576 pixels (1024 x 1024, 2048 x 512, 1000 x 700 ...).  Anything larger is refused with
 *     ROSY_TEX_ERR_ARG and a message that says so - use rosy_tex_mips_device with your own scratch for those;
 *   - always the one-kernel-per-level path, never the fused TMA kernel (whose set-up probe allocates).  On extents
 *     where rosy_tex_mips_device takes that path too (a side that is not a multiple of 64) the bytes are identical;
 *     where it takes the fused kernel, each result is within 1 of the CPU oracle in every 8-bit channel (the bar
 *     of the test suite), so the two can differ from one another by a unit or two in a few texels;
 *   - the FIRST call on each device does a one-time set-up for that device (looks up the array's device address, loads
 *     one 1 KB colour table, both synchronous).  Make that first call during start-up, outside any stream capture;
 *     later calls only enqueue kernels on `stream`;
 *   - all calls on a device share the array.  Nothing in it needs to be zero between calls (each level is written in
 *     full before it is read, so stale contents of an earlier call are harmless), but two calls whose kernels can be
 *     in flight on the same device at the same time - different streams, different threads - overwrite each other's
 *     data and produce garbage.  Keep them on one stream, or order them with events. */
#define ROSY_TEX_STATIC_SCRATCH_BYTES (5u * 1024u * 1024u + 256u)
ROSY_TEX_API int rosy_tex_mips_device_static(const rosy_tex_desc* desc, const void* d_rgba, void* d_chain,
                                             void* stream);

/* A batch of device-resident textures ("batched small textures per launch").
 * The mip chains of every texture whose extents are multiples of 64 are built by ONE kernel
 * launch per image kind (colour / normal map), so that small textures run at HBM speed instead
 * of launch latency; the block encoders then run per texture on the same stream.  Pointers as in
 * rosy_tex_encode_device.  d_batch_scratch: rosy_tex_batch_scratch_size(count) bytes of device
 * memory, not to be reused before the enqueued work has run.  do_mips / do_blocks select the
 * stages (both non-zero: the whole hot path). */
typedef struct rosy_tex_batch_item {
    rosy_tex_desc desc;
    const void* d_rgba;
    void* d_chain;
    void* d_scratch;
    void* d_blocks;
} rosy_tex_batch_item;
ROSY_TEX_API size_t rosy_tex_batch_scratch_size(int count);
ROSY_TEX_API int rosy_tex_encode_batch_device(const rosy_tex_batch_item* items, int count,
                                              void* d_batch_scratch, void* stream, int do_mips,
                                              int do_blocks);

/* Debug aid: a library built with `make -C rosy_b200/csrc checked` tests every global store of the warp-specialised mip
 * (+ BC5 / BC4) kernel against the extent of its buffer and counts the violations; this returns the count over the
 * devices in use.  The product build returns -1. */
ROSY_TEX_API long long rosy_tex_debug_faults(void);

/* number of kernels this library has launched since load (all threads, all devices) */
ROSY_TEX_API uint64_t rosy_tex_kernel_launches(void);

/* Health of the work queue's device workers (no reference counterpart).  A worker whose DEVICE fails (not merely one
 * texture) retires and its textures are finished by the others, so rosy_tex_wait_all can return 0 on a degraded
 * pool; this is how to see it.  *live_workers: workers still serving the queue (-1: the queue has not started yet);
 * *requeued: textures handed to another worker so far.  Either pointer may be NULL. */
ROSY_TEX_API void rosy_tex_queue_status(int* live_workers, int* requeued);

#ifdef __cplusplus
}
#endif
#endif /* ROSY_TEX_CUDA_H */
