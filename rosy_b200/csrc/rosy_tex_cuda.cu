// rosy_tex_cuda.cu - C-ABI implementation (include/rosy_tex_cuda.h): device selection, buffer
// pools, the DDS container writer, the single-texture entry points that replace the NVTT calls of
// /root/reference/src/Packager/Packager.cpp:216-308, and the per-texture multi-GPU work queue.
//
// No CPU fallback exists anywhere in this file: without an sm_100 device the compute entry
// points return ROSY_TEX_ERR_NO_DEVICE.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <condition_variable>
#include <deque>
#include <map>
#include <algorithm>
#include <mutex>
#include <new>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "tex_internal.h"

#include <nvtx3/nvToolsExt.h>

namespace rosy_tex {

// One NVTX range per texture (or per batch of small textures) on the worker thread: lets a timeline show
// copy / kernel overlap per slot.  Costs nothing when no tool is attached.
struct NvtxRange {
    NvtxRange(const char* what, uint32_t w, uint32_t h, int count) {
        char name[96];
        snprintf(name, sizeof name, "%s %ux%u x%d", what, w, h, count);
        nvtxRangePushA(name);
    }
    ~NvtxRange() { nvtxRangePop(); }
};

int load_image(const char* path, std::vector<uint8_t>& rgba, uint32_t& w, uint32_t& h, std::string& err);
int mips_upload_tables();

namespace {

thread_local std::string t_error = "";
std::atomic<uint64_t> g_launches{0};

int fail(int status, const std::string& msg) {
    t_error = msg;
    return status;
}
int cuda_fail(cudaError_t e, const char* what) {
    return fail(ROSY_TEX_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
// No exception may cross the C ABI or escape a worker thread (std::bad_alloc from a std::vector sized by a
// hostile image header would otherwise reach std::terminate through a bare std::thread).
template <class F>
int guarded(F&& f) {
    try {
        return f();
    } catch (const std::bad_alloc&) {
        return fail(ROSY_TEX_ERR_NOMEM, "out of host memory");
    } catch (const std::exception& e) {
        return fail(ROSY_TEX_ERR_NOMEM, std::string("internal error: ") + e.what());
    } catch (...) {
        return fail(ROSY_TEX_ERR_NOMEM, "internal error");
    }
}
#define CU_TRY(call)                                          \
    do {                                                      \
        cudaError_t e_ = (call);                              \
        if (e_ != cudaSuccess) return cuda_fail(e_, #call);   \
    } while (0)

// One in-flight texture: a stream plus grow-only device buffers.
struct Slot {
    int device = 0;
    cudaStream_t stream = nullptr;
    uint8_t *d_in = nullptr, *d_chain = nullptr, *d_scratch = nullptr, *d_blocks = nullptr, *d_desc = nullptr;
    size_t cap_in = 0, cap_chain = 0, cap_scratch = 0, cap_blocks = 0, cap_desc = 0;
    bool busy = false;
    bool ticket_dirty = false;  // a pass on this slot failed after its launches: the fused mip kernel's ticket may be non-zero
};

// A texture for the device workers: host pixels in, DDS bytes out.
struct Job {
    rosy_tex_desc desc;
    const uint8_t* rgba = nullptr;
    void* out = nullptr;
    size_t cap = 0;
    size_t* len = nullptr;
    // file jobs (rosy_tex_submit): rgba / out are pinned buffers of the pool, the .dds is written afterwards
    bool from_file = false;
    std::string dds_path;
    size_t staged = 0;  // pinned bytes this job holds
    int attempts = 0;   // times the job went back to the queue after a device fault
};

// Host-side stage of a file job, run by the I/O threads: decode before the device, write after it.
struct IoTask {
    enum Type { DECODE, WRITE } type = DECODE;
    std::string path, dds_path;
    int kind = 0;
    void *pin_in = nullptr, *pin_out = nullptr;  // WRITE: buffers to give back to the pool
    size_t len = 0, staged = 0;
};

struct PinBuf {
    void* p = nullptr;
    size_t cap = 0;
    bool used = false;
};

struct State {
    std::mutex mu;
    std::condition_variable cv;
    bool inited = false;
    rosy_tex_config cfg;
    std::vector<int> devices;
    std::deque<Slot> slots;  // deque: references stay valid when the queue adds its own slots
    std::vector<int> tables_ready;  // device ordinals with constant tables uploaded
    // work queue
    // ordered largest texture first, submission order among equals: the pop below is O(log n) however long the backlog
    std::map<std::pair<uint64_t, uint64_t>, Job> queue;  // key: (~pixels, sequence number)
    uint64_t queue_seq = 0;
    std::vector<std::thread> workers;
    bool queue_batching = true;  // ROSY_TEX_QUEUE_BATCH=0: one texture per pass whatever its size
    int running = 0;
    int live_workers = 0;  // device workers that have not retired after a device fault
    int requeued = 0;      // textures handed to another worker after a device fault (diagnostic)
    int pending = 0;  // submitted textures that are not completely done (decode, device, write)
    // file pipeline: decode / write threads, decoded images waiting for a device slot, pinned staging pool
    std::deque<IoTask> io_queue;
    std::vector<std::thread> io_workers;
    int decoded_inflight = 0, decoded_limit = 4;
    size_t staged_bytes = 0, staged_limit = (size_t)4 << 30;  // pinned bytes of decoded images not yet written out
    std::vector<PinBuf> pinned;
    int first_error = 0;
    std::string first_error_msg;
    bool stop = false;
};
State& st() {
    // Leaked on purpose: worker threads may still be parked on the condition variable when the
    // process exits without rosy_tex_shutdown(); destroying joinable std::threads would terminate().
    static State* s = new State;
    return *s;
}

int env_enum(const char* name, const char* const* names, int n, int dflt) {
    const char* e = getenv(name);
    if (!e) return dflt;
    for (int i = 0; i < n; i++)
        if (!strcasecmp(e, names[i])) return i;
    return dflt;
}

int ensure_tables_locked(State& s, int device) {
    for (int d : s.tables_ready)
        if (d == device) return ROSY_TEX_OK;
    if (bc7_upload_tables() != 0 || mips_upload_tables() != 0)
        return fail(ROSY_TEX_ERR_CUDA, std::string("table upload: ") + cudaGetErrorString(cudaGetLastError()));
    s.tables_ready.push_back(device);
    return ROSY_TEX_OK;
}

int init_locked(State& s, const rosy_tex_config* user) {
    if (s.inited) return ROSY_TEX_OK;
    rosy_tex_config c;
    memset(&c, 0, sizeof c);
    c.quality = ROSY_TEX_QUALITY_NORMAL;
    c.normalize = ROSY_TEX_NORMALIZE_UNPACK;
    c.streams_per_device = 2;
    if (user) {
        size_t n = user->struct_size ? user->struct_size : sizeof c;
        if (n > sizeof c) n = sizeof c;
        rosy_tex_config u;
        memset(&u, 0, sizeof u);
        memcpy(&u, user, n);
        c.num_devices = u.num_devices;
        memcpy(c.devices, u.devices, sizeof c.devices);
        c.quality = u.quality;
        memcpy(c.format_for_kind, u.format_for_kind, sizeof c.format_for_kind);
        c.normalize = u.normalize;
        if (u.streams_per_device > 0) c.streams_per_device = u.streams_per_device;
        c.linear_data_maps = u.linear_data_maps != 0;
        c.srgb_flag = u.srgb_flag != 0;
    }
    if (const char* e = getenv("ROSY_TEX_LINEAR_DATA")) c.linear_data_maps = atoi(e) != 0;
    if (const char* e = getenv("ROSY_TEX_SRGB_FLAG")) c.srgb_flag = atoi(e) != 0;
    if (const char* e = getenv("ROSY_TEX_QUEUE_BATCH")) s.queue_batching = atoi(e) != 0;
    static const char* const qn[] = {"fast", "normal", "high"};
    static const char* const fn[] = {"bc7", "bc5", "bc4"};
    static const char* const nn[] = {"off", "unpack"};
    c.quality = env_enum("ROSY_TEX_QUALITY", qn, 3, c.quality);
    c.format_for_kind[ROSY_TEX_KIND_COLOR] = env_enum("ROSY_TEX_FORMAT_COLOR", fn, 3, c.format_for_kind[0]);
    c.format_for_kind[ROSY_TEX_KIND_NORMAL_MAP] = env_enum("ROSY_TEX_FORMAT_NORMAL", fn, 3, c.format_for_kind[1]);
    c.format_for_kind[ROSY_TEX_KIND_METALLIC_ROUGHNESS] = env_enum("ROSY_TEX_FORMAT_MR", fn, 3, c.format_for_kind[2]);
    c.format_for_kind[ROSY_TEX_KIND_MIXMAP] = env_enum("ROSY_TEX_FORMAT_MIXMAP", fn, 3, c.format_for_kind[3]);
    c.normalize = env_enum("ROSY_TEX_NORMALIZE", nn, 2, c.normalize);
    if (const char* e = getenv("ROSY_TEX_DEVICES")) {
        c.num_devices = 0;
        for (const char* p = e; *p && c.num_devices < ROSY_TEX_MAX_DEVICES;) {
            c.devices[c.num_devices++] = atoi(p);
            while (*p && *p != ',') p++;
            if (*p == ',') p++;
        }
    }
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
        cudaGetLastError();
        return fail(ROSY_TEX_ERR_NO_DEVICE, "no CUDA device is visible (there is no CPU fallback)");
    }
    std::vector<int> want;
    if (c.num_devices > 0)
        want.assign(c.devices, c.devices + c.num_devices);
    else
        for (int d = 0; d < count; d++) want.push_back(d);
    s.devices.clear();
    for (int d : want) {
        int major = 0;
        if (d < 0 || d >= count) continue;
        if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, d) == cudaSuccess && major == 10)
            s.devices.push_back(d);
    }
    if (s.devices.empty())
        return fail(ROSY_TEX_ERR_NO_DEVICE, "no sm_100 (Blackwell B200) device among the visible devices (there is no CPU fallback)");
    int prev = 0;
    cudaGetDevice(&prev);
    for (int d : s.devices) {
        CU_TRY(cudaSetDevice(d));
        int rc = ensure_tables_locked(s, d);
        if (rc) return rc;
        for (int k = 0; k < c.streams_per_device; k++) {
            Slot sl;
            sl.device = d;
            CU_TRY(cudaStreamCreateWithFlags(&sl.stream, cudaStreamNonBlocking));
            s.slots.push_back(sl);
        }
    }
    cudaSetDevice(prev);
    s.cfg = c;
    s.inited = true;
    return ROSY_TEX_OK;
}

int ensure_init() {
    State& s = st();
    std::lock_guard<std::mutex> g(s.mu);
    return init_locked(s, nullptr);
}

struct Resolved {
    uint32_t w, h;
    bool normal_map;  // the chain is a plain box filter of the stored values (+ optional re-normalisation)
    int format, quality, normalize;
    bool srgb_header;  // dxgiFormat 99 instead of 98
};

int resolve(const rosy_tex_desc* d, Resolved& r) {
    if (!d || d->width == 0 || d->height == 0) return fail(ROSY_TEX_ERR_ARG, "null descriptor or zero extent");
    if (d->width > 32768 || d->height > 32768) return fail(ROSY_TEX_ERR_ARG, "extent above 32768");
    if (d->kind < 0 || d->kind > 3) return fail(ROSY_TEX_ERR_ARG, "bad kind");
    const State& s = st();
    r.w = d->width;
    r.h = d->height;
    r.normal_map = d->kind == ROSY_TEX_KIND_NORMAL_MAP;
    r.format = d->format >= 0 ? d->format : (s.inited ? s.cfg.format_for_kind[d->kind] : ROSY_TEX_FORMAT_BC7);
    r.quality = d->quality >= 0 ? d->quality : (s.inited ? s.cfg.quality : ROSY_TEX_QUALITY_NORMAL);
    r.normalize = d->normalize >= 0 ? d->normalize : (s.inited ? s.cfg.normalize : ROSY_TEX_NORMALIZE_UNPACK);
    if (r.format > 2 || r.quality > 2 || r.normalize > 1) return fail(ROSY_TEX_ERR_ARG, "bad format / quality / normalize");
    const bool data_kind = d->kind == ROSY_TEX_KIND_METALLIC_ROUGHNESS || d->kind == ROSY_TEX_KIND_MIXMAP;
    if (data_kind && s.inited && s.cfg.linear_data_maps) {
        // opt-in (N4): data maps skip the sRGB round trip: the normal-map chain without the re-normalisation
        r.normal_map = true;
        r.normalize = ROSY_TEX_NORMALIZE_OFF;
    }
    r.srgb_header = s.inited && s.cfg.srgb_flag && r.format == ROSY_TEX_FORMAT_BC7 && !r.normal_map;
    return ROSY_TEX_OK;
}

size_t block_bytes(int format) { return format == ROSY_TEX_FORMAT_BC4 ? 8 : 16; }

void put32(uint8_t* p, uint32_t v) {
    p[0] = (uint8_t)v;
    p[1] = (uint8_t)(v >> 8);
    p[2] = (uint8_t)(v >> 16);
    p[3] = (uint8_t)(v >> 24);
}

// Context::outputHeader (Packager.cpp:237 / :286): 'DDS ' + DDS_HEADER + DDS_HEADER_DXT10, laid out
// as the Engine's reader expects (src/libs/dds_formats.hpp:251-311, src/libs/dds.hpp:415-440).
void write_dds_header(const Resolved& r, int levels, uint8_t* h) {
    memset(h, 0, ROSY_TEX_DDS_HEADER_BYTES);
    memcpy(h, "DDS ", 4);
    put32(h + 4, 124);                                                // dwSize
    put32(h + 8, 0x1 | 0x2 | 0x4 | 0x1000 | 0x20000 | 0x80000);       // CAPS|HEIGHT|WIDTH|PIXELFORMAT|MIPMAPCOUNT|LINEARSIZE
    put32(h + 12, r.h);
    put32(h + 16, r.w);
    put32(h + 20, (uint32_t)(((r.w + 3) / 4) * ((r.h + 3) / 4) * block_bytes(r.format)));
    put32(h + 28, (uint32_t)levels);
    put32(h + 76, 32);                                                // ddspf.dwSize
    put32(h + 80, 0x4);                                               // DDPF_FOURCC
    memcpy(h + 84, "DX10", 4);
    put32(h + 108, 0x1000 | 0x400000 | 0x8);                          // TEXTURE | MIPMAP | COMPLEX
    put32(h + 128, r.format == ROSY_TEX_FORMAT_BC7 ? (r.srgb_header ? 99u : 98u) : (r.format == ROSY_TEX_FORMAT_BC5 ? 83u : 80u));
    put32(h + 132, 3);                                                // D3D10_RESOURCE_DIMENSION_TEXTURE2D
    put32(h + 140, 1);                                                // arraySize
}

int grow(uint8_t*& p, size_t& cap, size_t need) {
    if (need <= cap) return ROSY_TEX_OK;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    cudaError_t e = cudaMalloc(&p, need);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(ROSY_TEX_ERR_NOMEM, std::string("cudaMalloc: ") + cudaGetErrorString(e));
    }
    cap = need;
    return ROSY_TEX_OK;
}

int encode_on_stream(const Chain& c, const Resolved& r, const uint8_t* d_rgba, uint8_t* d_chain, void* d_scratch,
                     uint8_t* d_blocks, cudaStream_t stream, bool do_mips, bool do_blocks) {
    // BC5 / BC4 with both stages: the fused mip kernel encodes levels 0..4 from the tiles it holds (they never reach
    // d_chain); the stand-alone encoder takes the remaining small levels
    const int fused_levels = do_mips && do_blocks ? mips_fused_encode_levels(c, r.format) : 0;
    if (do_mips && c.levels > 1 || (do_mips && r.normal_map)) {
        int n = launch_mips(c, r.normal_map, r.normalize, d_rgba, d_chain, d_scratch, stream, fused_levels ? r.format : -1,
                            fused_levels ? d_blocks : nullptr);
        if (n < 0) return cuda_fail(cudaGetLastError(), "mip kernels");
        g_launches += (uint64_t)n;
    }
    if (do_blocks) {
        const uint8_t* level0 = r.normal_map ? nullptr : d_rgba;
        int n;
        if (r.format == ROSY_TEX_FORMAT_BC7)
            n = launch_bc7(c, 0, c.levels, r.quality, level0, d_chain, d_blocks, stream);
        else if (fused_levels >= c.levels)
            n = 0;
        else
            n = launch_bc45(c, fused_levels, c.levels - fused_levels, r.format == ROSY_TEX_FORMAT_BC5, level0, d_chain, d_blocks, stream);
        if (n < 0) return cuda_fail(cudaGetLastError(), "encoder kernel");
        g_launches += (uint64_t)n;
    }
    return ROSY_TEX_OK;
}

Slot* acquire_slot() {
    State& s = st();
    std::unique_lock<std::mutex> g(s.mu);
    for (;;) {
        for (Slot& sl : s.slots)
            if (!sl.busy) {
                sl.busy = true;
                return &sl;
            }
        s.cv.wait(g);
    }
}
void release_slot(Slot* sl) {
    State& s = st();
    {
        std::lock_guard<std::mutex> g(s.mu);
        sl->busy = false;
    }
    s.cv.notify_all();
}

enum What { WHAT_DDS, WHAT_MIPS, WHAT_LEVEL };

// host buffers in, host buffers out, on one slot
int run_host_impl(Slot* sl, const Resolved& r, What what, const uint8_t* rgba, uint8_t* out, size_t cap, size_t* len);
constexpr int RC_LAYOUT_RETRY = -100;  // internal: the fused mip kernel refused to run; the device now uses the per-level kernels
int run_host(Slot* sl, const Resolved& r, What what, const uint8_t* rgba, uint8_t* out, size_t cap, size_t* len) {
    int rc = run_host_impl(sl, r, what, rgba, out, cap, len);
    if (rc == RC_LAYOUT_RETRY) {  // once more, on the generic path (mips_take_layout_fault switched the device over)
        sl->ticket_dirty = true;
        rc = run_host_impl(sl, r, what, rgba, out, cap, len);
        if (rc == RC_LAYOUT_RETRY) rc = fail(ROSY_TEX_ERR_CUDA, "fused mip kernel met an unexpected shared-memory window layout");
    }
    if (rc && rc != ROSY_TEX_ERR_ARG) sl->ticket_dirty = true;
    return rc;
}
int run_host_impl(Slot* sl, const Resolved& r, What what, const uint8_t* rgba, uint8_t* out, size_t cap, size_t* len) {
    Chain c = make_chain(r.w, r.h);
    if (what == WHAT_LEVEL) {  // a single level, no mip chain
        c.levels = 1;
        c.total_px = (uint64_t)r.w * r.h;
        c.total_blocks = (uint64_t)c.lv[0].bw * c.lv[0].bh;
    }
    const size_t in_bytes = (size_t)r.w * r.h * 4, chain_bytes = (size_t)c.total_px * 4;
    const size_t blk_bytes = (size_t)c.total_blocks * block_bytes(r.format);
    const size_t need = what == WHAT_DDS ? ROSY_TEX_DDS_HEADER_BYTES + blk_bytes : (what == WHAT_MIPS ? chain_bytes : blk_bytes);
    if (len) *len = need;
    if (!rgba || !out || cap < need) return fail(ROSY_TEX_ERR_ARG, "null buffer or output capacity too small");
    CU_TRY(cudaSetDevice(sl->device));
    int rc;
    if ((rc = grow(sl->d_in, sl->cap_in, in_bytes))) return rc;
    if ((rc = grow(sl->d_chain, sl->cap_chain, chain_bytes))) return rc;
    if ((rc = grow(sl->d_blocks, sl->cap_blocks, blk_bytes))) return rc;
    const size_t sc = mips_scratch_bytes(c);
    if (sc > sl->cap_scratch || sl->ticket_dirty) {
        if ((rc = grow(sl->d_scratch, sl->cap_scratch, sc))) return rc;
        CU_TRY(cudaMemsetAsync(sl->d_scratch, 0, 256, sl->stream));  // the fused kernel's ticket
        sl->ticket_dirty = false;
    }
    CU_TRY(cudaMemcpyAsync(sl->d_in, rgba, in_bytes, cudaMemcpyHostToDevice, sl->stream));
    if (what == WHAT_LEVEL) {
        Resolved rl = r;
        rl.normal_map = false;  // the level is taken as is
        if ((rc = encode_on_stream(c, rl, sl->d_in, sl->d_chain, sl->d_scratch, sl->d_blocks, sl->stream, false, true)))
            return rc;
        CU_TRY(cudaMemcpyAsync(out, sl->d_blocks, blk_bytes, cudaMemcpyDeviceToHost, sl->stream));
    } else if (what == WHAT_MIPS) {
        if ((rc = encode_on_stream(c, r, sl->d_in, sl->d_chain, sl->d_scratch, sl->d_blocks, sl->stream, true, false)))
            return rc;
        if (r.normal_map) {
            CU_TRY(cudaMemcpyAsync(out, sl->d_chain, chain_bytes, cudaMemcpyDeviceToHost, sl->stream));
        } else {  // colour level 0 is the input itself (Packager.cpp:245 at mip 0)
            memcpy(out, rgba, in_bytes);
            if (chain_bytes > in_bytes)
                CU_TRY(cudaMemcpyAsync(out + in_bytes, sl->d_chain + in_bytes, chain_bytes - in_bytes,
                                       cudaMemcpyDeviceToHost, sl->stream));
        }
    } else {
        if ((rc = encode_on_stream(c, r, sl->d_in, sl->d_chain, sl->d_scratch, sl->d_blocks, sl->stream, true, true)))
            return rc;
        write_dds_header(r, c.levels, out);
        CU_TRY(cudaMemcpyAsync(out + ROSY_TEX_DDS_HEADER_BYTES, sl->d_blocks, blk_bytes, cudaMemcpyDeviceToHost, sl->stream));
    }
    CU_TRY(cudaStreamSynchronize(sl->stream));
    if (mips_take_layout_fault()) return RC_LAYOUT_RETRY;
    return ROSY_TEX_OK;
}

// Mip chains and blocks of several device-resident textures on one stream: one mip launch per image
// kind, one BC7 launch per quality tier, one BC5 and one BC4 launch.  d_desc: batch_desc_bytes(n).
size_t batch_desc_bytes(int count) {
    return count > 0 ? mips_batch_desc_bytes(count) + bc7_batch_desc_bytes(count) + bc45_batch_desc_bytes(count) : 0;
}
int encode_batch_on_stream(int count, const Resolved* rs, const MipBatchItem* mi, uint8_t* const* d_blocks, void* d_desc,
                           cudaStream_t s, bool do_mips, bool do_blocks) {
    if (!d_desc) return fail(ROSY_TEX_ERR_ARG, "null batch scratch");
    std::vector<int> fused_levels((size_t)count, 0);
    std::vector<MipBatchItem> mi_enc;
    if (do_mips && do_blocks) {
        mi_enc.assign(mi, mi + count);
        for (int i = 0; i < count; i++) {
            fused_levels[(size_t)i] = mips_fused_encode_levels(mi[i].chain, rs[i].format);
            if (fused_levels[(size_t)i]) {
                mi_enc[(size_t)i].enc_format = rs[i].format;
                mi_enc[(size_t)i].d_blocks = d_blocks[i];
            }
        }
        mi = mi_enc.data();
    }
    if (do_mips) {
        const int n = launch_mips_batch(count, mi, d_desc, s);
        if (n < 0) return cuda_fail(cudaGetLastError(), "batched mip kernels");
        g_launches += (uint64_t)n;
    }
    if (do_blocks) {
        std::vector<Bc7BatchItem> bi;
        std::vector<Bc45BatchItem> ai;
        for (int i = 0; i < count; i++) {
            const uint8_t* level0 = rs[i].normal_map ? nullptr : mi[i].d_rgba;
            if (rs[i].format == ROSY_TEX_FORMAT_BC7) {
                Bc7BatchItem b;
                b.chain = mi[i].chain;
                b.quality = rs[i].quality;
                b.d_level0 = level0;
                b.d_chain = mi[i].d_chain;
                b.d_blocks = d_blocks[i];
                bi.push_back(b);
            } else {
                Bc45BatchItem a;
                a.chain = mi[i].chain;
                a.bc5 = rs[i].format == ROSY_TEX_FORMAT_BC5;
                a.first_level = fused_levels[(size_t)i];
                a.d_level0 = level0;
                a.d_chain = mi[i].d_chain;
                a.d_blocks = d_blocks[i];
                ai.push_back(a);
            }
        }
        uint8_t* desc = static_cast<uint8_t*>(d_desc) + mips_batch_desc_bytes(count);
        if (!bi.empty()) {
            const int n = launch_bc7_batch((int)bi.size(), bi.data(), desc, s);
            if (n < 0) return cuda_fail(cudaGetLastError(), "batched BC7 kernel");
            g_launches += (uint64_t)n;
        }
        if (!ai.empty()) {
            const int n = launch_bc45_batch((int)ai.size(), ai.data(), desc + bc7_batch_desc_bytes(count), s);
            if (n < 0) return cuda_fail(cudaGetLastError(), "batched BC5 / BC4 kernel");
            g_launches += (uint64_t)n;
        }
    }
    return ROSY_TEX_OK;
}

// Several small textures on one slot (host buffers in, DDS bytes out): one copy in and one copy out
// per texture, ONE set of launches and ONE synchronisation for all of them.  Per-texture launch and
// synchronisation overhead is what limits small textures (256^2: 0.16 ms each against 0.02 ms of work).
struct HostJob {
    Resolved r;
    const uint8_t* rgba;
    uint8_t* out;
    size_t cap;
    size_t* len;
};
int run_host_batch(Slot* sl, std::vector<HostJob>& jobs) {
    const int n = (int)jobs.size();
    std::vector<MipBatchItem> mi((size_t)n);
    std::vector<Resolved> rs((size_t)n);
    std::vector<size_t> off_in((size_t)n), off_chain((size_t)n), off_scr((size_t)n), off_blk((size_t)n), blk_bytes((size_t)n);
    size_t tot_in = 0, tot_chain = 0, tot_scr = 0, tot_blk = 0;
    auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
    for (int i = 0; i < n; i++) {
        rs[i] = jobs[(size_t)i].r;
        const Chain c = make_chain(rs[i].w, rs[i].h);
        mi[i].chain = c;
        blk_bytes[i] = (size_t)c.total_blocks * block_bytes(rs[i].format);
        const size_t need = ROSY_TEX_DDS_HEADER_BYTES + blk_bytes[i];
        if (jobs[(size_t)i].len) *jobs[(size_t)i].len = need;
        if (!jobs[(size_t)i].rgba || !jobs[(size_t)i].out || jobs[(size_t)i].cap < need)
            return fail(ROSY_TEX_ERR_ARG, "null buffer or output capacity too small");
        off_in[i] = tot_in;
        tot_in += al((size_t)rs[i].w * rs[i].h * 4);
        off_chain[i] = tot_chain;
        tot_chain += al((size_t)c.total_px * 4);
        off_scr[i] = tot_scr;
        tot_scr += al(mips_scratch_bytes(c));
        off_blk[i] = tot_blk;
        tot_blk += al(blk_bytes[i]);
    }
    CU_TRY(cudaSetDevice(sl->device));
    int rc;
    if ((rc = grow(sl->d_in, sl->cap_in, tot_in))) return rc;
    if ((rc = grow(sl->d_chain, sl->cap_chain, tot_chain))) return rc;
    if ((rc = grow(sl->d_blocks, sl->cap_blocks, tot_blk))) return rc;
    if ((rc = grow(sl->d_scratch, sl->cap_scratch, tot_scr))) return rc;
    if ((rc = grow(sl->d_desc, sl->cap_desc, batch_desc_bytes(n)))) return rc;
    // the sub-buffers move from batch to batch: the tickets (first 256 bytes of each scratch) are zeroed every time
    CU_TRY(cudaMemsetAsync(sl->d_scratch, 0, tot_scr, sl->stream));
    std::vector<uint8_t*> d_blk((size_t)n);
    for (int i = 0; i < n; i++) {
        CU_TRY(cudaMemcpyAsync(sl->d_in + off_in[i], jobs[(size_t)i].rgba, (size_t)rs[i].w * rs[i].h * 4, cudaMemcpyHostToDevice, sl->stream));
        mi[i].normal_map = rs[i].normal_map;
        mi[i].normalize = rs[i].normalize;
        mi[i].d_rgba = sl->d_in + off_in[i];
        mi[i].d_chain = sl->d_chain + off_chain[i];
        mi[i].d_scratch = sl->d_scratch + off_scr[i];
        d_blk[i] = sl->d_blocks + off_blk[i];
    }
    if ((rc = encode_batch_on_stream(n, rs.data(), mi.data(), d_blk.data(), sl->d_desc, sl->stream, true, true))) return rc;
    for (int i = 0; i < n; i++) {
        write_dds_header(rs[i], mi[i].chain.levels, jobs[(size_t)i].out);
        CU_TRY(cudaMemcpyAsync(jobs[(size_t)i].out + ROSY_TEX_DDS_HEADER_BYTES, d_blk[i], blk_bytes[i], cudaMemcpyDeviceToHost, sl->stream));
    }
    sl->ticket_dirty = true;  // until the batch is known to have run to its end
    CU_TRY(cudaStreamSynchronize(sl->stream));
    // a refused fused launch: the caller (worker_main) runs the members one by one, which now take the per-level kernels
    if (mips_take_layout_fault()) return fail(ROSY_TEX_ERR_CUDA, "fused mip kernel met an unexpected shared-memory window layout");
    // the single-texture path relies on a zero ticket at the start of d_scratch: still true (tickets return to zero)
    sl->ticket_dirty = false;
    return ROSY_TEX_OK;
}

int run_host_any(const rosy_tex_desc* desc, What what, const uint8_t* rgba, void* out, size_t cap, size_t* len) {
    int rc = ensure_init();
    if (rc) return rc;
    Resolved r;
    if ((rc = resolve(desc, r))) return rc;
    int prev = 0;
    cudaGetDevice(&prev);
    Slot* sl = acquire_slot();
    rc = guarded([&] { return run_host(sl, r, what, rgba, static_cast<uint8_t*>(out), cap, len); });
    release_slot(sl);
    cudaSetDevice(prev);
    return rc;
}

std::string replace_extension(const std::string& path, const char* ext) {
    const size_t slash = path.find_last_of("/\\");
    const size_t dot = path.find_last_of('.');
    if (dot == std::string::npos || (slash != std::string::npos && dot < slash)) return path + ext;
    return path.substr(0, dot) + ext;
}

int write_dds_file(const std::string& dds_path, const void* data, size_t len);

// the whole of generate_srgb_texture / generate_normal_map_texture for one file, on one slot
int run_file(Slot* sl, const std::string& path, int kind, const std::string& dds_path_in) {
    std::vector<uint8_t> rgba;
    uint32_t w = 0, h = 0;
    std::string err;
    int rc = load_image(path.c_str(), rgba, w, h, err);  // nvtt::Surface::load, Packager.cpp:222
    if (rc) return fail(rc, err);
    rosy_tex_desc d = {w, h, kind, -1, -1, -1};
    Resolved r;
    if ((rc = resolve(&d, r))) return rc;
    std::vector<uint8_t> dds(ROSY_TEX_DDS_HEADER_BYTES + (size_t)make_chain(w, h).total_blocks * block_bytes(r.format));
    size_t len = 0;
    if ((rc = run_host(sl, r, WHAT_DDS, rgba.data(), dds.data(), dds.size(), &len))) return rc;
    const std::string dds_path = dds_path_in.empty() ? replace_extension(path, ".dds") : dds_path_in;  // :231
    return write_dds_file(dds_path, dds.data(), len);
}

// ---- pinned staging pool (file pipeline) ----
void* pin_acquire(State& s, size_t bytes) {
    {
        std::lock_guard<std::mutex> g(s.mu);
        PinBuf* best = nullptr;
        for (PinBuf& b : s.pinned)
            if (!b.used && b.cap >= bytes && (!best || b.cap < best->cap)) best = &b;
        if (best) {
            best->used = true;
            return best->p;
        }
    }
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    std::lock_guard<std::mutex> g(s.mu);
    PinBuf b;
    b.p = p;
    b.cap = bytes;
    b.used = true;
    s.pinned.push_back(b);
    return p;
}
void pin_release_locked(State& s, void* p) {
    for (PinBuf& b : s.pinned)
        if (b.p == p) b.used = false;
}

void note_error_locked(State& s, int rc) {
    if (rc && !s.first_error) {
        s.first_error = rc;
        s.first_error_msg = t_error;
    }
}

// device worker: one per queue slot.  Small textures (<= 512^2 pixels each) are taken from the queue
// several at a time - a fair share of the backlog, at most 64 textures / 4 Mpix - and go through one
// set of launches (run_host_batch); anything larger is one texture per pass.  Measured on one B200
// through submit_rgba8 (Mpix/s, batched vs one by one): 256^2 2282 vs 408, 512^2 2703 vs 1519,
// 1024^2 2786 vs 2997 - hence the member limit.
constexpr uint64_t BATCH_MEMBER_PX = 1u << 18, BATCH_TOTAL_PX = 4u << 20;
constexpr size_t BATCH_MAX = 64;
inline uint64_t job_px(const Job& j) { return (uint64_t)j.desc.width * j.desc.height; }
// call with s.mu held
inline void queue_push_locked(State& s, Job&& j) {
    const std::pair<uint64_t, uint64_t> key{~job_px(j), s.queue_seq++};
    s.queue.emplace(key, std::move(j));
}

// Pop order: largest texture first, so that the long jobs start early and the small ones fill
// the gaps at the end (LPT).  Small textures are then taken together as described above.
// Failure handling: after a ROSY_TEX_ERR_CUDA the worker probes its device (cudaDeviceSynchronize).
// A healthy device means the failure belongs to that texture (a refused launch, a tensor map that cannot be
// built): the error is reported on it and the worker carries on.  A probe that fails too means the context is
// gone (lost device, illegal address, ECC): the worker retires, frees its device buffers, says so on stderr, and
// its textures go back to the queue once to be finished by the remaining workers (other devices).  When no
// worker is left, or a texture fails twice, the error is reported by wait_all, which is the reference's
// behaviour: any failure aborts the asset (Gltf.cpp:225-229).  rosy_tex_queue_status() shows a degraded pool.
void worker_main(Slot* sl, int slot_index) {
    State& s = st();
    std::vector<Job> jobs;
    std::vector<HostJob> hj;
    std::vector<size_t> lens;
    std::vector<int> rcs;
    int fault_left = 0;
#ifdef ROSY_TEX_TEST_HOOKS  // librosy_tex_cuda_hooks.so only: ROSY_TEX_FAULT_SLOT=<slot index> makes that worker's first pass fail like a lost device
    if (const char* e = getenv("ROSY_TEX_FAULT_SLOT")) fault_left = atoi(e) == slot_index ? 1 : 0;
#endif
    for (;;) {
        jobs.clear();
        {
            std::unique_lock<std::mutex> g(s.mu);
            s.cv.wait(g, [&] { return s.stop || !s.queue.empty(); });
            if (s.stop && s.queue.empty()) return;
            const size_t live = s.live_workers > 0 ? (size_t)s.live_workers : 1;
            const size_t share = (s.queue.size() + live - 1) / live;
            jobs.push_back(std::move(s.queue.begin()->second));  // the largest texture waiting
            s.queue.erase(s.queue.begin());
            uint64_t px = job_px(jobs[0]);
            // the largest one is small: so is everything else in the queue
            if (s.queue_batching && px <= BATCH_MEMBER_PX)
                while (!s.queue.empty() && jobs.size() < share && jobs.size() < BATCH_MAX &&
                       px + job_px(s.queue.begin()->second) <= BATCH_TOTAL_PX) {
                    px += job_px(s.queue.begin()->second);
                    jobs.push_back(std::move(s.queue.begin()->second));
                    s.queue.erase(s.queue.begin());
                }
            s.running += (int)jobs.size();
        }
        const size_t n = jobs.size();
        lens.assign(n, 0);
        rcs.assign(n, 0);
        bool batched = false, injected = false;
        const int guard_rc = guarded([&] {
            if (fault_left > 0) {
                fault_left--;
                injected = true;
                for (size_t i = 0; i < n; i++) rcs[i] = fail(ROSY_TEX_ERR_CUDA, "injected device fault (ROSY_TEX_FAULT_SLOT)");
                return (int)ROSY_TEX_OK;
            }
            NvtxRange range(n > 1 ? "rosy_tex batch" : "rosy_tex texture", jobs[0].desc.width, jobs[0].desc.height, (int)n);
            if (n > 1) {
                hj.resize(n);
                int rc = 0;
                for (size_t i = 0; i < n && !rc; i++) {
                    rc = resolve(&jobs[i].desc, hj[i].r);
                    hj[i].rgba = jobs[i].rgba;
                    hj[i].out = static_cast<uint8_t*>(jobs[i].out);
                    hj[i].cap = jobs[i].cap;
                    hj[i].len = jobs[i].len ? jobs[i].len : &lens[i];
                }
                if (!rc) rc = run_host_batch(sl, hj);
                batched = rc == 0;  // a failing member: one by one below, so that the error lands on that texture alone
            }
            if (!batched)
                for (size_t i = 0; i < n; i++) {
                    Resolved r;
                    rcs[i] = resolve(&jobs[i].desc, r);
                    if (!rcs[i])
                        rcs[i] = run_host(sl, r, WHAT_DDS, jobs[i].rgba, static_cast<uint8_t*>(jobs[i].out), jobs[i].cap,
                                          jobs[i].len ? jobs[i].len : &lens[i]);
                    if (rcs[i] && n > 1) {  // keep the message of the first failing member
                        std::lock_guard<std::mutex> g(s.mu);
                        note_error_locked(s, rcs[i]);
                    }
                }
            return (int)ROSY_TEX_OK;
        });
        if (guard_rc)
            for (size_t i = 0; i < n; i++) rcs[i] = rcs[i] ? rcs[i] : guard_rc;
        bool retire = false;
        // is the device itself gone, or did only this texture fail?  (probed outside the lock: it waits for the device)
        bool cuda_error = false, device_lost = false;
        for (size_t i = 0; i < n; i++) cuda_error = cuda_error || rcs[i] == ROSY_TEX_ERR_CUDA;
        if (cuda_error) {
            device_lost = injected;
            if (!injected) {
                cudaSetDevice(sl->device);
                device_lost = cudaDeviceSynchronize() != cudaSuccess;
            }
            cudaGetLastError();
        }
        {
            std::lock_guard<std::mutex> g(s.mu);
            // a lost device: hand the textures to the other workers (once per texture) and retire
            if (device_lost && s.live_workers > 1) {
                retire = true;
                s.live_workers--;
            }
            for (size_t i = 0; i < n; i++) {
                Job& job = jobs[i];
                int rc = rcs[i];
                s.running--;
                if (retire && rc == ROSY_TEX_ERR_CUDA && job.attempts == 0) {
                    job.attempts++;
                    s.requeued++;
                    queue_push_locked(s, std::move(job));
                    continue;
                }
                note_error_locked(s, rc);
                if (job.from_file) {
                    s.decoded_inflight--;
                    if (rc) {
                        pin_release_locked(s, const_cast<uint8_t*>(job.rgba));
                        pin_release_locked(s, job.out);
                        s.staged_bytes -= job.staged;
                        s.pending--;
                    } else {  // the .dds is written by an I/O thread; the slot goes on with the next texture
                        IoTask t;
                        t.type = IoTask::WRITE;
                        t.dds_path = job.dds_path;
                        t.pin_in = const_cast<uint8_t*>(job.rgba);
                        t.pin_out = job.out;
                        t.len = lens[i];
                        t.staged = job.staged;
                        s.io_queue.push_front(std::move(t));
                    }
                } else {
                    s.pending--;
                }
            }
        }
        s.cv.notify_all();
        if (retire) {
            // give the device memory back (on a dead context these calls fail, which is fine) and make the loss visible
            for (uint8_t** p : {&sl->d_in, &sl->d_chain, &sl->d_scratch, &sl->d_blocks, &sl->d_desc}) {
                cudaFree(*p);
                *p = nullptr;
            }
            sl->cap_in = sl->cap_chain = sl->cap_scratch = sl->cap_blocks = sl->cap_desc = 0;
            cudaGetLastError();
            int left;
            {
                std::lock_guard<std::mutex> g(s.mu);
                left = s.live_workers;
            }
            fprintf(stderr, "rosy_tex: worker %d on device %d retired after a device fault; %d worker(s) left\n", slot_index,
                    sl->device, left);
            return;
        }
    }
}

// the device workers exist once the first texture is queued; call with s.mu held
void start_workers_locked(State& s) {
    if (!s.workers.empty()) return;
    // one worker thread per slot; the queue's slots are separate from the synchronous ones
    const size_t n = s.slots.size();
    for (size_t i = 0; i < n; i++) {
        Slot sl;
        sl.device = s.slots[i].device;
        sl.busy = true;  // owned by its worker
        cudaSetDevice(sl.device);
        cudaStreamCreateWithFlags(&sl.stream, cudaStreamNonBlocking);
        s.slots.push_back(sl);
    }
    s.live_workers = (int)n;
    for (size_t i = 0; i < n; i++) s.workers.emplace_back(worker_main, &s.slots[n + i], (int)i);
}

int write_dds_file(const std::string& dds_path, const void* data, size_t len) {
    const std::string tmp = dds_path + ".tmp";
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) return fail(ROSY_TEX_ERR_IO, "cannot create '" + tmp + "'");
    const bool ok = fwrite(data, 1, len, f) == len;
    if (fclose(f) != 0 || !ok || rename(tmp.c_str(), dds_path.c_str()) != 0) {
        remove(tmp.c_str());
        return fail(ROSY_TEX_ERR_IO, "cannot write '" + dds_path + "'");
    }
    return ROSY_TEX_OK;
}

// I/O worker: decodes image files into pinned buffers ahead of the device (nvtt::Surface::load,
// Packager.cpp:222 / :269, for many files at once) and writes finished .dds files behind it.
void io_main() {
    State& s = st();
    for (;;) {
        IoTask t;
        {
            std::unique_lock<std::mutex> g(s.mu);
            auto runnable = [&]() -> int {
                for (size_t i = 0; i < s.io_queue.size(); i++)
                    if (s.io_queue[i].type == IoTask::WRITE ||
                        (s.decoded_inflight < s.decoded_limit && s.staged_bytes < s.staged_limit))
                        return (int)i;
                return -1;
            };
            s.cv.wait(g, [&] { return s.stop || runnable() >= 0; });
            const int i = runnable();
            if (i < 0) return;  // stop
            t = std::move(s.io_queue[(size_t)i]);
            s.io_queue.erase(s.io_queue.begin() + i);
            if (t.type == IoTask::DECODE) s.decoded_inflight++;
        }
        if (t.type == IoTask::WRITE) {
            const int rc = write_dds_file(t.dds_path, t.pin_out, t.len);
            std::lock_guard<std::mutex> g(s.mu);
            note_error_locked(s, rc);
            pin_release_locked(s, t.pin_in);
            pin_release_locked(s, t.pin_out);
            s.staged_bytes -= t.staged;
            s.pending--;
        } else {
            std::vector<uint8_t> rgba;
            uint32_t w = 0, h = 0;
            std::string err;
            int rc = guarded([&] {
                const int r = load_image(t.path.c_str(), rgba, w, h, err);
                return r ? fail(r, err) : r;
            });
            Job j;
            j.desc = {w, h, t.kind, -1, -1, -1};
            Resolved r;
            if (!rc) rc = resolve(&j.desc, r);
            if (!rc) {
                j.cap = ROSY_TEX_DDS_HEADER_BYTES + (size_t)make_chain(w, h).total_blocks * block_bytes(r.format);
                void* pin_in = pin_acquire(s, rgba.size());
                j.out = pin_acquire(s, j.cap);
                if (!pin_in || !j.out) {
                    rc = fail(ROSY_TEX_ERR_NOMEM, "pinned host allocation failed");
                    std::lock_guard<std::mutex> g(s.mu);
                    if (pin_in) pin_release_locked(s, pin_in);
                    if (j.out) pin_release_locked(s, j.out);
                } else {
                    memcpy(pin_in, rgba.data(), rgba.size());
                    j.rgba = static_cast<const uint8_t*>(pin_in);
                }
            }
            std::lock_guard<std::mutex> g(s.mu);
            if (rc) {
                note_error_locked(s, rc);
                s.decoded_inflight--;
                s.pending--;
            } else {
                j.from_file = true;
                j.staged = rgba.size() + j.cap;
                s.staged_bytes += j.staged;
                j.dds_path = t.dds_path.empty() ? replace_extension(t.path, ".dds") : t.dds_path;  // Packager.cpp:231
                queue_push_locked(s, std::move(j));
            }
        }
        s.cv.notify_all();
    }
}

int io_thread_count() {
    const char* e = getenv("ROSY_TEX_IO_THREADS");
    int n = e ? atoi(e) : (int)std::thread::hardware_concurrency();
    if (n < 1) n = 1;
    if (n > 32) n = 32;
    return n;
}

int enqueue(Job&& job) {
    int rc = ensure_init();
    if (rc) return rc;
    State& s = st();
    {
        std::lock_guard<std::mutex> g(s.mu);
        start_workers_locked(s);
        s.pending++;
        queue_push_locked(s, std::move(job));
    }
    s.cv.notify_all();
    return ROSY_TEX_OK;
}

int enqueue_file(const char* path, int kind) {
    int rc = ensure_init();
    if (rc) return rc;
    State& s = st();
    {
        std::lock_guard<std::mutex> g(s.mu);
        start_workers_locked(s);
        if (s.io_workers.empty()) {
            const int n = io_thread_count();
            for (int i = 0; i < n; i++) s.io_workers.emplace_back(io_main);
            // images being decoded or waiting for a device slot: every I/O thread busy plus two per slot
            s.decoded_limit = n + 2 * (int)s.workers.size();
            // and a byte budget for the pinned staging (sizes are known only after decode, so it can be
            // overshot by one image per I/O thread): ROSY_TEX_STAGING_MB, default 4096
            if (const char* e = getenv("ROSY_TEX_STAGING_MB")) {
                const long mb = atol(e);
                if (mb > 0) s.staged_limit = (size_t)mb << 20;
            }
        }
        IoTask t;
        t.type = IoTask::DECODE;
        t.path = path;
        t.kind = kind;
        s.pending++;
        s.io_queue.push_back(std::move(t));
    }
    s.cv.notify_all();
    return ROSY_TEX_OK;
}

}  // namespace

Chain make_chain(uint32_t w, uint32_t h) {
    Chain c;
    memset(&c, 0, sizeof c);
    uint64_t px = 0, blk = 0;
    int l = 0;
    for (;; l++) {
        Level& L = c.lv[l];
        L.w = w;
        L.h = h;
        L.bw = (w + 3) / 4;
        L.bh = (h + 3) / 4;
        L.px_offset = px;
        L.blk_offset = blk;
        px += (uint64_t)w * h;
        blk += (uint64_t)L.bw * L.bh;
        if (w == 1 && h == 1) break;
        w = w / 2 > 1 ? w / 2 : 1;  // nvtt::Surface::countMipmaps / buildNextMipmap extents
        h = h / 2 > 1 ? h / 2 : 1;
    }
    c.levels = l + 1;
    c.total_px = px;
    c.total_blocks = blk;
    return c;
}

}  // namespace rosy_tex

using namespace rosy_tex;

extern "C" {

int rosy_tex_init(const rosy_tex_config* config) {
    State& s = st();
    std::lock_guard<std::mutex> g(s.mu);
    return init_locked(s, config);
}

void rosy_tex_shutdown(void) {
    State& s = st();
    {
        std::lock_guard<std::mutex> g(s.mu);
        s.stop = true;
    }
    s.cv.notify_all();
    for (std::thread& t : s.workers) t.join();
    for (std::thread& t : s.io_workers) t.join();
    std::lock_guard<std::mutex> g(s.mu);
    s.workers.clear();
    s.live_workers = 0;
    s.io_workers.clear();
    s.io_queue.clear();
    for (PinBuf& b : s.pinned) cudaFreeHost(b.p);
    s.pinned.clear();
    s.pending = 0;
    s.decoded_inflight = 0;
    s.staged_bytes = 0;
    for (Slot& sl : s.slots) {
        cudaSetDevice(sl.device);
        if (sl.stream) cudaStreamDestroy(sl.stream);
        cudaFree(sl.d_in);
        cudaFree(sl.d_chain);
        cudaFree(sl.d_scratch);
        cudaFree(sl.d_blocks);
        cudaFree(sl.d_desc);
    }
    s.slots.clear();
    s.devices.clear();
    s.queue.clear();
    s.stop = false;
    s.inited = false;
    s.first_error = 0;
}

int rosy_tex_device_count(void) {
    if (ensure_init()) return 0;
    return (int)st().devices.size();
}

const char* rosy_tex_last_error(void) { return t_error.c_str(); }

const char* rosy_tex_status_string(int status) {
    switch (status) {
        case ROSY_TEX_OK: return "ok";
        case ROSY_TEX_ERR_IO: return "io error";
        case ROSY_TEX_ERR_DECODE: return "image decode error";
        case ROSY_TEX_ERR_ARG: return "invalid argument";
        case ROSY_TEX_ERR_CUDA: return "cuda error";
        case ROSY_TEX_ERR_NO_DEVICE: return "no sm_100 device (no CPU fallback)";
        case ROSY_TEX_ERR_NOMEM: return "out of device memory";
        default: return "unknown status";
    }
}

int rosy_tex_generate_texture_file(const char* image_path, int kind, const char* dds_path) {
    if (!image_path) return fail(ROSY_TEX_ERR_ARG, "null path");
    if (kind < 0 || kind > 3) return fail(ROSY_TEX_ERR_ARG, "bad kind");
    int rc = ensure_init();
    if (rc) return rc;
    int prev = 0;
    cudaGetDevice(&prev);
    Slot* sl = acquire_slot();
    rc = guarded([&] { return run_file(sl, image_path, kind, dds_path ? dds_path : ""); });
    release_slot(sl);
    cudaSetDevice(prev);
    return rc;
}

int rosy_tex_generate_srgb_texture(const char* image_path) {
    return rosy_tex_generate_texture_file(image_path, ROSY_TEX_KIND_COLOR, nullptr);
}

int rosy_tex_generate_normal_map_texture(const char* image_path) {
    return rosy_tex_generate_texture_file(image_path, ROSY_TEX_KIND_NORMAL_MAP, nullptr);
}

int rosy_tex_load_image(const char* image_path, uint8_t** rgba, uint32_t* width, uint32_t* height) {
    if (!image_path || !rgba || !width || !height) return fail(ROSY_TEX_ERR_ARG, "null argument");
    return guarded([&] {
        std::vector<uint8_t> px;
        std::string err;
        int rc = load_image(image_path, px, *width, *height, err);
        if (rc) return fail(rc, err);
        *rgba = static_cast<uint8_t*>(malloc(px.size()));
        if (!*rgba) return fail(ROSY_TEX_ERR_NOMEM, "malloc");
        memcpy(*rgba, px.data(), px.size());
        return (int)ROSY_TEX_OK;
    });
}

void rosy_tex_free_image(uint8_t* rgba) { free(rgba); }

int rosy_tex_count_mipmaps(uint32_t width, uint32_t height) {
    if (!width || !height) return 0;
    return make_chain(width, height).levels;
}

void rosy_tex_level_extent(uint32_t width, uint32_t height, int level, uint32_t* w, uint32_t* h) {
    for (int i = 0; i < level; i++) {
        width = width / 2 > 1 ? width / 2 : 1;
        height = height / 2 > 1 ? height / 2 : 1;
    }
    if (w) *w = width;
    if (h) *h = height;
}

size_t rosy_tex_dds_size(const rosy_tex_desc* desc) {
    Resolved r;
    if (resolve(desc, r)) return 0;
    return ROSY_TEX_DDS_HEADER_BYTES + (size_t)make_chain(r.w, r.h).total_blocks * block_bytes(r.format);
}

size_t rosy_tex_chain_size(uint32_t width, uint32_t height) {
    if (!width || !height) return 0;
    return (size_t)make_chain(width, height).total_px * 4;
}

int rosy_tex_dds_header(const rosy_tex_desc* desc, uint8_t* out148) {
    Resolved r;
    int rc = resolve(desc, r);
    if (rc) return rc;
    if (!out148) return fail(ROSY_TEX_ERR_ARG, "null buffer");
    write_dds_header(r, make_chain(r.w, r.h).levels, out148);
    return ROSY_TEX_OK;
}

int rosy_tex_encode_rgba8(const rosy_tex_desc* desc, const uint8_t* rgba, void* out, size_t cap, size_t* len) {
    return run_host_any(desc, WHAT_DDS, rgba, out, cap, len);
}

int rosy_tex_mips_rgba8(const rosy_tex_desc* desc, const uint8_t* rgba, uint8_t* out, size_t cap, size_t* len) {
    return run_host_any(desc, WHAT_MIPS, rgba, out, cap, len);
}

int rosy_tex_compress_level_rgba8(const rosy_tex_desc* desc, const uint8_t* rgba, void* out, size_t cap, size_t* len) {
    return run_host_any(desc, WHAT_LEVEL, rgba, out, cap, len);
}

int rosy_tex_submit(const char* image_path, int kind) {
    if (!image_path) return fail(ROSY_TEX_ERR_ARG, "null path");
    if (kind < 0 || kind > 3) return fail(ROSY_TEX_ERR_ARG, "bad kind");
    return guarded([&] { return enqueue_file(image_path, kind); });
}

int rosy_tex_submit_rgba8(const rosy_tex_desc* desc, const uint8_t* rgba, void* out, size_t cap, size_t* len) {
    Resolved r;
    int rc = ensure_init();
    if (rc) return rc;
    if ((rc = resolve(desc, r))) return rc;
    if (!rgba || !out) return fail(ROSY_TEX_ERR_ARG, "null buffer");
    return guarded([&] {
        Job j;
        j.desc = *desc;
        j.rgba = rgba;
        j.out = out;
        j.cap = cap;
        j.len = len;
        return enqueue(std::move(j));
    });
}

int rosy_tex_wait_all(void) {
    State& s = st();
    std::unique_lock<std::mutex> g(s.mu);
    s.cv.wait(g, [&] { return s.pending == 0; });
    const int rc = s.first_error;
    if (rc) t_error = s.first_error_msg;
    s.first_error = 0;
    return rc;
}

void* rosy_tex_alloc_host(size_t bytes) {
    if (ensure_init()) return nullptr;
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}

void rosy_tex_free_host(void* p) {
    if (p) cudaFreeHost(p);
}

size_t rosy_tex_device_scratch_size(const rosy_tex_desc* desc) {
    Resolved r;
    if (resolve(desc, r)) return 0;
    return mips_scratch_bytes(make_chain(r.w, r.h));
}

static bool misaligned(const void* a, const void* b, const void* c, const void* d) {
    return (((uintptr_t)a | (uintptr_t)b | (uintptr_t)c | (uintptr_t)d) & 15u) != 0;
}

static int device_common(const rosy_tex_desc* desc, Resolved& r) {
    int rc = ensure_init();
    if (rc) return rc;
    if ((rc = resolve(desc, r))) return rc;
    int dev = 0;
    CU_TRY(cudaGetDevice(&dev));
    State& s = st();
    std::lock_guard<std::mutex> g(s.mu);
    return ensure_tables_locked(s, dev);
}

int rosy_tex_mips_device(const rosy_tex_desc* desc, const void* d_rgba, void* d_chain, void* d_scratch, void* stream) {
    Resolved r;
    int rc = device_common(desc, r);
    if (rc) return rc;
    if (!d_rgba || !d_chain || !d_scratch) return fail(ROSY_TEX_ERR_ARG, "null device pointer");
    if (misaligned(d_rgba, d_chain, d_scratch, nullptr)) return fail(ROSY_TEX_ERR_ARG, "device pointers must be 16-byte aligned");
    return encode_on_stream(make_chain(r.w, r.h), r, static_cast<const uint8_t*>(d_rgba), static_cast<uint8_t*>(d_chain),
                            d_scratch, nullptr, static_cast<cudaStream_t>(stream), true, false);
}

int rosy_tex_blocks_device(const rosy_tex_desc* desc, const void* d_rgba, const void* d_chain, void* d_blocks, void* stream) {
    Resolved r;
    int rc = device_common(desc, r);
    if (rc) return rc;
    if (!d_rgba || !d_chain || !d_blocks) return fail(ROSY_TEX_ERR_ARG, "null device pointer");
    if (misaligned(d_rgba, d_chain, d_blocks, nullptr)) return fail(ROSY_TEX_ERR_ARG, "device pointers must be 16-byte aligned");
    return encode_on_stream(make_chain(r.w, r.h), r, static_cast<const uint8_t*>(d_rgba),
                            const_cast<uint8_t*>(static_cast<const uint8_t*>(d_chain)), nullptr,
                            static_cast<uint8_t*>(d_blocks), static_cast<cudaStream_t>(stream), false, true);
}

int rosy_tex_encode_device(const rosy_tex_desc* desc, const void* d_rgba, void* d_chain, void* d_scratch, void* d_blocks,
                           void* stream) {
    Resolved r;
    int rc = device_common(desc, r);
    if (rc) return rc;
    if (!d_rgba || !d_chain || !d_scratch || !d_blocks) return fail(ROSY_TEX_ERR_ARG, "null device pointer");
    if (misaligned(d_rgba, d_chain, d_scratch, d_blocks)) return fail(ROSY_TEX_ERR_ARG, "device pointers must be 16-byte aligned");
    return encode_on_stream(make_chain(r.w, r.h), r, static_cast<const uint8_t*>(d_rgba), static_cast<uint8_t*>(d_chain),
                            d_scratch, static_cast<uint8_t*>(d_blocks), static_cast<cudaStream_t>(stream), true, true);
}

// The allocation-free variant of rosy_tex_mips_device (see the header).  Deliberately does NOT go through
// ensure_init(): that creates streams and probes the device with a cudaMalloc.  The descriptor is resolved against
// the configuration if rosy_tex_init() has run, else against the documented defaults.
int rosy_tex_mips_device_static(const rosy_tex_desc* desc, const void* d_rgba, void* d_chain, void* stream) {
    Resolved r;
    int rc = resolve(desc, r);
    if (rc) return rc;
    if (!d_rgba || !d_chain) return fail(ROSY_TEX_ERR_ARG, "null device pointer");
    if (misaligned(d_rgba, d_chain, nullptr, nullptr)) return fail(ROSY_TEX_ERR_ARG, "device pointers must be 16-byte aligned");
    int launches = 0;
    const char* err = "";
    rc = mips_static_launch(make_chain(r.w, r.h), r.normal_map, r.normalize, static_cast<const uint8_t*>(d_rgba),
                            static_cast<uint8_t*>(d_chain), static_cast<cudaStream_t>(stream), &launches, &err);
    if (rc) return fail(rc, std::string("rosy_tex_mips_device_static: ") + err);
    g_launches += (uint64_t)launches;
    return ROSY_TEX_OK;
}

size_t rosy_tex_batch_scratch_size(int count) { return batch_desc_bytes(count); }

}  // extern "C"
static int encode_batch_device_impl(const rosy_tex_batch_item* items, int count, void* d_batch_scratch, void* stream,
                                    int do_mips, int do_blocks);
extern "C" {

int rosy_tex_encode_batch_device(const rosy_tex_batch_item* items, int count, void* d_batch_scratch, void* stream,
                                 int do_mips, int do_blocks) {
    if (!items || count <= 0) return fail(ROSY_TEX_ERR_ARG, "empty batch");
    return guarded([&] { return encode_batch_device_impl(items, count, d_batch_scratch, stream, do_mips, do_blocks); });
}
}  // extern "C"

static int encode_batch_device_impl(const rosy_tex_batch_item* items, int count, void* d_batch_scratch, void* stream,
                                    int do_mips, int do_blocks) {
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    std::vector<Resolved> rs((size_t)count);
    std::vector<MipBatchItem> mi((size_t)count);
    for (int i = 0; i < count; i++) {
        int rc = device_common(&items[i].desc, rs[i]);
        if (rc) return rc;
        const rosy_tex_batch_item& it = items[i];
        if (!it.d_rgba || !it.d_chain || (do_mips && !it.d_scratch) || (do_blocks && !it.d_blocks))
            return fail(ROSY_TEX_ERR_ARG, "null device pointer in batch item");
        if (misaligned(it.d_rgba, it.d_chain, it.d_scratch, it.d_blocks))
            return fail(ROSY_TEX_ERR_ARG, "device pointers must be 16-byte aligned");
        mi[i].chain = make_chain(rs[i].w, rs[i].h);
        mi[i].normal_map = rs[i].normal_map;
        mi[i].normalize = rs[i].normalize;
        mi[i].d_rgba = static_cast<const uint8_t*>(it.d_rgba);
        mi[i].d_chain = static_cast<uint8_t*>(it.d_chain);
        mi[i].d_scratch = it.d_scratch;
    }
    std::vector<uint8_t*> blk((size_t)count);
    for (int i = 0; i < count; i++) blk[(size_t)i] = static_cast<uint8_t*>(items[i].d_blocks);
    return encode_batch_on_stream(count, rs.data(), mi.data(), blk.data(), d_batch_scratch, s, do_mips != 0, do_blocks != 0);
}

extern "C" {
long long rosy_tex_debug_faults(void) {
    // stores outside their buffers counted by a ROSY_TEX_CHECKED build on every device in use; -1: not a checked build
    State& s = st();
    long long total = 0;
    int prev = 0;
    cudaGetDevice(&prev);
    std::vector<int> devs;
    {
        std::lock_guard<std::mutex> g(s.mu);
        for (const Slot& sl : s.slots)
            if (std::find(devs.begin(), devs.end(), sl.device) == devs.end()) devs.push_back(sl.device);
    }
    if (devs.empty()) devs.push_back(prev);
    for (int d : devs) {
        cudaSetDevice(d);
        cudaDeviceSynchronize();
        const long long n = mips_check_faults();
        if (n < 0) {
            cudaSetDevice(prev);
            return n;
        }
        total += n;
    }
    cudaSetDevice(prev);
    return total;
}

uint64_t rosy_tex_kernel_launches(void) { return g_launches.load(); }

void rosy_tex_queue_status(int* live_workers, int* requeued) {
    State& s = st();
    std::lock_guard<std::mutex> g(s.mu);
    if (live_workers) *live_workers = s.workers.empty() ? -1 : s.live_workers;
    if (requeued) *requeued = s.requeued;
}

}  // extern "C"
