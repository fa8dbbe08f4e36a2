// h2d_ceiling.cu - what the host can feed: plain concurrent cudaMemcpyAsync between pinned host memory and the GPUs,
// nothing else.  SURVEY.md 8(d) "Scaling" asks for this ceiling (the texture stage moves 4 B/px host->device and
// <= 1.33 B/px device->host) so that an end-to-end shortfall at 8 GPUs can be attributed: if the bare copies stop
// scaling at the same aggregate rate as bench.py's e2e leg, the wall is the host (PCIe root / memory), not this library.
//
// One host thread per GPU, cudaHostAlloc(Portable) buffers like rosy_tex_alloc_host, 16 MiB per copy (one 2048^2
// texture), two streams per GPU.  Three legs per GPU count: H2D alone, D2H alone, and the texture stage's mix
// (16 MiB up + 5.33 MiB down per texture, both directions in flight).  One JSON object per line on stdout.
//   nvcc -O2 -std=c++17 -o tools/h2d_ceiling tools/h2d_ceiling.cu        usage: tools/h2d_ceiling [seconds per leg]
#include <cuda_runtime.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

static const size_t UP = 16u << 20, DOWN = (16u << 20) / 3;  // bytes per texture: level 0 in, all BC7 levels out
static const int DEPTH = 4;                                  // copies in flight per stream

#define CK(x)                                                                        \
    do {                                                                             \
        cudaError_t e_ = (x);                                                        \
        if (e_ != cudaSuccess) {                                                     \
            fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_));                 \
            exit(1);                                                                 \
        }                                                                            \
    } while (0)

struct Gpu {
    int dev;
    uint8_t *h_up[DEPTH], *h_down[DEPTH], *d_up[DEPTH], *d_down[DEPTH];
    cudaStream_t s_up, s_down;
    double up_bytes = 0, down_bytes = 0, seconds = 0;
};

static std::string slurp(const std::string& path) {
    FILE* f = fopen(path.c_str(), "r");
    if (!f) return "?";
    char buf[256];
    size_t n = fread(buf, 1, sizeof buf - 1, f);
    fclose(f);
    buf[n] = 0;
    while (n && (buf[n - 1] == '\n' || buf[n - 1] == ' ')) buf[--n] = 0;
    return buf;
}

// mode: 0 H2D only, 1 D2H only, 2 both (the texture stage's ratio)
static void run(Gpu* g, int mode, double seconds, std::atomic<int>* ready, int ngpu) {
    CK(cudaSetDevice(g->dev));
    ready->fetch_add(1);
    while (ready->load() < ngpu) {}  // all GPUs start together
    const auto t0 = std::chrono::steady_clock::now();
    double up = 0, down = 0;
    for (;;) {
        for (int k = 0; k < DEPTH; k++) {
            if (mode != 1) CK(cudaMemcpyAsync(g->d_up[k], g->h_up[k], UP, cudaMemcpyHostToDevice, g->s_up));
            if (mode != 0) CK(cudaMemcpyAsync(g->h_down[k], g->d_down[k], mode == 2 ? DOWN : UP, cudaMemcpyDeviceToHost, g->s_down));
        }
        CK(cudaStreamSynchronize(g->s_up));
        CK(cudaStreamSynchronize(g->s_down));
        if (mode != 1) up += (double)DEPTH * UP;
        if (mode != 0) down += (double)DEPTH * (mode == 2 ? DOWN : UP);
        const double el = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (el >= seconds) {
            g->seconds = el;
            break;
        }
    }
    g->up_bytes = up;
    g->down_bytes = down;
}

int main(int argc, char** argv) {
    const double seconds = argc > 1 ? atof(argv[1]) : 1.5;
    int count = 0;
    CK(cudaGetDeviceCount(&count));
    printf("{\"what\": \"host\", \"gpus_visible\": %d, \"numa_nodes_online\": \"%s\", \"bytes_up_per_copy\": %zu, \"bytes_down_mix\": %zu}\n", count,
           slurp("/sys/devices/system/node/online").c_str(), UP, DOWN);
    {   // Mems_allowed_list / Cpus_allowed_list of this process
        FILE* f = fopen("/proc/self/status", "r");
        char line[512];
        while (f && fgets(line, sizeof line, f))
            if (!strncmp(line, "Mems_allowed_list", 17) || !strncmp(line, "Cpus_allowed_list", 17)) {
                line[strcspn(line, "\n")] = 0;
                for (char* p = line; *p; p++)
                    if (*p == '\t') *p = ' ';
                printf("{\"what\": \"proc_status\", \"line\": \"%s\"}\n", line);
            }
        if (f) fclose(f);
    }
    std::vector<Gpu> gpus((size_t)count);
    for (int d = 0; d < count; d++) {
        Gpu& g = gpus[(size_t)d];
        g.dev = d;
        CK(cudaSetDevice(d));
        char bus[32] = "";
        CK(cudaDeviceGetPCIBusId(bus, sizeof bus, d));
        for (char* p = bus; *p; p++) *p = (char)tolower(*p);
        const std::string base = std::string("/sys/bus/pci/devices/") + bus;
        printf("{\"what\": \"gpu\", \"device\": %d, \"pci\": \"%s\", \"numa_node\": \"%s\", \"link_speed\": \"%s\", \"link_width\": \"%s\"}\n", d, bus,
               slurp(base + "/numa_node").c_str(), slurp(base + "/current_link_speed").c_str(), slurp(base + "/current_link_width").c_str());
        for (int k = 0; k < DEPTH; k++) {
            CK(cudaHostAlloc(&g.h_up[k], UP, cudaHostAllocPortable));
            CK(cudaHostAlloc(&g.h_down[k], UP, cudaHostAllocPortable));
            memset(g.h_up[k], 0x5a, UP);
            memset(g.h_down[k], 0, UP);
            CK(cudaMalloc(&g.d_up[k], UP));
            CK(cudaMalloc(&g.d_down[k], UP));
            CK(cudaMemset(g.d_down[k], 1, UP));
        }
        CK(cudaStreamCreateWithFlags(&g.s_up, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&g.s_down, cudaStreamNonBlocking));
    }
    static const char* const names[3] = {"h2d_only", "d2h_only", "texture_mix_16MiB_up_5.33MiB_down"};
    for (int n = 1; n <= count; n *= 2) {
        for (int mode = 0; mode < 3; mode++) {
            std::atomic<int> ready{0};
            std::vector<std::thread> th;
            for (int d = 0; d < n; d++) th.emplace_back(run, &gpus[(size_t)d], mode, seconds, &ready, n);
            for (auto& t : th) t.join();
            double up = 0, down = 0, lo_up = 1e30, lo_down = 1e30;
            for (int d = 0; d < n; d++) {
                const double u = gpus[(size_t)d].up_bytes / gpus[(size_t)d].seconds / 1e9, w = gpus[(size_t)d].down_bytes / gpus[(size_t)d].seconds / 1e9;
                up += u;
                down += w;
                lo_up = u < lo_up ? u : lo_up;
                lo_down = w < lo_down ? w : lo_down;
            }
            // Mpix/s the stage could reach if the copies were all there is: 4 B/px up
            printf("{\"what\": \"copy\", \"gpus\": %d, \"leg\": \"%s\", \"h2d_gbs_total\": %.2f, \"d2h_gbs_total\": %.2f, \"h2d_gbs_per_gpu\": %.2f, "
                   "\"d2h_gbs_per_gpu\": %.2f, \"h2d_gbs_slowest_gpu\": %.2f, \"d2h_gbs_slowest_gpu\": %.2f, \"mpix_s_ceiling\": %.0f, \"seconds\": %.2f}\n",
                   n, names[mode], up, down, up / n, down / n, mode == 1 ? 0.0 : lo_up, mode == 0 ? 0.0 : lo_down, mode == 1 ? 0.0 : up * 1e9 / 4 / 1e6, seconds);
            fflush(stdout);
        }
    }
    return 0;
}
