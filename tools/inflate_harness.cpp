// inflate_harness.cpp - test tooling: rosy_b200/csrc/inflate_fast.cpp against zlib on seeded inputs (noise, few symbols, ramps,
// RGBA-like repeats, zero runs, long matches; every deflate level and strategy, zlib and raw framing), wrong expected sizes,
// truncated and bit-flipped streams (must fail or succeed, never crash: build with -fsanitize=address,undefined).
//   g++ -O2 -std=c++17 -o inflate_harness tools/inflate_harness.cpp rosy_b200/csrc/inflate_fast.cpp -lz && ./inflate_harness [iterations] [notiming]
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include <chrono>
#include <zlib.h>
namespace rosy_tex { bool inflate_fast(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_len, bool raw); }
static uint64_t rng = 88172645463325252ull;
static uint32_t rnd() { rng ^= rng << 13; rng ^= rng >> 7; rng ^= rng << 17; return (uint32_t)(rng >> 11); }
int main(int argc, char** argv) {
    int fails = 0, cases = 0, fastok = 0;
    const int iters = argc > 1 ? atoi(argv[1]) : 3000;
    for (int iter = 0; iter < iters; iter++) {
        const size_t n = iter < 20 ? iter : (rnd() % (iter % 50 == 0 ? 3000000 : 70000));
        std::vector<uint8_t> src(n);
        const int kind = rnd() % 6;
        for (size_t i = 0; i < n; i++) {
            switch (kind) {
                case 0: src[i] = (uint8_t)rnd(); break;                       // noise
                case 1: src[i] = (uint8_t)(rnd() % 4); break;                 // few symbols
                case 2: src[i] = (uint8_t)(i % 7 == 0 ? rnd() : i); break;    // ramps
                case 3: src[i] = i >= 4 && rnd() % 5 ? src[i - 4] : (uint8_t)rnd(); break;  // RGBA-like repeats (distance 4)
                case 4: src[i] = (uint8_t)(rnd() % 3 ? 0 : rnd()); break;     // long zero runs
                default: src[i] = i >= 300 && rnd() % 3 ? src[i - 1 - rnd() % 300] : (uint8_t)(rnd() % 32); break;
            }
        }
        const int level = rnd() % 10, strategy = (rnd() % 8 == 0) ? Z_FIXED : ((rnd() % 8 == 0) ? Z_HUFFMAN_ONLY : Z_DEFAULT_STRATEGY);
        const bool raw = rnd() % 4 == 0;
        z_stream zs; memset(&zs, 0, sizeof zs);
        deflateInit2(&zs, level, Z_DEFLATED, raw ? -15 : 15, 1 + rnd() % 9, strategy);
        std::vector<uint8_t> comp(deflateBound(&zs, n) + 64 + 32);
        zs.next_in = src.data(); zs.avail_in = (uInt)n; zs.next_out = comp.data(); zs.avail_out = (uInt)comp.size() - 32;
        if (deflate(&zs, Z_FINISH) != Z_STREAM_END) { printf("deflate failed\n"); return 2; }
        const size_t clen = zs.total_out; deflateEnd(&zs);
        memset(comp.data() + clen, 0, 32);
        std::vector<uint8_t> out(n + 16, 0xCD);
        cases++;
        const bool ok = rosy_tex::inflate_fast(comp.data(), clen, out.data(), n, raw);
        if (!ok || (n && memcmp(out.data(), src.data(), n))) { fails++; if (fails < 10) printf("FAIL iter %d n %zu kind %d level %d strat %d raw %d ok %d\n", iter, n, kind, level, strategy, raw, ok); }
        else fastok++;
        // wrong expected size must fail, not crash
        if (n > 2) { std::vector<uint8_t> o2(n + 16); if (rosy_tex::inflate_fast(comp.data(), clen, o2.data(), n - 1, raw)) { printf("accepted short output\n"); fails++; }
                     std::vector<uint8_t> o3(n + 17 + 16); if (rosy_tex::inflate_fast(comp.data(), clen, o3.data(), n + 1, raw)) { printf("accepted long output\n"); fails++; } }
        // truncated and corrupted streams must not crash (result may be either)
        if (clen > 8) { std::vector<uint8_t> c2(comp); const size_t cut = rnd() % clen; memset(c2.data() + cut, 0, c2.size() - cut); std::vector<uint8_t> o4(n + 16); rosy_tex::inflate_fast(c2.data(), cut, o4.data(), n, raw);
                        std::vector<uint8_t> c3(comp); for (int k = 0; k < 3; k++) c3[rnd() % clen] ^= (uint8_t)(1u << (rnd() % 8)); std::vector<uint8_t> o5(n + 16); rosy_tex::inflate_fast(c3.data(), clen, o5.data(), n, raw); }
    }
    printf("cases %d fast-ok %d fails %d\n", cases, fastok, fails);
    if (argc > 2) return fails ? 1 : 0;  // any second argument: skip the timing
    // speed on noise-like PNG data
    { const size_t n = 16u << 20; std::vector<uint8_t> src(n); for (size_t i = 0; i < n; i++) src[i] = (uint8_t)((rnd() % 64) + (i % 4) * 40);
      uLongf cl = compressBound(n); std::vector<uint8_t> comp(cl + 32); compress2(comp.data(), &cl, src.data(), n, 6); memset(comp.data() + cl, 0, 32);
      std::vector<uint8_t> out(n + 16);
      auto t0 = std::chrono::steady_clock::now(); bool ok = true; for (int r = 0; r < 3; r++) ok &= rosy_tex::inflate_fast(comp.data(), cl, out.data(), n, false);
      auto t1 = std::chrono::steady_clock::now(); uLongf got = n; for (int r = 0; r < 3; r++) { got = n; uncompress(out.data(), &got, comp.data(), cl); }
      auto t2 = std::chrono::steady_clock::now();
      printf("16 MiB noise-like (ratio %.2f): fast %.1f ms (ok %d), zlib %.1f ms\n", (double)n / cl, std::chrono::duration<double, std::milli>(t1 - t0).count() / 3, ok, std::chrono::duration<double, std::milli>(t2 - t1).count() / 3); }
    return fails ? 1 : 0;
}
