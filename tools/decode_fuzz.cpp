// decode_fuzz.cpp - test tooling: the image decoders of librosy_tex_cuda (image_io.cpp, jpeg_decode.cpp, inflate_fast.cpp)
// on mutated files.  Seeds are the PNG / JPEG / TGA files of tests/golden/*.npz (written out by the test); every seed is
// decoded as is and then under seeded mutations: bit flips, byte splats, truncation, 16-bit / 32-bit header field
// overwrites, chunk duplication.  A mutated file may decode or fail; it must never read or write memory it does not own,
// so build with -fsanitize=address,undefined ("run the decoders under ASAN and a fuzzer").
//   g++ -O1 -g -fsanitize=address,undefined -fwrapv -std=c++17 -Iinclude -o decode_fuzz tools/decode_fuzz.cpp \
//       rosy_b200/csrc/image_io.cpp rosy_b200/csrc/jpeg_decode.cpp rosy_b200/csrc/inflate_fast.cpp -lz
//   ./decode_fuzz <mutations per seed> file...
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

namespace rosy_tex {
int decode_image_memory(const uint8_t* data, size_t len, const char* name_hint, std::vector<uint8_t>& rgba, uint32_t& w,
                        uint32_t& h, std::string& err);
}

static uint64_t rng = 0x524F5359ull;
static uint32_t rnd() {
    rng ^= rng << 13;
    rng ^= rng >> 7;
    rng ^= rng << 17;
    return (uint32_t)(rng >> 11);
}

int main(int argc, char** argv) {
    if (argc < 3) return 2;
    const int per_seed = atoi(argv[1]);
    long decoded = 0, failed = 0, seeds_ok = 0;
    for (int a = 2; a < argc; a++) {
        FILE* fp = fopen(argv[a], "rb");
        if (!fp) return 2;
        std::vector<uint8_t> seed;
        uint8_t buf[65536];
        size_t n;
        while ((n = fread(buf, 1, sizeof buf, fp)) > 0) seed.insert(seed.end(), buf, buf + n);
        fclose(fp);
        std::vector<uint8_t> rgba;
        uint32_t w = 0, h = 0;
        std::string err;
        if (rosy_tex::decode_image_memory(seed.data(), seed.size(), argv[a], rgba, w, h, err) == 0) seeds_ok++;
        for (int m = 0; m < per_seed && seed.size() > 4; m++) {
            std::vector<uint8_t> f(seed);
            const int kind = rnd() % 6, reps = 1 + rnd() % 4;
            for (int r = 0; r < reps; r++) {
                const size_t at = rnd() % f.size();
                switch (kind) {
                    case 0: f[at] ^= (uint8_t)(1u << (rnd() % 8)); break;
                    case 1: f[at] = (uint8_t)rnd(); break;
                    case 2: f.resize(1 + rnd() % f.size()); break;
                    case 3: if (at + 2 <= f.size()) { const uint16_t v = (uint16_t)(rnd() % 3 ? rnd() : 0xFFFF); memcpy(&f[at], &v, 2); } break;
                    case 4: if (at + 4 <= f.size()) { const uint32_t v = rnd() % 3 ? rnd() : 0xFFFFFFFFu; memcpy(&f[at], &v, 4); } break;
                    default: {  // duplicate a slice somewhere else
                        const size_t len = 1 + rnd() % 64, to = rnd() % f.size();
                        for (size_t i = 0; i < len && at + i < f.size() && to + i < f.size(); i++) f[to + i] = f[at + i];
                    }
                }
                if (f.empty()) f.push_back(0);
            }
            // a mutated header can ask for a huge image: the decoders bound allocations by the file size / 32768^2; skip
            // decodes that would legitimately need more than ~1 GB so that the test stays small
            rgba.clear();
            const int rc = rosy_tex::decode_image_memory(f.data(), f.size(), argv[a], rgba, w, h, err);
            if (rc == 0) {
                decoded++;
                if (rgba.size() != (size_t)w * h * 4) {
                    printf("size mismatch on %s\n", argv[a]);
                    return 1;
                }
            } else {
                failed++;
            }
        }
    }
    printf("decode_fuzz: %d seeds (%ld decode), %ld mutated files decoded, %ld rejected\n", argc - 2, seeds_ok, decoded, failed);
    return 0;
}
