/* ref_dds.cpp - build recipe glue for oracle/_ref/libdds_ref.so (test infrastructure).
 * Wraps the REFERENCE's own DDS reader, /root/reference/src/libs/dds.hpp:409-574
 * (dds::readImage), which is what the Engine calls at src/Engine/Graphics.cpp:3610.
 * The header is compiled where it lies (Makefile -I/root/reference/src/libs); not copied. */
#include <cstdint>
#include <cstring>
#include <vector>

#include "dds.hpp"

extern "C" int ref_dds_read(const uint8_t* data, size_t size, uint32_t* out /*[8+32]*/) {
    std::vector<uint8_t> copy(data, data + size);
    dds::Image img;
    auto res = dds::readImage(copy.data(), copy.size(), &img);
    if (res != dds::ReadResult::Success) return 1 + static_cast<int>(res);
    out[0] = img.numMips;
    out[1] = img.arraySize;
    out[2] = img.width;
    out[3] = img.height;
    out[4] = static_cast<uint32_t>(img.dimension);
    out[5] = static_cast<uint32_t>(img.format);
    out[6] = img.supportsAlpha ? 1u : 0u;
    out[7] = static_cast<uint32_t>(img.mipmaps.size());
    for (size_t i = 0; i < img.mipmaps.size() && i < 16; i++) {
        out[8 + 2 * i] = static_cast<uint32_t>(img.mipmaps[i].data() - copy.data());
        out[9 + 2 * i] = static_cast<uint32_t>(img.mipmaps[i].size());
    }
    return 0;
}
