/* ref_asset.cpp - build recipe glue for oracle/_ref/libasset_ref.so (test infrastructure).
 * Wraps the REFERENCE's own .rsy writer and reader, /root/reference/src/Asset/Asset.cpp:39-113 + :296-340
 * (asset::write) and :409-848 (asset::read), compiled where they lie together with src/Logger/Logger.cpp
 * (Makefile: -I/root/reference/src, pch.h from ref_asset_shim/); nothing is copied.  Pins the image table that
 * rosy_b200/host/asset_images.cpp writes. */
#include "pch.h"

#include "Asset/Asset.h"

/* writes <path> with asset::write for an asset that holds `count` images {types[i], names[i]} and nothing else */
extern "C" int ref_asset_write_images(const char* path, int count, const uint32_t* types, const char* const* names) {
    auto l = std::make_shared<rosy_logger::log>();
    l->level = rosy_logger::log_level::disabled;
    rosy_asset::asset a{};
    a.asset_path = path;
    for (int i = 0; i < count; i++) {
        rosy_asset::image img{};
        img.image_type = types[i];
        img.name.assign(names[i], names[i] + std::strlen(names[i]));
        a.images.push_back(img);
    }
    return a.write(l) == rosy::result::ok ? 0 : 1;
}

/* reads <path> with asset::read; returns the image count (< 0: read failed) and the counts of the other sections;
 * image i's type goes to types[i] and its name, NUL-terminated, to names + i * name_stride */
extern "C" int ref_asset_read_images(const char* path, int max_images, uint32_t* types, char* names, int name_stride,
                                     uint64_t* other_counts /*[5]: materials, samplers, scenes, nodes, meshes*/) {
    auto l = std::make_shared<rosy_logger::log>();
    l->level = rosy_logger::log_level::disabled;
    rosy_asset::asset a{};
    a.asset_path = path;
    if (a.read(l) != rosy::result::ok) return -1;
    other_counts[0] = a.materials.size();
    other_counts[1] = a.samplers.size();
    other_counts[2] = a.scenes.size();
    other_counts[3] = a.nodes.size();
    other_counts[4] = a.meshes.size();
    const int n = (int)a.images.size();
    for (int i = 0; i < n && i < max_images; i++) {
        types[i] = a.images[i].image_type;
        const size_t len = std::min<size_t>(a.images[i].name.size(), (size_t)name_stride - 1);
        std::memcpy(names + (size_t)i * name_stride, a.images[i].name.data(), len);
        names[(size_t)i * name_stride + len] = 0;
    }
    return n;
}
