// asset_images.cpp - see asset_images.h.
#include "asset_images.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <format>

namespace rosy_packager {

namespace {
void put_u32(std::vector<uint8_t>& o, uint32_t v) {
    for (int i = 0; i < 4; i++) o.push_back((uint8_t)(v >> (8 * i)));
}
void put_u64(std::vector<uint8_t>& o, uint64_t v) {
    for (int i = 0; i < 8; i++) o.push_back((uint8_t)(v >> (8 * i)));
}
void put_f32(std::vector<uint8_t>& o, float f) {
    uint32_t u;
    std::memcpy(&u, &f, 4);
    put_u32(o, u);
}
bool ends_with(const std::string& s, const char* suffix) {
    const size_t n = std::strlen(suffix);
    return s.size() >= n && s.compare(s.size() - n, n, suffix) == 0;
}
}  // namespace

std::vector<uint8_t> rsy_image_table_bytes(const std::vector<asset_image>& images) {
    std::vector<uint8_t> o;
    // file_header, Asset.h:13-26 / Asset.cpp:56-62
    put_u32(o, 0x52535946u);  // rosy_format "RSYF"
    put_u32(o, 1);            // current_version
    put_u32(o, 1);            // endianness: little
    for (int i = 0; i < 16; i++) put_f32(o, (i % 5) == 0 ? 1.0f : 0.0f);  // asset_coordinate_system: identity
    put_u32(o, 0);            // root_scene
    // sizes, Asset.cpp:83-99: materials, samplers, scenes, nodes, images, meshes
    for (int i = 0; i < 6; i++) put_u64(o, i == 4 ? (uint64_t)images.size() : 0u);
    // images, Asset.cpp:296-340
    for (const asset_image& im : images) {
        put_u64(o, 1);  // num_image_types
        put_u64(o, (uint64_t)im.name.size());
        put_u32(o, im.image_type);
        o.insert(o.end(), im.name.begin(), im.name.end());
    }
    return o;
}

rosy::result write_rsy_image_table(const std::shared_ptr<rosy_logger::log>& l, const std::filesystem::path& asset_path,
                                   const std::vector<asset_image>& images) {
    const std::vector<uint8_t> bytes = rsy_image_table_bytes(images);
    // Never replace a complete asset: if the target already is a .rsy that holds anything besides images (written by
    // the reference's importer: materials, samplers, scenes, nodes or meshes, Asset.cpp:60-113), leave it alone.
    if (FILE* old = std::fopen(asset_path.string().c_str(), "rb")) {
        uint8_t head[128];
        const size_t got = std::fread(head, 1, sizeof head, old);
        std::fclose(old);
        if (got == sizeof head && std::memcmp(head, "FYSR", 4) == 0) {
            for (int k = 0; k < 6; k++) {
                if (k == 4) continue;  // the image count
                uint64_t n = 0;
                std::memcpy(&n, head + 80 + 8 * k, 8);
                if (n != 0) {
                    l->error(std::format("refusing to overwrite {}: it holds a complete asset, not only an image table",
                                         asset_path.string()));
                    return rosy::result::error;
                }
            }
        }
    }
    FILE* stream = std::fopen(asset_path.string().c_str(), "wb");  // fopen_s in the reference (Windows only)
    if (!stream) {
        l->error(std::format("failed to open for writing {}", asset_path.string()));
        return rosy::result::error;
    }
    const bool ok = std::fwrite(bytes.data(), 1, bytes.size(), stream) == bytes.size();
    if (std::fclose(stream) != 0 || !ok) {
        l->error(std::format("failed to write {} image records", images.size()));
        return rosy::result::error;
    }
    l->debug(std::format("wrote {} images", images.size()));
    return rosy::result::ok;
}

rosy::result plan_fbx_images(const std::shared_ptr<rosy_logger::log>& l, const std::filesystem::path& fbx_path,
                             std::vector<fbx_image_job>& jobs) {
    jobs.clear();
    std::filesystem::path parent_dir = fbx_path.parent_path();
    if (parent_dir.empty()) parent_dir = ".";
    std::error_code ec;
    std::vector<std::filesystem::path> entries;
    for (std::filesystem::directory_iterator it(parent_dir, ec), end; !ec && it != end; it.increment(ec)) entries.push_back(it->path());
    if (ec) {
        l->error(std::format("cannot list {}: {}", parent_dir.string(), ec.message()));
        return rosy::result::error;
    }
    std::sort(entries.begin(), entries.end());
    for (const std::filesystem::path& entry_path : entries) {
        if (entry_path.extension() != ".tga") continue;  // FBX import only supports .tga images (Fbx.cpp:424)
        const std::string fn = entry_path.filename().string();
        fbx_image_job job;
        if (ends_with(fn, "normal.tga"))
            job.image_type = 1;  // image_type_normal_map
        else if (ends_with(fn, "mixmap.tga"))
            job.image_type = 3;  // image_type_mixmap
        else if (ends_with(fn, "albedo.tga"))
            job.image_type = 0;  // image_type_color
        else
            continue;
        l->info(std::format("image found: {}", entry_path.string()));
        job.source_path = entry_path;
        job.dds_path = entry_path;
        job.dds_path.replace_extension(".dds");
        jobs.push_back(job);
    }
    return rosy::result::ok;
}

rosy::result condition_fbx_images(const std::shared_ptr<rosy_logger::log>& l, const std::vector<fbx_image_job>& jobs,
                                  bool serial, std::vector<asset_image>& images) {
    images.clear();
    for (const fbx_image_job& job : jobs) {
        const char* what = job.image_type == 1 ? "normal" : (job.image_type == 3 ? "mixmap" : "albedo");
        rosy::result res;
        if (serial)
            res = job.image_type == 1 ? generate_normal_map_texture(l, job.source_path) : generate_srgb_texture(l, job.source_path);
        else
            res = submit_texture(l, job.source_path, (int)job.image_type);
        if (res != rosy::result::ok) {
            l->error(std::format("error creating {} fbx image: {} for {}", what, static_cast<uint8_t>(res),
                                 job.source_path.filename().string()));
            if (!serial) (void)wait_for_textures(l);
            return res;
        }
        images.push_back({job.image_type, job.dds_path.string()});
    }
    if (!serial) return wait_for_textures(l);
    return rosy::result::ok;
}

}  // namespace rosy_packager
