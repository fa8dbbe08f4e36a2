// asset_images.h - the image table of a .rsy asset and the FBX importer's image routing.
//
//  * rosy_asset::image (src/Asset/Asset.h:115-119) and the part of asset::write that stores the images
//    (src/Asset/Asset.cpp:39-113 header + sizes, :296-340 one record per image).  write_rsy_image_table emits
//    exactly the bytes asset::write produces for an asset that holds these images and nothing else
//    (no materials, samplers, scenes, nodes, meshes): the mesh / material sections stay with the reference's
//    importers, which need fastgltf / the FBX SDK / MikkTSpace.
//  * fbx::import's IMAGES section (src/Packager/Fbx.cpp:421-474): a directory scan next to the .fbx for
//    `*normal.tga`, `*mixmap.tga`, `*albedo.tga`; it needs no FBX SDK.
#pragma once
#include <cstdint>
#include <filesystem>
#include <memory>
#include <string>
#include <vector>

#include "packager_texture.h"

namespace rosy_packager {

// rosy_asset::image: image_type per Asset.h:109-112, name = path of the .dds as the importer built it
struct asset_image {
    uint32_t image_type{0};
    std::string name;
};

// The .rsy the Packager would write for an asset with these images only (Asset.cpp:39-340).
// Layout: file_header {magic 'RSYF', version 1, endianness 1, 16 floats coordinate system, root_scene} (80 bytes);
// six size_t counts {materials, samplers, scenes, nodes, images, meshes}; per image: size_t[2] {1, name chars},
// uint32_t image_type, the name's chars.  Little endian, size_t = 8 bytes (the reference targets x64).
std::vector<uint8_t> rsy_image_table_bytes(const std::vector<asset_image>& images);
// Refuses (error) to replace an existing .rsy that holds anything besides images: that is a complete asset.
[[nodiscard]] rosy::result write_rsy_image_table(const std::shared_ptr<rosy_logger::log>& l,
                                                 const std::filesystem::path& asset_path,
                                                 const std::vector<asset_image>& images);

struct fbx_image_job {
    std::filesystem::path source_path;  // <dir>/<x>normal.tga | mixmap.tga | albedo.tga
    std::filesystem::path dds_path;     // source_path with .dds (what generate_*_texture writes, Packager.cpp:231)
    uint32_t image_type{0};             // 1 normal map, 3 mixmap, 0 colour (Fbx.cpp:450,459,468)
};

// Fbx.cpp:421-441: every `.tga` in the .fbx's directory whose file name ends with normal.tga, mixmap.tga or
// albedo.tga (tested in that order, first match wins).  directory_iterator's order is unspecified; entries are
// visited in file-name order here so that the image table is reproducible.
[[nodiscard]] rosy::result plan_fbx_images(const std::shared_ptr<rosy_logger::log>& l,
                                           const std::filesystem::path& fbx_path, std::vector<fbx_image_job>& jobs);

// Fbx.cpp:443-473: normal -> generate_normal_map_texture, mixmap / albedo -> generate_srgb_texture; fills the
// asset's image list (name = the .dds path).  serial = true: one call after the other and stop at the first
// failure with the reference's error line; otherwise every image goes through the work queue and one wait.
[[nodiscard]] rosy::result condition_fbx_images(const std::shared_ptr<rosy_logger::log>& l,
                                                const std::vector<fbx_image_job>& jobs, bool serial,
                                                std::vector<asset_image>& images);

}  // namespace rosy_packager
