// gltf_images.h - the image part of the Packager's glTF import, without fastgltf.
//
// Follows /root/reference/src/Packager/Gltf.cpp:
//   :97-111   one rosy_asset::image per glTF image; name = <asset dir>/<uri up to the first '.'>.dds
//   :113-200  per material: baseColorTexture / metallicRoughnessTexture / normalTexture ->
//             textures[].source -> image index, collected into three lists
//   :202-286  the three loops: colour images -> generate_srgb_texture (image_type_color),
//             metallic-roughness images -> generate_srgb_texture (image_type_metallic_roughness),
//             normal maps -> generate_normal_map_texture (image_type_normal_map); an image whose
//             source is not a URI is an error ("not a URI datasource")
// Emissive and occlusion textures are not read.  The reference's sort + unique on
// (image index, ordinal) tuples never removes anything (F14): here every image index is processed
// once per list, and an image that sits in several lists is processed by the LAST list that names
// it, which is the file the reference's serial loops leave on disk.
#pragma once
#include <cstdint>
#include <filesystem>
#include <memory>
#include <string>
#include <vector>

#include "packager_texture.h"

namespace rosy_packager {

// Asset.h:109-112
constexpr uint8_t image_type_color = 0;
constexpr uint8_t image_type_normal_map = 1;
constexpr uint8_t image_type_metallic_roughness = 2;
constexpr uint8_t image_type_mixmap = 3;

struct gltf_image_job {
    uint32_t image_index{0};            // index into the glTF "images" array = index of rosy_asset::image
    uint8_t image_type{image_type_color};
    std::string uri;                    // as written in the glTF
    std::filesystem::path source_path;  // <asset dir>/<uri>, what generate_*_texture receives
    std::filesystem::path dds_path;     // what generate_*_texture writes: source_path.replace_extension(".dds")
    std::string asset_image_name;       // what the .rsy image table stores (Gltf.cpp:102-108)
    uint32_t uses{0};                   // how many material slots referenced it (the reference compresses it that often)
};

struct gltf_image_plan {
    std::vector<std::string> image_names;  // one per glTF image, in order (rosy_asset::image::name)
    std::vector<gltf_image_job> jobs;      // colour first, then metallic-roughness, then normal maps
    uint32_t num_color{0}, num_metallic{0}, num_normal{0};  // list sizes after de-duplication
    uint32_t reference_calls{0};           // generate_* calls the reference's loops would make
};

// Parse `gltf_path` (JSON .gltf) and fill `plan`.  Errors are logged like the importer does and
// return rosy::result::error.
// all_textures (opt-in): also route emissiveTexture (as a colour image) and
// occlusionTexture (as a metallic-roughness / data image), which the reference never reads (F4).
[[nodiscard]] rosy::result plan_gltf_images(const std::shared_ptr<rosy_logger::log>& l,
                                            const std::filesystem::path& gltf_path, gltf_image_plan& plan,
                                            bool all_textures = false);

// The conditioning step (Gltf.cpp:202-286 with cfg.condition_images): every job through the work
// queue, one wait at the end.  serial = true calls generate_*_texture one after the other instead,
// exactly like the reference.
[[nodiscard]] rosy::result condition_gltf_images(const std::shared_ptr<rosy_logger::log>& l,
                                                 const gltf_image_plan& plan, bool serial);

// JSON manifest of the plan: what the .rsy image table would hold for these images.
std::string plan_to_json(const gltf_image_plan& plan);

}  // namespace rosy_packager
