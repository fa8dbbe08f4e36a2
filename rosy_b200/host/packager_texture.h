// packager_texture.h - host-side C++ mirror of the reference's texture-stage interface.
//
// Same names, argument meaning and error behaviour as /root/reference/src/Packager/Packager.h:8-9:
//   rosy::result generate_srgb_texture(const std::shared_ptr<rosy_logger::log>&, const std::filesystem::path&);
//   rosy::result generate_normal_map_texture(...same...);
// In the reference tree these two definitions replace Packager.cpp:216-308 verbatim;
// here the reference's `rosy::result` (src/Engine/Types.h:13) and `rosy_logger::log`
// (src/Logger/Logger.h:17-26) are re-declared minimally so the mirror builds without the Engine.
#pragma once
#include <cstdint>
#include <filesystem>
#include <memory>
#include <string_view>

namespace rosy {
enum class result : uint8_t { ok, error };  // the texture stage only ever returns these two
}

namespace rosy_logger {
enum class log_level : uint8_t { debug, info, warn, error, disabled };
class log {
public:
    log_level level{log_level::info};
    void debug(std::string_view log_message) const;
    void info(std::string_view log_message) const;
    void warn(std::string_view log_message) const;
    void error(std::string_view log_message) const;
};
}  // namespace rosy_logger

namespace rosy_packager {
[[nodiscard]] rosy::result generate_srgb_texture(const std::shared_ptr<rosy_logger::log>& l,
                                                 const std::filesystem::path& image_path);
[[nodiscard]] rosy::result generate_normal_map_texture(const std::shared_ptr<rosy_logger::log>& l,
                                                       const std::filesystem::path& image_path);

// Asynchronous variant for importers that want every image of an asset in flight at once
// (Gltf.cpp:209-285 walks them serially today): enqueue each image, then wait once before
// asset.write (Main.cpp:32).  image_type follows Asset.h:109-112.
[[nodiscard]] rosy::result submit_texture(const std::shared_ptr<rosy_logger::log>& l,
                                          const std::filesystem::path& image_path, int image_type);
[[nodiscard]] rosy::result wait_for_textures(const std::shared_ptr<rosy_logger::log>& l);
}  // namespace rosy_packager
