// packager_texture.cpp - the two texture-stage functions of the Rosy Packager on top of the C ABI.
//
// Drop-in for /root/reference/src/Packager/Packager.cpp:216-261 and :263-308.  Every NVTT call of
// those bodies (Surface::load, Context::outputHeader, Context::compress, the Surface mip ops) now
// happens inside ONE call into librosy_tex_cuda.so; what remains here is what the reference's
// bodies did around NVTT: path -> string, status -> log line + rosy::result.
#include "packager_texture.h"

#include <format>
#include <iostream>
#include <string>

#include "../../include/rosy_tex_cuda.h"

namespace rosy_logger {
// src/Logger/Logger.cpp:8-30 writes timestamped lines to stdout / stderr; the mirror keeps the split.
void log::debug(std::string_view m) const { if (level <= log_level::debug) std::cout << "[debug] " << m << '\n'; }
void log::info(std::string_view m) const { if (level <= log_level::info) std::cout << "[info] " << m << '\n'; }
void log::warn(std::string_view m) const { if (level <= log_level::warn) std::cout << "[warn] " << m << '\n'; }
void log::error(std::string_view m) const { if (level <= log_level::error) std::cerr << "[error] " << m << '\n'; }
}  // namespace rosy_logger

namespace {
rosy::result report(const std::shared_ptr<rosy_logger::log>& l, int status, const std::string& input_filename) {
    if (status == ROSY_TEX_OK) return rosy::result::ok;
    // the reference logs one of three messages (Packager.cpp:224, :239, :247) and returns error
    const char* what = status == ROSY_TEX_ERR_IO || status == ROSY_TEX_ERR_DECODE
                           ? "Failed to load file  for  {}"
                           : "Compressing and writing the dds file failed for  {}";
    l->error(std::vformat(what, std::make_format_args(input_filename)));
    l->error(std::format("rosy_tex: {} ({})", rosy_tex_status_string(status), rosy_tex_last_error()));
    return rosy::result::error;
}
}  // namespace

rosy::result rosy_packager::generate_srgb_texture(const std::shared_ptr<rosy_logger::log>& l,
                                                  const std::filesystem::path& image_path) {
    const std::string input_filename{image_path.string()};
    return report(l, rosy_tex_generate_srgb_texture(input_filename.c_str()), input_filename);
}

rosy::result rosy_packager::generate_normal_map_texture(const std::shared_ptr<rosy_logger::log>& l,
                                                        const std::filesystem::path& image_path) {
    const std::string input_filename{image_path.string()};
    return report(l, rosy_tex_generate_normal_map_texture(input_filename.c_str()), input_filename);
}

rosy::result rosy_packager::submit_texture(const std::shared_ptr<rosy_logger::log>& l,
                                           const std::filesystem::path& image_path, int image_type) {
    const std::string input_filename{image_path.string()};
    return report(l, rosy_tex_submit(input_filename.c_str(), image_type), input_filename);
}

rosy::result rosy_packager::wait_for_textures(const std::shared_ptr<rosy_logger::log>& l) {
    return report(l, rosy_tex_wait_all(), "<texture queue>");
}
