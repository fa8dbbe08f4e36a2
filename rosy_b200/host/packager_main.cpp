// packager_main.cpp - a small CLI over the mirror, for exercising the boundary by hand:
//   rosy_packager_tex [--normal|--srgb|--mr] <image> [...]      one generate_* call per image, serial
//   rosy_packager_tex --queue [--normal|--srgb|--mr] <image> ... every image through the work queue
// Exit code 0 / EXIT_FAILURE like Packager.exe (src/Packager/Main.cpp:143-173).  The glTF walk that
// decides which function an image goes to (Gltf.cpp:97-286) stays in the reference's importer.
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <format>
#include <memory>

#include "packager_texture.h"

int main(int argc, char** argv) {
    auto l = std::make_shared<rosy_logger::log>();
    if (argc < 2) {
        l->error("Need to provide an image path");
        return EXIT_FAILURE;
    }
    bool queue = false;
    int image_type = 0;  // Asset.h:109-112: 0 colour, 1 normal map, 2 metallic-roughness, 3 mixmap
    int done = 0;
    const auto start = std::chrono::system_clock::now();
    for (int i = 1; i < argc; i++) {
        if (!std::strcmp(argv[i], "--queue")) { queue = true; continue; }
        if (!std::strcmp(argv[i], "--normal")) { image_type = 1; continue; }
        if (!std::strcmp(argv[i], "--srgb")) { image_type = 0; continue; }
        if (!std::strcmp(argv[i], "--mr")) { image_type = 2; continue; }
        rosy::result r;
        if (queue)
            r = rosy_packager::submit_texture(l, argv[i], image_type);
        else if (image_type == 1)
            r = rosy_packager::generate_normal_map_texture(l, argv[i]);
        else
            r = rosy_packager::generate_srgb_texture(l, argv[i]);
        if (r != rosy::result::ok) return EXIT_FAILURE;
        done++;
    }
    if (queue && rosy_packager::wait_for_textures(l) != rosy::result::ok) return EXIT_FAILURE;
    const auto ms = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::system_clock::now() - start);
    l->info(std::format("Finished packaging {} texture(s). Took {}ms", done, ms.count()));  // Main.cpp:61-63
    return 0;
}
