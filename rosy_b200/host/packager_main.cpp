// packager_main.cpp - a small CLI over the mirror, for exercising the boundary by hand:
//   rosy_packager_tex [--normal|--srgb|--mr] <image> [...]      one generate_* call per image, serial
//   rosy_packager_tex --queue [--normal|--srgb|--mr] <image> ... every image through the work queue
//   rosy_packager_tex [--serial] [--dry-run|--image-table] [--all-textures] <file.gltf>   the image part of `Packager.exe <file.gltf>`
//       (Main.cpp:126-174 -> Gltf.cpp:97-286): every material image routed to its function, all in
//       flight through the work queue (--serial: one after the other like the reference), and a
//       <file>.textures.json manifest; --dry-run prints it only.
//   rosy_packager_tex [--serial] [--dry-run|--image-table] <file.fbx>   the image part of `Packager.exe <file.fbx>`
//       (Main.cpp:67-124 -> Fbx.cpp:421-474): the *normal.tga / *mixmap.tga / *albedo.tga next to the .fbx.
// Both asset routes then write <file>.images.rsy holding the asset's IMAGE TABLE exactly as asset::write stores it
// (Asset.cpp:39-113 header + sizes, :296-340 image records; asset_images.h).  The name is deliberately NOT <file>.rsy:
// that is where the reference's importer puts the complete asset (Main.cpp:13-14) and a table with zero materials,
// nodes and meshes must never replace it.  --dry-run writes nothing at all; --image-table writes the table only
// (no image is conditioned, no GPU needed).  N1 ends there: materials, nodes and meshes are written by the reference's
// importers, which need fastgltf / the FBX SDK / MikkTSpace / meshoptimizer (absent here) and are outside the
// texture stage.
// Exit code 0 / EXIT_FAILURE like Packager.exe (src/Packager/Main.cpp:143-173).
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <format>
#include <memory>

#include <fstream>
#include <iostream>

#include "asset_images.h"
#include "gltf_images.h"
#include "packager_texture.h"

int main(int argc, char** argv) {
    auto l = std::make_shared<rosy_logger::log>();
    if (argc < 2) {
        l->error("Need to provide an image path");
        return EXIT_FAILURE;
    }
    bool queue = false, serial = false, dry_run = false, table_only = false, all_textures = false;
    int image_type = 0;  // Asset.h:109-112: 0 colour, 1 normal map, 2 metallic-roughness, 3 mixmap
    int done = 0;
    const auto start = std::chrono::system_clock::now();
    for (int i = 1; i < argc; i++) {
        if (!std::strcmp(argv[i], "--queue")) { queue = true; continue; }
        if (!std::strcmp(argv[i], "--normal")) { image_type = 1; continue; }
        if (!std::strcmp(argv[i], "--srgb")) { image_type = 0; continue; }
        if (!std::strcmp(argv[i], "--mr")) { image_type = 2; continue; }
        if (!std::strcmp(argv[i], "--serial")) { serial = true; continue; }
        if (!std::strcmp(argv[i], "--dry-run")) { dry_run = true; continue; }          // plan to stdout, no file is touched
        if (!std::strcmp(argv[i], "--image-table")) { table_only = true; continue; }  // write <file>.images.rsy only
        if (!std::strcmp(argv[i], "--all-textures")) { all_textures = true; continue; }  // + emissive, occlusion (opt-in)
        if (std::filesystem::path(argv[i]).extension() == ".gltf") {
            // Main.cpp:10-65 load_gltf, image part
            rosy_packager::gltf_image_plan plan;
            if (rosy_packager::plan_gltf_images(l, argv[i], plan, all_textures) != rosy::result::ok) return EXIT_FAILURE;
            const std::string manifest = rosy_packager::plan_to_json(plan);
            // rosy_asset::asset::images: one per glTF image (Gltf.cpp:97-111), typed by the list that conditions it (:202-286)
            std::vector<rosy_packager::asset_image> images;
            for (const std::string& name : plan.image_names) images.push_back({rosy_packager::image_type_color, name});
            for (const auto& job : plan.jobs)
                if (job.image_index < images.size()) images[job.image_index].image_type = job.image_type;
            std::filesystem::path rsy{argv[i]};
            rsy.replace_extension(".images.rsy");  // NOT Main.cpp:13-14's <file>.rsy: that one is the importer's
            if (dry_run) {
                std::cout << manifest;
                continue;
            }
            if (table_only) {
                if (rosy_packager::write_rsy_image_table(l, rsy, images) != rosy::result::ok) return EXIT_FAILURE;
                continue;
            }
            if (rosy_packager::condition_gltf_images(l, plan, serial) != rosy::result::ok) return EXIT_FAILURE;
            std::filesystem::path out{argv[i]};
            out.replace_extension(".textures.json");
            std::ofstream(out, std::ios::binary) << manifest;
            if (rosy_packager::write_rsy_image_table(l, rsy, images) != rosy::result::ok) return EXIT_FAILURE;
            done += (int)plan.jobs.size();
            continue;
        }
        if (std::filesystem::path(argv[i]).extension() == ".fbx") {
            // Main.cpp:67-124 load_fbx, image part (Fbx.cpp:421-474)
            std::filesystem::path rsy{argv[i]};
            rsy.replace_extension(".images.rsy");
            l->info(std::format("Parsing {} as {}", argv[i], rsy.string()));
            std::vector<rosy_packager::fbx_image_job> jobs;
            if (rosy_packager::plan_fbx_images(l, argv[i], jobs) != rosy::result::ok) return EXIT_FAILURE;
            std::vector<rosy_packager::asset_image> images;
            if (dry_run || table_only) {
                for (const auto& job : jobs) {
                    images.push_back({job.image_type, job.dds_path.string()});
                    if (dry_run) std::cout << job.image_type << " " << job.source_path.string() << " -> " << job.dds_path.string() << "\n";
                }
                if (dry_run) continue;
            } else if (rosy_packager::condition_fbx_images(l, jobs, serial, images) != rosy::result::ok) {
                return EXIT_FAILURE;
            }
            if (rosy_packager::write_rsy_image_table(l, rsy, images) != rosy::result::ok) return EXIT_FAILURE;
            done += (int)jobs.size();
            continue;
        }
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
