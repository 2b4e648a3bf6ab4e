// crt_renderer [<scene-file>] [<output-file>] -- the reference CLI contract (src/standalone/main.cpp:11-48):
// same defaults, same messages, same exit codes, the same "Execution time" bracket around render_image only.
#include <chrono>
#include <cstdlib>
#include <filesystem>
#include <fstream>
#include <iostream>
#include <string_view>

#include "core/crt_image_ppm.h"
#include "core/crt_json.h"
#include "core/crt_renderer.h"

int main(int argc, char *argv[]) {
    const std::filesystem::path scene_path = argc > 1 ? argv[1] : "../scenes/15-01-conclusion/scene2.crtscene";
    std::ifstream scene_file{ scene_path, std::ios::in | std::ios::binary };
    if (!scene_file.is_open()) {
        std::cerr << "Error: Could not open input file: " << scene_path << '\n';
        return 1;
    }
    std::optional<crt::Scene> scene = crt::json::read_scene_from_istream(scene_file, scene_path.parent_path());
    if (!scene) {
        std::cerr << "Error: Could not parse JSON file: " << scene_path << '\n';
        return 1;
    }
    const char *output_path = argc > 2 ? argv[2] : "output.ppm";
    std::ofstream output{ output_path, std::ios::out | std::ios::binary };
    if (!output.is_open()) {
        std::cerr << "Error: Could not open output file: " << output_path << '\n';
        return 1;
    }

    // Device bring-up (CUDA context, flatten, upload) belongs to "the scene is loaded", like the reference's tree
    // build, which also sits outside its timed region (main.cpp:22 vs :37-43).
    try {
        crt::prepare_device(*scene);
    } catch (const std::exception &e) {
        std::cerr << "Error: " << e.what() << '\n';
        return 1;
    }

    const crt::RendererSettings settings;
    const auto start = std::chrono::high_resolution_clock::now();
    crt::Image image = crt::render_image(*scene, settings);
    const auto stop = std::chrono::high_resolution_clock::now();
    const long double seconds = std::chrono::duration_cast<std::chrono::microseconds>(stop - start).count() / 1'000'000.0l;
    std::cout << "Execution time: " << seconds << " seconds.\n";
    if (const char *v = std::getenv("CRT_STATS"); v && *v == '1') {
        const crt::RenderBreakdown b = crt::last_render_breakdown();
        std::cerr << "{\"gpu_ms\": " << b.gpu_ms << ", \"kernel_launches\": " << b.kernel_launches << ", \"render_image_s\": " << b.total_s << "}\n";
    }

    // the frame is still on the device: the GPU formats the P3 text (same bytes as crt::write_ppm); CRT_PPM=host keeps
    // the host encoder, which is also what any caller-modified Image goes through
    const char *ppm = std::getenv("CRT_PPM");
    if (ppm && std::string_view(ppm) == "host") {
        crt::write_ppm(image, output);
    } else if (!crt::write_ppm_device(*scene, output)) {
        std::cerr << "Error: Could not encode the frame on the device\n";
        return 1;
    }
    return 0;
}
