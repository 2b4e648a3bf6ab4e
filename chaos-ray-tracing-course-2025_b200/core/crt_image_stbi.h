#pragma once
#include <filesystem>
#include <optional>

#include "crt_scene.h"

namespace crt {
// JPEG/PNG -> float rgb Image, 3 channels forced (reference: src/core/crt_image_stbi.cpp:16-40)
std::optional<Image> read_stb(const std::filesystem::path &filename);
}
