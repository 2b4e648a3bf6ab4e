#pragma once
#include <ostream>

#include "crt_scene.h"

namespace crt {
// P3 text, same bytes as the reference writer (src/core/crt_image_ppm.cpp:9-23)
void write_ppm(const Image &image, std::ostream &os, int max_color_component = 255);
}
