// Host P3 encoder.  Byte-for-byte the reference's output format ("P3\nW H\nmax\n", then "r g b\t" per pixel and
// "\n" per row; each channel = clamp(int(c * max), 0, max), C truncation -- src/core/crt_image_ppm.cpp:9-23), but
// formatted into one buffer with a hand-rolled integer printer instead of three operator<< calls per pixel.
#include "crt_image_ppm.h"

#include <algorithm>
#include <string>

namespace crt {

namespace {
inline char *put_int(char *p, int v) {
    char tmp[12];
    int n = 0;
    unsigned u = unsigned(v);
    do {
        tmp[n++] = char('0' + u % 10);
        u /= 10;
    } while (u);
    while (n)
        *p++ = tmp[--n];
    return p;
}
} // namespace

void write_ppm(const Image &image, std::ostream &os, int max_color_component) {
    os << "P3\n" << image.width << ' ' << image.height << '\n' << max_color_component << '\n';
    std::string row;
    row.resize(size_t(image.width) * 36 + 2);
    const float scale = float(max_color_component);
    for (int y = 0; y < image.height; ++y) {
        char *p = row.data();
        const Color *src = image.buffer.data() + size_t(y) * size_t(image.width);
        for (int x = 0; x < image.width; ++x) {
            const Color c = src[x] * scale;
            p = put_int(p, std::clamp(static_cast<int>(c.x), 0, max_color_component));
            *p++ = ' ';
            p = put_int(p, std::clamp(static_cast<int>(c.y), 0, max_color_component));
            *p++ = ' ';
            p = put_int(p, std::clamp(static_cast<int>(c.z), 0, max_color_component));
            *p++ = '\t';
        }
        *p++ = '\n';
        os.write(row.data(), p - row.data());
    }
}

} // namespace crt
