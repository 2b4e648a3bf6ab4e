// Bitmap texture decode.  The decoder itself is third-party (stb_image v2.30, the copy the reference vendors
// under vendor/stb); it is an input decoder, not part of the hot path, and is compiled from where it lies
// (-I$(STB_INCLUDE)).  When the header is not available at build time, bitmap textures fail to load, which the
// loader reports exactly like an unreadable file (reference: read_stb returns nullopt, crt_image_stbi.cpp:19-20).
#include "crt_image_stbi.h"

#include <cstdint>

#if __has_include(<stb_image.h>)
#define CRT_HAVE_STB 1
#define STB_IMAGE_IMPLEMENTATION
#define STBI_NO_HDR
#define STBI_NO_LINEAR
#include <stb_image.h>
#else
#define CRT_HAVE_STB 0
#endif

namespace crt {

std::optional<Image> read_stb(const std::filesystem::path &filename) {
#if CRT_HAVE_STB
    int width = 0, height = 0, components = 0;
    uint8_t *bytes = stbi_load(filename.c_str(), &width, &height, &components, STBI_rgb);
    if (!bytes)
        return std::nullopt;
    // the reference rejects files whose NATIVE channel count is not 3 (crt_image_stbi.cpp:24-25)
    if (components != 3) {
        stbi_image_free(bytes);
        return std::nullopt;
    }
    std::vector<Color> texels(size_t(width) * size_t(height));
    for (size_t i = 0; i < texels.size(); ++i)
        texels[i] = Color{ bytes[3 * i] / 255.0f, bytes[3 * i + 1] / 255.0f, bytes[3 * i + 2] / 255.0f };
    stbi_image_free(bytes);
    return Image{ width, height, std::move(texels) };
#else
    (void)filename;
    return std::nullopt;
#endif
}

} // namespace crt
