#pragma once
// The operator boundary, same signature as the reference (src/core/crt_renderer.h:10-27).
#include <cstdint>
#include <ostream>

#include "crt_scene.h"

namespace crt {

inline constexpr int DEFAULT_MAX_RAY_DEPTH = 3;
inline constexpr int DEFAULT_DIFFUSE_REFLECTION_RAY_COUNT = 4;
inline constexpr float DEFAULT_SHADOW_BIAS = 1e-2f;
inline constexpr float DEFAULT_REFLECTION_BIAS = 1e-2f;
inline constexpr float DEFAULT_DIFFUSE_REFLECTION_BIAS = 1e-2f;
inline constexpr float DEFAULT_REFRACTION_BIAS = 1e-2f;

struct RendererSettings {
    uint32_t max_ray_depth{ DEFAULT_MAX_RAY_DEPTH };
    uint32_t diffuse_reflection_ray_count{ DEFAULT_DIFFUSE_REFLECTION_RAY_COUNT };
    float shadow_bias{ DEFAULT_SHADOW_BIAS };
    float reflection_bias{ DEFAULT_REFLECTION_BIAS };
    float diffuse_reflection_bias{ DEFAULT_DIFFUSE_REFLECTION_BIAS };
    float refraction_bias{ DEFAULT_REFRACTION_BIAS };
};

// Renders on the GPU (device from $CRT_DEVICE, default 0).  Throws std::runtime_error when no CUDA device is
// usable: there is deliberately no CPU path.  The flattened scene is uploaded on first use and stays resident
// on the Scene (Scene::device), so repeated calls only pay launch + read-back.
Image render_image(const Scene &scene, const RendererSettings &settings);

// Optional: flatten + upload now (e.g. right after loading) so the first render_image carries no setup cost.
void prepare_device(const Scene &scene);

// P3 text (crt::write_ppm's bytes, src/core/crt_image_ppm.cpp:9-23) of the frame the last render_image of `scene`
// left on the device, formatted by the GPU (csrc/crt_ppm.cuh).  False when the scene has no rendered frame.
bool write_ppm_device(const Scene &scene, std::ostream &os);

// timings of the last render_image on this thread, for the CLI's optional breakdown ($CRT_STATS=1)
struct RenderBreakdown {
    double flatten_upload_s{ 0 }, gpu_ms{ 0 }, total_s{ 0 };
    int kernel_launches{ 0 };
};
RenderBreakdown last_render_breakdown();

} // namespace crt
