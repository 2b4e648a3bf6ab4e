// crt::render_image on the GPU: flatten (once) -> upload (once) -> kernels -> read back into Image.
// Replaces the bucket queue + thread pool of the reference (src/core/crt_renderer.cpp:147-199); bucket_size has
// no semantic effect (pixels are independent) and is ignored.
#include "crt_renderer.h"

#include <chrono>
#include <cstdlib>
#include <stdexcept>
#include <string>

#include "../../include/crt_cuda.h"
#include "crt_flatten.h"

namespace crt {

struct DeviceScene {
    FlatScene flat;
    crt_cuda_ctx *ctx{ nullptr };
    ~DeviceScene() { crt_cuda_destroy(ctx); }
};

namespace {

thread_local RenderBreakdown g_breakdown;

int env_int(const char *name, int fallback) {
    const char *v = std::getenv(name);
    return v && *v ? std::atoi(v) : fallback;
}

DeviceScene &device_scene(const Scene &scene) {
    if (scene.device)
        return *scene.device;
    auto dev = std::make_shared<DeviceScene>();
    const int rc = crt_cuda_create(env_int("CRT_DEVICE", 0), &dev->ctx);
    if (rc != CRT_OK)
        throw std::runtime_error("crt::render_image: cannot create a CUDA context (code " + std::to_string(rc) +
                                 "); this build renders on the GPU only");
    flatten_scene(scene, dev->flat);
    if (crt_cuda_upload_scene(dev->ctx, &dev->flat.view) != CRT_OK)
        throw std::runtime_error(std::string("crt::render_image: scene upload failed: ") + crt_cuda_last_error(dev->ctx));
    scene.device = std::move(dev);
    return *scene.device;
}

} // namespace

void prepare_device(const Scene &scene) { (void)device_scene(scene); }

RenderBreakdown last_render_breakdown() { return g_breakdown; }

bool write_ppm_device(const Scene &scene, std::ostream &os) {
    if (!scene.device)
        return false;
    crt_cuda_ctx *ctx = scene.device->ctx;
    size_t length = 0;
    if (crt_cuda_encode_ppm(ctx, nullptr, 0, &length) != CRT_OK)
        return false;
    std::string text(length, '\0');
    if (crt_cuda_encode_ppm(ctx, text.data(), text.size(), &length) != CRT_OK)
        return false;
    os.write(text.data(), std::streamsize(length));
    return true;
}

Image render_image(const Scene &scene, const RendererSettings &settings) {
    using clock = std::chrono::steady_clock;
    const auto t0 = clock::now();
    DeviceScene &dev = device_scene(scene);
    const auto t1 = clock::now();

    Image result{ scene.camera.resolution_x(), scene.camera.resolution_y() };

    crt_render_params params;
    crt_render_params_default(&params);
    params.max_ray_depth = settings.max_ray_depth;
    params.diffuse_reflection_ray_count = settings.diffuse_reflection_ray_count;
    params.shadow_bias = settings.shadow_bias;
    params.reflection_bias = settings.reflection_bias;
    params.diffuse_reflection_bias = settings.diffuse_reflection_bias;
    params.refraction_bias = settings.refraction_bias;
    // new knobs come from the environment so the CLI / _crt signatures stay the reference's
    params.traversal = env_int("CRT_TRAVERSAL", CRT_TRAVERSAL_KD);
    params.pipeline = env_int("CRT_PIPELINE", CRT_PIPELINE_AUTO);
    params.shadows = env_int("CRT_SHADOWS", 0);

    // the camera may have been moved since the upload (Camera::pan_around & co.): resend the 72 B block
    const crt_camera cam = flatten_camera(scene.camera);
    crt_render_stats stats{};
    if (crt_cuda_set_camera(dev.ctx, &cam) != CRT_OK ||
        crt_cuda_render(dev.ctx, &params, reinterpret_cast<float *>(result.buffer.data()), nullptr, &stats) != CRT_OK)
        throw std::runtime_error(std::string("crt::render_image: ") + crt_cuda_last_error(dev.ctx));

    const auto t2 = clock::now();
    g_breakdown.flatten_upload_s = std::chrono::duration<double>(t1 - t0).count();
    g_breakdown.total_s = std::chrono::duration<double>(t2 - t0).count();
    g_breakdown.gpu_ms = stats.gpu_ms;
    g_breakdown.kernel_launches = stats.kernel_launches;
    return result;
}

} // namespace crt
