// crt::render_image on the GPU: flatten (once) -> upload (once) -> kernels -> read back into Image.
// Replaces the bucket queue + thread pool of the reference (src/core/crt_renderer.cpp:147-199).  bucket_size only
// decides whether the frame gets any bucket at all (an image smaller than half a bucket stays black, as there).
//
// What is cached and what is not: the GEOMETRY (vertices, triangles, both trees, bitmap texels) is flattened and
// uploaded on the first render and kept in Scene::device; the camera, the lights, materials, texture records,
// background colour, gi / reflections / refractions flags and bucket_size are re-read from the Scene on EVERY call,
// like the reference does (it reads the Scene on every call).  After editing vertices / triangles / bitmap images, or
// changing the NUMBER of lights, materials or textures, drop the cache: scene.device.reset().
#include "crt_renderer.h"

#include <chrono>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/crt_cuda.h"
#include "crt_flatten.h"

namespace crt {

struct DeviceScene {
    FlatScene flat;
    crt_cuda_ctx *ctx{ nullptr };     // one GPU ($CRT_DEVICE)
    crt_cuda_group *group{ nullptr }; // several GPUs ($CRT_DEVICES=0,1,...): row bands, frame assembled on the first
    ~DeviceScene() {
        crt_cuda_destroy(ctx);
        crt_cuda_group_destroy(group);
    }
};

namespace {

thread_local RenderBreakdown g_breakdown;

int env_int(const char *name, int fallback) {
    const char *v = std::getenv(name);
    return v && *v ? std::atoi(v) : fallback;
}

std::vector<int> env_devices() {
    std::vector<int> out;
    const char *v = std::getenv("CRT_DEVICES");
    for (const char *p = v; p && *p;) {
        char *end = nullptr;
        const long d = std::strtol(p, &end, 10);
        if (end == p)
            break;
        out.push_back(int(d));
        p = *end == ',' ? end + 1 : end;
    }
    return out;
}

// the reference's tree is built, flattened and uploaded only when the literal traversal is asked for
bool wants_reference_tree() { return env_int("CRT_TRAVERSAL", CRT_TRAVERSAL_KD) == CRT_TRAVERSAL_LITERAL; }

DeviceScene &device_scene(const Scene &scene) {
    if (scene.device)
        return *scene.device;
    auto dev = std::make_shared<DeviceScene>();
    if (const std::vector<int> devices = env_devices(); devices.size() > 1) {
        const int rc = crt_cuda_group_create(devices.data(), int(devices.size()), &dev->group);
        if (rc != CRT_OK)
            throw std::runtime_error("crt::render_image: cannot create CUDA contexts with peer access for $CRT_DEVICES (code " + std::to_string(rc) +
                                     "); this build renders on the GPU only");
        flatten_scene(scene, dev->flat, wants_reference_tree());
        if (crt_cuda_group_upload_scene(dev->group, &dev->flat.view) != CRT_OK)
            throw std::runtime_error(std::string("crt::render_image: scene upload failed: ") + crt_cuda_group_last_error(dev->group));
        scene.device = std::move(dev);
        return *scene.device;
    }
    const int rc = crt_cuda_create(env_int("CRT_DEVICE", 0), &dev->ctx);
    if (rc != CRT_OK)
        throw std::runtime_error("crt::render_image: cannot create a CUDA context (code " + std::to_string(rc) +
                                 "); this build renders on the GPU only");
    flatten_scene(scene, dev->flat, wants_reference_tree());
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
    crt_cuda_group *group = scene.device->group;
    size_t length = 0;
    if ((group ? crt_cuda_group_encode_ppm(group, nullptr, 0, &length) : crt_cuda_encode_ppm(ctx, nullptr, 0, &length)) != CRT_OK)
        return false;
    std::string text(length, '\0');
    if ((group ? crt_cuda_group_encode_ppm(group, text.data(), text.size(), &length) : crt_cuda_encode_ppm(ctx, text.data(), text.size(), &length)) != CRT_OK)
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

    // lights, materials, texture records, background, flags: whatever the caller changed since the last call
    flatten_shading(scene, dev.flat, false);
    if ((dev.group ? crt_cuda_group_update_shading(dev.group, &dev.flat.view) : crt_cuda_update_shading(dev.ctx, &dev.flat.view)) != CRT_OK) {
        // a count changed: the cached upload no longer describes this Scene
        scene.device.reset();
        return render_image(scene, settings);
    }
    // the camera may have been moved since the upload (Camera::pan_around & co.): resend the 72 B block
    const crt_camera cam = flatten_camera(scene.camera);
    crt_render_stats stats{};
    if (dev.group) {
        // the quantised frame is assembled on the first device as well, so that write_ppm_device can format it there
        std::vector<uint8_t> rgb8(size_t(result.width) * size_t(result.height) * 3);
        if (crt_cuda_group_set_camera(dev.group, &cam) != CRT_OK ||
            crt_cuda_group_render(dev.group, &params, reinterpret_cast<float *>(result.buffer.data()), rgb8.data(), &stats) != CRT_OK)
            throw std::runtime_error(std::string("crt::render_image: ") + crt_cuda_group_last_error(dev.group));
    } else if (crt_cuda_set_camera(dev.ctx, &cam) != CRT_OK ||
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
