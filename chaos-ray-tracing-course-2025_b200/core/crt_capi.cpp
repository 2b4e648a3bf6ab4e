// C-ABI, host group (include/crt_cuda.h: crt_scene_*): load + tree build + flatten behind an opaque handle.
#include <filesystem>
#include <fstream>
#include <iterator>
#include <new>
#include <string>

#include "../../include/crt_cuda.h"
#include "crt_flatten.h"
#include "crt_json.h"

struct crt_scene {
    crt::Scene scene;
    crt::FlatScene flat;
    bool flat_valid{ false };
    explicit crt_scene(crt::Scene &&s) : scene(std::move(s)) {}
};

namespace {
thread_local std::string g_error;

int load(std::string_view text, const std::filesystem::path &root, crt_scene **out) {
    std::optional<crt::Scene> scene = crt::json::read_scene_from_memory(text, root);
    if (!scene) {
        g_error = "scene failed validation";
        return CRT_ERR_PARSE;
    }
    *out = new (std::nothrow) crt_scene{ std::move(*scene) };
    return *out ? CRT_OK : CRT_ERR_INVALID;
}
} // namespace

extern "C" {

const char *crt_scene_last_error(void) { return g_error.c_str(); }

int crt_scene_load_file(const char *path, crt_scene **out) {
    if (!path || !out)
        return CRT_ERR_INVALID;
    *out = nullptr;
    std::ifstream in{ path, std::ios::in | std::ios::binary };
    if (!in.is_open()) {
        g_error = std::string("could not open ") + path;
        return CRT_ERR_IO;
    }
    const std::string text{ std::istreambuf_iterator<char>{ in }, std::istreambuf_iterator<char>{} };
    return load(text, std::filesystem::path{ path }.parent_path(), out);
}

int crt_scene_load_json(const char *json, size_t length, const char *asset_root, crt_scene **out) {
    if (!json || !out)
        return CRT_ERR_INVALID;
    *out = nullptr;
    return load(std::string_view{ json, length }, std::filesystem::path{ asset_root ? asset_root : "" }, out);
}

int crt_scene_flatten_ex(const crt_scene *scene, int flags, crt_flat_scene *out) {
    if (!scene || !out)
        return CRT_ERR_INVALID;
    crt_scene *mut = const_cast<crt_scene *>(scene);
    const bool with_tree = (flags & CRT_FLATTEN_REFERENCE_TREE) != 0;
    if (!mut->flat_valid) {
        crt::flatten_scene(mut->scene, mut->flat, with_tree);
        mut->flat_valid = true;
    } else if (with_tree) {
        crt::add_reference_tree(mut->scene, mut->flat); // earlier views stay valid: only the tree arrays are filled in
    }
    mut->flat.view.camera = crt::flatten_camera(mut->scene.camera);
    *out = mut->flat.view;
    if (!with_tree) { // a view without the reference tree, whatever an earlier call built
        out->n_nodes = 0, out->n_refs = 0, out->max_depth = 0;
        out->box_nodes = nullptr, out->kd_nodes = nullptr, out->refs = nullptr;
    }
    return CRT_OK;
}

int crt_scene_flatten(const crt_scene *scene, crt_flat_scene *out) { return crt_scene_flatten_ex(scene, CRT_FLATTEN_REFERENCE_TREE, out); }

int crt_scene_set_camera(crt_scene *scene, const float position[3], const float rotation[9], float fov_degrees) {
    if (!scene || !position || !rotation)
        return CRT_ERR_INVALID;
    crt::Transform t;
    t.location = crt::Vector{ position[0], position[1], position[2] };
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            t.rotation.data[i][j] = rotation[i * 3 + j];
    scene->scene.camera.set_transform(t);
    if (fov_degrees > 0.0f)
        scene->scene.camera.set_fov_degrees(fov_degrees);
    return CRT_OK;
}

int crt_scene_set_resolution(crt_scene *scene, int width, int height) {
    if (!scene || width < 0 || height < 0)
        return CRT_ERR_INVALID;
    scene->scene.camera.set_resolution(width, height);
    return CRT_OK;
}

void crt_scene_free(crt_scene *scene) { delete scene; }

} // extern "C"
