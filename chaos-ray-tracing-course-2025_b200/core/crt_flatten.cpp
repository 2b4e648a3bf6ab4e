#include "crt_flatten.h"

#include <cmath>
#include <cstring>

namespace crt {

crt_camera flatten_camera(const Camera &camera) {
    crt_camera c{};
    const Transform &t = camera.transform();
    std::memcpy(c.position, t.location.data, sizeof c.position);
    std::memcpy(c.rotation, t.rotation.data, sizeof c.rotation);
    // per-frame constants of Camera::generate_ray (src/core/crt_camera.cpp:21-26), evaluated once with the host
    // libm so the kernel multiplies by exactly the value the reference multiplies by at every pixel
    c.tan_half_fov = std::tan(camera.fov_radians() * 0.5f);
    c.aspect = float(camera.resolution_x()) / camera.resolution_y();
    c.width = camera.resolution_x();
    c.height = camera.resolution_y();
    return c;
}

void flatten_scene(const Scene &scene, FlatScene &out) {
    out = FlatScene{};
    const size_t n_tri = scene.triangles.size(), n_vert = scene.vertices.size();

    out.tri_verts.resize(3 * n_tri);
    out.tri_meta.resize(4 * n_tri);
    for (size_t i = 0; i < n_tri; ++i) {
        const Triangle &t = scene.triangles[i];
        const Vector &a = scene.vertices[size_t(t.v0)].position, &b = scene.vertices[size_t(t.v1)].position,
                     &c = scene.vertices[size_t(t.v2)].position;
        out.tri_verts[3 * i + 0] = { a.x, a.y, a.z, t.face_normal.x };
        out.tri_verts[3 * i + 1] = { b.x, b.y, b.z, t.face_normal.y };
        out.tri_verts[3 * i + 2] = { c.x, c.y, c.z, t.face_normal.z };
        out.tri_meta[4 * i + 0] = t.v0;
        out.tri_meta[4 * i + 1] = t.v1;
        out.tri_meta[4 * i + 2] = t.v2;
        out.tri_meta[4 * i + 3] = (t.material_index & CRT_META_MATERIAL_MASK) | (t.flags.smooth_shading ? CRT_META_SMOOTH_BIT : 0) |
                                  (t.flags.back_face_culling ? CRT_META_CULL_BIT : 0);
    }

    out.vert_normal.resize(n_vert);
    out.vert_uv.resize(2 * n_vert);
    for (size_t i = 0; i < n_vert; ++i) {
        const Vertex &v = scene.vertices[i];
        out.vert_normal[i] = { v.normal.x, v.normal.y, v.normal.z, 0.0f };
        out.vert_uv[2 * i] = v.uv.x; // uv.z is interpolated by the reference but never read (crt_texture.cpp:29-44)
        out.vert_uv[2 * i + 1] = v.uv.y;
    }

    const AccelerationTree &tree = scene.acceleration_tree;
    out.box_nodes.resize(tree.size());
    out.kd_nodes.resize(tree.size());
    for (size_t i = 0; i < tree.size(); ++i) {
        const AccelerationTreeNode &n = tree[i];
        crt_box_node &b = out.box_nodes[i];
        std::memcpy(b.min, n.bounds.min.data, 12);
        std::memcpy(b.max, n.bounds.max.data, 12);
        crt_kd_node &k = out.kd_nodes[i];
        if (n.is_leaf()) {
            b.left = -n.first_reference - 2;
            b.right = n.reference_count;
            k = { 0.0f, n.first_reference, n.reference_count, 3 };
        } else {
            b.left = n.children_indices[0];
            b.right = n.children_indices[1];
            const int axis = n.depth % 3;
            k = { (n.bounds.min.data[axis] + n.bounds.max.data[axis]) * 0.5f, n.children_indices[0], n.children_indices[1], axis };
        }
    }
    out.refs.resize(tree.references.size());
    for (size_t i = 0; i < out.refs.size(); ++i) {
        const uint32_t t = tree.references[i];
        out.refs[i] = t | (scene.triangles[t].flags.back_face_culling ? CRT_REF_CULL_BIT : 0u);
    }

    for (const Light &l : scene.lights)
        out.lights.push_back({ l.position.x, l.position.y, l.position.z, l.intensity });
    for (const Material &m : scene.materials)
        out.materials.push_back({ int32_t(m.type), m.albedo_map_texture_index, m.ior, 0 });
    for (const Texture &t : scene.textures) {
        crt_texture f{};
        f.type = int32_t(t.type);
        std::memcpy(f.color0, t.color_a.data, 12);
        std::memcpy(f.color1, t.color_b.data, 12);
        f.param = t.scalar;
        if (t.type == TextureType::Bitmap && t.image) {
            f.width = t.image->width;
            f.height = t.image->height;
            f.texel_offset = int32_t(out.texels.size());
            for (const Color &c : t.image->buffer)
                out.texels.push_back({ c.x, c.y, c.z, 0.0f });
        }
        out.textures.push_back(f);
    }

    crt_flat_scene &v = out.view;
    v.camera = flatten_camera(scene.camera);
    std::memcpy(v.background, scene.background_color.data, 12);
    v.gi_on = scene.gi_on;
    v.reflections_on = scene.reflections_on;
    v.refractions_on = scene.refractions_on;
    v.bucket_size = scene.bucket_size;
    if (!tree.empty()) {
        std::memcpy(v.root_min, tree[0].bounds.min.data, 12);
        std::memcpy(v.root_max, tree[0].bounds.max.data, 12);
    }
    v.n_triangles = int32_t(n_tri);
    v.n_vertices = int32_t(n_vert);
    v.n_nodes = int32_t(tree.size());
    v.n_refs = int32_t(out.refs.size());
    v.n_lights = int32_t(out.lights.size());
    v.n_materials = int32_t(out.materials.size());
    v.n_textures = int32_t(out.textures.size());
    v.n_texels = int32_t(out.texels.size());
    v.max_depth = tree.max_depth;
    v.tri_verts = out.tri_verts.data();
    v.tri_meta = out.tri_meta.data();
    v.vert_normal = out.vert_normal.data();
    v.vert_uv = out.vert_uv.data();
    v.box_nodes = out.box_nodes.data();
    v.kd_nodes = out.kd_nodes.data();
    v.refs = out.refs.data();
    v.lights = out.lights.data();
    v.materials = out.materials.data();
    v.textures = out.textures.data();
    v.texels = out.texels.data();
}

} // namespace crt
