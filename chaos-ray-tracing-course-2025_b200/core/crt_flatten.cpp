#include "crt_flatten.h"
#include "crt_kdbuild.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <unordered_map>

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

// ---------------------------------------------------------------------------------------------------------------
// The tree the k-d walk traverses.
//
// The reference builder (src/core/crt_acceleration_tree.cpp:31-85) splits at the spatial median of axis depth % 3
// until a node holds <= 16 triangles, whatever the split achieves.  Where triangles are larger than the cells (every
// bundled mesh; the 1 M-triangle height field ends with cells of 0.04 x 0.003 x 0.04 around triangles of 0.03) most
// splits separate nothing: both children keep nearly all of the parent's triangles, 56 % of the height field's leaves
// hold a list identical to another leaf's, and a ray walks ten nodes per leaf to re-find triangles it has tested.
// The closest hit does not depend on how space is cut (the walk keeps min (t, parse index) over everything it tests),
// so the flattener is free to UNDO splits that do not pay:
//   * bottom-up, a subtree is replaced by one leaf holding the union of its triangles when the expected cost of the
//     leaf (fixed + per-triangle) is not above the expected cost of walking the subtree (node + children weighted by
//     the surface-area ratio of their boxes: the standard SAH estimate, evaluated on the reference's own boxes);
//   * identical leaf lists are stored once.
// Topology, split planes and boxes of the surviving nodes are the reference's; CRT_TREE_COLLAPSE=0 keeps every node.
// ---------------------------------------------------------------------------------------------------------------
namespace {

float env_float(const char *name, float fallback) {
    const char *v = std::getenv(name);
    return v && *v ? float(std::atof(v)) : fallback;
}

float half_area(const AABB &b) {
    const float dx = b.max.x - b.min.x, dy = b.max.y - b.min.y, dz = b.max.z - b.min.z;
    return dx * dy + dy * dz + dz * dx;
}

struct ListKey {
    const uint32_t *data;
    uint32_t count;
    bool operator==(const ListKey &o) const { return count == o.count && std::memcmp(data, o.data, size_t(count) * 4) == 0; }
};
struct ListHash {
    size_t operator()(const ListKey &k) const {
        uint64_t h = 1469598103934665603ull ^ k.count;
        for (uint32_t i = 0; i < k.count; ++i)
            h = (h ^ k.data[i]) * 1099511628211ull;
        return size_t(h);
    }
};

} // namespace

FastTreeStats build_fast_tree(const AccelerationTree &tree, const std::vector<Triangle> &triangles, FlatScene &out) {
    FastTreeStats stats;
    out.fast_nodes.clear();
    out.fast_refs.clear();
    const size_t n = tree.size();
    if (n == 0)
        return stats;
    const bool collapse = env_float("CRT_TREE_COLLAPSE", 1.0f) != 0.0f;
    // costs in units of one node step of the trace kernel (instruction counts of the stages)
    const float c_node = env_float("CRT_SAH_NODE", 1.0f), c_leaf = env_float("CRT_SAH_LEAF", 2.0f), c_tri = env_float("CRT_SAH_TRI", 2.8f);
    const uint32_t max_list = uint32_t(env_float("CRT_SAH_MAX_LEAF", 64.0f));

    // per node: expected cost of its subtree as decided so far, and (when small) the sorted union of its triangles
    std::vector<float> cost(n, 0.0f);
    std::vector<uint8_t> is_leaf(n, 0);
    std::vector<uint32_t> arena;           // unions, appended as they are formed
    std::vector<uint64_t> span(n, 0);      // offset << 8 | count ; count 255 = "too large, not kept"
    arena.reserve(tree.references.size() * 2);
    auto set_span = [&](size_t i, size_t offset, size_t count) { span[i] = (uint64_t(offset) << 8) | uint64_t(count); };
    auto span_count = [&](size_t i) { return uint32_t(span[i] & 0xff); };
    auto span_data = [&](size_t i) { return arena.data() + (span[i] >> 8); };
    std::vector<uint32_t> merged;
    for (size_t k = n; k-- > 0;) { // children have larger indices than their parent (DFS pre-order)
        const AccelerationTreeNode &node = tree[k];
        if (node.is_leaf()) {
            const size_t offset = arena.size();
            arena.insert(arena.end(), tree.references.begin() + node.first_reference, tree.references.begin() + node.first_reference + node.reference_count);
            std::sort(arena.begin() + offset, arena.end());
            const size_t count = size_t(node.reference_count);
            is_leaf[k] = 1;
            cost[k] = c_leaf + c_tri * float(count);
            if (count < 255)
                set_span(k, offset, count);
            else { // cannot take part in a merge; emitted from the reference's own list
                arena.resize(offset);
                set_span(k, 0, 255);
            }
            continue;
        }
        const int c0 = node.children_indices[0], c1 = node.children_indices[1];
        const float area = half_area(node.bounds);
        float inner = c_node;
        bool small = true;
        merged.clear();
        for (int c : { c0, c1 }) {
            if (c < 0)
                continue;
            const float p = area > 0.0f ? std::min(1.0f, half_area(tree[size_t(c)].bounds) / area) : 1.0f;
            inner += p * cost[size_t(c)];
            if (span_count(size_t(c)) == 255)
                small = false;
        }
        if (small) {
            if (c0 >= 0 && c1 >= 0) {
                merged.resize(span_count(size_t(c0)) + span_count(size_t(c1)));
                merged.resize(size_t(std::set_union(span_data(size_t(c0)), span_data(size_t(c0)) + span_count(size_t(c0)), span_data(size_t(c1)),
                                                    span_data(size_t(c1)) + span_count(size_t(c1)), merged.begin()) -
                                     merged.begin()));
            } else if (c0 >= 0 || c1 >= 0) {
                const size_t c = size_t(c0 >= 0 ? c0 : c1);
                merged.assign(span_data(c), span_data(c) + span_count(c));
            }
            small = merged.size() <= max_list && merged.size() < 255;
        }
        cost[k] = inner;
        set_span(k, 0, 255);
        if (small) {
            const size_t offset = arena.size();
            arena.insert(arena.end(), merged.begin(), merged.end());
            set_span(k, offset, merged.size());
            const float as_leaf = c_leaf + c_tri * float(merged.size());
            if (collapse && !merged.empty() && as_leaf <= inner) {
                is_leaf[k] = 1;
                cost[k] = as_leaf;
            }
        }
    }

    // emit: DFS pre-order from the root, child0's subtree first (the reference's order), leaf lists deduplicated
    std::unordered_map<ListKey, uint32_t, ListHash> seen;
    std::vector<uint32_t> list;
    struct Item {
        int node, parent_slot, which, depth;
    };
    std::vector<Item> todo{ { 0, -1, 0, 0 } };
    while (!todo.empty()) {
        const Item it = todo.back();
        todo.pop_back();
        const size_t k = size_t(it.node);
        const AccelerationTreeNode &node = tree[k];
        const int self = int(out.fast_nodes.size());
        if (it.parent_slot >= 0)
            (it.which == 0 ? out.fast_nodes[size_t(it.parent_slot)].a : out.fast_nodes[size_t(it.parent_slot)].b) = self;
        stats.max_depth = std::max(stats.max_depth, it.depth);
        if (is_leaf[k]) {
            const uint32_t *data;
            uint32_t count;
            if (span_count(k) == 255) {
                data = tree.references.data() + node.first_reference, count = uint32_t(node.reference_count);
            } else {
                data = span_data(k), count = span_count(k);
            }
            auto found = seen.find(ListKey{ data, count });
            uint32_t first;
            if (found != seen.end()) {
                first = found->second;
            } else {
                first = uint32_t(out.fast_refs.size());
                for (uint32_t i = 0; i < count; ++i)
                    out.fast_refs.push_back(data[i] | (triangles[data[i]].flags.back_face_culling ? CRT_REF_CULL_BIT : 0u));
                seen.emplace(ListKey{ data, count }, first); // key points into arena / tree.references: both stay put
            }
            out.fast_nodes.push_back({ 0.0f, int32_t(first), int32_t(count), 3 });
            ++stats.leaves;
            stats.max_leaf = std::max(stats.max_leaf, int(count));
        } else {
            const int axis = node.depth % 3;
            out.fast_nodes.push_back({ (node.bounds.min.data[axis] + node.bounds.max.data[axis]) * 0.5f, -1, -1, axis });
            // pushed in reverse so that child0 is emitted first
            if (node.children_indices[1] >= 0)
                todo.push_back({ node.children_indices[1], self, 1, it.depth + 1 });
            if (node.children_indices[0] >= 0)
                todo.push_back({ node.children_indices[0], self, 0, it.depth + 1 });
        }
    }
    stats.nodes = int(out.fast_nodes.size());
    stats.refs = int(out.fast_refs.size());
    return stats;
}

// lights, materials, texture records (+ bitmap texels when with_texels) and the scalar block: everything of a Scene that
// can change between two render_image calls without touching the geometry
void flatten_shading(const Scene &scene, FlatScene &out, bool with_texels) {
    out.lights.clear(), out.materials.clear(), out.textures.clear();
    if (with_texels)
        out.texels.clear();
    for (const Light &l : scene.lights)
        out.lights.push_back({ l.position.x, l.position.y, l.position.z, l.intensity });
    for (const Material &m : scene.materials)
        out.materials.push_back({ int32_t(m.type), m.albedo_map_texture_index, m.ior, 0 });
    size_t texel_cursor = 0;
    for (const Texture &t : scene.textures) {
        crt_texture f{};
        f.type = int32_t(t.type);
        std::memcpy(f.color0, t.color_a.data, 12);
        std::memcpy(f.color1, t.color_b.data, 12);
        f.param = t.scalar;
        if (t.type == TextureType::Bitmap && t.image) {
            f.width = t.image->width;
            f.height = t.image->height;
            f.texel_offset = int32_t(texel_cursor);
            texel_cursor += t.image->buffer.size();
            if (with_texels)
                for (const Color &c : t.image->buffer)
                    out.texels.push_back({ c.x, c.y, c.z, 0.0f });
        }
        out.textures.push_back(f);
    }
    crt_flat_scene &v = out.view;
    std::memcpy(v.background, scene.background_color.data, 12);
    v.gi_on = scene.gi_on;
    v.reflections_on = scene.reflections_on;
    v.refractions_on = scene.refractions_on;
    v.bucket_size = scene.bucket_size;
    v.n_lights = int32_t(out.lights.size());
    v.n_materials = int32_t(out.materials.size());
    v.n_textures = int32_t(out.textures.size());
    v.lights = out.lights.data();
    v.materials = out.materials.data();
    v.textures = out.textures.data();
}

void flatten_scene(const Scene &scene, FlatScene &out, bool with_reference_tree) {
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

    flatten_shading(scene, out, true);

    crt_flat_scene &v = out.view;
    v.camera = flatten_camera(scene.camera);
    // root box = min / max over every triangle vertex, starting from the empty box (crt_acceleration_tree.cpp:89-94)
    AABB root = AABB::vacuum();
    for (size_t i = 0; i < 3 * n_tri; ++i) {
        const crt_float4 &p = out.tri_verts[i];
        root.min.x = std::min(root.min.x, p.x), root.min.y = std::min(root.min.y, p.y), root.min.z = std::min(root.min.z, p.z);
        root.max.x = std::max(root.max.x, p.x), root.max.y = std::max(root.max.y, p.y), root.max.z = std::max(root.max.z, p.z);
    }
    std::memcpy(v.root_min, root.min.data, 12);
    std::memcpy(v.root_max, root.max.data, 12);
    v.n_triangles = int32_t(n_tri);
    v.n_vertices = int32_t(n_vert);
    v.n_lights = int32_t(out.lights.size());
    v.n_materials = int32_t(out.materials.size());
    v.n_textures = int32_t(out.textures.size());
    v.n_texels = int32_t(out.texels.size());
    // the walk tree: an SAH k-d tree of our own (crt_kdbuild.cpp) unless CRT_WALK_TREE=reference asks for the
    // reference's tree with the non-paying splits merged (build_fast_tree above)
    const char *which = std::getenv("CRT_WALK_TREE");
    const bool walk_reference = which && std::strcmp(which, "reference") == 0;
    if (with_reference_tree || walk_reference)
        add_reference_tree(scene, out);
    if (walk_reference) {
        const FastTreeStats fast = build_fast_tree(scene.reference_tree(), scene.triangles, out);
        v.fast_max_depth = fast.max_depth;
        v.n_fast_nodes = fast.nodes;
        v.n_fast_refs = fast.refs;
    } else {
        std::vector<uint8_t> cull(n_tri);
        for (size_t i = 0; i < n_tri; ++i)
            cull[i] = scene.triangles[i].flags.back_face_culling ? 1 : 0;
        const KdBuildStats fast = build_sah_tree(out.tri_verts.data(), n_tri, cull.data(), v.root_min, v.root_max, KdBuildParams::from_environment(),
                                                 out.fast_nodes, out.fast_refs);
        v.fast_max_depth = fast.max_depth;
        v.n_fast_nodes = fast.nodes;
        v.n_fast_refs = fast.refs;
    }
    v.fast_nodes = out.fast_nodes.data();
    v.fast_refs = out.fast_refs.data();
    v.tri_verts = out.tri_verts.data();
    v.tri_meta = out.tri_meta.data();
    v.vert_normal = out.vert_normal.data();
    v.vert_uv = out.vert_uv.data();
    v.lights = out.lights.data();
    v.materials = out.materials.data();
    v.textures = out.textures.data();
    v.texels = out.texels.data();
}

void add_reference_tree(const Scene &scene, FlatScene &out) {
    if (!out.box_nodes.empty())
        return;
    const AccelerationTree &tree = scene.reference_tree();
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

    crt_flat_scene &v = out.view;
    v.n_nodes = int32_t(tree.size());
    v.n_refs = int32_t(out.refs.size());
    v.max_depth = tree.max_depth;
    v.box_nodes = out.box_nodes.data();
    v.kd_nodes = out.kd_nodes.data();
    v.refs = out.refs.data();
}

} // namespace crt
