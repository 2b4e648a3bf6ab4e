#pragma once
// Public scene types of the render API, same names and fields as the reference
// (src/core/crt_scene.h:16-30, crt_vertex.h:7-11, crt_triangle.h:14-46, crt_material.h:5-16, crt_light.h:10-16,
// crt_texture.h:8-45, crt_image.h:9-27, crt_camera.h:11-70, crt_acceleration_tree.h:12-26, crt_aabb.h:11-45),
// re-laid-out for a GPU consumer: triangles reference vertices by INDEX (the reference stores raw pointers and
// copies every triangle into each leaf it straddles), the tree is one node array plus one leaf-reference
// array, and a Scene carries a handle to its flattened, device-resident twin.

#include <array>
#include <cstdint>
#include <limits>
#include <memory>
#include <numbers>
#include <utility>
#include <vector>

#include "crt_math.h"

namespace crt {

using Color = Vector;

struct Image {
    int width, height;
    std::vector<Color> buffer; // row-major, row 0 = top

    Image(int w, int h) : width(w), height(h), buffer(size_t(w) * size_t(h)) {}
    Image(int w, int h, std::vector<Color> &&pixels) : width(w), height(h), buffer(std::move(pixels)) {}
};

struct Vertex {
    Vector position{};
    Vector normal{};
    Vector uv{};
};

struct TriangleFlags {
    uint8_t smooth_shading : 1;
    uint8_t back_face_culling : 1;
};

// counter-clockwise; v0..v2 index Scene::vertices
struct Triangle {
    int v0, v1, v2;
    Vector face_normal;
    int material_index;
    TriangleFlags flags;
};

enum class MaterialType { Diffuse, Reflective, Refractive, Constant };

struct Material {
    MaterialType type;
    int albedo_map_texture_index;
    float ior;
};

struct Light {
    float intensity;
    Vector position;
};

enum class TextureType { Albedo, Edges, Checker, Bitmap };

struct Texture {
    TextureType type{ TextureType::Albedo };
    Color color_a{};   // albedo | edge_color | color_A
    Color color_b{};   // inner_color | color_B
    float scalar{};    // edge_width | square_size
    std::shared_ptr<const Image> image; // bitmap texels as float rgb = byte / 255.0f
};

struct AABB {
    Vector min, max;

    static constexpr AABB vacuum() {
        constexpr float inf = std::numeric_limits<float>::infinity();
        return { { inf, inf, inf }, { -inf, -inf, -inf } };
    }
    constexpr std::pair<AABB, AABB> split(unsigned axis) const {
        AABB lo = *this, hi = *this;
        const float mid = (min.data[axis] + max.data[axis]) * 0.5f;
        lo.max.data[axis] = mid;
        hi.min.data[axis] = mid;
        return { lo, hi };
    }
    constexpr bool intersects(const AABB &o) const {
        for (int a = 0; a < 3; ++a)
            if (o.min.data[a] > max.data[a] || o.max.data[a] < min.data[a])
                return false;
        return true;
    }
};

inline constexpr int MAX_ACCELERATION_TREE_DEPTH = 39;
inline constexpr int MAX_BOX_TRIANGLE_COUNT = 16;

struct AccelerationTreeNode {
    AABB bounds;
    std::array<int, 2> children_indices{ -1, -1 };
    int parent_index{ -1 };
    int first_reference{ 0 }; // into AccelerationTree::references
    int reference_count{ 0 };
    int depth{ 0 };
    constexpr bool is_leaf() const noexcept { return reference_count > 0; }
};

struct AccelerationTree {
    std::vector<AccelerationTreeNode> nodes; // DFS pre-order, child0's subtree before child1's node
    std::vector<uint32_t> references;        // triangle indices, leaf by leaf, parse order within a leaf
    int max_depth{ 0 };

    size_t size() const { return nodes.size(); }
    bool empty() const { return nodes.empty(); }
    const AccelerationTreeNode &operator[](size_t i) const { return nodes[i]; }
};

namespace acceleration_tree {
AccelerationTree build(const std::vector<Vertex> &vertices, const std::vector<Triangle> &triangles);
}

class Camera {
public:
    Camera(int resolution_x, int resolution_y, const Transform &transform = {}) : Camera(resolution_x, resolution_y, 90.0f, transform) {}
    Camera(int resolution_x, int resolution_y, float fov_degrees, const Transform &transform = {})
        : m_resolution_x(resolution_x), m_resolution_y(resolution_y), m_fov_radians(fov_degrees * std::numbers::pi_v<float> / 180.0f), m_transform(transform) {}

    void dolly(float d) { m_transform.translate_local({ 0.0f, 0.0f, d }); }
    void truck(float d) { m_transform.translate_local({ d, 0.0f, 0.0f }); }
    void pedestal(float d) { m_transform.translate_local({ 0.0f, d, 0.0f }); }
    void pan(float a) { m_transform.rotate_y(a); }
    void tilt(float a) { m_transform.rotate_x(a); }
    void roll(float a) { m_transform.rotate_z(a); }
    void pan_around(float a, const Vector &anchor) { m_transform.rotate_y_around(a, anchor); }
    void tilt_around(float a, const Vector &anchor) { m_transform.rotate_x_around(a, anchor); }

    constexpr int resolution_x() const { return m_resolution_x; }
    constexpr int resolution_y() const { return m_resolution_y; }
    constexpr float fov_radians() const { return m_fov_radians; }
    constexpr const Transform &transform() const { return m_transform; }
    void set_resolution(int w, int h) { m_resolution_x = w, m_resolution_y = h; }
    void set_transform(const Transform &t) { m_transform = t; }
    void set_fov_degrees(float deg) { m_fov_radians = deg * std::numbers::pi_v<float> / 180.0f; }

private:
    int m_resolution_x, m_resolution_y;
    float m_fov_radians;
    Transform m_transform;
};

inline constexpr int DEFAULT_SCENE_BUCKET_SIZE = 24;

struct DeviceScene; // flattened arrays + CUDA context (crt_device_scene.h); created on first use

struct Scene {
    Color background_color;
    Camera camera;
    std::vector<Vertex> vertices;
    std::vector<Triangle> triangles; // parse order (the reference moves these into the tree)
    // the reference's acceleration tree (crt_acceleration_tree.cpp), built on first use: only CRT_TRAVERSAL_LITERAL
    // and CRT_WALK_TREE=reference read it, the default walk uses the SAH tree of crt_kdbuild.cpp
    const AccelerationTree &reference_tree() const {
        if (!m_reference_tree)
            m_reference_tree = std::make_shared<const AccelerationTree>(acceleration_tree::build(vertices, triangles));
        return *m_reference_tree;
    }
    mutable std::shared_ptr<const AccelerationTree> m_reference_tree;
    std::vector<Light> lights;
    std::vector<Texture> textures;
    std::vector<Material> materials;
    int bucket_size{ DEFAULT_SCENE_BUCKET_SIZE };
    uint8_t gi_on : 1 { 0 };
    uint8_t reflections_on : 1 { 1 };
    uint8_t refractions_on : 1 { 1 };
    mutable std::shared_ptr<DeviceScene> device; // lazily created by render_image / prepare_device
};

} // namespace crt
