// TEST INFRASTRUCTURE ONLY -- never linked into, imported by, or executed from the product path.
//
// ref_harness: a small driver of OURS that is compiled against the *unmodified* reference headers
// (include root /root/reference/src) and linked with the reference's own objects (oracle/Makefile).
// It never re-implements reference arithmetic: every number it emits is produced by calling the
// reference's public functions
//   crt::json::read_scene_from_istream   (src/core/crt_json.h:11)
//   crt::render_image                    (src/core/crt_renderer.h:27)
//   crt::Camera::generate_ray            (src/core/crt_camera.h:24)
//   crt::intersection::ray_intersect_aabb_p / ray_intersect_triangle   (src/core/crt_intersection.h:24-25)
// The only logic restated here is the *control flow* of ray_intersect_acceleration_tree
// (src/core/crt_intersection.cpp:109-136) and of ray_intersect_triangle_span (:95-107), because the
// reference's Intersection struct carries no triangle identity (crt_intersection.h:13-20): we walk the
// tree in the same order with the same exported predicates so that we can record WHICH triangle won,
// plus the number of box / triangle tests per ray.
//
// usage: ref_harness <scene.crtscene> [--asset-root DIR] [--render OUT.f32] [--repeat K]
//                    [--max-depth N] [--gi-rays N] [--hits OUT.bin] [--rays OUT.f32] [--stride S]
//                    [--dump OUT.bin] [--brute]
// prints one JSON object on stdout.

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include <rapidjson/document.h>

#include "core/crt_intersection.h"
#include "core/crt_json.h"
#include "core/crt_renderer.h"
#include "core/crt_scene.h"

namespace {

struct TriKey {
    int a, b, c;
    bool operator==(const TriKey &o) const { return a == o.a && b == o.b && c == o.c; }
};
struct TriKeyHash {
    size_t operator()(const TriKey &k) const {
        uint64_t h = (uint64_t(uint32_t(k.a)) * 0x9E3779B97F4A7C15ull) ^ (uint64_t(uint32_t(k.b)) * 0xC2B2AE3D27D4EB4Full) ^
                     (uint64_t(uint32_t(k.c)) * 0x165667B19E3779F9ull);
        return size_t(h ^ (h >> 29));
    }
};
using TriIndexMap = std::unordered_map<TriKey, int, TriKeyHash>;

// Parse-order triangle identity: objects in file order, triangles in order within the object
// (crt_json.cpp:176-215 feeds crt_mesh.cpp:19-21 in exactly that order). Key = global vertex indices.
TriIndexMap build_parse_order_map(const std::string &json_text, int &triangle_count) {
    TriIndexMap map;
    rapidjson::Document doc;
    doc.Parse(json_text.c_str(), json_text.size());
    triangle_count = 0;
    if (doc.HasParseError() || !doc.IsObject() || !doc.HasMember("objects"))
        return map;
    int base = 0;
    for (const auto &obj : doc["objects"].GetArray()) {
        const auto &tris = obj["triangles"];
        for (rapidjson::SizeType i = 0; i + 2 < tris.Size(); i += 3) {
            TriKey key{ base + tris[i].GetInt(), base + tris[i + 1].GetInt(), base + tris[i + 2].GetInt() };
            map.emplace(key, triangle_count); // first occurrence wins, like the strict '<' in the span loop
            ++triangle_count;
        }
        base += int(obj["vertices"].Size() / 3);
    }
    return map;
}

int triangle_id(const crt::Scene &scene, const TriIndexMap &map, const crt::Triangle &t) {
    const crt::Vertex *base = scene.vertices.data();
    auto it = map.find(TriKey{ int(t.v0 - base), int(t.v1 - base), int(t.v2 - base) });
    return it == map.end() ? -2 : it->second;
}

struct HitRecord {
    int tri = -1;
    float t = 0.0f;
    int n_node = 0, n_tri = 0;
};

// Same control flow as crt_intersection.cpp:109-136 / :95-107, using the reference's own predicates.
HitRecord walk_tree(const crt::Ray &ray, const crt::Scene &scene, const TriIndexMap &map) {
    using namespace crt::intersection;
    HitRecord rec;
    bool have = false;
    std::vector<int> stack{ 0 };
    while (!stack.empty()) {
        const int node_index = stack.back();
        stack.pop_back();
        const crt::AccelerationTreeNode &node = scene.acceleration_tree[node_index];
        ++rec.n_node;
        if (!ray_intersect_aabb_p(ray, node.bounds))
            continue;
        if (node.is_leaf()) {
            // span loop: first strictly-closer hit inside the leaf ...
            bool leaf_have = false;
            float leaf_t = 0.0f;
            int leaf_tri = -1;
            for (const auto &triangle : node.triangles) {
                ++rec.n_tri;
                if (auto hit = ray_intersect_triangle(ray, triangle)) {
                    if (!leaf_have || hit->distance < leaf_t) {
                        leaf_have = true;
                        leaf_t = hit->distance;
                        leaf_tri = triangle_id(scene, map, triangle);
                    }
                }
            }
            // ... then strictly closer than the best so far (:124)
            if (leaf_have && (!have || leaf_t < rec.t)) {
                have = true;
                rec.t = leaf_t;
                rec.tri = leaf_tri;
            }
        } else {
            if (node.children_indices[0] != -1)
                stack.push_back(node.children_indices[0]);
            if (node.children_indices[1] != -1)
                stack.push_back(node.children_indices[1]);
        }
    }
    return rec;
}

// brute force over the de-duplicated triangles in parse order (the pre-tree path, crt_intersection.cpp:95-107)
HitRecord walk_brute(const crt::Ray &ray, const std::vector<const crt::Triangle *> &unique) {
    using namespace crt::intersection;
    HitRecord rec;
    bool have = false;
    for (size_t i = 0; i < unique.size(); ++i) {
        if (!unique[i])
            continue;
        ++rec.n_tri;
        if (auto hit = ray_intersect_triangle(ray, *unique[i])) {
            if (!have || hit->distance < rec.t) {
                have = true;
                rec.t = hit->distance;
                rec.tri = int(i);
            }
        }
    }
    return rec;
}

template <class T> void put(std::ofstream &os, const T &v) { os.write(reinterpret_cast<const char *>(&v), sizeof(T)); }

void dump_scene(const crt::Scene &scene, const TriIndexMap &map, const std::string &path) {
    std::ofstream os{ path, std::ios::binary };
    const auto &tree = scene.acceleration_tree;
    int32_t n_refs = 0;
    for (const auto &n : tree)
        n_refs += int32_t(n.triangles.size());
    int32_t header[12] = { int32_t(scene.vertices.size()), int32_t(tree.size()), n_refs, int32_t(scene.lights.size()),
                           int32_t(scene.textures.size()), int32_t(scene.materials.size()), scene.bucket_size,
                           int32_t(scene.gi_on), int32_t(scene.reflections_on), int32_t(scene.refractions_on),
                           scene.camera.resolution_x(), scene.camera.resolution_y() };
    os.write(reinterpret_cast<const char *>(header), sizeof(header));
    put(os, scene.background_color);
    for (const auto &v : scene.vertices) {
        put(os, v.position);
        put(os, v.normal);
        put(os, v.uv);
    }
    int32_t first = 0;
    for (const auto &n : tree) {
        put(os, n.bounds.min);
        put(os, n.bounds.max);
        int32_t rec[5] = { n.children_indices[0], n.children_indices[1], n.parent_index, first, int32_t(n.triangles.size()) };
        os.write(reinterpret_cast<const char *>(rec), sizeof(rec));
        first += int32_t(n.triangles.size());
    }
    const crt::Vertex *base = scene.vertices.data();
    for (const auto &n : tree) {
        for (const auto &t : n.triangles) {
            int32_t rec[6] = { int32_t(t.v0 - base), int32_t(t.v1 - base), int32_t(t.v2 - base), t.material_index,
                               int32_t(t.flags.smooth_shading) | (int32_t(t.flags.back_face_culling) << 1),
                               triangle_id(scene, map, t) };
            os.write(reinterpret_cast<const char *>(rec), sizeof(rec));
            put(os, t.face_normal);
        }
    }
    for (const auto &l : scene.lights) {
        put(os, l.intensity);
        put(os, l.position);
    }
    for (const auto &m : scene.materials) {
        int32_t rec[2] = { int32_t(m.type), m.albedo_map_texture_index };
        os.write(reinterpret_cast<const char *>(rec), sizeof(rec));
        put(os, m.ior);
    }
    for (const auto &t : scene.textures) {
        int32_t type = int32_t(t.type), w = 0, h = 0;
        float p[7] = { 0, 0, 0, 0, 0, 0, 0 };
        switch (t.type) {
            case crt::TextureType::Albedo:
                std::memcpy(p, &t.as_albedo_tex.albedo, 12);
                break;
            case crt::TextureType::Edges:
                std::memcpy(p, &t.as_edges_tex.edge_color, 12);
                std::memcpy(p + 3, &t.as_edges_tex.inner_color, 12);
                p[6] = t.as_edges_tex.edge_width;
                break;
            case crt::TextureType::Checker:
                std::memcpy(p, &t.as_checker_tex.color_a, 12);
                std::memcpy(p + 3, &t.as_checker_tex.color_b, 12);
                p[6] = t.as_checker_tex.square_size;
                break;
            case crt::TextureType::Bitmap:
                w = t.as_bitmap_tex.image->width;
                h = t.as_bitmap_tex.image->height;
                break;
        }
        put(os, type);
        os.write(reinterpret_cast<const char *>(p), sizeof(p));
        put(os, w);
        put(os, h);
    }
    for (const auto &t : scene.textures) {
        if (t.type == crt::TextureType::Bitmap) {
            const auto &buf = t.as_bitmap_tex.image->buffer;
            os.write(reinterpret_cast<const char *>(buf.data()), std::streamsize(buf.size() * sizeof(crt::Color)));
        }
    }
}

} // namespace

int main(int argc, char **argv) {
    if (argc < 2) {
        std::fprintf(stderr, "usage: ref_harness <scene> [options]\n");
        return 2;
    }
    std::string scene_path = argv[1], asset_root, render_out, hits_out, rays_out, dump_out;
    int repeat = 1, stride = 1;
    bool brute = false;
    crt::RendererSettings settings;
    for (int i = 2; i < argc; ++i) {
        std::string a = argv[i];
        auto next = [&]() -> std::string { return i + 1 < argc ? argv[++i] : ""; };
        if (a == "--asset-root") asset_root = next();
        else if (a == "--render") render_out = next();
        else if (a == "--hits") hits_out = next();
        else if (a == "--rays") rays_out = next();
        else if (a == "--dump") dump_out = next();
        else if (a == "--repeat") repeat = std::stoi(next());
        else if (a == "--stride") stride = std::stoi(next());
        else if (a == "--max-depth") settings.max_ray_depth = uint32_t(std::stoul(next()));
        else if (a == "--gi-rays") settings.diffuse_reflection_ray_count = uint32_t(std::stoul(next()));
        else if (a == "--brute") brute = true;
        else {
            std::fprintf(stderr, "unknown option %s\n", a.c_str());
            return 2;
        }
    }

    std::ifstream in{ scene_path, std::ios::binary };
    if (!in.is_open()) {
        std::fprintf(stderr, "cannot open %s\n", scene_path.c_str());
        return 1;
    }
    std::stringstream ss;
    ss << in.rdbuf();
    const std::string text = ss.str();
    std::filesystem::path root = asset_root.empty() ? std::filesystem::path{ scene_path }.parent_path() : std::filesystem::path{ asset_root };

    using clock = std::chrono::steady_clock;
    auto t0 = clock::now();
    std::istringstream is{ text };
    std::optional<crt::Scene> scene = crt::json::read_scene_from_istream(is, root);
    auto t1 = clock::now();
    if (!scene) {
        std::printf("{\"loaded\": false}\n");
        return 3;
    }
    const int W = scene->camera.resolution_x(), H = scene->camera.resolution_y();

    int triangle_count = 0;
    TriIndexMap map;
    if (!hits_out.empty() || !dump_out.empty())
        map = build_parse_order_map(text, triangle_count);

    std::ostringstream js;
    js << "{\"loaded\": true, \"width\": " << W << ", \"height\": " << H << ", \"nodes\": " << scene->acceleration_tree.size()
       << ", \"vertices\": " << scene->vertices.size() << ", \"load_s\": " << std::chrono::duration<double>(t1 - t0).count()
       << ", \"threads\": " << std::thread::hardware_concurrency() << ", \"max_ray_depth\": " << settings.max_ray_depth
       << ", \"gi_rays\": " << settings.diffuse_reflection_ray_count;

    if (!dump_out.empty())
        dump_scene(*scene, map, dump_out);

    if (!render_out.empty() || repeat > 1) {
        std::vector<double> times;
        std::optional<crt::Image> image;
        for (int k = 0; k < repeat; ++k) {
            auto a = clock::now();
            image.emplace(crt::render_image(*scene, settings)); // the reference's own timed region (main.cpp:37-43)
            auto b = clock::now();
            times.push_back(std::chrono::duration<double>(b - a).count());
        }
        js << ", \"render_s\": [";
        for (size_t k = 0; k < times.size(); ++k)
            js << (k ? ", " : "") << times[k];
        js << "]";
        if (!render_out.empty()) {
            std::ofstream os{ render_out, std::ios::binary };
            os.write(reinterpret_cast<const char *>(image->buffer.data()), std::streamsize(image->buffer.size() * sizeof(crt::Color)));
        }
    }

    if (!rays_out.empty()) {
        std::ofstream os{ rays_out, std::ios::binary };
        for (int y = 0; y < H; y += stride)
            for (int x = 0; x < W; x += stride) {
                crt::Ray r = scene->camera.generate_ray(x, y);
                put(os, r.origin);
                put(os, r.direction);
            }
    }

    if (!hits_out.empty()) {
        std::vector<const crt::Triangle *> unique;
        if (brute) {
            unique.assign(size_t(triangle_count), nullptr);
            for (const auto &n : scene->acceleration_tree)
                for (const auto &t : n.triangles) {
                    int id = triangle_id(*scene, map, t);
                    if (id >= 0 && !unique[size_t(id)])
                        unique[size_t(id)] = &t;
                }
        }
        const int SW = (W + stride - 1) / stride, SH = (H + stride - 1) / stride;
        std::vector<int32_t> tri(size_t(SW) * SH), nn(size_t(SW) * SH), nt(size_t(SW) * SH);
        std::vector<float> dist(size_t(SW) * SH);
        const unsigned n_threads = std::max(1u, std::thread::hardware_concurrency());
        std::vector<std::thread> pool;
        for (unsigned w = 0; w < n_threads; ++w)
            pool.emplace_back([&, w]() {
                for (int sy = int(w); sy < SH; sy += int(n_threads))
                    for (int sx = 0; sx < SW; ++sx) {
                        crt::Ray r = scene->camera.generate_ray(sx * stride, sy * stride);
                        HitRecord h = brute ? walk_brute(r, unique) : walk_tree(r, *scene, map);
                        size_t i = size_t(sy) * SW + sx;
                        tri[i] = h.tri;
                        dist[i] = h.t;
                        nn[i] = h.n_node;
                        nt[i] = h.n_tri;
                    }
            });
        for (auto &t : pool)
            t.join();
        std::ofstream os{ hits_out, std::ios::binary };
        os.write(reinterpret_cast<const char *>(tri.data()), std::streamsize(tri.size() * 4));
        os.write(reinterpret_cast<const char *>(dist.data()), std::streamsize(dist.size() * 4));
        os.write(reinterpret_cast<const char *>(nn.data()), std::streamsize(nn.size() * 4));
        os.write(reinterpret_cast<const char *>(nt.data()), std::streamsize(nt.size() * 4));
        long long tot_n = 0, tot_t = 0, hit = 0;
        for (size_t i = 0; i < tri.size(); ++i) {
            tot_n += nn[i];
            tot_t += nt[i];
            hit += tri[i] >= 0;
        }
        js << ", \"triangles\": " << triangle_count << ", \"hits_w\": " << SW << ", \"hits_h\": " << SH << ", \"n_node_total\": " << tot_n
           << ", \"n_tri_total\": " << tot_t << ", \"hit_pixels\": " << hit;
    }

    js << "}";
    std::cout << js.str() << std::endl;
    return 0;
}
