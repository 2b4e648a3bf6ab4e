// TEST INFRASTRUCTURE ONLY -- the CPU restatement ("port") of the reference render path.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm may load this library; it is
// the checker, never the thing measured or shipped.  It restates, as plain scalar recursive C++ operating on the
// flattened scene of include/crt_cuda.h, the algorithm of bvpav/chaos-ray-tracing-course-2025 at HEAD, each
// function citing the reference lines it follows.  It is PINNED: tests/test_oracle_pinned.py checks that its float
// framebuffers, hit triangles and box/triangle test counters are bit-identical to those of the unmodified reference
// compiled into oracle/_ref (same compiler, no FMA, same libm), on every loadable bundled scene, and against the
// golden vectors under tests/golden/ that were generated from that reference build.
//
// Build: g++ -O2 -std=c++17 -ffp-contract=off (oracle/Makefile, target `port`).

#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <optional>
#include <thread>
#include <utility>
#include <vector>

#include "crt_cuda.h"

namespace {

struct Vec {
    float x, y, z;
};
inline Vec operator+(Vec a, Vec b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
inline Vec operator-(Vec a, Vec b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
inline Vec operator-(Vec a) { return { -a.x, -a.y, -a.z }; }
inline Vec operator*(Vec a, float s) { return { a.x * s, a.y * s, a.z * s }; }
inline Vec operator/(Vec a, float s) { return { a.x / s, a.y / s, a.z / s }; }
inline float dot(Vec a, Vec b) { return a.x * b.x + a.y * b.y + a.z * b.z; }                                      // crt_vector.h:115-117
inline Vec cross(Vec a, Vec b) { return { a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; } // :107-113
inline float length(Vec a) { return std::sqrt(dot(a, a)); }                                                        // crt_vector.cpp:7-9
inline Vec normalized(Vec a) { return a / length(a); }
inline float comp(Vec v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }
inline Vec xyz(const crt_float4 &v) { return { v.x, v.y, v.z }; }

// Vector * Matrix with the matrix given by rows (crt_matrix.h:66-74): result[i] = sum_j v[j] * m[j][i], from 0
inline Vec mul_rows(Vec v, Vec r0, Vec r1, Vec r2) {
    Vec out{ 0.0f, 0.0f, 0.0f };
    const float vv[3] = { v.x, v.y, v.z };
    const Vec rows[3] = { r0, r1, r2 };
    float *o[3] = { &out.x, &out.y, &out.z };
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            *o[i] += vv[j] * comp(rows[j], i);
    return out;
}

struct Ray {
    Vec origin, direction;
    int depth;
};

struct Hit { // crt_intersection.h:13-20 plus the triangle identity
    float distance;
    Vec point, normal;
    float u, v; // uv.x, uv.y
    float bary_u, bary_v;
    int material;
    int triangle;
};

struct Stats {
    uint64_t rays = 0, shadow_rays = 0, node_tests = 0, tri_tests = 0;
    uint64_t by_depth[16] = {};
};

struct Ctx {
    const crt_flat_scene &s;
    const crt_render_params &p;
};

// Camera::generate_ray, crt_camera.cpp:7-35
Ray generate_ray(const crt_camera &c, int raster_x, int raster_y) {
    float dx = raster_x + 0.5f, dy = raster_y + 0.5f;
    dx /= c.width;
    dy /= c.height;
    dx = (2.0f * dx) - 1.0f;
    dy = 1.0f - (2.0f * dy);
    dx *= c.aspect;       // float(m_resolution_x) / m_resolution_y
    dx *= c.tan_half_fov; // std::tan(m_fov_radians * 0.5f)
    dy *= c.tan_half_fov;
    Vec d = mul_rows({ dx, dy, -1.0f }, { c.rotation[0], c.rotation[1], c.rotation[2] }, { c.rotation[3], c.rotation[4], c.rotation[5] },
                     { c.rotation[6], c.rotation[7], c.rotation[8] });
    return { { c.position[0], c.position[1], c.position[2] }, normalized(d), 0 };
}

// ray_intersect_aabb_p, crt_intersection.cpp:14-45
bool box_test(const Ray &ray, const crt_box_node &b) {
    const Vec extents[2] = { { b.min[0], b.min[1], b.min[2] }, { b.max[0], b.max[1], b.max[2] } };
    static const int u_table[3] = { 1, 2, 0 }, v_table[3] = { 2, 0, 1 };
    for (int e = 0; e < 2; ++e)
        for (int axis = 0; axis < 3; ++axis) {
            const float da = comp(ray.direction, axis);
            if (std::fabs(da) < 1e-6f)
                continue;
            const float t = (comp(extents[e], axis) - comp(ray.origin, axis)) / da;
            if (t < 0.0f)
                continue;
            const Vec p = ray.origin + ray.direction * t;
            const int u = u_table[axis], v = v_table[axis];
            if (comp(p, u) >= b.min[u] && comp(p, u) <= b.max[u] && comp(p, v) >= b.min[v] && comp(p, v) <= b.max[v])
                return true;
        }
    return false;
}

// ray_intersect_triangle, crt_intersection.cpp:47-93
std::optional<Hit> triangle_test(const crt_flat_scene &s, const Ray &ray, int tri) {
    const crt_float4 &A = s.tri_verts[3 * tri], &B = s.tri_verts[3 * tri + 1], &C = s.tri_verts[3 * tri + 2];
    const int32_t *meta = &s.tri_meta[4 * tri];
    const Vec v0 = xyz(A), v1 = xyz(B), v2 = xyz(C), N{ A.w, B.w, C.w };
    const Vec e0 = v1 - v0, e1 = v2 - v1, e2 = v0 - v2; // crt_triangle.h:39-45
    const float ray_normal_dist = dot(N, ray.direction);
    if (std::fabs(ray_normal_dist) < 1e-6f)
        return std::nullopt;
    const float origin_plane_dist = dot(N, v0 - ray.origin);
    const bool is_front_face = origin_plane_dist < 0.0f;
    if (!(is_front_face || !(meta[3] & CRT_META_CULL_BIT)))
        return std::nullopt;
    const float t = origin_plane_dist / ray_normal_dist;
    if (t < 0.0f)
        return std::nullopt;
    const Vec P = ray.origin + ray.direction * t;
    const Vec v0p = P - v0, v1p = P - v1, v2p = P - v2;
    if (!(dot(N, cross(e0, v0p)) >= 0.0f && dot(N, cross(e1, v1p)) >= 0.0f && dot(N, cross(e2, v2p)) >= 0.0f))
        return std::nullopt;
    const Vec v0v1 = e0, v0v2 = -e2;
    const float bary_u = length(cross(v0p, v0v2)) / length(cross(v0v1, v0v2));
    const float bary_v = length(cross(v0v1, v0p)) / length(cross(v0v1, v0v2));
    Hit h;
    h.distance = t;
    h.point = P;
    if (meta[3] & CRT_META_SMOOTH_BIT) {
        const Vec n0 = xyz(s.vert_normal[meta[0]]), n1 = xyz(s.vert_normal[meta[1]]), n2 = xyz(s.vert_normal[meta[2]]);
        h.normal = n1 * bary_u + n2 * bary_v + n0 * (1 - bary_u - bary_v);
    } else {
        h.normal = N;
    }
    const float *t0 = &s.vert_uv[2 * meta[0]], *t1 = &s.vert_uv[2 * meta[1]], *t2 = &s.vert_uv[2 * meta[2]];
    h.u = t1[0] * bary_u + t2[0] * bary_v + t0[0] * (1.0f - bary_u - bary_v);
    h.v = t1[1] * bary_u + t2[1] * bary_v + t0[1] * (1.0f - bary_u - bary_v);
    h.bary_u = bary_u;
    h.bary_v = bary_v;
    h.material = meta[3] & CRT_META_MATERIAL_MASK;
    h.triangle = tri;
    return h;
}

// ray_intersect_triangle_span over leaf references, crt_intersection.cpp:95-107
std::optional<Hit> span_test(const crt_flat_scene &s, const Ray &ray, const uint32_t *refs, int count, Stats &st) {
    std::optional<Hit> closest;
    for (int i = 0; i < count; ++i) {
        ++st.tri_tests;
        if (auto h = triangle_test(s, ray, int(refs[i] & 0x7FFFFFFFu)))
            if (!closest || h->distance < closest->distance)
                closest = h;
    }
    return closest;
}

// ray_intersect_acceleration_tree, crt_intersection.cpp:109-136
std::optional<Hit> trace_ray(const crt_flat_scene &s, const Ray &ray, Stats &st) {
    std::optional<Hit> closest;
    std::vector<int> stack{ 0 };
    while (!stack.empty()) {
        const int node_index = stack.back();
        stack.pop_back();
        const crt_box_node &node = s.box_nodes[node_index];
        ++st.node_tests;
        if (!box_test(ray, node))
            continue;
        if (node.left <= -2) { // leaf
            auto h = span_test(s, ray, s.refs + (-node.left - 2), node.right, st);
            if (h && (!closest || h->distance < closest->distance))
                closest = h;
        } else {
            if (node.left != -1)
                stack.push_back(node.left);
            if (node.right != -1)
                stack.push_back(node.right);
        }
    }
    return closest;
}

// the pre-tree path: one span over every triangle in parse order (config 2's checker)
std::optional<Hit> trace_brute(const crt_flat_scene &s, const Ray &ray, Stats &st) {
    std::optional<Hit> closest;
    for (int i = 0; i < s.n_triangles; ++i) {
        ++st.tri_tests;
        if (auto h = triangle_test(s, ray, i))
            if (!closest || h->distance < closest->distance)
                closest = h;
    }
    return closest;
}

// Texture::sample, crt_texture.cpp:9-49
Vec sample_texture(const crt_flat_scene &s, int index, const Hit &h) {
    const crt_texture &t = s.textures[index];
    const Vec c0{ t.color0[0], t.color0[1], t.color0[2] }, c1{ t.color1[0], t.color1[1], t.color1[2] };
    switch (t.type) {
        case 1: // edges
            if (h.bary_u <= t.param || h.bary_v <= t.param || (1.0f - h.bary_u - h.bary_v) <= t.param)
                return c0;
            return c1;
        case 2: { // checker
            const int row = static_cast<int>(h.u / t.param);
            const int col = static_cast<int>(h.v / t.param);
            return ((row + col) & 1) ? c1 : c0;
        }
        case 3: { // bitmap
            int raster_x = static_cast<int>(h.u * t.width) % t.width;
            int raster_y = static_cast<int>((1.0f - h.v) * t.height) % t.height;
            // negative uv is an out-of-bounds read in the reference (undefined); both sides of the parity test clamp
            raster_x = raster_x < 0 ? 0 : raster_x;
            raster_y = raster_y < 0 ? 0 : raster_y;
            return xyz(s.texels[t.texel_offset + raster_y * t.width + raster_x]);
        }
        default:
            return c0;
    }
}

// PCG32 + make_pcg, crt_random.h:10-43
struct Pcg32 {
    uint64_t state, inc;
    uint32_t next() {
        const uint64_t old = state;
        state = old * UINT64_C(6364136223846793005) + inc;
        const uint32_t xorshifted = uint32_t(((old >> 18u) ^ old) >> 27u);
        const uint32_t rot = uint32_t(old >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((-rot) & 31));
    }
    float uniform() {
        union {
            uint32_t i;
            float f;
        } u;
        u.i = 0x3f800000u | (next() >> 9);
        return u.f - 1.0f;
    }
};
Pcg32 make_pcg(uint32_t raster_x, uint32_t raster_y) {
    const uint64_t seed = (uint64_t(raster_x) << 32) | raster_y;
    Pcg32 rng;
    rng.state = 0;
    rng.inc = (seed << 1) | 1;
    (void)rng.next();
    rng.state += seed;
    (void)rng.next();
    return rng;
}

constexpr float kPi = 3.14159274101257324219f; // std::numbers::pi_v<float>

// Vector::refract, crt_vector.cpp:11-27
bool refract(Vec &d, Vec normal, float outside_ior, float inside_ior) {
    const float cos_alpha = -dot(d, normal);
    const float sin_alpha = std::sqrt(1.0f - cos_alpha * cos_alpha);
    if (sin_alpha > inside_ior / outside_ior)
        return false;
    const float sin_beta = sin_alpha * outside_ior / inside_ior;
    const float cos_beta = std::sqrt(1.0f - sin_beta * sin_beta);
    d = d + normal * cos_alpha;
    d = normalized(d);
    d = d * sin_beta;
    d = d + (-normal) * cos_beta;
    return true;
}

// Vector::reflected, crt_vector.h:119-127
Vec reflected(Vec d, Vec normal) { return d - normal * 2.0f * dot(d, normal); }

// shade_ray, crt_renderer.cpp:46-145.  p.shadows == 0 is HEAD: trace_ray_with_refractions (:29-44) never enters
// its loop, so every light is unoccluded.  p.shadows == 1 is the PATCHED oracle for the flagged corrected-shadows
// mode: an opaque-occluder query, lit <=> no hit with distance^2 <= r^2.
Vec shade_ray(const Ctx &c, const Ray &ray, Pcg32 &rng, Stats &st) {
    const crt_flat_scene &s = c.s;
    const crt_render_params &p = c.p;
    if (uint32_t(ray.depth) > p.max_ray_depth)
        return { 0.0f, 0.0f, 0.0f };
    ++st.rays;
    if (ray.depth < 16)
        ++st.by_depth[ray.depth];
    const std::optional<Hit> hit = trace_ray(s, ray, st);
    if (!hit)
        return { s.background[0], s.background[1], s.background[2] };

    const crt_material &material = s.materials[hit->material];
    Vec normal = hit->normal;
    switch (material.type) {
        case 0: { // Diffuse
            Vec final_color{ 0.0f, 0.0f, 0.0f };
            if (s.gi_on) {
                for (uint32_t i = 0; i < p.diffuse_reflection_ray_count; ++i) {
                    const Vec right = normalized(cross(ray.direction, normal));
                    const Vec up = normal;
                    const Vec forward = cross(right, up);
                    const float angle_xy = kPi * rng.uniform();
                    Vec direction{ std::cos(angle_xy), std::sin(angle_xy), 0.0f };
                    const float angle_xz = 2.0f * kPi * rng.uniform();
                    // Matrix::rotation_y, crt_matrix.cpp:15-21
                    direction = mul_rows(direction, { std::cos(angle_xz), 0.0f, -std::sin(angle_xz) }, { 0.0f, 1.0f, 0.0f },
                                         { std::sin(angle_xz), 0.0f, std::cos(angle_xz) });
                    direction = mul_rows(direction, right, up, forward); // Matrix::from_axes, crt_matrix.h:25-31
                    const Ray child{ hit->point + normal * p.diffuse_reflection_bias, direction, ray.depth + 1 };
                    final_color = final_color + shade_ray(c, child, rng, st);
                }
            }
            const Vec albedo = sample_texture(s, material.texture_index, *hit);
            for (int l = 0; l < s.n_lights; ++l) {
                Vec light_dir = xyz(s.lights[l]) - hit->point;
                const float r2 = dot(light_dir, light_dir);
                light_dir = light_dir / std::sqrt(r2);
                const float cos_law = std::max(0.0f, dot(light_dir, normal));
                const float sphere_area = 4 * kPi * r2;
                bool lit = true;
                if (p.shadows) {
                    ++st.shadow_rays;
                    const Ray shadow{ hit->point + normal * p.shadow_bias, light_dir, 0 };
                    const std::optional<Hit> blocker = trace_ray(s, shadow, st);
                    lit = !blocker || blocker->distance * blocker->distance > r2;
                }
                if (lit)
                    final_color = final_color + albedo * s.lights[l].w / sphere_area * cos_law;
            }
            return final_color / float(p.diffuse_reflection_ray_count + 1);
        }
        case 1: { // Reflective
            const Ray reflection{ hit->point + normal * p.reflection_bias, reflected(ray.direction, normal), ray.depth + 1 };
            const Vec a = sample_texture(s, material.texture_index, *hit);
            if (!s.reflections_on)
                return a;
            const Vec r = shade_ray(c, reflection, rng, st);
            return { a.x * r.x, a.y * r.y * a.y, a.z * r.z }; // Vector::operator*(Vector), crt_vector.h:76-78
        }
        case 2: { // Refractive
            if (!s.refractions_on)
                return { 0.0f, 0.0f, 0.0f };
            float outside_ior = 1.0f, inside_ior = material.ior;
            if (dot(ray.direction, normal) > 0.0f) {
                normal = -normal;
                std::swap(inside_ior, outside_ior);
            }
            // Ray::refracted_at / refract_at, crt_ray.h:31-50: the bias parameter is not forwarded (always 1e-2f)
            Ray refraction = ray;
            const bool refracted = refract(refraction.direction, normal, outside_ior, inside_ior);
            if (refracted) {
                refraction.origin = hit->point + (-normal) * 1e-2f;
                refraction.depth = ray.depth + 1;
            }
            const Ray reflection{ hit->point + normal * p.reflection_bias, reflected(ray.direction, normal), ray.depth + 1 };
            const Vec reflection_color = shade_ray(c, reflection, rng, st);
            if (!refracted)
                return reflection_color;
            const Vec refraction_color = shade_ray(c, refraction, rng, st);
            const float fresnel = 0.5f * std::pow((1.0f + dot(ray.direction, normal)), 5.0f);
            return reflection_color * fresnel + refraction_color * (1.0f - fresnel);
        }
        default: // Constant
            return sample_texture(s, material.texture_index, *hit);
    }
}

// write_ppm's channel rule, crt_image_ppm.cpp:16-19
uint8_t quantize(float c) {
    const float v = c * 255;
    int q;
    if (!(v < 2147483648.0f) || v < -2147483648.0f)
        q = INT32_MIN; // what the x86 conversion the reference compiles to produces for out-of-range / NaN
    else
        q = static_cast<int>(v);
    return uint8_t(q < 0 ? 0 : (q > 255 ? 255 : q));
}

std::vector<int> owned_rows(const crt_render_params &p, int height) {
    std::vector<int> rows;
    if (p.world <= 1) {
        for (int y = 0; y < height; ++y)
            rows.push_back(y);
        return rows;
    }
    const int band = p.band_rows > 0 ? p.band_rows : 8;
    const int bands = (height + band - 1) / band;
    for (int b = p.rank; b < bands; b += p.world)
        for (int y = b * band; y < std::min(height, (b + 1) * band); ++y)
            rows.push_back(y);
    return rows;
}

template <class Fn>
void parallel_rows(int n_rows, int threads, Fn fn) {
    unsigned n = threads > 0 ? unsigned(threads) : std::max(1u, std::thread::hardware_concurrency());
    std::atomic<int> next{ 0 };
    std::vector<std::thread> pool;
    for (unsigned w = 0; w < n; ++w)
        pool.emplace_back([&, w]() {
            for (;;) {
                const int row = next.fetch_add(1);
                if (row >= n_rows)
                    return;
                fn(row, int(w));
            }
        });
    for (auto &t : pool)
        t.join();
}

} // namespace

extern "C" {

// render_image / render_region, crt_renderer.cpp:147-199 (pixels are independent; rows are handed to threads)
int crt_oracle_render(const crt_flat_scene *s, const crt_render_params *p, int threads, float *rgb, uint8_t *rgb8, crt_render_stats *stats) {
    if (!s || !p)
        return -1;
    const int W = s->camera.width;
    const std::vector<int> rows = owned_rows(*p, s->camera.height);
    const unsigned n = threads > 0 ? unsigned(threads) : std::max(1u, std::thread::hardware_concurrency());
    std::vector<Stats> per_thread(n);
    const Ctx ctx{ *s, *p };
    // bucket counts per axis, crt_renderer.cpp:160-161: int(float(extent) / bucket_size + 0.5); the last bucket of a row
    // or column absorbs the remainder (:165-170), so the only effect on the pixels is that an image narrower or lower
    // than half a bucket gets no bucket at all and stays the zero-initialised Image
    const auto bucket_count = [&](int extent) { return s->bucket_size > 0 ? static_cast<int>(float(extent) / s->bucket_size + 0.5) : 0; };
    if (bucket_count(W) <= 0 || bucket_count(s->camera.height) <= 0) {
        if (rgb)
            std::memset(rgb, 0, rows.size() * size_t(W) * 3 * sizeof(float));
        if (rgb8)
            std::memset(rgb8, 0, rows.size() * size_t(W) * 3);
        if (stats) {
            std::memset(stats, 0, sizeof *stats);
            stats->rows_rendered = int(rows.size());
        }
        return 0;
    }
    parallel_rows(int(rows.size()), int(n), [&](int local_row, int worker) {
        const int y = rows[size_t(local_row)];
        for (int x = 0; x < W; ++x) {
            Pcg32 rng = make_pcg(uint32_t(x), uint32_t(y));
            const Ray camera_ray = generate_ray(s->camera, x, y);
            const Vec c = shade_ray(ctx, camera_ray, rng, per_thread[size_t(worker)]);
            const size_t i = (size_t(local_row) * W + x) * 3;
            if (rgb)
                rgb[i] = c.x, rgb[i + 1] = c.y, rgb[i + 2] = c.z;
            if (rgb8)
                rgb8[i] = quantize(c.x), rgb8[i + 1] = quantize(c.y), rgb8[i + 2] = quantize(c.z);
        }
    });
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        stats->rows_rendered = int(rows.size());
        for (const Stats &t : per_thread) {
            stats->rays += t.rays;
            stats->shadow_rays += t.shadow_rays;
            stats->node_tests += t.node_tests;
            stats->tri_tests += t.tri_tests;
            for (int d = 0; d < 16; ++d)
                stats->rays_by_depth[d] += t.by_depth[d];
        }
    }
    return 0;
}

int crt_oracle_primary_hits(const crt_flat_scene *s, int brute, int threads, int32_t *tri, float *t, int32_t *node_tests, int32_t *tri_tests) {
    if (!s)
        return -1;
    const int W = s->camera.width, H = s->camera.height;
    parallel_rows(H, threads, [&](int y, int) {
        for (int x = 0; x < W; ++x) {
            Stats st;
            const Ray r = generate_ray(s->camera, x, y);
            const std::optional<Hit> h = brute ? trace_brute(*s, r, st) : trace_ray(*s, r, st);
            const size_t i = size_t(y) * W + x;
            if (tri)
                tri[i] = h ? h->triangle : -1;
            if (t)
                t[i] = h ? h->distance : 0.0f;
            if (node_tests)
                node_tests[i] = int32_t(st.node_tests);
            if (tri_tests)
                tri_tests[i] = int32_t(st.tri_tests);
        }
    });
    return 0;
}

int crt_oracle_primary_rays(const crt_flat_scene *s, float *rays) {
    if (!s || !rays)
        return -1;
    for (int y = 0; y < s->camera.height; ++y)
        for (int x = 0; x < s->camera.width; ++x) {
            const Ray r = generate_ray(s->camera, x, y);
            float *o = rays + (size_t(y) * s->camera.width + x) * 6;
            o[0] = r.origin.x, o[1] = r.origin.y, o[2] = r.origin.z;
            o[3] = r.direction.x, o[4] = r.direction.y, o[5] = r.direction.z;
        }
    return 0;
}

} // extern "C"
