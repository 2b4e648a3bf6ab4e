// Kernels of the render path.  See crt_device.cuh for the arithmetic and DESIGN.md for the roofline of each.
#pragma once

#include "crt_device.cuh"

namespace crtdev {

constexpr int kTileW = 16, kTileH = 8, kBlock = kTileW * kTileH; // 4 warps, each an 8x4 pixel patch

// thread -> pixel inside a 16x8 tile: warp w covers the 8x4 patch at ((w&1)*8, (w>>1)*4)
__device__ __forceinline__ void tile_pixel(int &x, int &ly) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    x = blockIdx.x * kTileW + (warp & 1) * 8 + (lane & 7);
    ly = blockIdx.y * kTileH + (warp >> 1) * 4 + (lane >> 3);
}

template <int TRAV, bool STATS>
__device__ __forceinline__ Hit trace(const DScene &S, const Ray &r, Counters *cn) {
    if (TRAV == CRT_TRAVERSAL_LITERAL)
        return trace_literal<STATS>(S, r, cn);
    return trace_kd<STATS>(S, r, cn);
}

__device__ __forceinline__ void flush_counters(const Counters &cn, Counters *global) {
    const unsigned long long *src = reinterpret_cast<const unsigned long long *>(&cn);
    unsigned long long *dst = reinterpret_cast<unsigned long long *>(global);
    for (int k = 0; k < int(sizeof(Counters) / 8); ++k) {
        unsigned long long v = src[k];
        for (int o = 16; o > 0; o >>= 1)
            v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0 && v)
            atomicAdd(&dst[k], v);
    }
}

// ------------------------------------------------------------------------------------------------ parity hooks
__global__ void __launch_bounds__(kBlock) k_primary_rays(const __grid_constant__ DScene S, float *__restrict__ out) {
    int x, y;
    tile_pixel(x, y);
    if (x >= S.cam.width || y >= S.cam.height)
        return;
    const Ray r = generate_ray(S.cam, x, y);
    float *o = out + size_t(y * S.cam.width + x) * 6;
    o[0] = r.o.x, o[1] = r.o.y, o[2] = r.o.z, o[3] = r.d.x, o[4] = r.d.y, o[5] = r.d.z;
}

template <int TRAV>
__global__ void __launch_bounds__(kBlock) k_primary_hits(const __grid_constant__ DScene S, int *__restrict__ tri_out, float *__restrict__ t_out,
                                                         int *__restrict__ node_tests, int *__restrict__ tri_tests, Counters *__restrict__ counters) {
    __shared__ float4 stage[TRAV == CRT_TRAVERSAL_BRUTE ? 3 * kBruteChunk : 1];
    int x, y;
    tile_pixel(x, y);
    const bool active = x < S.cam.width && y < S.cam.height;
    Counters cn{};
    Hit h{ 0.0f, -1 };
    Ray r{};
    if (active)
        r = generate_ray(S.cam, x, y);
    if (TRAV == CRT_TRAVERSAL_BRUTE)
        h = trace_brute<true>(S, r, active, &cn, stage);
    else if (active)
        h = trace<TRAV, true>(S, r, &cn);
    if (active) {
        cn.rays = 1;
        cn.rays_by_depth[0] = 1;
        const size_t i = size_t(y) * S.cam.width + x;
        if (tri_out)
            tri_out[i] = h.tri;
        if (t_out)
            t_out[i] = h.tri >= 0 ? h.t : 0.0f;
        if (node_tests)
            node_tests[i] = int(cn.node_tests);
        if (tri_tests)
            tri_tests[i] = int(cn.tri_tests);
    }
    flush_counters(cn, counters);
}

// ------------------------------------------------------------------------------------------------ megakernel
// One thread evaluates shade_ray (src/core/crt_renderer.cpp:46-145) for one pixel, depth-first, with an explicit
// frame stack instead of recursion.  Evaluation order is the reference's: GI children in index order before the
// light loop, reflection subtree before refraction subtree, one PCG32 stream per pixel consumed in that order --
// so float results are combined in exactly the nesting the CPU uses.
enum FrameKind : int { kDiffuseGi = 0, kReflect = 1, kRefractWaitReflection = 2, kRefractWaitRefraction = 3, kRefractTir = 4,
                       kDirectShadow = 5 /* k_render_persistent: light loop with one shadow query per light */ };

struct Frame {
    int kind, i;
    float f;
    V3 a, b, c, e, g;
};

constexpr int kMaxDepthCap = 15; // frames[depth] for depth 0..15

template <int TRAV, bool STATS>
__global__ void __launch_bounds__(kBlock) k_render_mega(const __grid_constant__ DScene S, const __grid_constant__ DParams P, float *__restrict__ rgb,
                                                        uint8_t *__restrict__ rgb8, Counters *__restrict__ counters) {
    int x, ly;
    tile_pixel(x, ly);
    const bool active = x < S.cam.width && ly < P.rows;
    Counters cn{};
    if (active) {
        const int y = global_row(P, ly);
        const V3 bg = mk(S.bg[0], S.bg[1], S.bg[2]);
        const float gi_div = float(P.gi_rays + 1u);
        const bool gi = S.gi_on && P.gi_rays > 0;
        Pcg32 rng = make_pcg(uint32_t(x), uint32_t(y));
        Frame frames[kMaxDepthCap + 1];
        Ray ray = generate_ray(S.cam, x, y);
        int depth = 0;
        V3 c = mk(0.0f, 0.0f, 0.0f);

        auto no_shadow = [](V3, float) { return false; };
        auto shadow_query = [&](const Surface &s) {
            return [&, s](V3 dir, float r2) {
                if (STATS)
                    ++cn.shadow_rays;
                const Hit sh = trace<TRAV, STATS>(S, Ray{ s.P + s.n * P.shadow_bias, dir }, &cn);
                return sh.tri >= 0 && !(sh.t * sh.t > r2);
            };
        };
        auto direct = [&](V3 acc, const Surface &s, V3 albedo) {
            return P.shadows ? add_direct_light(S, acc, s, albedo, shadow_query(s)) : add_direct_light(S, acc, s, albedo, no_shadow);
        };

        for (;;) {
            // ---- shade_ray(ray, depth)
            bool descend = false;
            if (uint32_t(depth) > P.max_ray_depth) {
                c = mk(0.0f, 0.0f, 0.0f);
            } else {
                if (STATS) {
                    ++cn.rays;
                    ++cn.rays_by_depth[depth];
                }
                const Hit h = trace<TRAV, STATS>(S, ray, &cn);
                if (h.tri < 0) {
                    c = bg;
                } else {
                    const Surface s = make_surface(S, ray, h);
                    const int4 m = __ldg(&S.materials[s.material]);
                    Frame &f = frames[depth];
                    if (m.x == 0) { // Diffuse
                        const V3 albedo = sample_texture(S, m.y, s);
                        if (gi) {
                            f.kind = kDiffuseGi, f.i = 0;
                            f.a = mk(0.0f, 0.0f, 0.0f), f.b = s.P, f.c = s.n, f.e = ray.d, f.g = albedo;
                            ray = Ray{ s.P + s.n * P.diffuse_reflection_bias, gi_direction(ray.d, s.n, rng) };
                            descend = true;
                        } else {
                            c = direct(mk(0.0f, 0.0f, 0.0f), s, albedo) / gi_div;
                        }
                    } else if (m.x == 1) { // Reflective
                        const V3 albedo = sample_texture(S, m.y, s);
                        if (!S.reflections_on) {
                            c = albedo;
                        } else {
                            f.kind = kReflect, f.a = albedo;
                            ray = Ray{ s.P + s.n * P.reflection_bias, reflect(ray.d, s.n) };
                            descend = true;
                        }
                    } else if (m.x == 2) { // Refractive
                        if (!S.refractions_on) {
                            c = mk(0.0f, 0.0f, 0.0f);
                        } else {
                            V3 n = s.n;
                            float outside = 1.0f, inside = __int_as_float(m.z);
                            if (dot(ray.d, n) > 0.0f) { // leaving the volume
                                n = -n;
                                const float tmp = outside;
                                outside = inside, inside = tmp;
                            }
                            V3 rd = ray.d;
                            const bool refracted = refract(rd, n, outside, inside);
                            f.kind = refracted ? kRefractWaitReflection : kRefractTir;
                            f.f = fresnel_term(1.0f + dot(ray.d, n));
                            f.b = s.P + (-n) * 1e-2f; // the refraction_bias argument is dropped by the reference (crt_ray.h:33)
                            f.c = rd;
                            ray = Ray{ s.P + n * P.reflection_bias, reflect(ray.d, n) };
                            descend = true;
                        }
                    } else { // Constant
                        c = sample_texture(S, m.y, s);
                    }
                }
            }
            if (descend) {
                ++depth;
                continue;
            }
            // ---- return c to the callers
            bool done = false;
            for (;;) {
                if (depth == 0) {
                    done = true;
                    break;
                }
                Frame &f = frames[depth - 1];
                if (f.kind == kReflect) {
                    c = tint(f.a, c);
                } else if (f.kind == kRefractTir) {
                    // reflection colour passes through
                } else if (f.kind == kRefractWaitReflection) {
                    f.a = c;
                    f.kind = kRefractWaitRefraction;
                    ray = Ray{ f.b, f.c };
                    break; // evaluate the refraction ray at the same depth
                } else if (f.kind == kRefractWaitRefraction) {
                    c = f.a * f.f + c * (1.0f - f.f);
                } else { // kDiffuseGi
                    f.a = f.a + c;
                    if (uint32_t(++f.i) < P.gi_rays) {
                        ray = Ray{ f.b + f.c * P.diffuse_reflection_bias, gi_direction(f.e, f.c, rng) };
                        break;
                    }
                    Surface s{};
                    s.P = f.b, s.n = f.c;
                    c = direct(f.a, s, f.g) / gi_div;
                }
                --depth;
            }
            if (done)
                break;
        }

        const size_t i = (size_t(ly) * S.cam.width + x) * 3;
        if (rgb)
            rgb[i] = c.x, rgb[i + 1] = c.y, rgb[i + 2] = c.z;
        if (rgb8)
            rgb8[i] = quantize(c.x), rgb8[i + 1] = quantize(c.y), rgb8[i + 2] = quantize(c.z);
    }
    if (STATS)
        flush_counters(cn, counters);
}

} // namespace crtdev
