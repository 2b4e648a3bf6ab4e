// Wavefront pipeline: per bounce level one TRACE kernel (persistent warps, finished rays are replaced lane by lane
// from the level's queue) and one SHADE kernel (one thread per ray, emits the next level's rays with warp-aggregated
// ballot/shfl compaction), then a bottom-up RESOLVE that combines child colours in exactly the nesting of the
// reference's recursion (src/core/crt_renderer.cpp:103-135), so float results stay bit-identical to the depth-first
// evaluation.  Handles every material except GI (one PCG stream per pixel consumed depth-first -> megakernel) and the
// flagged shadow mode.
//
// Node ids: level-0 nodes are pixel SLOTS s in [0, P) (P = padded pixel count, 8x4-patch-major order so consecutive
// slots are image neighbours); level >= 1 nodes are P + q, q in [0, cap), allocated level by level, so a level's
// nodes are contiguous: [P + start(L), P + start(L) + count[L]).
#pragma once

#include "crt_kernels.cuh"

namespace crtdev {

constexpr bool kWavefrontReady = true;

// Frame buffers and ray / hit queues are written once and read once (or never, on the device): streamed through L2
// with the evict-first policy they do not push the walk tree and the triangles out of it.  (ncu r2b: one C4 frame reads
// 0.9 GB from DRAM for a 104 MB scene, and nearly every warp-wide node visit waits for at least one DRAM access.)
#ifndef CRT_STREAM_IO
#define CRT_STREAM_IO 0
#endif
template <class T>
__device__ __forceinline__ void put_stream(T *p, const T &v) {
#if CRT_STREAM_IO
    __stcs(p, v);
#else
    *p = v;
#endif
}
template <class T>
__device__ __forceinline__ T get_stream(const T *p) {
#if CRT_STREAM_IO
    return __ldcs(p);
#else
    return *p;
#endif
}
constexpr int kMaxLevels = kMaxDepthCap + 2;

struct WfCtl {                 // zeroed at the start of every frame
    int count[kMaxLevels];     // rays per level (level 0: primary rays that enter the root box)
    int fetch[kMaxLevels];     // trace kernels' queue cursors
    int pending0;              // non-terminal pixels (entries of `pend`)
    int overflow;              // a level outgrew the pool: the host retries with a larger one
    int pend_hi;               // 1 + highest local row holding a non-terminal pixel (0: none)
    int pend_lo_inv;           // INT_MAX - lowest such row: both grow under atomicMax from the zeroed state
};

struct WfBuffers {
    // ray records of levels >= 1, [cap]: {origin, root entry t} {direction, root exit t} {1/direction, -}
    float4 *ray_o, *ray_d, *ray_i;
    // primary rays that enter the root box, compacted, [P]: {direction, entry t} {1/direction, exit t} slot
    // (their origin is the camera position)
    float4 *ray0_d, *ray0_i;
    int *ray0_slot;
    float2 *hit;           // [P + cap]  {t, tri as int bits}
    int4 *rec;             // [P + cap]  {kind, first child node id, fresnel bits, 0}
    float4 *alb;           // [P + cap]  reflective albedo
    float4 *col;           // [P + cap]  resolved colour of a node
    int *pend;             // [P]        slots of non-terminal pixels
    WfCtl *ctl;
    int P, cap, patches_x;
};

struct WavefrontState {
    WfBuffers b{};
    size_t node_capacity = 0; // allocated entries of hit
    size_t tree_capacity = 0; // allocated entries of rec/alb/col
    size_t ray_capacity = 0;  // allocated entries of ray_o/ray_d/ray_i
    size_t ray0_capacity = 0; // allocated entries of ray0_*
    size_t pend_capacity = 0;
    int sm_count = 0;
    int *h_flags = nullptr;   // pinned: {overflow, pend_hi, pend_lo_inv} of the last frame ...
    int *h_count = nullptr;   // ... and its rays per level (WfCtl::count): what a pool that overflowed is re-sized from
    double grow = 1.0;        // pool = grow * pixels
};

enum NodeKind : int { kTerminal = 0, kNodeReflect = 1, kNodeRefract = 2, kNodeTir = 3 };

// rays of a level (level >= 1) that really have a record: a level that outgrew the pool (ctl->overflow; the host
// discards the frame and renders it again with a larger pool) counted more rays than fit -- never read past the pool
__device__ __forceinline__ int level_count(const WfCtl *ctl, int cap, int level, int start) {
    return max(0, min(ctl->count[level], cap - start));
}

__device__ __forceinline__ int level_start(const WfCtl *ctl, int level) { // q-offset of a level (level >= 1)
    int s = 0;
    for (int l = 1; l < level; ++l)
        s += ctl->count[l];
    return s;
}

// slot -> pixel.  Slots run patch by patch (8 wide, 4 tall), patches row-major.
__device__ __forceinline__ bool slot_pixel(const WfBuffers &b, const DScene &S, const DParams &P, int slot, int &x, int &ly) {
    const int patch = slot >> 5, lane = slot & 31;
    x = (patch % b.patches_x) * 8 + (lane & 7);
    ly = (patch / b.patches_x) * 4 + (lane >> 3);
    return x < S.cam.width && ly < P.rows;
}

// ------------------------------------------------------------------------------------------------ RAYGEN
// Camera::generate_ray (crt_camera.cpp:7-35) for every pixel slot + entry into the root box, at full SIMD width.
// Rays that enter the box are appended to the level-0 queue (ballot compaction, one atomicAdd per warp; order stays
// patch-major so neighbouring queue entries are neighbouring pixels); the others get their miss record right here.
constexpr int kRaygenBlock = 256;

__device__ __forceinline__ void write_pixel(const DScene &S, float *__restrict__ rgb, uint8_t *__restrict__ rgb8, int x, int ly, V3 c);

template <bool STATS>
__global__ void __launch_bounds__(kRaygenBlock) k_wf_raygen(const __grid_constant__ DScene S, const __grid_constant__ DParams P,
                                                             const __grid_constant__ WfBuffers b, Counters *__restrict__ counters,
                                                             float *__restrict__ rgb, uint8_t *__restrict__ rgb8) {
    const unsigned lane = threadIdx.x & 31u;
    Counters cn{};
    for (int first = blockIdx.x * kRaygenBlock + (threadIdx.x & ~31); first < b.P; first += gridDim.x * kRaygenBlock) {
        const int slot = first + int(lane);
        int x, ly;
        bool have = slot_pixel(b, S, P, slot, x, ly);
        if (P.coherent) { // a fully coherent ray batch: what the non-divergent instruction count per visit is measured on
            x = __shfl_sync(0xffffffffu, x, 0), ly = __shfl_sync(0xffffffffu, ly, 0);
            have = __shfl_sync(0xffffffffu, int(have), 0) != 0;
        }
        Ray r{};
        RaySetup q{};
        q.live = false;
        if (have) {
            r = generate_ray(S.cam, x, global_row(P, ly));
            q = ray_setup(S, r);
            if (STATS) {
                ++cn.rays;
                ++cn.rays_by_depth[0];
                ++cn.node_tests;
                cn.entered += q.live ? 1 : 0;
            }
        }
        const unsigned live = __ballot_sync(0xffffffffu, q.live);
        int base = 0;
        if (live) {
            if (lane == 0)
                base = atomicAdd(&b.ctl->count[0], __popc(live));
            base = __shfl_sync(0xffffffffu, base, 0);
        }
        if (q.live) {
            const int k = base + __popc(live & ((1u << lane) - 1u));
            put_stream(&b.ray0_d[k], make_float4(r.d.x, r.d.y, r.d.z, q.tmin));
            put_stream(&b.ray0_i[k], make_float4(q.inv.x, q.inv.y, q.inv.z, q.tmax));
            put_stream(&b.ray0_slot[k], slot);
        }
        // a ray that misses the root box hits nothing (crt_renderer.cpp:142-144): its pixel is final here, and the
        // level-0 shade only visits the queued rays
        if (have && !q.live)
            write_pixel(S, rgb, rgb8, x, ly, mk(S.bg[0], S.bg[1], S.bg[2]));
    }
    if (STATS)
        flush_counters(cn, counters);
}

// record of a ray of a level >= 1 (written by the shade stage)
__device__ __forceinline__ void write_ray(const DScene &S, const WfBuffers &b, int q, V3 o, V3 d) {
    const RaySetup u = ray_setup(S, Ray{ o, d });
    b.ray_o[q] = make_float4(o.x, o.y, o.z, u.tmin);
    b.ray_d[q] = make_float4(d.x, d.y, d.z, u.live ? u.tmax : -1.0f); // exit < entry marks "nothing to walk"
    b.ray_i[q] = make_float4(u.inv.x, u.inv.y, u.inv.z, 0.0f);
}

// ------------------------------------------------------------------------------------------------ TRACE
} // namespace crtdev
#include "crt_trace.cuh"
#include "crt_trace_lean.cuh"
#include "crt_trace_packet.cuh"
namespace crtdev {

// ------------------------------------------------------------------------------------------------ SHADE
__device__ __forceinline__ void write_pixel(const DScene &S, float *__restrict__ rgb, uint8_t *__restrict__ rgb8, int x, int ly, V3 c) {
    const size_t i = (size_t(ly) * S.cam.width + x) * 3;
    if (rgb)
        put_stream(&rgb[i], c.x), put_stream(&rgb[i + 1], c.y), put_stream(&rgb[i + 2], c.z);
    if (rgb8)
        put_stream(&rgb8[i], quantize(c.x)), put_stream(&rgb8[i + 1], quantize(c.y)), put_stream(&rgb8[i + 2], quantize(c.z));
}

constexpr int kShadeBlock = 128;

__global__ void __launch_bounds__(kShadeBlock) k_wf_shade(const __grid_constant__ DScene S, const __grid_constant__ DParams P,
                                                          const __grid_constant__ WfBuffers b, int level, float *__restrict__ rgb, uint8_t *__restrict__ rgb8,
                                                          bool track_rows, bool queued0) {
    // queued0 (level 0 after k_wf_raygen): visit the rays of the level-0 queue only; the pixels whose ray missed the root
    // box were finished by the ray generation.  Otherwise (literal traversal) level 0 visits every pixel slot.
    WfCtl *ctl = b.ctl;
    const int start = level == 0 ? 0 : level_start(ctl, level);
    const int total = level == 0 ? (queued0 ? ctl->count[0] : b.P) : level_count(ctl, b.cap, level, start);
    const int next_start = level == 0 ? 0 : start + total;
    const unsigned lane = threadIdx.x & 31u;
    const V3 zero = mk(0.0f, 0.0f, 0.0f);
    // warp-uniform trip count: every lane of a warp runs the same iterations, so the collectives below are safe
    for (int first = (blockIdx.x * kShadeBlock + (threadIdx.x & ~31)); first < total; first += gridDim.x * kShadeBlock) {
        int i = first + int(lane);
        int x = 0, ly = 0;
        bool have = i < total;
        if (have && level == 0) {
            if (queued0)
                i = b.ray0_slot[i];
            have = slot_pixel(b, S, P, i, x, ly);
            if (P.coherent) { // the ray this slot was traced with (k_wf_raygen): lane 0 of its patch
                int x0, ly0;
                have = slot_pixel(b, S, P, i & ~31, x0, ly0);
                x = x0, ly = ly0;
            }
        }
        const int node_id = level == 0 ? i : b.P + start + i;
        int kind = kTerminal, children = 0;
        V3 c = zero, albedo = zero, o0 = zero, d0 = zero, o1 = zero, d1 = zero;
        float fres = 0.0f;
        if (have) {
            Ray ray;
            if (level == 0)
                ray = generate_ray(S.cam, x, global_row(P, ly));
            else
                ray = Ray{ xyz(b.ray_o[start + i]), xyz(b.ray_d[start + i]) };
            const float2 hrec = b.hit[node_id];
            const Hit h{ hrec.x, __float_as_int(hrec.y) };
            const bool last = uint32_t(level) >= P.max_ray_depth; // children would be beyond max depth: black (crt_renderer.cpp:47-48)
            if (h.tri < 0) {
                c = mk(S.bg[0], S.bg[1], S.bg[2]);
            } else {
                const Surface s = make_surface(S, ray, h);
                const int4 m = __ldg(&S.materials[s.material]);
                if (m.x == 0) {
                    c = add_direct_light(S, zero, s, sample_texture(S, m.y, s), [](V3, float) { return false; }) / float(P.gi_rays + 1u);
                } else if (m.x == 1) {
                    albedo = sample_texture(S, m.y, s);
                    if (!S.reflections_on) {
                        c = albedo;
                    } else if (last) {
                        c = tint(albedo, zero);
                    } else {
                        kind = kNodeReflect, children = 1;
                        o0 = s.P + s.n * P.reflection_bias, d0 = reflect(ray.d, s.n);
                    }
                } else if (m.x == 2) {
                    if (S.refractions_on) {
                        V3 n = s.n;
                        float outside = 1.0f, inside = __int_as_float(m.z);
                        if (dot(ray.d, n) > 0.0f) {
                            n = -n;
                            const float tmp = outside;
                            outside = inside, inside = tmp;
                        }
                        V3 rd = ray.d;
                        const bool refracted = refract(rd, n, outside, inside);
                        fres = fresnel_term(1.0f + dot(ray.d, n));
                        if (last) {
                            c = refracted ? zero * fres + zero * (1.0f - fres) : zero;
                        } else {
                            kind = refracted ? kNodeRefract : kNodeTir;
                            children = refracted ? 2 : 1;
                            o0 = s.P + n * P.reflection_bias, d0 = reflect(ray.d, n);
                            o1 = s.P + (-n) * 1e-2f, d1 = rd; // refraction bias is always the default (crt_ray.h:33)
                        }
                    }
                } else {
                    c = sample_texture(S, m.y, s);
                }
            }
        }
        // ---- warp-aggregated allocation of the children in the next level's queue (ballot/shfl compaction)
        int prefix = children;
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, prefix, o);
            if (int(lane) >= o)
                prefix += v;
        }
        const int warp_total = __shfl_sync(0xffffffffu, prefix, 31);
        int base = 0;
        if (warp_total) {
            if (lane == 0)
                base = atomicAdd(&ctl->count[level + 1], warp_total);
            base = __shfl_sync(0xffffffffu, base, 0);
        }
        const unsigned pend_mask = __ballot_sync(0xffffffffu, level == 0 && children > 0);
        int pend_base = 0;
        if (pend_mask) {
            // rows that resolve will write again after this kernel: the host copies the frame out early and only
            // these rows a second time (crt_cuda_render)
            // (only when the host asked for it: two atomics per warp on two hot words cost 0.06 ms of C4's frame)
            if (track_rows) {
                const bool pending = level == 0 && children > 0;
                const int hi = __reduce_max_sync(0xffffffffu, pending ? ly + 1 : 0);
                const int lo_inv = __reduce_max_sync(0xffffffffu, pending ? 0x7fffffff - ly : 0);
                if (lane == 0) {
                    atomicMax(&ctl->pend_hi, hi);
                    atomicMax(&ctl->pend_lo_inv, lo_inv);
                }
            }
            if (lane == 0)
                pend_base = atomicAdd(&ctl->pending0, __popc(pend_mask));
            pend_base = __shfl_sync(0xffffffffu, pend_base, 0);
        }
        if (!have)
            continue;
        if (children) {
            const int q = next_start + base + (prefix - children);
            if (q + children > b.cap) {
                ctl->overflow = 1;
                kind = kTerminal, children = 0, c = zero; // frame is discarded and re-rendered with a larger pool
            } else {
                write_ray(S, b, q, o0, d0);
                if (children == 2)
                    write_ray(S, b, q + 1, o1, d1);
                b.rec[node_id] = make_int4(kind, b.P + q, __float_as_int(fres), 0);
                b.alb[node_id] = make_float4(albedo.x, albedo.y, albedo.z, 0.0f);
                if (level == 0)
                    b.pend[pend_base + __popc(pend_mask & ((1u << lane) - 1u))] = i;
            }
        }
        if (!children) {
            if (level == 0) {
                write_pixel(S, rgb, rgb8, x, ly, c);
            } else {
                b.rec[node_id].x = kTerminal;
                b.col[node_id] = make_float4(c.x, c.y, c.z, 0.0f);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ RESOLVE
__device__ __forceinline__ V3 combine(const WfBuffers &b, const int4 rec, int node_id) {
    const V3 c0 = xyz(b.col[rec.y]);
    if (rec.x == kNodeReflect)
        return tint(xyz(b.alb[node_id]), c0); // albedo * shade(reflection), crt_renderer.cpp:106
    if (rec.x == kNodeTir)
        return c0; // :133
    const float f = __int_as_float(rec.z);
    return c0 * f + xyz(b.col[rec.y + 1]) * (1.0f - f); // :131
}

__global__ void __launch_bounds__(256) k_wf_resolve(const __grid_constant__ DScene S, const __grid_constant__ DParams P, const __grid_constant__ WfBuffers b,
                                                    int level, float *__restrict__ rgb, uint8_t *__restrict__ rgb8) {
    const WfCtl *ctl = b.ctl;
    // a level outgrew the pool: the host discards this frame and renders it again.  Nothing to resolve -- and `pend`
    // has holes (pending0 counts the pixels whose children did not fit, their entries were never written)
    if (ctl->overflow)
        return;
    if (level == 0) {
        const int total = ctl->pending0;
        for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < total; k += gridDim.x * blockDim.x) {
            const int slot = b.pend[k];
            int x, ly;
            slot_pixel(b, S, P, slot, x, ly);
            write_pixel(S, rgb, rgb8, x, ly, combine(b, b.rec[slot], slot));
        }
        return;
    }
    const int start = level_start(ctl, level), total = level_count(ctl, b.cap, level, start);
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < total; k += gridDim.x * blockDim.x) {
        const int node_id = b.P + start + k;
        const int4 rec = b.rec[node_id];
        if (rec.x != kTerminal) {
            const V3 c = combine(b, rec, node_id);
            b.col[node_id] = make_float4(c.x, c.y, c.z, 0.0f);
        }
    }
}

// ------------------------------------------------------------------------------------------------ host driver
inline void wavefront_free(WavefrontState &w) {
    cudaFree(w.b.ray_o), cudaFree(w.b.ray_d), cudaFree(w.b.ray_i), cudaFree(w.b.ray0_d), cudaFree(w.b.ray0_i), cudaFree(w.b.ray0_slot), cudaFree(w.b.hit), cudaFree(w.b.rec), cudaFree(w.b.alb), cudaFree(w.b.col), cudaFree(w.b.pend),
        cudaFree(w.b.ctl);
    cudaFreeHost(w.h_flags);
    w = WavefrontState{};
}

inline cudaError_t wavefront_reserve(WavefrontState &w, int P, bool secondary) {
    cudaError_t e = cudaSuccess;
    if (!w.b.ctl) {
        if ((e = cudaMalloc(&w.b.ctl, sizeof(WfCtl))) != cudaSuccess)
            return e;
        if ((e = cudaHostAlloc(&w.h_flags, (3 + kMaxLevels) * sizeof(int), cudaHostAllocDefault)) != cudaSuccess)
            return e;
        w.h_count = w.h_flags + 3;
        std::memset(w.h_flags, 0, (3 + kMaxLevels) * sizeof(int));
        cudaDeviceProp prop;
        int dev = 0;
        cudaGetDevice(&dev);
        cudaGetDeviceProperties(&prop, dev);
        w.sm_count = prop.multiProcessorCount;
    }
    const size_t cap = secondary ? size_t(double(P) * w.grow) + 1024 : 0;
    const size_t nodes = size_t(P) + cap;
    if (nodes > w.node_capacity) {
        cudaFree(w.b.hit);
        w.b.hit = nullptr, w.node_capacity = 0;
        if ((e = cudaMalloc(&w.b.hit, nodes * sizeof(float2))) != cudaSuccess)
            return e;
        w.node_capacity = nodes;
    }
    if (secondary && nodes > w.tree_capacity) { // rec/alb/col exist only once a scene with secondary rays was seen
        cudaFree(w.b.rec), cudaFree(w.b.alb), cudaFree(w.b.col);
        w.b.rec = nullptr, w.b.alb = nullptr, w.b.col = nullptr, w.tree_capacity = 0;
        if ((e = cudaMalloc(&w.b.rec, nodes * sizeof(int4))) != cudaSuccess)
            return e;
        if ((e = cudaMalloc(&w.b.alb, nodes * sizeof(float4))) != cudaSuccess)
            return e;
        if ((e = cudaMalloc(&w.b.col, nodes * sizeof(float4))) != cudaSuccess)
            return e;
        w.tree_capacity = nodes;
    }
    if (cap > w.ray_capacity) {
        cudaFree(w.b.ray_o), cudaFree(w.b.ray_d), cudaFree(w.b.ray_i);
        w.b.ray_o = nullptr, w.b.ray_d = nullptr, w.b.ray_i = nullptr, w.ray_capacity = 0;
        if ((e = cudaMalloc(&w.b.ray_i, cap * sizeof(float4))) != cudaSuccess)
            return e;
        if ((e = cudaMalloc(&w.b.ray_o, cap * sizeof(float4))) != cudaSuccess)
            return e;
        if ((e = cudaMalloc(&w.b.ray_d, cap * sizeof(float4))) != cudaSuccess)
            return e;
        w.ray_capacity = cap;
    }
    if (size_t(P) > w.ray0_capacity) {
        cudaFree(w.b.ray0_d), cudaFree(w.b.ray0_i), cudaFree(w.b.ray0_slot);
        w.b.ray0_d = nullptr, w.b.ray0_i = nullptr, w.b.ray0_slot = nullptr, w.ray0_capacity = 0;
        if ((e = cudaMalloc(&w.b.ray0_d, size_t(P) * sizeof(float4))) != cudaSuccess)
            return e;
        if ((e = cudaMalloc(&w.b.ray0_i, size_t(P) * sizeof(float4))) != cudaSuccess)
            return e;
        if ((e = cudaMalloc(&w.b.ray0_slot, size_t(P) * sizeof(int))) != cudaSuccess)
            return e;
        w.ray0_capacity = size_t(P);
    }
    if (secondary && size_t(P) > w.pend_capacity) {
        cudaFree(w.b.pend);
        w.b.pend = nullptr, w.pend_capacity = 0;
        if ((e = cudaMalloc(&w.b.pend, size_t(P) * sizeof(int))) != cudaSuccess)
            return e;
        w.pend_capacity = size_t(P);
    }
    w.b.P = P;
    w.b.cap = int(cap);
    return cudaSuccess;
}

template <int TRAV>
inline void wavefront_launch_trace(const WavefrontState &w, const DScene &S, const DParams &dp, bool stats, int level, int blocks, Counters *counters,
                                   cudaStream_t stream, bool light) {
    if (TRAV == CRT_TRAVERSAL_KD && light) {
        if (stats)
            k_wf_trace_lean<true><<<blocks, kTraceBlock, 0, stream>>>(S, dp, w.b, level, counters);
        else
            k_wf_trace_lean<false><<<blocks, kTraceBlock, 0, stream>>>(S, dp, w.b, level, counters);
        return;
    }
    if (stats)
        k_wf_trace<TRAV, true><<<blocks, kTraceBlock, 0, stream>>>(S, dp, w.b, level, counters);
    else
        k_wf_trace<TRAV, false><<<blocks, kTraceBlock, 0, stream>>>(S, dp, w.b, level, counters);
}

// `secondary` = the scene can spawn reflection / refraction rays at all (known from its materials on the host)
// `light_leaves` = the walk tree's leaves hold a few triangles each (SAH tree): k_wf_trace_lean instead of k_wf_trace
inline cudaError_t wavefront_render(WavefrontState &w, const DScene &S, const DParams &dp, int traversal, bool stats, bool secondary, float *rgb,
                                    uint8_t *rgb8, Counters *counters, cudaStream_t stream, int *launches, bool light_leaves,
                                    int first_levels_only = 0, cudaEvent_t after_primary = nullptr, bool packet_primary = false) {
    const int patches_x = (S.cam.width + 7) / 8, patches_y = (dp.rows + 3) / 4;
    const int P = patches_x * patches_y * 32;
    secondary = secondary && dp.max_ray_depth > 0;
    cudaError_t e = wavefront_reserve(w, P, secondary);
    if (e != cudaSuccess)
        return e;
    w.b.patches_x = patches_x;
    if ((e = cudaMemsetAsync(w.b.ctl, 0, sizeof(WfCtl), stream)) != cudaSuccess)
        return e;
    // first_levels_only = L > 0 (hybrid pipeline): raygen, then trace and shade of levels 0 .. L-1 only; the caller
    // evaluates the subtrees of the queued level-L rays (k_render_persistent<SEEDED>) and runs the resolves of levels L-1 .. 0
    const int all_levels = secondary ? int(dp.max_ray_depth) + 1 : 1;
    const int levels = first_levels_only > 0 ? min(first_levels_only, all_levels) : all_levels;
    const int trace_blocks = w.sm_count * (light_leaves && traversal == CRT_TRAVERSAL_KD ? CRT_N_MINBLOCKS : 8);
    int n = 0;
    // primary level of the k-d walk as warp packets (crt_trace_packet.cuh): ray generation, root-box entry and the walk
    // of a whole 8x4 patch in one kernel, no ray records
    const bool packet = packet_primary && traversal == CRT_TRAVERSAL_KD;
    const bool queued0 = traversal != CRT_TRAVERSAL_LITERAL; // k_wf_raygen / k_wf_trace_packet queue the rays that enter the root box
    if (packet) {
        const int pb = (P / 32 + kTraceWarps - 1) / kTraceWarps; // one patch per warp
        if (stats)
            k_wf_trace_packet<true><<<pb, kTraceBlock, 0, stream>>>(S, dp, w.b, counters, rgb, rgb8);
        else
            k_wf_trace_packet<false><<<pb, kTraceBlock, 0, stream>>>(S, dp, w.b, counters, rgb, rgb8);
        ++n;
    } else if (traversal != CRT_TRAVERSAL_LITERAL) {
        const int rb = min(w.sm_count * 16, (P + kRaygenBlock - 1) / kRaygenBlock);
        if (stats)
            k_wf_raygen<true><<<rb, kRaygenBlock, 0, stream>>>(S, dp, w.b, counters, rgb, rgb8);
        else
            k_wf_raygen<false><<<rb, kRaygenBlock, 0, stream>>>(S, dp, w.b, counters, rgb, rgb8);
        ++n;
    }
    for (int level = 0; level < levels; ++level) {
        const int tb = level == 0 ? min(trace_blocks, (P + kTraceBlock - 1) / kTraceBlock) : trace_blocks;
        if (packet && level == 0)
            --n; // traced above
        else if (traversal == CRT_TRAVERSAL_LITERAL)
            wavefront_launch_trace<CRT_TRAVERSAL_LITERAL>(w, S, dp, stats, level, tb, counters, stream, false);
        else
            wavefront_launch_trace<CRT_TRAVERSAL_KD>(w, S, dp, stats, level, tb, counters, stream, light_leaves);
        const int sb = level == 0 ? min(w.sm_count * 32, (P + kShadeBlock - 1) / kShadeBlock) : w.sm_count * 8;
        k_wf_shade<<<sb, kShadeBlock, 0, stream>>>(S, dp, w.b, level, rgb, rgb8, level == 0 && after_primary != nullptr, level == 0 && queued0);
        n += 2;
        // every pixel whose primary ray ends in a terminal surface is final now; the rest are rewritten by resolve
        if (level == 0 && after_primary && (e = cudaEventRecord(after_primary, stream)) != cudaSuccess)
            return e;
    }
    for (int level = levels - 2; level >= 0 && first_levels_only <= 0; --level) {
        k_wf_resolve<<<w.sm_count * 4, 256, 0, stream>>>(S, dp, w.b, level, rgb, rgb8);
        ++n;
    }
    if (secondary) {
        e = cudaMemcpyAsync(w.h_flags, &w.b.ctl->overflow, 3 * sizeof(int), cudaMemcpyDeviceToHost, stream);
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(w.h_count, w.b.ctl->count, kMaxLevels * sizeof(int), cudaMemcpyDeviceToHost, stream);
    } else if (w.h_flags) {
        w.h_flags[0] = 0; // a frame without secondary rays cannot overflow: never inherit an older frame's flag
    }
    if (launches)
        *launches = n;
    return e != cudaSuccess ? e : cudaGetLastError();
}

} // namespace crtdev
