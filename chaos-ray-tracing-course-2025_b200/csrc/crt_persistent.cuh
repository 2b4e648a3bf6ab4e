// PERSISTENT per-pixel pipeline: one kernel renders the frame.  A lane owns a PIXEL and keeps it through the whole
// recursion of shade_ray (src/core/crt_renderer.cpp:46-145), depth-first with the explicit frame stack of
// k_render_mega (crt_kernels.cuh) -- so evaluation order, RNG order (one PCG32 stream per pixel, crt_random.h) and
// the nesting in which colours are combined are the reference's -- while the closest-hit queries are served by the
// warp-cooperative walk of k_wf_trace_lean: speculative walk, leaves parked and tested warp-wide, work stealing.
//
// Why (ncu r1p): the level-synchronous wavefront pays one tail per bounce level (levels 2-8 of C4 hold
// 18 % of the rays and take 55 % of the frame) and cannot run GI at all; the per-pixel megakernel can, but its lanes
// sit in unrelated phases of trace / shade code (300 Mrays/s on the GI scene).  Here a warp alternates between three
// phases, each entered only when enough lanes want it:
//   WALK   every lane with an unfinished query visits up to CRT_P_STEPS nodes (CRT_P_STEPS_GI with GI);
//   TEST   parked leaves of all lanes, one (ray, triangle) pair per lane and iteration;
//   SHADE  lanes whose query is finished run the continuation of their pixel's recursion: shade the hit, then either
//          start the next query (reflection, refraction, GI child, sibling) or finish the pixel and take a new one.
#pragma once

namespace crtdev {

#ifndef CRT_P_MINBLOCKS
#define CRT_P_MINBLOCKS 8 // 64 registers (r2b: 2.92 -> 2.84 ms on C4 together with 9 blocks of the trace kernel)
#endif
#ifndef CRT_P_MINBLOCKS_GI
#define CRT_P_MINBLOCKS_GI 7 // 72 registers, nothing spilled (r2j: 8 blocks = 64 registers spill ~100 B in the GI continuation: 177.6 -> 163.6 ms)
#endif
#ifndef CRT_P_PATIENCE_GI
#define CRT_P_PATIENCE_GI 16
#endif
#ifndef CRT_P_PATIENCE
#define CRT_P_PATIENCE 8
#endif
#ifndef CRT_P_STEPS
#define CRT_P_STEPS 10 // (r2y: swept together with CRT_N_STEPS; r3g, with levels 1-8 of C4 in this kernel: 6 / 8 / 10 / 12 -> 2.29 / 2.27 / 2.25 / 2.28 ms)
#endif
#ifndef CRT_P_STEPS_GI
#define CRT_P_STEPS_GI 2 // GI queries are short (a diffuse bounce ends at the next surface): more frequent phase changes pay
#endif
#ifndef CRT_P_SHADE_TRIGGER_GI
#define CRT_P_SHADE_TRIGGER_GI 20 // the GI continuation is long (sinf / cosf, basis, PCG): run it for well-filled warps only (r2k: 28 -> 164.0 ms, 24 -> 159.9, 20 -> 157.8)
#endif
#ifndef CRT_P_SHADE_TRIGGER
#define CRT_P_SHADE_TRIGGER 12
#endif
#ifndef CRT_P_EAGER_MISS
#define CRT_P_EAGER_MISS 0 // 1: queries without a hit and freshly taken work items continue at once, hits wait for the batch (step 29:
                           // parity-green, measured, slower -- C4 2.27 -> 2.30 ms, one rank's share of an 8-rank frame 0.66 -> 0.70 ms)
#endif
#ifndef CRT_P_FAST_UNWIND
#define CRT_P_FAST_UNWIND 1 // return path of reflection chains: the frames' loads four at a time (step 25)
#endif

struct PersistCtl {
    int next_slot; // pixel slots handed out so far
};

// slot -> pixel, 8x4 patches, patches row-major (neighbouring slots are neighbouring pixels)
__device__ __forceinline__ bool persist_slot_pixel(int slot, int patches_x, int width, int rows, int &x, int &ly) {
    const int patch = slot >> 5, l = slot & 31;
    x = (patch % patches_x) * 8 + (l & 7);
    ly = (patch / patches_x) * 4 + (l >> 3);
    return x < width && ly < rows;
}

// one light of add_direct_light (crt_device.cuh), same operation sequence
struct LightTerm {
    V3 dir;
    float r2, cos_law, area, intensity;
};
__device__ __forceinline__ LightTerm light_term(const DScene &S, int i, V3 Pnt, V3 n) {
    const float four_pi = 4.0f * 3.14159274101257324219f;
    const float4 L = __ldg(&S.lights[i]);
    LightTerm t;
    V3 dir = xyz(L) - Pnt;
    t.r2 = dot(dir, dir);
    t.dir = dir / sqrtf(t.r2);
    const float ln = dot(t.dir, n);
    t.cos_law = (0.0f < ln) ? ln : 0.0f;
    t.area = four_pi * t.r2;
    t.intensity = L.w;
    return t;
}

// SEEDED: the work items are not pixels but the level-1 rays the wavefront's level-0 shade stage queued (hybrid
// pipeline: the dense, coherent primary level runs as raygen / trace / shade kernels at full SIMD width, everything
// deeper runs here).  A lane takes a level-1 ray, evaluates its whole subtree depth-first and stores the subtree's
// colour in the level-1 node record, which the level-0 resolve kernel combines as before.
template <bool STATS, bool GI, bool SEEDED, bool SHADOWS>
__global__ void __launch_bounds__(kTraceBlock, GI ? CRT_P_MINBLOCKS_GI : CRT_P_MINBLOCKS) k_render_persistent(const __grid_constant__ DScene S, const __grid_constant__ DParams P,
                                                                                    PersistCtl *__restrict__ ctl, int patches_x, int n_slots_arg,
                                                                                    float *__restrict__ rgb, uint8_t *__restrict__ rgb8,
                                                                                    Counters *__restrict__ counters, const __grid_constant__ WfBuffers b) {
    // SEEDED: `patches_x` carries the bounce level whose queued rays are the seeds (1 = everything below the primary level)
    const int base_depth = SEEDED ? patches_x : 0;
    const int seed_start = SEEDED ? level_start(b.ctl, base_depth) : 0;
    const int n_slots = SEEDED ? level_count(b.ctl, b.cap, base_depth, seed_start) : n_slots_arg;
    Counters cn{};
    __shared__ LightWarpShared shared[kTraceWarps];
    LightWarpShared &ws = shared[threadIdx.x >> 5];
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    const unsigned full = 0xffffffffu;
    const V3 bg = mk(S.bg[0], S.bg[1], S.bg[2]);
    const float gi_div = float(P.gi_rays + 1u);
    auto no_shadow = [](V3, float) { return false; };

    crtwalk::LeanWalker w;
    uint2 spill[kStack]; // stack entries below the shared-memory window
    w.node = crtwalk::kDone, w.sp = 1, w.lo = 0, w.flagged = 0, w.tmin = 0.0f, w.tmax = 0.0f;
    const uint32_t empty_leaf = uint32_t(S.kd_empty_leaf);
    // ---- pixel state
    int px = 0, ply = -1; // my pixel (local row), ply < 0 = no pixel
    int depth = 0;
    bool live = false;    // the current query entered the root box (its result is in ws.best[lane])
    bool fresh = false;   // pixel just taken: its primary ray has not been generated yet (done in the SHADE phase)
    bool shadow_q = false; // the running query is a shadow ray of frames[depth]'s light loop, not a shade_ray call
    Pcg32 rng{};
    Frame frames[kMaxDepthCap + 1];
    // ---- walk state (see k_wf_trace_lean)
    int parked0 = -1, parked1 = -1; // parked leaves (node ids)
    bool exhausted = false;
    unsigned src = lane;
    bool helper = false;
    int stolen = 1;   // entries [1, stolen) of my stack are tombstones (helpers took them over)
    int patience = 0; // rounds some lane has been waiting for a SHADE phase (warp-uniform)

    for (;;) {
        // ---- refill: idle lanes take the next pixel slots
        const unsigned busy = __ballot_sync(full, ply >= 0 || helper);
        if (!exhausted && __popc(busy) < CRT_N_REFILL) {
            const unsigned idle = ~busy;
            int base = 0;
            if (lane == 0)
                base = atomicAdd(&ctl->next_slot, __popc(idle));
            base = __shfl_sync(full, base, 0);
            if (base + __popc(idle) >= n_slots)
                exhausted = true;
            if (ply < 0 && !helper) {
                const int slot = base + __popc(idle & lt_mask);
                int x, ly;
                if (slot < n_slots && (SEEDED || persist_slot_pixel(slot, patches_x, S.cam.width, P.rows, x, ly))) {
                    if (SEEDED)
                        x = slot, ly = 0; // px = index of my seed ray within its level
                    px = x, ply = ly, depth = base_depth;
                    fresh = true, live = false;
                    src = lane, stolen = 1;
                    ws.helpers[lane] = 0;
                    w.node = crtwalk::kDone;
                    parked0 = -1, parked1 = -1;
                }
            }
            __syncwarp();
        }
        if (__ballot_sync(full, ply >= 0 || helper) == 0u) {
            if (exhausted)
                break;
            continue;
        }
        // ---- the queue is empty and some lanes have nothing to do: they take over pending subtrees of the lanes that
        //      are still walking (crt_trace_lean.cuh: lean_steal)
        if (exhausted)
            lean_steal<STATS>(ws, w, lane, lt_mask, ply >= 0, helper, src, stolen, empty_leaf, cn);
        // ---- walk: a few node visits per lane; a leaf is parked and the walk goes on (crt_trace_lean.cuh: lean_walk)
        bool blocked = false;
        if ((ply >= 0 || helper) && !w.done()) // fresh pixels have a finished walk
            blocked = lean_walk<STATS, (GI ? CRT_P_STEPS_GI : CRT_P_STEPS)>(S, ws, w, spill, lane, src, parked0, parked1, cn);
        // ---- test phase: all parked leaves of all lanes at once
        const unsigned parked = __ballot_sync(full, parked0 >= 0);
        const unsigned walkers = __ballot_sync(full, (ply >= 0 || helper) && !w.done() && !blocked);
        const unsigned waiting = __ballot_sync(full, (ply >= 0 || helper) && w.done() && parked0 >= 0);
        // (once the queue is empty nothing is gained by batching: see k_wf_trace_lean)
        const bool fire = parked != 0u && (__popc(parked) >= CRT_N_TRIGGER || __any_sync(full, blocked) || walkers == 0u ||
                                           __popc(waiting) >= (exhausted ? 1 : CRT_N_WAITERS));
        if (fire) {
            lean_test_parked<STATS>(S, ws, lane, src, parked0, parked1, cn);
            parked0 = -1, parked1 = -1;
        }
        // ---- helpers that finished hand their subtree's result back (it is already merged into the owner's best)
        if (helper && w.done() && parked0 < 0) {
            atomicSub(&ws.helpers[src], 1);
            helper = false;
            src = lane;
        }
        __syncwarp();
        // ---- SHADE phase: lanes whose query is finished (or whose pixel is new) continue their pixel's recursion
        const bool ready = ply >= 0 && w.done() && parked0 < 0 && (!live || *(volatile int *)&ws.helpers[lane] == 0);
#if CRT_P_EAGER_MISS
        // (experiment, off -- step 29.)  What a lane waits for a well-filled phase with is the LONG continuation: surface,
        // material, the next ray.  A query that hit nothing only returns the background colour through its frames, and a
        // work item just taken only needs its ray set up: here those go at once and the batch holds hits only.  Measured
        // slower: the short phases interrupt the walking lanes more often than the freed lanes help.
        const bool at_once = !GI && ready && !(SHADOWS && shadow_q) && (fresh || !live || unsigned(ws.best[lane] & 0xffffffffull) == 0xffffffffu);
        const unsigned ready_m = __ballot_sync(full, ready && !at_once);
        const bool any_at_once = __any_sync(full, at_once);
#else
        const bool at_once = false, any_at_once = false;
        const unsigned ready_m = __ballot_sync(full, ready);
#endif
        const unsigned walking_m = __ballot_sync(full, (ply >= 0 || helper) && !w.done());
        patience = ready_m != 0u ? patience + 1 : 0;
        // (with the queue empty a ready lane continues at once: its pixel's remaining bounces are the critical path)
        const bool batch = ready_m != 0u && (__popc(ready_m) >= (GI ? CRT_P_SHADE_TRIGGER_GI : CRT_P_SHADE_TRIGGER) || walking_m == 0u ||
                                             patience >= (exhausted && !GI ? 1 : (GI ? CRT_P_PATIENCE_GI : CRT_P_PATIENCE)));
        if (batch || any_at_once) {
            patience = batch ? 0 : patience;
            if (ready && (batch || at_once)) {
                int act = 0; // 0: consume the finished query, 1: shade_ray(ray, depth) entry, 2: return c to the callers
                Ray ray{};
                V3 c = mk(0.0f, 0.0f, 0.0f);
                if (fresh) {
                    fresh = false;
                    if (SEEDED) {
                        ray = Ray{ xyz(b.ray_o[seed_start + px]), xyz(b.ray_d[seed_start + px]) };
                    } else {
                        const int y = global_row(P, ply);
                        if (GI)
                            rng = make_pcg(uint32_t(px), uint32_t(y));
                        ray = generate_ray(S.cam, px, y);
                    }
                    depth = base_depth;
                    act = 1;
                }
                for (;;) {
                    if (act == 1) {
                        if (uint32_t(depth) > P.max_ray_depth) { // crt_renderer.cpp:47-48: black before tracing
                            c = mk(0.0f, 0.0f, 0.0f);
                            act = 2;
                            continue;
                        }
                        if (STATS) {
                            ++cn.rays;
                            ++cn.rays_by_depth[depth];
                            ++cn.node_tests;
                        }
                        const RaySetup q = ray_setup(S, ray);
                        if (STATS && q.live)
                            ++cn.entered;
                        if (q.live) { // leave the phase: the walk answers this query
                            w.start(ws, lane, ray.o.x, ray.o.y, ray.o.z, ray.d.x, ray.d.y, ray.d.z, q.inv.x, q.inv.y, q.inv.z, q.tmin, q.tmax);
                            ws.best[lane] = kNoHitKey;
                            live = true;
                            stolen = 1;
                            break;
                        }
                        live = false; // misses the root box (or NaN direction): a finished query without a hit
                        act = 0;
                        continue;
                    }
                    if (act == 3) { // shadow query of the light loop (flagged mode, crt_renderer.cpp:88-92 as intended)
                        if (STATS) {
                            ++cn.shadow_rays;
                            ++cn.node_tests;
                        }
                        shadow_q = true;
                        const RaySetup q = ray_setup(S, ray);
                        if (q.live) {
                            w.start(ws, lane, ray.o.x, ray.o.y, ray.o.z, ray.d.x, ray.d.y, ray.d.z, q.inv.x, q.inv.y, q.inv.z, q.tmin, q.tmax);
                            ws.best[lane] = kNoHitKey;
                            live = true;
                            stolen = 1;
                            break;
                        }
                        live = false;
                        act = 0;
                        continue;
                    }
                    if (act == 0) {
                        Hit h{ kInf, -1 };
                        if (live) {
                            const unsigned long long key = ws.best[lane];
                            h.t = __uint_as_float(unsigned(key >> 32));
                            h.tri = int(unsigned(key & 0xffffffffull));
                            ray = Ray{ mk(ws.ray[0][lane], ws.ray[1][lane], ws.ray[2][lane]), mk(ws.ray[3][lane], ws.ray[4][lane], ws.ray[5][lane]) };
                        }
                        live = false;
                        if (SHADOWS && shadow_q) { // lit iff nothing is hit within the light's distance; then the next light
                            shadow_q = false;
                            Frame &f = frames[depth];
                            const LightTerm lt = light_term(S, f.i, f.b, f.c);
                            if (!(h.tri >= 0 && !(h.t * h.t > lt.r2)))
                                f.a = f.a + ((f.g * lt.intensity) / lt.area) * lt.cos_law;
                            ++f.i;
                            if (f.i < S.n_lights) {
                                ray = Ray{ f.b + f.c * P.shadow_bias, light_term(S, f.i, f.b, f.c).dir };
                                act = 3;
                            } else {
                                c = f.a / gi_div;
                                act = 2;
                            }
                            continue;
                        }
                        act = 2;
                        if (h.tri < 0) {
                            c = bg;
                            continue;
                        }
                        const Surface s = make_surface(S, ray, h);
                        const int4 m = __ldg(&S.materials[s.material]);
                        Frame &f = frames[depth];
                        if (m.x == 0) { // Diffuse
                            const V3 albedo = sample_texture(S, m.y, s);
                            if (GI) {
                                f.kind = kDiffuseGi, f.i = 0;
                                f.a = mk(0.0f, 0.0f, 0.0f), f.b = s.P, f.c = s.n, f.e = ray.d, f.g = albedo;
                                if (uint32_t(depth) + 1u > P.max_ray_depth) {
                                    // every bounce ray of this surface lies beyond max depth: shade_ray returns black
                                    // before tracing (crt_renderer.cpp:47-48), so their directions are never used and
                                    // only the pixel's generator moves on, two draws per ray (:66-70).  This is the
                                    // last level, i.e. most of the rays of a GI frame (74 % in 15-01/scene2): skipping
                                    // 4 x (two binary64 sincos, a cross product, two normalisations) per hit.
                                    for (uint32_t k = 0; k < 2u * P.gi_rays; ++k)
                                        (void)rng.next();
                                    f.i = int(P.gi_rays) - 1; // as if the last bounce ray had just returned black
                                    c = mk(0.0f, 0.0f, 0.0f);
                                    ++depth, act = 2;
                                } else {
                                    ray = Ray{ s.P + s.n * P.diffuse_reflection_bias, gi_direction(ray.d, s.n, rng) };
                                    ++depth, act = 1;
                                }
                            } else if (SHADOWS && P.shadows && S.n_lights > 0) {
                                f.kind = kDirectShadow, f.i = 0;
                                f.a = mk(0.0f, 0.0f, 0.0f), f.b = s.P, f.c = s.n, f.g = albedo;
                                ray = Ray{ s.P + s.n * P.shadow_bias, light_term(S, 0, s.P, s.n).dir };
                                act = 3;
                            } else {
                                c = add_direct_light(S, mk(0.0f, 0.0f, 0.0f), s, albedo, no_shadow) / gi_div;
                            }
                        } else if (m.x == 1) { // Reflective
                            const V3 albedo = sample_texture(S, m.y, s);
                            if (!S.reflections_on) {
                                c = albedo;
                            } else {
                                f.kind = kReflect, f.a = albedo;
                                ray = Ray{ s.P + s.n * P.reflection_bias, reflect(ray.d, s.n) };
                                ++depth, act = 1;
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
                                ++depth, act = 1;
                            }
                        } else { // Constant
                            c = sample_texture(S, m.y, s);
                        }
                        continue;
                    }
                    // act == 2: return c to the callers (the frame stack is the recursion's return path)
                    if (CRT_P_FAST_UNWIND && !GI) {
                        // A chain of mirrors returns through one kReflect frame per bounce: c = tint(albedo, c), innermost
                        // first (crt_renderer.cpp:106).  Frame by frame that is one local-memory load after the other,
                        // each waited for by the whole warp while one or two lanes unwind (ncu r2s: 12 % of the seeded
                        // kernel's warp time).  Here the four frames on top are loaded TOGETHER (frames below the base
                        // are read and ignored) and as many of them as are kReflect are applied, in the same order.
                        while (depth > base_depth) {
                            const int avail = depth - base_depth;
                            const int d1 = depth - 1, d2 = max(depth - 2, 0), d3 = max(depth - 3, 0), d4 = max(depth - 4, 0);
                            const int k1 = frames[d1].kind, k2 = frames[d2].kind, k3 = frames[d3].kind, k4 = frames[d4].kind;
                            const V3 a1 = frames[d1].a, a2 = frames[d2].a, a3 = frames[d3].a, a4 = frames[d4].a;
                            const bool r1 = k1 == kReflect, r2 = r1 & (avail > 1) & (k2 == kReflect), r3 = r2 & (avail > 2) & (k3 == kReflect),
                                       r4 = r3 & (avail > 3) & (k4 == kReflect);
                            if (r1)
                                c = tint(a1, c);
                            if (r2)
                                c = tint(a2, c);
                            if (r3)
                                c = tint(a3, c);
                            if (r4)
                                c = tint(a4, c);
                            depth -= int(r1) + int(r2) + int(r3) + int(r4);
                            if (!r4)
                                break;
                        }
                    }
                    if (depth == base_depth) {
                        if (SEEDED)
                            b.col[b.P + seed_start + px] = make_float4(c.x, c.y, c.z, 0.0f);
                        else
                            write_pixel(S, rgb, rgb8, px, ply, c);
                        ply = -1;
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
                        act = 1; // evaluate the refraction ray at the same depth
                        continue;
                    } else if (f.kind == kRefractWaitRefraction) {
                        c = f.a * f.f + c * (1.0f - f.f);
                    } else { // kDiffuseGi
                        f.a = f.a + c;
                        if (uint32_t(++f.i) < P.gi_rays) {
                            ray = Ray{ f.b + f.c * P.diffuse_reflection_bias, gi_direction(f.e, f.c, rng) };
                            act = 1;
                            continue;
                        }
                        if (SHADOWS && P.shadows && S.n_lights > 0) { // the light loop of this surface, one shadow query per light
                            f.kind = kDirectShadow, f.i = 0; // a (GI sum so far), b, c, g stay
                            --depth;
                            ray = Ray{ f.b + f.c * P.shadow_bias, light_term(S, 0, f.b, f.c).dir };
                            act = 3;
                            continue;
                        }
                        Surface s{};
                        s.P = f.b, s.n = f.c;
                        c = add_direct_light(S, f.a, s, f.g, no_shadow) / gi_div;
                    }
                    --depth;
                }
            }
            __syncwarp();
        }
    }
    if (STATS)
        flush_counters(cn, counters);
}

} // namespace crtdev
