// PRIMARY level as warp packets: one warp = one 8x4 pixel patch = 32 camera rays that share their origin and walk
// the tree TOGETHER.  Included by crt_wavefront.cuh after crt_trace_lean.cuh; replaces k_wf_raygen + the level-0
// launch of k_wf_trace_lean (CRT_TRAVERSAL_KD; $CRT_PACKET=0 keeps those).
//
// Why: the per-lane walk treats the 32 rays of a warp as unrelated -- every lane loads its own node, keeps its own
// stack in shared memory, and the lanes drift apart (ncu r1z, level 0: 19.5 of 32 threads per instruction, ~120 thread
// instructions per node visit).  Camera rays of one patch are nearly parallel and visit nearly the same nodes, so here
// the NODE and the STACK OF NODES are the warp's (one uniform load per visit, node ids in shared memory once per
// warp); per lane there is only the ray, its interval [tmin, tmax] in the current cell, and its best hit.  A lane takes
// part in a node when its own walk would have visited it: the per-lane decisions are TraceWalker::step's, word for
// word (same slack, same in-plane and absent-child rules), so every lane tests a superset of the leaves its own walk
// tests and the winner rule min (t, parse index) gives the same hit.  The per-lane intervals of pushed subtrees go to a
// local-memory array indexed by the warp-uniform stack pointer (one coalesced 256 B access per push / pop).
// Triangles of a leaf are loaded once per warp (uniform) and tested by every lane that is active in the leaf.
// Ray generation (Camera::generate_ray, crt_camera.cpp:7-35) and the root-box entry are done here too: no ray record
// is written or read for the primary level (only the slots of the rays that enter the box, for the level-0 shade).
#pragma once

namespace crtdev {

#ifndef CRT_K_MINBLOCKS
#define CRT_K_MINBLOCKS 10 // (48 registers: ten blocks per SM; with 8 the compiler settles at 52 = nine blocks and C1 / C3 lose 3-6 %)
#endif

template <bool STATS>
__global__ void __launch_bounds__(kTraceBlock, CRT_K_MINBLOCKS) k_wf_trace_packet(const __grid_constant__ DScene S, const __grid_constant__ DParams P,
                                                                                  const __grid_constant__ WfBuffers b, Counters *__restrict__ counters,
                                                                                  float *__restrict__ rgb, uint8_t *__restrict__ rgb8) {
    __shared__ int node_stack[kTraceWarps][kStack];
    int *ns = node_stack[threadIdx.x >> 5];
    float2 ivl[kStack]; // my [tmin, tmax] in the cell of each pending node; tmax = -inf: that subtree is not mine to visit
    const unsigned lane = threadIdx.x & 31u;
    const unsigned full = 0xffffffffu;
    const int patches = b.P >> 5;
    const float kOut = -kInf;
    const V3 o = mk(S.cam.position[0], S.cam.position[1], S.cam.position[2]);
    Counters cn{};
    // one patch per warp, handed out by the block scheduler (a shared cursor costs one atomic on a hot word per patch:
    // 15 % of a 0.2 ms frame)
    for (int patch = blockIdx.x * kTraceWarps + int(threadIdx.x >> 5); patch < patches; patch += gridDim.x * kTraceWarps) {
        const int slot = patch * 32 + int(lane);
        int x, ly;
        const bool have = slot_pixel(b, S, P, slot, x, ly);
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
        // a ray that misses the root box hits nothing: its pixel is final here; the others are queued for the level-0
        // shade (slot only -- the shade stage regenerates the camera ray, as it does after k_wf_raygen)
        if (have && !q.live)
            write_pixel(S, rgb, rgb8, x, ly, mk(S.bg[0], S.bg[1], S.bg[2]));
        const unsigned live_mask = __ballot_sync(full, q.live);
        if (live_mask == 0u)
            continue;
        int queue_base = 0;
        if (lane == 0)
            queue_base = atomicAdd(&b.ctl->count[0], __popc(live_mask));
        queue_base = __shfl_sync(full, queue_base, 0);
        if (q.live)
            b.ray0_slot[queue_base + __popc(live_mask & ((1u << lane) - 1u))] = slot;
        const V3 d = r.d, inv = q.inv;
        float tmin = q.tmin, tmax = q.live ? q.tmax : kOut;
        unsigned long long best = kNoHitKey;
        float best_t = kInf;
        int node = 0, sp = 0;
        for (;;) {
            // TraceWalker::prune_at_visit: a cell entered clearly beyond the best hit cannot beat or tie it
            const bool act = tmax != kOut && !(tmin - (fabsf(tmin) * 1e-5f + 1e-6f) > best_t);
            bool descend = false;
            if (__any_sync(full, act)) {
                const int4 n = kd_node_raw(S, node); // (a leaf's {first, count} are stored complemented: crt_walk_lean.h, CRT_LEAF_NEG)
                if (STATS && act && node != 0)
                    ++cn.node_tests;
                if (n.w == 3) {
                    const int leaf_first = crtwalk::kLeafNeg ? ~n.y : n.y, leaf_count = crtwalk::kLeafNeg ? ~n.z : n.z;
                    for (int k = 0; k < leaf_count; ++k) {
                        const uint32_t ref = __ldg(&S.kd_refs[leaf_first + k]);
                        const int tri = int(ref & 0x7FFFFFFFu);
                        const float4 A = __ldg(&S.tri_verts[3 * tri]), B = __ldg(&S.tri_verts[3 * tri + 1]), C = __ldg(&S.tri_verts[3 * tri + 2]);
                        if (act) {
                            if (STATS)
                                ++cn.tri_tests;
                            float t;
                            if (triangle_test_flat(A, B, C, (ref >> 31) != 0, o, d, t)) {
                                t = t + 0.0f; // -0 -> +0 so the bit pattern orders like the value
                                const unsigned long long key = ((unsigned long long)__float_as_uint(t) << 32) | (unsigned long long)uint32_t(tri);
                                if (key < best)
                                    best = key, best_t = t;
                            }
                        }
                    }
                } else {
                    // ---- per lane: TraceWalker::step, with the two children's intervals made explicit
                    const int axis = n.w;
                    const float split = __int_as_float(n.x);
                    const float oa = axis == 0 ? o.x : (axis == 1 ? o.y : o.z);
                    const float ia = axis == 0 ? inv.x : (axis == 1 ? inv.y : inv.z);
                    const bool below = (oa < split) | ((oa == split) & (ia <= 0.0f));
                    const int nearc = below ? n.y : n.z, farc = below ? n.z : n.y;
                    const float tp = (split - oa) * ia;
                    const float slack = fabsf(tp) * 1e-5f + 1e-6f;
                    const bool in_plane = tp != tp;
                    const bool near_only = !in_plane & ((tp <= 0.0f) | (tp > tmax + slack));
                    const bool far_only = !in_plane & !near_only & (tp < tmin - slack);
                    const bool both = !near_only & !far_only;
                    const bool has_near = nearc >= 0, has_far = farc >= 0;
                    const bool do_push = both & has_near & has_far;
                    const bool go_far = far_only | (both & !has_near);
                    // near child: visited next unless the walk goes straight to the far one
                    const bool want_n = act & !go_far & has_near;
                    const float n0 = tmin, n1 = (do_push & !in_plane) ? fminf(tmax, tp) : tmax;
                    // far child: visited next (go_far) or pushed (do_push; an in-plane entry restarts at 0, see pop_fast)
                    const bool want_f = act & has_far & (go_far | do_push);
                    const float f0 = go_far ? ((both & !in_plane) ? fmaxf(tmin, tp) : tmin) : (in_plane ? 0.0f : tp);
                    const float f1 = tmax;
                    // ---- per warp: which of the node's two children (n.y, n.z) does anybody want; the side the common
                    //      origin lies on goes first (the order only matters for how early best_t prunes)
                    const bool want_l = below ? want_n : want_f, want_r = below ? want_f : want_n;
                    const float l0 = below ? n0 : f0, l1 = below ? n1 : f1, r0 = below ? f0 : n0, r1 = below ? f1 : n1;
                    const bool left_first = oa < split;
                    const bool want_a = left_first ? want_l : want_r, want_b = left_first ? want_r : want_l;
                    const float a0 = left_first ? l0 : r0, a1 = left_first ? l1 : r1, b0 = left_first ? r0 : l0, b1 = left_first ? r1 : l1;
                    const int child_a = left_first ? n.y : n.z, child_b = left_first ? n.z : n.y;
                    const bool any_a = __any_sync(full, want_a), any_b = __any_sync(full, want_b);
                    if (any_a & any_b) {
                        ns[sp] = child_b; // every lane stores the same word
                        ivl[sp] = make_float2(b0, want_b ? b1 : kOut);
                        ++sp;
                    }
                    if (any_a) {
                        node = child_a, tmin = a0, tmax = want_a ? a1 : kOut;
                        descend = true;
                    } else if (any_b) {
                        node = child_b, tmin = b0, tmax = want_b ? b1 : kOut;
                        descend = true;
                    }
                }
            }
            if (!descend) {
                if (sp == 0)
                    break;
                --sp;
                node = ns[sp];
                const float2 f = ivl[sp];
                tmin = f.x, tmax = f.y;
            }
        }
        if (q.live)
            b.hit[slot] = make_float2(__uint_as_float(unsigned(best >> 32)), __int_as_float(int(unsigned(best & 0xffffffffull))));
    }
    if (STATS)
        flush_counters(cn, counters);
}

} // namespace crtdev
