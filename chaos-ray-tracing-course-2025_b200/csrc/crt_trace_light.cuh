// TRACE stage for walk trees with LIGHT leaves (the SAH tree of core/crt_kdbuild.cpp: ~2 triangles per leaf, hardly
// any triangle referenced twice).  Included by crt_wavefront.cuh after crt_trace.cuh, whose TraceWalker it reuses.
//
// With such a tree a ray takes ~25-65 node steps but only 1-4 leaf visits and 1-6 triangle tests (ncu capture r1l:
// the candidate ring, the mailbox and the per-round batch bookkeeping of k_wf_trace were 40 % of the issued
// instructions while the triangle tests themselves were 5 %).  This kernel keeps what paid off there -- persistent
// warps with lane-level ray replacement, the speculative walk, the shared-memory stack window, warp-wide branch-free
// triangle tests merged by a 64-bit atomicMin on (t, triangle) -- and drops the rest:
//   * a lane that reaches a leaf parks it in one of two per-lane slots (registers) and walks on;
//   * only when enough leaves are parked (or a lane has both slots full, or nobody can walk) does the warp run ONE
//     test phase: shfl prefix sum of the parked triangle counts, owner lane by binary search, every lane tests
//     whichever (ray, triangle) pair comes next.  After a phase every parked leaf of every lane has been tested, so a
//     ray retires as soon as its walk is over and its slots are empty: no marks, no ring, no mailbox (a triangle
//     referenced from two leaves is simply tested twice; the winner rule min (t, parse index) does not care);
//   * the pop is bare (TraceWalker::pop_fast): whether a popped subtree can still hold a closer hit is decided when
//     it is visited (prune_at_visit), in the instruction stream all lanes share;
//   * when the level's queue is empty, idle lanes take over the bottom stack entries of lanes that still walk.
#pragma once

namespace crtdev {

#ifndef CRT_L_STEPS
#define CRT_L_STEPS 8
#endif
#ifndef CRT_L_TRIGGER
#define CRT_L_TRIGGER 16
#endif
#ifndef CRT_L_WAITERS
#define CRT_L_WAITERS 6
#endif
#ifndef CRT_L_MINBLOCKS
#define CRT_L_MINBLOCKS 9
#endif
#ifndef CRT_L_REFILL
#define CRT_L_REFILL 25
#endif

struct __align__(16) LightWarpShared {
    float ray[9][32];            // origin xyz, direction xyz, 1/direction xyz of each lane's ray
    unsigned long long best[32]; // (float bits of t) << 32 | triangle index
    uint2 stack[kWin][32];       // {node | kFlagBit, entry distance bits}
    int helpers[32];             // lanes currently walking a stolen subtree of this lane's ray
};

template <bool STATS>
__global__ void __launch_bounds__(kTraceBlock, CRT_L_MINBLOCKS) k_wf_trace_light(const __grid_constant__ DScene S, const __grid_constant__ DParams P,
                                                                                 const __grid_constant__ WfBuffers b, int level,
                                                                                 Counters *__restrict__ counters) {
    WfCtl *ctl = b.ctl;
    const int start = level == 0 ? 0 : level_start(ctl, level);
    const int total = level == 0 ? ctl->count[0] : level_count(ctl, b.cap, level, start);
    Counters cn{};
    __shared__ LightWarpShared shared[kTraceWarps];
    LightWarpShared &ws = shared[threadIdx.x >> 5];
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    const unsigned full = 0xffffffffu;
    TraceWalker w;
    uint2 spill[kStack]; // stack entries below the shared-memory window
    w.node = -1, w.sp = 0, w.lo = 0, w.flagged = 0;
    int my = -1;      // index of my ray within the level, -1 = idle
    int node_id = 0;  // where my ray's hit record goes
    bool live = false;
    int p0_first = 0, p0_count = 0, p1_first = 0, p1_count = 0; // parked leaves
    bool exhausted = false;
    unsigned src = lane;  // lane whose ray I am walking: myself, or the owner of the subtree I took over
    bool helper = false;  // I own no ray, I walk a stolen subtree of lane `src`'s ray
    int stolen = 0;       // bottom entries of my stack that helpers have taken over (tombstones)
#ifdef CRT_DEBUG_TAIL
    unsigned rounds = 0;
    unsigned long long t_start, t_exhausted = 0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
#endif
    for (;;) {
        // ---- refill idle lanes from the level's queue
        const unsigned busy = __ballot_sync(full, my >= 0 || helper);
        if (!exhausted && __popc(busy) < CRT_L_REFILL) {
            const unsigned idle = ~busy;
            int base = 0;
            if (lane == 0)
                base = atomicAdd(&ctl->fetch[level], __popc(idle));
            base = __shfl_sync(full, base, 0);
            if (base + __popc(idle) >= total) {
                exhausted = true;
#ifdef CRT_DEBUG_TAIL
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_exhausted));
#endif
            }
            if (my < 0 && !helper) {
                const int i = base + __popc(idle & lt_mask);
                if (i < total) {
                    V3 o, d, inv;
                    float t_enter, t_exit;
                    if (level == 0) {
                        const float4 rd = b.ray0_d[i], ri = b.ray0_i[i];
                        node_id = b.ray0_slot[i];
                        o = mk(S.cam.position[0], S.cam.position[1], S.cam.position[2]);
                        d = xyz(rd), inv = xyz(ri), t_enter = rd.w, t_exit = ri.w;
                    } else {
                        const float4 ro = b.ray_o[start + i], rd = b.ray_d[start + i], ri = b.ray_i[start + i];
                        node_id = b.P + start + i;
                        o = xyz(ro), d = xyz(rd), inv = xyz(ri), t_enter = ro.w, t_exit = rd.w;
                        if (STATS) {
                            ++cn.rays;
                            ++cn.rays_by_depth[level];
                            ++cn.node_tests;
                        }
                    }
                    my = i;
                    src = lane;
                    stolen = 0;
                    ws.helpers[lane] = 0;
                    live = t_enter <= t_exit;
                    if (STATS && level > 0 && live)
                        ++cn.entered; // (level 0: counted by the ray generation)
                    w.node = -1;
                    if (live) {
                        w.start(ws, lane, o, d, inv, t_enter, t_exit);
                        ws.best[lane] = kNoHitKey;
                    }
                }
            }
            __syncwarp();
        }
        if (__ballot_sync(full, my >= 0 || helper) == 0u) {
            if (exhausted)
                break;
            continue;
        }
        // ---- the queue is empty and some lanes have nothing to do: they take over pending subtrees of the lanes that
        //      are still walking (the BOTTOM stack entry = the farthest, usually largest piece of the remaining path).
        //      A bounce level ends with its longest ray; a skimming ray crosses ~1000 cells one dependent load at a
        //      time, and this spreads that walk over the warp.  The victim's entry becomes a tombstone (its entry
        //      distance still bounds the entry above it); hits merge through the owner's ws.best as always.
        if (exhausted) {
            const unsigned thieves = __ballot_sync(full, my < 0 && !helper);
            stolen = min(stolen, w.sp); // entries below `stolen` are tombstones; popping past them frees the slots
            const uint32_t lowest = ws.stack[stolen & (kWin - 1)][lane].x;
            const bool can_give = (my >= 0 || helper) && w.node != -1 && w.lo == 0 && w.sp - stolen >= 2 && lowest < uint32_t(kStolenNode);
            const unsigned victims = __ballot_sync(full, can_give);
            if (thieves != 0u && victims != 0u) {
                const int pairs = min(__popc(thieves), __popc(victims));
                const bool thief = my < 0 && !helper;
                const int rank = thief ? __popc(thieves & lt_mask) : (can_give ? __popc(victims & lt_mask) : 32);
                const bool paired = rank < pairs;
                const unsigned v = (paired && thief) ? __fns(victims, 0, rank + 1) : lane; // the rank-th victim
                const unsigned v_src = __shfl_sync(full, src, v);
                const float v_root_exit = __shfl_sync(full, w.root_tmax, v);
                const int v_stolen = __shfl_sync(full, stolen, v);
                if (paired && thief) {
                    const uint2 e = ws.stack[v_stolen & (kWin - 1)][v];
                    // exit distance of that entry = entry distance of the one below it (tombstones keep theirs)
                    const float exit_t = v_stolen == 0 ? v_root_exit : __uint_as_float(ws.stack[(v_stolen - 1) & (kWin - 1)][v].y);
                    helper = true;
                    src = v_src;
                    stolen = 0;
                    w.sp = 0, w.lo = 0, w.flagged = 0;
                    w.node = int(e.x);
                    w.tmin = __uint_as_float(e.y);
                    w.tmax = exit_t, w.root_tmax = exit_t;
                    atomicAdd(&ws.helpers[v_src], 1);
                    if (STATS)
                        ++cn.node_tests;
                }
                __syncwarp();
                if (paired && !thief) {
                    ws.stack[stolen & (kWin - 1)][lane].x = uint32_t(kStolenNode);
                    ++stolen;
                }
                __syncwarp();
            }
        }
        // ---- walk: up to CRT_L_STEPS node visits per lane; a leaf is parked and the walk goes on (speculatively: the
        //      leaf's triangles have not been tested yet, so the distance used for pruning may be stale -- never wrong)
#ifdef CRT_DEBUG_TAIL
        if (my >= 0)
            ++rounds;
#endif
        bool blocked = false;
        if ((my >= 0 || helper) && w.node != -1) {
            const float best_t = __uint_as_float(unsigned(ws.best[src] >> 32));
#pragma unroll
            for (int k = 0; k < CRT_L_STEPS; ++k) {
                if (w.node == -1 || blocked)
                    break;
                bool need_pop = w.prune_at_visit(best_t);
                if (!need_pop) {
                    const int4 n = kd_node_at(S, w.node);
                    if (n.w == 3) {
                        if (p1_count == 0) {
                            if (p0_count == 0)
                                p0_first = n.y, p0_count = n.z;
                            else
                                p1_first = n.y, p1_count = n.z;
                            need_pop = true;
                        } else {
                            blocked = true; // both slots taken: wait for the test phase below
                        }
                    } else {
                        w.step<STATS>(ws, spill, lane, src, n, best_t, cn);
                        need_pop = w.node < 0;
                    }
                }
                if (need_pop)
                    w.pop_fast<STATS>(ws, spill, lane, cn);
            }
        }
        // ---- test phase: all parked leaves of all lanes at once
        const unsigned parked = __ballot_sync(full, p0_count > 0);
        const unsigned walkers = __ballot_sync(full, (my >= 0 || helper) && w.node != -1 && !blocked);
        const unsigned waiting = __ballot_sync(full, (my >= 0 || helper) && w.node == -1 && p0_count > 0);
        const bool fire = parked != 0u && (__popc(parked) >= CRT_L_TRIGGER || __any_sync(full, blocked) || walkers == 0u || __popc(waiting) >= CRT_L_WAITERS);
        if (fire) {
            const int mine = p0_count + p1_count;
            int incl = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(full, incl, o);
                if (int(lane) >= o)
                    incl += v;
            }
            const int items = __shfl_sync(full, incl, 31);
            const int excl = incl - mine;
            for (int base_item = 0; base_item < items; base_item += 32) {
                const int item = base_item + int(lane);
                // owner = last lane whose exclusive prefix is <= item; lanes with equal prefixes have nothing parked
                // except the last of the run, which owns the item
                int owner = 0;
#pragma unroll
                for (int stride = 16; stride > 0; stride >>= 1) {
                    const int probe = __shfl_sync(full, excl, owner + stride);
                    if (probe <= item)
                        owner += stride;
                }
                const int o_excl = __shfl_sync(full, excl, owner);
                const int o_first0 = __shfl_sync(full, p0_first, owner), o_count0 = __shfl_sync(full, p0_count, owner);
                const int o_first1 = __shfl_sync(full, p1_first, owner);
                owner = int(__shfl_sync(full, src, owner)); // a helper's leaves are tested against the ray it helps
                if (item < items) {
                    const int k = item - o_excl;
                    const uint32_t ref = __ldg(&S.kd_refs[k < o_count0 ? o_first0 + k : o_first1 + (k - o_count0)]);
                    const int tri = int(ref & 0x7FFFFFFFu);
                    const float4 A = __ldg(&S.tri_verts[3 * tri]), B = __ldg(&S.tri_verts[3 * tri + 1]), C = __ldg(&S.tri_verts[3 * tri + 2]);
                    const V3 o = mk(ws.ray[0][owner], ws.ray[1][owner], ws.ray[2][owner]), d = mk(ws.ray[3][owner], ws.ray[4][owner], ws.ray[5][owner]);
                    if (STATS)
                        ++cn.tri_tests;
                    float t;
                    if (triangle_test_flat(A, B, C, (ref >> 31) != 0, o, d, t)) {
                        t = t + 0.0f; // -0 -> +0 so the bit pattern orders like the value
                        atomicMin(&ws.best[owner], ((unsigned long long)__float_as_uint(t) << 32) | (unsigned long long)uint32_t(tri));
                    }
                }
            }
            __syncwarp();
            p0_count = 0, p1_count = 0;
        }
        // ---- retire: walk over and nothing parked; helpers first, then the rays nobody is helping any more
        if (helper && w.node == -1 && p0_count == 0) {
            atomicSub(&ws.helpers[src], 1);
            helper = false;
            src = lane;
        }
        __syncwarp();
        if (my >= 0 && w.node == -1 && p0_count == 0 && (!live || *(volatile int *)&ws.helpers[lane] == 0)) {
            Hit h{ kInf, -1 };
            if (live) {
                const unsigned long long key = ws.best[lane];
                h.t = __uint_as_float(unsigned(key >> 32));
                h.tri = int(unsigned(key & 0xffffffffull));
            }
            b.hit[node_id] = make_float2(h.t, __int_as_float(h.tri));
#ifdef CRT_DEBUG_TAIL
            atomicMax(&g_tail[level][0], (unsigned long long)rounds);
            atomicAdd(&g_tail[level][1], (unsigned long long)rounds);
            atomicAdd(&g_tail[level][2 + (rounds ? 32 - __clz(rounds) : 0)], 1ull);
            rounds = 0;
#endif
            my = -1;
        }
    }
#ifdef CRT_DEBUG_TAIL
    if (lane == 0) {
        unsigned long long t_end;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
        atomicMax(&g_tail[level][30], t_exhausted - t_start); // ns until this warp found the queue empty
        atomicMax(&g_tail[level][31], t_end - t_start);       // ns until this warp left
        atomicAdd(&g_tail[level][32], t_end - t_start);       // sum over warps
        atomicAdd(&g_tail[level][33], 1ull);
    }
#endif
    if (STATS)
        flush_counters(cn, counters);
}

} // namespace crtdev
