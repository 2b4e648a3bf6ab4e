// TRACE stage for walk trees with LIGHT leaves (the SAH tree of core/crt_kdbuild.cpp: ~2 triangles per leaf, hardly any
// triangle referenced twice): closest hit of every ray of one bounce level.  Included by crt_wavefront.cuh after
// crt_trace.cuh (whose k_wf_trace serves walk trees with heavy leaves, $CRT_WALK_TREE=reference).
// Replaces ray_intersect_acceleration_tree + ray_intersect_triangle (src/core/crt_intersection.cpp:109-136, :47-93).
//
// With such a tree a ray takes ~25-65 node visits but only 1-4 leaf visits and 1-6 triangle tests.  Organisation:
//   * persistent warps; a lane owns one ray at a time, idle lanes take the next rays of the level's queue (one atomicAdd
//     per warp) when fewer than CRT_N_REFILL lanes are busy;
//   * the walk is crt_walk_lean.h, one straight-line iteration per node visit for all walking lanes (round 2: 58 SASS
//     instructions per visit instead of 138); before every visit the lane asks whether
//     the subtree it stands in starts beyond the best hit it knows (plus a margin): then its walk is over;
//   * a lane that reaches a leaf PARKS it (its node id, two register slots) and walks on -- speculatively: the leaf's
//     triangles have not been tested yet, so the distance used for pruning may be stale, never wrong;
//   * when enough leaves are parked (or a lane has both slots full, or nobody can walk) the warp runs ONE test phase:
//     shfl prefix sum of the parked triangle counts, owner lane by binary search, every lane tests whichever
//     (ray, triangle) pair comes next, branch-free; hits merge through a 64-bit shared atomicMin on (t bits, triangle) =
//     the reference's winner rule (closest, ties to the lowest parse index).  A triangle referenced from two leaves is
//     simply tested twice;
//   * when the level's queue is empty, idle lanes take over the bottom stack entries (the farthest pending subtrees) of
//     lanes that still walk; the victim finds the scene's EMPTY LEAF in that entry's place (a real node the upload
//     appends to the walk tree), so a visit never asks whether its node was stolen.  From then on tests run as soon as a
//     lane waits for them (step 22): what is left is a chain of dependent work per lane;
//   * node visits are counted where the node is loaded (STATS builds).
#pragma once

#include "crt_walk_lean.h"

namespace crtdev {

#ifndef CRT_N_STEPS
#define CRT_N_STEPS 10 // (r2y, with the 62-instruction visit of step 27: 8 -> 10 steps per round C4 2.36 -> 2.33 ms; 6: 2.43, 12: 2.35)
#endif
#ifndef CRT_N_TRIGGER
#define CRT_N_TRIGGER 16
#endif
#ifndef CRT_N_WAITERS
#define CRT_N_WAITERS 6
#endif
#ifndef CRT_N_MINBLOCKS
#define CRT_N_MINBLOCKS 9 // 56 registers, nothing spilled (r2b: 10 blocks = 48 registers spill ~100 B around the refill and cost 2 %)
#endif
#ifndef CRT_N_REFILL
#define CRT_N_REFILL 25
#endif
#ifndef CRT_PARK_SHIFT
#define CRT_PARK_SHIFT 1 // (two selects under one predicate instead of two predicates: 62 -> 60 instructions per visit)
#endif
#ifndef CRT_PREFETCH_PUSH
#define CRT_PREFETCH_PUSH 0
#endif

// per-warp state in shared memory ([entry][lane] layouts: conflict-free)
struct __align__(16) LightWarpShared {
    float ray[9][32];            // origin xyz, direction xyz, 1/direction xyz of each lane's ray
    unsigned long long best[32]; // (float bits of t) << 32 | triangle index
    uint2 stack[kWin][32];       // {node, exit distance bits}
    int helpers[32];             // lanes currently walking a stolen subtree of this lane's ray
};

static_assert(crtwalk::kWin == kWin, "stack window of the lean walker and of LightWarpShared");

__device__ __forceinline__ crtwalk::Node as_walk_node(const int4 n) { return crtwalk::Node{ n.x, n.y, n.z, n.w }; }

// ---- pieces shared with k_render_persistent -----------------------------------------------------------------------

// Idle lanes take over the bottom stack entries (the farthest pending subtrees) of lanes that still walk.  A level ends
// with its longest ray; a skimming ray crosses ~1000 cells one dependent load at a time, and this spreads that walk over
// the warp.  Thieves and victims are paired by rank among their ballots; the thief reads the victim's entry {node, exit}
// (its interval starts where the entry above it ends) and walks it against the victim's ray (`src`); the victim's entry
// becomes a tombstone -- the empty leaf -- that keeps its exit distance for the entries above it; hits merge through the
// owner's ws.best as always, and the owner retires only when ws.helpers says nobody walks for it any more.
// owns_work = this lane owns a ray (or pixel); results come back through the references.
template <bool STATS>
__device__ __forceinline__ void lean_steal(LightWarpShared &ws, crtwalk::LeanWalker &w, unsigned lane, unsigned lt_mask, bool owns_work, bool &helper,
                                           unsigned &src, int &stolen, uint32_t empty_leaf, Counters &cn) {
    const unsigned full = 0xffffffffu;
    const bool thief = !owns_work && !helper;
    const unsigned thieves = __ballot_sync(full, thief);
    stolen = max(1, min(stolen, w.sp)); // entries in [1, stolen) are tombstones; popping past them frees the slots
    const uint32_t lowest = ws.stack[stolen & (kWin - 1)][lane].x;
    const bool can_give = (owns_work || helper) && !w.done() && w.node >= 0 && w.lo == 0 && w.sp - stolen >= 2 && w.flagged == 0 &&
                          lowest != empty_leaf;
    const unsigned victims = __ballot_sync(full, can_give);
    if (thieves != 0u && victims != 0u) {
        const int pairs = min(__popc(thieves), __popc(victims));
        const int rank = thief ? __popc(thieves & lt_mask) : (can_give ? __popc(victims & lt_mask) : 32);
        const bool paired = rank < pairs;
        const unsigned v = (paired && thief) ? __fns(victims, 0, rank + 1) : lane; // the rank-th victim
        const unsigned v_src = __shfl_sync(full, src, v);
        const int v_stolen = __shfl_sync(full, stolen, v);
        if (paired && thief) {
            // the victim's lowest live entry {node, exit}; its interval starts where the entry above it ends
            const uint2 e = ws.stack[v_stolen & (kWin - 1)][v];
            const float enter_t = __uint_as_float(ws.stack[(v_stolen + 1) & (kWin - 1)][v].y);
            helper = true;
            src = v_src;
            stolen = 1;
            w.begin(ws, lane, int(e.x), enter_t, __uint_as_float(e.y));
            atomicAdd(&ws.helpers[v_src], 1);
        }
        __syncwarp();
        if (paired && !thief) {
            ws.stack[stolen & (kWin - 1)][lane].x = empty_leaf;
            ++stolen;
        }
        __syncwarp();
    }
}

// up to STEPS node visits of this lane's walk; a reached leaf is PARKED (its node id kept in one of two registers) and
// the walk goes on -- speculatively: its triangles have not been tested yet, so the distance used for pruning may be
// stale, never wrong.  Returns true when the lane is blocked on a leaf with both slots taken.
template <bool STATS, int STEPS>
__device__ __forceinline__ bool lean_walk(const DScene &S, LightWarpShared &ws, crtwalk::LeanWalker &w, uint2 *spill, unsigned lane, unsigned src,
                                          int &parked0, int &parked1, Counters &cn) {
    // where this round's visits end the walk: beyond the best hit KNOWN now (parked leaves are not tested yet)
    const float cut = crtwalk::walk_cut(__uint_as_float(unsigned(ws.best[src] >> 32)), w.flagged);
    bool stopped = false;
#pragma unroll
    for (int k = 0; k < STEPS; ++k) {
        // (beyond the cut the walk is over: the lane leaves the round here and is marked after it -- a lane that leaves
        //  because it is stopped stands where its last check found it within the cut)
        if ((w.tmin > cut) | w.done() | stopped)
            break;
        const int at = w.node;
        const int4 n = kd_node_raw(S, at);
        if (STATS)
            ++cn.node_tests;
        bool took_leaf;
#if CRT_PREFETCH_PUSH
        int pushed;
        stopped = w.visit(ws, lane, src, as_walk_node(n), parked1 < 0, took_leaf, &pushed);
        if (pushed >= 0) { // the walk will come back to this node: have its record on the way by then (1: to L2; 2: to L1 -- r3b: neither pays, 2.36-2.41 against 2.32-2.37 ms)
#if CRT_PREFETCH_PUSH >= 2
            asm volatile("prefetch.global.L1 [%0];" ::"l"(&S.kd_nodes[pushed]));
#else
            asm volatile("prefetch.global.L2 [%0];" ::"l"(&S.kd_nodes[pushed]));
#endif
        }
#else
        stopped = w.visit(ws, lane, src, as_walk_node(n), parked1 < 0, took_leaf);
#endif
        // (a leaf without triangles -- the tombstone of a stolen subtree -- is parked like any other: the test phase
        //  finds nothing in it)
#if CRT_PARK_SHIFT
        parked1 = took_leaf ? parked0 : parked1; // (the newest leaf in slot 0, the one before it moves to slot 1)
        parked0 = took_leaf ? at : parked0;
#else
        const bool to0 = took_leaf & (parked0 < 0), to1 = took_leaf & (parked0 >= 0);
        parked0 = to0 ? at : parked0;
        parked1 = to1 ? at : parked1;
#endif
    }
    w.end_if_beyond(cut);
    if (!stopped)
        return false;
    w.make_room(ws, spill, lane); // (rare: stack window full / popped entry spilled; nothing to do for a blocked leaf)
    return !w.done() && parked1 >= 0; // blocked: both slots taken -- the test phase must run before this lane walks on
}

// test phase: every parked leaf of every lane at once.  shfl prefix sum of the parked counts, owner lane by binary
// search, every lane tests whichever (ray, triangle) pair comes next; hits merge through a 64-bit atomicMin on
// (t bits, triangle) = the reference's winner rule (closest, ties to the lowest parse index).
template <bool STATS>
__device__ __forceinline__ void lean_test_parked(const DScene &S, LightWarpShared &ws, unsigned lane, unsigned src, int parked0, int parked1, Counters &cn) {
    const unsigned full = 0xffffffffu;
    // the parked leaves' records again (they were loaded a few visits ago: L1)
    int p0_first = 0, p0_count = 0, p1_first = 0, p1_count = 0;
    if (parked0 >= 0) {
        const int4 n = kd_node_at(S, parked0);
        p0_first = n.y, p0_count = n.z;
    }
    if (parked1 >= 0) {
        const int4 n = kd_node_at(S, parked1);
        p1_first = n.y, p1_count = n.z;
    }
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
        // owner = last lane whose exclusive prefix is <= item; lanes with equal prefixes have nothing parked except the
        // last of the run, which owns the item
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
}

// -------------------------------------------------------------------------------------------------------------------

template <bool STATS>
__global__ void __launch_bounds__(kTraceBlock, CRT_N_MINBLOCKS) k_wf_trace_lean(const __grid_constant__ DScene S, const __grid_constant__ DParams P,
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
    const uint32_t empty_leaf = uint32_t(S.kd_empty_leaf);
    crtwalk::LeanWalker w;
    uint2 spill[kStack]; // stack entries below the shared-memory window
    w.node = crtwalk::kDone, w.sp = 1, w.lo = 0, w.flagged = 0, w.tmin = 0.0f, w.tmax = 0.0f;
    int my = -1;      // index of my ray within the level, -1 = idle
    int node_id = 0;  // where my ray's hit record goes
    bool live = false;
    int parked0 = -1, parked1 = -1; // parked leaves (node ids)
    bool exhausted = false;
    unsigned src = lane;  // lane whose ray I am walking: myself, or the owner of the subtree I took over
    bool helper = false;  // I own no ray, I walk a stolen subtree of lane `src`'s ray
    int stolen = 1;       // entries [1, stolen) of my stack are tombstones (helpers took them over)
    for (;;) {
        // ---- refill idle lanes from the level's queue
        const unsigned busy = __ballot_sync(full, my >= 0 || helper);
        if (!exhausted && __popc(busy) < CRT_N_REFILL) {
            const unsigned idle = ~busy;
            int base = 0;
            if (lane == 0)
                base = atomicAdd(&ctl->fetch[level], __popc(idle));
            base = __shfl_sync(full, base, 0);
            if (base + __popc(idle) >= total)
                exhausted = true;
            if (my < 0 && !helper) {
                const int i = base + __popc(idle & lt_mask);
                if (i < total) {
                    V3 o, d, inv;
                    float t_enter, t_exit;
                    if (level == 0) {
                        const float4 rd = get_stream(&b.ray0_d[i]), ri = get_stream(&b.ray0_i[i]);
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
                    stolen = 1;
                    ws.helpers[lane] = 0;
                    live = t_enter <= t_exit;
                    if (STATS && level > 0 && live)
                        ++cn.entered; // (level 0: counted by the ray generation)
                    w.node = crtwalk::kDone;
                    if (live) {
                        w.start(ws, lane, o.x, o.y, o.z, d.x, d.y, d.z, inv.x, inv.y, inv.z, t_enter, t_exit);
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
        if (exhausted)
            lean_steal<STATS>(ws, w, lane, lt_mask, my >= 0, helper, src, stolen, empty_leaf, cn);
        // ---- walk
        bool blocked = false;
        if ((my >= 0 || helper) && !w.done())
            blocked = lean_walk<STATS, CRT_N_STEPS>(S, ws, w, spill, lane, src, parked0, parked1, cn);
        // ---- test phase: all parked leaves of all lanes at once
        const unsigned parked = __ballot_sync(full, parked0 >= 0);
        const unsigned walkers = __ballot_sync(full, (my >= 0 || helper) && !w.done() && !blocked);
        const unsigned waiting = __ballot_sync(full, (my >= 0 || helper) && w.done() && parked0 >= 0);
        // batching the tests buys SIMD width while rays keep coming; once the queue is empty what is left is a chain of
        // dependent work per lane (r2d: one rank's share of an 8-rank frame is all tail), and a lane that only waits
        // for its tests is tested at once
        const bool fire = parked != 0u && (__popc(parked) >= CRT_N_TRIGGER || __any_sync(full, blocked) || walkers == 0u ||
                                           __popc(waiting) >= (exhausted ? 1 : CRT_N_WAITERS));
        if (fire) {
            lean_test_parked<STATS>(S, ws, lane, src, parked0, parked1, cn);
            parked0 = -1, parked1 = -1;
        }
        // ---- retire: walk over and nothing parked; helpers first, then the rays nobody is helping any more
        if (helper && w.done() && parked0 < 0) {
            atomicSub(&ws.helpers[src], 1);
            helper = false;
            src = lane;
        }
        __syncwarp();
        if (my >= 0 && w.done() && parked0 < 0 && (!live || *(volatile int *)&ws.helpers[lane] == 0)) {
            Hit h{ kInf, -1 };
            if (live) {
                const unsigned long long key = ws.best[lane];
                h.t = __uint_as_float(unsigned(key >> 32));
                h.tri = int(unsigned(key & 0xffffffffull));
            }
            put_stream(&b.hit[node_id], make_float2(h.t, __int_as_float(h.tri)));
            my = -1;
        }
    }
    if (STATS)
        flush_counters(cn, counters);
}

} // namespace crtdev
