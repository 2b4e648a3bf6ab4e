// TRACE stage of the wavefront pipeline (included by crt_wavefront.cuh): closest hit of every ray of one bounce level.
//
// Persistent warps over the level's ray queue.  Each lane owns one ray at a time; when fewer than kRefillBelow lanes
// still have work the idle lanes take the next rays of the queue (one atomicAdd per warp, ballot + popc offsets).
//
// What shaped this kernel (ncu captures under profiles/, history in DESIGN.md):
//   * per-lane leaf loops with early-outs ran with 2.4 of 32 threads active per instruction and were 75 % of all issued
//     instructions -> triangle tests are a WARP-wide job list: every lane tests whichever (ray, triangle) pair comes
//     next, branch-free, and merges hits with a 64-bit shared-memory atomicMin on (t bits << 32 | triangle index),
//     which is exactly the reference's winner rule "closest, ties to the lowest parse index";
//   * the walk is a chain of dependent loads (node -> node -> ... -> leaf references -> triangles), a skimming ray
//     needs ~1000 of them, and a bounce level cannot end before its longest ray -> the walk never waits for the
//     tests: a lane that reaches a leaf hands the leaf's references to the job list and keeps walking with the best
//     distance known so far (SPECULATIVE: it may visit nodes a finished test would have pruned; extra candidates
//     cannot change the closest hit), and every round issues its three kinds of loads (next node of every walker,
//     leaf references of the leaves reached last round, triangles of the 32 oldest candidates) before consuming any.
//
// One round of a warp:
//   a. walkers load their next k-d node (16 B);            b. the 32 oldest candidates load their triangle (48 B);
//   c. leaves reached last round are gathered: shfl prefix sum of the leaf sizes, owner lane by binary search over the
//      prefix registers, reference load, per-ray mailbox filter (the builder references a triangle from every leaf it
//      overlaps, 38x in the 1 M-triangle scene), ballot compaction into the candidate ring in shared memory;
//   d. the triangle tests of (b) run and merge;            e. walkers take one step with the node of (a);
//   f. rays whose walk is over and whose candidates have all been tested retire (hit record written) and free the lane.
#pragma once

namespace crtdev {

constexpr int kTraceBlock = 128;
constexpr int kTraceWarps = kTraceBlock / 32;
#ifndef CRT_REFILL
#define CRT_REFILL 25
#endif
#ifndef CRT_STEPS
#define CRT_STEPS 2
#endif
#ifndef CRT_MBITS
#define CRT_MBITS 3
#endif
#ifndef CRT_WIN
#define CRT_WIN 8
#endif
#ifndef CRT_TEST_FIRST
#define CRT_TEST_FIRST 1
#endif
#ifndef CRT_MINBLOCKS
#define CRT_MINBLOCKS 7
#endif
constexpr int kRefillBelow = CRT_REFILL;
constexpr int kRing = 128;      // candidates: < 32 carried over + one round's gather, drained 32 at a time
constexpr int kWin = CRT_WIN;         // top-of-stack window kept in shared memory (older entries spill to local memory)
constexpr int kMailboxBits = CRT_MBITS;
constexpr int kTraceMailbox = 1 << kMailboxBits;
constexpr int kStepsPerRound = CRT_STEPS;

struct __align__(16) TraceWarpShared {
    float ray[9][32];                  // origin xyz, direction xyz, 1/direction xyz of each lane's ray
    unsigned long long best[32];       // (float bits of t) << 32 | triangle index; t >= 0 so the order is numeric
    int mailbox[kTraceMailbox][32];    // per-ray direct-mapped table of triangles already handed to the job list
    uint2 stack[kWin][32];             // {node | kFlagBit, entry distance bits}
    uint2 ring[kRing];                 // candidates: {leaf reference (triangle | cull bit), owner lane}
};

// The k-d walk of KdWalker (crt_device.cuh) with its state laid out for this kernel: the ray lives in shared memory
// (indexed by the split axis instead of select chains), the stack is {node, entry distance} with the EXIT distance
// implied -- the exit distance of a pending far child is the entry distance of the stack entry below it (or the root
// exit), because every push clips the current interval at the pushed plane.  Implied exits are never smaller than the
// true ones, so the walk visits a superset of KdWalker's nodes; results are unchanged.  The top kWin entries sit in
// shared memory ([entry][lane]: conflict-free), older ones are spilled to local memory and rarely read back.
constexpr uint32_t kFlagBit = 0x80000000u; // entry pushed for a ray lying IN the split plane: its interval starts at 0

struct TraceWalker {
    float tmin, tmax, root_tmax;
    int node; // node to visit next, or -1 when the walk is finished
    int sp, lo, flagged;
    // the spilled entries live in a local array of the KERNEL (`spill`, kStack entries), not in this struct: a
    // dynamically indexed member would force the whole walker -- node, sp, tmin, tmax -- into local memory

    // start a walk from a ray record (entry/exit of the root box and 1/direction come with the record)
    template <class WS>
    __device__ __forceinline__ void start(WS &ws, unsigned lane, V3 o, V3 d, V3 inv, float t_enter, float t_exit) {
        sp = 0, lo = 0, flagged = 0;
        tmin = t_enter, tmax = t_exit, root_tmax = t_exit;
        node = 0;
        ws.ray[0][lane] = o.x, ws.ray[1][lane] = o.y, ws.ray[2][lane] = o.z;
        ws.ray[3][lane] = d.x, ws.ray[4][lane] = d.y, ws.ray[5][lane] = d.z;
        ws.ray[6][lane] = inv.x, ws.ray[7][lane] = inv.y, ws.ray[8][lane] = inv.z;
    }

    __device__ __forceinline__ bool done() const { return node < 0; }

    // the store happens whether or not anything is pushed (slot sp is free), only the count moves: no branch per push
    template <class WS>
    __device__ __forceinline__ void push_if(WS &ws, uint2 *spill, unsigned lane, bool doit, uint32_t tagged_node, float entry) {
        const bool full = sp - lo == kWin;
        if (__builtin_expect(doit & full, 0)) { // window full (rare): the oldest entry goes to local memory
            spill[lo] = ws.stack[lo & (kWin - 1)][lane];
            ++lo;
        }
        if (doit | !full) // a full window's slot sp IS the oldest entry: leave it alone unless it was just spilled
            ws.stack[sp & (kWin - 1)][lane] = make_uint2(tagged_node, __float_as_uint(entry));
        sp += doit ? 1 : 0;
    }

    template <class WS>
    __device__ __forceinline__ uint2 entry_at(const WS &ws, const uint2 *spill, unsigned lane, int i) const {
        return i >= lo ? ws.stack[i & (kWin - 1)][lane] : spill[i];
    }

    // next pending subtree that can still hold a closer (or equal) hit; node = -1 when none is left
    template <bool STATS, class WS>
    __device__ __forceinline__ void pop(WS &ws, uint2 *spill, unsigned lane, float best_t, Counters &cn) {
        for (;;) {
            if (sp == 0) {
                node = -1;
                return;
            }
            --sp;
            uint2 e = ws.stack[sp & (kWin - 1)][lane];
            float below = __uint_as_float(ws.stack[(sp - 1) & (kWin - 1)][lane].y);
            if (sp <= lo) { // rare: the entry or the one below it was spilled
                if (sp < lo)
                    e = spill[sp], lo = sp;
                if (sp > 0)
                    below = __uint_as_float(entry_at(ws, spill, lane, sp - 1).y);
            }
            tmax = sp == 0 ? root_tmax : below;
            node = int(e.x & ~kFlagBit);
            if (e.x & kFlagBit) {
                --flagged;
                tmin = 0.0f;
                break;
            }
            tmin = __uint_as_float(e.y);
            // strict pruning with slack: entries that start clearly beyond the best hit cannot beat or tie it
            if (!(tmin - (fabsf(tmin) * 1e-5f + 1e-6f) > best_t))
                break;
            // entry distances grow towards the bottom of the stack (each push is at most one slack above the entry
            // below it), so an entry pruned with this much margin means everything below is pruned as well
            if (flagged == 0 && tmin - (fabsf(tmin) * 1e-3f + 1e-4f) > best_t) {
                sp = 0, lo = 0;
                node = -1;
                return;
            }
        }
        if (STATS)
            ++cn.node_tests;
    }

    // ONE node of the walk: true when `n` is a leaf.  Same decisions as KdWalker::step, written as selects so that
    // the lanes of a warp, which sit at unrelated nodes, share one instruction stream; only the push and the pop stay
    // conditional.
    template <bool STATS, class WS>
    __device__ __forceinline__ bool step(WS &ws, uint2 *spill, unsigned lane, const int4 n, float best_t, Counters &cn) {
        return step<STATS>(ws, spill, lane, lane, n, best_t, cn);
    }

    // `src` = the lane whose ray this walk belongs to (a helper lane walks a subtree of another lane's ray)
    template <bool STATS, class WS>
    __device__ __forceinline__ bool step(WS &ws, uint2 *spill, unsigned lane, unsigned src, const int4 n, float best_t, Counters &cn) {
        if (n.w == 3)
            return true;
        const int axis = n.w;
        const float split = __int_as_float(n.x);
        // an origin exactly ON the plane takes the side the ray leaves towards; the sign of 1/d is the sign of d (d = +-0
        // with the origin on the plane is the in-plane case below, where both sides are visited whatever `below` says)
        const float oa = ws.ray[axis][src], ia = ws.ray[6 + axis][src];
        const bool below = (oa < split) | ((oa == split) & (ia <= 0.0f));
        const int near = below ? n.y : n.z, far = below ? n.z : n.y;
        const float tp = (split - oa) * ia;
        const float slack = fabsf(tp) * 1e-5f + 1e-6f;
        const bool in_plane = tp != tp; // origin on the plane and direction parallel to it: the ray lies in both boxes
        const bool near_only = !in_plane & ((tp <= 0.0f) | (tp > tmax + slack));
        const bool far_only = !in_plane & !near_only & (tp < tmin - slack);
        const bool both = !near_only & !far_only;
        const bool has_near = near >= 0, has_far = far >= 0;
        const bool do_push = both & has_near & has_far;
        const bool go_far = far_only | (both & !has_near);
        const int next = go_far ? far : near;
        push_if(ws, spill, lane, do_push, uint32_t(far) | (in_plane ? kFlagBit : 0u), in_plane ? tmax : tp);
        flagged += (do_push & in_plane) ? 1 : 0;
        tmax = (do_push & !in_plane) ? fminf(tmax, tp) : tmax;
        tmin = (both & !has_near & has_far & !in_plane) ? fmaxf(tmin, tp) : tmin;
        node = next; // -1: the child is absent, the caller pops
        if (STATS && next >= 0)
            ++cn.node_tests;
        return false;
    }
};

#ifdef CRT_DEBUG_TAIL
// debugging aid (never in the product build): per level, the longest walk (in rounds of the trace loop), the sum of
// all rounds, and a log2 histogram of rounds per ray
__device__ unsigned long long g_tail[16][40];
#endif

constexpr unsigned long long kNoHitKey = (0x7f800000ull << 32) | 0xffffffffull; // t = +inf, tri = -1

struct Candidate {
    float4 A, B, C;
    uint32_t ref;
    int owner;
};

// (b) the lane's candidate of the batch at the ring's head: triangle loads are issued here, consumed in test_candidate
__device__ __forceinline__ Candidate load_candidate(const DScene &S, const TraceWarpShared &ws, int head, int n, unsigned lane) {
    Candidate c{};
    c.owner = -1;
    if (int(lane) < n) {
        const uint2 e = ws.ring[(head + int(lane)) & (kRing - 1)];
        const int tri = int(e.x & 0x7FFFFFFFu);
        c.ref = e.x;
        c.owner = int(e.y);
        c.A = __ldg(&S.tri_verts[3 * tri]);
        c.B = __ldg(&S.tri_verts[3 * tri + 1]);
        c.C = __ldg(&S.tri_verts[3 * tri + 2]);
    }
    return c;
}

// (d) branch-free test and merge into the owner's best
template <bool STATS>
__device__ __forceinline__ void test_candidate(TraceWarpShared &ws, const Candidate &c, Counters &cn) {
    if (c.owner >= 0) {
        const int owner = c.owner;
        const V3 o = mk(ws.ray[0][owner], ws.ray[1][owner], ws.ray[2][owner]), d = mk(ws.ray[3][owner], ws.ray[4][owner], ws.ray[5][owner]);
        if (STATS)
            ++cn.tri_tests;
        float t;
        if (triangle_test_flat(c.A, c.B, c.C, (c.ref >> 31) != 0, o, d, t)) {
            t = t + 0.0f; // -0 -> +0 so the bit pattern orders like the value
            atomicMin(&ws.best[owner], ((unsigned long long)__float_as_uint(t) << 32) | (unsigned long long)(c.ref & 0x7FFFFFFFu));
        }
    }
}

// verification traversal (reference order, crt_intersection.cpp:109-136): plain lane-per-ray loop, no speculation
template <bool STATS>
__device__ __forceinline__ void trace_level_literal(const DScene &S, const DParams &P, const WfBuffers &b, int level, int total, int start, Counters &cn) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        Ray r;
        if (level == 0) {
            int x, ly;
            if (!slot_pixel(b, S, P, i, x, ly))
                continue;
            r = generate_ray(S.cam, x, global_row(P, ly));
        } else {
            r = Ray{ xyz(b.ray_o[start + i]), xyz(b.ray_d[start + i]) };
        }
        if (STATS) {
            ++cn.rays;
            ++cn.rays_by_depth[level];
        }
        const Hit h = trace_literal<STATS>(S, r, &cn); // counts the ray as entered
        b.hit[level == 0 ? i : b.P + start + i] = make_float2(h.t, __int_as_float(h.tri));
    }
}

template <int TRAV, bool STATS>
__global__ void __launch_bounds__(kTraceBlock, CRT_MINBLOCKS) k_wf_trace(const __grid_constant__ DScene S, const __grid_constant__ DParams P,
                                                          const __grid_constant__ WfBuffers b, int level, Counters *__restrict__ counters) {
    WfCtl *ctl = b.ctl;
    const int start = level == 0 ? 0 : level_start(ctl, level);
    const int total = level == 0 ? (TRAV == CRT_TRAVERSAL_KD ? ctl->count[0] : b.P) : level_count(ctl, b.cap, level, start);
    Counters cn{};
    if (TRAV != CRT_TRAVERSAL_KD) {
        trace_level_literal<STATS>(S, P, b, level, total, start, cn);
        if (STATS)
            flush_counters(cn, counters);
        return;
    }
    __shared__ TraceWarpShared shared[kTraceWarps];
    TraceWarpShared &ws = shared[threadIdx.x >> 5];
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    const unsigned full = 0xffffffffu;
    TraceWalker w;
    uint2 spill[kStack];
    w.node = -1, w.sp = 0, w.lo = 0, w.flagged = 0;
    int my = -1;                        // index of my ray within the level, -1 = idle
    int node_id = 0;                    // where my ray's hit record goes
    bool live = false;                  // my ray entered the root box: its best lives in ws.best[lane]
    int leaf_first = 0, leaf_count = 0; // leaf reached in the previous round, gathered in this one
    // warp-uniform candidate bookkeeping: ring[head .. head+fill) are untested; `queued` / `tested` count candidates
    // ever queued / tested; a ray is settled once tested >= the value of `queued` right after its last gather
    int head = 0, fill = 0;
    unsigned queued = 0, tested = 0, my_mark = 0;
    bool exhausted = false;
#ifdef CRT_DEBUG_TAIL
    unsigned rounds = 0;
#endif
    for (;;) {
        // ---- refill idle lanes
        const unsigned busy = __ballot_sync(full, my >= 0);
        if (!exhausted && __popc(busy) < kRefillBelow) {
            const unsigned idle = ~busy;
            int base = 0;
            if (lane == 0)
                base = atomicAdd(&ctl->fetch[level], __popc(idle));
            base = __shfl_sync(full, base, 0);
            if (base + __popc(idle) >= total)
                exhausted = true;
            if (my < 0) {
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
                    my_mark = tested;
                    live = t_enter <= t_exit;
                    if (STATS && level > 0 && live)
                        ++cn.entered; // (level 0: counted by the ray generation)
                    w.node = -1;
                    if (live) {
                        w.start(ws, lane, o, d, inv, t_enter, t_exit);
                        ws.best[lane] = kNoHitKey;
#pragma unroll
                        for (int k = 0; k < kTraceMailbox; ++k)
                            ws.mailbox[k][lane] = -1;
                    }
                }
            }
            __syncwarp();
        }
        const unsigned active = __ballot_sync(full, my >= 0);
        if (active == 0u) {
            if (exhausted)
                break;
            continue;
        }
#ifdef CRT_DEBUG_TAIL
        if (my >= 0)
            ++rounds;
#endif
        // ---- a. walkers load their next node
        const bool walking = my >= 0 && !w.done();
        int4 n = make_int4(0, 0, 0, 0);
        if (walking)
            n = kd_node_at(S, w.node);
        const unsigned walkers = __ballot_sync(full, walking);
        const unsigned leaves = __ballot_sync(full, leaf_count > 0);
        // ---- b. the oldest candidates load their triangles: a full batch, or whatever is queued once too few lanes
        //         are still walking (the others wait for their candidates to be tested)
        int batch = 0;
        if (fill >= 32 || (fill > 0 && leaves == 0u && __popc(walkers) < kRefillBelow))
            batch = min(fill, 32);
        const Candidate cand = load_candidate(S, ws, head, batch, lane);
        head = (head + batch) & (kRing - 1);
        fill -= batch;
#if CRT_TEST_FIRST
        test_candidate<STATS>(ws, cand, cn);
        tested += batch;
        __syncwarp();
#endif
        // ---- c. gather the leaves reached last round into the ring
        if (leaves != 0u) {
            int incl = leaf_count;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(full, incl, o);
                if (int(lane) >= o)
                    incl += v;
            }
            const int items = __shfl_sync(full, incl, 31);
            const int excl = incl - leaf_count;
            for (int base_item = 0; base_item < items; base_item += 32) {
                const int item = base_item + int(lane);
                // owner = last lane whose exclusive prefix is <= item (binary search over the lanes' registers); lanes
                // with equal prefixes have empty leaves except the last of the run, which owns the item
                int owner = 0;
#pragma unroll
                for (int stride = 16; stride > 0; stride >>= 1) {
                    const int probe = __shfl_sync(full, excl, owner + stride);
                    if (probe <= item)
                        owner += stride;
                }
                const int owner_first = __shfl_sync(full, leaf_first, owner);
                const int owner_excl = __shfl_sync(full, excl, owner);
                bool keep = false;
                uint32_t ref = 0;
                if (item < items) {
                    ref = __ldg(&S.kd_refs[owner_first + (item - owner_excl)]);
                    const int tri = int(ref & 0x7FFFFFFFu);
                    int *slot = &ws.mailbox[(uint32_t(tri) * 0x9E3779B1u) >> (32 - kMailboxBits)][owner];
                    keep = *slot != tri;
                    if (keep)
                        *slot = tri;
                }
                if (fill > kRing - 32) { // make room: test one batch right here (rare: more than 96 candidates pending)
                    __syncwarp();
                    const Candidate extra = load_candidate(S, ws, head, 32, lane);
                    test_candidate<STATS>(ws, extra, cn);
                    head = (head + 32) & (kRing - 1);
                    fill -= 32;
                    tested += 32;
                    __syncwarp();
                }
                const unsigned kmask = __ballot_sync(full, keep);
                if (keep)
                    ws.ring[(head + fill + __popc(kmask & lt_mask)) & (kRing - 1)] = make_uint2(ref, unsigned(owner));
                fill += __popc(kmask);
                queued += __popc(kmask);
            }
            if (leaf_count > 0)
                my_mark = queued;
            leaf_count = 0;
        }
        // ---- d. test the batch loaded in (b)
#if !CRT_TEST_FIRST
        test_candidate<STATS>(ws, cand, cn);
        tested += batch;
        __syncwarp();
#endif
        // ---- e. steps of the walk, the first with the node loaded in (a); a leaf is handed over to the next round's
        //         gather and the walk goes on at once (speculatively: its tests have not run yet)
        if (walking) {
            const float best_t = live ? __uint_as_float(unsigned(ws.best[lane] >> 32)) : kInf;
#pragma unroll
            for (int k = 0; k < kStepsPerRound; ++k) {
                if (k > 0) {
                    if (w.done() || leaf_count > 0)
                        break;
                    n = kd_node_at(S, w.node);
                }
                if (w.step<STATS>(ws, spill, lane, n, best_t, cn))
                    leaf_first = n.y, leaf_count = n.z, w.node = -1;
                if (w.node < 0) // leaf handed over, or an absent child: one pop site for both
                    w.pop<STATS>(ws, spill, lane, best_t, cn);
            }
        }
        // ---- f. retire: walk over, nothing left to gather, every candidate tested
        if (my >= 0 && w.done() && leaf_count == 0 && (!live || int(tested - my_mark) >= 0)) {
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
    if (STATS)
        flush_counters(cn, counters);
}

} // namespace crtdev
