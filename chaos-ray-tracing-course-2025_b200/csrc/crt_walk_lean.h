// The k-d walk of the render path, one node visit at a time (host + device: the device kernels in crt_trace_lean.cuh
// and crt_persistent.cuh drive it warp-wide, tests/host/walk_lean_host.cpp drives the very same code one ray at a time
// on the CPU and compares every closest hit with brute force).
//
// Replaces: the box tests of ray_intersect_acceleration_tree (src/core/crt_intersection.cpp:109-136) and
// ray_intersect_aabb_p (:14-45) for the walk tree of core/crt_kdbuild.cpp.  A ray only needs its parametric interval
// [tmin, tmax] inside the current cell; children are taken in the order the ray's LINE passes them.
//
// Why a second walker (ncu r1zz): TraceWalker::step + prune_at_visit + pop_fast cost 138 SASS
// instructions per node visit on kernels that sit at 77-79 % issue-slot utilisation -- the walk kernels are bound by
// instructions per visit, not by bytes.  This one makes the same decisions (a superset in degenerate cases, see below)
// with about a third of the instructions:
//   * children ordered by the SIGN of the direction component (first = the child the line passes first) instead of by
//     the side the origin lies on: no `below` predicate, no `tp <= 0` case.  For tmin >= 0 this visits exactly the
//     children TraceWalker visits, plus the far side when the origin lies on the plane itself;
//   * no "absent child x near/far x both" case table: visit-first / visit-second are two comparisons, an absent child
//     simply masks its side;
//   * stack entries hold {node, EXIT distance}; the entry distance of a popped subtree is the tmax the walk has at that
//     moment, so a pop is ONE load (before: the entry and the one below it), and a SENTINEL entry {kDone, root exit} at
//     the bottom means pop never asks whether the stack is empty;
//   * a ray lying IN a split plane (tp = 0 * inf = NaN) takes both children with the whole interval through the NaN
//     semantics of the comparisons and of fminf / fmaxf; only its stack entry needs the flag bit (rare path);
//   * pruning against the best hit is one comparison per visit against best + slack, computed once per round.
// Distances carry the same slack as before (|tp| * 1e-5 + 1e-6): grazing cases visit a SUPERSET of the cells the
// reference's six-plane box test accepts; extra leaves cannot change min (t, parse index).
#pragma once

#include <cstdint>
#include <cstring>
#include <cmath>

#if defined(__CUDACC__)
#define CRT_WALK_FN __host__ __device__ __forceinline__
#else
#define CRT_WALK_FN inline
#endif

namespace crtwalk {

constexpr uint32_t kFlagBit = 0x80000000u; // stack entry pushed for a ray lying IN the split plane: its interval starts at 0
constexpr int kDone = 0x7FFFFFFE;          // node id of the sentinel entry: the walk is over
#ifndef CRT_WIN
#define CRT_WIN 8
#endif
constexpr int kWin = CRT_WIN;              // top-of-stack window per lane in shared memory; older entries spill

// INLINE LEAVES (round 2, step 24; an experiment, OFF: measured on B200 it costs more than it saves): a
// child pointer with kLeafTag set is not a node index but the leaf itself -- {count:4 | first reference:26} -- so a
// walk never fetches a leaf's record: the parent's visit hands the leaf over and pops in the same iteration (9 % fewer
// iterations, 10-12 % fewer records fetched, 13 more instructions per visit).  The upload rewrites the pointers
// (inline_leaf_children below); leaves that do not fit (more than 15 triangles, a reference index beyond 2^26) and a
// root that is a leaf keep their records and are recognised by axis == 3 as before.  Absent children stay -1.
// With CRT_LEAF_INLINE=0 every term below that mentions it folds away.
#ifndef CRT_LEAF_INLINE
#define CRT_LEAF_INLINE 0
#endif
constexpr bool kInlineLeaves = CRT_LEAF_INLINE != 0;
constexpr uint32_t kLeafTag = 0x40000000u;
constexpr int kLeafCountShift = 26, kLeafCountMax = 15;
constexpr uint32_t kLeafFirstMask = (1u << kLeafCountShift) - 1u;

struct Node {   // crt_kd_node as the kernels load it: {split bits, a, b, axis} / {0, first, count, 3}
    int32_t x, y, z, w;
};
struct Entry {  // {node | kFlagBit, exit distance bits}
    uint32_t x, y;
};

CRT_WALK_FN float bits_to_float(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f;
    std::memcpy(&f, &u, 4);
    return f;
#endif
}
CRT_WALK_FN uint32_t float_to_bits(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u;
    std::memcpy(&u, &f, 4);
    return u;
#endif
}

// id >= 0 (a node to visit or a parked leaf): is it an inline leaf?  (-1 = absent has the bit as well: ask only about ids)
CRT_WALK_FN bool is_inline_leaf(int id) { return kInlineLeaves && (uint32_t(id) & kLeafTag) != 0u; }
CRT_WALK_FN int inline_leaf_first(int id) { return int(uint32_t(id) & kLeafFirstMask); }
CRT_WALK_FN int inline_leaf_count(int id) { return int((uint32_t(id) >> kLeafCountShift) & uint32_t(kLeafCountMax)); }
CRT_WALK_FN int make_inline_leaf(int first, int count) { return int(kLeafTag | (uint32_t(count) << kLeafCountShift) | uint32_t(first)); }

// pruning threshold for a best hit distance: a subtree whose interval starts beyond it cannot beat or tie the hit
CRT_WALK_FN float prune_threshold(float best_t) { return fmaf(fabsf(best_t), 1e-5f, best_t) + 1e-6f; }

// WS: per-warp storage with  float ray[9][32]  (origin, direction, 1/direction)  and  Entry-like stack[kWin][32]
// SPILL: Entry-like array of kStack entries owned by the caller (local memory on the device)
//
// Stack entry = {node | kFlagBit, EXIT distance of the pushed subtree}.  Its ENTRY distance need not be stored: when an
// entry is popped, everything pushed after it has been popped too, so the current tmax is again what the push left
// behind -- the plane distance the entry's interval starts at (or less, within the slack: a superset).  A pop is one
// 8-byte load: tmin = tmax, tmax = entry.exit.
struct LeanWalker {
    float tmin, tmax;
    int node; // node to visit next; kDone when the walk is over
    int sp, lo, flagged;

    CRT_WALK_FN bool done() const { return node == kDone; }

    template <class WS>
    CRT_WALK_FN void begin(WS &ws, unsigned lane, int first_node, float t_enter, float t_exit) {
        sp = 1, lo = 0, flagged = 0;
        ws.stack[0][lane].x = uint32_t(kDone); // the sentinel: popping it ends the walk
        ws.stack[0][lane].y = float_to_bits(t_exit);
        tmin = t_enter, tmax = t_exit;
        node = first_node;
    }

    // start a walk from a ray record (entry / exit of the root box and 1/direction come with the record)
    template <class WS>
    CRT_WALK_FN void start(WS &ws, unsigned lane, float ox, float oy, float oz, float dx, float dy, float dz, float ix, float iy, float iz,
                           float t_enter, float t_exit) {
        begin(ws, lane, 0, t_enter, t_exit);
        ws.ray[0][lane] = ox, ws.ray[1][lane] = oy, ws.ray[2][lane] = oz;
        ws.ray[3][lane] = dx, ws.ray[4][lane] = dy, ws.ray[5][lane] = dz;
        ws.ray[6][lane] = ix, ws.ray[7][lane] = iy, ws.ray[8][lane] = iz;
    }

    // ONE iteration of the walk of lane `src`'s ray at node `n` (= the record of `node`; anything when the subtree is
    // pruned), written as selects: the lanes of a warp sit at unrelated nodes and must share one instruction stream
    // (ncu r2a: the same decisions written with branches ran at 14 instead of 19 active threads per instruction and
    // issued as many warp instructions as the walker they replaced).  Returns true when the walk could NOT move on and
    // the caller has something to do first: a leaf and no room for it (leaf_room was false), or one of the two rare
    // events -- a push found the stack window full, a pop found its entry in the spill array -- which are not
    // handled here (if-converted they would cost their issue slots on every visit): call make_room(), visit again.
    //   threshold = prune_threshold (best_t): a subtree whose interval starts beyond it cannot beat or tie the best hit.
    //   With a wide margin and no in-plane entry pending, everything on the stack is beyond it as well (exit distances
    //   grow towards the bottom of the stack, and each interval starts where the one above it ends): the walk is over.
    //   took_leaf: the caller takes over leaf `leaf_id` (meaningless otherwise): the id of the record leaf just visited
    //   (`n` is its record), or an inline leaf -- the visited node itself (popped from the stack / refused before) or the
    //   CHILD this visit would have moved to: the walk has already popped past it.
    template <class WS>
    CRT_WALK_FN bool visit(WS &ws, unsigned lane, unsigned src, const Node n, float threshold, float best_t, bool leaf_room, bool &took_leaf,
                           int &leaf_id, int *pushed = nullptr) {
        const int at = node;
        const bool pruned = tmin > threshold;
        const bool leaf = is_inline_leaf(at) | (n.w == 3); // (an inline leaf comes with n = zeros)
        const bool inner = !pruned & !leaf;
        // ---- the plane (a leaf's axis 3 reads direction.x and the word after the ray block: harmless, unused)
        const int axis = n.w;
        const float split = bits_to_float(uint32_t(n.x));
        const float oa = ws.ray[axis][src], ia = ws.ray[6 + axis][src];
        const float tp = (split - oa) * ia; // NaN: origin on the plane, direction parallel to it
        const bool backwards = ia < 0.0f;   // sign of the direction component (1/-0 = -inf keeps it)
        const int first = backwards ? n.z : n.y, second = backwards ? n.y : n.z;
        // slack |tp| * 1e-5 + 1e-6 on both comparisons, folded into the plane distance (one FFMA each); written so
        // that tp = -inf / +inf (direction parallel to the plane, origin off it) still means "behind" / "never
        // reached" -- a slack of inf would turn both comparisons into NaN tests -- and tp = NaN means both
        const bool want_first = !(fmaf(tp, 1.00001f, 1e-6f) < tmin), want_second = !(fmaf(tp, 0.99999f, -1e-6f) > tmax);
        const bool go_first = inner & want_first & (first >= 0), go_second = inner & want_second & (second >= 0);
        const bool to_child = go_first | go_second, both = go_first & go_second;
        const int child = go_first ? first : second;
        // the child is an inline leaf and the caller has room: handed over right here, and the walk pops past it (the
        // pop below takes the entry this very visit pushes when both children are wanted)
        const bool absorb = kInlineLeaves & to_child & is_inline_leaf(child) & leaf_room;
        // ---- rare events: nothing happens in this call
        const bool full = both & (sp - lo == kWin);
        const bool over = pruned & (flagged == 0) & (fmaf(tmin, 0.999f, -1e-4f) > best_t);
        const bool wants_pop = (pruned & !over) | leaf | (inner & !to_child) | absorb;
        const bool top_spilled = wants_pop & !both & (sp - 1 < lo);
        const bool frozen = full | top_spilled;
        // ---- push the second child (a predicated store), go to the first
        const bool push = both & !frozen;
        const bool in_plane = tp != tp;
        if (push) {
            // an in-plane entry starts at 0 instead of at the current tmax (which the NaN leaves unclipped)
            ws.stack[sp & (kWin - 1)][lane].x = uint32_t(second) | (in_plane ? kFlagBit : 0u);
            ws.stack[sp & (kWin - 1)][lane].y = float_to_bits(tmax);
        }
        sp += push ? 1 : 0;
        if (pushed)
            *pushed = push ? second : -1; // (for a prefetch of the node the walk will come back to)
        flagged += (push & in_plane) ? 1 : 0;
        tmax = push ? fminf(tmax, tp) : tmax; // NaN: unchanged
        const bool moved = to_child & !frozen;
        tmin = (moved & !go_first) ? fmaxf(tmin, tp) : tmin; // NaN: unchanged
        node = moved ? child : node;
        // ---- or hand the leaf over / pop the next pending subtree
        const bool take_here = !pruned & leaf & leaf_room & !frozen, no_room = !pruned & leaf & !leaf_room;
        const bool take_child = absorb & !frozen;
        const bool pop = wants_pop & !no_room & !frozen;
        sp -= pop ? 1 : 0;
        const uint32_t ex = ws.stack[sp & (kWin - 1)][lane].x, ey = ws.stack[sp & (kWin - 1)][lane].y;
        const bool flag = pop & ((ex & kFlagBit) != 0u);
        flagged -= flag ? 1 : 0;
        tmin = pop ? (flag ? 0.0f : tmax) : tmin;
        tmax = pop ? bits_to_float(ey) : tmax;
        node = pop ? int(ex & ~kFlagBit) : node;
        node = over ? kDone : node;
        took_leaf = take_here | take_child;
        leaf_id = leaf ? at : child;
        return no_room | frozen;
    }

    // after visit() returned true: the oldest window entry goes to the spill array when the window is full, the top
    // entry comes back from it when it is there (both are valid whenever their condition holds, whatever stopped the walk)
    template <class WS, class SPILL>
    CRT_WALK_FN void make_room(WS &ws, SPILL *spill, unsigned lane) {
        if (sp - lo == kWin) {
            spill[lo].x = ws.stack[lo & (kWin - 1)][lane].x, spill[lo].y = ws.stack[lo & (kWin - 1)][lane].y;
            ++lo;
        } else if (sp - 1 < lo && sp >= 1) {
            lo = sp - 1;
            ws.stack[lo & (kWin - 1)][lane].x = spill[lo].x, ws.stack[lo & (kWin - 1)][lane].y = spill[lo].y;
        }
    }
};

// HOST: rewrite the child pointers of a walk tree ({split, a, b, axis} / {0, first, count, 3} records with members
// x, y, z, w) that lead to leaves small enough into inline leaves.  The leaves' records stay where they are (nothing
// points to them any more, so they drop out of the walk's working set).  Returns how many pointers were rewritten.
template <class NodeRec>
inline long long inline_leaf_children(NodeRec *nodes, int n_nodes) {
    long long rewritten = 0;
    for (int i = 0; kInlineLeaves && i < n_nodes; ++i) {
        if (nodes[i].w == 3)
            continue;
        for (int side = 0; side < 2; ++side) {
            auto &child = side ? nodes[i].z : nodes[i].y;
            if (child < 0 || child >= n_nodes || nodes[child].w != 3)
                continue;
            const long long first = nodes[child].y, count = nodes[child].z;
            if (count < 0 || count > kLeafCountMax || first < 0 || first > (long long)kLeafFirstMask)
                continue;
            child = make_inline_leaf(int(first), int(count));
            ++rewritten;
        }
    }
    return rewritten;
}

} // namespace crtwalk
