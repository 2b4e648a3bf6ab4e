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
//     semantics of the comparisons and of fminf / fmaxf; the subtree it pushes there starts where the cell starts,
//     which an entry cannot say: the ray is FLAGGED from then on -- every pop starts at 0, no early end (a superset);
//   * the best hit ENDS a walk, it does not prune it (step 27): before a visit the walk asks whether the subtree it
//     stands in starts beyond best + margin (one comparison against a number computed once per round); if so,
//     everything still on the stack does too.  The visit itself knows nothing about hits.
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

constexpr int kDone = 0x7FFFFFFE;          // node id of the sentinel entry: the walk is over
#ifndef CRT_WIN
#define CRT_WIN 8
#endif
constexpr int kWin = CRT_WIN;              // top-of-stack window per lane in shared memory; older entries spill

// CRT_LEAF_NEG: the device's copy of a LEAF record is {0, ~first, ~count, 3}: both "children" of a leaf are negative =
// absent, so the visit needs no "and not a leaf" on its two child decisions (crt_cuda_upload_scene rewrites the records,
// kd_node_at hands the plain form to the kernels that do not use this walker)
#ifndef CRT_LEAF_NEG
#define CRT_LEAF_NEG 1
#endif
constexpr bool kLeafNeg = CRT_LEAF_NEG != 0;

struct Node {   // crt_kd_node as the kernels load it: {split bits, a, b, axis} / {0, first, count, 3} (see CRT_LEAF_NEG)
    int32_t x, y, z, w;
};
struct Entry {  // {node, exit distance bits}
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

// Where a walk ENDS for a best hit distance: a subtree whose interval starts beyond best_t cannot beat or tie the hit,
// and when it starts beyond it by a margin wider than the slack of the comparisons below, so does everything still on
// the stack -- exit distances grow towards the bottom of the stack, each interval starts where the one above it ends,
// and the walk's tmin can run ahead of those starts only by the slack a second-child move allows (1e-5 relative +
// 1e-6; the margin is ten times that).  A FLAGGED ray (visit: it lay in a split plane at some push) has entries that
// start earlier: no end before the stack is empty.  Computed once per round of visits.
CRT_WALK_FN float walk_cut(float best_t, int flagged) { return flagged != 0 ? INFINITY : fmaf(best_t, 1.0001f, 1e-5f); }

// WS: per-warp storage with  float ray[9][32]  (origin, direction, 1/direction)  and  Entry-like stack[kWin][32]
// SPILL: Entry-like array of kStack entries owned by the caller (local memory on the device)
//
// Stack entry = {node, EXIT distance of the pushed subtree}.  Its ENTRY distance need not be stored: when an
// entry is popped, everything pushed after it has been popped too, so the current tmax is again what the push left
// behind -- the plane distance the entry's interval starts at (or less, within the slack: a superset).  A pop is one
// 8-byte load: tmin = tmax, tmax = entry.exit.
struct LeanWalker {
    float tmin, tmax;
    int node; // node to visit next; kDone when the walk is over
    int sp, lo;
    int flagged; // 1 once this ray has pushed an entry while lying IN a split plane (see visit): from then on every popped
                 // interval starts at 0 and the walk never ends early -- a superset, for rays that all but never occur

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

    // the subtree about to be visited starts beyond the cut (walk_cut): the walk is over.  Asked before every visit.
    CRT_WALK_FN void end_if_beyond(float cut) { node = tmin > cut ? kDone : node; }

    // ONE iteration of the walk of lane `src`'s ray at node `n` (= the record of `node`), written as selects: the lanes
    // of a warp sit at unrelated nodes and must share one instruction stream (ncu r2a: the same decisions written with
    // branches ran at 14 instead of 19 active threads per instruction and issued as many warp instructions as the walker
    // they replaced).  Returns true when the caller has something to do before the next visit: a leaf and no room for it
    // (leaf_room was false: nothing happened), one of the two rare events -- a push found the stack window full, a pop
    // found its entry in the spill array: nothing happened, call make_room() and visit again; if-converted they would
    // cost their issue slots on every visit -- or the ray has just been flagged (compute the cut again).
    //   took_leaf: `n` is a leaf, the caller takes it, the walk has moved on.
    // Pruning is not in here (step 27): a subtree is either visited or, beyond the cut, the end of the walk; the narrow
    // band in between (starts beyond the best hit, but by less than the margin) is walked -- 1.5 % more visits on C4 for
    // a dozen instructions less on every one of them.
    template <class WS>
    CRT_WALK_FN bool visit(WS &ws, unsigned lane, unsigned src, const Node n, bool leaf_room, bool &took_leaf, int *pushed = nullptr) {
        const bool leaf = n.w == 3;
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
        const bool inner = kLeafNeg | !leaf; // (kLeafNeg: a leaf's "children" are negative)
        const bool go_first = inner & want_first & (first >= 0), go_second = inner & want_second & (second >= 0);
        const bool to_child = go_first | go_second, both = go_first & go_second;
        // ---- rare events: nothing happens in this call
        const bool full = both & (sp - lo == kWin);
        const bool top_spilled = !to_child & (sp - 1 < lo); // (a leaf, or a node none of whose children the ray enters: pop)
        const bool frozen = full | top_spilled;
        // ---- push the second child (a predicated store), go to the first
        const bool push = both & !frozen;
        if (push) {
            ws.stack[sp & (kWin - 1)][lane].x = uint32_t(second);
            ws.stack[sp & (kWin - 1)][lane].y = float_to_bits(tmax);
        }
        sp += push ? 1 : 0;
        // an entry pushed in-plane starts at the current tmin, not at the tmax of the moment it is popped (which the NaN
        // left unclipped): entries do not say where they start, so from here on EVERY pop of this ray starts at 0
        const bool flag_now = push & (tp != tp);
        flagged = flag_now ? 1 : flagged;
        tmax = push ? fminf(tmax, tp) : tmax; // NaN: unchanged
        const bool moved = to_child & !frozen;
        tmin = (moved & !go_first) ? fmaxf(tmin, tp) : tmin; // NaN: unchanged
        node = moved ? (go_first ? first : second) : node;
        // ---- or hand the leaf over / pop the next pending subtree
        const bool no_room = leaf & !leaf_room;
        const bool pop = !to_child & !no_room & !frozen;
        sp -= pop ? 1 : 0;
        const uint32_t ex = ws.stack[sp & (kWin - 1)][lane].x, ey = ws.stack[sp & (kWin - 1)][lane].y;
        tmin = pop ? (flagged != 0 ? 0.0f : tmax) : tmin;
        tmax = pop ? bits_to_float(ey) : tmax;
        node = pop ? int(ex) : node;
        took_leaf = leaf & pop;
        if (pushed)
            *pushed = push ? second : -1; // (experiment CRT_PREFETCH_PUSH: a prefetch of the node the walk will come back to)
        return no_room | frozen | flag_now;
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

} // namespace crtwalk
