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
//   * the stack holds a SENTINEL entry {kDone, root exit} at the bottom: pop never asks whether the stack is empty or
//     whether the entry below exists;
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
constexpr int kWin = 8;                    // top-of-stack window per lane in shared memory; older entries spill

struct Node {   // crt_kd_node as the kernels load it: {split bits, a, b, axis} / {0, first, count, 3}
    int32_t x, y, z, w;
};
struct Entry {  // {node | kFlagBit, entry distance bits}; the EXIT distance of an entry is the entry distance of the
    uint32_t x, y; // entry below it (every push clips the current interval at the pushed plane)
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

// pruning threshold for a best hit distance: a subtree whose interval starts beyond it cannot beat or tie the hit
CRT_WALK_FN float prune_threshold(float best_t) { return fmaf(fabsf(best_t), 1e-5f, best_t) + 1e-6f; }

// WS: per-warp storage with  float ray[9][32]  (origin, direction, 1/direction)  and  Entry-like stack[kWin][32]
// SPILL: Entry-like array of kStack entries owned by the caller (local memory on the device)
struct LeanWalker {
    float tmin, tmax;
    int node; // node to visit next; kDone when the walk is over; < 0 right after step() when the child is absent
    int sp, lo, flagged;

    CRT_WALK_FN bool done() const { return node == kDone; }

    template <class WS>
    CRT_WALK_FN void begin(WS &ws, unsigned lane, int first_node, float t_enter, float t_exit) {
        sp = 1, lo = 0, flagged = 0;
        ws.stack[0][lane].x = uint32_t(kDone);
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

    // ONE inner node `n` of the walk of lane `src`'s ray.  Leaves node = the child to visit next, or -1 when neither
    // child is wanted (the caller pops).
    template <class WS, class SPILL>
    CRT_WALK_FN void step(WS &ws, SPILL *spill, unsigned lane, unsigned src, const Node n) {
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
        const bool go_first = want_first & (first >= 0), go_second = want_second & (second >= 0);
        if (go_first & go_second) {
            const bool in_plane = tp != tp;
            if (sp - lo == kWin) { // window full (rare): the oldest entry goes to the spill array
                spill[lo].x = ws.stack[lo & (kWin - 1)][lane].x, spill[lo].y = ws.stack[lo & (kWin - 1)][lane].y;
                ++lo;
            }
            // an in-plane entry keeps the exit distance for the entries pushed above it and starts at 0 itself
            ws.stack[sp & (kWin - 1)][lane].x = uint32_t(second) | (in_plane ? kFlagBit : 0u);
            ws.stack[sp & (kWin - 1)][lane].y = float_to_bits(in_plane ? tmax : tp);
            ++sp;
            flagged += in_plane ? 1 : 0;
            tmax = fminf(tmax, tp); // NaN: unchanged
        }
        tmin = (!go_first & go_second) ? fmaxf(tmin, tp) : tmin; // NaN: unchanged
        node = go_first ? first : (go_second ? second : -1);
    }

    // next pending subtree; the sentinel at the bottom ends the walk
    template <class WS, class SPILL>
    CRT_WALK_FN void pop(const WS &ws, const SPILL *spill, unsigned lane) {
        --sp;
        uint32_t ex = ws.stack[sp & (kWin - 1)][lane].x, ey = ws.stack[sp & (kWin - 1)][lane].y;
        uint32_t below = ws.stack[(sp - 1) & (kWin - 1)][lane].y;
        if (sp <= lo) { // rare: the entry or the one below it was spilled (or this is the sentinel)
            if (sp < lo)
                ex = spill[sp].x, ey = spill[sp].y, lo = sp;
            if (sp > 0)
                below = sp - 1 >= lo ? ws.stack[(sp - 1) & (kWin - 1)][lane].y : spill[sp - 1].y;
        }
        const bool flag = (ex & kFlagBit) != 0u;
        flagged -= flag ? 1 : 0;
        tmax = bits_to_float(below);
        tmin = flag ? 0.0f : bits_to_float(ey);
        node = int(ex & ~kFlagBit);
    }

    // the node about to be visited starts beyond the best hit (tmin > threshold): with a wide margin and no in-plane
    // entry pending, everything on the stack is beyond it as well (entry distances grow towards the bottom of the
    // stack) and the walk is over; otherwise the caller pops
    CRT_WALK_FN bool pruned_everything(float best_t) {
        if (flagged == 0 && tmin - (fabsf(tmin) * 1e-3f + 1e-4f) > best_t) {
            node = kDone;
            return true;
        }
        return false;
    }
};

} // namespace crtwalk
