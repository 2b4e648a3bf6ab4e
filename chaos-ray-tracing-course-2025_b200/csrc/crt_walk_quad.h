// The k-d walk two levels at a time (host + device: crt_trace_lean.cuh / crt_persistent.cuh drive it warp-wide,
// tests/host/walk_lean_host.cpp drives the very same code one ray at a time on the CPU against brute force).
//
// Why (ncu r2b): the lean binary walker cut the warp instructions of the C4 walk kernels by 13 %
// and their time by 1 % -- issue-slot utilisation fell from 79 to 70 %, `long scoreboard` rose to 5.4 warps per issue:
// the walk is a chain of DEPENDENT node fetches (53 per ray, each an L2 or DRAM round trip for at least one of the 32
// lanes), and what bounds it is the length of that chain, not the instructions between two fetches.  A QUAD node holds
// a k-d node together with its two children -- three split planes, four grandchildren -- in 32 bytes (one sector,
// two 16-byte loads issued together): one fetch decides two levels, the chain is half as long.  Leaves are no nodes at
// all any more: a child reference names either a quad node (>= 0), nothing (-1), or a LEAF (<= -2, index into a table of
// {first reference, count}), so reaching a leaf costs no fetch either.
//
// Replaces: the box tests of ray_intersect_acceleration_tree (src/core/crt_intersection.cpp:109-136) and
// ray_intersect_aabb_p (:14-45) for the walk tree of core/crt_kdbuild.cpp, as crt_walk_lean.h did one level at a time.
// Same decisions per plane (children in the order the ray's line passes them, slack |tp| * 1e-5 + 1e-6, a ray lying in a
// plane takes both sides through the NaN semantics of the comparisons): a SUPERSET of the cells the reference's
// six-plane test accepts; extra leaves cannot change min (t, parse index).
//
// Stack entry = {child reference, half2 (start rounded DOWN, exit rounded UP)}: 8 bytes with both ends of the interval
// explicit (binary16 keeps 11 bits: the interval grows by < 0.05 %, never shrinks).  With up to four children per visit
// the "a popped interval starts where the walk stands" shortcut of crt_walk_lean.h would stretch intervals over the
// gaps that absent and leaf children leave, and a stretched interval walks cells the ray never enters.
#pragma once

#include <cstdint>
#include <cstring>
#include <cmath>

#if defined(__CUDACC__)
#include <cuda_fp16.h>
#define CRT_QUAD_FN __host__ __device__ __forceinline__
#else
#define CRT_QUAD_FN inline
#endif

namespace crtquad {

constexpr int kWin = 8;     // top-of-stack window per lane in shared memory; older entries spill
constexpr int kDone = -1;   // `node` of a finished walk; also the sentinel entry at the bottom of the stack and an absent child
// child references: >= 0 quad node, -1 absent, <= -2 leaf (leaf id = -ref - 2)
CRT_QUAD_FN int leaf_ref(int leaf_id) { return -leaf_id - 2; }
CRT_QUAD_FN int leaf_id(int ref) { return -ref - 2; }

struct Quad {       // 32 bytes: two 16-byte loads
    int32_t split_r, split_a, split_b; // float bits of the three planes: the node's own, child a's, child b's
    int32_t meta;                      // axis of r | axis of a << 2 | axis of b << 4; axis 3 = "no plane here": the side is
                                       // a single child (a leaf or nothing) kept in its first slot
    int32_t c[4];                      // grandchildren: a's two, b's two (below the plane first)
};
struct Entry {
    int32_t ref;
    uint32_t interval; // binary16 start | binary16 exit << 16
};

CRT_QUAD_FN float as_float(int32_t u) {
#if defined(__CUDA_ARCH__)
    return __int_as_float(u);
#else
    float f;
    std::memcpy(&f, &u, 4);
    return f;
#endif
}

// ---- binary16 with directed rounding (start down, exit up); the host versions exist for the CPU harness
#if !defined(__CUDA_ARCH__)
inline float half_bits_to_float(uint16_t h) {
    const uint32_t sign = uint32_t(h & 0x8000u) << 16, exp = (h >> 10) & 31u, man = h & 1023u;
    float out;
    if (exp == 31u) {
        const uint32_t u = sign | 0x7f800000u | (man << 13);
        std::memcpy(&out, &u, 4);
    } else if (exp == 0u) {
        out = std::ldexp(float(man), -24);
        if (sign)
            out = -out;
    } else {
        const uint32_t u = sign | ((exp + 112u) << 23) | (man << 13);
        std::memcpy(&out, &u, 4);
    }
    return out;
}
inline uint16_t float_to_half_bits_directed(float f, bool up) {
    if (f != f)
        return 0x7e00u;
    // candidates around f: search the monotone map of non-negative / negative halves
    const bool neg = std::signbit(f);
    const float a = std::fabs(f);
    // largest half magnitude <= a (round towards zero in magnitude)
    uint16_t lo = 0, hi = 0x7c00u; // 0 .. +inf
    while (hi - lo > 1) {
        const uint16_t mid = uint16_t((lo + hi) / 2);
        if (half_bits_to_float(mid) <= a)
            lo = mid;
        else
            hi = mid;
    }
    const bool exact = half_bits_to_float(lo) == a;
    uint16_t mag = lo;
    // towards +inf for positive values / towards -inf for negative ones means a LARGER magnitude
    if (!exact && ((up && !neg) || (!up && neg)))
        mag = uint16_t(lo + 1); // may become 0x7c00 = inf
    if (a == INFINITY)
        mag = 0x7c00u;
    return uint16_t(mag | (neg ? 0x8000u : 0u));
}
#endif

CRT_QUAD_FN uint32_t pack_interval(float start, float exit) {
#if defined(__CUDA_ARCH__)
    return uint32_t(__half_as_ushort(__float2half_rd(start))) | (uint32_t(__half_as_ushort(__float2half_ru(exit))) << 16);
#else
    return uint32_t(float_to_half_bits_directed(start, false)) | (uint32_t(float_to_half_bits_directed(exit, true)) << 16);
#endif
}
CRT_QUAD_FN float interval_start(uint32_t v) {
#if defined(__CUDA_ARCH__)
    return __half2float(__ushort_as_half((unsigned short)(v & 0xffffu)));
#else
    return half_bits_to_float(uint16_t(v & 0xffffu));
#endif
}
CRT_QUAD_FN float interval_exit(uint32_t v) {
#if defined(__CUDA_ARCH__)
    return __half2float(__ushort_as_half((unsigned short)(v >> 16)));
#else
    return half_bits_to_float(uint16_t(v >> 16));
#endif
}

// pruning threshold for a best hit distance: a subtree whose interval starts beyond it cannot beat or tie the hit
CRT_QUAD_FN float prune_threshold(float best_t) { return fmaf(fabsf(best_t), 1e-5f, best_t) + 1e-6f; }

// WS: per-warp storage with  float ray[9][32]  (origin, direction, 1/direction)  and  uint2-like stack[kWin][32]
struct QuadWalker {
    float tmin, tmax;
    int node; // quad node (>= 0) or leaf (<= -2) the walk stands at; kDone when it is over
    int sp, lo;
    int flagged; // pending entries that were pushed for a ray lying IN a split plane (see visit)

    CRT_QUAD_FN bool done() const { return node == kDone; }

    template <class WS>
    CRT_QUAD_FN void begin(WS &ws, unsigned lane, int first_node, float t_enter, float t_exit) {
        sp = 1, lo = 0, flagged = 0;
        ws.stack[0][lane].x = uint32_t(kDone); // the sentinel: popping it ends the walk
        ws.stack[0][lane].y = 0u;
        tmin = t_enter, tmax = t_exit;
        node = first_node;
    }

    template <class WS>
    CRT_QUAD_FN void start(WS &ws, unsigned lane, float ox, float oy, float oz, float dx, float dy, float dz, float ix, float iy, float iz,
                           float t_enter, float t_exit) {
        begin(ws, lane, 0, t_enter, t_exit);
        ws.ray[0][lane] = ox, ws.ray[1][lane] = oy, ws.ray[2][lane] = oz;
        ws.ray[3][lane] = dx, ws.ray[4][lane] = dy, ws.ray[5][lane] = dz;
        ws.ray[6][lane] = ix, ws.ray[7][lane] = iy, ws.ray[8][lane] = iz;
    }

    // one plane against the interval [lo_t, hi_t]: which of its two sides the ray's line passes first, which of them
    // the interval reaches, and where they meet.  axis 3 = no plane: one "first" side with the whole interval.
    struct Cut {
        float tp;
        bool backwards, want_first, want_second;
    };
    template <class WS>
    CRT_QUAD_FN Cut cut(const WS &ws, unsigned src, int axis, float split, float lo_t, float hi_t) const {
        const bool plane = axis != 3;
        const float oa = ws.ray[axis][src], ia = ws.ray[6 + axis][src]; // (axis 3 reads direction.x / the word after the ray block: unused)
        Cut c;
        c.tp = (split - oa) * ia;                // NaN: origin on the plane, direction parallel to it
        c.backwards = plane & (ia < 0.0f);       // sign of the direction component (1/-0 = -inf keeps it)
        // slack |tp| * 1e-5 + 1e-6 folded into the plane distance; tp = -inf / +inf (parallel, origin off the plane)
        // still mean "behind" / "never reached", NaN means both
        c.want_first = !plane | !(fmaf(c.tp, 1.00001f, 1e-6f) < lo_t);
        c.want_second = plane & !(fmaf(c.tp, 0.99999f, -1e-6f) > hi_t);
        c.tp = plane ? c.tp : hi_t; // no plane: the first side keeps the whole interval (fminf (hi, hi))
        return c;
    }

    // ONE iteration of the walk of lane `src`'s ray, written as selects (the lanes of a warp sit at unrelated nodes and
    // share one instruction stream).  `q` = the record of `node` when the walk stands at a quad node whose interval is
    // not pruned (anything otherwise).  Returns true when the walk could NOT move on and the caller has something to do
    // first: it stands at a leaf and leaf_room is false, or a rare event -- the pushes do not fit the stack window, the
    // entry to pop is in the spill array -- for which the caller calls make_room() and visits again.
    // took_leaf: the leaf (id) the walk stood at and the caller must take now, or -1.
    //   threshold = prune_threshold (best_t): an interval that starts beyond it cannot beat or tie the best hit; with a
    //   wide margin everything on the stack starts beyond it as well (front to back), and the walk is over -- unless an
    //   entry is pending that was pushed for a ray lying IN a split plane (tp = NaN): both sides of such a plane get the
    //   same interval, so that entry starts no later than the side visited first.  Those entries carry a flag (the sign
    //   bit of their start, which is never negative) and are counted in `flagged`.
    template <class WS>
    CRT_QUAD_FN bool visit(WS &ws, unsigned lane, unsigned src, const Quad &q, float threshold, float best_t, bool leaf_room, int &took_leaf) {
        const bool pruned = tmin > threshold;
        const bool at_leaf = node <= -2;
        const bool inner = !pruned & !at_leaf;
        // ---- stage 1: the node's own plane
        const Cut r = cut(ws, src, q.meta & 3, as_float(q.split_r), tmin, tmax);
        const float f_lo = tmin, f_hi = fminf(tmax, r.tp), s_lo = fmaxf(tmin, r.tp), s_hi = tmax; // NaN: whole interval both
        // the side passed first is a unless the ray runs backwards; its plane and its two slots
        const int f_axis = r.backwards ? (q.meta >> 4) & 3 : (q.meta >> 2) & 3, s_axis = r.backwards ? (q.meta >> 2) & 3 : (q.meta >> 4) & 3;
        const float f_split = as_float(r.backwards ? q.split_b : q.split_a), s_split = as_float(r.backwards ? q.split_a : q.split_b);
        const int f_c0 = r.backwards ? q.c[2] : q.c[0], f_c1 = r.backwards ? q.c[3] : q.c[1];
        const int s_c0 = r.backwards ? q.c[0] : q.c[2], s_c1 = r.backwards ? q.c[1] : q.c[3];
        // ---- stage 2: the plane of either side against that side's interval
        const Cut f = cut(ws, src, f_axis, f_split, f_lo, f_hi), s = cut(ws, src, s_axis, s_split, s_lo, s_hi);
        // the four grandchildren in the order the line passes them, with their intervals
        const int k0 = f.backwards ? f_c1 : f_c0, k1 = f.backwards ? f_c0 : f_c1, k2 = s.backwards ? s_c1 : s_c0, k3 = s.backwards ? s_c0 : s_c1;
        const float lo0 = f_lo, hi0 = fminf(f_hi, f.tp), lo1 = fmaxf(f_lo, f.tp), hi1 = f_hi;
        const float lo2 = s_lo, hi2 = fminf(s_hi, s.tp), lo3 = fmaxf(s_lo, s.tp), hi3 = s_hi;
        const bool w0 = inner & r.want_first & f.want_first & (k0 != -1) & !(lo0 > threshold);
        const bool w1 = inner & r.want_first & f.want_second & (k1 != -1) & !(lo1 > threshold);
        const bool w2 = inner & r.want_second & s.want_first & (k2 != -1) & !(lo2 > threshold);
        const bool w3 = inner & r.want_second & s.want_second & (k3 != -1) & !(lo3 > threshold);
        // ---- the first wanted child is visited next, the others are pushed, farthest first
        const bool p1 = w1 & w0, p2 = w2 & (w0 | w1), p3 = w3 & (w0 | w1 | w2);
        const int n_push = (p1 ? 1 : 0) + (p2 ? 1 : 0) + (p3 ? 1 : 0);
        const bool any = w0 | w1 | w2 | w3;
        // ---- rare events: nothing happens in this call
        const bool full = sp - lo + n_push > kWin;
        const bool over = pruned & (flagged == 0) & (fmaf(tmin, 0.999f, -1e-4f) > best_t);
        const bool nan_r = r.tp != r.tp, nan_f = f.tp != f.tp, nan_s = s.tp != s.tp;
        const bool flag1 = nan_f, flag2 = nan_r, flag3 = nan_r | nan_s;
        const bool leaf_here = at_leaf & !pruned;
        const bool wants_pop = (pruned & !over) | (inner & !any) | (leaf_here & leaf_room);
        const bool top_spilled = wants_pop & (sp - 1 < lo);
        const bool frozen = full | top_spilled;
        if (p3 & !frozen) {
            ws.stack[sp & (kWin - 1)][lane].x = uint32_t(k3), ws.stack[sp & (kWin - 1)][lane].y = pack_interval(lo3, hi3) | (flag3 ? 0x8000u : 0u);
        }
        sp += (p3 & !frozen) ? 1 : 0;
        if (p2 & !frozen) {
            ws.stack[sp & (kWin - 1)][lane].x = uint32_t(k2), ws.stack[sp & (kWin - 1)][lane].y = pack_interval(lo2, hi2) | (flag2 ? 0x8000u : 0u);
        }
        sp += (p2 & !frozen) ? 1 : 0;
        if (p1 & !frozen) {
            ws.stack[sp & (kWin - 1)][lane].x = uint32_t(k1), ws.stack[sp & (kWin - 1)][lane].y = pack_interval(lo1, hi1) | (flag1 ? 0x8000u : 0u);
        }
        sp += (p1 & !frozen) ? 1 : 0;
        flagged += frozen ? 0 : ((p1 & flag1) ? 1 : 0) + ((p2 & flag2) ? 1 : 0) + ((p3 & flag3) ? 1 : 0);
        const bool moved = any & !frozen;
        const int here = node;
        node = moved ? (w0 ? k0 : (w1 ? k1 : (w2 ? k2 : k3))) : node;
        const float next_lo = w0 ? lo0 : (w1 ? lo1 : (w2 ? lo2 : lo3)), next_hi = w0 ? hi0 : (w1 ? hi1 : (w2 ? hi2 : hi3));
        tmin = moved ? next_lo : tmin;
        tmax = moved ? next_hi : tmax;
        // ---- or pop the next pending child
        const bool pop = wants_pop & !frozen;
        sp -= pop ? 1 : 0;
        const uint32_t ex = ws.stack[sp & (kWin - 1)][lane].x, ey = ws.stack[sp & (kWin - 1)][lane].y;
        node = pop ? int(ex) : node;
        flagged -= (pop & ((ey & 0x8000u) != 0u)) ? 1 : 0;
        tmin = pop ? interval_start(ey & ~0x8000u) : tmin;
        tmax = pop ? interval_exit(ey) : tmax;
        node = over ? kDone : node;
        took_leaf = (leaf_here & leaf_room & !frozen) ? leaf_id(here) : -1;
        return (leaf_here & !leaf_room) | frozen;
    }

    // after visit() returned true: the oldest window entries go to the spill array while the window is too full for
    // three more, the top entry comes back from it when it is there (both are valid whenever their condition holds)
    template <class WS, class SPILL>
    CRT_QUAD_FN void make_room(WS &ws, SPILL *spill, unsigned lane) {
        if (sp - 1 < lo && sp >= 1) {
            lo = sp - 1;
            ws.stack[lo & (kWin - 1)][lane].x = spill[lo].x, ws.stack[lo & (kWin - 1)][lane].y = spill[lo].y;
        }
        while (node >= 0 && sp - lo + 3 > kWin) { // (only a walk that stands at a quad node pushes)
            spill[lo].x = ws.stack[lo & (kWin - 1)][lane].x, spill[lo].y = ws.stack[lo & (kWin - 1)][lane].y;
            ++lo;
        }
    }
};

} // namespace crtquad
