// TEST INFRASTRUCTURE: drives csrc/crt_walk_lean.h -- the very code the device kernels run per lane -- one ray at a time
// on the CPU, next to a brute-force loop over every triangle with the same triangle test, so that the walk's decisions
// (which cells a ray visits) are checked on every scene without a GPU.  The walk must find min (t, parse index) over
// everything the ray hits: ray_intersect_acceleration_tree's winner, src/core/crt_intersection.cpp:109-136, :95-107.
//
//   g++ -O2 -std=c++17 -fPIC -shared -ffp-contract=off -I../../include tests/host/walk_lean_host.cpp -o tests/host/_build/libwalk_lean_host.so
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <limits>
#include <vector>

#include "../../include/crt_cuda.h"
#include "../../chaos-ray-tracing-course-2025_b200/csrc/crt_walk_lean.h"

namespace {

using crtwalk::Entry;
using crtwalk::LeanWalker;
using crtwalk::Node;

struct HostWarp { // one lane's worth of LightWarpShared
    float ray[9][1];
    Entry stack[crtwalk::kWin][1];
};

struct V {
    float x, y, z;
};
inline V operator-(V a, V b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
inline V operator+(V a, V b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
inline V operator*(V a, float s) { return { a.x * s, a.y * s, a.z * s }; }
inline float dot(V a, V b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V cross(V a, V b) { return { a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; }

// ray_intersect_triangle's accept / reject and distance (crt_intersection.cpp:47-70), as crt_device.cuh states it
bool triangle_test(const crt_flat_scene &s, uint32_t ref, V o, V d, float &t_out) {
    const int tri = int(ref & 0x7FFFFFFFu);
    const bool cull = (ref >> 31) != 0;
    const crt_float4 A = s.tri_verts[3 * tri], B = s.tri_verts[3 * tri + 1], C = s.tri_verts[3 * tri + 2];
    const V v0{ A.x, A.y, A.z }, v1{ B.x, B.y, B.z }, v2{ C.x, C.y, C.z }, N{ A.w, B.w, C.w };
    const float rn = dot(N, d);
    if (std::fabs(rn) < 1e-6f)
        return false;
    const float dist = dot(N, v0 - o);
    if (!(dist < 0.0f) && cull)
        return false;
    const float t = dist / rn;
    if (t < 0.0f)
        return false;
    const V P = o + d * t;
    const V e0 = v1 - v0, e1 = v2 - v1, e2 = v0 - v2;
    if (dot(N, cross(e0, P - v0)) >= 0.0f && dot(N, cross(e1, P - v1)) >= 0.0f && dot(N, cross(e2, P - v2)) >= 0.0f) {
        t_out = t + 0.0f; // -0 -> +0, as the device's (t bits, triangle) key needs it
        return true;
    }
    return false;
}

// ray_setup of crt_device.cuh: entry / exit of the root box, slack on the exit
bool root_interval(const crt_flat_scene &s, V o, V d, V &inv, float &tmin, float &tmax) {
    inv = { 1.0f / d.x, 1.0f / d.y, 1.0f / d.z };
    tmin = 0.0f, tmax = std::numeric_limits<float>::infinity();
    bool live = d.x == d.x && d.y == d.y && d.z == d.z;
    const float oa[3] = { o.x, o.y, o.z }, da[3] = { d.x, d.y, d.z }, ia[3] = { inv.x, inv.y, inv.z };
    for (int k = 0; k < 3; ++k) {
        if (da[k] != 0.0f) {
            const float a = (s.root_min[k] - oa[k]) * ia[k], b = (s.root_max[k] - oa[k]) * ia[k];
            tmin = std::fmax(tmin, std::fmin(a, b));
            tmax = std::fmin(tmax, std::fmax(a, b));
        } else if (oa[k] < s.root_min[k] || oa[k] > s.root_max[k])
            live = false;
    }
    tmax = tmax + (std::fabs(tmax) * 1e-5f + 1e-6f);
    if (!(tmin <= tmax))
        live = false;
    return live;
}

struct Best {
    float t = std::numeric_limits<float>::infinity();
    int tri = -1;
    void offer(float t_new, int tri_new) {
        if (t_new < t || (t_new == t && tri_new < tri))
            t = t_new, tri = tri_new;
    }
};

} // namespace

extern "C" {

// closest hit of every ray through LeanWalker; rays = n x {origin xyz, direction xyz}.  visits_out (optional): nodes
// loaded per ray.  `lazy_best`: how many leaf visits pass before the pruning threshold sees a new hit (the device tests
// parked leaves late, so its walk prunes with a stale best distance: 0 = immediately, k = after k more leaves).
int walk_lean_closest(const crt_flat_scene *s, const float *rays, int n, int lazy_best, int32_t *tri_out, float *t_out, int32_t *visits_out,
                      long long *rare_events) {
    if (!s || !rays || !s->fast_nodes)
        return -1;
    std::vector<Entry> spill(64);
    for (int r = 0; r < n; ++r) {
        const V o{ rays[6 * r], rays[6 * r + 1], rays[6 * r + 2] }, d{ rays[6 * r + 3], rays[6 * r + 4], rays[6 * r + 5] };
        V inv;
        float t_enter, t_exit;
        Best best;
        int visits = 0;
        if (root_interval(*s, o, d, inv, t_enter, t_exit)) {
            HostWarp ws{};
            LeanWalker w;
            w.start(ws, 0u, o.x, o.y, o.z, d.x, d.y, d.z, inv.x, inv.y, inv.z, t_enter, t_exit);
            float seen_t = best.t; // the distance the walk's end is decided with
            int pending = 0;
            bool refuse_next = false;
            for (;;) {
                // the device computes the cut once per round of visits: here a round is one visit
                w.end_if_beyond(crtwalk::walk_cut(seen_t, w.flagged));
                if (w.done())
                    break;
                if (w.sp - w.lo > crtwalk::kWin || w.sp > 60)
                    return -2; // stack discipline broken
                if (w.node < 0 || w.node >= s->n_fast_nodes)
                    return -4; // the walk asked for a record that does not exist
                const crt_kd_node &k = s->fast_nodes[w.node];
                Node node;
                std::memcpy(&node.x, &k.split, 4);
                node.y = k.a, node.z = k.b, node.w = k.axis;
                const int leaf_first = k.a, leaf_count = k.b;
                if (crtwalk::kLeafNeg && node.w == 3)
                    node.y = ~node.y, node.z = ~node.z; // as crt_cuda_upload_scene stores leaves on the device
                ++visits;
                // every other leaf is refused once (no room) and must be offered again: the device's blocked lanes
                const bool room = !(refuse_next && node.w == 3);
                bool took_leaf = false;
                const bool stopped = w.visit(ws, 0u, 0u, node, room, took_leaf);
                if (stopped && node.w == 3 && !room)
                    refuse_next = false; // refused once: the next offer of this leaf is taken
                if (took_leaf) {
                    refuse_next = true;
                    for (int i = 0; i < leaf_count; ++i) {
                        const uint32_t ref = s->fast_refs[leaf_first + i];
                        float t;
                        if (triangle_test(*s, ref, o, d, t))
                            best.offer(t, int(ref & 0x7FFFFFFFu));
                    }
                    if (++pending > lazy_best)
                        seen_t = best.t, pending = 0;
                }
                if (stopped) {
                    const int lo_before = w.lo;
                    w.make_room(ws, spill.data(), 0u);
                    *rare_events += w.lo != lo_before;
                }
                if (visits > 50000000)
                    return -3; // runaway walk
            }
        }
        tri_out[r] = best.tri;
        t_out[r] = best.tri >= 0 ? best.t : 0.0f;
        if (visits_out)
            visits_out[r] = visits;
    }
    return 0;
}

// every triangle in parse order, min (t, index): what any correct traversal must return
int brute_closest(const crt_flat_scene *s, const float *rays, int n, int32_t *tri_out, float *t_out) {
    if (!s || !rays)
        return -1;
    for (int r = 0; r < n; ++r) {
        const V o{ rays[6 * r], rays[6 * r + 1], rays[6 * r + 2] }, d{ rays[6 * r + 3], rays[6 * r + 4], rays[6 * r + 5] };
        Best best;
        for (int tri = 0; tri < s->n_triangles; ++tri) {
            const bool cull = (s->tri_meta[4 * tri + 3] & CRT_META_CULL_BIT) != 0;
            float t;
            if (triangle_test(*s, uint32_t(tri) | (cull ? 0x80000000u : 0u), o, d, t))
                best.offer(t, tri);
        }
        tri_out[r] = best.tri;
        t_out[r] = best.tri >= 0 ? best.t : 0.0f;
    }
    return 0;
}

} // extern "C"
