// Device-side building blocks of the render path (sm_100a): float3 algebra, ray generation, the two tree
// traversals, the triangle test, surface reconstruction, texture sampling and the material combinators.
//
// PARITY RULES: every expression below is written with the SAME operations in the SAME
// order as the reference's scalar C++ (IEEE binary32, one rounding per operation).  The translation unit is
// compiled with -fmad=false (no FMA contraction: the reference binary contains none), -prec-div=true,
// -prec-sqrt=true and -ftz=false (nvcc defaults, stated explicitly in the Makefile), so accept/reject decisions,
// hit distances and colours are bit-identical to the CPU's wherever libm is not involved.
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/crt_cuda.h"
#include "crt_sincos.h"

namespace crtdev {

// ------------------------------------------------------------------------------------------------ scene view
// Walk-tree node as the kernels read it.  CRT_NODE8 = 1: 8 bytes {split | first, axis:2 has_a:1 has_b:1 child:28} --
// crt_cuda_upload_scene re-numbers the nodes so that the children of a node are neighbours (child a at `child`, child b
// right after it), which makes one index do for both; a leaf is {first reference, axis = 3 | count << 2}.  Half the
// bytes of the 16-byte crt_kd_node: four nodes per 32-byte sector, siblings in the same sector, and a C4 walk tree of
// 24 MB instead of 49 MB in L2.  kd_node_at() hands the kernels the {split, a, b, axis} / {0, first, count, 3} int4
// they were written against.  OFF by default: measured on B200 the better hit rates do not pay for
// the 16 extra instructions per node step on kernels that are bound by instruction issue (C4 2.97 -> 3.09 ms).
#ifndef CRT_NODE8
#define CRT_NODE8 0
#endif
#if CRT_NODE8 && !defined(CRT_LEAF_NEG)
#define CRT_LEAF_NEG 0 // (the packed records of that experiment keep their own leaf form)
#endif
} // namespace crtdev
#include "crt_walk_lean.h"
namespace crtdev {
#if CRT_NODE8
typedef int2 KdNodeRec;
#else
typedef int4 KdNodeRec;
#endif
constexpr int kNode8MaxNodes = 1 << 28; // child index field
constexpr int kNode8MaxCount = 1 << 29; // leaf count field (a leaf is marked by the low two bits = 3)

struct DScene {
    crt_camera cam;
    float bg[3];
    int gi_on, reflections_on, refractions_on;
    float root_min[3], root_max[3];
    int n_triangles, n_nodes, n_lights, n_materials, n_textures;
    int kd_empty_leaf; // index of a leaf without triangles that the upload appends to the walk tree (tombstone of stolen subtrees)
    const float4 *__restrict__ tri_verts;   // 3 per triangle
    const int4 *__restrict__ tri_meta;      // 1 per triangle
    const float4 *__restrict__ vert_normal; // 1 per vertex
    const float2 *__restrict__ vert_uv;     // 1 per vertex
    const float4 *__restrict__ box_nodes;   // 2 per node
    const KdNodeRec *__restrict__ kd_nodes; // 1 per node of the walk tree (crt_flat_scene.fast_nodes), see kd_node_at
    const uint32_t *__restrict__ kd_refs;   // its leaf lists (crt_flat_scene.fast_refs)
    const uint32_t *__restrict__ refs;      // leaf lists of the reference tree (box_nodes)
    const float4 *__restrict__ lights;
    const int4 *__restrict__ materials;
    const float4 *__restrict__ textures; // 3 float4 per texture (crt_texture is 48 B)
    const float4 *__restrict__ texels;
};

// the record as the device holds it (crt_walk_lean.h, CRT_LEAF_NEG: a leaf is {0, ~first, ~count, 3})
__device__ __forceinline__ int4 kd_node_raw(const DScene &S, int i);

// node `i` as {split, a, b, axis} / {0, first, count, 3}
__device__ __forceinline__ int4 kd_node_at(const DScene &S, int i) {
    int4 n = kd_node_raw(S, i);
    if (crtwalk::kLeafNeg && n.w == 3)
        n.y = ~n.y, n.z = ~n.z;
    return n;
}

__device__ __forceinline__ int4 kd_node_raw(const DScene &S, int i) {
#if CRT_NODE8
    const int2 r = __ldg(&S.kd_nodes[i]);
    const int axis = r.y & 3;
    const unsigned u = unsigned(r.y);
    const int has_a = int(u >> 2) & 1, has_b = int(u >> 3) & 1, child = int(u >> 4);
    const bool leaf = axis == 3;
    return make_int4(r.x, leaf ? r.x : (has_a ? child : -1), leaf ? int(u >> 2) : (has_b ? child + has_a : -1), axis);
#else
    return __ldg(&S.kd_nodes[i]);
#endif
}

struct DParams {
    uint32_t max_ray_depth, gi_rays;
    float shadow_bias, reflection_bias, diffuse_reflection_bias, refraction_bias;
    int shadows;
    int rank, world, band_rows; // row-band ownership
    int rows;                   // rows this launch renders (owned rows)
    int coherent;               // measurement aid ($CRT_COHERENT=1, tools/inst_alg.py): the 32 pixels of a patch all take lane 0's ray
};

struct Counters {
    unsigned long long rays, shadow_rays, node_tests, tri_tests;
    unsigned long long rays_by_depth[16];
    unsigned long long entered; // rays that enter the root box (the rays a walk is started for); the others end in ray setup
};

constexpr int kStack = 48; // tree depth <= 40 (leaf rule depth > 39), stack need <= depth + 2; checked at upload
constexpr float kInf = __builtin_huge_valf();

// ------------------------------------------------------------------------------------------------ float3
struct V3 {
    float x, y, z;
};
__device__ __forceinline__ V3 mk(float x, float y, float z) { return V3{ x, y, z }; }
__device__ __forceinline__ V3 xyz(const float4 &v) { return V3{ v.x, v.y, v.z }; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return V3{ a.x + b.x, a.y + b.y, a.z + b.z }; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return V3{ a.x - b.x, a.y - b.y, a.z - b.z }; }
__device__ __forceinline__ V3 operator-(V3 a) { return V3{ -a.x, -a.y, -a.z }; }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return V3{ a.x * s, a.y * s, a.z * s }; }
// a / s, three IEEE binary32 divisions by one denominator (normalisations, colour / area), two ways with the same bits:
//   * binary32: three divisions.  Their fast path refuses a ZERO NUMERATOR (FCHK) and calls a ~50-instruction fix-up;
//     with axis-aligned walls most normalisations have one (ncu r1zz, GI kernel: 13.7 M calls of the fix-up, 4.2 % of
//     the issued instructions at 3 threads per instruction);
//   * CRT_DIV_VIA_DOUBLE: each quotient as RN32 (RN64 (a * RN64 (1 / s))).  A quotient of two binary32 numbers lies at
//     least 2^-48 (relative) away from every rounding boundary of binary32 unless it is exact, the binary64 product is
//     within 2^-52 of it, so the final rounding lands on the correctly rounded quotient -- zeros, infinities, NaNs and
//     subnormals included (tools/div_check.cpp: 3e8 operand pairs of every class; the GI parity tests compare the
//     two against each other bit for bit).  No fix-up calls, but binary64 work: 15-01/scene2 with GI 211 -> 186 ms,
//     the 1 M-triangle height field (no zero components anywhere) 2.95 -> 3.13 ms.
// So it is a per-kernel choice made at compile time: the GI instantiations of k_render_persistent are built in their own
// translation unit with CRT_DIV_VIA_DOUBLE = 1 (csrc/crt_gi.cu), everything else divides in binary32.  (Both paths in
// one kernel behind a runtime flag were measured too: +16 % code, and the GI kernel, whose top stall is instruction
// fetch, lost more than the division gained: 212 ms.)
#ifndef CRT_DIV_VIA_DOUBLE
#define CRT_DIV_VIA_DOUBLE 0
#endif
__device__ __forceinline__ V3 operator/(V3 a, float s) {
#if CRT_DIV_VIA_DOUBLE
    const double r = __drcp_rn(double(s));
    return V3{ float(double(a.x) * r), float(double(a.y) * r), float(double(a.z) * r) };
#else
    return V3{ a.x / s, a.y / s, a.z / s };
#endif
}
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) { return V3{ a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; }
__device__ __forceinline__ float length(V3 a) { return sqrtf(dot(a, a)); }
__device__ __forceinline__ V3 normalized(V3 a) { return a / length(a); }
// row vector * row-major 3x3, accumulating from 0 over j = 0,1,2 (crt_matrix.h:66-74)
__device__ __forceinline__ V3 mul_rows(V3 v, V3 r0, V3 r1, V3 r2) {
    return V3{ ((0.0f + v.x * r0.x) + v.y * r1.x) + v.z * r2.x, ((0.0f + v.x * r0.y) + v.y * r1.y) + v.z * r2.y,
               ((0.0f + v.x * r0.z) + v.y * r1.z) + v.z * r2.z };
}
__device__ __forceinline__ float pick(V3 v, int axis) { return axis == 0 ? v.x : (axis == 1 ? v.y : v.z); }

struct Ray {
    V3 o, d;
};
struct Hit {
    float t;
    int tri; // -1 = miss
};

// ------------------------------------------------------------------------------------------------ camera (a1)
// Camera::generate_ray, src/core/crt_camera.cpp:7-35.  tanf(fov/2) and float(W)/H come from the host.
__device__ __forceinline__ Ray generate_ray(const crt_camera &c, int px, int py) {
    float dx = float(px) + 0.5f, dy = float(py) + 0.5f;
    dx = dx / float(c.width);
    dy = dy / float(c.height);
    dx = (2.0f * dx) - 1.0f;
    dy = 1.0f - (2.0f * dy);
    dx = dx * c.aspect;
    dx = dx * c.tan_half_fov;
    dy = dy * c.tan_half_fov;
    V3 d = mul_rows(mk(dx, dy, -1.0f), mk(c.rotation[0], c.rotation[1], c.rotation[2]), mk(c.rotation[3], c.rotation[4], c.rotation[5]),
                    mk(c.rotation[6], c.rotation[7], c.rotation[8]));
    return Ray{ mk(c.position[0], c.position[1], c.position[2]), normalized(d) };
}

// ------------------------------------------------------------------------------------------------ triangle (a4)
// ray_intersect_triangle, src/core/crt_intersection.cpp:47-70 (the accept/reject part and the distance).
__device__ __forceinline__ bool triangle_test(const float4 &A, const float4 &B, const float4 &C, bool cull, const Ray &r, float &t_out) {
    const V3 v0 = xyz(A), v1 = xyz(B), v2 = xyz(C), N = mk(A.w, B.w, C.w);
    const float rn = dot(N, r.d);
    if (fabsf(rn) < 1e-6f)
        return false;
    const float dist = dot(N, v0 - r.o);
    if (!(dist < 0.0f) && cull)
        return false;
    const float t = dist / rn;
    if (t < 0.0f)
        return false;
    const V3 P = r.o + r.d * t;
    const V3 e0 = v1 - v0, e1 = v2 - v1, e2 = v0 - v2;
    if (dot(N, cross(e0, P - v0)) >= 0.0f && dot(N, cross(e1, P - v1)) >= 0.0f && dot(N, cross(e2, P - v2)) >= 0.0f) {
        t_out = t;
        return true;
    }
    return false;
}

// Same decisions and the same t as triangle_test, evaluated without branches (every lane of a warp runs the same
// instruction stream; rejected candidates simply yield false).  The arithmetic of each quantity is the same
// operation sequence, so accept/reject and t are bit-identical to the branching form: a quantity that the
// branching form would not have computed (e.g. t for rn ~ 0) only feeds comparisons whose result is masked.
__device__ __forceinline__ bool triangle_test_flat(const float4 &A, const float4 &B, const float4 &C, bool cull, V3 o, V3 d, float &t_out) {
    const V3 v0 = xyz(A), v1 = xyz(B), v2 = xyz(C), N = mk(A.w, B.w, C.w);
    const float rn = dot(N, d);
    const float dist = dot(N, v0 - o);
    const float t = dist / rn;
    const V3 P = o + d * t;
    const V3 e0 = v1 - v0, e1 = v2 - v1, e2 = v0 - v2;
    const float s0 = dot(N, cross(e0, P - v0)), s1 = dot(N, cross(e1, P - v1)), s2 = dot(N, cross(e2, P - v2));
    const bool ok = !(fabsf(rn) < 1e-6f) & !(!(dist < 0.0f) & cull) & !(t < 0.0f) & (s0 >= 0.0f) & (s1 >= 0.0f) & (s2 >= 0.0f);
    t_out = t;
    return ok;
}

// ------------------------------------------------------------------------------------------------ box (a3)
// ray_intersect_aabb_p, src/core/crt_intersection.cpp:14-45: six face planes, min faces first, x,y,z order,
// hit point must lie inside the face rectangle (inclusive).  NOT a slab test; kept literal for verification.
__device__ __forceinline__ bool box_test_literal(const Ray &r, V3 mn, V3 mx) {
#define CRT_FACE(PLANE, OA, DA, U, V)                                                      \
    if (!(fabsf(DA) < 1e-6f)) {                                                            \
        const float t = ((PLANE)-OA) / DA;                                                 \
        if (!(t < 0.0f)) {                                                                 \
            const V3 p = r.o + r.d * t;                                                    \
            if (p.U >= mn.U && p.U <= mx.U && p.V >= mn.V && p.V <= mx.V)                  \
                return true;                                                               \
        }                                                                                  \
    }
    CRT_FACE(mn.x, r.o.x, r.d.x, y, z)
    CRT_FACE(mn.y, r.o.y, r.d.y, z, x)
    CRT_FACE(mn.z, r.o.z, r.d.z, x, y)
    CRT_FACE(mx.x, r.o.x, r.d.x, y, z)
    CRT_FACE(mx.y, r.o.y, r.d.y, z, x)
    CRT_FACE(mx.z, r.o.z, r.d.z, x, y)
#undef CRT_FACE
    return false;
}

// ------------------------------------------------------------------------------------------------ traversal (a2)
// Reference order: ray_intersect_acceleration_tree, src/core/crt_intersection.cpp:109-136.  Stack seeded with
// node 0; pop; box test; leaf -> every triangle, strict '<' keeps the first of equal distances; inner -> push
// child0 then child1 (so child1's subtree is walked first); no pruning, no ordering.
template <bool STATS>
__device__ __forceinline__ Hit trace_literal(const DScene &S, const Ray &r, Counters *cn) {
    Hit best{ 0.0f, -1 };
    int stack[kStack];
    int sp = 0;
    stack[sp++] = 0;
    if (STATS)
        ++cn->entered; // the reference-order traversal has no entry test: every ray starts at the root node
    while (sp > 0) {
        const int node = stack[--sp];
        const float4 lo = __ldg(&S.box_nodes[2 * node]), hi = __ldg(&S.box_nodes[2 * node + 1]);
        if (STATS)
            ++cn->node_tests;
        if (!box_test_literal(r, xyz(lo), xyz(hi)))
            continue;
        const int left = __float_as_int(lo.w), right = __float_as_int(hi.w);
        if (left <= -2) { // leaf: first_ref = -left - 2, count = right
            const int first = -left - 2;
            for (int i = 0; i < right; ++i) {
                const uint32_t ref = __ldg(&S.refs[first + i]);
                const int tri = int(ref & 0x7FFFFFFFu);
                const float4 A = __ldg(&S.tri_verts[3 * tri]), B = __ldg(&S.tri_verts[3 * tri + 1]), C = __ldg(&S.tri_verts[3 * tri + 2]);
                if (STATS)
                    ++cn->tri_tests;
                float t;
                if (triangle_test(A, B, C, (ref >> 31) != 0, r, t) && (best.tri < 0 || t < best.t)) {
                    best.t = t;
                    best.tri = tri;
                }
            }
        } else {
            if (left != -1)
                stack[sp++] = left;
            if (right != -1)
                stack[sp++] = right;
        }
    }
    return best;
}

// Fast traversal.  Child boxes are the parent box cut at one plane (crt_aabb.h:24-35), so the structure is a
// k-d tree and a ray only needs its parametric interval inside the current box: no per-node box test at all.
// Front-to-back, far side pushed with its entry distance, popped entries skipped when they start beyond the best
// hit.  The reference keeps "the first triangle, in its own visiting order, among equal distances"; since a
// (ray, triangle) pair always yields the same t, that winner is reproduced with an explicit tie-break on the
// parse-order index (always the lowest index in practice).  Decisions carry a small
// slack so that grazing cases visit a SUPERSET of the reference's leaves (extra leaves cannot change the winner:
// any triangle hit found there is also referenced by the leaf that contains the hit point).
//
// The walk is a resumable state machine (KdWalker) so that the wavefront trace kernel can interleave "replace a
// finished ray" with "advance to the next leaf"; trace_kd() simply runs it to completion.
//
// Mailbox: the reference's builder references a straddling triangle from every leaf it overlaps (38 references
// per triangle in the 1 M-triangle scene), so consecutive leaves along a ray hold mostly the same triangles.
// A small direct-mapped per-ray table of already-tested triangle ids (shared memory, [slot][thread] layout =
// conflict-free) skips the repeats; a repeat test would return the same t, so results are unchanged.
// Entry of a ray into the root box (what KdWalker::init computes), as a value: used by the kernels that CREATE rays
// (primary ray generation, the shade stage) so that the trace stage starts a walk from a ready-made record.
struct RaySetup {
    V3 inv;
    float tmin, tmax;
    bool live; // false: NaN direction or the ray misses the root box -- nothing to walk
};
__device__ __forceinline__ RaySetup ray_setup(const DScene &S, const Ray &r) {
    RaySetup q;
    const V3 o = r.o, d = r.d;
    q.inv = mk(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    q.tmin = 0.0f, q.tmax = kInf;
    q.live = d.x == d.x && d.y == d.y && d.z == d.z; // NaN direction (refraction at normal incidence, crt_vector.cpp:11-27)
    const float lx = S.root_min[0], ly = S.root_min[1], lz = S.root_min[2], hx = S.root_max[0], hy = S.root_max[1], hz = S.root_max[2];
    if (d.x != 0.0f) {
        const float a = (lx - o.x) * q.inv.x, b = (hx - o.x) * q.inv.x;
        q.tmin = fmaxf(q.tmin, fminf(a, b));
        q.tmax = fminf(q.tmax, fmaxf(a, b));
    } else if (o.x < lx || o.x > hx)
        q.live = false;
    if (d.y != 0.0f) {
        const float a = (ly - o.y) * q.inv.y, b = (hy - o.y) * q.inv.y;
        q.tmin = fmaxf(q.tmin, fminf(a, b));
        q.tmax = fminf(q.tmax, fmaxf(a, b));
    } else if (o.y < ly || o.y > hy)
        q.live = false;
    if (d.z != 0.0f) {
        const float a = (lz - o.z) * q.inv.z, b = (hz - o.z) * q.inv.z;
        q.tmin = fmaxf(q.tmin, fminf(a, b));
        q.tmax = fminf(q.tmax, fmaxf(a, b));
    } else if (o.z < lz || o.z > hz)
        q.live = false;
    q.tmax = q.tmax + (fabsf(q.tmax) * 1e-5f + 1e-6f);
    if (!(q.tmin <= q.tmax))
        q.live = false;
    return q;
}

constexpr int kMailboxSlots = 16;

struct KdWalker {
    V3 o, d, inv;
    float tmin, tmax;
    int node; // node to visit next, or -1 when the walk is finished
    int sp;
    Hit best;
    int snode[kStack];
    float smin[kStack], smax[kStack];

    // false: the ray misses the root box (or has a NaN direction) and the walk is already finished
    template <bool STATS>
    __device__ __forceinline__ bool init(const DScene &S, const Ray &r, Counters *cn) {
        best = Hit{ kInf, -1 };
        node = -1;
        sp = 0;
        o = r.o, d = r.d;
        if (!(d.x == d.x && d.y == d.y && d.z == d.z))
            return false; // NaN direction (refraction at normal incidence, crt_vector.cpp:11-27): every compare is false
        inv = mk(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
        tmin = 0.0f, tmax = kInf;
        const float lx = S.root_min[0], ly = S.root_min[1], lz = S.root_min[2], hx = S.root_max[0], hy = S.root_max[1], hz = S.root_max[2];
        if (d.x != 0.0f) {
            const float a = (lx - o.x) * inv.x, b = (hx - o.x) * inv.x;
            tmin = fmaxf(tmin, fminf(a, b));
            tmax = fminf(tmax, fmaxf(a, b));
        } else if (o.x < lx || o.x > hx)
            return false;
        if (d.y != 0.0f) {
            const float a = (ly - o.y) * inv.y, b = (hy - o.y) * inv.y;
            tmin = fmaxf(tmin, fminf(a, b));
            tmax = fminf(tmax, fmaxf(a, b));
        } else if (o.y < ly || o.y > hy)
            return false;
        if (d.z != 0.0f) {
            const float a = (lz - o.z) * inv.z, b = (hz - o.z) * inv.z;
            tmin = fmaxf(tmin, fminf(a, b));
            tmax = fminf(tmax, fmaxf(a, b));
        } else if (o.z < lz || o.z > hz)
            return false;
        if (STATS)
            ++cn->node_tests;
        tmax = tmax + (fabsf(tmax) * 1e-5f + 1e-6f);
        if (tmin > tmax)
            return false;
        node = 0;
        return true;
    }

    __device__ __forceinline__ bool done() const { return node < 0; }

    // pop the next pending subtree that can still hold a closer (or equal) hit; node = -1 when none is left
    template <bool STATS>
    __device__ __forceinline__ void pop(Counters *cn) {
        for (;;) {
            if (sp == 0) {
                node = -1;
                return;
            }
            --sp;
            node = snode[sp];
            tmin = smin[sp];
            tmax = smax[sp];
            // strict pruning with slack: entries that start clearly beyond the best hit cannot beat or tie it
            if (!(tmin - (fabsf(tmin) * 1e-5f + 1e-6f) > best.t))
                break;
        }
        if (STATS)
            ++cn->node_tests;
    }

    // ONE node of the walk, given the already loaded node `n` of `node`: returns true when `n` is a leaf (the caller
    // tests n.y .. n.y + n.z and then calls pop); otherwise moves to the next node (pushing the far child, or
    // popping when a child is absent), leaving node = -1 when the walk has run out of nodes.
    template <bool STATS>
    __device__ __forceinline__ bool step(const int4 n, Counters *cn) {
        if (n.w == 3)
            return true;
        const int axis = n.w;
        const float split = __int_as_float(n.x);
        const float oa = pick(o, axis), da = pick(d, axis), ia = pick(inv, axis);
        const bool below = (oa < split) || (oa == split && da <= 0.0f);
        const int near = below ? n.y : n.z, far = below ? n.z : n.y;
        const float tp = (split - oa) * ia;
        const float slack = fabsf(tp) * 1e-5f + 1e-6f;
        int next;
        bool push = false;
        float push_min = tp;
        if (tp != tp) { // origin on the plane and direction parallel to it: the ray lies in both boxes
            push = far >= 0 && near >= 0;
            push_min = tmin;
            next = near >= 0 ? near : far;
        } else if (tp <= 0.0f || tp > tmax + slack) {
            next = near;
        } else if (tp < tmin - slack) {
            next = far;
        } else if (far >= 0 && near >= 0) {
            push = true;
            next = near;
        } else if (far >= 0) {
            next = far;
            tmin = fmaxf(tmin, tp);
        } else {
            next = near;
        }
        if (push) {
            snode[sp] = far, smin[sp] = push_min, smax[sp] = tmax, ++sp;
            if (tp == tp)
                tmax = fminf(tmax, tp);
        }
        if (next >= 0) {
            node = next;
            if (STATS)
                ++cn->node_tests;
        } else {
            pop<STATS>(cn);
        }
        return false;
    }

    // descend from `node` to the next leaf (pushing far children); returns the leaf's kd node, or w = -1 if the
    // walk ran out of nodes on the way
    template <bool STATS>
    __device__ __forceinline__ int4 descend(const DScene &S, Counters *cn) {
        for (;;) {
            const int4 n = kd_node_at(S, node);
            if (n.w == 3)
                return n;
            const int axis = n.w;
            const float split = __int_as_float(n.x);
            const float oa = pick(o, axis), da = pick(d, axis), ia = pick(inv, axis);
            const bool below = (oa < split) || (oa == split && da <= 0.0f);
            const int near = below ? n.y : n.z, far = below ? n.z : n.y;
            const float tp = (split - oa) * ia;
            const float slack = fabsf(tp) * 1e-5f + 1e-6f;
            int next;
            if (tp != tp) { // origin on the plane and direction parallel to it: the ray lies in both boxes
                if (far >= 0 && near >= 0) {
                    snode[sp] = far, smin[sp] = tmin, smax[sp] = tmax, ++sp;
                    next = near;
                } else
                    next = near >= 0 ? near : far;
            } else if (tp <= 0.0f || tp > tmax + slack) {
                next = near;
            } else if (tp < tmin - slack) {
                next = far;
            } else {
                if (far >= 0) {
                    if (near >= 0) {
                        snode[sp] = far, smin[sp] = tp, smax[sp] = tmax, ++sp;
                        next = near;
                        tmax = fminf(tmax, tp);
                    } else {
                        next = far;
                        tmin = fmaxf(tmin, tp);
                    }
                } else
                    next = near;
            }
            if (next >= 0) {
                node = next;
                if (STATS)
                    ++cn->node_tests;
            } else {
                pop<STATS>(cn);
                if (node < 0)
                    return make_int4(0, 0, 0, -1);
            }
        }
    }

    // test one leaf reference against the ray (mailbox = this thread's column of the shared table, or nullptr)
    template <bool STATS>
    __device__ __forceinline__ void test_ref(const DScene &S, uint32_t ref, Counters *cn, int *mailbox, int mailbox_stride) {
        const int tri = int(ref & 0x7FFFFFFFu);
        if (mailbox) {
            int *slot = mailbox + ((uint32_t(tri) * 0x9E3779B1u) >> 28) * mailbox_stride;
            if (*slot == tri)
                return;
            *slot = tri;
        }
        const float4 A = __ldg(&S.tri_verts[3 * tri]), B = __ldg(&S.tri_verts[3 * tri + 1]), C = __ldg(&S.tri_verts[3 * tri + 2]);
        if (STATS)
            ++cn->tri_tests;
        float t;
        if (triangle_test(A, B, C, (ref >> 31) != 0, Ray{ o, d }, t) && (t < best.t || (t == best.t && tri < best.tri))) {
            best.t = t;
            best.tri = tri;
        }
    }

    // one unit of work: descend to the next leaf, test its triangles, pop.  Call while !done().
    template <bool STATS>
    __device__ __forceinline__ void advance(const DScene &S, Counters *cn, int *mailbox, int mailbox_stride) {
        const int4 leaf = descend<STATS>(S, cn);
        if (leaf.w < 0)
            return;
        for (int i = 0; i < leaf.z; ++i)
            test_ref<STATS>(S, __ldg(&S.kd_refs[leaf.y + i]), cn, mailbox, mailbox_stride);
        pop<STATS>(cn);
    }
};

__device__ __forceinline__ void mailbox_clear(int *mailbox, int mailbox_stride) {
    if (mailbox)
        for (int k = 0; k < kMailboxSlots; ++k)
            mailbox[k * mailbox_stride] = -1;
}

template <bool STATS>
__device__ __forceinline__ Hit trace_kd(const DScene &S, const Ray &r, Counters *cn, int *mailbox = nullptr, int mailbox_stride = 0) {
    KdWalker w;
    if (w.init<STATS>(S, r, cn)) {
        if (STATS)
            ++cn->entered;
        mailbox_clear(mailbox, mailbox_stride);
        while (!w.done())
            w.advance<STATS>(S, cn, mailbox, mailbox_stride);
    }
    return w.best;
}

// Brute force over every triangle in parse order with strict '<' (ray_intersect_triangle_span over the whole
// scene, src/core/crt_intersection.cpp:95-107): config 2's checker.  Triangles are staged through shared memory
// by the whole CTA, so every thread of the block must call this (inactive pixels pass active = false).
constexpr int kBruteChunk = 256;
template <bool STATS>
__device__ __forceinline__ Hit trace_brute(const DScene &S, const Ray &r, bool active, Counters *cn, float4 *stage /* 3*kBruteChunk */) {
    Hit best{ 0.0f, -1 };
    for (int base = 0; base < S.n_triangles; base += kBruteChunk) {
        const int count = min(kBruteChunk, S.n_triangles - base);
        __syncthreads();
        for (int i = threadIdx.x; i < 3 * count; i += blockDim.x)
            stage[i] = __ldg(&S.tri_verts[3 * base + i]);
        __syncthreads();
        if (active) {
            for (int i = 0; i < count; ++i) {
                const bool cull = (__ldg(&S.tri_meta[base + i]).w & CRT_META_CULL_BIT) != 0;
                if (STATS)
                    ++cn->tri_tests;
                float t;
                if (triangle_test(stage[3 * i], stage[3 * i + 1], stage[3 * i + 2], cull, r, t) && (best.tri < 0 || t < best.t)) {
                    best.t = t;
                    best.tri = base + i;
                }
            }
        }
    }
    return best;
}

// ------------------------------------------------------------------------------------------------ surface (a4/a5)
struct Surface {
    V3 P, n;
    float u, v;     // interpolated texture coordinates (uv.x, uv.y)
    float bu, bv;   // barycentrics
    int material;
};

// The second half of ray_intersect_triangle (src/core/crt_intersection.cpp:64-88) for the winning triangle:
// barycentrics from cross-product LENGTHS, un-normalised interpolated normal, interpolated uv.
__device__ __forceinline__ Surface make_surface(const DScene &S, const Ray &r, const Hit &h) {
    Surface s;
    const float4 A = __ldg(&S.tri_verts[3 * h.tri]), B = __ldg(&S.tri_verts[3 * h.tri + 1]), C = __ldg(&S.tri_verts[3 * h.tri + 2]);
    const int4 meta = __ldg(&S.tri_meta[h.tri]);
    const V3 v0 = xyz(A), v1 = xyz(B), v2 = xyz(C);
    s.P = r.o + r.d * h.t;
    const V3 v0p = s.P - v0;
    const V3 v0v1 = v1 - v0;
    const V3 v0v2 = -(v0 - v2);
    const float bu = length(cross(v0p, v0v2)) / length(cross(v0v1, v0v2));
    const float bv = length(cross(v0v1, v0p)) / length(cross(v0v1, v0v2));
    const float bw = 1.0f - bu - bv;
    s.bu = bu;
    s.bv = bv;
    s.material = meta.w & CRT_META_MATERIAL_MASK;
    if (meta.w & CRT_META_SMOOTH_BIT) {
        const V3 n0 = xyz(__ldg(&S.vert_normal[meta.x])), n1 = xyz(__ldg(&S.vert_normal[meta.y])), n2 = xyz(__ldg(&S.vert_normal[meta.z]));
        s.n = n1 * bu + n2 * bv + n0 * bw;
    } else {
        s.n = mk(A.w, B.w, C.w);
    }
    const float2 t0 = __ldg(&S.vert_uv[meta.x]), t1 = __ldg(&S.vert_uv[meta.y]), t2 = __ldg(&S.vert_uv[meta.z]);
    s.u = t1.x * bu + t2.x * bv + t0.x * bw;
    s.v = t1.y * bu + t2.y * bv + t0.y * bw;
    return s;
}

// ------------------------------------------------------------------------------------------------ textures (a13)
// Texture::sample, src/core/crt_texture.cpp:9-49.
__device__ __forceinline__ V3 sample_texture(const DScene &S, int index, const Surface &s) {
    const float4 q0 = __ldg(&S.textures[3 * index]), q1 = __ldg(&S.textures[3 * index + 1]), q2 = __ldg(&S.textures[3 * index + 2]);
    const int type = __float_as_int(q0.x);
    const V3 c0 = mk(q1.x, q1.y, q1.z), c1 = mk(q2.x, q2.y, q2.z);
    const float param = q1.w;
    switch (type) {
        case 1: // edges
            return (s.bu <= param || s.bv <= param || (1.0f - s.bu - s.bv) <= param) ? c0 : c1;
        case 2: { // checker: C truncation toward zero
            const int row = int(s.u / param), col = int(s.v / param);
            return ((row + col) & 1) ? c1 : c0;
        }
        case 3: { // bitmap: nearest texel, C '%' wrap
            const int w = __float_as_int(q0.y), h = __float_as_int(q0.z), offset = __float_as_int(q0.w);
            int x = int(s.u * float(w)) % w;
            int y = int((1.0f - s.v) * float(h)) % h;
            // negative uv makes the reference read outside the image (undefined); we clamp instead
            x = max(0, x);
            y = max(0, y);
            return xyz(__ldg(&S.texels[offset + y * w + x]));
        }
        default: // albedo
            return c0;
    }
}

// ------------------------------------------------------------------------------------------------ shading pieces
// Diffuse direct light, src/core/crt_renderer.cpp:81-96 (the shadow query at :90-92 is dead code at HEAD:
// trace_ray_with_refractions never enters its loop, :32-33, so every light is unoccluded).
template <class OccludedFn>
__device__ __forceinline__ V3 add_direct_light(const DScene &S, V3 acc, const Surface &s, V3 albedo, OccludedFn occluded) {
    const float four_pi = 4.0f * 3.14159274101257324219f; // 4 * std::numbers::pi_v<float>, exact in binary32
    for (int i = 0; i < S.n_lights; ++i) {
        const float4 L = __ldg(&S.lights[i]);
        V3 dir = xyz(L) - s.P;
        const float r2 = dot(dir, dir);
        dir = dir / sqrtf(r2);
        const float ln = dot(dir, s.n);
        const float cos_law = (0.0f < ln) ? ln : 0.0f; // std::max(0.0f, x)
        const float area = four_pi * r2;
        if (!occluded(dir, r2))
            acc = acc + ((albedo * L.w) / area) * cos_law;
    }
    return acc;
}

// Vector::reflect, src/core/crt_vector.h:119-122: d - (n * 2) * (d . n)
__device__ __forceinline__ V3 reflect(V3 d, V3 n) { return d - (n * 2.0f) * dot(d, n); }

// Vector::refract, src/core/crt_vector.cpp:11-27; false = total internal reflection
__device__ __forceinline__ bool refract(V3 &d, V3 n, float outside_ior, float inside_ior) {
    const float cos_a = -dot(d, n);
    const float sin_a = sqrtf(1.0f - cos_a * cos_a);
    if (sin_a > inside_ior / outside_ior)
        return false;
    const float sin_b = sin_a * outside_ior / inside_ior;
    const float cos_b = sqrtf(1.0f - sin_b * sin_b);
    d = d + n * cos_a;
    d = normalized(d);
    d = d * sin_b;
    d = d + (-n) * cos_b;
    return true;
}

// the component-wise colour product with the reference's quirk (crt_vector.h:76-78): green is a.y*c.y*a.y
__device__ __forceinline__ V3 tint(V3 a, V3 c) { return V3{ a.x * c.x, a.y * c.y * a.y, a.z * c.z }; }

// 0.5f * powf(1 + d.n, 5), src/core/crt_renderer.cpp:130.  glibc's powf is correctly rounded in all but
// astronomically rare cases; x^5 evaluated in binary64 and rounded once is too, so the two agree (documented
// FP difference: a last-bit disagreement is possible in principle).
__device__ __forceinline__ float fresnel_term(float base) {
    const double b = double(base);
    const double b2 = b * b;
    return 0.5f * float(b2 * b2 * b);
}

// ------------------------------------------------------------------------------------------------ RNG (a12)
// PCG-XSH-RR 64/32 and the per-pixel seeding, src/core/crt_random.h:10-43.
struct Pcg32 {
    unsigned long long state, inc;
    __device__ __forceinline__ uint32_t next() {
        const unsigned long long old = state;
        state = old * 6364136223846793005ull + inc;
        const uint32_t xorshifted = uint32_t(((old >> 18u) ^ old) >> 27u);
        const uint32_t rot = uint32_t(old >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((0u - rot) & 31u));
    }
    __device__ __forceinline__ float uniform() { return __uint_as_float(0x3f800000u | (next() >> 9)) - 1.0f; }
};
__device__ __forceinline__ Pcg32 make_pcg(uint32_t px, uint32_t py) {
    const unsigned long long seed = ((unsigned long long)px << 32) | py;
    Pcg32 rng;
    rng.state = 0;
    rng.inc = (seed << 1) | 1ull;
    (void)rng.next();
    rng.state += seed;
    (void)rng.next();
    return rng;
}

// cosf/sinf of the reference go through glibc, which is not correctly rounded: over the 2 x 2^23 arguments the
// renderer can produce it differs from binary64 sin/cos rounded once to binary32 on 1.3 % of them, in the last bit
// (documented FP difference; only GI directions are affected, crt_renderer.cpp:68-74).
// CRT_SINCOS_LEAN (default): crt_sincos.h, the same values as the library call below, without its large-argument path
#ifndef CRT_SINCOS_LEAN
#define CRT_SINCOS_LEAN 1
#endif
__device__ __forceinline__ void sincos_ref(float a, float &s, float &c) {
#if CRT_SINCOS_LEAN
    sincos_0_2pi(a, s, c);
#else
    double sd, cd;
    sincos(double(a), &sd, &cd);
    s = float(sd);
    c = float(cd);
#endif
}

// One GI bounce direction, src/core/crt_renderer.cpp:62-74.
// CRT_GI_OUTLINE = 1: a real function, ONE copy of the two binary64 sine / cosine evaluations, the cross products and the
// normalisation instead of one per call site (the GI kernel is ~3 400 instructions and its top stall is instruction fetch).
#ifndef CRT_GI_OUTLINE
#define CRT_GI_OUTLINE 0
#endif
#if CRT_GI_OUTLINE
__device__ __noinline__ V3 gi_direction(V3 ray_dir, V3 normal, Pcg32 &rng) {
#else
__device__ __forceinline__ V3 gi_direction(V3 ray_dir, V3 normal, Pcg32 &rng) {
#endif
    const float pi = 3.14159274101257324219f;
    const V3 right = normalized(cross(ray_dir, normal));
    const V3 up = normal;
    const V3 forward = cross(right, up);
    const float a = pi * rng.uniform();
    float sa, ca;
    sincos_ref(a, sa, ca);
    V3 dir = mk(ca, sa, 0.0f);
    const float b = 2.0f * pi * rng.uniform();
    float sb, cb;
    sincos_ref(b, sb, cb);
    dir = mul_rows(dir, mk(cb, 0.0f, -sb), mk(0.0f, 1.0f, 0.0f), mk(sb, 0.0f, cb)); // Matrix::rotation_y, crt_matrix.cpp:15-21
    dir = mul_rows(dir, right, up, forward);                                           // Matrix::from_axes, crt_matrix.h:25-31
    return dir;
}

// ------------------------------------------------------------------------------------------------ output (a15)
// write_ppm's quantisation, src/core/crt_image_ppm.cpp:16-19: c * 255, C cast (x86 cvttss2si: out-of-range and
// NaN give INT_MIN), clamp to [0, 255].
__device__ __forceinline__ uint8_t quantize(float c) {
    const float v = c * 255.0f;
    int q;
    if (!(v < 2147483648.0f) || v < -2147483648.0f)
        q = -2147483647 - 1;
    else
        q = int(v);
    return uint8_t(min(255, max(0, q)));
}

// image row owned by local row `ly` under band ownership (crt_render_params.rank/world/band_rows)
__device__ __forceinline__ int global_row(const DParams &p, int ly) {
    if (p.world <= 1)
        return ly;
    const int local_band = ly / p.band_rows;
    return (local_band * p.world + p.rank) * p.band_rows + (ly - local_band * p.band_rows);
}

} // namespace crtdev
