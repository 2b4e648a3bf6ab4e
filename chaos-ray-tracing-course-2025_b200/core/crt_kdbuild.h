#pragma once
// Surface-area-heuristic k-d tree over the scene's triangles: the tree CRT_TRAVERSAL_KD walks on the device.
// No counterpart in the reference, whose builder (src/core/crt_acceleration_tree.cpp:31-85) cuts at the spatial
// median of axis depth % 3 until <= 16 triangles remain; the closest hit does not depend on how space is cut (the
// walk keeps min (t, parse index) over every candidate it tests), so the device may walk a better tree
// (SURVEY.md §8 f1).  The reference's own tree is still built and uploaded node for node: CRT_TRAVERSAL_LITERAL
// walks it, and the parity tests compare the two traversals.
#include <cstdint>
#include <vector>

#include "../../include/crt_cuda.h"

namespace crt {

struct KdBuildStats {
    int nodes{ 0 }, leaves{ 0 }, refs{ 0 }, max_depth{ 0 }, max_leaf{ 0 };
};

struct KdBuildParams {
    float c_node{ 1.0f };  // cost of one node step of the trace kernel (the unit)
    float c_leaf{ 1.0f };  // fixed cost of parking a leaf for the next test phase
    float c_tri{ 1.0f };   // cost per tested triangle.  Swept on B200 with every generation of the kernels (tools/gpu_sah.sh,
                           // profiles/r2p_*, r3g_*, r3h_*): node visits are what the walk pays for, tests run warp-wide --
                           // 1.5 -> 1.0 (r2p): C4 2.82 -> 2.77 ms, GI 156.4 -> 154.8 ms.  0.75 with the 58-instruction
                           // visit: C4 -1 % (r3g), but C3 0.61 -> 0.71 ms and C5 0.31 -> 0.35 ms (r3h): the packet kernel of
                           // the primary level tests a leaf's triangles one after the other -- stays 1.0
    float empty_bonus{ 1.0f }; // < 1 favours splits that cut away empty space; the sweep on B200 found no gain from it
    int max_depth{ 44 };   // device stack: fast_max_depth + 2 <= kStack (48)
    int exact_below{ 96 }; // nodes with at most this many triangles sweep every triangle bound; larger ones use bins
    int bins{ 64 };
    int threads{ 0 };      // 0 = hardware_concurrency
    static KdBuildParams from_environment();
};

// tri_verts: 3 float4 per triangle ({v0,.}{v1,.}{v2,.}); cull[i] != 0 sets CRT_REF_CULL_BIT on triangle i's references.
// Membership: a triangle is listed in every leaf whose cell its bounding box reaches into (touching a cell wall from
// outside does not count; a triangle lying IN a split plane goes to both sides).  Leaf lists ascend by triangle index.
KdBuildStats build_sah_tree(const crt_float4 *tri_verts, size_t n_triangles, const uint8_t *cull, const float root_min[3], const float root_max[3],
                            const KdBuildParams &params, std::vector<crt_kd_node> &nodes, std::vector<uint32_t> &refs);

} // namespace crt
