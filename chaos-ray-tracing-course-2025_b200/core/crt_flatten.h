#pragma once
// Scene -> 16-byte-aligned POD arrays (include/crt_cuda.h: crt_flat_scene).  This is the "scene upload and tree
// flattening" subsystem: it has no counterpart in the reference, whose renderer walks std::vector<Node> with
// per-node std::vector<Triangle> (src/core/crt_acceleration_tree.h:15-26).
#include <vector>

#include "../../include/crt_cuda.h"
#include "crt_scene.h"

namespace crt {

struct FlatScene {
    std::vector<crt_float4> tri_verts;
    std::vector<int32_t> tri_meta;
    std::vector<crt_float4> vert_normal;
    std::vector<float> vert_uv;
    std::vector<crt_box_node> box_nodes;
    std::vector<crt_kd_node> kd_nodes;
    std::vector<uint32_t> refs;
    std::vector<crt_kd_node> fast_nodes; // the tree the k-d walk uses: same topology, subtrees merged by cost
    std::vector<uint32_t> fast_refs;     // its leaf lists, identical lists stored once
    std::vector<crt_float4> lights;
    std::vector<crt_material> materials;
    std::vector<crt_texture> textures;
    std::vector<crt_float4> texels;
    crt_flat_scene view{}; // pointers into the vectors above
};

crt_camera flatten_camera(const Camera &camera);
// with_reference_tree = false leaves box_nodes / kd_nodes / refs empty (n_nodes == 0): enough for CRT_TRAVERSAL_KD, and
// it skips the build and the upload of the reference's tree (6.6 M nodes / 38 M references for the 1 M-triangle
// scene); $CRT_WALK_TREE=reference needs that tree and forces it.  add_reference_tree fills it in later without
// touching the other arrays.
void flatten_scene(const Scene &scene, FlatScene &out, bool with_reference_tree = true);
// only what shading reads and a caller may change between two render_image calls: lights, materials, texture records,
// background, the gi / reflections / refractions flags, bucket_size (bitmap texels stay as they are unless with_texels)
void flatten_shading(const Scene &scene, FlatScene &out, bool with_texels = false);
void add_reference_tree(const Scene &scene, FlatScene &out);

// Derives the walk tree from the reference tree (crt_flatten.cpp); exposed for tests.
struct FastTreeStats {
    int nodes{ 0 }, leaves{ 0 }, refs{ 0 }, max_depth{ 0 }, max_leaf{ 0 };
};
FastTreeStats build_fast_tree(const AccelerationTree &tree, const std::vector<Triangle> &triangles, FlatScene &out);

} // namespace crt
