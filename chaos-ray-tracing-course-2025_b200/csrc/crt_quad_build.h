// Host side of crt_walk_quad.h: folds the binary walk tree (crt_flat_scene.fast_nodes, core/crt_kdbuild.cpp) into quad
// nodes -- a node together with its two children -- and a leaf table.  Pure re-arrangement: every split plane, every
// leaf list and the order of the children are the binary tree's; used by crt_cuda_upload_scene and by the CPU harness
// of the walker (tests/host/walk_lean_host.cpp).
#pragma once

#include <cstdint>
#include <cstring>
#include <utility>
#include <vector>

#include "../../include/crt_cuda.h"
#include "crt_walk_quad.h"

namespace crtquad {

struct LeafRec {
    int32_t first, count;
};

// Returns NULL, or what is wrong with the tree.  quads[0] is the root; leaves get one extra, empty record at the end
// (what a stolen stack entry is replaced with, crt_trace_lean.cuh).  Node order: depth first, the subtree below the
// plane before the one above it, like the builder's own numbering.
inline const char *build_quads(const crt_kd_node *nodes, int n, std::vector<Quad> &quads, std::vector<LeafRec> &leaves, int *max_pending = nullptr) {
    quads.clear(), leaves.clear();
    if (!nodes || n < 1)
        return "walk tree is empty";
    auto bits = [](float f) {
        int32_t u;
        std::memcpy(&u, &f, 4);
        return u;
    };
    auto ok = [&](int i) { return i >= -1 && i < n; };
    auto add_leaf = [&](const crt_kd_node &k) {
        leaves.push_back({ k.a, k.b });
        return leaf_ref(int(leaves.size()) - 1);
    };
    struct Todo {
        int node, quad, depth;
    };
    std::vector<Todo> todo;
    quads.push_back(Quad{});
    int deepest = 0;
    if (nodes[0].axis == 3) { // the whole tree is one leaf: a quad without planes holding it
        Quad &q = quads[0];
        q.meta = 3 | 3 << 2 | 3 << 4;
        q.c[0] = add_leaf(nodes[0]), q.c[1] = q.c[2] = q.c[3] = -1;
    } else {
        todo.push_back({ 0, 0, 1 });
    }
    size_t visited = 0;
    while (!todo.empty()) {
        const Todo t = todo.back();
        todo.pop_back();
        if (++visited > size_t(n))
            return "walk tree is not a tree (a node is reachable twice)";
        deepest = t.depth > deepest ? t.depth : deepest;
        const crt_kd_node &r = nodes[t.node];
        if (r.axis < 0 || r.axis > 2 || !ok(r.a) || !ok(r.b))
            return "walk tree node with a broken axis or child index";
        Quad q{};
        q.split_r = bits(r.split);
        q.meta = r.axis;
        std::pair<int, int> later[4];
        int n_later = 0;
        for (int side = 0; side < 2; ++side) {
            const int s = side == 0 ? r.a : r.b;
            int axis = 3, c0 = -1, c1 = -1;
            int32_t split = 0;
            if (s >= 0 && nodes[s].axis == 3) {
                c0 = add_leaf(nodes[s]);
            } else if (s >= 0) {
                const crt_kd_node &k = nodes[s];
                if (k.axis < 0 || k.axis > 2 || !ok(k.a) || !ok(k.b))
                    return "walk tree node with a broken axis or child index";
                ++visited;
                axis = k.axis, split = bits(k.split);
                int *slot[2] = { &c0, &c1 };
                for (int g = 0; g < 2; ++g) {
                    const int child = g == 0 ? k.a : k.b;
                    if (child < 0)
                        continue;
                    if (nodes[child].axis == 3) {
                        *slot[g] = add_leaf(nodes[child]);
                    } else {
                        *slot[g] = int(quads.size());
                        quads.push_back(Quad{});
                        later[n_later++] = { child, *slot[g] };
                    }
                }
            }
            (side == 0 ? q.split_a : q.split_b) = split;
            q.meta |= axis << (side == 0 ? 2 : 4);
            q.c[2 * side] = c0, q.c[2 * side + 1] = c1;
        }
        quads[t.quad] = q;
        for (int k = n_later - 1; k >= 0; --k) // depth first, first grandchild next
            todo.push_back({ later[k].first, later[k].second, t.depth + 1 });
    }
    leaves.push_back({ 0, 0 }); // the empty leaf
    if (max_pending)
        *max_pending = 3 * deepest + 2; // stack entries a walk can have pending: three per quad level, and the sentinel
    return nullptr;
}

} // namespace crtquad
