// Host build of the reference's acceleration tree, emitted directly in flat form.
//
// The RULES are the reference's (src/core/crt_acceleration_tree.cpp:31-106, crt_aabb.h:17-45) because the tree is
// an input of the parity contract: same boxes, same membership, same node numbering, same in-leaf order.
//   * root box  = min/max over every triangle vertex, starting from the empty box (:89-94, :13-23)
//   * leaf      if depth > 39 or count <= 16 (:32)
//   * cut       axis = depth % 3 at (min + max) * 0.5f (:38, crt_aabb.h:30)
//   * member    triangle AABB overlaps the child box, inclusive; straddlers go to both children (:44-58)
//   * order     relative order is preserved; child0's whole subtree is numbered before child1's node (:63-84)
//   * an empty child is absent (-1)
// The MECHANICS are ours: triangles are indices into one array (no per-leaf copies: the reference duplicates a
// 48 B Triangle per straddle, 38 M copies for the 1 M-triangle scene), per-triangle boxes are computed once,
// and leaves append to a single reference array that is uploaded as is.

#include <algorithm>

#include "crt_scene.h"

namespace crt::acceleration_tree {

namespace {

struct Builder {
    const std::vector<AABB> &boxes;
    AccelerationTree &tree;

    void branch(int node_index, std::vector<uint32_t> &&members, int depth) {
        tree.max_depth = std::max(tree.max_depth, depth);
        if (depth > MAX_ACCELERATION_TREE_DEPTH || members.size() <= size_t(MAX_BOX_TRIANGLE_COUNT)) {
            AccelerationTreeNode &leaf = tree.nodes[size_t(node_index)];
            leaf.first_reference = int(tree.references.size());
            leaf.reference_count = int(members.size());
            tree.references.insert(tree.references.end(), members.begin(), members.end());
            return;
        }

        const auto [lo_box, hi_box] = tree.nodes[size_t(node_index)].bounds.split(unsigned(depth % 3));

        // stable partition in place: `members` keeps the low side, `hi` receives the high side
        std::vector<uint32_t> hi;
        hi.reserve(members.size() / 2);
        size_t kept = 0;
        for (size_t i = 0; i < members.size(); ++i) {
            const uint32_t t = members[i];
            const bool in_lo = lo_box.intersects(boxes[t]);
            const bool in_hi = hi_box.intersects(boxes[t]);
            if (in_hi)
                hi.push_back(t);
            if (in_lo)
                members[kept++] = t;
        }
        members.resize(kept);

        if (!members.empty()) {
            const int child = int(tree.nodes.size());
            tree.nodes.push_back(AccelerationTreeNode{ lo_box, { -1, -1 }, node_index, 0, 0, depth + 1 });
            tree.nodes[size_t(node_index)].children_indices[0] = child;
            branch(child, std::move(members), depth + 1);
        }
        std::vector<uint32_t>().swap(members); // release before descending into the other side
        if (!hi.empty()) {
            const int child = int(tree.nodes.size());
            tree.nodes.push_back(AccelerationTreeNode{ hi_box, { -1, -1 }, node_index, 0, 0, depth + 1 });
            tree.nodes[size_t(node_index)].children_indices[1] = child;
            branch(child, std::move(hi), depth + 1);
        }
    }
};

void grow(AABB &box, const Vector &p) {
    box.min.x = std::min(box.min.x, p.x);
    box.min.y = std::min(box.min.y, p.y);
    box.min.z = std::min(box.min.z, p.z);
    box.max.x = std::max(box.max.x, p.x);
    box.max.y = std::max(box.max.y, p.y);
    box.max.z = std::max(box.max.z, p.z);
}

} // namespace

AccelerationTree build(const std::vector<Vertex> &vertices, const std::vector<Triangle> &triangles) {
    std::vector<AABB> boxes(triangles.size(), AABB::vacuum());
    AABB root = AABB::vacuum();
    for (size_t i = 0; i < triangles.size(); ++i) {
        const Triangle &t = triangles[i];
        for (int v : { t.v0, t.v1, t.v2 }) {
            grow(boxes[i], vertices[size_t(v)].position);
            grow(root, vertices[size_t(v)].position);
        }
    }

    AccelerationTree tree;
    tree.nodes.push_back(AccelerationTreeNode{ root, { -1, -1 }, -1, 0, 0, 0 });
    std::vector<uint32_t> all(triangles.size());
    for (size_t i = 0; i < all.size(); ++i)
        all[i] = uint32_t(i);
    Builder{ boxes, tree }.branch(0, std::move(all), 0);
    return tree;
}

} // namespace crt::acceleration_tree
