#include "crt_kdbuild.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <limits>
#include <thread>

namespace crt {

namespace {

constexpr float kInf = std::numeric_limits<float>::infinity();

float env_float(const char *name, float fallback) {
    const char *v = std::getenv(name);
    return v && *v ? float(std::atof(v)) : fallback;
}

struct TriBox {
    float lo[3], hi[3];
};

struct Cell {
    float lo[3], hi[3];
};

struct Split {
    int axis{ -1 };
    float plane{ 0.0f };
    float cost{ kInf };
    bool valid() const { return axis >= 0; }
};

struct Subtree {
    std::vector<crt_kd_node> nodes; // indices relative to this subtree
    std::vector<uint32_t> refs;
    int max_depth{ 0 }, leaves{ 0 }, max_leaf{ 0 };
};

struct Builder {
    const std::vector<TriBox> &boxes;
    const uint8_t *cull;
    KdBuildParams p;

    float leaf_cost(size_t n) const { return n ? p.c_leaf + p.c_tri * float(n) : 0.0f; }

    // expected cost of cutting `cell` at `plane`: one node step plus the children weighted by the chance that a ray
    // which crosses the cell crosses the child (ratio of surface areas)
    float split_cost(const Cell &cell, int axis, float plane, size_t nl, size_t nr) const {
        const int u = (axis + 1) % 3, v = (axis + 2) % 3;
        const float du = cell.hi[u] - cell.lo[u], dv = cell.hi[v] - cell.lo[v], da = cell.hi[axis] - cell.lo[axis];
        const float whole = du * dv + da * (du + dv);
        if (!(whole > 0.0f))
            return kInf;
        const float pl = (du * dv + (plane - cell.lo[axis]) * (du + dv)) / whole;
        const float pr = (du * dv + (cell.hi[axis] - plane) * (du + dv)) / whole;
        float c = p.c_node + pl * leaf_cost(nl) + pr * leaf_cost(nr);
        if (nl == 0 || nr == 0)
            c *= p.empty_bonus;
        return c;
    }

    // every triangle bound inside the cell is a candidate plane (exact counts under the membership rule)
    void sweep_exact(const Cell &cell, const std::vector<uint32_t> &refs, int axis, Split &best) const {
        const size_t n = refs.size();
        const float clo = cell.lo[axis], chi = cell.hi[axis];
        // scratch per thread: this runs three times per small node, ~4 M times for a 1 M-triangle scene
        static thread_local std::vector<float> los, his, flat, cand;
        los.resize(n), his.resize(n), flat.clear();
        for (size_t i = 0; i < n; ++i) {
            const TriBox &b = boxes[refs[i]];
            los[i] = std::max(b.lo[axis], clo);
            his[i] = std::min(b.hi[axis], chi);
            if (b.lo[axis] == b.hi[axis])
                flat.push_back(b.lo[axis]);
        }
        cand.assign(los.begin(), los.end());
        cand.insert(cand.end(), his.begin(), his.end());
        std::sort(los.begin(), los.end());
        std::sort(his.begin(), his.end());
        std::sort(flat.begin(), flat.end());
        std::sort(cand.begin(), cand.end());
        cand.erase(std::unique(cand.begin(), cand.end()), cand.end());
        for (float plane : cand) {
            if (!(plane > clo && plane < chi))
                continue;
            const auto in_plane = std::equal_range(flat.begin(), flat.end(), plane);
            const size_t both = size_t(in_plane.second - in_plane.first);
            const size_t nl = size_t(std::lower_bound(los.begin(), los.end(), plane) - los.begin()) + both;
            const size_t nr = size_t(his.end() - std::upper_bound(his.begin(), his.end(), plane)) + both;
            const float c = split_cost(cell, axis, plane, nl, nr);
            if (c < best.cost && !(nl == n && nr == n))
                best = Split{ axis, plane, c };
        }
    }

    // large nodes: counts per bin of where the triangles' extents start and end; planes at the bin walls
    void sweep_binned(const Cell &cell, const std::vector<uint32_t> &refs, int axis, Split &best) const {
        const size_t n = refs.size();
        const int K = p.bins;
        const float clo = cell.lo[axis], chi = cell.hi[axis], da = chi - clo;
        if (!(da > 0.0f))
            return;
        std::vector<uint32_t> starts(size_t(K), 0), ends(size_t(K), 0);
        const float scale = float(K) / da;
        for (size_t i = 0; i < n; ++i) {
            const TriBox &b = boxes[refs[i]];
            const int s = std::clamp(int((std::max(b.lo[axis], clo) - clo) * scale), 0, K - 1);
            const int e = std::clamp(int((std::min(b.hi[axis], chi) - clo) * scale), 0, K - 1);
            ++starts[size_t(s)], ++ends[size_t(e)];
        }
        size_t nl = 0, ended = 0;
        for (int k = 1; k < K; ++k) {
            nl += starts[size_t(k - 1)];
            ended += ends[size_t(k - 1)];
            const size_t nr = n - ended;
            const float plane = clo + da * (float(k) / float(K));
            if (!(plane > clo && plane < chi))
                continue;
            const float c = split_cost(cell, axis, plane, nl, nr);
            if (c < best.cost && !(nl == n && nr == n))
                best = Split{ axis, plane, c };
        }
    }

    Split find_split(const Cell &cell, const std::vector<uint32_t> &refs, int depth) const {
        Split best;
        const size_t n = refs.size();
        if (n == 0 || depth >= p.max_depth)
            return best;
        for (int axis = 0; axis < 3; ++axis) {
            if (n <= size_t(p.exact_below))
                sweep_exact(cell, refs, axis, best);
            else
                sweep_binned(cell, refs, axis, best);
        }
        // a split has to pay for itself, except that very large nodes are always cut while a cut makes progress
        if (best.valid() && best.cost >= leaf_cost(n) && n <= 64)
            best = Split{};
        return best;
    }

    void partition(const std::vector<uint32_t> &refs, const Split &s, std::vector<uint32_t> &left, std::vector<uint32_t> &right) const {
        left.clear(), right.clear();
        for (uint32_t t : refs) {
            const float lo = boxes[t].lo[s.axis], hi = boxes[t].hi[s.axis];
            const bool in_plane = lo == s.plane && hi == s.plane;
            if (lo < s.plane || in_plane)
                left.push_back(t);
            if (hi > s.plane || in_plane)
                right.push_back(t);
        }
    }

    void emit_leaf(Subtree &out, int self, const std::vector<uint32_t> &refs) const {
        const int first = int(out.refs.size());
        for (uint32_t t : refs)
            out.refs.push_back(t | (cull && cull[t] ? CRT_REF_CULL_BIT : 0u));
        out.nodes[size_t(self)] = crt_kd_node{ 0.0f, first, int32_t(refs.size()), 3 };
        ++out.leaves;
        out.max_leaf = std::max(out.max_leaf, int(refs.size()));
    }

    // appends the subtree of (cell, refs) to `out` in DFS pre-order; returns its root's index within `out`
    int build_serial(Subtree &out, const Cell &cell, std::vector<uint32_t> &refs, int depth) const {
        const int self = int(out.nodes.size());
        out.nodes.push_back(crt_kd_node{ 0.0f, -1, -1, 0 });
        out.max_depth = std::max(out.max_depth, depth);
        const Split s = find_split(cell, refs, depth);
        if (!s.valid()) {
            if (!refs.empty())
                emit_leaf(out, self, refs);
            return self;
        }
        std::vector<uint32_t> left, right;
        partition(refs, s, left, right);
        std::vector<uint32_t>().swap(refs);
        Cell cl = cell, cr = cell;
        cl.hi[s.axis] = s.plane, cr.lo[s.axis] = s.plane;
        const int a = left.empty() ? -1 : build_serial(out, cl, left, depth + 1);
        const int b = right.empty() ? -1 : build_serial(out, cr, right, depth + 1);
        out.nodes[size_t(self)] = crt_kd_node{ s.plane, a, b, s.axis };
        return self;
    }

    static void append(Subtree &dst, const Subtree &src, int &root_index) {
        const int node_off = int(dst.nodes.size()), ref_off = int(dst.refs.size());
        root_index = node_off;
        for (crt_kd_node n : src.nodes) {
            if (n.axis == 3) {
                n.a += ref_off;
            } else {
                if (n.a >= 0)
                    n.a += node_off;
                if (n.b >= 0)
                    n.b += node_off;
            }
            dst.nodes.push_back(n);
        }
        dst.refs.insert(dst.refs.end(), src.refs.begin(), src.refs.end());
        dst.max_depth = std::max(dst.max_depth, src.max_depth);
        dst.leaves += src.leaves;
        dst.max_leaf = std::max(dst.max_leaf, src.max_leaf);
    }

    // fn(chunk, begin, end) over `workers` contiguous ranges of [0, n), the last one on the calling thread
    template <class Fn>
    static void for_chunks(size_t n, int workers, Fn fn) {
        std::vector<std::thread> pool;
        for (int c = 0; c < workers; ++c) {
            const size_t begin = n * size_t(c) / size_t(workers), end = n * size_t(c + 1) / size_t(workers);
            if (c + 1 < workers)
                pool.emplace_back([=] { fn(c, begin, end); });
            else
                fn(c, begin, end);
        }
        for (std::thread &t : pool)
            t.join();
    }

    // find_split for the large nodes at the top of the tree, with the threads that have nothing else to do yet: the
    // three axes are binned in one pass over the references, each worker counting its own range; the counts are
    // integers, so the sums -- and with them the chosen plane -- are the serial ones
    Split find_split_wide(const Cell &cell, const std::vector<uint32_t> &refs, int depth, int workers) const {
        Split best;
        const size_t n = refs.size();
        if (n == 0 || depth >= p.max_depth)
            return best;
        const int K = p.bins;
        std::vector<uint32_t> counts(size_t(workers) * 6 * size_t(K), 0); // [worker][axis][starts | ends][bin]
        float scale[3];
        for (int a = 0; a < 3; ++a)
            scale[a] = float(K) / (cell.hi[a] - cell.lo[a]);
        for_chunks(n, workers, [&](int c, size_t begin, size_t end) {
            uint32_t *mine = counts.data() + size_t(c) * 6 * size_t(K);
            for (size_t i = begin; i < end; ++i) {
                const TriBox &b = boxes[refs[i]];
                for (int a = 0; a < 3; ++a) {
                    const float clo = cell.lo[a], chi = cell.hi[a];
                    if (!(chi - clo > 0.0f))
                        continue;
                    const int st = std::clamp(int((std::max(b.lo[a], clo) - clo) * scale[a]), 0, K - 1);
                    const int en = std::clamp(int((std::min(b.hi[a], chi) - clo) * scale[a]), 0, K - 1);
                    ++mine[size_t(2 * a) * size_t(K) + size_t(st)], ++mine[size_t(2 * a + 1) * size_t(K) + size_t(en)];
                }
            }
        });
        for (int c = 1; c < workers; ++c)
            for (size_t k = 0; k < 6 * size_t(K); ++k)
                counts[k] += counts[size_t(c) * 6 * size_t(K) + k];
        for (int axis = 0; axis < 3; ++axis) {
            const float clo = cell.lo[axis], chi = cell.hi[axis], da = chi - clo;
            if (!(da > 0.0f))
                continue;
            const uint32_t *starts = counts.data() + size_t(2 * axis) * size_t(K), *ends = starts + K;
            size_t nl = 0, ended = 0;
            for (int k = 1; k < K; ++k) { // as in sweep_binned
                nl += starts[k - 1];
                ended += ends[k - 1];
                const size_t nr = n - ended;
                const float plane = clo + da * (float(k) / float(K));
                if (!(plane > clo && plane < chi))
                    continue;
                const float c = split_cost(cell, axis, plane, nl, nr);
                if (c < best.cost && !(nl == n && nr == n))
                    best = Split{ axis, plane, c };
            }
        }
        if (best.valid() && best.cost >= leaf_cost(n) && n <= 64)
            best = Split{};
        return best;
    }

    // partition() with every worker filling its own stretch of the two lists: same lists, same order
    void partition_wide(const std::vector<uint32_t> &refs, const Split &s, std::vector<uint32_t> &left, std::vector<uint32_t> &right, int workers) const {
        const auto sides = [&](uint32_t t, bool &l, bool &r) {
            const float lo = boxes[t].lo[s.axis], hi = boxes[t].hi[s.axis];
            const bool in_plane = lo == s.plane && hi == s.plane;
            l = lo < s.plane || in_plane, r = hi > s.plane || in_plane;
        };
        std::vector<size_t> nl(size_t(workers) + 1, 0), nr(size_t(workers) + 1, 0);
        for_chunks(refs.size(), workers, [&](int c, size_t begin, size_t end) {
            size_t a = 0, b = 0;
            for (size_t i = begin; i < end; ++i) {
                bool l, r;
                sides(refs[i], l, r);
                a += l, b += r;
            }
            nl[size_t(c) + 1] = a, nr[size_t(c) + 1] = b;
        });
        for (int c = 0; c < workers; ++c)
            nl[size_t(c) + 1] += nl[size_t(c)], nr[size_t(c) + 1] += nr[size_t(c)];
        left.resize(nl[size_t(workers)]), right.resize(nr[size_t(workers)]);
        for_chunks(refs.size(), workers, [&](int c, size_t begin, size_t end) {
            size_t a = nl[size_t(c)], b = nr[size_t(c)];
            for (size_t i = begin; i < end; ++i) {
                bool l, r;
                sides(refs[i], l, r);
                if (l)
                    left[a++] = refs[i];
                if (r)
                    right[b++] = refs[i];
            }
        });
    }

    // top of the tree: the two children of a large node are built concurrently, and while there are fewer such nodes
    // than threads (`workers` of them are free at this node) the node's own binning and partition are spread out too
    Subtree build_parallel(const Cell &cell, std::vector<uint32_t> &refs, int depth, int fork_levels, int workers) const {
        Subtree out;
        if (fork_levels <= 0 || refs.size() < 16384) {
            build_serial(out, cell, refs, depth);
            return out;
        }
        const bool wide = workers > 1 && refs.size() > size_t(p.exact_below);
        const Split s = wide ? find_split_wide(cell, refs, depth, workers) : find_split(cell, refs, depth);
        if (!s.valid()) {
            build_serial(out, cell, refs, depth);
            return out;
        }
        std::vector<uint32_t> left, right;
        if (wide)
            partition_wide(refs, s, left, right, workers);
        else
            partition(refs, s, left, right);
        std::vector<uint32_t>().swap(refs);
        Cell cl = cell, cr = cell;
        cl.hi[s.axis] = s.plane, cr.lo[s.axis] = s.plane;
        Subtree sub_l, sub_r;
        std::thread worker([&] {
            if (!left.empty())
                sub_l = build_parallel(cl, left, depth + 1, fork_levels - 1, (workers + 1) / 2);
        });
        if (!right.empty())
            sub_r = build_parallel(cr, right, depth + 1, fork_levels - 1, std::max(1, workers / 2));
        worker.join();
        out.nodes.push_back(crt_kd_node{ s.plane, -1, -1, s.axis });
        out.max_depth = depth;
        int a = -1, b = -1;
        if (!sub_l.nodes.empty())
            append(out, sub_l, a);
        if (!sub_r.nodes.empty())
            append(out, sub_r, b);
        out.nodes[0].a = a, out.nodes[0].b = b;
        return out;
    }
};

} // namespace

KdBuildParams KdBuildParams::from_environment() {
    KdBuildParams p;
    p.c_node = env_float("CRT_SAH_NODE", p.c_node);
    p.c_leaf = env_float("CRT_SAH_LEAF", p.c_leaf);
    p.c_tri = env_float("CRT_SAH_TRI", p.c_tri);
    p.empty_bonus = env_float("CRT_SAH_EMPTY", p.empty_bonus);
    p.max_depth = std::clamp(int(env_float("CRT_SAH_DEPTH", float(p.max_depth))), 1, 44);
    p.exact_below = std::max(1, int(env_float("CRT_SAH_EXACT", float(p.exact_below))));
    p.bins = std::clamp(int(env_float("CRT_SAH_BINS", float(p.bins))), 4, 1024);
    p.threads = int(env_float("CRT_SAH_THREADS", 0.0f));
    return p;
}

KdBuildStats build_sah_tree(const crt_float4 *tri_verts, size_t n_triangles, const uint8_t *cull, const float root_min[3], const float root_max[3],
                            const KdBuildParams &params, std::vector<crt_kd_node> &nodes, std::vector<uint32_t> &refs) {
    std::vector<TriBox> boxes(n_triangles);
    for (size_t i = 0; i < n_triangles; ++i) {
        const crt_float4 *v = tri_verts + 3 * i;
        const float xs[3][3] = { { v[0].x, v[1].x, v[2].x }, { v[0].y, v[1].y, v[2].y }, { v[0].z, v[1].z, v[2].z } };
        for (int a = 0; a < 3; ++a) {
            boxes[i].lo[a] = std::min(xs[a][0], std::min(xs[a][1], xs[a][2]));
            boxes[i].hi[a] = std::max(xs[a][0], std::max(xs[a][1], xs[a][2]));
        }
    }
    Cell root{};
    for (int a = 0; a < 3; ++a)
        root.lo[a] = root_min[a], root.hi[a] = root_max[a];
    std::vector<uint32_t> all(n_triangles);
    for (size_t i = 0; i < n_triangles; ++i)
        all[i] = uint32_t(i);

    const Builder builder{ boxes, cull, params };
    int threads = params.threads > 0 ? params.threads : int(std::thread::hardware_concurrency());
    int fork_levels = 0;
    while ((1 << fork_levels) < std::max(1, threads) * 2 && fork_levels < 8)
        ++fork_levels;
    Subtree tree = builder.build_parallel(root, all, 0, threads > 1 ? fork_levels : 0, std::max(1, threads));
    nodes = std::move(tree.nodes);
    refs = std::move(tree.refs);
    KdBuildStats stats;
    stats.nodes = int(nodes.size());
    stats.leaves = tree.leaves;
    stats.refs = int(refs.size());
    stats.max_depth = tree.max_depth;
    stats.max_leaf = tree.max_leaf;
    return stats;
}

} // namespace crt
