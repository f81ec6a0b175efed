// bvh_build.cpp — BLAS construction that reproduces the reference's tree exactly
// (rayca-soft/src/bvh/blas.rs:64-127,245-316) without its O(189 * N) rescans per node.
//
// The reference evaluates, for each of 3 axes and 63 plane positions, a full pass over the
// node's primitives: "centroid[axis] < pos ? grow left : grow right".  Because the 63 positions
// of an axis are non-decreasing, the set of candidates for which a primitive falls LEFT is a
// suffix {k, ..., 63}; k is found by binary search with the very same `<` comparisons.  Binning
// primitives by k and taking prefix/suffix unions of the bins yields, for every candidate, the
// identical count and the identical box (min/max are exact and associative), hence the identical
// f32 cost, the identical winning plane and the identical in-place partition.
#include "host_scene.h"

#include <algorithm>
#include <cstring>

#ifdef _OPENMP
#include <omp.h>
#endif

namespace rb {
namespace {

constexpr int kPlanes = 64; // AREA_COUNT (blas.rs:108)

struct Box {
    float mn[3], mx[3];
};
inline void box_origin(Box& b) { // AABB::default(): both corners at the origin (blas.rs:66-67, A-21)
    for (int k = 0; k < 3; ++k) b.mn[k] = b.mx[k] = 0.0f;
}
inline void box_grow(Box& b, const V3& lo, const V3& hi) {
    b.mn[0] = fminf(b.mn[0], lo.x); b.mn[1] = fminf(b.mn[1], lo.y); b.mn[2] = fminf(b.mn[2], lo.z);
    b.mx[0] = fmaxf(b.mx[0], hi.x); b.mx[1] = fmaxf(b.mx[1], hi.y); b.mx[2] = fmaxf(b.mx[2], hi.z);
}
inline void box_merge(Box& b, const Box& o) {
    for (int k = 0; k < 3; ++k) {
        b.mn[k] = fminf(b.mn[k], o.mn[k]);
        b.mx[k] = fmaxf(b.mx[k], o.mx[k]);
    }
}
// aabb.rs:20-23
inline float box_area(const Box& b) {
    float ex = b.mx[0] - b.mn[0], ey = b.mx[1] - b.mn[1], ez = b.mx[2] - b.mn[2];
    return ex * ey + ey * ez + ez * ex;
}
inline float node_area(const HostNode& n) {
    float ex = n.bmax.x - n.bmin.x, ey = n.bmax.y - n.bmin.y, ez = n.bmax.z - n.bmin.z;
    return ex * ey + ey * ez + ez * ex;
}
inline float axis_of(const V3& v, int a) { return a == 0 ? v.x : (a == 1 ? v.y : v.z); }

struct Bins {
    // bin k in [1, 64]: the primitive is LEFT for candidates i >= k (k == 64: never left)
    uint32_t count[3][kPlanes + 1];
    Box box[3][kPlanes + 1];
    bool used[3][kPlanes + 1];
    void clear() {
        std::memset(count, 0, sizeof(count));
        std::memset(used, 0, sizeof(used));
    }
    inline void add(int axis, int k, const HostTri& t) {
        if (!used[axis][k]) {
            used[axis][k] = true;
            Box& b = box[axis][k];
            b.mn[0] = t.bmin.x; b.mn[1] = t.bmin.y; b.mn[2] = t.bmin.z;
            b.mx[0] = t.bmax.x; b.mx[1] = t.bmax.y; b.mx[2] = t.bmax.z;
        } else {
            box_grow(box[axis][k], t.bmin, t.bmax);
        }
        count[axis][k]++;
    }
    void merge(const Bins& o) {
        for (int a = 0; a < 3; ++a)
            for (int k = 1; k <= kPlanes; ++k) {
                if (!o.used[a][k]) continue;
                if (!used[a][k]) {
                    used[a][k] = true;
                    box[a][k] = o.box[a][k];
                } else {
                    box_merge(box[a][k], o.box[a][k]);
                }
                count[a][k] += o.count[a][k];
            }
    }
};

struct Builder {
    HostBlas& blas;
    uint32_t max_depth;
    uint32_t deepest = 0;

    // BvhNode::new (blas.rs:27-36): bounds start at (MAX, MIN) and grow over the range
    HostNode make_node(uint32_t offset, uint32_t count) const {
        HostNode n;
        n.bmin = V3{kF32Max, kF32Max, kF32Max};
        n.bmax = V3{-kF32Max, -kF32Max, -kF32Max};
        for (uint32_t i = offset; i < offset + count; ++i) {
            const HostTri& t = blas.tris[i];
            n.bmin = V3{fminf(n.bmin.x, t.bmin.x), fminf(n.bmin.y, t.bmin.y), fminf(n.bmin.z, t.bmin.z)};
            n.bmax = V3{fmaxf(n.bmax.x, t.bmax.x), fmaxf(n.bmax.y, t.bmax.y), fmaxf(n.bmax.z, t.bmax.z)};
        }
        n.offset = offset;
        n.count = count;
        return n;
    }

    // find_best_split_plane (blas.rs:93-123)
    void best_split(const HostNode& node, int& best_axis, float& split_pos, float& best_cost) const {
        best_cost = kF32Max;
        best_axis = 0;
        split_pos = 0.0f;
        float pos[3][kPlanes];
        bool active[3];
        for (int a = 0; a < 3; ++a) {
            float lo = axis_of(node.bmin, a), hi = axis_of(node.bmax, a);
            active[a] = !(lo == hi);
            if (!active[a]) continue;
            float scale = (hi - lo) / (float)kPlanes;
            for (int i = 1; i < kPlanes; ++i) pos[a][i] = lo + (float)i * scale;
        }
        if (!active[0] && !active[1] && !active[2]) return;

        auto bin_of = [&](int a, float c) -> int {
            // first i in [1, 63] with c < pos[a][i]; 64 when none (also for NaN)
            int lo = 1, hi = kPlanes;
            while (lo < hi) {
                int mid = (lo + hi) >> 1;
                if (c < pos[a][mid]) hi = mid; else lo = mid + 1;
            }
            return lo;
        };

        Bins bins;
        bins.clear();
        const uint32_t begin = node.offset, end = node.offset + node.count;
#ifdef _OPENMP
        if (node.count >= (1u << 16)) {
#pragma omp parallel
            {
                Bins local;
                local.clear();
#pragma omp for schedule(static) nowait
                for (int64_t i = begin; i < (int64_t)end; ++i) {
                    const HostTri& t = blas.tris[(size_t)i];
                    for (int a = 0; a < 3; ++a)
                        if (active[a]) local.add(a, bin_of(a, axis_of(t.centroid, a)), t);
                }
#pragma omp critical
                bins.merge(local);
            }
        } else
#endif
        {
            for (uint32_t i = begin; i < end; ++i) {
                const HostTri& t = blas.tris[i];
                for (int a = 0; a < 3; ++a)
                    if (active[a]) bins.add(a, bin_of(a, axis_of(t.centroid, a)), t);
            }
        }

        for (int a = 0; a < 3; ++a) {
            if (!active[a]) continue;
            // suffix unions: right(i) = origin U bins k > i
            Box rbox[kPlanes + 1];
            uint32_t rcount[kPlanes + 1];
            Box acc;
            box_origin(acc);
            uint32_t cnt = 0;
            for (int i = kPlanes - 1; i >= 1; --i) {
                if (bins.used[a][i + 1]) box_merge(acc, bins.box[a][i + 1]);
                cnt += bins.count[a][i + 1];
                rbox[i] = acc;
                rcount[i] = cnt;
            }
            Box lacc;
            box_origin(lacc);
            uint32_t lcnt = 0;
            for (int i = 1; i < kPlanes; ++i) {
                if (bins.used[a][i]) box_merge(lacc, bins.box[a][i]);
                lcnt += bins.count[a][i];
                // evaluate_sah (blas.rs:83-88)
                float cost = (float)(int)lcnt * box_area(lacc) + (float)(int)rcount[i] * box_area(rbox[i]);
                if (!(cost > 0.0f)) cost = kF32Max;
                if (cost < best_cost) {
                    best_cost = cost;
                    best_axis = a;
                    split_pos = pos[a][i];
                }
            }
        }
    }

    // set_primitives_recursive (blas.rs:261-316)
    void recurse(size_t node_index, uint32_t level) {
        if (level > deepest) deepest = level;
        if (level >= max_depth) return;
        HostNode node = blas.nodes[node_index];
        int axis;
        float pos, cost;
        best_split(node, axis, pos, cost);
        float no_split = (float)node.count * node_area(node); // calculate_cost (blas.rs:125-127)
        if (cost > no_split) return;
        size_t i = node.offset, j = (size_t)node.offset + node.count;
        while (i < j) {
            if (axis_of(blas.tris[i].centroid, axis) < pos) {
                ++i;
            } else {
                std::swap(blas.tris[i], blas.tris[j - 1]);
                --j;
            }
        }
        uint32_t left = (uint32_t)i - node.offset;
        uint32_t right = node.count - left;
        if (left > 0 && right > 0) {
            uint32_t li = (uint32_t)blas.nodes.size();
            HostNode l = make_node(node.offset, left);
            HostNode r = make_node(node.offset + left, right);
            node.offset = li;
            node.count = 0;
            blas.nodes.push_back(l);
            blas.nodes.push_back(r);
            recurse(li, level + 1);
            recurse(li + 1, level + 1);
        }
        blas.nodes[node_index] = node;
    }
};

} // namespace

void build_blas(HostBlas& blas, uint32_t max_depth) {
    Builder b{blas, max_depth};
    blas.nodes.clear();
    blas.nodes.reserve(blas.tris.size() * 2 + 2);
    blas.nodes.push_back(b.make_node(0, (uint32_t)blas.tris.size()));
    HostNode dummy{}; // node 1 is never referenced (blas.rs:254-256)
    dummy.bmin = dummy.bmax = V3{0, 0, 0};
    dummy.offset = dummy.count = 0;
    blas.nodes.push_back(dummy);
    b.recurse(0, 0);
    blas.depth = b.deepest;
}

} // namespace rb
