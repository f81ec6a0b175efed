// fast_bvh.cpp — the performance layer: one high-quality binary BVH over ALL triangles of the scene.
//
// Why a second tree: the reference builder (bvh/blas.rs:64-127) starts every candidate box at the
// origin and has no traversal-cost term, so it stops splitting very early — its leaves hold hundreds
// to tens of thousands of triangles (scene7: 73 leaves for 100,000 triangles).  Walking that tree is
// what rayca-soft does on the CPU; on the GPU it would spend all its time in 48-byte triangle fetches.
//
// Exactness is kept by construction (see trace.cuh): the reference semantics are "a triangle is tested
// iff the slab test of its REFERENCE LEAF box accepts the ray; closest depth wins; ties go to the lowest
// (blas, primitive) index".  This tree only decides which triangles get looked at: its slab test is
// made conservative by a relative tolerance, every candidate hit is re-validated against the exact
// reference leaf box, and every triangle carries its reference index for the tie-break.
//
// Build: top-down binned SAH (32 bins, 3 axes) over bounding-box centroids, leaves of <= 4 triangles,
// OpenMP tasks for independent subtrees; children are stored as adjacent 32-byte records (one 64-byte
// fetch per visited inner node) and the array is laid out in depth-first order.
#include "host_scene.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>

#ifdef _OPENMP
#include <omp.h>
#endif

namespace rb {
namespace {

constexpr int kMaxBins = 64;
// Tunables (defaults picked by measurement on B200, profiles/; the RAYCA_FAST_* environment variables
// exist for such A/B runs only).
int kBins = 32;
uint32_t kLeafMax = 4;            // <= 8: a packed leaf reference holds count-1 in 3 bits
float kTraversalCost = 1.0f;      // cost of visiting one node relative to one triangle test

void read_tunables() {
    if (const char* e = std::getenv("RAYCA_FAST_BINS")) kBins = std::max(4, std::min(kMaxBins, std::atoi(e)));
    if (const char* e = std::getenv("RAYCA_FAST_LEAF_MAX")) kLeafMax = (uint32_t)std::max(1, std::min(8, std::atoi(e)));
    if (const char* e = std::getenv("RAYCA_FAST_TRAV_COST")) kTraversalCost = (float)std::atof(e);
}

struct Item {
    float mn[3], mx[3];
    float c[3];
    uint32_t id; // reference-global triangle index
    uint32_t sphere; // 0: triangle, else 1 (a leaf never mixes the two kinds)
};

struct Bx {
    float mn[3], mx[3];
    void reset() {
        for (int k = 0; k < 3; ++k) { mn[k] = kF32Max; mx[k] = -kF32Max; }
    }
    void grow(const float* lo, const float* hi) {
        for (int k = 0; k < 3; ++k) { mn[k] = fminf(mn[k], lo[k]); mx[k] = fmaxf(mx[k], hi[k]); }
    }
    void merge(const Bx& o) { grow(o.mn, o.mx); }
    float area() const {
        float ex = mx[0] - mn[0], ey = mx[1] - mn[1], ez = mx[2] - mn[2];
        if (ex < 0.0f) return 0.0f;
        return ex * ey + ey * ez + ez * ex;
    }
};

struct Rec { // one 32-byte record under construction
    Bx box;
    uint32_t link0, link1;
};

struct FastBuilder {
    std::vector<Item>& items;
    std::vector<Rec> slots; // pair for the inner node split at position m lives in slots[2m], slots[2m+1]
    uint32_t max_depth = 0;

    explicit FastBuilder(std::vector<Item>& it) : items(it), slots(it.size() * 2 + 2) {}

    // builds the subtree over [b, e); returns its record (box + links)
    Rec build(uint32_t b, uint32_t e, uint32_t depth) {
        Rec self;
        self.box.reset();
        Bx cb;
        cb.reset();
        for (uint32_t i = b; i < e; ++i) {
            self.box.grow(items[i].mn, items[i].mx);
            cb.grow(items[i].c, items[i].c);
        }
        const uint32_t n = e - b;
        auto make_leaf = [&]() {
            self.link0 = b;
            self.link1 = n;
            if (depth > max_depth) {
#pragma omp critical(fast_depth)
                if (depth > max_depth) max_depth = depth;
            }
            return self;
        };
        if (n <= 1) return make_leaf();
        bool mixed = false; // triangles and spheres have different 64-byte records: keep leaves pure
        for (uint32_t i = b + 1; i < e && !mixed; ++i) mixed = items[i].sphere != items[b].sphere;

        // binned SAH
        float best_cost = kF32Max;
        int best_axis = -1, best_bin = 0;
        for (int a = 0; a < 3; ++a) {
            float lo = cb.mn[a], hi = cb.mx[a];
            if (!(hi > lo)) continue;
            float scale = (float)kBins / (hi - lo);
            Bx bbox[kMaxBins];
            uint32_t bcnt[kMaxBins];
            for (int k = 0; k < kBins; ++k) { bbox[k].reset(); bcnt[k] = 0; }
            for (uint32_t i = b; i < e; ++i) {
                int k = (int)((items[i].c[a] - lo) * scale);
                k = k < 0 ? 0 : (k >= kBins ? kBins - 1 : k);
                bbox[k].grow(items[i].mn, items[i].mx);
                bcnt[k]++;
            }
            float rarea[kMaxBins];
            uint32_t rcnt[kMaxBins];
            Bx acc;
            acc.reset();
            uint32_t cnt = 0;
            for (int k = kBins - 1; k >= 1; --k) {
                if (bcnt[k]) acc.merge(bbox[k]);
                cnt += bcnt[k];
                rarea[k] = acc.area();
                rcnt[k] = cnt;
            }
            acc.reset();
            cnt = 0;
            for (int k = 1; k < kBins; ++k) { // split between bin k-1 and k
                if (bcnt[k - 1]) acc.merge(bbox[k - 1]);
                cnt += bcnt[k - 1];
                if (cnt == 0 || rcnt[k] == 0) continue;
                float cost = acc.area() * (float)cnt + rarea[k] * (float)rcnt[k];
                if (cost < best_cost) { best_cost = cost; best_axis = a; best_bin = k; }
            }
        }
        float parent_area = self.box.area();
        if (n <= kLeafMax) {
            float leaf_cost = (float)n * parent_area;
            float split_cost = kTraversalCost * parent_area + best_cost;
            if (!mixed && (best_axis < 0 || leaf_cost <= split_cost)) return make_leaf();
        }
        uint32_t mid;
        if (best_axis < 0) {
            mid = b + n / 2; // all centroids coincide: split by position
        } else {
            float lo = cb.mn[best_axis], scale = (float)kBins / (cb.mx[best_axis] - lo);
            Item* first = items.data() + b;
            Item* last = items.data() + e;
            Item* m = std::partition(first, last, [&](const Item& it) {
                int k = (int)((it.c[best_axis] - lo) * scale);
                k = k < 0 ? 0 : (k >= kBins ? kBins - 1 : k);
                return k < best_bin;
            });
            mid = (uint32_t)(m - items.data());
            if (mid == b || mid == e) mid = b + n / 2;
        }
        Rec l, r;
        if (n > 8192) {
#pragma omp task shared(l) firstprivate(b, mid, depth)
            l = build(b, mid, depth + 1);
#pragma omp task shared(r) firstprivate(mid, e, depth)
            r = build(mid, e, depth + 1);
#pragma omp taskwait
        } else {
            l = build(b, mid, depth + 1);
            r = build(mid, e, depth + 1);
        }
        slots[2 * (size_t)mid] = l;
        slots[2 * (size_t)mid + 1] = r;
        self.link0 = mid; // pair slot; rewritten by the depth-first relayout
        self.link1 = 0;
        return self;
    }
};

GpuNode to_gpu(const Rec& r) {
    GpuNode g;
    g.min_x = r.box.mn[0]; g.min_y = r.box.mn[1]; g.min_z = r.box.mn[2];
    g.max_x = r.box.mx[0]; g.max_y = r.box.mx[1]; g.max_z = r.box.mx[2];
    g.link0 = r.link0;
    g.link1 = r.link1;
    return g;
}

} // namespace

void fast_tree_tunables(int& bins, uint32_t& leaf_max, float& traversal_cost) {
    read_tunables();
    bins = kBins;
    leaf_max = kLeafMax;
    traversal_cost = kTraversalCost;
}

void build_fast_tree(HostScene& hs) {
    read_tunables();
    // gather triangles in reference-global order with their reference leaf
    const size_t nb = hs.blas.size();
    const uint32_t total = hs.blas_prim_offset[nb];
    std::vector<uint32_t> ref_leaf(total, 0);
    for (size_t bi = 0; bi < nb; ++bi) {
        const HostBlas& b = hs.blas[bi];
        for (size_t k = 0; k < b.nodes.size(); ++k) {
            const HostNode& n = b.nodes[k];
            if (k == 1 || n.count == 0) continue;
            for (uint32_t i = 0; i < n.count; ++i) ref_leaf[hs.blas_prim_offset[bi] + n.offset + i] = hs.blas_node_offset[bi] + (uint32_t)k;
        }
    }
    std::vector<Item> items(total);
    for (size_t bi = 0; bi < nb; ++bi) {
        const HostBlas& b = hs.blas[bi];
        const uint32_t po = hs.blas_prim_offset[bi];
#pragma omp parallel for schedule(static) if (b.tris.size() > 8192)
        for (int64_t k = 0; k < (int64_t)b.tris.size(); ++k) {
            const HostTri& t = b.tris[(size_t)k];
            Item& it = items[po + (size_t)k];
            it.mn[0] = t.bmin.x; it.mn[1] = t.bmin.y; it.mn[2] = t.bmin.z;
            it.mx[0] = t.bmax.x; it.mx[1] = t.bmax.y; it.mx[2] = t.bmax.z;
            for (int a = 0; a < 3; ++a) it.c[a] = 0.5f * (it.mn[a] + it.mx[a]);
            it.id = po + (uint32_t)k;
            it.sphere = t.sphere ? 1u : 0u;
        }
    }
    FastBuilder fb(items);
    Rec root;
#pragma omp parallel
    {
#pragma omp single
        root = fb.build(0, total, 0);
    }
    // depth-first relayout of the pairs (left subtree right after its parent pair).  A child
    // reference (GpuNode::link0) is either a pair index or, with kLeafRef set, a packed leaf
    // {bits 0..27 first triangle, bits 28..30 count-1}.
    constexpr uint32_t kLeafRef = 0x80000000u, kSphereRef = 0x08000000u; // first index: 27 bits
    auto leaf_ref = [&](const Rec& r) {
        return kLeafRef | ((r.link1 - 1u) << 28) | (items[r.link0].sphere ? kSphereRef : 0u) | r.link0;
    };
    hs.fast_nodes.clear();
    hs.fast_nodes.reserve(items.size());
    struct Todo { uint32_t slot; uint32_t dst; };
    std::vector<Todo> stack;
    GpuNode groot = to_gpu(root);
    if (root.link1 == 0) {
        hs.fast_nodes.resize(2);
        stack.push_back({root.link0, 0});
        groot.link0 = 0;
    } else {
        groot.link0 = leaf_ref(root);
    }
    while (!stack.empty()) {
        Todo t = stack.back();
        stack.pop_back();
        Rec l = fb.slots[2 * (size_t)t.slot], r = fb.slots[2 * (size_t)t.slot + 1];
        GpuNode gl = to_gpu(l), gr = to_gpu(r);
        // allocate the right child's pair first so that the LEFT child's pair is popped (and laid out) next
        if (r.link1 == 0) {
            uint32_t dst = (uint32_t)hs.fast_nodes.size();
            hs.fast_nodes.resize(hs.fast_nodes.size() + 2);
            gr.link0 = dst;
            stack.push_back({r.link0, dst});
        } else {
            gr.link0 = leaf_ref(r);
        }
        if (l.link1 == 0) {
            uint32_t dst = (uint32_t)hs.fast_nodes.size();
            hs.fast_nodes.resize(hs.fast_nodes.size() + 2);
            gl.link0 = dst;
            stack.push_back({l.link0, dst});
        } else {
            gl.link0 = leaf_ref(l);
        }
        hs.fast_nodes[t.dst] = gl;
        hs.fast_nodes[t.dst + 1] = gr;
    }
    hs.fast_root = groot;
    hs.fast_depth = fb.max_depth;
    // 64-byte triangle records in leaf order: {n.xyz, d | v0, reference-global index | v1, reference
    // leaf node | v2, dot(n,n)}.  n = (v1-v0) x (v2-v0), d = -dot(n, v0) and dot(n,n) are evaluated with
    // the operation order of Triangle::intersects (rayca-geometry/src/triangle.rs:90-94,110,152), so the
    // kernel reads the bits the reference would recompute for every test.
    hs.fast_tri.resize((size_t)total * 4);
#pragma omp parallel for schedule(static) if (total > 8192)
    for (int64_t i = 0; i < (int64_t)total; ++i) {
        uint32_t id = items[(size_t)i].id;
        if (items[(size_t)i].sphere) {
            // sphere record: what `&Inversed(trs) * ray` applies (translate by -t, rotate by q^-1, scale by 1/s;
            // trs.rs:301-312,405-414) | {centre, radius^2} of the model-space sphere
            size_t bi = 0;
            while (bi + 1 < nb && id >= hs.blas_prim_offset[bi + 1]) ++bi;
            const HostBlas& hb = hs.blas[bi];
            const HostSphere& sp = hb.spheres[hb.tris[id - hs.blas_prim_offset[bi]].sphere - 1];
            const V3 rs = reciprocal(sp.world.s);
            const Q4 qi = conj(sp.world.q);
            float4* dst = &hs.fast_tri[(size_t)i * 4];
            dst[0] = make_float4(-sp.world.t.x, -sp.world.t.y, -sp.world.t.z, sp.radius2);
            dst[1] = make_float4(qi.x, qi.y, qi.z, qi.w);
            dst[2] = make_float4(rs.x, rs.y, rs.z, __builtin_bit_cast(float, id));
            dst[3] = make_float4(sp.center.x, sp.center.y, sp.center.z, __builtin_bit_cast(float, ref_leaf[id]));
            continue;
        }
        const float4* src = &hs.tri_pos[(size_t)id * 3];
        const V3 v0{src[0].x, src[0].y, src[0].z}, v1{src[1].x, src[1].y, src[1].z}, v2{src[2].x, src[2].y, src[2].z};
        const V3 n = cross(v1 - v0, v2 - v0);
        const float d = -dot(n, v0);
        const float den = dot(n, n);
        float4* dst = &hs.fast_tri[(size_t)i * 4];
        dst[0] = make_float4(n.x, n.y, n.z, d);
        dst[1] = make_float4(v0.x, v0.y, v0.z, __builtin_bit_cast(float, id));
        dst[2] = make_float4(v1.x, v1.y, v1.z, __builtin_bit_cast(float, ref_leaf[id]));
        dst[3] = make_float4(v2.x, v2.y, v2.z, den);
    }
}

} // namespace rb
