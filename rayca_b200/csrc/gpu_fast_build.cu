// gpu_fast_build.cu — the performance tree (fast_bvh.cpp) built on the GPU, emitted directly in the
// 32-byte record layout the walk reads.  SURVEY 8(f) rank 1: `draw` rebuilds its BVH on every call
// (rayca-soft/src/scene.rs:92-99), so once tracing takes a millisecond the build is the frame time.
//
// Same algorithm as the host builder: top-down binned SAH, 32 bins x 3 axes over bounding-box centroids,
// leaves of <= kLeafMax primitives, triangles and spheres never share a leaf.  Two tiers:
//   * BIG nodes (> 32 primitives) are processed level by level, all nodes of a level at once: one thread per
//     primitive adds its box to the (node, axis, bin) accumulators with ordered-integer atomicMin / atomicMax,
//     one thread per (node, axis) sweeps the 31 planes, a scan over the left/right flags gives every primitive
//     its place in a stable partition, and the scatter pass accumulates the boxes of the big children.
//   * a node of <= 32 primitives is finished by ONE WARP: the primitives live in shared memory, lane k
//     evaluates plane k of each axis, the subtree is built depth-first with a small stack, and its pairs go into
//     a block of pair slots reserved for it (n - 1 pairs at most).
// Pair numbering is deterministic (level base + node slot, scanned reservations): no atomics decide layout.
// The tree may differ from the host builder's in the order of primitives inside a leaf (stable partition here,
// std::partition there); the walk's result does not depend on it (every hit is validated against its reference
// leaf box and ties go to the lowest reference index, trace.cuh).
#include "gpu_fast_build.h"

#include <algorithm>
#include <cstdio>
#include <chrono>
#include <cstring>

namespace rb {
namespace {

constexpr uint32_t kSmall = 32;   // nodes of <= kSmall primitives are finished by one warp
constexpr int kGpuBins = 32;
constexpr uint32_t kNone = 0xFFFFFFFFu;
constexpr uint32_t kRootRec = 0xFFFFFFFEu;
constexpr uint32_t kLeafRef = 0x80000000u, kSphereRef = 0x08000000u;
constexpr float kMaxF = 3.402823466e+38f;

// float <-> unsigned with the same order (so that integer atomicMin / atomicMax order floats)
__device__ __forceinline__ uint32_t f2o(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float o2f(uint32_t o) {
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o);
}
constexpr uint32_t kOrdMaxInit = 0u;          // below every float
constexpr uint32_t kOrdMinInit = 0xFFFFFFFFu; // above every float

__device__ __forceinline__ void atomic_box(uint32_t* b, const float mn[3], const float mx[3]) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        if (mn[k] == mn[k]) atomicMin(b + k, f2o(mn[k])); // fminf / fmaxf ignore NaN operands
        if (mx[k] == mx[k]) atomicMax(b + 3 + k, f2o(mx[k]));
    }
}

struct BoxF {
    float mn[3], mx[3];
};
__device__ __forceinline__ void box_reset(BoxF& b) {
    for (int k = 0; k < 3; ++k) { b.mn[k] = kMaxF; b.mx[k] = -kMaxF; }
}
__device__ __forceinline__ void box_grow(BoxF& b, const float* lo, const float* hi) {
    for (int k = 0; k < 3; ++k) { b.mn[k] = fminf(b.mn[k], lo[k]); b.mx[k] = fmaxf(b.mx[k], hi[k]); }
}
__device__ __forceinline__ float box_area(const BoxF& b) { // Bx::area of fast_bvh.cpp
    float ex = b.mx[0] - b.mn[0], ey = b.mx[1] - b.mn[1], ez = b.mx[2] - b.mn[2];
    if (ex < 0.0f) return 0.0f;
    return ex * ey + ey * ez + ez * ex;
}
__device__ __forceinline__ int bin_of(float c, float lo, float scale) {
    int k = (int)((c - lo) * scale);
    return k < 0 ? 0 : (k >= kGpuBins ? kGpuBins - 1 : k);
}

// ---- exclusive scan of u32 (out has n + 1 entries), three kernels ---------------------------------------
constexpr int kScanBlock = 1024;
__device__ __forceinline__ uint32_t block_exclusive(uint32_t v, uint32_t* total) {
    __shared__ uint32_t warp_sums[32];
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
        if (lane >= (uint32_t)o) inc += t;
    }
    if (lane == 31) warp_sums[w] = inc;
    __syncthreads();
    if (w == 0) {
        uint32_t s = warp_sums[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xFFFFFFFFu, s, o);
            if (lane >= (uint32_t)o) s += t;
        }
        warp_sums[lane] = s;
    }
    __syncthreads();
    const uint32_t base = w ? warp_sums[w - 1] : 0u;
    if (total) *total = warp_sums[31];
    __syncthreads();
    return base + inc - v;
}
__global__ void __launch_bounds__(kScanBlock) scan_partials(const uint32_t* in, uint32_t n, uint32_t* partial) {
    const uint32_t i = blockIdx.x * kScanBlock + threadIdx.x;
    uint32_t total;
    block_exclusive(i < n ? in[i] : 0u, &total);
    if (threadIdx.x == 0) partial[blockIdx.x] = total;
}
__global__ void __launch_bounds__(kScanBlock) scan_partials_scan(uint32_t* partial, uint32_t blocks) {
    // one block walks the partial sums in chunks of 1024
    uint32_t carry = 0;
    for (uint32_t base = 0; base < blocks; base += kScanBlock) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < blocks ? partial[i] : 0u;
        uint32_t total;
        const uint32_t ex = block_exclusive(v, &total);
        if (i < blocks) partial[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) partial[blocks] = carry;
}
__global__ void __launch_bounds__(kScanBlock) scan_final(const uint32_t* in, uint32_t n, const uint32_t* partial, uint32_t* out) {
    const uint32_t i = blockIdx.x * kScanBlock + threadIdx.x;
    const uint32_t ex = block_exclusive(i < n ? in[i] : 0u, nullptr);
    if (i < n) out[i] = partial[blockIdx.x] + ex;
    if (i == 0) out[n] = partial[gridDim.x];
}
// short inputs (the per-node arrays of a level): one block walks them in chunks of 1024 — one launch instead of three
__global__ void __launch_bounds__(kScanBlock) scan_small(const uint32_t* in, uint32_t n, uint32_t* out) {
    uint32_t carry = 0;
    for (uint32_t base = 0; base < n; base += kScanBlock) {
        const uint32_t i = base + threadIdx.x;
        uint32_t total;
        const uint32_t ex = block_exclusive(i < n ? in[i] : 0u, &total);
        if (i < n) out[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) out[n] = carry;
}
void exclusive_scan(const uint32_t* in, uint32_t n, uint32_t* out, uint32_t* partial, cudaStream_t st) {
    if (n <= 16u * kScanBlock) {
        scan_small<<<1, kScanBlock, 0, st>>>(in, n, out);
        return;
    }
    const uint32_t blocks = (n + kScanBlock - 1) / kScanBlock;
    scan_partials<<<blocks, kScanBlock, 0, st>>>(in, n, partial);
    scan_partials_scan<<<1, kScanBlock, 0, st>>>(partial, blocks);
    scan_final<<<blocks, kScanBlock, 0, st>>>(in, n, partial, out);
}

// ---- tier 1: big nodes, level by level -------------------------------------------------------------------
struct WorkNode {
    uint32_t start, n, rec, depth; // rec: index of this node's 32-byte record (kRootRec: the root record)
};
struct NodeInfo { // decoded once per node per level
    float box_mn[3], box_mx[3];
    float lo[3], scale[3];
    uint32_t active; // bit a: the centroids spread along axis a
};
struct Decision {
    uint32_t axis, bin;  // primitive goes left iff bin_of(axis) < bin
    float lo, scale;
    uint32_t nl;         // primitives on the left
    uint32_t positional; // 1: centroids coincide (or the plane separates nothing): split by position at n / 2
    uint32_t pair;       // pair slot of the two children
    uint32_t child_slot[2]; // slot in the next level's big list, or kNone
};
constexpr int kAcc = 12;                 // box min, box max, centroid min, centroid max
constexpr int kBinWords = 7;             // box min, box max, count
constexpr int kNodeBinWords = 3 * kGpuBins * kBinWords;

__global__ void init_acc(uint32_t* acc, uint32_t nodes) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nodes * kAcc) return;
    const uint32_t k = i % kAcc;
    acc[i] = (k % 6) < 3 ? kOrdMinInit : kOrdMaxInit;
}
// Box and centroid box of every big node: positions of a node are contiguous, so a warp first reduces each run of
// equal slots with shuffles and only the first lane of a run touches memory (same-address atomics from different
// warps serialise at ~15 ns each: 2 ms per level for the top levels without this).
__device__ __forceinline__ uint32_t run_min(uint32_t v, uint32_t key, uint32_t lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t ov = __shfl_down_sync(0xFFFFFFFFu, v, o), ok = __shfl_down_sync(0xFFFFFFFFu, key, o);
        if (lane + o < 32 && ok == key) v = min(v, ov);
    }
    return v;
}
__device__ __forceinline__ uint32_t run_max(uint32_t v, uint32_t key, uint32_t lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t ov = __shfl_down_sync(0xFFFFFFFFu, v, o), ok = __shfl_down_sync(0xFFFFFFFFu, key, o);
        if (lane + o < 32 && ok == key) v = max(v, ov);
    }
    return v;
}
__global__ void accumulate_boxes(const float4* __restrict__ bmn, const float4* __restrict__ bmx, const uint32_t* __restrict__ order,
                                 const uint32_t* __restrict__ node_of, uint32_t n, uint32_t* acc) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31u;
    const uint32_t s = p < n ? node_of[p] : kNone;
    uint32_t lo[6], hi[6]; // ordered-integer images; NaN bounds are skipped like fminf / fmaxf skip them
#pragma unroll
    for (int k = 0; k < 6; ++k) { lo[k] = kOrdMinInit; hi[k] = kOrdMaxInit; }
    if (s != kNone) {
        const uint32_t id = order[p];
        const float4 a = bmn[id], b = bmx[id];
        const float mn[3] = {a.x, a.y, a.z}, mx[3] = {b.x, b.y, b.z};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float c = 0.5f * (mn[k] + mx[k]);
            if (mn[k] == mn[k]) lo[k] = f2o(mn[k]);
            if (mx[k] == mx[k]) hi[k] = f2o(mx[k]);
            if (c == c) { lo[3 + k] = f2o(c); hi[3 + k] = f2o(c); }
        }
    }
    const uint32_t prev = __shfl_up_sync(0xFFFFFFFFu, s, 1);
    const bool head = lane == 0 || prev != s;
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        lo[k] = run_min(lo[k], s, lane);
        hi[k] = run_max(hi[k], s, lane);
    }
    // A block whose primitives all belong to ONE node (every block of the top levels) reduces in shared memory first:
    // at level 0 the 12 words of the root took one atomic per warp each, 98k same-address atomics, 68 us.
    __shared__ uint32_t s_acc[kAcc];
    const uint32_t s0 = node_of[min(blockIdx.x * blockDim.x, n - 1u)];
    const bool uniform = __syncthreads_and(s == s0 || p >= n) && s0 != kNone;
    if (uniform) {
        if (threadIdx.x < (uint32_t)kAcc) s_acc[threadIdx.x] = (threadIdx.x % 6u) < 3u ? kOrdMinInit : kOrdMaxInit;
        __syncthreads();
    }
    if (s != kNone && head) {
        uint32_t* a = uniform ? s_acc : acc + (size_t)s * kAcc;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            atomicMin(a + k, lo[k]);
            atomicMax(a + 3 + k, hi[k]);
            atomicMin(a + 6 + k, lo[3 + k]);
            atomicMax(a + 9 + k, hi[3 + k]);
        }
    }
    if (!uniform) return;
    __syncthreads();
    if (threadIdx.x < (uint32_t)kAcc) {
        uint32_t* a = acc + (size_t)s0 * kAcc + threadIdx.x;
        if ((threadIdx.x % 6u) < 3u) atomicMin(a, s_acc[threadIdx.x]);
        else atomicMax(a, s_acc[threadIdx.x]);
    }
}
// one block per node: the node's bins are initialised here as well (was a launch of its own)
__global__ void prepare_nodes(const uint32_t* __restrict__ acc, uint32_t nodes, NodeInfo* info, uint32_t* bins) {
    const uint32_t s = blockIdx.x;
    if (s >= nodes) return;
    for (uint32_t i = threadIdx.x; i < (uint32_t)kNodeBinWords; i += blockDim.x) {
        const uint32_t k = i % kBinWords;
        bins[(size_t)s * kNodeBinWords + i] = k < 3 ? kOrdMinInit : (k < 6 ? kOrdMaxInit : 0u);
    }
    if (threadIdx.x != 0) return;
    const uint32_t* a = acc + (size_t)s * kAcc;
    NodeInfo ni;
    ni.active = 0;
    for (int k = 0; k < 3; ++k) {
        ni.box_mn[k] = o2f(a[k]);
        ni.box_mx[k] = o2f(a[3 + k]);
        const float lo = o2f(a[6 + k]), hi = o2f(a[9 + k]);
        ni.lo[k] = lo;
        ni.scale[k] = 0.0f;
        if (hi > lo) {
            ni.active |= 1u << k;
            ni.scale[k] = (float)kGpuBins / (hi - lo);
        }
    }
    info[s] = ni;
}
// One thread per primitive.  A block whose 256 positions all belong to ONE node (every block of the top levels)
// bins into shared memory and flushes its non-empty bins once; other blocks go to the global bins directly.
constexpr uint32_t kBinThreads = 256;
__global__ void __launch_bounds__(kBinThreads) bin_prims(const float4* __restrict__ bmn, const float4* __restrict__ bmx,
                                                       const uint32_t* __restrict__ order, const uint32_t* __restrict__ node_of, uint32_t n,
                                                       const NodeInfo* __restrict__ info, uint32_t* bins) {
    __shared__ uint32_t local[kNodeBinWords];
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t s = p < n ? node_of[p] : kNone;
    const uint32_t s0 = node_of[blockIdx.x * blockDim.x];
    const bool uniform = __syncthreads_and(s == s0 || p >= n) && s0 != kNone;
    if (uniform) {
        for (uint32_t i = threadIdx.x; i < (uint32_t)kNodeBinWords; i += kBinThreads) {
            const uint32_t k = i % kBinWords;
            local[i] = k < 3 ? kOrdMinInit : (k < 6 ? kOrdMaxInit : 0u);
        }
        __syncthreads();
    }
    if (s != kNone) {
        const uint32_t id = order[p];
        const float4 a = bmn[id], b = bmx[id];
        const float mn[3] = {a.x, a.y, a.z}, mx[3] = {b.x, b.y, b.z};
        const NodeInfo& ni = info[s];
        const uint32_t active = ni.active;
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) {
            if (!(active >> ax & 1u)) continue;
            const int k = bin_of(0.5f * (mn[ax] + mx[ax]), ni.lo[ax], ni.scale[ax]);
            uint32_t* w = (uniform ? local : bins + (size_t)s * kNodeBinWords) + ((size_t)ax * kGpuBins + k) * kBinWords;
            atomic_box(w, mn, mx);
            atomicAdd(w + 6, 1u);
        }
    }
    if (!uniform) return;
    __syncthreads();
    uint32_t* g = bins + (size_t)s0 * kNodeBinWords;
    for (uint32_t i = threadIdx.x; i < (uint32_t)kNodeBinWords; i += kBinThreads) {
        const uint32_t bin = i / kBinWords, k = i % kBinWords;
        if (local[bin * kBinWords + 6] == 0) continue;
        if (k < 3) atomicMin(g + i, local[i]);
        else if (k < 6) atomicMax(g + i, local[i]);
        else atomicAdd(g + i, local[i]);
    }
}
// One block per node, one warp per axis, one bin per lane: the boxes and counts left and right of the 31 planes are
// warp scans, lane k prices plane k (between bin k-1 and bin k), a butterfly keeps the cheapest (the lowest plane on
// equal cost, like an ascending sweep with strict `<`); thread 0 decides.  (One thread per (node, axis) sweeping the
// bins twice in global memory took 34 us per level whatever the node count.)
__device__ __forceinline__ BoxF fast_box_union(const BoxF& a, const BoxF& b) {
    BoxF r;
    for (int k = 0; k < 3; ++k) { r.mn[k] = fminf(a.mn[k], b.mn[k]); r.mx[k] = fmaxf(a.mx[k], b.mx[k]); }
    return r;
}
__device__ __forceinline__ BoxF fast_box_shfl(const BoxF& a, int d, bool up) {
    BoxF r;
    for (int k = 0; k < 3; ++k) {
        r.mn[k] = up ? __shfl_up_sync(0xFFFFFFFFu, a.mn[k], d) : __shfl_down_sync(0xFFFFFFFFu, a.mn[k], d);
        r.mx[k] = up ? __shfl_up_sync(0xFFFFFFFFu, a.mx[k], d) : __shfl_down_sync(0xFFFFFFFFu, a.mx[k], d);
    }
    return r;
}
__global__ void __launch_bounds__(96) decide_splits(const WorkNode* __restrict__ work, uint32_t nodes, const NodeInfo* __restrict__ info,
                                                    const uint32_t* __restrict__ bins, uint32_t pair_base, Decision* dec,
                                                    GpuNode* records, GpuNode* root_record, uint32_t* child_big, uint32_t* child_small) {
    static_assert(kGpuBins == 32, "one bin per lane");
    __shared__ float s_cost[3];
    __shared__ uint32_t s_bin[3], s_nl[3];
    const uint32_t s = blockIdx.x, ax = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    if (s >= nodes) return;
    float best_cost = kMaxF;
    uint32_t best_bin = 0, best_nl = 0;
    if (info[s].active >> ax & 1u) { // the same for the whole warp
        const uint32_t* b = bins + (((size_t)s * 3 + ax) * kGpuBins + lane) * kBinWords;
        BoxF mine;
        box_reset(mine); // an empty bin is the identity of the union
        const uint32_t cmine = b[6];
        if (cmine) {
            const float lo[3] = {o2f(b[0]), o2f(b[1]), o2f(b[2])}, hi[3] = {o2f(b[3]), o2f(b[4]), o2f(b[5])};
            box_grow(mine, lo, hi);
        }
        BoxF up = mine, down = mine; // inclusive prefix (bins <= lane) and suffix (bins >= lane)
        uint32_t cup = cmine, cdown = cmine;
        for (int d = 1; d < 32; d <<= 1) {
            const BoxF ou = fast_box_shfl(up, d, true), od = fast_box_shfl(down, d, false);
            const uint32_t cu = __shfl_up_sync(0xFFFFFFFFu, cup, d), cd = __shfl_down_sync(0xFFFFFFFFu, cdown, d);
            if ((int)lane >= d) { up = fast_box_union(up, ou); cup += cu; }
            if ((int)lane + d < 32) { down = fast_box_union(down, od); cdown += cd; }
        }
        const BoxF left = fast_box_shfl(up, 1, true); // bins < lane
        const uint32_t cnt = __shfl_up_sync(0xFFFFFFFFu, cup, 1);
        if (lane >= 1 && cnt != 0 && cdown != 0) { // plane `lane`: left = bins 0..lane-1, right = bins lane..31
            const float cost = box_area(left) * (float)cnt + box_area(down) * (float)cdown;
            if (cost < best_cost) { best_cost = cost; best_bin = lane; best_nl = cnt; }
        }
        for (int d = 16; d > 0; d >>= 1) {
            const float oc = __shfl_xor_sync(0xFFFFFFFFu, best_cost, d);
            const uint32_t ob = __shfl_xor_sync(0xFFFFFFFFu, best_bin, d), on = __shfl_xor_sync(0xFFFFFFFFu, best_nl, d);
            if (oc < best_cost || (oc == best_cost && ob < best_bin)) { best_cost = oc; best_bin = ob; best_nl = on; }
        }
    }
    if (lane == 0) {
        s_cost[ax] = best_cost;
        s_bin[ax] = best_bin;
        s_nl[ax] = best_nl;
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    constexpr uint32_t local = 0;
    int axis = -1;
    float cost = kMaxF;
    for (int a = 0; a < 3; ++a)
        if (s_cost[local * 3 + a] < cost) { cost = s_cost[local * 3 + a]; axis = a; }
    const WorkNode wn = work[s];
    const NodeInfo& ni = info[s];
    Decision d;
    d.positional = 0;
    d.axis = 0;
    d.bin = 0;
    d.lo = 0.0f;
    d.scale = 0.0f;
    d.nl = wn.n / 2;
    if (axis < 0) {
        d.positional = 1;
    } else {
        d.axis = (uint32_t)axis;
        d.bin = s_bin[local * 3 + axis];
        d.lo = ni.lo[axis];
        d.scale = ni.scale[axis];
        d.nl = s_nl[local * 3 + axis];
        if (d.nl == 0 || d.nl == wn.n) { // cannot happen (both sides counted), kept like the host builder
            d.positional = 1;
            d.nl = wn.n / 2;
        }
    }
    d.pair = pair_base + s;
    d.child_slot[0] = d.child_slot[1] = kNone;
    dec[s] = d;
    GpuNode g;
    g.min_x = ni.box_mn[0]; g.min_y = ni.box_mn[1]; g.min_z = ni.box_mn[2];
    g.max_x = ni.box_mx[0]; g.max_y = ni.box_mx[1]; g.max_z = ni.box_mx[2];
    g.link0 = 2 * d.pair; // record index of the left child; the right child follows it
    g.link1 = 0;
    *(wn.rec == kRootRec ? root_record : records + wn.rec) = g;
    const uint32_t nc[2] = {d.nl, wn.n - d.nl};
    for (int c = 0; c < 2; ++c) {
        child_big[2 * s + c] = nc[c] > kSmall ? 1u : 0u;
        child_small[2 * s + c] = nc[c] > kSmall ? 0u : 1u;
    }
}
// one thread per child: append it to the next level's big list or to the list of small subtrees
__global__ void make_children(const WorkNode* __restrict__ work, uint32_t nodes, Decision* dec, const uint32_t* __restrict__ big_index,
                              const uint32_t* __restrict__ small_index, WorkNode* next, WorkNode* small, uint32_t small_base, uint32_t* next_acc) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nodes * 2) return;
    const uint32_t s = i >> 1, c = i & 1u;
    const WorkNode wn = work[s];
    Decision& d = dec[s];
    WorkNode child;
    child.start = c ? wn.start + d.nl : wn.start;
    child.n = c ? wn.n - d.nl : d.nl;
    child.rec = 2 * d.pair + c;
    child.depth = wn.depth + 1;
    if (child.n > kSmall) {
        const uint32_t slot = big_index[i];
        next[slot] = child;
        d.child_slot[c] = slot;
        for (uint32_t k = 0; k < (uint32_t)kAcc; ++k) next_acc[(size_t)slot * kAcc + k] = (k % 6u) < 3u ? kOrdMinInit : kOrdMaxInit; // was init_acc
    } else {
        small[small_base + small_index[i]] = child;
    }
}
__global__ void left_flags(const float4* __restrict__ bmn, const float4* __restrict__ bmx, const uint32_t* __restrict__ order,
                           const uint32_t* __restrict__ node_of, uint32_t n, const WorkNode* __restrict__ work,
                           const Decision* __restrict__ dec, uint32_t* flags) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const uint32_t s = node_of[p];
    uint32_t f = 0;
    if (s != kNone) {
        const Decision& d = dec[s];
        if (d.positional) {
            f = (p - work[s].start) < d.nl;
        } else {
            const uint32_t id = order[p];
            const float4 a = bmn[id], b = bmx[id];
            const float mn = d.axis == 0 ? a.x : (d.axis == 1 ? a.y : a.z), mx = d.axis == 0 ? b.x : (d.axis == 1 ? b.y : b.z);
            f = (uint32_t)bin_of(0.5f * (mn + mx), d.lo, d.scale) < d.bin;
        }
    }
    flags[p] = f;
}
// stable partition of every big node's range + next level's node_of
__global__ void scatter_prims(const uint32_t* __restrict__ order, const uint32_t* __restrict__ node_of, uint32_t n,
                              const WorkNode* __restrict__ work, const Decision* __restrict__ dec, const uint32_t* __restrict__ flags,
                              const uint32_t* __restrict__ scan, uint32_t* order_out, uint32_t* node_of_out) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const uint32_t s = node_of[p];
    const uint32_t id = order[p];
    if (s == kNone) {
        order_out[p] = id;
        node_of_out[p] = kNone;
        return;
    }
    const WorkNode wn = work[s];
    const Decision& d = dec[s];
    const uint32_t lefts_before = scan[p] - scan[wn.start];
    const uint32_t q = p - wn.start;
    const uint32_t left = flags[p];
    const uint32_t dst = wn.start + (left ? lefts_before : d.nl + (q - lefts_before));
    const uint32_t slot = d.child_slot[left ? 0 : 1];
    order_out[dst] = id;
    node_of_out[dst] = slot;
}

// ---- tier 2: one warp finishes a subtree of <= 32 primitives ------------------------------------------------
struct SmallShared {
    float mn[3][32], mx[3][32], c[3][32];
    uint32_t id[32]; // bit 31: sphere
    uint32_t st_range[34], st_rec[34], st_depth[34]; // range = b | e << 8
};
__device__ __forceinline__ float warp_min_f(float v) { return o2f(__reduce_min_sync(0xFFFFFFFFu, f2o(v))); }
__device__ __forceinline__ float warp_max_f(float v) { return o2f(__reduce_max_sync(0xFFFFFFFFu, f2o(v))); }

constexpr int kSmallWarps = 4;
__global__ void __launch_bounds__(32 * kSmallWarps) build_small(const float4* __restrict__ bmn, const float4* __restrict__ bmx,
                                                                uint32_t* order, const WorkNode* __restrict__ small, uint32_t count,
                                                                const uint32_t* __restrict__ pair_offset, uint32_t pair_base,
                                                                GpuNode* records, GpuNode* root_record, uint32_t leaf_max, float trav_cost,
                                                                uint32_t* max_depth) {
    __shared__ SmallShared shared[kSmallWarps];
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    const uint32_t item = blockIdx.x * kSmallWarps + w;
    if (item >= count) return;
    SmallShared& S = shared[w];
    const WorkNode wn = small[item];
    if (lane < wn.n) {
        const uint32_t id = order[wn.start + lane];
        const float4 a = bmn[id], b = bmx[id];
        S.mn[0][lane] = a.x; S.mn[1][lane] = a.y; S.mn[2][lane] = a.z;
        S.mx[0][lane] = b.x; S.mx[1][lane] = b.y; S.mx[2][lane] = b.z;
        S.c[0][lane] = 0.5f * (a.x + b.x); S.c[1][lane] = 0.5f * (a.y + b.y); S.c[2][lane] = 0.5f * (a.z + b.z);
        S.id[lane] = id | (__float_as_uint(b.w) ? 0x80000000u : 0u);
    }
    uint32_t next_pair = pair_base + pair_offset[item];
    uint32_t deepest = 0;
    int sp = 0;
    if (lane == 0) { S.st_range[0] = 0u | (wn.n << 8); S.st_rec[0] = wn.rec; S.st_depth[0] = wn.depth; }
    sp = 1;
    __syncwarp();
    while (sp > 0) {
        --sp;
        const uint32_t range = S.st_range[sp], rec = S.st_rec[sp], depth = S.st_depth[sp];
        const uint32_t b = range & 0xFFu, e = range >> 8, m = e - b;
        const bool in = lane >= b && lane < e;
        // node box and centroid box
        float bx_mn[3], bx_mx[3], cb_mn[3], cb_mx[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            // NaN bounds are skipped like fminf / fmaxf skip them
            const float lo = in ? S.mn[k][lane] : kMaxF, hi = in ? S.mx[k][lane] : -kMaxF, cc = in ? S.c[k][lane] : kMaxF;
            bx_mn[k] = warp_min_f(lo == lo ? lo : kMaxF);
            bx_mx[k] = warp_max_f(hi == hi ? hi : -kMaxF);
            cb_mn[k] = warp_min_f(cc == cc ? cc : kMaxF);
            const float c2 = in ? S.c[k][lane] : -kMaxF;
            cb_mx[k] = warp_max_f(c2 == c2 ? c2 : -kMaxF);
        }
        GpuNode g;
        g.min_x = bx_mn[0]; g.min_y = bx_mn[1]; g.min_z = bx_mn[2];
        g.max_x = bx_mx[0]; g.max_y = bx_mx[1]; g.max_z = bx_mx[2];
        const uint32_t first_kind = S.id[b] >> 31;
        const bool mixed = __any_sync(0xFFFFFFFFu, in && (S.id[lane] >> 31) != first_kind);
        bool leaf = m <= 1;
        int best_axis = -1;
        uint32_t best_bin = 0;
        float best_cost = kMaxF;
        if (!leaf) {
            // lane k (1..31) evaluates the plane between bin k-1 and bin k on each axis
#pragma unroll 1
            for (int ax = 0; ax < 3; ++ax) {
                const float lo = cb_mn[ax], hi = cb_mx[ax];
                if (!(hi > lo)) continue;
                const float scale = (float)kGpuBins / (hi - lo);
                BoxF L, R;
                box_reset(L);
                box_reset(R);
                uint32_t nl = 0, nr = 0;
                for (uint32_t j = b; j < e; ++j) {
                    const int k = bin_of(S.c[ax][j], lo, scale);
                    const float pmn[3] = {S.mn[0][j], S.mn[1][j], S.mn[2][j]}, pmx[3] = {S.mx[0][j], S.mx[1][j], S.mx[2][j]};
                    if (k < (int)lane) { box_grow(L, pmn, pmx); ++nl; } else { box_grow(R, pmn, pmx); ++nr; }
                }
                float cost = kMaxF;
                if (lane >= 1 && nl != 0 && nr != 0) cost = box_area(L) * (float)nl + box_area(R) * (float)nr;
                // lowest cost, ties to the lowest plane (the host sweep keeps the first minimum)
                const uint32_t key = f2o(cost);
                const uint32_t mk = __reduce_min_sync(0xFFFFFFFFu, key);
                const uint32_t who = __ffs(__ballot_sync(0xFFFFFFFFu, key == mk)) - 1u;
                const float c = o2f(mk);
                if (c < best_cost) { best_cost = c; best_axis = ax; best_bin = who; }
            }
            if (m <= leaf_max) {
                BoxF pb;
                for (int k = 0; k < 3; ++k) { pb.mn[k] = bx_mn[k]; pb.mx[k] = bx_mx[k]; }
                const float parent_area = box_area(pb);
                const float leaf_cost = (float)m * parent_area;
                const float split_cost = trav_cost * parent_area + best_cost;
                if (!mixed && (best_axis < 0 || leaf_cost <= split_cost)) leaf = true;
            }
        }
        if (leaf) {
            g.link0 = kLeafRef | ((m - 1u) << 28) | (first_kind ? kSphereRef : 0u) | (wn.start + b);
            g.link1 = m;
            deepest = max(deepest, depth);
        } else {
            // stable partition of [b, e) inside the warp
            uint32_t mid = b + m / 2;
            if (best_axis >= 0) {
                const float lo = cb_mn[best_axis], scale = (float)kGpuBins / (cb_mx[best_axis] - lo);
                const bool left = in && (uint32_t)bin_of(S.c[best_axis][lane], lo, scale) < best_bin;
                const uint32_t lmask = __ballot_sync(0xFFFFFFFFu, left), imask = __ballot_sync(0xFFFFFFFFu, in);
                const uint32_t nl = __popc(lmask);
                if (nl != 0 && nl != m) {
                    const uint32_t below = (1u << lane) - 1u;
                    const uint32_t dst = left ? b + __popc(lmask & below) : b + nl + __popc(imask & ~lmask & below);
                    float v[9];
                    uint32_t vid = 0;
                    if (in) {
                        for (int k = 0; k < 3; ++k) { v[k] = S.mn[k][lane]; v[3 + k] = S.mx[k][lane]; v[6 + k] = S.c[k][lane]; }
                        vid = S.id[lane];
                    }
                    __syncwarp();
                    if (in) {
                        for (int k = 0; k < 3; ++k) { S.mn[k][dst] = v[k]; S.mx[k][dst] = v[3 + k]; S.c[k][dst] = v[6 + k]; }
                        S.id[dst] = vid;
                    }
                    mid = b + nl;
                }
            }
            const uint32_t pair = next_pair++;
            g.link0 = 2 * pair;
            g.link1 = 0;
            if (lane == 0) { // right below left: the left child is popped first (depth-first)
                S.st_range[sp] = mid | (e << 8); S.st_rec[sp] = 2 * pair + 1; S.st_depth[sp] = depth + 1;
                S.st_range[sp + 1] = b | (mid << 8); S.st_rec[sp + 1] = 2 * pair; S.st_depth[sp + 1] = depth + 1;
            }
            sp += 2;
        }
        if (lane == 0) *(rec == kRootRec ? root_record : records + rec) = g;
        __syncwarp();
    }
    if (lane < wn.n) order[wn.start + lane] = S.id[lane] & 0x7FFFFFFFu;
    if (lane == 0) atomicMax(max_depth, deepest);
}

// 64-byte records in leaf order (see fast_bvh.cpp): triangles {n, d | v0, id | v1, reference leaf | v2, n.n} with the
// operation order of Triangle::intersects (rayca-geometry/src/triangle.rs:90-94,110,152); spheres are copied
// from the host-made records.
__global__ void emit_records(const uint32_t* __restrict__ order, uint32_t n, const float4* __restrict__ tri_pos,
                             const uint32_t* __restrict__ ref_leaf, const float4* __restrict__ bmx,
                             const float4* __restrict__ sphere_records, float4* out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t id = order[i];
    float4* dst = out + (size_t)i * 4;
    if (__float_as_uint(bmx[id].w)) {
        const uint32_t k = __float_as_uint(bmx[id].w) - 1u;
        for (int j = 0; j < 4; ++j) dst[j] = sphere_records[(size_t)k * 4 + j];
        return;
    }
    const float4 p0 = tri_pos[(size_t)id * 3], p1 = tri_pos[(size_t)id * 3 + 1], p2 = tri_pos[(size_t)id * 3 + 2];
    const V3 v0{p0.x, p0.y, p0.z}, v1{p1.x, p1.y, p1.z}, v2{p2.x, p2.y, p2.z};
    const V3 nn = cross(v1 - v0, v2 - v0); // rmath.h: the reference's operation order
    const float d = -dot(nn, v0);
    const float den = dot(nn, nn);
    const float nx = nn.x, ny = nn.y, nz = nn.z;
    dst[0] = make_float4(nx, ny, nz, d);
    dst[1] = make_float4(p0.x, p0.y, p0.z, __uint_as_float(id));
    dst[2] = make_float4(p1.x, p1.y, p1.z, __uint_as_float(ref_leaf[id]));
    dst[3] = make_float4(p2.x, p2.y, p2.z, den);
}
// =====================================================================================================
// Reference BLAS on the device: the tree of rayca-soft/src/bvh/blas.rs:64-127,245-316, node for node and
// primitive order for primitive order (see bvh_build.cpp for the host version and why binning by candidate
// index is exact).  The reference builder stops early (its candidate boxes start at the origin, A-21), so
// its trees have a few hundred nodes over very large ranges: every level is a handful of passes over all
// primitives, which is what the host spends 110 ms on for 262k triangles.
//   * 63 planes x 3 axes: a primitive is LEFT of plane i iff centroid < lo + i * scale; k = first such i (binary
//     search with the same comparisons), bins k in [1, 64]; prefix / suffix unions give every candidate's boxes.
//   * the in-place swap partition (blas.rs:283-291) is a fixed permutation of the range given the left/right
//     flags: lefts below nL stay, the k-th right at or below nL (ascending) goes to t(k-1) - 1, the k-th left
//     above nL (descending, positions t1 > t2 > ...) goes into the k-th hole, every other right moves down by
//     one (t0 = n).  Checked against the sequential loop in tests/test_partition_permutation.py.
//   * nodes get temporary ids level by level; the host renumbers them in the reference's depth-first push order.
constexpr int kRefPlanes = 64;
constexpr int kRefBins = kRefPlanes + 1;                  // k in [1, 64]; entry 0 unused
constexpr int kRefNodeBinWords = 3 * kRefBins * kBinWords;
struct RefNode {
    float mn[3], mx[3];
    uint32_t start, n, left, level; // left: temporary id of the left child (right = left + 1), kNone for a leaf
};
struct RefInfo {
    float lo[3], scale[3];
    uint32_t active;
};
struct RefDecision {
    uint32_t axis;
    float pos;
    uint32_t nl;
    uint32_t partition; // the swap loop runs (cost <= no-split cost)
    uint32_t split;     // both sides non-empty: two children
    uint32_t child_slot[2];
};
__device__ __forceinline__ int ref_bin_of(float c, float lo, float scale) {
    int a = 1, b = kRefPlanes;
    while (a < b) {
        const int mid = (a + b) >> 1;
        if (c < lo + (float)mid * scale) b = mid; else a = mid + 1;
    }
    return a;
}
// one block per node; `bins` (when given) are initialised here as well (was a launch of its own)
__global__ void ref_prepare(const uint32_t* __restrict__ acc, const uint32_t* __restrict__ active_ids, uint32_t count, RefNode* nodes,
                            RefInfo* info, uint32_t* bins) {
    const uint32_t s = blockIdx.x;
    if (s >= count) return;
    if (bins)
        for (uint32_t i = threadIdx.x; i < (uint32_t)kRefNodeBinWords; i += blockDim.x) {
            const uint32_t k = i % kBinWords;
            bins[(size_t)s * kRefNodeBinWords + i] = k < 3 ? kOrdMinInit : (k < 6 ? kOrdMaxInit : 0u);
        }
    if (threadIdx.x != 0) return;
    const uint32_t* a = acc + (size_t)s * kAcc;
    RefNode& nd = nodes[active_ids[s]];
    RefInfo ri;
    ri.active = 0;
    for (int k = 0; k < 3; ++k) {
        const float lo = o2f(a[k]), hi = o2f(a[3 + k]);
        nd.mn[k] = lo;
        nd.mx[k] = hi;
        ri.lo[k] = lo;
        ri.scale[k] = 0.0f;
        if (!(lo == hi)) {
            ri.active |= 1u << k;
            ri.scale[k] = (hi - lo) / (float)kRefPlanes;
        }
    }
    info[s] = ri;
}
__global__ void __launch_bounds__(kBinThreads) ref_bin_prims(const float4* __restrict__ bmn, const float4* __restrict__ bmx,
                                                           const float4* __restrict__ cen, const uint32_t* __restrict__ order,
                                                           const uint32_t* __restrict__ node_of, uint32_t n,
                                                           const RefInfo* __restrict__ info, uint32_t* bins) {
    __shared__ uint32_t local[kRefNodeBinWords];
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t s = p < n ? node_of[p] : kNone;
    const uint32_t s0 = node_of[blockIdx.x * blockDim.x];
    const bool uniform = __syncthreads_and(s == s0 || p >= n) && s0 != kNone;
    if (uniform) {
        for (uint32_t i = threadIdx.x; i < (uint32_t)kRefNodeBinWords; i += kBinThreads) {
            const uint32_t k = i % kBinWords;
            local[i] = k < 3 ? kOrdMinInit : (k < 6 ? kOrdMaxInit : 0u);
        }
        __syncthreads();
    }
    if (s != kNone) {
        const uint32_t id = order[p];
        const float4 a = bmn[id], b = bmx[id], c4 = cen[id];
        const float mn[3] = {a.x, a.y, a.z}, mx[3] = {b.x, b.y, b.z}, c[3] = {c4.x, c4.y, c4.z};
        const RefInfo& ri = info[s];
        const uint32_t active = ri.active;
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) {
            if (!(active >> ax & 1u)) continue;
            const int k = ref_bin_of(c[ax], ri.lo[ax], ri.scale[ax]);
            uint32_t* w = (uniform ? local : bins + (size_t)s * kRefNodeBinWords) + ((size_t)ax * kRefBins + k) * kBinWords;
            atomic_box(w, mn, mx);
            atomicAdd(w + 6, 1u);
        }
    }
    if (!uniform) return;
    __syncthreads();
    uint32_t* g = bins + (size_t)s0 * kRefNodeBinWords;
    for (uint32_t i = threadIdx.x; i < (uint32_t)kRefNodeBinWords; i += kBinThreads) {
        const uint32_t bin = i / kBinWords, k = i % kBinWords;
        if (local[bin * kBinWords + 6] == 0) continue;
        if (k < 3) atomicMin(g + i, local[i]);
        else if (k < 6) atomicMax(g + i, local[i]);
        else atomicAdd(g + i, local[i]);
    }
}
__device__ __forceinline__ float ref_area(const BoxF& b) { // aabb.rs:20-23
    const float ex = b.mx[0] - b.mn[0], ey = b.mx[1] - b.mn[1], ez = b.mx[2] - b.mn[2];
    return ex * ey + ey * ez + ez * ex;
}
__device__ __forceinline__ void ref_origin(BoxF& b) { // AABB::default(): both corners at the origin (blas.rs:66-67)
    for (int k = 0; k < 3; ++k) b.mn[k] = b.mx[k] = 0.0f;
}
__device__ __forceinline__ void ref_merge_bin(BoxF& acc, const uint32_t* b) {
    const float lo[3] = {o2f(b[0]), o2f(b[1]), o2f(b[2])}, hi[3] = {o2f(b[3]), o2f(b[4]), o2f(b[5])};
    box_grow(acc, lo, hi);
}
// find_best_split_plane + the no-split test (blas.rs:93-127,270-276).  One block per node, one WARP per axis: lane l
// holds bins 2l+1 and 2l+2, prefix / suffix unions of boxes and counts are warp scans, every lane prices its two planes
// and a butterfly picks the cheapest (lowest plane on equal cost: the reference keeps the first minimum of its
// ascending sweep, strict `<`).  Unions of boxes and sums of counts are exact whatever the association, so the costs
// are the bits the sequential sweep gets.  (The first version ran one THREAD per (node, axis) over two dependent
// sweeps of 63 bins in global memory: 65 us per level whatever the node count, 1.2 of the 2.7 ms of a 262k-triangle
// build.)
__device__ __forceinline__ BoxF box_union(const BoxF& a, const BoxF& b) {
    BoxF r;
    for (int k = 0; k < 3; ++k) { r.mn[k] = fminf(a.mn[k], b.mn[k]); r.mx[k] = fmaxf(a.mx[k], b.mx[k]); }
    return r;
}
__device__ __forceinline__ BoxF box_shfl_up(const BoxF& a, int d) {
    BoxF r;
    for (int k = 0; k < 3; ++k) { r.mn[k] = __shfl_up_sync(0xFFFFFFFFu, a.mn[k], d); r.mx[k] = __shfl_up_sync(0xFFFFFFFFu, a.mx[k], d); }
    return r;
}
__device__ __forceinline__ BoxF box_shfl_down(const BoxF& a, int d) {
    BoxF r;
    for (int k = 0; k < 3; ++k) { r.mn[k] = __shfl_down_sync(0xFFFFFFFFu, a.mn[k], d); r.mx[k] = __shfl_down_sync(0xFFFFFFFFu, a.mx[k], d); }
    return r;
}
__device__ __forceinline__ void ref_load_bin(const uint32_t* w, int k, BoxF& b, uint32_t& c) {
    const uint32_t* p = w + k * kBinWords;
    c = p[6];
    box_reset(b); // an empty bin is the identity of the union (the sequential sweep skips it)
    if (c) ref_merge_bin(b, p);
}
__device__ __forceinline__ float ref_candidate_cost(const BoxF& left, uint32_t nl, const BoxF& right, uint32_t nr) {
    BoxF l, r;
    ref_origin(l); // candidate boxes start at the origin (AABB::default(), blas.rs:66-67)
    ref_origin(r);
    l = box_union(l, left);
    r = box_union(r, right);
    float cost = (float)(int)nl * ref_area(l) + (float)(int)nr * ref_area(r); // evaluate_sah (blas.rs:83-88)
    if (!(cost > 0.0f)) cost = kMaxF;
    return cost;
}
__global__ void __launch_bounds__(96) ref_decide(const uint32_t* __restrict__ active_ids, uint32_t count, const RefNode* __restrict__ nodes,
                                                 const RefInfo* __restrict__ info, const uint32_t* __restrict__ bins, RefDecision* dec,
                                                 uint32_t* split_flags) {
    __shared__ float s_cost[3];
    __shared__ uint32_t s_plane[3], s_nl[3];
    const uint32_t s = blockIdx.x, ax = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    if (s >= count) return;
    float best_cost = kMaxF;
    uint32_t best_plane = 0, best_nl = 0;
    if (info[s].active >> ax & 1u) { // the same for the whole warp
        const uint32_t* w = bins + ((size_t)s * 3 + ax) * kRefBins * kBinWords;
        BoxF b1, b2;
        uint32_t c1, c2;
        ref_load_bin(w, 2 * (int)lane + 1, b1, c1);
        ref_load_bin(w, 2 * (int)lane + 2, b2, c2);
        const BoxF both = box_union(b1, b2);
        const uint32_t cboth = c1 + c2;
        // bins of the lanes below (exclusive prefix) and above (exclusive suffix)
        BoxF up = both, down = both;
        uint32_t cup = cboth, cdown = cboth;
        for (int d = 1; d < 32; d <<= 1) {
            const BoxF ou = box_shfl_up(up, d), od = box_shfl_down(down, d);
            const uint32_t cu = __shfl_up_sync(0xFFFFFFFFu, cup, d), cd = __shfl_down_sync(0xFFFFFFFFu, cdown, d);
            if ((int)lane >= d) { up = box_union(up, ou); cup += cu; }
            if ((int)lane + d < 32) { down = box_union(down, od); cdown += cd; }
        }
        BoxF below = box_shfl_up(up, 1), above = box_shfl_down(down, 1);
        uint32_t cbelow = __shfl_up_sync(0xFFFFFFFFu, cup, 1), cabove = __shfl_down_sync(0xFFFFFFFFu, cdown, 1);
        if (lane == 0) { box_reset(below); cbelow = 0; }
        if (lane == 31) { box_reset(above); cabove = 0; }
        // plane i: left = bins 1..i, right = bins i+1..64
        // i = 2l+1: left = below U b1, right = b2 U above;  i = 2l+2 (l < 31): left = below U both, right = above
        {
            const uint32_t nl = cbelow + c1;
            const float cost = ref_candidate_cost(box_union(below, b1), nl, box_union(b2, above), c2 + cabove);
            if (cost < best_cost) { best_cost = cost; best_plane = 2 * lane + 1; best_nl = nl; }
        }
        if (lane < 31) {
            const uint32_t nl = cbelow + cboth;
            const float cost = ref_candidate_cost(box_union(below, both), nl, above, cabove);
            if (cost < best_cost) { best_cost = cost; best_plane = 2 * lane + 2; best_nl = nl; }
        }
        for (int d = 16; d > 0; d >>= 1) { // cheapest plane, the lowest one on equal cost
            const float oc = __shfl_xor_sync(0xFFFFFFFFu, best_cost, d);
            const uint32_t op = __shfl_xor_sync(0xFFFFFFFFu, best_plane, d), on = __shfl_xor_sync(0xFFFFFFFFu, best_nl, d);
            if (oc < best_cost || (oc == best_cost && op < best_plane)) { best_cost = oc; best_plane = op; best_nl = on; }
        }
    }
    if (lane == 0) {
        s_cost[ax] = best_cost;
        s_plane[ax] = best_plane;
        s_nl[ax] = best_nl;
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    float cost = kMaxF;
    int axis = 0;
    uint32_t plane = 0, nl = 0;
    bool any = false;
    for (int a = 0; a < 3; ++a)
        if (s_cost[a] < cost) { cost = s_cost[a]; axis = a; plane = s_plane[a]; nl = s_nl[a]; any = true; }
    const RefNode& nd = nodes[active_ids[s]];
    const RefInfo& ri = info[s];
    BoxF nb;
    for (int k = 0; k < 3; ++k) { nb.mn[k] = nd.mn[k]; nb.mx[k] = nd.mx[k]; }
    const float no_split = (float)nd.n * ref_area(nb); // calculate_cost (blas.rs:125-127)
    RefDecision d;
    d.axis = (uint32_t)axis;
    d.pos = any ? ri.lo[axis] + (float)plane * ri.scale[axis] : 0.0f;
    d.partition = cost > no_split ? 0u : 1u;
    // without a winning plane every centroid compares `< 0.0` (split_pos stays 0.0 on axis 0): count the lefts later
    d.nl = any ? nl : kNone;
    d.split = 0;
    d.child_slot[0] = d.child_slot[1] = kNone;
    dec[s] = d;
    split_flags[s] = 0;
}
// left / right flag of every primitive of a partitioned node: centroid[axis] < pos (blas.rs:284)
__global__ void ref_flags(const float4* __restrict__ cen, const uint32_t* __restrict__ order, const uint32_t* __restrict__ node_of,
                          uint32_t n, const RefDecision* __restrict__ dec, uint32_t* flags) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const uint32_t s = node_of[p];
    uint32_t f = 0;
    if (s != kNone && dec[s].partition) {
        const float4 c = cen[order[p]];
        const uint32_t ax = dec[s].axis;
        f = (ax == 0 ? c.x : (ax == 1 ? c.y : c.z)) < dec[s].pos;
    }
    flags[p] = f;
}
// after the scan: left counts, which nodes really split
__global__ void ref_counts(const uint32_t* __restrict__ active_ids, uint32_t count, const RefNode* __restrict__ nodes,
                           const uint32_t* __restrict__ scan, RefDecision* dec, uint32_t* split_flags) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= count) return;
    const RefNode& nd = nodes[active_ids[s]];
    RefDecision& d = dec[s];
    if (!d.partition) { split_flags[s] = 0; return; }
    d.nl = scan[nd.start + nd.n] - scan[nd.start];
    d.split = d.nl > 0 && d.nl < nd.n;
    split_flags[s] = d.split;
}
__global__ void ref_children(const uint32_t* __restrict__ active_ids, uint32_t count, RefNode* nodes, RefDecision* dec,
                             const uint32_t* __restrict__ split_index, uint32_t next_id, uint32_t level, uint32_t* next_active, uint32_t* next_acc) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= count) return;
    RefDecision& d = dec[s];
    RefNode& nd = nodes[active_ids[s]];
    if (!d.split) return;
    const uint32_t k = split_index[s];
    const uint32_t left = next_id + 2 * k;
    nd.left = left;
    for (uint32_t c = 0; c < 2; ++c) {
        RefNode ch;
        for (int j = 0; j < 3; ++j) { ch.mn[j] = kMaxF; ch.mx[j] = -kMaxF; }
        ch.start = c ? nd.start + d.nl : nd.start;
        ch.n = c ? nd.n - d.nl : d.nl;
        ch.left = kNone;
        ch.level = level + 1;
        nodes[left + c] = ch;
        next_active[2 * k + c] = left + c;
        d.child_slot[c] = 2 * k + c;
        for (uint32_t j = 0; j < (uint32_t)kAcc; ++j) next_acc[(size_t)(2 * k + c) * kAcc + j] = (j % 6u) < 3u ? kOrdMinInit : kOrdMaxInit; // was init_acc
    }
}
// the permutation of the swap loop, part 1: where the holes and the tail lefts are
__global__ void ref_tables(const uint32_t* __restrict__ node_of, uint32_t n, const uint32_t* __restrict__ active_ids,
                           const RefNode* __restrict__ nodes, const RefDecision* __restrict__ dec, const uint32_t* __restrict__ flags,
                           const uint32_t* __restrict__ scan, uint32_t* hole_pos, uint32_t* tail_pos) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const uint32_t s = node_of[p];
    if (s == kNone || !dec[s].partition) return;
    const RefNode& nd = nodes[active_ids[s]];
    const uint32_t nl = dec[s].nl, q = p - nd.start, lb = scan[p] - scan[nd.start];
    if (flags[p]) {
        if (q >= nl) tail_pos[nd.start + (nl - lb) - 1] = q; // k-th left above nL, counted from the end
    } else if (q <= nl) {
        hole_pos[nd.start + (q - lb)] = q; // k-th right at or below nL, k - 1 = rights before it
    }
}
__global__ void ref_scatter(const uint32_t* __restrict__ order, const uint32_t* __restrict__ node_of, uint32_t n,
                            const uint32_t* __restrict__ active_ids, const RefNode* __restrict__ nodes,
                            const RefDecision* __restrict__ dec, const uint32_t* __restrict__ flags, const uint32_t* __restrict__ scan,
                            const uint32_t* __restrict__ hole_pos, const uint32_t* __restrict__ tail_pos, uint32_t* order_out,
                            uint32_t* node_of_out) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const uint32_t s = node_of[p];
    const uint32_t id = order[p];
    if (s == kNone || !dec[s].partition) {
        order_out[p] = id;
        node_of_out[p] = kNone;
        return;
    }
    const RefNode& nd = nodes[active_ids[s]];
    const RefDecision& d = dec[s];
    const uint32_t nl = d.nl, q = p - nd.start, lb = scan[p] - scan[nd.start];
    uint32_t dst;
    if (flags[p]) {
        dst = q < nl ? q : hole_pos[nd.start + (nl - lb) - 1];
    } else if (q <= nl) {
        const uint32_t k = q - lb; // rights before this one
        dst = (k == 0 ? nd.n : tail_pos[nd.start + k - 1]) - 1u;
    } else {
        dst = q - 1u;
    }
    order_out[nd.start + dst] = id;
    node_of_out[nd.start + dst] = d.split ? d.child_slot[dst < nl ? 0 : 1] : kNone;
}

__global__ void small_sizes(const WorkNode* __restrict__ small, uint32_t count, uint32_t* sizes) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) sizes[i] = small[i].n - 1u; // pairs a subtree over n primitives can need
}

// One allocation for all temporaries (35 separate allocations cost more than the build), taken from the device's
// stream-ordered pool on the build's stream: `Draw::draw` rebuilds per call, and the pool (release threshold raised in
// rayca_cuda_init) hands the same memory back without a cudaMalloc / cudaFree round trip (3 ms per draw before).
struct DeviceBuffers {
    struct Want { void** p; size_t offset; };
    std::vector<Want> wants;
    size_t bytes = 0;
    void* base = nullptr;
    cudaStream_t st;
    std::vector<void*> kept; // separately allocated outputs, freed unless released
    explicit DeviceBuffers(cudaStream_t stream) : st(stream) {}
    ~DeviceBuffers() { // the builders synchronise `st` before returning: nothing is in flight here
        if (base) cudaFreeAsync(base, st);
        for (void* p : kept) cudaFreeAsync(p, st);
    }
    template <class T> cudaError_t alloc(T** out, size_t count) { // temporaries: address assigned by commit()
        wants.push_back({(void**)out, bytes});
        bytes += (std::max<size_t>(count, 1) * sizeof(T) + 255) & ~size_t(255);
        return cudaSuccess;
    }
    cudaError_t commit() {
        cudaError_t e = cudaMallocAsync(&base, bytes, st);
        if (e != cudaSuccess) return e;
        for (const Want& w : wants) *w.p = (char*)base + w.offset;
        return cudaSuccess;
    }
    template <class T> cudaError_t alloc_output(T** out, size_t count) {
        void* p = nullptr;
        cudaError_t e = cudaMallocAsync(&p, std::max<size_t>(count, 1) * sizeof(T), st);
        if (e == cudaSuccess) kept.push_back(p);
        *out = (T*)p;
        return e;
    }
    void release(void* p) { kept.erase(std::remove(kept.begin(), kept.end(), p), kept.end()); }
};

inline uint32_t blocks_for(size_t n, uint32_t threads) { return (uint32_t)((n + threads - 1) / threads); }

} // namespace

#define FB_TRY(expr)                                                                 \
    do {                                                                             \
        cudaError_t e__ = (expr);                                                    \
        if (e__ != cudaSuccess) {                                                    \
            err = std::string("device tree build: ") + cudaGetErrorString(e__);     \
            return e__ == cudaErrorMemoryAllocation ? RAYCA_E_OOM : RAYCA_E_CUDA;    \
        }                                                                            \
    } while (0)

// Triangle-only scenes: primitive bounds from the world-space vertices already on the device (the same fminf / fmaxf
// chain as the host's emit, host_scene.cpp: min and max are exact, so the bits are the host's) and the identity order —
// instead of building both on the host and uploading 36 bytes per primitive.
__global__ void prep_bounds_kernel(const float4* __restrict__ tri_pos, uint32_t n, float4* __restrict__ mn, float4* __restrict__ mx,
                                   uint32_t* __restrict__ order) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = tri_pos[3 * (size_t)i], b = tri_pos[3 * (size_t)i + 1], c = tri_pos[3 * (size_t)i + 2];
    mn[i] = make_float4(fminf(fminf(fminf(kF32Max, a.x), b.x), c.x), fminf(fminf(fminf(kF32Max, a.y), b.y), c.y),
                        fminf(fminf(fminf(kF32Max, a.z), b.z), c.z), 0.0f);
    mx[i] = make_float4(fmaxf(fmaxf(fmaxf(-kF32Max, a.x), b.x), c.x), fmaxf(fmaxf(fmaxf(-kF32Max, a.y), b.y), c.y),
                        fmaxf(fmaxf(fmaxf(-kF32Max, a.z), b.z), c.z), 0.0f); // w: sphere id, none here
    order[i] = i;
}

// reference leaf of every primitive from the packed reference nodes already on the device (link0 = first primitive,
// link1 = count for a leaf): one warp per node record of the BLAS range [0, blas_nodes) — the TLAS pairs behind it hold
// COPIES of BLAS roots and are skipped
__global__ void ref_leaf_kernel(const GpuNode* __restrict__ nodes, uint32_t blas_nodes, uint32_t* __restrict__ ref_leaf) {
    const uint32_t k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (k >= blas_nodes) return;
    const uint32_t first = nodes[k].link0, count = nodes[k].link1;
    for (uint32_t i = lane; i < count; i += 32u) ref_leaf[first + i] = k;
}

// the reference leaf (index into HostScene::nodes) of every primitive, reference-global primitive order
void reference_leaves(const HostScene& hs, PodVector<uint32_t>& ref_leaf) {
    const size_t nb = hs.blas.size();
    ref_leaf.resize(hs.blas_prim_offset[nb]); // every primitive lies in exactly one reference leaf: all entries are written
    for (size_t bi = 0; bi < nb; ++bi) {
        const HostBlas& b = hs.blas[bi];
        const uint32_t po = hs.blas_prim_offset[bi];
        // leaves own disjoint primitive ranges: the nodes are spread over the threads (a BLAS built without a tree is
        // one leaf holding everything, hence the inner loop is the parallel one then)
        if (b.nodes.size() < 64) {
            for (size_t k = 0; k < b.nodes.size(); ++k) {
                const HostNode& n = b.nodes[k];
                if (k == 1 || n.count == 0) continue;
                const uint32_t leaf = hs.blas_node_offset[bi] + (uint32_t)k;
#pragma omp parallel for schedule(static) if (n.count > 8192)
                for (int64_t i = 0; i < (int64_t)n.count; ++i) ref_leaf[po + n.offset + (size_t)i] = leaf;
            }
        } else {
#pragma omp parallel for schedule(static, 4096)
            for (int64_t k = 0; k < (int64_t)b.nodes.size(); ++k) {
                const HostNode& n = b.nodes[(size_t)k];
                if (k == 1 || n.count == 0) continue;
                const uint32_t leaf = hs.blas_node_offset[bi] + (uint32_t)k;
                for (uint32_t i = 0; i < n.count; ++i) ref_leaf[po + n.offset + (size_t)i] = leaf;
            }
        }
    }
}

int build_fast_tree_device(const HostScene& hs, const float4* d_tri_pos, cudaStream_t st, FastTreeDevice& out, std::string& err,
                           const void* d_ref_nodes) {
    int bins_unused;
    uint32_t leaf_max;
    float trav_cost;
    fast_tree_tunables(bins_unused, leaf_max, trav_cost);
    const size_t nb = hs.blas.size();
    const uint32_t N = hs.blas_prim_offset[nb];
    if (N == 0) { err = "no primitives"; return RAYCA_E_INVALID_ARG; }
    const bool timing = std::getenv("RAYCA_TIMING") != nullptr;
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    double t_lap = now();
    auto lap = [&](const char* what) {
        if (!timing) return;
        const double t = now();
        std::fprintf(stderr, "[rayca]     %-26s %8.2f ms\n", what, t - t_lap);
        t_lap = t;
    };

    // ---- host side: bounds in reference-global order, reference leaf of every primitive, sphere records
    bool any_sphere = false;
    for (size_t bi = 0; bi < nb; ++bi) any_sphere = any_sphere || !hs.blas[bi].spheres.empty();
    std::vector<float4> hmn(any_sphere ? N : 0), hmx(any_sphere ? N : 0); // sphere scenes: bounds from the host (a sphere's are not its record's)
    // triangle-only scenes with the reference nodes on the device: the table is made there (ref_leaf_kernel); sphere
    // scenes need it on the host for their sphere records
    const bool leaves_on_device = d_ref_nodes != nullptr && !any_sphere;
    PodVector<uint32_t> ref_leaf;
    if (!leaves_on_device) reference_leaves(hs, ref_leaf);
    std::vector<float4> sphere_records;
    std::vector<uint32_t> sphere_base(nb + 1, 0);
    for (size_t bi = 0; bi < nb; ++bi) sphere_base[bi + 1] = sphere_base[bi] + (uint32_t)hs.blas[bi].spheres.size();
    sphere_records.resize((size_t)sphere_base[nb] * 4);
    for (size_t bi = 0; bi < nb; ++bi) {
        const HostBlas& b = hs.blas[bi];
        const uint32_t po = hs.blas_prim_offset[bi];
#pragma omp parallel for schedule(static) if (any_sphere && b.tris.size() > 8192)
        for (int64_t k = 0; k < (int64_t)(any_sphere ? b.tris.size() : 0); ++k) {
            const HostTri& t = b.tris[(size_t)k];
            const uint32_t id = po + (uint32_t)k;
            uint32_t sphere = 0;
            if (t.sphere) {
                sphere = sphere_base[bi] + t.sphere; // 1-based, scene-global
                const HostSphere& sp = b.spheres[t.sphere - 1];
                const V3 rs = reciprocal(sp.world.s);
                const Q4 qi = conj(sp.world.q);
                float4* dst = &sphere_records[(size_t)(sphere - 1) * 4];
                dst[0] = make_float4(-sp.world.t.x, -sp.world.t.y, -sp.world.t.z, sp.radius2);
                dst[1] = make_float4(qi.x, qi.y, qi.z, qi.w);
                dst[2] = make_float4(rs.x, rs.y, rs.z, __builtin_bit_cast(float, id));
                dst[3] = make_float4(sp.center.x, sp.center.y, sp.center.z, 0.0f); // w: reference leaf, below
            }
            hmn[id] = make_float4(t.bmin.x, t.bmin.y, t.bmin.z, 0.0f);
            hmx[id] = make_float4(t.bmax.x, t.bmax.y, t.bmax.z, __builtin_bit_cast(float, sphere));
        }
        for (size_t k = 0; k < (b.spheres.empty() ? 0 : b.tris.size()); ++k) // (a serial scan of every primitive: sphere models only)
            if (b.tris[k].sphere) {
                const uint32_t sphere = sphere_base[bi] + b.tris[k].sphere;
                sphere_records[(size_t)(sphere - 1) * 4 + 3].w = __builtin_bit_cast(float, ref_leaf[po + k]);
            }
    }

    lap("fast: host prep");
    // ---- device buffers
    DeviceBuffers buf(st);
    const uint32_t cap_big = N / (kSmall + 1) + 1;
    float4 *d_mn, *d_mx, *d_sphere, *d_tri;
    uint32_t *d_order[2], *d_node_of[2], *d_flags, *d_scan, *d_partial, *d_ref_leaf, *d_acc[2], *d_bins, *d_child_big, *d_child_small,
        *d_big_index, *d_small_index, *d_sizes, *d_pair_offset, *d_depth;
    WorkNode *d_work[2], *d_small;
    NodeInfo* d_info;
    Decision* d_dec;
    GpuNode *d_records, *d_root;
    FB_TRY(buf.alloc(&d_mn, N));
    FB_TRY(buf.alloc(&d_mx, N));
    FB_TRY(buf.alloc(&d_sphere, sphere_records.size()));
    FB_TRY(buf.alloc_output(&d_tri, (size_t)N * 4));
    for (int k = 0; k < 2; ++k) {
        FB_TRY(buf.alloc(&d_order[k], N));
        FB_TRY(buf.alloc(&d_node_of[k], N));
        FB_TRY(buf.alloc(&d_acc[k], (size_t)cap_big * kAcc));
        FB_TRY(buf.alloc(&d_work[k], cap_big));
    }
    FB_TRY(buf.alloc(&d_flags, N));
    FB_TRY(buf.alloc(&d_scan, (size_t)N + 1));
    FB_TRY(buf.alloc(&d_partial, (size_t)N / kScanBlock + 4));
    FB_TRY(buf.alloc(&d_ref_leaf, N));
    FB_TRY(buf.alloc(&d_bins, (size_t)cap_big * kNodeBinWords));
    FB_TRY(buf.alloc(&d_child_big, (size_t)cap_big * 2));
    FB_TRY(buf.alloc(&d_child_small, (size_t)cap_big * 2));
    FB_TRY(buf.alloc(&d_big_index, (size_t)cap_big * 2 + 1));
    FB_TRY(buf.alloc(&d_small_index, (size_t)cap_big * 2 + 1));
    FB_TRY(buf.alloc(&d_small, N));
    FB_TRY(buf.alloc(&d_sizes, N));
    FB_TRY(buf.alloc(&d_pair_offset, (size_t)N + 1));
    FB_TRY(buf.alloc(&d_depth, 1));
    FB_TRY(buf.alloc(&d_info, cap_big));
    FB_TRY(buf.alloc(&d_dec, cap_big));
    FB_TRY(buf.alloc_output(&d_records, (size_t)N * 2 + 2));
    FB_TRY(buf.alloc(&d_root, 1));
    FB_TRY(buf.commit());
    lap("fast: alloc");

    cudaEvent_t e0, e1;
    FB_TRY(cudaEventCreate(&e0));
    FB_TRY(cudaEventCreate(&e1));
    if (any_sphere) {
        FB_TRY(cudaMemcpyAsync(d_mn, hmn.data(), (size_t)N * sizeof(float4), cudaMemcpyHostToDevice, st));
        FB_TRY(cudaMemcpyAsync(d_mx, hmx.data(), (size_t)N * sizeof(float4), cudaMemcpyHostToDevice, st));
    }
    if (leaves_on_device) {
        const uint32_t blas_nodes = hs.blas_node_offset[nb];
        ref_leaf_kernel<<<blocks_for((size_t)blas_nodes * 32, 256), 256, 0, st>>>(static_cast<const GpuNode*>(d_ref_nodes), blas_nodes, d_ref_leaf);
    } else {
        FB_TRY(cudaMemcpyAsync(d_ref_leaf, ref_leaf.data(), (size_t)N * 4, cudaMemcpyHostToDevice, st));
    }
    if (!sphere_records.empty())
        FB_TRY(cudaMemcpyAsync(d_sphere, sphere_records.data(), sphere_records.size() * sizeof(float4), cudaMemcpyHostToDevice, st));
    FB_TRY(cudaEventRecord(e0, st));
    FB_TRY(cudaMemsetAsync(d_depth, 0, 4, st));
    FB_TRY(cudaMemsetAsync(d_records, 0, ((size_t)N * 2 + 2) * sizeof(GpuNode), st));
    if (any_sphere) { // order = identity
        std::vector<uint32_t> iota(N);
        for (uint32_t i = 0; i < N; ++i) iota[i] = i;
        FB_TRY(cudaMemcpyAsync(d_order[0], iota.data(), (size_t)N * 4, cudaMemcpyHostToDevice, st));
        FB_TRY(cudaStreamSynchronize(st)); // iota goes out of scope
    } else {
        prep_bounds_kernel<<<(N + 255) / 256, 256, 0, st>>>(d_tri_pos, N, d_mn, d_mx, d_order[0]);
    }
    const uint32_t T = 256;
    uint32_t nbig = 0, nsmall = 0, pair_base = 0, levels = 0;
    int cur = 0;
    WorkNode root_work{0u, N, kRootRec, 0u};
    if (N > kSmall) {
        nbig = 1;
        FB_TRY(cudaMemcpyAsync(d_work[0], &root_work, sizeof(root_work), cudaMemcpyHostToDevice, st));
        FB_TRY(cudaMemsetAsync(d_node_of[0], 0, (size_t)N * 4, st));
        init_acc<<<1, 32, 0, st>>>(d_acc[0], 1);
        accumulate_boxes<<<blocks_for(N, T), T, 0, st>>>(d_mn, d_mx, d_order[0], d_node_of[0], N, d_acc[0]);
    } else {
        nsmall = 1;
        FB_TRY(cudaMemcpyAsync(d_small, &root_work, sizeof(root_work), cudaMemcpyHostToDevice, st));
    }
    while (nbig > 0) {
        const int nxt = cur ^ 1;
        prepare_nodes<<<nbig, 224, 0, st>>>(d_acc[cur], nbig, d_info, d_bins);
        bin_prims<<<blocks_for(N, kBinThreads), kBinThreads, 0, st>>>(d_mn, d_mx, d_order[cur], d_node_of[cur], N, d_info, d_bins);
        decide_splits<<<nbig, 96, 0, st>>>(d_work[cur], nbig, d_info, d_bins, pair_base, d_dec, d_records, d_root,
                                                           d_child_big, d_child_small);
        exclusive_scan(d_child_big, nbig * 2, d_big_index, d_partial, st);
        exclusive_scan(d_child_small, nbig * 2, d_small_index, d_partial, st);
        uint32_t counts[2];
        FB_TRY(cudaMemcpyAsync(&counts[0], d_big_index + nbig * 2, 4, cudaMemcpyDeviceToHost, st));
        FB_TRY(cudaMemcpyAsync(&counts[1], d_small_index + nbig * 2, 4, cudaMemcpyDeviceToHost, st));
        FB_TRY(cudaStreamSynchronize(st));
        if (counts[0] > cap_big || (size_t)nsmall + counts[1] > N) { err = "device tree build: work list overflow"; return RAYCA_E_CUDA; }
        make_children<<<blocks_for((size_t)nbig * 2, T), T, 0, st>>>(d_work[cur], nbig, d_dec, d_big_index, d_small_index, d_work[nxt],
                                                                  d_small, nsmall, d_acc[nxt]);
        left_flags<<<blocks_for(N, T), T, 0, st>>>(d_mn, d_mx, d_order[cur], d_node_of[cur], N, d_work[cur], d_dec, d_flags);
        exclusive_scan(d_flags, N, d_scan, d_partial, st);
        scatter_prims<<<blocks_for(N, T), T, 0, st>>>(d_order[cur], d_node_of[cur], N, d_work[cur], d_dec, d_flags, d_scan, d_order[nxt],
                                                      d_node_of[nxt]);
        if (counts[0]) accumulate_boxes<<<blocks_for(N, T), T, 0, st>>>(d_mn, d_mx, d_order[nxt], d_node_of[nxt], N, d_acc[nxt]);
        pair_base += nbig;
        nsmall += counts[1];
        nbig = counts[0];
        cur = nxt;
        ++levels;
        if (levels > 4096) { err = "device tree build: runaway depth"; return RAYCA_E_CUDA; }
    }
    lap("fast: level loop");
    if (nsmall) {
        small_sizes<<<blocks_for(nsmall, T), T, 0, st>>>(d_small, nsmall, d_sizes);
        exclusive_scan(d_sizes, nsmall, d_pair_offset, d_partial, st);
        build_small<<<blocks_for(nsmall, kSmallWarps), 32 * kSmallWarps, 0, st>>>(d_mn, d_mx, d_order[cur], d_small, nsmall, d_pair_offset,
                                                                              pair_base, d_records, d_root, leaf_max, trav_cost, d_depth);
    }
    emit_records<<<blocks_for(N, T), T, 0, st>>>(d_order[cur], N, d_tri_pos, d_ref_leaf, d_mx, d_sphere, d_tri);
    FB_TRY(cudaEventRecord(e1, st));
    FB_TRY(cudaMemcpyAsync(&out.root, d_root, sizeof(GpuNode), cudaMemcpyDeviceToHost, st));
    FB_TRY(cudaMemcpyAsync(&out.depth, d_depth, 4, cudaMemcpyDeviceToHost, st));
    FB_TRY(cudaStreamSynchronize(st));
    FB_TRY(cudaGetLastError());
    cudaEventElapsedTime(&out.build_ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    out.levels = levels;
    lap("fast: subtrees + records");
    out.small_subtrees = nsmall;
    out.nodes = d_records;
    out.tri = d_tri;
    buf.release(d_records);
    buf.release(d_tri);
    return RAYCA_OK;
}


// AoS sibling pairs {lo0.xyz ref0 | hi0.xyz - | lo1.xyz ref1 | hi1.xyz -} -> plane-major
// {lo0.x lo1.x hi0.x hi1.x | lo0.y lo1.y hi0.y hi1.y | lo0.z lo1.z hi0.z hi1.z | ref0 ref1 w0 w1}, in place (trace.cuh,
// RAYCA_SOA_PAIRS): one thread per pair reads its 64 bytes and writes them back.
__global__ void transpose_pairs_kernel(float4* __restrict__ rec, size_t pairs) {
    const size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= pairs) return;
    float4* q = rec + 4 * p;
    const float4 a0 = q[0], a1 = q[1], b0 = q[2], b1 = q[3];
    q[0] = make_float4(a0.x, b0.x, a1.x, b1.x);
    q[1] = make_float4(a0.y, b0.y, a1.y, b1.y);
    q[2] = make_float4(a0.z, b0.z, a1.z, b1.z);
    q[3] = make_float4(a0.w, b0.w, a1.w, b1.w);
}

int transpose_pairs_device(void* d_nodes, size_t node_count, cudaStream_t st) {
    const size_t pairs = node_count / 2;
    if (pairs == 0) return RAYCA_OK;
    transpose_pairs_kernel<<<(unsigned)((pairs + 255) / 256), 256, 0, st>>>(reinterpret_cast<float4*>(d_nodes), pairs);
    return cudaGetLastError() == cudaSuccess ? RAYCA_OK : RAYCA_E_CUDA;
}

__global__ void iota_kernel(uint32_t* order, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) order[i] = i;
}

int build_blas_device(HostBlas& blas, uint32_t max_depth, cudaStream_t st, std::string& err, float* device_ms, const DeviceBlasInput* dev) {
    const uint32_t N = dev ? dev->n : (uint32_t)blas.tris.size();
    if (N == 0) { err = "no primitives"; return RAYCA_E_INVALID_ARG; }
    const bool timing = std::getenv("RAYCA_TIMING") != nullptr;
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    double t_lap = now();
    auto lap = [&](const char* what) {
        if (!timing) return;
        const double t = now();
        std::fprintf(stderr, "[rayca]     %-26s %8.2f ms\n", what, t - t_lap);
        t_lap = t;
    };
    // PodVector: recycled, page-locked blocks without value initialisation (host_scene.h) — three std::vectors of N
    // float4 cost 1.4 ms of page faults and memset per build, and their uploads were pageable
    PodVector<float4> hmn(dev ? 0 : N), hmx(dev ? 0 : N), hcen(dev ? 0 : N);
    PodVector<uint32_t> order(N);
#pragma omp parallel for schedule(static) if (N > 8192)
    for (int64_t i = 0; i < (int64_t)(dev ? 0 : N); ++i) {
        order[(size_t)i] = (uint32_t)i;
        const HostTri& t = blas.tris[(size_t)i];
        hmn[(size_t)i] = make_float4(t.bmin.x, t.bmin.y, t.bmin.z, 0.0f);
        hmx[(size_t)i] = make_float4(t.bmax.x, t.bmax.y, t.bmax.z, 0.0f);
        hcen[(size_t)i] = make_float4(t.centroid.x, t.centroid.y, t.centroid.z, 0.0f);
    }
    DeviceBuffers buf(st);
    const uint32_t cap = std::min<uint32_t>(N, 16384u); // nodes of one level; broader levels fall back to the host builder
    float4 *d_mn, *d_mx, *d_cen;
    uint32_t *d_order[2], *d_node_of[2], *d_flags, *d_scan, *d_partial, *d_hole, *d_tail, *d_acc[2], *d_bins, *d_split, *d_split_index,
        *d_active[2];
    RefNode* d_nodes;
    RefInfo* d_info;
    RefDecision* d_dec;
    const size_t node_cap = (size_t)N * 2 + 2;
    if (!dev) {
        FB_TRY(buf.alloc(&d_mn, N));
        FB_TRY(buf.alloc(&d_mx, N));
        FB_TRY(buf.alloc(&d_cen, N));
    }
    for (int k = 0; k < 2; ++k) {
        FB_TRY(buf.alloc(&d_order[k], N));
        FB_TRY(buf.alloc(&d_node_of[k], N));
        FB_TRY(buf.alloc(&d_acc[k], (size_t)cap * 2 * kAcc));
        FB_TRY(buf.alloc(&d_active[k], (size_t)cap * 2));
    }
    FB_TRY(buf.alloc(&d_flags, N));
    FB_TRY(buf.alloc(&d_scan, (size_t)N + 1));
    FB_TRY(buf.alloc(&d_partial, (size_t)N / kScanBlock + 4));
    FB_TRY(buf.alloc(&d_hole, N));
    FB_TRY(buf.alloc(&d_tail, N));
    FB_TRY(buf.alloc(&d_bins, (size_t)cap * kRefNodeBinWords));
    FB_TRY(buf.alloc(&d_split, cap));
    FB_TRY(buf.alloc(&d_split_index, (size_t)cap + 1));
    FB_TRY(buf.alloc(&d_nodes, node_cap));
    FB_TRY(buf.alloc(&d_info, cap));
    FB_TRY(buf.alloc(&d_dec, cap));
    FB_TRY(buf.commit());
    lap("blas: host prep + alloc");

    cudaEvent_t e0, e1;
    FB_TRY(cudaEventCreate(&e0));
    FB_TRY(cudaEventCreate(&e1));
    if (dev) {
        d_mn = const_cast<float4*>(dev->mn);
        d_mx = const_cast<float4*>(dev->mx);
        d_cen = const_cast<float4*>(dev->cen);
        iota_kernel<<<blocks_for(N, 256), 256, 0, st>>>(d_order[0], N);
    } else {
        FB_TRY(cudaMemcpyAsync(d_mn, hmn.data(), (size_t)N * sizeof(float4), cudaMemcpyHostToDevice, st));
        FB_TRY(cudaMemcpyAsync(d_mx, hmx.data(), (size_t)N * sizeof(float4), cudaMemcpyHostToDevice, st));
        FB_TRY(cudaMemcpyAsync(d_cen, hcen.data(), (size_t)N * sizeof(float4), cudaMemcpyHostToDevice, st));
        FB_TRY(cudaMemcpyAsync(d_order[0], order.data(), (size_t)N * 4, cudaMemcpyHostToDevice, st));
    }
    FB_TRY(cudaEventRecord(e0, st));
    RefNode init[2];
    std::memset(init, 0, sizeof(init));
    init[0].start = 0;
    init[0].n = N;
    init[0].left = kNone;
    init[0].level = 0;
    init[1].left = kNone;
    FB_TRY(cudaMemcpyAsync(d_nodes, init, sizeof(init), cudaMemcpyHostToDevice, st));
    const uint32_t zero = 0;
    FB_TRY(cudaMemcpyAsync(d_active[0], &zero, 4, cudaMemcpyHostToDevice, st));
    FB_TRY(cudaMemsetAsync(d_node_of[0], 0, (size_t)N * 4, st));
    const uint32_t T = 256;
    init_acc<<<1, 32, 0, st>>>(d_acc[0], 1);
    accumulate_boxes<<<blocks_for(N, T), T, 0, st>>>(d_mn, d_mx, d_order[0], d_node_of[0], N, d_acc[0]);
    uint32_t count = 1, next_id = 2, level = 0;
    int cur = 0;
    for (;;) {
        ref_prepare<<<count, 256, 0, st>>>(d_acc[cur], d_active[cur], count, d_nodes, d_info, level >= max_depth ? nullptr : d_bins);
        if (level >= max_depth) break;
        const int nxt = cur ^ 1;
        ref_bin_prims<<<blocks_for(N, kBinThreads), kBinThreads, 0, st>>>(d_mn, d_mx, d_cen, d_order[cur], d_node_of[cur], N, d_info, d_bins);
        ref_decide<<<count, 96, 0, st>>>(d_active[cur], count, d_nodes, d_info, d_bins, d_dec, d_split);
        ref_flags<<<blocks_for(N, T), T, 0, st>>>(d_cen, d_order[cur], d_node_of[cur], N, d_dec, d_flags);
        exclusive_scan(d_flags, N, d_scan, d_partial, st);
        ref_counts<<<blocks_for(count, T), T, 0, st>>>(d_active[cur], count, d_nodes, d_scan, d_dec, d_split);
        exclusive_scan(d_split, count, d_split_index, d_partial, st);
        uint32_t splits = 0;
        FB_TRY(cudaMemcpyAsync(&splits, d_split_index + count, 4, cudaMemcpyDeviceToHost, st));
        FB_TRY(cudaStreamSynchronize(st));
        if (splits * 2 > cap || (size_t)next_id + 2 * (size_t)splits > node_cap) {
            err = "device BLAS build: level too broad";
            return RAYCA_E_UNSUPPORTED;
        }
        ref_children<<<blocks_for(count, T), T, 0, st>>>(d_active[cur], count, d_nodes, d_dec, d_split_index, next_id, level, d_active[nxt], d_acc[nxt]);
        ref_tables<<<blocks_for(N, T), T, 0, st>>>(d_node_of[cur], N, d_active[cur], d_nodes, d_dec, d_flags, d_scan, d_hole, d_tail);
        ref_scatter<<<blocks_for(N, T), T, 0, st>>>(d_order[cur], d_node_of[cur], N, d_active[cur], d_nodes, d_dec, d_flags, d_scan, d_hole,
                                                    d_tail, d_order[nxt], d_node_of[nxt]);
        cur = nxt; // the partition may have reordered ranges even when nothing split
        if (splits == 0) break;
        accumulate_boxes<<<blocks_for(N, T), T, 0, st>>>(d_mn, d_mx, d_order[cur], d_node_of[cur], N, d_acc[cur]);
        next_id += 2 * splits;
        count = 2 * splits;
        ++level;
    }
    FB_TRY(cudaEventRecord(e1, st));
    lap("blas: uploads + level loop");
    std::vector<RefNode> tmp(next_id);
    FB_TRY(cudaMemcpyAsync(tmp.data(), d_nodes, (size_t)next_id * sizeof(RefNode), cudaMemcpyDeviceToHost, st));
    FB_TRY(cudaMemcpyAsync(order.data(), d_order[cur], (size_t)N * 4, cudaMemcpyDeviceToHost, st));
    if (dev) FB_TRY(cudaMemcpyAsync(dev->order_out, d_order[cur], (size_t)N * 4, cudaMemcpyDeviceToDevice, st));
    FB_TRY(cudaStreamSynchronize(st));
    FB_TRY(cudaGetLastError());
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (device_ms) *device_ms = ms;
    lap("blas: read back");

    // renumber in the reference's push order (blas.rs:296-311): a node's children are appended when the node is
    // processed, then the left subtree is processed completely before the right one
    blas.nodes.clear();
    blas.nodes.reserve(next_id);
    auto to_host = [](const RefNode& r) {
        HostNode h;
        h.bmin = V3{r.mn[0], r.mn[1], r.mn[2]};
        h.bmax = V3{r.mx[0], r.mx[1], r.mx[2]};
        h.offset = r.start;
        h.count = r.n;
        return h;
    };
    blas.nodes.push_back(to_host(tmp[0]));
    HostNode dummy{};
    dummy.bmin = dummy.bmax = V3{0, 0, 0};
    dummy.offset = dummy.count = 0;
    blas.nodes.push_back(dummy);
    uint32_t deepest = 0;
    struct Todo { uint32_t tmp_id, out_index; };
    std::vector<Todo> stack{{0u, 0u}};
    while (!stack.empty()) {
        const Todo t = stack.back();
        stack.pop_back();
        const RefNode& r = tmp[t.tmp_id];
        deepest = std::max(deepest, r.level);
        HostNode h = to_host(r);
        if (r.left != kNone) {
            const uint32_t li = (uint32_t)blas.nodes.size();
            blas.nodes.resize(blas.nodes.size() + 2);
            h.offset = li;
            h.count = 0;
            stack.push_back({r.left + 1, li + 1}); // popped after the whole left subtree
            stack.push_back({r.left, li});
        }
        blas.nodes[t.out_index] = h;
    }
    blas.depth = deepest;
    lap("blas: renumber nodes");
    if (dev) { // the primitives live on the device: the host keeps the order only (rayca_cuda_scene_bvh_read)
        blas.order.swap(order);
        blas.device_prims = N;
        return RAYCA_OK;
    }
    PodVector<HostTri> sorted(N);
#pragma omp parallel for schedule(static) if (N > 8192)
    for (int64_t i = 0; i < (int64_t)N; ++i) sorted[(size_t)i] = blas.tris[order[(size_t)i]];
    blas.tris.swap(sorted);
    lap("blas: permute primitives");
    return RAYCA_OK;
}

// ---- mesh triangles emitted on the device ---------------------------------------------------------------------------
__device__ __forceinline__ const DeviceMeshJob& job_of_emit(const DeviceMeshJob* jobs, uint32_t job_count, uint32_t t) {
    uint32_t lo = 0, hi = job_count;
    while (lo < hi) { // first job whose base exceeds t (t lies in a job: the jobs tile [0, n))
        const uint32_t mid = (lo + hi) >> 1;
        if (jobs[mid].base <= t) lo = mid + 1;
        else hi = mid;
    }
    return jobs[lo - 1];
}

__global__ void emit_mesh_kernel(const DeviceMeshJob* __restrict__ jobs, uint32_t job_count, uint32_t n, const rayca_vertex* __restrict__ vertices,
                                 const uint32_t* __restrict__ indices, float4* __restrict__ pos, float4* __restrict__ mn, float4* __restrict__ mx,
                                 float4* __restrict__ cen, uint32_t* bad) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const DeviceMeshJob& job = job_of_emit(jobs, job_count, t);
    const uint32_t local = t - job.base;
    const TrsF world{V3{job.t[0], job.t[1], job.t[2]}, Q4{job.q[0], job.q[1], job.q[2], job.q[3]}, V3{job.s[0], job.s[1], job.s[2]}};
    V3 mv[3], wv[3];
    for (int k = 0; k < 3; ++k) {
        uint32_t vi = indices[job.first_index + (size_t)local * 3 + k];
        if (vi >= job.vertex_count) { *bad = 1u; vi = 0; }
        const rayca_vertex& v = vertices[job.first_vertex + vi];
        mv[k] = V3{v.pos[0], v.pos[1], v.pos[2]};
        wv[k] = trs_point(world, mv[k]); // Triangle::get_vertex (triangle.rs:67-69)
        pos[3 * (size_t)t + k] = make_float4(wv[k].x, wv[k].y, wv[k].z, 0.0f);
    }
    // Triangle::new centroid quirk (triangle.rs:59-62) then `trs * Vec3` (trs.rs:253-262)
    const V3 c = trs_vec(world, ((mv[0] + mv[1]) + mv[2]) * 0.3333f);
    cen[t] = make_float4(c.x, c.y, c.z, 0.0f);
    mn[t] = make_float4(fminf(fminf(fminf(kF32Max, wv[0].x), wv[1].x), wv[2].x), fminf(fminf(fminf(kF32Max, wv[0].y), wv[1].y), wv[2].y),
                        fminf(fminf(fminf(kF32Max, wv[0].z), wv[1].z), wv[2].z), 0.0f);
    mx[t] = make_float4(fmaxf(fmaxf(fmaxf(-kF32Max, wv[0].x), wv[1].x), wv[2].x), fmaxf(fmaxf(fmaxf(-kF32Max, wv[0].y), wv[1].y), wv[2].y),
                        fmaxf(fmaxf(fmaxf(-kF32Max, wv[0].z), wv[1].z), wv[2].z), 0.0f);
}

__global__ void gather_tri_pos_kernel(const float4* __restrict__ pos, const uint32_t* __restrict__ order, const DeviceMeshJob* __restrict__ jobs,
                                      uint32_t job_count, uint32_t n, uint32_t blas_index, float4* __restrict__ tri_pos) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint32_t src = order[k];
    const DeviceMeshJob& job = job_of_emit(jobs, job_count, src);
    const float4 a = pos[3 * (size_t)src], b = pos[3 * (size_t)src + 1], c = pos[3 * (size_t)src + 2];
    tri_pos[3 * (size_t)k] = make_float4(a.x, a.y, a.z, __uint_as_float(src));
    tri_pos[3 * (size_t)k + 1] = make_float4(b.x, b.y, b.z, __uint_as_float(job.material));
    tri_pos[3 * (size_t)k + 2] = make_float4(c.x, c.y, c.z, __uint_as_float(blas_index));
}

void free_device_blas_mesh(DeviceBlasMesh& m, cudaStream_t st) {
    void* ptrs[] = {m.pos, m.mn, m.mx, m.cen, m.order, m.jobs, m.bad};
    for (void* p : ptrs)
        if (p) cudaFreeAsync(p, st);
    m = DeviceBlasMesh{};
}

int emit_mesh_device(const std::vector<DeviceMeshJob>& jobs, uint32_t n, const void* d_vertices, const void* d_indices, cudaStream_t st,
                     DeviceBlasMesh& out, std::string& err) {
    out = DeviceBlasMesh{};
    out.n = n;
    out.job_count = (uint32_t)jobs.size();
    cudaError_t ce = cudaMallocAsync((void**)&out.pos, (size_t)n * 3 * sizeof(float4), st);
    if (ce == cudaSuccess) ce = cudaMallocAsync((void**)&out.mn, (size_t)n * sizeof(float4), st);
    if (ce == cudaSuccess) ce = cudaMallocAsync((void**)&out.mx, (size_t)n * sizeof(float4), st);
    if (ce == cudaSuccess) ce = cudaMallocAsync((void**)&out.cen, (size_t)n * sizeof(float4), st);
    if (ce == cudaSuccess) ce = cudaMallocAsync((void**)&out.order, (size_t)n * 4, st);
    if (ce == cudaSuccess) ce = cudaMallocAsync((void**)&out.jobs, jobs.size() * sizeof(DeviceMeshJob), st);
    if (ce == cudaSuccess) ce = cudaMallocAsync((void**)&out.bad, 4, st);
    if (ce == cudaSuccess) ce = cudaMemsetAsync(out.bad, 0, 4, st);
    // `jobs` is a pageable vector of the caller: this copy is staged before the call returns
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(out.jobs, jobs.data(), jobs.size() * sizeof(DeviceMeshJob), cudaMemcpyHostToDevice, st);
    if (ce == cudaSuccess) {
        emit_mesh_kernel<<<blocks_for(n, 256), 256, 0, st>>>(out.jobs, out.job_count, n, static_cast<const rayca_vertex*>(d_vertices),
                                                            static_cast<const uint32_t*>(d_indices), out.pos, out.mn, out.mx, out.cen, out.bad);
        ce = cudaGetLastError();
    }
    if (ce != cudaSuccess) {
        err = std::string("device emit: ") + cudaGetErrorString(ce);
        free_device_blas_mesh(out, st);
        return ce == cudaErrorMemoryAllocation ? RAYCA_E_OOM : RAYCA_E_CUDA;
    }
    return RAYCA_OK;
}

int gather_tri_pos_device(const DeviceBlasMesh& m, uint32_t blas_index, float4* d_tri_pos, cudaStream_t st) {
    gather_tri_pos_kernel<<<blocks_for(m.n, 256), 256, 0, st>>>(m.pos, m.order, m.jobs, m.job_count, m.n, blas_index, d_tri_pos);
    return cudaGetLastError() == cudaSuccess ? RAYCA_OK : RAYCA_E_CUDA;
}

} // namespace rb
