// Binned-SAH BVH builder used behind the C ABI ("better BVH").
//
// The tree that crosses the boundary is the reference's: a full-sweep SAH over triangles sorted by ONE
// vertex (hw8/src/include/bvh.h:63-65), or -- in hw6 -- over an arbitrary order, because hw6 sorts by
// Figure::position, which is (0,0,0) for every glTF triangle (hw6/src/include/bvh.h:63-65; depth ~60,
// two orders of magnitude slower to traverse).  Closest-hit results do not depend on the tree
// (SURVEY Q7), so the device path may traverse its own: top-down, 16 centroid bins per axis, SAH with a
// traversal-cost term, leaves of up to 4 triangles.  On the benchmark scene it tests 20 % fewer boxes
// per ray than the reference tree (profiles/r01_sweep_bvh_rebuild.log).
//
// Parallel build: the top of the tree is split node by node with the binning spread over the host
// cores; once every open node is small, the subtrees are built independently, one task each.
// Output: reference-shaped nodes (left == 0 marks a leaf holding [first, last) of the PERMUTED triangle
// order returned in `order`; children always have larger indices than their parent).
#pragma once
#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cstdint>
#include <chrono>
#include <cstdio>
#include <condition_variable>
#include <cstdlib>
#include <functional>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>
#include <xmmintrin.h>      // host-side SSE for the binning loop (x86-64 hosts only, like the rest of the build)
#include "b200rt.h"
#include "../common/host_pool.h"

namespace b200rt {

struct RebuiltBvh {
    std::vector<b200rt_bvh_node> nodes;
    std::vector<uint32_t> order;     // order[i] = index of the input triangle stored at position i
    // scratch kept between builds (a context re-uploading a scene does not page-fault fresh memory again)
    std::vector<float> scratch_box, scratch_cen;
};

namespace bvh_detail {

constexpr int BINS = 16;
// Nodes of up to Ctx::sweep_max triangles skip the bins: three 16-bin tables cost ~3000 operations per node whatever its
// size, and most nodes of a tree are tiny (the rebuild spent 70 % of its time there).  They get the exact object-split
// SAH instead -- sort by centroid on each axis, sweep prefix / suffix areas -- which is cheaper AND never worse.
constexpr int SWEEP_MAX = 64;     // capacity; the threshold in use is Ctx::sweep_max (12: profiles/r01_sweep_bvh_sweep.log)

struct B3 {
    float mn[3], mx[3];
    void reset() { for (int a = 0; a < 3; a++) { mn[a] = FLT_MAX; mx[a] = -FLT_MAX; } }
    void grow(const B3 &o) { for (int a = 0; a < 3; a++) { mn[a] = std::min(mn[a], o.mn[a]); mx[a] = std::max(mx[a], o.mx[a]); } }
    void grow_point(const float *p) { for (int a = 0; a < 3; a++) { mn[a] = std::min(mn[a], p[a]); mx[a] = std::max(mx[a], p[a]); } }
    float area() const {
        const float dx = mx[0] - mn[0], dy = mx[1] - mn[1], dz = mx[2] - mn[2];
        return dx < 0 ? 0.f : 2.f * (dx * dy + dx * dz + dy * dz);
    }
};

// Bin tables of one node.  Minima / maxima are kept as 4-float vectors (lane 3 unused) so that binning one triangle
// on one axis is four SSE min / max on memory instead of eighteen scalar ones: the binning passes are the bulk of
// the build once the small nodes take the sweep path.
struct Bins {
    // plain float storage (nvcc rejects vector types inside std::vector's element type); accessed with SSE loads
    alignas(16) float mn[3][BINS][4], mx[3][BINS][4];       // triangle bounds per bin
    alignas(16) float cmn[3][BINS][4], cmx[3][BINS][4];     // centroid bounds per bin (become the children's binning ranges)
    uint32_t cnt[3][BINS];
    void reset() {
        const __m128 hi = _mm_set1_ps(FLT_MAX), lo = _mm_set1_ps(-FLT_MAX);
        for (int a = 0; a < 3; a++) for (int b = 0; b < BINS; b++) {
            _mm_store_ps(mn[a][b], hi); _mm_store_ps(cmn[a][b], hi); _mm_store_ps(mx[a][b], lo); _mm_store_ps(cmx[a][b], lo);
            cnt[a][b] = 0;
        }
    }
    static void vmin(float *dst, __m128 v) { _mm_store_ps(dst, _mm_min_ps(_mm_load_ps(dst), v)); }
    static void vmax(float *dst, __m128 v) { _mm_store_ps(dst, _mm_max_ps(_mm_load_ps(dst), v)); }
    void merge(const Bins &o) {
        for (int a = 0; a < 3; a++) for (int b = 0; b < BINS; b++) {
            vmin(mn[a][b], _mm_load_ps(o.mn[a][b])); vmax(mx[a][b], _mm_load_ps(o.mx[a][b]));
            vmin(cmn[a][b], _mm_load_ps(o.cmn[a][b])); vmax(cmx[a][b], _mm_load_ps(o.cmx[a][b]));
            cnt[a][b] += o.cnt[a][b];
        }
    }
    static B3 unpack(const float *l, const float *h) {
        B3 r;
        for (int k = 0; k < 3; k++) { r.mn[k] = l[k]; r.mx[k] = h[k]; }
        return r;
    }
    B3 box(int a, int b) const { return unpack(mn[a][b], mx[a][b]); }
    B3 cbox(int a, int b) const { return unpack(cmn[a][b], cmx[a][b]); }
};

// Bounds a node is created with: its triangles' box and its centroids' box (both known from the parent's bins).
struct NodeBounds { B3 nb, cb; };

struct Ctx {
    const B3 *box;
    const float *cen;      // 3 per triangle
    uint32_t *order;
    int leaf_max;
    int sweep_max;         // <= SWEEP_MAX
    float trav;
    unsigned threads;
    int sah_depth;         // nodes at this depth or below are split at the median (bounds the tree depth, see split_node)
};

inline int bin_of(float c, float lo, float k) { return std::min(BINS - 1, (int)((c - lo) * k)); }

// Bounds of the triangles and of their centroids over order[first, last).
inline void bounds(const Ctx &c, uint32_t first, uint32_t last, B3 &nb, B3 &cb) {
    nb.reset(); cb.reset();
    for (uint32_t i = first; i < last; i++) {
        const uint32_t t = c.order[i];
        nb.grow(c.box[t]);
        cb.grow_point(c.cen + (size_t)t * 3);
    }
}

// The 4-float loads below read one float past a triangle's 3-float record: box[] and cen[] are allocated with a
// float of slack after the last element (rebuild_bvh_binned_sah).
inline void bin_range(const Ctx &c, uint32_t first, uint32_t last, const B3 &cb, const float *k, Bins &bins) {
    for (uint32_t i = first; i < last; i++) {
        const uint32_t t = c.order[i];
        const __m128 lo = _mm_loadu_ps(c.box[t].mn), hi = _mm_loadu_ps(c.box[t].mx);
        const float *cp = c.cen + (size_t)t * 3;
        const __m128 cv = _mm_loadu_ps(cp);
        for (int a = 0; a < 3; a++) {
            if (!(k[a] > 0.f)) continue;
            const int b = bin_of(cp[a], cb.mn[a], k[a]);
            Bins::vmin(bins.mn[a][b], lo);
            Bins::vmax(bins.mx[a][b], hi);
            Bins::vmin(bins.cmn[a][b], cv);
            Bins::vmax(bins.cmx[a][b], cv);
            bins.cnt[a][b]++;
        }
    }
}

// Chooses the split of one node and partitions order[first, last).  Returns mid, or 0 for "make a leaf".
// `in` holds the node's bounds (from its parent's bins); `kids` receives the children's.
// `parallel` spreads the binning over the host cores (used for the large nodes at the top).
// `depth` (root = 0): SAH splits can be arbitrarily lopsided, and the traversal kernels keep a fixed stack of
// B200RT_STACK entries without bound checks.  From depth Ctx::sah_depth on, a node is therefore cut at the median of
// its centroids on their widest axis: at most ceil(log2 N) more levels, so any triangle soup uploads (the
// reference's recursive traversal accepts any depth, hw8/src/include/bvh.h:111-142).
inline uint32_t split_node(const Ctx &c, b200rt_bvh_node &node, const NodeBounds &in, NodeBounds kids[2], bool parallel, int depth) {
    const uint32_t first = node.first, last = node.last, count = last - first;
    const B3 &nb = in.nb, &cb = in.cb;
    const unsigned nt = parallel ? c.threads : 1;
    for (int a = 0; a < 3; a++) { node.aabb_min[a] = nb.mn[a]; node.aabb_max[a] = nb.mx[a]; }
    if (count <= 1) return 0;
    float k[3];
    bool any_axis = false;
    for (int a = 0; a < 3; a++) {
        const float ext = cb.mx[a] - cb.mn[a];
        k[a] = ext > 0.f ? (float)BINS / ext : 0.f;
        any_axis = any_axis || k[a] > 0.f;
    }
    auto halves = [&](uint32_t m) {     // no usable bins: split by index and measure the halves
        bounds(c, first, m, kids[0].nb, kids[0].cb);
        bounds(c, m, last, kids[1].nb, kids[1].cb);
        return m;
    };
    if (!any_axis) return count <= 15 ? 0 : halves(first + count / 2);      // identical centroids
    if (depth >= c.sah_depth) {
        if (count <= (uint32_t)c.leaf_max) return 0;
        int axis = 0;
        for (int a = 1; a < 3; a++) if (cb.mx[a] - cb.mn[a] > cb.mx[axis] - cb.mn[axis]) axis = a;
        const uint32_t m = first + count / 2;
        std::nth_element(c.order + first, c.order + m, c.order + last, [&](uint32_t x, uint32_t y) {
            const float kx = c.cen[(size_t)x * 3 + axis], ky = c.cen[(size_t)y * 3 + axis];
            return kx < ky || (kx == ky && x < y);
        });
        return halves(m);
    }
    if (count <= (uint32_t)c.sweep_max) {
        uint32_t idx[3][SWEEP_MAX];
        float best_cost = FLT_MAX;
        int best_axis = -1;
        uint32_t best_i = 0;
        float left_area[SWEEP_MAX];
        for (int a = 0; a < 3; a++) {
            uint32_t *id = idx[a];
            if (count <= 12) {
                for (uint32_t i = 0; i < count; i++) {          // insertion sort by centroid
                    const uint32_t t = c.order[first + i];
                    const float key = c.cen[(size_t)t * 3 + a];
                    uint32_t j = i;
                    while (j > 0 && c.cen[(size_t)id[j - 1] * 3 + a] > key) { id[j] = id[j - 1]; j--; }
                    id[j] = t;
                }
            } else {
                for (uint32_t i = 0; i < count; i++) id[i] = c.order[first + i];
                // ties broken by triangle index: the result does not depend on the sort algorithm
                std::sort(id, id + count, [&](uint32_t x, uint32_t y) {
                    const float kx = c.cen[(size_t)x * 3 + a], ky = c.cen[(size_t)y * 3 + a];
                    return kx < ky || (kx == ky && x < y);
                });
            }
            B3 acc;
            acc.reset();
            for (uint32_t i = 0; i + 1 < count; i++) { acc.grow(c.box[id[i]]); left_area[i + 1] = acc.area(); }
            acc.reset();
            for (uint32_t i = count - 1; i >= 1; i--) {
                acc.grow(c.box[id[i]]);
                const float cost = left_area[i] * (float)i + acc.area() * (float)(count - i);
                if (cost < best_cost) { best_cost = cost; best_axis = a; best_i = i; }
            }
        }
        if (count <= (uint32_t)c.leaf_max && best_cost + c.trav * nb.area() >= nb.area() * (float)count) return 0;
        for (uint32_t i = 0; i < count; i++) c.order[first + i] = idx[best_axis][i];
        return halves(first + best_i);
    }
    Bins bins;
    bins.reset();
    if (nt > 1) {
        std::vector<Bins> part(nt);
        const uint32_t chunk = (count + nt - 1) / nt;
        HostPool::get().run(nt, [&](unsigned t) {
            part[t].reset();
            bin_range(c, std::min(last, first + t * chunk), std::min(last, first + (t + 1) * chunk), cb, k, part[t]);
        });
        for (unsigned t = 0; t < nt; t++) bins.merge(part[t]);
    } else {
        bin_range(c, first, last, cb, k, bins);
    }
    float best_cost = FLT_MAX;
    int best_axis = -1, best_bin = 0;
    for (int a = 0; a < 3; a++) {
        if (!(k[a] > 0.f)) continue;
        float right_area[BINS];
        uint32_t right_cnt[BINS];
        B3 acc;
        acc.reset();
        uint32_t n = 0;
        for (int b = BINS - 1; b > 0; b--) { if (bins.cnt[a][b]) acc.grow(bins.box(a, b)); n += bins.cnt[a][b]; right_area[b] = acc.area(); right_cnt[b] = n; }
        acc.reset(); n = 0;
        for (int b = 0; b < BINS - 1; b++) {
            if (bins.cnt[a][b]) acc.grow(bins.box(a, b));
            n += bins.cnt[a][b];
            if (n == 0 || right_cnt[b + 1] == 0) continue;
            const float cost = acc.area() * (float)n + right_area[b + 1] * (float)right_cnt[b + 1];
            if (cost < best_cost) { best_cost = cost; best_axis = a; best_bin = b; }
        }
    }
    if (best_axis < 0) return count <= 15 ? 0 : halves(first + count / 2);
    // SAH: one box step costs `trav` triangle tests
    if (count <= (uint32_t)c.leaf_max && best_cost + c.trav * nb.area() >= nb.area() * (float)count) return 0;
    const float lo = cb.mn[best_axis], kk = k[best_axis];
    auto it = std::partition(c.order + first, c.order + last,
                             [&](uint32_t t) { return bin_of(c.cen[(size_t)t * 3 + best_axis], lo, kk) <= best_bin; });
    const uint32_t mid = (uint32_t)(it - c.order);
    if (mid == first || mid == last) return halves(first + count / 2);
    for (int h = 0; h < 2; h++) { kids[h].nb.reset(); kids[h].cb.reset(); }
    for (int b = 0; b < BINS; b++) {
        NodeBounds &k2 = kids[b <= best_bin ? 0 : 1];
        if (!bins.cnt[best_axis][b]) continue;
        k2.nb.grow(bins.box(best_axis, b));
        k2.cb.grow(bins.cbox(best_axis, b));
    }
    return mid;
}

// Serial build of the subtree under nodes[root_id] (first/last set, bounds in `root`).  Descendants are
// written to nodes[slot, slot + 2 * count): the caller reserves that range, so workers never touch the
// same elements; unused slots stay zero (empty leaves nobody points to).
inline void build_subtree(const Ctx &c, b200rt_bvh_node *nodes, uint32_t root_id, uint32_t slot, const NodeBounds &root, int root_depth) {
    struct Item { uint32_t id; NodeBounds b; int depth; };
    std::vector<Item> todo;
    todo.push_back({root_id, root, root_depth});
    while (!todo.empty()) {
        const Item it = todo.back();
        todo.pop_back();
        const uint32_t first = nodes[it.id].first, last = nodes[it.id].last;
        NodeBounds kids[2];
        const uint32_t mid = split_node(c, nodes[it.id], it.b, kids, false, it.depth);
        if (!mid) continue;
        const uint32_t l = slot;
        slot += 2;
        nodes[l] = b200rt_bvh_node{};
        nodes[l].first = first; nodes[l].last = mid;
        nodes[l + 1] = b200rt_bvh_node{};
        nodes[l + 1].first = mid; nodes[l + 1].last = last;
        nodes[it.id].left = l; nodes[it.id].right = l + 1;
        todo.push_back({l + 1, kids[1], it.depth + 1});
        todo.push_back({l, kids[0], it.depth + 1});
    }
}

}  // namespace bvh_detail

inline void rebuild_bvh_binned_sah(const b200rt_triangle *tris, int n, RebuiltBvh &out, bool light_set = false) {
    using namespace bvh_detail;
    out.order.resize((size_t)std::max(n, 0));
    if (n <= 0) { out.nodes.clear(); return; }      // (otherwise the node array keeps its size: no 21 MB value-initialisation per build)
    Ctx c{};
    c.leaf_max = 4; c.trav = 1.f;        // sweep: profiles/r01_sweep_bvh_rebuild.log
    c.sweep_max = 12;
    c.sah_depth = 32;      // + ceil(log2(2^27 / leaf)) median levels < B200RT_STACK (64)
#ifdef B200RT_DEV      // tools/sweep_*.sh
    if (const char *e = getenv("B200RT_REBUILD_SWEEP")) c.sweep_max = std::max(0, std::min(SWEEP_MAX, atoi(e)));
    if (const char *e = getenv(light_set ? "B200RT_REBUILD_LEAF_LIGHT" : "B200RT_REBUILD_LEAF")) c.leaf_max = std::max(1, std::min(15, atoi(e)));
    if (const char *e = getenv(light_set ? "B200RT_REBUILD_TRAV_LIGHT" : "B200RT_REBUILD_TRAV")) c.trav = (float)atof(e);
#endif
    c.threads = host_threads();
    if (n < 65536) c.threads = 1;
    static_assert(sizeof(B3) == 6 * sizeof(float), "B3 is six floats");
    out.scratch_box.resize((size_t)n * 6 + 4);     // + slack for the 4-float loads of bin_range
    out.scratch_cen.resize((size_t)n * 3 + 4);
    B3 *box = reinterpret_cast<B3 *>(out.scratch_box.data());
    float *cen = out.scratch_cen.data();
    {
        auto prep = [&](size_t a, size_t b) {
            for (size_t i = a; i < b; i++) {
                B3 &bx = box[i];
                for (int k = 0; k < 3; k++) {
                    bx.mn[k] = std::min(tris[i].p[2][k], std::min(tris[i].p[0][k], tris[i].p[1][k]));
                    bx.mx[k] = std::max(tris[i].p[2][k], std::max(tris[i].p[0][k], tris[i].p[1][k]));
                    cen[i * 3 + k] = 0.5f * (bx.mn[k] + bx.mx[k]);
                }
                out.order[i] = (uint32_t)i;
            }
        };
        if (c.threads > 1) {
            const size_t chunk = ((size_t)n + c.threads - 1) / c.threads;
            HostPool::get().run(c.threads, [&](unsigned t) { prep(std::min((size_t)n, t * chunk), std::min((size_t)n, (t + 1) * chunk)); });
        } else prep(0, (size_t)n);
    }
    c.box = box; c.cen = cen; c.order = out.order.data();
#ifdef B200RT_DEV
    const bool timing = !light_set && getenv("B200RT_UPLOAD_TIMING") != nullptr;
    auto t_last = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) {
        if (!timing) return;
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[rebuild] %-20s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
        t_last = now;
    };
#else
    auto lap = [](const char *) {};
#endif
    lap("boxes, centroids");

    // Node slots: 0 = root; a node with `count` triangles owns at most 2 * count - 2 descendants.  The top of
    // the tree takes slots from a shared cursor; a subtree task gets a private range up front.
    std::vector<b200rt_bvh_node> &nodes = out.nodes;
    const uint32_t top_cap = c.threads > 1 ? 4096u : 1u;      // slots [0, top_cap): root and the nodes split in phases A / B
    {
        // Unused slots must read as empty leaves.  Clearing all 21 MB (268 k triangles) up front cost 2 ms of a 12 ms
        // build; instead the top slots are cleared here, every subtree task clears its own range right before it
        // writes into it (parallel, and the lines are then warm), and the gap between the two is cleared at the end.
        const size_t need = (size_t)top_cap + (size_t)n * 2 + 2;
        if (nodes.capacity() < need) { nodes.clear(); nodes.reserve(need); }
        nodes.resize(need);
        memset((void *)nodes.data(), 0, (c.threads > 1 ? (size_t)top_cap : need) * sizeof(b200rt_bvh_node));
    }
    lap("zero node slots");
    nodes[0].first = 0; nodes[0].last = (uint32_t)n;
    // root bounds: the only full bounds pass (children get theirs from their parent's bins)
    NodeBounds rb;
    if (c.threads > 1) {
        std::vector<NodeBounds> part(c.threads);
        const uint32_t chunk = ((uint32_t)n + c.threads - 1) / c.threads;
        HostPool::get().run(c.threads, [&](unsigned t) { bounds(c, std::min((uint32_t)n, t * chunk), std::min((uint32_t)n, (t + 1) * chunk), part[t].nb, part[t].cb); });
        rb.nb.reset(); rb.cb.reset();
        for (auto &pb : part) { rb.nb.grow(pb.nb); rb.cb.grow(pb.cb); }
    } else {
        bounds(c, 0, (uint32_t)n, rb.nb, rb.cb);
    }
    auto shrink = [&] {        // drop the unused tail (slots in the middle stay as unreferenced empty leaves)
        size_t used = nodes.size();
        while (used > 1 && nodes[used - 1].left == 0 && nodes[used - 1].first == nodes[used - 1].last) used--;
        nodes.resize(used);
    };
    lap("root bounds");
    if (c.threads == 1) { build_subtree(c, nodes.data(), 0, top_cap, rb, 0); shrink(); return; }

    // phase A: the few very large nodes at the top, one by one, with the binning spread over the cores
    const uint32_t big = 65536, task_size = (uint32_t)std::max(1024, n / (int)(c.threads * 16));
    // the bounds of the nodes created in phases A and B live in `nbs` (same index as `nodes`)
    uint32_t cursor = 1;
    std::vector<NodeBounds> nbs((size_t)top_cap, rb);
    std::vector<int> ndepth((size_t)top_cap, 0);
    std::vector<uint32_t> queue(1, 0u);
    auto add_children = [&](uint32_t id, uint32_t first, uint32_t mid, uint32_t last, const NodeBounds kids[2]) {
        const uint32_t l = cursor;
        cursor += 2;
        nodes[l].first = first; nodes[l].last = mid; nbs[l] = kids[0];
        nodes[l + 1].first = mid; nodes[l + 1].last = last; nbs[l + 1] = kids[1];
        ndepth[l] = ndepth[l + 1] = ndepth[id] + 1;
        nodes[id].left = l; nodes[id].right = l + 1;
        queue.push_back(l + 1);
        queue.push_back(l);
    };
    for (bool again = true; again;) {
        again = false;
        for (size_t q = 0; q < queue.size(); q++) {
            const uint32_t id = queue[q];
            if (nodes[id].last - nodes[id].first <= big || cursor + 2 > top_cap) continue;
            queue.erase(queue.begin() + (long)q);
            const uint32_t first = nodes[id].first, last = nodes[id].last;
            NodeBounds kids[2];
            const uint32_t mid = split_node(c, nodes[id], nbs[id], kids, true, ndepth[id]);
            if (mid) add_children(id, first, mid, last, kids);
            again = true;
            break;
        }
    }
    lap("phase A (top nodes)");
    // phase B: a shared work queue; a worker either splits a medium node (children go back to the queue)
    // or builds a small node's whole subtree into a slot range of its own.  Subtree ranges are handed out
    // from the END of the array downwards so that they never meet the cursor of the top nodes.
    std::mutex mu;
    int inflight = 0;
    uint32_t tail = (uint32_t)nodes.size();
    {
        HostPool::get().run(c.threads, [&](unsigned) {
                for (;;) {
                    uint32_t id, slot = 0;
                    NodeBounds mine;
                    int my_depth;
                    bool whole;
                    {
                        std::unique_lock<std::mutex> lk(mu);
                        if (queue.empty()) {
                            if (inflight == 0) return;
                            lk.unlock();
                            std::this_thread::yield();
                            continue;
                        }
                        id = queue.back();
                        queue.pop_back();
                        mine = nbs[id];
                        my_depth = ndepth[id];
                        const uint32_t count = nodes[id].last - nodes[id].first;
                        whole = count <= task_size || cursor + 2 > top_cap;     // top slots exhausted: finish serially
                        if (whole) { tail -= 2 * count; slot = tail; }
                        inflight++;
                    }
                    const uint32_t first = nodes[id].first, last = nodes[id].last;
                    if (whole) {
                        memset((void *)(nodes.data() + slot), 0, (size_t)2 * (last - first) * sizeof(b200rt_bvh_node));
                        build_subtree(c, nodes.data(), id, slot, mine, my_depth);
                        std::lock_guard<std::mutex> lk(mu);
                        inflight--;
                    } else {
                        b200rt_bvh_node nd = nodes[id];
                        NodeBounds kids[2];
                        const uint32_t mid = split_node(c, nd, mine, kids, false, my_depth);
                        std::lock_guard<std::mutex> lk(mu);
                        for (int a = 0; a < 3; a++) { nodes[id].aabb_min[a] = nd.aabb_min[a]; nodes[id].aabb_max[a] = nd.aabb_max[a]; }
                        if (mid) add_children(id, first, mid, last, kids);
                        inflight--;
                    }
                }
            });
    }
    if (tail > top_cap) memset((void *)(nodes.data() + top_cap), 0, (size_t)(tail - top_cap) * sizeof(b200rt_bvh_node));      // never handed out
    lap("phase B (subtrees)");
    shrink();
    lap("shrink");
}

}  // namespace b200rt
