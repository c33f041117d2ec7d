// Binned-SAH BVH builder used behind the C ABI when the tree that crosses the boundary is unusable
// for traversal: hw6 sorts its triangles by Figure::position, which is (0,0,0) for every glTF
// triangle (hw6/src/include/bvh.h:63-65), so the reference tree is a SAH split of an arbitrary
// order -- depth ~60 and two orders of magnitude slower to traverse.
// Closest-hit results do not depend on the tree, so the device path builds its own over the
// same triangles: top-down, 16 centroid bins per axis, reference-shaped output nodes (left == 0 marks
// a leaf holding [first, last) of the PERMUTED triangle order returned in `order`).
#pragma once
#include <algorithm>
#include <cfloat>
#include <cstdint>
#include <vector>
#include "b200rt.h"

namespace b200rt {

struct RebuiltBvh {
    std::vector<b200rt_bvh_node> nodes;
    std::vector<uint32_t> order;     // order[i] = index of the input triangle stored at position i
};

inline void rebuild_bvh_binned_sah(const b200rt_triangle *tris, int n, RebuiltBvh &out) {
    constexpr int BINS = 16, LEAF_MAX = 4;
    struct B3 {
        float mn[3], mx[3];
        void reset() { for (int a = 0; a < 3; a++) { mn[a] = FLT_MAX; mx[a] = -FLT_MAX; } }
        void grow(const B3 &o) { for (int a = 0; a < 3; a++) { mn[a] = std::min(mn[a], o.mn[a]); mx[a] = std::max(mx[a], o.mx[a]); } }
        float area() const {
            const float dx = mx[0] - mn[0], dy = mx[1] - mn[1], dz = mx[2] - mn[2];
            return dx < 0 ? 0.f : 2.f * (dx * dy + dx * dz + dy * dz);
        }
    };
    out.nodes.clear();
    out.order.resize((size_t)std::max(n, 0));
    if (n <= 0) return;
    std::vector<B3> box((size_t)n);
    std::vector<float> cen((size_t)n * 3);
    for (int i = 0; i < n; i++) {
        B3 &b = box[(size_t)i];
        for (int a = 0; a < 3; a++) {
            b.mn[a] = std::min(tris[i].p[2][a], std::min(tris[i].p[0][a], tris[i].p[1][a]));
            b.mx[a] = std::max(tris[i].p[2][a], std::max(tris[i].p[0][a], tris[i].p[1][a]));
            cen[(size_t)i * 3 + a] = 0.5f * (b.mn[a] + b.mx[a]);
        }
        out.order[(size_t)i] = (uint32_t)i;
    }
    out.nodes.reserve((size_t)n * 2);
    b200rt_bvh_node root{};
    root.first = 0; root.last = (uint32_t)n;
    out.nodes.push_back(root);
    std::vector<uint32_t> todo(1, 0u);
    while (!todo.empty()) {
        const uint32_t id = todo.back();
        todo.pop_back();
        const uint32_t first = out.nodes[id].first, last = out.nodes[id].last, count = last - first;
        B3 nb, cb;
        nb.reset(); cb.reset();
        for (uint32_t i = first; i < last; i++) {
            const uint32_t t = out.order[i];
            nb.grow(box[t]);
            for (int a = 0; a < 3; a++) { cb.mn[a] = std::min(cb.mn[a], cen[(size_t)t * 3 + a]); cb.mx[a] = std::max(cb.mx[a], cen[(size_t)t * 3 + a]); }
        }
        for (int a = 0; a < 3; a++) { out.nodes[id].aabb_min[a] = nb.mn[a]; out.nodes[id].aabb_max[a] = nb.mx[a]; }
        if (count <= 1) continue;
        // best binned split over the three axes
        float best_cost = FLT_MAX;
        int best_axis = -1, best_bin = 0;
        for (int a = 0; a < 3; a++) {
            const float ext = cb.mx[a] - cb.mn[a];
            if (!(ext > 0.f)) continue;
            const float k = (float)BINS / ext;
            B3 bb[BINS];
            uint32_t bc[BINS] = {};
            for (int b = 0; b < BINS; b++) bb[b].reset();
            for (uint32_t i = first; i < last; i++) {
                const uint32_t t = out.order[i];
                const int b = std::min(BINS - 1, (int)((cen[(size_t)t * 3 + a] - cb.mn[a]) * k));
                bb[b].grow(box[t]); bc[b]++;
            }
            float right_area[BINS];
            uint32_t right_cnt[BINS];
            B3 acc; acc.reset();
            uint32_t c = 0;
            for (int b = BINS - 1; b > 0; b--) { acc.grow(bb[b]); c += bc[b]; right_area[b] = acc.area(); right_cnt[b] = c; }
            acc.reset(); c = 0;
            for (int b = 0; b < BINS - 1; b++) {
                acc.grow(bb[b]); c += bc[b];
                if (c == 0 || right_cnt[b + 1] == 0) continue;
                const float cost = acc.area() * (float)c + right_area[b + 1] * (float)right_cnt[b + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = a; best_bin = b; }
            }
        }
        uint32_t mid;
        if (best_axis < 0) {
            if (count <= 15) continue;                 // identical centroids: a leaf the flattener can store
            mid = first + count / 2;
        } else {
            if (count <= LEAF_MAX && best_cost >= nb.area() * (float)count) continue;
            const float k = (float)BINS / (cb.mx[best_axis] - cb.mn[best_axis]);
            const float lo = cb.mn[best_axis];
            auto it = std::partition(out.order.begin() + first, out.order.begin() + last, [&](uint32_t t) {
                return std::min(BINS - 1, (int)((cen[(size_t)t * 3 + best_axis] - lo) * k)) <= best_bin;
            });
            mid = (uint32_t)(it - out.order.begin());
            if (mid == first || mid == last) mid = first + count / 2;
        }
        const uint32_t l = (uint32_t)out.nodes.size();
        b200rt_bvh_node child{};
        child.first = first; child.last = mid;
        out.nodes.push_back(child);
        child.first = mid; child.last = last;
        out.nodes.push_back(child);
        out.nodes[id].left = l; out.nodes[id].right = l + 1;
        todo.push_back(l + 1);
        todo.push_back(l);
    }
}

}  // namespace b200rt
