// SAH BVH builder following the reference ALGORITHM (hw8/src/include/bvh.h:34-109):
//   per node: sort the range by one key coordinate on x, y, z in turn; for each, sweep
//   prefix/suffix box areas, score = S_left * n_left + S_right * n_right; take the best axis;
//   stop when best >= S * n or <= 1 item.  Nodes are appended parent-first (pre-order).
// The sort key is ONE vertex (`data3.coords`, hw7/hw8) -- the caller supplies it per item.
// Items are permuted in place; `order` returns the permutation.  std::sort on the same
// comparison outcomes yields the same permutation as the reference's sort of Figures.
#pragma once
#include <algorithm>
#include <cstdint>
#include <vector>
#include "b200rt.h"

namespace b200rt_host {

struct BvhItem {
    float key[3];
    float mn[3], mx[3];
    uint32_t index;  // original position
};

struct Box3 {
    float mn[3], mx[3];
    void set(const BvhItem &it) {
        for (int a = 0; a < 3; a++) { mn[a] = it.mn[a]; mx[a] = it.mx[a]; }
    }
    void extend(const BvhItem &it) {
        // AABB::extend(const AABB&) = extend(min) then extend(max)  (hw8/src/primitives.cpp:145-157)
        for (int a = 0; a < 3; a++) {
            mx[a] = std::max(mx[a], it.mn[a]); mn[a] = std::min(mn[a], it.mn[a]);
        }
        for (int a = 0; a < 3; a++) {
            mx[a] = std::max(mx[a], it.mx[a]); mn[a] = std::min(mn[a], it.mx[a]);
        }
    }
    float area() const {
        float dx = mx[0] - mn[0], dy = mx[1] - mn[1], dz = mx[2] - mn[2];
        return 2 * (dx * dy + dx * dz + dy * dz);
    }
};

class BvhBuilder {
public:
    std::vector<b200rt_bvh_node> nodes;
    uint32_t root = 0;

    void build(std::vector<BvhItem> &items, uint32_t n) {
        nodes.clear();
        nodes.reserve(2 * (size_t)n + 1);
        root = build_node(items, 0, n);
    }

private:
    std::vector<float> scores_;

    static void sort_axis(std::vector<BvhItem> &it, uint32_t first, uint32_t last, int axis) {
        std::sort(it.begin() + first, it.begin() + last,
                  [axis](const BvhItem &l, const BvhItem &r) { return l.key[axis] < r.key[axis]; });
    }

    std::pair<float, uint32_t> best_split(const std::vector<BvhItem> &it, uint32_t first, uint32_t last) {
        const size_t n = last - first;
        scores_.assign(n, 0.f);
        Box3 pre; pre.set(it[first]);
        for (size_t i = 1; i < n; i++) {
            scores_[i] = pre.area() * i;
            pre.extend(it[first + i]);
        }
        Box3 suf; suf.set(it[last - 1]);
        for (size_t i = n - 1; i >= 1; i--) {
            scores_[i] += suf.area() * (n - i);
            suf.extend(it[first + i - 1]);
        }
        std::pair<float, uint32_t> best = {scores_[1], first + 1};
        for (size_t i = 2; i < n; i++)
            if (scores_[i] < best.first) best = {scores_[i], (uint32_t)(first + i)};
        return best;
    }

    uint32_t build_node(std::vector<BvhItem> &it, uint32_t first, uint32_t last) {
        b200rt_bvh_node cur{};
        cur.first = first; cur.last = last; cur.left = cur.right = 0;
        Box3 box{};
        if (first < last) box.set(it[first]);
        else for (int a = 0; a < 3; a++) box.mn[a] = box.mx[a] = 0.f;
        for (uint32_t i = first + 1; i < last; i++) box.extend(it[i]);
        for (int a = 0; a < 3; a++) { cur.aabb_min[a] = box.mn[a]; cur.aabb_max[a] = box.mx[a]; }
        uint32_t pos = (uint32_t)nodes.size();
        nodes.push_back(cur);
        if (last - first <= 1) return pos;

        sort_axis(it, first, last, 0); auto sx = best_split(it, first, last);
        sort_axis(it, first, last, 1); auto sy = best_split(it, first, last);
        sort_axis(it, first, last, 2); auto sz = best_split(it, first, last);
        float best = std::min(sx.first, std::min(sy.first, sz.first));
        if (best >= box.area() * (last - first)) return pos;
        uint32_t mid;
        if (best == sx.first) { mid = sx.second; sort_axis(it, first, last, 0); }
        else if (best == sy.first) { mid = sy.second; sort_axis(it, first, last, 1); }
        else { mid = sz.second; /* already sorted by z */ sort_axis(it, first, last, 2); }
        uint32_t l = build_node(it, first, mid);
        nodes[pos].left = l;
        uint32_t r = build_node(it, mid, last);
        nodes[pos].right = r;
        return pos;
    }
};

}  // namespace b200rt_host
