// Host-only check of csrc/cuda/bvh_rebuild.h: builds over a seeded triangle soup (serial path below 65536
// triangles, parallel path above) and verifies the tree invariants the flattener and the kernels rely on.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>
#include "bvh_rebuild.h"

using namespace b200rt;

static int check(int n, unsigned seed) {
    std::mt19937 rng(seed);
    std::uniform_real_distribution<float> u(-10.f, 10.f), e(-0.2f, 0.2f);
    std::vector<b200rt_triangle> tris((size_t)n);
    for (auto &t : tris) {
        float c[3] = {u(rng), u(rng), u(rng)};
        for (int v = 0; v < 3; v++) for (int a = 0; a < 3; a++) t.p[v][a] = c[a] + e(rng);
    }
    for (int i = 0; i < n / 50; i++) tris[(size_t)i] = tris[0];      // many identical centroids
    RebuiltBvh b;
    rebuild_bvh_binned_sah(tris.data(), n, b);
    if ((int)b.order.size() != n || b.nodes.empty()) return 1;
    std::vector<int> seen((size_t)n, 0), covered((size_t)n, 0);
    for (uint32_t o : b.order) { if (o >= (uint32_t)n) return 2; seen[o]++; }
    for (int s : seen) if (s != 1) return 3;                          // `order` is a permutation
    std::vector<uint32_t> stack(1, 0u);
    size_t visited = 0, max_leaf = 0;
    while (!stack.empty()) {
        const uint32_t id = stack.back();
        stack.pop_back();
        visited++;
        const b200rt_bvh_node &nd = b.nodes[id];
        if (nd.first > nd.last || nd.last > (uint32_t)n) return 4;
        if (nd.left == 0) {
            if (nd.last - nd.first > 15) return 5;                    // the flattener stores at most 15 per leaf
            max_leaf = std::max<size_t>(max_leaf, nd.last - nd.first);
            for (uint32_t i = nd.first; i < nd.last; i++) {
                covered[i]++;
                const b200rt_triangle &t = tris[b.order[i]];
                for (int v = 0; v < 3; v++) for (int a = 0; a < 3; a++)
                    if (t.p[v][a] < nd.aabb_min[a] || t.p[v][a] > nd.aabb_max[a]) return 6;
            }
            continue;
        }
        if (nd.left <= id || nd.right <= id || nd.left >= b.nodes.size() || nd.right >= b.nodes.size()) return 7;
        const b200rt_bvh_node &l = b.nodes[nd.left], &r = b.nodes[nd.right];
        if (l.first != nd.first || l.last != r.first || r.last != nd.last || l.first == l.last || r.first == r.last) return 8;
        for (int a = 0; a < 3; a++)
            if (l.aabb_min[a] < nd.aabb_min[a] || r.aabb_min[a] < nd.aabb_min[a] || l.aabb_max[a] > nd.aabb_max[a] || r.aabb_max[a] > nd.aabb_max[a]) return 9;
        stack.push_back(nd.left); stack.push_back(nd.right);
    }
    for (int cvr : covered) if (cvr != 1) return 10;                  // every triangle in exactly one leaf
    // slots nobody points to must be empty leaves (the parallel build leaves gaps between subtree ranges)
    size_t empty = 0;
    for (const auto &nd : b.nodes) if (nd.left == 0 && nd.first == nd.last) empty++;
    if (visited + empty < b.nodes.size()) return 11;
    printf("n=%d nodes=%zu reachable=%zu max_leaf=%zu ok\n", n, b.nodes.size(), visited, max_leaf);
    return 0;
}

// A soup that drives SAH into maximally lopsided splits (sizes and positions in a geometric series: peeling off the
// largest triangle is always the cheapest cut): without the depth cap of split_node the tree is a 4000-level comb.
static int check_depth(int n) {
    std::vector<b200rt_triangle> tris((size_t)n);
    double x = 0, s = 1;
    for (int i = 0; i < n; i++) {
        b200rt_triangle &t = tris[(size_t)i];
        memset(&t, 0, sizeof(t));
        t.p[0][0] = (float)x; t.p[1][0] = (float)(x + s); t.p[2][0] = (float)x;
        t.p[1][1] = (float)s; t.p[2][2] = (float)s;
        x += 1.5 * s; s *= 0.98;
    }
    RebuiltBvh b;
    rebuild_bvh_binned_sah(tris.data(), n, b);
    struct It { uint32_t id; int d; };
    std::vector<It> st(1, It{0u, 1});
    int max_depth = 0;
    while (!st.empty()) {
        const It it = st.back();
        st.pop_back();
        max_depth = std::max(max_depth, it.d);
        const b200rt_bvh_node &nd = b.nodes[it.id];
        if (nd.left) { st.push_back({nd.left, it.d + 1}); st.push_back({nd.right, it.d + 1}); }
    }
    printf("comb n=%d depth=%d %s\n", n, max_depth, max_depth <= 60 ? "ok" : "TOO DEEP");
    return max_depth <= 60 ? 0 : 12;
}

int main() {
    if (int rc = check_depth(4000)) return rc;
    const int sizes[] = {1, 2, 17, 5000, 70000, 300000};
    for (int n : sizes) {
        const int rc = check(n, 1234u + (unsigned)n);
        if (rc) { printf("n=%d FAILED rule %d\n", n, rc); return rc; }
    }
    return 0;
}
