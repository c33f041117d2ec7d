// Development tool (CPU only): how many traversal steps would a 4-wide collapse of the library's binned-SAH tree save?
// Loads a glTF scene through libb200rt_host, builds the binary tree with csrc/cuda/bvh_rebuild.h (the tree the device
// traverses), walks a seeded ray set (camera rays + random segments inside the scene box) with the device's ordered
// closest-hit traversal, then collapses the tree to BVH4 (repeatedly open the child of largest area) and walks it
// nearest-child-first with the other hit children pushed in slot order.  Prints steps / boxes / leaf visits per ray.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>
#include "b200rt_host.h"
#include "bvh_rebuild.h"
using namespace b200rt;

struct Ray { float o[3], d[3], inv[3]; };
static bool slab(const float *mn, const float *mx, const Ray &r, float tbest, float &tn) {
    float t0 = -1e30f, t1 = 1e30f;
    for (int a = 0; a < 3; a++) {
        float x0 = (mn[a] - r.o[a]) * r.inv[a], x1 = (mx[a] - r.o[a]) * r.inv[a];
        t0 = std::max(t0, std::min(x0, x1)); t1 = std::min(t1, std::max(x0, x1));
    }
    tn = t0;
    return t0 <= t1 && t1 >= 0.f && t0 <= tbest;
}
static bool tri_hit(const b200rt_triangle &T, const Ray &r, float tmax, float &t) {
    float e1[3], e2[3], tv[3], pv[3], qv[3];
    for (int a = 0; a < 3; a++) { e1[a] = T.p[0][a] - T.p[2][a]; e2[a] = T.p[1][a] - T.p[2][a]; tv[a] = r.o[a] - T.p[2][a]; }
    pv[0] = r.d[1] * e2[2] - r.d[2] * e2[1]; pv[1] = r.d[2] * e2[0] - r.d[0] * e2[2]; pv[2] = r.d[0] * e2[1] - r.d[1] * e2[0];
    float det = e1[0] * pv[0] + e1[1] * pv[1] + e1[2] * pv[2];
    if (det == 0.f) return false;
    float id = 1.f / det, u = (tv[0] * pv[0] + tv[1] * pv[1] + tv[2] * pv[2]) * id;
    qv[0] = tv[1] * e1[2] - tv[2] * e1[1]; qv[1] = tv[2] * e1[0] - tv[0] * e1[2]; qv[2] = tv[0] * e1[1] - tv[1] * e1[0];
    float v = (r.d[0] * qv[0] + r.d[1] * qv[1] + r.d[2] * qv[2]) * id;
    t = (e2[0] * qv[0] + e2[1] * qv[1] + e2[2] * qv[2]) * id;
    return u >= 0 && v >= 0 && u + v <= 1 && t > 0 && t < tmax;
}
#define MAX_WIDE 16
struct Q4 { int n; uint32_t child[MAX_WIDE]; };   // child = node id of the binary tree (leaf or inner)
int main(int argc, char **argv) {
    if (argc < 2) return 2;
    b200rt_host_scene *hs = nullptr;
    if (b200rt_host_load_gltf(argv[1], 0, 64, 64, 1, &hs)) { printf("load failed\n"); return 1; }
    const b200rt_scene_desc *d = b200rt_host_desc(hs);
    RebuiltBvh bvh;
    rebuild_bvh_binned_sah(d->tris, d->n_tris, bvh);
    const auto &N = bvh.nodes;
    std::vector<b200rt_triangle> tris(d->n_tris);
    for (int i = 0; i < d->n_tris; i++) tris[i] = d->tris[bvh.order[i]];
    auto area = [&](uint32_t i) { float x = N[i].aabb_max[0] - N[i].aabb_min[0], y = N[i].aabb_max[1] - N[i].aabb_min[1], z = N[i].aabb_max[2] - N[i].aabb_min[2]; return x * y + y * z + x * z; };
    const int wide = argc > 2 ? atoi(argv[2]) : 4;
    if (wide < 2 || wide > MAX_WIDE) { fprintf(stderr, "width must be 2 .. %d\n", MAX_WIDE); return 2; }
    std::vector<Q4> q4(N.size());
    for (uint32_t i = 0; i < N.size(); i++) {
        Q4 &q = q4[i]; q.n = 0;
        if (N[i].left == 0) continue;
        q.child[q.n++] = N[i].left; q.child[q.n++] = N[i].right;
        while (q.n < wide) {
            int best = -1;
            for (int k = 0; k < q.n; k++) if (N[q.child[k]].left && (best < 0 || area(q.child[k]) > area(q.child[best]))) best = k;
            if (best < 0) break;
            uint32_t c = q.child[best];
            q.child[best] = N[c].left; q.child[q.n++] = N[c].right;
        }
    }
    // rays
    std::mt19937 rng(7);
    std::uniform_real_distribution<float> U(0, 1);
    std::vector<Ray> rays;
    float lo[3], hi[3];
    for (int a = 0; a < 3; a++) { lo[a] = N[0].aabb_min[a]; hi[a] = N[0].aabb_max[a]; }
    const int n_cam = 20000, n_rand = 60000;
    float ty = std::tan(d->cam_fov / 2), tx = ty * 16.f / 9.f;
    for (int i = 0; i < n_cam + n_rand; i++) {
        Ray r;
        if (i < n_cam) {
            float u = 2 * U(rng) - 1, v = 2 * U(rng) - 1, l = 0;
            for (int a = 0; a < 3; a++) { r.o[a] = d->cam_pos[a]; r.d[a] = u * tx * d->cam_right[a] - v * ty * d->cam_up[a] + d->cam_forward[a]; l += r.d[a] * r.d[a]; }
            for (int a = 0; a < 3; a++) r.d[a] /= std::sqrt(l);
        } else {
            float l = 0, b[3];
            for (int a = 0; a < 3; a++) { r.o[a] = lo[a] + U(rng) * (hi[a] - lo[a]); b[a] = lo[a] + U(rng) * (hi[a] - lo[a]); r.d[a] = b[a] - r.o[a]; l += r.d[a] * r.d[a]; }
            for (int a = 0; a < 3; a++) r.d[a] /= std::sqrt(l);
        }
        for (int a = 0; a < 3; a++) r.inv[a] = 1.f / r.d[a];
        rays.push_back(r);
    }
    double steps2 = 0, boxes2 = 0, leaves2 = 0, tris2 = 0, push2 = 0, steps4 = 0, boxes4 = 0, leaves4 = 0, tris4 = 0, push4 = 0, diff = 0;
    for (const Ray &r : rays) {
        float tb2 = 1e4f, tb4 = 1e4f;
        {   // binary, near child first
            uint32_t st[128]; int sp = 0; uint32_t cur = 0;
            for (;;) {
                if (N[cur].left) {
                    steps2++; boxes2 += 2;
                    float tl, tr; uint32_t L = N[cur].left, R = N[cur].right;
                    bool hl = slab(N[L].aabb_min, N[L].aabb_max, r, tb2, tl), hr = slab(N[R].aabb_min, N[R].aabb_max, r, tb2, tr);
                    if (hl && hr) { bool lf = tl <= tr; st[sp++] = lf ? R : L; push2++; cur = lf ? L : R; continue; }
                    if (hl) { cur = L; continue; } if (hr) { cur = R; continue; }
                } else { leaves2++; for (uint32_t i = N[cur].first; i < N[cur].last; i++) { float t; tris2++; if (tri_hit(tris[i], r, tb2, t)) tb2 = t; } }
                if (!sp) break; cur = st[--sp];
            }
        }
        {   // wide: nearest first, others pushed in slot order
            uint32_t st[256]; int sp = 0; uint32_t cur = 0;
            for (;;) {
                if (N[cur].left) {
                    const Q4 &q = q4[cur];
                    steps4++; boxes4 += wide;
                    float t[MAX_WIDE]; bool h[MAX_WIDE]; int best = -1;
                    for (int k = 0; k < q.n; k++) { h[k] = slab(N[q.child[k]].aabb_min, N[q.child[k]].aabb_max, r, tb4, t[k]); if (h[k] && (best < 0 || t[k] < t[best])) best = k; }
                    if (best >= 0) {
                        for (int k = q.n - 1; k >= 0; k--) if (h[k] && k != best) { st[sp++] = q.child[k]; push4++; }
                        cur = q.child[best]; continue;
                    }
                } else { leaves4++; for (uint32_t i = N[cur].first; i < N[cur].last; i++) { float t; tris4++; if (tri_hit(tris[i], r, tb4, t)) tb4 = t; } }
                if (!sp) break; cur = st[--sp];
            }
        }
        if (tb2 != tb4) diff++;
    }
    double n = (double)rays.size();
    printf("tris %d nodes %zu | binary: steps %.2f boxes %.2f leaf visits %.2f tri tests %.2f pushes %.2f\n", d->n_tris, N.size(), steps2 / n, boxes2 / n, leaves2 / n, tris2 / n, push2 / n);
    printf("wide %d: steps %.2f boxes %.2f leaf visits %.2f tri tests %.2f pushes %.2f | different hits %.0f\n", wide, steps4 / n, boxes4 / n, leaves4 / n, tris4 / n, push4 / n, diff);
    return 0;
}
