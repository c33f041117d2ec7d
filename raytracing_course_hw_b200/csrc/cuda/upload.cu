// Scene upload: flattening of the reference-shaped description to the device layout, the one
// arena + one copy, the hw5 split into analytic figures and baked triangles, and scene replication between contexts.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <type_traits>
#include "analytic_kernels.h"
#include "context.h"

int ensure_arena(b200rt_ctx *ctx, size_t bytes) {
    if (bytes <= ctx->arena_cap) return B200RT_OK;
    CK(cudaStreamSynchronize(ctx->stream));
    cudaFree(ctx->arena_dev); cudaFreeHost(ctx->arena_pin);
    ctx->arena_dev = ctx->arena_pin = nullptr; ctx->arena_cap = 0;
    const size_t cap = bytes + bytes / 8;
    CK(cudaMalloc((void **)&ctx->arena_dev, cap));
    CK(cudaHostAlloc((void **)&ctx->arena_pin, cap, cudaHostAllocDefault));
    ctx->arena_cap = cap;
    return B200RT_OK;
}

// L2 residency of the scene was measured and dropped: declaring the arena a persisting access-policy window left
// extend unchanged and made shade 16 - 90 % slower (profiles/r01_sweep_l2_state.log: the stages gain more from finding
// the path state the previous stage just wrote in L2 than from scene hits).
// The arena stays allocated for the next upload; b200rt_destroy frees it.
void release_scene(b200rt_ctx *c) { c->has_scene = false; }

void free_scene_memory(b200rt_ctx *c) {
    cudaFree(c->arena_dev); cudaFreeHost(c->arena_pin);
    c->arena_dev = c->arena_pin = nullptr;
    c->arena_cap = 0;
    c->has_scene = false;
}

namespace {

// ---- BVH flattening: reference-shaped 40-byte nodes -> 32-byte quantised sibling-pair records ------
struct FBox { float mn[3], mx[3]; };

struct Flattener {
    const b200rt_bvh_node *ref = nullptr;
    const b200rt_triangle *tris = nullptr;
    std::vector<DPair> out;
    DQuant q{};
    FBox root_box{};
    uint32_t root_word = B200RT_LEAF;
    int max_depth = 0;

    static void tri_box(const b200rt_triangle &t, float *mn, float *mx) {
        for (int a = 0; a < 3; a++) {
            mn[a] = std::min(t.p[0][a], std::min(t.p[1][a], t.p[2][a]));
            mx[a] = std::max(t.p[0][a], std::max(t.p[1][a], t.p[2][a]));
        }
    }
    FBox range_box(uint32_t first, uint32_t last) const {
        FBox b{};
        for (uint32_t i = first; i < last; i++) {
            float a[3], c[3];
            tri_box(tris[i], a, c);
            for (int k = 0; k < 3; k++) {
                b.mn[k] = i == first ? a[k] : std::min(b.mn[k], a[k]);
                b.mx[k] = i == first ? c[k] : std::max(b.mx[k], c[k]);
            }
        }
        return b;
    }
    FBox ref_box(uint32_t r) const {
        FBox b;
        for (int k = 0; k < 3; k++) { b.mn[k] = ref[r].aabb_min[k]; b.mx[k] = ref[r].aabb_max[k]; }
        return b;
    }
    // Conservative quantisation: lo rounded down, hi rounded up (the float rounding of the
    // ray's own transform to grid space is covered by the guard band B200RT_QEPS in slab_q).
    void quantise(const FBox &b, uint16_t *lo, uint16_t *hi) const {
        for (int k = 0; k < 3; k++) {
            double l = std::floor(((double)b.mn[k] - q.base[k]) / q.scale[k]);
            double h = std::ceil(((double)b.mx[k] - q.base[k]) / q.scale[k]);
            lo[k] = (uint16_t)std::min((double)B200RT_QMAX, std::max(0.0, l));
            hi[k] = (uint16_t)std::min((double)B200RT_QMAX, std::max(0.0, h));
        }
    }
    static DPair pack(const uint16_t *l0, const uint16_t *h0, const uint16_t *l1, const uint16_t *h1, uint32_t m0, uint32_t m1) {
        DPair p;
        p.a.x = l0[0] | ((uint32_t)h0[0] << 16); p.a.y = l0[1] | ((uint32_t)h0[1] << 16); p.a.z = l0[2] | ((uint32_t)h0[2] << 16);
        p.a.w = l1[0] | ((uint32_t)h1[0] << 16); p.b.x = l1[1] | ((uint32_t)h1[1] << 16); p.b.y = l1[2] | ((uint32_t)h1[2] << 16);
        p.b.z = m0; p.b.w = m1;
        return p;
    }
    // A node is either a reference node (r >= 0) or a virtual range produced by splitting a
    // leaf with more than 15 triangles (closest-hit results are tree-independent).
    uint32_t emit(long r, uint32_t first, uint32_t last, int depth) {
        max_depth = std::max(max_depth, depth);
        const bool leaf = r < 0 || ref[r].left == 0;
        if (r >= 0) { first = ref[r].first; last = ref[r].last; }
        if (leaf && last - first <= 15) return B200RT_LEAF | ((last - first) << 27) | first;
        const uint32_t idx = (uint32_t)out.size();
        out.emplace_back();
        FBox bl, br;
        uint32_t ml, mr;
        if (leaf) {
            const uint32_t mid = first + (last - first) / 2;
            bl = range_box(first, mid); br = range_box(mid, last);
            ml = emit(-1, first, mid, depth + 1); mr = emit(-1, mid, last, depth + 1);
        } else {
            bl = ref_box(ref[r].left); br = ref_box(ref[r].right);
            ml = emit(ref[r].left, 0, 0, depth + 1); mr = emit(ref[r].right, 0, 0, depth + 1);
        }
        uint16_t l0[3], h0[3], l1[3], h1[3];
        quantise(bl, l0, h0); quantise(br, l1, h1);
        out[idx] = pack(l0, h0, l1, h1, ml, mr);
        return idx;
    }
    void set_grid(const FBox &rb) {
        root_box = rb;
        for (int k = 0; k < 3; k++) {
            const double ext = (double)root_box.mx[k] - (double)root_box.mn[k];
            const double sc = ext > 0 ? ext / (B200RT_QMAX - 2.0) : 1.0;   // one spare grid unit on each side
            q.scale[k] = (float)sc;
            q.base[k] = (float)((double)root_box.mn[k] - sc);
            q.inv_scale[k] = (float)(1.0 / (double)q.scale[k]);
        }
    }
    // Non-recursive form for the common case (no leaf above 15 triangles): record index = rank of the
    // inner node in array order (the reference builder emits parents before children, so this is the
    // depth-first order too).  plan() is one serial pass; fill() is independent per node.
    std::vector<uint32_t> rank;
    std::vector<uint8_t> depth;      // 0 = not reachable from the root
    uint32_t n_inner = 0;
    bool plan(const b200rt_bvh_node *nodes, int n_nodes, int root, int n_tris) {
        ref = nodes;
        for (int k = 0; k < 3; k++) { q.base[k] = 0.f; q.scale[k] = 1.f; q.inv_scale[k] = 1.f; root_box.mn[k] = root_box.mx[k] = 0.f; }
        root_word = B200RT_LEAF; n_inner = 0; max_depth = 0;
        if (n_nodes <= 0 || n_tris <= 0) return true;
        rank.assign((size_t)n_nodes, 0u);
        depth.assign((size_t)n_nodes, 0);
        depth[(size_t)root] = 1;
        for (int i = 0; i < n_nodes; i++) {
            const b200rt_bvh_node &n = nodes[i];
            if (!depth[(size_t)i]) continue;
            if (n.left == 0) {
                if (n.last - n.first > 15) return false;
            } else {
                rank[(size_t)i] = n_inner++;
                const uint8_t dd = (uint8_t)std::min(255, depth[(size_t)i] + 1);
                depth[n.left] = dd; depth[n.right] = dd;
            }
            max_depth = std::max(max_depth, (int)depth[(size_t)i]);
        }
        set_grid(ref_box((uint32_t)root));
        root_word = word_of((uint32_t)root);
        return true;
    }
    uint32_t word_of(uint32_t r) const {
        const b200rt_bvh_node &n = ref[r];
        return n.left == 0 ? (B200RT_LEAF | ((n.last - n.first) << 27) | n.first) : rank[r];
    }
    void fill(int n_nodes, DPair *dst) const {
        parallel_for((size_t)n_nodes, [&](size_t a, size_t b) {
            for (size_t i = a; i < b; i++) {
                const b200rt_bvh_node &n = ref[i];
                if (n.left == 0 || !depth[i]) continue;
                uint16_t l0[3], h0[3], l1[3], h1[3];
                quantise(ref_box(n.left), l0, h0); quantise(ref_box(n.right), l1, h1);
                dst[rank[i]] = pack(l0, h0, l1, h1, word_of(n.left), word_of(n.right));
            }
        });
    }
    void run(const b200rt_bvh_node *nodes, int n_nodes, int root, const b200rt_triangle *t, int n_tris) {
        ref = nodes; tris = t;
        out.clear();
        for (int k = 0; k < 3; k++) { q.base[k] = 0.f; q.scale[k] = 1.f; q.inv_scale[k] = 1.f; root_box.mn[k] = root_box.mx[k] = 0.f; }
        root_word = B200RT_LEAF;
        if (n_nodes <= 0 || n_tris <= 0) { out.emplace_back(); return; }
        set_grid(ref_box((uint32_t)root));
        max_depth = 0;
        root_word = emit(root, 0, 0, 1);
        if (out.empty()) out.emplace_back();
    }
};

float __int_as_float_host(int v) { float f; memcpy(&f, &v, 4); return f; }

DTri make_tri(const b200rt_triangle &t) {
    DTri d;
    d.v0 = make_float4(t.p[2][0], t.p[2][1], t.p[2][2], 0.f);
    d.e1 = make_float4(t.p[0][0] - t.p[2][0], t.p[0][1] - t.p[2][1], t.p[0][2] - t.p[2][2], t.tan[0][3]);
    d.e2 = make_float4(t.p[1][0] - t.p[2][0], t.p[1][1] - t.p[2][1], t.p[1][2] - t.p[2][2], 0.f);
    return d;
}

bool validate_tree(const b200rt_bvh_node *nodes, int n_nodes, int root, int n_tris) {
    if (n_tris == 0) return true;
    if (!nodes || root < 0 || root >= n_nodes) return false;
    for (int i = 0; i < n_nodes; i++) {
        const auto &n = nodes[i];
        if (n.left == 0) { if (n.first > n.last || n.last > (uint32_t)n_tris) return false; }
        else if (n.left >= (uint32_t)n_nodes || n.right >= (uint32_t)n_nodes || n.left <= (uint32_t)i || n.right <= (uint32_t)i) return false;
    }
    return true;
}
}  // namespace

namespace {
// Cut through the top of the light BVH: start at the root and keep replacing the inner node with the largest
// surface area by its children until `want` boxes are open (or only leaves remain).
void light_top_cut(const b200rt_bvh_node *nodes, int n_nodes, int n_light_tris, int want, DGltfScene &g, std::vector<int> &cut_nodes) {
    g.n_light_top = 0;
    cut_nodes.clear();
    if (n_light_tris <= 0 || n_nodes <= 0 || want < 2) return;
    want = std::min(want, B200RT_LIGHT_TOP);
    auto area = [&](int i) {
        const float dx = nodes[i].aabb_max[0] - nodes[i].aabb_min[0], dy = nodes[i].aabb_max[1] - nodes[i].aabb_min[1], dz = nodes[i].aabb_max[2] - nodes[i].aabb_min[2];
        return dx * dy + dx * dz + dy * dz;
    };
    std::vector<int> open(1, 0);
    while ((int)open.size() < want) {
        int best = -1;
        for (int k = 0; k < (int)open.size(); k++)
            if (nodes[open[k]].left != 0 && (best < 0 || area(open[k]) > area(open[best]))) best = k;
        if (best < 0) break;
        const int id = open[best];
        open[best] = (int)nodes[id].left;
        open.push_back((int)nodes[id].right);
    }
    if (open.size() < 2) return;
    cut_nodes = open;
    for (int id : open) {
        float *b = g.light_top[g.n_light_top++];
        for (int k = 0; k < 3; k++) { b[k] = nodes[id].aabb_min[k]; b[3 + k] = nodes[id].aabb_max[k]; }
    }
}

}  // namespace

int upload_gltf(b200rt_ctx *ctx, const b200rt_scene_desc *s) {
    if (s->n_tris < 0 || s->n_tris >= (1 << 27)) return fail(ctx, B200RT_ERR_INVALID, "triangle count out of range");
    if (s->n_tris > 0 && (!s->tris || !s->materials)) return fail(ctx, B200RT_ERR_INVALID, "missing triangle/material arrays");
    if (s->n_light_tris < 0 || s->n_materials < 0) return fail(ctx, B200RT_ERR_INVALID, "negative array size");
    if (s->n_light_tris > 0 && !s->light_tris) return fail(ctx, B200RT_ERR_INVALID, "missing light triangle array");
    // The description's trees are optional (b200rt_host_load_gltf delivers none): by default the library builds the
    // trees it traverses itself.  They are validated only when they are actually used (B200RT_FLAG_REFERENCE_BVH;
    // hw6 always rebuilds, its reference tree is a SAH split of an arbitrary order).
    const bool use_ref_trees = (ctx->flags & B200RT_FLAG_REFERENCE_BVH) && s->mode != B200RT_MODE_HW6;
    if (use_ref_trees) {
        if (s->n_tris > 0 && !s->nodes) return fail(ctx, B200RT_ERR_INVALID, "B200RT_FLAG_REFERENCE_BVH needs the description's scene BVH");
        if (!validate_tree(s->nodes, s->n_nodes, s->root, s->n_tris)) return fail(ctx, B200RT_ERR_INVALID, "malformed scene BVH");
        if (s->n_light_tris > 0 && !validate_tree(s->light_nodes, s->n_light_nodes, 0, s->n_light_tris)) return fail(ctx, B200RT_ERR_INVALID, "malformed light BVH");
    }

    DGltfScene g{};
    g.hw8 = s->mode == B200RT_MODE_HW8;
    g.hw6 = s->mode == B200RT_MODE_HW6;
    g.weight_clamp = (ctx->flags & B200RT_FLAG_NO_FIREFLY_CLAMP) ? 3.402823466e38f : 6.f;
    for (int k = 0; k < 3; k++) g.bg[k] = s->bg_color[k];
    int rc;
#ifdef B200RT_DEV      // development builds only: per-phase times on stderr
    const bool timing = getenv("B200RT_UPLOAD_TIMING") != nullptr;
    auto t_last = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) {
        if (!timing) return;
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[upload] %-22s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
        t_last = now;
    };
#else
    auto lap = [](const char *) {};
#endif

    // The scene tree is rebuilt behind the ABI (bvh_rebuild.h) unless the caller asks for the tree of the
    // description; hw6 always rebuilds (its reference tree is a SAH split of an arbitrary order).
    RebuiltBvh &own = ctx->own_bvh;                        // buffers reused by the next upload
    std::vector<b200rt_triangle> &own_tris = ctx->own_tris;
    b200rt_scene_desc local;
    const bool rebuild = !use_ref_trees;
    if (rebuild && s->n_tris > 0) {
        rebuild_bvh_binned_sah(s->tris, s->n_tris, own);
        own_tris.resize((size_t)s->n_tris);
        {
            const b200rt_triangle *src = s->tris;
            parallel_for((size_t)s->n_tris, [&](size_t a, size_t b) { for (size_t i = a; i < b; i++) own_tris[i] = src[own.order[i]]; });
        }
        local = *s;
        local.tris = own_tris.data();
        local.nodes = own.nodes.data(); local.n_nodes = (int32_t)own.nodes.size(); local.root = 0;
        s = &local;
        lap("rebuild BVH");
    }

    // the light set keeps the reference's ORDER for sampling (light_sample_tris); its traversal tree is
    // rebuilt like the scene's, over a permuted copy (the pdf sum visits every pierced box: tree quality = cost)
    RebuiltBvh light_own;
    std::vector<b200rt_triangle> light_perm;
    const b200rt_triangle *ltri_src = s->light_tris;
    const b200rt_bvh_node *lnode_src = s->light_nodes;
    int n_lnodes = s->n_light_nodes;
    if (rebuild && s->n_light_tris > 0) {
        rebuild_bvh_binned_sah(s->light_tris, s->n_light_tris, light_own, true);
        light_perm.resize((size_t)s->n_light_tris);
        for (int i = 0; i < s->n_light_tris; i++) light_perm[(size_t)i] = s->light_tris[light_own.order[(size_t)i]];
        ltri_src = light_perm.data();
        lnode_src = light_own.nodes.data(); n_lnodes = (int)light_own.nodes.size();
    }

    // 1. plan both trees (serial pass each); a leaf above 15 triangles needs the recursive splitter
    Flattener fl, ll;
    const bool fast_scene = fl.plan(s->nodes, s->n_nodes, s->root, s->n_tris);
    if (!fast_scene) fl.run(s->nodes, s->n_nodes, s->root, s->tris, s->n_tris);
    const bool fast_light = ll.plan(lnode_src, n_lnodes, 0, s->n_light_tris);
    if (!fast_light) ll.run(lnode_src, n_lnodes, 0, ltri_src, s->n_light_tris);
    if (fl.max_depth > B200RT_STACK) return fail(ctx, B200RT_ERR_STACK, "scene BVH deeper than the traversal stack");
    if (ll.max_depth > B200RT_STACK) return fail(ctx, B200RT_ERR_STACK, "light BVH deeper than the traversal stack");
    const size_t n_pairs = std::max<size_t>(1, fast_scene ? fl.n_inner : fl.out.size());
    const size_t n_lpairs = std::max<size_t>(1, fast_light ? ll.n_inner : ll.out.size());
    lap("plan");

    // 2. arena layout
    const size_t nt = (size_t)s->n_tris, nl = (size_t)s->n_light_tris, nm = (size_t)s->n_materials;
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t o = off; off = align256(off + std::max<size_t>(bytes, 16)); return o; };
    const size_t o_nodes = take(n_pairs * sizeof(DPair)), o_tris = take(nt * sizeof(DTri)), o_shade = take(nt * sizeof(DShade));
    const size_t o_lnodes = take(n_lpairs * sizeof(DPair)), o_ltris = take(nl * sizeof(DTri)), o_lnorm = take(nl * sizeof(DLightNormals));
    const size_t o_lsample = take(nl * sizeof(DTri));
    const size_t o_mats = take(nm * sizeof(DMaterial)), o_lut = take(256 * sizeof(float));
    // images (hw8 only; the environment map is the last one): RGBA8 words, one array per image, in the same arena
    std::vector<const b200rt_image *> imgs;
    if (g.hw8) {
        for (int i = 0; i < s->n_images; i++) imgs.push_back(&s->images[i]);
        if (s->env_map) imgs.push_back(s->env_map);
    }
    std::vector<size_t> o_img(imgs.size());
    for (size_t i = 0; i < imgs.size(); i++) {
        if (!imgs[i]->rgb || imgs[i]->width <= 0 || imgs[i]->height <= 0) return fail(ctx, B200RT_ERR_INVALID, "empty image");
        if ((long long)imgs[i]->width * imgs[i]->height > (1ll << 30)) return fail(ctx, B200RT_ERR_UNSUPPORTED, "image above 2^30 texels");
        o_img[i] = take((size_t)imgs[i]->width * imgs[i]->height * sizeof(uint32_t));
    }
    if ((rc = ensure_arena(ctx, off))) return rc;
    char *pin = ctx->arena_pin, *dev = ctx->arena_dev;
    // the staging buffer may still feed the copy of the previous upload
    CK(cudaStreamSynchronize(ctx->stream));
    lap("arena");

    // 3. records, built in the staging buffer
    DPair *pairs = reinterpret_cast<DPair *>(pin + o_nodes);
    if (fast_scene) { if (fl.n_inner) fl.fill(s->n_nodes, pairs); else pairs[0] = DPair{}; }
    else memcpy(pairs, fl.out.data(), fl.out.size() * sizeof(DPair));
    g.root = fl.root_word; g.q = fl.q;
    lap("scene BVH records");

    DTri *tris = reinterpret_cast<DTri *>(pin + o_tris);
    DShade *shade = reinterpret_cast<DShade *>(pin + o_shade);
    std::vector<int> bad_material(1, 0);
    parallel_for(nt, [&](size_t a, size_t b) {
        for (size_t i = a; i < b; i++) {
            const b200rt_triangle &t = s->tris[i];
            if (t.material >= (uint32_t)s->n_materials) { bad_material[0] = 1; continue; }
            tris[i] = make_tri(t);
            tris[i].v0.w = __int_as_float_host((int)t.material);
            DShade &q = shade[i];
            q.q[0] = make_float4(t.n[0][0], t.n[0][1], t.n[0][2], t.uv[0][0]);
            q.q[1] = make_float4(t.n[1][0], t.n[1][1], t.n[1][2], t.uv[0][1]);
            q.q[2] = make_float4(t.n[2][0], t.n[2][1], t.n[2][2], t.uv[1][0]);
            q.q[3] = make_float4(t.tan[0][0], t.tan[0][1], t.tan[0][2], t.uv[1][1]);
            q.q[4] = make_float4(t.tan[1][0], t.tan[1][1], t.tan[1][2], t.uv[2][0]);
            q.q[5] = make_float4(t.tan[2][0], t.tan[2][1], t.tan[2][2], t.uv[2][1]);
        }
    });
    if (bad_material[0]) return fail(ctx, B200RT_ERR_INVALID, "material index out of range");
    lap("triangle records");

    // lights: FiguresMix::figures_ + its BVH (hw8/src/include/distributions.h:99-115)
    g.n_light_tris = s->n_light_tris;
    DPair *lpairs = reinterpret_cast<DPair *>(pin + o_lnodes);
    if (fast_light) { if (ll.n_inner) ll.fill(n_lnodes, lpairs); else lpairs[0] = DPair{}; }
    else memcpy(lpairs, ll.out.data(), ll.out.size() * sizeof(DPair));
    g.light_root = ll.root_word; g.lq = ll.q;
    for (int k = 0; k < 3; k++) { g.light_min[k] = ll.root_box.mn[k]; g.light_max[k] = ll.root_box.mx[k]; }
    std::vector<int> cut_nodes;
    light_top_cut(lnode_src, n_lnodes, s->n_light_tris, ctx->light_top_max, g, cut_nodes);
    // the traversal may start at the cut's nodes when the flattener kept the tree's shape (one record per inner node)
    g.light_top_start = fast_light && !cut_nodes.empty() ? 1 : 0;
    for (size_t k = 0; k < cut_nodes.size() && g.light_top_start; k++) g.light_top_word[k] = ll.word_of((uint32_t)cut_nodes[k]);
    DTri *ltris = reinterpret_cast<DTri *>(pin + o_ltris);
    DLightNormals *lnorm = reinterpret_cast<DLightNormals *>(pin + o_lnorm);
    DTri *lsample = reinterpret_cast<DTri *>(pin + o_lsample);
    auto light_record = [](const b200rt_triangle &t, DTri &d) {
        d = make_tri(t);
        // pointProb = 1.0 / (0.5 * |b.cross(c)|) in the reference's float/double mix (distributions.h:71-76)
        float bx = d.e1.x, by = d.e1.y, bz = d.e1.z, cx = d.e2.x, cy = d.e2.y, cz = d.e2.z;
        float nx = bz * cy - by * cz, ny = bx * cz - bz * cx, nz = by * cx - bx * cy;
        float len = (float)std::sqrt((double)(nx * nx + ny * ny + nz * nz));
        d.v0.w = (float)(1.0 / (0.5 * len));
    };
    for (size_t i = 0; i < nl; i++) {
        const b200rt_triangle &t = ltri_src[i];
        light_record(t, ltris[i]);
        light_record(s->light_tris[i], lsample[i]);
        lnorm[i].n0 = make_float4(t.n[0][0], t.n[0][1], t.n[0][2], 0);
        lnorm[i].n1 = make_float4(t.n[1][0], t.n[1][1], t.n[1][2], 0);
        lnorm[i].n2 = make_float4(t.n[2][0], t.n[2][1], t.n[2][2], 0);
    }
    lap("light set");

    // textures: RGB8 -> RGBA8 words straight into the staging arena, on all host cores
    std::vector<DTexels> tex(imgs.size());
    for (size_t i = 0; i < imgs.size(); i++) {
        const size_t n = (size_t)imgs[i]->width * imgs[i]->height;
        uint32_t *dst = reinterpret_cast<uint32_t *>(pin + o_img[i]);
        const uint8_t *src = imgs[i]->rgb;
        parallel_for(n, [&](size_t a, size_t b) {
            for (size_t k = a; k < b; k++) dst[k] = (uint32_t)src[3 * k] | ((uint32_t)src[3 * k + 1] << 8) | ((uint32_t)src[3 * k + 2] << 16) | 0xFF000000u;
        });
        tex[i] = reinterpret_cast<DTexels>(dev + o_img[i]);
    }
    if (g.hw8 && s->env_map) { g.env = tex.back(); g.env_w = s->env_map->width; g.env_h = s->env_map->height; }
    lap("textures");

    DMaterial *mats = reinterpret_cast<DMaterial *>(pin + o_mats);
    for (size_t i = 0; i < nm; i++) {
        const b200rt_material &m = s->materials[i];
        DMaterial &d = mats[i];
        memset(&d, 0, sizeof(d));
        for (int k = 0; k < 3; k++) { d.base_color[k] = m.base_color[k]; d.emission[k] = m.emission[k]; }
        d.metallic_factor = m.metallic_factor; d.roughness_factor = m.roughness_factor;
        d.kind = m.kind; d.ior = m.ior;
        auto bind = [&](int32_t idx, DTexels &t, int32_t &w, int32_t &h) -> bool {
            if (!g.hw8 || idx < 0) return true;
            if (idx >= s->n_images) return false;
            t = tex[(size_t)idx]; w = s->images[idx].width; h = s->images[idx].height;
            return true;
        };
        if (!bind(m.base_color_image, d.tex_base, d.w_base, d.h_base) || !bind(m.emissive_image, d.tex_emissive, d.w_emissive, d.h_emissive) ||
            !bind(m.metallic_roughness_image, d.tex_mr, d.w_mr, d.h_mr) || !bind(m.normal_image, d.tex_normal, d.w_normal, d.h_normal))
            return fail(ctx, B200RT_ERR_INVALID, "image index out of range");
    }
    // sRGB decode table with the reference's own arithmetic: powf(float(1./255) * c, 2.2f) (scene.cpp:11-13)
    float *lut = reinterpret_cast<float *>(pin + o_lut);
    for (int c = 0; c < 256; c++) lut[c] = std::pow((float)(1. / 255) * (1.f * (float)c), 2.2f);

    // 4. one copy for everything
    CK(cudaMemcpyAsync(dev, pin, off, cudaMemcpyHostToDevice, ctx->stream));
    g.nodes = reinterpret_cast<const DPair *>(dev + o_nodes);
    g.tris = reinterpret_cast<const DTri *>(dev + o_tris);
    g.shade = reinterpret_cast<const DShade *>(dev + o_shade);
    g.light_nodes = reinterpret_cast<const DPair *>(dev + o_lnodes);
    g.light_tris = reinterpret_cast<const DTri *>(dev + o_ltris);
    g.light_normals = reinterpret_cast<const DLightNormals *>(dev + o_lnorm);
    g.light_sample_tris = reinterpret_cast<const DTri *>(dev + o_lsample);
    g.materials = reinterpret_cast<const DMaterial *>(dev + o_mats);
    ctx->lut_srgb = reinterpret_cast<float *>(dev + o_lut);
    ctx->gs = g;
    ctx->arena_used = off; ctx->n_materials = (int)nm;
    ctx->n_pairs = n_pairs; ctx->n_dtris = nt;
    lap("materials, lut, copy");
    return B200RT_OK;
}


namespace {
// hw5 (text scene with TRIANGLE figures): analytic figures -> constant memory like hw4; triangles -> baked to world
// space (position + conj(rotation) applied, hw5/src/primitives.cpp:13-33 runs the same transform per ray) and uploaded
// through the glTF path as an hw6-style scene with one material per triangle; the trees of the description (which
// mix figure kinds) are not used: closest hits and light-pdf sums are tree-independent, the library builds its own.
struct HQuat { float x, y, z, w; };
inline HQuat hq_mul(const HQuat &a, const HQuat &b) {
    // Quaternion::operator* with the reference's negated cross (hw5/src/include/quaternion.h:36-38)
    return {a.w * b.x + b.w * a.x + (a.z * b.y - a.y * b.z), a.w * b.y + b.w * a.y + (a.x * b.z - a.z * b.x),
            a.w * b.z + b.w * a.z + (a.y * b.x - a.x * b.y), a.w * b.w - (a.x * b.x + a.y * b.y + a.z * b.z)};
}
inline void hq_to_world(const b200rt_primitive &f, const float *v, float *out) {
    const HQuat q{-f.rotation[0], -f.rotation[1], -f.rotation[2], f.rotation[3]};      // rotation.conjugate()
    const HQuat r = hq_mul(hq_mul(q, HQuat{v[0], v[1], v[2], 0.f}), HQuat{-q.x, -q.y, -q.z, q.w});
    out[0] = f.position[0] + r.x; out[1] = f.position[1] + r.y; out[2] = f.position[2] + r.z;
}

}  // namespace

int upload_hw5(b200rt_ctx *ctx, const b200rt_scene_desc *s) {
    if (s->n_prims < 0 || (s->n_prims && !s->prims) || s->n_light_prims < 0 || (s->n_light_prims && !s->light_prims))
        return fail(ctx, B200RT_ERR_INVALID, "missing figure / light arrays");
    static DAnalyticScene a;
    memset(&a, 0, sizeof(a));
    std::vector<b200rt_triangle> tris, ltris;
    std::vector<b200rt_material> mats;
    auto bake = [](const b200rt_primitive &f, uint32_t material) {
        b200rt_triangle t{};
        hq_to_world(f, f.data, t.p[0]); hq_to_world(f, f.data2, t.p[1]); hq_to_world(f, f.data3, t.p[2]);
        for (int v = 0; v < 3; v++) { t.n[v][2] = 1.f; t.tan[v][0] = 1.f; t.tan[v][3] = 1.f; }   // unused by the hw5 / hw6 integrator
        t.material = material;
        return t;
    };
    for (int i = 0; i < s->n_prims; i++) {
        const b200rt_primitive &f = s->prims[i];
        if (f.type == B200RT_PRIM_TRIANGLE) {
            b200rt_material m{};
            for (int k = 0; k < 3; k++) { m.base_color[k] = f.color[k]; m.emission[k] = f.emission[k]; }
            m.base_color_image = m.emissive_image = m.metallic_roughness_image = m.normal_image = -1;
            m.kind = f.material; m.ior = f.ior;
            tris.push_back(bake(f, (uint32_t)mats.size()));
            mats.push_back(m);
        } else {
            if (a.n_prims >= B200RT_MAX_PRIMS) return fail(ctx, B200RT_ERR_UNSUPPORTED, "too many analytic figures (planes / ellipsoids / boxes) for the constant-memory scan");
            a.prims[a.n_prims++] = f;
        }
    }
    // FiguresMix::figures_ order; analytic lights are found among the scanned figures by value (the list holds copies)
    std::vector<int32_t> order((size_t)s->n_light_prims);
    for (int i = 0; i < s->n_light_prims; i++) {
        const b200rt_primitive &f = s->light_prims[i];
        if (f.type == B200RT_PRIM_TRIANGLE) {
            order[(size_t)i] = ~(int32_t)ltris.size();
            ltris.push_back(bake(f, 0));
        } else {
            int found = -1;
            for (int k = 0; k < a.n_prims && found < 0; k++)
                if (memcmp(&a.prims[k], &f, sizeof(f)) == 0) found = k;
            if (found < 0 || (f.type != B200RT_PRIM_BOX && f.type != B200RT_PRIM_ELLIPSOID)) return fail(ctx, B200RT_ERR_INVALID, "light figure not among the scene figures");
            order[(size_t)i] = found;
            bool seen = false;
            for (int k = 0; k < a.n_light_prims; k++) seen = seen || a.light_prims[k] == found;
            if (!seen) a.light_prims[a.n_light_prims++] = found;
        }
    }
    a.mode = s->mode;
    for (int k = 0; k < 3; k++) a.bg[k] = s->bg_color[k];
    CK(analytic_upload(a, ctx->stream));
    ctx->hw5_has_tris = !tris.empty();
    ctx->gs = DGltfScene{};
    for (int k = 0; k < 3; k++) ctx->gs.bg[k] = s->bg_color[k];
    if (!tris.empty()) {
        b200rt_scene_desc t = *s;
        t.mode = B200RT_MODE_HW6;
        t.prims = nullptr; t.n_prims = 0; t.light_prims = nullptr; t.n_light_prims = 0;
        t.tris = tris.data(); t.n_tris = (int32_t)tris.size();
        t.nodes = nullptr; t.n_nodes = 0; t.root = 0;
        t.light_tris = ltris.data(); t.n_light_tris = (int32_t)ltris.size();
        t.light_nodes = nullptr; t.n_light_nodes = 0;
        t.materials = mats.data(); t.n_materials = (int32_t)mats.size();
        t.images = nullptr; t.n_images = 0; t.env_map = nullptr;
        const int rc = upload_gltf(ctx, &t);
        if (rc) return rc;
    }
    if (s->n_light_prims > ctx->hw5_light_cap) {
        cudaFree(ctx->hw5_light_order); ctx->hw5_light_order = nullptr; ctx->hw5_light_cap = 0;
        CK(cudaMalloc(&ctx->hw5_light_order, (size_t)s->n_light_prims * sizeof(int32_t)));
        ctx->hw5_light_cap = s->n_light_prims;
    }
    if (s->n_light_prims > 0) CK(cudaMemcpyAsync(ctx->hw5_light_order, order.data(), order.size() * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));      // `order` and the staging vectors die with this frame
    ctx->hw5_n_lights = s->n_light_prims;
    return B200RT_OK;
}


// ---- scene replication (include/b200rt.h) ----------------------------------------------------------------
namespace {
struct SceneImage {
    uint64_t magic, base, bytes, lut;
    int32_t n_materials, pad;
    uint64_t n_pairs, n_dtris;
    DGltfScene gs;
    b200rt_scene_desc desc;      // scalars only (pointers nulled)
};
static_assert(sizeof(SceneImage) <= sizeof(b200rt_scene_image), "b200rt_scene_image too small");
const uint64_t kSceneMagic = 0x6232303072743031ull;     // "b200rt01"

// texel pointers live INSIDE the blob (DMaterial records): shift them by the distance between the two arenas
__global__ void rebase_materials_kernel(DMaterial *m, int n, long long delta) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    auto fix = [delta](DTexels &t) { if (t) t = reinterpret_cast<DTexels>(reinterpret_cast<const char *>(t) + delta); };
    fix(m[i].tex_base); fix(m[i].tex_emissive); fix(m[i].tex_mr); fix(m[i].tex_normal);
}
}  // namespace

extern "C" {

int b200rt_export_scene(b200rt_ctx *ctx, b200rt_scene_image *hdr, void **device_blob, uint64_t *bytes) {
    if (!ctx || !hdr || !device_blob || !bytes) return B200RT_ERR_INVALID;
    if (!ctx->has_scene) return fail(ctx, B200RT_ERR_NO_SCENE, "export before upload_scene");
    if (ctx->desc.mode < B200RT_MODE_HW6 || !ctx->arena_dev) return fail(ctx, B200RT_ERR_UNSUPPORTED, "scene export exists for the glTF modes only");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    memset(hdr, 0, sizeof(*hdr));
    SceneImage im{};
    im.magic = kSceneMagic; im.base = (uint64_t)(uintptr_t)ctx->arena_dev; im.bytes = ctx->arena_used;
    im.lut = (uint64_t)(uintptr_t)ctx->lut_srgb; im.n_materials = ctx->n_materials;
    im.n_pairs = ctx->n_pairs; im.n_dtris = ctx->n_dtris;
    im.gs = ctx->gs; im.desc = ctx->desc;
    memcpy(hdr, &im, sizeof(im));
    *device_blob = ctx->arena_dev;
    *bytes = ctx->arena_used;
    return B200RT_OK;
}

int b200rt_import_scene_begin(b200rt_ctx *ctx, const b200rt_scene_image *hdr, void **device_dst) {
    if (!ctx || !hdr || !device_dst) return B200RT_ERR_INVALID;
    SceneImage im;
    memcpy(&im, hdr, sizeof(im));
    if (im.magic != kSceneMagic || im.bytes == 0) return fail(ctx, B200RT_ERR_INVALID, "not a scene image");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    release_scene(ctx);
    int rc = ensure_arena(ctx, (size_t)im.bytes);
    if (rc) return rc;
    *device_dst = ctx->arena_dev;
    return B200RT_OK;
}

int b200rt_import_scene_end(b200rt_ctx *ctx, const b200rt_scene_image *hdr) {
    if (!ctx || !hdr) return B200RT_ERR_INVALID;
    SceneImage im;
    memcpy(&im, hdr, sizeof(im));
    if (im.magic != kSceneMagic || !ctx->arena_dev || ctx->arena_cap < im.bytes) return fail(ctx, B200RT_ERR_INVALID, "import_scene_end without a matching begin");
    CK(cudaSetDevice(ctx->device));
    const long long delta = (long long)((intptr_t)ctx->arena_dev - (intptr_t)im.base);
    auto move = [delta](auto *&p) { if (p) p = reinterpret_cast<std::remove_reference_t<decltype(p)>>(reinterpret_cast<const char *>(p) + delta); };
    DGltfScene g = im.gs;
    move(g.nodes); move(g.tris); move(g.shade); move(g.light_nodes); move(g.light_tris); move(g.light_sample_tris);
    move(g.light_normals); move(g.materials); move(g.env);
    if (im.n_materials > 0) {
        rebase_materials_kernel<<<(im.n_materials + 63) / 64, 64, 0, ctx->stream>>>(const_cast<DMaterial *>(g.materials), im.n_materials, delta);
        CK(cudaGetLastError());
    }
    g.weight_clamp = (ctx->flags & B200RT_FLAG_NO_FIREFLY_CLAMP) ? 3.402823466e38f : 6.f;      // a property of THIS context
    ctx->gs = g;
    ctx->lut_srgb = reinterpret_cast<float *>(reinterpret_cast<char *>((uintptr_t)im.lut) + delta);
    ctx->arena_used = (size_t)im.bytes; ctx->n_materials = im.n_materials;
    ctx->n_pairs = im.n_pairs; ctx->n_dtris = im.n_dtris;
    ctx->hw5_has_tris = false; ctx->hw5_n_lights = 0;
    ctx->desc = im.desc;
    ctx->has_scene = true;
    return b200rt_clear(ctx);
}

int b200rt_replicate_scene(b200rt_ctx *src, b200rt_ctx *dst) {
    if (!src || !dst || src == dst) return B200RT_ERR_INVALID;
    b200rt_scene_image hdr;
    void *from = nullptr, *to = nullptr;
    uint64_t bytes = 0;
    int rc = b200rt_export_scene(src, &hdr, &from, &bytes);
    if (rc) { dst->last_error = src->last_error; return rc; }
    if ((rc = b200rt_import_scene_begin(dst, &hdr, &to))) return rc;
    b200rt_ctx *ctx = dst;
    CK(cudaMemcpyPeerAsync(to, dst->device, from, src->device, (size_t)bytes, dst->stream));
    return b200rt_import_scene_end(dst, &hdr);
}


}  // extern "C"
