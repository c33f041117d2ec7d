// C ABI of the B200 render core: context, scene flattening + upload, launches, resolve.
// See include/b200rt.h for the contract and the reference call sites each entry replaces.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <type_traits>
#include <vector>
#include "analytic_kernels.h"
#include "b200rt.h"
#include "device_scene.h"
#include "gltf_kernel.cuh"
#include "wavefront_trace.cuh"
#include "hw6_kernel.cuh"
#include "bvh_rebuild.h"
#include <chrono>
#include <thread>
#ifdef B200RT_DBG_SORT
// development experiment only (not in the product build): host-synchronised sort of a bounce's ray queue,
// to measure what ray coherence is worth to the traversal before building a device-side counting sort
#include <thrust/device_ptr.h>
#include <thrust/sort.h>
#include <thrust/execution_policy.h>
namespace b200rt {
__device__ __forceinline__ uint32_t dbg_spread3(uint32_t x) {   // 10 bits -> every third bit
    x &= 0x3FFu;
    x = (x | (x << 16)) & 0x030000FFu;
    x = (x | (x << 8)) & 0x0300F00Fu;
    x = (x | (x << 4)) & 0x030C30C3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}
__global__ void dbg_sort_keys(const DGltfScene sc, const WaveBuffers wb, const int *q, int n, uint32_t *keys, int mode, int obits) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int p = q[i];
    const float4 ro = wb.ray_o[p], rd = wb.ray_d[p];
    const float gmax = B200RT_QBITS == 15 ? 32768.f : 65536.f;
    const float s = (float)(1 << obits) / gmax;
    const int m = (1 << obits) - 1;
    const uint32_t gx = (uint32_t)min(max((int)((ro.x - sc.q.base[0]) * sc.q.inv_scale[0] * s), 0), m);
    const uint32_t gy = (uint32_t)min(max((int)((ro.y - sc.q.base[1]) * sc.q.inv_scale[1] * s), 0), m);
    const uint32_t gz = (uint32_t)min(max((int)((ro.z - sc.q.base[2]) * sc.q.inv_scale[2] * s), 0), m);
    const uint32_t mort = dbg_spread3(gx) | (dbg_spread3(gy) << 1) | (dbg_spread3(gz) << 2);
    const uint32_t oct = (rd.x < 0.f ? 1u : 0u) | (rd.y < 0.f ? 2u : 0u) | (rd.z < 0.f ? 4u : 0u);
    keys[i] = mode == 0 ? oct : mode == 1 ? ((oct << (3 * obits)) | mort) : mode == 2 ? ((mort << 3) | oct) : mort;
}
}
#endif

using namespace b200rt;

#define B200RT_PIPE 3

struct b200rt_ctx {
    int device = 0;
    uint32_t seed = 0;
    int flags = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::string last_error;

    bool has_scene = false;
    b200rt_scene_desc desc{};     // scalar fields only (pointers are not retained)
    DGltfScene gs{};
    // Scene arena: ONE device allocation and ONE pinned staging buffer of the same layout hold every
    // scene array (BVH records, triangle / shading records, light set, materials, sRGB table); an upload
    // builds the records in the staging buffer on all host cores and moves them with one async copy.
    // Both grow on demand and are reused by later uploads (no cudaMalloc / cudaFree per frame).
    char *arena_dev = nullptr, *arena_pin = nullptr;
    size_t arena_cap = 0, arena_used = 0;     // arena_used / n_materials: what b200rt_export_scene hands out
    int n_materials = 0;
    float *lut_srgb = nullptr;
    RebuiltBvh own_bvh;                      // host scratch of the BVH rebuild (bvh_rebuild.h)
    std::vector<b200rt_triangle> own_tris;   // triangles in the rebuilt tree's order

    // hw5: FiguresMix order over the mixed light kinds (>= 0 analytic figure, < 0 ~emissive-triangle index)
    int32_t *hw5_light_order = nullptr;
    int hw5_n_lights = 0, hw5_light_cap = 0;
    bool hw5_has_tris = false;

    int acc_w = 0, acc_h = 0, n_slabs = 0;
    float *total = nullptr, *slabs = nullptr;
    uint8_t *rgb8 = nullptr;
    int32_t *ids = nullptr;
    DCounters *counters = nullptr;
    // wavefront pipeline state (glTF modes)
    // Up to B200RT_PIPE waves are in flight, each on its own stream with its own buffers (pipe 0 uses
    // the main stream): while one wave's persistent traversal drains, the next wave's kernels fill the SMs.
    WaveBuffers wave[B200RT_PIPE]{};
    cudaStream_t pipe_stream[B200RT_PIPE] = {};
    cudaEvent_t ev_fork = nullptr, ev_join[B200RT_PIPE] = {}, ev_acc[B200RT_PIPE] = {};
    int wf_pipe = 2;
    size_t wave_capacity = 0;
    int num_sms = 0;
    int blocks_extend[2] = {0, 0}, blocks_light[2] = {0, 0}, blocks_trace[2] = {0, 0};
    long long wf_wave_paths = 4ll << 20;
    int light_top_max = B200RT_LIGHT_TOP;  // boxes of the light-BVH cut used as the light-pdf pre-filter (0/1 = root box only)
    int wf_flat_per_sm = 16;         // blocks per SM of the grid-stride kernels (generate, shade); B200RT_WF_FLAT_PER_SM
    bool wf_first_implicit = true;   // camera paths: only the direction is stored by generate (B200RT_WF_FIRST_IMPLICIT=0: whole record)
    bool wf_drop_dark = true;    // paths whose last bounce cannot reach an emitter end one bounce early (black background only)
    bool wf_direct = true;       // shade finishes bounces whose ray misses the light root box (no light-stage entry)
    bool wf_fused = false;       // one traversal kernel per bounce (wf_trace) instead of extend + light
    int wf_batch = B200RT_WF_BATCH, wf_leaf_t = B200RT_WF_LEAF_T, wf_leaf_t_light = B200RT_WF_LEAF_T_LIGHT, wf_blocks_cap = 0;
    int wf_batch_light = B200RT_WF_BATCH_LIGHT;
    b200rt_stats stats{};
    bool timing_pending = false;
    // per-stage timing of the wavefront: event pairs (begin, end, stage) recorded around launches
    std::vector<cudaEvent_t> ev_pool;
    std::vector<int> ev_stage;      // stage of pair i (0 extend, 1 shade, 2 light)
    size_t ev_used = 0;
    bool stage_timing = true;

    int stamp(bool begin, int stage, cudaStream_t on) {
        if (!stage_timing) return 0;
        if (ev_used == ev_pool.size()) {
            cudaEvent_t e;
            if (cudaEventCreate(&e) != cudaSuccess) return -1;
            ev_pool.push_back(e);
        }
        if (begin) ev_stage.push_back(stage);
        return cudaEventRecord(ev_pool[ev_used++], on) == cudaSuccess ? 0 : -1;
    }
};

namespace {

int fail(b200rt_ctx *c, int code, const char *what, cudaError_t e = cudaSuccess) {
    if (c) {
        c->last_error = what;
        if (e != cudaSuccess) { c->last_error += ": "; c->last_error += cudaGetErrorString(e); }
    }
    return code;
}
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(ctx, B200RT_ERR_CUDA, #call, e_); } while (0)

// Runs f(begin, end) over [0, n) on the host cores (record building is embarrassingly parallel).
template <class F>
void parallel_for(size_t n, const F &f) {
    const unsigned nt = (unsigned)std::min<size_t>((size_t)host_threads(), (n + 16383) / 16384);
    if (nt <= 1) { f((size_t)0, n); return; }
    const size_t chunk = (n + nt - 1) / nt;
    HostPool::get().run(nt, [&](unsigned t) {           // persistent workers (bvh_rebuild.h)
        const size_t a = t * chunk, b = std::min(n, a + chunk);
        if (a < b) f(a, b);
    });
}

inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

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

// L2 residency of the scene (measured, OFF by default; B200RT_L2_PERSIST=1 turns it on).  One stage of a wave streams
// ~0.5 GB of path state through the 126 MB L2 and ncu showed the shade stage's L2 hit rate at 32 % although nodes +
// triangles + shading records (the arena) are only ~50 MB, so the arena was declared a PERSISTING access-policy
// window on every stream of the context.  Result on B200 (1080p probe, serial waves): extend unchanged, shade 8.2 ->
// 9.5 ms with the whole arena persisting, 15.8 ms with a 64 MB carve-out: the stages gain more from finding the
// state the previous stage just wrote in L2 than from scene hits (profiles/r01_sweep_l2_state.log).
int apply_l2_policy(b200rt_ctx *ctx, size_t bytes) {
    static const bool on = getenv("B200RT_L2_PERSIST") && atoi(getenv("B200RT_L2_PERSIST")) != 0;
    int max_persist = 0, max_window = 0;
    cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, ctx->device);
    cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, ctx->device);
    if (getenv("B200RT_DEBUG_L2")) fprintf(stderr, "b200rt: L2 persisting max %d MB, window max %d MB, arena %zu MB, policy %s\n",
                                           max_persist >> 20, max_window >> 20, bytes >> 20, on ? "on" : "off");
    if (!on || max_persist <= 0 || max_window <= 0 || bytes == 0) return B200RT_OK;
    size_t carve = std::min<size_t>((size_t)max_persist, bytes);
    if (const char *e = getenv("B200RT_L2_PERSIST_MB")) carve = std::min<size_t>((size_t)max_persist, (size_t)atoll(e) << 20);
    if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve) != cudaSuccess) { cudaGetLastError(); return B200RT_OK; }
    cudaStreamAttrValue v{};
    v.accessPolicyWindow.base_ptr = ctx->arena_dev;
    v.accessPolicyWindow.num_bytes = std::min<size_t>(bytes, (size_t)max_window);
    v.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)carve / (double)v.accessPolicyWindow.num_bytes);
    v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    for (int k = 0; k < B200RT_PIPE; k++) {
        cudaStream_t st = k == 0 ? ctx->stream : ctx->pipe_stream[k];
        if (st && cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &v) != cudaSuccess) cudaGetLastError();
    }
    return B200RT_OK;
}

// The arena stays allocated for the next upload; b200rt_destroy frees it.
void release_scene(b200rt_ctx *c) { c->has_scene = false; }

void free_scene_memory(b200rt_ctx *c) {
    cudaFree(c->arena_dev); cudaFreeHost(c->arena_pin);
    c->arena_dev = c->arena_pin = nullptr;
    c->arena_cap = 0;
    c->has_scene = false;
}

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

// ---- small kernels ---------------------------------------------------------------------------
// Adds the per-chunk slabs to the running total in a fixed order (deterministic sum).
__global__ void fold_slabs_kernel(float *__restrict__ total, const float *__restrict__ slabs, int n_slabs, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float acc = total[i];
    for (int s = 0; s < n_slabs; s++) acc += slabs[(size_t)s * n + i];
    total[i] = acc;
}

// K5 resolve: mean, ACES fit, clamp, gamma 1/2.2, round to RGB8 (hw8/src/color.cpp:4-31,
// hw8/src/scene.cpp:176).  hw1 quantises the raw colour (hw1/src/color.cpp:13-19).
// The gamma pow is evaluated in double: the reference calls double pow() in hw4-hw8 and powf in
// hw2; rounding the double result to float reproduces both except on exact rounding ties.
__global__ void resolve_kernel(const float *__restrict__ total, uint8_t *__restrict__ rgb, size_t n_px, float inv_spp, int tonemap) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_px) return;
    const float gamma = (float)(1. / 2.2);
#pragma unroll
    for (int c = 0; c < 3; c++) {
        float x = inv_spp * total[3 * i + c];
        if (tonemap) {
            const float a = 2.51f, b = 0.03f, cc = 2.43f, d = 0.59f, e = 0.14f;
            float num = __fmul_rn(x, __fadd_rn(__fmul_rn(a, x), b));
            float den = __fadd_rn(__fmul_rn(x, __fadd_rn(__fmul_rn(cc, x), d)), e);
            float y = __fdiv_rn(num, den);
            y = fminf(1.f, fmaxf(0.f, y));
            x = (float)pow((double)y, (double)gamma);
        }
        float q = roundf(__fmul_rn(255.f, x));
        rgb[3 * i + c] = (uint8_t)q;
    }
}

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
    for (int k = 0; k < 3; k++) g.bg[k] = s->bg_color[k];
    int rc;
    const bool timing = getenv("B200RT_UPLOAD_TIMING") != nullptr;
    auto t_last = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) {
        if (!timing) return;
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[upload] %-22s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
        t_last = now;
    };

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
    apply_l2_policy(ctx, off);
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
    lap("materials, lut, copy");
    return B200RT_OK;
}

int ensure_buffers(b200rt_ctx *ctx, int w, int h, int n_slabs) {
    size_t n = (size_t)w * h * 3;
    if (w != ctx->acc_w || h != ctx->acc_h) {
        cudaFree(ctx->total); cudaFree(ctx->rgb8); cudaFree(ctx->ids); cudaFree(ctx->slabs);
        ctx->total = nullptr; ctx->rgb8 = nullptr; ctx->ids = nullptr; ctx->slabs = nullptr; ctx->n_slabs = 0;
        CK(cudaMalloc(&ctx->total, n * sizeof(float)));
        CK(cudaMalloc(&ctx->rgb8, n));
        CK(cudaMalloc(&ctx->ids, (size_t)w * h * sizeof(int32_t)));
        CK(cudaMemsetAsync(ctx->total, 0, n * sizeof(float), ctx->stream));
        ctx->acc_w = w; ctx->acc_h = h;
    }
    if (n_slabs > ctx->n_slabs) {
        cudaFree(ctx->slabs); ctx->slabs = nullptr;
        CK(cudaMalloc(&ctx->slabs, n * sizeof(float) * (size_t)n_slabs));
        ctx->n_slabs = n_slabs;
    }
    return B200RT_OK;
}

void free_wave(b200rt_ctx *c) {
    for (WaveBuffers &w : c->wave) {
        cudaFree(w.state); if (w.ctr) cudaFree(w.ctr - 4);
        for (int i = 0; i < 3; i++) cudaFree(w.queue[i]);
        w = WaveBuffers{};
    }
    c->wave_capacity = 0;
}

#define B200RT_MAX_DEPTH 64
int ensure_wave(b200rt_ctx *ctx, size_t slots) {
    if (slots <= ctx->wave_capacity) return B200RT_OK;
    free_wave(ctx);
    for (int k = 0; k < ctx->wf_pipe; k++) {
        WaveBuffers &w = ctx->wave[k];
        // one allocation: 128-byte records (AoS), or six arrays back to back (B200RT_STATE_AOS=0)
        CK(cudaMalloc(&w.state, slots * sizeof(float4) * 8));
        StateField *f[6] = {&w.ray_o, &w.ray_d, &w.hit, &w.thr, &w.rad, &w.bsdf};
        for (int k = 0; k < 6; k++) f[k]->base = w.state + (B200RT_STATE_AOS ? (size_t)k : (size_t)k * slots);
        for (int i = 0; i < 3; i++) CK(cudaMalloc(&w.queue[i], slots * sizeof(int)));
        CK(cudaMalloc(&w.ctr, 4 * (B200RT_MAX_DEPTH + 2) * sizeof(int)));
        w.ctr += 4;      // ctr[-1] = rays entering bounce 0 (layout: wavefront.cuh, WaveBuffers::ctr)
    }
    ctx->wave_capacity = slots;
    return B200RT_OK;
}

// Wavefront render of sample indices [spp_begin, spp_end): waves of (all pixels) x S_w samples.
template <bool HW8, bool COUNT>
int render_wavefront(b200rt_ctx *ctx, const DRenderParams &rp, int *launches) {
    const int W = rp.cam.width, H = rp.cam.height;
    WaveParams wp{};
    wp.cam = rp.cam;
    wp.tiles_x = (W + 7) / 8;
    wp.n_tiles = wp.tiles_x * ((H + 3) / 4);
    wp.seed = rp.seed;
    const long long per = (long long)wp.n_tiles * 32;
    const int total = rp.spp_end - rp.spp_begin;
    long long s_w = std::max<long long>(1, ctx->wf_wave_paths / per);    // aim at >= wf_wave_paths paths per wave
    s_w = std::min<long long>({s_w, (long long)total, std::max<long long>(1, (32ll << 20) / per)});
    // enough waves to keep the pipes busy, as long as a wave stays above 2 M paths
    if (ctx->wf_pipe > 1 && (total + s_w - 1) / s_w < ctx->wf_pipe) {
        const long long split = (total + ctx->wf_pipe - 1) / ctx->wf_pipe;
        if (split * per >= (2ll << 20)) s_w = split;
    }
    const int n_waves = (int)((total + s_w - 1) / s_w);
    const int n_pipe = std::max(1, std::min(ctx->wf_pipe, n_waves));
    int rc = ensure_wave(ctx, (size_t)(per * s_w));
    if (rc) return rc;
    if (rp.ray_depth > B200RT_MAX_DEPTH) return fail(ctx, B200RT_ERR_UNSUPPORTED, "ray depth above 64");
    if (!ctx->blocks_extend[COUNT]) {
        int dev = ctx->device, nb = 0;
        CK(cudaDeviceGetAttribute(&ctx->num_sms, cudaDevAttrMultiProcessorCount, dev));
        // L1 / shared-memory split (percent of shared memory): development knobs, default = the driver's choice.
        // Asking for the largest L1 outright is SLOWER here (shade +54 %, extend +5 %: profiles/r01_sweep_carveout.log).
        if (const char *e = getenv("B200RT_CARVEOUT_TRAVERSE")) {
            cudaFuncSetAttribute(wf_extend<COUNT>, cudaFuncAttributePreferredSharedMemoryCarveout, atoi(e));
            cudaFuncSetAttribute(wf_light<COUNT>, cudaFuncAttributePreferredSharedMemoryCarveout, atoi(e));
        }
        if (const char *e = getenv("B200RT_CARVEOUT_SHADE")) {
            cudaFuncSetAttribute(wf_shade<HW8, COUNT, false>, cudaFuncAttributePreferredSharedMemoryCarveout, atoi(e));
            cudaFuncSetAttribute(wf_shade<HW8, COUNT, true>, cudaFuncAttributePreferredSharedMemoryCarveout, atoi(e));
        }
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, wf_extend<COUNT>, 128, 0));
        if (ctx->wf_blocks_cap) nb = std::min(nb, ctx->wf_blocks_cap);
        ctx->blocks_extend[COUNT] = std::max(1, nb) * ctx->num_sms;
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, wf_light<COUNT>, 128, 0));
        if (ctx->wf_blocks_cap) nb = std::min(nb, ctx->wf_blocks_cap);
        ctx->blocks_light[COUNT] = std::max(1, nb) * ctx->num_sms;
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, wf_trace<COUNT>, 128, 0));
        if (ctx->wf_blocks_cap) nb = std::min(nb, ctx->wf_blocks_cap);
        ctx->blocks_trace[COUNT] = std::max(1, nb) * ctx->num_sms;
    }
    const bool lights = ctx->gs.n_light_tris > 0;
    // a light-queue entry carries the light-cut mask above bit 26: waves of 2^26 slots or more (frames above 67 M
    // pixels) leave it out and the light stage tests the cut boxes itself
    const int pack_mask = per * s_w < (1ll << B200RT_WF_MASK_SHIFT) ? 1 : 0;
    const int flat_blocks = ctx->num_sms * ctx->wf_flat_per_sm;
    ctx->pipe_stream[0] = ctx->stream;
    if (n_pipe > 1) {      // fork: the side streams start after everything already queued on the main stream
        CK(cudaEventRecord(ctx->ev_fork, ctx->stream));
        for (int k = 1; k < n_pipe; k++) CK(cudaStreamWaitEvent(ctx->pipe_stream[k], ctx->ev_fork, 0));
    }
    int wave = 0;
    for (int s0 = rp.spp_begin; s0 < rp.spp_end; s0 += (int)s_w, wave++) {
        const int k = wave % n_pipe;
        const WaveBuffers &wb = ctx->wave[k];
        cudaStream_t st = ctx->pipe_stream[k];
        wp.s_begin = s0;
        wp.s_count = std::min<int>((int)s_w, rp.spp_end - s0);
        CK(cudaMemsetAsync(wb.ctr - 4, 0, 4 * (B200RT_MAX_DEPTH + 2) * sizeof(int), st));
        // camera paths carry implicit state through their first extend / shade unless the path is one bounce long
        // or the fused scheduling (which reads the whole record) is on
        const int first_implicit = (rp.ray_depth > 1 && !(ctx->wf_fused && lights) && ctx->wf_first_implicit) ? 1 : 0;
        wf_generate<<<flat_blocks, 128, 0, st>>>(wp, wb, ctx->counters, W % 8 == 0 && H % 4 == 0, !first_implicit);
        (*launches)++;
        int *qa = wb.queue[0], *qb = wb.queue[2];
        for (int b = 0; b < rp.ray_depth; b++) {
#ifdef B200RT_DBG_SORT
            if (b > 0 && getenv("B200RT_DBG_SORT_MODE")) {
                static uint32_t *keys = nullptr;
                if (!keys) CK(cudaMalloc(&keys, ctx->wave_capacity * sizeof(uint32_t)));
                int n_rays = 0;
                CK(cudaMemcpyAsync(&n_rays, wb.ctr + 4 * b - 1, sizeof(int), cudaMemcpyDeviceToHost, st));
                CK(cudaStreamSynchronize(st));
                if (n_rays > 0) {
                    const char *e = getenv("B200RT_DBG_SORT_BITS");
                    dbg_sort_keys<<<(n_rays + 255) / 256, 256, 0, st>>>(ctx->gs, wb, qa, n_rays, keys, atoi(getenv("B200RT_DBG_SORT_MODE")), e ? atoi(e) : 5);
                    thrust::sort_by_key(thrust::cuda::par.on(st), thrust::device_pointer_cast(keys), thrust::device_pointer_cast(keys) + n_rays, thrust::device_pointer_cast(qa));
                }
            }
#endif
            int *c = wb.ctr + 4 * b;     // c[-1] rays in, c[0] extend cursor, c[1] light cursor, c[2] light queries, c[3] rays into b+1
            if (ctx->wf_fused && lights) {
                // fused scheduling: the traversal of bounce b also answers the light-pdf query of
                // bounce b-1 for the same ray and applies its weight / clamp (wavefront_trace.cuh)
                ctx->stamp(true, 0, st);
                wf_trace<COUNT><<<ctx->blocks_trace[COUNT], 128, 0, st>>>(ctx->gs, wb, qa, c - 1, c + 0, ctx->counters, ctx->wf_batch, ctx->wf_leaf_t, b > 0 ? 1 : 0);
                ctx->stamp(false, 0, st);
                ctx->stamp(true, 1, st);
                wf_shade<HW8, COUNT><<<flat_blocks, 128, 0, st>>>(ctx->gs, wb, wp, qa, c - 1, qb, c + 3, nullptr, nullptr, b, ctx->lut_srgb, ctx->counters);
                ctx->stamp(false, 1, st);
                (*launches) += 2;
                std::swap(qa, qb);
                continue;
            }
            ctx->stamp(true, 0, st);
            const int first = (b == 0 && first_implicit) ? 1 : 0;
            wf_extend<COUNT><<<ctx->blocks_extend[COUNT], 128, 0, st>>>(ctx->gs, wb, qa, c - 1, c + 0, ctx->counters, ctx->wf_batch, ctx->wf_leaf_t, first, wp.cam.pos[0], wp.cam.pos[1], wp.cam.pos[2]);
            ctx->stamp(false, 0, st);
            const bool last = b == rp.ray_depth - 1;
            // the bounce after this one is the last and a miss adds nothing: rays that cannot reach an emitter are dropped
            const int drop = (ctx->wf_drop_dark && b == rp.ray_depth - 2 && !ctx->gs.env &&
                              ctx->gs.bg[0] == 0.f && ctx->gs.bg[1] == 0.f && ctx->gs.bg[2] == 0.f) ? 1 : 0;
            if (lights && !last) {
                ctx->stamp(true, 1, st);
                wf_shade<HW8, COUNT><<<flat_blocks, 128, 0, st>>>(ctx->gs, wb, wp, qa, c - 1, wb.queue[1], c + 2, ctx->wf_direct ? qb : nullptr, c + 3, b, ctx->lut_srgb, ctx->counters, drop, first, pack_mask);
                ctx->stamp(false, 1, st);
                ctx->stamp(true, 2, st);
                wf_light<COUNT><<<ctx->blocks_light[COUNT], 128, 0, st>>>(ctx->gs, wb, wb.queue[1], c + 2, c + 1, qb, c + 3, ctx->counters, ctx->wf_batch_light, ctx->wf_leaf_t_light, ctx->wf_direct && pack_mask, drop);
                ctx->stamp(false, 2, st);
                (*launches) += 3;
            } else {
                // last bounce: the next ray is never traced (recLimit == 0), so its pdf is not needed
                ctx->stamp(true, 1, st);
                if (last) wf_shade<HW8, COUNT, true><<<flat_blocks, 128, 0, st>>>(ctx->gs, wb, wp, qa, c - 1, qb, lights ? c + 2 : c + 3, nullptr, nullptr, b, ctx->lut_srgb, ctx->counters, 0, first);
                else wf_shade<HW8, COUNT><<<flat_blocks, 128, 0, st>>>(ctx->gs, wb, wp, qa, c - 1, qb, lights ? c + 2 : c + 3, nullptr, nullptr, b, ctx->lut_srgb, ctx->counters, 0, first);
                ctx->stamp(false, 1, st);
                (*launches) += 2;
            }
            std::swap(qa, qb);
        }
        // total[] is summed wave after wave in sample order, whatever stream a wave ran on
        if (n_pipe > 1 && wave > 0) CK(cudaStreamWaitEvent(st, ctx->ev_acc[(wave - 1) % n_pipe], 0));
        wf_accumulate<<<(unsigned)((per + 255) / 256), 256, 0, st>>>(wp, wb, ctx->total);
        if (n_pipe > 1) CK(cudaEventRecord(ctx->ev_acc[k], st));
        (*launches)++;
        CK(cudaGetLastError());
    }
    for (int k = 1; k < n_pipe; k++) {     // join
        CK(cudaEventRecord(ctx->ev_join[k], ctx->pipe_stream[k]));
        CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_join[k], 0));
    }
    return B200RT_OK;
}

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

// tan of the half field of view with the libm call the reference's TU binds to (hw1/hw2 include
// <math.h> => tanf; hw3..hw8 include only <cmath> => double tan rounded to float).
void camera_of(const b200rt_scene_desc &d, int w, int h, DCamera &c) {
    for (int k = 0; k < 3; k++) { c.pos[k] = d.cam_pos[k]; c.right[k] = d.cam_right[k]; c.up[k] = d.cam_up[k]; c.forward[k] = d.cam_forward[k]; }
    float half = d.cam_fov / 2;
    float t = d.mode <= B200RT_MODE_HW2 ? tanf(half) : (float)std::tan((double)half);
    if (d.mode <= B200RT_MODE_HW5) { c.tan_x = t; c.tan_y = t * h / w; }   // fovX based (hw4/src/scene.cpp:121-122)
    else { c.tan_y = t; c.tan_x = t * w / h; }                             // fovY based (hw8/src/scene.cpp:180-181)
    c.width = w; c.height = h;
}

}  // namespace

// ================================== C ABI =====================================================
extern "C" {

int b200rt_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int b200rt_create(b200rt_ctx **out, const b200rt_options *opt) {
    if (!out) return B200RT_ERR_INVALID;
    *out = nullptr;
    int n = b200rt_device_count();
    if (n <= 0) return B200RT_ERR_NO_DEVICE;       // no CPU fallback
    int dev = opt ? opt->device : 0;
    if (dev < 0 || dev >= n) return B200RT_ERR_INVALID;
    b200rt_ctx *ctx = new (std::nothrow) b200rt_ctx();
    if (!ctx) return B200RT_ERR_INVALID;
    ctx->device = dev;
    ctx->seed = opt ? opt->seed : 0;
    ctx->flags = opt ? opt->flags : 0;
    // tuning knobs (development only; defaults are the measured optimum)
    if (const char *e = getenv("B200RT_WF_BATCH")) ctx->wf_batch = std::max(1, atoi(e));
    if (const char *e = getenv("B200RT_WF_BATCH_LIGHT")) ctx->wf_batch_light = std::max(1, atoi(e));
    if (const char *e = getenv("B200RT_WF_LEAF_T")) ctx->wf_leaf_t = std::max(1, atoi(e));
    if (const char *e = getenv("B200RT_WF_LEAF_T_LIGHT")) ctx->wf_leaf_t_light = std::max(1, atoi(e));
    if (const char *e = getenv("B200RT_STAGE_TIMING")) ctx->stage_timing = atoi(e) != 0;
    if (const char *e = getenv("B200RT_WF_WAVE_PATHS")) ctx->wf_wave_paths = std::max(1ll, atoll(e));
    if (const char *e = getenv("B200RT_WF_FUSED")) ctx->wf_fused = atoi(e) != 0;
    if (const char *e = getenv("B200RT_WF_DIRECT")) ctx->wf_direct = atoi(e) != 0;
    if (const char *e = getenv("B200RT_WF_FLAT_PER_SM")) ctx->wf_flat_per_sm = std::max(1, atoi(e));
    if (const char *e = getenv("B200RT_WF_DROP_DARK")) ctx->wf_drop_dark = atoi(e) != 0;
    if (const char *e = getenv("B200RT_WF_FIRST_IMPLICIT")) ctx->wf_first_implicit = atoi(e) != 0;
    if (const char *e = getenv("B200RT_LIGHT_TOP")) ctx->light_top_max = std::max(0, atoi(e));
    if (const char *e = getenv("B200RT_WF_PIPE")) ctx->wf_pipe = std::min(B200RT_PIPE, std::max(1, atoi(e)));
    if (ctx->flags & B200RT_FLAG_SERIAL_WAVES) ctx->wf_pipe = 1;
    if (const char *e = getenv("B200RT_WF_BLOCKS_PER_SM")) ctx->wf_blocks_cap = std::max(0, atoi(e));
    cudaError_t e;
    if ((e = cudaSetDevice(dev)) != cudaSuccess || (e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaEventCreate(&ctx->ev0)) != cudaSuccess || (e = cudaEventCreate(&ctx->ev1)) != cudaSuccess ||
        (e = cudaMalloc(&ctx->counters, sizeof(DCounters))) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming)) != cudaSuccess) {
        delete ctx;
        return B200RT_ERR_CUDA;
    }
    for (int k = 0; k < B200RT_PIPE; k++) {
        if ((k > 0 && (e = cudaStreamCreateWithFlags(&ctx->pipe_stream[k], cudaStreamNonBlocking)) != cudaSuccess) ||
            (e = cudaEventCreateWithFlags(&ctx->ev_join[k], cudaEventDisableTiming)) != cudaSuccess ||
            (e = cudaEventCreateWithFlags(&ctx->ev_acc[k], cudaEventDisableTiming)) != cudaSuccess) {
            delete ctx;
            return B200RT_ERR_CUDA;
        }
    }
    *out = ctx;
    return B200RT_OK;
}

void b200rt_destroy(b200rt_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    free_scene_memory(ctx);
    free_wave(ctx);
    cudaFree(ctx->total); cudaFree(ctx->slabs); cudaFree(ctx->rgb8); cudaFree(ctx->ids); cudaFree(ctx->counters);
    cudaEventDestroy(ctx->ev0); cudaEventDestroy(ctx->ev1); cudaEventDestroy(ctx->ev_fork);
    for (int k = 0; k < B200RT_PIPE; k++) {
        if (k > 0 && ctx->pipe_stream[k]) cudaStreamDestroy(ctx->pipe_stream[k]);
        cudaEventDestroy(ctx->ev_join[k]); cudaEventDestroy(ctx->ev_acc[k]);
    }
    for (auto e : ctx->ev_pool) cudaEventDestroy(e);
    cudaFree(ctx->hw5_light_order);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int b200rt_upload_scene(b200rt_ctx *ctx, const b200rt_scene_desc *s) {
    if (!ctx || !s) return B200RT_ERR_INVALID;
    if (s->width <= 0 || s->height <= 0 || s->samples <= 0) return fail(ctx, B200RT_ERR_INVALID, "bad frame size / sample count");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    release_scene(ctx);
    int rc = B200RT_OK;
    switch (s->mode) {
        case B200RT_MODE_HW1: case B200RT_MODE_HW2: case B200RT_MODE_HW3: case B200RT_MODE_HW4: {
            if (s->n_prims < 0 || s->n_prims > B200RT_MAX_PRIMS || s->n_lights < 0 || s->n_lights > B200RT_MAX_LIGHTS)
                return fail(ctx, B200RT_ERR_UNSUPPORTED, "too many primitives/lights for the analytic kernels");
            if ((s->n_prims && !s->prims) || (s->n_lights && !s->lights)) return fail(ctx, B200RT_ERR_INVALID, "missing primitive/light arrays");
            static DAnalyticScene a;   // 9 KB: keep off the stack; a ctx is driven by one thread
            memset(&a, 0, sizeof(a));
            for (int i = 0; i < s->n_prims; i++) {
                a.prims[i] = s->prims[i];
                if (a.prims[i].type == B200RT_PRIM_TRIANGLE) return fail(ctx, B200RT_ERR_UNSUPPORTED, "TRIANGLE primitives (hw5) not implemented");
                // hw4 light set: emissive BOX / ELLIPSOID (hw4/src/scene.cpp:12-19)
                const float *e = s->prims[i].emission;
                if ((e[0] > 0 || e[1] > 0 || e[2] > 0) && (s->prims[i].type == B200RT_PRIM_BOX || s->prims[i].type == B200RT_PRIM_ELLIPSOID))
                    a.light_prims[a.n_light_prims++] = i;
            }
            for (int i = 0; i < s->n_lights; i++) a.lights[i] = s->lights[i];
            a.n_prims = s->n_prims; a.n_lights = s->n_lights; a.mode = s->mode;
            for (int k = 0; k < 3; k++) { a.bg[k] = s->bg_color[k]; a.ambient[k] = s->ambient[k]; }
            CK(analytic_upload(a, ctx->stream));
            break;
        }
        case B200RT_MODE_HW5:
            rc = upload_hw5(ctx, s);
            break;
        case B200RT_MODE_HW6: case B200RT_MODE_HW7: case B200RT_MODE_HW8:
            rc = upload_gltf(ctx, s);
            break;
        default:
            return fail(ctx, B200RT_ERR_UNSUPPORTED, "unknown scene mode");
    }
    if (rc) { release_scene(ctx); return rc; }
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->desc = *s;
    ctx->desc.prims = nullptr; ctx->desc.lights = nullptr; ctx->desc.tris = nullptr; ctx->desc.nodes = nullptr;
    ctx->desc.light_tris = nullptr; ctx->desc.light_nodes = nullptr; ctx->desc.light_prims = nullptr; ctx->desc.materials = nullptr; ctx->desc.images = nullptr; ctx->desc.env_map = nullptr;
    ctx->has_scene = true;
    return b200rt_clear(ctx);
}

int b200rt_clear(b200rt_ctx *ctx) {
    if (!ctx) return B200RT_ERR_INVALID;
    CK(cudaSetDevice(ctx->device));
    if (ctx->total) CK(cudaMemsetAsync(ctx->total, 0, (size_t)ctx->acc_w * ctx->acc_h * 3 * sizeof(float), ctx->stream));
    return B200RT_OK;
}

int b200rt_render(b200rt_ctx *ctx, int32_t width, int32_t height, int32_t spp_begin, int32_t spp_end, int32_t ray_depth) {
    if (!ctx) return B200RT_ERR_INVALID;
    if (!ctx->has_scene) return fail(ctx, B200RT_ERR_NO_SCENE, "render before upload_scene");
    const b200rt_scene_desc &d = ctx->desc;
    if (width <= 0) width = d.width;
    if (height <= 0) height = d.height;
    if (ray_depth <= 0) ray_depth = d.ray_depth;
    const bool deterministic = d.mode == B200RT_MODE_HW1 || d.mode == B200RT_MODE_HW2;
    if (deterministic) { spp_begin = 0; spp_end = 1; }
    if (spp_begin < 0 || spp_end <= spp_begin) return fail(ctx, B200RT_ERR_INVALID, "empty sample range");
    if ((long long)width * height > (1ll << 30)) return fail(ctx, B200RT_ERR_INVALID, "frame too large");
    CK(cudaSetDevice(ctx->device));

    // enough warps to fill 148 SMs several times over even for small frames
    const long long n_px = (long long)width * height;
    int n_chunks = 1;
    if (!deterministic) {
        const long long want = 148ll * 2048 * 4;
        n_chunks = (int)std::min<long long>({(want + n_px - 1) / n_px, (long long)(spp_end - spp_begin), 256ll});
        n_chunks = std::max(n_chunks, 1);
    }
    int rc = ensure_buffers(ctx, width, height, n_chunks);
    if (rc) return rc;

    DRenderParams rp{};
    camera_of(d, width, height, rp.cam);
    rp.spp_begin = spp_begin; rp.spp_end = spp_end; rp.n_chunks = n_chunks; rp.ray_depth = ray_depth;
    rp.seed = ctx->seed; rp.accum = ctx->slabs; rp.counters = ctx->counters;

    CK(cudaMemsetAsync(ctx->counters, 0, sizeof(DCounters), ctx->stream));
    ctx->ev_used = 0;
    ctx->ev_stage.clear();
    CK(cudaEventRecord(ctx->ev0, ctx->stream));
    int launches = 0;
    if (d.mode == B200RT_MODE_HW1) { CK(launch_cast(rp.cam, ctx->slabs, ctx->ids, ctx->stream)); launches++; }
    else if (d.mode == B200RT_MODE_HW2) { CK(launch_whitted(rp.cam, ray_depth, ctx->slabs, ctx->ids, ctx->stream)); launches++; }
    else if (d.mode == B200RT_MODE_HW3 || d.mode == B200RT_MODE_HW4) { CK(launch_pt_analytic(rp, d.mode == B200RT_MODE_HW3, ctx->stream)); launches++; }
    else if (d.mode == B200RT_MODE_HW5) {
        CK(launch_hw5(ctx->gs, rp, ctx->hw5_light_order, ctx->hw5_n_lights, ctx->hw5_has_tris, ctx->stream));
        launches++;
    }
    else if (d.mode == B200RT_MODE_HW6) {
        const long long warps = (long long)((width + 7) / 8) * ((height + 3) / 4) * n_chunks;
        const long long blocks = (warps * 32 + 127) / 128;
        if (blocks > 0x7fffffffll) return fail(ctx, B200RT_ERR_INVALID, "grid too large");
        if (ctx->flags & B200RT_FLAG_COUNTERS) hw6_path_kernel<true><<<(unsigned)blocks, 128, 0, ctx->stream>>>(ctx->gs, rp);
        else hw6_path_kernel<false><<<(unsigned)blocks, 128, 0, ctx->stream>>>(ctx->gs, rp);
        CK(cudaGetLastError());
        launches++;
    }
    else if (!(ctx->flags & B200RT_FLAG_MEGAKERNEL)) {
        const bool count = ctx->flags & B200RT_FLAG_COUNTERS;
        int wrc;
        if (ctx->gs.hw8) wrc = count ? render_wavefront<true, true>(ctx, rp, &launches) : render_wavefront<true, false>(ctx, rp, &launches);
        else wrc = count ? render_wavefront<false, true>(ctx, rp, &launches) : render_wavefront<false, false>(ctx, rp, &launches);
        if (wrc) return wrc;
        n_chunks = 0;   // the wavefront accumulates into `total` itself
    }
    else {
        const int tiles = ((width + 7) / 8) * ((height + 3) / 4);
        const long long warps = (long long)tiles * n_chunks;
        const int block = 128;
        const long long blocks = (warps * 32 + block - 1) / block;
        if (blocks > 0x7fffffffll) return fail(ctx, B200RT_ERR_INVALID, "grid too large");
        const bool count = ctx->flags & B200RT_FLAG_COUNTERS;
        dim3 g((unsigned)blocks), b(block);
        if (ctx->gs.hw8) {
            if (count) gltf_path_kernel<true, true><<<g, b, 0, ctx->stream>>>(ctx->gs, rp, ctx->lut_srgb);
            else gltf_path_kernel<true, false><<<g, b, 0, ctx->stream>>>(ctx->gs, rp, ctx->lut_srgb);
        } else {
            if (count) gltf_path_kernel<false, true><<<g, b, 0, ctx->stream>>>(ctx->gs, rp, ctx->lut_srgb);
            else gltf_path_kernel<false, false><<<g, b, 0, ctx->stream>>>(ctx->gs, rp, ctx->lut_srgb);
        }
        CK(cudaGetLastError());
        launches++;
    }
    CK(cudaEventRecord(ctx->ev1, ctx->stream));
    if (n_chunks > 0) {
        const size_t n = (size_t)n_px * 3;
        fold_slabs_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->total, ctx->slabs, n_chunks, n);
        CK(cudaGetLastError());
        launches++;
    }
    ctx->stats.launches = launches;
    ctx->timing_pending = true;
    return B200RT_OK;
}

int b200rt_synchronize(b200rt_ctx *ctx) {
    if (!ctx) return B200RT_ERR_INVALID;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    return B200RT_OK;
}

int b200rt_resolve(b200rt_ctx *ctx, int32_t total_spp, uint8_t *rgb_out) {
    if (!ctx || !rgb_out) return B200RT_ERR_INVALID;
    if (!ctx->has_scene || !ctx->total) return fail(ctx, B200RT_ERR_NO_SCENE, "resolve before render");
    if (total_spp <= 0) return fail(ctx, B200RT_ERR_INVALID, "total_spp must be positive");
    CK(cudaSetDevice(ctx->device));
    const size_t n_px = (size_t)ctx->acc_w * ctx->acc_h;
    const bool hw1 = ctx->desc.mode == B200RT_MODE_HW1;
    const float inv = (ctx->desc.mode <= B200RT_MODE_HW2) ? 1.f : (float)(1.0 / total_spp);   // `1.0 / samples * color`
    resolve_kernel<<<(unsigned)((n_px + 255) / 256), 256, 0, ctx->stream>>>(ctx->total, ctx->rgb8, n_px, inv, hw1 ? 0 : 1);
    CK(cudaGetLastError());
    ctx->stats.launches += 1;
    CK(cudaMemcpyAsync(rgb_out, ctx->rgb8, n_px * 3, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return B200RT_OK;
}

int b200rt_download_accum(b200rt_ctx *ctx, float *rgb_out) {
    if (!ctx || !rgb_out) return B200RT_ERR_INVALID;
    if (!ctx->total) return fail(ctx, B200RT_ERR_NO_SCENE, "no accumulation buffer yet");
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpyAsync(rgb_out, ctx->total, (size_t)ctx->acc_w * ctx->acc_h * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return B200RT_OK;
}

int b200rt_upload_accum(b200rt_ctx *ctx, const float *rgb_in) {
    if (!ctx || !rgb_in) return B200RT_ERR_INVALID;
    if (!ctx->total) return fail(ctx, B200RT_ERR_NO_SCENE, "no accumulation buffer yet");
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpyAsync(ctx->total, rgb_in, (size_t)ctx->acc_w * ctx->acc_h * 3 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return B200RT_OK;
}

int b200rt_accum_device_ptr(b200rt_ctx *ctx, void **ptr, size_t *n_floats) {
    if (!ctx || !ptr || !n_floats) return B200RT_ERR_INVALID;
    if (!ctx->total) return fail(ctx, B200RT_ERR_NO_SCENE, "no accumulation buffer yet");
    *ptr = ctx->total;
    *n_floats = (size_t)ctx->acc_w * ctx->acc_h * 3;
    return B200RT_OK;
}

// ---- scene replication (include/b200rt.h) ----------------------------------------------------------------
namespace {
struct SceneImage {
    uint64_t magic, base, bytes, lut;
    int32_t n_materials, pad;
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
    ctx->gs = g;
    ctx->lut_srgb = reinterpret_cast<float *>(reinterpret_cast<char *>((uintptr_t)im.lut) + delta);
    ctx->arena_used = (size_t)im.bytes; ctx->n_materials = im.n_materials;
    ctx->hw5_has_tris = false; ctx->hw5_n_lights = 0;
    apply_l2_policy(ctx, (size_t)im.bytes);
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

int b200rt_primary_ids(b200rt_ctx *ctx, int32_t width, int32_t height, int32_t *ids_out) {
    if (!ctx || !ids_out) return B200RT_ERR_INVALID;
    if (!ctx->has_scene) return fail(ctx, B200RT_ERR_NO_SCENE, "primary_ids before upload_scene");
    if (ctx->desc.mode != B200RT_MODE_HW1 && ctx->desc.mode != B200RT_MODE_HW2) return fail(ctx, B200RT_ERR_UNSUPPORTED, "primary ids exist for hw1/hw2 only");
    if (!ctx->ids || (width > 0 && width != ctx->acc_w) || (height > 0 && height != ctx->acc_h)) return fail(ctx, B200RT_ERR_INVALID, "render the frame first");
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpyAsync(ids_out, ctx->ids, (size_t)ctx->acc_w * ctx->acc_h * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return B200RT_OK;
}

int b200rt_get_stats(b200rt_ctx *ctx, b200rt_stats *out) {
    if (!ctx || !out) return B200RT_ERR_INVALID;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    if (ctx->timing_pending) {
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
        DCounters c;
        CK(cudaMemcpy(&c, ctx->counters, sizeof(c), cudaMemcpyDeviceToHost));
        int launches = ctx->stats.launches;
        ctx->stats = b200rt_stats{};
        ctx->stats.launches = launches;
        ctx->stats.kernel_ms = ms;
        ctx->stats.paths = c.paths; ctx->stats.rays = c.rays; ctx->stats.light_queries = c.light_queries;
        ctx->stats.node_tests = c.node_tests; ctx->stats.tri_tests = c.tri_tests;
        ctx->stats.light_node_tests = c.light_node_tests; ctx->stats.light_tri_tests = c.light_tri_tests;
        ctx->stats.shaded_hits = c.shaded_hits; ctx->stats.tex_lookups = c.tex_lookups;
        for (size_t i = 0; i + 1 < ctx->ev_used; i += 2) {
            float t = 0;
            CK(cudaEventElapsedTime(&t, ctx->ev_pool[i], ctx->ev_pool[i + 1]));
            const int stage = ctx->ev_stage[i / 2];
            if (stage == 0) { ctx->stats.extend_ms += t; ctx->stats.extend_launches++; }
            else if (stage == 1) ctx->stats.shade_ms += t;
            else ctx->stats.light_ms += t;
        }
        ctx->timing_pending = false;
        if (c.stack_overflow) return fail(ctx, B200RT_ERR_STACK, "traversal stack overflow during render");
#ifdef B200RT_DBG_PHASE
        {
            unsigned long long ph[8];
            if (cudaMemcpyFromSymbol(ph, g_phase, sizeof(ph)) == cudaSuccess) {
                fprintf(stderr, "[phase] inner steps %llu lanes %.2f (waiting at leaf %.2f, idle %.2f) | leaf steps %llu lanes %.2f | refills %llu\n",
                        ph[0], ph[0] ? (double)ph[1] / ph[0] : 0., ph[0] ? (double)ph[6] / ph[0] : 0., ph[0] ? (double)ph[7] / ph[0] : 0.,
                        ph[2], ph[2] ? (double)ph[3] / ph[2] : 0., ph[5]);
                memset(ph, 0, sizeof(ph));
                cudaMemcpyToSymbol(g_phase, ph, sizeof(ph));
            }
        }
#endif
    }
    *out = ctx->stats;
    return B200RT_OK;
}

int b200rt_stream(b200rt_ctx *ctx, void **stream_out) {
    if (!ctx || !stream_out) return B200RT_ERR_INVALID;
    *stream_out = (void *)ctx->stream;
    return B200RT_OK;
}

const char *b200rt_strerror(int code) {
    switch (code) {
        case B200RT_OK: return "ok";
        case B200RT_ERR_INVALID: return "invalid argument";
        case B200RT_ERR_NO_DEVICE: return "no CUDA device (this library has no CPU fallback)";
        case B200RT_ERR_CUDA: return "CUDA runtime error";
        case B200RT_ERR_NO_SCENE: return "no scene uploaded / nothing rendered";
        case B200RT_ERR_UNSUPPORTED: return "not implemented on the device path";
        case B200RT_ERR_STACK: return "BVH deeper than the traversal stack";
        default: return "unknown error";
    }
}

const char *b200rt_last_error(b200rt_ctx *ctx) { return ctx ? ctx->last_error.c_str() : ""; }

}  // extern "C"
