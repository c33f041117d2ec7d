// Internal header of libb200rt.so: the context behind the opaque b200rt_ctx handle and the entry points the
// translation units share (upload.cu: scene flattening + arena; scheduler.cu: launches; b200rt.cu: the C ABI;
// multi_gpu.cu: groups, peer reduce).
#pragma once
#include <cuda_runtime.h>
#include <new>
#include <string>
#include <vector>
#include "b200rt.h"
#include "device_scene.h"
#include "wave_state.h"
#include "bvh_rebuild.h"

using namespace b200rt;

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
    uint64_t n_pairs = 0, n_dtris = 0;        // records of the scene BVH / triangle records in the arena (b200rt_scene_bvh_*)
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
    int blocks_extend[2] = {0, 0}, blocks_light[2] = {0, 0};
    long long wf_wave_paths = 4ll << 20;
    int light_top_max = B200RT_LIGHT_TOP;  // boxes of the light-BVH cut used as the light-pdf pre-filter (0/1 = root box only)
    int wf_flat_per_sm = 16;         // blocks per SM of the grid-stride kernels (generate, shade); B200RT_WF_FLAT_PER_SM
    bool wf_first_implicit = true;   // camera paths: only the direction is stored by generate (B200RT_WF_FIRST_IMPLICIT=0: whole record)
    bool wf_drop_dark = true;    // paths whose last bounce cannot reach an emitter end one bounce early (black background only)
    bool wf_direct = true;       // shade finishes bounces whose ray misses the light root box (no light-stage entry)
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

static inline int fail(b200rt_ctx *c, int code, const char *what, cudaError_t e = cudaSuccess) {
    if (c) {
        c->last_error = what;
        if (e != cudaSuccess) { c->last_error += ": "; c->last_error += cudaGetErrorString(e); }
    }
    return code;
}
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(ctx, B200RT_ERR_CUDA, #call, e_); } while (0)
// The C ABI never throws (include/b200rt.h): every extern "C" body that can allocate runs inside this pair.
#define B200RT_TRY try {
#define B200RT_CATCH(ctx_) } catch (const std::bad_alloc &) { return fail(ctx_, B200RT_ERR_INVALID, "out of host memory"); } \
                           catch (...) { return fail(ctx_, B200RT_ERR_INVALID, "unexpected C++ exception"); }

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// upload.cu
int upload_gltf(b200rt_ctx *ctx, const b200rt_scene_desc *s);
int upload_hw5(b200rt_ctx *ctx, const b200rt_scene_desc *s);
int ensure_arena(b200rt_ctx *ctx, size_t bytes);
void release_scene(b200rt_ctx *c);
void free_scene_memory(b200rt_ctx *c);
// scheduler.cu
void free_wave(b200rt_ctx *c);
int ensure_buffers(b200rt_ctx *ctx, int w, int h, int n_slabs);
int ensure_wave(b200rt_ctx *ctx, size_t slots);
