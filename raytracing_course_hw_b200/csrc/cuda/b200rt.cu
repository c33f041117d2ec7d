// C ABI of the B200 render core: context life cycle, scene upload dispatch, resolve and the small accessors.
// See include/b200rt.h for the contract and the reference call sites each entry replaces.  The other units of the
// library: upload.cu (flattening + arena + replication), scheduler.cu (launches, b200rt_render, statistics),
// multi_gpu.cu (groups, fused peer reduce + resolve), analytic_kernels.cu (hw1 - hw5 kernels, -fmad=false).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <stdexcept>
#include "analytic_kernels.h"
#include "context.h"
#include "resolve.cuh"

// K5 resolve (arithmetic: resolve.cuh), one thread per pixel.
__global__ void resolve_kernel(const float *__restrict__ total, uint8_t *__restrict__ rgb, size_t n_px, float inv_spp, int tonemap) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_px) return;
#pragma unroll
    for (int c = 0; c < 3; c++) rgb[3 * i + c] = resolve_channel(total[3 * i + c], inv_spp, tonemap);
}

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
    if (ctx->flags & B200RT_FLAG_SERIAL_WAVES) ctx->wf_pipe = 1;
    if (ctx->flags & B200RT_FLAG_KEEP_DARK_RAYS) ctx->wf_drop_dark = false;
    if (ctx->flags & B200RT_FLAG_EXPLICIT_CAMERA_STATE) ctx->wf_first_implicit = false;
#ifdef B200RT_DEV
    // tuning knobs of development builds (tools/sweep_*.sh); the product reads no environment variable: its
    // defaults are the measured optimum
    if (const char *e = getenv("B200RT_WF_BATCH")) ctx->wf_batch = std::max(1, atoi(e));
    if (const char *e = getenv("B200RT_WF_BATCH_LIGHT")) ctx->wf_batch_light = std::max(1, atoi(e));
    if (const char *e = getenv("B200RT_WF_LEAF_T")) ctx->wf_leaf_t = std::max(1, atoi(e));
    if (const char *e = getenv("B200RT_WF_LEAF_T_LIGHT")) ctx->wf_leaf_t_light = std::max(1, atoi(e));
    if (const char *e = getenv("B200RT_STAGE_TIMING")) ctx->stage_timing = atoi(e) != 0;
    if (const char *e = getenv("B200RT_WF_WAVE_PATHS")) ctx->wf_wave_paths = std::max(1ll, atoll(e));
    if (const char *e = getenv("B200RT_WF_DIRECT")) ctx->wf_direct = atoi(e) != 0;
    if (const char *e = getenv("B200RT_WF_FLAT_PER_SM")) ctx->wf_flat_per_sm = std::max(1, atoi(e));
    if (const char *e = getenv("B200RT_LIGHT_TOP")) ctx->light_top_max = std::max(0, atoi(e));
    if (const char *e = getenv("B200RT_WF_PIPE")) ctx->wf_pipe = std::min(B200RT_PIPE, std::max(1, atoi(e)));
    if (const char *e = getenv("B200RT_WF_BLOCKS_PER_SM")) ctx->wf_blocks_cap = std::max(0, atoi(e));
#endif
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

static int upload_scene_impl(b200rt_ctx *ctx, const b200rt_scene_desc *s);
// "Never throws" (include/b200rt.h): the flattener and the BVH rebuild allocate std::vectors
int b200rt_upload_scene(b200rt_ctx *ctx, const b200rt_scene_desc *s) {
    B200RT_TRY
    return upload_scene_impl(ctx, s);
    B200RT_CATCH(ctx)
}
static int upload_scene_impl(b200rt_ctx *ctx, const b200rt_scene_desc *s) {
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
        case B200RT_ERR_TOO_LARGE: return "frame above B200RT_NOALLOC_MAX_PIXELS: too large for the static scratch of b200rt_resolve_noalloc";
        default: return "unknown error";
    }
}

const char *b200rt_last_error(b200rt_ctx *ctx) { return ctx ? ctx->last_error.c_str() : ""; }

}  // extern "C"

