// Scheduler: frame buffers, wave buffers, the wavefront pipeline's launches (two waves in flight), the per-mode
// dispatch behind b200rt_render and the statistics read-back.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include "analytic_kernels.h"
#include "context.h"
#include "gltf_kernel.cuh"
#include "wavefront.cuh"
#include "hw6_kernel.cuh"

// Adds the per-chunk slabs to the running total in a fixed order (deterministic sum).
__global__ void fold_slabs_kernel(float *__restrict__ total, const float *__restrict__ slabs, int n_slabs, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float acc = total[i];
    for (int s = 0; s < n_slabs; s++) acc += slabs[(size_t)s * n + i];
    total[i] = acc;
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
template <bool HW8, bool COUNT, bool HW6 = false>
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
        // (the L1 / shared-memory carve-out stays the driver's choice: asking for the largest L1 made shade 54 % and
        // extend 5 % slower, profiles/r01_sweep_carveout.log)
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, wf_extend<COUNT>, 128, 0));
        if (ctx->wf_blocks_cap) nb = std::min(nb, ctx->wf_blocks_cap);
        ctx->blocks_extend[COUNT] = std::max(1, nb) * ctx->num_sms;
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, wf_light<COUNT>, 128, 0));
        if (ctx->wf_blocks_cap) nb = std::min(nb, ctx->wf_blocks_cap);
        ctx->blocks_light[COUNT] = std::max(1, nb) * ctx->num_sms;
    }
    const bool lights = ctx->gs.n_light_tris > 0;
    // paths that need no light-pdf query skip the light stage; hw6's mirror / dielectric bounces never have one, so
    // there the shortcut is not optional
    const bool direct_on = ctx->wf_direct || HW6;
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
        const int first_implicit = (rp.ray_depth > 1 && ctx->wf_first_implicit) ? 1 : 0;
        wf_generate<<<flat_blocks, 128, 0, st>>>(wp, wb, ctx->counters, W % 8 == 0 && H % 4 == 0, !first_implicit, !HW6);
        (*launches)++;
        int *qa = wb.queue[0], *qb = wb.queue[2];
        for (int b = 0; b < rp.ray_depth; b++) {
            int *c = wb.ctr + 4 * b;     // c[-1] rays in, c[0] extend cursor, c[1] light cursor, c[2] light queries, c[3] rays into b+1
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
                wf_shade<HW8, COUNT, false, HW6><<<flat_blocks, 128, 0, st>>>(ctx->gs, wb, wp, qa, c - 1, wb.queue[1], c + 2, direct_on ? qb : nullptr, c + 3, b, ctx->lut_srgb, ctx->counters, drop, first, pack_mask);
                ctx->stamp(false, 1, st);
                ctx->stamp(true, 2, st);
                wf_light<COUNT><<<ctx->blocks_light[COUNT], 128, 0, st>>>(ctx->gs, wb, wb.queue[1], c + 2, c + 1, qb, c + 3, ctx->counters, ctx->wf_batch_light, ctx->wf_leaf_t_light, direct_on && pack_mask, drop);
                ctx->stamp(false, 2, st);
                (*launches) += 3;
            } else {
                // last bounce: the next ray is never traced (recLimit == 0), so its pdf is not needed
                ctx->stamp(true, 1, st);
                if (last) wf_shade<HW8, COUNT, true, HW6><<<flat_blocks, 128, 0, st>>>(ctx->gs, wb, wp, qa, c - 1, qb, lights ? c + 2 : c + 3, nullptr, nullptr, b, ctx->lut_srgb, ctx->counters, 0, first);
                else wf_shade<HW8, COUNT, false, HW6><<<flat_blocks, 128, 0, st>>>(ctx->gs, wb, wp, qa, c - 1, qb, lights ? c + 2 : c + 3, nullptr, nullptr, b, ctx->lut_srgb, ctx->counters, 0, first);
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

namespace {
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

extern "C" {

static int render_impl(b200rt_ctx *ctx, int32_t width, int32_t height, int32_t spp_begin, int32_t spp_end, int32_t ray_depth);
int b200rt_render(b200rt_ctx *ctx, int32_t width, int32_t height, int32_t spp_begin, int32_t spp_end, int32_t ray_depth) {
    B200RT_TRY                 // (the per-stage event pool is a std::vector)
    return render_impl(ctx, width, height, spp_begin, spp_end, ray_depth);
    B200RT_CATCH(ctx)
}
static int render_impl(b200rt_ctx *ctx, int32_t width, int32_t height, int32_t spp_begin, int32_t spp_end, int32_t ray_depth) {
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
    else if (d.mode == B200RT_MODE_HW6 && !(ctx->flags & B200RT_FLAG_MEGAKERNEL) &&
             ((ctx->flags & B200RT_FLAG_WAVEFRONT) || ctx->n_dtris > B200RT_HW6_PIXEL_KERNEL_MAX_TRIS)) {
        // the hw6 integrator as the shade stage of the wavefront pipeline (persistent traversal kernels of hw7 / hw8):
        // practice6_1 (17 k triangles) 1.78 -> 2.68, practice6_2 (145 k) 0.55 -> 1.13 Gpaths/s; a scene of a few dozen
        // triangles is 15 % faster in the per-pixel kernel (profiles/r02_hw6_wavefront.log), hence the size rule
        const int wrc = (ctx->flags & B200RT_FLAG_COUNTERS) ? render_wavefront<false, true, true>(ctx, rp, &launches)
                                                             : render_wavefront<false, false, true>(ctx, rp, &launches);
        if (wrc) return wrc;
        n_chunks = 0;
    }
    else if (d.mode == B200RT_MODE_HW6) {      // one thread per pixel (K6): small scenes, and B200RT_FLAG_MEGAKERNEL
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

// ---- SURVEY 8(d) cross-check aid: the production closest-hit kernel on a caller-supplied ray set ----------------
namespace {
__global__ void load_rays_kernel(const float *__restrict__ rays, int n, const WaveBuffers wb) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *r = rays + (size_t)i * 6;
    wb.ray_o[i] = make_float4(r[0], r[1], r[2], 0.f);
    wb.ray_d[i] = make_float4(r[3], r[4], r[5], 0.f);
    wb.queue[0][i] = i;
    if (i == 0) wb.ctr[-1] = n;
}
__global__ void store_hits_kernel(float4 *__restrict__ out, int n, const WaveBuffers wb) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = wb.hit[i];
}
}  // namespace

int b200rt_trace_rays(b200rt_ctx *ctx, const float *rays, int32_t n, float *hits, uint64_t *node_tests, uint64_t *tri_tests) {
    if (!ctx || !rays || !hits || n <= 0) return B200RT_ERR_INVALID;
    if (!ctx->has_scene) return fail(ctx, B200RT_ERR_NO_SCENE, "trace_rays before upload_scene");
    if (ctx->desc.mode < B200RT_MODE_HW6) return fail(ctx, B200RT_ERR_UNSUPPORTED, "ray queries exist for the glTF modes only");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    int rc = ensure_wave(ctx, (size_t)n);
    if (rc) return rc;
    const WaveBuffers &wb = ctx->wave[0];
    // device staging for the rays and the hits (a test / measurement entry point, not a hot path: plain cudaMalloc)
    float *d_rays = nullptr;
    float4 *d_hits = nullptr;
    CK(cudaMalloc(&d_rays, (size_t)n * 6 * sizeof(float)));
    if (cudaMalloc(&d_hits, (size_t)n * sizeof(float4)) != cudaSuccess) { cudaFree(d_rays); return fail(ctx, B200RT_ERR_CUDA, "cudaMalloc"); }
    cudaStream_t st = ctx->stream;
    auto done = [&](int code) { cudaFree(d_rays); cudaFree(d_hits); return code; };
    if (cudaMemcpyAsync(d_rays, rays, (size_t)n * 6 * sizeof(float), cudaMemcpyHostToDevice, st) != cudaSuccess ||
        cudaMemsetAsync(ctx->counters, 0, sizeof(DCounters), st) != cudaSuccess ||
        cudaMemsetAsync(wb.ctr - 4, 0, 4 * 66 * sizeof(int), st) != cudaSuccess)
        return done(fail(ctx, B200RT_ERR_CUDA, "trace_rays staging"));
    load_rays_kernel<<<(n + 255) / 256, 256, 0, st>>>(d_rays, n, wb);
    int nb = 0;
    if (!ctx->num_sms) cudaDeviceGetAttribute(&ctx->num_sms, cudaDevAttrMultiProcessorCount, ctx->device);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, wf_extend<true>, 128, 0);
    const int blocks = std::max(1, nb) * std::max(1, ctx->num_sms);
    int *c = wb.ctr;
    wf_extend<true><<<blocks, 128, 0, st>>>(ctx->gs, wb, wb.queue[0], c - 1, c + 0, ctx->counters, ctx->wf_batch, ctx->wf_leaf_t, 0, 0.f, 0.f, 0.f);
    store_hits_kernel<<<(n + 255) / 256, 256, 0, st>>>(d_hits, n, wb);
    DCounters cnt;
    if (cudaGetLastError() != cudaSuccess ||
        cudaMemcpyAsync(hits, d_hits, (size_t)n * sizeof(float4), cudaMemcpyDeviceToHost, st) != cudaSuccess ||
        cudaMemcpyAsync(&cnt, ctx->counters, sizeof(cnt), cudaMemcpyDeviceToHost, st) != cudaSuccess ||
        cudaStreamSynchronize(st) != cudaSuccess)
        return done(fail(ctx, B200RT_ERR_CUDA, "trace_rays launch", cudaGetLastError()));
    if (node_tests) *node_tests = cnt.node_tests;
    if (tri_tests) *tri_tests = cnt.tri_tests;
    return done(B200RT_OK);
}

int b200rt_scene_bvh_layout(b200rt_ctx *ctx, b200rt_bvh_layout *out) {
    if (!ctx || !out) return B200RT_ERR_INVALID;
    if (!ctx->has_scene || ctx->desc.mode < B200RT_MODE_HW6) return fail(ctx, B200RT_ERR_NO_SCENE, "no glTF-mode scene uploaded");
    out->n_records = ctx->n_pairs; out->n_tris = ctx->n_dtris; out->root = ctx->gs.root;
    out->quant_bits = B200RT_QBITS; out->guard = B200RT_QEPS;
    for (int k = 0; k < 3; k++) { out->base[k] = ctx->gs.q.base[k]; out->scale[k] = ctx->gs.q.scale[k]; out->inv_scale[k] = ctx->gs.q.inv_scale[k]; }
    return B200RT_OK;
}

int b200rt_scene_bvh_download(b200rt_ctx *ctx, void *records, void *tris) {
    if (!ctx || !records || !tris) return B200RT_ERR_INVALID;
    if (!ctx->has_scene || ctx->desc.mode < B200RT_MODE_HW6) return fail(ctx, B200RT_ERR_NO_SCENE, "no glTF-mode scene uploaded");
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpyAsync(records, ctx->gs.nodes, ctx->n_pairs * sizeof(DPair), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(tris, ctx->gs.tris, ctx->n_dtris * sizeof(DTri), cudaMemcpyDeviceToHost, ctx->stream));
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

}  // extern "C"
