// Several GPUs behind the C ABI: sample split of one frame, and the one exchange step it needs --
// combining the per-GPU radiance buffers -- fused with the resolve into ONE kernel that reads the other GPUs'
// buffers through peer memory (NVLink / NVSwitch loads on mapped peer pointers).  Replaces the OpenMP pixel loop as
// the unit of parallelism (hw8/src/sceneio.cpp:386-396) and the host-side sum the CLI used in round 1.
//
// Why not ncclReduce + resolve: the reduce would write a 99.5 MB sum on rank 0 (4K) that the resolve reads right
// back; here every float crosses NVLink once, is summed in registers in rank order (deterministic) and leaves as one
// byte.  Bound: bytes over NVLink into the resolving GPU, (G - 1) * W * H * 12 B at the measured 770 GB/s per
// direction -> 0.9 ms for 8 GPUs at 4K.
#include <algorithm>
#include <cstring>
#include <new>
#include <string>
#include <vector>
#include "context.h"
#include "resolve.cuh"

namespace {

struct PeerList {
    const float *p[B200RT_MAX_PEERS];
    int n;
};

// One thread = four consecutive floats of the frame (channels are independent: resolve.cuh).  sum_out / rgb may be
// null.  Peer loads are plain 128-bit global loads on peer-mapped addresses; streaming hint: nothing is re-read.
__global__ void __launch_bounds__(256) reduce_resolve_peers_kernel(const float *__restrict__ own, const PeerList peers, float *sum_out,
                                                                    uint8_t *__restrict__ rgb, size_t n_floats, float inv_spp, int tonemap) {
    const size_t n4 = n_floats >> 2;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 acc = __ldcs(reinterpret_cast<const float4 *>(own) + i);
#pragma unroll 4
        for (int k = 0; k < peers.n; k++) {
            const float4 v = __ldcs(reinterpret_cast<const float4 *>(peers.p[k]) + i);
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        if (sum_out) reinterpret_cast<float4 *>(sum_out)[i] = acc;
        if (rgb) {
            const uint32_t w = (uint32_t)resolve_channel(acc.x, inv_spp, tonemap) | ((uint32_t)resolve_channel(acc.y, inv_spp, tonemap) << 8) |
                               ((uint32_t)resolve_channel(acc.z, inv_spp, tonemap) << 16) | ((uint32_t)resolve_channel(acc.w, inv_spp, tonemap) << 24);
            reinterpret_cast<uint32_t *>(rgb)[i] = w;
        }
    }
    // tail (frames whose float count is not a multiple of four)
    if (blockIdx.x == 0 && threadIdx.x < (n_floats & 3)) {
        const size_t i = (n4 << 2) + threadIdx.x;
        float acc = own[i];
        for (int k = 0; k < peers.n; k++) acc += peers.p[k][i];
        if (sum_out) sum_out[i] = acc;
        if (rgb) rgb[i] = resolve_channel(acc, inv_spp, tonemap);
    }
}

int launch_peers(b200rt_ctx *ctx, const void *const *peer_accum, int32_t n_peers, float *sum_out, uint8_t *rgb, float inv_spp, int tonemap) {
    if (n_peers < 0 || n_peers > B200RT_MAX_PEERS || (n_peers > 0 && !peer_accum)) return fail(ctx, B200RT_ERR_INVALID, "bad peer list");
    PeerList pl{};
    pl.n = n_peers;
    for (int k = 0; k < n_peers; k++) {
        if (!peer_accum[k]) return fail(ctx, B200RT_ERR_INVALID, "null peer buffer");
        pl.p[k] = static_cast<const float *>(peer_accum[k]);
    }
    const size_t n = (size_t)ctx->acc_w * ctx->acc_h * 3;
    int sms = ctx->num_sms;
    if (!sms) { CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device)); ctx->num_sms = sms; }
    // grid-stride over a grid that is a multiple of the SM count (8 resident blocks of 256 threads per SM)
    const unsigned blocks = (unsigned)std::max<size_t>(1, std::min<size_t>((size_t)sms * 8, ((n >> 2) + 255) / 256));
    reduce_resolve_peers_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->total, pl, sum_out, rgb, n, inv_spp, tonemap);
    CK(cudaGetLastError());
    return B200RT_OK;
}

}  // namespace

extern "C" {

int b200rt_enable_peer(b200rt_ctx *ctx, int32_t peer_device) {
    if (!ctx) return B200RT_ERR_INVALID;
    if (peer_device == ctx->device) return B200RT_OK;
    CK(cudaSetDevice(ctx->device));
    int can = 0;
    CK(cudaDeviceCanAccessPeer(&can, ctx->device, peer_device));
    if (!can) return fail(ctx, B200RT_ERR_UNSUPPORTED, "no peer access between the two devices");
    const cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return B200RT_OK; }
    if (e != cudaSuccess) return fail(ctx, B200RT_ERR_CUDA, "cudaDeviceEnablePeerAccess", e);
    return B200RT_OK;
}

int b200rt_resolve_peers(b200rt_ctx *ctx, int32_t total_spp, const void *const *peer_accum, int32_t n_peers, uint8_t *rgb_out) {
    if (!ctx || !rgb_out) return B200RT_ERR_INVALID;
    if (!ctx->has_scene || !ctx->total) return fail(ctx, B200RT_ERR_NO_SCENE, "resolve before render");
    if (total_spp <= 0) return fail(ctx, B200RT_ERR_INVALID, "total_spp must be positive");
    CK(cudaSetDevice(ctx->device));
    const bool hw1 = ctx->desc.mode == B200RT_MODE_HW1;
    const float inv = (ctx->desc.mode <= B200RT_MODE_HW2) ? 1.f : (float)(1.0 / total_spp);
    const int rc = launch_peers(ctx, peer_accum, n_peers, nullptr, ctx->rgb8, inv, hw1 ? 0 : 1);
    if (rc) return rc;
    ctx->stats.launches += 1;
    CK(cudaMemcpyAsync(rgb_out, ctx->rgb8, (size_t)ctx->acc_w * ctx->acc_h * 3, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return B200RT_OK;
}

int b200rt_reduce_peers(b200rt_ctx *ctx, const void *const *peer_accum, int32_t n_peers) {
    if (!ctx) return B200RT_ERR_INVALID;
    if (!ctx->total) return fail(ctx, B200RT_ERR_NO_SCENE, "no accumulation buffer yet");
    CK(cudaSetDevice(ctx->device));
    return launch_peers(ctx, peer_accum, n_peers, ctx->total, nullptr, 1.f, 0);
}

int b200rt_accum_ipc_export(b200rt_ctx *ctx, b200rt_ipc_handle *out) {
    if (!ctx || !out) return B200RT_ERR_INVALID;
    if (!ctx->total) return fail(ctx, B200RT_ERR_NO_SCENE, "no accumulation buffer yet (render or clear a frame first)");
    static_assert(sizeof(cudaIpcMemHandle_t) <= sizeof(b200rt_ipc_handle), "b200rt_ipc_handle too small");
    CK(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, ctx->total));
    memset(out, 0, sizeof(*out));
    memcpy(out, &h, sizeof(h));
    return B200RT_OK;
}

int b200rt_ipc_open(b200rt_ctx *ctx, const b200rt_ipc_handle *h, void **peer_ptr) {
    if (!ctx || !h || !peer_ptr) return B200RT_ERR_INVALID;
    CK(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t ch;
    memcpy(&ch, h, sizeof(ch));
    CK(cudaIpcOpenMemHandle(peer_ptr, ch, cudaIpcMemLazyEnablePeerAccess));
    return B200RT_OK;
}

int b200rt_ipc_close(b200rt_ctx *ctx, void *peer_ptr) {
    if (!ctx || !peer_ptr) return B200RT_ERR_INVALID;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaIpcCloseMemHandle(peer_ptr));
    return B200RT_OK;
}

}  // extern "C"

// ---- group level: one process drives all GPUs -----------------------------------------------------------------
struct b200rt_group {
    std::vector<b200rt_ctx *> m;
    std::vector<bool> direct;          // member k's buffer is readable from member 0's device
    float *scratch = nullptr;          // member 0's device: staging for members without peer access, frame-sum scratch
    size_t scratch_floats = 0;
    std::string last_error;
};

namespace {

int gfail(b200rt_group *g, int code, const char *what, b200rt_ctx *from = nullptr) {
    if (g) {
        g->last_error = what;
        if (from) { g->last_error += ": "; g->last_error += from->last_error; }
    }
    return code;
}

int group_scratch(b200rt_group *g, size_t floats) {
    b200rt_ctx *ctx = g->m[0];
    if (floats <= g->scratch_floats) return B200RT_OK;
    CK(cudaSetDevice(ctx->device));
    cudaFree(g->scratch); g->scratch = nullptr; g->scratch_floats = 0;
    CK(cudaMalloc(&g->scratch, floats * sizeof(float)));
    g->scratch_floats = floats;
    return B200RT_OK;
}

// Waits for every member and collects the buffers member 0 adds to its own, in member order.  Members whose device
// member 0 cannot read directly are copied (cudaMemcpyPeerAsync) into scratch slices first.
int group_peers(b200rt_group *g, std::vector<const void *> &peers) {
    b200rt_ctx *c0 = g->m[0];
    const size_t n = (size_t)c0->acc_w * c0->acc_h * 3;
    size_t staged = 0;
    for (size_t k = 1; k < g->m.size(); k++) if (!g->direct[k]) staged++;
    if (staged) { const int rc = group_scratch(g, n * (staged + 1)); if (rc) return gfail(g, rc, "scratch", c0); }
    peers.clear();
    size_t slot = 1;      // slice 0 is the frame-sum scratch of group_download_accum
    for (size_t k = 1; k < g->m.size(); k++) {
        b200rt_ctx *c = g->m[k];
        if (c->acc_w != c0->acc_w || c->acc_h != c0->acc_h || !c->total) return gfail(g, B200RT_ERR_INVALID, "members hold different frames");
        int rc = b200rt_synchronize(c);
        if (rc) return gfail(g, rc, "member synchronize", c);
        if (g->direct[k]) peers.push_back(c->total);
        else {
            b200rt_ctx *ctx = c0;
            float *dst = g->scratch + n * slot++;
            CK(cudaSetDevice(c0->device));
            CK(cudaMemcpyPeerAsync(dst, c0->device, c->total, c->device, n * sizeof(float), c0->stream));
            peers.push_back(dst);
        }
    }
    return B200RT_OK;
}

}  // namespace

extern "C" {

int b200rt_group_create(b200rt_group **out, const int32_t *devices, int32_t n_devices, const b200rt_options *opt) {
    if (!out) return B200RT_ERR_INVALID;
    *out = nullptr;
    if (!devices || n_devices < 1 || n_devices > B200RT_MAX_PEERS + 1) return B200RT_ERR_INVALID;
    b200rt_group *g = new (std::nothrow) b200rt_group();
    if (!g) return B200RT_ERR_INVALID;
    try {
        for (int i = 0; i < n_devices; i++) {
            b200rt_options o = opt ? *opt : b200rt_options{};
            o.device = devices[i];
            b200rt_ctx *c = nullptr;
            const int rc = b200rt_create(&c, &o);
            if (rc) { b200rt_group_destroy(g); return rc; }
            g->m.push_back(c);
            // member 0 reads the others' buffers in place when the devices allow it
            g->direct.push_back(i == 0 || b200rt_enable_peer(g->m[0], devices[i]) == B200RT_OK);
        }
    } catch (...) { b200rt_group_destroy(g); return B200RT_ERR_INVALID; }
    *out = g;
    return B200RT_OK;
}

void b200rt_group_destroy(b200rt_group *g) {
    if (!g) return;
    if (g->scratch && !g->m.empty()) { cudaSetDevice(g->m[0]->device); cudaFree(g->scratch); }
    for (b200rt_ctx *c : g->m) b200rt_destroy(c);
    delete g;
}

int32_t b200rt_group_size(const b200rt_group *g) { return g ? (int32_t)g->m.size() : 0; }
b200rt_ctx *b200rt_group_member(b200rt_group *g, int32_t i) { return (g && i >= 0 && i < (int32_t)g->m.size()) ? g->m[(size_t)i] : nullptr; }
const char *b200rt_group_last_error(b200rt_group *g) { return g ? g->last_error.c_str() : ""; }

int b200rt_group_upload_scene(b200rt_group *g, const b200rt_scene_desc *scene) {
    if (!g || !scene) return B200RT_ERR_INVALID;
    int rc = b200rt_upload_scene(g->m[0], scene);
    if (rc) return gfail(g, rc, "upload on member 0", g->m[0]);
    for (size_t k = 1; k < g->m.size(); k++) {
        // glTF modes: device-to-device copy of the resident scene; text scenes are a few hundred bytes per member
        rc = scene->mode >= B200RT_MODE_HW6 ? b200rt_replicate_scene(g->m[0], g->m[k]) : b200rt_upload_scene(g->m[k], scene);
        if (rc) return gfail(g, rc, "scene replication", g->m[k]);
    }
    return B200RT_OK;
}

int b200rt_group_clear(b200rt_group *g) {
    if (!g) return B200RT_ERR_INVALID;
    for (b200rt_ctx *c : g->m) { const int rc = b200rt_clear(c); if (rc) return gfail(g, rc, "clear", c); }
    return B200RT_OK;
}

int b200rt_group_render(b200rt_group *g, int32_t width, int32_t height, int32_t spp_begin, int32_t spp_end, int32_t ray_depth) {
    if (!g) return B200RT_ERR_INVALID;
    b200rt_ctx *c0 = g->m[0];
    if (!c0->has_scene) return gfail(g, B200RT_ERR_NO_SCENE, "render before upload_scene");
    if (width <= 0) width = c0->desc.width;
    if (height <= 0) height = c0->desc.height;
    const bool deterministic = c0->desc.mode <= B200RT_MODE_HW2;       // one sample per pixel: nothing to split
    if (deterministic) { spp_begin = 0; spp_end = 1; }
    if (spp_begin < 0 || spp_end <= spp_begin) return gfail(g, B200RT_ERR_INVALID, "empty sample range");
    const long long S = spp_end - spp_begin, G = (long long)g->m.size();
    for (long long k = 0; k < G; k++) {
        b200rt_ctx *ctx = g->m[(size_t)k];
        // hw1 / hw2 hold ONE sample per pixel: member 0 renders it, nothing is split
        const int32_t b = deterministic ? 0 : spp_begin + (int32_t)(S * k / G);
        const int32_t e = deterministic ? (k == 0 ? 1 : 0) : spp_begin + (int32_t)(S * (k + 1) / G);
        int rc;
        if (e > b) rc = b200rt_render(ctx, width, height, b, e, ray_depth);
        else {
            // no samples for this member: it still contributes a buffer of the frame's size (zeroed when the frame
            // size changes, by b200rt_group_clear / upload otherwise)
            CK(cudaSetDevice(ctx->device));
            rc = ensure_buffers(ctx, width, height, 0);
        }
        if (rc) return gfail(g, rc, "member render", ctx);
    }
    return B200RT_OK;
}

int b200rt_group_resolve(b200rt_group *g, int32_t total_spp, uint8_t *rgb_out) {
    if (!g || !rgb_out) return B200RT_ERR_INVALID;
    std::vector<const void *> peers;
    int rc;
    try { rc = group_peers(g, peers); } catch (...) { return gfail(g, B200RT_ERR_INVALID, "out of host memory"); }
    if (rc) return rc;
    rc = b200rt_resolve_peers(g->m[0], total_spp, peers.data(), (int32_t)peers.size(), rgb_out);
    return rc ? gfail(g, rc, "peer resolve", g->m[0]) : B200RT_OK;
}

int b200rt_group_download_accum(b200rt_group *g, float *rgb_out) {
    if (!g || !rgb_out) return B200RT_ERR_INVALID;
    b200rt_ctx *ctx = g->m[0];
    if (!ctx->total) return gfail(g, B200RT_ERR_NO_SCENE, "no accumulation buffer yet");
    std::vector<const void *> peers;
    int rc;
    try { rc = group_peers(g, peers); } catch (...) { return gfail(g, B200RT_ERR_INVALID, "out of host memory"); }
    if (rc) return rc;
    const size_t n = (size_t)ctx->acc_w * ctx->acc_h * 3;
    if ((rc = group_scratch(g, n))) return gfail(g, rc, "scratch", ctx);
    CK(cudaSetDevice(ctx->device));
    if ((rc = launch_peers(ctx, peers.data(), (int32_t)peers.size(), g->scratch, nullptr, 1.f, 0))) return gfail(g, rc, "peer reduce", ctx);
    CK(cudaMemcpyAsync(rgb_out, g->scratch, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return B200RT_OK;
}

int b200rt_group_get_stats(b200rt_group *g, b200rt_stats *out) {
    if (!g || !out) return B200RT_ERR_INVALID;
    b200rt_stats sum{};
    for (b200rt_ctx *c : g->m) {
        b200rt_stats s{};
        const int rc = b200rt_get_stats(c, &s);
        if (rc) return gfail(g, rc, "member stats", c);
        sum.paths += s.paths; sum.rays += s.rays; sum.light_queries += s.light_queries; sum.node_tests += s.node_tests;
        sum.tri_tests += s.tri_tests; sum.light_node_tests += s.light_node_tests; sum.light_tri_tests += s.light_tri_tests;
        sum.shaded_hits += s.shaded_hits; sum.tex_lookups += s.tex_lookups; sum.launches += s.launches;
        sum.extend_launches += s.extend_launches;
        sum.kernel_ms = std::max(sum.kernel_ms, s.kernel_ms); sum.extend_ms = std::max(sum.extend_ms, s.extend_ms);
        sum.shade_ms = std::max(sum.shade_ms, s.shade_ms); sum.light_ms = std::max(sum.light_ms, s.light_ms);
    }
    *out = sum;
    return B200RT_OK;
}

}  // extern "C"
