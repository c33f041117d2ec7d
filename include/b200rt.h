/*
 * b200rt.h -- C ABI of the B200 (sm_100a) render core.
 *
 * This is the drop-in boundary for ONE hot path of volivan239/raytracing-course-hw: the
 * per-pixel, per-sample ray-scene intersection and shading loop.  In the reference that
 * path is the body of the pixel loop in sceneio::renderScene
 *     hw8/src/sceneio.cpp:386-396  (OpenMP loop: rng(i), Scene::getPixel, tonemap, gamma, quantise)
 *     hw4/src/sceneio.cpp:114-122, hw2/src/sceneio.cpp:156-163, hw1/src/sceneio.cpp:103-108 (serial)
 * and everything below Scene::getPixel (hw8/src/scene.cpp:167-177, hw4/src/scene.cpp:110-118,
 * hw2/src/scene.cpp:87-89, hw1/src/scene.cpp:7-20).  The reference has no FFI; a maintainer
 * would replace that loop with:  b200rt_upload_scene(once) ; b200rt_render ; b200rt_resolve.
 * INTEGRATION.md shows that binding.
 *
 * Conventions: plain C structs, plain pointers and sizes.  The caller owns every host buffer
 * (they are copied during the call); the library owns all device memory.  Every entry point
 * returns 0 (B200RT_OK) or a negative error code, never throws, never exits.  A context is
 * driven by one host thread.  There is NO CPU fallback: without a CUDA device every compute
 * entry point returns B200RT_ERR_NO_DEVICE.
 */
#ifndef B200RT_H
#define B200RT_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- error codes ------------------------------------------------------------------- */
#define B200RT_OK                0
#define B200RT_ERR_INVALID      -1  /* bad argument / malformed scene description          */
#define B200RT_ERR_NO_DEVICE    -2  /* no CUDA device (there is no CPU fallback)            */
#define B200RT_ERR_CUDA         -3  /* a CUDA runtime call failed; see b200rt_last_error    */
#define B200RT_ERR_NO_SCENE     -4  /* render/resolve before upload_scene                   */
#define B200RT_ERR_UNSUPPORTED  -5  /* mode / feature not implemented on the device path    */
#define B200RT_ERR_STACK        -6  /* BVH deeper than the traversal stack                  */
#define B200RT_ERR_TOO_LARGE    -7  /* frame above B200RT_NOALLOC_MAX_PIXELS (b200rt_resolve_noalloc) */

/* ---- scene semantics: which homework's integrator is reproduced -- */
#define B200RT_MODE_HW1 1  /* flat-colour primary cast          hw1/src/scene.cpp:7-20      */
#define B200RT_MODE_HW2 2  /* Whitted: lights, shadows, branch  hw2/src/scene.cpp:29-85     */
#define B200RT_MODE_HW3 3  /* MC, uniform hemisphere            hw3/src/scene.cpp:31-87     */
#define B200RT_MODE_HW4 4  /* MC, cosine + light mixture        hw4/src/scene.cpp:50-108    */
#define B200RT_MODE_HW5 5  /* hw4 + triangles + BVH (text)      hw5/src/scene.cpp:47-105    */
#define B200RT_MODE_HW6 6  /* hw5 integrator on glTF triangles  hw6/src/scene.cpp           */
#define B200RT_MODE_HW7 7  /* GGX metal-rough path tracer       hw7/src/scene.cpp:29-61     */
#define B200RT_MODE_HW8 8  /* hw7 + textures/normal map/env map hw8/src/scene.cpp:84-165    */

/* analytic primitive types (hw4/src/include/primitives.h FigureType; hw5 adds TRIANGLE) */
#define B200RT_PRIM_ELLIPSOID 0
#define B200RT_PRIM_PLANE     1
#define B200RT_PRIM_BOX       2
#define B200RT_PRIM_TRIANGLE  3

/* text-scene materials (hw2/src/include/material.h) */
#define B200RT_MAT_DIFFUSE    0
#define B200RT_MAT_METALLIC   1
#define B200RT_MAT_DIELECTRIC 2

/* hw2 lights (hw2/src/include/light_source.h:12-31) */
#define B200RT_LIGHT_POINT       0
#define B200RT_LIGHT_DIRECTIONAL 1

/* ---- reference-shaped scene description (what Scene holds after loadScene) ------------ */

/* hw1-hw5 `Figure` (hw4/src/include/primitives.h:33-52): world position + rotation quaternion
 * (x,y,z,w, never normalised -- SURVEY Q3), shape parameter `data` (ellipsoid radii, plane
 * normal, box half extents; hw5 triangles also use data2/data3). */
typedef struct b200rt_primitive {
    int32_t type;
    int32_t material;
    float   data[3];
    float   data2[3];
    float   data3[3];
    float   position[3];
    float   rotation[4];
    float   color[3];
    float   emission[3];
    float   ior;
} b200rt_primitive;

/* hw2 light (hw2/src/light_source.cpp:9-23): `v` is the position (point) or direction. */
typedef struct b200rt_light {
    int32_t type;
    float   intensity[3];
    float   v[3];
    float   attenuation[3];
} b200rt_light;

/* hw7/hw8 `Figure` (hw8/src/include/primitives.h:58-75): three `Vertex` in the reference's
 * member order data, data2, data3 (= glTF p1, p3, p2 after the loader's swizzle,
 * hw8/src/sceneio.cpp:289); data3 is the barycentric origin.  World space. */
typedef struct b200rt_triangle {
    float    p[3][3];    /* coords   of data, data2, data3 */
    float    uv[3][2];   /* texcoords                       */
    float    n[3][3];    /* normals (unit)                  */
    float    tan[3][4];  /* tangents xyz (unit) + w         */
    uint32_t material;   /* index into materials            */
} b200rt_triangle;

/* `BvhNode` (hw8/src/include/bvh.h:9-16): left==0 marks a leaf holding [first,last). */
typedef struct b200rt_bvh_node {
    float    aabb_min[3];
    float    aabb_max[3];
    uint32_t left, right, first, last;
} b200rt_bvh_node;

/* `GltfMaterial` + `MaterialModel` (hw8/src/include/gltf_structs.h:38-47, material.h:5-9).
 * Texture fields are IMAGE indices (textureDescs[..].source already applied) or -1. */
typedef struct b200rt_material {
    float   base_color[3];
    float   emission[3];      /* emissiveFactor x KHR emissive strength */
    float   metallic_factor;
    float   roughness_factor;
    int32_t base_color_image;
    int32_t emissive_image;
    int32_t metallic_roughness_image;
    int32_t normal_image;
    /* hw6 only (hw6/src/include/gltf_structs.h:28-36, hw6/src/sceneio.cpp:179-183): the loader maps
     * baseColorFactor alpha < 1 to DIELECTRIC (ior 1.5), else metallicFactor > 0 to METALLIC, else DIFFUSE. */
    int32_t kind;             /* B200RT_MAT_* */
    float   ior;
} b200rt_material;

/* `Texture` (hw8/src/include/gltf_structs.h:10-13): tightly packed RGB8, row 0 first. */
typedef struct b200rt_image {
    int32_t        width, height;
    const uint8_t *rgb;
} b200rt_image;

typedef struct b200rt_scene_desc {
    int32_t mode;                 /* B200RT_MODE_HWn */
    int32_t width, height;        /* DIMENSIONS / argv */
    int32_t samples;              /* SAMPLES / argv (1 for hw1, hw2) */
    int32_t ray_depth;            /* RAY_DEPTH; 6 for glTF modes (hw8/src/include/scene.h:32) */
    float   bg_color[3];
    float   ambient[3];           /* hw2 AMBIENT_LIGHT */
    float   cam_pos[3], cam_right[3], cam_up[3], cam_forward[3];
    float   cam_fov;              /* CAMERA_FOV_X (text modes) or yfov (glTF modes) */

    const b200rt_primitive *prims;        int32_t n_prims;        /* hw1-hw5 */
    const b200rt_light     *lights;       int32_t n_lights;       /* hw2     */

    const b200rt_triangle  *tris;         int32_t n_tris;         /* hw6-hw8, in BVH order */
    const b200rt_bvh_node  *nodes;        int32_t n_nodes;        /* Scene::bvh.nodes      */
    int32_t                 root;
    const b200rt_triangle  *light_tris;   int32_t n_light_tris;   /* FiguresMix::figures_  */
    const b200rt_bvh_node  *light_nodes;  int32_t n_light_nodes;  /* FiguresMix::bvh.nodes */
    const b200rt_material  *materials;    int32_t n_materials;
    const b200rt_image     *images;       int32_t n_images;
    const b200rt_image     *env_map;      /* NULL when absent (hw8 argv[6]) */

    /* hw5 (text scenes with TRIANGLE figures, hw5/src/scene.cpp:18-45, hw5/src/include/distributions.h:172-199):
     * `prims` holds Scene::figures after initBVH -- the n_bvh_prims non-plane figures in BVH order first, then the
     * planes (scanned linearly); `nodes`/`root` are Scene::bvh over prims[0, n_bvh_prims); `light_prims` are
     * FiguresMix::figures_ (copies of the emissive BOX / ELLIPSOID / TRIANGLE figures in the light BVH's order) and
     * `light_nodes` FiguresMix::bvh over them.  Zero / NULL in every other mode. */
    const b200rt_primitive *light_prims;  int32_t n_light_prims;
    int32_t                 n_bvh_prims;
} b200rt_scene_desc;

/* ---- context ------------------------------------------------------------------------ */
typedef struct b200rt_ctx b200rt_ctx;

typedef struct b200rt_options {
    int32_t  device;        /* CUDA device ordinal */
    uint32_t seed;          /* key of the counter-based RNG (0 = default stream) */
    int32_t  flags;         /* B200RT_FLAG_* */
} b200rt_options;

#define B200RT_FLAG_COUNTERS   1  /* also count node fetches / triangle tests / texture lookups */
#define B200RT_FLAG_MEGAKERNEL 2  /* glTF modes (hw6 - hw8): one-kernel-per-path variant instead of the wavefront pipeline */
#define B200RT_FLAG_REFERENCE_BVH 8 /* glTF modes: traverse the scene BVH exactly as given in the description (the reference's
                                      SAH tree) instead of the binned-SAH tree the library builds over the same triangles;
                                      results are identical up to closest-hit ties, the rebuilt tree tests ~20 % fewer boxes.
                                      hw6 always rebuilds (its reference tree is degenerate) */
#define B200RT_FLAG_SERIAL_WAVES 4 /* wavefront: one wave in flight on one stream (default: two waves overlap on two
                                      streams); the per-stage times in b200rt_stats are then exclusive kernel times */

#define B200RT_FLAG_KEEP_DARK_RAYS 16 /* wavefront: do not end a path one bounce early when its last ray cannot reach an
                                       emitter (black background only; the image is bit-identical either way -- test aid) */
#define B200RT_FLAG_EXPLICIT_CAMERA_STATE 32 /* wavefront: wf_generate stores the whole state record of a camera path
                                       instead of the direction only (bit-identical image -- test aid) */

#define B200RT_FLAG_WAVEFRONT 128 /* hw6: always the wavefront pipeline.  By default hw6 scenes of up to
                                       B200RT_HW6_PIXEL_KERNEL_MAX_TRIS triangles run the one-thread-per-pixel kernel (faster on
                                       a handful of triangles: no path state to stream), larger ones the wavefront pipeline
                                       (2x on the reference's practice6_2); both trace the same paths */
#define B200RT_HW6_PIXEL_KERNEL_MAX_TRIS 2048

#define B200RT_FLAG_NO_FIREFLY_CLAMP 64 /* hw7 / hw8: do not end a path when a bounce weight exceeds 6 in some channel (the
                                       reference's DIRTY_DENOISE_HACK_BUBEN, hw8/src/scene.cpp:7,161 -- biased, darkens).  The
                                       one-sample MIS estimator is then unbiased; NaN / inf weights still end the path.
                                       Default (flag off) = the reference's image */

typedef struct b200rt_stats {
    double   kernel_ms;        /* device time of the last b200rt_render (CUDA events) */
    uint64_t paths;            /* camera samples started                                */
    uint64_t rays;             /* closest-hit scene queries (Scene::intersect calls)    */
    uint64_t light_queries;    /* FiguresMix::pdf traversals                            */
    uint64_t node_tests;       /* scene-BVH node boxes tested    (COUNTERS flag only)   */
    uint64_t tri_tests;        /* scene triangles tested          (COUNTERS flag only)   */
    uint64_t light_node_tests; /* light-BVH node boxes tested     (COUNTERS flag only)   */
    uint64_t light_tri_tests;  /* light triangles tested          (COUNTERS flag only)   */
    uint64_t shaded_hits;      /* hits that fetched a shading record (COUNTERS only)    */
    uint64_t tex_lookups;      /* bilinear texture lookups           (COUNTERS only)    */
    int32_t  launches;         /* kernels launched by the last render+resolve           */
    int32_t  extend_launches;  /* wavefront: closest-hit traversal launches in the last render */
    double   extend_ms;        /* wavefront: summed device time of those launches (CUDA events) */
    double   shade_ms;         /* wavefront: summed device time of the shade launches   */
    double   light_ms;         /* wavefront: summed device time of the light-pdf launches */
} b200rt_stats;

/* Replaces: constructing the render state inside main() (hw8/src/main.cpp:8-15). */
int b200rt_create(b200rt_ctx **out, const b200rt_options *opt);
void b200rt_destroy(b200rt_ctx *ctx);

/* Replaces: Scene being resident in host memory for getPixel (hw8/src/include/scene.h:23-41).
 * Flattens the reference-shaped arrays to the device layout and uploads. */
int b200rt_upload_scene(b200rt_ctx *ctx, const b200rt_scene_desc *scene);

/* Replaces: the pixel loop + Scene::getPixel (hw8/src/sceneio.cpp:387-392, scene.cpp:167-177).
 * Renders global sample indices [spp_begin, spp_end) of every pixel and ADDS them to the
 * float accumulation buffer (zeroed by b200rt_upload_scene, b200rt_clear or a frame-size change).
 * width/height/ray_depth <= 0 take the scene's values.  Asynchronous on the context stream. */
int b200rt_render(b200rt_ctx *ctx, int32_t width, int32_t height,
                  int32_t spp_begin, int32_t spp_end, int32_t ray_depth);
int b200rt_clear(b200rt_ctx *ctx);

/* Replaces: `1.0/samples * color`, aces_tonemap, gamma_corrected, toExternColorFormat
 * (hw8/src/scene.cpp:176, hw8/src/color.cpp:4-31; hw1: round(255 c) only, hw1/src/color.cpp:13-19).
 * total_spp = number of samples the accumulation buffer holds (all ranks).  Blocks; writes
 * width*height*3 bytes, top row first (the PPM payload). */
int b200rt_resolve(b200rt_ctx *ctx, int32_t total_spp, uint8_t *rgb_out);

/* Accumulated radiance SUM, width*height*3 floats (for RMSE tests and multi-GPU reduce). */
int b200rt_download_accum(b200rt_ctx *ctx, float *rgb_out);
int b200rt_upload_accum(b200rt_ctx *ctx, const float *rgb_in);
/* Device pointer + element count of the accumulation buffer (for NCCL reduce by the host). */
int b200rt_accum_device_ptr(b200rt_ctx *ctx, void **ptr, size_t *n_floats);

/* Scene replication for multi-GPU runs (the scene is replicated on every GPU).  Instead of repeating the host
 * side of b200rt_upload_scene (BVH rebuild, record building) once per GPU, the context that uploaded EXPORTS its
 * device-resident scene -- one contiguous device blob + a fixed-size header -- and every peer IMPORTS a copy the
 * caller delivers device to device (ncclBroadcast over NVLink in raytracing_course_hw_b200/dist.py, or a peer copy):
 *     b200rt_export_scene(ctx0, &hdr, &src, &bytes);
 *     b200rt_import_scene_begin(ctxK, &hdr, &dst);     -- sizes ctxK's arena, returns where the blob goes
 *     <copy `bytes` bytes src -> dst on ctxK's stream (b200rt_stream)>
 *     b200rt_import_scene_end(ctxK, &hdr);             -- rebases the pointers inside the copy; ctxK can render
 * glTF modes (hw6-hw8) only; text scenes are a few hundred bytes, upload them per GPU. */
typedef struct b200rt_scene_image { uint64_t opaque[256]; } b200rt_scene_image;
int b200rt_export_scene(b200rt_ctx *ctx, b200rt_scene_image *hdr, void **device_blob, uint64_t *bytes);
int b200rt_import_scene_begin(b200rt_ctx *ctx, const b200rt_scene_image *hdr, void **device_dst);
int b200rt_import_scene_end(b200rt_ctx *ctx, const b200rt_scene_image *hdr);
/* The three steps in one call for two contexts of the same process (cudaMemcpyPeerAsync between their devices). */
int b200rt_replicate_scene(b200rt_ctx *src, b200rt_ctx *dst);

/* ---- several GPUs (the sample split of ONE frame, reduce inside the library) -------------------
 * Replaces: the OpenMP pixel loop as the unit of parallelism (hw8/src/sceneio.cpp:386-396).  The scene is replicated,
 * GPU g renders the global sample indices [b + g*S/G, b + (g+1)*S/G) of every pixel into its own float buffer, and the
 * buffers are combined by ONE kernel on the first GPU that READS the other GPUs' buffers through peer memory (NVLink /
 * NVSwitch loads), sums them in rank order and tonemaps in the same pass -- no host copy, no separate collective.
 *
 * Peer level (any mix of contexts, also of other processes through CUDA IPC): */
#define B200RT_MAX_PEERS 15
/* rgb_out = resolve(own accumulation buffer + peer_accum[0] + ... + peer_accum[n_peers-1]) (summed in that order);
 * every peer pointer must be readable from ctx's device (same device, peer access, or b200rt_ipc_open) and hold the
 * same frame size; the peers' renders must have finished (b200rt_synchronize on their contexts).  Blocks. */
int b200rt_resolve_peers(b200rt_ctx *ctx, int32_t total_spp, const void *const *peer_accum, int32_t n_peers, uint8_t *rgb_out);
/* own accumulation buffer += the peers' (same order; for b200rt_download_accum of the whole frame).  Asynchronous. */
int b200rt_reduce_peers(b200rt_ctx *ctx, const void *const *peer_accum, int32_t n_peers);
/* Lets ctx's device read memory of `peer_device` directly (cudaDeviceEnablePeerAccess; a no-op on the same device). */
int b200rt_enable_peer(b200rt_ctx *ctx, int32_t peer_device);
/* One process per GPU: a 64-byte handle of the accumulation buffer to hand to the reducing process ... */
typedef struct b200rt_ipc_handle { uint8_t opaque[64]; } b200rt_ipc_handle;
int b200rt_accum_ipc_export(b200rt_ctx *ctx, b200rt_ipc_handle *out);
/* ... which maps it into its own address space (and unmaps it when done). */
int b200rt_ipc_open(b200rt_ctx *ctx, const b200rt_ipc_handle *h, void **peer_ptr);
int b200rt_ipc_close(b200rt_ctx *ctx, void *peer_ptr);

/* Group level (one process drives all GPUs; what the CLI uses).  devices[] may repeat a device (two contexts on one
 * GPU take the same path: that is how a single-GPU box tests it). */
typedef struct b200rt_group b200rt_group;
int b200rt_group_create(b200rt_group **out, const int32_t *devices, int32_t n_devices, const b200rt_options *opt);
void b200rt_group_destroy(b200rt_group *g);
int32_t b200rt_group_size(const b200rt_group *g);
b200rt_ctx *b200rt_group_member(b200rt_group *g, int32_t i);
/* One host-side upload on member 0, device-to-device replication to the others (glTF modes; text scenes upload per member). */
int b200rt_group_upload_scene(b200rt_group *g, const b200rt_scene_desc *scene);
/* Splits [spp_begin, spp_end) contiguously over the members (hw1 / hw2: member 0 alone) and launches all of them. */
int b200rt_group_render(b200rt_group *g, int32_t width, int32_t height, int32_t spp_begin, int32_t spp_end, int32_t ray_depth);
int b200rt_group_clear(b200rt_group *g);
/* Waits for every member, then the fused peer reduce + resolve on member 0. */
int b200rt_group_resolve(b200rt_group *g, int32_t total_spp, uint8_t *rgb_out);
/* Radiance SUM of the whole frame (peer reduce into a scratch copy on member 0; the members' buffers are untouched). */
int b200rt_group_download_accum(b200rt_group *g, float *rgb_out);
/* Counters summed over the members, times = the slowest member's. */
int b200rt_group_get_stats(b200rt_group *g, b200rt_stats *out);
const char *b200rt_group_last_error(b200rt_group *g);

/* ---- measurement cross-check ("N_* counted by the kernel and cross-checked against a CPU traversal of the
 * same flattened BVH on a fixed ray set") ---------------------------------------------------------------------------
 * Replaces nothing in the reference by itself: it exposes the traversal that replaces BVH::intersect_
 * (hw8/src/include/bvh.h:111-142) to a checker.  b200rt_trace_rays runs the PRODUCTION closest-hit kernel (wf_extend,
 * counting build) over n caller-supplied rays (host array, 6 floats each: origin xyz, direction xyz; t in (0, 1e4)) and
 * returns per ray {t, u, v, bits} with bits = triangle index | inside << 31, or 0x7FFFFFFF for a miss, plus the number of
 * child boxes and triangle records it tested.  b200rt_scene_bvh_layout / _download hand out the flattened tree exactly as
 * it sits in device memory (32-byte sibling-pair records, 48-byte triangle records) so that
 * oracle/pairs_traverse.c can walk the same records on the CPU.  glTF modes only. */
typedef struct b200rt_bvh_layout {
    uint64_t n_records, n_tris;
    uint32_t root;            /* child word of the root (leaf flag 0x80000000 | count << 27 | first) */
    int32_t  quant_bits;      /* 15 */
    float    guard;           /* conservative guard band of the box test, in grid units */
    float    base[3], scale[3], inv_scale[3];   /* world = base + q * scale */
} b200rt_bvh_layout;
int b200rt_trace_rays(b200rt_ctx *ctx, const float *rays, int32_t n, float *hits, uint64_t *node_tests, uint64_t *tri_tests);
int b200rt_scene_bvh_layout(b200rt_ctx *ctx, b200rt_bvh_layout *out);
int b200rt_scene_bvh_download(b200rt_ctx *ctx, void *records, void *tris);

/* ---- no-allocation resolve ------------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arithmetic and bytes as b200rt_resolve (mean, ACES, gamma, quantise) but context-free and allocation-free: the
 * RGB8 staging buffer between the kernel and the device-to-host copy is a static __device__ array inside the library
 * (B200RT_NOALLOC_MAX_PIXELS * 3 bytes = 6 MiB on every GPU that loads libb200rt.so, until exit) instead of a
 * cudaMalloc'ed buffer of a context.  The call performs no cudaMalloc / cudaFree / stream or event creation.
 *   device            CUDA device ordinal that owns accum_sum_device
 *   stream            cudaStream_t as void* (NULL = the default stream); the call blocks until the bytes are on the host
 *   accum_sum_device  DEVICE pointer: width * height * 3 floats, the radiance SUM (e.g. b200rt_accum_device_ptr)
 *   total_spp, mode   as in b200rt_resolve: hw1 / hw2 hold one sample, hw1 is not tonemapped (B200RT_MODE_HWn)
 *   rgb_out           HOST pointer, width * height * 3 bytes, top row first
 * Frames above B200RT_NOALLOC_MAX_PIXELS pixels are rejected with B200RT_ERR_TOO_LARGE (use b200rt_resolve).
 * The static array is shared by all calls on a device and keeps the last call's contents; no part of it has to be
 * cleared (a call overwrites every byte it later reads), but calls that overlap on one GPU race on it: serialise them. */
#define B200RT_NOALLOC_MAX_PIXELS (2048 * 1024)
int b200rt_resolve_noalloc(int32_t device, void *stream, const float *accum_sum_device, int32_t width, int32_t height,
                           int32_t total_spp, int32_t mode, uint8_t *rgb_out);

/* hw1/hw2 parity aid: primitive index of the primary hit per pixel (-1 = miss), int32 W*H. */
int b200rt_primary_ids(b200rt_ctx *ctx, int32_t width, int32_t height, int32_t *ids_out);

int b200rt_synchronize(b200rt_ctx *ctx);
int b200rt_get_stats(b200rt_ctx *ctx, b200rt_stats *out);
/* cudaStream_t of the context as void* (so callers can record events on it). */
int b200rt_stream(b200rt_ctx *ctx, void **stream_out);

const char *b200rt_strerror(int code);
const char *b200rt_last_error(b200rt_ctx *ctx);
int b200rt_device_count(void);

#ifdef __cplusplus
}
#endif
#endif /* B200RT_H */
