// Device-side scene layout and launch parameters.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "b200rt.h"

namespace b200rt {

// 32-byte BVH record = ONE traversal step: the boxes of BOTH children of an inner node, quantised
// to B200RT_QBITS (15) bits per coordinate on the grid spanned by the tree's root box (conservatively:
// lo rounded down, hi rounded up), plus the two child words.  Fetched with one 256-bit read-only load.
// (ncu on the first wavefront showed the L1/TEX pipe at 91 % of peak with 64-byte float pairs --
// profiles/r01_ncu_full_summary.json -- so the record was halved.)  lo and hi of one axis share a
// word so that a per-ray PRMT selector picks the near / far plane without a select instruction:
//   a.x = lo0.x | hi0.x << 16   a.y = lo0.y | hi0.y << 16   a.z = lo0.z | hi0.z << 16
//   a.w = lo1.x | hi1.x << 16   b.x = lo1.y | hi1.y << 16   b.y = lo1.z | hi1.z << 16
//   b.z = child word 0          b.w = child word 1
// child word: inner node -> index of its own record; leaf -> B200RT_LEAF | count << 27 | first triangle
#ifndef B200RT_QBITS
#define B200RT_QBITS 15      /* 15: the PRMT result is used as a float directly (no FADD); 16: finer grid, one more op per plane */
#endif
#define B200RT_QMAX ((1 << B200RT_QBITS) - 1)
#define B200RT_LEAF 0x80000000u   /* child word: leaf flag */
#define B200RT_STACK 64           /* traversal stack entries per thread; the upload bounds the tree depth by it */
struct DPair {
    uint4 a, b;
};
static_assert(sizeof(DPair) == 32, "DPair must be 32 bytes");

// Quantisation grid of one tree: world = base + q * scale.
struct DQuant {
    float base[3], scale[3], inv_scale[3];
};

// 48-byte Moeller-Trumbore record: v0 = data3 (barycentric origin), e1 = data - data3 (weight u),
// e2 = data2 - data3 (weight v)  (hw8/src/primitives.cpp:86-88).
//   scene triangles: v0.w = material id (int bits), e1.w = tangent w of `data` (Q14), e2.w = 0
//   light triangles: v0.w = pointProb = 1 / (0.5 |e1 x e2|) (distributions.h:76)
struct DTri {
    float4 v0, e1, e2;
};
static_assert(sizeof(DTri) == 48, "DTri must be 48 bytes");

// 96-byte shading record of a triangle, six float4:
//   q[0] = n(data).xyz,  uv(data).x     q[3] = tan(data).xyz,  uv(data2).y
//   q[1] = n(data2).xyz, uv(data).y     q[4] = tan(data2).xyz, uv(data3).x
//   q[2] = n(data3).xyz, uv(data2).x    q[5] = tan(data3).xyz, uv(data3).y
struct DShade {
    float4 q[6];
};
static_assert(sizeof(DShade) == 96, "DShade must be 96 bytes");

// Light triangles carry their three vertex normals for the hw8 pdf (distributions.h:140-145).
struct DLightNormals {
    float4 n0, n1, n2;  // data, data2, data3
};

// An image on the device: row-major 32-bit texels (R | G << 8 | B << 16 | A << 24) in the scene arena, top row first
// like the reference's Texture (hw8/src/include/gltf_structs.h).  Plain read-only loads, not texture objects: the
// kernel needs the four integer taps of the reference's own bilinear blend, and per-lane texture HANDLES made every
// fetch a loop over the distinct handles of a warp (28 % of the shade stage's stall samples, ncu).
typedef const uint32_t *DTexels;

struct DMaterial {
    float base_color[3];
    float emission[3];
    float metallic_factor;
    float roughness_factor;
    DTexels tex_base, tex_emissive, tex_mr, tex_normal;  // null = absent
    int32_t w_base, h_base, w_emissive, h_emissive, w_mr, h_mr, w_normal, h_normal;
    int32_t kind;        // hw6: B200RT_MAT_DIFFUSE / METALLIC / DIELECTRIC
    float ior;           // hw6 dielectric
};

struct DCamera {
    float pos[3], right[3], up[3], forward[3];
    float tan_x, tan_y;   // computed on the host with the reference's own libm call
    int32_t width, height;
};

#define B200RT_LIGHT_TOP 6      /* <= 6: the wavefront packs the hit mask into the top bits of a light-queue entry */
struct DGltfScene {
    const DPair *nodes;
    const DTri *tris;
    const DShade *shade;
    const DPair *light_nodes;
    const DTri *light_tris;          // emissive triangles in the order of the light BVH's leaves (traversal)
    const DTri *light_sample_tris;   // the same triangles in FiguresMix::figures_ order (the index a random number picks)
    const DLightNormals *light_normals;
    const DMaterial *materials;
    int32_t n_light_tris;
    uint32_t root, light_root;   // child word of each tree's root
    DQuant q, lq;                // quantisation grids of the scene BVH and the light BVH
    float light_min[3], light_max[3];   // root box of the light BVH (a real filter, unlike the scene's)
    // a cut through the top of the light BVH (<= B200RT_LIGHT_TOP boxes covering every emitter): a ray that
    // misses all of them has a light pdf of 0 -- the cheap pre-filter shade uses to skip the light stage
    float light_top[B200RT_LIGHT_TOP][6];
    int32_t n_light_top;
    // child words of the cut's nodes: the light-pdf traversal starts at the boxes the pre-filter found pierced
    // instead of walking down from the root again (light_top_start = 0: words unavailable, start at the root)
    uint32_t light_top_word[B200RT_LIGHT_TOP];
    int32_t light_top_start;
    int32_t hw8;                 // 1 = hw8 semantics (textures, shading-normal pdf), 0 = hw7
    int32_t hw6;                 // 1 = hw6: the hw5 integrator (diffuse / mirror / dielectric) on these triangles (hw6_kernel.cuh)
    float weight_clamp;          // a bounce weight above this in any channel ends the path: 6 (reference), or FLT_MAX
    float bg[3];
    DTexels env;                 // null = absent
    int32_t env_w, env_h;
};

struct DCounters {
    unsigned long long paths, rays, light_queries, node_tests, tri_tests, light_node_tests, light_tri_tests,
        shaded_hits, tex_lookups, stack_overflow;
};

struct DRenderParams {
    DCamera cam;
    int32_t spp_begin, spp_end;   // global sample indices rendered by this launch
    int32_t n_chunks;             // the sample range is cut into n_chunks slabs of accum
    int32_t ray_depth;
    uint32_t seed;
    float *accum;                 // [n_chunks][H*W*3] partial sums (slab 0 is the running total)
    DCounters *counters;
};

// analytic scenes (hw1, hw2, hw4): primitives + lights live in __constant__ memory
#define B200RT_MAX_PRIMS 64
#define B200RT_MAX_LIGHTS 16
struct DAnalyticScene {
    b200rt_primitive prims[B200RT_MAX_PRIMS];
    b200rt_light lights[B200RT_MAX_LIGHTS];
    int32_t n_prims, n_lights;
    int32_t light_prims[B200RT_MAX_PRIMS];  // hw4: indices of emissive BOX/ELLIPSOID figures
    int32_t n_light_prims;
    float bg[3], ambient[3];
    int32_t mode;
    // Monte-Carlo modes (hw3 - hw5) only: the world -> local map p -> q p conj(q) of each figure as a row-major 3x3
    // matrix (built at upload by applying the reference's quaternion arithmetic to the basis vectors: it carries the
    // |q|^2 scale of non-unit quaternions), frame[k][9] != 0 marks an identity rotation (no transform at
    // all); frame[k][12..14] = 1 / data[0..2] (ellipsoid radii, box half sizes): the Monte-Carlo modes multiply where
    // the reference divides.  hw1 / hw2 keep the two quaternion sandwiches per ray and figure: their images are bit-exact.
    float frame[B200RT_MAX_PRIMS][16];
};

}  // namespace b200rt
