// K4: glTF path-tracing megakernel (hw7 / hw8 semantics) for sm_100a.
// Replaces Scene::getPixel / getColor / intersect and everything below them:
//   hw8/src/scene.cpp:84-187, hw8/src/include/bvh.h:111-142, hw8/src/primitives.cpp:18-125,
//   hw8/src/include/distributions.h:35-280, hw8/src/include/material.h:11-65
//   (hw7 deltas: hw7/src/scene.cpp:29-61, hw7/src/include/material.h:44-59, distributions.h:140-145).
// Not a port: iterative path loop instead of recursion, while-while stack traversal over 32-byte
// sibling-pair nodes with near-child-first ordering and distance pruning, Moeller-Trumbore
// triangle records, Philox counter RNG, texels through point-sampled texture objects with a
// shared-memory sRGB decode table.
#pragma once
#include "device_scene.h"
#include "rt_math.cuh"

namespace b200rt {

#define B200RT_TMAX 1e4f
#define B200RT_EPS 1e-4f

struct TraceHit {
    float t, u, v;
    int tri;
    float det;
};

struct ThreadCounters {
    unsigned paths, rays, light_queries, node_tests, tri_tests, light_node_tests, light_tri_tests, shaded, tex, overflow;
};

// Slab test.  AABB::intersect accepts t1 <= t2 && t2 >= 0 (hw8/src/primitives.cpp:38); tbest adds the
// distance prune of BVH::intersect_ (bvh.h:118) -- conservative (<=) so ties still descend.
__device__ __forceinline__ bool slab6(float lx, float ly, float lz, float hx, float hy, float hz,
                                      const V3 o, const V3 inv, float tbest, float &tnear) {
    const float x0 = (lx - o.x) * inv.x, x1 = (hx - o.x) * inv.x;
    const float y0 = (ly - o.y) * inv.y, y1 = (hy - o.y) * inv.y;
    const float z0 = (lz - o.z) * inv.z, z1 = (hz - o.z) * inv.z;
    const float tmin = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fminf(z0, z1));
    const float tmax = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fmaxf(z0, z1));
    tnear = tmin;
    return tmin <= tmax && tmax >= 0.f && tmin <= tbest;
}

// Light-pdf pre-filter: false when the ray (o, d) misses the light BVH's root box or every box of the cut
// through its top levels (then FiguresMix::getTotalPdf is exactly 0: no pierced leaf box, distributions.h:122-165).
// Every light-pdf evaluation (megakernel, hw6, wavefront shade / light stages) starts with this same test, so all
// schedulings see the same set of non-zero queries.
// Returns the set of pierced cut boxes as a bit mask (bit k = box k); with no cut, bit 0 stands for the root box.
__device__ __forceinline__ uint32_t light_boxes_mask(const DGltfScene &sc, const V3 o, const V3 d) {
    const V3 inv = mk(1.f / d.x, 1.f / d.y, 1.f / d.z);
    float tn;
    if (!slab6(sc.light_min[0], sc.light_min[1], sc.light_min[2], sc.light_max[0], sc.light_max[1], sc.light_max[2], o, inv, INFINITY, tn)) return 0u;
    if (sc.n_light_top == 0) return 1u;
    uint32_t mask = 0u;
#pragma unroll
    for (int k = 0; k < B200RT_LIGHT_TOP; k++)
        if (k < sc.n_light_top &&
            slab6(sc.light_top[k][0], sc.light_top[k][1], sc.light_top[k][2], sc.light_top[k][3], sc.light_top[k][4], sc.light_top[k][5], o, inv, INFINITY, tn))
            mask |= 1u << k;
    return mask;
}
__device__ __forceinline__ bool light_boxes_hit(const DGltfScene &sc, const V3 o, const V3 d) { return light_boxes_mask(sc, o, d) != 0u; }
// Start of a light-pdf traversal for a non-zero mask: the pierced cut nodes in ascending order (the first becomes
// `cur`, the others are stacked so that they pop in order) -- every scheduling sums in this order.
__device__ __forceinline__ void light_start(const DGltfScene &sc, uint32_t mask, uint32_t *stack, int &sp, uint32_t &cur) {
    sp = 0;
    if (!sc.light_top_start || sc.n_light_top == 0) { cur = sc.light_root; return; }
#pragma unroll
    for (int k = B200RT_LIGHT_TOP - 1; k >= 0; k--)
        if (mask & (1u << k)) stack[sp++] = sc.light_top_word[k];
    cur = stack[--sp];
}

// A ray expressed on the 16-bit grid of a tree.  With o' = (o - base) / scale and inv' = scale / d the
// ray parameter of a grid plane q is t = q * inv' - o' * inv' (same t as in world space): one FFMA.
// The near/far plane of each axis is picked by the sign of d, and the constant terms carry a
// +-B200RT_QEPS grid-unit guard band that absorbs the float rounding of o' (boxes stay conservative).
// Zero direction components get a huge finite reciprocal so that no 0 * inf NaN can hide a miss.
#ifndef B200RT_QEPS
#define B200RT_QEPS 0.03f
#endif
struct QRay {
    V3 inv;        // scale / d (times 2^16 for the 15-bit encoding)
    V3 bn, bf;     // -(o' * inv') -/+ eps * |inv'|  (minus the encoding offset)
    uint32_t sx, sy, sz;   // PRMT selector of the NEAR plane of each axis (far = near ^ B200RT_SEL_FLIP)
};
__device__ __forceinline__ float safe_rcp(float d) {
    const float r = __fdiv_rn(1.f, d);
    return fabsf(r) < 1e30f ? r : copysignf(1e30f, d);
}
// u16 -> float without the conversion pipe: one PRMT splices a half of the record word into a float.
//   16-bit grid: into the low mantissa of 2^23 -> 2^23 + q, then one FADD removes the 2^23.
//   15-bit grid: into bits 8..22 under the exponent of 0.5 -> m = 0.5 + q * 2^-16, used AS IS: the plane
//   parameter q * inv + b becomes m * (2^16 inv) + (b - 2^15 inv), one FFMA; the rounding of the folded
//   constant moves a plane by < 2^-9 grid units (covered by B200RT_QEPS).
#if B200RT_QBITS == 15
#define B200RT_SEL_LO 0x7104u
#define B200RT_SEL_HI 0x7324u
#define B200RT_SEL_FLIP 0x0220u
__device__ __forceinline__ float q16(uint32_t w, uint32_t sel) { return __uint_as_float(__byte_perm(w, 0x3F000000u, sel)); }
#else
#define B200RT_SEL_LO 0x7610u
#define B200RT_SEL_HI 0x7632u
#define B200RT_SEL_FLIP 0x0022u
__device__ __forceinline__ float q16(uint32_t w, uint32_t sel) { return __uint_as_float(__byte_perm(w, 0x4B000000u, sel)) - 8388608.f; }
#endif
// Every operation is an explicitly rounded one (no FMA contraction left to the compiler): oracle/pairs_traverse.c
// restates this function on the CPU and must get the same planes bit for bit (SURVEY 8d counter cross-check).
__device__ __forceinline__ float qray_axis(float o, float d, float base, float scale, float inv_scale, float &bn, float &bf) {
    const float oq = __fmul_rn(__fsub_rn(o, base), inv_scale);
    const float ig = __fmul_rn(scale, safe_rcp(d));
    const float c = -__fmul_rn(oq, ig);
    // guard band: B200RT_QEPS grid units for the plane arithmetic + 2^-21 |o'| for the rounding of the origin's own
    // transform, which grows with its distance from the grid (a ray origin far outside the LIGHT tree's small grid --
    // a lamp in a large room -- would otherwise see boxes that are no longer conservative)
    const float g = __fmul_rn(fmaf(fabsf(oq), 4.76837158203125e-7f, B200RT_QEPS), fabsf(ig));
    bn = __fsub_rn(c, g); bf = __fadd_rn(c, g);
#if B200RT_QBITS == 15
    const float off = __fmul_rn(32768.f, ig);      // exact (power of two)
    bn = __fsub_rn(bn, off); bf = __fsub_rn(bf, off);
    return __fmul_rn(65536.f, ig);
#else
    return ig;
#endif
}
__device__ __forceinline__ QRay make_qray(const DQuant &q, const V3 o, const V3 d) {
    QRay r;
    r.inv.x = qray_axis(o.x, d.x, q.base[0], q.scale[0], q.inv_scale[0], r.bn.x, r.bf.x);
    r.inv.y = qray_axis(o.y, d.y, q.base[1], q.scale[1], q.inv_scale[1], r.bn.y, r.bf.y);
    r.inv.z = qray_axis(o.z, d.z, q.base[2], q.scale[2], q.inv_scale[2], r.bn.z, r.bf.z);
    r.sx = d.x < 0.f ? B200RT_SEL_HI : B200RT_SEL_LO;
    r.sy = d.y < 0.f ? B200RT_SEL_HI : B200RT_SEL_LO;
    r.sz = d.z < 0.f ? B200RT_SEL_HI : B200RT_SEL_LO;
    return r;
}

// AABB::intersect accepts t1 <= t2 && t2 >= 0 (hw8/src/primitives.cpp:38); tbest adds the distance
// prune of BVH::intersect_ (bvh.h:118) -- conservative (<=) so ties still descend.
__device__ __forceinline__ bool slab_q(uint32_t wx, uint32_t wy, uint32_t wz, const QRay &r, float tbest, float &tnear) {
    const float tx0 = fmaf(q16(wx, r.sx), r.inv.x, r.bn.x), tx1 = fmaf(q16(wx, r.sx ^ B200RT_SEL_FLIP), r.inv.x, r.bf.x);
    const float ty0 = fmaf(q16(wy, r.sy), r.inv.y, r.bn.y), ty1 = fmaf(q16(wy, r.sy ^ B200RT_SEL_FLIP), r.inv.y, r.bf.y);
    const float tz0 = fmaf(q16(wz, r.sz), r.inv.z, r.bn.z), tz1 = fmaf(q16(wz, r.sz ^ B200RT_SEL_FLIP), r.inv.z, r.bf.z);
    const float tmin = fmaxf(fmaxf(tx0, ty0), tz0);
    const float tmax = fminf(fminf(tx1, ty1), tz1);
    tnear = tmin;
    return tmin <= tmax && tmax >= 0.f && tmin <= tbest;
}

// sm_100 256-bit read-only load: one 32-byte record = one L1 request per lane.  ncu on the two-LDG.128
// version showed the traversal bound by L1 tag throughput (one wavefront per distinct sector per
// instruction, profiles/r01_ncu_full_summary.json); a single LDG.E.256 halves the wavefronts.
#ifndef B200RT_LDG256
#define B200RT_LDG256 1
#endif
__device__ __forceinline__ void ldg32B(const void *ptr, uint4 &a, uint4 &b) {
#if B200RT_LDG256
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)
                 : "l"(ptr));
#else
    const uint4 *p = reinterpret_cast<const uint4 *>(ptr);
    a = __ldg(p); b = __ldg(p + 1);
#endif
}

// One traversal step: fetch a 32-byte record and test both child boxes.
__device__ __forceinline__ void pair_test(const DPair *nodes, uint32_t idx, const QRay &r, float tbest,
                                          bool &hl, bool &hr, float &tl, float &tr, uint32_t &ml, uint32_t &mr) {
    uint4 a, b;
    ldg32B(nodes + idx, a, b);
    hl = slab_q(a.x, a.y, a.z, r, tbest, tl);
    hr = slab_q(a.w, b.x, b.y, r, tbest, tr);
    ml = b.z; mr = b.w;
}

// Moeller-Trumbore on a 48-byte record with the division deferred to the accepted hits: the barycentric and
// distance tests run on the numerators against |det| (signs folded in with one XOR each), so a rejected triangle
// -- three out of four at the leaves -- costs no reciprocal.  Accepts 0 <= u, 0 <= v, u + v <= 1, 0 < t < tmax
// (intersectPlaneAndRay + the inside test of hw8/src/primitives.cpp:18-27, 85-108, up to rounding at the edges);
// det = 0 or NaN rejects.  `det` keeps its sign: det > 0 means the ray runs along the reference's normal ("inside").
// Every operation below is explicitly rounded / explicitly fused: the same bits come out of every kernel this is inlined
// into (the compiler's own FMA contraction differed between kernels) and out of oracle/pairs_traverse.c on the CPU.
__device__ __forceinline__ float dot_p(const V3 a, const V3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, __fmul_rn(a.x, b.x))); }
__device__ __forceinline__ V3 cross_p(const V3 a, const V3 b) {
    return mk(fmaf(a.y, b.z, -__fmul_rn(a.z, b.y)), fmaf(a.z, b.x, -__fmul_rn(a.x, b.z)), fmaf(a.x, b.y, -__fmul_rn(a.y, b.x)));
}
__device__ __forceinline__ bool tri_hit(const float4 v0, const float4 e1, const float4 e2, const V3 o, const V3 d, const float tmax,
                                        float &t, float &u, float &v, float &det) {
    const V3 E1 = mk(e1.x, e1.y, e1.z), E2 = mk(e2.x, e2.y, e2.z);
    const V3 pv = cross_p(d, E2);
    det = dot_p(E1, pv);
    const V3 tv = mk(__fsub_rn(o.x, v0.x), __fsub_rn(o.y, v0.y), __fsub_rn(o.z, v0.z));
    const V3 qv = cross_p(tv, E1);
    const uint32_t sb = __float_as_uint(det) & 0x80000000u;
    const float ad = fabsf(det);
    const float un = __uint_as_float(__float_as_uint(dot_p(tv, pv)) ^ sb);
    const float vn = __uint_as_float(__float_as_uint(dot_p(d, qv)) ^ sb);
    const float tn = __uint_as_float(__float_as_uint(dot_p(E2, qv)) ^ sb);
    if (!(un >= 0.f && vn >= 0.f && __fadd_rn(un, vn) <= ad && tn > 0.f && tn < __fmul_rn(tmax, ad))) return false;
    const float idet = __fdiv_rn(1.f, ad);
    t = __fmul_rn(tn, idet); u = __fmul_rn(un, idet); v = __fmul_rn(vn, idet);
    return true;
}

// Closest hit over the scene BVH.  t in (0, 1e4) like intersectPlaneAndRay (primitives.cpp:20).
template <bool COUNT>
__device__ __forceinline__ void trace_closest(const DGltfScene &sc, const V3 o, const V3 d, TraceHit &h, ThreadCounters &tc) {
    h.t = B200RT_TMAX; h.tri = -1; h.u = h.v = 0.f; h.det = 0.f;
    const QRay qr = make_qray(sc.q, o, d);
    uint32_t stack[B200RT_STACK];
    int sp = 0;
    uint32_t cur = sc.root;   // the root's own box is not tested (it is the whole scene)
    while (true) {
        if (!(cur & B200RT_LEAF)) {
            float tl, tr;
            bool hl, hr;
            uint32_t ml, mr;
            pair_test(sc.nodes, cur, qr, h.t, hl, hr, tl, tr, ml, mr);
            if (COUNT) tc.node_tests += 2;
            if (hl && hr) {
                const bool left_first = tl <= tr;
                if (sp < B200RT_STACK) stack[sp++] = left_first ? mr : ml; else tc.overflow = 1;
                cur = left_first ? ml : mr;
                continue;
            }
            if (hl) { cur = ml; continue; }
            if (hr) { cur = mr; continue; }
        } else {
            const uint32_t first = cur & 0x07FFFFFFu, count = (cur >> 27) & 15u;
            for (uint32_t i = first; i < first + count; i++) {
                const float4 *q = reinterpret_cast<const float4 *>(sc.tris + i);
                const float4 v0 = __ldg(q), e1 = __ldg(q + 1), e2 = __ldg(q + 2);
                if (COUNT) tc.tri_tests++;
                float t, u, v, det;
                if (tri_hit(v0, e1, e2, o, d, h.t, t, u, v, det)) { h.t = t; h.u = u; h.v = v; h.tri = (int)i; h.det = det; }
            }
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
}

// One emitter's term of FiguresMix::getTotalPdf for a ray that hits it at parameter t with barycentrics (u, v):
// pointProb * |x - y|^2 / |d . n_y|, n_y = the interpolated shading normal (hw8 distributions.h:140-145) or the geometric
// normal (hw7 :145).  Written ONCE with explicitly rounded / fused operations: wf_light and the megakernel's
// light_pdf_sum must add the same bits (their images are compared bit for bit).
__device__ __forceinline__ float light_term(const DGltfScene &sc, uint32_t i, const float4 v0, const float4 e1, const float4 e2,
                                            const V3 d, float t, float u, float v) {
    V3 yn;
    if (sc.hw8) {
        const float4 *nq = reinterpret_cast<const float4 *>(sc.light_normals + i);
        const float4 n0 = __ldg(nq), n1 = __ldg(nq + 1), n2 = __ldg(nq + 2);
        yn = mk(fmaf(v, __fsub_rn(n1.x, n2.x), fmaf(u, __fsub_rn(n0.x, n2.x), n2.x)),
                fmaf(v, __fsub_rn(n1.y, n2.y), fmaf(u, __fsub_rn(n0.y, n2.y), n2.y)),
                fmaf(v, __fsub_rn(n1.z, n2.z), fmaf(u, __fsub_rn(n0.z, n2.z), n2.z)));
    } else {
        yn = cross_p(mk(e2.x, e2.y, e2.z), mk(e1.x, e1.y, e1.z));
    }
    const float il = rsqrtf(dot_p(yn, yn));
    yn = mk(__fmul_rn(il, yn.x), __fmul_rn(il, yn.y), __fmul_rn(il, yn.z));
    const V3 td = mk(__fmul_rn(t, d.x), __fmul_rn(t, d.y), __fmul_rn(t, d.z));      // y - x
    return __fdiv_rn(__fmul_rn(v0.w, dot_p(td, td)), fabsf(dot_p(d, yn)));
}

// FiguresMix::getTotalPdf (distributions.h:122-165): every emissive triangle whose box the ray
// (x, d) pierces contributes pointProb * |x-y|^2 / |d . n_y|; NO distance pruning.
template <bool COUNT>
__device__ __forceinline__ float light_pdf_sum(const DGltfScene &sc, const V3 o, const V3 d, ThreadCounters &tc) {
    uint32_t stack[B200RT_STACK];
    int sp = 0;
    float total = 0.f;
    // unlike the scene root, the light root's box IS a real filter
    if (COUNT) tc.light_node_tests++;
    const uint32_t mask = light_boxes_mask(sc, o, d);
    if (!mask) return 0.f;
    const QRay qr = make_qray(sc.lq, o, d);
    uint32_t cur;
    light_start(sc, mask, stack, sp, cur);
    while (true) {
        if (!(cur & B200RT_LEAF)) {
            float tl, tr;
            bool hl, hr;
            uint32_t ml, mr;
            pair_test(sc.light_nodes, cur, qr, INFINITY, hl, hr, tl, tr, ml, mr);
            if (COUNT) tc.light_node_tests += 2;
            if (hl && hr) {
                if (sp < B200RT_STACK) stack[sp++] = mr; else tc.overflow = 1;
                cur = ml;
                continue;
            }
            if (hl) { cur = ml; continue; }
            if (hr) { cur = mr; continue; }
        } else {
            const uint32_t first = cur & 0x07FFFFFFu, count = (cur >> 27) & 15u;
            for (uint32_t i = first; i < first + count; i++) {
                const float4 *q = reinterpret_cast<const float4 *>(sc.light_tris + i);
                const float4 v0 = __ldg(q), e1 = __ldg(q + 1), e2 = __ldg(q + 2);
                if (COUNT) tc.light_tri_tests++;
                float t, u, v, det;
                if (tri_hit(v0, e1, e2, o, d, B200RT_TMAX, t, u, v, det)) total = __fadd_rn(total, light_term(sc, i, v0, e1, e2, d, t, u, v));
            }
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
    return total;
}

// ---- textures (hw8/src/scene.cpp:9-35): wrap, bilinear on the integer texel grid, sRGB decode
// per texel BEFORE blending.  Texels are 32-bit RGBA8 words in linear memory (DTexels); decode goes
// through a 256-entry table (host-computed with the reference's powf) in shared memory.
// A sample is split into a fetch (4 point taps issued back to back) and a resolve, so that the taps
// of all textures of a material are in flight together: ncu attributed half of the shade stage's
// stall samples to four serial fetch round trips executed by a handful of lanes.
struct TexTaps {
    uint32_t c11, c12, c21, c22;      // packed RGBA8 texels (one register per tap)
    float dx, dy;
};
__device__ __forceinline__ void tex_fetch(DTexels tex, int w, int h, float tx, float ty, TexTaps &k) {
    tx -= floorf(tx);
    ty -= floorf(ty);
    tx *= (float)w;
    ty *= (float)h;
    int ix1 = (int)floorf(tx), iy1 = (int)floorf(ty);
    // (i + 1) % n for i in [0, n] without the integer division
    int ix2 = ix1 + 1, iy2 = iy1 + 1;
    if (ix2 >= w) ix2 -= w;
    if (iy2 >= h) iy2 -= h;
    ix1 = min(ix1, w - 1); iy1 = min(iy1, h - 1);
    k.dx = tx - (float)ix1; k.dy = ty - (float)iy1;
    const uint32_t *r1 = tex + iy1 * w, *r2 = tex + iy2 * w;
    k.c11 = __ldg(r1 + ix1); k.c12 = __ldg(r2 + ix1);
    k.c21 = __ldg(r1 + ix2); k.c22 = __ldg(r2 + ix2);
}
__device__ __forceinline__ V3 texel(const uint32_t c, const float *lut) { return mk(lut[c & 255u], lut[(c >> 8) & 255u], lut[(c >> 16) & 255u]); }
__device__ __forceinline__ V3 tex_resolve(const TexTaps &k, const float *lut) {
    const V3 p11 = texel(k.c11, lut), p12 = texel(k.c12, lut), p21 = texel(k.c21, lut), p22 = texel(k.c22, lut);
    return (1.f - k.dx) * ((1.f - k.dy) * p11 + k.dy * p12) + k.dx * ((1.f - k.dy) * p21 + k.dy * p22);
}
__device__ __forceinline__ V3 sample_texture(DTexels tex, int w, int h, float tx, float ty, const float *lut) {
    TexTaps k;
    tex_fetch(tex, w, h, tx, ty, k);
    return tex_resolve(k, lut);
}

// ---- BRDF (hw8/src/include/material.h:11-65) ----
__device__ __forceinline__ float ggx_d(float hn, float a2) {
    if (hn <= 0.f) return 0.f;
    const float k = fmaxf(0.f, (a2 - 1.f) * hn * hn + 1.f);
    return a2 / (3.14159265358979f * k * k);
}
__device__ __forceinline__ float smith_v1(float nx, float a2) {
    return 1.f / (fabsf(nx) + sqrtf(fmaxf(0.f, a2 + (1.f - a2) * nx * nx)));
}
__device__ __forceinline__ float pow5(float x) { const float x2 = x * x; return x2 * x2 * x; }

template <bool HW8>
__device__ __forceinline__ V3 eval_brdf(const V3 base, float metallic_factor, const V3 l, const V3 v, const V3 n,
                                        const V3 color, float metallic, float alpha) {
    const V3 h = normalize(l + v);
    const float a2 = alpha * alpha;
    float specular = 0.f;
    if (!(dot(h, l) < 1e-4f || dot(h, v) < 1e-4f)) specular = ggx_d(dot(h, n), a2) * smith_v1(dot(n, l), a2) * smith_v1(dot(n, v), a2);
    const float fw = pow5(fmaxf(0.f, 1.f - fabsf(dot(v, h))));
    V3 metal = mk(0, 0, 0), diel = mk(0, 0, 0);
    const V3 albedo = HW8 ? base * color : base;
    metallic = HW8 ? metallic * metallic_factor : metallic_factor;
    const float ln = dot(l, n), vn = dot(v, n);
    if (metallic > 0.f && (!HW8 || (vn >= 0.f && ln >= 0.f))) {
        const V3 ft = albedo + fw * (mk(1, 1, 1) - albedo);
        metal = specular * ft;
    }
    if (metallic < 1.f) {
        V3 diffuse = mk(0, 0, 0);
        if (!HW8 || ln >= 0.f) diffuse = 0.318309886183791f * albedo;
        const float f = 0.04f + fw * 0.96f;
        diel = (1.f - f) * diffuse + mk(specular * f, specular * f, specular * f);
    }
    return (1.f - metallic) * diel + metallic * metal;
}

// ---- VNDF (Heitz 2018) in the frame that maps the shading normal to +z
// (hw8/src/include/distributions.h:168-246).  The reference builds the frame with a quaternion
// (getQ); any rotation taking n to +z gives the same distribution, but the port comparison wants
// the same frame, so the same half-angle quaternion is used here.
struct Frame {
    float qx, qy, qz, qw;   // Hamilton quaternion q such that rotate(q, n) = +z
};
__device__ __forceinline__ Frame frame_of(const V3 n) {
    Frame f;
    if (n.z > 0.9999f) { f.qx = f.qy = f.qz = 0.f; f.qw = 1.f; return f; }
    if (n.z < -0.9999f) { f.qx = f.qy = f.qz = 0.f; f.qw = -1.f; return f; }
    // reference: a = n.cross(newN) (= newN x n in standard terms), w = |n| + n.z, normalised;
    // its transform() is the inverse rotation of the Hamilton convention => conjugate axis.
    const V3 a = cross(mk(0, 0, 1), n);
    const float w = sqrtf(len2(n)) + n.z;
    const float il = rsqrtf(len2(a) + w * w);
    f.qx = -a.x * il; f.qy = -a.y * il; f.qz = -a.z * il; f.qw = w * il;
    return f;
}
__device__ __forceinline__ V3 rotate(float qx, float qy, float qz, float qw, const V3 p) {
    // p' = p + 2 w (q x p) + 2 q x (q x p)
    const V3 q = mk(qx, qy, qz);
    const V3 t = 2.f * cross(q, p);
    return p + qw * t + cross(q, t);
}
__device__ __forceinline__ V3 to_local(const Frame &f, const V3 p) { return rotate(f.qx, f.qy, f.qz, f.qw, p); }
__device__ __forceinline__ V3 to_world(const Frame &f, const V3 p) { return rotate(-f.qx, -f.qy, -f.qz, f.qw, p); }

__device__ __forceinline__ V3 vndf_sample_local(const V3 v, float a, float u1, float u2) {
    const V3 vh = normalize(mk(a * v.x, a * v.y, v.z));
    const float lensq = vh.x * vh.x + vh.y * vh.y;
    const V3 T1 = lensq > 0.f ? rsqrtf(lensq) * mk(-vh.y, vh.x, 0.f) : mk(1, 0, 0);
    const V3 T2 = cross(vh, T1);   // reference: T1.cross(vh) with the negated cross
    const float r = sqrtf(u1);
    float sn, cs;
    sincosf(6.283185307179586f * u2, &sn, &cs);
    const float t1 = r * cs;
    float t2 = r * sn;
    const float s = 0.5f * (1.f + vh.z);
    t2 = (1.f - s) * sqrtf(1.f - t1 * t1) + s * t2;
    const V3 nh = t1 * T1 + t2 * T2 + sqrtf(fmaxf(0.f, 1.f - t1 * t1 - t2 * t2)) * vh;
    const V3 ne = normalize(mk(a * nh.x, a * nh.y, fmaxf(0.f, nh.z)));
    return 2.f * dot(ne, v) * ne - v;
}
__device__ __forceinline__ float vndf_pdf_local(const V3 d, const V3 v, float a) {
    const V3 ni = normalize(v + d);
    const float a2 = a * a;
    const float k = ni.x * ni.x / a2 + ni.y * ni.y / a2 + ni.z * ni.z;
    const float D = 1.f / (3.14159265358979f * a2 * k * k);
    const float lambda = 0.5f * (-1.f + sqrtf(1.f + (a2 * v.x * v.x + a2 * v.y * v.y) / (v.z * v.z)));
    const float G1 = 1.f / (1.f + lambda);
    const float vni = dot(v, ni);
    const float dv = G1 * fmaxf(0.f, vni) * D / fabsf(v.z);
    return dv / (4.f * vni);
}

}  // namespace b200rt

namespace b200rt {

// One bounce of Scene::getColor (hw8/src/scene.cpp:84-165), split in two so that the light-BVH
// query between them can run either inline (megakernel) or as its own wavefront stage.
struct BounceOut {
    V3 xo, nd, brdf;        // next ray origin / direction, BRDF value
    float pdf_partial;      // cosine pdf + VNDF pdf
    float absdot;           // |nd . n_shading|
};

// The per-triangle records shade_surface reads: intersection record (geometric normal, material id, tangent sign)
// and shading record (normals, uv, tangents).  The megakernel fetches them from global memory; the wavefront's
// shade stage has them staged in shared memory one item ahead (wavefront.cuh).
struct SurfRec {
    float4 tv0, te1, te2;
    float4 s0, s1, s2, s3, s4, s5;     // s3..s5 only in hw8 mode
};
template <bool HW8>
__device__ __forceinline__ void load_surf_rec(const DGltfScene &sc, int tri, SurfRec &r) {
    const float4 *tq = reinterpret_cast<const float4 *>(sc.tris + tri);
    r.tv0 = __ldg(tq); r.te1 = __ldg(tq + 1); r.te2 = __ldg(tq + 2);
    const float4 *sq = reinterpret_cast<const float4 *>(sc.shade + tri);
    r.s0 = __ldg(sq); r.s1 = __ldg(sq + 1); r.s2 = __ldg(sq + 2);
    if (HW8) { r.s3 = __ldg(sq + 3); r.s4 = __ldg(sq + 4); r.s5 = __ldg(sq + 5); }
}

// Surface part: fetch the shading record, material and textures, add the emission to L, draw the
// next direction from Mix{Cosine, Vndf, [lights]} and evaluate the BRDF.  Returns false when the
// path ends here (BRDF below 1e-4 in every channel, scene.cpp:154-156).
template <bool HW8, bool COUNT>
__device__ __forceinline__ bool shade_surface(const DGltfScene &sc, const float *lut_srgb, const float *lut_lin,
                                              const V3 o, const V3 d, float t, float u, float v, const SurfRec &rec, bool inside,
                                              V3 &L, const V3 T, const uint4 rnd, ThreadCounters &tc, BounceOut &out) {
    const float4 tv0 = rec.tv0, te1 = rec.te1, te2 = rec.te2;
    if (COUNT) tc.shaded++;

    const float4 s0 = rec.s0, s1 = rec.s1, s2 = rec.s2;
    V3 ns = normalize(mk(s2.x + u * (s0.x - s2.x) + v * (s1.x - s2.x),
                         s2.y + u * (s0.y - s2.y) + v * (s1.y - s2.y),
                         s2.z + u * (s0.z - s2.z) + v * (s1.z - s2.z)));
    if (inside) ns = -ns;
    V3 ng = normalize(cross(mk(te2.x, te2.y, te2.z), mk(te1.x, te1.y, te1.z)));
    if (inside) ng = -ng;

    const DMaterial *m = sc.materials + __float_as_int(tv0.w);
    V3 emission = mk(__ldg(&m->emission[0]), __ldg(&m->emission[1]), __ldg(&m->emission[2]));
    const V3 base = mk(__ldg(&m->base_color[0]), __ldg(&m->base_color[1]), __ldg(&m->base_color[2]));
    const float metallic_factor = __ldg(&m->metallic_factor);
    const float roughness_factor = __ldg(&m->roughness_factor);
    V3 color = mk(1, 1, 1);
    float alpha, metallic = 1.f;
    if (HW8) {
        const float4 s3 = rec.s3, s4 = rec.s4, s5 = rec.s5;
        // uv: data = (s0.w, s1.w), data2 = (s2.w, s3.w), data3 = (s4.w, s5.w)
        const float tx = s4.w + u * (s0.w - s4.w) + v * (s2.w - s4.w);
        const float ty = s5.w + u * (s1.w - s5.w) + v * (s3.w - s5.w);
        const DTexels tb = m->tex_base, te = m->tex_emissive, tm = m->tex_mr, tn = m->tex_normal;
        TexTaps kb, ke, km, kn;
        if (tb) tex_fetch(tb, m->w_base, m->h_base, tx, ty, kb);
        if (te) tex_fetch(te, m->w_emissive, m->h_emissive, tx, ty, ke);
        if (tm) tex_fetch(tm, m->w_mr, m->h_mr, tx, ty, km);
        if (tn) tex_fetch(tn, m->w_normal, m->h_normal, tx, ty, kn);
        if (tb) { color = tex_resolve(kb, lut_srgb); if (COUNT) tc.tex++; }
        if (te) { emission = emission * tex_resolve(ke, lut_srgb); if (COUNT) tc.tex++; }
        V3 mr = mk(1, 1, 1);
        if (tm) { mr = tex_resolve(km, lut_lin); if (COUNT) tc.tex++; }
        if (tn) {
            // applyNormalMaps (scene.cpp:37-52): X = tangent, Z = shading normal, Y = w * X.cross(Z) = w * (Z x X)
            const V3 smp = tex_resolve(kn, lut_lin);
            if (COUNT) tc.tex++;
            const V3 X = normalize(mk(s5.x + u * (s3.x - s5.x) + v * (s4.x - s5.x),
                                      s5.y + u * (s3.y - s5.y) + v * (s4.y - s5.y),
                                      s5.z + u * (s3.z - s5.z) + v * (s4.z - s5.z)));
            const V3 Y = te1.w * cross(ns, X);
            const V3 ln = mk(2.f * smp.x - 1.f, 2.f * smp.y - 1.f, 2.f * smp.z - 1.f);
            ns = normalize(ln.x * X + ln.y * Y + ln.z * ns);
        }
        const float rgh = fmaxf(0.08f, roughness_factor * mr.y);
        alpha = rgh * rgh;
        metallic = mr.z;
    } else {
        alpha = roughness_factor * roughness_factor;
    }

    const V3 x = o + t * d;
    const V3 xo = x + B200RT_EPS * ng;     // origin offset along the GEOMETRIC normal (scene.cpp:151-152)
    const V3 vdir = -d;                    // towards the viewer

    // Mix::sample (distributions.h:256-265): uniform choice among {Cosine, Vndf, [lights]}
    const int k = sc.n_light_tris > 0 ? 3 : 2;
    const int which = (int)(u01(rnd.x) * (float)k);
    const float r1 = u01(rnd.y), r2 = u01(rnd.z), r3 = u01(rnd.w);
    const Frame fr = frame_of(ns);
    const V3 vloc = to_local(fr, vdir);
    V3 nd;
    if (which == 0) {
        V3 c = sphere_dir(r1, r2) + ns;
        const float l = sqrtf(len2(c));
        nd = (l <= 1e-9f || dot(c, ns) <= 1e-9f || isnan(l)) ? ns : (1.f / l) * c;
    } else if (which == 2) {
        // FiguresMix::sample (:117-120) + TriangleLight::sample (:81-94): draws = index, u, v
        const int li = min((int)(r1 * (float)sc.n_light_tris), sc.n_light_tris - 1);
        const float4 *lq = reinterpret_cast<const float4 *>(sc.light_sample_tris + li);
        const float4 a = __ldg(lq), b = __ldg(lq + 1), c = __ldg(lq + 2);
        float lu = r2, lv = r3;
        if (lu + lv > 1.f) { lu = 1.f - lu; lv = 1.f - lv; }
        const V3 point = mk(a.x + lu * b.x + lv * c.x, a.y + lu * b.y + lv * c.y, a.z + lu * b.z + lv * c.z);
        nd = normalize(point - xo);
    } else {
        nd = to_world(fr, vndf_sample_local(vloc, alpha, r1, r2));
    }

    L = L + T * emission;
    const V3 brdf = eval_brdf<HW8>(base, metallic_factor, nd, vdir, ns, color, metallic, alpha);
    if (brdf.x < B200RT_EPS && brdf.y < B200RT_EPS && brdf.z < B200RT_EPS) return false;

    // first two terms of Mix::pdf (:267-279)
    float pdf = fmaxf(0.f, dot(nd, ns) * 0.318309886183791f);
    pdf += vndf_pdf_local(to_local(fr, nd), vloc, alpha);
    out.xo = xo; out.nd = nd; out.brdf = brdf; out.pdf_partial = pdf; out.absdot = fabsf(dot(nd, ns));
    return true;
}

// The emission term of shade_surface alone (same arithmetic, same bits).  The last bounce of a path needs nothing
// else: Scene::getColor(ray, 1) returns emission + weight * getColor(next, 0) and getColor(.., 0) is black
// (hw8/src/scene.cpp:85-87), so the sampled direction, the BRDF and three of the four texture lookups are dead work.
template <bool HW8, bool COUNT>
__device__ __forceinline__ V3 surface_emission(const DGltfScene &sc, const float *lut_srgb, int tri, float u, float v, ThreadCounters &tc) {
    if (COUNT) tc.shaded++;
    const float mat = __ldg(&sc.tris[tri].v0.w);
    const DMaterial *m = sc.materials + __float_as_int(mat);
    V3 emission = mk(__ldg(&m->emission[0]), __ldg(&m->emission[1]), __ldg(&m->emission[2]));
    if (HW8) {
        const DTexels te = m->tex_emissive;
        if (te) {
            const float4 *sq = reinterpret_cast<const float4 *>(sc.shade + tri);
            const float s0w = __ldg(&sq[0].w), s1w = __ldg(&sq[1].w), s2w = __ldg(&sq[2].w), s3w = __ldg(&sq[3].w), s4w = __ldg(&sq[4].w), s5w = __ldg(&sq[5].w);
            const float tx = s4w + u * (s0w - s4w) + v * (s2w - s4w);
            const float ty = s5w + u * (s1w - s5w) + v * (s3w - s5w);
            TexTaps ke;
            tex_fetch(te, m->w_emissive, m->h_emissive, tx, ty, ke);
            emission = emission * tex_resolve(ke, lut_srgb);
            if (COUNT) tc.tex++;
        }
    }
    return emission;
}

// Second part: mean of the component pdfs, path weight, firefly clamp (scene.cpp:158-163).
// light_sum = FiguresMix::getTotalPdf for the ray (xo, nd); ignored when there are no emitters.
__device__ __forceinline__ bool finish_bounce(const V3 brdf, float pdf_partial, float absdot, float light_sum,
                                              int n_light_tris, V3 &T, const float clamp_at) {
    float pdf = pdf_partial;
    float k = 2.f;
    if (n_light_tris > 0) { pdf += light_sum / (float)n_light_tris; k = 3.f; }
    pdf /= k;
    const float w = (1.f / pdf) * absdot;
    const V3 mult = w * brdf;
    // any channel > 6 or NaN ends the path (the reference's DIRTY_DENOISE_HACK_BUBEN, hw8/src/scene.cpp:7,161: biased,
    // darkens).  clamp_at = DGltfScene::weight_clamp: 6 by default; B200RT_FLAG_NO_FIREFLY_CLAMP raises it to FLT_MAX,
    // which leaves the one-sample MIS estimator of Mix::pdf unbiased and still ends paths on NaN / inf weights.
    if (!(mult.x <= clamp_at) || !(mult.y <= clamp_at) || !(mult.z <= clamp_at)) return false;
    T = T * mult;
    return true;
}

template <bool HW8, bool COUNT>
__device__ __forceinline__ bool shade_bounce(const DGltfScene &sc, const float *lut_srgb, const float *lut_lin,
                                             const TraceHit &hit, V3 &o, V3 &d, V3 &L, V3 &T,
                                             const uint4 rnd, ThreadCounters &tc) {
    BounceOut bo;
    SurfRec rec;
    load_surf_rec<HW8>(sc, hit.tri, rec);
    if (!shade_surface<HW8, COUNT>(sc, lut_srgb, lut_lin, o, d, hit.t, hit.u, hit.v, rec, hit.det > 0.f, L, T, rnd, tc, bo)) return false;
    float ls = 0.f;
    if (sc.n_light_tris > 0) {
        tc.light_queries++;
        ls = light_pdf_sum<COUNT>(sc, bo.xo, bo.nd, tc);
    }
    if (!finish_bounce(bo.brdf, bo.pdf_partial, bo.absdot, ls, sc.n_light_tris, T, sc.weight_clamp)) return false;
    o = bo.xo;
    d = bo.nd;
    return true;
}

__device__ __forceinline__ void flush_counters(DCounters *g, const ThreadCounters &tc) {
    unsigned vals[10] = {tc.paths, tc.rays, tc.light_queries, tc.node_tests, tc.tri_tests,
                         tc.light_node_tests, tc.light_tri_tests, tc.shaded, tc.tex, tc.overflow};
    unsigned long long *dst = reinterpret_cast<unsigned long long *>(g);
#pragma unroll
    for (int i = 0; i < 10; i++) {
        unsigned v = vals[i];
        // 32-bit per-thread counts: at most ~1e5 per thread, the warp sum fits in 32 bits
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(dst + i, (unsigned long long)v);
    }
}

// Work decomposition: the image is cut into 8x4-pixel tiles, one warp per (tile, sample chunk),
// so the 32 primary rays of a warp are coherent.  A thread renders samples [s0, s1) of its pixel
// into slab `chunk` of the accumulation buffer; fold_slabs() adds the slabs in a fixed order
// (deterministic, no float atomics).
template <bool HW8, bool COUNT>
__global__ void __launch_bounds__(128) gltf_path_kernel(const DGltfScene sc, const DRenderParams rp,
                                                        const float *__restrict__ lut_srgb_g) {
    __shared__ float lut_srgb[256];
    __shared__ float lut_lin[256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        lut_srgb[i] = lut_srgb_g[i];
        lut_lin[i] = 0.00392156862745098f * (float)i;   // (1./255) * c as float (scene.cpp:11)
    }
    __syncthreads();

    const int W = rp.cam.width, H = rp.cam.height;
    const int tiles_x = (W + 7) >> 3, tiles_y = (H + 3) >> 2;
    const long long n_tiles = (long long)tiles_x * tiles_y;
    const long long warp_id = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    ThreadCounters tc = {};
    if (warp_id < n_tiles * rp.n_chunks) {
        const int chunk = (int)(warp_id / n_tiles);
        const long long tile = warp_id % n_tiles;
        const int px = (int)(tile % tiles_x) * 8 + (lane & 7);
        const int py = (int)(tile / tiles_x) * 4 + (lane >> 3);
        if (px < W && py < H) {
            const int total = rp.spp_end - rp.spp_begin;
            const int per = (total + rp.n_chunks - 1) / rp.n_chunks;
            const int s0 = rp.spp_begin + chunk * per;
            const int s1 = min(rp.spp_end, s0 + per);
            const uint32_t pixel = (uint32_t)(py * W + px);
            const V3 cpos = ld3(rp.cam.pos), cright = ld3(rp.cam.right), cup = ld3(rp.cam.up), cfwd = ld3(rp.cam.forward);
            V3 sum = mk(0, 0, 0);
            for (int s = s0; s < s1; s++) {
                // camera sample (scene.cpp:172-173, 179-187): jitter in [0,1)^2, fovY-based, normalised
                const uint4 j = philox4x32_10((uint32_t)s, 0u, pixel, rp.seed);
                const float fx = (float)px + u01(j.x), fy = (float)py + u01(j.y);
                const float nx = rp.cam.tan_x * (2.f * fx / (float)W - 1.f);
                const float ny = rp.cam.tan_y * (2.f * fy / (float)H - 1.f);
                V3 o = cpos;
                V3 d = normalize(nx * cright - ny * cup + cfwd);
                V3 L = mk(0, 0, 0), T = mk(1, 1, 1);
                tc.paths++;
                for (int bounce = 0; bounce < rp.ray_depth; bounce++) {
                    TraceHit hit;
                    tc.rays++;
                    trace_closest<COUNT>(sc, o, d, hit, tc);
                    if (hit.tri < 0) {
                        V3 bg = ld3(sc.bg);
                        if (HW8 && sc.env) {
                            // equirect lookup (scene.cpp:94-96), decoded as sRGB
                            const float tx = 0.5f + 0.5f * atan2f(d.z, d.x) * 0.318309886183791f;
                            const float ty = 0.5f - asinf(d.y) * 0.318309886183791f;
                            bg = sample_texture(sc.env, sc.env_w, sc.env_h, tx, ty, lut_srgb);
                        }
                        L = L + T * bg;
                        break;
                    }
                    if (bounce == rp.ray_depth - 1) {      // final bounce: emission only (see surface_emission)
                        L = L + T * surface_emission<HW8, COUNT>(sc, lut_srgb, hit.tri, hit.u, hit.v, tc);
                        break;
                    }
                    const uint4 rnd = philox4x32_10((uint32_t)s, 16u * (uint32_t)(bounce + 1), pixel, rp.seed);
                    if (!shade_bounce<HW8, COUNT>(sc, lut_srgb, lut_lin, hit, o, d, L, T, rnd, tc)) break;
                }
                sum = sum + L;
            }
            float *dst = rp.accum + ((size_t)chunk * W * H + pixel) * 3;
            dst[0] = sum.x; dst[1] = sum.y; dst[2] = sum.z;
        }
    }
    flush_counters(rp.counters, tc);
}

}  // namespace b200rt
