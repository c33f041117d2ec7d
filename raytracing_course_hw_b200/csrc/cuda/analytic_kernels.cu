// K1 (hw1 primary cast), K2 (hw2 Whitted), K3 (hw3 / hw4 analytic path tracers), K7 (hw5 mixed figures) for sm_100a.
// This file is compiled TWICE (build.py):
//   * with -fmad=false: K1, K2 and the scene upload.  Every float operation is written in the reference's own order
//     and with its float/double promotions, so that the deterministic configs (hw1, hw2) reproduce primary-hit
//     indices and radiance bit for bit (north_star);
//   * with -DB200RT_ANALYTIC_MC_UNIT and the compiler's default FMA contraction: K3 and K7, the Monte-Carlo modes,
//     whose parity is statistical -- a third of their arithmetic fuses.  Each unit has its own __constant__ copy of
//     the scene; analytic_upload fills both.
//   hw1: hw1/src/scene.cpp:7-30, hw1/src/primitives.cpp:24-91
//   hw2: hw2/src/scene.cpp:8-99, hw2/src/primitives.cpp:25-137, hw2/src/light_source.cpp:9-23
//   hw4: hw4/src/scene.cpp:10-128, hw4/src/include/distributions.h:46-205, hw4/src/primitives.cpp:9-118
// Primitives and lights live in __constant__ memory (<= 64 + 16 entries, warp-uniform reads).
#include "analytic_kernels.h"
#include "rt_math.cuh"
#include "gltf_kernel.cuh"      // hw5: the triangles of a text scene go through the glTF modes' BVH traversal

namespace b200rt {

static __constant__ DAnalyticScene g_scene;

#ifdef B200RT_ANALYTIC_MC_UNIT
cudaError_t analytic_mc_upload(const DAnalyticScene &staged, cudaStream_t stream) {
    return cudaMemcpyToSymbolAsync(g_scene, &staged, sizeof(staged), 0, cudaMemcpyHostToDevice, stream);
}
#else
cudaError_t analytic_mc_upload(const DAnalyticScene &staged, cudaStream_t stream);

namespace {
// Host restatement of Quaternion::operator* / transform with the reference's operation order (negated cross,
// hw4/src/include/quaternion.h:36-48) -- the same float operations the kernels' qtransform performs.
struct HQ { float x, y, z, w; };
inline HQ hq_mul(const HQ &a, const HQ &b) {
    return {(a.w * b.x + b.w * a.x) + (a.z * b.y - a.y * b.z), (a.w * b.y + b.w * a.y) + (a.x * b.z - a.z * b.x),
            (a.w * b.z + b.w * a.z) + (a.y * b.x - a.x * b.y), a.w * b.w - (a.x * b.x + a.y * b.y + a.z * b.z)};
}
}  // namespace

cudaError_t analytic_upload(const DAnalyticScene &scene, cudaStream_t stream) {
    static DAnalyticScene staged;       // a context is driven by one thread; pageable source: staged during the call
    staged = scene;
    for (int k = 0; k < staged.n_prims && k < B200RT_MAX_PRIMS; k++) {
        const float *r = staged.prims[k].rotation;
        const HQ q{r[0], r[1], r[2], r[3]}, qc{-1.f * r[0], -1.f * r[1], -1.f * r[2], r[3]};
        float *m = staged.frame[k];
        for (int c = 0; c < 3; c++) {       // column c = image of basis vector e_c
            const HQ e{c == 0 ? 1.f : 0.f, c == 1 ? 1.f : 0.f, c == 2 ? 1.f : 0.f, 0.f};
            const HQ t = hq_mul(hq_mul(q, e), qc);
            m[0 + c] = t.x; m[3 + c] = t.y; m[6 + c] = t.z;
        }
        const bool ident = r[0] == 0.f && r[1] == 0.f && r[2] == 0.f && r[3] == 1.f;
        m[9] = ident ? 1.f : 0.f; m[10] = m[11] = m[15] = 0.f;
        for (int c = 0; c < 3; c++) m[12 + c] = 1.f / staged.prims[k].data[c];
    }
    const cudaError_t e = cudaMemcpyToSymbolAsync(g_scene, &staged, sizeof(staged), 0, cudaMemcpyHostToDevice, stream);
    return e != cudaSuccess ? e : analytic_mc_upload(staged, stream);
}
#endif

namespace {

// ---- reference-order float helpers --------------------------------------------------------
// Vec3::normalize = float(1.0 / double(len)) * v with len = sqrt(x*x + y*y + z*z) (vec3.h:77-79)
__device__ __forceinline__ V3 rnorm(V3 a) {
    const float k = (float)(1.0 / (double)sqrtf(dot(a, a)));
    return k * a;
}
// Vec3::cross of the reference = b x a in standard terms, written out in its own order.
__device__ __forceinline__ V3 rcross(V3 a, V3 b) { return mk(a.z * b.y - a.y * b.z, a.x * b.z - a.z * b.x, a.y * b.x - a.x * b.y); }

struct Quat { V3 v; float w; };
__device__ __forceinline__ Quat qmul(const Quat &a, const Quat &b) {
    Quat r;
    r.v = (a.w * b.v + b.w * a.v) + rcross(a.v, b.v);
    r.w = a.w * b.w - dot(a.v, b.v);
    return r;
}
__device__ __forceinline__ Quat qconj(const Quat &q) { return Quat{-1.f * q.v, q.w}; }
__device__ __forceinline__ V3 qtransform(const Quat &q, V3 p) { return qmul(qmul(q, Quat{p, 0.f}), qconj(q)).v; }

struct AHit { float t; V3 n; bool inside; };

__device__ __forceinline__ bool quadratic(float a, float b, float c, float &t, bool &inside) {
    const float d = b * b - 4 * a * c;
    if (d <= 0) return false;
    float x1 = (-b - sqrtf(d)) / (2 * a);
    float x2 = (-b + sqrtf(d)) / (2 * a);
    if (x1 > x2) { const float s = x1; x1 = x2; x2 = s; }
    if (x2 < 0) return false;
    if (x1 < 0) { t = x2; inside = true; } else { t = x1; inside = false; }
    return true;
}

// std::min(a, b) = (b < a) ? b : a ; std::max(a, b) = (a < b) ? b : a  (NaN behaviour)
__device__ __forceinline__ float smin(float a, float b) { return (b < a) ? b : a; }
__device__ __forceinline__ float smax(float a, float b) { return (a < b) ? b : a; }

__device__ __forceinline__ bool box_slabs(V3 s, V3 o, V3 d, float &t, bool &inside) {
    const V3 ts1 = (-1.f * s - o) / d;
    const V3 ts2 = (s - o) / d;
    const float t1x = smin(ts1.x, ts2.x), t2x = smax(ts1.x, ts2.x);
    const float t1y = smin(ts1.y, ts2.y), t2y = smax(ts1.y, ts2.y);
    const float t1z = smin(ts1.z, ts2.z), t2z = smax(ts1.z, ts2.z);
    const float t1 = smax(smax(t1x, t1y), t1z);
    const float t2 = smin(smin(t2x, t2y), t2z);
    if (t1 > t2 || t2 < 0) return false;
    if (t1 < 0) { inside = true; t = t2; } else { inside = false; t = t1; }
    return true;
}

// Monte-Carlo modes (FAST): the same tests with one approximate reciprocal (MUFU.RCP, 1 ulp) where the reference
// divides -- an IEEE division is ~10 instructions, and an ellipsoid test had eight of them.  Their parity is
// statistical; hw1 / hw2 never come here.
__device__ __forceinline__ float rcp_fast(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float sqrt_fast(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ bool quadratic_fast(float a, float b, float c, float &t, bool &inside) {
    const float d = b * b - 4 * a * c;
    if (d <= 0) return false;
    const float sq = sqrt_fast(d), inv = rcp_fast(2 * a);
    float x1 = (-b - sq) * inv, x2 = (-b + sq) * inv;
    if (x1 > x2) { const float s = x1; x1 = x2; x2 = s; }
    if (x2 < 0) return false;
    if (x1 < 0) { t = x2; inside = true; } else { t = x1; inside = false; }
    return true;
}
__device__ __forceinline__ bool box_slabs_fast(V3 s, V3 o, V3 d, float &t, bool &inside) {
    const V3 inv = mk(rcp_fast(d.x), rcp_fast(d.y), rcp_fast(d.z));
    const V3 ts1 = (-1.f * s - o) * inv;
    const V3 ts2 = (s - o) * inv;
    const float t1x = smin(ts1.x, ts2.x), t2x = smax(ts1.x, ts2.x);
    const float t1y = smin(ts1.y, ts2.y), t2y = smax(ts1.y, ts2.y);
    const float t1z = smin(ts1.z, ts2.z), t2z = smax(ts1.z, ts2.z);
    const float t1 = smax(smax(t1x, t1y), t1z);
    const float t2 = smin(smin(t2x, t2y), t2z);
    if (t1 > t2 || t2 < 0) return false;
    if (t1 < 0) { inside = true; t = t2; } else { inside = false; t = t1; }
    return true;
}

// World -> local of a ray for figure k.  FAST (Monte-Carlo modes): the precomputed 3x3 frame with explicit fused
// multiply-adds (this file is compiled -fmad=false for hw1 / hw2), nothing at all for identity rotations; otherwise the
// reference's two quaternion sandwiches in its own operation order (hw1 / hw2: bit-exact images).
__device__ __forceinline__ V3 frame_apply(const float *m, V3 p) {
    return mk(fmaf(m[2], p.z, fmaf(m[1], p.y, m[0] * p.x)), fmaf(m[5], p.z, fmaf(m[4], p.y, m[3] * p.x)), fmaf(m[8], p.z, fmaf(m[7], p.y, m[6] * p.x)));
}
__device__ __forceinline__ V3 frame_apply_t(const float *m, V3 p) {      // transpose: local -> world (up to the |q|^2 scale)
    return mk(fmaf(m[6], p.z, fmaf(m[3], p.y, m[0] * p.x)), fmaf(m[7], p.z, fmaf(m[4], p.y, m[1] * p.x)), fmaf(m[8], p.z, fmaf(m[5], p.y, m[2] * p.x)));
}
template <bool FAST>
__device__ __forceinline__ void to_local(int k, const b200rt_primitive &f, V3 o, V3 d, V3 &lo, V3 &ld) {
    if (FAST) {
        const float *m = g_scene.frame[k];
        const V3 p = o - ld3(f.position);
        if (m[9] != 0.f) { lo = p; ld = d; }
        else { lo = frame_apply(m, p); ld = frame_apply(m, d); }
    } else {
        const Quat q{ld3(f.rotation), f.rotation[3]};
        lo = qtransform(q, o - ld3(f.position));
        ld = qtransform(q, d);
    }
}
template <bool FAST>
__device__ __forceinline__ V3 normal_to_world(int k, const b200rt_primitive &f, V3 n) {
    if (FAST) {
        const float *m = g_scene.frame[k];
        const V3 w = m[9] != 0.f ? n : frame_apply_t(m, n);
        return rsqrtf(fmaf(w.z, w.z, fmaf(w.y, w.y, w.x * w.x))) * w;
    }
    const Quat q{ld3(f.rotation), f.rotation[3]};
    return rnorm(qtransform(qconj(q), n));
}

// Figure::intersect of hw1..hw5 for figure k of the constant-memory scene.  WANT_N = false skips the normal (hw1,
// shadow rays, and the candidates of the Monte-Carlo modes: there only the WINNER's normal is computed).
template <bool WANT_N, bool FAST>
__device__ __forceinline__ bool hit_prim(int k, V3 o, V3 d, bool plane_tmax, AHit &h) {
    const b200rt_primitive &f = g_scene.prims[k];
    V3 lo, ld;
    to_local<FAST>(k, f, o, d, lo, ld);
    V3 n = mk(0, 0, 0);
    if (f.type == B200RT_PRIM_ELLIPSOID) {
        if constexpr (FAST) {
            const V3 ir = ld3(g_scene.frame[k] + 12);
            const V3 orr = lo * ir, dr = ld * ir;
            if (!quadratic_fast(dot(dr, dr), 2.f * dot(orr, dr), dot(orr, orr) - 1, h.t, h.inside)) return false;
            if (WANT_N) {
                n = (lo + h.t * ld) * (ir * ir);      // (normalised once, in world space)
                if (h.inside) n = -1.f * n;
            }
        } else {
        const V3 r = ld3(f.data);
        const V3 orr = lo / r, dr = ld / r;
        const float c = dot(orr, orr) - 1;
        const float b = 2.f * dot(orr, dr);
        const float a = dot(dr, dr);
        if (!quadratic(a, b, c, h.t, h.inside)) return false;
        if (WANT_N) {
            const V3 point = lo + h.t * ld;
            n = point / (r * r);
            if (h.inside) n = -1.f * n;
            n = rnorm(n);
        }
        }
    } else if (f.type == B200RT_PRIM_PLANE) {
        const V3 pn = ld3(f.data);
        const float t = FAST ? -dot(lo, pn) * rcp_fast(dot(ld, pn)) : -dot(lo, pn) / dot(ld, pn);
        if (!(t > 0 && (!plane_tmax || t < 1e4f))) return false;
        h.t = t;
        if (dot(ld, pn) > 0) { n = -1.f * pn; h.inside = true; } else { n = pn; h.inside = false; }
    } else {
        const V3 s = ld3(f.data);
        if (!(FAST ? box_slabs_fast(s, lo, ld, h.t, h.inside) : box_slabs(s, lo, ld, h.t, h.inside))) return false;
        if (WANT_N) {
            const V3 p = lo + h.t * ld;
            n = FAST ? p * ld3(g_scene.frame[k] + 12) : p / s;
            const float mx = smax(smax(fabsf(n.x), fabsf(n.y)), fabsf(n.z));
            if (fabsf(n.x) != mx) n.x = 0;
            if (fabsf(n.y) != mx) n.y = 0;
            if (fabsf(n.z) != mx) n.z = 0;
            if (h.inside) n = -1.f * n;
        }
    }
    if (WANT_N) h.n = normal_to_world<FAST>(k, f, n);
    return true;
}

#ifndef B200RT_ANALYTIC_MC_UNIT
// Camera of hw1/hw2 (pixel centre +0.5, evaluated in double: hw1/src/scene.cpp:26-27).
__device__ __forceinline__ V3 camera_dir_center(const DCamera &c, int x, int y) {
    const float nx = (float)((double)c.tan_x * (2 * (x + 0.5) / c.width - 1));
    const float ny = (float)((double)c.tan_y * (2 * (y + 0.5) / c.height - 1));
    return (nx * ld3(c.right) - ny * ld3(c.up)) + ld3(c.forward);
}

// ---- K1 ------------------------------------------------------------------------------------
__global__ void cast_kernel(const DCamera cam, float *__restrict__ out, int32_t *__restrict__ ids) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cam.width * cam.height) return;
    const int x = i % cam.width, y = i / cam.width;
    const V3 o = ld3(cam.pos), d = camera_dir_center(cam, x, y);
    V3 ans = ld3(g_scene.bg);
    float best = -1;
    int id = -1;
    for (int k = 0; k < g_scene.n_prims; k++) {
        AHit h;
        // hw1 planes have no T_MAX and the nearest test is `best == -1 || t < best` (scene.cpp:13)
        if (hit_prim<false, false>(k, o, d, false, h) && (best == -1 || h.t < best)) {
            best = h.t; id = k; ans = ld3(g_scene.prims[k].color);
        }
    }
    out[3 * i] = ans.x; out[3 * i + 1] = ans.y; out[3 * i + 2] = ans.z;
    ids[i] = id;
}

#endif
// ---- K2 ------------------------------------------------------------------------------------
template <bool FAST>
__device__ __forceinline__ int closest_prim(V3 o, V3 d, float tmax, bool plane_tmax, AHit &best) {
    int pos = -1;
    if constexpr (FAST) {
        // candidates without normals; the winner's normal afterwards (one more test of ONE figure instead of a
        // local -> world transform + normalisation for every figure the ray pierces)
        for (int k = 0; k < g_scene.n_prims; k++) {
            AHit h;
            if (hit_prim<false, true>(k, o, d, plane_tmax, h) && h.t <= tmax && (pos == -1 || h.t < best.t)) { best.t = h.t; pos = k; }
        }
        if (pos >= 0) hit_prim<true, true>(pos, o, d, plane_tmax, best);
    } else {
        for (int k = 0; k < g_scene.n_prims; k++) {
            AHit h;
            if (hit_prim<true, false>(k, o, d, plane_tmax, h) && h.t <= tmax && (pos == -1 || h.t < best.t)) { best = h; pos = k; }
        }
    }
    return pos;
}
#ifndef B200RT_ANALYTIC_MC_UNIT
__device__ __forceinline__ bool occluded(V3 o, V3 d, float tmax) {
    for (int k = 0; k < g_scene.n_prims; k++) {
        AHit h;
        if (hit_prim<false, false>(k, o, d, false, h) && h.t <= tmax) return true;
    }
    return false;
}

#define B200RT_WHITTED_MAX_DEPTH 16
struct WFrame {
    V3 o, d;        // incoming ray
    V3 aux;         // stage 3: reflected colour
    V3 n;
    float t;
    int prim, depth, stage;
    bool inside;
};

// Explicit-stack evaluation of the branching recursion Scene::getColor (hw2/src/scene.cpp:29-85),
// post-order so every float operation happens in the recursion's order.
__global__ void whitted_kernel(const DCamera cam, int ray_depth, float *__restrict__ out, int32_t *__restrict__ ids) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cam.width * cam.height) return;
    const int x = i % cam.width, y = i / cam.width;
    WFrame st[B200RT_WHITTED_MAX_DEPTH + 1];
    int sp = 0;
    st[0].o = ld3(cam.pos); st[0].d = camera_dir_center(cam, x, y); st[0].depth = ray_depth; st[0].stage = 0;
    V3 ret = mk(0, 0, 0);
    int primary = -1;
    while (sp >= 0) {
        WFrame &f = st[sp];
        if (f.stage == 0) {
            if (f.depth == 0) { ret = mk(0.f, 0.f, 0.f); sp--; continue; }
            AHit h;
            const int pos = closest_prim<false>(f.o, f.d, INFINITY, false, h);
            if (sp == 0) primary = pos;
            if (pos < 0) { ret = ld3(g_scene.bg); sp--; continue; }
            const b200rt_primitive &fig = g_scene.prims[pos];
            f.t = h.t; f.n = h.n; f.inside = h.inside; f.prim = pos;
            if (fig.material == B200RT_MAT_DIFFUSE) {
                V3 color = ld3(g_scene.ambient);
                for (int li = 0; li < g_scene.n_lights; li++) {
                    const b200rt_light &L = g_scene.lights[li];
                    const V3 p = f.o + h.t * f.d;
                    V3 l, c; float tmax;
                    if (L.type == B200RT_LIGHT_POINT) {
                        const V3 direction = ld3(L.v) - p;
                        const float r = sqrtf(dot(direction, direction));
                        const float k = (float)(1.0 / (double)(L.attenuation[0] + L.attenuation[1] * r + L.attenuation[2] * r * r));
                        c = k * ld3(L.intensity);
                        l = rnorm(direction);
                        tmax = r;
                    } else {
                        l = rnorm(ld3(L.v)); c = ld3(L.intensity); tmax = INFINITY;
                    }
                    const float reflected = dot(l, h.n);
                    if (reflected >= 0 && !occluded(p + 0.0001f * l, l, tmax)) color = color + reflected * c;
                }
                ret = color * ld3(fig.color);
                sp--;
                continue;
            }
            // METALLIC and DIELECTRIC both start with the mirror direction (scene.cpp:53-56)
            const V3 dn = rnorm(f.d);
            const V3 rd = dn - (2.f * dot(h.n, dn)) * h.n;
            f.stage = fig.material == B200RT_MAT_METALLIC ? 1 : 2;
            WFrame &c = st[sp + 1];
            c.o = (f.o + h.t * f.d) + 0.0001f * rd; c.d = rd; c.depth = f.depth - 1; c.stage = 0;
            sp++;
            continue;
        }
        const b200rt_primitive &fig = g_scene.prims[f.prim];
        if (f.stage == 1) { ret = ld3(fig.color) * ret; sp--; continue; }
        if (f.stage == 2) {
            float eta1 = 1.f, eta2 = fig.ior;
            if (f.inside) { const float s = eta1; eta1 = eta2; eta2 = s; }
            const V3 l = -1.f * rnorm(f.d);
            const float nl = dot(f.n, l);
            const float sin2 = eta1 / eta2 * sqrtf(1 - nl * nl);
            if (fabsf(sin2) > 1.f) { sp--; continue; }      // total internal reflection: ret = reflected
            const float cos2 = sqrtf(1 - sin2 * sin2);
            const V3 rd = (eta1 / eta2) * (-1.f * l) + (eta1 / eta2 * nl - cos2) * f.n;
            f.aux = ret;
            f.stage = 3;
            WFrame &c = st[sp + 1];
            c.o = (f.o + f.t * f.d) + 0.0001f * rd; c.d = rd; c.depth = f.depth - 1; c.stage = 0;
            sp++;
            continue;
        }
        {   // stage 3: Schlick blend (scene.cpp:78-84)
            V3 refr = ret;
            if (!f.inside) refr = refr * ld3(fig.color);
            float eta1 = 1.f, eta2 = fig.ior;
            if (f.inside) { const float s = eta1; eta1 = eta2; eta2 = s; }
            const V3 l = -1.f * rnorm(f.d);
            const float q = (eta1 - eta2) / (eta1 + eta2);
            const float r0 = q * q;                                  // pow(q, 2.) rounds to the float product
            const double b1 = (double)(1 - dot(f.n, l));
            const double b2 = b1 * b1;
            const float r = (float)((double)r0 + (double)(1 - r0) * (b2 * b2 * b1));   // pow(b, 5.) in double
            ret = r * f.aux + (1 - r) * refr;
            sp--;
        }
    }
    out[3 * i] = ret.x; out[3 * i + 1] = ret.y; out[3 * i + 2] = ret.z;
    ids[i] = primary;
}

#else      // B200RT_ANALYTIC_MC_UNIT
// ---- K3: hw4 Monte-Carlo path tracer over analytic primitives ---------------------------------
// Iterative form of Scene::getColor (hw4/src/scene.cpp:50-108): L += T * emission per hit,
// T *= bounce weight.  Sampling = Mix{Cosine, Mix{BoxLight | EllipsoidLight ...}}
// (hw4/src/scene.cpp:10-28, distributions.h:46-205).  Statistical parity only (counter RNG);
// the stream layout equals oracle/port_mc.cpp's Philox policy draw for draw.
#define B200RT_PI_F 3.14159274f          /* const float PI = acos(-1) */
#define B200RT_MAX_REJECTION 8           /* the reference loops unbounded */

// FigureLight::pdf (distributions.h:85-106) = the density of the first hit plus that of the second hit, found from
// just behind the first.  Boxes and ellipsoids are convex: the two hits are the entry and the exit of ONE test (the
// exit counts when it lies at least 1e-4 behind the entry, which is when the reference's second test finds it; an
// origin inside the figure sees the exit only).  One world -> local transform and one root solve per emitter instead
// of two of each plus a world -> local transform of every hit point.
__device__ __forceinline__ float light_pdf(int k, V3 x, V3 d) {
    const b200rt_primitive &f = g_scene.prims[k];
    const float *m = g_scene.frame[k];
    const V3 ir = ld3(m + 12);
    V3 lo, ld;
    to_local<true>(k, f, x, d, lo, ld);
    const bool box = f.type == B200RT_PRIM_BOX;
    float t1, t2;
    if (box) {
        const V3 s = ld3(f.data);
        const V3 inv = mk(rcp_fast(ld.x), rcp_fast(ld.y), rcp_fast(ld.z));
        const V3 ts1 = (-1.f * s - lo) * inv, ts2 = (s - lo) * inv;
        t1 = smax(smax(smin(ts1.x, ts2.x), smin(ts1.y, ts2.y)), smin(ts1.z, ts2.z));
        t2 = smin(smin(smax(ts1.x, ts2.x), smax(ts1.y, ts2.y)), smax(ts1.z, ts2.z));
        if (t1 > t2 || t2 < 0) return 0.f;
    } else {
        const V3 orr = lo * ir, dr = ld * ir;
        const float qa = dot(dr, dr), qb = 2.f * dot(orr, dr), qc = dot(orr, orr) - 1;
        const float disc = qb * qb - 4 * qa * qc;
        if (disc <= 0) return 0.f;
        const float sq = sqrt_fast(disc), inv = rcp_fast(2 * qa);
        t1 = (-qb - sq) * inv; t2 = (-qb + sq) * inv;
        if (t1 > t2) { const float sw = t1; t1 = t2; t2 = sw; }
        if (t2 < 0) return 0.f;
    }
    const bool inside = t1 < 0;
    const float first = inside ? t2 : t1;
    if (isnan(first)) return INFINITY;
    const float dd = dot(d, d);
    float s_total = 0.f;
    if (box) { const float sx = f.data[0], sy = f.data[1], sz = f.data[2]; s_total = 8 * (sy * sz + sx * sz + sx * sy); }
    // density of the hit at ray parameter t: |x - y|^2 / (area density^-1 * |d . n_y|)
    auto term = [&](float t) {
        const V3 lp = lo + t * ld;
        V3 n = lp * ir;                    // box: p / s;  ellipsoid: p / r
        float point_prob_inv;
        if (box) {
            const float mx = smax(smax(fabsf(n.x), fabsf(n.y)), fabsf(n.z));
            if (fabsf(n.x) != mx) n.x = 0;
            if (fabsf(n.y) != mx) n.y = 0;
            if (fabsf(n.z) != mx) n.z = 0;
            point_prob_inv = s_total;
        } else {
            const V3 r = ld3(f.data);
            const V3 g = mk(n.x * r.y * r.z, r.x * n.y * r.z, r.x * r.y * n.z);
            point_prob_inv = 4 * B200RT_PI_F * sqrt_fast(dot(g, g));
            n = n * ir;                    // gradient direction p / r^2
        }
        const V3 yn = normal_to_world<true>(k, f, n);
        return t * t * dd * rcp_fast(point_prob_inv * fabsf(dot(d, yn)));
    };
    float ans = term(first);
    if (!inside && t2 - (t1 + 0.0001f) >= 0) ans += term(t2);
    return ans;
}

__device__ __forceinline__ V3 light_sample(int k, V3 x, Rng &rng) {
    const b200rt_primitive &f = g_scene.prims[k];
    const float *m = g_scene.frame[k];
    V3 dir = mk(0, 0, 1);
    for (int iter = 0; iter < B200RT_MAX_REJECTION; iter++) {
        V3 point;
        if (f.type == B200RT_PRIM_BOX) {
            const float sx = f.data[0], sy = f.data[1], sz = f.data[2];
            const float wx = sy * sz, wy = sx * sz, wz = sx * sy;
            const float u = rng.u() * (wx + wy + wz);
            const float flip = rng.u() > 0.5f ? 1.f : -1.f;
            // the reference's Vec3(...) constructor call evaluates its arguments right to left: the LATER coordinate
            // takes the earlier draw (oracle/port_mc.cpp, pinned against the hw4 / hw5 binaries)
            const float b = rng.u(), a = rng.u();
            if (u < wx) point = mk(flip * sx, (2 * a - 1) * sy, (2 * b - 1) * sz);
            else if (u < wx + wy) point = mk((2 * a - 1) * sx, flip * sy, (2 * b - 1) * sz);
            else point = mk((2 * a - 1) * sx, (2 * b - 1) * sy, flip * sz);
        } else {
            const float u1 = rng.u(), u2 = rng.u();
            point = ld3(f.data) * sphere_dir(u1, u2);
        }
        // local -> world: the transpose of the frame; for a non-unit quaternion both maps scale by |q|^2, so
        // the transpose alone stands for q-conjugate's sandwich
        const V3 actual = (m[9] != 0.f ? point : frame_apply_t(m, point)) + ld3(f.position);
        const V3 ax = actual - x;
        dir = rsqrtf(fmaf(ax.z, ax.z, fmaf(ax.y, ax.y, ax.x * ax.x))) * ax;
        AHit h;
        if (hit_prim<false, true>(k, x, dir, true, h)) break;
    }
    return dir;
}

// HW3 = the uniform-hemisphere integrator of hw3 (hw3/src/scene.cpp:31-106): the diffuse bounce draws a uniform
// direction on the hemisphere around the normal and weighs it by 2 (w . n) colour; planes have no T_MAX yet
// (hw3/src/primitives.cpp:72); the camera adds 0.5 to the jittered pixel position (hw3/src/scene.cpp:103-104).
// Resident blocks per SM asked of ptxas (profiles/r02_sweep_analytic_minb.log): these per-path kernels wait on their own
// dependent arithmetic and loads, and more resident warps pay for the spills up to 64 registers (K3) / 42 (K7).
#ifndef B200RT_K3_MINB
#define B200RT_K3_MINB 8
#endif
#ifndef B200RT_K7_MINB
#define B200RT_K7_MINB 12
#endif
template <bool HW3>
__global__ void __launch_bounds__(128, B200RT_K3_MINB) pt_analytic_kernel(const DRenderParams rp) {
    const int W = rp.cam.width, H = rp.cam.height;
    const int tiles_x = (W + 7) >> 3, tiles_y = (H + 3) >> 2;
    const long long n_tiles = (long long)tiles_x * tiles_y;
    const long long warp_id = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    unsigned paths = 0, rays = 0;
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
            Rng rng;
            rng.pixel = (uint32_t)(py * W + px); rng.seed = rp.seed;
            const int n_lights = g_scene.n_light_prims;
            V3 sum = mk(0, 0, 0);
            // Path regeneration: ONE loop over bounces; a lane whose path has ended starts its pixel's next sample in
            // the same pass instead of idling until the longest path of the warp's 32 pixels is through (ncu: 14.6 of
            // 32 lanes active with a loop over samples around a loop over bounces).  The counter-based generator makes
            // every path independent of when it is traced, and each pixel still adds its samples in order.
            int s = rp.ray_depth > 0 ? s0 : s1, bounce = 0;
            bool alive = false;
            V3 o = mk(0, 0, 0), d = mk(0, 0, 1), L = mk(0, 0, 0), T = mk(1, 1, 1);
            while (true) {
                if (!alive) {
                    if (s >= s1) break;
                    rng.begin_sample((uint32_t)s);
                    // hw4 camera: jitter only, no +0.5, fovX based, NOT normalised (scene.cpp:113-128)
                    const float fx = (float)px + rng.u();
                    const float fy = (float)py + rng.u();
                    // hw3: 2 * (x + 0.5) / width - 1 is evaluated in double (the 0.5 literal), then rounded
                    const float nx = HW3 ? rp.cam.tan_x * (float)(2 * ((double)fx + 0.5) / W - 1) : rp.cam.tan_x * (2 * fx / W - 1);
                    const float ny = HW3 ? rp.cam.tan_y * (float)(2 * ((double)fy + 0.5) / H - 1) : rp.cam.tan_y * (2 * fy / H - 1);
                    o = ld3(rp.cam.pos);
                    d = (nx * ld3(rp.cam.right) - ny * ld3(rp.cam.up)) + ld3(rp.cam.forward);
                    L = mk(0, 0, 0); T = mk(1, 1, 1);
                    paths++;
                    bounce = 0;
                    alive = true;
                }
                // one bounce of this lane's current path; `end`: the path stops here
                bool end = false;
                do {
                    AHit h;
                    rays++;
                    const int pos = closest_prim<true>(o, d, INFINITY, !HW3, h);
                    if (pos < 0) { L = L + T * ld3(g_scene.bg); end = true; break; }
                    rng.begin_bounce((uint32_t)bounce);
                    const b200rt_primitive &f = g_scene.prims[pos];
                    const V3 x = o + h.t * d;
                    L = L + T * ld3(f.emission);
                    if (HW3 && f.material == B200RT_MAT_DIFFUSE) {
                        // uniform direction on the sphere (the reference normalises three normals), mirrored into
                        // the hemisphere of the normal; weight 2 (w . n) colour = colour/pi * cos / (1 / 2pi)
                        const float u1 = rng.u(), u2 = rng.u();
                        V3 w = sphere_dir(u1, u2);
                        if (dot(w, h.n) < 0) w = -1.f * w;
                        T = T * ((2 * dot(w, h.n)) * ld3(f.color));
                        o = x + 0.0001f * w; d = w;
                    } else if (f.material == B200RT_MAT_DIFFUSE) {
                        const V3 xs = x + 0.0001f * h.n;
                        // Mix::sample: uniform choice between Cosine and the light mixture
                        const int k = n_lights > 0 ? 2 : 1;
                        const int which = (int)(rng.u() * k);
                        V3 nd;
                        if (which == 0) {
                            const float u1 = rng.u(), u2 = rng.u();
                            V3 c = sphere_dir(u1, u2) + h.n;
                            const float l = sqrtf(dot(c, c));
                            nd = (l <= 1e-9f || dot(c, h.n) <= 1e-9f || isnan(l)) ? h.n : (1.f / l) * c;
                        } else {
                            const int li = min((int)(rng.u() * n_lights), n_lights - 1);
                            nd = light_sample(g_scene.light_prims[li], xs, rng);
                        }
                        if (dot(nd, h.n) < 0) { end = true; break; }
                        float pdf = smax(0.f, dot(nd, h.n) / B200RT_PI_F);
                        if (n_lights > 0) {
                            float inner = 0;
                            for (int li = 0; li < n_lights; li++) inner += light_pdf(g_scene.light_prims[li], xs, nd);
                            pdf = (pdf + inner / n_lights) / 2;
                        }
                        const float w = (1.f / (B200RT_PI_F * pdf)) * dot(nd, h.n);
                        T = T * (w * ld3(f.color));
                        o = x + 0.0001f * nd; d = nd;
                    } else {
                        const V3 dn = rnorm(d);
                        const V3 rd = dn - (2.f * dot(h.n, dn)) * h.n;
                        bool reflect = true;
                        V3 nd = rd;
                        if (f.material == B200RT_MAT_METALLIC) {
                            T = T * ld3(f.color);
                        } else {
                            // Fresnel coin first (counter RNG: same estimator as tracing the reflected
                            // branch before the coin)
                            const float coin = rng.u();
                            float eta1 = 1.f, eta2 = f.ior;
                            if (h.inside) { const float sw = eta1; eta1 = eta2; eta2 = sw; }
                            const V3 l = -1.f * dn;
                            const float nl = dot(h.n, l);
                            const float sin2 = eta1 / eta2 * sqrtf(1 - nl * nl);
                            if (!(fabsf(sin2) > 1.f)) {
                                const float q = (eta1 - eta2) / (eta1 + eta2);
                                const float r0 = q * q;
                                const float b1 = 1 - nl, b2 = b1 * b1;
                                const float r = r0 + (1 - r0) * (b2 * b2 * b1);
                                if (!(coin < r)) {
                                    const float cos2 = sqrtf(1 - sin2 * sin2);
                                    nd = (eta1 / eta2) * (-1.f * l) + (eta1 / eta2 * nl - cos2) * h.n;
                                    reflect = false;
                                    if (!h.inside) T = T * ld3(f.color);
                                }
                            }
                        }
                        (void)reflect;
                        o = x + 0.0001f * nd; d = nd;
                    }
                } while (false);
                if (++bounce >= rp.ray_depth) end = true;
                if (end) { sum = sum + L; s++; alive = false; }
            }
            float *dst = rp.accum + ((size_t)chunk * W * H + (size_t)(py * W + px)) * 3;
            dst[0] = sum.x; dst[1] = sum.y; dst[2] = sum.z;
        }
    }
    // per-warp counter flush
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        paths += __shfl_down_sync(0xffffffffu, paths, off);
        rays += __shfl_down_sync(0xffffffffu, rays, off);
    }
    if (lane == 0 && paths) {
        atomicAdd(&rp.counters->paths, (unsigned long long)paths);
        atomicAdd(&rp.counters->rays, (unsigned long long)rays);
    }
}

// ---- K7: hw5 -- text scenes with TRIANGLE figures ------------------------------------------------
// hw5/src/scene.cpp:25-126 with hw5/src/include/distributions.h:61-302.  The reference keeps ellipsoids, boxes and
// triangles in ONE BVH of Figures and scans the planes; here the analytic figures (planes, ellipsoids, boxes: a
// handful) are scanned in constant memory like hw4's, and the triangles -- baked to world space at upload -- go
// through the glTF modes' tree and traversal (closest hit and light-pdf sum do not depend on how the acceleration
// structure is organised).  The light mixture is FiguresMix over all three emissive kinds in the reference's
// order: `light_order[i]` >= 0 names an analytic figure, < 0 the emissive triangle ~light_order[i].
// Integrator = hw4's / hw6's (diffuse with Mix{Cosine, lights}, mirror, dielectric with the Fresnel coin first).
// Stream layout = oracle/port_mc.cpp Hw5<RngPhilox>, draw for draw (sequential Rng view).
__global__ void __launch_bounds__(128, B200RT_K7_MINB) hw5_path_kernel(const DGltfScene sc, const DRenderParams rp, const int32_t *__restrict__ light_order,
                                                       const int n_lights, const int has_tris) {
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
            Rng rng;
            rng.pixel = (uint32_t)(py * W + px); rng.seed = rp.seed;
            V3 sum = mk(0, 0, 0);
            // path regeneration, as in pt_analytic_kernel: one loop over bounces, ended paths restart at once
            int s = rp.ray_depth > 0 ? s0 : s1, bounce = 0;
            bool alive = false;
            V3 o = mk(0, 0, 0), d = mk(0, 0, 1), L = mk(0, 0, 0), T = mk(1, 1, 1);
            while (true) {
                if (!alive) {
                    if (s >= s1) break;
                    rng.begin_sample((uint32_t)s);
                    // camera of hw4 / hw5: jitter only, fovX based, NOT normalised (hw5/src/scene.cpp:107-126)
                    const float fx = (float)px + rng.u();
                    const float fy = (float)py + rng.u();
                    const float nx = rp.cam.tan_x * (2 * fx / W - 1);
                    const float ny = rp.cam.tan_y * (2 * fy / H - 1);
                    o = ld3(rp.cam.pos);
                    d = (nx * ld3(rp.cam.right) - ny * ld3(rp.cam.up)) + ld3(rp.cam.forward);
                    L = mk(0, 0, 0); T = mk(1, 1, 1);
                    tc.paths++;
                    bounce = 0;
                    alive = true;
                }
                bool end = false;
                do {
                    // Scene::intersect: analytic figures by scan, triangles through the BVH, nearest wins
                    AHit h;
                    tc.rays++;
                    const int pos = closest_prim<true>(o, d, INFINITY, true, h);
                    TraceHit th;
                    th.tri = -1;
                    if (has_tris) trace_closest<false>(sc, o, d, th, tc);
                    const bool on_tri = th.tri >= 0 && (pos < 0 || th.t < h.t);
                    if (pos < 0 && !on_tri) { L = L + T * ld3(sc.bg); end = true; break; }
                    rng.begin_bounce((uint32_t)bounce);
                    V3 emission, color;
                    int material;
                    float ior;
                    if (on_tri) {
                        const float4 *tq = reinterpret_cast<const float4 *>(sc.tris + th.tri);
                        const float4 tv0 = __ldg(tq), te1 = __ldg(tq + 1), te2 = __ldg(tq + 2);
                        V3 n = normalize(cross(mk(te2.x, te2.y, te2.z), mk(te1.x, te1.y, te1.z)));
                        h.inside = th.det > 0.f;
                        h.n = h.inside ? -n : n;
                        h.t = th.t;
                        const DMaterial *m = sc.materials + __float_as_int(tv0.w);
                        emission = mk(__ldg(&m->emission[0]), __ldg(&m->emission[1]), __ldg(&m->emission[2]));
                        color = mk(__ldg(&m->base_color[0]), __ldg(&m->base_color[1]), __ldg(&m->base_color[2]));
                        material = __ldg(&m->kind);
                        ior = __ldg(&m->ior);
                    } else {
                        const b200rt_primitive &f = g_scene.prims[pos];
                        emission = ld3(f.emission); color = ld3(f.color); material = f.material; ior = f.ior;
                    }
                    const V3 x = o + h.t * d;
                    L = L + T * emission;
                    if (material == B200RT_MAT_DIFFUSE) {
                        const V3 xs = x + 0.0001f * h.n;
                        const int k = n_lights > 0 ? 2 : 1;
                        const int which = (int)(rng.u() * k);
                        V3 nd;
                        if (which == 0) {
                            const float u1 = rng.u(), u2 = rng.u();
                            V3 c = sphere_dir(u1, u2) + h.n;
                            const float l = sqrtf(dot(c, c));
                            nd = (l <= 1e-9f || dot(c, h.n) <= 1e-9f || isnan(l)) ? h.n : (1.f / l) * c;
                        } else {
                            // FiguresMix::sample: index, then the sampler of that light's kind
                            const int li = min((int)(rng.u() * n_lights), n_lights - 1);
                            const int32_t e = __ldg(light_order + li);
                            if (e >= 0) {
                                nd = light_sample(e, xs, rng);
                            } else {
                                const float4 *lq = reinterpret_cast<const float4 *>(sc.light_sample_tris + ~e);
                                const float4 a = __ldg(lq), b = __ldg(lq + 1), c = __ldg(lq + 2);
                                float lu = rng.u(), lv = rng.u();
                                if (lu + lv > 1.f) { lu = 1.f - lu; lv = 1.f - lv; }
                                const V3 point = mk(a.x + lu * b.x + lv * c.x, a.y + lu * b.y + lv * c.y, a.z + lu * b.z + lv * c.z);
                                nd = rnorm(point - xs);
                            }
                        }
                        if (dot(nd, h.n) < 0) { end = true; break; }
                        float pdf = smax(0.f, dot(nd, h.n) / B200RT_PI_F);
                        if (n_lights > 0) {
                            // FiguresMix::pdf: every emissive figure the ray pierces contributes (both hits for closed ones)
                            tc.light_queries++;
                            float inner = 0;
                            for (int li = 0; li < g_scene.n_light_prims; li++) inner += light_pdf(g_scene.light_prims[li], xs, nd);
                            if (sc.n_light_tris > 0) inner += light_pdf_sum<false>(sc, xs, nd, tc);
                            pdf = (pdf + inner / n_lights) / 2;
                        }
                        const float w = (1.f / (B200RT_PI_F * pdf)) * dot(nd, h.n);
                        T = T * (w * color);
                        o = x + 0.0001f * nd; d = nd;
                    } else {
                        const V3 dn = rnorm(d);
                        V3 nd = dn - (2.f * dot(h.n, dn)) * h.n;
                        if (material == B200RT_MAT_METALLIC) {
                            T = T * color;
                        } else {
                            const float coin = rng.u();      // Fresnel coin first (see K3)
                            float eta1 = 1.f, eta2 = ior;
                            if (h.inside) { const float sw = eta1; eta1 = eta2; eta2 = sw; }
                            const V3 l = -1.f * dn;
                            const float nl = dot(h.n, l);
                            const float sin2 = eta1 / eta2 * sqrtf(1 - nl * nl);
                            if (!(fabsf(sin2) > 1.f)) {
                                const float q = (eta1 - eta2) / (eta1 + eta2);
                                const float r0 = q * q;
                                const float b1 = 1 - nl, b2 = b1 * b1;
                                const float r = r0 + (1 - r0) * (b2 * b2 * b1);
                                if (!(coin < r)) {
                                    const float cos2 = sqrtf(1 - sin2 * sin2);
                                    nd = (eta1 / eta2) * (-1.f * l) + (eta1 / eta2 * nl - cos2) * h.n;
                                    if (!h.inside) T = T * color;
                                }
                            }
                        }
                        o = x + 0.0001f * nd; d = nd;
                    }
                } while (false);
                if (++bounce >= rp.ray_depth) end = true;
                if (end) { sum = sum + L; s++; alive = false; }
            }
            float *dst = rp.accum + ((size_t)chunk * W * H + (size_t)(py * W + px)) * 3;
            dst[0] = sum.x; dst[1] = sum.y; dst[2] = sum.z;
        }
    }
    flush_counters(rp.counters, tc);
}

#endif     // B200RT_ANALYTIC_MC_UNIT

}  // namespace

#ifndef B200RT_ANALYTIC_MC_UNIT
cudaError_t launch_cast(const DCamera &cam, float *out, int32_t *ids, cudaStream_t stream) {
    const int n = cam.width * cam.height;
    cast_kernel<<<(n + 127) / 128, 128, 0, stream>>>(cam, out, ids);
    return cudaGetLastError();
}

cudaError_t launch_whitted(const DCamera &cam, int ray_depth, float *out, int32_t *ids, cudaStream_t stream) {
    if (ray_depth > B200RT_WHITTED_MAX_DEPTH) return cudaErrorInvalidValue;
    const int n = cam.width * cam.height;
    whitted_kernel<<<(n + 63) / 64, 64, 0, stream>>>(cam, ray_depth, out, ids);
    return cudaGetLastError();
}
#else
cudaError_t launch_pt_analytic(const DRenderParams &rp, bool hw3, cudaStream_t stream) {
    const long long tiles = (long long)((rp.cam.width + 7) / 8) * ((rp.cam.height + 3) / 4);
    const long long warps = tiles * rp.n_chunks;
    const long long blocks = (warps * 32 + 127) / 128;
    if (hw3) pt_analytic_kernel<true><<<(unsigned)blocks, 128, 0, stream>>>(rp);
    else pt_analytic_kernel<false><<<(unsigned)blocks, 128, 0, stream>>>(rp);
    return cudaGetLastError();
}

cudaError_t launch_hw5(const DGltfScene &sc, const DRenderParams &rp, const int32_t *light_order, int n_lights, bool has_tris, cudaStream_t stream) {
    const long long tiles = (long long)((rp.cam.width + 7) / 8) * ((rp.cam.height + 3) / 4);
    const long long warps = tiles * rp.n_chunks;
    const long long blocks = (warps * 32 + 127) / 128;
    hw5_path_kernel<<<(unsigned)blocks, 128, 0, stream>>>(sc, rp, light_order, n_lights, has_tris ? 1 : 0);
    return cudaGetLastError();
}
#endif

}  // namespace b200rt
