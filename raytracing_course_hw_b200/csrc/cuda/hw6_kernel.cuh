// K6: hw6 -- the hw5 integrator (DIFFUSE / METALLIC mirror / DIELECTRIC with Schlick coin) on glTF
// triangles through the BVH (hw6/src/scene.cpp:47-126).  Sampling = Mix{Cosine, FiguresMix over the
// emissive triangles} (hw6/src/include/distributions.h:36-58, 107-141, 167-302).  Iterative form of
// Scene::getColor: L += T * emission per hit, T *= bounce weight.  One thread per (pixel, sample
// chunk), like gltf_path_kernel; closest hit and light-pdf sum are the shared traversal routines of
// gltf_kernel.cuh (geometric normals: hw6 triangles carry no shading normals).  Statistical parity
// (counter RNG); the stream layout equals oracle/port_mc.cpp's Hw6<RngPhilox> draw for draw: bounce b
// uses the four values of block 16 (b + 1) in order.
#pragma once
#include "gltf_kernel.cuh"

namespace b200rt {

#define B200RT_PI_F6 3.14159274f          /* const float PI = acos(-1) (distributions.h:11) */

// The integrator is written once, as two device functions shared by both schedulings of hw6: the wavefront pipeline
// (wavefront.cuh, the default: wf_shade<..., HW6> + wf_light) and the one-thread-per-pixel kernel below
// (B200RT_FLAG_MEGAKERNEL, kept as the cross-check build).

// One hit of hw6's Scene::getColor: adds the emission to L and makes the next ray.  A DIFFUSE hit leaves its bounce
// weight to hw6_finish (it needs the light-pdf sum of the new ray): out.brdf = the surface colour, out.pdf_partial =
// the cosine pdf, out.absdot = nd . n.  A mirror / dielectric hit updates T here and sets `specular`.  Returns
// false when the path ends (the sampled direction points below the surface).
template <bool COUNT>
__device__ __forceinline__ bool hw6_surface(const DGltfScene &sc, const V3 o, const V3 d, const float t, const int tri, const bool inside,
                                            const uint4 rnd, V3 &L, V3 &T, ThreadCounters &tc, BounceOut &out, bool &specular) {
    const float4 *tq = reinterpret_cast<const float4 *>(sc.tris + tri);
    const float4 tv0 = __ldg(tq), te1 = __ldg(tq + 1), te2 = __ldg(tq + 2);
    if (COUNT) tc.shaded++;
    // n = b.cross(c) of the reference (its cross is negated: c x b), flipped towards the ray
    V3 n = normalize(cross(mk(te2.x, te2.y, te2.z), mk(te1.x, te1.y, te1.z)));
    if (inside) n = -n;
    const DMaterial *m = sc.materials + __float_as_int(tv0.w);
    const V3 emission = mk(__ldg(&m->emission[0]), __ldg(&m->emission[1]), __ldg(&m->emission[2]));
    const V3 color = mk(__ldg(&m->base_color[0]), __ldg(&m->base_color[1]), __ldg(&m->base_color[2]));
    const int kind = __ldg(&m->kind);
    const int n_lights = sc.n_light_tris;
    const V3 x = o + t * d;
    L = L + T * emission;
    V3 nd;
    if (kind == B200RT_MAT_DIFFUSE) {
        specular = false;
        const V3 xs = x + B200RT_EPS * n;
        // Mix::sample: uniform choice between Cosine and the emissive triangles
        const int k = n_lights > 0 ? 2 : 1;
        const int which = (int)(u01(rnd.x) * (float)k);
        if (which == 0) {
            V3 c = sphere_dir(u01(rnd.y), u01(rnd.z)) + n;
            const float l = sqrtf(len2(c));
            nd = (l <= 1e-9f || dot(c, n) <= 1e-9f || isnan(l)) ? n : (1.f / l) * c;
        } else {
            // FiguresMix::sample + TriangleLight::sample: draws = index, u, v
            const int li = min((int)(u01(rnd.y) * (float)n_lights), n_lights - 1);
            const float4 *lq = reinterpret_cast<const float4 *>(sc.light_sample_tris + li);
            const float4 a = __ldg(lq), b = __ldg(lq + 1), c = __ldg(lq + 2);
            float lu = u01(rnd.z), lv = u01(rnd.w);
            if (lu + lv > 1.f) { lu = 1.f - lu; lv = 1.f - lv; }
            const V3 point = mk(a.x + lu * b.x + lv * c.x, a.y + lu * b.y + lv * c.y, a.z + lu * b.z + lv * c.z);
            nd = normalize(point - xs);
        }
        const float cs = dot(nd, n);
        if (cs < 0.f) return false;
        out.brdf = color;
        out.pdf_partial = fmaxf(0.f, cs / B200RT_PI_F6);
        out.absdot = cs;
    } else {
        specular = true;
        const V3 dn = normalize(d);
        nd = dn - (2.f * dot(n, dn)) * n;      // mirror direction
        if (kind == B200RT_MAT_METALLIC) {
            T = T * color;
        } else {
            // Fresnel coin first (counter RNG: same estimator as the reference, which traces the
            // reflected branch before the coin and discards it when refraction wins)
            const float coin = u01(rnd.x);
            float eta1 = 1.f, eta2 = __ldg(&m->ior);
            if (inside) { const float sw = eta1; eta1 = eta2; eta2 = sw; }
            const V3 l = -dn;
            const float nl = dot(n, l);
            const float sin2 = eta1 / eta2 * sqrtf(1.f - nl * nl);
            if (!(fabsf(sin2) > 1.f)) {
                const float q = (eta1 - eta2) / (eta1 + eta2);
                const float r0 = q * q;
                const float b1 = 1.f - nl, b2 = b1 * b1;
                const float r = r0 + (1.f - r0) * (b2 * b2 * b1);
                if (!(coin < r)) {
                    const float cos2 = sqrtf(1.f - sin2 * sin2);
                    nd = (eta1 / eta2) * (-l) + (eta1 / eta2 * nl - cos2) * n;
                    if (!inside) T = T * color;
                }
            }
        }
        out.brdf = mk(0, 0, 0); out.pdf_partial = 0.f; out.absdot = 0.f;
    }
    out.xo = x + B200RT_EPS * nd;
    out.nd = nd;
    return true;
}

// The weight of a DIFFUSE bounce: Mix::pdf = mean of the cosine pdf and FiguresMix::pdf (light_sum / n_lights) when
// the scene has emitters, the cosine pdf alone otherwise; T *= colour / (pi pdf) * (nd . n)  (hw6/src/scene.cpp:75-88).
// No firefly clamp in hw6.
__device__ __forceinline__ void hw6_finish(const V3 color, float pdf_partial, float absdot, float light_sum, int n_lights, V3 &T) {
    float pdf = pdf_partial;
    if (n_lights > 0) pdf = (pdf + light_sum / (float)n_lights) * 0.5f;
    const float w = (1.f / (B200RT_PI_F6 * pdf)) * absdot;
    T = T * (w * color);
}

#ifndef B200RT_K6_MINB
#define B200RT_K6_MINB 8        /* resident blocks per SM asked of ptxas (profiles/r02_sweep_analytic_minb.log: 3.01 -> 3.20 Gpaths/s) */
#endif
template <bool COUNT>
__global__ void __launch_bounds__(128, B200RT_K6_MINB) hw6_path_kernel(const DGltfScene sc, const DRenderParams rp) {
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
            const int n_lights = sc.n_light_tris;
            V3 sum = mk(0, 0, 0);
            for (int s = s0; s < s1; s++) {
                // camera sample (scene.cpp:107-126): jitter in [0,1)^2, fovY based, direction NOT normalised
                const uint4 j = philox4x32_10((uint32_t)s, 0u, pixel, rp.seed);
                const float fx = (float)px + u01(j.x), fy = (float)py + u01(j.y);
                const float nx = rp.cam.tan_x * (2.f * fx / (float)W - 1.f);
                const float ny = rp.cam.tan_y * (2.f * fy / (float)H - 1.f);
                V3 o = cpos;
                V3 d = nx * cright - ny * cup + cfwd;
                V3 L = mk(0, 0, 0), T = mk(1, 1, 1);
                tc.paths++;
                for (int bounce = 0; bounce < rp.ray_depth; bounce++) {
                    TraceHit hit;
                    tc.rays++;
                    trace_closest<COUNT>(sc, o, d, hit, tc);
                    if (hit.tri < 0) { L = L + T * ld3(sc.bg); break; }
                    const uint4 rnd = philox4x32_10((uint32_t)s, 16u * (uint32_t)(bounce + 1), pixel, rp.seed);
                    BounceOut bo;
                    bool specular;
                    if (!hw6_surface<COUNT>(sc, o, d, hit.t, hit.tri, hit.det > 0.f, rnd, L, T, tc, bo, specular)) break;
                    if (!specular) {
                        float light_sum = 0.f;
                        if (n_lights > 0) {
                            tc.light_queries++;
                            light_sum = light_pdf_sum<COUNT>(sc, bo.xo, bo.nd, tc);
                        }
                        hw6_finish(bo.brdf, bo.pdf_partial, bo.absdot, light_sum, n_lights, T);
                    }
                    o = bo.xo; d = bo.nd;
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
