// TEST INFRASTRUCTURE ONLY -- CPU restatement of the MONTE-CARLO configs:
//   hw4 analytic path tracer      hw4/src/scene.cpp:10-128, hw4/src/include/distributions.h:46-205
//   hw7 / hw8 glTF path tracer    hw8/src/scene.cpp:9-187, hw8/src/primitives.cpp:18-125,
//                                 hw8/src/include/bvh.h:111-142, distributions.h:35-280, material.h:11-65
//                                 (hw7 deltas: hw7/src/scene.cpp:29-61, hw7/src/include/material.h:44-59,
//                                  hw7/src/include/distributions.h:140-145)
// This TU sees only <cmath> (never <math.h>), so unqualified tan/pow/sqrt/cos/fabs on floats bind
// to the DOUBLE C functions exactly as in the reference's hw3..hw8 translation units.
// Pinned by tests/test_oracle.py against oracle/_ref/hw4, hw7, hw8 with the minstd policy.
#include <cmath>
#include <algorithm>
#include <type_traits>
#include <vector>
#include "port_common.h"

static_assert(std::is_same<decltype(tan(1.f)), double>::value, "<cmath>-only environment expected (hw3..hw8 TUs)");

namespace port {
namespace mc {

#include "port_figures.inl"

static const float PI = acos(-1);
static const long double eps = 1e-4;

// hw8/src/color.cpp:4-31 under <cmath>: gamma via double pow.
static V saturate(const V &c) {
    return V(std::min(1.f, std::max(0.f, c.x)), std::min(1.f, std::max(0.f, c.y)), std::min(1.f, std::max(0.f, c.z)));
}
static V aces(const V &x) {
    const float a = 2.51f, b = 0.03f, c = 2.43f, d = 0.59f, e = 0.14f;
    return saturate((x * (a * x + b)) / (x * (c * x + d) + e));
}
static V gamma_corrected(const V &x) {
    float gamma = 1. / 2.2;
    return V(pow(x.x, gamma), pow(x.y, gamma), pow(x.z, gamma));
}
void quantise(const V &radiance, uint8_t *out) {
    V o = gamma_corrected(aces(radiance));
    out[0] = round(255 * o.x);
    out[1] = round(255 * o.y);
    out[2] = round(255 * o.z);
}

// ======================================== hw3 / hw4 =========================================
// RNG plumbing of hw4: ONE global minstd_rand consumed in scan order, but every distribution
// object owns its own normal_distribution (with its own cached second variate):
// Cosine::n01 (hw4 distributions.h:48) and one per EllipsoidLight (:154).
struct Hw4Minstd {
    std::minstd_rand rng;
    std::uniform_real_distribution<float> u01{0.0, 1.0};
    std::normal_distribution<float> cosine_n01{0.f, 1.f};
    std::vector<std::normal_distribution<float>> light_n01;
    void begin_pixel(uint32_t, uint32_t) {}
    void begin_sample(uint32_t) {}
    void begin_bounce(uint32_t) {}
    float u() { return u01(rng); }
    V cosine_sphere() { return V{cosine_n01(rng), cosine_n01(rng), cosine_n01(rng)}.normalize(); }
    V light_sphere(size_t i) { auto &n = light_n01[i]; return V{n(rng), n(rng), n(rng)}.normalize(); }
    static constexpr bool kCounterBased = false;
};
struct Hw4Philox : RngPhilox {
    V cosine_sphere() { return sphere(); }
    V light_sphere(size_t) { return sphere(); }
};

static const int kMaxRejection = 8;  // counter-based policy only; the reference loops unbounded

template <class Rng>
struct Hw4 {
    const b200rt_scene_desc &s;
    Rng &rng;
    std::vector<int> lights;  // indices of emissive BOX / ELLIPSOID figures (hw4/src/scene.cpp:12-19)

    Hw4(const b200rt_scene_desc &s_, Rng &r) : s(s_), rng(r) {
        for (int i = 0; i < s.n_prims; i++) {
            const auto &f = s.prims[i];
            if ((f.emission[0] > 0 || f.emission[1] > 0 || f.emission[2] > 0) &&
                (f.type == B200RT_PRIM_BOX || f.type == B200RT_PRIM_ELLIPSOID)) lights.push_back(i);
        }
    }

    // hw3 (hw3/src/scene.cpp, hw3/src/primitives.cpp) is this integrator before importance sampling arrived:
    // uniform-hemisphere diffuse bounce from ONE global normal_distribution, +0.5 in the camera ray, planes
    // without T_MAX.  Everything else (metal, dielectric, scan-order intersect) is shared with hw4.
    bool hw3() const { return s.mode == B200RT_MODE_HW3; }

    bool intersect(const Ray &ray, Hit &best, int &pos) const {
        pos = -1;
        for (int i = 0; i < s.n_prims; i++) {
            Hit h = hit_figure(s.prims[i], ray, !hw3());
            if (h.valid && (pos == -1 || h.t < best.t)) { best = h; pos = i; }
        }
        return pos != -1;
    }

    // ---- distributions ----
    V cosine_sample(V n) {
        V d = rng.cosine_sphere();
        d = d + n;
        float len = d.len();
        if (len <= 1e-9f || d.dot(n) <= 1e-9f || std::isnan(len)) return n;
        return 1. / len * d;
    }
    float cosine_pdf(V n, V d) const { return std::max(0.f, d.dot(n) / PI); }

    float pdf_one(const b200rt_primitive &f, V x, V d, V y, V yn) const {
        if (f.type == B200RT_PRIM_BOX) {
            float sx = f.data[0], sy = f.data[1], sz = f.data[2];
            float sTotal = 8 * (sy * sz + sx * sz + sx * sy);
            return (x - y).len2() / (sTotal * fabs(d.dot(yn)));
        }
        V r(f.data);
        V n = Q(f.rotation).transform(y - V(f.position)) / r;
        float pointProb = 1. / (4 * PI * V{n.x * r.y * r.z, r.x * n.y * r.z, r.x * r.y * n.z}.len());
        return pointProb * (x - y).len2() / fabs(d.dot(yn));
    }
    // FigureLight::pdf (hw4 distributions.h:85-106): first hit + second hit from just behind it.
    float light_pdf(const b200rt_primitive &f, V x, V d) const {
        Hit h1 = hit_figure(f, Ray(x, d), true);
        if (!h1.valid) return 0.;
        if (std::isnan(h1.t)) return INFINITY;
        V y = x + h1.t * d;
        float ans = pdf_one(f, x, d, y, h1.norma);
        Hit h2 = hit_figure(f, Ray(x + (h1.t + 0.0001) * d, d), true);
        if (!h2.valid) return ans;
        V y2 = x + (h1.t + 0.0001 + h2.t) * d;
        return ans + pdf_one(f, x, d, y2, h2.norma);
    }
    V light_sample(size_t li, V x) {
        const b200rt_primitive &f = s.prims[lights[li]];
        Q rot(f.rotation);
        V dir;
        for (int iter = 0;; iter++) {
            V point;
            if (f.type == B200RT_PRIM_BOX) {
                float sx = f.data[0], sy = f.data[1], sz = f.data[2];
                float wx = sy * sz, wy = sx * sz, wz = sx * sy;
                float u = rng.u() * (wx + wy + wz);
                float flipSign = rng.u() > 0.5 ? 1 : -1;
                // Vec3(...) is a constructor call: g++ evaluates its arguments right to left, so the LATER coordinate
                // takes the earlier random number (pinned against the hw4 and hw5 binaries on box-light scenes)
                if (u < wx) { float b = rng.u(), a = rng.u(); point = V(flipSign * sx, (2 * a - 1) * sy, (2 * b - 1) * sz); }
                else if (u < wx + wy) { float b = rng.u(), a = rng.u(); point = V((2 * a - 1) * sx, flipSign * sy, (2 * b - 1) * sz); }
                else { float b = rng.u(), a = rng.u(); point = V((2 * a - 1) * sx, (2 * b - 1) * sy, flipSign * sz); }
            } else {
                point = V(f.data) * rng.light_sphere(li);
            }
            V actual = rot.conjugate().transform(point) + V(f.position);
            dir = (actual - x).normalize();
            if (hit_figure(f, Ray(x, dir), true).valid) return dir;
            if (Rng::kCounterBased && iter + 1 >= kMaxRejection) return dir;
        }
    }
    // Mix{Cosine, Mix{lights...}} (hw4/src/scene.cpp:21-27)
    V mix_sample(V x, V n) {
        size_t k = lights.empty() ? 1 : 2;
        int which = rng.u() * k;
        if (which == 0) return cosine_sample(n);
        int li = rng.u() * lights.size();
        return light_sample((size_t)li, x);
    }
    float mix_pdf(V x, V n, V d) const {
        float ans = 0;
        ans += cosine_pdf(n, d);
        size_t k = 1;
        if (!lights.empty()) {
            float inner = 0;
            for (int li : lights) inner += light_pdf(s.prims[li], x, d);
            ans += inner / lights.size();
            k = 2;
        }
        return ans / k;
    }

    V color(const Ray &ray, int rec) {
        if (rec == 0) return V(0., 0., 0.);
        Hit h; int pos;
        if (!intersect(ray, h, pos)) return V(s.bg_color);
        rng.begin_bounce((uint32_t)(s.ray_depth - rec));
        const b200rt_primitive &f = s.prims[pos];
        float t = h.t; V norma = h.norma; bool is_inside = h.inside;
        V x = ray.o + t * ray.d;
        V emission(f.emission), fcolor(f.color);
        if (hw3() && f.material == B200RT_MAT_DIFFUSE) {
            // hw3/src/scene.cpp:45-51; the global n01 (:7) is this policy's cosine_n01 (no other user in hw3)
            V w = rng.cosine_sphere();
            if (w.dot(norma) < 0) w = -1. * w;
            Ray wRay(ray.o + t * ray.d + 0.0001 * w, w);
            return emission + 2 * w.dot(norma) * fcolor * color(wRay, rec - 1);
        } else if (f.material == B200RT_MAT_DIFFUSE) {
            V d = mix_sample(x + 0.0001 * norma, norma);
            if (d.dot(norma) < 0) return emission;
            float pdf = mix_pdf(x + 0.0001 * norma, norma, d);
            Ray dRay(x + 0.0001 * d, d);
            return emission + 1. / (PI * pdf) * d.dot(norma) * fcolor * color(dRay, rec - 1);
        } else if (f.material == B200RT_MAT_METALLIC) {
            V reflectedDir = ray.d.normalize() - 2. * norma.dot(ray.d.normalize()) * norma;
            Ray reflected(ray.o + t * ray.d + 0.0001 * reflectedDir, reflectedDir);
            return emission + fcolor * color(reflected, rec - 1);
        } else {
            V reflectedDir = ray.d.normalize() - 2. * norma.dot(ray.d.normalize()) * norma;
            Ray reflected(ray.o + t * ray.d + 0.0001 * reflectedDir, reflectedDir);
            // Reference: the reflected branch is traced BEFORE the Fresnel coin and regardless of
            // it.  Counter-based policy: the coin comes from this bounce's own block,
            // so the unused branch can be skipped without changing the estimator.
            V reflectedColor;
            float coin = 0;
            if (Rng::kCounterBased) coin = rng.u();
            else reflectedColor = color(reflected, rec - 1);
            float eta1 = 1., eta2 = f.ior;
            if (is_inside) std::swap(eta1, eta2);
            V l = -1. * ray.d.normalize();
            float sinTheta2 = eta1 / eta2 * sqrt(1 - norma.dot(l) * norma.dot(l));
            if (fabs(sinTheta2) > 1.) {
                if (Rng::kCounterBased) reflectedColor = color(reflected, rec - 1);
                return emission + reflectedColor;
            }
            float r0 = pow((eta1 - eta2) / (eta1 + eta2), 2.);
            float r = r0 + (1 - r0) * pow(1 - norma.dot(l), 5.);
            if (!Rng::kCounterBased) coin = rng.u();
            if (coin < r) {
                if (Rng::kCounterBased) reflectedColor = color(reflected, rec - 1);
                return emission + reflectedColor;
            }
            float cosTheta2 = sqrt(1 - sinTheta2 * sinTheta2);
            V refractedDir = eta1 / eta2 * (-1. * l) + (eta1 / eta2 * norma.dot(l) - cosTheta2) * norma;
            Ray refracted(ray.o + t * ray.d + 0.0001 * refractedDir, refractedDir);
            V refractedColor = color(refracted, rec - 1);
            if (!is_inside) refractedColor = refractedColor * fcolor;
            return emission + refractedColor;
        }
    }

    Ray camera_ray(float x, float y) const {
        float tanFovX = tan(s.cam_fov / 2);
        float tanFovY = tanFovX * s.height / s.width;
        float nx = tanFovX * (2 * x / s.width - 1);
        float ny = tanFovY * (2 * y / s.height - 1);
        if (hw3()) {   // hw3/src/scene.cpp:103-104: the 0.5 literal makes the expression double
            nx = tanFovX * (2 * (x + 0.5) / s.width - 1);
            ny = tanFovY * (2 * (y + 0.5) / s.height - 1);
        }
        return {V(s.cam_pos), nx * V(s.cam_right) - ny * V(s.cam_up) + V(s.cam_forward)};
    }
    // Sum over samples [b, e) (hw4/src/scene.cpp:110-118 without the final 1/samples).
    V pixel_sum(int x, int y, int b, int e) {
        V c(0, 0, 0);
        for (int k = b; k < e; k++) {
            rng.begin_sample((uint32_t)k);
            float nx = x + rng.u();
            float ny = y + rng.u();
            c = c + color(camera_ray(nx, ny), s.ray_depth);
        }
        return c;
    }
};

// ======================================= hw7 / hw8 =========================================
struct TriHit {
    float t = 0, u = 0, v = 0;
    bool inside = false;
    V n;  // unnormalised plane normal, already flipped towards the ray origin
};

struct Surface {
    float t;
    V geomNorma, shadingNorma, tangent;
    float tangent_w, tx, ty;
    bool inside;
};

static const float magic1[] = {0.239, 0.419, 0.533};
static const float magic2[] = {0.35743, 0.66682, 0.69695};

// intersectPlaneAndRay + the projected 2x2 solve (hw8/src/primitives.cpp:18-27, 76-108).
static inline bool hit_triangle(const b200rt_triangle &f, const Ray &ray, TriHit &h) {
    V a(f.p[2]);
    V b = V(f.p[0]) - a;
    V c = V(f.p[1]) - a;
    V n = b.cross(c);
    Ray r = ray.shifted(a);
    float t = -r.o.dot(n) / r.d.dot(n);
    if (!(t > 0 && t < kTMax)) return false;
    if (r.d.dot(n) > 0) { h.n = -1. * n; h.inside = true; }
    else { h.n = n; h.inside = false; }
    V p = ray.o - a + t * ray.d;
    auto a1 = magic1[0] * b.x + magic1[1] * b.y + magic1[2] * b.z;
    auto b1 = magic1[0] * c.x + magic1[1] * c.y + magic1[2] * c.z;
    auto c1 = magic1[0] * p.x + magic1[1] * p.y + magic1[2] * p.z;
    auto a2 = magic2[0] * b.x + magic2[1] * b.y + magic2[2] * b.z;
    auto b2 = magic2[0] * c.x + magic2[1] * c.y + magic2[2] * c.z;
    auto c2 = magic2[0] * p.x + magic2[1] * p.y + magic2[2] * p.z;
    float y = (c1 * a2 - c2 * a1) / (b1 * a2 - a1 * b2);
    float x = a2 == 0 ? (c1 - b1 * y) / a1 : (c2 - b2 * y) / a2;
    if (x < 0 || y < 0 || x + y > 1) return false;
    h.t = t; h.u = x; h.v = y;
    return true;
}

// Attribute interpolation of the accepted hit (hw8/src/primitives.cpp:110-124).
static inline Surface surface_of(const b200rt_triangle &f, const TriHit &h) {
    Surface s;
    float u = h.u, v = h.v;
    V n0(f.n[0]), n1(f.n[1]), n2(f.n[2]);
    V shading = n2 + u * (n0 - n2) + v * (n1 - n2);
    s.tx = f.uv[2][0] + u * (f.uv[0][0] - f.uv[2][0]) + v * (f.uv[1][0] - f.uv[2][0]);
    s.ty = f.uv[2][1] + u * (f.uv[0][1] - f.uv[2][1]) + v * (f.uv[1][1] - f.uv[2][1]);
    V t0(f.tan[0]), t1(f.tan[1]), t2(f.tan[2]);
    V tangent = t2 + u * (t0 - t2) + v * (t1 - t2);
    s.tangent = tangent.normalize();
    s.tangent_w = f.tan[0][3];
    shading = shading.normalize();
    if (h.inside) shading = -1. * shading;
    s.shadingNorma = shading;
    s.geomNorma = h.n.normalize();
    s.t = h.t;
    s.inside = h.inside;
    return s;
}

// AABB::intersect -> intersectBoxAndRay(half, ray - centre, false) (hw8/src/primitives.cpp:29-53,163-165)
static inline bool hit_aabb(const b200rt_bvh_node &nd, const Ray &ray, float &t, bool &inside) {
    V mn(nd.aabb_min), mx(nd.aabb_max);
    V s = 0.5 * (mx - mn);
    Ray r = ray.shifted(0.5 * (mn + mx));
    V ts1 = (-1. * s - r.o) / r.d;
    V ts2 = (s - r.o) / r.d;
    float t1x = std::min(ts1.x, ts2.x), t2x = std::max(ts1.x, ts2.x);
    float t1y = std::min(ts1.y, ts2.y), t2y = std::max(ts1.y, ts2.y);
    float t1z = std::min(ts1.z, ts2.z), t2z = std::max(ts1.z, ts2.z);
    float t1 = std::max(std::max(t1x, t1y), t1z);
    float t2 = std::min(std::min(t2x, t2y), t2z);
    if (t1 > t2 || t2 < 0) return false;
    if (t1 < 0) { inside = true; t = t2; } else { inside = false; t = t1; }
    return true;
}

struct Counters { uint64_t rays = 0, node_tests = 0, tri_tests = 0, light_queries = 0, light_node_tests = 0, light_tri_tests = 0; };

// Descriptions WITHOUT trees (the product's default load: b200rt_host_load_gltf skips the reference's SAH build).
// render_mc then traverses private trees over permuted COPIES of the triangles (closest hits and light-pdf sums do
// not depend on the tree) while FiguresMix::sample keeps picking from the description's own light order:
// g_light_sample points at it for the duration of a render (test infrastructure, one render at a time).
static const b200rt_triangle *g_light_sample = nullptr;
static inline const b200rt_triangle &light_sample_tri(const b200rt_scene_desc &s, int which) {
    return (g_light_sample ? g_light_sample : s.light_tris)[which];
}

template <class Rng>
struct Gltf {
    const b200rt_scene_desc &s;
    Rng &rng;
    Counters *cnt;
    bool hw8;

    Gltf(const b200rt_scene_desc &s_, Rng &r, Counters *c) : s(s_), rng(r), cnt(c), hw8(s_.mode == B200RT_MODE_HW8) {}

    // BVH::intersect_ (hw8/src/include/bvh.h:111-142), restated with the same control flow:
    // each call returns the best hit of ITS subtree; the bound `cur_best` is passed by value and
    // tightened only by the left sibling's result; a node is pruned when a strictly closer hit is
    // known and the origin is outside its box; strict '<' everywhere keeps the first candidate.
    bool closest(uint32_t pos, const Ray &ray, bool have, float cur_best, TriHit &out, int &out_id) const {
        const b200rt_bvh_node &nd = s.nodes[pos];
        float tb; bool inside;
        if (cnt) cnt->node_tests++;
        if (!hit_aabb(nd, ray, tb, inside)) return false;
        if (have && cur_best < tb && !inside) return false;
        bool found = false;
        if (nd.left == 0) {
            for (uint32_t i = nd.first; i < nd.last; i++) {
                TriHit h;
                if (cnt) cnt->tri_tests++;
                if (hit_triangle(s.tris[i], ray, h) && (!found || h.t < out.t)) { out = h; out_id = (int)i; found = true; }
            }
            return found;
        }
        found = closest(nd.left, ray, have, cur_best, out, out_id);
        if (found && (!have || out.t < cur_best)) { have = true; cur_best = out.t; }
        TriHit rh; int rid = -1;
        if (closest(nd.right, ray, have, cur_best, rh, rid) && (!found || rh.t < out.t)) { out = rh; out_id = rid; found = true; }
        return found;
    }

    // ---- textures (hw8/src/scene.cpp:9-35) ----
    static V texel(int ix, int iy, const b200rt_image &tex, bool srgb) {
        size_t offset = 3 * (ix + (size_t)tex.width * iy);
        V res = (1. / 255) * V{1.f * tex.rgb[offset], 1.f * tex.rgb[offset + 1], 1.f * tex.rgb[offset + 2]};
        if (srgb) return V{std::pow(res.x, 2.2f), std::pow(res.y, 2.2f), std::pow(res.z, 2.2f)};
        return res;
    }
    static V sample_texture(float tx, float ty, const b200rt_image &tex, bool srgb) {
        tx -= std::floor(tx);
        ty -= std::floor(ty);
        tx *= tex.width;
        ty *= tex.height;
        int ix1 = std::floor(tx), ix2 = (ix1 + 1) % tex.width;
        int iy1 = std::floor(ty), iy2 = (iy1 + 1) % tex.height;
        // the reference reads out of bounds when frac(u) * width rounds up to width; clamp instead
        ix1 = std::min(ix1, tex.width - 1); iy1 = std::min(iy1, tex.height - 1);
        float dx = tx - ix1;
        float dy = ty - iy1;
        V p11 = texel(ix1, iy1, tex, srgb), p12 = texel(ix1, iy2, tex, srgb);
        V p21 = texel(ix2, iy1, tex, srgb), p22 = texel(ix2, iy2, tex, srgb);
        return (1 - dx) * ((1 - dy) * p11 + dy * p12) + dx * ((1 - dy) * p21 + dy * p22);
    }
    static V apply_normal_map(V shadingNorma, V tangent, float w, V sample) {
        V localX = tangent, localZ = shadingNorma;
        V localY = w * localX.cross(localZ);
        V localNorma = 2.f * sample - V{1., 1., 1.};
        V norma = localNorma.x * localX + localNorma.y * localY + localNorma.z * localZ;
        return norma.normalize();
    }

    // ---- BRDF (hw8/src/include/material.h:11-65; hw7/src/include/material.h:13-59) ----
    static float distribution_term(const V &h, const V &n, float alpha2) {
        float dotHN = h.dot(n);
        if (dotHN <= 0) return 0;
        return alpha2 / (M_PI * pow(std::max(0.f, (alpha2 - 1) * dotHN * dotHN + 1), 2.0));
    }
    static float v1(const V &n, const V &x, float alpha2) {
        return 1. / (fabs(n.dot(x)) + sqrt(std::max(0.f, alpha2 + (1 - alpha2) * n.dot(x) * n.dot(x))));
    }
    static float specular_brdf(const V &l, const V &v, const V &n, float alpha2) {
        V h = (l + v).normalize();
        if (h.dot(l) < 1e-4 || h.dot(v) < 1e-4) return 0;
        return distribution_term(h, n, alpha2) * v1(n, l, alpha2) * v1(n, v, alpha2);
    }
    static V fresnel(const V &f0, const V &f90, const V &v, const V &h) {
        return f0 + pow(std::max<float>(0.f, 1.f - fabs(v.dot(h))), 5.0) * (f90 - f0);
    }
    V brdf(const b200rt_material &m, const V &l, const V &v, const V &n, const V &color, float metallic, float alpha) const {
        V base(m.base_color);
        V h = (l + v).normalize();
        float specular = specular_brdf(l, v, n, alpha * alpha);
        V metalBrdf, dielectricBrdf;
        if (hw8) {
            metallic *= m.metallic_factor;
            if (metallic > 0 && v.dot(n) >= 0 && l.dot(n) >= 0) metalBrdf = specular * fresnel(base * color, {1, 1, 1}, v, h);
            if (metallic < 1) {
                V diffuse;
                if (l.dot(n) >= 0) diffuse = (1. / M_PI) * (base * color);
                V ft = fresnel({0.04, 0.04, 0.04}, {1, 1, 1}, v, h);
                dielectricBrdf = diffuse * (V{1, 1, 1} - ft) + specular * ft;
            }
        } else {
            metallic = m.metallic_factor;
            if (metallic > 0) metalBrdf = specular * fresnel(base, {1, 1, 1}, v, h);
            if (metallic < 1) {
                V diffuse = (1. / M_PI) * base;
                V ft = fresnel({0.04, 0.04, 0.04}, {1, 1, 1}, v, h);
                dielectricBrdf = diffuse * (V{1, 1, 1} - ft) + specular * ft;
            }
        }
        return (1.0 - metallic) * dielectricBrdf + metallic * metalBrdf;
    }

    // ---- distributions (hw8/src/include/distributions.h:35-280) ----
    V cosine_sample(V n) {
        V d = rng.sphere();
        d = d + n;
        float len = d.len();
        if (len <= 1e-9f || d.dot(n) <= 1e-9f || std::isnan(len)) return n;
        return 1. / len * d;
    }
    static float cosine_pdf(V n, V d) { return std::max(0.f, d.dot(n) / (float)M_PI); }

    V vndf_sample_local(V v, float alpha_) {
        V vh = V(alpha_ * v.x, alpha_ * v.y, v.z).normalize();
        float lensq = vh.x * vh.x + vh.y * vh.y;
        V T1 = lensq > 0 ? 1. / sqrt(lensq) * V(-vh.y, vh.x, 0) : V(1, 0, 0);
        V T2 = T1.cross(vh);
        float u1 = rng.u(), u2 = rng.u();
        float r = sqrt(u1);
        float phi = 2.0 * M_PI * u2;
        float t1 = r * cos(phi);
        float t2 = r * sin(phi);
        float sgm = 0.5 * (1.0 + vh.z);
        t2 = (1.0 - sgm) * sqrt(1.f - t1 * t1) + sgm * t2;
        V nh = t1 * T1 + t2 * T2 + sqrt(std::max<float>(0.f, 1.0 - t1 * t1 - t2 * t2)) * vh;
        V ne = V(alpha_ * nh.x, alpha_ * nh.y, std::max<float>(0.0, nh.z)).normalize();
        return 2 * ne.dot(v) * ne - v;
    }
    static float vndf_D(V n, float a) {
        return 1. / (M_PI * a * a * pow(n.x * n.x / (a * a) + n.y * n.y / (a * a) + n.z * n.z, 2.0));
    }
    static float vndf_G1(V v, float a) {
        float lambda = 0.5 * (-1 + sqrt(1 + (a * a * v.x * v.x + a * a * v.y * v.y) / (v.z * v.z)));
        return 1. / (1 + lambda);
    }
    static float vndf_pdf_local(V d, V v, float a) {
        V ni = (v + d).normalize();
        float dv = vndf_G1(v, a) * std::max(0.f, v.dot(ni)) * vndf_D(ni, a) / fabs(v.z);
        float res = dv / (4 * v.dot(ni));
        return res;
    }
    static Q frame_q(V n) {
        V newN = {0, 0, 1};
        if (n.dot(newN) > 0.9999) return {};
        if (n.dot(newN) < -0.9999) return Q{0, 0, 0, -1};
        V a = n.cross(newN);
        float w = sqrt(n.len2()) + n.dot(newN);
        float len = sqrt(a.len2() + w * w);
        return Q{1. / len * a, w / len};
    }
    V vndf_sample(V n, V v, float alpha_) {
        v = -1. * v;
        Q q = frame_q(n);
        V vT = q.transform(v);
        V dT = vndf_sample_local(vT, alpha_);
        return q.conjugate().transform(dT);
    }
    static float vndf_pdf(V n, V d, V v, float alpha_) {
        v = -1. * v;
        Q q = frame_q(n);
        return vndf_pdf_local(q.transform(d), q.transform(v), alpha_);
    }

    // TriangleLight / FiguresMix (distributions.h:60-166)
    static float point_prob(const b200rt_triangle &f) {
        V a(f.p[2]); V b = V(f.p[0]) - a; V c = V(f.p[1]) - a;
        V n = b.cross(c);
        return 1.0 / (0.5 * n.len());
    }
    V light_sample(V x) {
        int which = rng.u() * (size_t)s.n_light_tris;
        const b200rt_triangle &f = light_sample_tri(s, which);
        V a(f.p[2]); V b = V(f.p[0]) - a; V c = V(f.p[1]) - a;
        float u = rng.u();
        float v = rng.u();
        if (u + v > 1.) { u = 1 - u; v = 1 - v; }
        V point = a + u * b + v * c;
        return (point - x).normalize();
    }
    float light_pdf_one(const b200rt_triangle &f, V x, V d) const {
        TriHit h;
        if (cnt) cnt->light_tri_tests++;
        if (!hit_triangle(f, Ray(x, d), h)) return 0.;
        Surface sf = surface_of(f, h);
        if (std::isnan(h.t)) return INFINITY;
        V y = x + h.t * d;
        V yn = hw8 ? sf.shadingNorma : sf.geomNorma;  // hw7 uses the geometric normal (hw7 distributions.h:145)
        float pp = point_prob(f);
        return pp * (x - y).len2() / fabs(d.dot(yn));
    }
    float light_total(uint32_t pos, V x, V d) const {
        const b200rt_bvh_node &nd = s.light_nodes[pos];
        float tb; bool inside;
        if (cnt) cnt->light_node_tests++;
        if (!hit_aabb(nd, Ray(x, d), tb, inside)) return 0;
        if (nd.left == 0) {
            float result = 0;
            for (uint32_t i = nd.first; i < nd.last; i++) result += light_pdf_one(s.light_tris[i], x, d);
            return result;
        }
        return light_total(nd.left, x, d) + light_total(nd.right, x, d);
    }
    float light_pdf(V x, V d) const {
        if (cnt) cnt->light_queries++;
        return light_total(0, x, d) / (size_t)s.n_light_tris;
    }

    size_t components() const { return s.n_light_tris > 0 ? 3 : 2; }
    V mix_sample(V x, V n, V v, float alpha) {
        int which = rng.u() * components();
        if (which == 0) return cosine_sample(n);
        if (which == 2) return light_sample(x);
        return vndf_sample(n, v, alpha);
    }
    float mix_pdf(V x, V n, V d, V v, float alpha) const {
        float ans = 0;
        ans += cosine_pdf(n, d);
        ans += vndf_pdf(n, d, v, alpha);
        if (s.n_light_tris > 0) ans += light_pdf(x, d);
        return ans / components();
    }

    V color(const Ray &ray, int rec) {
        if (rec == 0) return V(0., 0., 0.);
        TriHit th; int id = -1;
        if (cnt) cnt->rays++;
        if (!closest((uint32_t)s.root, ray, false, 0.f, th, id)) {
            if (!hw8 || !s.env_map) return V(s.bg_color);
            float tx = 0.5 + 0.5 * std::atan2(ray.d.z, ray.d.x) / M_PI;
            float ty = 0.5 - std::asin(ray.d.y) / M_PI;
            return sample_texture(tx, ty, *s.env_map, true);
        }
        rng.begin_bounce((uint32_t)(s.ray_depth - rec));
        const b200rt_triangle &f = s.tris[id];
        const b200rt_material &m = s.materials[f.material];
        Surface sf = surface_of(f, th);
        float t = sf.t;
        V geomNorma = sf.geomNorma, shadingNorma = sf.shadingNorma;
        V x = ray.o + t * ray.d;
        V emission(m.emission);
        V col{1, 1, 1};
        float alpha, metallic = 1;
        if (hw8) {
            if (m.base_color_image >= 0) col = sample_texture(sf.tx, sf.ty, s.images[m.base_color_image], true);
            if (m.emissive_image >= 0) emission = emission * sample_texture(sf.tx, sf.ty, s.images[m.emissive_image], true);
            V mr = {1, 1, 1};
            if (m.metallic_roughness_image >= 0) mr = sample_texture(sf.tx, sf.ty, s.images[m.metallic_roughness_image], false);
            V smp{0.5, 0.5, 1};
            if (m.normal_image >= 0) smp = sample_texture(sf.tx, sf.ty, s.images[m.normal_image], false);
            shadingNorma = apply_normal_map(shadingNorma, sf.tangent, sf.tangent_w, smp);
            alpha = pow(std::max(0.08f, m.roughness_factor * mr.y), 2.0);
            metallic = mr.z;
        } else {
            alpha = pow(m.roughness_factor, 2.0);
        }
        V d = mix_sample(x + eps * geomNorma, shadingNorma, ray.d, alpha);
        Ray dRay(x + eps * geomNorma, d);
        V f_r = brdf(m, dRay.d, -1. * ray.d, shadingNorma, col, metallic, alpha);
        if (f_r.x < eps && f_r.y < eps && f_r.z < eps) return emission;
        float pdf = mix_pdf(x + eps * geomNorma, shadingNorma, d, ray.d, alpha);
        auto mult = 1. / pdf * fabs(d.dot(shadingNorma)) * f_r;
        if (mult.x > 6 || mult.y > 6 || mult.z > 6 || std::isnan(mult.x) || std::isnan(mult.y) || std::isnan(mult.z)) return emission;
        return emission + mult * color(dRay, rec - 1);
    }

    Ray camera_ray(float x, float y) const {
        float tanFovY = tan(s.cam_fov / 2);
        float tanFovX = tanFovY * s.width / s.height;
        float nx = tanFovX * (2 * x / s.width - 1);
        float ny = tanFovY * (2 * y / s.height - 1);
        return {V(s.cam_pos), (nx * V(s.cam_right) - ny * V(s.cam_up) + V(s.cam_forward)).normalize()};
    }
    V pixel_sum(int x, int y, int b, int e) {
        V c{0, 0, 0};
        for (int k = b; k < e; k++) {
            rng.begin_sample((uint32_t)k);
            float nx = x + rng.u();
            float ny = y + rng.u();
            c = c + color(camera_ray(nx, ny), s.ray_depth);
        }
        return c;
    }
};

// =========================================== hw6 ============================================
// hw6 = the hw5 integrator (DIFFUSE / METALLIC mirror / DIELECTRIC, Mix{Cosine, FiguresMix}) on glTF
// triangles through the BVH (hw6/src/scene.cpp:8-126, hw6/src/primitives.cpp:76-85,143-164,
// hw6/src/include/distributions.h:36-58,107-141,167-302).  Triangles carry no shading normals.
struct Hit6 { float t = 0; V norma; bool inside = false; };

// Figure::intersect (position 0, identity rotation) -> intersectAsTriangle: plane hit + three edge tests.
static inline bool hit_triangle6(const b200rt_triangle &f, const Ray &ray, Hit6 &h) {
    V a(f.p[2]);
    V b = V(f.p[0]) - a;
    V c = V(f.p[1]) - a;
    V n = b.cross(c);
    Ray r = ray.shifted(a);
    float t = -r.o.dot(n) / r.d.dot(n);
    if (!(t > 0 && t < kTMax)) return false;
    V norma; bool inside;
    if (r.d.dot(n) > 0) { norma = -1. * n; inside = true; }
    else { norma = n; inside = false; }
    V p = ray.o - a + t * ray.d;
    if (b.cross(p).dot(n) < 0) return false;
    if (p.cross(c).dot(n) < 0) return false;
    if ((c - b).cross(p - b).dot(n) < 0) return false;
    h.t = t; h.norma = norma.normalize(); h.inside = inside;
    return true;
}

template <class Rng>
struct Hw6 {
    const b200rt_scene_desc &s;
    Rng &rng;
    Counters *cnt;

    Hw6(const b200rt_scene_desc &s_, Rng &r, Counters *c) : s(s_), rng(r), cnt(c) {}

    // BVH::intersect_ (hw6/src/include/bvh.h:111-142): same control flow as hw8 (see Gltf::closest)
    bool closest(uint32_t pos, const Ray &ray, bool have, float cur_best, Hit6 &out, int &out_id) const {
        const b200rt_bvh_node &nd = s.nodes[pos];
        float tb; bool inside;
        if (cnt) cnt->node_tests++;
        if (!hit_aabb(nd, ray, tb, inside)) return false;
        if (have && cur_best < tb && !inside) return false;
        bool found = false;
        if (nd.left == 0) {
            for (uint32_t i = nd.first; i < nd.last; i++) {
                Hit6 h;
                if (cnt) cnt->tri_tests++;
                if (hit_triangle6(s.tris[i], ray, h) && (!found || h.t < out.t)) { out = h; out_id = (int)i; found = true; }
            }
            return found;
        }
        found = closest(nd.left, ray, have, cur_best, out, out_id);
        if (found && (!have || out.t < cur_best)) { have = true; cur_best = out.t; }
        Hit6 rh; int rid = -1;
        if (closest(nd.right, ray, have, cur_best, rh, rid) && (!found || rh.t < out.t)) { out = rh; out_id = rid; found = true; }
        return found;
    }

    V cosine_sample(V n) {
        V d = rng.sphere();
        d = d + n;
        float len = d.len();
        if (len <= 1e-9f || d.dot(n) <= 1e-9f || std::isnan(len)) return n;
        return 1. / len * d;
    }
    static float cosine_pdf(V n, V d) { return std::max(0.f, d.dot(n) / PI); }

    // FiguresMix::sample -> TriangleLight::sample (distributions.h:127-141, 222-231)
    V light_sample(V x) {
        int which = rng.u() * (size_t)s.n_light_tris;
        const b200rt_triangle &f = light_sample_tri(s, which);
        V a(f.p[2]); V b = V(f.p[0]) - a; V c = V(f.p[1]) - a;
        float u = rng.u();
        float v = rng.u();
        if (u + v > 1.) { u = 1 - u; v = 1 - v; }
        V point = a + u * b + v * c;
        return (point - x).normalize();
    }
    // pdfOneFigureLight for a triangle (:242-262): first hit only
    float light_pdf_one(const b200rt_triangle &f, V x, V d) const {
        Hit6 h;
        if (cnt) cnt->light_tri_tests++;
        if (!hit_triangle6(f, Ray(x, d), h)) return 0.;
        if (std::isnan(h.t)) return INFINITY;
        V y = x + h.t * d;
        V a(f.p[2]); V b = V(f.p[0]) - a; V c = V(f.p[1]) - a;
        V n = b.cross(c);
        float pointProb = 1.0 / (0.5 * n.len());
        return pointProb * (x - y).len2() / fabs(d.dot(h.norma));
    }
    float light_total(uint32_t pos, V x, V d) const {
        const b200rt_bvh_node &nd = s.light_nodes[pos];
        float tb; bool inside;
        if (cnt) cnt->light_node_tests++;
        if (!hit_aabb(nd, Ray(x, d), tb, inside)) return 0;
        if (nd.left == 0) {
            float result = 0;
            for (uint32_t i = nd.first; i < nd.last; i++) result += light_pdf_one(s.light_tris[i], x, d);
            return result;
        }
        return light_total(nd.left, x, d) + light_total(nd.right, x, d);
    }
    size_t components() const { return s.n_light_tris > 0 ? 2 : 1; }
    V mix_sample(V x, V n) {
        int which = rng.u() * components();
        if (which == 0) return cosine_sample(n);
        return light_sample(x);
    }
    float mix_pdf(V x, V n, V d) const {
        float ans = 0;
        ans += cosine_pdf(n, d);
        if (s.n_light_tris > 0) {
            if (cnt) cnt->light_queries++;
            ans += light_total(0, x, d) / (size_t)s.n_light_tris;
        }
        return ans / components();
    }

    V color(const Ray &ray, int rec) {
        if (rec == 0) return V(0., 0., 0.);
        Hit6 h; int pos = -1;
        if (cnt) cnt->rays++;
        if (!closest((uint32_t)s.root, ray, false, 0.f, h, pos)) return V(s.bg_color);
        rng.begin_bounce((uint32_t)(s.ray_depth - rec));
        const b200rt_material &m = s.materials[s.tris[pos].material];
        float t = h.t; V norma = h.norma; bool is_inside = h.inside;
        V x = ray.o + t * ray.d;
        V emission(m.emission), fcolor(m.base_color);
        if (m.kind == B200RT_MAT_DIFFUSE) {
            V d = mix_sample(x + eps * norma, norma);
            if (d.dot(norma) < 0) return emission;
            float pdf = mix_pdf(x + eps * norma, norma, d);
            Ray dRay(x + eps * d, d);
            return emission + 1. / (PI * pdf) * d.dot(norma) * fcolor * color(dRay, rec - 1);
        } else if (m.kind == B200RT_MAT_METALLIC) {
            V reflectedDir = ray.d.normalize() - 2. * norma.dot(ray.d.normalize()) * norma;
            Ray reflected(ray.o + t * ray.d + eps * reflectedDir, reflectedDir);
            return emission + fcolor * color(reflected, rec - 1);
        } else {
            V reflectedDir = ray.d.normalize() - 2. * norma.dot(ray.d.normalize()) * norma;
            Ray reflected(ray.o + t * ray.d + eps * reflectedDir, reflectedDir);
            // reference: reflected branch traced before, and regardless of, the Fresnel coin (see Hw4::color)
            V reflectedColor;
            float coin = 0;
            if (Rng::kCounterBased) coin = rng.u();
            else reflectedColor = color(reflected, rec - 1);
            float eta1 = 1., eta2 = m.ior;
            if (is_inside) std::swap(eta1, eta2);
            V l = -1. * ray.d.normalize();
            float sinTheta2 = eta1 / eta2 * sqrt(1 - norma.dot(l) * norma.dot(l));
            if (fabs(sinTheta2) > 1.) {
                if (Rng::kCounterBased) reflectedColor = color(reflected, rec - 1);
                return emission + reflectedColor;
            }
            float r0 = pow((eta1 - eta2) / (eta1 + eta2), 2.);
            float r = r0 + (1 - r0) * pow(1 - norma.dot(l), 5.);
            if (!Rng::kCounterBased) coin = rng.u();
            if (coin < r) {
                if (Rng::kCounterBased) reflectedColor = color(reflected, rec - 1);
                return emission + reflectedColor;
            }
            float cosTheta2 = sqrt(1 - sinTheta2 * sinTheta2);
            V refractedDir = eta1 / eta2 * (-1. * l) + (eta1 / eta2 * norma.dot(l) - cosTheta2) * norma;
            Ray refracted(ray.o + t * ray.d + eps * refractedDir, refractedDir);
            V refractedColor = color(refracted, rec - 1);
            if (!is_inside) refractedColor = refractedColor * fcolor;
            return emission + refractedColor;
        }
    }

    Ray camera_ray(float x, float y) const {   // fovY based, direction NOT normalised (hw6/src/scene.cpp:118-126)
        float tanFovY = tan(s.cam_fov / 2);
        float tanFovX = tanFovY * s.width / s.height;
        float nx = tanFovX * (2 * x / s.width - 1);
        float ny = tanFovY * (2 * y / s.height - 1);
        return {V(s.cam_pos), nx * V(s.cam_right) - ny * V(s.cam_up) + V(s.cam_forward)};
    }
    V pixel_sum(int x, int y, int b, int e) {
        V c{0, 0, 0};
        for (int k = b; k < e; k++) {
            rng.begin_sample((uint32_t)k);
            float nx = x + rng.u();
            float ny = y + rng.u();
            c = c + color(camera_ray(nx, ny), s.ray_depth);
        }
        return c;
    }
};

// =========================================== hw5 ============================================
// Text scenes with TRIANGLE figures (hw5/src/*): planes are scanned, every other figure (ELLIPSOID, BOX, TRIANGLE,
// each with its own position / rotation) sits in a BVH keyed by Figure::position; the light mixture takes all
// three emissive kinds.  hw6 is this program with a glTF loader, so the integrator is Hw6's and the analytic
// lights are Hw4's; only the plumbing over b200rt_primitive is new.
//   Figure::intersect / intersectAsTriangle   hw5/src/primitives.cpp:13-33, 143-164
//   Scene::intersect                          hw5/src/scene.cpp:25-45
//   BVH::intersect_                           hw5/src/include/bvh.h:111-142
//   Box / Triangle / Ellipsoid lights         hw5/src/include/distributions.h:61-172
//   FiguresMix                                hw5/src/include/distributions.h:174-268
static inline Hit hit_figure5(const b200rt_primitive &f, const Ray &ray) {
    if (f.type != B200RT_PRIM_TRIANGLE) return hit_figure(f, ray, true);
    Q rot(f.rotation);
    Ray local = ray.shifted(V(f.position)).rotated(rot);
    V a(f.data3);
    V b = V(f.data) - a;
    V c = V(f.data2) - a;
    V n = b.cross(c);
    Hit h = hit_plane(n, local.shifted(a), true);
    if (!h.valid) return h;
    V p = local.o - a + h.t * local.d;
    h.valid = false;
    if (b.cross(p).dot(n) < 0) return h;
    if (p.cross(c).dot(n) < 0) return h;
    if ((c - b).cross(p - b).dot(n) < 0) return h;
    h.norma = rot.conjugate().transform(h.norma).normalize();
    h.valid = true;
    return h;
}

template <class Rng>
struct Hw5 {
    const b200rt_scene_desc &s;
    Rng &rng;
    Counters *cnt;

    Hw5(const b200rt_scene_desc &s_, Rng &r, Counters *c) : s(s_), rng(r), cnt(c) {}

    static bool hit_box_of(const b200rt_bvh_node &nd, const Ray &ray, float &t, bool &inside) { return hit_aabb(nd, ray, t, inside); }

    bool closest_bvh(uint32_t pos, const Ray &ray, bool have, float cur_best, Hit &out, int &out_id) const {
        const b200rt_bvh_node &nd = s.nodes[pos];
        float tb; bool inside;
        if (cnt) cnt->node_tests++;
        if (!hit_aabb(nd, ray, tb, inside)) return false;
        if (have && cur_best < tb && !inside) return false;
        bool found = false;
        if (nd.left == 0) {
            for (uint32_t i = nd.first; i < nd.last; i++) {
                if (cnt) cnt->tri_tests++;
                Hit h = hit_figure5(s.prims[i], ray);
                if (h.valid && (!found || h.t < out.t)) { out = h; out_id = (int)i; found = true; }
            }
            return found;
        }
        found = closest_bvh(nd.left, ray, have, cur_best, out, out_id);
        if (found && (!have || out.t < cur_best)) { have = true; cur_best = out.t; }
        Hit rh; int rid = -1;
        if (closest_bvh(nd.right, ray, have, cur_best, rh, rid) && (!found || rh.t < out.t)) { out = rh; out_id = rid; found = true; }
        return found;
    }
    // Scene::intersect: the planes first, then the BVH with the planes' best distance as the prune bound
    bool intersect(const Ray &ray, Hit &best, int &pos) const {
        pos = -1;
        for (int i = s.n_bvh_prims; i < s.n_prims; i++) {
            Hit h = hit_figure5(s.prims[i], ray);
            if (h.valid && (pos == -1 || h.t < best.t)) { best = h; pos = i; }
        }
        Hit bh; int bid = -1;
        if (closest_bvh((uint32_t)s.root, ray, pos != -1, pos != -1 ? best.t : 0.f, bh, bid) && (pos == -1 || bh.t < best.t)) { best = bh; pos = bid; }
        return pos != -1;
    }

    V cosine_sample(V n) {
        V d = rng.sphere();
        d = d + n;
        float len = d.len();
        if (len <= 1e-9f || d.dot(n) <= 1e-9f || std::isnan(len)) return n;
        return 1. / len * d;
    }
    static float cosine_pdf(V n, V d) { return std::max(0.f, d.dot(n) / PI); }

    // pdfOne of the three light kinds
    static float pdf_one(const b200rt_primitive &f, V x, V d, V y, V yn) {
        if (f.type == B200RT_PRIM_BOX) {
            float sx = f.data[0], sy = f.data[1], sz = f.data[2];
            float sTotal = 8 * (sy * sz + sx * sz + sx * sy);
            return (x - y).len2() / (sTotal * fabs(d.dot(yn)));
        }
        if (f.type == B200RT_PRIM_ELLIPSOID) {
            V r(f.data);
            V n = Q(f.rotation).transform(y - V(f.position)) / r;
            float pointProb = 1. / (4 * PI * V{n.x * r.y * r.z, r.x * n.y * r.z, r.x * r.y * n.z}.len());
            return pointProb * (x - y).len2() / fabs(d.dot(yn));
        }
        V a(f.data3); V b = V(f.data) - a; V c = V(f.data2) - a;
        V n = b.cross(c);
        float pointProb = 1.0 / (0.5 * n.len());
        return pointProb * (x - y).len2() / fabs(d.dot(yn));
    }
    // FiguresMix::pdfOneFigureLight (distributions.h:218-250): triangles count their one hit, closed figures also
    // the second hit from just behind the first
    float light_pdf_one(const b200rt_primitive &f, V x, V d) const {
        if (cnt) cnt->light_tri_tests++;
        Hit h1 = hit_figure5(f, Ray(x, d));
        if (!h1.valid) return 0.;
        if (std::isnan(h1.t)) return INFINITY;
        V y = x + h1.t * d;
        float ans = pdf_one(f, x, d, y, h1.norma);
        if (f.type == B200RT_PRIM_TRIANGLE) return ans;
        Hit h2 = hit_figure5(f, Ray(x + (h1.t + eps) * d, d));
        if (!h2.valid) return ans;
        V y2 = x + (h1.t + eps + h2.t) * d;
        return ans + pdf_one(f, x, d, y2, h2.norma);
    }
    float light_total(uint32_t pos, V x, V d) const {
        const b200rt_bvh_node &nd = s.light_nodes[pos];
        float tb; bool inside;
        if (cnt) cnt->light_node_tests++;
        if (!hit_aabb(nd, Ray(x, d), tb, inside)) return 0;
        if (nd.left == 0) {
            float result = 0;
            for (uint32_t i = nd.first; i < nd.last; i++) result += light_pdf_one(s.light_prims[i], x, d);
            return result;
        }
        return light_total(nd.left, x, d) + light_total(nd.right, x, d);
    }
    // FiguresMix::sample: index, then the kind's own sampler
    V light_sample(V x) {
        int which = rng.u() * (size_t)s.n_light_prims;
        const b200rt_primitive &f = s.light_prims[which];
        Q rot(f.rotation);
        if (f.type == B200RT_PRIM_TRIANGLE) {
            V a(f.data3); V b = V(f.data) - a; V c = V(f.data2) - a;
            float u = rng.u();
            float v = rng.u();
            if (u + v > 1.) { u = 1 - u; v = 1 - v; }
            V point = V(f.position) + rot.conjugate().transform(a + u * b + v * c);
            return (point - x).normalize();
        }
        for (int iter = 0;; iter++) {
            V point;
            if (f.type == B200RT_PRIM_BOX) {
                float sx = f.data[0], sy = f.data[1], sz = f.data[2];
                float wx = sy * sz, wy = sx * sz, wz = sx * sy;
                float u = rng.u() * (wx + wy + wz);
                float flipSign = rng.u() > 0.5 ? 1 : -1;
                // Vec3(...) is a constructor call: g++ evaluates its arguments right to left, so the LATER coordinate
                // takes the earlier random number (pinned by the byte-equality test against the hw5 binary)
                if (u < wx) { float b = rng.u(), a = rng.u(); point = V(flipSign * sx, (2 * a - 1) * sy, (2 * b - 1) * sz); }
                else if (u < wx + wy) { float b = rng.u(), a = rng.u(); point = V((2 * a - 1) * sx, flipSign * sy, (2 * b - 1) * sz); }
                else { float b = rng.u(), a = rng.u(); point = V((2 * a - 1) * sx, (2 * b - 1) * sy, flipSign * sz); }
            } else {
                point = V(f.data) * rng.sphere();
            }
            V actual = rot.conjugate().transform(point) + V(f.position);
            V dir = (actual - x).normalize();
            if (hit_figure5(f, Ray(x, dir)).valid) return dir;
            if (Rng::kCounterBased && iter + 1 >= kMaxRejection) return dir;
        }
    }
    size_t components() const { return s.n_light_prims > 0 ? 2 : 1; }
    V mix_sample(V x, V n) {
        int which = rng.u() * components();
        if (which == 0) return cosine_sample(n);
        return light_sample(x);
    }
    float mix_pdf(V x, V n, V d) const {
        float ans = 0;
        ans += cosine_pdf(n, d);
        if (s.n_light_prims > 0) {
            if (cnt) cnt->light_queries++;
            ans += light_total(0, x, d) / (size_t)s.n_light_prims;
        }
        return ans / components();
    }

    // Scene::getColor (hw5/src/scene.cpp:47-105)
    V color(const Ray &ray, int rec) {
        if (rec == 0) return V(0., 0., 0.);
        Hit h; int pos = -1;
        if (cnt) cnt->rays++;
        if (!intersect(ray, h, pos)) return V(s.bg_color);
        rng.begin_bounce((uint32_t)(s.ray_depth - rec));
        const b200rt_primitive &f = s.prims[pos];
        float t = h.t; V norma = h.norma; bool is_inside = h.inside;
        V x = ray.o + t * ray.d;
        V emission(f.emission), fcolor(f.color);
        if (f.material == B200RT_MAT_DIFFUSE) {
            V d = mix_sample(x + eps * norma, norma);
            if (d.dot(norma) < 0) return emission;
            float pdf = mix_pdf(x + eps * norma, norma, d);
            Ray dRay(x + eps * d, d);
            return emission + 1. / (PI * pdf) * d.dot(norma) * fcolor * color(dRay, rec - 1);
        } else if (f.material == B200RT_MAT_METALLIC) {
            V reflectedDir = ray.d.normalize() - 2. * norma.dot(ray.d.normalize()) * norma;
            Ray reflected(ray.o + t * ray.d + eps * reflectedDir, reflectedDir);
            return emission + fcolor * color(reflected, rec - 1);
        } else {
            V reflectedDir = ray.d.normalize() - 2. * norma.dot(ray.d.normalize()) * norma;
            Ray reflected(ray.o + t * ray.d + eps * reflectedDir, reflectedDir);
            // reference: reflected branch traced before, and regardless of, the Fresnel coin (see Hw4::color)
            V reflectedColor;
            float coin = 0;
            if (Rng::kCounterBased) coin = rng.u();
            else reflectedColor = color(reflected, rec - 1);
            float eta1 = 1., eta2 = f.ior;
            if (is_inside) std::swap(eta1, eta2);
            V l = -1. * ray.d.normalize();
            float sinTheta2 = eta1 / eta2 * sqrt(1 - norma.dot(l) * norma.dot(l));
            if (fabs(sinTheta2) > 1.) {
                if (Rng::kCounterBased) reflectedColor = color(reflected, rec - 1);
                return emission + reflectedColor;
            }
            float r0 = pow((eta1 - eta2) / (eta1 + eta2), 2.);
            float r = r0 + (1 - r0) * pow(1 - norma.dot(l), 5.);
            if (!Rng::kCounterBased) coin = rng.u();
            if (coin < r) {
                if (Rng::kCounterBased) reflectedColor = color(reflected, rec - 1);
                return emission + reflectedColor;
            }
            float cosTheta2 = sqrt(1 - sinTheta2 * sinTheta2);
            V refractedDir = eta1 / eta2 * (-1. * l) + (eta1 / eta2 * norma.dot(l) - cosTheta2) * norma;
            Ray refracted(ray.o + t * ray.d + eps * refractedDir, refractedDir);
            V refractedColor = color(refracted, rec - 1);
            if (!is_inside) refractedColor = refractedColor * fcolor;
            return emission + refractedColor;
        }
    }

    Ray camera_ray(float x, float y) const {   // fovX based, jitter only (hw5/src/scene.cpp:119-126)
        float tanFovX = tan(s.cam_fov / 2);
        float tanFovY = tanFovX * s.height / s.width;
        float nx = tanFovX * (2 * x / s.width - 1);
        float ny = tanFovY * (2 * y / s.height - 1);
        return {V(s.cam_pos), nx * V(s.cam_right) - ny * V(s.cam_up) + V(s.cam_forward)};
    }
    V pixel_sum(int x, int y, int b, int e) {
        V c{0, 0, 0};
        for (int k = b; k < e; k++) {
            rng.begin_sample((uint32_t)k);
            float nx = x + rng.u();
            float ny = y + rng.u();
            c = c + color(camera_ray(nx, ny), s.ray_depth);
        }
        return c;
    }
};

}  // namespace mc

// ------------------------------------------------------------------------------------------
// Drivers.  `accum` receives the SUM over samples [spp_begin, spp_end) per pixel (what
// b200rt_download_accum returns); `rgb` the quantised image using mean = float(1.0/n) * sum
// (hw8/src/scene.cpp:176).
template <class Sc, class Rng>
static void render_per_pixel_rng(const b200rt_scene_desc &s, uint32_t seed, int b, int e, float *accum, uint8_t *rgb,
                                 mc::Counters *cnt) {
    const int n = s.width * s.height;
    mc::Counters total;
#pragma omp parallel
    {
        mc::Counters local;
#pragma omp for schedule(dynamic, 8)
        for (int i = 0; i < n; i++) {
            Rng rng;
            rng.begin_pixel((uint32_t)i, seed);   // hw8/src/sceneio.cpp:389: rng_type rng(i)
            Sc sc(s, rng, cnt ? &local : nullptr);
            V sum = sc.pixel_sum(i % s.width, i / s.width, b, e);
            if (accum) { accum[3 * i] = sum.x; accum[3 * i + 1] = sum.y; accum[3 * i + 2] = sum.z; }
            if (rgb) mc::quantise(1.0 / (e - b) * sum, rgb + 3 * (size_t)i);
        }
#pragma omp critical
        {
            total.rays += local.rays; total.node_tests += local.node_tests; total.tri_tests += local.tri_tests;
            total.light_queries += local.light_queries; total.light_node_tests += local.light_node_tests;
            total.light_tri_tests += local.light_tri_tests;
        }
    }
    if (cnt) *cnt = total;
}

template <class Rng>
struct Hw4Adapter : mc::Hw4<Rng> {
    Hw4Adapter(const b200rt_scene_desc &s, Rng &r, mc::Counters *) : mc::Hw4<Rng>(s, r) {}
};

// Median-split tree in the reference's node format (parents before children, leaves of <= 4 triangles) over
// tris[first, last), permuting them in place.  Only for descriptions that carry no tree of their own.
static uint32_t private_tree(std::vector<b200rt_triangle> &tris, std::vector<b200rt_bvh_node> &nodes, uint32_t first, uint32_t last) {
    b200rt_bvh_node nd{};
    nd.first = first; nd.last = last;
    float cmn[3] = {INFINITY, INFINITY, INFINITY}, cmx[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int a = 0; a < 3; a++) { nd.aabb_min[a] = INFINITY; nd.aabb_max[a] = -INFINITY; }
    auto centroid = [](const b200rt_triangle &t, int a) { return t.p[0][a] + t.p[1][a] + t.p[2][a]; };
    for (uint32_t i = first; i < last; i++)
        for (int a = 0; a < 3; a++) {
            for (int v = 0; v < 3; v++) { nd.aabb_min[a] = std::min(nd.aabb_min[a], tris[i].p[v][a]); nd.aabb_max[a] = std::max(nd.aabb_max[a], tris[i].p[v][a]); }
            cmn[a] = std::min(cmn[a], centroid(tris[i], a)); cmx[a] = std::max(cmx[a], centroid(tris[i], a));
        }
    const uint32_t pos = (uint32_t)nodes.size();
    nodes.push_back(nd);
    if (last - first <= 4) return pos;
    int axis = 0;
    for (int a = 1; a < 3; a++) if (cmx[a] - cmn[a] > cmx[axis] - cmn[axis]) axis = a;
    const uint32_t mid = first + (last - first) / 2;
    std::nth_element(tris.begin() + first, tris.begin() + mid, tris.begin() + last,
                     [&](const b200rt_triangle &l, const b200rt_triangle &r) { return centroid(l, axis) < centroid(r, axis); });
    const uint32_t l = private_tree(tris, nodes, first, mid);
    const uint32_t r = private_tree(tris, nodes, mid, last);
    nodes[pos].left = l; nodes[pos].right = r;
    return pos;
}

void render_mc(const b200rt_scene_desc &s_in, int rng_kind, uint32_t seed, int b, int e, float *accum, uint8_t *rgb,
               uint64_t *counters) {
    b200rt_scene_desc s = s_in;
    std::vector<b200rt_triangle> own_tris, own_ltris;
    std::vector<b200rt_bvh_node> own_nodes, own_lnodes;
    mc::g_light_sample = nullptr;
    if (s.mode >= B200RT_MODE_HW6) {
        if (s.n_tris > 0 && !s.nodes) {
            own_tris.assign(s.tris, s.tris + s.n_tris);
            s.root = (int32_t)private_tree(own_tris, own_nodes, 0, (uint32_t)s.n_tris);
            s.tris = own_tris.data(); s.nodes = own_nodes.data(); s.n_nodes = (int32_t)own_nodes.size();
        }
        if (s.n_light_tris > 0 && !s.light_nodes) {
            own_ltris.assign(s.light_tris, s.light_tris + s.n_light_tris);
            private_tree(own_ltris, own_lnodes, 0, (uint32_t)s.n_light_tris);      // root = node 0
            mc::g_light_sample = s_in.light_tris;
            s.light_tris = own_ltris.data(); s.light_nodes = own_lnodes.data(); s.n_light_nodes = (int32_t)own_lnodes.size();
        }
    }
    mc::Counters cnt;
    mc::Counters *cp = counters ? &cnt : nullptr;
    if (s.mode == B200RT_MODE_HW3 || s.mode == B200RT_MODE_HW4) {
        if (rng_kind == 0) {
            // hw4: one global generator consumed in scan order (hw4/src/scene.cpp:5-6) => serial.
            mc::Hw4Minstd rng;
            mc::Hw4<mc::Hw4Minstd> sc(s, rng);
            rng.light_n01.assign(sc.lights.size(), std::normal_distribution<float>(0.f, 1.f));
            for (int y = 0; y < s.height; y++)
                for (int x = 0; x < s.width; x++) {
                    size_t i = (size_t)y * s.width + x;
                    V sum = sc.pixel_sum(x, y, b, e);
                    if (accum) { accum[3 * i] = sum.x; accum[3 * i + 1] = sum.y; accum[3 * i + 2] = sum.z; }
                    if (rgb) mc::quantise(1.0 / (e - b) * sum, rgb + 3 * i);
                }
        } else {
            render_per_pixel_rng<Hw4Adapter<mc::Hw4Philox>, mc::Hw4Philox>(s, seed, b, e, accum, rgb, nullptr);
        }
    } else if (s.mode == B200RT_MODE_HW5) {
        if (rng_kind == 0) render_per_pixel_rng<mc::Hw5<RngMinstd>, RngMinstd>(s, seed, b, e, accum, rgb, cp);
        else render_per_pixel_rng<mc::Hw5<RngPhilox>, RngPhilox>(s, seed, b, e, accum, rgb, cp);
    } else if (s.mode == B200RT_MODE_HW6) {
        if (rng_kind == 0) render_per_pixel_rng<mc::Hw6<RngMinstd>, RngMinstd>(s, seed, b, e, accum, rgb, cp);
        else render_per_pixel_rng<mc::Hw6<RngPhilox>, RngPhilox>(s, seed, b, e, accum, rgb, cp);
    } else {
        if (rng_kind == 0) render_per_pixel_rng<mc::Gltf<RngMinstd>, RngMinstd>(s, seed, b, e, accum, rgb, cp);
        else render_per_pixel_rng<mc::Gltf<RngPhilox>, RngPhilox>(s, seed, b, e, accum, rgb, cp);
    }
    if (counters) {
        counters[0] = cnt.rays; counters[1] = cnt.node_tests; counters[2] = cnt.tri_tests;
        counters[3] = cnt.light_queries; counters[4] = cnt.light_node_tests; counters[5] = cnt.light_tri_tests;
    }
}

// Resolve a float SUM buffer the way the reference does per pixel (used to check K5).
void resolve_mc(const float *sum, int n_pixels, int total_spp, uint8_t *rgb) {
    for (int i = 0; i < n_pixels; i++)
        mc::quantise(1.0 / total_spp * V(sum[3 * i], sum[3 * i + 1], sum[3 * i + 2]), rgb + 3 * (size_t)i);
}

}  // namespace port
