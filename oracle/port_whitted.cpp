// TEST INFRASTRUCTURE ONLY -- CPU restatement of the DETERMINISTIC configs:
//   hw1 flat-colour primary cast   hw1/src/scene.cpp:7-30, hw1/src/primitives.cpp:24-91, hw1/src/color.cpp:13-19
//   hw2 Whitted recursion          hw2/src/scene.cpp:8-99, hw2/src/light_source.cpp:9-23, hw2/src/color.cpp:5-31
// This TU includes <math.h> like every hw1/hw2 source file does, so tan/pow/sqrt on floats
// bind to the float overloads (tanf, powf, ...) exactly as in those programs.
// Pinned by tests/test_oracle.py: the PPM payload must equal oracle/_ref/hw1 and hw2 byte for byte.
#include <math.h>
#include <algorithm>
#include <type_traits>
#include <vector>
#include "port_common.h"

static_assert(std::is_same<decltype(tan(1.f)), float>::value, "<math.h> environment expected (hw1/hw2 TUs)");

namespace port {
namespace whitted {

#include "port_figures.inl"

struct Cam {
    const b200rt_scene_desc &s;
    // hw1/src/scene.cpp:22-30, hw2/src/scene.cpp:91-99: pixel centre +0.5 in double, fovX based.
    Ray ray(int x, int y) const {
        float tanFovX = tan(s.cam_fov / 2);
        float tanFovY = tanFovX * s.height / s.width;
        float nx = tanFovX * (2 * (x + 0.5) / s.width - 1);
        float ny = tanFovY * (2 * (y + 0.5) / s.height - 1);
        return {V(s.cam_pos), nx * V(s.cam_right) - ny * V(s.cam_up) + V(s.cam_forward)};
    }
};

// ---- hw1 ---------------------------------------------------------------------------------
// hw1's intersectors return only t; its plane normal is not normalised and there is no
// normal transform at all (hw1/src/primitives.cpp:24-31).
static bool hw1_hit(const b200rt_primitive &f, const Ray &ray, float &t) {
    Ray local = ray.shifted(V(f.position)).rotated(Q(f.rotation));
    if (f.type == B200RT_PRIM_ELLIPSOID) {
        V r(f.data);
        float c = (local.o / r).len2() - 1;
        float b = 2 * (local.o / r).dot(local.d / r);
        float a = (local.d / r).len2();
        bool inside;
        return smallest_positive_root(a, b, c, t, inside);
    }
    if (f.type == B200RT_PRIM_PLANE) {
        V n(f.data);
        t = -(local.o.dot(n)) / (local.d.dot(n));
        return t > 0;
    }
    V s(f.data);
    V ts1 = (-1. * s - local.o) / local.d;
    V ts2 = (s - local.o) / local.d;
    float t1x = std::min(ts1.x, ts2.x), t2x = std::max(ts1.x, ts2.x);
    float t1y = std::min(ts1.y, ts2.y), t2y = std::max(ts1.y, ts2.y);
    float t1z = std::min(ts1.z, ts2.z), t2z = std::max(ts1.z, ts2.z);
    float t1 = std::max(std::max(t1x, t1y), t1z);
    float t2 = std::min(std::min(t2x, t2y), t2z);
    if (t1 > t2 || t2 < 0) return false;
    t = t1 < 0 ? t2 : t1;
    return true;
}

static void render_hw1(const b200rt_scene_desc &s, uint8_t *rgb, int32_t *ids, float *accum) {
    Cam cam{s};
    for (int y = 0; y < s.height; y++)
        for (int x = 0; x < s.width; x++) {
            V ans(s.bg_color);
            float best = -1;
            int best_id = -1;
            for (int i = 0; i < s.n_prims; i++) {
                float t;
                if (hw1_hit(s.prims[i], cam.ray(x, y), t) && (best == -1 || t < best)) {
                    best = t; best_id = i; ans = V(s.prims[i].color);
                }
            }
            size_t p = (size_t)y * s.width + x;
            if (ids) ids[p] = best_id;
            if (accum) { accum[3 * p] = ans.x; accum[3 * p + 1] = ans.y; accum[3 * p + 2] = ans.z; }
            if (rgb) {
                rgb[3 * p] = round(255 * ans.x);
                rgb[3 * p + 1] = round(255 * ans.y);
                rgb[3 * p + 2] = round(255 * ans.z);
            }
        }
}

// ---- hw2 ---------------------------------------------------------------------------------
struct Hw2 {
    const b200rt_scene_desc &s;

    // hw2/src/scene.cpp:8-27: keeps t <= tmax and the strictly smallest t (first wins ties).
    bool intersect(const Ray &ray, float tmax, Hit &best, int &pos) const {
        pos = -1;
        for (int i = 0; i < s.n_prims; i++) {
            Hit h = hit_figure(s.prims[i], ray, false);
            if (h.valid && h.t <= tmax && (pos == -1 || h.t < best.t)) { best = h; pos = i; }
        }
        return pos != -1;
    }

    // hw2/src/light_source.cpp:9-23
    void light_at(const b200rt_light &L, V point, V &l, V &c, float &tmax) const {
        if (L.type == B200RT_LIGHT_POINT) {
            V direction = V(L.v) - point;
            float r = direction.len();
            V att(L.attenuation);
            c = 1. / (att.x + att.y * r + att.z * r * r) * V(L.intensity);
            l = direction.normalize();
            tmax = r;
        } else {
            l = V(L.v).normalize();
            c = V(L.intensity);
            tmax = 1. / 0.;
        }
    }

    V color(const Ray &ray, int rec, int *primary_id = nullptr) const {
        if (primary_id) *primary_id = -1;
        if (rec == 0) return V(0., 0., 0.);
        Hit h; int pos;
        if (!intersect(ray, 1. / 0., h, pos)) return V(s.bg_color);
        if (primary_id) *primary_id = pos;
        const b200rt_primitive &f = s.prims[pos];
        float t = h.t; V norma = h.norma; bool is_inside = h.inside;
        if (f.material == B200RT_MAT_DIFFUSE) {
            V col(s.ambient);
            for (int i = 0; i < s.n_lights; i++) {
                V p = ray.o + t * ray.d;
                V l, c; float tmax;
                light_at(s.lights[i], p, l, c, tmax);
                float reflected = l.dot(norma);
                Hit sh; int sp;
                if (reflected >= 0 && !intersect(Ray(p + 0.0001 * l, l), tmax, sh, sp)) col = col + reflected * c;
            }
            return col * V(f.color);
        } else if (f.material == B200RT_MAT_METALLIC) {
            V reflectedDir = ray.d.normalize() - 2. * norma.dot(ray.d.normalize()) * norma;
            Ray reflected(ray.o + t * ray.d + 0.0001 * reflectedDir, reflectedDir);
            return V(f.color) * color(reflected, rec - 1);
        } else {
            V reflectedDir = ray.d.normalize() - 2. * norma.dot(ray.d.normalize()) * norma;
            Ray reflected(ray.o + t * ray.d + 0.0001 * reflectedDir, reflectedDir);
            V reflectedComponent = color(reflected, rec - 1);
            float eta1 = 1., eta2 = f.ior;
            if (is_inside) std::swap(eta1, eta2);
            V l = -1. * ray.d.normalize();
            float sinTheta2 = eta1 / eta2 * sqrt(1 - norma.dot(l) * norma.dot(l));
            if (fabs(sinTheta2) > 1.) return reflectedComponent;
            float cosTheta2 = sqrt(1 - sinTheta2 * sinTheta2);
            V refractedDir = eta1 / eta2 * (-1. * l) + (eta1 / eta2 * norma.dot(l) - cosTheta2) * norma;
            Ray refracted(ray.o + t * ray.d + 0.0001 * refractedDir, refractedDir);
            V refractedComponent = color(refracted, rec - 1);
            if (!is_inside) refractedComponent = refractedComponent * V(f.color);
            float r0 = pow((eta1 - eta2) / (eta1 + eta2), 2.);
            float r = r0 + (1 - r0) * pow(1 - norma.dot(l), 5.);
            return r * reflectedComponent + (1 - r) * refractedComponent;
        }
    }
};

// hw2/src/color.cpp:5-31 under <math.h>: gamma via powf.
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

static void render_hw2(const b200rt_scene_desc &s, uint8_t *rgb, int32_t *ids, float *accum) {
    Cam cam{s};
    Hw2 sc{s};
#pragma omp parallel for schedule(dynamic, 8)
    for (int i = 0; i < s.width * s.height; i++) {
        int x = i % s.width, y = i / s.width, id;
        V c = sc.color(cam.ray(x, y), s.ray_depth, &id);
        if (ids) ids[i] = id;
        if (accum) { accum[3 * i] = c.x; accum[3 * i + 1] = c.y; accum[3 * i + 2] = c.z; }
        if (rgb) {
            V o = gamma_corrected(aces(c));
            rgb[3 * i] = round(255 * o.x);
            rgb[3 * i + 1] = round(255 * o.y);
            rgb[3 * i + 2] = round(255 * o.z);
        }
    }
}

}  // namespace whitted

void render_deterministic(const b200rt_scene_desc &s, uint8_t *rgb, int32_t *ids, float *accum) {
    if (s.mode == B200RT_MODE_HW1) whitted::render_hw1(s, rgb, ids, accum);
    else whitted::render_hw2(s, rgb, ids, accum);
}

}  // namespace port
