// TEST INFRASTRUCTURE ONLY -- CPU restatement ("port") of the reference's hot path.
// Nothing in the product may include, link or execute this (see oracle/README.md).
//
// Shared pieces of the port: the float vector type with the reference's operator set and
// promotion behaviour (hw8/src/include/vec3.h:5-80 -- cross() is b x a, normalize() is
// float(1.0/len) * v), the quaternion (hw8/src/include/quaternion.h:26-51), and the two
// RNG policies:
//   * RngMinstd : std::minstd_rand + libstdc++ distributions, i.e. the reference's own
//                 generator (hw8/src/sceneio.cpp:389, hw8/src/scene.cpp:168-169).  With it the
//                 port must reproduce the reference binary's PPM => that is how the port is pinned.
//   * RngPhilox : Philox4x32-10 counter-based stream, the generator the CUDA kernels use
//                 (DESIGN.md section 5).  With it the port predicts the GPU image sample for sample.
// IMPORTANT: math calls are written UNQUALIFIED on purpose; which overload they bind to
// (float vs double) depends on whether the including TU saw <math.h> (hw1, hw2) or only
// <cmath> (hw3..hw8), exactly like the reference's translation units.
#pragma once
#include <cstdint>
#include <random>
#include "../include/b200rt.h"

namespace port {

struct V {
    float x = 0, y = 0, z = 0;
    V() {}
    V(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    explicit V(const float *p) : x(p[0]), y(p[1]), z(p[2]) {}
    V operator+(const V &o) const { return {x + o.x, y + o.y, z + o.z}; }
    V operator-(const V &o) const { return {x - o.x, y - o.y, z - o.z}; }
    V operator/(const V &o) const { return {x / o.x, y / o.y, z / o.z}; }
    V operator*(const V &o) const { return {x * o.x, y * o.y, z * o.z}; }
    V operator+(float v) const { return {x + v, y + v, z + v}; }
    float dot(const V &o) const { return x * o.x + y * o.y + z * o.z; }
    V cross(const V &o) const { return {z * o.y - y * o.z, x * o.z - z * o.x, y * o.x - x * o.y}; }
    float len2() const { return dot(*this); }
    float len() const { return sqrt(len2()); }
    V normalize() const;
};
inline V operator*(float k, const V &p) { return {k * p.x, k * p.y, k * p.z}; }
inline V V::normalize() const { return 1. / len() * (*this); }

struct Q {
    V v;
    float w = 1;
    Q() {}
    Q(float x, float y, float z, float w_) : v(x, y, z), w(w_) {}
    Q(V v_, float w_) : v(v_), w(w_) {}
    explicit Q(const float *p) : v(p[0], p[1], p[2]), w(p[3]) {}
    Q operator*(const Q &o) const { return {w * o.v + o.w * v + v.cross(o.v), w * o.w - v.dot(o.v)}; }
    Q conjugate() const { return {(-1.) * v, w}; }
    V transform(const V &p) const { return ((*this) * Q(p.x, p.y, p.z, 0.) * conjugate()).v; }
};

struct Ray {
    V o, d;
    Ray() {}
    Ray(V o_, V d_) : o(o_), d(d_) {}
    Ray shifted(const V &by) const { return {o - by, d}; }
    Ray rotated(const Q &q) const { return {q.transform(o), q.transform(d)}; }
};

// ---- Philox4x32-10 (Salmon et al. 2011), key = (pixel, seed), counter = (sample, block, 0, 0) ----
inline void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]) {
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// Stream layout shared with the CUDA kernels: per camera sample, block 0 holds the pixel
// jitter; bounce b draws from blocks 16*(b+1) + k, k = 0, 1, ... (k > 0 only for hw4's
// rejection loops).  u01 = (x >> 8) * 2^-24 in [0, 1).
struct RngPhilox {
    uint32_t pixel = 0, seed = 0, sample = 0, block = 0;
    uint32_t buf[4];
    int idx = 4;
    void begin_pixel(uint32_t pix, uint32_t seed_) { pixel = pix; seed = seed_; }
    void begin_sample(uint32_t s) { sample = s; block = 0; idx = 4; }
    void begin_bounce(uint32_t b) { block = 16 * (b + 1); idx = 4; }
    float u() {
        if (idx == 4) { philox4x32_10(sample, block, 0, 0, pixel, seed, buf); block++; idx = 0; }
        return (float)(buf[idx++] >> 8) * 5.9604644775390625e-8f;
    }
    // Uniform direction on the unit sphere from two uniforms (stands in for normalize(N(0,1)^3)).
    V sphere() {
        float u1 = u(), u2 = u();
        float z = 1.f - 2.f * u1;
        float r = sqrtf(fmaxf(0.f, 1.f - z * z));
        float phi = 6.283185307179586f * u2;
        return V(r * cosf(phi), r * sinf(phi), z);
    }
    static constexpr bool kCounterBased = true;
};

// The reference's generator.  One u01 and one n01 object per pixel in hw5..hw8; n01 caches its
// second variate across samples and bounces of the pixel.
struct RngMinstd {
    std::minstd_rand rng;
    std::uniform_real_distribution<float> u01{0.0, 1.0};
    std::normal_distribution<float> n01{0.0, 1.0};
    void begin_pixel(uint32_t pix, uint32_t) {
        rng.seed(pix);
        u01 = std::uniform_real_distribution<float>(0.0, 1.0);
        n01 = std::normal_distribution<float>(0.0, 1.0);
    }
    void begin_sample(uint32_t) {}
    void begin_bounce(uint32_t) {}
    float u() { return u01(rng); }
    V sphere() { return V{n01(rng), n01(rng), n01(rng)}.normalize(); }
    static constexpr bool kCounterBased = false;
};

// Output stage shared by hw2..hw8 (hw8/src/color.cpp:4-31); defined per TU because pow/round
// bind differently under <math.h> and <cmath>.
}  // namespace port
