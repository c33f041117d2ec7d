// Device math + RNG shared by all kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200rt {

struct V3 {
    float x, y, z;
};
__device__ __forceinline__ V3 mk(float x, float y, float z) { return V3{x, y, z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator*(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ V3 operator/(V3 a, V3 b) { return mk(a.x / b.x, a.y / b.y, a.z / b.z); }
__device__ __forceinline__ V3 operator*(float k, V3 a) { return mk(k * a.x, k * a.y, k * a.z); }
__device__ __forceinline__ V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// Standard right-handed cross product a x b.  The reference's Vec3::cross(a, b) is b x a
// (hw8/src/include/vec3.h:57-59); call sites below swap the operands accordingly.
__device__ __forceinline__ V3 cross(V3 a, V3 b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
__device__ __forceinline__ float len2(V3 a) { return dot(a, a); }
__device__ __forceinline__ V3 normalize(V3 a) { return rsqrtf(len2(a)) * a; }
__device__ __forceinline__ V3 ld3(const float *p) { return mk(p[0], p[1], p[2]); }

// ---- Philox4x32-10; key = (pixel, seed), counter = (sample, block, 0, 0) -----------------------
// Stream layout (mirrored by oracle/port_common.h RngPhilox): block 0 of a
// camera sample = pixel jitter; bounce b draws from blocks 16*(b+1)+k.
__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t k0, uint32_t k1) {
    uint32_t c2 = 0, c3 = 0;
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint32_t h0 = __umulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
        uint32_t h1 = __umulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
        uint32_t n0 = h1 ^ c1 ^ k0, n2 = h0 ^ c3 ^ k1;
        c0 = n0; c1 = l1; c2 = n2; c3 = l0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}
__device__ __forceinline__ float u01(uint32_t x) { return (float)(x >> 8) * 5.9604644775390625e-8f; }

// Sequential view of the stream for code with data-dependent draw counts (hw4).
struct Rng {
    uint32_t pixel, seed, sample, block;
    uint4 buf;
    int idx;
    __device__ __forceinline__ void begin_sample(uint32_t s) { sample = s; block = 0; idx = 4; }
    __device__ __forceinline__ void begin_bounce(uint32_t b) { block = 16u * (b + 1u); idx = 4; }
    __device__ __forceinline__ float u() {
        if (idx == 4) { buf = philox4x32_10(sample, block, pixel, seed); block++; idx = 0; }
        uint32_t v = idx == 0 ? buf.x : idx == 1 ? buf.y : idx == 2 ? buf.z : buf.w;
        idx++;
        return u01(v);
    }
};

// Uniform direction on the unit sphere from two uniforms (replaces normalize(N(0,1)^3),
// hw8/src/include/distributions.h:44).
__device__ __forceinline__ V3 sphere_dir(float u1, float u2) {
    float z = 1.f - 2.f * u1;
    float r = sqrtf(fmaxf(0.f, 1.f - z * z));
    float s, c;
    sincosf(6.283185307179586f * u2, &s, &c);
    return mk(r * c, r * s, z);
}

}  // namespace b200rt
