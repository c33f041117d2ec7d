// K5 arithmetic shared by the plain resolve (b200rt.cu) and the fused peer reduce + resolve (multi_gpu.cu):
// mean, ACES fit, clamp, gamma 1/2.2, round to 8 bits (hw8/src/color.cpp:4-31, hw8/src/scene.cpp:176).  hw1
// quantises the raw colour (hw1/src/color.cpp:13-19).  The gamma pow is evaluated in double: the reference calls
// double pow() in hw4-hw8 and powf in hw2; rounding the double result to float reproduces both except on exact
// rounding ties.  Channels are independent, so the callers may walk the buffer as a flat float array.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200rt {

__device__ __forceinline__ uint8_t resolve_channel(float sum, float inv_spp, int tonemap) {
    float x = inv_spp * sum;
    if (tonemap) {
        const float a = 2.51f, b = 0.03f, cc = 2.43f, d = 0.59f, e = 0.14f;
        const float num = __fmul_rn(x, __fadd_rn(__fmul_rn(a, x), b));
        const float den = __fadd_rn(__fmul_rn(x, __fadd_rn(__fmul_rn(cc, x), d)), e);
        float y = __fdiv_rn(num, den);
        y = fminf(1.f, fmaxf(0.f, y));
        x = (float)pow((double)y, (double)(float)(1. / 2.2));
    }
    return (uint8_t)roundf(__fmul_rn(255.f, x));
}

}  // namespace b200rt
