// Host-visible part of the wavefront pipeline's state: buffer layout and per-wave parameters (the kernels are in
// wavefront.cuh; the context and the scheduler only need these types).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "device_scene.h"

namespace b200rt {

// scheduling constants shared by the kernels and the scheduler
#define B200RT_WF_BATCH 16       /* traversal steps between refills (sweep: profiles/r01_sweep_wf.log) */
#define B200RT_WF_BATCH_LIGHT 16 /* same for the light-pdf traversal */
#define B200RT_WF_LEAF_T_LIGHT 24 /* light-pdf traversal: leaves are rare and dear, wait for a fuller warp (profiles/r01_sweep_light.log) */
#define B200RT_WF_LEAF_T 8       /* lanes waiting at a leaf that trigger a leaf phase */
#define B200RT_WF_MASK_SHIFT 26  /* light-queue entry = path slot | light-cut mask << 26 (slots < 2^26, <= 6 cut boxes) */
#define B200RT_PIPE 3            /* waves that can be in flight, each on its own stream with its own buffers */

// Path state is one 128-byte record per path slot (AoS): after the first bounce the stages visit slots in an
// order scrambled by the persistent kernels, and ncu showed every scattered 16-byte field access of the former
// six-array (SoA) layout costing a 64-byte DRAM burst (shade: 1.8 GB read per launch for 0.56 GB of state).  In
// the record, the fields a stage touches together share a 32-byte sector: {ray_o, ray_d} {hit, thr} {rad, bsdf}.
// B200RT_STATE_AOS=0 restores the separate arrays.
#ifndef B200RT_STATE_AOS
#define B200RT_STATE_AOS 1
#endif
#if B200RT_STATE_AOS
#define B200RT_STATE_STRIDE 8      /* float4 units: 128 bytes per path slot */
#else
#define B200RT_STATE_STRIDE 1
#endif
struct StateField {
    float4 *base;
    __host__ __device__ __forceinline__ float4 &operator[](size_t slot) const { return base[slot * B200RT_STATE_STRIDE]; }
};
struct WaveBuffers {
    float4 *state;      // the one allocation behind the fields below
    StateField ray_o;   // o.xyz, pixel index (int bits): the Philox key of the path
    StateField ray_d;   // d.xyz, |nd . n_s| of the bounce that produced this ray
    StateField hit;     // t, u, v, bits: triangle index | inside << 31 (B200RT_WF_MISS = no hit)
    StateField thr;     // throughput T.xyz, partial pdf (cosine + vndf)
    StateField rad;     // radiance L.xyz, global sample index (int bits): the Philox counter of the path
    StateField bsdf;    // BRDF value of the pending bounce
    int *queue[3];
    // per bounce b, c = ctr + 4b: c[-1] = rays entering bounce b, c[0] = extend cursor, c[1] = light cursor,
    //               c[2] = light queries of bounce b, c[3] = rays entering bounce b+1.  (c[2], c[3]) is an aligned
    //               64-bit word: the shade stage appends to both queues with ONE atomic (queue_push2).
    //               ctr points 4 ints into its allocation, so ctr[-1] exists.
    int *ctr;
};

struct WaveParams {
    DCamera cam;
    int32_t tiles_x, n_tiles;     // 8x4 pixel tiles
    int32_t s_begin, s_count;     // global sample indices [s_begin, s_begin + s_count) in this wave
    uint32_t seed;
};

}  // namespace b200rt
