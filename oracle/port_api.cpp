// TEST INFRASTRUCTURE ONLY -- C entry points of the oracle port (ctypes / bench cpu_baseline).
#include "../include/b200rt.h"
#include <cstdint>

namespace port {
void render_deterministic(const b200rt_scene_desc &s, uint8_t *rgb, int32_t *ids, float *accum);
void render_mc(const b200rt_scene_desc &s, int rng_kind, uint32_t seed, int b, int e, float *accum, uint8_t *rgb, uint64_t *counters);
void resolve_mc(const float *sum, int n_pixels, int total_spp, uint8_t *rgb);
}

extern "C" {

// rng_kind: 0 = the reference's minstd_rand stream (pins the port against oracle/_ref),
//           1 = the Philox stream of the CUDA kernels (predicts the GPU image).
// Any output pointer may be NULL.  counters (6 x u64): rays, node tests, tri tests, light
// queries, light node tests, light tri tests (glTF modes).
int oracle_render(const b200rt_scene_desc *s, int rng_kind, uint32_t seed, int spp_begin, int spp_end,
                  float *accum_sum, uint8_t *rgb, int32_t *primary_ids, uint64_t *counters) {
    if (!s || s->width <= 0 || s->height <= 0) return -1;
    switch (s->mode) {
        case B200RT_MODE_HW1:
        case B200RT_MODE_HW2:
            port::render_deterministic(*s, rgb, primary_ids, accum_sum);
            return 0;
        case B200RT_MODE_HW3:
        case B200RT_MODE_HW4:
        case B200RT_MODE_HW5:
        case B200RT_MODE_HW6:
        case B200RT_MODE_HW7:
        case B200RT_MODE_HW8:
            if (spp_end <= spp_begin) return -1;
            port::render_mc(*s, rng_kind, seed, spp_begin, spp_end, accum_sum, rgb, counters);
            return 0;
        default:
            return -5;
    }
}

int oracle_resolve(const float *sum, int n_pixels, int total_spp, uint8_t *rgb) {
    if (!sum || !rgb || total_spp <= 0) return -1;
    port::resolve_mc(sum, n_pixels, total_spp, rgb);
    return 0;
}

}  // extern "C"
