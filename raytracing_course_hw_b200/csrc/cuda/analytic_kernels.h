// Launchers of the analytic-scene kernels K1 (hw1 cast), K2 (hw2 Whitted), K3 (hw4 path tracer).
// They live in their own translation unit compiled with -fmad=false so that the deterministic
// configs reproduce the reference's float operation order (bit-exact primary IDs).
#pragma once
#include <cuda_runtime.h>
#include "device_scene.h"

namespace b200rt {
cudaError_t analytic_upload(const DAnalyticScene &scene, cudaStream_t stream);
// K1/K2: one thread per pixel; write raw radiance (3 floats) to `out` and the primary-hit
// primitive index (-1 = miss) to `ids`.
cudaError_t launch_cast(const DCamera &cam, float *out, int32_t *ids, cudaStream_t stream);
cudaError_t launch_whitted(const DCamera &cam, int ray_depth, float *out, int32_t *ids, cudaStream_t stream);
// K3: same work decomposition as the glTF kernel (8x4 tiles x sample chunks -> slabs).
cudaError_t launch_pt_analytic(const DRenderParams &rp, cudaStream_t stream);
}  // namespace b200rt
