// Launchers of the analytic-scene kernels K1 (hw1 cast), K2 (hw2 Whitted), K3 (hw3 / hw4 path tracers).
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
// hw3 = uniform-hemisphere integrator, +0.5 camera, planes without T_MAX (hw3/src/scene.cpp:31-106).
cudaError_t launch_pt_analytic(const DRenderParams &rp, bool hw3, cudaStream_t stream);
// K7 (hw5): analytic figures from the constant-memory scene + triangles of `sc`; light_order = FiguresMix order
// (>= 0 analytic figure index, < 0 ~index of an emissive triangle in sc.light_sample_tris), device pointer.
cudaError_t launch_hw5(const DGltfScene &sc, const DRenderParams &rp, const int32_t *light_order, int n_lights, bool has_tris, cudaStream_t stream);
}  // namespace b200rt
