/*
 * b200rt_host.h -- C ABI of the host-side scene library (no CUDA inside).
 *
 * Host mirror of the reference's `sceneio` namespace: same inputs (hw1-hw5 text scenes,
 * hw6-hw8 glTF + .bin + images), same post-load hooks (SAH BVH build, light set + light BVH),
 * same output format (binary P6 PPM).  It produces the reference-shaped b200rt_scene_desc that
 * b200rt_upload_scene consumes.
 *
 *   b200rt_host_load_text  <-> sceneio::loadScene(std::istream&)   hw1/src/sceneio.cpp:54-97,
 *                               hw2/src/sceneio.cpp:102-150, hw4/src/sceneio.cpp:62-108, hw5/src/sceneio.cpp
 *   b200rt_host_load_gltf  <-> sceneio::loadScene(std::string_view) hw8/src/sceneio.cpp:348-372 (hw6, hw7 alike)
 *   b200rt_host_set_env_map<-> sceneio::loadTexture + Scene::environmentMap  hw8/src/main.cpp:12-14
 *   b200rt_host_load_gltf_ex: + the reference's SAH BVH build  hw8/src/scene.cpp:76-78, include/bvh.h:34-109
 *   b200rt_host_write_ppm  <-> header + payload of renderScene      hw8/src/sceneio.cpp:382-385,397-401
 */
#ifndef B200RT_HOST_H
#define B200RT_HOST_H

#include "b200rt.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200rt_host_scene b200rt_host_scene;

#define B200RT_HOST_ERR_IO     -101  /* file missing / unreadable          */
#define B200RT_HOST_ERR_PARSE  -102  /* malformed JSON / missing required member */
#define B200RT_HOST_ERR_IMAGE  -103  /* image could not be decoded         */

/* mode = 0: infer from content (SAMPLES => hw4, NEW_LIGHT/AMBIENT_LIGHT => hw2, else hw1;
 * a TRIANGLE primitive => hw5); otherwise B200RT_MODE_HW1..HW5. */
int b200rt_host_load_text(const char *path, int32_t mode, b200rt_host_scene **out);

/* mode = B200RT_MODE_HW6 / HW7 / HW8 (0 => HW8 when the file has "images", else HW7).
 * The description comes WITHOUT trees (nodes = light_nodes = NULL): triangles in file order, light set = the emissive
 * triangles in file order; b200rt_upload_scene builds the trees it traverses.  Files without TEXCOORD_0 / TANGENT
 * load under hw8 rules with generated attributes (uv = 0, tangent from the uv mapping or the normal's frame). */
int b200rt_host_load_gltf(const char *path, int32_t mode, int32_t width, int32_t height,
                          int32_t samples, b200rt_host_scene **out);

/* Same with flags.  B200RT_HOST_FLAG_REFERENCE_BVH also runs the reference's own SAH build (Scene::initBVH and the
 * FiguresMix constructor, hw8/src/include/bvh.h:34-109, distributions.h:103-115): triangles in ITS order + both trees.
 * Only callers that must reproduce the reference's sample streams need it (the order of the light set decides which
 * emitter a random number picks); it is O(N log^2 N) on one thread -- 1.95 s for 268 k triangles. */
#define B200RT_HOST_FLAG_REFERENCE_BVH 1
int b200rt_host_load_gltf_ex(const char *path, int32_t mode, int32_t width, int32_t height,
                             int32_t samples, int32_t flags, b200rt_host_scene **out);

int b200rt_host_set_env_map(b200rt_host_scene *scene, const char *image_path);

/* Pointer stays valid until b200rt_host_free; fields width/height/samples may be edited
 * through b200rt_host_set_frame. */
const b200rt_scene_desc *b200rt_host_desc(const b200rt_host_scene *scene);
int b200rt_host_set_frame(b200rt_host_scene *scene, int32_t width, int32_t height, int32_t samples);

void b200rt_host_free(b200rt_host_scene *scene);

int b200rt_host_write_ppm(const char *path, int32_t width, int32_t height, const uint8_t *rgb);
/* Other containers for the same frame (reference: hw8/src/sceneio.cpp:397-401 writes PPM only):
 * lossless RGB8 PNG, and the float radiance mean (sum / total_spp, before tonemapping) as a binary PFM. */
int b200rt_host_write_png(const char *path, int32_t width, int32_t height, const uint8_t *rgb);
int b200rt_host_write_pfm(const char *path, int32_t width, int32_t height, const float *rgb_sum, int32_t total_spp);

/* Load-phase timings in ms (parse, bvh build) for reporting. */
int b200rt_host_timings(const b200rt_host_scene *scene, double *parse_ms, double *bvh_ms);

const char *b200rt_host_strerror(int code);

#ifdef __cplusplus
}
#endif
#endif
