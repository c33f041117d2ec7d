// Host scene container: owns every array the b200rt_scene_desc points at.
#pragma once
#include <string>
#include <vector>
#include "b200rt_host.h"

struct b200rt_host_scene {
    b200rt_scene_desc desc{};
    std::vector<b200rt_primitive> prims;
    std::vector<b200rt_light> lights;
    std::vector<b200rt_triangle> tris;
    std::vector<b200rt_bvh_node> nodes;
    std::vector<b200rt_triangle> light_tris;
    std::vector<b200rt_bvh_node> light_nodes;
    std::vector<b200rt_primitive> light_prims;   // hw5: FiguresMix::figures_
    std::vector<b200rt_material> materials;
    std::vector<std::vector<uint8_t>> image_data;
    std::vector<b200rt_image> images;
    std::vector<uint8_t> env_data;
    b200rt_image env{};
    bool has_env = false;
    double parse_ms = 0, bvh_ms = 0;

    // Re-point the desc at the (possibly reallocated) vectors.
    void sync() {
        desc.prims = prims.data();            desc.n_prims = (int32_t)prims.size();
        desc.lights = lights.data();          desc.n_lights = (int32_t)lights.size();
        desc.tris = tris.data();              desc.n_tris = (int32_t)tris.size();
        desc.nodes = nodes.data();            desc.n_nodes = (int32_t)nodes.size();
        desc.light_tris = light_tris.data();  desc.n_light_tris = (int32_t)light_tris.size();
        desc.light_nodes = light_nodes.data();desc.n_light_nodes = (int32_t)light_nodes.size();
        desc.light_prims = light_prims.data();desc.n_light_prims = (int32_t)light_prims.size();
        desc.materials = materials.data();    desc.n_materials = (int32_t)materials.size();
        for (size_t i = 0; i < images.size(); i++) images[i].rgb = image_data[i].data();
        desc.images = images.data();          desc.n_images = (int32_t)images.size();
        if (has_env) { env.rgb = env_data.data(); desc.env_map = &env; } else desc.env_map = nullptr;
    }
};

namespace b200rt_host {
int load_text_scene(const char *path, int32_t mode, b200rt_host_scene &s);
int load_gltf_scene(const char *path, int32_t mode, int32_t flags, b200rt_host_scene &s);
// Decodes to tightly packed RGB8 (3 channels forced, like stbi_load(..., 3)).
int load_image_rgb8(const std::string &path, int32_t &w, int32_t &h, std::vector<uint8_t> &rgb);
// Builds Scene::bvh over s.tris (reorders them) and the FiguresMix light set + light BVH.
void build_triangle_accel(b200rt_host_scene &s);
// The product default: no trees; light set = the emissive triangles in file order (b200rt_upload_scene builds the trees).
void collect_lights(b200rt_host_scene &s);
// hw5: Scene::initBVH over the non-plane figures of s.prims (reorders them, planes last) and the FiguresMix
// light set + light BVH (hw5/src/scene.cpp:18-23, hw5/src/include/distributions.h:172-199).
void build_figure_accel(b200rt_host_scene &s);
double now_ms();
}  // namespace b200rt_host
