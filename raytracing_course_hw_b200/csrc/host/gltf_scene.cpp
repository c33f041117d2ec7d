// hw6-hw8 glTF loader (host, load-time).  Accepts the same subset as the reference loader
// (hw8/src/sceneio.cpp:14-372, hw7/src/sceneio.cpp) and produces world-space
// triangles in the reference's vertex order (data, data2, data3) = (p1, p3, p2)
// (hw8/src/sceneio.cpp:289).  Differences, all on malformed input only: missing members
// fail cleanly instead of being UB; a missing buffer file is an error instead of zeros;
// hw6/hw7 files without TANGENT / images load with zero tangents and no textures.
#include <fstream>
#include <iostream>
#include <sstream>
#include "hmath.h"
#include "host_scene.h"
#include "json.h"

namespace b200rt_host {
namespace {

struct GNode {
    long mesh = -1, camera = -1, parent = -1;
    Mat4 local, total;
    std::vector<size_t> children;
};
struct GView { size_t buffer = 0, length = 0, offset = 0; };
struct GAccessor { size_t view = 0, count = 0, component = 0, offset = 0; std::string type; };

std::string dir_of(const std::string &p) {
    size_t k = p.find_last_of('/');
    return k == std::string::npos ? std::string() : p.substr(0, k + 1);
}

struct Loader {
    const Json &g;
    b200rt_host_scene &s;
    std::string base;
    int mode;
    std::vector<std::vector<char>> buffers;
    std::vector<GView> views;
    std::vector<GAccessor> accessors;
    std::vector<GNode> nodes;
    std::vector<size_t> texture_source;  // textureDescs[i].source

    int load_buffers() {
        for (const Json &b : g["buffers"].arr) {
            size_t sz = b["byteLength"].as_uint();
            std::vector<char> buf(sz);
            std::ifstream f(base + b["uri"].str, std::ios::binary);
            if (!f) return B200RT_HOST_ERR_IO;
            f.read(buf.data(), (std::streamsize)sz);
            buffers.push_back(std::move(buf));
        }
        return 0;
    }
    int load_images() {
        if (mode != B200RT_MODE_HW8) return 0;
        for (const Json &img : g["images"].arr) {
            int32_t w = 0, h = 0;
            std::vector<uint8_t> rgb;
            int rc = load_image_rgb8(base + img["uri"].str, w, h, rgb);
            if (rc) return rc;
            s.image_data.push_back(std::move(rgb));
            s.images.push_back(b200rt_image{w, h, nullptr});
        }
        for (const Json &t : g["textures"].arr) texture_source.push_back(t["source"].as_uint());
        return 0;
    }
    void load_views() {
        for (const Json &v : g["bufferViews"].arr)
            views.push_back(GView{v["buffer"].as_uint(), v["byteLength"].as_uint(), v["byteOffset"].as_uint()});
    }
    void load_accessors() {
        for (const Json &a : g["accessors"].arr) {
            GAccessor acc;
            acc.view = a["bufferView"].as_uint(); acc.count = a["count"].as_uint();
            acc.component = a["componentType"].as_uint(); acc.type = a["type"].str;
            acc.offset = a["byteOffset"].as_uint();
            accessors.push_back(acc);
        }
    }
    void load_nodes() {
        for (const Json &n : g["nodes"].arr) {
            GNode cur;
            if (n.has("mesh")) cur.mesh = (long)n["mesh"].as_uint();
            if (n.has("camera")) cur.camera = (long)n["camera"].as_uint();
            float q[4] = {0, 0, 0, 1}; V3 t(0, 0, 0), sc(1, 1, 1);
            if (n.has("rotation")) for (int i = 0; i < 4; i++) q[i] = n["rotation"][i].as_float();
            if (n.has("translation")) t = V3(n["translation"][0].as_float(), n["translation"][1].as_float(), n["translation"][2].as_float());
            if (n.has("scale")) sc = V3(n["scale"][0].as_float(), n["scale"][1].as_float(), n["scale"][2].as_float());
            if (n.has("children")) for (const Json &c : n["children"].arr) cur.children.push_back(c.as_uint());
            if (n.has("matrix")) {
                for (size_t i = 0; i < 16; i++) cur.local.m[i % 4][i / 4] = n["matrix"][i].as_float();  // column-major
            } else {
                cur.local = Mat4::translation(t).compose(Mat4::rotation(q[0], q[1], q[2], q[3])).compose(Mat4::scale(sc));
            }
            nodes.push_back(cur);
        }
        for (size_t i = 0; i < nodes.size(); i++)
            for (size_t c : nodes[i].children)
                if (c < nodes.size()) nodes[c].parent = (long)i;
        for (auto &n : nodes) {
            n.total = n.local;
            long par = n.parent; int guard = 0;
            while (par >= 0 && guard++ < 1024) {
                n.total = nodes[par].local.compose(n.total);
                par = nodes[par].parent;
            }
        }
    }
    void load_materials() {
        auto tex = [&](const Json &t) -> int32_t {
            if (mode != B200RT_MODE_HW8 || t.is_null()) return -1;
            size_t ti = t["index"].as_uint();
            if (ti >= texture_source.size() || texture_source[ti] >= s.images.size()) return -1;
            return (int32_t)texture_source[ti];
        };
        for (const Json &m : g["materials"].arr) {
            b200rt_material cur{};
            cur.base_color[0] = cur.base_color[1] = cur.base_color[2] = 1.f;
            cur.metallic_factor = 1.f; cur.roughness_factor = 1.f;
            cur.base_color_image = cur.emissive_image = cur.metallic_roughness_image = cur.normal_image = -1;
            cur.kind = B200RT_MAT_DIFFUSE; cur.ior = 1.5f;
            float alpha = 1.f;
            const Json &pbr = m["pbrMetallicRoughness"];
            if (!pbr.is_null()) {
                if (pbr.has("baseColorFactor")) {
                    for (int i = 0; i < 3; i++) cur.base_color[i] = pbr["baseColorFactor"][i].as_float();
                    alpha = pbr["baseColorFactor"][3].as_float();
                }
                cur.base_color_image = tex(pbr["baseColorTexture"]);
                if (pbr.has("metallicFactor")) cur.metallic_factor = pbr["metallicFactor"].as_float();
                if (pbr.has("roughnessFactor")) {
                    cur.roughness_factor = pbr["roughnessFactor"].as_float();
                    // hw7 clamps at load, only when the member is present (hw7/src/sceneio.cpp:168)
                    if (mode == B200RT_MODE_HW7) cur.roughness_factor = std::max(cur.roughness_factor, 0.04f);
                }
                cur.metallic_roughness_image = tex(pbr["metallicRoughnessTexture"]);
            }
            if (m.has("emissiveFactor")) for (int i = 0; i < 3; i++) cur.emission[i] = m["emissiveFactor"][i].as_float();
            cur.emissive_image = tex(m["emissiveTexture"]);
            cur.normal_image = tex(m["normalTexture"]);
            const Json &strength = m["extensions"]["KHR_materials_emissive_strength"];
            if (!strength.is_null()) {
                float k = strength["emissiveStrength"].as_float();
                for (int i = 0; i < 3; i++) cur.emission[i] = k * cur.emission[i];
            }
            if (mode == B200RT_MODE_HW6) {   // hw6/src/sceneio.cpp:179-183
                if (alpha < 1) cur.kind = B200RT_MAT_DIELECTRIC;
                else if (cur.metallic_factor > 0) cur.kind = B200RT_MAT_METALLIC;
            }
            s.materials.push_back(cur);
        }
    }

    // Tightly packed float accessor (byteStride ignored like the reference, hw8/src/sceneio.cpp:136-191).
    bool read_floats(long index, int comps, std::vector<float> &out) {
        out.clear();
        if (index < 0 || (size_t)index >= accessors.size()) return false;
        const GAccessor &a = accessors[index];
        if (a.view >= views.size() || views[a.view].buffer >= buffers.size()) return false;
        const auto &buf = buffers[views[a.view].buffer];
        size_t off = views[a.view].offset + a.offset, bytes = a.count * comps * 4;
        if (off + bytes > buf.size()) return false;
        out.resize(a.count * comps);
        memcpy(out.data(), buf.data() + off, bytes);
        return true;
    }

    int load_triangles() {
        for (const GNode &node : nodes) {
            if (node.mesh < 0) continue;
            const Json &prims = g["meshes"][(size_t)node.mesh]["primitives"];
            for (const Json &pr : prims.arr) {
                const Json &at = pr["attributes"];
                auto idx = [&](const Json &j) -> long { return j.is_number() ? (long)j.as_uint() : -1; };
                std::vector<float> pos, uv, nrm, tan;
                if (!read_floats(idx(at["POSITION"]), 3, pos)) return B200RT_HOST_ERR_PARSE;
                // hw6 reads POSITION only (hw6/src/sceneio.cpp:85-87); hw7/hw8 require NORMAL
                const bool has_nrm = read_floats(idx(at["NORMAL"]), 3, nrm) && nrm.size() == pos.size();
                if (!has_nrm && mode != B200RT_MODE_HW6) return B200RT_HOST_ERR_PARSE;
                bool has_uv = read_floats(idx(at["TEXCOORD_0"]), 2, uv);
                bool has_tan = read_floats(idx(at["TANGENT"]), 4, tan);
                if (mode == B200RT_MODE_HW8 && (!has_uv || !has_tan)) return B200RT_HOST_ERR_PARSE;
                long ia = idx(pr["indices"]);
                if (ia < 0 || (size_t)ia >= accessors.size() || !pr["material"].is_number()) return B200RT_HOST_ERR_PARSE;
                const GAccessor &acc = accessors[ia];
                if (acc.view >= views.size() || views[acc.view].buffer >= buffers.size()) return B200RT_HOST_ERR_PARSE;
                const auto &buf = buffers[views[acc.view].buffer];
                size_t base_off = views[acc.view].offset + acc.offset;
                size_t isz = acc.component == 5123 ? 2 : 4;
                if (acc.component != 5123 && acc.component != 5125)
                    std::cerr << "Unexpected accessor component type: " << acc.component << std::endl;
                if (base_off + acc.count * isz > buf.size()) return B200RT_HOST_ERR_PARSE;
                uint32_t material = (uint32_t)pr["material"].as_uint();
                if (material >= s.materials.size()) return B200RT_HOST_ERR_PARSE;
                const Mat4 &M = node.total;
                Mat4 N = M.inverted().transposed();
                V3 shift = M.apply(V3(0, 0, 0));
                size_t nv = pos.size() / 3;
                for (size_t i = 0; i + 2 < acc.count; i += 3) {
                    size_t id[3];
                    for (int k = 0; k < 3; k++) {
                        const char *p = buf.data() + base_off + isz * (i + k);
                        if (isz == 2) { uint16_t v; memcpy(&v, p, 2); id[k] = v; } else { uint32_t v; memcpy(&v, p, 4); id[k] = v; }
                        if (id[k] >= nv) return B200RT_HOST_ERR_PARSE;
                    }
                    b200rt_triangle t{};
                    const int slot[3] = {0, 2, 1};  // Figure(data=p1, data2=p3, data3=p2)
                    for (int k = 0; k < 3; k++) {
                        size_t v = id[k]; int d = slot[k];
                        V3 P = M.apply(V3(pos[3 * v], pos[3 * v + 1], pos[3 * v + 2]));
                        t.p[d][0] = P.x; t.p[d][1] = P.y; t.p[d][2] = P.z;
                        if (has_nrm) {
                            V3 Nn = N.apply(V3(nrm[3 * v], nrm[3 * v + 1], nrm[3 * v + 2])).normalized();
                            t.n[d][0] = Nn.x; t.n[d][1] = Nn.y; t.n[d][2] = Nn.z;
                        }
                        if (has_uv && v * 2 + 1 < uv.size()) { t.uv[d][0] = uv[2 * v]; t.uv[d][1] = uv[2 * v + 1]; }
                        if (has_tan && v * 4 + 3 < tan.size()) {
                            V3 T = (M.apply(V3(tan[4 * v], tan[4 * v + 1], tan[4 * v + 2])) - shift).normalized();
                            t.tan[d][0] = T.x; t.tan[d][1] = T.y; t.tan[d][2] = T.z; t.tan[d][3] = tan[4 * v + 3];
                        }
                    }
                    t.material = material;
                    s.tris.push_back(t);
                }
            }
        }
        return 0;
    }

    void load_camera() {
        b200rt_scene_desc &d = s.desc;
        V3 up(0, 1, 0), fwd(0, 0, -1), right(1, 0, 0), pos(0, 0, 0);
        for (const GNode &n : nodes) {
            if (n.camera < 0) continue;
            if (mode != B200RT_MODE_HW8) std::cerr << "Detected camera" << std::endl;  // hw6/hw7 chatter
            d.cam_fov = g["cameras"][(size_t)n.camera]["perspective"]["yfov"].as_float();
            pos = n.total.apply(V3(0, 0, 0));
            // the reference re-applies to the running vectors (last camera wins, hw8/src/sceneio.cpp:317-325)
            up = n.total.apply(up) - pos;
            right = n.total.apply(right) - pos;
            fwd = n.total.apply(fwd) - pos;
        }
        d.cam_pos[0] = pos.x; d.cam_pos[1] = pos.y; d.cam_pos[2] = pos.z;
        d.cam_up[0] = up.x; d.cam_up[1] = up.y; d.cam_up[2] = up.z;
        d.cam_right[0] = right.x; d.cam_right[1] = right.y; d.cam_right[2] = right.z;
        d.cam_forward[0] = fwd.x; d.cam_forward[1] = fwd.y; d.cam_forward[2] = fwd.z;
    }
};

}  // namespace

int load_gltf_scene(const char *path, int32_t mode, b200rt_host_scene &s) {
    double t0 = now_ms();
    std::ifstream in(path, std::ios::binary);
    if (!in) return B200RT_HOST_ERR_IO;
    std::stringstream ss; ss << in.rdbuf();
    Json g; size_t err = 0;
    if (!Json::parse(ss.str(), g, err)) return B200RT_HOST_ERR_PARSE;
    if (mode == 0) mode = g.has("images") ? B200RT_MODE_HW8 : B200RT_MODE_HW7;
    if (mode != B200RT_MODE_HW6 && mode != B200RT_MODE_HW7 && mode != B200RT_MODE_HW8) return B200RT_ERR_INVALID;
    s.desc.mode = mode;
    s.desc.ray_depth = 6;  // hw8/src/include/scene.h:32
    Loader L{g, s, dir_of(path), mode, {}, {}, {}, {}, {}};
    int rc;
    if ((rc = L.load_buffers())) return rc;
    if ((rc = L.load_images())) return rc;
    L.load_views();
    L.load_nodes();
    L.load_accessors();
    L.load_materials();
    if ((rc = L.load_triangles())) return rc;
    L.load_camera();
    s.parse_ms = now_ms() - t0;
    t0 = now_ms();
    build_triangle_accel(s);
    s.bvh_ms = now_ms() - t0;
    s.sync();
    return B200RT_OK;
}

}  // namespace b200rt_host
