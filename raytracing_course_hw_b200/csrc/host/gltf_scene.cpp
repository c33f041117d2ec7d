// hw6-hw8 glTF loader (host, load-time).  Accepts the same subset as the reference loader
// (hw8/src/sceneio.cpp:14-372, hw7/src/sceneio.cpp) and produces world-space
// triangles in the reference's vertex order (data, data2, data3) = (p1, p3, p2)
// (hw8/src/sceneio.cpp:289).  Differences, all on malformed input only: missing members
// fail cleanly instead of being UB; a missing buffer file is an error instead of zeros;
// files without TEXCOORD_0 / TANGENT / images load under hw8 rules with generated attributes and no textures.
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <sstream>
#include "hmath.h"
#include "host_scene.h"
#include "json.h"
#include "../common/host_pool.h"
#include <algorithm>

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
    // Images are decoded concurrently on the host pool (the inflate of one PNG is serial); the caller runs this
    // before the geometry so that nothing else waits on it.
    int load_images() {
        if (mode != B200RT_MODE_HW8) return 0;
        const auto &imgs = g["images"].arr;
        const size_t n = imgs.size();
        s.image_data.assign(n, {});
        s.images.assign(n, b200rt_image{0, 0, nullptr});
        std::vector<int> rcs(n, 0);
        b200rt::parallel_for(n, [&](size_t a, size_t b) {
            for (size_t i = a; i < b; i++) {
                int32_t w = 0, h = 0;
                try { rcs[i] = load_image_rgb8(base + imgs[i]["uri"].str, w, h, s.image_data[i]); }
                catch (...) { rcs[i] = B200RT_HOST_ERR_IMAGE; }
                s.images[i] = b200rt_image{w, h, nullptr};
            }
        }, 1);
        for (int rc : rcs) if (rc) return rc;
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

    // One mesh primitive of one node, validated and ready for the parallel triangle pass.
    struct PrimJob {
        const GNode *node;
        std::vector<float> pos, uv, nrm, tan;
        bool has_nrm = false, has_uv = false, has_tan = false;
        const char *indices = nullptr;
        size_t isz = 2, n_tris = 0, first = 0;
        uint32_t material = 0;
        Mat4 M, N;
        V3 shift;
    };

    // hw8 semantics need TEXCOORD_0 and TANGENT on every primitive (hw8/src/sceneio.cpp:88-93 reads them
    // unconditionally); hw6 / hw7 files carry neither.  Instead of rejecting such files, the missing attributes are
    // generated: uv = (0, 0); tangent = the triangle's dp/du when its uv mapping is not
    // degenerate, else a unit vector orthogonal to the vertex normal (branchless frame of Duff et al. 2017), w = +1.
    static V3 any_tangent(const V3 &n) {
        const float sg = std::copysign(1.f, n.z);
        const float a = -1.f / (sg + n.z), b = n.x * n.y * a;
        return V3(1.f + sg * n.x * n.x * a, sg * b, -sg * n.x);
    }

    void emit_triangle(const PrimJob &j, size_t i, b200rt_triangle &t) const {
        size_t id[3];
        for (int k = 0; k < 3; k++) {
            const char *p = j.indices + j.isz * (3 * i + k);
            if (j.isz == 2) { uint16_t v; memcpy(&v, p, 2); id[k] = v; } else { uint32_t v; memcpy(&v, p, 4); id[k] = v; }
        }
        t = b200rt_triangle{};
        const int slot[3] = {0, 2, 1};  // Figure(data=p1, data2=p3, data3=p2)
        const Mat4 &M = j.M, &N = j.N;
        for (int k = 0; k < 3; k++) {
            const size_t v = id[k]; const int d = slot[k];
            const V3 P = M.apply(V3(j.pos[3 * v], j.pos[3 * v + 1], j.pos[3 * v + 2]));
            t.p[d][0] = P.x; t.p[d][1] = P.y; t.p[d][2] = P.z;
            if (j.has_nrm) {
                const V3 Nn = N.apply(V3(j.nrm[3 * v], j.nrm[3 * v + 1], j.nrm[3 * v + 2])).normalized();
                t.n[d][0] = Nn.x; t.n[d][1] = Nn.y; t.n[d][2] = Nn.z;
            }
            if (j.has_uv && v * 2 + 1 < j.uv.size()) { t.uv[d][0] = j.uv[2 * v]; t.uv[d][1] = j.uv[2 * v + 1]; }
            if (j.has_tan && v * 4 + 3 < j.tan.size()) {
                const V3 T = (M.apply(V3(j.tan[4 * v], j.tan[4 * v + 1], j.tan[4 * v + 2])) - j.shift).normalized();
                t.tan[d][0] = T.x; t.tan[d][1] = T.y; t.tan[d][2] = T.z; t.tan[d][3] = j.tan[4 * v + 3];
            }
        }
        if (mode == B200RT_MODE_HW8 && !j.has_tan) {
            // generated tangents (see above): dp/du of the triangle, or a frame vector of the vertex normal
            const V3 p0(t.p[2][0], t.p[2][1], t.p[2][2]);
            const V3 e1 = V3(t.p[0][0], t.p[0][1], t.p[0][2]) - p0, e2 = V3(t.p[1][0], t.p[1][1], t.p[1][2]) - p0;
            const float du1 = t.uv[0][0] - t.uv[2][0], dv1 = t.uv[0][1] - t.uv[2][1];
            const float du2 = t.uv[1][0] - t.uv[2][0], dv2 = t.uv[1][1] - t.uv[2][1];
            const float det = du1 * dv2 - du2 * dv1;
            V3 face(0, 0, 0);
            bool have_face = false;
            if (j.has_uv && std::fabs(det) > 1e-12f) {
                face = V3(dv2 * e1.x - dv1 * e2.x, dv2 * e1.y - dv1 * e2.y, dv2 * e1.z - dv1 * e2.z);
                if (det < 0) face = V3(-face.x, -face.y, -face.z);
                have_face = face.len() > 0.f;
            }
            for (int d = 0; d < 3; d++) {
                const V3 n(t.n[d][0], t.n[d][1], t.n[d][2]);
                V3 T = any_tangent(n);
                if (have_face) {
                    const float k = face.dot(n);
                    const V3 g(face.x - k * n.x, face.y - k * n.y, face.z - k * n.z);      // Gram-Schmidt against the normal
                    if (g.len() > 1e-12f) T = g;
                }
                T = T.normalized();
                t.tan[d][0] = T.x; t.tan[d][1] = T.y; t.tan[d][2] = T.z; t.tan[d][3] = 1.f;
            }
        }
        t.material = j.material;
    }

    int load_triangles() {
        std::vector<PrimJob> jobs;
        size_t total = 0;
        for (const GNode &node : nodes) {
            if (node.mesh < 0) continue;
            const Json &prims = g["meshes"][(size_t)node.mesh]["primitives"];
            for (const Json &pr : prims.arr) {
                const Json &at = pr["attributes"];
                auto idx = [&](const Json &j) -> long { return j.is_number() ? (long)j.as_uint() : -1; };
                jobs.emplace_back();
                PrimJob &j = jobs.back();
                j.node = &node;
                if (!read_floats(idx(at["POSITION"]), 3, j.pos)) return B200RT_HOST_ERR_PARSE;
                // hw6 reads POSITION only (hw6/src/sceneio.cpp:85-87); hw7/hw8 require NORMAL
                j.has_nrm = read_floats(idx(at["NORMAL"]), 3, j.nrm) && j.nrm.size() == j.pos.size();
                if (!j.has_nrm && mode != B200RT_MODE_HW6) return B200RT_HOST_ERR_PARSE;
                j.has_uv = read_floats(idx(at["TEXCOORD_0"]), 2, j.uv);
                j.has_tan = read_floats(idx(at["TANGENT"]), 4, j.tan);
                const long ia = idx(pr["indices"]);
                if (ia < 0 || (size_t)ia >= accessors.size() || !pr["material"].is_number()) return B200RT_HOST_ERR_PARSE;
                const GAccessor &acc = accessors[ia];
                if (acc.view >= views.size() || views[acc.view].buffer >= buffers.size()) return B200RT_HOST_ERR_PARSE;
                const auto &buf = buffers[views[acc.view].buffer];
                const size_t base_off = views[acc.view].offset + acc.offset;
                j.isz = acc.component == 5123 ? 2 : 4;
                if (acc.component != 5123 && acc.component != 5125)
                    std::cerr << "Unexpected accessor component type: " << acc.component << std::endl;
                if (base_off + acc.count * j.isz > buf.size()) return B200RT_HOST_ERR_PARSE;
                j.indices = buf.data() + base_off;
                j.material = (uint32_t)pr["material"].as_uint();
                if (j.material >= s.materials.size()) return B200RT_HOST_ERR_PARSE;
                j.M = node.total;
                j.N = j.M.inverted().transposed();
                j.shift = j.M.apply(V3(0, 0, 0));
                j.n_tris = acc.count / 3;
                j.first = total;
                total += j.n_tris;
                // every index must address a vertex (checked once here, not per triangle in the parallel pass)
                const size_t nv = j.pos.size() / 3;
                bool ok = true;
                const size_t n_idx = j.n_tris * 3;
                if (j.isz == 2) { for (size_t k = 0; k < n_idx && ok; k++) { uint16_t v; memcpy(&v, j.indices + 2 * k, 2); ok = v < nv; } }
                else { for (size_t k = 0; k < n_idx && ok; k++) { uint32_t v; memcpy(&v, j.indices + 4 * k, 4); ok = v < nv; } }
                if (!ok) return B200RT_HOST_ERR_PARSE;
            }
        }
        s.tris.resize(total);
        // triangles of all primitives, on the host cores (file order is kept: job after job, triangle after triangle)
        std::vector<size_t> starts;
        for (const PrimJob &j : jobs) starts.push_back(j.first);
        b200rt::parallel_for(total, [&](size_t a, size_t b) {
            size_t ji = (size_t)(std::upper_bound(starts.begin(), starts.end(), a) - starts.begin()) - 1;
            for (size_t i = a; i < b; i++) {
                while (ji + 1 < jobs.size() && i >= jobs[ji + 1].first) ji++;
                emit_triangle(jobs[ji], i - jobs[ji].first, s.tris[i]);
            }
        }, 4096);
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

int load_gltf_scene(const char *path, int32_t mode, int32_t flags, b200rt_host_scene &s) {
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
#ifdef B200RT_DEV      // development builds only: per-phase times on stderr
    const bool timing = getenv("B200RT_LOAD_TIMING") != nullptr;
    double tl = now_ms();
    auto lap = [&](const char *what) { if (timing) { const double n = now_ms(); fprintf(stderr, "[load] %-12s %8.3f ms\n", what, n - tl); tl = n; } };
#else
    auto lap = [](const char *) {};
#endif
    lap("json");
    if ((rc = L.load_buffers())) return rc;
    lap("buffers");
    if ((rc = L.load_images())) return rc;
    lap("images");
    L.load_views();
    L.load_nodes();
    L.load_accessors();
    L.load_materials();
    lap("tables");
    if ((rc = L.load_triangles())) return rc;
    lap("triangles");
    L.load_camera();
    s.parse_ms = now_ms() - t0;
    t0 = now_ms();
    // Default: no trees.  The description carries the triangles in file order and the emissive ones (a filtered
    // copy, file order) as the light set; b200rt_upload_scene builds the trees it traverses.  The reference's own
    // O(N log^2 N) SAH build (Scene::initBVH + FiguresMix, hw8/src/include/bvh.h:34-109) took 1.95 s of a 2.24 s load
    // on the 268 k-triangle benchmark scene and its tree was then discarded by the device library; it now runs only
    // on request, for callers that need the reference's triangle / light ORDER (oracle tests: the order decides
    // which emitter a random number picks).
    if (flags & B200RT_HOST_FLAG_REFERENCE_BVH) build_triangle_accel(s);
    else collect_lights(s);
    s.bvh_ms = now_ms() - t0;
    s.sync();
    return B200RT_OK;
}

}  // namespace b200rt_host
