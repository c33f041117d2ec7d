// C ABI of the host scene library + acceleration-structure hooks + image/PPM I/O.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <new>
#include "bvh_build.h"
#include "hmath.h"
#include "host_scene.h"

#if defined(B200RT_HAVE_STB)
#define STB_IMAGE_IMPLEMENTATION
#define STBI_NO_HDR
#define STBI_NO_LINEAR
#pragma GCC diagnostic push
#pragma GCC diagnostic ignored "-Wunused-function"
#pragma GCC diagnostic ignored "-Wunused-but-set-variable"
#include "stb_image.h"   // third-party decoder the reference also uses (hw8/src/sceneio.cpp:374-379)
#pragma GCC diagnostic pop
#endif

namespace b200rt_host {

double now_ms() {
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

bool decode_png_rgb8(const std::vector<uint8_t> &file, int32_t &w, int32_t &h, std::vector<uint8_t> &rgb);
void encode_png_rgb8(int32_t w, int32_t h, const uint8_t *rgb, std::vector<uint8_t> &out);

int load_image_rgb8(const std::string &path, int32_t &w, int32_t &h, std::vector<uint8_t> &rgb) {
    std::ifstream f(path, std::ios::binary);
    if (!f) return B200RT_HOST_ERR_IO;
    std::vector<uint8_t> file((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    // Built-in PNG decoder first (lossless => identical texels to stb_image); other formats via stb.
    if (decode_png_rgb8(file, w, h, rgb)) return B200RT_OK;
#if defined(B200RT_HAVE_STB)
    int ww = 0, hh = 0, ch = 0;
    unsigned char *px = stbi_load_from_memory(file.data(), (int)file.size(), &ww, &hh, &ch, 3);
    if (!px) return B200RT_HOST_ERR_IMAGE;
    w = ww; h = hh;
    rgb.assign(px, px + (size_t)ww * hh * 3);
    stbi_image_free(px);
    return B200RT_OK;
#else
    return B200RT_HOST_ERR_IMAGE;
#endif
}

static BvhItem item_of(const b200rt_triangle &t, uint32_t index, bool hw6_key) {
    BvhItem it;
    for (int a = 0; a < 3; a++) {
        // hw7/hw8 sort by data3.coords (hw8/src/include/bvh.h:63-65); hw6 sorts by Figure::position, which is
        // (0,0,0) for every glTF triangle (hw6/src/include/bvh.h:63-65): the sorts then only permute equal keys
        it.key[a] = hw6_key ? 0.f : t.p[2][a];
        it.mn[a] = std::min(t.p[2][a], std::min(t.p[0][a], t.p[1][a]));
        it.mx[a] = std::max(t.p[2][a], std::max(t.p[0][a], t.p[1][a]));
    }
    it.index = index;
    return it;
}

static void build_over(std::vector<b200rt_triangle> &tris, uint32_t n, std::vector<b200rt_bvh_node> &nodes, int32_t *root, bool hw6_key) {
    std::vector<BvhItem> items(n);
    for (uint32_t i = 0; i < n; i++) items[i] = item_of(tris[i], i, hw6_key);
    BvhBuilder b;
    b.build(items, n);
    std::vector<b200rt_triangle> sorted(n);
    for (uint32_t i = 0; i < n; i++) sorted[i] = tris[items[i].index];
    std::copy(sorted.begin(), sorted.end(), tris.begin());
    nodes = std::move(b.nodes);
    if (root) *root = (int32_t)b.root;
}

// Scene::initBVH (hw8/src/scene.cpp:76-78) then Scene::initDistribution -> FiguresMix ctor
// (hw8/src/include/distributions.h:103-115): copy the BVH-ordered figures, std::partition the
// emissive ones (material emissiveFactor != 0; textures ignored) to the front, build a second
// BVH over them.
void build_triangle_accel(b200rt_host_scene &s) {
    const bool hw6_key = s.desc.mode == B200RT_MODE_HW6;
    build_over(s.tris, (uint32_t)s.tris.size(), s.nodes, &s.desc.root, hw6_key);
    std::vector<b200rt_triangle> copy = s.tris;
    auto emissive = [&](const b200rt_triangle &t) {
        const float *e = s.materials[t.material].emission;
        return !(e[0] == 0 && e[1] == 0 && e[2] == 0);
    };
    size_t n = (size_t)(std::partition(copy.begin(), copy.end(), emissive) - copy.begin());
    copy.resize(n);
    s.light_tris = std::move(copy);
    s.light_nodes.clear();
    if (n > 0) build_over(s.light_tris, (uint32_t)n, s.light_nodes, nullptr, hw6_key);
    // n == 0: the reference still builds a one-node BVH over nothing but never uses it.
}

void collect_lights(b200rt_host_scene &s) {
    s.nodes.clear(); s.light_nodes.clear();
    s.desc.root = 0;
    s.light_tris.clear();
    for (const b200rt_triangle &t : s.tris) {
        const float *e = s.materials[t.material].emission;      // emissiveFactor only; textures ignored
        if (!(e[0] == 0 && e[1] == 0 && e[2] == 0)) s.light_tris.push_back(t);
    }
}

// ---- hw5: figures of mixed type in the BVH -------------------------------------------------------------
// Quaternion product / transform with the reference's operation order (hw5/src/include/quaternion.h:36-48;
// its cross is the negated one, vec3.h:57-59) so that rotated boxes get bit-identical AABBs: the SAH build
// compares scores with == (bvh.h:96-104), a last-ulp difference could pick another axis and reorder the lights.
namespace {
struct Quat { V3 v; float w; };
inline V3 scale3(float k, const V3 &p) { return {k * p.x, k * p.y, k * p.z}; }
inline Quat qmul(const Quat &a, const Quat &b) {
    return {scale3(a.w, b.v) + scale3(b.w, a.v) + a.v.cross_ref(b.v), a.w * b.w - a.v.dot(b.v)};
}
inline Quat qconj(const Quat &q) { return {scale3(-1.f, q.v), q.w}; }
inline V3 qtransform(const Quat &q, const V3 &p) { return qmul(qmul(q, Quat{p, 0.f}), qconj(q)).v; }

// AABB::AABB(const Figure&) (hw5/src/primitives.cpp:166-196)
BvhItem item_of_figure(const b200rt_primitive &f, uint32_t index) {
    V3 mn, mx;
    if (f.type == B200RT_PRIM_TRIANGLE) {
        mn = V3(std::min(f.data3[0], std::min(f.data[0], f.data2[0])), std::min(f.data3[1], std::min(f.data[1], f.data2[1])),
                std::min(f.data3[2], std::min(f.data[2], f.data2[2])));
        mx = V3(std::max(f.data3[0], std::max(f.data[0], f.data2[0])), std::max(f.data3[1], std::max(f.data[1], f.data2[1])),
                std::max(f.data3[2], std::max(f.data[2], f.data2[2])));
    } else {
        mx = V3(f.data[0], f.data[1], f.data[2]);
        mn = scale3(-1.f, mx);
    }
    const Quat rot = qconj(Quat{V3(f.rotation[0], f.rotation[1], f.rotation[2]), f.rotation[3]});
    const V3 u_mn = mn, u_mx = mx;
    mn = mx = qtransform(rot, u_mn);
    auto extend = [&](const V3 &p) {
        mx.x = std::max(mx.x, p.x); mx.y = std::max(mx.y, p.y); mx.z = std::max(mx.z, p.z);
        mn.x = std::min(mn.x, p.x); mn.y = std::min(mn.y, p.y); mn.z = std::min(mn.z, p.z);
    };
    extend(qtransform(rot, V3(u_mn.x, u_mn.y, u_mx.z)));
    extend(qtransform(rot, V3(u_mn.x, u_mx.y, u_mn.z)));
    extend(qtransform(rot, V3(u_mn.x, u_mx.y, u_mx.z)));
    extend(qtransform(rot, V3(u_mx.x, u_mn.y, u_mn.z)));
    extend(qtransform(rot, V3(u_mx.x, u_mn.y, u_mx.z)));
    extend(qtransform(rot, V3(u_mx.x, u_mx.y, u_mn.z)));
    extend(qtransform(rot, V3(u_mx.x, u_mx.y, u_mx.z)));
    const V3 pos(f.position[0], f.position[1], f.position[2]);
    mn = mn + pos; mx = mx + pos;
    BvhItem it;
    it.key[0] = pos.x; it.key[1] = pos.y; it.key[2] = pos.z;      // halfSplit sorts by Figure::position (bvh.h:60-65)
    it.mn[0] = mn.x; it.mn[1] = mn.y; it.mn[2] = mn.z;
    it.mx[0] = mx.x; it.mx[1] = mx.y; it.mx[2] = mx.z;
    it.index = index;
    return it;
}

void build_over_figures(std::vector<b200rt_primitive> &figs, uint32_t n, std::vector<b200rt_bvh_node> &nodes, int32_t *root) {
    std::vector<BvhItem> items(n);
    for (uint32_t i = 0; i < n; i++) items[i] = item_of_figure(figs[i], i);
    BvhBuilder b;
    b.build(items, n);
    std::vector<b200rt_primitive> sorted(n);
    for (uint32_t i = 0; i < n; i++) sorted[i] = figs[items[i].index];
    std::copy(sorted.begin(), sorted.end(), figs.begin());
    nodes = std::move(b.nodes);
    if (root) *root = (int32_t)b.root;
}
}  // namespace

void build_figure_accel(b200rt_host_scene &s) {
    // Scene::initBVH: std::partition puts the non-planes first (same algorithm => same permutation)
    const size_t non_planes = (size_t)(std::partition(s.prims.begin(), s.prims.end(),
                                                      [](const b200rt_primitive &f) { return f.type != B200RT_PRIM_PLANE; }) - s.prims.begin());
    build_over_figures(s.prims, (uint32_t)non_planes, s.nodes, &s.desc.root);
    s.desc.n_bvh_prims = (int32_t)non_planes;
    // FiguresMix ctor: copy of ALL figures, emissive BOX / ELLIPSOID / TRIANGLE partitioned to the front, second BVH
    std::vector<b200rt_primitive> copy = s.prims;
    const size_t n = (size_t)(std::partition(copy.begin(), copy.end(), [](const b200rt_primitive &f) {
        if (f.emission[0] == 0 && f.emission[1] == 0 && f.emission[2] == 0) return false;
        return f.type == B200RT_PRIM_BOX || f.type == B200RT_PRIM_ELLIPSOID || f.type == B200RT_PRIM_TRIANGLE;
    }) - copy.begin());
    s.light_nodes.clear();
    // the reference builds this BVH over the partitioned full copy, i.e. the sorts run on the first n entries
    if (n > 0) build_over_figures(copy, (uint32_t)n, s.light_nodes, nullptr);
    copy.resize(n);
    s.light_prims = std::move(copy);
}

}  // namespace b200rt_host

using namespace b200rt_host;

extern "C" {

int b200rt_host_load_text(const char *path, int32_t mode, b200rt_host_scene **out) {
    if (!path || !out) return B200RT_ERR_INVALID;
    auto *s = new (std::nothrow) b200rt_host_scene();
    if (!s) return B200RT_ERR_INVALID;
    int rc = B200RT_ERR_INVALID;
    try { rc = load_text_scene(path, mode, *s); } catch (...) { rc = B200RT_HOST_ERR_PARSE; }
    if (rc) { delete s; return rc; }
    *out = s;
    return B200RT_OK;
}

int b200rt_host_load_gltf_ex(const char *path, int32_t mode, int32_t width, int32_t height, int32_t samples, int32_t flags,
                             b200rt_host_scene **out) {
    if (!path || !out || width <= 0 || height <= 0 || samples <= 0) return B200RT_ERR_INVALID;
    *out = nullptr;
    auto *s = new (std::nothrow) b200rt_host_scene();
    if (!s) return B200RT_ERR_INVALID;
    int rc = B200RT_ERR_INVALID;
    try { rc = load_gltf_scene(path, mode, flags, *s); }
    catch (const std::bad_alloc &) { rc = B200RT_HOST_ERR_IO; }
    catch (...) { rc = B200RT_HOST_ERR_PARSE; }
    if (rc) { delete s; return rc; }
    s->desc.width = width; s->desc.height = height; s->desc.samples = samples;
    *out = s;
    return B200RT_OK;
}

int b200rt_host_load_gltf(const char *path, int32_t mode, int32_t width, int32_t height, int32_t samples, b200rt_host_scene **out) {
    return b200rt_host_load_gltf_ex(path, mode, width, height, samples, 0, out);
}

int b200rt_host_set_env_map(b200rt_host_scene *s, const char *image_path) {
    if (!s || !image_path) return B200RT_ERR_INVALID;
    int32_t w, h;
    std::vector<uint8_t> rgb;
    int rc;
    try { rc = load_image_rgb8(image_path, w, h, rgb); } catch (...) { rc = B200RT_HOST_ERR_IMAGE; }
    if (rc) return rc;
    s->env_data = std::move(rgb);
    s->env.width = w; s->env.height = h;
    s->has_env = true;
    s->sync();
    return B200RT_OK;
}

const b200rt_scene_desc *b200rt_host_desc(const b200rt_host_scene *s) { return s ? &s->desc : nullptr; }

int b200rt_host_set_frame(b200rt_host_scene *s, int32_t width, int32_t height, int32_t samples) {
    if (!s) return B200RT_ERR_INVALID;
    if (width > 0) s->desc.width = width;
    if (height > 0) s->desc.height = height;
    if (samples > 0) s->desc.samples = samples;
    return B200RT_OK;
}

void b200rt_host_free(b200rt_host_scene *s) { delete s; }

int b200rt_host_write_ppm(const char *path, int32_t width, int32_t height, const uint8_t *rgb) {
    if (!path || !rgb || width <= 0 || height <= 0) return B200RT_ERR_INVALID;
    FILE *f = fopen(path, "wb");
    if (!f) return B200RT_HOST_ERR_IO;
    fprintf(f, "P6\n%d %d\n255\n", width, height);  // hw8/src/sceneio.cpp:383-385
    size_t n = (size_t)width * height * 3;
    size_t w = fwrite(rgb, 1, n, f);
    fclose(f);
    return w == n ? B200RT_OK : B200RT_HOST_ERR_IO;
}

// Lossless RGB8 PNG of the same frame (the reference's tree vendors stb_image_write for this and never calls it).
int b200rt_host_write_png(const char *path, int32_t width, int32_t height, const uint8_t *rgb) {
    if (!path || !rgb || width <= 0 || height <= 0) return B200RT_ERR_INVALID;
    try {
        std::vector<uint8_t> png;
        encode_png_rgb8(width, height, rgb, png);
        FILE *f = fopen(path, "wb");
        if (!f) return B200RT_HOST_ERR_IO;
        const size_t w = fwrite(png.data(), 1, png.size(), f);
        fclose(f);
        return w == png.size() ? B200RT_OK : B200RT_HOST_ERR_IO;
    } catch (...) { return B200RT_HOST_ERR_IO; }
}

// Float radiance (mean over `total_spp` samples of the accumulated SUM buffer, before tonemapping) as a binary PFM:
// "PF\n<W> <H>\n-1.0\n" + little-endian float RGB, BOTTOM row first (the format's convention).
int b200rt_host_write_pfm(const char *path, int32_t width, int32_t height, const float *rgb_sum, int32_t total_spp) {
    if (!path || !rgb_sum || width <= 0 || height <= 0 || total_spp <= 0) return B200RT_ERR_INVALID;
    FILE *f = fopen(path, "wb");
    if (!f) return B200RT_HOST_ERR_IO;
    fprintf(f, "PF\n%d %d\n-1.0\n", width, height);
    const float inv = (float)(1.0 / total_spp);
    std::vector<float> row((size_t)width * 3);
    bool ok = true;
    for (int32_t y = height - 1; y >= 0 && ok; y--) {
        const float *src = rgb_sum + (size_t)y * width * 3;
        for (size_t i = 0; i < row.size(); i++) row[i] = inv * src[i];
        ok = fwrite(row.data(), sizeof(float), row.size(), f) == row.size();
    }
    fclose(f);
    return ok ? B200RT_OK : B200RT_HOST_ERR_IO;
}

int b200rt_host_timings(const b200rt_host_scene *s, double *parse_ms, double *bvh_ms) {
    if (!s) return B200RT_ERR_INVALID;
    if (parse_ms) *parse_ms = s->parse_ms;
    if (bvh_ms) *bvh_ms = s->bvh_ms;
    return B200RT_OK;
}

const char *b200rt_host_strerror(int code) {
    switch (code) {
        case B200RT_OK: return "ok";
        case B200RT_ERR_INVALID: return "invalid argument";
        case B200RT_HOST_ERR_IO: return "file missing or unreadable";
        case B200RT_HOST_ERR_PARSE: return "malformed scene (JSON syntax or missing required member)";
        case B200RT_HOST_ERR_IMAGE: return "image could not be decoded";
        default: return "unknown error";
    }
}

}  // extern "C"
