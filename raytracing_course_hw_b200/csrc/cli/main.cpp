// b200rt -- command line with the reference's argv conventions (drop-in for hwN/build/main):
//   text scenes (hw1-hw5):  b200rt <scene.txt> <out.ppm>                      hw1/src/main.cpp:7-15
//   glTF scenes (hw6-hw8):  b200rt <scene.gltf> <width> <height> <samples> <out.ppm> [<env map>]
//                                                                             hw8/src/main.cpp:7-18
// Same stderr chatter ("UNKNOWN COMMAND: ..", "FINISH"), same binary P6 output.  Unlike the
// reference (which never checks argc) a wrong invocation prints a usage line and exits 2.
// Environment: B200RT_MODE=hw1..hw8 forces the semantics (default: inferred from the scene),
// B200RT_GPUS=N renders with N GPUs (samples split; the buffers are combined by one peer-memory kernel on GPU 0),
// output container by file name (.ppm / anything: binary P6 like the reference; .png; .pfm = float radiance),
// B200RT_NO_CLAMP=1 (hw7 / hw8: unbiased estimator instead of the reference's firefly clamp), B200RT_SEED, B200RT_STATS=1 (JSON with timings and ray counts on stderr).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "b200rt.h"
#include "b200rt_host.h"

static double now_ms() {
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

static int usage(const char *argv0) {
    fprintf(stderr, "usage: %s <scene.txt> <out.ppm>\n       %s <scene.gltf> <width> <height> <samples> <out.ppm> [<environment map>]\n", argv0, argv0);
    return 2;
}

static int parse_mode(const char *s) {
    if (!s || !*s) return 0;
    if ((s[0] == 'h' || s[0] == 'H') && (s[1] == 'w' || s[1] == 'W')) s += 2;
    int m = atoi(s);
    return (m >= 1 && m <= 8) ? m : 0;
}

int main(int argc, const char *argv[]) {
    const bool gltf = argc >= 6;
    if (!(argc == 3 || argc == 6 || argc == 7)) return usage(argv[0]);
    const int mode = parse_mode(getenv("B200RT_MODE"));
    const bool want_stats = getenv("B200RT_STATS") && atoi(getenv("B200RT_STATS"));
    int n_gpus = getenv("B200RT_GPUS") ? atoi(getenv("B200RT_GPUS")) : 1;

    double t0 = now_ms();
    b200rt_host_scene *scene = nullptr;
    const char *out_path;
    int rc;
    if (gltf) {
        long w = strtol(argv[2], nullptr, 10), h = strtol(argv[3], nullptr, 10), spp = strtol(argv[4], nullptr, 10);
        out_path = argv[5];
        rc = b200rt_host_load_gltf(argv[1], mode, (int32_t)w, (int32_t)h, (int32_t)spp, &scene);
        if (!rc && argc > 6) rc = b200rt_host_set_env_map(scene, argv[6]);
    } else {
        out_path = argv[2];
        rc = b200rt_host_load_text(argv[1], mode, &scene);
    }
    if (rc) { fprintf(stderr, "b200rt: cannot load %s: %s\n", argv[1], b200rt_host_strerror(rc)); return 1; }
    const b200rt_scene_desc *d = b200rt_host_desc(scene);
    double t_load = now_ms() - t0;

    int avail = b200rt_device_count();
    if (avail <= 0) { fprintf(stderr, "b200rt: no CUDA device (this renderer has no CPU fallback)\n"); return 1; }
    if (n_gpus < 1) n_gpus = 1;
    if (n_gpus > B200RT_MAX_PEERS + 1) n_gpus = B200RT_MAX_PEERS + 1;
    // more members than devices: they share devices round robin (same code path; how a one-GPU box exercises it)
    if (d->mode <= B200RT_MODE_HW2 || d->samples < n_gpus) n_gpus = 1;   // one sample per pixel: nothing to split

    t0 = now_ms();
    // All GPUs behind one group: one host-side upload, device-to-device replication, sample split
    // [g*S/G, (g+1)*S/G), and ONE fused peer reduce + resolve kernel on the first GPU (include/b200rt.h).
    b200rt_options opt{};
    opt.seed = getenv("B200RT_SEED") ? (uint32_t)strtoul(getenv("B200RT_SEED"), nullptr, 10) : 0u;
    if (getenv("B200RT_NO_CLAMP") && atoi(getenv("B200RT_NO_CLAMP"))) opt.flags |= B200RT_FLAG_NO_FIREFLY_CLAMP;
    std::vector<int32_t> devices((size_t)n_gpus);
    for (int g = 0; g < n_gpus; g++) devices[(size_t)g] = g % avail;
    b200rt_group *grp = nullptr;
    rc = b200rt_group_create(&grp, devices.data(), n_gpus, &opt);
    if (rc) { fprintf(stderr, "b200rt: cannot create the render contexts: %s\n", b200rt_strerror(rc)); b200rt_host_free(scene); return 1; }
    auto report = [&](const char *what) { fprintf(stderr, "b200rt: %s: %s (%s)\n", what, b200rt_strerror(rc), b200rt_group_last_error(grp)); };
    rc = b200rt_group_upload_scene(grp, d);
    if (rc) report("upload");
    double t_upload = now_ms() - t0;
    t0 = now_ms();
    if (!rc && (rc = b200rt_group_render(grp, 0, 0, 0, d->samples, 0))) report("render");
    const size_t n_bytes = (size_t)d->width * d->height * 3;
    const size_t len = strlen(out_path);
    const bool want_pfm = len > 4 && !strcmp(out_path + len - 4, ".pfm");
    const bool want_png = len > 4 && !strcmp(out_path + len - 4, ".png");
    std::vector<uint8_t> rgb;
    std::vector<float> sum;
    if (!rc) {
        try { if (want_pfm) sum.resize(n_bytes); else rgb.resize(n_bytes); }
        catch (...) { fprintf(stderr, "b200rt: out of memory for the frame\n"); rc = B200RT_ERR_INVALID; }
    }
    if (!rc) {
        // the output container follows the file name: .pfm = float radiance before tonemapping, .png, else the
        // reference's binary P6 (hw8/src/sceneio.cpp:382-401)
        rc = want_pfm ? b200rt_group_download_accum(grp, sum.data()) : b200rt_group_resolve(grp, d->samples, rgb.data());
        if (rc) report("resolve");
    }
    double t_render = now_ms() - t0;
    if (!rc) {
        rc = want_pfm ? b200rt_host_write_pfm(out_path, d->width, d->height, sum.data(), d->mode <= B200RT_MODE_HW2 ? 1 : d->samples)
           : want_png ? b200rt_host_write_png(out_path, d->width, d->height, rgb.data())
                      : b200rt_host_write_ppm(out_path, d->width, d->height, rgb.data());
        if (rc) fprintf(stderr, "b200rt: cannot write %s\n", out_path);
    }
    if (!rc && want_stats) {
        b200rt_stats st{};
        if (b200rt_group_get_stats(grp, &st) == 0)
            fprintf(stderr, "{\"mode\": %d, \"width\": %d, \"height\": %d, \"samples\": %d, \"gpus\": %d, \"load_ms\": %.3f, \"upload_ms\": %.3f, "
                            "\"render_ms\": %.3f, \"kernel_ms\": %.3f, \"paths\": %llu, \"rays\": %llu, \"mpaths_per_s\": %.3f}\n",
                    d->mode, d->width, d->height, d->samples, n_gpus, t_load, t_upload, t_render, st.kernel_ms,
                    (unsigned long long)st.paths, (unsigned long long)st.rays, st.kernel_ms > 0 ? st.paths / st.kernel_ms / 1e3 : 0.0);
    }
    b200rt_group_destroy(grp);
    b200rt_host_free(scene);
    if (!rc && gltf) fprintf(stderr, "FINISH\n");   // hw6-hw8 main.cpp print this on success
    return rc ? 1 : 0;
}
