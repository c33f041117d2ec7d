// b200rt -- command line with the reference's argv conventions (drop-in for hwN/build/main):
//   text scenes (hw1-hw5):  b200rt <scene.txt> <out.ppm>                      hw1/src/main.cpp:7-15
//   glTF scenes (hw6-hw8):  b200rt <scene.gltf> <width> <height> <samples> <out.ppm> [<env map>]
//                                                                             hw8/src/main.cpp:7-18
// Same stderr chatter ("UNKNOWN COMMAND: ..", "FINISH"), same binary P6 output.  Unlike the
// reference (which never checks argc) a wrong invocation prints a usage line and exits 2.
// Environment: B200RT_MODE=hw1..hw8 forces the semantics (default: inferred from the scene),
// B200RT_GPUS=N renders with N GPUs (samples split, partial sums added on the host in rank order),
// B200RT_SEED, B200RT_STATS=1 (JSON with timings and ray counts on stderr).
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
    if (n_gpus > avail) n_gpus = avail;
    if (d->mode <= B200RT_MODE_HW2 || d->samples < n_gpus) n_gpus = 1;   // one sample per pixel: nothing to split

    t0 = now_ms();
    std::vector<b200rt_ctx *> ctx((size_t)n_gpus, nullptr);
    b200rt_options opt{};
    opt.seed = getenv("B200RT_SEED") ? (uint32_t)strtoul(getenv("B200RT_SEED"), nullptr, 10) : 0u;
    for (int g = 0; g < n_gpus && !rc; g++) {
        opt.device = g;
        rc = b200rt_create(&ctx[g], &opt);
        // one host-side upload; the other GPUs get a device-to-device copy of the resident scene (glTF modes)
        if (!rc) rc = (g > 0 && d->mode >= B200RT_MODE_HW6) ? b200rt_replicate_scene(ctx[0], ctx[g]) : b200rt_upload_scene(ctx[g], d);
        if (rc) fprintf(stderr, "b200rt: GPU %d: %s (%s)\n", g, b200rt_strerror(rc), ctx[g] ? b200rt_last_error(ctx[g]) : "");
    }
    double t_upload = now_ms() - t0;
    t0 = now_ms();
    // sample-split: GPU g renders global sample indices [g*S/G, (g+1)*S/G)  (async on each context)
    for (int g = 0; g < n_gpus && !rc; g++) {
        int b = (int)((long long)d->samples * g / n_gpus), e = (int)((long long)d->samples * (g + 1) / n_gpus);
        rc = b200rt_render(ctx[g], 0, 0, b, e, 0);
        if (rc) fprintf(stderr, "b200rt: render on GPU %d: %s (%s)\n", g, b200rt_strerror(rc), b200rt_last_error(ctx[g]));
    }
    std::vector<uint8_t> rgb((size_t)d->width * d->height * 3);
    if (!rc && n_gpus > 1) {
        size_t n = rgb.size();
        std::vector<float> sum(n, 0.f), part(n);
        for (int g = 0; g < n_gpus && !rc; g++) {
            rc = b200rt_download_accum(ctx[g], part.data());
            for (size_t i = 0; i < n; i++) sum[i] += part[i];
        }
        if (!rc) rc = b200rt_upload_accum(ctx[0], sum.data());
    }
    if (!rc) rc = b200rt_resolve(ctx[0], d->samples, rgb.data());
    double t_render = now_ms() - t0;
    if (rc) fprintf(stderr, "b200rt: %s (%s)\n", b200rt_strerror(rc), ctx[0] ? b200rt_last_error(ctx[0]) : "");
    if (!rc) {
        rc = b200rt_host_write_ppm(out_path, d->width, d->height, rgb.data());
        if (rc) fprintf(stderr, "b200rt: cannot write %s\n", out_path);
    }
    if (!rc && want_stats) {
        b200rt_stats st{};
        unsigned long long paths = 0, rays = 0;
        double kernel_ms = 0;
        for (int g = 0; g < n_gpus; g++)
            if (b200rt_get_stats(ctx[g], &st) == 0) { paths += st.paths; rays += st.rays; if (st.kernel_ms > kernel_ms) kernel_ms = st.kernel_ms; }
        fprintf(stderr, "{\"mode\": %d, \"width\": %d, \"height\": %d, \"samples\": %d, \"gpus\": %d, \"load_ms\": %.3f, \"upload_ms\": %.3f, "
                        "\"render_ms\": %.3f, \"kernel_ms\": %.3f, \"paths\": %llu, \"rays\": %llu, \"mpaths_per_s\": %.3f}\n",
                d->mode, d->width, d->height, d->samples, n_gpus, t_load, t_upload, t_render, kernel_ms, paths, rays,
                kernel_ms > 0 ? paths / kernel_ms / 1e3 : 0.0);
    }
    for (auto c : ctx) b200rt_destroy(c);
    b200rt_host_free(scene);
    if (!rc && gltf) fprintf(stderr, "FINISH\n");   // hw6-hw8 main.cpp print this on success
    return rc ? 1 : 0;
}
