// Stress of the persistent host worker pool behind the upload path (csrc/cuda/bvh_rebuild.h: HostPool): thousands of
// short jobs of varying width, from two caller threads at once, every task must run exactly once per job.
#include <atomic>
#include <cstdio>
#include <thread>
#include <vector>
#include "bvh_rebuild.h"

using namespace b200rt;

static int hammer(unsigned seed, int rounds) {
    unsigned x = seed;
    for (int r = 0; r < rounds; r++) {
        x = x * 1664525u + 1013904223u;
        const unsigned n = 1 + (x >> 24) % 17;              // 1 .. 17 tasks (more than the 16 workers once in a while)
        std::vector<std::atomic<int>> hits(n);
        for (auto &h : hits) h = 0;
        std::atomic<long long> sum{0};
        HostPool::get().run(n, [&](unsigned t) { hits[t]++; sum += (long long)t + 1; });
        for (unsigned t = 0; t < n; t++) if (hits[t] != 1) return 1;
        if (sum != (long long)n * (n + 1) / 2) return 2;
    }
    return 0;
}

int main() {
    int rc_a = -1, rc_b = -1;
    std::thread a([&] { rc_a = hammer(1u, 4000); }), b([&] { rc_b = hammer(7u, 4000); });
    a.join(); b.join();
    if (rc_a || rc_b) { printf("pool check failed: %d %d\n", rc_a, rc_b); return 1; }
    HostPool::get().run(0, [](unsigned) {});
    printf("host pool ok\n");
    return 0;
}
