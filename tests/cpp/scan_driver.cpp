// Concurrent index scans, the way DuckDB connections issue them: T host threads, each calling gvss_search with ONE query at a time
// (HNSWIndex::InitializeScan under the shared side of rwlock, hnsw_index.cpp:285-311), against the compiled reference
// (oracle/_ref/libref_usearch.so, loaded with dlopen: test infrastructure) searching the SAME graph with T threads.
//   scan_driver <n> <dims> <M> <ef> <threads> <path to libref_usearch.so | ->
// Prints single-query latency, T-thread QPS of both, and whether the ids agree.
#include "gvss.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <random>
#include <thread>
#include <vector>

using clk = std::chrono::steady_clock;
static double secs(clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); }
#define CHECK(x) do { if (x) { std::printf("FAILED %s: %s\n", #x, gvss_last_error()); return 2; } } while (0)

static int write_cb(void *u, const void *p, size_t n) {
    auto *v = static_cast<std::vector<unsigned char> *>(u);
    v->insert(v->end(), (const unsigned char *)p, (const unsigned char *)p + n);
    return 1;
}

int main(int argc, char **argv) {
    const uint64_t n = argc > 1 ? strtoull(argv[1], nullptr, 10) : 200000;
    const uint32_t d = argc > 2 ? (uint32_t)atoi(argv[2]) : 128, M = argc > 3 ? (uint32_t)atoi(argv[3]) : 16;
    const uint32_t ef = argc > 4 ? (uint32_t)atoi(argv[4]) : 64, T = argc > 5 ? (uint32_t)atoi(argv[5]) : 16;
    const char *ref_path = argc > 6 ? argv[6] : "-";
    const uint32_t k = 10, nq = 2048;
    std::mt19937_64 rng(3);
    std::normal_distribution<float> N01(0.f, 1.f);
    const uint32_t C = (uint32_t)std::max<uint64_t>(16, n / 100), r = 16;
    std::vector<float> centres((size_t)C * d), proj((size_t)r * d), rows(n * d), q((size_t)nq * d);
    for (auto &v : centres) v = N01(rng);
    for (auto &v : proj) v = 0.35f * N01(rng);
    auto draw = [&](float *out) {
        const uint32_t c = (uint32_t)(rng() % C);
        float z[16];
        for (auto &v : z) v = N01(rng);
        for (uint32_t j = 0; j < d; j++) {
            float x = centres[(size_t)c * d + j] + 0.05f * N01(rng);
            for (uint32_t t = 0; t < r; t++) x += z[t] * proj[(size_t)t * d + j];
            out[j] = x;
        }
    };
    for (uint64_t i = 0; i < n; i++) draw(rows.data() + i * d);
    for (uint32_t i = 0; i < nq; i++) draw(q.data() + (size_t)i * d);

    gvss_config cfg{};
    cfg.dimensions = d; cfg.metric = 'e'; cfg.connectivity = M; cfg.connectivity_base = 2 * M; cfg.expansion_add = M == 16 ? 128 : 200; cfg.device = -1;
    gvss_index *ix = nullptr;
    CHECK(gvss_create(&cfg, &ix));
    CHECK(gvss_reserve(ix, n, 1));
    CHECK(gvss_add_batch(ix, nullptr, rows.data(), n));
    CHECK(gvss_flush(ix));

    // one thread, one query per call
    std::vector<int64_t> keys((size_t)nq * k);
    for (uint32_t i = 0; i < 32; i++) CHECK(gvss_search(ix, q.data() + (size_t)i * d, 1, k, ef, keys.data() + (size_t)i * k, nullptr, nullptr, nullptr));
    std::vector<double> lat;
    for (uint32_t i = 0; i < 512; i++) {
        const auto t0 = clk::now();
        CHECK(gvss_search(ix, q.data() + (size_t)i * d, 1, k, ef, keys.data() + (size_t)i * k, nullptr, nullptr, nullptr));
        lat.push_back(secs(t0, clk::now()));
    }
    std::sort(lat.begin(), lat.end());
    // T threads, one query per call each
    std::atomic<int> bad{0};
    auto t0 = clk::now();
    std::vector<std::thread> ts;
    for (uint32_t t = 0; t < T; t++)
        ts.emplace_back([&, t] {
            for (uint32_t i = t; i < nq; i += T)
                if (gvss_search(ix, q.data() + (size_t)i * d, 1, k, ef, keys.data() + (size_t)i * k, nullptr, nullptr, nullptr)) bad++;
        });
    for (auto &th : ts) th.join();
    const double t_gvss = secs(t0, clk::now());
    gvss_info info{};
    CHECK(gvss_get_info(ix, &info));
    std::printf("n %llu dims %u M %u ef %u threads %u\n", (unsigned long long)n, d, M, ef, T);
    std::printf("gvss single_query_median_ms %.3f p95_ms %.3f threads_qps %.0f failures %d concurrent_peak %u\n", 1e3 * lat[lat.size() / 2],
                1e3 * lat[lat.size() * 95 / 100], nq / t_gvss, bad.load(), info.concurrent_searches_peak);

    if (std::strcmp(ref_path, "-") != 0) {
        void *h = dlopen(ref_path, RTLD_NOW);
        if (!h) { std::printf("reference unavailable: %s\n", dlerror()); return 0; }
        auto ref_create = (void *(*)(uint32_t, char, uint32_t, uint32_t, uint32_t, uint32_t))dlsym(h, "ref_create");
        auto ref_load = (int (*)(void *, const uint8_t *, uint64_t))dlsym(h, "ref_load");
        auto ref_reserve = (int (*)(void *, uint64_t, uint32_t))dlsym(h, "ref_reserve");
        auto ref_search_batch = (double (*)(void *, const float *, uint64_t, uint32_t, uint32_t, uint32_t, int64_t *, float *, uint32_t *, uint64_t *, uint64_t *))dlsym(h, "ref_search_batch");
        if (!ref_create || !ref_load || !ref_search_batch || !ref_reserve) { std::printf("reference unavailable: symbols\n"); return 0; }
        std::vector<unsigned char> blob;
        CHECK(gvss_save(ix, write_cb, &blob));
        void *ref = ref_create(d, 'e', M, 2 * M, cfg.expansion_add, 64);
        if (ref_load(ref, blob.data(), blob.size())) { std::printf("reference load failed\n"); return 2; }
        ref_reserve(ref, n, T); // thread ids 0..T-1
        std::vector<int64_t> rkeys((size_t)nq * k);
        std::vector<float> rdist((size_t)nq * k);
        ref_search_batch(ref, q.data(), 256, k, ef, T, rkeys.data(), rdist.data(), nullptr, nullptr, nullptr); // warm
        const double t1 = ref_search_batch(ref, q.data(), 512, k, ef, 1, rkeys.data(), rdist.data(), nullptr, nullptr, nullptr);
        const double tT = ref_search_batch(ref, q.data(), nq, k, ef, T, rkeys.data(), rdist.data(), nullptr, nullptr, nullptr);
        std::printf("reference single_query_mean_ms %.3f threads_qps %.0f same_ids %d\n", 1e3 * t1 / 512, nq / tT, (int)(rkeys == keys));
    }
    gvss_destroy(ix);
    return 0;
}
