// SURVEY.md 8(f) N3 — replays PhysicalCreateHNSWIndex's feeding pattern (T threads pull 2048-row chunks from a shared cursor and
// call add(key, vec, thread), hnsw_index_physical_create.cpp:144-205) against the C ABI, with and without the background
// builder, and reports wall clock from the first add to a searchable index. `rows_per_s` throttles the feeders to emulate the
// table scan that produces the rows (0 = as fast as memory allows).
//   sink_driver <n> <dims> <threads> <rows_per_s>
// Prints: one_shot_s, streamed_s, resident_build_s (rows already in HBM), recall of both graphs against the exact top-k.
#include "gvss.h"

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <thread>
#include <vector>

using clk = std::chrono::steady_clock;
static double secs(clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); }
#define CHECK(x) do { if (x) { std::printf("FAILED %s: %s\n", #x, gvss_last_error()); return 2; } } while (0)

static double feed(gvss_index *ix, const std::vector<float> &rows, uint64_t n, uint32_t d, uint32_t T, double rows_per_s, int *bad) {
    std::atomic<uint64_t> cursor{0};
    std::atomic<int> failures{0};
    const auto t0 = clk::now();
    std::vector<std::thread> ts;
    for (uint32_t t = 0; t < T; t++)
        ts.emplace_back([&, t] {
            for (;;) {
                const uint64_t b = cursor.fetch_add(2048);
                if (b >= n) break;
                if (rows_per_s > 0) { // the scan has not produced this chunk yet
                    const double due = (double)b / rows_per_s;
                    while (secs(t0, clk::now()) < due) std::this_thread::sleep_for(std::chrono::microseconds(200));
                }
                const uint64_t e = b + 2048 < n ? b + 2048 : n;
                for (uint64_t i = b; i < e; i++)
                    if (gvss_add(ix, (int64_t)i, rows.data() + i * d, t)) failures++;
            }
        });
    for (auto &t : ts) t.join();
    if (gvss_flush(ix)) failures++;
    *bad = failures.load();
    return secs(t0, clk::now());
}

static double recall(gvss_index *ix, const std::vector<float> &q, uint64_t nq, uint32_t k, const std::vector<int64_t> &truth) {
    std::vector<int64_t> keys(nq * k);
    if (gvss_search(ix, q.data(), nq, k, 128, keys.data(), nullptr, nullptr, nullptr)) return -1;
    uint64_t hit = 0;
    for (uint64_t i = 0; i < nq; i++)
        for (uint32_t a = 0; a < k; a++)
            for (uint32_t b = 0; b < k; b++) hit += keys[i * k + a] == truth[i * k + b];
    return (double)hit / (double)(nq * k);
}

int main(int argc, char **argv) {
    const uint64_t n = argc > 1 ? strtoull(argv[1], nullptr, 10) : 200000;
    const uint32_t d = argc > 2 ? (uint32_t)atoi(argv[2]) : 128, T = argc > 3 ? (uint32_t)atoi(argv[3]) : 8;
    const double rate = argc > 4 ? atof(argv[4]) : 0.0;
    const uint64_t nq = 500;
    const uint32_t k = 10;
    // clustered rows: centre + low-rank latent + noise (same law as the bench generator, libstdc++ RNG)
    std::mt19937_64 rng(7);
    std::normal_distribution<float> N01(0.f, 1.f);
    const uint32_t C = (uint32_t)(n / 100 ? n / 100 : 1), r = 16;
    std::vector<float> centres((size_t)C * d), proj((size_t)r * d), rows(n * d), q(nq * d);
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
    for (uint64_t i = 0; i < nq; i++) draw(q.data() + i * d);
    std::vector<int64_t> truth(nq * k);
    CHECK(gvss_exact_topk('e', rows.data(), n, nullptr, q.data(), nq, d, k, truth.data(), nullptr));

    gvss_config cfg{};
    cfg.dimensions = d; cfg.metric = 'e'; cfg.connectivity = 16; cfg.connectivity_base = 32; cfg.expansion_add = 128; cfg.expansion_search = 64; cfg.device = -1;
    double t_shot = 0, t_stream = 0, r_shot = 0, r_stream = 0;
    int bad = 0;
    for (int mode = 0; mode < 2; mode++) {
        gvss_index *ix = nullptr;
        CHECK(gvss_create(&cfg, &ix));
        CHECK(gvss_reserve(ix, n, T));
        if (mode) CHECK(gvss_set_background_build(ix, 32768));
        const double t = feed(ix, rows, n, d, T, rate, &bad);
        if (bad) { std::printf("FAILED: %d adds/flush failed: %s\n", bad, gvss_last_error()); return 2; }
        gvss_info info{};
        CHECK(gvss_get_info(ix, &info));
        if (info.size != n) { std::printf("FAILED: size %llu != %llu\n", (unsigned long long)info.size, (unsigned long long)n); return 2; }
        (mode ? t_stream : t_shot) = t;
        (mode ? r_stream : r_shot) = recall(ix, q, nq, k, truth);
        gvss_destroy(ix);
    }
    std::printf("n %llu dims %u threads %u scan_rows_per_s %.0f\n", (unsigned long long)n, d, T, rate);
    std::printf("one_shot_s %.3f streamed_s %.3f scan_s %.3f recall_one_shot %.4f recall_streamed %.4f\n", t_shot, t_stream,
                rate > 0 ? (double)n / rate : 0.0, r_shot, r_stream);
    return 0;
}
