// Times gvss_save / gvss_load (usearch v2.12 stream through the same kind of callbacks HNSWIndex::PersistToDisk and the
// HNSWIndex constructor use, hnsw_index.cpp:193-206, 510-532) against an in-memory sink, and checks that the reloaded index
// answers like the saved one.   persist_driver <n> <dims>
#include "gvss.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

using clk = std::chrono::steady_clock;
static double secs(clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); }
#define CHECK(x) do { if (x) { std::printf("FAILED %s: %s\n", #x, gvss_last_error()); return 2; } } while (0)

struct Sink { std::vector<unsigned char> bytes; size_t pos = 0; };
static int write_cb(void *u, const void *p, size_t n) {
    auto *s = static_cast<Sink *>(u);
    s->bytes.insert(s->bytes.end(), (const unsigned char *)p, (const unsigned char *)p + n);
    return 1;
}
static int read_cb(void *u, void *p, size_t n) {
    auto *s = static_cast<Sink *>(u);
    if (s->pos + n > s->bytes.size()) return 0;
    std::memcpy(p, s->bytes.data() + s->pos, n);
    s->pos += n;
    return 1;
}

int main(int argc, char **argv) {
    const uint64_t n = argc > 1 ? strtoull(argv[1], nullptr, 10) : 200000;
    const uint32_t d = argc > 2 ? (uint32_t)atoi(argv[2]) : 128;
    std::mt19937_64 rng(11);
    std::normal_distribution<float> N01(0.f, 1.f);
    std::vector<float> rows(n * d), q(64 * d);
    for (auto &v : rows) v = N01(rng);
    for (auto &v : q) v = N01(rng);
    gvss_config cfg{};
    cfg.dimensions = d; cfg.metric = 'e'; cfg.device = -1;
    gvss_index *ix = nullptr, *iy = nullptr;
    CHECK(gvss_create(&cfg, &ix));
    CHECK(gvss_reserve(ix, n, 1));
    CHECK(gvss_add_batch(ix, nullptr, rows.data(), n));
    CHECK(gvss_flush(ix));
    uint64_t len = 0;
    CHECK(gvss_serialized_length(ix, &len));
    Sink sink;
    sink.bytes.reserve(len);
    CHECK(gvss_save(ix, write_cb, &sink)); // warm: pinned allocations, first-touch of the sink
    sink.bytes.clear();
    auto t0 = clk::now();
    CHECK(gvss_save(ix, write_cb, &sink));
    const double t_save = secs(t0, clk::now());
    CHECK(gvss_create(&cfg, &iy));
    t0 = clk::now();
    CHECK(gvss_load(iy, read_cb, &sink));
    const double t_load = secs(t0, clk::now());
    std::vector<int64_t> a(64 * 10), b(64 * 10);
    CHECK(gvss_search(ix, q.data(), 64, 10, 64, a.data(), nullptr, nullptr, nullptr));
    CHECK(gvss_search(iy, q.data(), 64, 10, 64, b.data(), nullptr, nullptr, nullptr));
    std::printf("n %llu dims %u stream_bytes %zu serialized_length %llu save_s %.3f (%.2f GB/s) load_s %.3f (%.2f GB/s) same_results %d\n",
                (unsigned long long)n, d, sink.bytes.size(), (unsigned long long)len, t_save, sink.bytes.size() / t_save / 1e9, t_load,
                sink.bytes.size() / t_load / 1e9, (int)(a == b));
    gvss_destroy(ix);
    gvss_destroy(iy);
    return 0;
}
