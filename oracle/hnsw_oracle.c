/* TEST INFRASTRUCTURE ONLY — plain-C restatement of the reference's HNSW hot path.
 *
 * This file is the parity checker for the CUDA library. It is never linked into, imported by or
 * called from the product path (duckdb_vss_b200/, include/): only tests/, __graft_entry__.smoke()
 * and bench.py's cpu-baseline leg use it. Parity status: PINNED — tests/test_oracle.py checks it
 * (a) byte-for-byte against streams written by the compiled reference (oracle/_ref, single-thread
 * builds are deterministic), (b) id-for-id / bit-for-bit on search results of the compiled
 * reference, (c) against the golden fixtures under tests/golden/ and the known answers of the
 * reference's own tests (hnsw_result.test:20-25; DuckDB array_distance.test:20-27 etc.).
 *
 * Every function cites the reference lines it restates (paths relative to /root/reference/src/include/usearch
 * unless stated). Must be compiled with -ffp-contract=off so that fp32 sums round like the reference's
 * x86-64 build (separate multiply and add).
 */
#include "hnsw_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <stdio.h>
static int g_trace = 0;
void orc_set_trace(int on) { g_trace = on; }
#define FREE_KEY INT64_MAX /* index_dense.hpp: free_key_ = default_free_value<vector_key_t>() */

typedef struct { float distance; uint32_t slot; } cand_t; /* index.hpp:1805-1809, ordered by distance only */

struct orc_index {
    uint32_t dims; char metric; uint32_t M, M0, ef_add, ef_search;
    uint64_t count, capacity; int32_t max_level; uint64_t entry;
    float *vectors; int64_t *keys; int16_t *levels;
    uint32_t *nbr0;    /* capacity × (1+M0): count, slots              (node tape level 0, index.hpp:3152-3166) */
    uint32_t **upper;  /* per node: level × (1+M) u32 or NULL          (levels 1..L) */
    uint64_t rng;      /* std::default_random_engine == minstd_rand0 state (index.hpp:1892) */
    uint64_t computed, cycles; /* context_t::computed_distances_count / iteration_cycles (index.hpp:1893-1914) */
    uint8_t *visits;   /* membership only: any exact set is equivalent to growing_hash_set_gt (index.hpp:852-970) */
    cand_t *top; uint64_t top_size;   /* sorted_buffer_gt (index.hpp:655-766) */
    cand_t *next; uint64_t next_size, next_cap; /* max_heap_gt on -distance (index.hpp:516-645) */
    uint32_t next_peak;
};

/* ---- metrics: index_plugins.hpp:981-993 (ip), 1006-1026 (cos), 1037-1052 (l2sq); strict left-to-right fp32 ---- */
static float m_ip(const float *a, const float *b, uint32_t d) {
    float ab = 0; for (uint32_t i = 0; i < d; i++) ab += a[i] * b[i]; return 1 - ab;
}
static float m_cos(const float *a, const float *b, uint32_t d) {
    float ab = 0, a2 = 0, b2 = 0;
    for (uint32_t i = 0; i < d; i++) { float ai = a[i], bi = b[i]; ab += ai * bi; a2 += ai * ai; b2 += bi * bi; }
    if (a2 == 0 && b2 == 0) return 0;
    if (a2 == 0 || b2 == 0) return 1;
    return 1 - ab / (sqrtf(a2) * sqrtf(b2));
}
static float m_l2sq(const float *a, const float *b, uint32_t d) {
    float s = 0; for (uint32_t i = 0; i < d; i++) { float t = a[i] - b[i]; s += t * t; } return s;
}
float orc_metric(char metric, const float *a, const float *b, uint32_t d) {
    return metric == 'c' ? m_cos(a, b, d) : metric == 'i' ? m_ip(a, b, d) : m_l2sq(a, b, d);
}

/* ---- DuckDB folds: duckdb/extension/core_functions/include/core_functions/array_kernels.hpp:11-105 ---- */
float orc_duckdb_distance(char metric, const float *a, const float *b, uint32_t d) {
    if (metric == 'i') { float r = 0; for (uint32_t i = 0; i < d; i++) r += a[i] * b[i]; return -r; }
    if (metric == 'c') {
        float dist = 0, nl = 0, nr = 0;
        for (uint32_t i = 0; i < d; i++) { float x = a[i], y = b[i]; dist += x * y; nl += x * x; nr += y * y; }
        float sim = dist / sqrtf(nl * nr);
        /* std::max(-1, std::min(sim, 1)) with their NaN behaviour: min(sim,1) = (1 < sim) ? 1 : sim ; max(-1,x) = (-1 < x) ? x : -1 */
        float mn = (1.0f < sim) ? 1.0f : sim;
        float mx = (-1.0f < mn) ? mn : -1.0f;
        return 1.0f - mx;
    }
    float s = 0; for (uint32_t i = 0; i < d; i++) { float t = a[i] - b[i]; s += t * t; } return sqrtf(s);
}

/* brute force ordered by (distance, rowid): SURVEY §8(c) comparator `ORDER BY array_distance(..), rowid LIMIT k` */
void orc_exact_topk(char metric, const float *base, uint64_t n, const float *qs, uint64_t nq, uint32_t d, uint32_t k,
                    int64_t *ids_out, float *dist_out) {
    for (uint64_t qi = 0; qi < nq; qi++) {
        int64_t *ids = ids_out + qi * k; float *ds = dist_out + qi * k; uint32_t m = 0;
        for (uint64_t i = 0; i < n; i++) {
            float dist = orc_duckdb_distance(metric, base + i * d, qs + qi * d, d);
            /* NaN sorts last (DuckDB orders NaN as the greatest float) */
            uint32_t pos = m;
            while (pos > 0) {
                float o = ds[pos - 1];
                int less = isnan(o) ? !isnan(dist) : (!isnan(dist) && dist < o);
                if (!less) break;
                pos--;
            }
            if (pos >= k) continue;
            uint32_t last = m < k ? m : k - 1;
            for (uint32_t j = last; j > pos; j--) { ds[j] = ds[j - 1]; ids[j] = ids[j - 1]; }
            ds[pos] = dist; ids[pos] = (int64_t)i; if (m < k) m++;
        }
        for (uint32_t j = m; j < k; j++) { ids[j] = -1; ds[j] = INFINITY; }
    }
}

/* ---- level law: index.hpp:3315-3319 over libstdc++'s minstd_rand0 + generate_canonical<double,53> (2 draws) ---- */
static uint32_t minstd_next(uint64_t *s) { *s = (*s * 16807ull) % 2147483647ull; return (uint32_t)*s; }
int16_t orc_level_from_state(uint64_t *state, uint32_t M) {
    const double R = 2147483646.0; /* max - min + 1 = (2^31-2) - 1 + 1 */
    double sum = 0, tmp = 1;
    for (int k = 0; k < 2; k++) { sum += (double)(minstd_next(state) - 1u) * tmp; tmp *= R; }
    double u = sum / tmp;
    if (u >= 1.0) u = nextafter(1.0, 0.0);
    double r = -log(u) * (1.0 / log((double)M));
    return (int16_t)r;
}

/* ---- container ---- */
orc_index *orc_create(uint32_t dims, char metric, uint32_t M, uint32_t M0, uint32_t ef_add, uint32_t ef_search) {
    orc_index *x = (orc_index *)calloc(1, sizeof(orc_index));
    x->dims = dims; x->metric = metric;
    x->M = M ? M : 16; x->M0 = M0 ? M0 : 2 * x->M;      /* index.hpp:1097-1107 defaults; hnsw_index.cpp:178-187 */
    x->ef_add = ef_add ? ef_add : 128; x->ef_search = ef_search ? ef_search : 64;
    x->max_level = -1; x->rng = 1;
    return x;
}
static void free_nodes(orc_index *x) {
    if (x->upper) for (uint64_t i = 0; i < x->capacity; i++) free(x->upper[i]);
    free(x->vectors); free(x->keys); free(x->levels); free(x->nbr0); free(x->upper); free(x->visits);
    x->vectors = NULL; x->keys = NULL; x->levels = NULL; x->nbr0 = NULL; x->upper = NULL; x->visits = NULL;
}
void orc_destroy(orc_index *x) { if (!x) return; free_nodes(x); free(x->top); free(x->next); free(x); }

/* index.hpp:2104-2127 + index_dense.hpp:753-765: grow only; fresh thread contexts => level RNG restarts */
int orc_reserve(orc_index *x, uint64_t members) {
    if (members < x->count) members = x->count;
    uint64_t old = x->capacity;
    x->vectors = (float *)realloc(x->vectors, (size_t)members * x->dims * sizeof(float) + 16);
    x->keys = (int64_t *)realloc(x->keys, (size_t)members * 8 + 8);
    x->levels = (int16_t *)realloc(x->levels, (size_t)members * 2 + 2);
    x->nbr0 = (uint32_t *)realloc(x->nbr0, (size_t)members * (1 + x->M0) * 4 + 4);
    x->upper = (uint32_t **)realloc(x->upper, (size_t)members * sizeof(uint32_t *) + 8);
    x->visits = (uint8_t *)realloc(x->visits, (size_t)members + 1);
    for (uint64_t i = old; i < members; i++) x->upper[i] = NULL;
    x->capacity = members; x->rng = 1;
    return 0;
}
uint64_t orc_node_count(const orc_index *x) { return x->count; }
uint64_t orc_size(const orc_index *x) { uint64_t n = 0; for (uint64_t i = 0; i < x->count; i++) n += x->keys[i] != FREE_KEY; return n; }
int32_t orc_max_level(const orc_index *x) { return x->max_level < 0 ? 0 : x->max_level; }
uint64_t orc_entry_slot(const orc_index *x) { return x->entry; }

static inline uint32_t *nbrs(orc_index *x, uint64_t slot, int level) { /* index.hpp:3204-3212 */
    return level ? x->upper[slot] + (size_t)(level - 1) * (1 + x->M) : x->nbr0 + slot * (1 + x->M0);
}
static inline const float *vec(const orc_index *x, uint64_t slot) { return x->vectors + slot * x->dims; }
static inline float measure(orc_index *x, const float *a, const float *b) { x->computed++; return orc_metric(x->metric, a, b, x->dims); }

/* sorted_buffer_gt::insert_reserved (index.hpp:725-733) / insert with limit (:738-749): lower_bound => equal goes first */
static uint64_t top_lower_bound(const orc_index *x, float d) {
    uint64_t lo = 0, hi = x->top_size;
    while (lo < hi) { uint64_t mid = (lo + hi) / 2; if (x->top[mid].distance < d) lo = mid + 1; else hi = mid; }
    return lo;
}
static void top_reserve(orc_index *x, uint64_t n) { x->top = (cand_t *)realloc(x->top, (n + 2) * sizeof(cand_t)); }
static void top_insert_reserved(orc_index *x, cand_t e) {
    uint64_t slot = top_lower_bound(x, e.distance);
    for (uint64_t j = x->top_size; j > slot; j--) x->top[j] = x->top[j - 1];
    x->top[slot] = e; x->top_size++;
}
static int top_insert(orc_index *x, cand_t e, uint64_t limit) {
    uint64_t slot = top_lower_bound(x, e.distance);
    if (slot == limit) return 0;
    int full = x->top_size == limit;
    for (uint64_t j = x->top_size - (full ? 1 : 0); j > slot; j--) x->top[j] = x->top[j - 1];
    x->top[slot] = e; if (!full) x->top_size++;
    return 1;
}
/* max_heap_gt of {-distance, slot}: shift_up/shift_down exactly as index.hpp:624-644 */
static void next_push(orc_index *x, cand_t e) {
    if (x->next_size + 1 > x->next_cap) { x->next_cap = x->next_cap ? x->next_cap * 2 : 64; x->next = (cand_t *)realloc(x->next, x->next_cap * sizeof(cand_t)); }
    uint64_t i = x->next_size++; x->next[i] = e;
    while (i) { uint64_t p = (i - 1) / 2; if (!(x->next[p].distance < x->next[i].distance)) break; cand_t t = x->next[p]; x->next[p] = x->next[i]; x->next[i] = t; i = p; }
    if (x->next_size > x->next_peak) x->next_peak = (uint32_t)x->next_size;
}
static cand_t next_pop(orc_index *x) {
    cand_t r = x->next[0]; cand_t t = x->next[0]; x->next[0] = x->next[x->next_size - 1]; x->next[x->next_size - 1] = t;
    x->next_size--;
    uint64_t i = 0;
    for (;;) {
        uint64_t mx = i, l = 2 * i + 1, rr = 2 * i + 2;
        if (l < x->next_size && x->next[mx].distance < x->next[l].distance) mx = l;
        if (rr < x->next_size && x->next[mx].distance < x->next[rr].distance) mx = rr;
        if (mx == i) break;
        t = x->next[i]; x->next[i] = x->next[mx]; x->next[mx] = t; i = mx;
    }
    return r;
}

/* search_for_one_: index.hpp:3384-3422 */
static uint64_t search_for_one(orc_index *x, const float *q, uint64_t closest, int begin_level, int end_level) {
    float cd = measure(x, q, vec(x, closest));
    for (int level = begin_level; level > end_level; --level) {
        int changed;
        do {
            changed = 0;
            uint32_t *nb = nbrs(x, closest, level); uint32_t cnt = nb[0];
            for (uint32_t j = 0; j < cnt; j++) {
                uint32_t c = nb[1 + j];
                float d = measure(x, q, vec(x, c));
                if (d < cd) { cd = d; closest = c; changed = 1; }
            }
            x->cycles++;
        } while (changed);
    }
    return closest;
}

/* search_to_find_in_base_: index.hpp:3504-3573 (predicate = key != free_key, index_dense.hpp:1815-1820).
 * Deviation, documented in SURVEY Appendix A(4): while `top` is empty the reference reads elements_[-1] (UB);
 * here radius is +inf while top is empty. */
static void search_base(orc_index *x, const float *q, uint64_t start, uint64_t limit, uint64_t *hops) {
    memset(x->visits, 0, x->count);
    x->next_size = 0; x->top_size = 0; top_reserve(x, limit);
    float radius = measure(x, q, vec(x, start));
    cand_t e = {-radius, (uint32_t)start}; next_push(x, e);
    x->visits[start] = 1;
    if (x->keys[start] != FREE_KEY) { cand_t t = {radius, (uint32_t)start}; top_insert_reserved(x, t); }
    else radius = INFINITY;
    while (x->next_size) {
        cand_t c = x->next[0];
        if ((-c.distance) > radius) break;
        next_pop(x); x->cycles++; if (hops) (*hops)++;
        if (g_trace) printf("pop %llu slot %u d %.9g radius %.9g heap %llu top %llu computed %llu\n", hops ? (unsigned long long)*hops : 0ull, c.slot, -c.distance, radius, (unsigned long long)x->next_size, (unsigned long long)x->top_size, (unsigned long long)x->computed);
        uint32_t *nb = nbrs(x, c.slot, 0); uint32_t cnt = nb[0];
        for (uint32_t j = 0; j < cnt; j++) {
            uint32_t s = nb[1 + j];
            if (x->visits[s]) continue;
            x->visits[s] = 1;
            float d = measure(x, q, vec(x, s));
            if (x->top_size < limit || d < radius) {
                cand_t n = {-d, s}; next_push(x, n);
                if (x->keys[s] != FREE_KEY) { cand_t t = {d, s}; top_insert(x, t, limit); }
                radius = x->top_size ? x->top[x->top_size - 1].distance : INFINITY;
            }
        }
    }
}

/* search_to_insert_: index.hpp:3430-3496 */
static void search_insert(orc_index *x, const float *q, uint64_t start, uint64_t new_slot, int level, uint64_t limit) {
    memset(x->visits, 0, x->count);
    x->next_size = 0; x->top_size = 0;
    float radius = measure(x, q, vec(x, start));
    cand_t e = {-radius, (uint32_t)start}; next_push(x, e);
    cand_t t0 = {radius, (uint32_t)start}; top_insert_reserved(x, t0);
    x->visits[start] = 1;
    while (x->next_size) {
        cand_t c = x->next[0];
        if ((-c.distance) > radius && x->top_size == limit) break;
        next_pop(x); x->cycles++;
        if (c.slot == new_slot) continue;
        uint32_t *nb = nbrs(x, c.slot, level); uint32_t cnt = nb[0];
        for (uint32_t j = 0; j < cnt; j++) {
            uint32_t s = nb[1 + j];
            if (x->visits[s]) continue;
            x->visits[s] = 1;
            float d = measure(x, q, vec(x, s));
            if (x->top_size < limit || d < radius) {
                cand_t n = {-d, s}; next_push(x, n);
                cand_t t = {d, s}; top_insert(x, t, limit);
                radius = x->top[x->top_size - 1].distance;
            }
        }
    }
}

/* refine_: index.hpp:3602-3638 — keep c iff for every kept s: d(c,s) >= d(c,q) */
static uint64_t refine(orc_index *x, uint64_t needed) {
    uint64_t n = x->top_size;
    if (n < needed) return n;
    uint64_t submitted = 1, consumed = 1;
    while (submitted < needed && consumed < n) {
        cand_t c = x->top[consumed]; int good = 1;
        for (uint64_t i = 0; i < submitted; i++) {
            float d = measure(x, vec(x, c.slot), vec(x, x->top[i].slot));
            if (d < c.distance) { good = 0; break; }
        }
        if (good) x->top[submitted++] = x->top[consumed];
        consumed++;
    }
    x->top_size = submitted;
    return submitted;
}

/* connect_node_across_levels_ (index.hpp:3227-3244) + connect_new_node_ (:3247-3267) + reconnect_neighbor_nodes_ (:3270-3313) */
static void connect_across_levels(orc_index *x, const float *value, uint64_t node, uint64_t entry, int max_level, int target) {
    uint64_t closest = search_for_one(x, value, entry, max_level, target);
    for (int level = target < max_level ? target : max_level; level >= 0; --level) {
        search_insert(x, value, closest, node, level, x->ef_add);
        uint32_t *mine = nbrs(x, node, level);
        uint64_t kept = refine(x, x->M);
        /* documented clamp: the reference writes up to M links into a level-0 list sized M0 even if M0 < M */
        uint32_t cap_here = level ? x->M : x->M0; if (kept > cap_here) kept = cap_here;
        for (uint64_t i = 0; i < kept; i++) mine[1 + mine[0]++] = x->top[i].slot;
        closest = mine[1];
        uint32_t cmax = level ? x->M : x->M0;
        uint32_t my_cnt = mine[0];
        for (uint32_t i = 0; i < my_cnt; i++) {
            uint32_t close = mine[1 + i];
            if (close == node) continue;
            uint32_t *hdr = nbrs(x, close, level);
            if (hdr[0] < cmax) { hdr[1 + hdr[0]++] = (uint32_t)node; continue; }
            x->top_size = 0;
            cand_t a = {measure(x, value, vec(x, close)), (uint32_t)node}; top_insert_reserved(x, a);
            for (uint32_t j = 0; j < hdr[0]; j++) {
                uint32_t succ = hdr[1 + j];
                cand_t b = {measure(x, vec(x, close), vec(x, succ)), succ}; top_insert_reserved(x, b);
            }
            uint32_t oldn = hdr[0]; hdr[0] = 0;
            uint64_t k2 = refine(x, cmax);
            for (uint64_t j = 0; j < k2; j++) hdr[1 + hdr[0]++] = x->top[j].slot;
            for (uint32_t j = hdr[0]; j < oldn; j++) hdr[1 + j] = 0; /* clear() + push_back leaves zeros? keep tail zeroed for byte parity */
        }
    }
}

/* index_gt::add (index.hpp:2297-2378) + update for a recycled slot (:2405-2463; index_dense.hpp:1766-1793) */
int orc_add(orc_index *x, int64_t key, const float *v, uint64_t *computed, uint64_t *visited) {
    uint64_t c0 = x->computed, v0 = x->cycles;
    uint64_t cmx = x->M0 > x->M ? x->M0 : x->M;
    uint64_t lim = cmx + 1 > x->ef_add ? cmx + 1 : x->ef_add;
    top_reserve(x, lim + 1);
    /* removed slot to recycle? (free-list order: ring FIFO, index_dense.hpp:1766-1771) — the port recycles the lowest free slot;
       only used by CRUD tests where at most one slot is free */
    int64_t reuse = -1;
    for (uint64_t i = 0; i < x->count; i++) if (x->keys[i] == FREE_KEY) { reuse = (int64_t)i; break; }
    if (reuse >= 0) {
        uint64_t s = (uint64_t)reuse; int lvl = x->levels[s];
        memset(nbrs(x, s, 0), 0, (1 + x->M0) * 4);
        if (lvl) memset(x->upper[s], 0, (size_t)lvl * (1 + x->M) * 4);
        memcpy(x->vectors + s * x->dims, v, x->dims * 4);
        connect_across_levels(x, v, s, x->entry, x->max_level, lvl);
        x->keys[s] = key;
    } else {
        int16_t level = orc_level_from_state(&x->rng, x->M);
        if (x->count >= x->capacity) return 1; /* "Reserve capacity ahead of insertions!" index.hpp:2332-2335 */
        uint64_t s = x->count++;
        x->keys[s] = key; x->levels[s] = level;
        memcpy(x->vectors + s * x->dims, v, x->dims * 4);
        memset(nbrs(x, s, 0), 0, (1 + x->M0) * 4);
        free(x->upper[s]); x->upper[s] = level ? (uint32_t *)calloc((size_t)level * (1 + x->M), 4) : NULL;
        if (s == 0) { x->entry = 0; x->max_level = level; }
        else {
            int ml = x->max_level; uint64_t en = x->entry;
            connect_across_levels(x, v, s, en, ml, level);
            if (level > ml) { x->entry = s; x->max_level = level; }
        }
    }
    if (computed) *computed = x->computed - c0;
    if (visited) *visited = x->cycles - v0;
    return 0;
}

/* index_gt::search (index.hpp:2480-2534) via index_dense_gt::search_ (index_dense.hpp:1797-1827) */
uint32_t orc_search(orc_index *x, const float *q, uint32_t k, uint32_t ef, int64_t *keys, float *dists,
                    uint64_t *computed, uint64_t *visited, uint64_t *base_hops, uint32_t *next_peak) {
    if (computed) *computed = 0;
    if (visited) *visited = 0;
    if (base_hops) *base_hops = 0;
    if (!k || !x->count) return 0;
    if (!ef) ef = 64; /* default_expansion_search, index.hpp:1125 */
    uint64_t c0 = x->computed, v0 = x->cycles;
    uint64_t expansion = ef > k ? ef : k;
    x->next_peak = 0;
    uint64_t closest = search_for_one(x, q, x->entry, x->max_level, 0);
    search_base(x, q, closest, expansion, base_hops);
    uint32_t n = (uint32_t)(x->top_size < k ? x->top_size : k);
    for (uint32_t i = 0; i < n; i++) { keys[i] = x->keys[x->top[i].slot]; if (dists) dists[i] = x->top[i].distance; }
    if (computed) *computed = x->computed - c0;
    if (visited) *visited = x->cycles - v0;
    if (next_peak) *next_peak = x->next_peak;
    return n;
}

/* index_dense_gt::remove (index_dense.hpp:1228-1255): lazy delete, links stay */
uint64_t orc_remove(orc_index *x, int64_t key) {
    uint64_t n = 0;
    for (uint64_t i = 0; i < x->count; i++) if (x->keys[i] == key && key != FREE_KEY) { x->keys[i] = FREE_KEY; n++; }
    return n;
}

/* index_gt::stats(level) (index.hpp:2612-2635) incl. the swapped max_edges quirk (:2628-2629) */
void orc_level_stats(const orc_index *x, uint32_t level, uint64_t *out4) {
    uint64_t nodes = 0, edges = 0, bytes = 0;
    orc_index *y = (orc_index *)x;
    for (uint64_t i = 0; i < x->count; i++) {
        if ((uint32_t)x->levels[i] < level) continue;
        nodes++; edges += nbrs(y, i, (int)level)[0];
        bytes += (level ? (1 + x->M) : (1 + x->M0)) * 4;
    }
    out4[0] = nodes; out4[1] = edges; out4[2] = nodes * (level ? x->M0 : x->M); out4[3] = bytes;
}

/* ---- stream: index_dense.hpp:811-878 / 900-973 and index.hpp:2709-2749 / 2755-2807 ---- */
static uint64_t node_bytes(const orc_index *x, int level) { return 10 + (4 + 4ull * x->M0) + (uint64_t)level * (4 + 4ull * x->M); }
uint64_t orc_serialized_length(const orc_index *x) {
    uint64_t n = 8 + x->count * x->dims * 4ull + 64 + 40 + 2 * x->count;
    for (uint64_t i = 0; i < x->count; i++) n += node_bytes(x, x->levels[i]);
    return n;
}
uint64_t orc_save(const orc_index *x, uint8_t *buf, uint64_t cap) {
    uint64_t need = orc_serialized_length(x); if (need > cap) return 0;
    uint8_t *p = buf; orc_index *y = (orc_index *)x;
    uint32_t rows = (uint32_t)x->count, cols = x->dims * 4;
    memcpy(p, &rows, 4); p += 4; memcpy(p, &cols, 4); p += 4;
    memcpy(p, x->vectors, (size_t)rows * cols); p += (size_t)rows * cols;
    uint8_t head[64]; memset(head, 0, 64);
    memcpy(head, "usearch", 7);
    uint16_t ver[3] = {2, 12, 0}; memcpy(head + 7, ver, 6);
    head[13] = (uint8_t)x->metric; head[14] = 11; head[15] = 20; head[16] = 15; /* f32_k, i64 key kind, u32 slot kind */
    uint64_t present = orc_size(x), deleted = x->count - present, dims = x->dims;
    memcpy(head + 17, &present, 8); memcpy(head + 25, &deleted, 8); memcpy(head + 33, &dims, 8); head[41] = 0;
    memcpy(p, head, 64); p += 64;
    uint64_t gh[5] = {x->count, x->M, x->M0, (uint64_t)orc_max_level(x), x->entry};
    memcpy(p, gh, 40); p += 40;
    for (uint64_t i = 0; i < x->count; i++) { memcpy(p, &x->levels[i], 2); p += 2; }
    for (uint64_t i = 0; i < x->count; i++) {
        memcpy(p, &x->keys[i], 8); p += 8; memcpy(p, &x->levels[i], 2); p += 2;
        memcpy(p, nbrs(y, i, 0), (1 + x->M0) * 4); p += (1 + x->M0) * 4;
        if (x->levels[i]) { memcpy(p, x->upper[i], (size_t)x->levels[i] * (1 + x->M) * 4); p += (size_t)x->levels[i] * (1 + x->M) * 4; }
    }
    return (uint64_t)(p - buf);
}
int orc_load(orc_index *x, const uint8_t *buf, uint64_t n) {
    const uint8_t *p = buf, *end = buf + n;
    if (n < 8) return 1;
    uint32_t rows, cols; memcpy(&rows, p, 4); memcpy(&cols, p + 4, 4); p += 8;
    if (cols != x->dims * 4) return 2;
    if ((uint64_t)(end - p) < (uint64_t)rows * cols + 104) return 3;
    const uint8_t *vecs = p; p += (size_t)rows * cols;
    if (memcmp(p, "usearch", 7)) return 4;
    uint16_t major; memcpy(&major, p + 7, 2); if (major != 2) return 5;           /* index_dense.hpp:949 */
    if (p[15] != 20 || p[16] != 15) return 6;                                       /* key / slot kind, :951-957 */
    x->metric = (char)p[13];
    p += 64;
    uint64_t gh[5]; memcpy(gh, p, 40); p += 40;
    if (gh[0] != rows) return 7;                                                     /* :965-969 */
    free_nodes(x); x->capacity = 0; x->count = 0;
    x->M = (uint32_t)gh[1]; x->M0 = (uint32_t)gh[2];
    orc_reserve(x, rows);
    x->count = rows; x->max_level = rows ? (int32_t)gh[3] : -1; x->entry = gh[4];
    memcpy(x->vectors, vecs, (size_t)rows * cols);
    if ((uint64_t)(end - p) < 2ull * rows) return 8;
    p += 2ull * rows;
    for (uint64_t i = 0; i < rows; i++) {
        if (end - p < 10) return 9;
        memcpy(&x->keys[i], p, 8); memcpy(&x->levels[i], p + 8, 2); p += 10;
        uint64_t nb = node_bytes(x, x->levels[i]) - 10; if ((uint64_t)(end - p) < nb) return 10;
        memcpy(nbrs(x, i, 0), p, (1 + x->M0) * 4); p += (1 + x->M0) * 4;
        if (x->levels[i]) {
            x->upper[i] = (uint32_t *)malloc((size_t)x->levels[i] * (1 + x->M) * 4);
            memcpy(x->upper[i], p, (size_t)x->levels[i] * (1 + x->M) * 4); p += (size_t)x->levels[i] * (1 + x->M) * 4;
        }
    }
    return 0;
}
