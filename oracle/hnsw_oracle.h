/* TEST INFRASTRUCTURE ONLY — CPU restatement of the reference HNSW hot path (see hnsw_oracle.c).
 * Never linked into or called from the product library. */
#ifndef HNSW_ORACLE_H
#define HNSW_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_index orc_index;

orc_index *orc_create(uint32_t dims, char metric, uint32_t M, uint32_t M0, uint32_t ef_add, uint32_t ef_search);
void orc_destroy(orc_index *);
int orc_reserve(orc_index *, uint64_t members);
int orc_add(orc_index *, int64_t key, const float *vec, uint64_t *computed, uint64_t *visited);
uint32_t orc_search(orc_index *, const float *q, uint32_t k, uint32_t ef, int64_t *keys, float *dists,
                    uint64_t *computed, uint64_t *visited, uint64_t *base_hops, uint32_t *next_peak);
uint64_t orc_remove(orc_index *, int64_t key);
uint64_t orc_size(const orc_index *);
uint64_t orc_node_count(const orc_index *);
int32_t orc_max_level(const orc_index *);
uint64_t orc_entry_slot(const orc_index *);
uint64_t orc_serialized_length(const orc_index *);
uint64_t orc_save(const orc_index *, uint8_t *buf, uint64_t cap);
int orc_load(orc_index *, const uint8_t *buf, uint64_t n);
void orc_level_stats(const orc_index *, uint32_t level, uint64_t *out4);

float orc_metric(char metric, const float *a, const float *b, uint32_t d);
float orc_duckdb_distance(char metric, const float *a, const float *b, uint32_t d);
void orc_exact_topk(char metric, const float *base, uint64_t n, const float *qs, uint64_t nq, uint32_t d, uint32_t k,
                    int64_t *ids_out, float *dist_out);
int16_t orc_level_from_state(uint64_t *minstd_state, uint32_t M);

#ifdef __cplusplus
}
#endif
#endif
