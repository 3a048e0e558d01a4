/* TEST INFRASTRUCTURE ONLY -- internals shared by dbt_oracle.c and dbt_oracle_thread.c. */
#ifndef DBT_ORACLE_INT_H_
#define DBT_ORACLE_INT_H_
#include "dbt_oracle.h"
#include <stdlib.h>
#include <string.h>

#define NAN32 0xFFFFFFFFu
#define NAN64 0xFFFFFFFFFFFFFFFFull

/* tiny open-addressing maps (stand-ins for the reference's unordered_map / unordered_set) */
typedef struct { uint64_t cap; uint64_t* k; uint64_t* k2; uint32_t* v; uint8_t* used; } omap;

static inline uint64_t mix64(uint64_t x) { x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33; return x; }

static inline int omap_init(omap* m, uint64_t n) {
	uint64_t cap = 16; while (cap < 2*n + 2) cap <<= 1;
	m->cap = cap;
	m->k = (uint64_t*)malloc(cap*8); m->k2 = (uint64_t*)malloc(cap*8);
	m->v = (uint32_t*)malloc(cap*4); m->used = (uint8_t*)calloc(cap, 1);
	return m->k && m->k2 && m->v && m->used;
}
static inline void omap_free(omap* m) { free(m->k); free(m->k2); free(m->v); free(m->used); }
/* last write wins, like operator[] assignment in the reference loaders */
static inline void omap_put(omap* m, uint64_t key, uint64_t key2, uint32_t val) {
	uint64_t i = mix64(key ^ mix64(key2 + 0x9e3779b97f4a7c15ull)) & (m->cap-1);
	while (m->used[i]) { if (m->k[i] == key && m->k2[i] == key2) { m->v[i] = val; return; } i = (i+1) & (m->cap-1); }
	m->used[i] = 1; m->k[i] = key; m->k2[i] = key2; m->v[i] = val;
}
static inline int omap_get(const omap* m, uint64_t key, uint64_t key2, uint32_t* val) {
	uint64_t i = mix64(key ^ mix64(key2 + 0x9e3779b97f4a7c15ull)) & (m->cap-1);
	while (m->used[i]) { if (m->k[i] == key && m->k2[i] == key2) { if (val) *val = m->v[i]; return 1; } i = (i+1) & (m->cap-1); }
	return 0;
}


struct dbo_index {
	uint64_t nloci, nk_tr, nvv;
	omap dbi;        /* kmerDBi: k-mer -> val (h:35, kio:59-74)                                   */
	uint32_t* vv;    /* kmerDBi_vv                                                               */
	omap tr;         /* trResults[locus]: (k-mer, locus) -> position in the count array (h:31)    */
	omap fl;         /* flankDB[locus]: (k-mer, locus) membership (h:47)                          */
	uint8_t* qc;     /* qcFilter or NULL                                                         */
	int has_graph;
	omap graph;      /* graphDB[locus]: (strand-specific k-mer, locus) -> 4-bit out-edge mask (h:32) */
};
#endif
