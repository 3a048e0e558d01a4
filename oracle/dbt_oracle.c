/*
 * TEST INFRASTRUCTURE ONLY -- see dbt_oracle.h.  CPU restatement of the reference algorithm.
 * Every function cites the reference lines it follows ("cpp" = src/aQueryFasta_thread.cpp,
 * "h" = src/aQueryFasta_thread.h under /root/reference).  Written for clarity, not speed.
 */
#include "dbt_oracle_int.h"


dbo_index* dbo_index_new(const dbt_index* in) {
	dbo_index* ix = (dbo_index*)calloc(1, sizeof(*ix));
	if (!ix) return NULL;
	ix->nloci = in->nloci; ix->nvv = in->nvv;
	uint64_t ntr = 0, nfl = 0;
	for (uint64_t l = 0; l < in->nloci; ++l) { ntr += in->tr_n[l]; nfl += in->fl_n ? in->fl_n[l] : 0; }
	ix->nk_tr = ntr;
	if (!omap_init(&ix->dbi, in->nk_dbi) || !omap_init(&ix->tr, ntr) || !omap_init(&ix->fl, nfl)) return NULL;
	/* readBinaryIndex, h:671: kmerDBi[key] = val in file order */
	for (uint64_t i = 0; i < in->nk_dbi; ++i) omap_put(&ix->dbi, in->dbi_keys[i], 0, in->dbi_vals[i]);
	ix->vv = (uint32_t*)malloc((in->nvv + 1) * 4);
	if (in->nvv) memcpy(ix->vv, in->vv, in->nvv * 4);
	uint64_t j = 0;
	for (uint64_t l = 0; l < in->nloci; ++l)
		for (uint64_t i = 0; i < in->tr_n[l]; ++i, ++j) omap_put(&ix->tr, in->tr_kmers[j], l, (uint32_t)j);
	j = 0;
	for (uint64_t l = 0; l < in->nloci; ++l)
		for (uint64_t i = 0; in->fl_n && i < in->fl_n[l]; ++i, ++j) omap_put(&ix->fl, in->fl_kmers[j], l, 1);
	if (in->qc) { ix->qc = (uint8_t*)malloc(in->nloci); memcpy(ix->qc, in->qc, in->nloci); }
	if (in->g_n && in->g_kmers && in->g_masks) {                 /* readGraphKmers, h:549-575: kmerDB[idx][kmer] |= c */
		uint64_t ng = 0;
		for (uint64_t l = 0; l < in->nloci; ++l) ng += in->g_n[l];
		if (!omap_init(&ix->graph, ng)) return NULL;
		j = 0;
		for (uint64_t l = 0; l < in->nloci; ++l)
			for (uint64_t i = 0; i < in->g_n[l]; ++i, ++j) {
				uint32_t old = 0;
				omap_get(&ix->graph, in->g_kmers[j], l, &old);
				omap_put(&ix->graph, in->g_kmers[j], l, old | in->g_masks[j]);
			}
		ix->has_graph = 1;
	}
	return ix;
}

void dbo_index_free(dbo_index* ix) {
	if (!ix) return;
	omap_free(&ix->dbi); omap_free(&ix->tr); omap_free(&ix->fl); if (ix->has_graph) omap_free(&ix->graph); free(ix->vv); free(ix->qc); free(ix);
}

/* ------------------------------------------------------------------------------------------ */
/* A2: k-mer primitives                                                                        */
/* ------------------------------------------------------------------------------------------ */
/* getNuRC, h:165-178, restated base by base: reverse the k 2-bit codes and complement (3-c). */
uint64_t dbo_revcomp(uint64_t x, uint32_t k) {
	uint64_t rc = 0;
	for (uint32_t i = 0; i < k; ++i) { rc = (rc << 2) | (3 - (x & 3)); x >>= 2; }
	return rc;
}

/* valid characters are exactly upper-case ACGT: h:52,144,297 */
static int base_code(uint8_t c) { return c == 'A' ? 0 : c == 'C' ? 1 : c == 'G' ? 2 : c == 'T' ? 3 : -1; }

/*
 * read2kmers_edges, h:274-311.  kmers[i] = canonical k-mer of window i or -1; edges[i] = canonical
 * (k+1)-mer joining windows i and i+1 when both are valid, else -1.  If no window is valid the
 * reference leaves both vectors EMPTY (h:282) -> return 0.
 */
uint32_t dbo_read2kmers_edges(const uint8_t* seq, uint32_t len, uint32_t k, uint64_t* kmers, uint64_t* edges) {
	if (len < k) return 0;
	uint32_t nk = len - k + 1, any = 0;
	uint64_t prev_f = NAN64, prev_r = NAN64;
	for (uint32_t i = 0; i < nk; ++i) {
		uint64_t f = 0; int ok = 1;
		for (uint32_t j = 0; j < k; ++j) { int c = base_code(seq[i+j]); if (c < 0) { ok = 0; break; } f = (f << 2) | (uint64_t)c; }
		if (!ok) { kmers[i] = NAN64; if (edges && i) edges[i-1] = NAN64; prev_f = prev_r = NAN64; continue; }
		uint64_t r = dbo_revcomp(f, k);
		kmers[i] = f < r ? f : r;                                   /* h:288 */
		any = 1;
		if (edges && i) {
			if (prev_f != NAN64) {                                   /* h:290-295 */
				uint64_t e = (prev_f << 2) + (f & 3), re = (r << 2) + (prev_r & 3);
				edges[i-1] = e < re ? e : re;
			} else edges[i-1] = NAN64;
		}
		prev_f = f; prev_r = r;
	}
	return any ? nk : 0;
}

/* ------------------------------------------------------------------------------------------ */
/* libstdc++ std::sort (bits/stl_algo.h:1942 __sort, GCC 13.3, __GLIBCXX__ 20240904) on an     */
/* index array with comparator key[a] < key[b] -- getSortedIndex, cpp:247-255.                  */
/* ------------------------------------------------------------------------------------------ */
#define LT(a, b) (key[(a)] < key[(b)])
static void swp(uint64_t* a, uint64_t* b) { uint64_t t = *a; *a = *b; *b = t; }

static void unguarded_linear_insert(uint64_t* last, const uint64_t* key) {      /* stl_algo.h:1789 */
	uint64_t val = *last; uint64_t* next = last - 1;
	while (LT(val, *next)) { *last = *next; last = next; --next; }
	*last = val;
}
static void insertion_sort(uint64_t* first, uint64_t* last, const uint64_t* key) { /* :1812 */
	if (first == last) return;
	for (uint64_t* i = first + 1; i != last; ++i) {
		if (LT(*i, *first)) { uint64_t val = *i; memmove(first + 1, first, (size_t)(i - first) * 8); *first = val; }
		else unguarded_linear_insert(i, key);
	}
}
static void push_heap_(uint64_t* first, long hole, long top, uint64_t value, const uint64_t* key) { /* stl_heap.h:135 */
	long parent = (hole - 1) / 2;
	while (hole > top && LT(first[parent], value)) { first[hole] = first[parent]; hole = parent; parent = (hole - 1) / 2; }
	first[hole] = value;
}
static void adjust_heap(uint64_t* first, long hole, long len, uint64_t value, const uint64_t* key) { /* stl_heap.h:224 */
	const long top = hole; long child = hole;
	while (child < (len - 1) / 2) {
		child = 2 * (child + 1);
		if (LT(first[child], first[child - 1])) child--;
		first[hole] = first[child]; hole = child;
	}
	if ((len & 1) == 0 && child == (len - 2) / 2) { child = 2 * (child + 1); first[hole] = first[child - 1]; hole = child - 1; }
	push_heap_(first, hole, top, value, key);
}
static void heap_sort(uint64_t* first, uint64_t* last, const uint64_t* key) {   /* __partial_sort(first,last,last) :1905 */
	long len = last - first;
	if (len >= 2) { long parent = (len - 2) / 2; for (;;) { adjust_heap(first, parent, len, first[parent], key); if (parent == 0) break; parent--; } } /* __make_heap */
	while (last - first > 1) { --last; uint64_t value = *last; *last = *first; adjust_heap(first, 0, last - first, value, key); }        /* __sort_heap */
}
static void move_median_to_first(uint64_t* result, uint64_t* a, uint64_t* b, uint64_t* c, const uint64_t* key) { /* :85 */
	if (LT(*a, *b)) { if (LT(*b, *c)) swp(result, b); else if (LT(*a, *c)) swp(result, c); else swp(result, a); }
	else if (LT(*a, *c)) swp(result, a);
	else if (LT(*b, *c)) swp(result, c);
	else swp(result, b);
}
static uint64_t* unguarded_partition(uint64_t* first, uint64_t* last, uint64_t* pivot, const uint64_t* key) { /* :1871 */
	for (;;) {
		while (LT(*first, *pivot)) ++first;
		--last;
		while (LT(*pivot, *last)) --last;
		if (!(first < last)) return first;
		swp(first, last);
		++first;
	}
}
static void introsort_loop(uint64_t* first, uint64_t* last, long depth, const uint64_t* key) { /* :1918 */
	while (last - first > 16) {
		if (depth == 0) { heap_sort(first, last, key); return; }
		--depth;
		uint64_t* mid = first + (last - first) / 2;
		move_median_to_first(first, first + 1, mid, last - 1, key);
		uint64_t* cut = unguarded_partition(first + 1, last, first, key);
		introsort_loop(cut, last, depth, key);
		last = cut;
	}
}
void dbo_introsort_by_key(uint64_t* idx, uint32_t n, const uint64_t* key) {
	if (n == 0) return;
	long lg = 0; for (uint32_t m = n; m > 1; m >>= 1) ++lg;                 /* std::__lg */
	introsort_loop(idx, idx + n, 2 * lg, key);
	if (n > 16) { insertion_sort(idx, idx + 16, key); for (uint64_t* i = idx + 16; i != idx + n; ++i) unguarded_linear_insert(i, key); } /* :1854 */
	else insertion_sort(idx, idx + n, key);
}
#undef LT

/* ------------------------------------------------------------------------------------------ */
/* per-pair pipeline                                                                            */
/* ------------------------------------------------------------------------------------------ */
typedef struct { uint64_t key; uint32_t val; } hit_t;          /* one kmerDBi iterator (cpp:207) */
typedef struct { uint64_t idx, fc, rc; } asgn_t;                /* cpp:146-149 */
typedef struct {                                                /* km_asgn_read_t, cpp:93-108 */
	int kf, hf, af, rm;
	int si, ei, nt, bs, ti, si_, ei_;
	int nas; int* as;
} mate_t;

/* subfilter, cpp:172-188.  Returns 1 = reject. */
static int subfilter(const dbt_params* p, const dbo_index* ix, const uint64_t* k1, uint64_t L1, const uint64_t* k2, uint64_t L2, uint64_t* nhash) {
	uint64_t N = p->n_filter, M = p->nm_filter;
	uint64_t S1 = L1 / (N-1), S2 = L2 / (N-1), h1 = 0, h2 = 0;
	for (uint64_t i = 0; i < N; ++i, ++*nhash) {
		uint64_t i1 = (i != N-1 ? i*S1 : L1-1);
		h1 += (uint64_t)omap_get(&ix->dbi, k1[i1], 0, NULL);
		if (h1 >= M) break;
	}
	if (h1 < M) return 1;
	for (uint64_t i = 0; i < N; ++i, ++*nhash) {
		uint64_t i2 = (i != N-1 ? i*S2 : L2-1);
		h2 += (uint64_t)omap_get(&ix->dbi, k2[i2], 0, NULL);
		if (h2 >= M) break;
	}
	return h2 < M;
}

/* kfilter for one mate, cpp:203-213 */
static void kfilter_mate(const dbo_index* ix, const uint64_t* ks, uint64_t nk, uint64_t cth, hit_t* its, uint64_t* nits, uint64_t* nhash, int* kf, int* rm) {
	uint64_t ns = 0, si = 0; const uint64_t MAX_NS = nk - cth;
	*nits = 0;
	for (; si < nk; ++si) {
		uint32_t v;
		++*nhash;
		if (!omap_get(&ix->dbi, ks[si], 0, &v)) { ++ns; if (ns > MAX_NS) { *nits = 0; break; } }
		else { its[*nits].key = ks[si]; its[*nits].val = v; ++*nits; }
	}
	*kf = (si != nk);
	*rm |= *kf;
}

/* updatetop2, cpp:331-347 */
static void updatetop2(uint64_t cf, uint32_t ind, uint64_t cr, asgn_t* top, asgn_t* second) {
	if (cf + cr > top->fc + top->rc) {
		if (top->idx != ind) { *second = *top; top->idx = ind; }
		top->fc = cf; top->rc = cr;
	} else if (cf + cr > second->fc + second->rc) {
		if (second->idx != ind) second->idx = ind;
		second->fc = cf; second->rc = cr;
	}
}
static int get_acm1(const asgn_t* top, uint64_t rem, uint64_t cth) { /* cpp:354-357 */
	return (top->fc < cth && cth - top->fc <= rem) || (top->rc < cth && cth - top->rc <= rem);
}
static int get_acm2(const asgn_t* top, const asgn_t* second, uint64_t rem) { /* cpp:359-362 */
	return (top->fc + top->rc - second->fc - second->rc) < rem;
}

/*
 * countHit = fillstats + find_matching_locus + threshold test, cpp:257-453.
 * its1/its2: hits of mate 1 / mate 2 (consumed).  hits1/hits2: zeroed scratch of nloci+1 entries.
 */
static uint64_t countHit(const dbo_index* ix, hit_t* its1, uint64_t n1, const hit_t* its2, uint64_t n2,
                         uint32_t* hits1, uint32_t* hits2, uint64_t cth, uint64_t* tri0, uint64_t* fc, uint64_t* rc,
                         int* hf1, int* hf2, int* rm1, int* rm2) {
	const uint32_t* vv = ix->vv;
	uint64_t n = n1 + n2;
	/* countDupRemove, cpp:257-296 */
	hit_t* its = (hit_t*)malloc(n * sizeof(hit_t));
	uint8_t* orient = (uint8_t*)malloc(n);
	uint64_t* ord = (uint64_t*)malloc(n * 8);
	uint64_t* keys = (uint64_t*)malloc(n * 8);
	for (uint64_t i = 0; i < n1; ++i) { its[i] = its1[i]; orient[i] = 0; }
	for (uint64_t i = 0; i < n2; ++i) { its[n1+i] = its2[i]; orient[n1+i] = 1; }
	for (uint64_t i = 0; i < n; ++i) { ord[i] = i; keys[i] = its[i].key; }
	dbo_introsort_by_key(ord, (uint32_t)n, keys);             /* std::sort by ->first, cpp:254 */
	hit_t* u = (hit_t*)malloc(n * sizeof(hit_t));               /* unique entries */
	uint8_t* df = (uint8_t*)malloc(n); uint8_t* dr = (uint8_t*)malloc(n);   /* PE_KMC: uint8 wrap-around is the reference's (cpp:42) */
	uint64_t U = 0;
	{
		uint8_t f = 0, r = 0; hit_t last = its[ord[0]];
		u[0] = last; U = 1;
		if (orient[ord[0]]) ++r; else ++f;
		for (uint64_t i = 1; i < n; ++i) {
			hit_t cur = its[ord[i]];
			if (cur.key != last.key) { df[U-1] = f; dr[U-1] = r; f = r = 0; u[U] = cur; last = cur; ++U; }
			if (orient[ord[i]]) ++r; else ++f;
		}
		df[U-1] = f; dr[U-1] = r;
	}
	/* fillstats, cpp:308-329: sort by number of mapped loci (unstable std::sort) */
	uint64_t* nml = (uint64_t*)malloc(U * 8);
	uint64_t* perm = (uint64_t*)malloc(U * 8);
	for (uint64_t i = 0; i < U; ++i) { uint32_t vi = u[i].val; nml[i] = (vi % 2 ? vv[vi>>1] : 1); perm[i] = i; }
	dbo_introsort_by_key(perm, (uint32_t)U, nml);
	hit_t* su = (hit_t*)malloc(U * sizeof(hit_t)); uint8_t* sf = (uint8_t*)malloc(U); uint8_t* sr = (uint8_t*)malloc(U);
	for (uint64_t i = 0; i < U; ++i) { su[i] = u[perm[i]]; sf[i] = df[perm[i]]; sr[i] = dr[perm[i]]; }
	/* countRemain, cpp:298-306 (std::accumulate with an int accumulator; last entry stays 0) */
	uint64_t* remain = (uint64_t*)calloc(U, 8);
	{
		int acc = 0;
		for (uint64_t i = 0; i < U; ++i) acc = (int)((uint64_t)acc + sf[i] + sr[i]);
		uint64_t dupsum = (uint64_t)acc;
		remain[0] = dupsum - sf[0] - sr[0];
		for (uint64_t i = 1; i + 1 < U; ++i) remain[i] = remain[i-1] - sf[i] - sr[i];
	}
	/* find_matching_locus, cpp:364-422 */
	asgn_t top = { NAN32, 0, 0 }, second = { NAN32, 0, 0 };
	uint32_t* touched = (uint32_t*)malloc(8); uint64_t ntouched = 0, captouched = 2;
#define TOUCH(l) do { if (ntouched == captouched) { captouched *= 2; touched = (uint32_t*)realloc(touched, captouched * 4); } touched[ntouched++] = (l); } while (0)
	for (uint64_t i = 0; i < U; ++i) {
		uint32_t vi = su[i].val;
		if (vi % 2) {
			uint64_t j0 = (vi>>1) + 1, j1 = j0 + vv[vi>>1];
			for (; j0 < j1; ++j0) {
				uint32_t locus = vv[j0];
				hits1[locus] += sf[i]; hits2[locus] += sr[i]; TOUCH(locus);
				updatetop2(hits1[locus], locus, hits2[locus], &top, &second);
			}
		} else {
			uint32_t locus = vi >> 1;
			hits1[locus] += sf[i]; hits2[locus] += sr[i]; TOUCH(locus);
			updatetop2(hits1[locus], locus, hits2[locus], &top, &second);
		}
		if (!get_acm2(&top, &second, remain[i])) {
			uint64_t j = i;
			while (get_acm1(&top, remain[j], cth)) {
				if (++j >= U) break;
				uint32_t vj = su[j].val;
				if (vj % 2) {
					uint64_t j0 = (vj>>1) + 1, j1 = j0 + vv[vj>>1];
					for (; j0 < j1; ++j0) if (vv[j0] == top.idx) { top.fc += sf[j]; top.rc += sr[j]; break; }
				} else if ((vj >> 1) == top.idx) { top.fc += sf[j]; top.rc += sr[j]; }
			}
			break;
		}
	}
	for (uint64_t i = 0; i < ntouched; ++i) { hits1[touched[i]] = 0; hits2[touched[i]] = 0; }
#undef TOUCH
	/* countHit tail, cpp:436-452 */
	*tri0 = top.idx; *fc = top.fc; *rc = top.rc;
	uint64_t tri;
	int test1 = (top.fc >= cth && top.rc >= cth), test2 = (top.fc + top.rc) >= 2*cth;
	if ((test1 || test2) && top.idx != NAN32) tri = top.idx;
	else { tri = ix->nloci; *hf1 = 1 & !*rm1; *hf2 = 1 & !*rm2; *rm1 = 1; *rm2 = 1; }
	free(its); free(orient); free(ord); free(keys); free(u); free(df); free(dr); free(nml); free(perm);
	free(su); free(sf); free(sr); free(remain); free(touched);
	return tri;
}

/* assignTRkmc, cpp:1450-1556.  its[i] = count-array position of TR k-mer i or -1. */
static void assignTRkmc(const dbt_params* p, const dbo_index* ix, const uint64_t* kmers, int nk, uint64_t locus,
                        int64_t* its, mate_t* r, int okam) {
	if (!okam && r->rm) return;
	uint8_t ntr = 0;
	int s = 0, s_ = 0, s__ = 0;
	int ti2 = -1, si1 = -1, ei1 = -1, si2 = -1, ei2 = -1;
	r->nas = nk;
	for (int i = 0; i < nk; ++i) {
		uint32_t pos; uint64_t km = kmers[i];
		r->as[i] = 0;
		its[i] = omap_get(&ix->tr, km, locus, &pos) ? (int64_t)pos : -1;
		if (omap_get(&ix->fl, km, locus, NULL)) r->as[i] = 1;
		else if (its[i] >= 0) { r->as[i] = 2; ++ntr; }
	}
	if (r->rm) { r->nt = -1; r->bs = -1; r->ti = -1; return; }
	for (int i = 0; i < nk; ++i) {
		s = r->as[i];
		if (s && s__) {
			if (s != s__) {
				++r->nt;
				if ((uint64_t)r->nt > p->max_nt) { r->af = 1; r->rm = 1; return; }
				if (r->nt == 1) { r->ti = i; if (s_) { si1 = -1; ei1 = -1; } }
				else if (r->nt == 2) {
					if (r->bs == 2) { r->af = 1; r->rm = 1; return; }
					ti2 = i; if (s_) { si2 = -1; ei2 = -1; }
				}
			}
		}
		if (!r->bs) { if (s) r->bs = s; }
		if (!s) {
			if (r->nt == 0) { if (!s_) ++ei1; else { si1 = i; ei1 = i+1; } }
			if (r->nt == 1) { if (!s_) ++ei2; else { si2 = i; ei2 = i+1; } }
		}
		s_ = s;
		if (s) s__ = s;
	}
	int ti1 = r->ti;
	if (r->nt == 0) {
		if (r->bs != 2) { r->af = 1; r->rm = 1; return; }
		r->si = 0; r->ei = nk; r->si_ = 0; r->ei_ = nk;
	} else if (r->nt == 1) {
		if (r->bs == 1) { r->si = si1 >= 0 ? (si1+ei1)/2 : ti1; r->ei = nk; r->si_ = si1 >= 0 ? ei1 : ti1; r->ei_ = nk; }
		else            { r->si = 0; r->ei = si1 >= 0 ? (si1+ei1)/2 : ti1; r->si_ = 0; r->ei_ = si1 >= 0 ? si1 : ti1; }
	} else {
		if (ntr < p->nm_tr) { r->af = 1; r->rm = 1; return; }
		r->si = (si1 >= 0 ? (si1+ei1)/2 : ti1);
		r->ei = (si2 >= 0 ? (si2+ei2)/2 : ti2);
		r->si_ = ei1 >= 0 ? ei1 : ti1;
		r->ei_ = si2 >= 0 ? si2 : ti2;
	}
}

static void mate_init(mate_t* r, int* asbuf) {
	r->kf = r->hf = r->af = r->rm = 0;
	r->si = r->ei = r->si_ = r->ei_ = r->ti = -1; r->nt = 0; r->bs = 0; r->nas = 0; r->as = asbuf;
}

static void write_record(uint8_t* rec, uint32_t stride, uint32_t as_stride, uint64_t pair_index, int status,
                         uint64_t dst, uint64_t dst0, uint64_t fc, uint64_t rc, uint64_t nh0, uint64_t nh1, const mate_t* m1, const mate_t* m2) {
	memset(rec, 0, stride);
	dbt_pair_record* h = (dbt_pair_record*)rec;
	h->pair_index = pair_index; h->dst = (uint32_t)dst; h->dst0 = (uint32_t)dst0; h->fc = (uint32_t)fc; h->rc = (uint32_t)rc;
	h->nhash0 = (uint16_t)nh0; h->nhash1 = (uint16_t)nh1; h->status = (uint8_t)status;
	const mate_t* ms[2] = { m1, m2 };
	for (int m = 0; m < 2; ++m) {
		const mate_t* r = ms[m]; dbt_mate_record* o = &h->mate[m];
		o->si = (int16_t)r->si; o->ei = (int16_t)r->ei; o->si_ = (int16_t)r->si_; o->ei_ = (int16_t)r->ei_; o->ti = (int16_t)r->ti;
		o->nt = (int8_t)r->nt; o->bs = (int8_t)r->bs;
		o->flags = (uint8_t)((r->kf ? DBT_MF_KF : 0) | (r->hf ? DBT_MF_HF : 0) | (r->af ? DBT_MF_AF : 0) | (r->rm ? DBT_MF_RM : 0));
		o->nk = (uint16_t)r->nas;
		uint8_t* as = rec + sizeof(dbt_pair_record) + (size_t)m * as_stride;
		for (int i = 0; i < r->nas; ++i) as[i >> 2] |= (uint8_t)(r->as[i] << (2 * (i & 3)));
	}
}

/* sam_t.r1 / r2 (cpp:53-75) of a pair that entered threading -> the two dbt_aln_mate blobs behind the header */
static void write_aln_blobs(uint8_t* rec, uint32_t stride, const dbt_params* p, const dbo_thread_out* th) {
	const uint32_t es_cap = DBT_ES_CAP_FOR(p->max_read_len, p->max_ncorrection);
	const uint32_t aln_stride = (stride - (uint32_t)sizeof(dbt_pair_record)) / 2;
	for (int m = 0; m < 2; ++m) {
		uint8_t* b = rec + sizeof(dbt_pair_record) + (size_t)m * aln_stride;
		memset(b, 0, aln_stride);
		dbt_aln_mate* h = (dbt_aln_mate*)b;
		h->alned = (uint8_t)th[m].rc; h->err = (uint8_t)th[m].err;
		h->nes = (uint16_t)th[m].nes; h->ntr = (uint16_t)th[m].ntr; h->nkm = (uint16_t)th[m].nkm;
		uint64_t hash = 0xcbf29ce484222325ull;
		for (uint32_t i = 0; i < th[m].nkm; ++i)
			for (int q = 0; q < 8; ++q) { hash ^= (th[m].kmers[i] >> (8*q)) & 0xFF; hash *= 0x100000001b3ull; }
		h->kmers_hash = hash;
		uint8_t* es = b + sizeof(dbt_aln_mate); uint8_t* tr = es + es_cap;
		for (uint32_t i = 0; i < th[m].nes && i < es_cap; ++i) {
			uint8_t t = th[m].es_t[i], g = th[m].es_g[i];
			uint8_t tc = t == '*' ? 0 : t == '=' ? 1 : t == 'X' ? 2 : t == 'D' ? 3 : 4;
			uint8_t gc = g == 0 ? 0 : g == 'A' ? 1 : g == 'C' ? 2 : g == 'G' ? 3 : g == 'T' ? 4 : 5;
			es[i] = (uint8_t)(tc | (gc << 3));
		}
		for (uint32_t i = 0; i < th[m].ntr && sizeof(dbt_aln_mate) + es_cap + i < aln_stride; ++i)
			tr[i] = th[m].tr[i] == '*' ? 0 : th[m].tr[i] == '.' ? 1 : 2;
	}
}

/* The pair loop of CountWords, cpp:2002-2176 (asgn mode: threading off, no bait/bubbles/sim/extract). */
void dbo_run(const dbo_index* ix, const dbt_params* p, uint64_t first_pair_index, uint64_t npairs,
             const uint8_t* bases, const uint64_t* offsets, uint8_t* records, uint32_t stride, uint32_t as_stride,
             uint64_t* trkmc, uint64_t* kmc, uint32_t* nmapread, uint64_t* counters) {
	const uint32_t k = p->k; const uint64_t cth = p->cth, nloci = ix->nloci;
	const int okam = (p->flags & DBT_F_KAM) != 0, qc = (p->flags & DBT_F_QC) != 0 && ix->qc;
	uint32_t maxlen = 1;
	for (uint64_t i = 0; i < 2*npairs; ++i) { uint64_t l = offsets[i+1] - offsets[i]; if (l > maxlen) maxlen = (uint32_t)l; }
	uint64_t* k1 = (uint64_t*)malloc(maxlen * 8); uint64_t* k2 = (uint64_t*)malloc(maxlen * 8);
	hit_t* its1 = (hit_t*)malloc(maxlen * sizeof(hit_t)); hit_t* its2 = (hit_t*)malloc(maxlen * sizeof(hit_t));
	int64_t* kits1 = (int64_t*)malloc(maxlen * 8); int64_t* kits2 = (int64_t*)malloc(maxlen * 8);
	int* as1 = (int*)malloc(maxlen * 4); int* as2 = (int*)malloc(maxlen * 4);
	uint32_t* hits1 = (uint32_t*)calloc(nloci + 1, 4); uint32_t* hits2 = (uint32_t*)calloc(nloci + 1, 4);
	dbo_thread_out* th = (dbo_thread_out*)calloc(2, sizeof(dbo_thread_out));

	for (uint64_t pi = 0; pi < npairs; ++pi) {
		const uint8_t* s1 = bases + offsets[2*pi]; uint32_t l1 = (uint32_t)(offsets[2*pi+1] - offsets[2*pi]);
		const uint8_t* s2 = bases + offsets[2*pi+1]; uint32_t l2 = (uint32_t)(offsets[2*pi+2] - offsets[2*pi+1]);
		mate_t r1, r2; mate_init(&r1, as1); mate_init(&r2, as2);
		uint64_t nhash0 = 0, nhash1 = 0, destLocus = nloci, destLocus0 = NAN32, fc = 0, rc = 0, n1 = 0, n2 = 0;
		int status = -1;
		counters[DBT_C_NREADS] += 2;

		uint32_t nk1 = dbo_read2kmers_edges(s1, l1, k, k1, NULL);
		uint32_t nk2 = dbo_read2kmers_edges(s2, l2, k, k2, NULL);
		if (!nk1 || !nk2) { ++counters[DBT_C_NSHORT]; status = DBT_ST_SHORT; }                     /* cpp:2037 */
		if (status < 0 && p->n_filter && p->nm_filter) {                                          /* cpp:2045 */
			if (subfilter(p, ix, k1, nk1, k2, nk2, &nhash0)) { counters[DBT_C_NSUBFILTERED] += 2; status = DBT_ST_SUBFILTERED; }
		}
		if (status < 0) {                                                                         /* kfilter, cpp:190-200 */
			++counters[DBT_C_NCANDIDATES];
			r1.kf = nk1 < cth; r2.kf = nk2 < cth; r1.rm |= r1.kf; r2.rm |= r2.kf;
			if (!(r1.rm && r2.rm)) {
				if (!r1.rm) kfilter_mate(ix, k1, nk1, cth, its1, &n1, &nhash1, &r1.kf, &r1.rm);
				if (!r2.rm) kfilter_mate(ix, k2, nk2, cth, its2, &n2, &nhash1, &r2.kf, &r2.rm);
			}
			counters[DBT_C_NKMERFILTERED] += (uint64_t)(r1.kf + r2.kf);
			if (r1.rm && r2.rm) status = DBT_ST_KFILTERED;
		}
		if (status < 0) {
			destLocus = countHit(ix, its1, n1, its2, n2, hits1, hits2, cth, &destLocus0, &fc, &rc, &r1.hf, &r2.hf, &r1.rm, &r2.rm);
			counters[DBT_C_NLOCUSASSIGNFILTERED] += (uint64_t)(r1.hf + r2.hf);
			if (destLocus == nloci) status = DBT_ST_UNASSIGNED;
		}
		if (status < 0 && qc && !ix->qc[destLocus]) { counters[DBT_C_NQCFILTERED] += (uint64_t)(2 - r1.rm - r2.rm); status = DBT_ST_QCFILTERED; }
		if (status < 0 && (p->flags & DBT_F_THREADING) && !ix->has_graph) {                         /* cpp:2070; 2072-2090: alned stays false */
			counters[DBT_C_NTHREADINGREADS] += 2; status = DBT_ST_NOT_THREADED;
		}
		int threaded = 0;
		if (status < 0 && (p->flags & DBT_F_THREADING)) {
			/* restored caller (SURVEY.md §8(c) decision D1): commented blocks cpp:2072-2088 and 2195-2209 */
			counters[DBT_C_NTHREADINGREADS] += 2;
			threaded = 1;
			dbo_thread_read(ix, destLocus, s1, l1, p, &th[0]);
			dbo_thread_read(ix, destLocus, s2, l2, p, &th[1]);
			if (th[0].rc || th[1].rc) {
				counters[DBT_C_NFEASIBLEREADS] += 2;
				for (int m = 0; m < 2; ++m)
					for (uint32_t i = 0; i < th[m].nkm; ++i) {
						uint64_t km = th[m].kmers[i]; uint32_t pos;
						if (km == NAN64) continue;
						uint64_t r = dbo_revcomp(km, k); if (r < km) km = r;
						if (omap_get(&ix->tr, km, destLocus, &pos)) { ++trkmc[pos]; ++counters[DBT_C_NINC]; }
					}
				status = DBT_ST_COUNTED;
			} else { destLocus = nloci; status = DBT_ST_THREAD_FAILED; }
		}
		if (status < 0) {
			counters[DBT_C_NTHREADINGREADS] += 2; counters[DBT_C_NFEASIBLEREADS] += 2;            /* cpp:2070, 2092 */
			if (okam || !r1.rm || !r2.rm) {
				assignTRkmc(p, ix, k1, (int)nk1, destLocus0, kits1, &r1, okam);
				assignTRkmc(p, ix, k2, (int)nk2, destLocus0, kits2, &r2, okam);
			}
			if (r1.rm && r2.rm) { destLocus = nloci; status = DBT_ST_ASGN_REMOVED; }
			else {
				int n = 2 - r1.rm - r2.rm;
				nmapread[destLocus] += (uint32_t)n; counters[DBT_C_NASGNREADS] += (uint64_t)n;
				kmc[destLocus] += (uint64_t)(int64_t)((r1.ei - r1.si) + (r2.ei - r2.si));          /* cpp:2152 */
				if (!r1.rm) for (int i = 0; i < r1.nas; ++i) if (r1.as[i] == 2) { ++trkmc[kits1[i]]; ++counters[DBT_C_NINC]; }
				if (!r2.rm) for (int i = 0; i < r2.nas; ++i) if (r2.as[i] == 2) { ++trkmc[kits2[i]]; ++counters[DBT_C_NINC]; }
				status = DBT_ST_COUNTED;
			}
		}
		counters[DBT_C_NHASH0] += nhash0; counters[DBT_C_NHASH1] += nhash1;
		if (records) {
			write_record(records + pi * (uint64_t)stride, stride, as_stride, first_pair_index + pi, status,
			             status >= DBT_ST_UNASSIGNED ? destLocus : nloci, status >= DBT_ST_UNASSIGNED ? destLocus0 : NAN32,
			             status >= DBT_ST_UNASSIGNED ? fc : 0, status >= DBT_ST_UNASSIGNED ? rc : 0, nhash0, nhash1, &r1, &r2);
			if (threaded) write_aln_blobs(records + pi * (uint64_t)stride, stride, p, th);
			++counters[DBT_C_NRECORDS];
		}
	}
	free(k1); free(k2); free(its1); free(its2); free(kits1); free(kits2); free(as1); free(as2); free(hits1); free(hits2); free(th);
}
