/*
 * TEST INFRASTRUCTURE ONLY -- see dbt_oracle.h.  CPU restatement of the reference's graph threading
 * (P1): isThreadFeasible and its helpers, src/aQueryFasta_thread.cpp:526-1260
 * ("cpp"), on fixed-capacity arrays.  The reference compiles these functions but never calls them at
 * HEAD (cpp:2072-2088); oracle/_ref/ref_harness calls the reference's own code per read and
 * tests/test_oracle_thread.py pins this restatement against it (return code, edited k-mer vector,
 * CIGAR `es`, annotation `tr`), so parity is pinned at function level.
 *
 * Where the reference has undefined behaviour (vector reads past the end, assert(false) on a node
 * missing from the graph) this restatement sets out->err and returns 0; the tests use closed graphs
 * (every successor is a node, as fa2kmers builds them, h:215-243) where none of these fire.
 */
#include "dbt_oracle_int.h"

#define TCAP DBO_TCAP
#define MSC 5ull                                   /* cpp:1120 */

typedef struct { uint8_t t, r, g; } edit_t;        /* cpp:46-51 */
typedef struct {                                   /* cigar_t, cpp:53-68 */
	int ni;
	int nes; edit_t es[TCAP];
	int ntr; char tr[TCAP];
} cigar_t;

typedef struct {                                   /* per-read context */
	const dbo_index* ix; uint64_t locus; uint32_t k; uint64_t rmask;
	uint64_t kmers[TCAP]; int nkm;
	cigar_t cg;
	int err;
} thr_t;

static const char ALPHABET[4] = { 'A', 'C', 'G', 'T' };
static int nuc_num(uint8_t c) { return c == 'A' ? 0 : c == 'C' ? 1 : c == 'G' ? 2 : c == 'T' ? 3 : 127; }  /* baseNumConversion, h:53-69 */
static uint8_t nuc_comp_char(uint8_t c) {          /* baseComplement on characters, h:71-87 */
	switch (c) { case 'A': return 'T'; case 'C': return 'G'; case 'G': return 'C'; case 'T': return 'A'; case 'I': return 'I'; case 'N': return 'N'; default: return 127; }
}

static int g_find(const thr_t* T, uint64_t node, uint8_t* mask) {
	uint32_t v;
	if (!omap_get(&T->ix->graph, node, T->locus, &v)) return 0;
	if (mask) *mask = (uint8_t)v;
	return 1;
}
static uint64_t to_ca(const thr_t* T, uint64_t km) { uint64_t r = dbo_revcomp(km, T->k); return km < r ? km : r; }   /* toCaKmer, h:180-183 */
static char tr_annot(const thr_t* T, uint64_t km) { return omap_get(&T->ix->tr, to_ca(T, km), T->locus, NULL) ? '=' : '.'; }

/* vector insert / erase on the fixed arrays */
static void km_insert(thr_t* T, int pos, uint64_t v) {
	if (T->nkm + 1 > TCAP || pos < 0 || pos > T->nkm) { T->err = 1; return; }
	memmove(T->kmers + pos + 1, T->kmers + pos, (size_t)(T->nkm - pos) * 8); T->kmers[pos] = v; ++T->nkm;
}
static void km_erase(thr_t* T, int a, int b) {
	if (a < 0 || b > T->nkm || a > b) { T->err = 1; return; }
	memmove(T->kmers + a, T->kmers + b, (size_t)(T->nkm - b) * 8); T->nkm -= (b - a);
}
static void tr_insert(thr_t* T, int pos, char v) {
	cigar_t* c = &T->cg;
	if (c->ntr + 1 > TCAP || pos < 0 || pos > c->ntr) { T->err = 1; return; }
	memmove(c->tr + pos + 1, c->tr + pos, (size_t)(c->ntr - pos)); c->tr[pos] = v; ++c->ntr;
}
static void tr_erase(thr_t* T, int a, int b) {
	cigar_t* c = &T->cg;
	if (a < 0 || b > c->ntr || a > b) { T->err = 1; return; }
	memmove(c->tr + a, c->tr + b, (size_t)(c->ntr - b)); c->ntr -= (b - a);
}
static void es_insert(thr_t* T, int pos, edit_t v) {
	cigar_t* c = &T->cg;
	if (c->nes + 1 > TCAP || pos < 0 || pos > c->nes) { T->err = 1; return; }
	memmove(c->es + pos + 1, c->es + pos, (size_t)(c->nes - pos) * sizeof(edit_t)); c->es[pos] = v; ++c->nes;
}
static void es_erase(thr_t* T, int pos) {
	cigar_t* c = &T->cg;
	if (pos < 0 || pos >= c->nes) { T->err = 1; return; }
	memmove(c->es + pos, c->es + pos + 1, (size_t)(c->nes - pos - 1) * sizeof(edit_t)); --c->nes;
}
/* bounds-checked element access: an out-of-range index is undefined behaviour in the reference */
static edit_t* ES(thr_t* T, int i) { static edit_t dummy; if (i < 0 || i >= T->cg.nes) { T->err = 1; dummy.t = '='; dummy.r = dummy.g = 0; return &dummy; } return &T->cg.es[i]; }
static char* TR(thr_t* T, int i) { static char dummy; if (i < 0 || i >= T->cg.ntr) { T->err = 1; dummy = '*'; return &dummy; } return &T->cg.tr[i]; }
static uint64_t KM(thr_t* T, int64_t i) { if (i < 0 || i >= T->nkm) { T->err = 1; return NAN64; } return T->kmers[i]; }

/* getOutNodes, cpp:526-540: successors of `node`; nnts is OR-accumulated.  Missing node: the reference aborts. */
static int out_nodes(thr_t* T, uint64_t node, uint64_t* nnds, int* nnts) {
	uint8_t bits = 0; int n = 0;
	if (!g_find(T, node, &bits)) { T->err = 1; return 0; }
	uint64_t nnd = (node & T->rmask) << 2;
	for (uint64_t i = 0; i < 4; ++i) {
		if (bits & 1) nnds[n++] = nnd + i;
		nnts[i] |= bits & 1;
		bits >>= 1;
	}
	return n;
}
/* getNextNucs, cpp:547-557: ASSIGNS nnts when the node exists, leaves it untouched otherwise */
static void next_nucs(const thr_t* T, uint64_t node, int* nnts) {
	uint8_t bits;
	if (g_find(T, node, &bits)) for (int i = 0; i < 4; ++i) { nnts[i] = bits & 1; bits >>= 1; }
}

/* thread_ext_t, cpp:596-625 */
typedef struct {
	int rv;
	uint64_t nem1[4], nem2[16], nemi[4], nemd[16], ned1[4], ned2[16], nei1, nei2;
	uint64_t mes, ms1, ms2, score, nrk, nm, nd, ni;
	int dt_km, dt_ki, dt_nti;
	int nedits; edit_t edits[2];
} txt_t;
static void txt_init(txt_t* x, uint64_t mes, int rv) { memset(x, 0, sizeof(*x)); x->ms1 = MSC; x->ms2 = 2 * MSC; x->mes = mes; x->rv = rv; }
static void set1(txt_t* x, uint64_t s, char t, char g) { x->score = s; x->nedits = 1; x->edits[0].t = (uint8_t)t; x->edits[0].r = 0; x->edits[0].g = (uint8_t)g; }
static void set2(txt_t* x, uint64_t s, char t0, char g0, char t1, char g1) {
	x->score = s; x->nedits = 2;
	x->edits[0].t = (uint8_t)t0; x->edits[0].r = 0; x->edits[0].g = (uint8_t)g0;
	x->edits[1].t = (uint8_t)t1; x->edits[1].r = 0; x->edits[1].g = (uint8_t)g1;
}
/* get_edit, cpp:628-647: first maximum in this scan order wins */
static int get_edit(txt_t* x) {
	for (int i = 0; i < 4; ++i) if (x->nem1[i] > x->score && x->nem1[i] >= x->ms1) set1(x, x->nem1[i], 'X', ALPHABET[i]);
	for (int i = 0; i < 4; ++i) if (x->ned1[i] > x->score && x->ned1[i] >= x->ms1) set1(x, x->ned1[i], 'D', ALPHABET[i]);
	if (x->nei1 > x->score && x->nei1 >= x->ms1) set1(x, x->nei1, 'I', 0);
	if (x->mes > 1) {
		for (int i = 0; i < 4; ++i) {
			for (int j = 0; j < 4; ++j) {
				uint64_t sm2 = x->nem2[i*4+j], smd = x->nemd[i*4+j], sd2 = x->ned2[i*4+j];
				if (sm2 > x->score && sm2 >= x->ms2) set2(x, sm2, 'X', ALPHABET[i], 'X', ALPHABET[j]);
				if (smd > x->score && smd >= x->ms2) set2(x, smd, 'X', ALPHABET[i], 'D', ALPHABET[j]);
				if (sd2 > x->score && sd2 >= x->ms2) set2(x, sd2, 'D', ALPHABET[i], 'D', ALPHABET[j]);
			}
			if (x->nemi[i] > x->score && x->nemi[i] >= x->ms2) set2(x, x->nemi[i], 'X', ALPHABET[i], 'I', 0);
		}
		if (x->nei2 > x->score && x->nei2 >= x->ms2) set2(x, x->nei2, 'I', 0, 'I', 0);
	}
	return x->score > 0;
}

/* shared extension walk of the eight edit families (cpp:931-940 and its seven siblings) */
static uint64_t ext_walk(const thr_t* T, uint64_t crkmer, int* nnts, const uint64_t* km, uint64_t ki, uint64_t jbeg, uint64_t jend, const int* good) {
	uint64_t n = 0;
	for (uint64_t j = jbeg; j < jend; ++j) {
		if (!good[j]) break;
		crkmer = ((crkmer & T->rmask) << 2) + km[ki+j] % 4;
		if (nnts[crkmer % 4]) { ++n; next_nucs(T, crkmer, nnts); }
		else break;
	}
	return n;
}
static uint64_t min64(uint64_t a, uint64_t b) { return a < b ? a : b; }

/*
 * errorCorrection_forward, cpp:898-1089.  km[0..nkmers) is the k-mer vector being corrected (the read's, or
 * the reverse-complemented prefix built by errorCorrection_backward), ki the first k-mer without a matching
 * node, nnds/nts0 the successors of the last threaded node.  Returns skip (1 = no edit found).
 */
static int ec_forward(thr_t* T, const uint64_t* nnds, int nn, const uint64_t* km, uint64_t nkmers, uint64_t ki, const int* nts0, txt_t* txt, uint64_t mes) {
	const uint64_t ksize = T->k;
	int nts1[4] = {0}, nts2[4] = {0};
	int mat[64] = {0};                                                   /* graph_triplet_t, cpp:865-875 */
	const uint64_t oldnt = km[ki] % 4;
	for (int a = 0; a < nn; ++a) {                                        /* cpp:907-920 */
		uint64_t node_i = nnds[a], nt0 = node_i % 4, n1[4];
		int c1 = out_nodes(T, node_i, n1, nts1);
		for (int b = 0; b < c1; ++b) {
			uint64_t nt1 = n1[b] % 4, n2[4];
			int c2 = out_nodes(T, n1[b], n2, nts2);
			for (int c = 0; c < c2; ++c) mat[nt0*16 + nt1*4 + n2[c] % 4] = 1;
		}
	}
	int good[40] = {0};                                                   /* bool good[ksize+2], cpp:922-923 */
	for (uint64_t i = 0; i < min64(ksize+2, nkmers-ki); ++i) good[i] = km[ki+i] != NAN64;
#define NNTS1(i, out) do { for (int j_ = 0; j_ < 4; ++j_) for (int k_ = 0; k_ < 4; ++k_) (out)[j_] |= mat[(i)*16 + j_*4 + k_]; } while (0)
#define NNTS2(i, j, out) do { for (int k_ = 0; k_ < 4; ++k_) (out)[k_] |= mat[(i)*16 + (j)*4 + k_]; } while (0)
	/* the reference evaluates nts?[kmers[ki+j] % 4] BEFORE good[j]; good[j] implies ki+j in range */
	/* one mismatch, cpp:925-942 */
	if (good[1] && nts1[km[ki+1] % 4]) {
		for (uint64_t nt0 = 0; nt0 < 4; ++nt0) {
			if (!nts0[nt0]) continue;
			int nnts[4] = {0}; NNTS1(nt0, nnts);
			txt->nem1[nt0] += ext_walk(T, km[ki] - oldnt + nt0, nnts, km, ki, 1, min64(ksize+1, nkmers-ki), good);
		}
	}
	/* two mismatches, cpp:944-967 (else-if of the above) */
	else if (good[2] && nts2[km[ki+2] % 4] && mes >= 2) {
		for (uint64_t nt0 = 0; nt0 < 4; ++nt0) {
			if (!nts0[nt0]) continue;
			uint64_t cr0 = km[ki] - oldnt + nt0;
			int nnt0[4] = {0}; NNTS1(nt0, nnt0);
			for (uint64_t nt1 = 0; nt1 < 4; ++nt1) {
				if (!nnt0[nt1]) continue;
				int nnt1[4] = {0}; NNTS2(nt0, nt1, nnt1);
				txt->nem2[nt0*4+nt1] += ext_walk(T, ((cr0 & T->rmask) << 2) + nt1, nnt1, km, ki, 2, min64(ksize+2, nkmers-ki), good);
			}
		}
	}
	/* 1 substitution + 1 insertion, cpp:969-986 */
	if (good[2] && nts1[km[ki+2] % 4] && mes >= 2) {
		for (uint64_t nt0 = 0; nt0 < 4; ++nt0) {
			if (!nts0[nt0]) continue;
			int nnt0[4] = {0}; NNTS1(nt0, nnt0);
			txt->nemi[nt0] += ext_walk(T, km[ki] - oldnt + nt0, nnt0, km, ki, 2, min64(ksize+2, nkmers-ki), good);
		}
	}
	/* 1 substitution + 1 deletion, cpp:988-1011 */
	if (good[1] && nts2[km[ki+1] % 4] && mes >= 2) {
		for (uint64_t nt0 = 0; nt0 < 4; ++nt0) {
			if (!nts0[nt0]) continue;
			uint64_t cr0 = km[ki] - oldnt + nt0;
			int nnt0[4] = {0}; NNTS1(nt0, nnt0);
			for (uint64_t nt1 = 0; nt1 < 4; ++nt1) {
				if (!nnt0[nt1]) continue;
				int nnt1[4] = {0}; NNTS2(nt0, nt1, nnt1);
				txt->nemd[nt0*4+nt1] += ext_walk(T, ((cr0 & T->rmask) << 2) + nt1, nnt1, km, ki, 1, min64(ksize+1, nkmers-ki), good);
			}
		}
	}
	/* 1 insertion, cpp:1013-1026 */
	if (good[1] && nts0[km[ki+1] % 4]) {
		int nnt0[4] = { nts0[0], nts0[1], nts0[2], nts0[3] };
		txt->nei1 += ext_walk(T, km[ki-1], nnt0, km, ki, 1, min64(ksize+1, nkmers-ki), good);
	}
	/* 1 deletion, cpp:1028-1045 */
	if (good[0] && nts1[km[ki] % 4]) {
		for (uint64_t nt0 = 0; nt0 < 4; ++nt0) {
			if (!nts0[nt0]) continue;
			int nnt0[4] = {0}; NNTS1(nt0, nnt0);
			txt->ned1[nt0] += ext_walk(T, km[ki] - oldnt + nt0, nnt0, km, ki, 0, min64(ksize, nkmers-ki), good);
		}
	}
	/* 2 insertions, cpp:1047-1060 */
	if (good[2] && nts0[km[ki+2] % 4] && mes >= 2) {
		int nnt0[4] = { nts0[0], nts0[1], nts0[2], nts0[3] };
		txt->nei2 += ext_walk(T, km[ki-1], nnt0, km, ki, 2, min64(ksize+2, nkmers-ki), good);
	}
	/* 2 deletions, cpp:1062-1085 */
	if (good[0] && nts2[km[ki] % 4] && mes >= 2) {
		for (uint64_t nt0 = 0; nt0 < 4; ++nt0) {
			if (!nts0[nt0]) continue;
			uint64_t cr0 = km[ki] - oldnt + nt0;
			int nnt0[4] = {0}; NNTS1(nt0, nnt0);
			for (uint64_t nt1 = 0; nt1 < 4; ++nt1) {
				if (!nnt0[nt1]) continue;
				int nnt1[4] = {0}; NNTS2(nt0, nt1, nnt1);
				txt->ned2[nt0*4+nt1] += ext_walk(T, ((cr0 & T->rmask) << 2) + nt1, nnt1, km, ki, 0, min64(ksize, nkmers-ki), good);
			}
		}
	}
#undef NNTS1
#undef NNTS2
	return !get_edit(txt);
}

/* errorCorrection_backward, cpp:1091-1106: the same scorer on the reverse-complemented prefix kmers[0..ki) */
static int ec_backward(thr_t* T, uint64_t node, uint64_t ki, txt_t* txt, uint64_t mes) {
	int nts0_rc[4] = {0};
	uint64_t nnds_rc[4], kmers_rc[TCAP];
	uint64_t node_rc = dbo_revcomp(node, T->k);
	int nn = out_nodes(T, node_rc, nnds_rc, nts0_rc);
	if (ki + 1 > TCAP) { T->err = 1; return 1; }
	kmers_rc[0] = node_rc;
	int64_t j; uint64_t k;
	for (j = (int64_t)ki - 1, k = 1; j >= 0; --j, ++k) kmers_rc[k] = T->kmers[j] != NAN64 ? dbo_revcomp(T->kmers[j], T->k) : NAN64;
	return ec_forward(T, nnds_rc, nn, kmers_rc, ki + 1, 1, nts0_rc, txt, mes);
}

/* thread_ext_t::edit_kmers_forward, cpp:828-862 */
static void edit_kmers_forward(thr_t* T, txt_t* x, uint64_t* ki_io, uint64_t* ncorrection) {
	cigar_t* cg = &T->cg;
	const uint64_t ksize = T->k;
	uint64_t ki = *ki_io;
	int n0 = T->nkm - (int)ki;
	int good[TCAP]; uint64_t nts[TCAP];
	for (int i = 0; i < n0; ++i) { good[i] = T->kmers[ki+i] != NAN64; nts[i] = T->kmers[ki+i] % 4; }
	for (int e = 0; e < x->nedits; ++e) {
		edit_t* ed = &x->edits[e];
		if      (ed->t == 'X') {                              T->kmers[ki] = ((KM(T, (int64_t)ki-1) & T->rmask) << 2) + (uint64_t)nuc_num(ed->g); ++ki; ++x->nm; }
		else if (ed->t == 'D') { km_insert(T, (int)ki, 0);    T->kmers[ki] = ((KM(T, (int64_t)ki-1) & T->rmask) << 2) + (uint64_t)nuc_num(ed->g); ++ki; ++x->nd; }
		else if (ed->t == 'I') { km_erase(T, (int)ki, (int)ki+1);                                                                          ++x->ni; }
	}
	x->dt_nti = (int)(x->nm + x->ni);
	x->dt_ki = (int)(x->nm + x->nd);
	x->dt_km = (int)x->nd - (int)x->ni;
	for (uint64_t i = ki; i < min64((uint64_t)T->nkm, ki + ksize); ++i) {
		if (x->dt_nti >= n0) { T->err = 1; break; }
		if (!good[x->dt_nti]) break;
		T->kmers[i] = ((T->kmers[i-1] & T->rmask) << 2) + nts[x->dt_nti++];
	}
	if (x->dt_km > 0) { for (int i = 0; i < x->dt_km; ++i) tr_insert(T, cg->ntr, '*'); }     /* cg.tr.resize(size + dt_km, '*') */
	else if (x->dt_km < 0) { tr_erase(T, cg->ntr + x->dt_km, cg->ntr); }
	for (uint64_t i = 0; i < x->nd; ++i) { edit_t d = { 'D', 0, '*' }; es_insert(T, cg->ni + (int)ksize - 1 + (int)x->nm, d); }
	int ki_ = (int)ki - x->dt_ki;
	for (int i = 0; i < x->dt_ki + (int)x->score; ++i) *TR(T, ki_+i) = tr_annot(T, KM(T, ki_+i));
	for (int i = 0; i < x->nedits; ++i, ++cg->ni) {
		edit_t* e0 = ES(T, cg->ni + (int)ksize - 1);
		e0->t = x->edits[i].t; e0->g = x->edits[i].g;
	}
	for (uint64_t i = 0; i < x->score; ++i, ++cg->ni) ES(T, cg->ni + (int)ksize - 1)->t = '=';
	--cg->ni;
	ki += (x->score - 1);
	*ncorrection += (uint64_t)x->nedits;
	*ki_io = ki;
}

static int is_edit(uint8_t t) { return t == 'X' || t == 'D' || t == 'I'; }

/* thread_ext_t::edit_kmers_backward, cpp:649-826 */
static void edit_kmers_backward(thr_t* T, txt_t* x, uint64_t* ki_io, uint64_t* ncorrection, uint64_t* nskip) {
	cigar_t* cg = &T->cg;
	const uint64_t ksize = T->k;
	uint64_t ki = *ki_io;
	const int ki0 = (int)ki;
	int good[TCAP]; uint64_t nts[TCAP];
	const uint64_t lmask = 3ULL << 2*(ksize-1), lbase = 1ULL << 2*(ksize-1);
	for (int i = 0; i < ki0; ++i) { good[i] = T->kmers[i] != NAN64; nts[i] = T->kmers[i] & lmask; }
	for (int e = 0; e < x->nedits; ++e) {
		uint8_t t = x->edits[e].t;
		if (t == 'X') ++x->nm; else if (t == 'D') ++x->nd; else if (t == 'I') ++x->ni;
	}
	x->dt_km = (int)x->nd - (int)x->ni;
	x->dt_nti = -(int)(x->nm + x->ni);
	cg->ni += (int)x->nd;
	if (x->dt_km > 0) { for (int i = 0; i < x->dt_km; ++i) { km_insert(T, (int)ki, 0); tr_insert(T, (int)ki, '*'); } }
	else if (x->dt_km < 0) { km_erase(T, (int)ki + x->dt_km, (int)ki); tr_erase(T, (int)ki + x->dt_km, (int)ki); }
	ki += (uint64_t)(int64_t)x->dt_km;
	/* corrected k-mers */
	int ki_ = (int)ki;
	for (int e = 0; e < x->nedits; ++e) {
		edit_t* ed = &x->edits[e];
		if (ed->t == 'X' || ed->t == 'D') {
			if (ki_ - 1 < 0 || ki_ >= T->nkm) { T->err = 1; break; }
			T->kmers[ki_-1] = (T->kmers[ki_] >> 2) + (uint64_t)(3 - nuc_num(ed->g)) * lbase;   /* baseComplement[baseNumConversion[e.g]] */
			--ki_;
		}
	}
	/* extended k-mers */
	for (int i = ki_; i > (0 > ki_-(int)ksize ? 0 : ki_-(int)ksize); --i) {
		if (i - 1 >= ki0) { T->err = 1; break; }
		if (!good[i-1]) break;
		T->kmers[i-1] = (T->kmers[i] >> 2) + nts[i-1];
	}
	int lb = (int)(ki - x->nm - x->nd - x->score);
	for (int i = (int)ki - 1; i >= lb; --i) {
		if (*TR(T, i) == '*') ++x->nrk;
		*TR(T, i) = tr_annot(T, KM(T, i));
	}
	x->nrk -= (x->nm + x->nd);
	*nskip -= x->nrk;
	*ncorrection += (uint64_t)x->nedits;
	/* CIGAR, cpp:696-823 */
	{
		int cni = 0;
		int nti_ = (int)ki - x->dt_km;
		for (int i = 0; i < nti_ + cni; ++i) if (ES(T, i)->t == 'I') ++cni;
		int nti = nti_ + cni - 1;
		int e0, e1;
		for (int i = 0; i < x->nedits; ++i, --nti) {
			edit_t ed1 = x->edits[i];
			if (ed1.t == 'D') { ++nti; edit_t d = { 'D', 0, '*' }; es_insert(T, nti, d); }
			edit_t* ed0 = ES(T, nti);
			if (ed0->t == 'D') {
				if (ed1.t == 'I') { es_erase(T, nti); --cg->ni; }
				else ed0->g = nuc_comp_char(ed1.g);
			} else {
				while (ES(T, nti)->t == 'I' && !T->err) --nti;
				ed0 = ES(T, nti);
				ed0->t = ed1.t;
				ed0->g = ed1.g ? nuc_comp_char(ed1.g) : 0;
			}
		}
		e0 = nti + 1;
		e1 = e0;
		for (uint64_t i = 0; i < x->score; ++i, --nti) {
			edit_t* e = ES(T, nti);
			if (e->t == '=') { }
			else if (e->t == '*') e->t = '=';
			else break;
		}
		{
			uint8_t t = ES(T, e1)->t;
			while (is_edit(t) && !T->err) { ++e1; t = ES(T, e1)->t; }
			t = ES(T, e0-1)->t;
			while (is_edit(t) && !T->err) { --e0; t = ES(T, e0-1)->t; }
		}
		/* merge edits if possible */
		uint8_t rnts[TCAP], gnts[TCAP]; int nets = 0, nr = 0, ng = 0;
		for (int i = e0; i < e1; ++i) {
			edit_t* e = ES(T, i);
			++nets;                                                           /* ets.push_back(e.t) */
			if (e->r) rnts[nr++] = e->r;
			if (e->g) gnts[ng++] = e->g;
		}
		if (nr == ng) {
			int no_edit = 1;
			for (int i = 0; i < nr; ++i) if (rnts[i] != gnts[i]) { no_edit = 0; break; }
			if (no_edit) {                                                            /* edits cancelled out, cpp:755-770 */
				int dt_es = 0;
				for (int i = e0; i < e1; ++i) {
					uint8_t t = ES(T, i+dt_es)->t;
					if (t == 'D') { es_erase(T, i+dt_es); --dt_es; }
					else { edit_t* e = ES(T, i+dt_es); e->t = '='; e->g = 0; }
				}
				cg->ni += dt_es;
				*ncorrection -= (uint64_t)(e1-e0);
				*nskip -= (uint64_t)(e1-e0);
			} else if (nets != nr) {                                                  /* D+I at one position -> X, cpp:772-798 */
				int dt_es = 0;
				int dt_es_ = nr - nets;
				int i = e0, j = 0, k = 0;
				for ( ; i < e1; ++i) {
					uint8_t t = ES(T, i+dt_es)->t;
					if (t == 'D' && dt_es != dt_es_) { es_erase(T, i+dt_es); --dt_es; }
					else {
						edit_t* e = ES(T, i+dt_es);
						if (k >= nr || j >= ng) { T->err = 1; break; }
						if (rnts[k] == gnts[k]) { e->t = '='; e->g = 0; }
						else { e->t = 'X'; e->g = gnts[j]; }
						++j; ++k;
					}
				}
				if (dt_es != dt_es_) T->err = 1;                                     /* assert, cpp:794 */
				cg->ni += dt_es;
				*ncorrection += (uint64_t)(int64_t)dt_es;
				*nskip += (uint64_t)(int64_t)dt_es;
			} else {                                                                  /* match / mismatch only, cpp:799-809 */
				for (int i = 0; i < nr; ++i) {
					if (rnts[i] == gnts[i]) { edit_t* e = ES(T, e0+i); e->t = '='; e->g = 0; --*ncorrection; --*nskip; }
				}
			}
		} else {                                                                      /* cpp:812-822 */
			for (int i = 0; i < nets; ++i) {
				edit_t* e = ES(T, e0+i);
				if (e->r == e->g) { e->t = '='; e->g = 0; --*ncorrection; --*nskip; }
			}
		}
	}
	*ki_io = ki;
}

/* find_anchor with CIGAR bookkeeping, cpp:878-888 */
static int find_anchor(thr_t* T, uint64_t* nskip, uint64_t* ki, uint64_t* node) {
	cigar_t* cg = &T->cg;
	while (!g_find(T, KM(T, (int64_t)*ki), NULL)) {
		++*nskip;
		++cg->ni;
		if (++*ki >= (uint64_t)T->nkm) return 0;
	}
	*node = T->kmers[*ki];
	*TR(T, (int)*ki) = tr_annot(T, *node);
	for (int i = cg->ni; i < cg->ni + (int)T->k; ++i) { edit_t* e = ES(T, i); if (e->t == '*') e->t = '='; }
	return 1;
}

/* annot_gap, cpp:1108-1111 */
static void annot_gap(thr_t* T, uint64_t gap, uint64_t ki, uint64_t* nskip) {
	for (uint64_t i = 0; i < gap && !T->err; ++i) *TR(T, (int)--ki) = '*';
	*nskip -= gap;
}

/* isThreadFeasible, cpp:1114-1260.  T->kmers / T->cg initialised by the caller (read2kmers keepN, cigar_t::init). */
static int is_thread_feasible(thr_t* T, uint64_t thread_cth, int correction, uint64_t maxncorrection) {
	cigar_t* cg = &T->cg;
	const uint64_t ksize = T->k;
	const uint64_t maxnskip = ((uint64_t)T->nkm >= thread_cth ? (uint64_t)T->nkm - thread_cth : 0);
	uint64_t ki = 0, nskip = 0, ncorrection = 0;
	uint64_t node = T->kmers[0];
	const uint64_t nkmers = (uint64_t)T->nkm;
	uint64_t mes;

	if (!find_anchor(T, &nskip, &ki, &node)) return 0;
	if (ki > 0 && correction && ncorrection < maxncorrection) {               /* cpp:1131-1141 */
		if (ki >= MSC+1) {
			mes = (ki >= 2*MSC + 2) ? 2 : 1;
			txt_t txtr; txt_init(&txtr, mes, 1);
			int skip = ec_backward(T, node, ki, &txtr, mes);
			if (!skip) edit_kmers_backward(T, &txtr, &ki, &ncorrection, &nskip);
		}
	}

#define NEXT continue
	for (ki = ki+1, cg->ni = cg->ni+1; ki < (uint64_t)T->nkm && !T->err; ++ki, ++cg->ni) {
		if (T->kmers[ki] == NAN64) {                                           /* "N" in read, cpp:1145-1151 */
			*TR(T, (int)ki) = '*';
			ES(T, cg->ni + (int)ksize - 1)->t = '*';
			++nskip;
			if (nskip > maxnskip) return 0;
			NEXT;
		}
		if (T->kmers[ki] == T->kmers[ki-1]) {                                  /* homopolymer run, cpp:1152-1158 */
			*TR(T, (int)ki) = '*';
			ES(T, cg->ni + (int)ksize - 1)->t = '*';
			++nskip;
			if (nskip > maxnskip) return 0;
			NEXT;
		}
		if (T->kmers[ki-1] == NAN64) {                                         /* after an N, cpp:1159-1165 */
			if (!find_anchor(T, &nskip, &ki, &node)) break;
			if (nskip > maxnskip) return 0;
			NEXT;
		}

		int skip = 1;
		int nts0[4] = {0};
		uint64_t nnds[4];
		int nn = out_nodes(T, node, nnds, nts0);
		for (int a = 0; a < nn; ++a) {
			if (T->kmers[ki] == nnds[a]) {                                       /* matching node, cpp:1172-1178 */
				node = nnds[a];
				skip = 0;
				*TR(T, (int)ki) = tr_annot(T, T->kmers[ki]);
				ES(T, cg->ni + (int)ksize - 1)->t = '=';
				break;
			}
		}
		if (!skip) NEXT;
		if (ki + MSC >= nkmers) {                                              /* not enough info, cpp:1182-1185 */
			nskip += (nkmers - ki);
			return (nskip <= maxnskip ? (ncorrection ? 2 : 1) : 0);
		}
		if (correction && ncorrection < maxncorrection) {
			mes = ((uint64_t)T->nkm - ki >= 2*MSC + 2) ? 2 : 1;
			txt_t txtf; txt_init(&txtf, mes, 0);
			skip = ec_forward(T, nnds, nn, T->kmers, (uint64_t)T->nkm, ki, nts0, &txtf, mes);
			if (!skip) {                                                         /* cpp:1192-1197 */
				nskip += (uint64_t)txtf.nedits;
				if (nskip > maxnskip) return 0;
				edit_kmers_forward(T, &txtf, &ki, &ncorrection);
				node = KM(T, (int64_t)ki);
			} else {                                                             /* cpp:1198-1248 */
				uint64_t gap;
				if (!find_anchor(T, &nskip, &ki, &node)) break;
				mes = 2;
				txt_t txtr; txt_init(&txtr, mes, 1);
				skip = ec_backward(T, node, ki, &txtr, mes);
				if (!skip) {
					edit_kmers_backward(T, &txtr, &ki, &ncorrection, &nskip);
					++ncorrection;
					gap = min64(ksize, ki - txtr.nm - txtr.nd) - txtr.score;
					uint64_t ki0 = ki, ki1 = ki;
					while (!skip && gap && !T->err) {
						ki0 = ki1;
						ki1 = ki0 - txtr.nm - txtr.nd - txtr.score;
						mes = (ki1 >= 2*MSC + 2) ? 2 : 1;
						if (ki1 < MSC+1) break;
						txt_init(&txtr, mes, 1);
						uint64_t node_ = KM(T, (int64_t)ki1);
						if (!g_find(T, node_, NULL)) { T->err = 1; break; }               /* assert(g.count(node_)), cpp:1222 */
						skip = ec_backward(T, node_, ki1, &txtr, mes);
						if (!skip) {
							edit_kmers_backward(T, &txtr, &ki1, &ncorrection, &nskip);
							ki += txtr.nd - txtr.ni;
							gap = min64(ksize, ki1 - txtr.nm - txtr.nd) - txtr.score;
						}
					}
					if (gap) annot_gap(T, gap, ki1, &nskip);
					if (nskip > maxnskip) return 0;
				}
				if (skip) {                                                        /* cpp:1234-1243 */
					if (!find_anchor(T, &nskip, &ki, &node)) break;
					if (nskip > maxnskip) return 0;
					NEXT;
				}
			}
		} else {                                                               /* cpp:1250-1256 */
			if (!find_anchor(T, &nskip, &ki, &node)) break;
			if (nskip > maxnskip) return 0;
			NEXT;
		}
	}
#undef NEXT
	if (T->err) return 0;
	return (nskip <= maxnskip && ncorrection <= maxncorrection ? (ncorrection ? 2 : 1) : 0);
}

/* read2kmers(.., canonical=false, keepN=true), h:246-271: strand-specific k-mers, -1 over invalid windows */
static int read2kmers_nonca(const uint8_t* seq, uint32_t len, uint32_t k, uint64_t* out) {
	if (len < k) return 0;
	int nk = (int)(len - k + 1), any = 0;
	for (int i = 0; i < nk; ++i) {
		uint64_t f = 0; int ok = 1;
		for (uint32_t j = 0; j < k; ++j) { int c = nuc_num(seq[i+j]); if (c > 3) { ok = 0; break; } f = (f << 2) | (uint64_t)c; }
		out[i] = ok ? f : NAN64;
		any |= ok;
	}
	return any ? nk : 0;
}

int dbo_thread_read(const dbo_index* ix, uint64_t locus, const uint8_t* seq, uint32_t len, const dbt_params* p, dbo_thread_out* out) {
	static thr_t Ts;                         /* single-threaded test code */
	thr_t* T = &Ts;
	memset(out, 0, sizeof(*out));
	if (!ix->has_graph || len > TCAP - 16) { out->err = 1; return 0; }
	T->ix = ix; T->locus = locus; T->k = p->k; T->rmask = (1ULL << 2*(p->k-1)) - 1; T->err = 0;
	/* cigar_t::init, cpp:62-67 */
	T->cg.ni = 0; T->cg.nes = (int)len;
	for (uint32_t i = 0; i < len; ++i) { T->cg.es[i].t = '*'; T->cg.es[i].r = seq[i]; T->cg.es[i].g = 0; }
	T->cg.ntr = len >= p->k ? (int)(len - p->k + 1) : 0;
	memset(T->cg.tr, '*', (size_t)T->cg.ntr);
	T->nkm = read2kmers_nonca(seq, len, p->k, T->kmers);
	int rc = 0;
	/* a read without a valid k-mer: the reference would read kmers[0] of an empty vector (cpp:1124); decision D1: not threadable */
	if (T->nkm) rc = is_thread_feasible(T, p->thread_cth, (p->flags & DBT_F_CORRECTION) != 0, p->max_ncorrection);
	out->rc = rc; out->err = T->err;
	out->nkm = (uint32_t)T->nkm; memcpy(out->kmers, T->kmers, (size_t)T->nkm * 8);
	out->nes = (uint32_t)T->cg.nes;
	for (int i = 0; i < T->cg.nes; ++i) { out->es_t[i] = T->cg.es[i].t; out->es_g[i] = T->cg.es[i].g; }
	out->ntr = (uint32_t)T->cg.ntr; memcpy(out->tr, T->cg.tr, (size_t)T->cg.ntr);
	return rc;
}
