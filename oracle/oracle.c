/* TEST INFRASTRUCTURE — see oracle.h. Plain-C restatement of the expANN query hot path.
 * Built by oracle/Makefile: gcc -O3 -mavx2 -mfma -ffp-contract=off (explicit fmaf only where the reference
 * has an explicit FMA intrinsic). Nothing under expann_b200/ may depend on this file. */
#define _GNU_SOURCE
#include "oracle.h"

#include <float.h>
#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ======================================================================================================
 * distance primitives
 * ====================================================================================================== */

/* _mm512_reduce_add_ps as GCC/clang expand it: halves folded 16->8->4->2->1 */
static inline float reduce16(const float* l) {
	float t8[8], t4[4], t2[2];
	for (int i = 0; i < 8; ++i) t8[i] = l[i] + l[i + 8];
	for (int i = 0; i < 4; ++i) t4[i] = t8[i] + t8[i + 4];
	for (int i = 0; i < 2; ++i) t2[i] = t4[i] + t4[i + 2];
	return t2[0] + t2[1];
}

/* src/distance.h:136-148 / :86-111 */
float orc_dist2_f32(const float* a, const float* b, size_t d) {
	float acc[16];
	for (int l = 0; l < 16; ++l) acc[l] = 0.0f;
	size_t i = 0;
	for (; i + 16 <= d; i += 16)
		for (int l = 0; l < 16; ++l) {
			float diff = a[i + l] - b[i + l];      /* _mm512_sub_ps   */
			acc[l] = fmaf(diff, diff, acc[l]);     /* _mm512_fmadd_ps */
		}
	float s = reduce16(acc);
	for (; i < d; ++i) { /* not reachable in the reference (size%16==0 required) */
		float diff = a[i] - b[i];
		s = fmaf(diff, diff, s);
	}
	return s;
}

/* src/distance.h:181-190 */
float orc_dot_f32(const float* a, const float* b, size_t d) {
	float acc[16];
	for (int l = 0; l < 16; ++l) acc[l] = 0.0f;
	size_t i = 0;
	for (; i + 16 <= d; i += 16)
		for (int l = 0; l < 16; ++l) acc[l] = fmaf(a[i + l], b[i + l], acc[l]);
	float s = reduce16(acc);
	for (; i < d; ++i) s = fmaf(a[i], b[i], s);
	return s;
}

/* src/antitopo_engine.h:38-61, query conversion :726-737 (uint32_t(q[..]) truncation) */
int32_t orc_dist2_u8(const float* q, const uint8_t* code, size_t d) {
	uint32_t s = 0;
	for (size_t i = 0; i < d; ++i) {
		uint32_t qa = (uint32_t)q[i];
		uint32_t diff = qa - (uint32_t)code[i]; /* _mm512_sub_epi32 (wrapping) */
		s += diff * diff;                       /* _mm512_mullo_epi32, _mm512_add_epi32 */
	}
	return (int32_t)s;
}

/* src/distance.h:29-53 as written */
int32_t orc_dist2_i8_asref(const int8_t* a, const int8_t* b, size_t d) {
	uint32_t s = 0;
	for (size_t i = 0; i < d; ++i) {
		uint8_t diff = (uint8_t)((uint8_t)a[i] - (uint8_t)b[i]); /* _mm512_sub_epi8 */
		uint32_t z = diff;                                       /* unpack with zero */
		s += z * z;                                              /* madd_epi16 */
	}
	return (int32_t)s;
}
int32_t orc_dist2_i8(const int8_t* a, const int8_t* b, size_t d) {
	int32_t s = 0;
	for (size_t i = 0; i < d; ++i) {
		int32_t diff = (int32_t)a[i] - (int32_t)b[i];
		s += diff * diff;
	}
	return s;
}
/* src/distance.h:14-27 as written */
int32_t orc_dist2_i16_asref(const int16_t* a, const int16_t* b, size_t d) {
	uint32_t s = 0;
	for (size_t i = 0; i < d; ++i) {
		int16_t diff = (int16_t)((uint16_t)a[i] - (uint16_t)b[i]);       /* sub_epi16 */
		int16_t sq = (int16_t)((uint32_t)((int32_t)diff * (int32_t)diff)); /* mullo_epi16 */
		s += (uint32_t)(int32_t)sq;                                      /* madd_epi16(1, sq) sign-extends */
	}
	return (int32_t)s;
}
int32_t orc_dist2_i16(const int16_t* a, const int16_t* b, size_t d) {
	int32_t s = 0;
	for (size_t i = 0; i < d; ++i) {
		int32_t diff = (int32_t)a[i] - (int32_t)b[i];
		s += diff * diff;
	}
	return s;
}

/* ======================================================================================================
 * quantizers
 * ====================================================================================================== */
void orc_quant_simple_u8(const float* rows, size_t n, size_t d, uint8_t* codes) {
	for (size_t i = 0; i < n * d; ++i) codes[i] = (uint8_t)rows[i]; /* src/quantizer.h:136-138 T(x) */
}
void orc_quant_ranged_q8_fit(const float* rows, size_t n, size_t d, float* scale, float* offset) {
	float max_val = FLT_MIN, min_val = FLT_MAX; /* src/quantizer.h:210-211 */
	for (size_t i = 0; i < n * d; ++i) {
		if (rows[i] > max_val) max_val = rows[i];
		if (rows[i] < min_val) min_val = rows[i];
	}
	float sf = (float)128 / (max_val - min_val); /* :219 q_range()=128 */
	*scale = sf;
	*offset = -sf * min_val - (float)0;          /* :220 q_min = 0 */
}
void orc_quant_ranged_q8_encode(const float* rows, size_t n, size_t d, float scale, float offset,
                                int8_t* codes) {
	for (size_t i = 0; i < n * d; ++i) {
		float r = roundf(scale * rows[i] + offset); /* :197-198 */
		int v;
		if (r < 0.0f) v = 127;       /* negative -> size_t wraps huge -> clamp hi (observed behaviour) */
		else if (r > 127.0f) v = 127;
		else v = (int)r;
		codes[i] = (int8_t)v;
	}
}

/* ======================================================================================================
 * std::priority_queue<std::pair<float,size_t>> restated (libstdc++ bits/stl_heap.h algorithms)
 * ====================================================================================================== */
typedef struct { float first; uint64_t second; } pr;
enum { CMP_PAIR_LESS = 0, /* std::less<pair>: max-heap on (dist,id) */
       CMP_FIRST_LESS = 1, /* worst_elem  a.first < b.first -> max-heap on dist */
       CMP_FIRST_GREATER = 2 /* best_elem a.first > b.first -> min-heap on dist */ };
typedef struct { pr* a; size_t n, cap; int cmp; } heap_t;

static inline int hcomp(int cmp, pr x, pr y) {
	switch (cmp) {
	case CMP_PAIR_LESS:
		return x.first < y.first || (!(y.first < x.first) && x.second < y.second);
	case CMP_FIRST_LESS:
		return x.first < y.first;
	default:
		return x.first > y.first;
	}
}
static void heap_init(heap_t* h, int cmp) { h->a = NULL; h->n = h->cap = 0; h->cmp = cmp; }
static void heap_free(heap_t* h) { free(h->a); h->a = NULL; h->n = h->cap = 0; }
static void heap_reserve(heap_t* h, size_t c) {
	if (c > h->cap) {
		size_t nc = h->cap ? h->cap * 2 : 16;
		if (nc < c) nc = c;
		h->a = (pr*)realloc(h->a, nc * sizeof(pr));
		h->cap = nc;
	}
}
static void push_heap_(pr* first, size_t hole, size_t top, pr value, int cmp) {
	size_t parent = (hole - 1) / 2;
	while (hole > top && hcomp(cmp, first[parent], value)) {
		first[hole] = first[parent];
		hole = parent;
		parent = (hole - 1) / 2;
	}
	first[hole] = value;
}
static void adjust_heap_(pr* first, size_t hole, size_t len, pr value, int cmp) {
	const size_t top = hole;
	size_t child = hole;
	while (len >= 1 && child < (len - 1) / 2) {
		child = 2 * (child + 1);
		if (hcomp(cmp, first[child], first[child - 1])) child--;
		first[hole] = first[child];
		hole = child;
	}
	if ((len & 1) == 0 && len >= 2 && child == (len - 2) / 2) {
		child = 2 * (child + 1);
		first[hole] = first[child - 1];
		hole = child - 1;
	}
	push_heap_(first, hole, top, value, cmp);
}
__attribute__((unused)) static void heap_make(heap_t* h) {
	size_t len = h->n;
	if (len < 2) return;
	size_t parent = (len - 2) / 2;
	for (;;) {
		pr v = h->a[parent];
		adjust_heap_(h->a, parent, len, v, h->cmp);
		if (parent == 0) return;
		parent--;
	}
}
static void heap_push(heap_t* h, float d, uint64_t id) {
	heap_reserve(h, h->n + 1);
	pr v = { d, id };
	h->a[h->n] = v;
	h->n++;
	push_heap_(h->a, h->n - 1, 0, v, h->cmp);
}
static void heap_pop(heap_t* h) {
	if (h->n > 1) {
		pr v = h->a[h->n - 1];
		h->a[h->n - 1] = h->a[0];
		adjust_heap_(h->a, 0, h->n - 1, v, h->cmp);
	}
	h->n--;
}
static inline pr heap_top(const heap_t* h) { return h->a[0]; }

/* ======================================================================================================
 * topk_t  (src/topk_t.h:9-67)
 * ====================================================================================================== */
struct orc_topk { heap_t h; size_t k; };
static int topk_known(const orc_topk* t, uint64_t v) { /* robin_hood set membership == "currently held" */
	for (size_t i = 0; i < t->h.n; ++i)
		if (t->h.a[i].second == v) return 1;
	return 0;
}
orc_topk* orc_topk_create(size_t k) {
	orc_topk* t = (orc_topk*)calloc(1, sizeof(orc_topk));
	heap_init(&t->h, CMP_PAIR_LESS);
	t->k = k;
	return t;
}
void orc_topk_destroy(orc_topk* t) { if (t) { heap_free(&t->h); free(t); } }
int orc_topk_consider(orc_topk* t, float d, uint64_t v) {
	int good = !topk_known(t, v) && (t->h.n < t->k || heap_top(&t->h).first > d);
	if (good) heap_push(&t->h, d, v);
	if (t->h.n > t->k) heap_pop(&t->h);
	return good;
}
void orc_topk_discard_until_size(orc_topk* t, size_t goal) { while (t->h.n > goal) heap_pop(&t->h); }
size_t orc_topk_size(const orc_topk* t) { return t->h.n; }
uint64_t orc_topk_worst(const orc_topk* t) { return heap_top(&t->h).second; }
float orc_topk_worst_val(const orc_topk* t) { return heap_top(&t->h).first; }
int orc_topk_at_capacity(const orc_topk* t) { return t->h.n == t->k; }
size_t orc_topk_to_combined(const orc_topk* t, float* d_out, uint64_t* id_out) {
	heap_t c;
	heap_init(&c, CMP_PAIR_LESS);
	heap_reserve(&c, t->h.n ? t->h.n : 1);
	memcpy(c.a, t->h.a, t->h.n * sizeof(pr));
	c.n = t->h.n;
	size_t n = c.n;
	for (size_t i = 0; i < n; ++i) { /* drain max-first, store reversed -> ascending */
		pr p = heap_top(&c);
		heap_pop(&c);
		if (d_out) d_out[n - 1 - i] = p.first;
		if (id_out) id_out[n - 1 - i] = p.second;
	}
	heap_free(&c);
	return n;
}

/* ======================================================================================================
 * brute force  (src/brute_force_engine.h:29-46)
 * ====================================================================================================== */
static void bf_one(const float* base, size_t n, size_t d, const float* qv, size_t k, int metric,
                   uint64_t* ids, float* dists, uint64_t* count, heap_t* h) {
	h->n = 0;
	for (size_t i = 0; i < n; ++i) {
		float key = metric == ORC_L2 ? orc_dist2_f32(qv, base + i * d, d) : -orc_dot_f32(qv, base + i * d, d);
		if (h->n < k || heap_top(h).first > key) heap_push(h, key, i); /* :33-35 */
		if (h->n > k) heap_pop(h);                                     /* :36-37 */
	}
	size_t c = h->n;
	if (count) *count = c;
	for (size_t j = k; j-- > c;) { ids[j] = UINT64_MAX; if (dists) dists[j] = INFINITY; }
	for (size_t j = c; j-- > 0;) { /* drain worst-first, fill from the back: nearest first (:39-45) */
		pr p = heap_top(h);
		heap_pop(h);
		ids[j] = p.second;
		if (dists) dists[j] = p.first;
	}
}

typedef struct {
	const float* base; size_t n, d; const float* q; size_t m, k; int metric;
	uint64_t* ids; float* dists; uint64_t* counts; size_t* next; pthread_mutex_t* mu;
} bf_job;
static size_t claim(size_t* next, pthread_mutex_t* mu, size_t chunk, size_t m) {
	pthread_mutex_lock(mu);
	size_t i = *next;
	*next = i + chunk;
	pthread_mutex_unlock(mu);
	return i < m ? i : m;
}
static void* bf_worker(void* arg) {
	bf_job* j = (bf_job*)arg;
	heap_t h;
	heap_init(&h, CMP_PAIR_LESS);
	heap_reserve(&h, j->k + 2);
	for (;;) {
		size_t i = claim(j->next, j->mu, 1, j->m);
		if (i >= j->m) break;
		bf_one(j->base, j->n, j->d, j->q + i * j->d, j->k, j->metric, j->ids + i * j->k,
		       j->dists ? j->dists + i * j->k : NULL, j->counts ? j->counts + i : NULL, &h);
	}
	heap_free(&h);
	return NULL;
}
static void run_threads(void* (*fn)(void*), void* arg, int threads) {
	if (threads < 1) threads = 1;
	pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)threads);
	for (int t = 1; t < threads; ++t) pthread_create(&th[t], NULL, fn, arg);
	fn(arg);
	for (int t = 1; t < threads; ++t) pthread_join(th[t], NULL);
	free(th);
}
void orc_bf_query(const float* base, size_t n, size_t d, const float* q, size_t m, size_t k, int metric,
                  uint64_t* ids_out, float* dists_out, uint64_t* counts_out, int threads) {
	size_t next = 0;
	pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
	bf_job j = { base, n, d, q, m, k, metric, ids_out, dists_out, counts_out, &next, &mu };
	run_threads(bf_worker, &j, threads);
}

/* ---- compressed exhaustive scan + fp32 rerank --------------------------------------------------------- */
static int32_t ip_u8(const float* q, const uint8_t* code, size_t d) {
	uint32_t s = 0;
	for (size_t i = 0; i < d; ++i) s += (uint32_t)q[i] * (uint32_t)code[i];
	return -(int32_t)s;
}
typedef struct { int64_t score; uint64_t id; } sc;
static int sc_less(sc a, sc b) { return a.score < b.score || (a.score == b.score && a.id < b.id); }
typedef struct {
	const float* base; const uint8_t* codes; size_t n, d; const float* q; size_t m, k, kprime; int metric;
	uint64_t* ids; float* dists; int32_t* scores; uint64_t* cand; size_t* next; pthread_mutex_t* mu;
} u8_job;
static int pr_cmp_qsort(const void* a, const void* b) {
	const pr* x = (const pr*)a; const pr* y = (const pr*)b;
	if (x->first < y->first) return -1;
	if (x->first > y->first) return 1;
	return x->second < y->second ? -1 : (x->second > y->second ? 1 : 0);
}
static void* u8_worker(void* arg) {
	u8_job* j = (u8_job*)arg;
	const size_t kp = j->kprime;
	sc* best = (sc*)malloc(sizeof(sc) * (kp + 1)); /* sorted ascending, insertion */
	pr* rr = (pr*)malloc(sizeof(pr) * (kp + 1));
	for (;;) {
		size_t qi = claim(j->next, j->mu, 1, j->m);
		if (qi >= j->m) break;
		const float* qv = j->q + qi * j->d;
		size_t cnt = 0;
		for (size_t i = 0; i < j->n; ++i) {
			const uint8_t* c = j->codes + i * j->d;
			sc e = { j->metric == ORC_L2 ? (int64_t)orc_dist2_u8(qv, c, j->d) : (int64_t)ip_u8(qv, c, j->d), i };
			if (cnt == kp && !sc_less(e, best[cnt - 1])) continue;
			size_t p = cnt < kp ? cnt++ : cnt - 1;
			while (p > 0 && sc_less(e, best[p - 1])) { best[p] = best[p - 1]; --p; }
			best[p] = e;
		}
		for (size_t i = 0; i < kp; ++i) {
			if (j->scores) j->scores[qi * kp + i] = i < cnt ? (int32_t)best[i].score : INT32_MAX;
			if (j->cand) j->cand[qi * kp + i] = i < cnt ? best[i].id : UINT64_MAX;
		}
		for (size_t i = 0; i < cnt; ++i) {
			const float* x = j->base + best[i].id * j->d;
			rr[i].first = j->metric == ORC_L2 ? orc_dist2_f32(qv, x, j->d) : -orc_dot_f32(qv, x, j->d);
			rr[i].second = best[i].id;
		}
		qsort(rr, cnt, sizeof(pr), pr_cmp_qsort);
		for (size_t i = 0; i < j->k; ++i) {
			j->ids[qi * j->k + i] = i < cnt ? rr[i].second : UINT64_MAX;
			if (j->dists) j->dists[qi * j->k + i] = i < cnt ? rr[i].first : INFINITY;
		}
	}
	free(best);
	free(rr);
	return NULL;
}
void orc_scan_u8_query(const float* base, const uint8_t* codes, size_t n, size_t d, const float* q, size_t m,
                       size_t k, size_t kprime, int metric, uint64_t* ids_out, float* dists_out,
                       int32_t* scores_out, uint64_t* cand_out, int threads) {
	size_t next = 0;
	pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
	if (kprime < k) kprime = k;
	u8_job j = { base, codes, n, d, q, m, k, kprime, metric, ids_out, dists_out, scores_out, cand_out, &next, &mu };
	run_threads(u8_worker, &j, threads);
}

/* ======================================================================================================
 * graph  (src/antitopo_engine.h)
 * ====================================================================================================== */
struct orc_graph {
	size_t n, d;
	float* vectors;      /* n*d */
	uint8_t* codes;      /* n*d, quantizer_simple<uint8_t> (built lazily) */
	uint32_t* nlayers;   /* n */
	uint64_t* vl_off;    /* n+1: index of (v,0) in the (v,layer) list */
	uint64_t* offsets;   /* total_layers+1 */
	uint32_t* edges;
	uint64_t starting_vertex, max_layer, M, M0;
};
typedef struct { char* visited; uint64_t* recent; size_t nrecent; heap_t cand, nearest; uint32_t* nbl; } gscratch;

static void gscratch_init(gscratch* s, const orc_graph* g) {
	s->visited = (char*)calloc(g->n ? g->n : 1, 1);
	s->recent = (uint64_t*)malloc(sizeof(uint64_t) * (g->n ? g->n : 1));
	s->nrecent = 0;
	heap_init(&s->cand, CMP_FIRST_GREATER);
	heap_init(&s->nearest, CMP_FIRST_LESS);
	s->nbl = (uint32_t*)malloc(sizeof(uint32_t) * 65536);
}
static void gscratch_free(gscratch* s) {
	free(s->visited); free(s->recent); heap_free(&s->cand); heap_free(&s->nearest); free(s->nbl);
}

static int rd(FILE* f, void* p, size_t sz) { return fread(p, 1, sz, f) == sz ? 0 : -1; }

orc_graph* orc_graph_load(const char* path) { /* :994-1074 */
	FILE* f = fopen(path, "rb");
	if (!f) return NULL;
	orc_graph* g = (orc_graph*)calloc(1, sizeof(orc_graph));
	uint64_t u64; uint8_t b; float fl;
	int bad = 0;
	bad |= rd(f, &g->starting_vertex, 8);
	bad |= rd(f, &g->M, 8);
	bad |= rd(f, &g->M0, 8);
	bad |= rd(f, &u64, 8);               /* ef_search_mult (ignored on load, :1000-1003) */
	bad |= rd(f, &b, 1);                 /* has_ef_search */
	if (!bad && b) bad |= rd(f, &u64, 8);
	bad |= rd(f, &u64, 8);               /* ef_construction */
	bad |= rd(f, &u64, 8);               /* ortho_count */
	bad |= rd(f, &fl, 4);                /* ortho_factor */
	bad |= rd(f, &fl, 4);                /* ortho_bias */
	bad |= rd(f, &u64, 8);               /* prune_overflow */
	bad |= rd(f, &b, 1);                 /* use_compression (ignored, :1021-1024) */
	bad |= rd(f, &b, 1);                 /* use_largest_direction_filtering */
	bad |= rd(f, &g->max_layer, 8);
	uint64_t n = 0;
	bad |= rd(f, &n, 8);
	if (bad) { fclose(f); free(g); return NULL; }
	g->n = n;
	for (uint64_t i = 0; i < n && !bad; ++i) {
		uint64_t sz;
		bad |= rd(f, &sz, 8);
		if (i == 0) { g->d = sz; g->vectors = (float*)malloc(sizeof(float) * n * sz); }
		if (sz != g->d) bad = 1;
		if (!bad) bad |= rd(f, g->vectors + i * g->d, 4 * g->d);
	}
	uint64_t nv = 0;
	bad |= rd(f, &nv, 8);
	if (bad || nv != n) { fclose(f); free(g->vectors); free(g); return NULL; }
	g->nlayers = (uint32_t*)malloc(4 * (n + 1));
	g->vl_off = (uint64_t*)malloc(8 * (n + 1));
	size_t ocap = 2 * n + 16, ecap = 64 * n + 16, no = 0, ne = 0;
	g->offsets = (uint64_t*)malloc(8 * ocap);
	g->edges = (uint32_t*)malloc(4 * ecap);
	for (uint64_t v = 0; v < n && !bad; ++v) {
		uint64_t nl;
		bad |= rd(f, &nl, 8);
		g->nlayers[v] = (uint32_t)nl;
		g->vl_off[v] = no;
		for (uint64_t l = 0; l < nl && !bad; ++l) {
			uint64_t deg;
			bad |= rd(f, &deg, 8);
			if (no + 2 > ocap) { ocap *= 2; g->offsets = (uint64_t*)realloc(g->offsets, 8 * ocap); }
			g->offsets[no++] = ne;
			if (ne + deg > ecap) { while (ne + deg > ecap) ecap *= 2; g->edges = (uint32_t*)realloc(g->edges, 4 * ecap); }
			for (uint64_t e = 0; e < deg && !bad; ++e) {
				float len; uint64_t id;
				bad |= rd(f, &len, 4);
				bad |= rd(f, &id, 8);
				g->edges[ne++] = (uint32_t)id;
			}
		}
	}
	g->vl_off[n] = no;
	g->offsets[no] = ne;
	fclose(f);
	if (bad) { orc_graph_destroy(g); return NULL; }
	return g;
}

orc_graph* orc_graph_from_arrays(const float* vectors, size_t n, size_t d, const uint32_t* nlayers,
                                 const uint64_t* offsets, const uint32_t* edges, uint64_t starting_vertex,
                                 uint64_t max_layer, uint64_t M, uint64_t M0) {
	orc_graph* g = (orc_graph*)calloc(1, sizeof(orc_graph));
	g->n = n; g->d = d; g->starting_vertex = starting_vertex; g->max_layer = max_layer; g->M = M; g->M0 = M0;
	g->vectors = (float*)malloc(sizeof(float) * n * d);
	memcpy(g->vectors, vectors, sizeof(float) * n * d);
	g->nlayers = (uint32_t*)malloc(4 * (n + 1));
	memcpy(g->nlayers, nlayers, 4 * n);
	g->vl_off = (uint64_t*)malloc(8 * (n + 1));
	uint64_t tot = 0;
	for (size_t v = 0; v < n; ++v) { g->vl_off[v] = tot; tot += nlayers[v]; }
	g->vl_off[n] = tot;
	g->offsets = (uint64_t*)malloc(8 * (tot + 1));
	memcpy(g->offsets, offsets, 8 * (tot + 1));
	g->edges = (uint32_t*)malloc(4 * (offsets[tot] ? offsets[tot] : 1));
	memcpy(g->edges, edges, 4 * offsets[tot]);
	return g;
}
void orc_graph_destroy(orc_graph* g) {
	if (!g) return;
	free(g->vectors); free(g->codes); free(g->nlayers); free(g->vl_off); free(g->offsets); free(g->edges);
	free(g);
}
uint64_t orc_graph_info(const orc_graph* g, int what) {
	switch (what) {
	case 0: return g->starting_vertex;
	case 1: return g->max_layer;
	case 2: return g->n;
	case 3: return g->M;
	case 4: return g->M0;
	case 5: return g->d;
	}
	return 0;
}
const float* orc_graph_vectors(const orc_graph* g) { return g->vectors; }
static inline const uint32_t* nbrs(const orc_graph* g, uint64_t v, uint32_t layer, size_t* deg) {
	uint64_t o = g->vl_off[v] + layer;
	*deg = (size_t)(g->offsets[o + 1] - g->offsets[o]);
	return g->edges + g->offsets[o];
}
size_t orc_graph_neighbours(const orc_graph* g, uint64_t v, uint32_t layer, uint32_t* out, size_t cap) {
	if (v >= g->n || layer >= g->nlayers[v]) return 0;
	size_t deg;
	const uint32_t* e = nbrs(g, v, layer, &deg);
	for (size_t i = 0; i < deg && i < cap; ++i) out[i] = e[i];
	return deg;
}
static void ensure_codes(orc_graph* g) { /* :485-486 quant->build */
	if (g->codes) return;
	uint8_t* c = (uint8_t*)malloc(g->n * g->d);
	orc_quant_simple_u8(g->vectors, g->n, g->d, c);
	g->codes = c;
}

/* one query; returns number of results written (<= k) */
static size_t graph_one(const orc_graph* g, gscratch* s, const float* q, size_t k, size_t ef, int compressed,
                        uint64_t* ids, float* dists, uint64_t* distcomps) {
	const size_t d = g->d;
	uint64_t dc = 0;
	/* ---- upper layers: greedy descent (:861-903); entry_points is empty so the ortho term is inert */
	uint64_t ep = g->starting_vertex;
	float ep_dist = orc_dist2_f32(g->vectors + ep * d, q, d);
	++dc;
	for (uint64_t layer = g->max_layer - 1; g->max_layer > 0 && layer > 0; --layer) {
		int changed = 1;
		while (changed) {
			changed = 0;
			size_t deg;
			/* NB: the reference iterates hadj_flat[entry_point][layer] by reference while entry_point is
			 * reassigned inside the loop (:879-889); the range-for has already bound the container of the OLD
			 * entry point, so the scan continues over the old list. Restated the same way. */
			const uint32_t* e = nbrs(g, ep, (uint32_t)layer, &deg);
			for (size_t i = 0; i < deg; ++i) {
				float nd = orc_dist2_f32(g->vectors + (uint64_t)e[i] * d, q, d);
				++dc;
				if (nd < ep_dist) { ep = e[i]; ep_dist = nd; changed = 1; }
			}
		}
	}
	/* ---- bottom layer beam search (:495-708 / :710-851) */
	heap_t* cand = &s->cand; heap_t* nearest = &s->nearest;
	cand->n = 0; nearest->n = 0;
	float sd;
	if (compressed) sd = (float)orc_dist2_u8(q, g->codes + ep * d, d);
	else sd = orc_dist2_f32(q, g->vectors + ep * d, d);
	++dc;
	heap_push(cand, sd, ep);    /* range ctor + make_heap of a single element */
	heap_push(nearest, sd, ep);
	while (nearest->n > ef) heap_pop(nearest);
	s->visited[ep] = 1;
	s->recent[s->nrecent++] = ep;
	while (cand->n) {
		pr cur = heap_top(cand);
		heap_pop(cand);
		if (cur.first > heap_top(nearest).first && nearest->n == ef) break; /* :588-590 */
		size_t deg, nn = 0;
		const uint32_t* e = nbrs(g, cur.second, 0, &deg);
		for (size_t i = 0; i < deg; ++i) /* :595-607 */
			if (!s->visited[e[i]]) {
				s->nbl[nn++] = e[i];
				s->visited[e[i]] = 1;
				s->recent[s->nrecent++] = e[i];
			}
		for (size_t i = 0; i < nn; ++i) { /* :658-677 */
			uint64_t nb = s->nbl[i];
			float dn = compressed ? (float)orc_dist2_u8(q, g->codes + nb * d, d)
			                      : orc_dist2_f32(q, g->vectors + nb * d, d);
			++dc;
			if (nearest->n < ef || dn < heap_top(nearest).first) {
				heap_push(cand, dn, nb);
				heap_push(nearest, dn, nb);
				if (nearest->n > ef) heap_pop(nearest);
			}
		}
	}
	for (size_t i = 0; i < s->nrecent; ++i) s->visited[s->recent[i]] = 0; /* :692-694 */
	s->nrecent = 0;
	size_t cnt = nearest->n;
	/* drain max-first then reverse -> ascending (:695-707); compressed: fp32 rerank without re-sort (:845-849) */
	pr* tmp = (pr*)malloc(sizeof(pr) * (cnt ? cnt : 1));
	for (size_t i = 0; i < cnt; ++i) { tmp[cnt - 1 - i] = heap_top(nearest); heap_pop(nearest); }
	if (compressed)
		for (size_t i = 0; i < cnt; ++i) tmp[i].first = orc_dist2_f32(g->vectors + tmp[i].second * d, q, d);
	size_t out = cnt < k ? cnt : k; /* :914-919 */
	for (size_t i = 0; i < k; ++i) {
		ids[i] = i < out ? tmp[i].second : UINT64_MAX;
		if (dists) dists[i] = i < out ? tmp[i].first : INFINITY;
	}
	free(tmp);
	if (distcomps) *distcomps += dc;
	return out;
}

void orc_graph_query(orc_graph* g, const float* q, size_t m, size_t k, size_t ef, int use_compression,
                     uint64_t* ids_out, float* dists_out, uint64_t* counts_out, uint64_t* distcomps) {
	if (use_compression) ensure_codes(g);
	gscratch s;
	gscratch_init(&s, g);
	for (size_t i = 0; i < m; ++i) {
		size_t c = graph_one(g, &s, q + i * g->d, k, ef, use_compression, ids_out + i * k,
		                     dists_out ? dists_out + i * k : NULL, distcomps);
		if (counts_out) counts_out[i] = c;
	}
	gscratch_free(&s);
}

typedef struct {
	orc_graph* g; const float* q; size_t m, k, ef; int comp; uint64_t* ids; uint64_t* counts;
	size_t* next; pthread_mutex_t* mu;
} g_job;
static void* g_worker(void* arg) {
	g_job* j = (g_job*)arg;
	gscratch s;
	gscratch_init(&s, j->g);
	for (;;) {
		size_t i0 = claim(j->next, j->mu, 16, j->m);
		if (i0 >= j->m) break;
		size_t i1 = i0 + 16 < j->m ? i0 + 16 : j->m;
		for (size_t i = i0; i < i1; ++i) {
			size_t c = graph_one(j->g, &s, j->q + i * j->g->d, j->k, j->ef, j->comp, j->ids + i * j->k, NULL, NULL);
			if (j->counts) j->counts[i] = c;
		}
	}
	gscratch_free(&s);
	return NULL;
}
void orc_graph_query_mt(orc_graph* g, const float* q, size_t m, size_t k, size_t ef, int use_compression,
                        uint64_t* ids_out, uint64_t* counts_out, int threads) {
	if (use_compression) ensure_codes(g);
	size_t next = 0;
	pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
	g_job j = { g, q, m, k, ef, use_compression, ids_out, counts_out, &next, &mu };
	run_threads(g_worker, &j, threads);
}
