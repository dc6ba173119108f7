/*
 * bvgen.c -- synthetic graph generators and a BVGraph *encoder*, used only to manufacture inputs
 * (.graph / .offsets / .properties triples) for the decode engine, its tests and its benchmark.
 *
 * There is no JVM on this image, so the files the reference's compressor would have written have to be
 * produced some other way.  The encoder follows the rules of the reference compressor (greedy reference
 * choice over the window with the max-ref-count limit, copy blocks, intervals, gap-coded residuals:
 * BVGraph.java:1595-1618, 1977-2159, 2218-2286 of webgraph-big) closely enough that
 * re-encoding the reference's cnr-2000 golden graph reproduces cnr-2000.graph bit for bit
 * (tests/test_datagen.py).  It is independent of oracle/ (own bit writer, own code-length tables), so an
 * encode -> oracle-decode round trip cross-checks both.
 *
 * Parallelism mirrors the reference's own (BVGraph.java:2408-2483): the node range is cut into contiguous
 * splits, every split is compressed starting from an empty window, and the bit streams are concatenated
 * bit-exactly.  The number of splits is a parameter of the data set (not of the machine) so that the output
 * is identical whatever the number of worker threads.
 */
#define _GNU_SOURCE
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { C_DELTA = 1, C_GAMMA = 2, C_GOLOMB = 3, C_UNARY = 5, C_ZETA = 6, C_NIBBLE = 7 };

typedef struct {
	int32_t window, maxref, minint, zetak;
	int32_t outdegree_coding, block_coding, residual_coding, reference_coding, blockcount_coding, offset_coding;
} bvg_params;

enum { BVG_MODEL_CSR = 0, BVG_MODEL_ER = 1, BVG_MODEL_COPY = 2, BVG_MODEL_BIG = 3 };
/* encoder statistics: copied / intervalised / residual arcs, rows with a reference, blocks, intervals, and the bits
 * spent on outdegrees, references, blocks (counts included), intervals (count included) and residuals; rows whose
 * reference row has more than 128 successors; sum of reference-chain lengths; rows with a chain of length 1, 2, 3, >3 */
enum { ST_COPIED = 0, ST_INTERVAL, ST_RESIDUAL, ST_REFROWS, ST_BLOCKS, ST_INTERVALS, ST_BITS_OUTDEG, ST_BITS_REF, ST_BITS_BLOCKS,
       ST_BITS_INTERVALS, ST_BITS_RESIDUALS, ST_BIGREF, ST_CHAIN_SUM, ST_CHAIN1, ST_CHAIN2, ST_CHAIN3, ST_CHAIN4, BVG_NSTATS = 20 };

typedef struct {
	int32_t model;
	int64_t n;            /* nodes generated: ids [first_node, first_node + n) */
	int64_t first_node;   /* id of the first node (parts of a larger graph are generated independently) */
	int64_t global_n;     /* nodes of the whole graph: successors lie in [0, global_n) */
	uint64_t seed;
	/* ER: p = arc probability.  COPY: mean_degree, p_ref, p_keep, p_run, far_frac, near_scale, max_degree.
	 * BIG: giant_every = one giant row every that many nodes (0: only node 0), giant_degree, giant_stride. */
	double p;
	double mean_degree, p_ref, p_keep, p_run, far_frac, near_scale;
	int64_t max_degree;
	int64_t giant_every, giant_degree, giant_stride;
	/* CSR model */
	const int64_t *rowptr, *col;
} bvg_model;

/* ------------------------------------------------------------------ bit writer (MSB first) -- */
typedef struct {
	uint8_t *buf;
	uint64_t cap;   /* bytes */
	uint64_t bits;  /* bits written */
	uint64_t acc;   /* pending bits, right aligned */
	int nacc;
} obs_t;

static void obs_reserve(obs_t *o, uint64_t more_bytes) {
	const uint64_t need = (o->bits >> 3) + more_bytes + 16;
	if (need <= o->cap) return;
	uint64_t c = o->cap ? o->cap : 1 << 16;
	while (c < need) c += c >> 1;
	o->buf = (uint8_t *)realloc(o->buf, c);
	o->cap = c;
}

static inline void obs_flush_bytes(obs_t *o) {
	while (o->nacc >= 8) {
		o->nacc -= 8;
		o->buf[(o->bits - o->nacc - 8) >> 3] = (uint8_t)(o->acc >> o->nacc);
	}
}

/* append the low `len` bits of v, len in [0,57] */
static inline void obs_put(obs_t *o, uint64_t v, int len) {
	if (len == 0) return;
	obs_reserve(o, 16);
	o->acc = (o->acc << len) | (v & (len == 64 ? ~0ull : ((1ull << len) - 1)));
	o->nacc += len;
	o->bits += len;
	obs_flush_bytes(o);
}

static inline void obs_put_long(obs_t *o, uint64_t v, int len) { /* len in [0,64] */
	if (len > 32) { obs_put(o, v >> 32, len - 32); obs_put(o, v & 0xffffffffu, 32); }
	else obs_put(o, v, len);
}

static void obs_finish(obs_t *o) { /* zero-pad the last byte */
	obs_reserve(o, 16);
	if (o->nacc) { o->buf[(o->bits - o->nacc) >> 3] = (uint8_t)(o->acc << (8 - o->nacc)); }
}

static inline int msb64(uint64_t x) { return 63 - __builtin_clzll(x); }

/* ------------------------------------------------------------------ codes: lengths + writers -- */
static inline uint64_t len_unary(uint64_t x) { return x + 1; }
static inline uint64_t len_gamma(uint64_t x) { return 2 * (uint64_t)msb64(x + 1) + 1; }
static inline uint64_t len_delta(uint64_t x) { const int m = msb64(x + 1); return (uint64_t)m + len_gamma((uint64_t)m); }
static inline uint64_t len_zeta(uint64_t x, int k) {
	const uint64_t y = x + 1;
	const int h = msb64(y) / k;
	const uint64_t left = 1ull << (h * k);
	return (uint64_t)h + 1 + (uint64_t)(h * k + k - 1) + (y - left < left ? 0 : 1);
}
static inline uint64_t len_nibble(uint64_t x) {
	const int bits = x ? msb64(x) + 1 : 1;
	return 4 * (uint64_t)((bits + 2) / 3);
}
static inline uint64_t len_golomb(uint64_t x, int b) {
	if (b == 0) return 0;
	const int s = msb64((uint64_t)b);
	const uint64_t M = (1ull << (s + 1)) - (uint64_t)b;
	return x / b + 1 + (uint64_t)s + ((x % b) < M ? 0 : 1);
}

static void put_unary(obs_t *o, uint64_t x) {
	while (x >= 32) { obs_put(o, 0, 32); x -= 32; }
	obs_put(o, 1, (int)x + 1);
}
static void put_gamma(obs_t *o, uint64_t x) {
	const int m = msb64(x + 1);
	put_unary(o, (uint64_t)m);
	obs_put_long(o, (x + 1) & ((1ull << m) - 1), m);
}
static void put_delta(obs_t *o, uint64_t x) {
	const int m = msb64(x + 1);
	put_gamma(o, (uint64_t)m);
	obs_put_long(o, (x + 1) & ((1ull << m) - 1), m);
}
static void put_zeta(obs_t *o, uint64_t x, int k) {
	const uint64_t y = x + 1;
	const int h = msb64(y) / k;
	const uint64_t left = 1ull << (h * k);
	put_unary(o, (uint64_t)h);
	if (y - left < left) obs_put_long(o, y - left, h * k + k - 1);
	else obs_put_long(o, y, h * k + k);
}
static void put_nibble(obs_t *o, uint64_t x) {
	const int bits = x ? msb64(x) + 1 : 1;
	int g = (bits + 2) / 3;
	while (g-- > 0) obs_put(o, ((g == 0 ? 8u : 0u) | ((x >> (3 * g)) & 7u)), 4);
}
static void put_golomb(obs_t *o, uint64_t x, int b) {
	if (b == 0) return;
	const int s = msb64((uint64_t)b);
	const uint64_t M = (1ull << (s + 1)) - (uint64_t)b, r = x % b;
	put_unary(o, x / b);
	if (r < M) obs_put(o, r, s);
	else obs_put(o, r + M, s + 1);
}

static inline uint64_t len_code(int coding, uint64_t x, int k) {
	switch (coding) {
	case C_GAMMA: return len_gamma(x);
	case C_DELTA: return len_delta(x);
	case C_UNARY: return len_unary(x);
	case C_ZETA: return len_zeta(x, k);
	case C_NIBBLE: return len_nibble(x);
	case C_GOLOMB: return len_golomb(x, k);
	default: return 1ull << 40;
	}
}
static inline void put_code(obs_t *o, int coding, uint64_t x, int k) {
	switch (coding) {
	case C_GAMMA: put_gamma(o, x); break;
	case C_DELTA: put_delta(o, x); break;
	case C_UNARY: put_unary(o, x); break;
	case C_ZETA: put_zeta(o, x, k); break;
	case C_NIBBLE: put_nibble(o, x); break;
	case C_GOLOMB: put_golomb(o, x, k); break;
	default: break;
	}
}

static inline uint64_t int2nat(int64_t v) { return v >= 0 ? (uint64_t)v << 1 : (((uint64_t)(-(v + 1))) << 1) + 1; }

/* ------------------------------------------------------------------ per-node encoder -- */
typedef struct { int64_t *a; int64_t n, cap; } lvec;
static inline void lv_push(lvec *v, int64_t x) {
	if (v->n == v->cap) { v->cap = v->cap ? v->cap * 2 : 64; v->a = (int64_t *)realloc(v->a, (size_t)v->cap * 8); }
	v->a[v->n++] = x;
}
static inline void lv_need(lvec *v, int64_t n) {
	if (n > v->cap) { while (v->cap < n) v->cap = v->cap ? v->cap * 2 : 64; v->a = (int64_t *)realloc(v->a, (size_t)v->cap * 8); }
}

typedef struct {
	const bvg_params *g;
	lvec blocks, extras, left, len, residuals;
	/* what the reference's compressor reports in .properties (BVG:2490-2567), accumulated while emitting */
	int64_t stats[BVG_NSTATS];
} enc_t;

/* The reference's differential compressor: emits (or, with o == NULL, only prices) everything after
 * the outdegree for node `x` whose list is cur[0..curlen) against the reference list (distance ref). */
static uint64_t diff_comp(enc_t *e, obs_t *o, int64_t x, int ref, const int64_t *refl, int64_t reflen,
                          const int64_t *cur, int64_t curlen) {
	const bvg_params *g = e->g;
	uint64_t bits = 0;
	int64_t j = 0, k = 0, run = 0;
	int copying = 1;
	if (ref == 0) reflen = 0;
	e->blocks.n = e->extras.n = 0;
	while (j < curlen && k < reflen) {
		if (copying) {
			if (cur[j] > refl[k]) { lv_push(&e->blocks, run); copying = 0; run = 0; }
			else if (cur[j] < refl[k]) lv_push(&e->extras, cur[j++]);
			else { j++; k++; run++; }
		} else {
			if (cur[j] < refl[k]) lv_push(&e->extras, cur[j++]);
			else if (cur[j] > refl[k]) { k++; run++; }
			else { lv_push(&e->blocks, run); copying = 1; run = 0; }
		}
	}
	if (copying && k < reflen) lv_push(&e->blocks, run);
	while (j < curlen) lv_push(&e->extras, cur[j++]);

	if (g->window > 0) {
		bits += len_code(g->reference_coding, (uint64_t)ref, 0);
		if (o) { put_code(o, g->reference_coding, (uint64_t)ref, 0); e->stats[ST_BITS_REF] += (int64_t)bits; }
	}
	if (ref != 0) {
		const int64_t bc = e->blocks.n;
		const uint64_t before = bits;
		bits += len_code(g->blockcount_coding, (uint64_t)bc, 0);
		if (o) put_code(o, g->blockcount_coding, (uint64_t)bc, 0);
		for (int64_t i = 0; i < bc; i++) {
			const uint64_t v = (uint64_t)(e->blocks.a[i] - (i ? 1 : 0));
			bits += len_code(g->block_coding, v, 0);
			if (o) put_code(o, g->block_coding, v, 0);
		}
		if (o) {
			e->stats[ST_REFROWS]++;
			e->stats[ST_BLOCKS] += bc;
			e->stats[ST_BITS_BLOCKS] += (int64_t)(bits - before);
			e->stats[ST_COPIED] += curlen - e->extras.n;
			if (reflen > 128) e->stats[ST_BIGREF]++;
		}
	}
	if (e->extras.n > 0) {
		const int64_t *res = e->extras.a;
		int64_t nres = e->extras.n;
		if (g->minint != 0) {
			/* intervalize: maximal runs of consecutive integers of length >= minint */
			const int64_t *v = e->extras.a; const int64_t vl = e->extras.n;
			e->left.n = e->len.n = e->residuals.n = 0;
			for (int64_t i = 0; i < vl; i++) {
				int64_t r = 0;
				if (i < vl - 1 && v[i] + 1 == v[i + 1]) {
					do r++; while (i + r < vl - 1 && v[i + r] + 1 == v[i + r + 1]);
					r++;
					if (r >= g->minint) { lv_push(&e->left, v[i]); lv_push(&e->len, r); i += r - 1; }
				}
				if (r < g->minint) lv_push(&e->residuals, v[i]);
			}
			const int64_t ic = e->left.n;
			const uint64_t before = bits;
			bits += len_gamma((uint64_t)ic);
			if (o) put_gamma(o, (uint64_t)ic);
			int64_t prev = 0;
			for (int64_t i = 0; i < ic; i++) {
				const uint64_t a = i == 0 ? int2nat(e->left.a[i] - x) : (uint64_t)(e->left.a[i] - prev - 1);
				const uint64_t b = (uint64_t)(e->len.a[i] - g->minint);
				prev = e->left.a[i] + e->len.a[i];
				bits += len_gamma(a) + len_gamma(b);
				if (o) { put_gamma(o, a); put_gamma(o, b); e->stats[ST_INTERVAL] += e->len.a[i]; }
			}
			if (o) { e->stats[ST_INTERVALS] += ic; e->stats[ST_BITS_INTERVALS] += (int64_t)(bits - before); }
			res = e->residuals.a; nres = e->residuals.n;
		}
		if (o) e->stats[ST_RESIDUAL] += nres;
		const uint64_t before_res = bits;
		if (nres != 0) {
			int64_t prev = res[0];
			uint64_t a = int2nat(prev - x);
			bits += len_code(g->residual_coding, a, g->zetak);
			if (o) put_code(o, g->residual_coding, a, g->zetak);
			for (int64_t i = 1; i < nres; i++) {
				a = (uint64_t)(res[i] - prev - 1);
				bits += len_code(g->residual_coding, a, g->zetak);
				if (o) put_code(o, g->residual_coding, a, g->zetak);
				prev = res[i];
			}
		}
		if (o) e->stats[ST_BITS_RESIDUALS] += (int64_t)(bits - before_res);
	}
	return bits;
}

/* ------------------------------------------------------------------ RNG -- */
static inline uint64_t splitmix64(uint64_t *s) {
	uint64_t z = (*s += 0x9e3779b97f4a7c15ull);
	z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
	z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
	return z ^ (z >> 31);
}
static inline double u01(uint64_t *s) { return (double)(splitmix64(s) >> 11) * (1.0 / 9007199254740992.0); }
static inline uint64_t node_seed(uint64_t seed, int64_t x) {
	uint64_t s = seed ^ (0xd1342543de82ef95ull * (uint64_t)(x + 1));
	splitmix64(&s);
	return s;
}

static int cmp_i64(const void *a, const void *b) {
	const int64_t x = *(const int64_t *)a, y = *(const int64_t *)b;
	return x < y ? -1 : x > y;
}
static void sort_unique(lvec *v) {
	if (v->n < 2) return;
	/* insertion sort for the short lists that dominate, qsort otherwise */
	if (v->n <= 48) {
		for (int64_t i = 1; i < v->n; i++) {
			const int64_t t = v->a[i]; int64_t j = i;
			while (j > 0 && v->a[j - 1] > t) { v->a[j] = v->a[j - 1]; j--; }
			v->a[j] = t;
		}
	} else qsort(v->a, (size_t)v->n, 8, cmp_i64);
	int64_t w = 1;
	for (int64_t i = 1; i < v->n; i++) if (v->a[i] != v->a[w - 1]) v->a[w++] = v->a[i];
	v->n = w;
}

/* ------------------------------------------------------------------ generators -- */
/* Erdos-Renyi G(n,p): successors by geometric gap sampling. */
static void gen_er(const bvg_model *m, int64_t x, lvec *out) {
	uint64_t s = node_seed(m->seed, x);
	out->n = 0;
	const double lq = log1p(-m->p);
	int64_t y = -1;
	for (;;) {
		const double u = 1.0 - u01(&s);
		y += 1 + (int64_t)floor(log(u) / lq);
		if (y >= m->global_n || y < 0) break;
		lv_push(out, y);
	}
}

/* Copy model (web-like): a log-normal degree; with probability p_ref the node copies runs of the list of a
 * recent predecessor; with probability p_run it links a run of consecutive ids; the rest are links near x
 * (two-sided geometric) plus a fraction of uniformly random far links. */
static void gen_copy(const bvg_model *m, int64_t x, int64_t split_start, lvec *const *ring, int cyc, lvec *out) {
	uint64_t s = node_seed(m->seed, x);
	out->n = 0;
	if (m->giant_every > 0 && x % m->giant_every == m->giant_every / 2) {
		/* a BigGraph-style giant row (slow/.../BVGraphSlowTest.java:30-96): giant_degree successors i * giant_stride */
		lv_need(out, m->giant_degree);
		for (int64_t i = 0; i < m->giant_degree; i++) {
			const int64_t y = i * m->giant_stride;
			if (y >= m->global_n) break;
			out->a[out->n++] = y;
		}
		return;
	}
	if (u01(&s) < 0.05) return;
	const double sigma = 1.0, mu = log(m->mean_degree) - 0.5 * sigma * sigma;
	const double u1 = 1.0 - u01(&s), u2 = u01(&s);
	const double z = sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2);
	int64_t d = (int64_t)floor(exp(mu + sigma * z) * (1.0 / 0.95));
	if (d > m->max_degree) d = m->max_degree;
	if (d <= 0) return;
	if (x - split_start >= 1 && cyc > 1 && u01(&s) < m->p_ref) {
		int64_t maxr = x - split_start < cyc - 1 ? x - split_start : cyc - 1;
		const int r = 1 + (int)(splitmix64(&s) % (uint64_t)maxr);
		const lvec *src = ring[(x - r) % cyc];
		int keep = u01(&s) < m->p_keep;
		/* run lengths: keep-runs average 8, skip-runs average 8 (1 - p_keep) / p_keep */
		const double pk = 1.0 / 8.0, ps = m->p_keep / (8.0 * (1.0 - m->p_keep) + 1e-9);
		for (int64_t i = 0; i < src->n && out->n < d; ) {
			const double pr = keep ? pk : (ps > 1 ? 1 : ps);
			int64_t run = 1 + (int64_t)floor(log(1.0 - u01(&s)) / log1p(-pr + 1e-12));
			if (run < 1) run = 1;
			if (keep) for (int64_t j = 0; j < run && i + j < src->n && out->n < d; j++) lv_push(out, src->a[i + j]);
			i += run;
			keep = !keep;
		}
	}
	int64_t rest = d - out->n;
	if (rest > 0 && u01(&s) < m->p_run) {
		int64_t len = 4 + (int64_t)(splitmix64(&s) % 13);
		if (len > rest) len = rest;
		int64_t a = x + (int64_t)(splitmix64(&s) % 2001) - 1000;
		if (a < 0) a = 0;
		for (int64_t j = 0; j < len && a + j < m->global_n; j++) lv_push(out, a + j);
		rest -= len;
	}
	const double lq = log1p(-1.0 / m->near_scale);
	for (; rest > 0; rest--) {
		int64_t y;
		if (u01(&s) < m->far_frac) y = (int64_t)(splitmix64(&s) % (uint64_t)m->global_n);
		else {
			const int64_t mag = (int64_t)floor(log(1.0 - u01(&s)) / lq);
			y = (splitmix64(&s) & 1) ? x + mag : x - mag;
			if (y < 0) y = -y;
			if (y >= m->global_n) y = 2 * (m->global_n - 1) - y;
			if (y < 0 || y >= m->global_n) y = (int64_t)(splitmix64(&s) % (uint64_t)m->global_n);
		}
		lv_push(out, y);
	}
	sort_unique(out);
}

/* BigGraph of the reference's slow test (slow/.../BVGraphSlowTest.java:30-85), generalised: node x has
 * successors {x-1, x} (node 0: {0}) except the giant rows, which have giant_degree successors
 * i * giant_stride. */
static void gen_big(const bvg_model *m, int64_t x, lvec *out) {
	out->n = 0;
	const int giant = m->giant_every > 0 ? (x % m->giant_every == 0) : (x == 0);
	if (giant) {
		lv_need(out, m->giant_degree);
		for (int64_t i = 0; i < m->giant_degree; i++) {
			const int64_t y = i * m->giant_stride;
			if (y >= m->global_n) break;
			out->a[out->n++] = y;
		}
		return;
	}
	if (x > 0) lv_push(out, x - 1);
	lv_push(out, x);
}

/* ------------------------------------------------------------------ splits -- */
typedef struct {
	int64_t from, to;
	obs_t obs;
	uint32_t *bitlen;   /* per node record length in bits (records are < 2^32 bits in every data set we make) */
	int error;          /* a record longer than 2^32-1 bits */
	int64_t arcs;
	uint64_t checksum;
	int64_t stats[BVG_NSTATS];
} split_t;

typedef struct {
	bvg_params g;
	bvg_model m;
	int nsplits;
	split_t *splits;
	int next; /* atomic ticket */
	/* results */
	uint8_t *graph; uint64_t graph_bytes, graph_bits;
	uint64_t *offsets;           /* n + 1 */
	uint8_t *offstream; uint64_t offstream_bytes;
	int64_t arcs;
	uint64_t checksum;
	uint64_t *seg_checksum; int64_t seg_nodes, nseg;
	int error;
	int64_t stats[BVG_NSTATS];
} bvg_job;

/* Order-sensitive hash of one successor list and its mix into a commutative graph checksum.  The CUDA side
 * recomputes exactly this from the decoded CSR (tests and bench use it as the "checksum of checksums"). */
static inline uint64_t row_hash(int64_t x, const int64_t *a, int64_t d) {
	uint64_t h = 0x9e3779b97f4a7c15ull ^ (uint64_t)d;
	for (int64_t i = 0; i < d; i++) h = (h ^ (uint64_t)a[i]) * 0x100000001b3ull + 0x632be59bd9b4e019ull;
	h ^= (uint64_t)x * 0xff51afd7ed558ccdull;
	h ^= h >> 32; h *= 0xd6e8feb86659fd93ull; h ^= h >> 32;
	return h;
}

static void encode_split(bvg_job *J, split_t *sp) {
	const bvg_params *g = &J->g;
	const int cyc = g->window + 1;
	lvec *lists = (lvec *)calloc((size_t)cyc, sizeof(lvec));
	lvec **ring = (lvec **)calloc((size_t)cyc, sizeof(lvec *));
	int *refcount = (int *)calloc((size_t)cyc, sizeof(int));
	for (int i = 0; i < cyc; i++) ring[i] = &lists[i];
	enc_t e; memset(&e, 0, sizeof e); e.g = g;
	obs_t *o = &sp->obs;
	const int64_t cnt = sp->to - sp->from;
	sp->bitlen = (uint32_t *)malloc((size_t)(cnt ? cnt : 1) * 4);
	for (int64_t x = sp->from; x < sp->to; x++) {
		const int ci = (int)(x % cyc);
		lvec *cur = &lists[ci];
		switch (J->m.model) {
		case BVG_MODEL_ER: gen_er(&J->m, x, cur); break;
		case BVG_MODEL_COPY: gen_copy(&J->m, x, sp->from, ring, cyc, cur); break;
		case BVG_MODEL_BIG: gen_big(&J->m, x, cur); break;
		default: {
			const int64_t a = J->m.rowptr[x - J->m.first_node], b = J->m.rowptr[x - J->m.first_node + 1];
			lv_need(cur, b - a); memcpy(cur->a, J->m.col + a, (size_t)(b - a) * 8); cur->n = b - a;
		}
		}
		const uint64_t start = o->bits;
		const int64_t d = cur->n;
		put_code(o, g->outdegree_coding, (uint64_t)d, 0);
		e.stats[ST_BITS_OUTDEG] += (int64_t)(o->bits - start);
		if (d > 0) {
			uint64_t best = ~0ull; int bestref = -1, bestcand = -1;
			refcount[ci] = -1;
			for (int ref = 0; ref < cyc; ref++) {
				if (ref > x - sp->from) break; /* slots older than the split start are empty */
				const int cand = (int)((x - ref) % cyc);
				if (refcount[cand] < g->maxref && lists[cand].n != 0) {
					const uint64_t c = diff_comp(&e, NULL, x, ref, lists[cand].a, lists[cand].n, cur->a, d);
					if (c < best) { best = c; bestref = ref; bestcand = cand; }
				}
			}
			refcount[ci] = refcount[bestcand] + 1;
			diff_comp(&e, o, x, bestref, lists[bestcand].a, lists[bestcand].n, cur->a, d);
			if (bestref) {
				const int c = refcount[ci];
				e.stats[ST_CHAIN_SUM] += c;
				e.stats[c >= 4 ? ST_CHAIN4 : ST_CHAIN1 + c - 1]++;
			}
		}
		if (o->bits - start > 0xffffffffull) sp->error = 1;
		sp->bitlen[x - sp->from] = (uint32_t)(o->bits - start);
		sp->arcs += d;
		{
			const uint64_t h = row_hash(x, cur->a, d);
			sp->checksum += h;
			if (J->seg_checksum) __atomic_fetch_add(&J->seg_checksum[(x - J->m.first_node) / J->seg_nodes], h, __ATOMIC_RELAXED);
		}
	}
	obs_finish(o);
	memcpy(sp->stats, e.stats, sizeof sp->stats);
	for (int i = 0; i < cyc; i++) free(lists[i].a);
	free(lists); free(ring); free(refcount);
	free(e.blocks.a); free(e.extras.a); free(e.left.a); free(e.len.a); free(e.residuals.a);
}

static void *worker_encode(void *p) {
	bvg_job *J = (bvg_job *)p;
	for (;;) {
		const int k = __atomic_fetch_add(&J->next, 1, __ATOMIC_RELAXED);
		if (k >= J->nsplits) break;
		encode_split(J, &J->splits[k]);
	}
	return NULL;
}

/* bit-exact concatenation: split k goes to bit offset dst[k] of the output */
typedef struct { bvg_job *J; uint64_t *dst; } cat_t;
static void *worker_concat(void *p) {
	cat_t *c = (cat_t *)p;
	bvg_job *J = c->J;
	for (;;) {
		const int k = __atomic_fetch_add(&J->next, 1, __ATOMIC_RELAXED);
		if (k >= J->nsplits) break;
		split_t *sp = &J->splits[k];
		const uint64_t L = sp->obs.bits, D = c->dst[k];
		if (L == 0) continue;
		const uint8_t *src = sp->obs.buf;
		uint8_t *out = J->graph + (D >> 3);
		const int s = (int)(D & 7);
		const uint64_t nsrc = (L + 7) >> 3;
		if (s == 0) {
			/* last byte may be shared with the next split: OR it in */
			if (nsrc > 1) memcpy(out, src, nsrc - 1);
			__atomic_fetch_or(&out[nsrc - 1], src[nsrc - 1], __ATOMIC_RELAXED);
		} else {
			const uint64_t nout = (s + L + 7) >> 3;
			__atomic_fetch_or(&out[0], (uint8_t)(src[0] >> s), __ATOMIC_RELAXED);
			for (uint64_t i = 1; i < nout; i++) {
				const uint8_t hi = (uint8_t)(src[i - 1] << (8 - s));
				const uint8_t lo = i < nsrc ? (uint8_t)(src[i] >> s) : 0;
				if (i + 1 < nout) out[i] = hi | lo;
				else __atomic_fetch_or(&out[i], (uint8_t)(hi | lo), __ATOMIC_RELAXED);
			}
		}
		/* absolute offsets of this split's nodes */
		uint64_t acc = D;
		for (int64_t x = sp->from; x < sp->to; x++) { J->offsets[x - J->m.first_node] = acc; acc += sp->bitlen[x - sp->from]; }
		free(sp->obs.buf); sp->obs.buf = NULL;
		free(sp->bitlen); sp->bitlen = NULL;
	}
	return NULL;
}

static void run_threads(void *(*fn)(void *), void *arg, int threads) {
	pthread_t *th = (pthread_t *)calloc((size_t)threads, sizeof(pthread_t));
	for (int i = 0; i < threads; i++) pthread_create(&th[i], NULL, fn, arg);
	for (int i = 0; i < threads; i++) pthread_join(th[i], NULL);
	free(th);
}

/* Generates (or takes, for BVG_MODEL_CSR) a graph and BV-encodes it.  nsplits fixes the split structure of the
 * output; seg_nodes > 0 also produces one checksum per segment of that many nodes. */
bvg_job *bvg_run(const bvg_params *g, const bvg_model *m, int nsplits, int threads, int64_t seg_nodes) {
	bvg_job *J = (bvg_job *)calloc(1, sizeof(bvg_job));
	J->g = *g; J->m = *m;
	if (J->m.global_n <= 0) J->m.global_n = J->m.first_node + J->m.n;
	m = &J->m;
	const int64_t n = m->n;
	/* nsplits < 0: every split has exactly -nsplits nodes (the last one what is left), so that parts of one large graph
	 * generated separately agree on where the compression splits are */
	int64_t per_exact = 0;
	if (nsplits < 0) {
		per_exact = -(int64_t)nsplits;
		nsplits = (int)((n + per_exact - 1) / per_exact);
	}
	if (nsplits < 1) nsplits = 1;
	if ((int64_t)nsplits > n && n > 0) nsplits = (int)n;
	if (n == 0) nsplits = 1;
	if (threads < 1) threads = 1;
	J->nsplits = nsplits;
	J->splits = (split_t *)calloc((size_t)nsplits, sizeof(split_t));
	const int64_t per = per_exact ? per_exact : (n + nsplits - 1) / nsplits;
	for (int k = 0; k < nsplits; k++) {
		J->splits[k].from = m->first_node + (per * k < n ? per * k : n);
		J->splits[k].to = m->first_node + (per * (k + 1) < n ? per * (k + 1) : n);
	}
	if (seg_nodes > 0) {
		J->seg_nodes = seg_nodes; J->nseg = (n + seg_nodes - 1) / seg_nodes;
		J->seg_checksum = (uint64_t *)calloc((size_t)(J->nseg ? J->nseg : 1), 8);
	}
	J->next = 0;
	run_threads(worker_encode, J, threads);

	uint64_t *dst = (uint64_t *)calloc((size_t)nsplits + 1, 8);
	for (int k = 0; k < nsplits; k++) {
		dst[k + 1] = dst[k] + J->splits[k].obs.bits;
		J->arcs += J->splits[k].arcs;
		for (int q = 0; q < BVG_NSTATS; q++) J->stats[q] += J->splits[k].stats[q];
		J->checksum += J->splits[k].checksum;
		J->error |= J->splits[k].error;
	}
	J->graph_bits = dst[nsplits];
	J->graph_bytes = (J->graph_bits + 7) >> 3;
	J->graph = (uint8_t *)calloc(J->graph_bytes + 16, 1);
	J->offsets = (uint64_t *)malloc((size_t)(n + 1) * 8);
	J->offsets[n] = J->graph_bits;
	cat_t c = { J, dst };
	J->next = 0;
	run_threads(worker_concat, &c, threads);
	free(dst);
	return J;
}

/* The .offsets stream: n+1 gamma (or delta) coded differences, the first being 0 (BVGraph.java:2228, 2311). */
static void build_offstream(bvg_job *J) {
	if (J->offstream) return;
	obs_t o; memset(&o, 0, sizeof o);
	uint64_t prev = 0;
	for (int64_t i = 0; i <= J->m.n; i++) {
		put_code(&o, J->g.offset_coding, J->offsets[i] - prev, 0);
		prev = J->offsets[i];
	}
	obs_finish(&o);
	J->offstream = o.buf; J->offstream_bytes = (o.bits + 7) >> 3;
}

const uint8_t *bvg_graph(bvg_job *J, uint64_t *bytes, uint64_t *bits) { *bytes = J->graph_bytes; *bits = J->graph_bits; return J->graph; }
const uint64_t *bvg_offsets(bvg_job *J) { return J->offsets; }
const uint8_t *bvg_offsets_stream(bvg_job *J, uint64_t *bytes) { build_offstream(J); *bytes = J->offstream_bytes; return J->offstream; }
int64_t bvg_arcs(bvg_job *J) { return J->arcs; }
/* out[BVG_NSTATS]: the encoder statistics (see the ST_ enumeration) */
void bvg_stats(bvg_job *J, int64_t *out) { memcpy(out, J->stats, sizeof J->stats); }
int bvg_error(bvg_job *J) { return J->error; }
uint64_t bvg_checksum(bvg_job *J) { return J->checksum; }
const uint64_t *bvg_seg_checksums(bvg_job *J, int64_t *nseg) { *nseg = J->nseg; return J->seg_checksum; }

void bvg_free(bvg_job *J) {
	if (!J) return;
	for (int k = 0; k < J->nsplits; k++) { free(J->splits[k].obs.buf); free(J->splits[k].bitlen); }
	free(J->splits); free(J->graph); free(J->offsets); free(J->offstream); free(J->seg_checksum);
	free(J);
}

/* The same row hash over a CSR held in host memory (used by tests to cross-check the CUDA checksum kernel). */
uint64_t bvg_csr_checksum(int64_t first_node, int64_t n, const int64_t *rowptr, const int64_t *col) {
	uint64_t acc = 0;
	for (int64_t i = 0; i < n; i++) acc += row_hash(first_node + i, col + (rowptr[i] - rowptr[0]), rowptr[i + 1] - rowptr[i]);
	return acc;
}

/* Arc labels as BitStreamArcLabelledImmutableGraph.store writes them (labelling/BitStreamArcLabelledImmutableGraph.java:
 * 654-681): the labels of every node's arcs one after the other in `out`; node_bits[x] = bits taken by node x (what
 * the .labeloffsets stream gamma-codes after a leading gamma(0)).  kind 0: `width`-bit integers (FixedWidthIntLabel),
 * kind 1: gamma-coded naturals (GammaCodedIntLabel).  Returns the bits written (nothing is copied if out_cap is short). */
uint64_t bvg_write_labels(int kind, int width, int64_t n, const int64_t *rowptr, const int64_t *values, uint8_t *out, uint64_t out_cap,
                          uint64_t *node_bits) {
	obs_t o; memset(&o, 0, sizeof o);
	for (int64_t x = 0; x < n; x++) {
		const uint64_t before = o.bits;
		for (int64_t j = rowptr[x]; j < rowptr[x + 1]; j++) {
			if (kind == 0) obs_put_long(&o, (uint64_t)values[j], width);
			else put_gamma(&o, (uint64_t)values[j]);
		}
		node_bits[x] = o.bits - before;
	}
	obs_finish(&o);
	const uint64_t bytes = (o.bits + 7) >> 3;
	if (bytes <= out_cap) memcpy(out, o.buf, bytes);
	const uint64_t bits = o.bits;
	free(o.buf);
	return bits;
}

/* List-valued labels (FixedWidthIntListLabel.toBitStream, labelling/FixedWidthIntListLabel.java:81-85, and
 * FixedWidthLongListLabel.java:87-92): per arc gamma(length) and then each element in `width` bits (0..64).
 * list_ptr[j] .. list_ptr[j + 1] are arc j's elements in `values`. */
uint64_t bvg_write_label_lists(int width, int64_t n, const int64_t *rowptr, const int64_t *list_ptr, const int64_t *values, uint8_t *out,
                               uint64_t out_cap, uint64_t *node_bits) {
	obs_t o; memset(&o, 0, sizeof o);
	const uint64_t mask = width >= 64 ? ~(uint64_t)0 : ((uint64_t)1 << width) - 1;
	for (int64_t x = 0; x < n; x++) {
		const uint64_t before = o.bits;
		for (int64_t j = rowptr[x]; j < rowptr[x + 1]; j++) {
			put_gamma(&o, (uint64_t)(list_ptr[j + 1] - list_ptr[j]));
			for (int64_t t = list_ptr[j]; t < list_ptr[j + 1]; t++) obs_put_long(&o, (uint64_t)values[t] & mask, width);
		}
		node_bits[x] = o.bits - before;
	}
	obs_finish(&o);
	const uint64_t bytes = (o.bits + 7) >> 3;
	if (bytes <= out_cap) memcpy(out, o.buf, bytes);
	const uint64_t bits = o.bits;
	free(o.buf);
	return bits;
}

/* The same for synthetic labels that need no value array (benchmark scale): arc number j (in CSR order) carries
 * label_of_arc(j, width) = the low `width` bits of j * 0x9E3779B97F4A7C15 >> 20.  Fixed-width labels only. */
static inline uint64_t label_of_arc(uint64_t j, int width) {
	const uint64_t h = (j * 0x9E3779B97F4A7C15ull) >> 20;
	return width >= 64 ? h : h & ((1ull << width) - 1);
}
uint64_t bvg_write_labels_synthetic(int width, int64_t n, const int64_t *rowptr, uint8_t *out, uint64_t out_cap, uint64_t *node_bits) {
	const uint64_t arcs = (uint64_t)(rowptr[n] - rowptr[0]);
	const uint64_t bytes = (arcs * (uint64_t)width + 7) >> 3;
	if (bytes + 16 > out_cap) return arcs * (uint64_t)width;
	obs_t o; memset(&o, 0, sizeof o);
	o.buf = out;             /* written in place: the caller's buffer is large enough, obs_reserve never grows it */
	o.cap = out_cap;
	uint64_t j = 0;
	for (int64_t x = 0; x < n; x++) {
		const int64_t d = rowptr[x + 1] - rowptr[x];
		for (int64_t k = 0; k < d; k++, j++) obs_put_long(&o, label_of_arc(j, width), width);
		node_bits[x] = (uint64_t)d * (uint64_t)width;
	}
	if (o.nacc) o.buf[(o.bits - o.nacc) >> 3] = (uint8_t)(o.acc << (8 - o.nacc));
	return o.bits;
}

/* Single-code writer exposed so that tests can round-trip every code through the oracle's readers. */
uint64_t bvg_write_codes(int coding, int k, const uint64_t *vals, int64_t n, uint8_t *out, uint64_t out_cap) {
	obs_t o; memset(&o, 0, sizeof o);
	for (int64_t i = 0; i < n; i++) put_code(&o, coding, vals[i], k);
	obs_finish(&o);
	const uint64_t bytes = (o.bits + 7) >> 3;
	if (bytes <= out_cap) memcpy(out, o.buf, bytes);
	const uint64_t bits = o.bits;
	free(o.buf);
	return bits;
}
