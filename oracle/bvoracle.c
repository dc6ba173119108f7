/*
 * bvoracle.c -- CPU restatement of webgraph-big's BVGraph decode path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this
 * library, and there only as the checker or as the timed CPU baseline.  The shipped decode
 * path is the CUDA library (libwgbv.so); it has no CPU fallback and never links this file.
 *
 * Parity status: PINNED for the default codings (gamma outdegrees/blocks/block counts/intervals/
 * offsets, unary references, zeta_3 residuals, nat2int, MSB-first bit order) by the reference's own
 * known-answer fixture cnr-2000 (the assertion of test/.../BVGraphTest.java:105-123): see
 * tests/test_oracle_golden.py.  UNPINNED for delta (any field), gamma/delta/nibble/Golomb residuals,
 * delta offsets and zeta_k with k != 3: the reference ships no fixture or test using those flags;
 * they follow the published dsiutils definitions.
 *
 * The bit-level code readers live in dsiutils' InputBitStream, a third-party dependency that is not
 * in /root/reference (ivy.xml:18, rev "latest.release"); they are restated from their published
 * definitions and anchored on the reference's call sites, cited function by function below.
 * "BVG" = src/it/unimi/dsi/big/webgraph/BVGraph.java of the reference.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

/* Code ids, CompressionFlags.java:26-46. */
enum { BVO_DELTA = 1, BVO_GAMMA = 2, BVO_GOLOMB = 3, BVO_SKEWED_GOLOMB = 4, BVO_UNARY = 5, BVO_ZETA = 6, BVO_NIBBLE = 7 };

enum { BVO_OK = 0, BVO_EINDEX = -1, BVO_ESTATE = -2, BVO_EIO = -3, BVO_EUNSUPPORTED = -4, BVO_ECAPACITY = -5 };

typedef struct {
	int64_t nodes, arcs;
	int32_t window, maxref, minint, zetak;
	/* BVG:526-542 defaults: gamma, gamma, zeta, unary, gamma, gamma. */
	int32_t outdegree_coding, block_coding, residual_coding, reference_coding, blockcount_coding, offset_coding;
} bvo_params;

/* ---------------------------------------------------------------------------------------------
 * Bit stream: bits are consumed MSB-first inside each byte (dsiutils InputBitStream convention).
 * Reads past the end return zero bits; `overrun` records it so that callers can report EIO.
 * --------------------------------------------------------------------------------------------- */
typedef struct {
	const uint8_t *p;
	uint64_t nbits;
	uint64_t pos;
	int overrun;
} ibs_t;

static inline void ibs_init(ibs_t *s, const uint8_t *p, uint64_t nbytes, uint64_t pos) {
	s->p = p; s->nbits = nbytes * 8; s->pos = pos; s->overrun = 0;
}

static inline int ibs_bit(ibs_t *s) {
	if (s->pos >= s->nbits) { s->overrun = 1; s->pos++; return 0; }
	const int b = (s->p[s->pos >> 3] >> (7 - (s->pos & 7))) & 1;
	s->pos++;
	return b;
}

/* At least 57 valid bits starting at the current position, left-aligned; only legal when 8 whole
 * bytes can be loaded (the callers fall back to the bit-serial path near the end of the buffer). */
static inline int ibs_can_peek(const ibs_t *s) { return (s->pos >> 3) + 8 <= (s->nbits >> 3); }
static inline uint64_t ibs_peek57(const ibs_t *s) {
	uint64_t w;
	memcpy(&w, s->p + (s->pos >> 3), 8);
	return __builtin_bswap64(w) << (s->pos & 7);
}

/* InputBitStream.readLong(len), len in [0,64]. */
static inline uint64_t ibs_bits(ibs_t *s, int len) {
	if (len == 0) return 0;
	if (len <= 57 && ibs_can_peek(s)) {
		const uint64_t v = ibs_peek57(s) >> (64 - len);
		s->pos += len;
		return v;
	}
	uint64_t v = 0;
	while (len > 0) {
		if (s->pos >= s->nbits) { s->overrun = 1; s->pos += len; return len < 64 ? v << len : 0; }
		const int avail = 8 - (int)(s->pos & 7);
		const int take = len < avail ? len : avail;
		const unsigned byte = s->p[s->pos >> 3];
		v = (v << take) | ((byte >> (avail - take)) & ((1u << take) - 1));
		s->pos += take;
		len -= take;
	}
	return v;
}

/* InputBitStream.readUnary: number of zero bits before the first one bit. */
static inline uint64_t ibs_unary(ibs_t *s) {
	uint64_t z = 0;
	while (ibs_can_peek(s)) {
		const uint64_t w = ibs_peek57(s) | ((uint64_t)1 << 6); /* sentinel below the 57 valid bits */
		const int lz = __builtin_clzll(w);
		if (lz < 57) { s->pos += lz + 1; return z + lz; }
		s->pos += 57; z += 57;
	}
	for (;;) {
		if (s->pos >= s->nbits) { s->overrun = 1; return z; }
		if (ibs_bit(s)) return z;
		z++;
	}
}

/* InputBitStream.readLongGamma: unary(msb) then msb low bits of x+1. */
static inline uint64_t ibs_gamma(ibs_t *s) {
	const int msb = (int)ibs_unary(s);
	if (msb > 63) { s->overrun = 1; return 0; }
	return (((uint64_t)1 << msb) | ibs_bits(s, msb)) - 1;
}

/* InputBitStream.readLongDelta: gamma(msb) then msb low bits of x+1. */
static inline uint64_t ibs_delta(ibs_t *s) {
	const uint64_t msb = ibs_gamma(s);
	if (msb > 63) { s->overrun = 1; return 0; }
	return (((uint64_t)1 << msb) | ibs_bits(s, (int)msb)) - 1;
}

/* InputBitStream.readLongZeta(k). */
static inline uint64_t ibs_zeta(ibs_t *s, int k) {
	const uint64_t h = ibs_unary(s);
	if (h * k + k > 64) { s->overrun = 1; return 0; }
	const uint64_t left = (uint64_t)1 << (h * k);
	const uint64_t m = ibs_bits(s, (int)(h * k + k - 1));
	if (m < left) return m + left - 1;
	return (m << 1) + ibs_bit(s) - 1;
}

/* InputBitStream.readLongNibble: 4-bit groups, stop flag first, most significant group first. */
static inline uint64_t ibs_nibble(ibs_t *s) {
	uint64_t x = 0;
	int b;
	do {
		x <<= 3;
		b = ibs_bit(s);
		x |= ibs_bits(s, 3);
	} while (!b && !s->overrun);
	return x;
}

/* InputBitStream.readLongGolomb(b): unary quotient, minimal-binary remainder. */
static inline uint64_t ibs_golomb(ibs_t *s, int b) {
	if (b == 0) return 0;
	const uint64_t q = ibs_unary(s);
	int log2b = 0;
	while (((uint64_t)2 << log2b) <= (uint64_t)b) log2b++; /* floor(log2 b) */
	const uint64_t m = ((uint64_t)1 << (log2b + 1)) - b;
	uint64_t x = ibs_bits(s, log2b);
	if (x >= m) x = (x << 1) + ibs_bit(s) - m;
	return q * b + x;
}

/* Fast.nat2int. */
static inline int64_t nat2int(uint64_t x) { return (x & 1) ? -(int64_t)((x + 1) >> 1) : (int64_t)(x >> 1); }

/* The switchboards: BVG:627-633 (offset), 654-661 (outdegree), 692-703 (reference), 728-735
 * (block count), 758-765 (block), 787-796 (residual).  *err is set to EUNSUPPORTED exactly where
 * the reference throws UnsupportedOperationException. */
static inline uint64_t rd_gd(ibs_t *s, int coding, int allow_unary, int *err) {
	switch (coding) {
	case BVO_GAMMA: return ibs_gamma(s);
	case BVO_DELTA: return ibs_delta(s);
	case BVO_UNARY: if (allow_unary) return ibs_unary(s); /* fallthrough */
	default: *err = BVO_EUNSUPPORTED; return 0;
	}
}
static inline uint64_t rd_offset(ibs_t *s, const bvo_params *g, int *err) { return rd_gd(s, g->offset_coding, 0, err); }
static inline uint64_t rd_outdegree(ibs_t *s, const bvo_params *g, int *err) { return rd_gd(s, g->outdegree_coding, 0, err); }
static inline uint64_t rd_reference(ibs_t *s, const bvo_params *g, int *err) {
	const uint64_t r = rd_gd(s, g->reference_coding, 1, err);
	if (r > (uint64_t)g->window && !*err) *err = BVO_ESTATE; /* BVG:701 IllegalStateException */
	return r;
}
static inline uint64_t rd_blockcount(ibs_t *s, const bvo_params *g, int *err) { return rd_gd(s, g->blockcount_coding, 1, err); }
static inline uint64_t rd_block(ibs_t *s, const bvo_params *g, int *err) { return rd_gd(s, g->block_coding, 1, err); }
static inline uint64_t rd_residual(ibs_t *s, const bvo_params *g, int *err) {
	switch (g->residual_coding) {
	case BVO_GAMMA: return ibs_gamma(s);
	case BVO_ZETA: return ibs_zeta(s, g->zetak);
	case BVO_DELTA: return ibs_delta(s);
	case BVO_GOLOMB: return ibs_golomb(s, g->zetak); /* BVG:792 passes zetaK as the modulus */
	case BVO_NIBBLE: return ibs_nibble(s);
	default: *err = BVO_EUNSUPPORTED; return 0;
	}
}

/* ---------------------------------------------------------------------------------------------
 * Offsets: OffsetsLongIterator, BVG:870-898 -- n+1 running sums of offset deltas.
 * --------------------------------------------------------------------------------------------- */
int bvo_read_offsets(const bvo_params *g, const uint8_t *off, uint64_t off_bytes, uint64_t *out) {
	ibs_t s; ibs_init(&s, off, off_bytes, 0);
	uint64_t acc = 0;
	int err = 0;
	for (int64_t i = 0; i <= g->nodes; i++) {
		acc += rd_offset(&s, g, &err);
		out[i] = acc;
		if (err) return err;
		if (s.overrun) return BVO_EIO;
	}
	return BVO_OK;
}

/* ---------------------------------------------------------------------------------------------
 * One node.  Mirrors BVGraph.successors(x, ibs, window, outd), BVG:995-1097, with the lazy
 * iterator tree flattened into arrays:
 *   MaskedLongIterator (MaskedLongIterator.java:67-100)   -> masked copy of the reference list
 *   LongIntervalSequenceIterator (:57-77)                 -> interval expansion
 *   ResidualLongIterator (BVG:902-954)                    -> gap-decoded residuals
 *   MergedLongIterator (MergedLongIterator.java:54-92)    -> sorted union, duplicates dropped,
 *                                                            at most d values (BVG:1090)
 * `ref_list`/`ref_len` is the finished successor list of x-ref (window slot or recursive decode).
 * On return *len_out is what the reference's drain loop (BVG:1173) would have stored before
 * running into -1; it equals d for every well-formed file.
 * --------------------------------------------------------------------------------------------- */
typedef struct { int64_t *a; int64_t cap; } vec_t;

static int vec_need(vec_t *v, int64_t n) {
	if (n <= v->cap) return 0;
	int64_t c = v->cap ? v->cap : 64;
	while (c < n) c <<= 1;
	int64_t *a = (int64_t *)realloc(v->a, (size_t)c * sizeof(int64_t));
	if (!a) return -1;
	v->a = a; v->cap = c;
	return 0;
}

typedef struct {
	vec_t blocks, left, len, extra, copied;
} scratch_t;

static void scratch_free(scratch_t *t) {
	free(t->blocks.a); free(t->left.a); free(t->len.a); free(t->extra.a); free(t->copied.a);
	memset(t, 0, sizeof *t);
}

/* Header part shared by sequential and random access: everything after the outdegree.
 * `dref` is the outdegree of x-ref, consulted only when the block count is even (BVG:1030). */
/* Optional per-row shape figures for analysis (bvo_row_shapes): ncopied, block count, interval count, interval
 * elements, residual count.  Not part of the decode. */
static __thread int64_t *tl_shape = NULL;

static int decode_tail(const bvo_params *g, ibs_t *s, int64_t x, int64_t d, int ref,
                       const int64_t *ref_list, int64_t dref, scratch_t *t, int64_t *out, int64_t *len_out) {
	int err = 0;
	int64_t extra_count, block_count = 0, ncopied = 0;

	if (ref > 0) {
		block_count = (int64_t)rd_blockcount(s, g, &err);
		if (err) return err;
		if (vec_need(&t->blocks, block_count)) return BVO_ECAPACITY;
		int64_t copied = 0, total = 0;
		for (int64_t i = 0; i < block_count; i++) {
			t->blocks.a[i] = (int64_t)rd_block(s, g, &err) + (i == 0 ? 0 : 1);
			total += t->blocks.a[i];
			if ((i & 1) == 0) copied += t->blocks.a[i];
			if (s->overrun) return BVO_EIO;
		}
		if (err) return err;
		if ((block_count & 1) == 0) copied += dref - total;
		extra_count = d - copied;

		/* MaskedLongIterator: keep B0, skip B1, keep B2 ...; the tail is kept iff the count is even. */
		if (vec_need(&t->copied, dref > 0 ? dref : 1)) return BVO_ECAPACITY;
		int64_t k = 0;
		for (int64_t i = 0; i < block_count && k < dref; i++) {
			int64_t b = t->blocks.a[i];
			if ((i & 1) == 0) { for (; b > 0 && k < dref; b--) t->copied.a[ncopied++] = ref_list[k++]; }
			else k += b;
		}
		if ((block_count & 1) == 0) while (k < dref) t->copied.a[ncopied++] = ref_list[k++];
	} else extra_count = d;

	int64_t interval_count = 0, nextra = 0;
	if (extra_count > 0) {
		/* Intervals, BVG:1037-1060. */
		if (g->minint != 0 && (interval_count = (int64_t)ibs_gamma(s)) != 0) {
			if (vec_need(&t->left, interval_count) || vec_need(&t->len, interval_count)) return BVO_ECAPACITY;
			int64_t prev;
			t->left.a[0] = prev = nat2int(ibs_gamma(s)) + x;
			t->len.a[0] = (int64_t)ibs_gamma(s) + g->minint;
			prev += t->len.a[0];
			extra_count -= t->len.a[0];
			for (int64_t i = 1; i < interval_count; i++) {
				t->left.a[i] = prev = (int64_t)ibs_gamma(s) + prev + 1;
				t->len.a[i] = (int64_t)ibs_gamma(s) + g->minint;
				prev += t->len.a[i];
				extra_count -= t->len.a[i];
				if (s->overrun) return BVO_EIO;
			}
		}
	}
	const int64_t residual_count = extra_count;
	if (residual_count < 0) return BVO_ESTATE; /* malformed: the reference would misbehave here */
	if (tl_shape) {
		int64_t ti = 0;
		for (int64_t i = 0; i < interval_count; i++) ti += t->len.a[i];
		tl_shape[0] = ncopied; tl_shape[1] = block_count; tl_shape[2] = interval_count; tl_shape[3] = ti; tl_shape[4] = residual_count;
	}

	/* extras = Merged(IntervalSequence, Residuals), BVG:1067-1073: a plain merge with duplicates dropped. */
	{
		int64_t total_int = 0;
		for (int64_t i = 0; i < interval_count; i++) total_int += t->len.a[i];
		if (vec_need(&t->extra, total_int + residual_count + 1)) return BVO_ECAPACITY;
		int64_t ii = 0, ij = 0; /* interval index, offset inside it */
		int64_t r = 0, rleft = residual_count;
		if (rleft) r = x + nat2int(rd_residual(s, g, &err)); /* BVG:917 */
		for (;;) {
			const int has_i = ii < interval_count, has_r = rleft > 0;
			if (!has_i && !has_r) break;
			const int64_t vi = has_i ? t->left.a[ii] + ij : 0;
			int take_i = 0, take_r = 0;
			if (!has_r) take_i = 1;
			else if (!has_i) take_r = 1;
			else if (vi < r) take_i = 1;
			else if (vi > r) take_r = 1;
			else take_i = take_r = 1;
			t->extra.a[nextra++] = take_i ? vi : r;
			if (take_i) { if (++ij == t->len.a[ii]) { ii++; ij = 0; } }
			if (take_r) { if (--rleft != 0) r += (int64_t)rd_residual(s, g, &err) + 1; /* BVG:929 */ }
			if (s->overrun) return BVO_EIO;
		}
		if (err) return err;
	}

	/* Merged(blockIterator, extraIterator, d), BVG:1088-1090. */
	int64_t i = 0, j = 0, o = 0;
	while (o < d && (i < ncopied || j < nextra)) {
		if (i >= ncopied) out[o++] = t->extra.a[j++];
		else if (j >= nextra) out[o++] = t->copied.a[i++];
		else if (t->copied.a[i] < t->extra.a[j]) out[o++] = t->copied.a[i++];
		else if (t->copied.a[i] > t->extra.a[j]) out[o++] = t->extra.a[j++];
		else { out[o++] = t->copied.a[i++]; j++; }
	}
	*len_out = o;
	return s->overrun ? BVO_EIO : BVO_OK;
}

/* ---------------------------------------------------------------------------------------------
 * Random access.  BVGraph.outdegree(x) BVG:821-851 and successors(x) BVG:860-867, whose reference
 * list is a RECURSIVE successors(x - ref) (BVG:1084).  `offsets` has n+1 entries.
 * --------------------------------------------------------------------------------------------- */
int64_t bvo_outdegree(const bvo_params *g, const uint8_t *graph, uint64_t graph_bytes, const uint64_t *offsets, int64_t x) {
	if (x < 0 || x >= g->nodes) return BVO_EINDEX;
	ibs_t s; ibs_init(&s, graph, graph_bytes, offsets[x]);
	int err = 0;
	const int64_t d = (int64_t)rd_outdegree(&s, g, &err);
	return err ? err : d;
}

static int successors_rec(const bvo_params *g, const uint8_t *graph, uint64_t graph_bytes, const uint64_t *offsets,
                          int64_t x, int64_t **out, int64_t *len_out, int depth) {
	int err = 0;
	ibs_t s; ibs_init(&s, graph, graph_bytes, offsets[x]);
	const int64_t d = (int64_t)rd_outdegree(&s, g, &err);
	if (err) return err;
	*out = (int64_t *)malloc((size_t)(d ? d : 1) * sizeof(int64_t));
	*len_out = 0;
	if (!*out) return BVO_ECAPACITY;
	if (d == 0) return BVO_OK;
	const int ref = g->window > 0 ? (int)rd_reference(&s, g, &err) : -1;
	if (err) return err;
	int64_t *ref_list = NULL, dref = 0;
	if (ref > 0) {
		if (x - ref < 0) return BVO_ESTATE;
		err = successors_rec(g, graph, graph_bytes, offsets, x - ref, &ref_list, &dref, depth + 1);
		if (err) { free(ref_list); return err; }
		/* outdegreeInternal(x - ref), BVG:1030; equals the decoded length for well-formed files. */
		dref = bvo_outdegree(g, graph, graph_bytes, offsets, x - ref);
	}
	scratch_t t; memset(&t, 0, sizeof t);
	err = decode_tail(g, &s, x, d, ref, ref_list, dref, &t, *out, len_out);
	scratch_free(&t);
	free(ref_list);
	return err;
}

/* successors(x) into a caller buffer of `cap` longs; returns the outdegree or a negative error. */
int64_t bvo_successors(const bvo_params *g, const uint8_t *graph, uint64_t graph_bytes, const uint64_t *offsets,
                       int64_t x, int64_t *out, int64_t cap) {
	if (x < 0 || x >= g->nodes) return BVO_EINDEX; /* BVG:863 */
	int64_t *l = NULL, len = 0;
	const int err = successors_rec(g, graph, graph_bytes, offsets, x, &l, &len, 0);
	if (err) { free(l); return err; }
	if (len > cap) { free(l); return BVO_ECAPACITY; }
	memcpy(out, l, (size_t)len * sizeof(int64_t));
	free(l);
	return len;
}

/* ---------------------------------------------------------------------------------------------
 * Sequential scan.  BVGraphNodeIterator, BVG:1100-1245: a ring of window+1 successor lists indexed
 * by x mod (window+1); nodeIterator(from != 0) pre-fills the ring with the `window` predecessors by
 * random access (BVG:1139-1143) and positions the stream at offsets[from] (BVG:1144).
 * Output: deg_out[i] for node from+i, successors appended to col_out (capacity col_cap longs, may be
 * NULL to count only).  Returns the number of arcs decoded or a negative error.
 * --------------------------------------------------------------------------------------------- */
int64_t bvo_decode_range(const bvo_params *g, const uint8_t *graph, uint64_t graph_bytes, const uint64_t *offsets,
                         int64_t from, int64_t to, int64_t *deg_out, int64_t *col_out, int64_t col_cap) {
	if (from < 0 || from > g->nodes || to < from || to > g->nodes) return BVO_EINDEX; /* BVG:1128 */
	const int cyc = g->window + 1;
	vec_t *win = (vec_t *)calloc((size_t)cyc, sizeof(vec_t));
	int64_t *outd = (int64_t *)calloc((size_t)cyc, sizeof(int64_t));
	scratch_t t; memset(&t, 0, sizeof t);
	int64_t arcs = 0, rc = BVO_OK;
	ibs_t s; ibs_init(&s, graph, graph_bytes, 0);

	if (from != 0) {
		if (!offsets) { rc = BVO_ESTATE; goto done; } /* BVG:1136 */
		for (int i = 1; i < (from + 1 < cyc ? from + 1 : cyc); i++) {
			const int pos = (int)((from - i) % cyc);
			const int64_t d = bvo_outdegree(g, graph, graph_bytes, offsets, from - i);
			if (d < 0) { rc = d; goto done; }
			if (vec_need(&win[pos], d)) { rc = BVO_ECAPACITY; goto done; }
			const int64_t l = bvo_successors(g, graph, graph_bytes, offsets, from - i, win[pos].a, win[pos].cap);
			if (l < 0) { rc = l; goto done; }
			outd[pos] = d;
		}
		s.pos = offsets[from];
	}

	for (int64_t x = from; x < to; x++) {
		int err = 0;
		const int cur = (int)(x % cyc);
		const int64_t d = (int64_t)rd_outdegree(&s, g, &err); /* BVG:1010 */
		if (err) { rc = err; goto done; }
		if (s.overrun) { rc = BVO_EIO; goto done; }
		outd[cur] = d;
		if (deg_out) deg_out[x - from] = d;
		if (tl_shape) { tl_shape += 6; tl_shape[0] = tl_shape[1] = tl_shape[2] = tl_shape[3] = tl_shape[4] = 0; tl_shape[5] = 0; }
		if (vec_need(&win[cur], d)) { rc = BVO_ECAPACITY; goto done; }
		int64_t len = 0;
		if (d != 0) {
			const int ref = g->window > 0 ? (int)rd_reference(&s, g, &err) : -1; /* BVG:1015 */
			if (tl_shape) tl_shape[5] = ref > 0 ? ref : 0;
			if (err) { rc = err; goto done; }
			if (ref > 0 && x - ref < 0) { rc = BVO_ESTATE; goto done; }
			const int ri = (int)((x - ref + cyc) % cyc); /* BVG:1018 */
			err = decode_tail(g, &s, x, d, ref, ref > 0 ? win[ri].a : NULL, ref > 0 ? outd[ri] : 0, &t, win[cur].a, &len);
			if (err) { rc = err; goto done; }
			if (len != d) { rc = BVO_ESTATE; goto done; } /* the reference would store -1s here */
		}
		if (col_out) {
			if (arcs + d > col_cap) { rc = BVO_ECAPACITY; goto done; }
			memcpy(col_out + arcs, win[cur].a, (size_t)d * sizeof(int64_t));
		}
		arcs += d;
	}
	rc = arcs;
done:
	for (int i = 0; i < cyc; i++) free(win[i].a);
	free(win); free(outd);
	scratch_free(&t);
	return rc;
}

/* Batched random access on all host threads, the way SpeedTest -r calls successors(x) for random x
 * (src/it/unimi/dsi/big/webgraph/test/SpeedTest.java:95-108): thread t takes ids[t], ids[t + threads], ...
 * Returns the number of arcs read (or a negative error). */
typedef struct {
	const bvo_params *g; const uint8_t *graph; uint64_t graph_bytes; const uint64_t *offsets;
	const int64_t *ids; int64_t k; int t, threads; int64_t arcs; int64_t rc;
} ra_t;

static void *ra_main(void *p) {
	ra_t *w = (ra_t *)p;
	for (int64_t q = w->t; q < w->k; q += w->threads) {
		const int64_t x = w->ids[q];
		if (x < 0 || x >= w->g->nodes) { w->rc = BVO_EINDEX; return NULL; }
		int64_t *l = NULL, len = 0;
		const int err = successors_rec(w->g, w->graph, w->graph_bytes, w->offsets, x, &l, &len, 0);
		free(l);
		if (err) { w->rc = err; return NULL; }
		w->arcs += len;
	}
	return NULL;
}

int64_t bvo_random_access(const bvo_params *g, const uint8_t *graph, uint64_t graph_bytes, const uint64_t *offsets,
                          const int64_t *ids, int64_t k, int threads) {
	if (threads < 1) threads = 1;
	if (threads > 256) threads = 256;
	pthread_t th[256];
	ra_t w[256];
	for (int t = 0; t < threads; t++) {
		w[t] = (ra_t){g, graph, graph_bytes, offsets, ids, k, t, threads, 0, BVO_OK};
		pthread_create(&th[t], NULL, ra_main, &w[t]);
	}
	int64_t arcs = 0, rc = BVO_OK;
	for (int t = 0; t < threads; t++) {
		pthread_join(th[t], NULL);
		if (w[t].rc) rc = w[t].rc;
		arcs += w[t].arcs;
	}
	return rc ? rc : arcs;
}

/* Analysis helper: shape[6 * (x - from) ..] = ncopied, blocks, intervals, interval elements, residuals, reference of
 * every node of [from, to) (sequential scan).  shape must hold 6 * (to - from + 1) entries. */
int64_t bvo_row_shapes(const bvo_params *g, const uint8_t *graph, uint64_t graph_bytes, const uint64_t *offsets,
                       int64_t from, int64_t to, int64_t *deg_out, int64_t *shape) {
	tl_shape = shape - 6;
	const int64_t rc = bvo_decode_range(g, graph, graph_bytes, offsets, from, to, deg_out, NULL, 0);
	tl_shape = NULL;
	return rc;
}

/* ---------------------------------------------------------------------------------------------
 * Parallel scan: ImmutableGraph.splitNodeIterators(k), IG:405-436 -- contiguous ranges of
 * ceil(n/k) nodes, each scanned by its own iterator (ring pre-filled by random access).
 * deg_out has to-from entries; col_out, if not NULL, must already be laid out by the caller
 * through rowptr (rowptr[i] = position of node from+i's list; length to-from+1).  With
 * col_out == NULL only degrees and the arc count are produced (what SpeedTest measures).
 * --------------------------------------------------------------------------------------------- */
typedef struct {
	const bvo_params *g; const uint8_t *graph; uint64_t graph_bytes; const uint64_t *offsets;
	int64_t base, from, to; int64_t *deg_out; int64_t *col_out; const int64_t *rowptr; int64_t rc;
} split_t;

static void *split_main(void *p) {
	split_t *w = (split_t *)p;
	int64_t *col = NULL, cap = 0;
	if (w->col_out) { col = w->col_out + w->rowptr[w->from - w->base]; cap = w->rowptr[w->to - w->base] - w->rowptr[w->from - w->base]; }
	w->rc = bvo_decode_range(w->g, w->graph, w->graph_bytes, w->offsets, w->from, w->to,
	                         w->deg_out ? w->deg_out + (w->from - w->base) : NULL, col, cap);
	return NULL;
}

int64_t bvo_decode_split(const bvo_params *g, const uint8_t *graph, uint64_t graph_bytes, const uint64_t *offsets,
                         int64_t from, int64_t to, int how_many, int64_t *deg_out, int64_t *col_out, const int64_t *rowptr) {
	if (how_many < 1) how_many = 1;
	const int64_t n = to - from;
	const int64_t m = (n + how_many - 1) / how_many; /* IG:415 */
	split_t *w = (split_t *)calloc((size_t)how_many, sizeof(split_t));
	pthread_t *th = (pthread_t *)calloc((size_t)how_many, sizeof(pthread_t));
	int k = 0;
	for (int64_t a = from; a < to && k < how_many; a += m, k++) {
		w[k] = (split_t){ g, graph, graph_bytes, offsets, from, a, a + m < to ? a + m : to, deg_out, col_out, rowptr, 0 };
		pthread_create(&th[k], NULL, split_main, &w[k]);
	}
	int64_t arcs = 0, rc = 0;
	for (int i = 0; i < k; i++) {
		pthread_join(th[i], NULL);
		if (w[i].rc < 0) rc = w[i].rc; else arcs += w[i].rc;
	}
	free(w); free(th);
	return rc < 0 ? rc : arcs;
}

/* Exposed single-code readers so that the tests can pin each code on hand-made vectors. */
/* Arc labels stored as a bit stream: what BitStreamArcLabelledNodeIterator does for every node of [lo, hi)
 * (labelling/BitStreamArcLabelledImmutableGraph.java:487-560: position the label stream at the node's label offset,
 * then one label.fromBitStream() per successor, in successor order).  kind 0 = FixedWidthIntLabel (readInt(width),
 * FixedWidthIntLabel.java:70-73), kind 1 = GammaCodedIntLabel (readGamma, GammaCodedIntLabel.java:60-64).
 * label_offsets: the n+1 running sums of the .labeloffsets stream (same gamma-delta format as .offsets,
 * LabelOffsetsLongIterator :339-363, read with bvo_read_offsets).  degrees[i] = outdegree of node lo+i.
 * Returns the number of labels written to out, or -1 if a node's labels do not end where the next node's begin. */
int64_t bvo_decode_labels(int kind, int width, const uint8_t *labels, uint64_t labels_bytes, const uint64_t *label_offsets,
                          int64_t lo, int64_t hi, const int64_t *degrees, int64_t *out) {
	int64_t w = 0;
	for (int64_t x = lo; x < hi; x++) {
		ibs_t s; ibs_init(&s, labels, labels_bytes, label_offsets[x]);
		for (int64_t j = 0; j < degrees[x - lo]; j++) {
			const uint64_t v = kind == 0 ? (width ? ibs_bits(&s, width) : 0) : ibs_gamma(&s);
			out[w++] = (int64_t)(int32_t)(uint32_t)v;  /* the Java field is an int */
		}
		if (s.pos != label_offsets[x + 1]) return -1;
	}
	return w;
}

/* List-valued labels (FixedWidthIntListLabel.fromBitStream, labelling/FixedWidthIntListLabel.java:73-78; is_long:
 * FixedWidthLongListLabel.java:78-84): per arc readGamma() elements of readInt(width) / readLong(width).  lens[j] receives
 * the length of arc j's list; values (may be NULL: lengths only) the elements, `cap` of them at most.  Returns the number
 * of elements, -1 if a node's labels do not end at the next offset, -2 if cap is too small. */
int64_t bvo_decode_label_lists(int is_long, int width, const uint8_t *labels, uint64_t labels_bytes, const uint64_t *label_offsets,
                               int64_t lo, int64_t hi, const int64_t *degrees, int64_t *lens, int64_t *values, int64_t cap) {
	int64_t arc = 0, w = 0;
	for (int64_t x = lo; x < hi; x++) {
		ibs_t s; ibs_init(&s, labels, labels_bytes, label_offsets[x]);
		for (int64_t j = 0; j < degrees[x - lo]; j++) {
			const uint64_t len = ibs_gamma(&s);
			lens[arc++] = (int64_t)len;
			for (uint64_t t = 0; t < len; t++) {
				const uint64_t v = ibs_bits(&s, width);
				if (values) {
					if (w >= cap) return -2;
					values[w] = is_long ? (int64_t)v : (int64_t)(int32_t)(uint32_t)v;  /* long[] / int[] */
				}
				w++;
			}
		}
		if (s.pos != label_offsets[x + 1]) return -1;
	}
	return w;
}

uint64_t bvo_read_code(const uint8_t *p, uint64_t nbytes, uint64_t *bitpos, int coding, int k) {
	ibs_t s; ibs_init(&s, p, nbytes, *bitpos);
	uint64_t v;
	switch (coding) {
	case BVO_GAMMA: v = ibs_gamma(&s); break;
	case BVO_DELTA: v = ibs_delta(&s); break;
	case BVO_UNARY: v = ibs_unary(&s); break;
	case BVO_ZETA: v = ibs_zeta(&s, k); break;
	case BVO_NIBBLE: v = ibs_nibble(&s); break;
	case BVO_GOLOMB: v = ibs_golomb(&s, k); break;
	default: v = ~(uint64_t)0;
	}
	*bitpos = s.pos;
	return v;
}
