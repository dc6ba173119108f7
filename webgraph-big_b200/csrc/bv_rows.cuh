// bv_rows.cuh -- per-row decode steps shared by the bulk (CSR) and the random-access kernels.
//
// A row record is (BVGraph.java:995-1097 of the reference):
//   outdegree d | reference r | block count b, blocks B0..Bb-1 | interval count, (left,len)* | residuals
// The reference materialises it as a tree of lazy iterators (Masked / IntervalSequence / Residual / Merged).
// Here a row is produced in two data-parallel steps that need no intermediate storage:
//   1. row_extras: intervals and residuals are merged while they are decoded and written to the TAIL of the
//      row's own slot (positions d-e .. d-1, e = number of extras).  A row without a reference is complete.
//   2. row_copy: the copy blocks select runs of the (finished) reference row, which are merged with the extras
//      waiting in the tail, writing from position 0.  The write position can never overtake the read position
//      of the extras (out = copied_so_far + extras_consumed <= ncopied + extras_consumed = read index), so the
//      merge is in place.
//
// Warp-synchronous style.  One lane decodes one row, and rows differ in length and shape, so left to itself a warp
// drifts apart: lanes that leave a data-dependent loop early run ahead and from then on the warp executes the same
// code several times over with a few lanes each (the first versions of these kernels ran at 4-12 active lanes per
// instruction).  The serial loops below are therefore written so that all lanes named in `mask` iterate together
// (the loop condition is a warp vote) and every iteration has ONE decoding site, i.e. one instruction stream.
// Rules that keep this deadlock-free: `mask` comes from a __ballot_sync executed by the whole warp, every lane of
// `mask` calls the function, and there is no early return between votes (errors raise a flag and zero the work).
#pragma once
#include "bv_device.cuh"

namespace wgbv {

static constexpr unsigned FULL_MASK = 0xffffffffu;

// a prefetch may name any address: it never faults
__device__ __forceinline__ void prefetch_l1(const void *p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

struct RowHead {
	uint32_t d;        // outdegree
	uint32_t rf;       // reference (0: none)
	uint32_t bc;       // block count
	uint32_t ncopied;  // successors copied from the reference row
	uint64_t blocks_pos;  // bit position of the first block (after the block count)
	uint64_t cmask[2];    // bit k set: position k of the reference row is copied (complete only when dref <= 128)
	bool ok;           // false: inconsistent counts (an error bit has been raised); d and ncopied are zeroed
};

// bits [0, t) of a 128-bit mask
__device__ __forceinline__ void mask_prefix(uint64_t t, uint64_t &lo, uint64_t &hi) {
	lo = t >= 64 ? ~0ull : (1ull << t) - 1ull;
	hi = t <= 64 ? 0ull : t >= 128 ? ~0ull : (1ull << (t - 64)) - 1ull;
}

// The part of a row's head that follows the reference: block count and blocks (summed, and turned into the mask of
// copied positions); leaves the reader after the last block (at the interval section).  d and rf are the row's
// outdegree and reference, dref the outdegree of the reference row (only used when rf > 0).  BVG:1017-1033.
// The blocks alternate copy / skip starting with copy, so the copied positions are the XOR of the prefixes [0, c_k)
// over the running block ends c_k -- plus [0, dref) when the count is even, i.e. when the tail after the last block
// is copied too.  (Positions from 128 on are dropped: the mask is complete only when dref <= 128.)
template <int DEF, class RD>
__device__ __forceinline__ void row_blocks(const GraphDev &g, RD &r, uint32_t d, uint32_t rf, uint32_t dref, RowHead &h) {
	h.d = d;
	h.rf = rf;
	h.bc = 0;
	h.ncopied = 0;
	h.blocks_pos = 0;
	h.cmask[0] = h.cmask[1] = 0;
	h.ok = true;
	if (h.rf > 0) {
		uint64_t bc = rd_blockcount<DEF>(r, g);
		if (bc > (uint64_t)dref + 1) { h.ok = false; bc = 0; }
		h.bc = (uint32_t)bc;
		h.blocks_pos = r.tell();
		uint64_t copied = 0, total = 0, m0 = 0, m1 = 0;
		for (uint32_t i = 0; i < h.bc; i++) {
			const uint64_t b = rd_block<DEF>(r, g) + (i ? 1 : 0);
			if (!(i & 1)) copied += b;
			total += b;
			uint64_t lo, hi;
			mask_prefix(total, lo, hi);
			m0 ^= lo;
			m1 ^= hi;
		}
		if (total > dref) h.ok = false;
		else if (!(h.bc & 1)) {
			copied += dref - total;
			uint64_t lo, hi;
			mask_prefix(dref, lo, hi);
			m0 ^= lo;
			m1 ^= hi;
		}
		if (copied > d) h.ok = false;
		h.ncopied = (uint32_t)copied;
		h.cmask[0] = m0;
		h.cmask[1] = m1;
	}
	if (!h.ok) {
		raise(g, ERR_DEGREE);
		h.d = 0;
		h.ncopied = 0;
		h.bc = 0;
		h.cmask[0] = h.cmask[1] = 0;
	}
}

// Reads outdegree, reference, block count and sums the blocks; leaves the reader after the last block (at the
// interval section).  dref = outdegree of the reference row (only used when rf > 0).  BVG:1010-1033.
template <int DEF, class RD>
__device__ __forceinline__ void row_head(const GraphDev &g, RD &r, uint32_t dref, RowHead &h) {
	const uint64_t d = rd_outdegree<DEF>(r, g);
	uint32_t rf = 0;
	if (d != 0 && g.window > 0) rf = (uint32_t)rd_reference<DEF>(r, g);
	row_blocks<DEF>(g, r, (uint32_t)d, rf, dref, h);
}

// No-op window policy of row_extras_core: the residual reader needs no refilling.
struct NoWindow {
	static constexpr bool enabled = false;
	template <class RD> __device__ __forceinline__ bool low(const RD &) const { return false; }
	template <class RD> __device__ __forceinline__ void refill(RD &, unsigned) {}
};

// First pass over the interval section (reader at the interval count): number of intervals, where they start, and
// how many of the `nextra` extras are residuals.  Leaves the reader at the first residual.  BVG:1037-1060.
template <class RD>
__device__ __forceinline__ void row_intervals(const GraphDev &g, RD &r, uint32_t &nextra, uint32_t &ic, uint64_t &int_pos, uint32_t &nres) {
	ic = 0;
	int_pos = 0;
	nres = nextra;
	if (nextra && g.min_interval != 0) {
		uint64_t c = r.gamma();
		if (c > nextra) { raise(g, ERR_DEGREE); c = 0; nextra = 0; }
		ic = (uint32_t)c;
		int_pos = r.tell();
		uint64_t covered = 0;
		for (uint32_t i = 0; i < ic; i++) {
			r.gamma();
			covered += r.gamma() + (uint64_t)g.min_interval;
		}
		if (covered > nextra) { raise(g, ERR_DEGREE); ic = 0; nextra = 0; covered = 0; }
		nres = nextra - (uint32_t)covered;
	}
}

// Element types of a row store.  Successors are int64 at the C ABI; the fused tile kernel keeps rows of graphs with
// fewer than 2^32 - 1 nodes as uint32 in shared memory (half the bytes, 32-bit arithmetic in the serial loops) and
// widens them on the way out.  ElemMax<E> is a sentinel larger than every node id representable in E's mode.
template <class E> struct ElemTraits;
template <> struct ElemTraits<int64_t> {
	using V = int64_t;
	static constexpr int64_t MAX = INT64_MAX;
};
template <> struct ElemTraits<uint32_t> {
	using V = uint32_t;
	static constexpr uint32_t MAX = 0xffffffffu;
};

template <class T> struct IsPosReader { static constexpr bool value = false; };
template <> struct IsPosReader<PosReader> { static constexpr bool value = true; };

// The pure gap loop of row_extras_core for zeta_3 codes in staged words: the stream runs through a 64-bit register
// window (the next 32..63 bits, left-aligned), so the chain from one code to the next is count-leading-zeros, two
// shifts, an add and the window shift -- no load on it -- and the word that tops the window up is loaded ahead.
// Codes of more than 32 bits (gaps of 2^24 and more) leave through the generic reader.  Lock step as the loop it
// replaces: every lane of `mask` calls it; on return r stands after the last code read.  (Measured: four codes per
// vote instead of one made the step slower, 8.9 -> 11.4 ms per 2e7 nodes: the lanes of a warp stop at different counts.)
template <class V, class E>
__device__ __forceinline__ void gap_loop_zeta3(PosReader &r, V &res, uint32_t &o, uint32_t nextra, E *out, unsigned mask) {
	const uint32_t *w = r.base;
	const uint32_t lastw = r.last + 1;
	uint32_t idx = r.p >> 5;
	idx = idx < r.last ? idx : r.last;
	uint32_t sh = r.p & 31u;
	uint64_t win = (((uint64_t)w[idx] << 32) | w[idx + 1]) << sh;
	uint32_t have = 64u - sh;  // valid bits in win: 32..64 at the top of every iteration
	uint32_t nx = idx + 2;     // next word to load
	for (; __any_sync(mask, o < nextra); o++) {
		if (o < nextra) {
			const uint32_t t = (uint32_t)(win >> 32);
			const int h = __clz((int)t);
			if (h <= 7) {
				const uint32_t x1 = t << (h + 1);
				const bool longer = (x1 >> 30) != 0;  // the two leading payload bits: not both zero = one more bit follows
				const uint32_t m2 = x1 >> (29 - 3 * h);
				const uint32_t val = longer ? m2 - 1u : (m2 >> 1) + (1u << (3 * h)) - 1u;
				const uint32_t len = 4u * (uint32_t)h + 3u + (longer ? 1u : 0u);
				res = (V)(res + 1 + (V)val);
				out[o] = (E)res;
				win <<= len;
				have -= len;
				const uint32_t wi = nx < lastw ? nx : lastw;
				const uint32_t nw32 = w[wi];
				if (have < 32u) {
					win |= (uint64_t)nw32 << (32u - have);
					have += 32u;
					nx++;
				}
			} else {
				r.p = nx * 32u - have;
				res = (V)(res + 1 + (V)r.zeta(3));
				out[o] = (E)res;
				idx = r.p >> 5;
				idx = idx < r.last ? idx : r.last;
				sh = r.p & 31u;
				win = (((uint64_t)w[idx] << 32) | w[idx + 1]) << sh;
				have = 64u - sh;
				nx = idx + 2;
			}
		}
	}
	r.p = nx * 32u - have;
}

// The merge of intervals (cursor ri, `ic` of them) and residuals (cursor r, `nres` of them) into out[0..nextra).
// Warp-synchronous: every lane of `mask` must call it (nextra == 0: no work).  Three parts: a merging loop that runs
// while some lane of the warp still has interval material, then (one site, all lanes together) the first residual of
// rows that have not decoded any yet -- its code is a signed offset from x and may need 64 bits even when node ids fit
// 32 -- and then a tight loop of pure gap decoding in E's arithmetic (the kernels sort rows so that most warps never
// enter the first loop).  WIN lets the residual reader run on a small window of the stream that the warp refills
// together (see k_extras_heavy); with NoWindow that code vanishes.
template <int DEF, class RDI, class RDR, class WIN, class E = int64_t>
__device__ __forceinline__ void row_extras_core(const GraphDev &g, RDI &ri, RDR &r, int64_t x, uint32_t nextra, uint32_t ic, uint32_t nres,
                                                E *out, unsigned mask, WIN &win) {
	using V = typename ElemTraits<E>::V;
	V cur = 0, end = 0;  // current interval [cur, end)
	V res = (V)x;        // last residual decoded; pending (not yet written) iff have_res
	bool have_res = false, first_res = true, first_int = true;
	uint32_t o = 0;
	for (; __any_sync(mask, o < nextra && (cur != end || ic != 0)); o++) {
		if (WIN::enabled && __any_sync(mask, win.low(r))) win.refill(r, mask);
		if (o < nextra) {
			if (cur == end && ic) {  // next interval (rare: once per interval)
				const uint64_t gap = ri.gamma();
				cur = first_int ? (V)(x + nat2int(gap)) : (V)(end + 1 + (V)gap);
				first_int = false;
				end = (V)(cur + (V)ri.gamma() + (V)g.min_interval);
				ic--;
			}
			if (!have_res && nres) {
				const uint64_t code = rd_residual<DEF>(r, g);
				res = first_res ? (V)(x + nat2int(code)) : (V)(res + 1 + (V)code);
				first_res = false;
				have_res = true;
				nres--;
			}
			const bool take_int = cur != end && (!have_res || cur < res);
			out[o] = (E)(take_int ? cur : res);
			if (take_int) cur++;
			else have_res = false;
		}
	}
	if (have_res && o < nextra) out[o++] = (E)res;  // a residual decoded above but not written yet
	if (__any_sync(mask, first_res && o < nextra)) {
		if (WIN::enabled && __any_sync(mask, win.low(r))) win.refill(r, mask);
		if (first_res && o < nextra) {
			res = (V)(x + nat2int(rd_residual<DEF>(r, g)));
			out[o++] = (E)res;
		}
	}
	if constexpr (DEF == 2 && IsPosReader<RDR>::value && !WIN::enabled) {
		gap_loop_zeta3(r, res, o, nextra, out, mask);
	} else {
		for (; __any_sync(mask, o < nextra); o++) {
			if (WIN::enabled && __any_sync(mask, win.low(r))) win.refill(r, mask);
			if (o < nextra) {
				res = (V)(res + 1 + (V)rd_residual<DEF>(r, g));
				out[o] = (E)res;
			}
		}
	}
}

// Decodes the interval section and the residuals of row x (reader positioned at the interval count, i.e. right
// after the blocks) and writes the `nextra` merged values, in increasing order, to out[0..nextra).
// BVG:1037-1073, 902-954.  Warp-synchronous: every lane of `mask` must call it (nextra == 0: no work).
template <int DEF, class RD, class E = int64_t>
__device__ __forceinline__ void row_extras(const GraphDev &g, RD &r, int64_t x, uint32_t nextra, E *out, unsigned mask) {
	uint32_t ic, nres;
	uint64_t int_pos;
	row_intervals(g, r, nextra, ic, int_pos, nres);
	RD ri = r;  // interval cursor; r stays the residual cursor
	if (ic) ri.seek(int_pos);
	NoWindow nw;
	row_extras_core<DEF>(g, ri, r, x, nextra, ic, nres, out, mask, nw);
	if (nextra && r.past_end()) raise(g, ERR_OVERRUN);
}

// Copy-block resolution of one row: src[0..dref) is the finished reference row, ext[0..d-ncopied) the row's extras
// (in increasing order) and dst[0..d) the row's slot.  In place operation is allowed and is the normal case:
// ext == dst + ncopied, i.e. the extras wait in the tail of the slot; the write position can never overtake the read
// position of the extras.  Extras that follow the last copied successor are NOT written: they are expected to sit in
// dst's tail already (true in place, and also when ext is a staged copy of that tail).  The reader is positioned at
// the first block.  MaskedLongIterator.java:67-100 + MergedLongIterator.java:54-92 of the reference.
// Warp-synchronous: every lane of `mask` must call it (ncopied == 0 for lanes without work).
template <int DEF, class RD>
__device__ __forceinline__ void row_copy(const GraphDev &g, RD &r, const int64_t *src, uint32_t dref, const int64_t *ext, int64_t *dst,
                                         uint32_t d, uint32_t bc, uint32_t ncopied, unsigned mask) {
	const uint32_t ne = d - ncopied;
	uint32_t ei = 0;  // extras consumed
	uint32_t out = 0, k = 0, run = 0, blk = 0, left = ncopied;
	int64_t e = (left && ne) ? ext[0] : INT64_MAX;
	// An iteration either starts the next block or emits exactly one successor (single block-decoding site).
	while (__any_sync(mask, left != 0)) {
		if (left) {
			if (run == 0) {
				if (blk > bc) left = 0;  // malformed: fewer successors to copy than announced
				else {
					// blocks alternate copy / skip starting with copy; after the last one the tail is copied iff bc is even
					uint32_t len = blk < bc ? (uint32_t)rd_block<DEF>(r, g) + (blk ? 1u : 0u) : dref - k;
					if (len > dref - k) len = dref - k;
					if (blk & 1) k += len;
					else run = len;
					blk++;
				}
			} else {
				const int64_t c = src[k];
				// both streams are read one element at a time and each new 32-byte sector would cost a full memory round trip:
				// ask for the sectors two ahead (no register, no wait)
				if ((k & 3u) == 0 && k + 8 < dref) prefetch_l1(src + k + 8);
				if (e < c) {
					dst[out++] = e;
					if ((ei & 3u) == 0 && ei + 8 < ne) prefetch_l1(ext + ei + 8);
					e = ++ei < ne ? ext[ei] : INT64_MAX;
				} else {
					if (e == c) raise(g, ERR_NOT_SORTED);
					dst[out++] = c;
					k++;
					run--;
					left--;
				}
			}
		}
	}
}

// The one-lane merge for a reference row of at most 128 successors, whose copy blocks stage 3a has already turned into
// a bit mask (m, m1) of copied positions (RowHead.cmask): no block is decoded here and positions that are not copied
// are never visited, so every iteration emits exactly one successor -- the smaller of the next extra and the next
// copied element -- through one instruction stream.  ext[0..ne) are the extras; in place operation as in row_copy.
// Warp-synchronous: every lane of `mask` must call it (m == 0 for lanes without work).
// cg (warp-uniform): read the reference row past the L1 (it was written by another block of the running kernel).
template <class E = int64_t>
__device__ __forceinline__ void row_copy_masked(const GraphDev &g, const E *src, uint64_t m, uint64_t m1, const E *ext, E *dst, uint32_t ne,
                                                unsigned mask, bool cg = false) {
	constexpr bool GLOBAL = sizeof(E) == 8;  // int64 rows live in global memory (prefetches, optional ld.cg); uint32 rows in shared memory
	const E EMAX = ElemTraits<E>::MAX;
	uint32_t ei = 0;
	if (!m) { m = m1; m1 = 0; src += 64; }
	int pos = m ? __ffsll((long long)m) - 1 : 0;
	E e = (m && ne) ? ext[0] : EMAX;
	E c = m ? ((GLOBAL && cg) ? __ldcg(src + pos) : src[pos]) : 0;
	while (__any_sync(mask, m != 0)) {
		if (m) {
			const bool take_e = e < c;
			if (e == c) raise(g, ERR_NOT_SORTED);
			*dst++ = take_e ? e : c;
			if (take_e) {
				// each new 32-byte sector of a stream would cost a memory round trip: ask for the one two ahead
				if (GLOBAL && (ei & 3u) == 0 && ei + 8 < ne) prefetch_l1(ext + ei + 8);
				e = ++ei < ne ? ext[ei] : EMAX;
			} else {
				m &= m - 1;
				if (!m && m1) { m = m1; m1 = 0; src += 64; pos -= 64; }
				if (m) {
					const int q = __ffsll((long long)m) - 1;
					if (GLOBAL && !cg && (q >> 2) != (pos >> 2)) prefetch_l1(src + (q & ~3) + 8);
					pos = q;
					c = (GLOBAL && cg) ? __ldcg(src + q) : src[q];
				}
			}
		}
	}
}

// The same merge for rows kept in shared memory, written so that an iteration is one straight instruction stream:
// both candidates are loaded (a shared-memory load costs less than a divergent branch), the smaller one is stored,
// and the cursor that supplied it advances by predication.  Ends when the copied successors are exhausted: extras
// after the last copied one are in place already.  Every lane of `mask` must call it (m == m1 == 0: no work).
template <class E>
__device__ __forceinline__ void row_merge_lean(const GraphDev &g, const E *src, uint64_t m, uint64_t m1, const E *ext, E *dst, uint32_t ne,
                                               unsigned mask) {
	const E EMAX = ElemTraits<E>::MAX;
	uint32_t ei = 0, base = 0;
	if (!m) { m = m1; m1 = 0; base = 64; }
	bool dup = false;
	while (__any_sync(mask, m != 0)) {
		if (m) {
			const uint32_t pos = base + (uint32_t)__ffsll((long long)m) - 1u;
			const E c = src[pos];
			const E e = ei < ne ? ext[ei] : EMAX;
			const bool take_e = e < c;
			dup |= e == c;
			*dst++ = take_e ? e : c;
			ei += take_e ? 1u : 0u;
			const uint64_t mn = take_e ? m : (m & (m - 1));
			const bool sw = mn == 0;
			m = sw ? m1 : mn;
			m1 = sw ? 0ull : m1;
			base = sw ? 64u : base;
		}
	}
	if (dup) raise(g, ERR_NOT_SORTED);
}

// The merge of a row's copied successors with its extras by G lanes (G = 8 or 32) through merge path: no lane walks
// the whole row.  The row lives in shared memory: src is the finished reference row (at most 128 successors, (m0, m1) the
// mask of its copied positions), dst[0..d) the row's slot with the extras waiting in its tail dst[nc..d).
//   1. the copied successors are compacted into dst[0..nc), every lane taking 128 / G positions of the mask;
//   2. lane `part` produces outputs [part * q, part * q + q), q = ceil(d / G) <= 8: a binary search finds how many of
//      the first part * q outputs come from the copied run, a short serial merge fills the lane's registers;
//   3. after all lanes have read, they store.
// The whole warp must call it together (lanes without a row: active = false); rows of different groups do not interact.
// MaskedLongIterator.java:67-100 + MergedLongIterator.java:54-92 of the reference, as one data-parallel step.
template <int G, class E>
__device__ __forceinline__ void group_merge(const GraphDev &g, bool active, const E *src, uint64_t m0, uint64_t m1, E *dst, uint32_t d, uint32_t nc,
                                            uint32_t part) {
	static_assert(G == 8 || G == 32, "group sizes with 128 / G positions inside one mask word");
	constexpr uint32_t W = 128 / G;
	const E EMAX = ElemTraits<E>::MAX;
	if (active) {
		const uint32_t lo_bit = part * W;
		const uint64_t word = lo_bit < 64 ? m0 : m1;
		uint32_t slice = (uint32_t)(word >> (lo_bit & 63u)) & ((1u << W) - 1u);
		const uint32_t before = lo_bit < 64 ? (uint32_t)__popcll(m0 & ((1ull << lo_bit) - 1ull))
		                                    : (uint32_t)(__popcll(m0) + __popcll(m1 & ((1ull << (lo_bit - 64)) - 1ull)));
		E *o = dst + before;
		while (slice) {
			const uint32_t b = (uint32_t)__ffs((int)slice) - 1u;
			slice &= slice - 1u;
			*o++ = src[lo_bit + b];
		}
	}
	__syncwarp();
	E outv[8];
	uint32_t k0 = 0, cnt = 0;
	if (active) {
		const uint32_t q = (d + G - 1) / G;
		const uint32_t ne = d - nc;
		k0 = part * q;
		if (k0 < d) {
			cnt = d - k0 < q ? d - k0 : q;
			const E *A = dst, *B = dst + nc;
			// i: outputs before k0 that come from A
			uint32_t lo = k0 > ne ? k0 - ne : 0u, hi = k0 < nc ? k0 : nc;
			while (lo < hi) {
				const uint32_t i = (lo + hi) >> 1, j = k0 - i;
				if (j > 0 && B[j - 1] > A[i]) lo = i + 1;
				else hi = i;
			}
			uint32_t i = lo, j = k0 - lo;
			E a = i < nc ? A[i] : EMAX, b = j < ne ? B[j] : EMAX;
			bool dup = false;
#pragma unroll
			for (int t = 0; t < 8; t++) {
				if ((uint32_t)t < cnt) {
					const bool ta = a < b;
					dup |= a == b;
					outv[t] = ta ? a : b;
					i += ta ? 1u : 0u;
					j += ta ? 0u : 1u;
					const E na = i < nc ? A[i] : EMAX, nb = j < ne ? B[j] : EMAX;
					a = ta ? na : a;
					b = ta ? b : nb;
				}
			}
			if (dup) raise(g, ERR_NOT_SORTED);
		}
	}
	__syncwarp();
#pragma unroll
	for (int t = 0; t < 8; t++)
		if ((uint32_t)t < cnt) dst[k0 + t] = outv[t];
	__syncwarp();
}

// position of the n-th (0-based) set bit of m; n < popc(m)
__device__ __forceinline__ uint32_t nth_set_bit(uint32_t m, uint32_t n) {
	uint32_t pos = 0, t;
	t = (uint32_t)__popc(m & 0xffffu);
	if (n >= t) { n -= t; pos += 16; m >>= 16; }
	t = (uint32_t)__popc(m & 0xffu);
	if (n >= t) { n -= t; pos += 8; m >>= 8; }
	t = (uint32_t)__popc(m & 0xfu);
	if (n >= t) { n -= t; pos += 4; m >>= 4; }
	t = (uint32_t)__popc(m & 0x3u);
	if (n >= t) { n -= t; pos += 2; m >>= 2; }
	if (n >= (m & 1u)) pos += 1;
	return pos;
}

// The same merge done by a whole warp on ONE row (all 32 lanes pass the same arguments and each keeps its own copy of
// the block reader, so the block codes are decoded redundantly but uniformly).  The reference row is read 32
// positions at a time and the extras 32 at a time, both coalesced; every round emits everything that is not larger
// than the last element of either window, each element going straight to its final position
// (rank among its own kind + rank among the other kind), so the serial dependency of the one-lane merge -- and with
// it the long tail of very long rows -- disappears.  In place operation (ext == dst + ncopied) is safe for the same
// reason as in row_copy: a round writes below the position of the first extra it has not loaded yet.
template <int DEF, class RD, class E = int64_t>
__device__ __forceinline__ void row_copy_warp(const GraphDev &g, RD &r, const E *src, uint32_t dref, const E *ext, E *dst, uint32_t d,
                                              uint32_t bc, uint32_t ncopied, bool cg = false) {
	constexpr bool GLOBAL = sizeof(E) == 8;
	const E EMAX = ElemTraits<E>::MAX;
	const unsigned lane = threadIdx.x & 31u;
	const uint32_t ne = d - ncopied;
	uint32_t kb = 0;           // first position of the current window of the reference row
	uint32_t ei = 0;           // extras emitted so far
	uint32_t done = 0;         // copied successors emitted so far
	// block state: the block [bs, be) of reference positions currently open, whether it copies, blocks decoded so far
	uint32_t bs = 0, be = 0, blk = 0;
	bool bcopy = false;
	unsigned cmask = 0;        // copied positions of the current window not emitted yet
	E c = 0;                   // my reference element of the current window
	bool fresh = true;         // the window at kb has not been loaded yet
	while (done < ncopied) {
		if (fresh) {
			const uint32_t p = kb + lane;
			c = p < dref ? ((GLOBAL && cg) ? __ldcg(src + p) : src[p]) : EMAX;
			bool f = false;
			// walk the blocks that overlap [kb, kb + 32)
			for (;;) {
				if (p >= bs && p < be && bcopy) f = true;
				if (be >= kb + 32 || be >= dref || blk > bc) break;
				const uint32_t len = blk < bc ? (uint32_t)rd_block<DEF>(r, g) + (blk ? 1u : 0u) : dref - be;
				bcopy = !(blk & 1);
				blk++;
				bs = be;
				be = len > dref - be ? dref : be + len;
			}
			cmask = __ballot_sync(FULL_MASK, f && p < dref);
			fresh = false;
			if (kb + 32 >= dref && cmask == 0) break;  // malformed: announced copies that do not exist
			if (cmask == 0) {
				// nothing copied in this window; a skip block that reaches far beyond it (a few successors taken from a giant
				// reference row) is jumped over in one step
				kb = (!bcopy && be > kb + 32) ? (be & ~31u) : kb + 32;
				fresh = true;
				continue;
			}
		}
		// A: the copied elements of the window still to emit, compacted to lanes 0..nA-1
		const unsigned nA = (unsigned)__popc(cmask);
		const int srcl = lane < nA ? (int)nth_set_bit(cmask, lane) : 0;
		E A = __shfl_sync(FULL_MASK, c, srcl);
		if (lane >= nA) A = EMAX;
		// B: the next extras
		const uint32_t nB = ne - ei < 32u ? ne - ei : 32u;
		const E B = lane < nB ? ext[ei + lane] : EMAX;
		const bool src_last = kb + 32 >= dref;   // nothing copied follows this window
		const bool ext_last = ei + nB >= ne;     // no extra follows this window
		const E a_last = __shfl_sync(FULL_MASK, A, (int)nA - 1);
		const E b_last = nB ? __shfl_sync(FULL_MASK, B, (int)nB - 1) : (E)0;
		const E lim_a = src_last ? EMAX : a_last;
		const E lim_b = ext_last ? EMAX : b_last;  // (no extras at all: ext_last holds, the limit is not used)
		const E limit = lim_a < lim_b ? lim_a : lim_b;
		// ranks in the other window by binary search over lanes (both windows are sorted, padded with +inf)
		uint32_t a_in_b = 0, b_in_a = 0;  // #B < A_lane, #A < B_lane
#pragma unroll
		for (int step = 16; step; step >>= 1) {
			const E vb = __shfl_sync(FULL_MASK, B, (int)(a_in_b + step - 1));
			const E va = __shfl_sync(FULL_MASK, A, (int)(b_in_a + step - 1));
			if (vb < A) a_in_b += step;
			if (va < B) b_in_a += step;
		}
		if (__shfl_sync(FULL_MASK, B, (int)a_in_b) < A && a_in_b < 32) a_in_b++;
		if (__shfl_sync(FULL_MASK, A, (int)b_in_a) < B && b_in_a < 32) b_in_a++;
		const bool emitA = lane < nA && A <= limit;
		const bool emitB = lane < nB && B <= limit;
		const uint32_t out = done + ei;
		if (emitA) dst[out + lane + a_in_b] = A;
		if (emitB) dst[out + lane + b_in_a] = B;
		const unsigned eA = (unsigned)__popc(__ballot_sync(FULL_MASK, emitA));
		const unsigned eB = (unsigned)__popc(__ballot_sync(FULL_MASK, emitB));
		if (eA == 0 && eB == 0) break;  // cannot happen on well-formed input
		done += eA;
		ei += eB;
		// drop the emitted copied positions (the eA lowest set bits of cmask): lane eA-1 knows where the last of them is
		if (eA) cmask &= ~((2u << __shfl_sync(FULL_MASK, srcl, (int)eA - 1)) - 1u);
		if (cmask == 0) { kb += 32; fresh = true; }
		__syncwarp();
	}
}


// ------------------------------------------------------------------------------------------------ whole-warp rows
// A long residual list decoded by a whole warp.  The codes of a row are a serial chain -- where a code starts is only
// known once the one before it has been read -- and a row of hundreds (or, in the reference's slow test,
// BVGraphSlowTest.java:30-96, a billion) residuals would keep one lane busy while everything that depends on the
// row waits.  But the bit range of the list is known exactly (it ends where the next record starts), and the codes
// resynchronise: a parse started at a wrong position falls in step with the true one within a few codes.  So the
// range is cut into segments of COOP_SEG bits, one per lane.  Every lane parses its segment from a guessed entry
// (the segment's first bit; lane 0 knows the truth) and reports where its parse leaves the segment; that exit is
// the next lane's entry, so lanes whose entry changed parse again, until no entry changes (two or three rounds in
// practice, lane i is certainly right after round i + 1).  A scan over the lanes' code counts and gap sums then gives
// every lane the index and the value its first code starts from, and a last parse writes the successors.  All 32
// lanes pass the same arguments; proto is a reader on the stream (any position), [p0, p1) the bit range of the n codes.
// BVG:902-954 (ResidualLongIterator) is what this replaces.
#ifndef WGBV_COOP_SEG
#define WGBV_COOP_SEG 512
#endif
#ifndef WGBV_COOP_LONG
#define WGBV_COOP_LONG 128
#endif
// Shortest segment: a misaligned zeta_3 parse of 16-bit codes falls in step after 112 bits on average (90 % within 260,
// 99 % within 500: measured on the benchmark graph), and a lane that does not resynchronise inside its segment costs
// the lanes after it another round.
static constexpr uint32_t COOP_SEG_MIN = WGBV_COOP_SEG, COOP_SEG_MAX = 4096;
// rows of at least this many residuals are decoded by a whole warp (rows with intervals only if they have at most 32)
static constexpr uint32_t COOP_LONG = WGBV_COOP_LONG;

// The general form: decodes the codes that START in [p0, p1) -- p0 a true code start -- as the codes idx0, idx0 + 1, ... of
// the row, the first of them building on the value `base` (first_code: p0 is the start of the whole list, whose
// first code is a signed offset from x; base is then worked out here).  WRITE = false only counts: `count` codes, the
// sum `total` of (gap + 1) over them, and `exit`, the start of the first code at or after p1.  nmax bounds the indices
// written.  Giant rows are cut into chunks that whole warps scan, and later write, side by side (k_giant_*).
template <int DEF, bool WRITE, class RD, class E>
__device__ __forceinline__ void coop_residuals_ex(const GraphDev &g, const RD &proto, typename RD::pos_t p0, typename RD::pos_t p1, uint32_t nmax,
                                                  int64_t x, E *out, uint32_t idx0, typename ElemTraits<E>::V base, bool first_code, uint32_t &count,
                                                  typename ElemTraits<E>::V &total_out, typename RD::pos_t &exit_out) {
	using V = typename ElemTraits<E>::V;
	using P = typename RD::pos_t;
	const unsigned lane = threadIdx.x & 31u;
	P wstart = p0;         // true position of the first code not handled yet
	uint32_t done = idx0;  // index of that code
	V acc = 0;
	// segment length: the whole range in one window when it is short, long segments (fewer rounds per bit) when it is huge
	const uint64_t span = p1 > p0 ? (uint64_t)(p1 - p0) : 0;
	uint64_t sg = (span + 31) / 32;
	sg = (sg + 31) & ~31ull;
	const P seg = (P)(sg < COOP_SEG_MIN ? COOP_SEG_MIN : sg > COOP_SEG_MAX ? COOP_SEG_MAX : sg);
	while (wstart < p1 && (!WRITE || done < nmax)) {
		const P seg_lo = wstart + seg * lane, seg_hi = seg_lo + seg;
		const bool active = seg_lo < p1;
		const P lim = seg_hi < p1 ? seg_hi : p1;
		P entry = seg_lo, exit = seg_lo;
		uint32_t cnt = 0;
		V total = 0;  // sum of (gap + 1) over my codes
		bool dirty = active;
		for (int round = 0; round < 34; round++) {
			if (dirty) {
				RD r = proto;
				r.seek(entry);
				P p = entry;
				cnt = 0;
				total = 0;
				while (p < lim) {
					const uint64_t code = rd_residual<DEF>(r, g);
					total += (V)code + 1;
					cnt++;
					const P q = (P)r.tell();
					if (q <= p) { p = lim; break; }  // (cannot happen: every code has at least one bit)
					p = q;
				}
				exit = p > entry ? p : entry;
			}
			P prev = __shfl_up_sync(FULL_MASK, exit, 1);
			if (lane == 0) prev = wstart;
			dirty = active && prev != entry;
			if (dirty) entry = prev;
			if (!__any_sync(FULL_MASK, dirty)) break;
		}
		if (!active) { cnt = 0; total = 0; }
		// The row's first residual is a signed offset from x (BVG:917), the others are gaps (BVG:929).  The scans below
		// treat every code as a gap; the starting value is chosen so that the first one comes out right all the same.
		if (first_code && done == 0) {
			V b0 = 0;
			if (lane == 0) {
				RD r = proto;
				r.seek(wstart);
				const uint64_t code = rd_residual<DEF>(r, g);
				b0 = (V)(x + nat2int(code)) - ((V)code + 1);
			}
			base = __shfl_sync(FULL_MASK, b0, 0);
		}
		// index and value every lane starts from (inclusive scans over the lanes)
		uint32_t ci = cnt;
		V ti = total;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			const uint32_t c2 = __shfl_up_sync(FULL_MASK, ci, o);
			const V t2 = __shfl_up_sync(FULL_MASK, ti, o);
			if ((int)lane >= o) { ci += c2; ti += t2; }
		}
		const uint32_t wcount = __shfl_sync(FULL_MASK, ci, 31);
		const V wtotal = __shfl_sync(FULL_MASK, ti, 31);
		// the last parse: write
		if (WRITE && cnt) {
			uint32_t idx = done + ci - cnt;
			V val = base + (ti - total);
			RD r = proto;
			r.seek(entry);
			for (uint32_t k = 0; k < cnt; k++) {
				val += (V)rd_residual<DEF>(r, g) + 1;
				if (idx < nmax) out[idx] = (E)val;
				idx++;
			}
		}
		base += wtotal;
		acc += wtotal;
		done += wcount;
		// the window ends where its last active lane's parse ended
		const unsigned am = __ballot_sync(FULL_MASK, active);
		const int last = 31 - __clz((int)am);
		const P wend = __shfl_sync(FULL_MASK, exit, last);
		if (wcount == 0 || wend <= wstart) { wstart = p1; break; }  // malformed: no progress
		wstart = wend;
	}
	count = done - idx0;
	total_out = acc;
	exit_out = wstart;
}

template <int DEF, class RD, class E>
__device__ __forceinline__ void coop_residuals(const GraphDev &g, const RD &proto, typename RD::pos_t p0, typename RD::pos_t p1, uint32_t n,
                                               int64_t x, E *out) {
	uint32_t count;
	typename ElemTraits<E>::V total;
	typename RD::pos_t exit;
	coop_residuals_ex<DEF, true>(g, proto, p0, p1, n, x, out, 0u, (typename ElemTraits<E>::V)x, true, count, total, exit);
	if (count < n) raise(g, ERR_DEGREE);  // the record holds fewer codes than announced
}

// The intervals of a row merged into its residuals by a whole warp (all lanes pass the same arguments).  ext[0..ne) is
// the extras part of the row's slot; its last nres entries hold the residuals, in increasing order; the reader is
// positioned at the interval count and the row has at most 32 intervals.  Lane k keeps interval k.  Residuals move
// down to their final places first (in chunks, in increasing order: a target never lies above its source), then the
// interval elements fill the holes.  In files the reference writes no residual lies inside an interval
// (intervalize, BVG:1595-1618), so an interval goes in as one piece.  BVG:1037-1073, LongIntervalSequenceIterator.java:57-77.
template <class RD, class E>
__device__ __forceinline__ void coop_intervals_merge(const GraphDev &g, const RD &proto, typename RD::pos_t ipos, int64_t x, E *ext, uint32_t ne,
                                                     uint32_t nres) {
	using V = typename ElemTraits<E>::V;
	const unsigned lane = threadIdx.x & 31u;
	RD r = proto;
	r.seek(ipos);
	uint64_t ic64 = r.gamma();
	const uint32_t ic = ic64 > 32 ? 32u : (uint32_t)ic64;
	V left = 0, prev_end = 0;
	uint32_t len = 0;
	for (uint32_t i = 0; i < ic; i++) {  // every lane reads the same codes; lane i keeps interval i
		const uint64_t gap = r.gamma();
		const V l = i == 0 ? (V)(x + nat2int(gap)) : (V)(prev_end + 1 + (V)gap);
		const uint32_t ln = (uint32_t)r.gamma() + (uint32_t)g.min_interval;
		prev_end = l + (V)ln;
		if (i == lane) { left = l; len = ln; }
	}
	// s: interval elements before mine
	uint32_t s = len;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const uint32_t t = __shfl_up_sync(FULL_MASK, s, o);
		if ((int)lane >= o) s += t;
	}
	const uint32_t total = __shfl_sync(FULL_MASK, s, 31);
	s -= len;
	if (total != ne - nres) { raise(g, ERR_DEGREE); return; }
	E *res = ext + total;
	// c: residuals smaller than my interval
	uint32_t c = 0;
	if (lane < ic) {
		uint32_t lo = 0, hi = nres;
		while (lo < hi) {
			const uint32_t mid = (lo + hi) >> 1;
			if ((V)res[mid] < left) lo = mid + 1;
			else hi = mid;
		}
		c = lo;
	}
	__syncwarp();
	for (uint32_t u0 = 0; u0 < nres; u0 += 32) {
		const uint32_t u = u0 + lane;
		const bool a = u < nres;
		const V rv = a ? (V)res[u] : 0;
		uint32_t below = 0;
		for (uint32_t k = 0; k < ic; k++) {
			const V lk = __shfl_sync(FULL_MASK, left, k);
			const uint32_t nk = __shfl_sync(FULL_MASK, len, k);
			if (lk < rv) below += nk;
		}
		__syncwarp();
		if (a) ext[u + below] = (E)rv;
		__syncwarp();
	}
	for (uint32_t e0 = 0; e0 < total; e0 += 32) {
		const uint32_t e = e0 + lane;
		int kk = 0;
		for (uint32_t k = 1; k < ic; k++) {
			const uint32_t sk = __shfl_sync(FULL_MASK, s, k);
			if (sk <= e) kk = (int)k;
		}
		const V lk = __shfl_sync(FULL_MASK, left, kk);
		const uint32_t sk = __shfl_sync(FULL_MASK, s, kk);
		const uint32_t ck = __shfl_sync(FULL_MASK, c, kk);
		if (e < total) ext[e + ck] = (E)(lk + (V)(e - sk));
	}
	__syncwarp();
}

}  // namespace wgbv
