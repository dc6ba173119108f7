// bv_device.cuh -- device-side building blocks of the BVGraph decoder: the bit reader, the
// instantaneous codes and the Elias-Fano offsets list.
//
// The codes are the ones webgraph-big reads through dsiutils' InputBitStream at the call sites
// BVGraph.java:627-633, 654-661, 692-703, 728-735, 758-765, 787-796 (reference tree); definitions in
// SURVEY.md Appendix A.  Nothing here is a translation of that Java: the stream is kept in device memory
// as 64-bit words whose most significant bit is the first bit of the stream (the loader byte-swaps once at
// upload), so a code is decoded from a 64-bit window with one count-leading-zeros and a couple of shifts.
#pragma once
#include <stdint.h>

namespace wgbv {

enum : int { C_DELTA = 1, C_GAMMA = 2, C_GOLOMB = 3, C_SKEWED_GOLOMB = 4, C_UNARY = 5, C_ZETA = 6, C_NIBBLE = 7 };

// error bits raised by kernels (sticky, OR-ed into one device word)
enum : uint32_t {
	ERR_REF_WINDOW = 1u,      // reference larger than the window or than the node id (BVG:701)
	ERR_OVERRUN = 2u,         // a code ran past the end of the bit stream
	ERR_DEGREE = 4u,          // outdegree >= 2^31 or counts that do not add up (copied / intervals > outdegree)
	ERR_REF_UNRESOLVED = 8u,  // a needed row references a row outside the decodable range
	ERR_NOT_SORTED = 16u,     // duplicate between copied successors and extras (malformed file)
	ERR_CAPACITY = 32u,       // output capacity exceeded
	ERR_INDEX = 64u,          // a queried node id is outside the range the handle can decode (BVG:863)
	ERR_FUSED_FALLBACK = 128u,  // not an error: the fused tile kernel met something only the general path handles (the host reruns the decode)
	ERR_FUSED_DEEP = 256u,      // same, for a reference chain longer than the properties promise: the rerun must not trust them either
};

// Everything a kernel needs to know about an opened graph; passed by value.
struct GraphDev {
	const uint64_t *words;  // bit stream as 32-bit words, most significant bit first; read through (const uint32_t *)words.
	                        // Word 0 holds absolute bits [bit_base, bit_base + 32)
	uint64_t nwords;        // 64-bit units that may be loaded (includes zero padding): 2 * nwords 32-bit words
	uint64_t bit_base;      // absolute bit offset of words[0] (multiple of 64)
	uint64_t bit_end;       // absolute end of the stream held (exclusive)
	// Elias-Fano list of the offsets of nodes [ef_first, ef_first + ef_count); values are absolute offsets minus bit_base
	const uint64_t *ef_lower;   // l-bit lower parts, packed LSB-first in 64-bit words
	const uint64_t *ef_upper;   // unary upper parts: bit (v >> l) + i set for element i
	const uint64_t *ef_hints;   // for every 64th element: bit position in ef_upper, or (1<<63 | index) into ef_sparse
	const uint64_t *ef_sparse;  // explicit bit positions (64 per sparse block)
	int64_t ef_first;
	int64_t ef_count;
	int32_t ef_l;
	int64_t nodes;          // of the whole graph
	int32_t window, min_interval, zeta_k;
	int32_t outdegree_coding, block_coding, residual_coding, reference_coding, block_count_coding;
	uint32_t *err;          // sticky error word
};

__device__ __forceinline__ void raise(const GraphDev &g, uint32_t bit) { atomicOr(g.err, bit); }

// ------------------------------------------------------------------------------------------------ bit reader
// The stream lives in device memory as 32-bit words whose most significant bit comes first (the loader byte-swaps
// once at upload).  The reader holds three consecutive words in registers -- hi, lo and the prefetched nxt -- and
// the number of bits of hi already consumed, so that
//   * the next 32 bits of the stream are ONE funnel shift:  peek() = (hi:lo << sh) >> 32
//   * consuming bits is an add, and crossing a word boundary is three register moves plus the load of the word
//     after next (predicated, no divergent branch; its latency is off the critical path).
// Every code of up to 32 bits -- all but a vanishing fraction in practice -- is decoded from peek() with 32-bit
// arithmetic: one count-leading-zeros and a few shifts.  Longer codes take a slow path built from unary() and
// bits().  All positions are relative to g.bit_base.
// LDG = true: the words are in global memory and read through the read-only path; false: plain (generic) loads, so
// that the same decoder runs on a tile of the stream staged in shared memory.
template <bool LDG>
struct WordCore {
	static constexpr bool IS_GLOBAL = LDG;
	using pos_t = uint64_t;
	const uint32_t *base;  // address of the word the last seek() started from
	uint64_t j0;    // stream index of that word
	uint32_t lim;   // words that may be loaded from base (saturated)
	uint32_t idx;   // index (from base) of the word held in nxt
	uint32_t hi, lo, nxt;
	uint32_t sh;    // bits of hi already consumed, 0..31

	__device__ __forceinline__ uint32_t load(uint32_t i) const { return i < lim ? (LDG ? __ldg(base + i) : base[i]) : 0u; }
	// The word after next is loaded with one refill of lead time only (the compiler waits for it at the head of the
	// decoding loop), so a miss there costs a full DRAM round trip per lane.  Every 8 words (one 32-byte sector) the
	// reader therefore asks for the sector PF_AHEAD words further on; the request carries no register, hence no wait.
	static constexpr uint32_t PF_AHEAD = 16;
	__device__ __forceinline__ void prefetch(uint32_t i) const {
		if (LDG && (i & 7u) == 0 && i + PF_AHEAD < lim) asm volatile("prefetch.global.L1 [%0];" ::"l"(base + i + PF_AHEAD));
	}

	// positions the reader on an arbitrary word array (e.g. a staged tile): `words` holds `nwords` 32-bit words
	__device__ __forceinline__ void init_words(const uint32_t *words, uint32_t nwords, uint32_t rel_pos) {
		j0 = rel_pos >> 5;
		if (j0 > nwords) j0 = nwords;
		base = words + j0;
		lim = nwords - (uint32_t)j0;
		sh = rel_pos & 31u;
		hi = load(0);
		lo = load(1);
		idx = 2;
		nxt = load(2);
	}

	__device__ __forceinline__ void init(const GraphDev &g, uint64_t rel_pos) {
		const uint64_t nw = g.nwords * 2;
		j0 = rel_pos >> 5;
		if (j0 > nw) j0 = nw;
		base = (const uint32_t *)g.words + j0;
		const uint64_t room = nw - j0;
		lim = room > 0xffffffffull ? 0xffffffffu : (uint32_t)room;
		sh = (uint32_t)rel_pos & 31u;
		hi = load(0);
		lo = load(1);
		idx = 2;
		nxt = load(2);
	}
	// re-positions inside the same stream (any position at or after the one given to init)
	__device__ __forceinline__ void seek(uint64_t p) {
		const uint64_t j = p >> 5;
		const uint32_t i = (uint32_t)(j - j0);
		sh = (uint32_t)p & 31u;
		hi = load(i);
		lo = load(i + 1);
		idx = i + 2;
		nxt = load(idx);
	}
	__device__ __forceinline__ uint64_t tell() const { return (j0 + idx - 2) * 32 + sh; }
	__device__ __forceinline__ bool past_end() const { return idx >= 3 && idx - 3 > lim; }

	// the next 32 bits of the stream, first bit in the most significant position
	__device__ __forceinline__ uint32_t peek() const { return __funnelshift_l(lo, hi, sh); }
	__device__ __forceinline__ void consume(uint32_t n) {  // n in [0, 32]
		sh += n;
		if (LDG) {
			if (sh >= 32) {
				sh -= 32;
				hi = lo;
				lo = nxt;
				nxt = load(++idx);
				prefetch(idx);
			}
		} else {
			// words in shared memory: load the candidate unconditionally (cheap, uniform) and select, so that the lanes
			// of a warp -- which cross word boundaries at different times -- do not split into two instruction streams
			const uint32_t cand = lim ? base[idx + 1 < lim ? idx + 1 : lim - 1] : 0u;  // (lim == 0: a position at the very end)
			const bool adv = sh >= 32;
			sh = adv ? sh - 32 : sh;
			hi = adv ? lo : hi;
			lo = adv ? nxt : lo;
			nxt = adv ? cand : nxt;
			idx += adv ? 1u : 0u;
		}
	}
};

// A reader that keeps nothing but its bit position: every peek loads the two words under the position again.  Meant for
// streams staged in shared memory, where a load costs about as much as the register shuffling it replaces: fewer
// instructions per code and one live register instead of six.  `words` must be followed by at least one loadable
// word past the last one holding data (the staging code pads); positions beyond the staged span read the last pair.
struct PosCore {
	static constexpr bool IS_GLOBAL = false;
	using pos_t = uint32_t;
	const uint32_t *base;
	uint32_t last;  // index of the last word that may be the first of a loaded pair
	uint32_t p;     // bit position relative to base[0]
	__device__ __forceinline__ void init_words(const uint32_t *words, uint32_t nwords, uint32_t rel_pos) {
		base = words;
		last = nwords >= 2 ? nwords - 2 : 0;
		p = rel_pos;
	}
	__device__ __forceinline__ void seek(uint64_t pos) { p = (uint32_t)pos; }
	__device__ __forceinline__ uint64_t tell() const { return p; }
	__device__ __forceinline__ bool past_end() const { return (p >> 5) > last + 1; }
	__device__ __forceinline__ uint32_t peek() const {
		uint32_t w = p >> 5;
		w = w < last ? w : last;
		return __funnelshift_l(base[w + 1], base[w], p);
	}
	__device__ __forceinline__ void consume(uint32_t n) { p += n; }
};

// The instantaneous codes on top of a core that provides peek() / consume() / past_end().
template <class Core>
struct CodesT : Core {
	__device__ __forceinline__ uint64_t bits(unsigned n) {  // n in [0, 64]
		uint64_t v = 0;
		if (n > 32) {
			v = this->peek();
			this->consume(32);
			n -= 32;
		}
		if (n) {
			v = (v << n) | (this->peek() >> (32u - n));
			this->consume(n);
		}
		return v;
	}

	__device__ __forceinline__ uint64_t unary() {
		uint32_t t = this->peek();
		uint64_t total = 0;
		while (t == 0) {
			total += 32;
			this->consume(32);
			if (this->past_end()) return total;  // the kernels raise ERR_OVERRUN on their own checks
			t = this->peek();
		}
		const unsigned z = (unsigned)__clz((int)t);
		this->consume(z + 1);
		return total + z;
	}

	__device__ __forceinline__ uint64_t gamma() {
		const uint32_t t = this->peek();
		if (t >> 16) {  // at most 15 leading zeros: the whole code (<= 31 bits) is in the window
			const unsigned z = (unsigned)__clz((int)t);
			const uint32_t v = (t << z) >> (31u - z);
			this->consume(2 * z + 1);
			return (uint64_t)v - 1;
		}
		const uint64_t m = this->unary();
		if (m > 63) return 0;
		return ((1ull << m) | this->bits((unsigned)m)) - 1;
	}

	__device__ __forceinline__ uint64_t delta() {
		const uint64_t m = this->gamma();
		if (m > 63) return 0;
		return ((1ull << m) | this->bits((unsigned)m)) - 1;
	}

	__device__ __forceinline__ uint64_t zeta(int k) {
		const uint32_t t = this->peek();
		const int h = __clz((int)t);  // (32 for t == 0)
		const int hk = h * k;
		if (hk + h + k + 1 <= 32) {
			// unary part (h + 1 bits), then hk + k - 1 payload bits m and possibly one more bit b: m2 = m:b
			const uint32_t x = t << (h + 1);
			const uint32_t m2 = x >> (32 - (hk + k));
			const uint32_t two_left = 2u << hk;
			const bool shorter = m2 < two_left;  // m < left = 2^hk
			this->consume((uint32_t)(hk + h + k) + (shorter ? 0u : 1u));
			return (uint64_t)(shorter ? (m2 >> 1) + (two_left >> 1) - 1u : m2 - 1u);
		}
		const uint64_t hh = this->unary();
		if (hh * (uint64_t)k + (uint64_t)k > 64) return 0;
		const uint64_t left = 1ull << (hh * k);
		const uint64_t m = this->bits((unsigned)(hh * k + k - 1));
		if (m < left) return m + left - 1;
		return ((m << 1) | this->bits(1)) - 1;
	}

	__device__ __forceinline__ uint64_t nibble() {
		uint64_t x = 0, stop;
		int guard = 0;
		do {
			const uint64_t q = this->bits(4);
			stop = q >> 3;
			x = (x << 3) | (q & 7);
		} while (!stop && ++guard < 22);
		return x;
	}

	__device__ __forceinline__ uint64_t golomb(int b) {
		if (b == 0) return 0;
		const uint64_t q = this->unary();
		const int s = 31 - __clz(b);
		const uint64_t M = (1ull << (s + 1)) - (uint64_t)b;
		uint64_t x = this->bits((unsigned)s);
		if (x >= M) x = ((x << 1) | this->bits(1)) - M;
		return q * (uint64_t)b + x;
	}

	// gamma | delta | unary switchboard used by outdegrees, references, block counts and blocks
	__device__ __forceinline__ uint64_t code(int coding) {
		switch (coding) {
		case C_GAMMA: return this->gamma();
		case C_DELTA: return this->delta();
		default: return this->unary();
		}
	}
	__device__ __forceinline__ uint64_t residual(int coding, int k) {
		switch (coding) {
		case C_ZETA: return this->zeta(k);
		case C_GAMMA: return this->gamma();
		case C_DELTA: return this->delta();
		case C_NIBBLE: return this->nibble();
		default: return this->golomb(k);
		}
	}
};
template <bool LDG> using BitReaderT = CodesT<WordCore<LDG>>;
using BitReader = BitReaderT<true>;
using TileReader = BitReaderT<false>;
using PosReader = CodesT<PosCore>;

// Fast.nat2int
__device__ __forceinline__ int64_t nat2int(uint64_t x) { return (x & 1) ? -(int64_t)((x + 1) >> 1) : (int64_t)(x >> 1); }

// The field readers.  DEF selects how the codings are resolved: 0 = looked up at run time; 1 = the default codings
// (gamma outdegrees / blocks / block counts, unary references, zeta residuals, BVG:526-542) compiled in, zeta_k at
// run time; 2 = the default codings with zeta_k = 3 (the default k, BVG:473) compiled in.
template <int DEF, class RD> __device__ __forceinline__ uint64_t rd_outdegree(RD &r, const GraphDev &g) { return DEF ? r.gamma() : r.code(g.outdegree_coding); }
template <int DEF, class RD> __device__ __forceinline__ uint64_t rd_reference(RD &r, const GraphDev &g) { return DEF ? r.unary() : r.code(g.reference_coding); }
template <int DEF, class RD> __device__ __forceinline__ uint64_t rd_blockcount(RD &r, const GraphDev &g) { return DEF ? r.gamma() : r.code(g.block_count_coding); }
template <int DEF, class RD> __device__ __forceinline__ uint64_t rd_block(RD &r, const GraphDev &g) { return DEF ? r.gamma() : r.code(g.block_coding); }
template <int DEF, class RD> __device__ __forceinline__ uint64_t rd_residual(RD &r, const GraphDev &g) {
	return DEF == 2 ? r.zeta(3) : DEF == 1 ? r.zeta(g.zeta_k) : r.residual(g.residual_coding, g.zeta_k);
}

// ------------------------------------------------------------------------------------------------ Elias-Fano
// position of the k-th (0-based) set bit of x; x must have more than k set bits
__device__ __forceinline__ int select64(uint64_t x, int k) {
	uint32_t v = (uint32_t)x;
	int base = 0;
	int c = __popc(v);
	if (k >= c) { k -= c; v = (uint32_t)(x >> 32); base = 32; }
	c = __popc(v & 0xffffu);
	if (k >= c) { k -= c; v >>= 16; base += 16; }
	c = __popc(v & 0xffu);
	if (k >= c) { k -= c; v >>= 8; base += 8; }
	c = __popc(v & 0xfu);
	if (k >= c) { k -= c; v >>= 4; base += 4; }
	c = __popc(v & 0x3u);
	if (k >= c) { k -= c; v >>= 2; base += 2; }
	if (k >= (int)(v & 1u)) base += 1;
	return base;
}

// Bit offset (relative to g.bit_base) of node x's record; x in [ef_first, ef_first + ef_count).
// The role of EliasFanoMonotoneLongBigList.getLong at BVG:834, 847, 1144.
__device__ __forceinline__ uint64_t ef_get(const GraphDev &g, int64_t x) {
	const uint64_t i = (uint64_t)(x - g.ef_first);
	const uint64_t hint = __ldg(g.ef_hints + (i >> 6));
	const int r = (int)(i & 63);
	uint64_t p;
	if (hint >> 63) {
		p = __ldg(g.ef_sparse + ((hint & ~(1ull << 63)) << 6) + r);
	} else {
		uint64_t wi = hint >> 6;
		uint64_t word = __ldg(g.ef_upper + wi) & (~0ull << (hint & 63));
		int need = r;
		for (;;) {
			const int c = __popcll(word);
			if (need < c) break;
			need -= c;
			word = __ldg(g.ef_upper + ++wi);
		}
		p = (wi << 6) + select64(word, need);
	}
	const uint64_t high = p - i;
	uint64_t low = 0;
	if (g.ef_l) {
		const uint64_t lb = i * (uint64_t)g.ef_l;
		const uint64_t w0 = __ldg(g.ef_lower + (lb >> 6));
		const unsigned s = (unsigned)lb & 63u;
		low = w0 >> s;
		if (s + g.ef_l > 64) low |= __ldg(g.ef_lower + (lb >> 6) + 1) << (64u - s);
		low &= (1ull << g.ef_l) - 1;
	}
	return (high << g.ef_l) | low;
}

}  // namespace wgbv
