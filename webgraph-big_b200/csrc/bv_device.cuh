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
};

// Everything a kernel needs to know about an opened graph; passed by value.
struct GraphDev {
	const uint64_t *words;  // bit stream, big-endian 64-bit words; words[0] holds absolute bits [bit_base, bit_base+64)
	uint64_t nwords;        // words that may be loaded (includes zero padding)
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
// Position-based reader with the current and the next word cached in registers: one 8-byte load per 64 bits
// consumed.  All positions are relative to g.bit_base.
struct BitReader {
	const uint64_t *__restrict__ w;
	uint64_t nwords;
	uint64_t pos;   // relative bit position
	uint64_t widx;  // index of `cur`
	uint64_t cur, nxt;

	__device__ __forceinline__ uint64_t load(uint64_t i) const { return i < nwords ? __ldg(w + i) : 0ull; }

	__device__ __forceinline__ void seek(uint64_t p) {
		pos = p;
		widx = p >> 6;
		cur = load(widx);
		nxt = load(widx + 1);
	}
	__device__ __forceinline__ void init(const GraphDev &g, uint64_t rel_pos) {
		w = g.words;
		nwords = g.nwords;
		seek(rel_pos);
	}
	// the next 64 bits of the stream, first bit in the most significant position
	__device__ __forceinline__ uint64_t peek() const {
		const unsigned s = (unsigned)pos & 63u;
		return s ? (cur << s) | (nxt >> (64u - s)) : cur;
	}
	__device__ __forceinline__ void skip(uint64_t n) {
		pos += n;
		const uint64_t nw = pos >> 6;
		if (nw != widx) {
			cur = (nw == widx + 1) ? nxt : load(nw);
			widx = nw;
			nxt = load(nw + 1);
		}
	}
	__device__ __forceinline__ bool past_end() const { return widx >= nwords; }

	__device__ __forceinline__ uint64_t bits(unsigned n) {  // n in [0, 64]
		if (n == 0) return 0;
		const uint64_t v = peek() >> (64u - n);
		skip(n);
		return v;
	}

	__device__ __forceinline__ uint64_t unary() {
		uint64_t win = peek();
		if (win) {
			const int z = __clzll((long long)win);
			skip((uint64_t)z + 1);
			return (uint64_t)z;
		}
		uint64_t total = 0;
		do {
			total += 64;
			skip(64);
			if (past_end()) return total;  // callers see garbage; the kernels raise ERR_OVERRUN on their own checks
			win = peek();
		} while (!win);
		const int z = __clzll((long long)win);
		skip((uint64_t)z + 1);
		return total + (uint64_t)z;
	}

	__device__ __forceinline__ uint64_t gamma() {
		const uint64_t win = peek();
		if (win >> 32) {  // at most 31 leading zeros: the whole code (<= 63 bits) is inside the window
			const int z = __clzll((long long)win);
			const uint64_t v = (win << z) >> (63 - z);
			skip(2ull * z + 1);
			return v - 1;
		}
		const uint64_t m = unary();
		if (m > 63) return 0;
		return ((1ull << m) | bits((unsigned)m)) - 1;
	}

	__device__ __forceinline__ uint64_t delta() {
		const uint64_t m = gamma();
		if (m > 63) return 0;
		return ((1ull << m) | bits((unsigned)m)) - 1;
	}

	__device__ __forceinline__ uint64_t zeta(int k) {
		const uint64_t win = peek();
		const int h = win ? __clzll((long long)win) : 64;
		const int nb = h * k + k - 1;  // payload bits before the optional extra one
		if (h + 1 + nb + 1 <= 64) {
			const uint64_t x = win << (h + 1);
			const uint64_t left = 1ull << (h * k);
			const uint64_t m = nb ? x >> (64 - nb) : 0;
			if (m < left) {
				skip((uint64_t)(h + 1 + nb));
				return m + left - 1;
			}
			const uint64_t b = (x >> (63 - nb)) & 1ull;
			skip((uint64_t)(h + 2 + nb));
			return ((m << 1) | b) - 1;
		}
		const uint64_t hh = unary();
		if (hh * (uint64_t)k + (uint64_t)k > 64) return 0;
		const uint64_t left = 1ull << (hh * k);
		const uint64_t m = bits((unsigned)(hh * k + k - 1));
		if (m < left) return m + left - 1;
		return ((m << 1) | bits(1)) - 1;
	}

	__device__ __forceinline__ uint64_t nibble() {
		uint64_t x = 0, stop;
		int guard = 0;
		do {
			const uint64_t q = bits(4);
			stop = q >> 3;
			x = (x << 3) | (q & 7);
		} while (!stop && ++guard < 22);
		return x;
	}

	__device__ __forceinline__ uint64_t golomb(int b) {
		if (b == 0) return 0;
		const uint64_t q = unary();
		const int s = 31 - __clz(b);
		const uint64_t M = (1ull << (s + 1)) - (uint64_t)b;
		uint64_t x = bits((unsigned)s);
		if (x >= M) x = ((x << 1) | bits(1)) - M;
		return q * (uint64_t)b + x;
	}

	// gamma | delta | unary switchboard used by outdegrees, references, block counts and blocks
	__device__ __forceinline__ uint64_t code(int coding) {
		switch (coding) {
		case C_GAMMA: return gamma();
		case C_DELTA: return delta();
		default: return unary();
		}
	}
	__device__ __forceinline__ uint64_t residual(int coding, int k) {
		switch (coding) {
		case C_ZETA: return zeta(k);
		case C_GAMMA: return gamma();
		case C_DELTA: return delta();
		case C_NIBBLE: return nibble();
		default: return golomb(k);
		}
	}
};

// Fast.nat2int
__device__ __forceinline__ int64_t nat2int(uint64_t x) { return (x & 1) ? -(int64_t)((x + 1) >> 1) : (int64_t)(x >> 1); }

// The field readers.  DEF = the default codings (gamma outdegrees / blocks / block counts, unary references,
// zeta residuals, BVG:526-542) compiled in; otherwise the codings are looked up at run time.
template <bool DEF> __device__ __forceinline__ uint64_t rd_outdegree(BitReader &r, const GraphDev &g) { return DEF ? r.gamma() : r.code(g.outdegree_coding); }
template <bool DEF> __device__ __forceinline__ uint64_t rd_reference(BitReader &r, const GraphDev &g) { return DEF ? r.unary() : r.code(g.reference_coding); }
template <bool DEF> __device__ __forceinline__ uint64_t rd_blockcount(BitReader &r, const GraphDev &g) { return DEF ? r.gamma() : r.code(g.block_count_coding); }
template <bool DEF> __device__ __forceinline__ uint64_t rd_block(BitReader &r, const GraphDev &g) { return DEF ? r.gamma() : r.code(g.block_coding); }
template <bool DEF> __device__ __forceinline__ uint64_t rd_residual(BitReader &r, const GraphDev &g) { return DEF ? r.zeta(g.zeta_k) : r.residual(g.residual_coding, g.zeta_k); }

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
