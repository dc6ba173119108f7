// bv_encode.cu -- BVGraph compression of a CSR held on the device.
//
// What BVGraph.store does per node (reference tree, BVGraph.java): try the node itself ("no reference") and each of
// the window_size predecessors as the reference list, price the differential encoding of the successor list against
// every candidate (diffComp on a null stream, BVG:1977-2159, 2254-2266), keep the strictly cheapest -- the smallest
// reference wins ties -- among those whose reference chain is still shorter than max_ref_count, and write it
// (BVG:2269-2270).  Pricing depends only on the lists, so it is done for all (node, candidate) pairs at once; the
// chain-length rule is the only serial part (a node's choice depends on the chain lengths of its predecessors) and
// runs as one thread per compression split over the table of prices, a few instructions per node; emission is
// parallel again once a prefix sum has turned record lengths into bit offsets.  Splits start with an empty window,
// exactly like the threads of the reference's parallel compressor (BVG:2408-2457), so split_nodes = 0 (one run)
// reproduces what a single-threaded BVGraph.store writes: tests/test_gpu_parity.py re-encodes the reference's
// cnr-2000 fixture on the device and compares the stream byte for byte.
//
//   k_enc_cost    thread per (node, candidate): bits of everything after the outdegree
//   k_enc_choose  thread per split: the reference of every node, its record length
//   k_scan        record lengths -> bit offsets (bv_kernels.cu)
//   k_enc_emit    thread per node: the record, OR-ed into a zeroed stream of 32-bit words (most significant bit first)
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>

#include "bv_block.cuh"
#include "bv_kernels.cuh"

namespace wgbv {

__device__ __forceinline__ int e_msb(uint64_t x) { return 63 - __clzll((long long)x); }
__device__ __forceinline__ uint64_t e_int2nat(int64_t v) { return v >= 0 ? (uint64_t)v << 1 : (((uint64_t)(-(v + 1))) << 1) + 1; }

// lengths of the instantaneous codes
__device__ __forceinline__ uint64_t e_len_gamma(uint64_t x) { return 2 * (uint64_t)e_msb(x + 1) + 1; }
__device__ __forceinline__ uint64_t e_len(int coding, uint64_t x, int k) {
	switch (coding) {
	case C_GAMMA: return e_len_gamma(x);
	case C_DELTA: {
		const int m = e_msb(x + 1);
		return (uint64_t)m + e_len_gamma((uint64_t)m);
	}
	case C_UNARY: return x + 1;
	case C_ZETA: {
		const uint64_t y = x + 1;
		const int h = e_msb(y) / k;
		const uint64_t left = 1ull << (h * k);
		return (uint64_t)h + 1 + (uint64_t)(h * k + k - 1) + (y - left < left ? 0 : 1);
	}
	case C_NIBBLE: {
		const int bits = x ? e_msb(x) + 1 : 1;
		return 4 * (uint64_t)((bits + 2) / 3);
	}
	default: {  // Golomb with modulus k
		if (k == 0) return 0;
		const int s = e_msb((uint64_t)k);
		const uint64_t M = (1ull << (s + 1)) - (uint64_t)k;
		return x / k + 1 + (uint64_t)s + ((x % k) < M ? 0 : 1);
	}
	}
}

// Bits OR-ed into a zeroed array of 32-bit words, most significant bit first (neighbouring records share words).
struct BitSink {
	uint32_t *words;
	uint64_t pos;
	__device__ __forceinline__ void put(uint64_t v, int len) {  // the low len bits of v, len in [0, 32]
		if (len == 0) return;
		const uint32_t x = len == 32 ? (uint32_t)v : (uint32_t)v & ((1u << len) - 1u);
		const uint64_t w = pos >> 5;
		const int o = (int)(pos & 31);
		if (o + len <= 32) atomicOr(words + w, x << (32 - o - len));
		else {
			const int l2 = o + len - 32;
			atomicOr(words + w, x >> l2);
			atomicOr(words + w + 1, x << (32 - l2));
		}
		pos += (uint64_t)len;
	}
	__device__ __forceinline__ void put_long(uint64_t v, int len) {  // len in [0, 64]
		if (len > 32) {
			put(v >> 32, len - 32);
			put(v & 0xffffffffull, 32);
		} else put(v, len);
	}
	__device__ __forceinline__ void unary(uint64_t x) {
		pos += x;  // x zeros (the stream is zeroed)
		put(1, 1);
	}
	__device__ __forceinline__ void gamma(uint64_t x) {
		const int m = e_msb(x + 1);
		unary((uint64_t)m);
		put_long((x + 1) & ((1ull << m) - 1), m);
	}
	__device__ __forceinline__ void code(int coding, uint64_t x, int k) {
		switch (coding) {
		case C_GAMMA: gamma(x); break;
		case C_DELTA: {
			const int m = e_msb(x + 1);
			gamma((uint64_t)m);
			put_long((x + 1) & ((1ull << m) - 1), m);
			break;
		}
		case C_UNARY: unary(x); break;
		case C_ZETA: {
			const uint64_t y = x + 1;
			const int h = e_msb(y) / k;
			const uint64_t left = 1ull << (h * k);
			unary((uint64_t)h);
			if (y - left < left) put_long(y - left, h * k + k - 1);
			else put_long(y, h * k + k);
			break;
		}
		case C_NIBBLE: {
			const int bits = x ? e_msb(x) + 1 : 1;
			int gq = (bits + 2) / 3;
			while (gq-- > 0) put(((gq == 0 ? 8u : 0u) | ((x >> (3 * gq)) & 7u)), 4);
			break;
		}
		default: {
			if (k == 0) break;
			const int s = e_msb((uint64_t)k);
			const uint64_t M = (1ull << (s + 1)) - (uint64_t)k, r = x % k;
			unary(x / k);
			if (r < M) put(r, s);
			else put(r + M, s + 1);
		}
		}
	}
};

// The two-pointer walk of diffComp (BVG:1996-2051): alternating copy / skip runs over the reference list; what the
// current list has and the reference list lacks goes to `fe` in increasing order.  A trailing copy run is reported
// only if it stops before the end of the reference list.
template <class FB, class FE>
__device__ __forceinline__ void enc_walk(const int64_t *cur, int64_t dc, const int64_t *ref, int64_t dr, FB &fb, FE &fe) {
	int64_t j = 0, k = 0, run = 0;
	bool copying = true;
	while (j < dc && k < dr) {
		const int64_t a = cur[j], b = ref[k];
		if (copying) {
			if (a > b) { fb(run); copying = false; run = 0; }
			else if (a < b) { fe(a); j++; }
			else { j++; k++; run++; }
		} else {
			if (a < b) { fe(a); j++; }
			else if (a > b) { k++; run++; }
			else { fb(run); copying = true; run = 0; }
		}
	}
	if (copying && k < dr) fb(run);
	while (j < dc) fe(cur[j++]);
}

struct NoBlocks {
	__device__ __forceinline__ void operator()(int64_t) {}
};

// blocks: priced, or written
struct PriceBlocks {
	const EncDev &E;
	uint64_t bits = 0;
	int64_t nb = 0;
	__device__ __forceinline__ PriceBlocks(const EncDev &e) : E(e) {}
	__device__ __forceinline__ void operator()(int64_t run) {
		bits += e_len(E.block_coding, (uint64_t)(run - (nb ? 1 : 0)), 0);
		nb++;
	}
};
struct EmitBlocks {
	const EncDev &E;
	BitSink &S;
	int64_t nb = 0;
	__device__ __forceinline__ EmitBlocks(const EncDev &e, BitSink &s) : E(e), S(s) {}
	__device__ __forceinline__ void operator()(int64_t run) {
		S.code(E.block_coding, (uint64_t)(run - (nb ? 1 : 0)), 0);
		nb++;
	}
};

// extras: maximal runs of consecutive integers of at least minint elements are intervals, everything else residuals
// (intervalize, BVG:1595-1618).  MODE 0 prices both, 1 writes the intervals, 2 writes the residuals.
template <int MODE>
struct Extras {
	const EncDev &E;
	BitSink *S;
	int64_t x;
	uint64_t bits = 0;
	int64_t nextra = 0, ic = 0;
	int64_t rs = 0, rl = 0;  // the run of consecutive integers being collected
	bool have_res = false, have_int = false;
	int64_t prev_res = 0, prev_right = 0;
	__device__ __forceinline__ Extras(const EncDev &e, BitSink *s, int64_t node) : E(e), S(s), x(node) {}
	__device__ __forceinline__ void residual(int64_t v) {
		const uint64_t a = have_res ? (uint64_t)(v - prev_res - 1) : e_int2nat(v - x);
		if (MODE == 0) bits += e_len(E.residual_coding, a, E.zetak);
		if (MODE == 2) S->code(E.residual_coding, a, E.zetak);
		prev_res = v;
		have_res = true;
	}
	__device__ __forceinline__ void interval(int64_t left, int64_t len) {
		const uint64_t a = have_int ? (uint64_t)(left - prev_right - 1) : e_int2nat(left - x);
		const uint64_t b = (uint64_t)(len - E.minint);
		if (MODE == 0) bits += e_len_gamma(a) + e_len_gamma(b);
		if (MODE == 1) { S->gamma(a); S->gamma(b); }
		prev_right = left + len;
		have_int = true;
		ic++;
	}
	__device__ __forceinline__ void flush() {
		if (rl) {
			// (a lone integer is never an interval, even for min_interval_length 1: BVG:1605 only looks at runs of two and more)
			if (E.minint != 0 && rl >= E.minint && rl >= 2) interval(rs, rl);
			else
				for (int64_t i = 0; i < rl; i++) residual(rs + i);
			rl = 0;
		}
	}
	__device__ __forceinline__ void operator()(int64_t v) {
		nextra++;
		if (E.minint == 0) { residual(v); return; }
		if (rl && v == rs + rl) rl++;
		else {
			flush();
			rs = v;
			rl = 1;
		}
	}
};

// bits of everything after the outdegree when row i is coded against the row `ref` before it (0: no reference)
__device__ __forceinline__ uint64_t enc_price(const EncDev &E, int64_t i, int ref) {
	const int64_t a = E.rowptr[i], dc = E.rowptr[i + 1] - a;
	const int64_t *cur = E.col + (a - E.rowptr[0]);
	const int64_t *rl = nullptr;
	int64_t dr = 0;
	if (ref) {
		const int64_t b = E.rowptr[i - ref];
		rl = E.col + (b - E.rowptr[0]);
		dr = E.rowptr[i - ref + 1] - b;
	}
	PriceBlocks pb(E);
	Extras<0> pe(E, nullptr, E.first_node + i);
	enc_walk(cur, dc, rl, dr, pb, pe);
	pe.flush();
	uint64_t bits = 0;
	if (E.window > 0) bits += e_len(E.reference_coding, (uint64_t)ref, 0);
	if (ref) bits += e_len(E.blockcount_coding, (uint64_t)pb.nb, 0) + pb.bits;
	if (pe.nextra > 0) {
		if (E.minint != 0) bits += e_len_gamma((uint64_t)pe.ic);
		bits += pe.bits;
	}
	return bits;
}

static constexpr uint32_t COST_NONE = 0xffffffffu;

__global__ void __launch_bounds__(TPB) k_enc_cost(EncDev E, uint32_t *__restrict__ cost) {
	const int cyc = E.window + 1;
	const int64_t t = (int64_t)blockIdx.x * TPB + threadIdx.x;
	if (t >= E.n * cyc) return;
	const int64_t i = t / cyc;
	const int ref = (int)(t % cyc);
	uint32_t c = COST_NONE;
	const int64_t d = E.rowptr[i + 1] - E.rowptr[i];
	const int64_t split_start = (i / E.split) * E.split;
	if (d > 0 && i - ref >= split_start && (ref == 0 || E.rowptr[i - ref + 1] - E.rowptr[i - ref] > 0)) {
		const uint64_t bits = enc_price(E, i, ref);
		c = bits >= (uint64_t)COST_NONE ? COST_NONE - 1 : (uint32_t)bits;
	}
	cost[t] = c;
}

// The only serial rule of the compressor: a candidate is eligible while its own chain of references is shorter than
// max_ref_count; the node then continues that chain (BVG:2254-2270).  One thread per compression split.
__global__ void k_enc_choose(EncDev E, const uint32_t *__restrict__ cost, uint8_t *__restrict__ chosen, unsigned long long *__restrict__ bits) {
	const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	const int64_t lo = s * E.split;
	if (lo >= E.n) return;
	const int64_t hi = lo + E.split < E.n ? lo + E.split : E.n;
	const int cyc = E.window + 1;
	int refcount[256];  // (window <= 255)
	for (int q = 0; q < cyc; q++) refcount[q] = 0;
	// The decisions are a serial chain, the loads are not: row pointers and prices of a batch of nodes are fetched
	// first (independent loads, all in flight together), then the batch is decided from registers / local memory.
	constexpr int BATCH = 8;
	for (int64_t b0 = lo; b0 < hi; b0 += BATCH) {
		int64_t dd[BATCH];
		uint32_t cc[BATCH][9];
		const bool small = cyc <= 9;
#pragma unroll
		for (int u = 0; u < BATCH; u++) {
			const int64_t i = b0 + u;
			dd[u] = i < hi ? E.rowptr[i + 1] - E.rowptr[i] : 0;
			if (small && i < hi) {
#pragma unroll
				for (int r = 0; r < 9; r++) cc[u][r] = r < cyc ? cost[i * cyc + r] : COST_NONE;
			}
		}
#pragma unroll
		for (int u = 0; u < BATCH; u++) {
			const int64_t i = b0 + u;
			if (i >= hi) break;
			const int64_t d = dd[u];
			uint64_t len = e_len(E.outdegree_coding, (uint64_t)d, 0);
			int best_ref = 0;
			if (d > 0) {
				const int ci = (int)(i % cyc);
				refcount[ci] = -1;
				uint32_t best = COST_NONE;
				int best_cand = ci;
				for (int ref = 0; ref < cyc; ref++) {
					if (ref > i - lo) break;
					const int cand = (int)((i - ref) % cyc);
					const uint32_t c = (small && ref < 9) ? cc[u][ref] : cost[i * cyc + ref];
					if (refcount[cand] < E.maxref && c != COST_NONE && c < best) {
						best = c;
						best_ref = ref;
						best_cand = cand;
					}
				}
				refcount[ci] = refcount[best_cand] + 1;
				len += best;
			}
			chosen[i] = (uint8_t)best_ref;
			bits[i] = len;
		}
	}
}

__global__ void __launch_bounds__(TPB) k_enc_emit(EncDev E, const uint8_t *__restrict__ chosen, const int64_t *__restrict__ off, uint32_t *words) {
	const int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
	if (i >= E.n) return;
	const int64_t a = E.rowptr[i], dc = E.rowptr[i + 1] - a;
	const int64_t *cur = E.col + (a - E.rowptr[0]);
	BitSink S{words, (uint64_t)off[i]};
	S.code(E.outdegree_coding, (uint64_t)dc, 0);
	if (dc == 0) return;
	const int ref = chosen[i];
	if (E.window > 0) S.code(E.reference_coding, (uint64_t)ref, 0);
	const int64_t *rl = nullptr;
	int64_t dr = 0;
	if (ref) {
		const int64_t b = E.rowptr[i - ref];
		rl = E.col + (b - E.rowptr[0]);
		dr = E.rowptr[i - ref + 1] - b;
	}
	const int64_t x = E.first_node + i;
	// counts first (the stream wants them before what they count), then one walk per section
	PriceBlocks pb(E);
	Extras<0> pe(E, nullptr, x);
	enc_walk(cur, dc, rl, dr, pb, pe);
	pe.flush();
	if (ref) {
		S.code(E.blockcount_coding, (uint64_t)pb.nb, 0);
		EmitBlocks eb(E, S);
		Extras<3> none(E, nullptr, x);
		enc_walk(cur, dc, rl, dr, eb, none);
	}
	if (pe.nextra > 0) {
		NoBlocks nb;
		if (E.minint != 0) {
			S.gamma((uint64_t)pe.ic);
			Extras<1> ei(E, &S, x);
			enc_walk(cur, dc, rl, dr, nb, ei);
			ei.flush();
		}
		Extras<2> er(E, &S, x);
		enc_walk(cur, dc, rl, dr, nb, er);
		er.flush();
	}
}

// ---- host side
int encode_csr_device(cudaStream_t stream, const EncDev &E, uint32_t *d_cost, uint8_t *d_chosen, unsigned long long *d_bits, int64_t *d_off,
                      uint64_t *scan_state, uint32_t *ticket, uint32_t *d_words, uint64_t words_capacity, uint64_t *total_bits_out) {
	if (E.n <= 0) { *total_bits_out = 0; return 0; }
	Launch L{stream, nullptr};
	const int cyc = E.window + 1;
	const bool timing = getenv("WGBV_TIMING") != nullptr;
	cudaEvent_t ev[5];
	if (timing) {
		for (auto &e : ev) cudaEventCreate(&e);
		cudaEventRecord(ev[0], stream);
	}
	k_enc_cost<<<grid_for(E.n * cyc, TPB), TPB, 0, stream>>>(E, d_cost);
	if (timing) cudaEventRecord(ev[1], stream);
	const int64_t splits = (E.n + E.split - 1) / E.split;
	k_enc_choose<<<grid_for(splits, 64), 64, 0, stream>>>(E, d_cost, d_chosen, d_bits);
	if (timing) cudaEventRecord(ev[2], stream);
	launch_scan64(L, d_bits, d_off, E.n, scan_state, ticket);
	int64_t total = 0;
	if (cudaMemcpyAsync(&total, d_off + E.n, 8, cudaMemcpyDeviceToHost, stream) != cudaSuccess || cudaStreamSynchronize(stream) != cudaSuccess) return -1;
	*total_bits_out = (uint64_t)total;
	const uint64_t need = ((uint64_t)total + 63) / 64 * 2;  // 32-bit words, rounded to 64-bit units
	if (need > words_capacity) return -2;
	if (cudaMemsetAsync(d_words, 0, need * 4, stream) != cudaSuccess) return -1;
	if (timing) cudaEventRecord(ev[3], stream);
	k_enc_emit<<<grid_for(E.n, TPB), TPB, 0, stream>>>(E, d_chosen, d_off, d_words);
	// words -> bytes of the file: every 32-bit word holds its bits most significant first
	launch_bswap(L, reinterpret_cast<uint64_t *>(d_words), need / 2);
	if (timing) {
		cudaEventRecord(ev[4], stream);
		cudaStreamSynchronize(stream);
		float a = 0, b = 0, c = 0;
		cudaEventElapsedTime(&a, ev[0], ev[1]);
		cudaEventElapsedTime(&b, ev[1], ev[2]);
		cudaEventElapsedTime(&c, ev[3], ev[4]);
		fprintf(stderr, "[wgbv] compress %lld nodes: pricing %.1f ms, reference choice %.1f ms, emission %.1f ms\n", (long long)E.n, a, b, c);
		for (auto &e : ev) cudaEventDestroy(e);
	}
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

}  // namespace wgbv
