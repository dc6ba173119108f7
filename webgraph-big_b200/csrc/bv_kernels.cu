// bv_kernels.cu -- the CUDA kernels of the BVGraph -> CSR decoder (sm_100a).
//
// General ("level-synchronous") pipeline, valid for every graph including unbounded reference chains:
//   k_header      stage 1   thread per node: Elias-Fano offset -> outdegree + reference
//   k_scan        stage 2   single-pass decoupled look-back prefix sum -> CSR row pointers
//   k_depth       stage 2b  reference-chain depth of every node
//   k_extras      stage 3a  intervals + residuals of every row, merged, into the tail of the row's CSR slot
//   k_copy_level  stage 3b  one launch per depth level: copy blocks of the reference row merged in place
// All of it is integer/bit work bound by HBM and the integer pipes; no tensor cores are involved.
#include <stdio.h>

#include <algorithm>

#include "bv_kernels.cuh"
#include "bv_rows.cuh"
#include "bv_block.cuh"

namespace wgbv {



// ------------------------------------------------------------------------------------------------ load-time
__global__ void k_bswap(uint64_t *words, uint64_t nwords) {
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= nwords) return;
	const uint64_t v = words[i];
	const uint32_t lo = (uint32_t)v, hi = (uint32_t)(v >> 32);
	// each 32-bit half is byte-swapped in place: 32-bit word j of the result holds stream bits [32j, 32j+32), MSB first
	words[i] = ((uint64_t)__byte_perm(hi, 0, 0x0123) << 32) | __byte_perm(lo, 0, 0x0123);
}

void launch_bswap(const Launch &L, uint64_t *words, uint64_t nwords) {
	if (!nwords) return;
	k_bswap<<<grid_for((int64_t)nwords, TPB), TPB, 0, L.stream>>>(words, nwords);
	WGBV_LAUNCHED(L);
}

__global__ void k_peek(PeekArgs a, uint64_t *dst) {
	const unsigned t = threadIdx.x;
	if (t < a.n) dst[t] = a.bytes[t] == 8 ? *static_cast<const uint64_t *>(a.src[t]) : (uint64_t) * static_cast<const uint32_t *>(a.src[t]);
}

void launch_peek(const Launch &L, const PeekArgs &a, uint64_t *mapped_dst) {
	if (!a.n) return;
	k_peek<<<1, 32, 0, L.stream>>>(a, mapped_dst);
	WGBV_LAUNCHED(L);
}

// successors as uint32 for narrow delivery: four per thread (two 16-byte loads, one 16-byte store)
__global__ void __launch_bounds__(TPB) k_narrow(const int64_t *__restrict__ in, uint32_t *__restrict__ out, uint64_t count) {
	const uint64_t i = ((uint64_t)blockIdx.x * TPB + threadIdx.x) * 4;
	if (i + 4 <= count) {
		const longlong2 a = *reinterpret_cast<const longlong2 *>(in + i), b = *reinterpret_cast<const longlong2 *>(in + i + 2);
		*reinterpret_cast<uint4 *>(out + i) = make_uint4((uint32_t)a.x, (uint32_t)a.y, (uint32_t)b.x, (uint32_t)b.y);
	} else
		for (uint64_t j = i; j < count; j++) out[j] = (uint32_t)in[j];
}

void launch_narrow(const Launch &L, const int64_t *in, uint32_t *out, uint64_t count) {
	if (!count) return;
	k_narrow<<<grid_for((int64_t)((count + 3) / 4), TPB), TPB, 0, L.stream>>>(in, out, count);
	WGBV_LAUNCHED(L);
}

// ------------------------------------------------------------------------------------------------ stage 1
// Outdegree and reference of every node: readOutdegree / readReference at the node's offset
// (BVG:654-661, 692-703; the offset lookup is offsets.getLong(x), BVG:834).
template <int DEF>
__global__ void __launch_bounds__(TPB) k_header(GraphDev g, int64_t lo, int64_t count, uint32_t *__restrict__ deg, uint8_t *__restrict__ ref) {
	const int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
	if (i >= count) return;
	const int64_t x = lo + i;
	BitReader r;
	r.init(g, ef_get(g, x));
	const uint64_t d = rd_outdegree<DEF>(r, g);
	uint64_t rf = 0;
	if (d > 0x7fffffffull) { raise(g, ERR_DEGREE); deg[i] = 0; ref[i] = 0; return; }
	if (d && g.window > 0) {
		rf = rd_reference<DEF>(r, g);
		if (rf > (uint64_t)g.window || (int64_t)rf > x) { raise(g, ERR_REF_WINDOW); rf = 0; }
	}
	if (r.past_end()) raise(g, ERR_OVERRUN);
	deg[i] = (uint32_t)d;
	ref[i] = (uint8_t)rf;
}

void launch_header(const Launch &L, const GraphDev &g, int def, int64_t lo, int64_t count, uint32_t *deg, uint8_t *ref) {
	if (count <= 0) return;
	switch (def) {
	case 2: k_header<2><<<grid_for(count, TPB), TPB, 0, L.stream>>>(g, lo, count, deg, ref); break;
	case 1: k_header<1><<<grid_for(count, TPB), TPB, 0, L.stream>>>(g, lo, count, deg, ref); break;
	default: k_header<0><<<grid_for(count, TPB), TPB, 0, L.stream>>>(g, lo, count, deg, ref);
	}
	WGBV_LAUNCHED(L);
}

// ------------------------------------------------------------------------------------------------ stage 2
// Exclusive prefix sum, one pass, decoupled look-back.  Tiles are handed out through a ticket so that a tile
// only ever waits for tiles that are already running.  state[t] = flag (2 top bits) | value (62 bits):
// 1 = tile aggregate available, 2 = inclusive prefix available.
static constexpr int SCAN_ITEMS = 8;
static constexpr int SCAN_TILE = TPB * SCAN_ITEMS;
enum : uint64_t { FLAG_AGG = 1ull << 62, FLAG_PFX = 2ull << 62, FLAG_MASK = 3ull << 62 };

template <typename T>
__global__ void __launch_bounds__(TPB) k_scan(const T *__restrict__ in, int64_t *__restrict__ out, int64_t count,
                                              volatile uint64_t *state, uint32_t *ticket) {
	__shared__ uint32_t s_tile;
	__shared__ uint64_t s_warp[TPB / 32];
	__shared__ uint64_t s_base;
	if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
	__syncthreads();
	const uint32_t tile = s_tile;
	const int64_t base = (int64_t)tile * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;

	uint64_t v[SCAN_ITEMS];
#pragma unroll
	for (int j = 0; j < SCAN_ITEMS; j++) v[j] = base + j < count ? (uint64_t)in[base + j] : 0;
	uint64_t sum = 0;
#pragma unroll
	for (int j = 0; j < SCAN_ITEMS; j++) sum += v[j];

	// block-wide exclusive scan of the per-thread sums
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint64_t inc = sum;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const uint64_t t = __shfl_up_sync(0xffffffffu, inc, o);
		if (lane >= o) inc += t;
	}
	if (lane == 31) s_warp[warp] = inc;
	__syncthreads();
	if (warp == 0) {
		uint64_t w = lane < TPB / 32 ? s_warp[lane] : 0;
#pragma unroll
		for (int o = 1; o < TPB / 32; o <<= 1) {
			const uint64_t t = __shfl_up_sync(0xffffffffu, w, o);
			if (lane >= o) w += t;
		}
		const uint64_t tile_total = __shfl_sync(0xffffffffu, w, TPB / 32 - 1);
		if (lane < TPB / 32) s_warp[lane] = w;  // inclusive per-warp
		// look-back by the first warp
		uint64_t exclusive = 0;
		if (tile == 0) {
			if (lane == 0) { state[0] = FLAG_PFX | tile_total; }
		} else {
			if (lane == 0) { state[tile] = FLAG_AGG | tile_total; }
			int64_t look = (int64_t)tile - 1;
			for (;;) {
				const int64_t idx = look - lane;
				uint64_t s = idx >= 0 ? state[idx] : FLAG_PFX;
				// wait until every inspected predecessor has published something
				while (__any_sync(0xffffffffu, (s & FLAG_MASK) == 0)) {
					if ((s & FLAG_MASK) == 0) s = state[idx];
				}
				const unsigned pfx = __ballot_sync(0xffffffffu, (s & FLAG_MASK) == FLAG_PFX);
				const int first = pfx ? __ffs(pfx) - 1 : 32;  // nearest predecessor holding an inclusive prefix
				uint64_t contrib = (lane <= first && idx >= 0) ? (s & ~FLAG_MASK) : 0;
#pragma unroll
				for (int o = 16; o; o >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, o);
				exclusive += contrib;
				if (pfx) break;
				look -= 32;
			}
			if (lane == 0) { state[tile] = FLAG_PFX | (exclusive + tile_total); }
		}
		if (lane == 0) s_base = exclusive;
	}
	__syncthreads();
	uint64_t run = s_base + (warp ? s_warp[warp - 1] : 0) + (inc - sum);
#pragma unroll
	for (int j = 0; j < SCAN_ITEMS; j++) {
		if (base + j <= count) out[base + j] = (int64_t)run;
		run += v[j];
	}
}

template <typename T>
static void launch_scan_t(const Launch &L, const T *in, int64_t *out, int64_t count, uint64_t *state, uint32_t *ticket) {
	const int64_t items = count + 1;  // out has count+1 entries
	const unsigned tiles = grid_for(items, SCAN_TILE);
	cudaMemsetAsync(state, 0, (size_t)tiles * sizeof(uint64_t), L.stream);
	cudaMemsetAsync(ticket, 0, sizeof(uint32_t), L.stream);
	k_scan<T><<<tiles, TPB, 0, L.stream>>>(in, out, count, state, ticket);
	WGBV_LAUNCHED(L);
}
void launch_scan(const Launch &L, const uint32_t *deg, int64_t *rp, int64_t count, uint64_t *state, uint32_t *ticket) {
	launch_scan_t<uint32_t>(L, deg, rp, count, state, ticket);
}
void launch_scan64(const Launch &L, const unsigned long long *in, int64_t *out, int64_t count, uint64_t *state, uint32_t *ticket) {
	launch_scan_t<unsigned long long>(L, in, out, count, state, ticket);
}

__global__ void k_emit_rowptr(const int64_t *__restrict__ rp, int64_t skip, int64_t count, int64_t *__restrict__ out) {
	const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i > count) return;
	out[i] = rp[skip + i] - rp[skip];
}
void launch_emit_rowptr(const Launch &L, const int64_t *rp, int64_t skip, int64_t count, int64_t *rowptr_out) {
	k_emit_rowptr<<<grid_for(count + 1, TPB), TPB, 0, L.stream>>>(rp, skip, count, rowptr_out);
	WGBV_LAUNCHED(L);
}

// ------------------------------------------------------------------------------------------------ stage 2b
// depth[x] = 0 if x has no reference, else depth[x - ref] + 1.  Each pass follows a chain for at most
// DEPTH_HOPS hops or until it meets a node resolved by an earlier pass, so bounded graphs (maxrefcount small)
// finish in one pass and unbounded ones in ceil(longest chain / DEPTH_HOPS) passes.  Chains that leave the
// decodable range (possible for halo rows only) are marked invalid.
static constexpr int DEPTH_HOPS = 64;

__global__ void __launch_bounds__(TPB) k_depth(GraphDev g, const uint8_t *__restrict__ ref, uint32_t *depth, int64_t lo,
                                               int64_t from, int64_t count, int first_pass, uint32_t *stats) {
	const int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
	volatile uint32_t *vdepth = depth;
	uint32_t result = 0;
	if (i < count) {
		result = first_pass ? DEPTH_UNRESOLVED : vdepth[i];
		if (result == DEPTH_UNRESOLVED) {
			int64_t y = i;
			uint32_t hops = 0;
			for (int h = 0; h <= DEPTH_HOPS; h++) {
				const uint32_t rf = ref[y];
				if (rf == 0) { result = hops; break; }
				if (!first_pass && hops) {
					// resolved by an earlier pass, or earlier in this one: a 32-bit store is either seen whole or not at all
					const uint32_t dy = vdepth[y];
					if (dy != DEPTH_UNRESOLVED) { result = dy == DEPTH_INVALID ? DEPTH_INVALID : dy + hops; break; }
				}
				if (y - (int64_t)rf < 0) { result = DEPTH_INVALID; break; }
				y -= rf;
				hops++;
			}
			if (result == DEPTH_INVALID && lo + i >= from) raise(g, ERR_REF_UNRESOLVED);
			if (first_pass || result != DEPTH_UNRESOLVED) depth[i] = result;
		}
	}
	const bool unresolved = i < count && result == DEPTH_UNRESOLVED;
	const uint32_t m = (i < count && result < DEPTH_INVALID) ? result : 0;
	const uint32_t wm = __reduce_max_sync(0xffffffffu, m);
	const unsigned un = __ballot_sync(0xffffffffu, unresolved);
	if ((threadIdx.x & 31) == 0) {
		if (un) atomicAdd(&stats[0], (uint32_t)__popc(un));
		if (wm) atomicMax(&stats[1], wm);
	}
}

void launch_depth(const Launch &L, const GraphDev &g, const uint8_t *ref, uint32_t *depth, int64_t lo, int64_t from,
                  int64_t count, int first_pass, uint32_t *stats) {
	if (count <= 0) return;
	k_depth<<<grid_for(count, TPB), TPB, 0, L.stream>>>(g, ref, depth, lo, from, count, first_pass, stats);
	WGBV_LAUNCHED(L);
}

// ------------------------------------------------------------------------------------------------ stage 3
// Ordered level lists for shallow chains (at most LEVELS_MAX levels): rows of one level, in node order, so that the
// copy kernels find dense, memory-ordered work.  k_level_block_counts counts the rows of every level in blocks of
// LTILE nodes (layout: level-major), a scan turns the counts into positions, k_level_fill writes the row indices.
static constexpr int LEVELS_MAX = 8;
static constexpr int LTILE = TPB * 8;

// With `ref` given the kernel also COMPUTES the depths (chains of at most nlevels hops: a bounded walk over the
// reference array, which stage 1 left in memory) and stores them for the kernels that follow, so that the separate
// depth pass is not needed; a chain longer than nlevels sets *deep and the host falls back to k_depth.
__global__ void __launch_bounds__(TPB) k_level_block_counts(GraphDev g, const uint8_t *__restrict__ ref, uint32_t *depth, int64_t lo, int64_t from,
                                                            int64_t count, uint32_t level0, uint32_t nlevels, uint32_t nblocks,
                                                            uint32_t *__restrict__ counts, uint32_t *deep) {
	__shared__ uint32_t s_cnt[LEVELS_MAX + 1];
	if (threadIdx.x <= LEVELS_MAX) s_cnt[threadIdx.x] = 0;
	__syncthreads();
	const int64_t base = (int64_t)blockIdx.x * LTILE;
	for (int q = 0; q < 8; q++) {
		const int64_t i = base + q * TPB + threadIdx.x;
		uint32_t d = 0;
		if (i < count) {
			if (ref) {
				int64_t y = i;
				for (;;) {
					const uint32_t rf = ref[y];
					if (rf == 0) break;
					if (y < (int64_t)rf) { d = DEPTH_INVALID; break; }
					if (d == nlevels) { d = DEPTH_UNRESOLVED; atomicExch(deep, 1u); break; }
					y -= rf;
					d++;
				}
				if (d == DEPTH_INVALID && lo + i >= from) raise(g, ERR_REF_UNRESOLVED);
				depth[i] = d;
			} else d = depth[i];
		}
		d = (d > level0 && d < DEPTH_INVALID) ? d - level0 : 0u;  // levels level0+1 .. level0+nlevels are listed in this pass
		for (uint32_t l = 1; l <= nlevels; l++) {
			const unsigned m = __ballot_sync(0xffffffffu, d == l);
			if ((threadIdx.x & 31) == 0 && m) atomicAdd(&s_cnt[l], (uint32_t)__popc(m));
		}
	}
	__syncthreads();
	if (threadIdx.x >= 1 && threadIdx.x <= nlevels) counts[(size_t)(threadIdx.x - 1) * nblocks + blockIdx.x] = s_cnt[threadIdx.x];
}

__global__ void __launch_bounds__(TPB) k_level_fill(const uint32_t *__restrict__ depth, int64_t count, uint32_t level0, uint32_t nlevels,
                                                    uint32_t nblocks, const int64_t *__restrict__ pos, uint32_t *__restrict__ list) {
	__shared__ uint32_t s_cnt[8][TPB / 32][LEVELS_MAX + 1];  // rows of each level per (iteration, warp), then their prefix
	const int64_t base = (int64_t)blockIdx.x * LTILE;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	uint32_t myd[8], myrank[8];
	for (int q = 0; q < 8; q++) {
		const int64_t i = base + q * TPB + threadIdx.x;
		uint32_t d = i < count ? depth[i] : 0u;
		d = (d > level0 && d < DEPTH_INVALID) ? d - level0 : 0u;
		myd[q] = d;
		myrank[q] = 0;
		for (uint32_t l = 1; l <= nlevels; l++) {
			const unsigned m = __ballot_sync(0xffffffffu, d == l);
			if (d == l) myrank[q] = (uint32_t)__popc(m & ((1u << lane) - 1));
			if (lane == 0) s_cnt[q][warp][l] = (uint32_t)__popc(m);
		}
	}
	__syncthreads();
	if (threadIdx.x >= 1 && threadIdx.x <= nlevels) {
		const uint32_t l = threadIdx.x;
		uint32_t run = 0;
		for (int q = 0; q < 8; q++)
			for (int w = 0; w < TPB / 32; w++) {
				const uint32_t c = s_cnt[q][w][l];
				s_cnt[q][w][l] = run;
				run += c;
			}
	}
	__syncthreads();
	for (int q = 0; q < 8; q++) {
		const uint32_t d = myd[q];
		if (d >= 1 && d <= nlevels) {
			const int64_t i = base + q * TPB + threadIdx.x;
			list[pos[(size_t)(d - 1) * nblocks + blockIdx.x] + s_cnt[q][warp][d] + myrank[q]] = (uint32_t)i;
		}
	}
}

void launch_level_lists(const Launch &L, const GraphDev &g, const uint8_t *ref, uint32_t *depth, int64_t lo, int64_t from, int64_t count,
                        uint32_t level0, uint32_t nlevels, uint32_t *counts, int64_t *pos, uint32_t *list, uint64_t *scan_state, uint32_t *ticket,
                        uint32_t *deep) {
	const uint32_t nblocks = grid_for(count, LTILE);
	k_level_block_counts<<<nblocks, TPB, 0, L.stream>>>(g, ref, depth, lo, from, count, level0, nlevels, nblocks, counts, deep);
	WGBV_LAUNCHED(L);
	launch_scan(L, counts, pos, (int64_t)nlevels * nblocks, scan_state, ticket);
	k_level_fill<<<nblocks, TPB, 0, L.stream>>>(depth, count, level0, nlevels, nblocks, pos, list);
	WGBV_LAUNCHED(L);
}

// Deep chains (depths from k_depth, any number of levels): the same lists, eight levels per pass, over the rows that
// are still to do -- `ids` (increasing row indices; nullptr: all rows of the range).  Bucket l = 1..nlevels takes the rows
// of depth level0 + l, bucket nlevels + 1 the rows that lie deeper: that list is the `ids` of the next pass, so that a
// pass costs what its rows cost and not a sweep over the whole range (most rows are within the first eight levels).
__global__ void __launch_bounds__(TPB) k_deep_block_counts(const uint32_t *__restrict__ depth, const uint32_t *__restrict__ ids, int64_t count,
                                                           uint32_t level0, uint32_t nlevels, uint32_t nblocks, uint32_t *__restrict__ counts) {
	__shared__ uint32_t s_cnt[LEVELS_MAX + 2];
	if (threadIdx.x <= LEVELS_MAX + 1) s_cnt[threadIdx.x] = 0;
	__syncthreads();
	const int64_t base = (int64_t)blockIdx.x * LTILE;
	for (int q = 0; q < 8; q++) {
		const int64_t t = base + q * TPB + threadIdx.x;
		uint32_t d = 0;
		if (t < count) d = depth[ids ? ids[t] : (uint32_t)t];
		d = (d > level0 && d < DEPTH_INVALID) ? (d - level0 <= nlevels ? d - level0 : nlevels + 1) : 0u;
		for (uint32_t l = 1; l <= nlevels + 1; l++) {
			const unsigned m = __ballot_sync(0xffffffffu, d == l);
			if ((threadIdx.x & 31) == 0 && m) atomicAdd(&s_cnt[l], (uint32_t)__popc(m));
		}
	}
	__syncthreads();
	if (threadIdx.x >= 1 && threadIdx.x <= nlevels + 1) counts[(size_t)(threadIdx.x - 1) * nblocks + blockIdx.x] = s_cnt[threadIdx.x];
}

__global__ void __launch_bounds__(TPB) k_deep_fill(const uint32_t *__restrict__ depth, const uint32_t *__restrict__ ids, int64_t count, uint32_t level0,
                                                   uint32_t nlevels, uint32_t nblocks, const int64_t *__restrict__ pos, uint32_t *__restrict__ list) {
	__shared__ uint32_t s_cnt[8][TPB / 32][LEVELS_MAX + 2];
	const int64_t base = (int64_t)blockIdx.x * LTILE;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	uint32_t myd[8], myrank[8], myrow[8];
	for (int q = 0; q < 8; q++) {
		const int64_t t = base + q * TPB + threadIdx.x;
		uint32_t d = 0;
		myrow[q] = 0;
		if (t < count) {
			myrow[q] = ids ? ids[t] : (uint32_t)t;
			d = depth[myrow[q]];
		}
		d = (d > level0 && d < DEPTH_INVALID) ? (d - level0 <= nlevels ? d - level0 : nlevels + 1) : 0u;
		myd[q] = d;
		myrank[q] = 0;
		for (uint32_t l = 1; l <= nlevels + 1; l++) {
			const unsigned m = __ballot_sync(0xffffffffu, d == l);
			if (d == l) myrank[q] = (uint32_t)__popc(m & ((1u << lane) - 1));
			if (lane == 0) s_cnt[q][warp][l] = (uint32_t)__popc(m);
		}
	}
	__syncthreads();
	if (threadIdx.x >= 1 && threadIdx.x <= nlevels + 1) {
		const uint32_t l = threadIdx.x;
		uint32_t run = 0;
		for (int q = 0; q < 8; q++)
			for (int w = 0; w < TPB / 32; w++) {
				const uint32_t c = s_cnt[q][w][l];
				s_cnt[q][w][l] = run;
				run += c;
			}
	}
	__syncthreads();
	for (int q = 0; q < 8; q++) {
		const uint32_t d = myd[q];
		if (d >= 1) list[pos[(size_t)(d - 1) * nblocks + blockIdx.x] + s_cnt[q][warp][d] + myrank[q]] = myrow[q];
	}
}

// counts: (nlevels + 1) * nblocks words, pos: one more; nblocks = deep_list_blocks(count)
uint32_t deep_list_blocks(int64_t count) { return grid_for(count, LTILE); }
void launch_deep_lists(const Launch &L, const uint32_t *depth, const uint32_t *ids, int64_t count, uint32_t level0, uint32_t nlevels, uint32_t *counts,
                       int64_t *pos, uint32_t *list, uint64_t *scan_state, uint32_t *ticket) {
	const uint32_t nblocks = deep_list_blocks(count);
	if (!nblocks) return;
	k_deep_block_counts<<<nblocks, TPB, 0, L.stream>>>(depth, ids, count, level0, nlevels, nblocks, counts);
	WGBV_LAUNCHED(L);
	launch_scan(L, counts, pos, (int64_t)(nlevels + 1) * nblocks, scan_state, ticket);
	k_deep_fill<<<nblocks, TPB, 0, L.stream>>>(depth, ids, count, level0, nlevels, nblocks, pos, list);
	WGBV_LAUNCHED(L);
}

__device__ __forceinline__ int64_t *row_slot(const RangeDev &R, int64_t i, int64_t rp_from) {
	const int64_t p = R.rp[i];
	return (R.lo + i >= R.from) ? R.col + (p - rp_from) : R.halo + p;
}

static constexpr int RPT = 8;            // rows per thread
static constexpr int TILE = TPB * RPT;   // rows per block: the sort domain (see the comment on work binning)

// Rows with at least HEAVY (128) extras are not decoded inside their tile: a tile holds only a handful of them, of very
// different lengths, and whichever warp gets them runs nearly empty.  They are appended to a global list, sorted
// by class over the whole range (k_heavy_scatter), and decoded by k_extras_heavy where every warp finds 32 rows of
// one class.
#ifndef WGBV_HEAVY
#define WGBV_HEAVY 128
#endif
static constexpr uint32_t HEAVY = WGBV_HEAVY;
uint32_t heavy_threshold() { return HEAVY; }

// stage 3a: one block per ETILE consecutive nodes.  Rows without a reference come out complete.
// The records of consecutive rows are contiguous in the bit stream, so the block first copies the tile's whole span
// into shared memory with coalesced loads (STAGED) and every decoder of the block then reads words with ~30-cycle
// shared-memory latency instead of waiting for L2/DRAM at every word boundary of every lane; a tile whose span does
// not fit (a very long row in it) reads global memory instead.
#ifndef WGBV_ERPT
#define WGBV_ERPT 3      // 768 rows per tile, five blocks per SM: 28.59 ms per 1e8 nodes against 28.81 with 1024 rows and four
#define WGBV_EBITS 6912
#endif
static constexpr int ERPT = WGBV_ERPT;
static constexpr int ETILE = TPB * ERPT;
static constexpr int EBITS = WGBV_EBITS;  // 32-bit words of the stream a tile may stage (27 KiB: 9 words per row)

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src) {
	const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src));
}

struct ExtrasSmem {
	uint32_t bits[EBITS];
	uint64_t pos[ETILE];
	uint32_t nextra[ETILE], ncopied[ETILE];
	uint32_t hist[NBUCKETS + 1];
	uint32_t count[BIN_CLASSES];
	uint16_t order[ETILE];
	uint8_t cls[ETILE];
	unsigned long long gbase;
	uint64_t span[2];
	uint64_t wbase;  // stream word index of bits[0]
	uint32_t nw;     // words staged
	// one-launch front end.  (The outdegrees of the rows just before the tile, reference rows of its first rows, are
	// kept in `order`, which is not needed before the sort: 256 uint32 -- see hdeg().  Five blocks of this size fit an SM.)
	uint64_t warp_sum[TPB / 32];
	uint64_t tile_base;          // arcs before the tile
	uint64_t tile_total;         // arcs of the tile
	uint32_t tile;
};

static_assert(sizeof(ExtrasSmem) <= 44 * 1024, "five blocks of k_extras per SM");
static_assert(ETILE * sizeof(uint16_t) >= 256 * sizeof(uint32_t), "hdeg() borrows ExtrasSmem::order");
__device__ __forceinline__ uint32_t *hdeg(ExtrasSmem &S) { return reinterpret_cast<uint32_t *>(S.order); }

template <class RD>
__device__ __forceinline__ void tile_reader(RD &r, const GraphDev &g, const ExtrasSmem &S, uint64_t pos);
template <>
__device__ __forceinline__ void tile_reader<BitReader>(BitReader &r, const GraphDev &g, const ExtrasSmem &, uint64_t pos) { r.init(g, pos); }
template <>
__device__ __forceinline__ void tile_reader<TileReader>(TileReader &r, const GraphDev &, const ExtrasSmem &S, uint64_t pos) {
	r.init_words(S.bits, S.nw, (uint32_t)(pos - (S.wbase << 5)));
}
#ifndef WGBV_STAGED_READER
#define WGBV_STAGED_READER PosReader  // the reader of a staged tile: position only, the words are loaded at every peek
#endif
using StagedReader = WGBV_STAGED_READER;
template <>
__device__ __forceinline__ void tile_reader<PosReader>(PosReader &r, const GraphDev &, const ExtrasSmem &S, uint64_t pos) {
	r.init_words(S.bits, S.nw, (uint32_t)(pos - (S.wbase << 5)));
}

// Stages 1 and 2 for the rows of a tile, inside k_extras (RangeDev::rp_w): outdegree and reference of every row, then
// the exclusive scan of the outdegrees -- inside the tile through shared memory, across tiles by decoupled look-back
// on tile_state (tiles are handed out by a ticket, so a tile only waits for tiles that are running).  Leaves deg, ref
// and rp of the tile's rows in global memory, the rows' offsets in S.pos and the outdegrees of the window before the
// tile in S.hdeg.  Same checks as k_header.
template <int DEF, class RD>
__device__ __forceinline__ void tile_front_end(const GraphDev &g, const RangeDev &R, ExtrasSmem &S, int64_t base, int64_t count, uint64_t &rel) {
	const int64_t tend = base + ETILE < count ? base + ETILE : count;
	// outdegrees of the rows before the tile: not in the staged bits, a handful of global reads
	if ((int)threadIdx.x < g.window && (int64_t)threadIdx.x < base && threadIdx.x < 256) {
		BitReader r;
		r.init(g, ef_get(g, R.lo + base - 1 - threadIdx.x));
		hdeg(S)[threadIdx.x] = (uint32_t)rd_outdegree<DEF>(r, g);
	}
#pragma unroll 1
	for (int q = 0; q < ERPT; q++) {
		const int id = q * TPB + threadIdx.x;
		const int64_t i = base + id;
		uint32_t dd = 0;
		if (i < count) {
			const int64_t x = R.lo + i;
			const uint64_t off = ef_get(g, x);
			RD r;
			tile_reader<RD>(r, g, S, off);
			const uint64_t d = rd_outdegree<DEF>(r, g);
			uint64_t rf = 0;
			if (d > 0x7fffffffull) raise(g, ERR_DEGREE);
			else {
				dd = (uint32_t)d;
				if (d && g.window > 0) {
					rf = rd_reference<DEF>(r, g);
					if (rf > (uint64_t)g.window || (int64_t)rf > x) { raise(g, ERR_REF_WINDOW); rf = 0; }
				}
				if (r.past_end()) raise(g, ERR_OVERRUN);
			}
			R.deg_w[i] = dd;
			R.ref_w[i] = (uint8_t)rf;
			S.pos[id] = r.tell() + (RD::IS_GLOBAL ? 0ull : (S.wbase << 5));  // after the reference: where the head loop resumes
		}
		S.nextra[id] = dd;  // (nextra is free until the heads are parsed)
		__syncwarp();
	}
	__syncthreads();
	// scan in row order: thread t owns rows ERPT * t .. ERPT * t + ERPT - 1 of the tile
	uint64_t v[ERPT], sum = 0;
#pragma unroll
	for (int j = 0; j < ERPT; j++) { v[j] = S.nextra[ERPT * threadIdx.x + j]; sum += v[j]; }
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint64_t inc = sum;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const uint64_t t = __shfl_up_sync(FULL_MASK, inc, o);
		if (lane >= o) inc += t;
	}
	if (lane == 31) S.warp_sum[warp] = inc;
	__syncthreads();
	if (warp == 0) {
		uint64_t w = lane < TPB / 32 ? S.warp_sum[lane] : 0;
#pragma unroll
		for (int o = 1; o < TPB / 32; o <<= 1) {
			const uint64_t t = __shfl_up_sync(FULL_MASK, w, o);
			if (lane >= o) w += t;
		}
		const uint64_t tile_total = __shfl_sync(FULL_MASK, w, TPB / 32 - 1);
		if (lane < TPB / 32) S.warp_sum[lane] = w;  // inclusive per warp
		// the tile's total is published at once; the arcs before the tile are looked up later (tile_finish), when the
		// heads of the tile have been parsed: nothing before that needs them, and by then the tiles before this one have
		// published theirs -- a block that waits here waits for the slowest of its 32 predecessors
		volatile uint64_t *state = R.tile_state;
		if (lane == 0) {
			S.tile_total = tile_total;
			state[S.tile] = (S.tile == 0 ? FLAG_PFX : FLAG_AGG) | tile_total;
		}
	}
	__syncthreads();
	rel = (warp ? S.warp_sum[warp - 1] : 0) + (inc - sum);
}

// Second half of the front end: arcs before the tile by decoupled look-back (warp 0; `st` = the states of the 32 tiles
// before this one, read before the sort so that the latency is spent there), then the row pointers of the tile.
__device__ __forceinline__ void tile_finish(const GraphDev &g, const RangeDev &R, ExtrasSmem &S, int64_t base, int64_t count, uint64_t rel, uint64_t st) {
	const int64_t tend = base + ETILE < count ? base + ETILE : count;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	if (warp == 0) {
		volatile uint64_t *state = R.tile_state;
		const uint32_t tile = S.tile;
		uint64_t exclusive = 0;
		if (tile != 0) {
			int64_t look = (int64_t)tile - 1;
			for (;;) {
				const int64_t idx = look - lane;
				while (__any_sync(FULL_MASK, (st & FLAG_MASK) == 0))
					if ((st & FLAG_MASK) == 0) st = state[idx];
				const unsigned pfx = __ballot_sync(FULL_MASK, (st & FLAG_MASK) == FLAG_PFX);
				const int first = pfx ? __ffs(pfx) - 1 : 32;
				uint64_t contrib = (lane <= first && idx >= 0) ? (st & ~FLAG_MASK) : 0;
#pragma unroll
				for (int o = 16; o; o >>= 1) contrib += __shfl_xor_sync(FULL_MASK, contrib, o);
				exclusive += contrib;
				if (pfx) break;
				look -= 32;
				st = look - lane >= 0 ? state[look - lane] : FLAG_PFX;
			}
			if (lane == 0) state[tile] = FLAG_PFX | (exclusive + S.tile_total);
		}
		if (lane == 0) S.tile_base = exclusive;
	}
	__syncthreads();
	uint64_t run = S.tile_base + rel;
	const int64_t i0 = base + ERPT * (int64_t)threadIdx.x;
#pragma unroll
	for (int j = 0; j < ERPT; j++) {
		const int64_t i = i0 + j;
		if (i < tend) {
			R.rp_w[i] = (int64_t)run;
			run += R.deg[i];
		}
	}
	// rp[tend]: the end of the tile's last row (the next tile writes the same value there; for the last tile, the total)
	if (i0 < tend && i0 + ERPT >= tend) {
		R.rp_w[tend] = (int64_t)run;
		// the arcs up to the end of this tile do not fit the caller's array: reported here for the whole tile (rows without
		// extras never reach the per-row check of the decode loop; no kernel writes behind the capacity)
		if (R.lo + tend > R.from && (int64_t)run > R.rp_from_w && run - (uint64_t)R.rp_from_w > R.col_capacity) raise(g, ERR_CAPACITY);
	}
	__syncthreads();
}

template <int DEF, class RD, bool FUSED>
__device__ __forceinline__ void extras_tile(const GraphDev &g, const RangeDev &R, const HeavyDev &H, ExtrasSmem &S, int64_t base, int64_t count,
                                            int64_t rp_from) {
	uint64_t rel = 0;
	if (FUSED) tile_front_end<DEF, RD>(g, R, S, base, count, rel);
	// heads of my rows (uniform loop: the warp reconverges at the end of every iteration)
	int bucket[ERPT];
#pragma unroll 1
	for (int q = 0; q < ERPT; q++) {
		const int id = q * TPB + threadIdx.x;
		const int64_t i = base + id;
		bucket[q] = -1;
		// a halo row whose reference lies before the decoded range is never needed (and cannot be decoded): skip it
		if (i < count && R.deg[i] != 0 && (R.depth ? R.depth[i] < DEPTH_INVALID : (int64_t)R.ref[i] <= i)) {
			const int64_t x = R.lo + i;
			// (front end inside this kernel: the row pointers are not known yet, the rows are checked before they are decoded)
			if (!FUSED && x >= R.from && (uint64_t)(R.rp[i + 1] - rp_from) > R.col_capacity) raise(g, ERR_CAPACITY);
			else {
				RD r;
				tile_reader<RD>(r, g, S, FUSED ? S.pos[id] : ef_get(g, x));
				RowHead h;
				const uint32_t rf = R.ref[i];
				// (front end inside this kernel: rows before the tile belong to another block, their outdegrees are in S.hdeg;
				// outdegree and reference have been read already and the reader resumes after them)
				const uint32_t dref = !rf ? 0 : (FUSED && i - rf < base) ? hdeg(S)[base - 1 - (i - rf)] : R.deg[i - rf];
				if (FUSED) row_blocks<DEF>(g, r, R.deg[i], rf, dref, h);
				else row_head<DEF>(g, r, dref, h);
				if (rf && R.info) {
					RowInfo ri;
					if (row_is_masked(h.d, dref)) {
						ri.a = h.cmask[0];
						ri.b = h.cmask[1];
					} else {
						ri.a = h.blocks_pos + (RD::IS_GLOBAL ? 0ull : (S.wbase << 5));
						ri.b = (uint64_t)h.ncopied | (uint64_t)h.bc << 32;
					}
					R.info[i] = ri;
				}
				const uint32_t nextra = h.d - h.ncopied;
				if (nextra) {
					// rows with intervals first inside each class: warps without any skip the merging loop of row_extras
					bool has_int = false;
					if (g.min_interval != 0) { RD t = r; has_int = t.gamma() != 0; }
					// sort key: ALL rows with intervals first, then by work -- warps whose rows have no intervals (four out of five)
					// never enter the merging loop of row_extras_core, which costs twice the plain gap loop per successor
					bucket[q] = (has_int ? 0 : BIN_CLASSES) + work_bucket(nextra);
					S.pos[id] = r.tell() + (RD::IS_GLOBAL ? 0ull : (S.wbase << 5));  // absolute: tile_reader() converts back
					S.nextra[id] = nextra;
					S.ncopied[id] = h.ncopied;
					S.cls[id] = (uint8_t)(2 * work_bucket(nextra) + (has_int ? 0 : 1));  // class in the global heavy-row list
				}
			}
		}
		__syncwarp();
	}
	uint64_t st = FLAG_PFX;
	if (FUSED && threadIdx.x < 32 && S.tile != 0) {
		const int64_t idx = (int64_t)S.tile - 1 - (int64_t)threadIdx.x;
		if (idx >= 0) st = const_cast<volatile uint64_t *>(R.tile_state)[idx];
	}
	uint32_t total;
	block_bin_sort<ERPT>(bucket, S.hist, S.order, &total);
	if (FUSED) tile_finish(g, R, S, base, count, rel, st);
	// after the sort S.hist[b] is the first position of bucket b: rows with intervals in buckets [0, 64), the others in
	// [64, 128), each half heaviest first; the first fl buckets of either half hold the rows of >= HEAVY extras
	const int fl = H.rows ? work_bucket(HEAVY - 1) : 0;
	const uint32_t a0 = S.hist[fl], a1 = S.hist[BIN_CLASSES], b0 = S.hist[BIN_CLASSES + fl];
	const uint32_t nheavy_a = a0, nheavy = a0 + (b0 - a1);
	if (nheavy) {
		// per-class counts of the global list: class 2c (with intervals) = bucket c, class 2c + 1 = bucket 64 + c
		if ((int)threadIdx.x < 2 * fl) {
			const int bk = (threadIdx.x & 1 ? BIN_CLASSES : 0) + (int)(threadIdx.x >> 1);
			S.count[threadIdx.x] = S.hist[bk + 1] - S.hist[bk];
		}
		if (threadIdx.x == 0) S.gbase = atomicAdd(&H.counters[0], (unsigned long long)nheavy);
		__syncthreads();
		if ((int)threadIdx.x < 2 * fl && S.count[threadIdx.x]) atomicAdd(&H.counters[1 + threadIdx.x], (unsigned long long)S.count[threadIdx.x]);
		// (the list is sized from the arc count of the .properties: a file that holds more than it promises is reported,
		// never written past the list; the heavy-row kernels then do nothing)
		if (S.gbase + nheavy > H.capacity && threadIdx.x == 0) raise(g, ERR_DEGREE);
		for (uint32_t t = threadIdx.x; t < nheavy; t += TPB) {
			const int id = S.order[t < nheavy_a ? t : a1 + (t - nheavy_a)];
			if (S.gbase + t < H.capacity) {
				H.rows[S.gbase + t] = (uint32_t)(base + id);
				H.cls[S.gbase + t] = S.cls[id];
			}
		}
	}
	// serial part: the lanes of a warp take rows of the same class and run them in lock step (first the light rows with
	// intervals, positions [a0, a1) of the order, then those without, [b0, total))
	const uint32_t la = a1 - a0, nlight = la + (total - b0);
#pragma unroll 1
	for (uint32_t tb = (threadIdx.x & ~31u); tb < nlight; tb += TPB) {
		const uint32_t t = tb + (threadIdx.x & 31u);
		const bool active = t < nlight;
		const unsigned amask = __ballot_sync(FULL_MASK, active);
		int id = 0;
		bool ok = active;
		if (active) {
			id = S.order[t < la ? a0 + t : b0 + (t - la)];
			if (FUSED && R.lo + base + id >= R.from && (uint64_t)(R.rp[base + id + 1] - rp_from) > R.col_capacity) {
				raise(g, ERR_CAPACITY);
				ok = false;
			}
		}
		const unsigned mask = FUSED ? __ballot_sync(FULL_MASK, ok) : amask;
		if (ok) {
			const int64_t k = base + id;
			RD r;
			tile_reader<RD>(r, g, S, S.pos[id]);
			row_extras<DEF>(g, r, R.lo + k, S.nextra[id], row_slot(R, k, rp_from) + S.ncopied[id], mask);
		}
	}
}

template <int DEF, bool FUSED>
__global__ void __launch_bounds__(TPB) k_extras(GraphDev g, RangeDev R, HeavyDev H) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	ExtrasSmem &S = *reinterpret_cast<ExtrasSmem *>(smem_raw);
	if (FUSED) {
		if (threadIdx.x == 0) S.tile = atomicAdd(R.tile_ticket, 1u);
		__syncthreads();
	}
	const int64_t base = (int64_t)(FUSED ? S.tile : blockIdx.x) * ETILE;
	const int64_t count = R.hi - R.lo;
	const int64_t rp_from = FUSED ? R.rp_from_w : R.rp[R.from - R.lo];  // (front end inside this kernel: rp is not written yet)
	const int64_t tend = base + ETILE < count ? base + ETILE : count;
	if (threadIdx.x == 0) S.span[0] = ef_get(g, R.lo + base);
	if (threadIdx.x == 32) S.span[1] = ef_get(g, R.lo + tend);
	// everything the heads are going to read, asked for up front (the reads themselves are dependent chains):
	// degrees, references and row pointers of the tile, and the tile's part of the Elias-Fano arrays
	{
		const char *p;
		const int t = threadIdx.x;
		if (!FUSED) {
			p = (const char *)(R.deg + base) + (size_t)t * 128;
			if (p < (const char *)(R.deg + tend)) prefetch_l1(p);
			p = (const char *)(R.rp + base) + (size_t)t * 128;
			if (p < (const char *)(R.rp + tend + 1)) prefetch_l1(p);
			p = (const char *)(R.ref + base) + (size_t)t * 128;
			if (p < (const char *)(R.ref + tend)) prefetch_l1(p);
		}
		const uint64_t e0 = (uint64_t)(R.lo + base - g.ef_first), e1 = (uint64_t)(R.lo + tend - g.ef_first);
		p = (const char *)(g.ef_lower + (e0 * (uint64_t)g.ef_l >> 6)) + (size_t)t * 128;
		if (p <= (const char *)(g.ef_lower + (e1 * (uint64_t)g.ef_l >> 6))) prefetch_l1(p);
		p = (const char *)(g.ef_hints + (e0 >> 6)) + (size_t)t * 128;
		if (p <= (const char *)(g.ef_hints + (e1 >> 6))) prefetch_l1(p);
	}
	__syncthreads();
	// the upper-bits words of the tile lie between the positions of its first and last element
	{
		const uint64_t e0 = (uint64_t)(R.lo + base - g.ef_first), e1 = (uint64_t)(R.lo + tend - g.ef_first);
		const uint64_t u0 = ((S.span[0] >> g.ef_l) + e0) >> 6, u1 = ((S.span[1] >> g.ef_l) + e1) >> 6;
		const char *p = (const char *)(g.ef_upper + u0) + (size_t)threadIdx.x * 128;
		if (p <= (const char *)(g.ef_upper + u1)) prefetch_l1(p);
	}
	const uint64_t w0 = (S.span[0] >> 5) & ~3ull;  // 16-byte aligned: the staging copies are 16 bytes each
	const uint64_t nw = ((((S.span[1] + 31) >> 5) - w0 + 2) + 3) & ~3ull;
	const bool staged = nw <= (uint64_t)EBITS;
	if (staged) {
		const uint64_t avail = g.nwords * 2 > w0 ? g.nwords * 2 - w0 : 0;
		const uint32_t *src = (const uint32_t *)g.words + w0;
		// asynchronous 16-byte copies (LDGSTS): all of them in flight before anything waits
		for (uint32_t w = threadIdx.x * 4; w < (uint32_t)nw; w += TPB * 4) {
			if (w + 4 <= avail) cp_async16(&S.bits[w], src + w);
			else
				for (int j = 0; j < 4; j++) S.bits[w + j] = w + j < avail ? __ldg(src + w + j) : 0u;
		}
		asm volatile("cp.async.wait_all;" ::: "memory");
		if (threadIdx.x == 0) { S.wbase = w0; S.nw = (uint32_t)nw; }
	}
	__syncthreads();
	if (staged) extras_tile<DEF, StagedReader, FUSED>(g, R, H, S, base, count, rp_from);
	else extras_tile<DEF, BitReader, FUSED>(g, R, H, S, base, count, rp_from);
}

uint32_t extras_tiles(int64_t count) { return grid_for(count, ETILE); }

// class starts of the global heavy list: counters[1..64] = counts -> counters[65..128] = running cursors
__global__ void k_heavy_offsets(HeavyDev H) {
	const unsigned lane = threadIdx.x;
	const unsigned long long a = H.counters[1 + 2 * lane], b = H.counters[2 + 2 * lane];
	unsigned long long inc = a + b;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const unsigned long long t = __shfl_up_sync(0xffffffffu, inc, o);
		if ((int)lane >= o) inc += t;
	}
	const unsigned long long ex = inc - a - b;
	H.counters[65 + 2 * lane] = ex;
	H.counters[66 + 2 * lane] = ex + a;
}

// counting-sort scatter of the heavy list by class (block-aggregated reservations)
__global__ void __launch_bounds__(TPB) k_heavy_scatter(HeavyDev H) {
	__shared__ uint32_t s_cnt[BIN_CLASSES];
	__shared__ unsigned long long s_start[BIN_CLASSES];
	const unsigned long long total = H.counters[0];
	if (total > H.capacity) return;  // reported by k_extras
	for (unsigned long long tile = (unsigned long long)blockIdx.x * TILE; tile < total; tile += (unsigned long long)gridDim.x * TILE) {
		if (threadIdx.x < BIN_CLASSES) s_cnt[threadIdx.x] = 0;
		__syncthreads();
		uint32_t rank[RPT];
		int cls[RPT];
#pragma unroll
		for (int q = 0; q < RPT; q++) {
			const unsigned long long t = tile + q * TPB + threadIdx.x;
			cls[q] = t < total ? (int)H.cls[t] : -1;
			rank[q] = cls[q] >= 0 ? atomicAdd(&s_cnt[cls[q]], 1u) : 0u;
		}
		__syncthreads();
		if (threadIdx.x < BIN_CLASSES && s_cnt[threadIdx.x])
			s_start[threadIdx.x] = atomicAdd(&H.counters[65 + threadIdx.x], (unsigned long long)s_cnt[threadIdx.x]);
		__syncthreads();
#pragma unroll
		for (int q = 0; q < RPT; q++)
			if (cls[q] >= 0) H.sorted[s_start[cls[q]] + rank[q]] = H.rows[tile + q * TPB + threadIdx.x];
		__syncthreads();
	}
}

// stage 3a for heavy rows: every warp takes 32 consecutive entries of the class-sorted list.
// Each lane walks its own long stream of residual codes, 32 different places of the bit stream per warp: read
// directly, every lane would fetch and re-fetch its own cache lines (L1 cannot hold 32 streams per warp) and the
// warp would wait for the slowest lane at every word boundary.  Instead the warp keeps a WINDOW of the next
// HW_WORDS words of every lane's stream in shared memory, filled by all 32 lanes together with one coalesced
// 128-byte load per row; the decoders read the window and the warp refills it whenever some lane runs low.
#ifndef WGBV_HW_WORDS
#define WGBV_HW_WORDS 32
#endif
static constexpr int HW_WORDS = WGBV_HW_WORDS;   // words of stream per lane and refill (a multiple of 32)
static constexpr int HW_PITCH = HW_WORDS + 3;    // row pitch in words (odd: the 32 lanes hit 32 banks)
static constexpr uint32_t HW_LOW = HW_WORDS - 3; // refill when a reader's prefetched word index reaches this
#ifndef WGBV_HEAVY_BLOCKS
#define WGBV_HEAVY_BLOCKS 2
#endif
#ifndef WGBV_GIANT_ROW
#define WGBV_GIANT_ROW 4096
#endif
static constexpr uint32_t GIANT_ROW = WGBV_GIANT_ROW;   // residuals from which a row is decoded by a whole warp
static constexpr uint32_t GIANT_ROW2 = 32768;  // ... and from which by GIANT_CHUNKS warps (k_giant_*)

__device__ __forceinline__ void cp_async4(void *smem_dst, const void *gmem_src) {
	const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
	asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(gmem_src));
}

struct HeavyWindow {
	static constexpr bool enabled = true;
	uint32_t *win;          // this warp's HW_PITCH * 32 words
	const uint32_t *words;  // the whole stream
	uint64_t nwords32;
	uint64_t wbase;         // stream word index of my row's window[0]
	__device__ __forceinline__ bool low(const TileReader &r) const { return r.idx >= HW_LOW; }
	// all lanes of the warp call this together; lanes outside `mask` take part in the loads only
	__device__ __forceinline__ void fill(TileReader &r, uint64_t pos, bool mine) {
		const unsigned lane = threadIdx.x & 31u;
		const uint64_t my_w = pos >> 5;
		// asynchronous copies (LDGSTS): the loads of all 32 rows are in flight together and nothing waits for a register --
		// with plain loads the stores to shared memory made every group of four wait out a memory latency of its own
#pragma unroll 4
		for (int rr = 0; rr < 32; rr++) {
			const uint64_t w = __shfl_sync(FULL_MASK, my_w, rr);
#pragma unroll
			for (int k = 0; k < HW_WORDS; k += 32) {
				const uint64_t j = w + k + lane;
				if (j < nwords32) cp_async4(&win[rr * HW_PITCH + k + lane], words + j);
				else win[rr * HW_PITCH + k + lane] = 0u;
			}
		}
		asm volatile("cp.async.wait_all;" ::: "memory");
		__syncwarp();
		if (mine) {
			wbase = my_w;
			r.init_words(win + lane * HW_PITCH, HW_WORDS, (uint32_t)(pos & 31u));
		}
	}
	__device__ __forceinline__ uint64_t tell(const TileReader &r) const { return (wbase << 5) + r.tell(); }
	__device__ __forceinline__ void refill(TileReader &r, unsigned mask) {
		// only lanes of `mask` are here; the others idle in the caller, so the loads are done by the lanes present
		const unsigned lane = threadIdx.x & 31u;
		const uint64_t pos = tell(r);
		const uint64_t my_w = pos >> 5;
		__syncwarp(mask);
		for (unsigned m = mask; m; m &= m - 1) {
			const int rr = __ffs(m) - 1;
			const uint64_t w = __shfl_sync(mask, my_w, rr);
			for (unsigned l = (unsigned)__popc(mask & ((1u << lane) - 1u)); l < (unsigned)HW_WORDS; l += (unsigned)__popc(mask)) {
				const uint64_t j = w + l;
				if (j < nwords32) cp_async4(&win[rr * HW_PITCH + l], words + j);
				else win[rr * HW_PITCH + l] = 0u;
			}
		}
		asm volatile("cp.async.wait_all;" ::: "memory");
		__syncwarp(mask);
		wbase = my_w;
		r.init_words(win + lane * HW_PITCH, HW_WORDS, (uint32_t)(pos & 31u));
	}
};

template <int DEF>
__global__ void __launch_bounds__(TPB, WGBV_HEAVY_BLOCKS) k_extras_heavy(GraphDev g, RangeDev R, HeavyDev H) {
	extern __shared__ __align__(16) uint32_t s_win[];  // [TPB / 32][32 * HW_PITCH]
	const unsigned long long total = H.counters[0];
	if (total > H.capacity) return;  // reported by k_extras: the list is incomplete
	const int64_t rp_from = R.rp[R.from - R.lo];
	const unsigned long long warps = (unsigned long long)gridDim.x * (TPB / 32);
	const unsigned lane = threadIdx.x & 31;
	HeavyWindow win;
	win.win = s_win + (threadIdx.x >> 5) * (32 * HW_PITCH);
	win.words = (const uint32_t *)g.words;
	win.nwords32 = g.nwords * 2;
	win.wbase = 0;
	for (unsigned long long c = (unsigned long long)blockIdx.x * (TPB / 32) + (threadIdx.x >> 5); c * 32 < total; c += warps) {
		const unsigned long long t = c * 32 + lane;
		// (a row that does not fit the caller's capacity has been reported already, by k_extras or, with the front end inside
		// k_extras, right here: it is skipped)
		bool active = t < total;
		if (active) {
			const int64_t ih = H.sorted[t];
			if (R.lo + ih >= R.from && (uint64_t)(R.rp[ih + 1] - rp_from) > R.col_capacity) {
				raise(g, ERR_CAPACITY);
				active = false;
			}
		}
		const unsigned mask = __ballot_sync(FULL_MASK, active);
		// heads and interval sections straight from global memory (short), then the residuals through the window
		int64_t i = 0;
		uint32_t nextra = 0, ic = 0, ic0 = 0, nres = 0, ncopied = 0;
		uint64_t int_pos = 0, res_pos = 0, cnt_pos = 0;
		BitReader rh;
		if (active) {
			i = H.sorted[t];
			rh.init(g, ef_get(g, R.lo + i));
			RowHead h;
			const uint32_t rf = R.ref[i];
			row_head<DEF>(g, rh, rf ? R.deg[i - rf] : 0, h);
			nextra = h.d - h.ncopied;
			ncopied = h.ncopied;
			cnt_pos = rh.tell();
			row_intervals(g, rh, nextra, ic, int_pos, nres);
			ic0 = ic;
			res_pos = rh.tell();
			if (ic) rh.seek(int_pos);  // rh becomes the interval cursor
		}
		__syncwarp();
		// Giant rows (the reference's slow test has rows of 2^30 successors, BVGraphSlowTest.java:30-96): one lane would need
		// seconds, so the whole warp decodes them one after the other, every lane a segment of the list (coop_residuals).
		{
			bool giant = active && nres >= GIANT_ROW && ic <= 32;
			if (giant && nres >= GIANT_ROW2 && H.giants) {
				// the largest ones: many warps each, after this kernel
				const unsigned long long slot = atomicAdd(&H.counters[129], 1ull);
				if (slot < (unsigned long long)GIANT_MAX) {
					GiantRow gr;
					gr.x = R.lo + i;
					gr.p0 = res_pos;
					gr.p1 = ef_get(g, R.lo + i + 1);
					gr.cnt_pos = cnt_pos;
					gr.ext = row_slot(R, i, rp_from) + ncopied;
					gr.nres = nres;
					gr.ne = nextra;
					gr.ic = ic0;
					gr.pad = 0;
					H.giants[slot] = gr;
					giant = false;
					nextra = 0;
				}
			}
			unsigned gm = __ballot_sync(FULL_MASK, giant);
			while (gm) {
				const int src = __ffs((int)gm) - 1;
				gm &= gm - 1;
				const int64_t gi = __shfl_sync(FULL_MASK, i, src);
				const uint64_t g_res = __shfl_sync(FULL_MASK, res_pos, src);
				const uint32_t g_nres = __shfl_sync(FULL_MASK, nres, src), g_nextra = __shfl_sync(FULL_MASK, nextra, src);
				const uint32_t g_ncop = __shfl_sync(FULL_MASK, ncopied, src), g_ic = __shfl_sync(FULL_MASK, ic0, src);
				const int64_t gx = R.lo + gi;
				int64_t *ext = row_slot(R, gi, rp_from) + g_ncop;
				BitReader proto;
				proto.init(g, ef_get(g, gx));
				coop_residuals<DEF>(g, proto, g_res, ef_get(g, gx + 1), g_nres, gx, ext + (g_nextra - g_nres));
				if (g_ic) {
					__threadfence_block();
					__syncwarp();
					// (coop_intervals_merge starts at the interval count: one gamma code before the first interval)
					coop_intervals_merge(g, proto, __shfl_sync(FULL_MASK, cnt_pos, src), gx, ext, g_nextra, g_nres);
				}
				__syncwarp();
			}
			if (giant) nextra = 0;  // done: nothing left for the lock-step loops below
		}
		if (DEF == 0) {
			// run-time codings include Golomb / unary residuals, whose codes can be longer than any window: read them
			// straight from global memory
			if (active) {
				BitReader r = rh;
				r.seek(res_pos);
				NoWindow nw;
				row_extras_core<DEF>(g, rh, r, R.lo + i, nextra, ic, nres, row_slot(R, i, rp_from) + ncopied, mask, nw);
				if (nextra && r.past_end()) raise(g, ERR_OVERRUN);
			}
		} else {
			TileReader r;
			win.fill(r, res_pos, active);
			if (active) {
				row_extras_core<DEF>(g, rh, r, R.lo + i, nextra, ic, nres, row_slot(R, i, rp_from) + ncopied, mask, win);
				if (nextra && win.tell(r) > ((g.nwords * 2) << 5)) raise(g, ERR_OVERRUN);
			}
		}
		__syncwarp();
	}
}

// ---- giant rows: GIANT_CHUNKS warps per row (see GiantRow)
__device__ __forceinline__ void giant_chunk_range(const GiantRow &gr, int c, uint64_t &cs, uint64_t &ce) {
	const uint64_t span = gr.p1 - gr.p0;
	const uint64_t cb = ((span + GIANT_CHUNKS - 1) / GIANT_CHUNKS + 63) & ~63ull;
	cs = gr.p0 + cb * (uint64_t)c;
	ce = cs + cb;
	if (cs > gr.p1) cs = gr.p1;
	if (ce > gr.p1) ce = gr.p1;
}

template <int DEF>
__global__ void __launch_bounds__(TPB) k_giant_scan(GraphDev g, HeavyDev H) {
	const unsigned long long total = H.counters[129] < (unsigned long long)GIANT_MAX ? H.counters[129] : (unsigned long long)GIANT_MAX;
	const int c = (int)(blockIdx.x * (TPB / 32) + (threadIdx.x >> 5));
	for (unsigned long long gi = blockIdx.y; gi < total; gi += gridDim.y) {
		const GiantRow gr = H.giants[gi];
		uint64_t cs, ce;
		giant_chunk_range(gr, c, cs, ce);
		GiantChunk ch;
		ch.entry = ch.exit = cs;
		ch.total = ch.base = 0;
		ch.count = ch.idx0 = 0;
		if (cs < ce) {
			BitReader proto;
			proto.init(g, gr.p0);
			// a code start at or after cs: parse from GIANT_LEAD bits before it (every lane the same codes)
			uint64_t p = cs - gr.p0 > (uint64_t)GIANT_LEAD ? cs - GIANT_LEAD : gr.p0;
			if (c != 0) {
				BitReader r = proto;
				r.seek(p);
				while (p < cs) {
					rd_residual<DEF>(r, g);
					const uint64_t q = r.tell();
					if (q <= p) break;
					p = q;
				}
			} else p = gr.p0;
			ch.entry = p;
			uint32_t count;
			int64_t tot;
			uint64_t ex;
			coop_residuals_ex<DEF, false>(g, proto, p, ce, 0xffffffffu, gr.x, (int64_t *)nullptr, 0u, (int64_t)0, false, count, tot, ex);
			ch.count = count;
			ch.total = tot;
			ch.exit = ex > p ? ex : p;
		}
		if ((threadIdx.x & 31) == 0) H.gchunks[gi * GIANT_CHUNKS + c] = ch;
	}
}

template <int DEF>
__global__ void k_giant_fix(GraphDev g, HeavyDev H) {
	const unsigned long long total = H.counters[129] < (unsigned long long)GIANT_MAX ? H.counters[129] : (unsigned long long)GIANT_MAX;
	const unsigned lane = threadIdx.x & 31u;
	for (unsigned long long gi = blockIdx.x; gi < total; gi += gridDim.x) {
		const GiantRow gr = H.giants[gi];
		BitReader proto;
		proto.init(g, gr.p0);
		// value the first gap builds on: the first code is a signed offset from x (BVG:917)
		int64_t base;
		{
			BitReader r = proto;
			const uint64_t code = rd_residual<DEF>(r, g);
			base = gr.x + nat2int(code) - ((int64_t)code + 1);
		}
		uint64_t expect = gr.p0;
		uint32_t idx = 0;
		for (int c = 0; c < GIANT_CHUNKS; c++) {
			GiantChunk ch = H.gchunks[gi * GIANT_CHUNKS + c];
			uint64_t cs, ce;
			giant_chunk_range(gr, c, cs, ce);
			if (expect >= ce) {
				// (nothing starts in this chunk)
				ch.entry = ch.exit = expect;
				ch.count = 0;
				ch.total = 0;
			} else if (ch.entry != expect || cs >= ce) {
				// the lead-in had not fallen in step (or the chunk before ended late): scan again from the true start
				uint32_t count;
				int64_t tot;
				uint64_t ex;
				coop_residuals_ex<DEF, false>(g, proto, expect, ce, 0xffffffffu, gr.x, (int64_t *)nullptr, 0u, (int64_t)0, false, count, tot, ex);
				ch.entry = expect;
				ch.count = count;
				ch.total = tot;
				ch.exit = ex > expect ? ex : expect;
			}
			ch.idx0 = idx;
			ch.base = base;
			idx += ch.count;
			base += ch.total;
			expect = ch.exit;
			if (lane == 0) H.gchunks[gi * GIANT_CHUNKS + c] = ch;
		}
		if (lane == 0 && idx != gr.nres) raise(g, ERR_DEGREE);
	}
}

template <int DEF>
__global__ void __launch_bounds__(TPB) k_giant_write(GraphDev g, HeavyDev H) {
	const unsigned long long total = H.counters[129] < (unsigned long long)GIANT_MAX ? H.counters[129] : (unsigned long long)GIANT_MAX;
	const int c = (int)(blockIdx.x * (TPB / 32) + (threadIdx.x >> 5));
	for (unsigned long long gi = blockIdx.y; gi < total; gi += gridDim.y) {
		const GiantRow gr = H.giants[gi];
		const GiantChunk ch = H.gchunks[gi * GIANT_CHUNKS + c];
		if (ch.count == 0) continue;
		uint64_t cs, ce;
		giant_chunk_range(gr, c, cs, ce);
		BitReader proto;
		proto.init(g, gr.p0);
		uint32_t count;
		int64_t tot;
		uint64_t ex;
		coop_residuals_ex<DEF, true>(g, proto, ch.entry, ce, gr.nres, gr.x, gr.ext + (gr.ne - gr.nres), ch.idx0, ch.base, false, count, tot, ex);
	}
}

// the intervals of giant rows that have both (rare): one warp per row
__global__ void k_giant_merge(GraphDev g, HeavyDev H) {
	const unsigned long long total = H.counters[129] < (unsigned long long)GIANT_MAX ? H.counters[129] : (unsigned long long)GIANT_MAX;
	for (unsigned long long gi = blockIdx.x; gi < total; gi += gridDim.x) {
		const GiantRow gr = H.giants[gi];
		if (!gr.ic) continue;
		BitReader proto;
		proto.init(g, gr.cnt_pos);
		coop_intervals_merge(g, proto, gr.cnt_pos, gr.x, gr.ext, gr.ne, gr.nres);
	}
}

void launch_extras(const Launch &L, const GraphDev &g, int def, const RangeDev &r, const HeavyDev &h) {
	const int64_t count = r.hi - r.lo;
	if (count <= 0) return;
	if (h.rows) cudaMemsetAsync(h.counters, 0, 130 * sizeof(unsigned long long), L.stream);
	const size_t smem = sizeof(ExtrasSmem);
#define WGBV_LAUNCH_EXTRAS(D, F)                                                                              \
	do {                                                                                                     \
		cudaFuncSetAttribute(k_extras<D, F>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);        \
		k_extras<D, F><<<grid_for(count, ETILE), TPB, smem, L.stream>>>(g, r, h);                            \
	} while (0)
	const bool front = r.rp_w != nullptr;  // stages 1 and 2 inside the kernel
	if (front) {
		cudaMemsetAsync(r.tile_state, 0, (size_t)extras_tiles(count) * sizeof(uint64_t), L.stream);
		cudaMemsetAsync(r.tile_ticket, 0, sizeof(uint32_t), L.stream);
	}
	switch (def) {
	case 2:
		if (front) WGBV_LAUNCH_EXTRAS(2, true);
		else WGBV_LAUNCH_EXTRAS(2, false);
		break;
	case 1:
		if (front) WGBV_LAUNCH_EXTRAS(1, true);
		else WGBV_LAUNCH_EXTRAS(1, false);
		break;
	default:
		if (front) WGBV_LAUNCH_EXTRAS(0, true);
		else WGBV_LAUNCH_EXTRAS(0, false);
	}
#undef WGBV_LAUNCH_EXTRAS
	WGBV_LAUNCHED(L);
	if (!h.rows) return;
	k_heavy_offsets<<<1, 32, 0, L.stream>>>(h);
	WGBV_LAUNCHED(L);
	const unsigned grid = (unsigned)std::min<int64_t>(grid_for(count / 8 + 1, TILE), 148 * 8);
	k_heavy_scatter<<<grid, TPB, 0, L.stream>>>(h);
	WGBV_LAUNCHED(L);
	const size_t hsmem = (size_t)(TPB / 32) * 32 * HW_PITCH * sizeof(uint32_t);
	switch (def) {
	case 2:
		cudaFuncSetAttribute(k_extras_heavy<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hsmem);
		k_extras_heavy<2><<<148 * 8, TPB, hsmem, L.stream>>>(g, r, h);
		break;
	case 1:
		cudaFuncSetAttribute(k_extras_heavy<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hsmem);
		k_extras_heavy<1><<<148 * 8, TPB, hsmem, L.stream>>>(g, r, h);
		break;
	default:
		cudaFuncSetAttribute(k_extras_heavy<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hsmem);
		k_extras_heavy<0><<<148 * 8, TPB, hsmem, L.stream>>>(g, r, h);
	}
	WGBV_LAUNCHED(L);
	if (!h.giants) return;
	// giant rows, if any (the kernels read the count on the device and leave at once when there is none)
	const dim3 gg(GIANT_CHUNKS / (TPB / 32), 32);
	switch (def) {
	case 2: k_giant_scan<2><<<gg, TPB, 0, L.stream>>>(g, h); k_giant_fix<2><<<32, 32, 0, L.stream>>>(g, h); k_giant_write<2><<<gg, TPB, 0, L.stream>>>(g, h); break;
	case 1: k_giant_scan<1><<<gg, TPB, 0, L.stream>>>(g, h); k_giant_fix<1><<<32, 32, 0, L.stream>>>(g, h); k_giant_write<1><<<gg, TPB, 0, L.stream>>>(g, h); break;
	default: k_giant_scan<0><<<gg, TPB, 0, L.stream>>>(g, h); k_giant_fix<0><<<32, 32, 0, L.stream>>>(g, h); k_giant_write<0><<<gg, TPB, 0, L.stream>>>(g, h);
	}
	k_giant_merge<<<32, 32, 0, L.stream>>>(g, h);
	for (int q = 0; q < 4; q++) WGBV_LAUNCHED(L);
}

// stage 3b: the rows of one depth level.  With a list, a block takes CTILE list entries; without, it inspects CTILE
// consecutive nodes and keeps those at this level.  Long rows are merged by a whole warp each (row_copy_warp),
// short ones by one lane each, 32 rows of one class per warp (row_copy_masked).
#ifndef WGBV_CRPT
#define WGBV_CRPT 2
#endif
static constexpr int CRPT = WGBV_CRPT;  // 512 rows per block measured best (256: +0.45 ms, 768: +0.1 ms, 1024: +0.3 ms, 2048: +2.9 ms per 10^8 nodes)
static constexpr int CTILE = TPB * CRPT;

__device__ __forceinline__ void cp_async8(void *smem_dst, const void *gmem_src) {
	const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
	asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(gmem_src));
}

struct CopySmem {
	uint64_t a[CTILE], b[CTILE];  // the rows' RowInfo
	uint32_t row[CTILE];
	uint32_t hist[NBUCKETS + 1];
	uint16_t order[CTILE];
};

#ifndef WGBV_COPY_BLOCKS
#define WGBV_COPY_BLOCKS 5  // 48 registers, five blocks per SM: 27.9 ms per 1e8 nodes (four blocks 28.6, six 28.3, eight 28.7)
#endif
template <int DEF>
__global__ void __launch_bounds__(TPB, WGBV_COPY_BLOCKS) k_copy_level(GraphDev g, RangeDev R, uint32_t level, const uint32_t *__restrict__ list,
                                                    uint64_t list_count) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	CopySmem &S = *reinterpret_cast<CopySmem *>(smem_raw);
	const uint64_t base = (uint64_t)blockIdx.x * CTILE;
	const int64_t rp_from = R.rp[R.from - R.lo];
	int bucket[CRPT];
#pragma unroll 1
	for (int q = 0; q < CRPT; q++) {
		const int id = q * TPB + threadIdx.x;
		const uint64_t t = base + id;
		bucket[q] = -1;
		int64_t i = -1;
		if (list) {
			if (t < list_count) i = list[t];
		} else if ((int64_t)t < R.hi - R.lo && R.depth[t] == level) i = (int64_t)t;
		if (i >= 0 && R.deg[i] != 0 && !(R.lo + i >= R.from && (uint64_t)(R.rp[i + 1] - rp_from) > R.col_capacity)) {
			RowInfo ri;
			const uint32_t d = R.deg[i], dref = R.deg[i - R.ref[i]];
			const bool masked = row_is_masked(d, dref);
			if (R.info) ri = R.info[i];  // stage 3a left what it parsed
			else {
				BitReader r;
				r.init(g, ef_get(g, R.lo + i));
				RowHead h;
				row_head<DEF>(g, r, dref, h);
				ri.a = masked ? h.cmask[0] : h.blocks_pos;
				ri.b = masked ? h.cmask[1] : (uint64_t)h.ncopied | (uint64_t)h.bc << 32;
			}
			if (masked ? (ri.a | ri.b) != 0 : (uint32_t)ri.b != 0) {
				// the merge loop runs up to d times (copied + extras moved forward); rows without a mask go to the warps
				bucket[q] = 2 * work_bucket(masked || d >= WARP_ROW ? d : WARP_ROW);
				S.a[id] = ri.a;
				S.b[id] = ri.b;
				S.row[id] = (uint32_t)i;
			}
		}
		__syncwarp();
	}
	uint32_t total;
	block_bin_sort<CRPT>(bucket, S.hist, S.order, &total);
	const unsigned lane = threadIdx.x & 31u;
	// rows of at least WARP_ROW successors or without a mask (they come first in the sorted order): one warp per row
	const uint32_t nbig = S.hist[2 * work_bucket(WARP_ROW - 1)];
#pragma unroll 1
	for (uint32_t t = threadIdx.x >> 5; t < nbig; t += TPB / 32) {
		const int id = S.order[t];
		const int64_t k = S.row[id];
		const uint32_t rf = R.ref[k];
		const uint32_t ncopied = (uint32_t)S.b[id];
		int64_t *dst = row_slot(R, k, rp_from);
		BitReader r;
		r.init(g, S.a[id]);
		row_copy_warp<DEF>(g, r, row_slot(R, k - rf, rp_from), R.deg[k - rf], dst + ncopied, dst, R.deg[k], (uint32_t)(S.b[id] >> 32), ncopied);
	}
	__syncwarp();
	// the other rows: one lane per row, 32 rows of one class per warp
#pragma unroll 1
	for (uint32_t tb = nbig + (threadIdx.x & ~31u); tb < total; tb += TPB) {
		const uint32_t t = tb + lane;
		const bool active = t < total;
		const unsigned mask = __ballot_sync(FULL_MASK, active);
		if (active) {
			const int id = S.order[t];
			const int64_t k = S.row[id];
			const uint32_t rf = R.ref[k];
			const uint64_t m0 = S.a[id], m1 = S.b[id];
			const uint32_t ncopied = (uint32_t)(__popcll(m0) + __popcll(m1));
			int64_t *dst = row_slot(R, k, rp_from);
#ifdef WGBV_LEAN_COPY
			row_merge_lean<int64_t>(g, row_slot(R, k - rf, rp_from), m0, m1, dst + ncopied, dst, R.deg[k] - ncopied, mask);
#else
			row_copy_masked(g, row_slot(R, k - rf, rp_from), m0, m1, dst + ncopied, dst, R.deg[k] - ncopied, mask);
#endif
		}
	}
}

// A level of few rows (the deep end of long reference chains): one warp per row and no sorting -- the life of a block of
// k_copy_level (first pass, sort, two rounds of serial merges: 85 us) is what such a launch costs, whatever its size.
// Rows with a mask are merged by the warp's first lane, the others by the whole warp.
static constexpr uint64_t SMALL_LEVEL = 8192;  // (at 40 000 rows the sorted kernel is the faster one: 0.67 vs 0.72 ms for a 250 000-node range)

template <int DEF>
__global__ void __launch_bounds__(TPB) k_copy_small(GraphDev g, RangeDev R, const uint32_t *__restrict__ list, uint64_t list_count) {
	const uint64_t w = (uint64_t)blockIdx.x * (TPB / 32) + (threadIdx.x >> 5);
	if (w >= list_count) return;
	const unsigned lane = threadIdx.x & 31u;
	const int64_t rp_from = R.rp[R.from - R.lo];
	const int64_t i = list[w];
	const uint32_t d = R.deg[i];
	if (d == 0 || (R.lo + i >= R.from && (uint64_t)(R.rp[i + 1] - rp_from) > R.col_capacity)) return;
	const uint32_t rf = R.ref[i], dref = R.deg[i - rf];
	const RowInfo ri = R.info[i];
	int64_t *dst = row_slot(R, i, rp_from);
	const int64_t *src = row_slot(R, i - rf, rp_from);
	if (row_is_masked(d, dref)) {
		if (lane == 0 && (ri.a | ri.b) != 0) {
			const uint32_t ncopied = (uint32_t)(__popcll(ri.a) + __popcll(ri.b));
			row_copy_masked(g, src, ri.a, ri.b, dst + ncopied, dst, d - ncopied, 1u);
		}
	} else if ((uint32_t)ri.b != 0) {
		const uint32_t ncopied = (uint32_t)ri.b;
		BitReader r;
		r.init(g, ri.a);
		row_copy_warp<DEF>(g, r, src, dref, dst + ncopied, dst, d, (uint32_t)(ri.b >> 32), ncopied);
	}
}

void launch_copy_level(const Launch &L, const GraphDev &g, int def, const RangeDev &r, uint32_t level, const uint32_t *list,
                       uint64_t list_count) {
	const int64_t items = list ? (int64_t)list_count : r.hi - r.lo;
	if (items <= 0) return;
	if (list && r.info && list_count < SMALL_LEVEL) {
		const unsigned grid = grid_for((int64_t)list_count, TPB / 32);
		switch (def) {
		case 2: k_copy_small<2><<<grid, TPB, 0, L.stream>>>(g, r, list, list_count); break;
		case 1: k_copy_small<1><<<grid, TPB, 0, L.stream>>>(g, r, list, list_count); break;
		default: k_copy_small<0><<<grid, TPB, 0, L.stream>>>(g, r, list, list_count);
		}
		WGBV_LAUNCHED(L);
		return;
	}
	const size_t smem = sizeof(CopySmem);
	const unsigned grid = grid_for(items, CTILE);
	switch (def) {
	case 2:
		cudaFuncSetAttribute(k_copy_level<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		k_copy_level<2><<<grid, TPB, smem, L.stream>>>(g, r, level, list, list_count);
		break;
	case 1:
		cudaFuncSetAttribute(k_copy_level<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		k_copy_level<1><<<grid, TPB, smem, L.stream>>>(g, r, level, list, list_count);
		break;
	default:
		cudaFuncSetAttribute(k_copy_level<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		k_copy_level<0><<<grid, TPB, smem, L.stream>>>(g, r, level, list, list_count);
	}
	WGBV_LAUNCHED(L);
}

// ------------------------------------------------------------------------------------------------ checksum
// Same row hash as datagen/bvgen.c (row_hash) and, through it, what the generator computed before encoding.
__device__ __forceinline__ uint64_t row_hash(int64_t x, const int64_t *a, int64_t d) {
	uint64_t h = 0x9e3779b97f4a7c15ull ^ (uint64_t)d;
	for (int64_t i = 0; i < d; i++) h = (h ^ (uint64_t)a[i]) * 0x100000001b3ull + 0x632be59bd9b4e019ull;
	h ^= (uint64_t)x * 0xff51afd7ed558ccdull;
	h ^= h >> 32;
	h *= 0xd6e8feb86659fd93ull;
	h ^= h >> 32;
	return h;
}

__global__ void __launch_bounds__(TPB) k_checksum(int64_t first_node, int64_t rows, const int64_t *__restrict__ rowptr,
                                                  const int64_t *__restrict__ col, unsigned long long *out) {
	const int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
	uint64_t h = 0;
	if (i < rows) {
		const int64_t a = rowptr[i], b = rowptr[i + 1];
		h = row_hash(first_node + i, col + (a - rowptr[0]), b - a);
	}
#pragma unroll
	for (int o = 16; o; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
	if ((threadIdx.x & 31) == 0 && h) atomicAdd(out, (unsigned long long)h);
}

void launch_checksum(const Launch &L, int64_t first_node, int64_t rows, const int64_t *rowptr, const int64_t *col,
                     unsigned long long *out) {
	if (rows <= 0) return;
	k_checksum<<<grid_for(rows, TPB), TPB, 0, L.stream>>>(first_node, rows, rowptr, col, out);
	WGBV_LAUNCHED(L);
}

// ------------------------------------------------------------------------------------------------ random access
// outdegree(x) for arbitrary x (BVG:821-851).  Ids outside [lo, hi) raise the index error.
template <int DEF>
__global__ void __launch_bounds__(TPB) k_batch_degrees(GraphDev g, const int64_t *__restrict__ ids, uint64_t k, int64_t lo,
                                                       int64_t hi, uint32_t *__restrict__ deg) {
	const uint64_t q = (uint64_t)blockIdx.x * TPB + threadIdx.x;
	if (q >= k) return;
	const int64_t x = ids[q];
	if (x < lo || x >= hi) { raise(g, ERR_INDEX); deg[q] = 0; return; }
	BitReader r;
	r.init(g, ef_get(g, x));
	const uint64_t d = rd_outdegree<DEF>(r, g);
	if (d > 0x7fffffffull) { raise(g, ERR_DEGREE); deg[q] = 0; return; }
	deg[q] = (uint32_t)d;
}
void launch_batch_degrees(const Launch &L, const GraphDev &g, int def, const int64_t *ids, uint64_t k, int64_t lo, int64_t hi,
                          uint32_t *deg) {
	if (!k) return;
	switch (def) {
	case 2: k_batch_degrees<2><<<grid_for((int64_t)k, TPB), TPB, 0, L.stream>>>(g, ids, k, lo, hi, deg); break;
	case 1: k_batch_degrees<1><<<grid_for((int64_t)k, TPB), TPB, 0, L.stream>>>(g, ids, k, lo, hi, deg); break;
	default: k_batch_degrees<0><<<grid_for((int64_t)k, TPB), TPB, 0, L.stream>>>(g, ids, k, lo, hi, deg);
	}
	WGBV_LAUNCHED(L);
}

// The reference resolves successors(x) recursively through x - ref (BVG:1084).  Here every query's chain
// x, x - r1, x - r1 - r2, ... is laid out as explicit rows: the query row goes to the output, its ancestors to a
// scratch store; rows are then decoded with the same two steps as the bulk path, root level first.
template <int DEF>
__device__ __forceinline__ void chain_step(const GraphDev &g, int64_t y, uint32_t &d, uint32_t &rf) {
	BitReader r;
	r.init(g, ef_get(g, y));
	const uint64_t dd = rd_outdegree<DEF>(r, g);
	d = dd > 0x7fffffffull ? 0u : (uint32_t)dd;
	rf = 0;
	if (d && g.window > 0) {
		const uint64_t f = rd_reference<DEF>(r, g);
		if (f > (uint64_t)g.window || (int64_t)f > y) raise(g, ERR_REF_WINDOW);
		else rf = (uint32_t)f;
	}
}

template <int DEF>
__global__ void __launch_bounds__(TPB) k_chain_walk(GraphDev g, const int64_t *__restrict__ ids, uint64_t k, int64_t lo,
                                                    uint32_t *__restrict__ chain_len, unsigned long long *__restrict__ chain_arcs,
                                                    uint32_t *stats) {
	const uint64_t q = (uint64_t)blockIdx.x * TPB + threadIdx.x;
	if (q >= k) return;
	int64_t y = ids[q];
	uint32_t len = 0;
	unsigned long long arcs = 0;
	for (;;) {
		uint32_t d, rf;
		chain_step<DEF>(g, y, d, rf);
		if (len) arcs += d;
		len++;
		if (!rf) break;
		if (y - (int64_t)rf < lo) { raise(g, ERR_REF_UNRESOLVED); break; }
		y -= rf;
	}
	chain_len[q] = len;
	chain_arcs[q] = arcs;
	atomicMax(&stats[1], len);
}
void launch_chain_walk(const Launch &L, const GraphDev &g, int def, const int64_t *ids, uint64_t k, int64_t lo,
                       uint32_t *chain_len, unsigned long long *chain_arcs, uint32_t *stats) {
	if (!k) return;
	switch (def) {
	case 2: k_chain_walk<2><<<grid_for((int64_t)k, TPB), TPB, 0, L.stream>>>(g, ids, k, lo, chain_len, chain_arcs, stats); break;
	case 1: k_chain_walk<1><<<grid_for((int64_t)k, TPB), TPB, 0, L.stream>>>(g, ids, k, lo, chain_len, chain_arcs, stats); break;
	default: k_chain_walk<0><<<grid_for((int64_t)k, TPB), TPB, 0, L.stream>>>(g, ids, k, lo, chain_len, chain_arcs, stats);
	}
	WGBV_LAUNCHED(L);
}

template <int DEF>
__global__ void __launch_bounds__(TPB) k_chain_fill(GraphDev g, const int64_t *__restrict__ ids, uint64_t k,
                                                    const int64_t *__restrict__ chain_row0, const int64_t *__restrict__ store_off0,
                                                    const int64_t *__restrict__ rowptr, int64_t *node, int64_t *dst_off,
                                                    int32_t *parent, uint32_t *deg, uint8_t *is_query, uint32_t *level) {
	const uint64_t q = (uint64_t)blockIdx.x * TPB + threadIdx.x;
	if (q >= k) return;
	const int64_t row0 = chain_row0[q];
	const uint32_t len = (uint32_t)(chain_row0[q + 1] - row0);
	int64_t y = ids[q];
	int64_t soff = store_off0[q];
	for (uint32_t j = 0; j < len; j++) {
		uint32_t d, rf;
		chain_step<DEF>(g, y, d, rf);
		const int64_t row = row0 + j;
		node[row] = y;
		deg[row] = d;
		is_query[row] = j == 0;
		level[row] = len - 1 - j;
		parent[row] = (j + 1 < len) ? (int32_t)(row + 1) : -1;
		if (j == 0) dst_off[row] = rowptr[q];
		else { dst_off[row] = soff; soff += d; }
		y -= rf;
	}
}
void launch_chain_fill(const Launch &L, const GraphDev &g, int def, const int64_t *ids, uint64_t k, const int64_t *chain_row0,
                       const int64_t *store_off0, const int64_t *rowptr, int64_t *node, int64_t *dst_off, int32_t *parent,
                       uint32_t *deg, uint8_t *is_query, uint32_t *level) {
	if (!k) return;
	switch (def) {
	case 2: k_chain_fill<2><<<grid_for((int64_t)k, TPB), TPB, 0, L.stream>>>(g, ids, k, chain_row0, store_off0, rowptr, node, dst_off, parent, deg, is_query, level); break;
	case 1: k_chain_fill<1><<<grid_for((int64_t)k, TPB), TPB, 0, L.stream>>>(g, ids, k, chain_row0, store_off0, rowptr, node, dst_off, parent, deg, is_query, level); break;
	default: k_chain_fill<0><<<grid_for((int64_t)k, TPB), TPB, 0, L.stream>>>(g, ids, k, chain_row0, store_off0, rowptr, node, dst_off, parent, deg, is_query, level);
	}
	WGBV_LAUNCHED(L);
}

__device__ __forceinline__ int64_t *chain_slot(const ChainDev &c, uint64_t row) {
	return (c.is_query[row] ? c.col : c.store) + c.dst_off[row];
}

// The rows of a batch come in no order and differ wildly in length, so both chain kernels do what the bulk kernels do:
// a block parses the heads of CHT rows, sorts them by work class in shared memory and only then runs the serial part,
// 32 rows of one class per warp (unsorted, one row per thread, they ran at 9 and 2-4 active lanes per instruction).
static constexpr int CHRPT = 4;
static constexpr int CHT = TPB * CHRPT;

struct ChainSmem {
	uint64_t a[CHT], b[CHT];  // extras: reader position, (extras, copied); copy: the row's RowInfo
	uint32_t hist[NBUCKETS + 1];
	uint16_t order[CHT];
};

template <int DEF>
__global__ void __launch_bounds__(TPB) k_chain_extras(GraphDev g, ChainDev c, uint64_t rows) {
	__shared__ ChainSmem S;
	const uint64_t base = (uint64_t)blockIdx.x * CHT;
	int bucket[CHRPT];
#pragma unroll 1
	for (int q = 0; q < CHRPT; q++) {
		const int id = q * TPB + (int)threadIdx.x;
		const uint64_t row = base + id;
		bucket[q] = -1;
		if (row < rows && c.deg[row] != 0) {
			const int32_t p = c.parent[row];
			BitReader r;
			r.init(g, ef_get(g, c.node[row]));
			RowHead h;
			const uint32_t dref = p >= 0 ? c.deg[p] : 0;
			row_head<DEF>(g, r, dref, h);
			if (p >= 0) {
				// (left for the level that merges this row: it does not parse the head again)
				const bool masked = row_is_masked(h.d, dref);
				RowInfo ri;
				ri.a = masked ? h.cmask[0] : h.blocks_pos;
				ri.b = masked ? h.cmask[1] : (uint64_t)h.ncopied | (uint64_t)h.bc << 32;
				c.info[row] = ri;
			}
			const uint32_t nextra = h.d - h.ncopied;
			if (nextra) {
				bucket[q] = work_bucket(nextra);
				S.a[id] = r.tell();
				S.b[id] = (uint64_t)nextra | (uint64_t)h.ncopied << 32;
			}
		}
		__syncwarp();
	}
	uint32_t total;
	block_bin_sort<CHRPT>(bucket, S.hist, S.order, &total);
#pragma unroll 1
	for (uint32_t tb = (threadIdx.x & ~31u); tb < total; tb += TPB) {
		const uint32_t t = tb + (threadIdx.x & 31u);
		const bool active = t < total;
		const unsigned mask = __ballot_sync(FULL_MASK, active);
		if (active) {
			const int id = S.order[t];
			const uint64_t row = base + id;
			BitReader r;
			r.init(g, S.a[id]);
			row_extras<DEF>(g, r, c.node[row], (uint32_t)S.b[id], chain_slot(c, row) + (uint32_t)(S.b[id] >> 32), mask);
		}
	}
}
void launch_chain_extras(const Launch &L, const GraphDev &g, int def, const ChainDev &c, uint64_t rows) {
	if (!rows) return;
	switch (def) {
	case 2: k_chain_extras<2><<<grid_for((int64_t)rows, CHT), TPB, 0, L.stream>>>(g, c, rows); break;
	case 1: k_chain_extras<1><<<grid_for((int64_t)rows, CHT), TPB, 0, L.stream>>>(g, c, rows); break;
	default: k_chain_extras<0><<<grid_for((int64_t)rows, CHT), TPB, 0, L.stream>>>(g, c, rows);
	}
	WGBV_LAUNCHED(L);
}

// One level of the chains: the rows whose parent (the row holding their reference list) is complete.  As in
// k_copy_level: rows of fewer than WARP_ROW successors whose parent has at most 128 are merged one lane per row from the
// mask of copied positions, the others by a whole warp each from the block codes.
template <int DEF>
__global__ void __launch_bounds__(TPB) k_chain_copy(GraphDev g, ChainDev c, uint64_t rows, const uint32_t *__restrict__ level,
                                                    uint32_t this_level) {
	__shared__ ChainSmem S;
	const uint64_t base = (uint64_t)blockIdx.x * CHT;
	int bucket[CHRPT];
#pragma unroll 1
	for (int q = 0; q < CHRPT; q++) {
		const int id = q * TPB + (int)threadIdx.x;
		const uint64_t row = base + id;
		bucket[q] = -1;
		if (row < rows && level[row] == this_level && c.deg[row] != 0 && c.parent[row] >= 0) {
			const uint32_t d = c.deg[row], dref = c.deg[c.parent[row]];
			const bool masked = row_is_masked(d, dref);
			const RowInfo ri = c.info[row];  // k_chain_extras left what it parsed
			const uint64_t ra = ri.a, rb = ri.b;
			if (masked ? (ra | rb) != 0 : (uint32_t)rb != 0) {
				bucket[q] = 2 * work_bucket(masked || d >= WARP_ROW ? d : WARP_ROW);
				S.a[id] = ra;
				S.b[id] = rb;
			}
		}
		__syncwarp();
	}
	uint32_t total;
	block_bin_sort<CHRPT>(bucket, S.hist, S.order, &total);
	const unsigned lane = threadIdx.x & 31u;
	const uint32_t nbig = S.hist[2 * work_bucket(WARP_ROW - 1)];
#pragma unroll 1
	for (uint32_t t = threadIdx.x >> 5; t < nbig; t += TPB / 32) {
		const int id = S.order[t];
		const uint64_t row = base + id;
		const uint64_t p = (uint64_t)c.parent[row];
		const uint32_t ncopied = (uint32_t)S.b[id];
		int64_t *dst = chain_slot(c, row);
		BitReader r;
		r.init(g, S.a[id]);
		row_copy_warp<DEF>(g, r, chain_slot(c, p), c.deg[p], dst + ncopied, dst, c.deg[row], (uint32_t)(S.b[id] >> 32), ncopied);
	}
	__syncwarp();
#pragma unroll 1
	for (uint32_t tb = nbig + (threadIdx.x & ~31u); tb < total; tb += TPB) {
		const uint32_t t = tb + lane;
		const bool active = t < total;
		const unsigned mask = __ballot_sync(FULL_MASK, active);
		if (active) {
			const int id = S.order[t];
			const uint64_t row = base + id;
			const uint64_t m0 = S.a[id], m1 = S.b[id];
			const uint32_t ncopied = (uint32_t)(__popcll(m0) + __popcll(m1));
			int64_t *dst = chain_slot(c, row);
			row_copy_masked(g, chain_slot(c, (uint64_t)c.parent[row]), m0, m1, dst + ncopied, dst, c.deg[row] - ncopied, mask);
		}
	}
}
void launch_chain_copy(const Launch &L, const GraphDev &g, int def, const ChainDev &c, uint64_t rows, const uint32_t *level,
                       uint32_t this_level) {
	if (!rows) return;
	switch (def) {
	case 2: k_chain_copy<2><<<grid_for((int64_t)rows, CHT), TPB, 0, L.stream>>>(g, c, rows, level, this_level); break;
	case 1: k_chain_copy<1><<<grid_for((int64_t)rows, CHT), TPB, 0, L.stream>>>(g, c, rows, level, this_level); break;
	default: k_chain_copy<0><<<grid_for((int64_t)rows, CHT), TPB, 0, L.stream>>>(g, c, rows, level, this_level);
	}
	WGBV_LAUNCHED(L);
}

}  // namespace wgbv
