// bv_transpose.cu -- transpose of a CSR held in device memory.
//
// The most common consumer of a full sequential decode in the reference is Transform.transposeOffline
// (Transform.java:1098-1144): scan the graph, turn every arc (x, y) into (y, x), sort batches by (target, source),
// merge them.  With the CSR already in HBM the same result -- for every y the increasing list of the x with
// y in successors(x) -- is a stable sort of the arcs by target: the arcs of a CSR come in increasing source order, so
// a STABLE sort by target alone leaves every target's sources increasing.
//
//   k_t_pack      arc j -> one 64-bit word  target << 32 | source   (node ids below 2^32), counts in-degrees and the
//                 digits of the first pass (its blocks cover the same tiles as the radix kernels)
//   k_rs_hist     per pass: digit histogram of every tile of RS_TILE words
//   (k_scan)      digit-major scan of the tile histograms = first output position of every (digit, tile)
//   k_rs_scatter  per pass: stable scatter of the tile's words; the last pass writes the sources straight to t_col
//
// LSD radix sort, RS_BITS bits per pass over the target half only: ceil(log2(n) / RS_BITS) passes (3 for 10^8 nodes).
// Ranks are computed without atomics on the data path: a warp owns 512 consecutive words of the tile, finds the lanes
// that hold the same digit with one ballot per digit bit (rank in the warp = number of lower lanes with that digit) and
// keeps one running count per (warp, digit) in shared memory, which is what makes the pass stable.
#include "bv_block.cuh"
#include "bv_kernels.cuh"

namespace wgbv {

static constexpr unsigned FULL_MASK_T = 0xffffffffu;
static constexpr int RS_BITS = 9;
static constexpr int RS_DIGITS = 1 << RS_BITS;
static constexpr int RS_ITEMS = 16;               // words per thread
static constexpr int RS_TILE = TPB * RS_ITEMS;    // words per block
static constexpr int RS_WARP_ITEMS = 32 * RS_ITEMS;
static_assert(RS_DIGITS == 2 * TPB, "the scatter kernel scans two digits per thread");

uint64_t transpose_tiles(uint64_t arcs) { return arcs ? (arcs + RS_TILE - 1) / RS_TILE : 0; }
int transpose_digits() { return RS_DIGITS; }
int transpose_passes(uint64_t nodes) {
	int bits = 1;
	while (bits < 32 && (1ull << bits) < nodes) bits++;
	return (bits + RS_BITS - 1) / RS_BITS;
}

// One block per PACK_TILE consecutive arcs; the rows that hold them are found by binary search in the row pointers
// (first the block's range of rows, then every arc inside that range: a handful of steps that hit L1).
static constexpr int PACK_ITEMS = RS_ITEMS;
static constexpr int PACK_TILE = TPB * PACK_ITEMS;  // = RS_TILE: the block also counts the digits of the first pass for its tile

__device__ __forceinline__ int64_t row_of(const int64_t *__restrict__ rp, int64_t lo, int64_t hi, int64_t a) {
	// the last row r in [lo, hi] with rp[r] <= a
	while (lo < hi) {
		const int64_t mid = lo + (hi - lo + 1) / 2;
		if (rp[mid] <= a) lo = mid;
		else hi = mid - 1;
	}
	return lo;
}

__global__ void __launch_bounds__(TPB) k_t_pack(GraphDev g, int64_t n, const int64_t *__restrict__ rp, const int64_t *__restrict__ col,
                                                uint64_t arcs, uint64_t *__restrict__ pairs, uint32_t *indeg, uint32_t nblocks,
                                                uint32_t *__restrict__ counts) {
	__shared__ int64_t s_row[2];
	__shared__ uint32_t hist[RS_DIGITS];  // digits of radix pass 0 (bits 32 .. 32 + RS_BITS of the packed word)
	// rowmap[p]: the row of arc j0 + p, relative to the tile's first row.  Every row of the tile marks the position of
	// its first arc (atomicMax: of several rows starting at one position -- empty ones and the row that follows -- the
	// last owns the arc), an inclusive max-scan spreads the marks to the right.
	__shared__ uint32_t rowmap[PACK_TILE];
	__shared__ uint32_t wmax[TPB / 32];
	for (int d = threadIdx.x; d < RS_DIGITS; d += TPB) hist[d] = 0;
	for (int p = threadIdx.x; p < PACK_TILE; p += TPB) rowmap[p] = 0;
	const uint64_t j0 = (uint64_t)blockIdx.x * PACK_TILE;
	const uint64_t j1 = j0 + PACK_TILE < arcs ? j0 + PACK_TILE : arcs;
	const int64_t base = rp[0];
	if (threadIdx.x == 0) s_row[0] = row_of(rp, 0, n - 1, base + (int64_t)j0);
	if (threadIdx.x == 32) s_row[1] = row_of(rp, 0, n - 1, base + (int64_t)j1 - 1);
	__syncthreads();
	const int64_t r0 = s_row[0], r1 = s_row[1];
	for (int64_t r = r0 + 1 + threadIdx.x; r <= r1; r += TPB) {
		const int64_t p = rp[r] - base - (int64_t)j0;  // in (0, PACK_TILE): rows after r0 start inside the tile
		atomicMax(&rowmap[p], (uint32_t)(r - r0));
	}
	__syncthreads();
	{
		// every warp scans its 512 positions, 32 at a time; the warps' maxima are combined when the map is read
		const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
		uint32_t carry = 0;
#pragma unroll 4
		for (int q = 0; q < PACK_ITEMS; q++) {
			const int p = (int)warp * (32 * PACK_ITEMS) + q * 32 + (int)lane;
			uint32_t v = rowmap[p];
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) {
				const uint32_t t = __shfl_up_sync(FULL_MASK_T, v, o);
				if ((int)lane >= o && t > v) v = t;
			}
			if (carry > v) v = carry;
			rowmap[p] = v;
			carry = __shfl_sync(FULL_MASK_T, v, 31);
		}
		if (lane == 0) wmax[warp] = carry;
	}
	__syncthreads();
#pragma unroll
	for (int q = 0; q < PACK_ITEMS; q++) {
		const uint32_t p = (uint32_t)q * TPB + threadIdx.x;
		const uint64_t j = j0 + p;
		if (j < j1) {
			const int64_t y = col[j];
			uint32_t rel = rowmap[p];
			for (uint32_t w = 0; w < p / (32 * PACK_ITEMS); w++) rel = wmax[w] > rel ? wmax[w] : rel;
			const int64_t x = r0 + rel;
			if (y < 0 || y >= n) {
				raise(g, ERR_INDEX);
				pairs[j] = (uint64_t)x;  // keeps the sort well defined; the call fails anyway
				atomicAdd(&hist[0], 1u);
			} else {
				pairs[j] = (uint64_t)y << 32 | (uint64_t)x;
				atomicAdd(&indeg[y], 1u);
				atomicAdd(&hist[(uint32_t)y & (RS_DIGITS - 1)], 1u);
			}
		}
	}
	__syncthreads();
	for (int d = threadIdx.x; d < RS_DIGITS; d += TPB) counts[(size_t)d * nblocks + blockIdx.x] = hist[d];
}

__global__ void __launch_bounds__(TPB) k_rs_hist(const uint64_t *__restrict__ in, uint64_t count, int shift, uint32_t nblocks,
                                                 uint32_t *__restrict__ counts) {
	__shared__ uint32_t hist[RS_DIGITS];
	for (int d = threadIdx.x; d < RS_DIGITS; d += TPB) hist[d] = 0;
	__syncthreads();
	const uint64_t t0 = (uint64_t)blockIdx.x * RS_TILE;
#pragma unroll
	for (int q = 0; q < RS_ITEMS; q++) {
		const uint64_t j = t0 + (uint64_t)q * TPB + threadIdx.x;
		if (j < count) atomicAdd(&hist[(uint32_t)(in[j] >> shift) & (RS_DIGITS - 1)], 1u);
	}
	__syncthreads();
	for (int d = threadIdx.x; d < RS_DIGITS; d += TPB) counts[(size_t)d * nblocks + blockIdx.x] = hist[d];
}

// FINAL: the last pass; the sorted words are not needed any more, only their source halves, as int64, in t_col.
// The tile is first put in digit order in shared memory and then written out by consecutive threads, so that the words
// of one digit -- a run of consecutive output positions -- leave in whole sectors instead of 8 bytes at a time.
struct ScatterSmem {
	uint64_t sorted[RS_TILE];                // the tile's words in digit order
	uint32_t wbase[TPB / 32][RS_DIGITS];     // per warp: count of the digit, then the running base inside the tile's digit run
	int64_t gbase[RS_DIGITS];                // first output position of the tile's words with that digit, minus dstart
	uint32_t dstart[RS_DIGITS];              // first slot of the digit in sorted[]
	uint32_t wsum[TPB / 32];
};

template <bool FINAL>
__global__ void __launch_bounds__(TPB, 4) k_rs_scatter(const uint64_t *__restrict__ in, uint64_t count, int shift, uint32_t nblocks,
                                                       const int64_t *__restrict__ pos, uint64_t *__restrict__ out, int64_t *__restrict__ t_col) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	ScatterSmem &S = *reinterpret_cast<ScatterSmem *>(smem_raw);
	const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	const unsigned lt = (1u << lane) - 1u;
	for (int i = threadIdx.x; i < (TPB / 32) * RS_DIGITS; i += TPB) (&S.wbase[0][0])[i] = 0;
	__syncthreads();
	const uint64_t t0 = (uint64_t)blockIdx.x * RS_TILE;
	const uint64_t w0 = t0 + (uint64_t)warp * RS_WARP_ITEMS;
	uint64_t key[RS_ITEMS];
	uint32_t offp[RS_ITEMS / 2];  // per word: its index among the warp's words with the same digit (10 bits), two per register
	// counts per (warp, digit).  The lanes holding the same digit are found with one ballot per digit bit (match.any
	// costs the address-divergence unit a step per distinct value: with it this kernel ran at 95 % of that pipe and
	// nothing else); rank in the warp = number of lower lanes with the digit, and the count so far of the (warp, digit)
	// pair -- read before the first lane of the group adds the group -- makes the index stable across the rounds.
#pragma unroll
	for (int r = 0; r < RS_ITEMS; r++) {
		const uint64_t j = w0 + (uint64_t)r * 32 + lane;
		key[r] = j < count ? in[j] : 0;
	}
#pragma unroll
	for (int r = 0; r < RS_ITEMS; r++) {
		const uint64_t j = w0 + (uint64_t)r * 32 + lane;
		const bool valid = j < count;
		const uint32_t d = (uint32_t)(key[r] >> shift) & (RS_DIGITS - 1);
		unsigned peers = __ballot_sync(FULL_MASK_T, valid);
#pragma unroll
		for (int b = 0; b < RS_BITS; b++) {
			const bool bit = (d >> b) & 1u;
			const unsigned bal = __ballot_sync(FULL_MASK_T, bit);
			peers &= bit ? bal : ~bal;
		}
		const uint32_t rank = (uint32_t)__popc(peers & lt);
		const uint32_t b0 = valid ? S.wbase[warp][d] : 0u;
		__syncwarp();
		if (valid && rank == 0) S.wbase[warp][d] = b0 + (uint32_t)__popc(peers);
		__syncwarp();
		if (r & 1) offp[r >> 1] |= (b0 + rank) << 16;
		else offp[r >> 1] = b0 + rank;
	}
	__syncthreads();
	// per digit: bases of the warps inside the digit's run; then the runs' first slots (exclusive scan over the digits,
	// two digits per thread)
	uint32_t tot[2];
#pragma unroll
	for (int k = 0; k < 2; k++) {
		const int d = 2 * threadIdx.x + k;
		uint32_t run = 0;
#pragma unroll
		for (int w = 0; w < TPB / 32; w++) {
			const uint32_t c = S.wbase[w][d];
			S.wbase[w][d] = run;
			run += c;
		}
		tot[k] = run;
	}
	uint32_t inc = tot[0] + tot[1];
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const uint32_t t = __shfl_up_sync(FULL_MASK_T, inc, o);
		if ((int)lane >= o) inc += t;
	}
	if (lane == 31) S.wsum[warp] = inc;
	__syncthreads();
	uint32_t before = 0;
	for (unsigned w = 0; w < warp; w++) before += S.wsum[w];
	const uint32_t ex = before + inc - (tot[0] + tot[1]);
	S.dstart[2 * threadIdx.x] = ex;
	S.dstart[2 * threadIdx.x + 1] = ex + tot[0];
	S.gbase[2 * threadIdx.x] = pos[(size_t)(2 * threadIdx.x) * nblocks + blockIdx.x] - (int64_t)ex;
	S.gbase[2 * threadIdx.x + 1] = pos[(size_t)(2 * threadIdx.x + 1) * nblocks + blockIdx.x] - (int64_t)(ex + tot[0]);
	__syncthreads();
#pragma unroll
	for (int r = 0; r < RS_ITEMS; r++) {
		const uint64_t j = w0 + (uint64_t)r * 32 + lane;
		const uint32_t d = (uint32_t)(key[r] >> shift) & (RS_DIGITS - 1);
		const uint32_t off = (offp[r >> 1] >> ((r & 1) * 16)) & 0xffffu;
		if (j < count) S.sorted[S.dstart[d] + S.wbase[warp][d] + off] = key[r];
	}
	__syncthreads();
	const uint32_t here = count - t0 < (uint64_t)RS_TILE ? (uint32_t)(count - t0) : (uint32_t)RS_TILE;
	for (uint32_t i = threadIdx.x; i < here; i += TPB) {
		const uint64_t k = S.sorted[i];
		const int64_t p = S.gbase[(uint32_t)(k >> shift) & (RS_DIGITS - 1)] + i;
		if (FINAL) t_col[p] = (int64_t)(k & 0xffffffffull);
		else out[p] = k;
	}
}

void launch_transpose_pack(const Launch &L, const GraphDev &g, int64_t n, const int64_t *rowptr, const int64_t *col, uint64_t arcs,
                           uint64_t *pairs, uint32_t *indeg, uint32_t *counts) {
	if (!arcs) return;
	const uint32_t nblocks = (uint32_t)transpose_tiles(arcs);
	k_t_pack<<<nblocks, TPB, 0, L.stream>>>(g, n, rowptr, col, arcs, pairs, indeg, nblocks, counts);
	WGBV_LAUNCHED(L);
}

void launch_radix_hist(const Launch &L, const uint64_t *in, uint64_t count, int pass, uint32_t *counts) {
	const uint32_t nblocks = (uint32_t)transpose_tiles(count);
	if (!nblocks) return;
	k_rs_hist<<<nblocks, TPB, 0, L.stream>>>(in, count, 32 + pass * RS_BITS, nblocks, counts);
	WGBV_LAUNCHED(L);
}

void launch_radix_scatter(const Launch &L, const uint64_t *in, uint64_t count, int pass, const int64_t *pos, uint64_t *out, int64_t *t_col) {
	const uint32_t nblocks = (uint32_t)transpose_tiles(count);
	if (!nblocks) return;
	const size_t smem = sizeof(ScatterSmem);
	if (t_col) {
		cudaFuncSetAttribute(k_rs_scatter<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		k_rs_scatter<true><<<nblocks, TPB, smem, L.stream>>>(in, count, 32 + pass * RS_BITS, nblocks, pos, nullptr, t_col);
	} else {
		cudaFuncSetAttribute(k_rs_scatter<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		k_rs_scatter<false><<<nblocks, TPB, smem, L.stream>>>(in, count, 32 + pass * RS_BITS, nblocks, pos, out, nullptr);
	}
	WGBV_LAUNCHED(L);
}

// ------------------------------------------------------------------------------------------------ union / symmetrize
// Transform.union(g0, g1) for two CSRs over the same node set (UnionImmutableGraph: per node the sorted union of the two
// successor lists, duplicates dropped), which with g1 = transpose(g0) is Transform.symmetrizeOffline
// (Transform.java:573-574).  One thread per row: count, scan, fill.
template <bool FILL>
__global__ void __launch_bounds__(TPB) k_union(int64_t n, const int64_t *__restrict__ arp, const int64_t *__restrict__ acol,
                                               const int64_t *__restrict__ brp, const int64_t *__restrict__ bcol, uint32_t *__restrict__ cnt,
                                               const int64_t *__restrict__ urp, int64_t *__restrict__ ucol, uint64_t capacity, uint32_t *err) {
	const int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
	if (i >= n) return;
	const int64_t *a = acol + (arp[i] - arp[0]), *b = bcol + (brp[i] - brp[0]);
	const int64_t na = arp[i + 1] - arp[i], nb = brp[i + 1] - brp[i];
	int64_t *out = nullptr;
	if (FILL) {
		if ((uint64_t)urp[i + 1] > capacity) { atomicOr(err, ERR_CAPACITY); return; }
		out = ucol + urp[i];
	}
	int64_t j = 0, k = 0, o = 0;
	while (j < na && k < nb) {
		const int64_t x = a[j], y = b[k];
		const int64_t v = x < y ? x : y;
		if (FILL) out[o] = v;
		o++;
		j += x <= y;
		k += y <= x;
	}
	if (FILL) {
		for (; j < na; j++) out[o++] = a[j];
		for (; k < nb; k++) out[o++] = b[k];
	} else {
		o += (na - j) + (nb - k);
		if (o > 0x7fffffffll) { atomicOr(err, ERR_DEGREE); o = 0; }
		cnt[i] = (uint32_t)o;
	}
}

void launch_union_count(const Launch &L, int64_t n, const int64_t *arp, const int64_t *acol, const int64_t *brp, const int64_t *bcol, uint32_t *cnt,
                        uint32_t *err) {
	if (n <= 0) return;
	k_union<false><<<grid_for(n, TPB), TPB, 0, L.stream>>>(n, arp, acol, brp, bcol, cnt, nullptr, nullptr, 0, err);
	WGBV_LAUNCHED(L);
}
void launch_union_fill(const Launch &L, int64_t n, const int64_t *arp, const int64_t *acol, const int64_t *brp, const int64_t *bcol,
                       const int64_t *urp, int64_t *ucol, uint64_t capacity, uint32_t *err) {
	if (n <= 0) return;
	k_union<true><<<grid_for(n, TPB), TPB, 0, L.stream>>>(n, arp, acol, brp, bcol, nullptr, urp, ucol, capacity, err);
	WGBV_LAUNCHED(L);
}

}  // namespace wgbv
