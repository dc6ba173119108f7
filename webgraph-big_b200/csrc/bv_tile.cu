// bv_tile.cu -- stage 3 as ONE kernel for graphs with bounded reference chains (window_size * max_ref_count small):
// every block decodes a tile of TT consecutive rows completely on chip.
//
//   global memory            shared memory (per block)                                   global memory
//   compressed bits   --->   staged words of the tile  -> heads -> extras -> copy levels  --->  CSR rows
//   (coalesced, once)        rows of the tile (+ the few halo rows its references need)        (coalesced, once)
//
// A row may reference one of the window_size rows before it, and so on for at most max_ref_count hops
// (BVGraph.java:701 and :2258 of the reference), so a tile depends on at most HALO = window * maxref rows before
// its first one.  The block decodes the halo rows its tile actually needs redundantly (the CPU reference does the
// same when an iterator starts in the middle of the graph: window pre-fill, BVG:1139-1143) and therefore never
// waits for another block.  Traffic is the algorithmic minimum: the bit stream is read once (plus the halo), every
// successor is written once, and all the data-dependent accesses (bit reader, reference rows, merges) hit shared
// memory instead of L2/DRAM.
//
// A tile whose rows or bits do not fit the shared-memory budget keeps the same code path with generic pointers
// into global memory (its own rows at their final place, halo rows in a per-block scratch area); if even that
// scratch is too small the kernel reports it and the host falls back to the level-synchronous kernels.
#include <algorithm>

#include "bv_block.cuh"
#include "bv_rows.cuh"

namespace wgbv {

static constexpr int TT = 256;              // rows per tile
static constexpr int TH = TILE_MAX_HALO;    // halo rows a tile may look back at
static constexpr int TN = TT + TH;          // rows a block handles at most
static constexpr int ROWS_CAP = 5632;       // successors (int64) a tile may hold in shared memory
static constexpr int BITS_CAP = 3584;       // 32-bit words of the bit stream a tile may stage

struct TileSmem {
	int64_t rows[ROWS_CAP];
	uint32_t bits[BITS_CAP];
	uint32_t off[TN + 1];     // bit offset of each row's record, relative to the tile's first staged word
	uint32_t deg[TN];
	uint32_t rowoff[TN + 1];  // first slot of each row inside the tile's row storage
	uint32_t nextra[TN], ncopied[TN], ipos[TN], bpos[TN], bc[TN];
	uint32_t hist[NBUCKETS + 1];
	uint32_t warp_sum[TPB / 32];
	uint16_t order[TN];
	uint8_t ref[TN], depth[TN], need[TN];
	uint32_t max_depth, halo_arcs, fits_rows, fits_bits, bad;
	uint64_t bit_first, bit_last;  // span of the bit stream the tile needs (relative to g.bit_base)
};

// exclusive scan of one value per row over up to TN rows (two per thread); returns the total
__device__ __forceinline__ uint32_t tile_scan(TileSmem &S, const uint32_t *val, uint32_t *out, int nn) {
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int j0 = 2 * threadIdx.x, j1 = j0 + 1;
	const uint32_t a = j0 < nn ? val[j0] : 0u, b = j1 < nn ? val[j1] : 0u;
	uint32_t inc = a + b;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const uint32_t t = __shfl_up_sync(FULL_MASK, inc, o);
		if (lane >= o) inc += t;
	}
	if (lane == 31) S.warp_sum[warp] = inc;
	__syncthreads();
	uint32_t base = 0, total = 0;
#pragma unroll
	for (int w = 0; w < TPB / 32; w++) {
		const uint32_t v = S.warp_sum[w];
		if (w < warp) base += v;
		total += v;
	}
	const uint32_t ex = base + inc - a - b;
	if (j0 < nn) out[j0] = ex;
	if (j1 < nn) out[j1] = ex + a;
	if (threadIdx.x == 0) out[nn] = total;
	__syncthreads();
	return total;
}

template <int DEF>
__global__ void __launch_bounds__(TPB) k_tile(GraphDev g, RangeDev R, TileDev T) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	TileSmem &S = *reinterpret_cast<TileSmem *>(smem_raw);
	const int64_t rp_from = R.rp[R.from - R.lo];
	int64_t *const scratch = T.scratch + (size_t)blockIdx.x * T.scratch_cap;
	const int64_t ntiles = (R.hi - R.from + TT - 1) / TT;
	const int tid = threadIdx.x;

	for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
		const int64_t t0 = R.from + tile * TT;                           // first row of the tile (node id)
		const int64_t t1 = t0 + TT < R.hi ? t0 + TT : R.hi;
		const int64_t h0 = t0 - T.halo > R.lo ? t0 - T.halo : R.lo;      // first halo row
		const int nh = (int)(t0 - h0), nn = (int)(t1 - h0);
		const int64_t i0 = h0 - R.lo;                                     // index of row 0 of the block in R's arrays

		// ---- degrees, references, record offsets
		if (tid == 0) { S.max_depth = 0; S.bad = 0; }
		for (int j = tid; j < nn; j += TPB) {
			S.deg[j] = R.deg[i0 + j];
			S.ref[j] = R.ref[i0 + j];
			S.need[j] = j >= nh;
		}
		__syncthreads();
		// ---- reference chains: depth of every row, and which halo rows the tile needs
		for (int j = tid; j < nn; j += TPB) {
			int y = j, dep = 0;
			bool ok = true;
			while (S.ref[y]) {
				y -= S.ref[y];
				if (y < 0) { ok = false; break; }
				dep++;
				if (j >= nh && y < nh) S.need[y] = 1;
			}
			S.depth[j] = ok ? (uint8_t)(dep > 254 ? 254 : dep) : 255;
			if (j >= nh) {
				if (!ok) { raise(g, ERR_REF_UNRESOLVED); S.deg[j] = 0; }
				else atomicMax(&S.max_depth, (uint32_t)dep);
			}
		}
		__syncthreads();
		// a needed halo row always resolves inside the window (it lies on the chain of a tile row that does)
		for (int j = tid; j < nn; j += TPB) {
			const bool need = S.need[j] && S.depth[j] != 255;
			S.nextra[j] = need ? S.deg[j] : 0u;  // borrowed as the scan input
		}
		__syncthreads();
		const uint32_t total = tile_scan(S, S.nextra, S.rowoff, nn);
		const uint32_t halo_arcs = S.rowoff[nh];
		const bool fits_rows = total <= (uint32_t)ROWS_CAP;
		if (!fits_rows && halo_arcs > T.scratch_cap) {
			// not even the halo fits the scratch area: leave the tile to the level-synchronous path
			if (tid == 0) atomicExch(T.fallback, 1u);
			__syncthreads();
			continue;
		}
		// record offsets (Elias-Fano) of the rows that matter, and the span of the bit stream to stage
		int first = 0;
		while (first < nh && !(S.need[first] && S.depth[first] != 255)) first++;  // uniform: same smem for all threads
		if (tid == 0) S.bit_first = ef_get(g, h0 + first);
		if (tid == 32) S.bit_last = ef_get(g, t1);
		__syncthreads();
		const uint64_t bit_first = S.bit_first, bit_last = S.bit_last;
		const uint64_t w_first = bit_first >> 5;
		const uint64_t nwords = ((bit_last + 31) >> 5) - w_first + 2;
		const bool fits_bits = nwords <= (uint64_t)BITS_CAP && (bit_last - (w_first << 5)) < 0xffffff00ull;
		if (!fits_bits && (bit_last - (w_first << 5)) >= 0xffffff00ull) {
			if (tid == 0) atomicExch(T.fallback, 1u);  // offsets do not fit 32 bits: a multi-gigabit tile
			__syncthreads();
			continue;
		}
		for (int j = tid; j <= nn; j += TPB) {
			const bool need = j == nn || (j >= first && S.need[j] && S.depth[j] != 255);
			S.off[j] = need ? (uint32_t)(ef_get(g, h0 + j) - (w_first << 5)) : 0u;
		}
		if (fits_bits) {
			const uint32_t *src = (const uint32_t *)g.words + w_first;
			const uint64_t avail = g.nwords * 2 > w_first ? g.nwords * 2 - w_first : 0;
			for (uint32_t w = tid; w < (uint32_t)nwords; w += TPB) S.bits[w] = w < avail ? __ldg(src + w) : 0u;
		}
		__syncthreads();
		const uint32_t *const words = fits_bits ? S.bits : (const uint32_t *)g.words + w_first;
		const uint64_t words_avail = g.nwords * 2 > w_first ? g.nwords * 2 - w_first : 0;
		const uint32_t nw = fits_bits ? (uint32_t)nwords : (uint32_t)(words_avail > 0xffffffffull ? 0xffffffffull : words_avail);
		// where row j lives
		auto row_ptr = [&](int j) -> int64_t * {
			if (fits_rows) return S.rows + S.rowoff[j];
			return j >= nh ? R.col + (R.rp[i0 + j] - rp_from) : scratch + S.rowoff[j];
		};

		// ---- heads
		int bucket[2];
#pragma unroll
		for (int q = 0; q < 2; q++) {
			const int j = q * TPB + tid;
			bucket[q] = -1;
			if (j < nn) { S.nextra[j] = 0; S.ncopied[j] = 0; S.bc[j] = 0; }
			if (j < nn && S.need[j] && S.depth[j] != 255 && S.deg[j] != 0) {
				TileReader r;
				r.init_words(words, nw, S.off[j]);
				RowHead h;
				const uint32_t rf = S.ref[j];
				row_head<DEF>(g, r, rf ? S.deg[j - rf] : 0u, h);
				if (h.d != S.deg[j] || h.rf != rf) { h.d = 0; h.ncopied = 0; h.bc = 0; S.bad = 1; }
				const uint32_t nextra = h.d - h.ncopied;
				S.nextra[j] = nextra;
				S.ncopied[j] = h.ncopied;
				S.bc[j] = h.bc;
				S.bpos[j] = (uint32_t)h.blocks_pos;
				S.ipos[j] = (uint32_t)r.tell();
				if (nextra) {
					bool has_int = false;
					if (g.min_interval != 0) { TileReader t = r; has_int = t.gamma() != 0; }
					bucket[q] = 2 * work_bucket(nextra) + (has_int ? 0 : 1);
				}
			}
			__syncwarp();
		}
		uint32_t nwork;
		block_bin_sort<2>(bucket, S.hist, S.order, &nwork);
		// ---- extras: intervals and residuals of every row into the tail of its slot
#pragma unroll 1
		for (uint32_t tb = tid & ~31u; tb < nwork; tb += TPB) {
			const uint32_t t = tb + (tid & 31u);
			const bool active = t < nwork;
			const unsigned mask = __ballot_sync(FULL_MASK, active);
			if (active) {
				const int j = S.order[t];
				TileReader r;
				r.init_words(words, nw, S.ipos[j]);
				row_extras<DEF>(g, r, h0 + j, S.nextra[j], row_ptr(j) + S.ncopied[j], mask);
			}
		}
		if (!fits_rows) __threadfence_block();
		__syncthreads();
		// ---- copy blocks, one depth level after the other
		const uint32_t max_depth = S.max_depth;
		for (uint32_t level = 1; level <= max_depth; level++) {
#pragma unroll
			for (int q = 0; q < 2; q++) {
				const int j = q * TPB + tid;
				bucket[q] = (j < nn && S.depth[j] == level && S.need[j] && S.ncopied[j]) ? 2 * work_bucket(S.deg[j]) : -1;
			}
			uint32_t ncopy;
			block_bin_sort<2>(bucket, S.hist, S.order, &ncopy);
#pragma unroll 1
			for (uint32_t tb = tid & ~31u; tb < ncopy; tb += TPB) {
				const uint32_t t = tb + (tid & 31u);
				const bool active = t < ncopy;
				const unsigned mask = __ballot_sync(FULL_MASK, active);
				if (active) {
					const int j = S.order[t];
					const int y = j - S.ref[j];
					TileReader r;
					r.init_words(words, nw, S.bpos[j]);
					row_copy<DEF>(g, r, row_ptr(y), S.deg[y], row_ptr(j) + S.ncopied[j], row_ptr(j), S.deg[j], S.bc[j], S.ncopied[j], mask);
				}
			}
			if (!fits_rows) __threadfence_block();
			__syncthreads();
		}
		// ---- write the tile's rows out
		if (fits_rows) {
			const uint32_t own0 = S.rowoff[nh];
			int64_t *dst = R.col + (R.rp[i0 + nh] - rp_from);
			for (uint32_t e = tid; e < total - own0; e += TPB) dst[e] = S.rows[own0 + e];
		}
		if (tid == 0 && S.bad) raise(g, ERR_DEGREE);
		__syncthreads();
	}
}

size_t tile_smem_bytes() { return sizeof(TileSmem); }

int launch_tile(const Launch &L, const GraphDev &g, int def, const RangeDev &r, const TileDev &t, int blocks) {
	if (r.hi <= r.from) return 0;
	const size_t smem = sizeof(TileSmem);
	cudaError_t e;
	switch (def) {
	case 2: e = cudaFuncSetAttribute(k_tile<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); break;
	case 1: e = cudaFuncSetAttribute(k_tile<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); break;
	default: e = cudaFuncSetAttribute(k_tile<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
	}
	if (e != cudaSuccess) return (int)e;
	const int64_t ntiles = (r.hi - r.from + TT - 1) / TT;
	const unsigned grid = (unsigned)std::min<int64_t>(ntiles, blocks);
	switch (def) {
	case 2: k_tile<2><<<grid, TPB, smem, L.stream>>>(g, r, t); break;
	case 1: k_tile<1><<<grid, TPB, smem, L.stream>>>(g, r, t); break;
	default: k_tile<0><<<grid, TPB, smem, L.stream>>>(g, r, t);
	}
	WGBV_LAUNCHED(L);
	return 0;
}

}  // namespace wgbv
