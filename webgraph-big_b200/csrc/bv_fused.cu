// bv_fused.cu -- the whole BVGraph -> CSR decode of a node range in ONE pass over the bit stream (sm_100a).
//
//   global memory          shared memory of one block (one tile)                                    global memory
//   Elias-Fano offsets -->  record positions
//   compressed bits    -->  staged words -> outdegree/reference -> row slots (scan) -> blocks/masks
//   (coalesced, once)       -> intervals + residuals into the slots' tails -> copy levels, in place  --> rowptr, col
//                                                                                                    (coalesced, once)
//
// What the reference does per node with a tree of lazy iterators (BVGraph.java:995-1097: outdegree, reference, copy
// blocks, intervals, residuals; MaskedLongIterator.java:67-100, MergedLongIterator.java:54-92) a block does here for a
// tile of consecutive nodes, entirely on chip: every thread owns one row of the tile, rows are kept as uint32 node
// ids in shared memory, the serial bit decoding of a row reads staged words only, reference rows are read from
// shared memory, and the finished tile leaves as whole 128-byte lines.  Nothing is written that is not output:
// no degree / reference / depth arrays, no per-level lists, no row descriptors in global memory.
//
// A row may reference one of the window_size rows before it, for at most max_ref_count hops (BVG:701, :2258), so a tile
// depends on at most halo = window_size * max_ref_count rows before its first one.  The block decodes those of them
// its rows really need, redundantly (the CPU reference does the same when an iterator starts in the middle of a
// graph: window pre-fill, BVG:1139-1143), and therefore never waits for another block.  A block has 256 threads and
// handles 256 - halo tile rows plus the halo rows: one row per thread throughout.  The same mechanism gives shards and
// chunked scans their halo for free: the first tile of a range simply looks back into rows before the range.
//
// CSR row pointers need the number of arcs before the tile: tiles are handed out in node order by a ticket, publish
// the arcs of their rows as soon as the outdegrees are parsed, and resolve their base by decoupled look-back just
// before they write (by then the predecessors' aggregates have long been published).
//
// Tiles that do not fit the shared-memory budget (a very long row, unusually many arcs or bits) are only counted by
// this kernel -- their row pointers are written, their successors are not -- and decoded afterwards by k_slow, which
// works in global memory at the rows' final places, reads rows of earlier tiles (complete by then) where this
// kernel decodes halo rows, and waits for the previous tile only when that one is slow as well.
#include <algorithm>

#include "bv_block.cuh"
#include "bv_rows.cuh"

namespace wgbv {

static constexpr int FTH = 256;  // threads per block = rows (halo + tile) per block
static_assert(FTH == TPB, "block_bin_sort numbers items by TPB");
#ifndef WGBV_FRC
#define WGBV_FRC 7168  // successors a tile (with its needed halo rows) may hold in shared memory
#endif
#ifndef WGBV_FBW
#define WGBV_FBW 2560  // 32-bit words of the bit stream a tile may stage
#endif
#ifndef WGBV_FBLOCKS
#define WGBV_FBLOCKS 4  // resident blocks per SM the register allocation aims at
#endif
static constexpr int FRC = WGBV_FRC;
static constexpr int FBW = WGBV_FBW;
static constexpr uint32_t FLEVELS = 8;  // longest reference chain the kernel follows
static constexpr int FCLS = 16;         // sort classes per level, see fused_class()
static_assert(FLEVELS * FCLS <= NBUCKETS, "level x class must fit the sort buckets");
static_assert(FRC < 65536, "copied / extra counts of a tile row are kept in 16 bits");

enum : uint64_t { FLAG_AGG = 1ull << 62, FLAG_PFX = 2ull << 62, FLAG_MASK = 3ull << 62 };

template <class E>
struct FusedSmem {
	uint32_t bits[FBW];
	E rows[FRC];
	uint64_t m0[FTH], m1[FTH];  // rows with a reference: mask of copied positions, or (first block's position, ncopied | bc << 32)
	uint32_t pos[FTH];          // bit positions inside the staged words: of the interval section,
	uint32_t rpos[FTH];         // of the first residual,
	uint32_t roff[FTH + 1];     // of the record itself (entry j + 1: where record j ends)
	uint32_t rowoff[FTH + 1];   // first slot of each row in rows[]
	uint32_t deg[FTH];
	uint16_t ncopied[FTH], nres[FTH];
	uint16_t order[FTH];
	uint8_t ref[FTH], need[FTH];
	uint32_t hist[NBUCKETS + 1];
	uint64_t warp_sum[FTH / 32];
	uint64_t span[2];
	uint64_t tile_base;
	uint32_t tile, max_depth;
};

__device__ __forceinline__ void cp_async16_f(void *smem_dst, const void *gmem_src) {
	const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src));
}

// How a row with a reference is merged, by its outdegree d (masked: its reference row has at most 128 successors and
// d <= 256, so the copy blocks are held as a mask of copied positions):
//   class 0       not masked: a whole warp, from the block codes (row_copy_warp)
//   class 1       64 <= d <= 256: 32 lanes by merge path (group_merge<32>)
//   class 2       16 <= d < 64: 8 lanes by merge path (group_merge<8>), four rows per warp
//   classes 3..15 d < 16: one lane per row, rows of (nearly) one length per warp (row_merge_lean)
__device__ __forceinline__ int fused_class(uint32_t d, bool masked) {
	if (!masked) return 0;
	if (d >= 64) return 1;
	if (d >= 16) return 2;
	return 3 + (d >= 3 ? 15 - (int)d : 12);
}

// Outdegree and reference of row x (readOutdegree / readReference, BVG:654-661, 692-703) with k_header's checks.
template <int DEF, class RD>
__device__ __forceinline__ void fused_head(const GraphDev &g, RD &r, int64_t x, uint32_t &d, uint32_t &rf) {
	d = 0;
	rf = 0;
	const uint64_t dd = rd_outdegree<DEF>(r, g);
	if (dd > 0x7fffffffull) { raise(g, ERR_DEGREE); return; }
	d = (uint32_t)dd;
	if (d && g.window > 0) {
		const uint64_t f = rd_reference<DEF>(r, g);
		if (f > (uint64_t)g.window || (int64_t)f > x) raise(g, ERR_REF_WINDOW);
		else rf = (uint32_t)f;
	}
	if (r.past_end()) raise(g, ERR_OVERRUN);
}

// Exclusive prefix of this tile's arcs over all earlier tiles (decoupled look-back, first warp only; the tile's own
// aggregate has been published).  Returns the prefix in every lane and publishes the inclusive prefix.
__device__ __forceinline__ uint64_t fused_lookback(volatile uint64_t *state, uint32_t tile, uint64_t own_total, unsigned lane) {
	uint64_t exclusive = 0;
	if (tile == 0) return 0;  // (tile 0 published its inclusive prefix at once)
	int64_t look = (int64_t)tile - 1;
	for (;;) {
		const int64_t idx = look - lane;
		uint64_t st = idx >= 0 ? state[idx] : FLAG_PFX;
		while (__any_sync(FULL_MASK, (st & FLAG_MASK) == 0))
			if ((st & FLAG_MASK) == 0) st = state[idx];
		const unsigned pfx = __ballot_sync(FULL_MASK, (st & FLAG_MASK) == FLAG_PFX);
		const int first = pfx ? __ffs(pfx) - 1 : 32;
		uint64_t contrib = ((int)lane <= first && idx >= 0) ? (st & ~FLAG_MASK) : 0;
#pragma unroll
		for (int o = 16; o; o >>= 1) contrib += __shfl_xor_sync(FULL_MASK, contrib, o);
		exclusive += contrib;
		if (pfx) break;
		look -= 32;
	}
	if (lane == 0) state[tile] = FLAG_PFX | (exclusive + own_total);
	return exclusive;
}

// Writes count successors from a shared-memory row store to their final place, widening to the ABI's int64.
template <class E>
__device__ __forceinline__ void fused_flush(const E *src, int64_t *dst, uint32_t count) {
	const uint32_t tid = threadIdx.x;
	if (sizeof(E) == 4 && ((uintptr_t)dst & 7u) == 0) {
		// 16-byte stores (two successors each); a leading / trailing odd element goes alone
		const uint32_t k0 = (uint32_t)(((uintptr_t)dst >> 3) & 1u);  // 1: dst[0] is the upper half of a 16-byte unit
		if (k0 && tid == 0 && count) dst[0] = (int64_t)src[0];
		const uint32_t pairs = count > k0 ? (count - k0) >> 1 : 0;
		for (uint32_t q = tid; q < pairs; q += FTH) {
			const uint32_t k = k0 + 2 * q;
			*reinterpret_cast<longlong2 *>(dst + k) = make_longlong2((long long)src[k], (long long)src[k + 1]);
		}
		if (tid == 32 && count > k0 && ((count - k0) & 1u)) dst[count - 1] = (int64_t)src[count - 1];
	} else {
		for (uint32_t k = tid; k < count; k += FTH) dst[k] = (int64_t)src[k];
	}
}

template <int DEF, class E>
__global__ void __launch_bounds__(FTH, WGBV_FBLOCKS) k_fused(GraphDev g, FusedDev F) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	FusedSmem<E> &S = *reinterpret_cast<FusedSmem<E> *>(smem_raw);
	const int tid = threadIdx.x;
	const unsigned lane = tid & 31u, warp = tid >> 5;
	if (tid == 0) {
		S.tile = atomicAdd(F.ticket, 1u);
		S.max_depth = 0;
	}
	__syncthreads();
	const uint32_t tile = S.tile;
	const int64_t t0 = F.from + (int64_t)tile * F.tile_rows;  // first row of the tile (node id)
	const int64_t t1 = t0 + F.tile_rows < F.to ? t0 + F.tile_rows : F.to;
	const int64_t h0 = t0 - F.halo > g.ef_first ? t0 - F.halo : g.ef_first;  // first halo row
	const int nh = (int)(t0 - h0), nn = (int)(t1 - h0);                      // nn <= FTH: thread j owns row h0 + j
	volatile uint64_t *state = F.tile_state;

	// ---- record offsets (Elias-Fano), span of the bit stream
	uint64_t off = 0;
	if (tid < nn) off = ef_get(g, h0 + tid);
	if (tid == 0) S.span[0] = off;
	if (tid == nn - 1) S.span[1] = ef_get(g, t1);
	__syncthreads();
	const uint64_t w0 = (S.span[0] >> 5) & ~3ull;  // 16-byte aligned: the staging copies are 16 bytes each
	const uint64_t nw64 = ((((S.span[1] + 31) >> 5) - w0 + 2) + 3) & ~3ull;
	const bool staged = nw64 <= (uint64_t)FBW;
	const uint32_t nw = staged ? (uint32_t)nw64 : 0u;
	if (staged) {
		const uint64_t avail = g.nwords * 2 > w0 ? g.nwords * 2 - w0 : 0;
		const uint32_t *src = (const uint32_t *)g.words + w0;
		for (uint32_t w = tid * 4; w < nw; w += FTH * 4) {
			if (w + 4 <= avail) cp_async16_f(&S.bits[w], src + w);
			else
				for (int j = 0; j < 4; j++) S.bits[w + j] = w + j < avail ? __ldg(src + w + j) : 0u;
		}
		asm volatile("cp.async.wait_all;" ::: "memory");
	}
	__syncthreads();

	// ---- outdegree and reference of every row
	uint32_t d = 0, rf = 0, after = 0;
	if (staged) {
		if (tid < nn) S.roff[tid] = (uint32_t)(off - (w0 << 5));
		if (tid == 0) S.roff[nn] = (uint32_t)(S.span[1] - (w0 << 5));
	}
	if (tid < nn) {
		if (staged) {
			PosReader r;
			r.init_words(S.bits, nw, (uint32_t)(off - (w0 << 5)));
			fused_head<DEF>(g, r, h0 + tid, d, rf);
			after = (uint32_t)r.tell();
		} else {
			BitReader r;
			r.init(g, off);
			fused_head<DEF>(g, r, h0 + tid, d, rf);
		}
	}
	S.deg[tid] = d;
	S.ref[tid] = (uint8_t)rf;
	S.need[tid] = tid >= nh && tid < nn;
	__syncthreads();

	// ---- reference chains: depth of every row; the tile's rows mark the halo rows they need
	uint32_t depth = 0;
	bool valid = true;
	if (tid < nn && rf) {
		const bool own = tid >= nh;
		int y = tid;
		uint32_t f = rf;
		while (f) {
			y -= (int)f;
			if (y < 0 || depth == FLEVELS) { valid = false; break; }
			depth++;
			if (own && y < nh) S.need[y] = 1;
			f = S.ref[y];
		}
		// a chain that leaves the rows at hand or is longer than the properties promise: the general path takes over
		if (!valid && own) raise(g, ERR_FUSED_DEEP);
	}
	__syncthreads();
	const bool needed = tid < nn && S.need[tid] != 0 && valid;

	// ---- row slots: exclusive scan of the outdegrees of the needed rows (saturated: anything that large is slow anyway)
	const uint32_t v = needed ? (d < (1u << 23) ? d : (1u << 23)) : 0u;
	uint32_t inc = v;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const uint32_t t = __shfl_up_sync(FULL_MASK, inc, o);
		if ((int)lane >= o) inc += t;
	}
	if (lane == 31) S.warp_sum[warp] = inc;
	__syncthreads();
	uint32_t wbase = 0, total = 0;
#pragma unroll
	for (int w = 0; w < FTH / 32; w++) {
		const uint32_t ws = (uint32_t)S.warp_sum[w];
		if (w < (int)warp) wbase += ws;
		total += ws;
	}
	const uint32_t myoff = wbase + inc - v;
	S.rowoff[tid] = myoff;
	if (tid == FTH - 1) S.rowoff[FTH] = total;
	const bool fast = staged && total <= (uint32_t)FRC;
	__syncthreads();

	if (!fast) {
		// ---- a slow tile: exact row pointers now, successors by k_slow
		const bool own = tid >= nh && tid < nn;
		const uint64_t dv = own ? d : 0;
		uint64_t inc64 = dv;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			const uint64_t t = __shfl_up_sync(FULL_MASK, inc64, o);
			if ((int)lane >= o) inc64 += t;
		}
		if (lane == 31) S.warp_sum[warp] = inc64;
		__syncthreads();
		uint64_t base64 = 0, total64 = 0;
#pragma unroll
		for (int w = 0; w < FTH / 32; w++) {
			const uint64_t ws = S.warp_sum[w];
			if (w < (int)warp) base64 += ws;
			total64 += ws;
		}
		if (warp == 0) {
			if (lane == 0) state[tile] = (tile == 0 ? FLAG_PFX : FLAG_AGG) | total64;
			const uint64_t ex = fused_lookback(state, tile, total64, lane);
			if (lane == 0) S.tile_base = ex;
		}
		__syncthreads();
		const uint64_t tb = S.tile_base;
		if (own) F.rowptr[t0 - F.from + (tid - nh)] = (int64_t)(tb + base64 + inc64 - dv);
		if (tid == 0) {
			F.rowptr[t1 - F.from] = (int64_t)(tb + total64);
			F.tile_done[tile] = 0;
			atomicAdd(F.slow_count, 1u);
		}
		return;
	}

	// ---- publish the arcs of the tile's own rows (look-back happens at the end)
	const uint32_t own0 = S.rowoff[nh];
	const uint32_t own_total = total - own0;
	if (tid == 0) state[tile] = (tile == 0 ? FLAG_PFX : FLAG_AGG) | (uint64_t)own_total;

	// ---- blocks (as a mask of copied positions), counts, the interval section, where the residuals start
	int bucket[1] = {-1};
	uint32_t nextra = 0, ncop = 0, nres = 0;
	bool masked = false;
	if (needed && d) {
		PosReader r;
		r.init_words(S.bits, nw, after);
		const uint32_t dref = rf ? S.deg[tid - (int)rf] : 0u;
		RowHead h;
		row_blocks<DEF>(g, r, d, rf, dref, h);
		ncop = h.ncopied;
		nextra = h.d - h.ncopied;
		if (rf) {
			masked = d <= 256 && dref <= 128;
			S.m0[tid] = masked ? h.cmask[0] : h.blocks_pos;
			S.m1[tid] = masked ? h.cmask[1] : ((uint64_t)h.ncopied | (uint64_t)h.bc << 32);
		}
		if (nextra) {
			S.pos[tid] = (uint32_t)r.tell();
			uint32_t ic;
			uint64_t int_pos;
			row_intervals(g, r, nextra, ic, int_pos, nres);
			S.rpos[tid] = (uint32_t)r.tell();
			// Long residual lists go to a whole warp each (buckets 0 and 1: first in the order); of the others, ALL rows with
			// intervals come first (buckets 16..63), then the rest (80..127), each part sorted by work: warps without interval
			// rows never enter the merging loop of row_extras_core.  (A tile holds fewer than 8192 successors, so the work
			// classes of its rows start at 16.)
			if (nres >= COOP_LONG && ic <= 32) bucket[0] = ic ? 0 : 1;
			else if (nextra) bucket[0] = (ic ? 0 : BIN_CLASSES) + work_bucket(nextra);
		}
	}
	S.ncopied[tid] = (uint16_t)ncop;
	S.nres[tid] = (uint16_t)nres;
	uint32_t nwork;
	block_bin_sort<1>(bucket, S.hist, S.order, &nwork);
	const uint32_t nlong = S.hist[2];

	// ---- extras: intervals and residuals of every row into the tail of its slot
	{
		PosReader proto;
		proto.init_words(S.bits, nw, 0);
		// long residual lists: a warp each
#pragma unroll 1
		for (uint32_t t = warp; t < nlong; t += FTH / 32) {
			const int id = S.order[t];
			const uint32_t nc = S.ncopied[id], nr = S.nres[id], ne = S.deg[id] - nc;
			E *ext = S.rows + S.rowoff[id] + nc;
			coop_residuals<DEF>(g, proto, S.rpos[id], S.roff[id + 1], nr, h0 + id, ext + (ne - nr));
			if (ne != nr) {
				__syncwarp();
				coop_intervals_merge(g, proto, S.pos[id], h0 + id, ext, ne, nr);
			}
		}
		__syncwarp();
		// the others: rows of one work class per warp, in lock step; the heaviest group goes to the warp with the fewest long rows
		const uint32_t t = nlong + (FTH / 32 - 1 - warp) * 32 + lane;
		const bool active = t < nwork;
		const unsigned mask = __ballot_sync(FULL_MASK, active);
		if (active) {
			const int id = S.order[t];
			const uint32_t nc = S.ncopied[id], nr = S.nres[id], ne = S.deg[id] - nc;
			PosReader r = proto, ri = proto;
			r.seek(S.rpos[id]);
			uint32_t ic = 0;
			if (ne != nr) {
				ri.seek(S.pos[id]);
				ic = (uint32_t)ri.gamma();
			}
			NoWindow nwin;
			row_extras_core<DEF>(g, ri, r, h0 + id, ne, ic, nr, S.rows + S.rowoff[id] + nc, mask, nwin);
			if (r.past_end()) raise(g, ERR_OVERRUN);
		}
	}
	// ---- copy blocks, one depth level after the other (rows sorted by level, then work class)
	bucket[0] = -1;
	if (needed && rf && ncop) {
		bucket[0] = (int)(depth - 1) * FCLS + fused_class(d, masked);
		atomicMax(&S.max_depth, depth);
	}
	uint32_t ncopy;
	block_bin_sort<1>(bucket, S.hist, S.order, &ncopy);  // (its first barrier also ends the extras phase)
	const uint32_t max_depth = S.max_depth;
	for (uint32_t l = 1; l <= max_depth; l++) {
		const uint32_t *hb = S.hist + (l - 1) * FCLS;
		const uint32_t c0 = hb[0], c1 = hb[1], c2 = hb[2], c3 = hb[3], c16 = hb[FCLS];
		// rows without a mask: one warp per row, from the block codes
#pragma unroll 1
		for (uint32_t t = c0 + warp; t < c1; t += FTH / 32) {
			const int id = S.order[t];
			const int y = id - (int)S.ref[id];
			const uint32_t nc = (uint32_t)S.m1[id];
			E *dst = S.rows + S.rowoff[id];
			PosReader r;
			r.init_words(S.bits, nw, (uint32_t)S.m0[id]);
			row_copy_warp<DEF>(g, r, (const E *)(S.rows + S.rowoff[y]), S.deg[y], (const E *)(dst + nc), dst, S.deg[id], (uint32_t)(S.m1[id] >> 32), nc);
		}
		__syncwarp();
		// long masked rows: 32 lanes per row
#pragma unroll 1
		for (uint32_t t = c1 + warp; t < c2; t += FTH / 32) {
			const int id = S.order[t];
			const int y = id - (int)S.ref[id];
			group_merge<32, E>(g, true, (const E *)(S.rows + S.rowoff[y]), S.m0[id], S.m1[id], S.rows + S.rowoff[id], S.deg[id], S.ncopied[id], lane);
		}
		// medium rows: 8 lanes per row, four rows per warp
#pragma unroll 1
		for (uint32_t tb = c2 + warp * 4; tb < c3; tb += (FTH / 32) * 4) {
			const uint32_t t = tb + (lane >> 3);
			const bool active = t < c3;
			const int id = active ? S.order[t] : 0;
			const int y = id - (int)S.ref[id];
			group_merge<8, E>(g, active, (const E *)(S.rows + S.rowoff[y]), S.m0[id], S.m1[id], S.rows + S.rowoff[id], S.deg[id], S.ncopied[id], lane & 7u);
		}
		// short rows: one lane per row
#pragma unroll 1
		for (uint32_t tb = c3 + (tid & ~31u); tb < c16; tb += FTH) {
			const uint32_t t = tb + lane;
			const bool active = t < c16;
			const unsigned mask = __ballot_sync(FULL_MASK, active);
			if (active) {
				const int id = S.order[t];
				const int y = id - (int)S.ref[id];
				const uint32_t nc = S.ncopied[id];
				E *dst = S.rows + S.rowoff[id];
				row_merge_lean<E>(g, (const E *)(S.rows + S.rowoff[y]), S.m0[id], S.m1[id], (const E *)(dst + nc), dst, S.deg[id] - nc, mask);
			}
		}
		__syncthreads();
	}

	// ---- where the tile goes, then out in whole lines
	if (warp == 0) {
		const uint64_t ex = fused_lookback(state, tile, own_total, lane);
		if (lane == 0) S.tile_base = ex;
	}
	__syncthreads();
	const uint64_t tb = S.tile_base;
	if (tb + own_total > F.col_capacity) {
		if (tid == 0) raise(g, ERR_CAPACITY);
	} else fused_flush<E>(S.rows + own0, F.col + tb, own_total);
	if (tid >= nh && tid < nn) F.rowptr[t0 - F.from + (tid - nh)] = (int64_t)(tb + (myoff - own0));
	if (tid == 0) {
		F.rowptr[t1 - F.from] = (int64_t)(tb + own_total);
		F.tile_done[tile] = 1;
	}
}

// ------------------------------------------------------------------------------------------------ slow tiles
// Tiles k_fused left undone, decoded in global memory at the rows' final places with the row functions of the
// level-synchronous path.  Their row pointers exist, every fast tile is complete, so a reference to a row before the
// tile reads final data; only when the previous tile is slow as well does a tile wait for it (tile_done).  Blocks take
// chunks of SCH consecutive tiles by ticket and decode the slow ones of their chunk in order, so whatever a tile
// waits for belongs to the same block (done already) or to a block with an earlier ticket (running): no deadlock.
static constexpr int SCH = 64;

struct SlowSmem {
	uint64_t a[FTH], b[FTH];  // RowInfo of the rows with a reference
	uint64_t pos[FTH];        // bit position of the interval section
	uint64_t rpos[FTH];       // bit position of the first residual
	uint32_t deg[FTH], nres[FTH], ncopied[FTH];
	uint32_t hist[NBUCKETS + 1];
	uint16_t order[FTH];
	uint16_t list[SCH];
	uint8_t ref[FTH], dl[FTH];
	uint32_t chunk, nlist, max_dl, needs_prev;
};

__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t *p) {
	uint32_t v;
	asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
	return v;
}

template <int DEF>
__device__ __forceinline__ void slow_tile(const GraphDev &g, const FusedDev &F, SlowSmem &S, uint32_t tile) {
	const int tid = threadIdx.x;
	const unsigned lane = tid & 31u, warp = tid >> 5;
	const int64_t t0 = F.from + (int64_t)tile * F.tile_rows;
	const int64_t t1 = t0 + F.tile_rows < F.to ? t0 + F.tile_rows : F.to;
	const int own = (int)(t1 - t0);
	const int64_t i0 = t0 - F.from;  // index of the tile's first row in rowptr
	if (tid == 0) {
		S.max_dl = 0;
		S.needs_prev = 0;
	}
	__syncthreads();
	// ---- heads
	uint32_t d = 0, rf = 0, nextra = 0, ncop = 0, nres = 0;
	bool masked = false;
	int bucket[1] = {-1};
	if (tid < own) {
		const int64_t x = t0 + tid;
		BitReader r;
		r.init(g, ef_get(g, x));
		fused_head<DEF>(g, r, x, d, rf);
		if ((uint64_t)F.rowptr[i0 + tid + 1] > F.col_capacity) {
			raise(g, ERR_CAPACITY);
			d = 0;
			rf = 0;
		}
		uint32_t dref = 0;
		if (rf) {
			const int64_t yi = i0 + tid - (int64_t)rf;
			if (yi < 0) {
				// the reference row lies before the range: only the general path keeps halo rows in memory
				raise(g, ERR_FUSED_FALLBACK);
				d = 0;
				rf = 0;
			} else dref = (uint32_t)(F.rowptr[yi + 1] - F.rowptr[yi]);
		}
		if (d) {
			RowHead h;
			row_blocks<DEF>(g, r, d, rf, dref, h);
			ncop = h.ncopied;
			nextra = h.d - h.ncopied;
			if (rf) {
				masked = row_is_masked(d, dref);
				S.a[tid] = masked ? h.cmask[0] : h.blocks_pos;
				S.b[tid] = masked ? h.cmask[1] : ((uint64_t)h.ncopied | (uint64_t)h.bc << 32);
				if (tid < (int)rf) S.needs_prev = 1;
			}
			if (nextra) {
				S.pos[tid] = r.tell();
				uint32_t ic;
				uint64_t int_pos;
				row_intervals(g, r, nextra, ic, int_pos, nres);
				S.rpos[tid] = r.tell();
				if (nres >= COOP_LONG && ic <= 32) bucket[0] = ic ? 0 : 1;  // (a whole warp each, see k_fused)
				else if (nextra) {
					const int wb = 2 * work_bucket(nextra);
					bucket[0] = (wb < 2 ? 2 : wb) + (ic ? 0 : 1);
				}
			}
		}
	}
	S.deg[tid] = d;
	S.ref[tid] = (uint8_t)rf;
	S.nres[tid] = nres;
	S.ncopied[tid] = ncop;
	uint32_t nwork;
	block_bin_sort<1>(bucket, S.hist, S.order, &nwork);
	const uint32_t nlong = S.hist[2];
	// ---- depth inside the tile (rows before the tile are final: depth 0)
	uint32_t dl = 0;
	if (rf) {
		int y = tid;
		uint32_t f = rf;
		while (f) {
			dl++;
			y -= (int)f;
			if (y < 0) break;
			f = S.ref[y];
		}
		if (ncop) atomicMax(&S.max_dl, dl);
	}
	// ---- extras into the tails of the rows' slots: long residual lists by a warp each, the others in lock step
#pragma unroll 1
	for (uint32_t t = warp; t < nlong; t += FTH / 32) {
		const int id = S.order[t];
		const int64_t x = t0 + id;
		const uint32_t nc = S.ncopied[id], nr = S.nres[id], ne = S.deg[id] - nc;
		int64_t *ext = F.col + F.rowptr[i0 + id] + nc;
		BitReader proto;
		proto.init(g, ef_get(g, x));
		coop_residuals<DEF>(g, proto, S.rpos[id], ef_get(g, x + 1), nr, x, ext + (ne - nr));
		if (ne != nr) {
			__threadfence_block();
			__syncwarp();
			coop_intervals_merge(g, proto, S.pos[id], x, ext, ne, nr);
		}
	}
	__syncwarp();
#pragma unroll 1
	for (uint32_t tb = nlong + (tid & ~31u); tb < nwork; tb += FTH) {
		const uint32_t t = tb + lane;
		const bool active = t < nwork;
		const unsigned mask = __ballot_sync(FULL_MASK, active);
		if (active) {
			const int id = S.order[t];
			const uint32_t nc = S.ncopied[id], nr = S.nres[id], ne = S.deg[id] - nc;
			BitReader r, ri;
			r.init(g, S.rpos[id]);
			ri = r;
			uint32_t ic = 0;
			if (ne != nr) {
				ri.init(g, S.pos[id]);
				ic = (uint32_t)ri.gamma();
			}
			NoWindow nwin;
			row_extras_core<DEF>(g, ri, r, t0 + id, ne, ic, nr, F.col + F.rowptr[i0 + id] + nc, mask, nwin);
			if (r.past_end()) raise(g, ERR_OVERRUN);
		}
	}
	__threadfence_block();
	__syncthreads();
	if (S.needs_prev && tile > 0) {
		if (tid == 0)
			while (ld_acquire_u32(F.tile_done + tile - 1) == 0) __nanosleep(200);
		__syncthreads();
	}
	// ---- copy blocks, level by level
	const uint32_t max_dl = S.max_dl;
	for (uint32_t l = 1; l <= max_dl; l++) {
		bucket[0] = (dl == l && ncop) ? 2 * work_bucket(masked || d >= WARP_ROW ? d : WARP_ROW) : -1;
		uint32_t ncopy;
		block_bin_sort<1>(bucket, S.hist, S.order, &ncopy);
		const uint32_t nbig = S.hist[2 * work_bucket(WARP_ROW - 1)];
#pragma unroll 1
		for (uint32_t t = warp; t < nbig; t += FTH / 32) {
			const int id = S.order[t];
			const int64_t k = i0 + id, y = k - S.ref[id];
			const uint32_t nc = (uint32_t)S.b[id];
			int64_t *dst = F.col + F.rowptr[k];
			BitReader r;
			r.init(g, S.a[id]);
			row_copy_warp<DEF>(g, r, (const int64_t *)(F.col + F.rowptr[y]), (uint32_t)(F.rowptr[y + 1] - F.rowptr[y]), (const int64_t *)(dst + nc), dst,
			                   S.deg[id], (uint32_t)(S.b[id] >> 32), nc, true);
		}
		__syncwarp();
#pragma unroll 1
		for (uint32_t tb = nbig + (tid & ~31u); tb < ncopy; tb += FTH) {
			const uint32_t t = tb + lane;
			const bool active = t < ncopy;
			const unsigned mask = __ballot_sync(FULL_MASK, active);
			if (active) {
				const int id = S.order[t];
				const int64_t k = i0 + id, y = k - S.ref[id];
				const uint32_t nc = S.ncopied[id];
				int64_t *dst = F.col + F.rowptr[k];
				row_copy_masked<int64_t>(g, (const int64_t *)(F.col + F.rowptr[y]), S.a[id], S.b[id], (const int64_t *)(dst + nc), dst, S.deg[id] - nc, mask, true);
			}
		}
		__threadfence_block();
		__syncthreads();
	}
	__threadfence();
	__syncthreads();
	if (tid == 0) atomicExch(F.tile_done + tile, 1u);
}

template <int DEF>
__global__ void __launch_bounds__(FTH) k_slow(GraphDev g, FusedDev F) {
	__shared__ SlowSmem S;
	__shared__ uint32_t s_wcnt[FTH / 32];
	const int tid = threadIdx.x;
	const unsigned lane = tid & 31u, warp = tid >> 5;
	if (*F.slow_count == 0) return;  // no slow tile at all
	for (;;) {
		__syncthreads();
		if (tid == 0) S.chunk = atomicAdd(F.ticket + 1, 1u);
		__syncthreads();
		const uint32_t first = S.chunk * (uint32_t)SCH;
		if (first >= F.ntiles) return;
		const uint32_t tl = first + tid;
		const bool slow = tid < SCH && tl < F.ntiles && F.tile_done[tl] == 0;
		const unsigned bal = __ballot_sync(FULL_MASK, slow);
		if (lane == 0) s_wcnt[warp] = (uint32_t)__popc(bal);
		__syncthreads();
		uint32_t base = 0, n = 0;
#pragma unroll
		for (int w = 0; w < FTH / 32; w++) {
			if (w < (int)warp) base += s_wcnt[w];
			n += s_wcnt[w];
		}
		if (slow) S.list[base + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)tid;
		__syncthreads();
		for (uint32_t q = 0; q < n; q++) slow_tile<DEF>(g, F, S, first + S.list[q]);
	}
}

// Rows per tile for a graph with the given mean outdegree and mean record length: as many as the block has threads for
// (256 minus the halo), fewer when an average tile of that many rows would not fit the shared-memory budget with
// room to spare (dense graphs).  0: the tile kernel is not worth using (fewer than 24 rows would fit).
int fused_tile_rows(int halo, double mean_degree, double mean_record_bits) {
	int rows = FTH - halo;
	const double by_arcs = 0.7 * FRC / (mean_degree > 1 ? mean_degree : 1) - halo * 0.3;
	const double by_bits = 0.7 * FBW * 32 / (mean_record_bits > 1 ? mean_record_bits : 1) - halo;
	if (by_arcs < rows) rows = (int)by_arcs;
	if (by_bits < rows) rows = (int)by_bits;
	return rows < 24 ? 0 : rows;
}

int launch_fused(const Launch &L, const GraphDev &g, int def, const FusedDev &F, int sm_count) {
	if (F.to <= F.from) return 0;
	using E = uint32_t;
	const size_t smem = sizeof(FusedSmem<E>);
	cudaError_t e;
#define WGBV_LAUNCH_FUSED(D)                                                                                    \
	do {                                                                                                        \
		e = cudaFuncSetAttribute(k_fused<D, E>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);        \
		if (e == cudaSuccess) k_fused<D, E><<<F.ntiles, FTH, smem, L.stream>>>(g, F);                           \
	} while (0)
	switch (def) {
	case 2: WGBV_LAUNCH_FUSED(2); break;
	case 1: WGBV_LAUNCH_FUSED(1); break;
	default: WGBV_LAUNCH_FUSED(0);
	}
#undef WGBV_LAUNCH_FUSED
	if (e != cudaSuccess) return (int)e;
	WGBV_LAUNCHED(L);
	const unsigned chunks = (F.ntiles + SCH - 1) / SCH;
	const unsigned grid = std::min<unsigned>(chunks, (unsigned)sm_count * 4u);
	switch (def) {
	case 2: k_slow<2><<<grid, FTH, 0, L.stream>>>(g, F); break;
	case 1: k_slow<1><<<grid, FTH, 0, L.stream>>>(g, F); break;
	default: k_slow<0><<<grid, FTH, 0, L.stream>>>(g, F);
	}
	WGBV_LAUNCHED(L);
	return 0;
}

}  // namespace wgbv
