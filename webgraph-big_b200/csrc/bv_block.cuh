// bv_block.cuh -- block-level helpers shared by the kernels: launch geometry and the in-block work binning.
#pragma once
#include <stdint.h>

#include "bv_kernels.cuh"

namespace wgbv {

static constexpr int TPB = 256;

static inline unsigned grid_for(int64_t items, int per_block) {
	int64_t b = (items + per_block - 1) / per_block;
	if (b < 1) b = 1;
	return (unsigned)b;
}

#define WGBV_LAUNCHED(L) do { if ((L).launches) ++*(L).launches; } while (0)

// Work binning.  Rows differ wildly in length, and a warp that mixes a 3-successor row with a 300-successor row
// spends most of its issue slots on a handful of lanes (the first version of these kernels ran at ~4 active lanes
// per instruction).  So every block first parses the heads of its 256 rows, then sorts the rows by the amount of
// serial work they carry (a counting sort over 64 classes in shared memory, heaviest first), and only then does
// the serial part: the 32 lanes of a warp now run rows of the same class.
static constexpr int BIN_CLASSES = 64;        // work classes
static constexpr int NBUCKETS = 2 * BIN_CLASSES;  // sort buckets: (class, kind) pairs

// class of a work amount w >= 1: exact below 32, then two classes per octave; returned heaviest-first
__device__ __forceinline__ int work_bucket(uint32_t w) {
	int c;
	if (w < 32) c = (int)w;
	else {
		const int lg = 31 - __clz(w);                 // >= 5
		c = 32 + 2 * (lg - 5) + (int)((w >> (lg - 1)) & 1);
		if (c > BIN_CLASSES - 1) c = BIN_CLASSES - 1;
	}
	return BIN_CLASSES - 1 - c;
}

// Block-wide counting sort of up to TPB * RPT items, RPT per thread (bucket[q] < 0: no item in slot q; the id of
// slot q of thread t is q * TPB + t).  On return order[0..*total) lists the item ids by ascending bucket, i.e.
// heaviest class first.  All threads of the block must call it.
template <int RPT>
__device__ __forceinline__ void block_bin_sort(const int (&bucket)[RPT], uint32_t *hist /*[NBUCKETS+1]*/, uint16_t *order,
                                               uint32_t *total) {
	if (threadIdx.x <= NBUCKETS) hist[threadIdx.x] = 0;
	__syncthreads();
	uint32_t rank[RPT];
#pragma unroll
	for (int q = 0; q < RPT; q++) rank[q] = bucket[q] >= 0 ? atomicAdd(&hist[bucket[q]], 1u) : 0u;
	__syncthreads();
	if (threadIdx.x < 32) {
		// exclusive scan of the 128 counters by one warp, four per lane
		uint32_t v[4];
		uint32_t sum = 0;
#pragma unroll
		for (int j = 0; j < 4; j++) { v[j] = hist[4 * threadIdx.x + j]; sum += v[j]; }
		uint32_t inc = sum;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
			if ((int)threadIdx.x >= o) inc += t;
		}
		uint32_t run = inc - sum;
#pragma unroll
		for (int j = 0; j < 4; j++) { hist[4 * threadIdx.x + j] = run; run += v[j]; }
		if (threadIdx.x == 31) hist[NBUCKETS] = inc;
	}
	__syncthreads();
#pragma unroll
	for (int q = 0; q < RPT; q++)
		if (bucket[q] >= 0) order[hist[bucket[q]] + rank[q]] = (uint16_t)(q * TPB + threadIdx.x);
	*total = hist[NBUCKETS];
	__syncthreads();
}


}  // namespace wgbv
