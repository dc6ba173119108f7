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
#pragma once
#include "bv_device.cuh"

namespace wgbv {

struct RowHead {
	uint32_t d;        // outdegree
	uint32_t rf;       // reference (0: none)
	uint32_t bc;       // block count
	uint32_t ncopied;  // successors copied from the reference row
	uint64_t blocks_pos;  // bit position of the first block (after the block count)
};

// Reads outdegree, reference, block count and sums the blocks; leaves the reader after the last block (at the
// interval section).  dref = outdegree of the reference row (only used when rf > 0).  BVG:1010-1033.
template <bool DEF>
__device__ __forceinline__ bool row_head(const GraphDev &g, BitReader &r, uint32_t dref, RowHead &h) {
	const uint64_t d = rd_outdegree<DEF>(r, g);
	h.d = (uint32_t)d;
	h.rf = 0;
	h.bc = 0;
	h.ncopied = 0;
	h.blocks_pos = 0;
	if (d == 0) return true;
	if (g.window > 0) h.rf = (uint32_t)rd_reference<DEF>(r, g);
	if (h.rf > 0) {
		const uint64_t bc = rd_blockcount<DEF>(r, g);
		if (bc > (uint64_t)dref + 1) { raise(g, ERR_DEGREE); return false; }
		h.bc = (uint32_t)bc;
		h.blocks_pos = r.pos;
		uint64_t copied = 0, total = 0;
		for (uint32_t i = 0; i < h.bc; i++) {
			const uint64_t b = rd_block<DEF>(r, g) + (i ? 1 : 0);
			total += b;
			if (!(i & 1)) copied += b;
		}
		if (total > dref) { raise(g, ERR_DEGREE); return false; }
		if (!(h.bc & 1)) copied += dref - total;
		if (copied > d) { raise(g, ERR_DEGREE); return false; }
		h.ncopied = (uint32_t)copied;
	}
	return true;
}

// Decodes the interval section and the residuals of row x (reader positioned right after the blocks) and writes
// the `nextra` merged values, in increasing order, to out[0..nextra).  BVG:1037-1073, 902-954.
template <bool DEF>
__device__ __forceinline__ bool row_extras(const GraphDev &g, BitReader &r, int64_t x, uint32_t nextra, int64_t *out) {
	if (nextra == 0) return true;
	uint64_t ic = 0;
	uint64_t int_pos = 0;
	uint64_t nres = nextra;
	if (g.min_interval != 0) {
		ic = r.gamma();
		if (ic > nextra) { raise(g, ERR_DEGREE); return false; }
		int_pos = r.pos;
		// first pass over the intervals: only their total length, to find the number of residuals and where they start
		uint64_t covered = 0;
		for (uint64_t i = 0; i < ic; i++) {
			r.gamma();
			covered += r.gamma() + (uint64_t)g.min_interval;
			if (covered > nextra) { raise(g, ERR_DEGREE); return false; }
		}
		nres = nextra - covered;
	}
	BitReader ri = r;  // interval cursor (rewound below); r stays the residual cursor
	if (ic) ri.seek(int_pos);

	uint32_t o = 0;
	int64_t res = 0;
	if (nres) res = x + nat2int(rd_residual<DEF>(r, g));
	int64_t prev = 0;
	for (uint64_t i = 0; i < ic; i++) {
		int64_t left;
		if (i == 0) left = x + nat2int(ri.gamma());
		else left = prev + 1 + (int64_t)ri.gamma();
		const int64_t len = (int64_t)ri.gamma() + g.min_interval;
		prev = left + len;
		// residuals below the interval
		while (nres && res < left) {
			out[o++] = res;
			if (--nres) res += 1 + (int64_t)rd_residual<DEF>(r, g);
		}
		for (int64_t v = left; v < prev; v++) out[o++] = v;
	}
	while (nres) {
		out[o++] = res;
		if (--nres) res += 1 + (int64_t)rd_residual<DEF>(r, g);
	}
	if (r.past_end()) { raise(g, ERR_OVERRUN); return false; }
	return true;
}

// Copy-block resolution of one row: src[0..dref) is the finished reference row, dst[0..d) the row's slot whose
// last d - ncopied positions hold the extras.  The reader is positioned at the first block.
// MaskedLongIterator.java:67-100 + MergedLongIterator.java:54-92 of the reference.
template <bool DEF>
__device__ __forceinline__ void row_copy(const GraphDev &g, BitReader &r, const int64_t *__restrict__ src, uint32_t dref,
                                         int64_t *dst, uint32_t d, uint32_t bc, uint32_t ncopied) {
	uint32_t ei = ncopied;  // read index of the waiting extras
	uint32_t out = 0;
	uint32_t k = 0;
	int64_t e = ei < d ? dst[ei] : INT64_MAX;
	for (uint32_t i = 0; i <= bc; i++) {
		uint32_t len;
		if (i < bc) len = (uint32_t)rd_block<DEF>(r, g) + (i ? 1u : 0u);
		else len = (bc & 1) ? 0u : dref - k;  // implicit last block: copied iff the block count is even
		if (len > dref - k) len = dref - k;
		if (!(i & 1)) {
			for (uint32_t j = 0; j < len; j++) {
				const int64_t c = src[k + j];
				while (e < c) {
					dst[out++] = e;
					e = ++ei < d ? dst[ei] : INT64_MAX;
				}
				if (e == c) raise(g, ERR_NOT_SORTED);
				if (out < d) dst[out++] = c;
			}
		}
		k += len;
	}
	// the remaining extras already sit in their final positions (out == ei here)
}

}  // namespace wgbv
