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
	bool ok;           // false: inconsistent counts (an error bit has been raised); d and ncopied are zeroed
};

// Reads outdegree, reference, block count and sums the blocks; leaves the reader after the last block (at the
// interval section).  dref = outdegree of the reference row (only used when rf > 0).  BVG:1010-1033.
template <int DEF, class RD>
__device__ __forceinline__ void row_head(const GraphDev &g, RD &r, uint32_t dref, RowHead &h) {
	const uint64_t d = rd_outdegree<DEF>(r, g);
	h.d = (uint32_t)d;
	h.rf = 0;
	h.bc = 0;
	h.ncopied = 0;
	h.blocks_pos = 0;
	h.ok = true;
	if (d != 0 && g.window > 0) h.rf = (uint32_t)rd_reference<DEF>(r, g);
	if (h.rf > 0) {
		uint64_t bc = rd_blockcount<DEF>(r, g);
		if (bc > (uint64_t)dref + 1) { h.ok = false; bc = 0; }
		h.bc = (uint32_t)bc;
		h.blocks_pos = r.tell();
		uint64_t copied = 0, total = 0;
		for (uint32_t i = 0; i < h.bc; i++) {
			const uint64_t b = rd_block<DEF>(r, g) + (i ? 1 : 0);
			total += b;
			if (!(i & 1)) copied += b;
		}
		if (total > dref) h.ok = false;
		else if (!(h.bc & 1)) copied += dref - total;
		if (copied > d) h.ok = false;
		h.ncopied = (uint32_t)copied;
	}
	if (!h.ok) {
		raise(g, ERR_DEGREE);
		h.d = 0;
		h.ncopied = 0;
		h.bc = 0;
	}
}

// Decodes the interval section and the residuals of row x (reader positioned at the interval count, i.e. right
// after the blocks) and writes the `nextra` merged values, in increasing order, to out[0..nextra).
// BVG:1037-1073, 902-954.  Warp-synchronous: every lane of `mask` must call it (nextra == 0: no work).
// Two loops: a merging one that runs while some lane of the warp still has interval material, then a tight one
// for pure gap decoding (the kernels sort rows so that most warps never enter the first).
template <int DEF, class RD>
__device__ __forceinline__ void row_extras(const GraphDev &g, RD &r, int64_t x, uint32_t nextra, int64_t *out, unsigned mask) {
	uint32_t ic = 0, nres = nextra;
	uint64_t int_pos = 0;
	if (nextra && g.min_interval != 0) {
		uint64_t c = r.gamma();
		if (c > nextra) { raise(g, ERR_DEGREE); c = 0; nextra = 0; }
		ic = (uint32_t)c;
		int_pos = r.tell();
		// first pass over the intervals: only their total length, to find the number of residuals and where they start
		uint64_t covered = 0;
		for (uint32_t i = 0; i < ic; i++) {
			r.gamma();
			covered += r.gamma() + (uint64_t)g.min_interval;
		}
		if (covered > nextra) { raise(g, ERR_DEGREE); ic = 0; nextra = 0; covered = 0; }
		nres = nextra - (uint32_t)covered;
	}
	RD ri = r;  // interval cursor; r stays the residual cursor
	if (ic) ri.seek(int_pos);

	int64_t cur = 0, end = 0;  // current interval [cur, end)
	int64_t res = x;           // last residual decoded; pending (not yet written) iff have_res
	bool have_res = false, first_res = true, first_int = true;
	uint32_t o = 0;
	for (; __any_sync(mask, o < nextra && (cur != end || ic != 0)); o++) {
		if (o < nextra) {
			if (cur == end && ic) {  // next interval (rare: once per interval)
				const uint64_t gap = ri.gamma();
				cur = first_int ? x + nat2int(gap) : end + 1 + (int64_t)gap;
				first_int = false;
				end = cur + (int64_t)ri.gamma() + g.min_interval;
				ic--;
			}
			if (!have_res && nres) {
				const uint64_t code = rd_residual<DEF>(r, g);
				res = first_res ? x + nat2int(code) : res + 1 + (int64_t)code;
				first_res = false;
				have_res = true;
				nres--;
			}
			const bool take_int = cur < end && (!have_res || cur < res);
			out[o] = take_int ? cur : res;
			if (take_int) cur++;
			else have_res = false;
		}
	}
	if (have_res && o < nextra) out[o++] = res;  // a residual decoded above but not written yet
	for (; __any_sync(mask, o < nextra); o++) {
		if (o < nextra) {
			const uint64_t code = rd_residual<DEF>(r, g);
			res = first_res ? x + nat2int(code) : res + 1 + (int64_t)code;
			first_res = false;
			out[o] = res;
		}
	}
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
				if ((k & 3u) == 0) prefetch_l1(src + k + 8);
				if (e < c) {
					dst[out++] = e;
					if ((ei & 3u) == 0) prefetch_l1(ext + ei + 8);
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

}  // namespace wgbv
