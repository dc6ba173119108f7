// bv_labels.cu -- arc labels stored as a bit stream, decoded into an array parallel to the CSR successors.
//
// The reference keeps the labels of a BitStreamArcLabelledImmutableGraph in a `.labels` bit stream -- the labels of the
// arcs of node 0 in successor order, then those of node 1, ... -- and the bit position of every node's first label in
// `.labeloffsets`, gamma-coded differences exactly like `.offsets` (labelling/BitStreamArcLabelledImmutableGraph.java:
// 65-100, 339-363, 654-681).  Its node iterator positions the stream at the node's offset and calls
// label.fromBitStream() once per successor (:487-560).  Here the stream is held like the graph's (32-bit MSB-first
// words + Elias-Fano offsets, both built by the loader the graph uses); gamma-coded labels are read one node per lane,
// fixed-width ones one label per thread (see k_labels_fixed):
//   kind WGBV_LABEL_FIXED_INT  FixedWidthIntLabel: readInt(width)   (FixedWidthIntLabel.java:70-73)
//   kind WGBV_LABEL_GAMMA_INT  GammaCodedIntLabel: readGamma()      (GammaCodedIntLabel.java:60-64)
//   kind WGBV_LABEL_FIXED_INT_LIST / _LONG_LIST  FixedWidthIntListLabel / FixedWidthLongListLabel: readGamma() elements of
//        readInt(width) / readLong(width) each (FixedWidthIntListLabel.java:73-78, FixedWidthLongListLabel.java:78-84):
//        a ragged result, list_ptr[arc] .. list_ptr[arc + 1] into a values array (k_label_lists: lengths, scan, values)
// A node whose labels do not end where the next node's begin raises ERR_DEGREE (the stream, the offsets and the
// outdegrees do not belong together).
#include "bv_block.cuh"
#include "bv_kernels.cuh"

namespace wgbv {

template <int KIND>
__global__ void __launch_bounds__(TPB) k_labels(GraphDev lab, int64_t from, int64_t count, const int64_t *__restrict__ rp, int width,
                                                int64_t *__restrict__ out, uint64_t capacity) {
	const int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
	if (i >= count) return;
	const int64_t a = rp[i] - rp[0], b = rp[i + 1] - rp[0];
	if ((uint64_t)b > capacity) { raise(lab, ERR_CAPACITY); return; }
	const uint64_t pos = ef_get(lab, from + i), end = ef_get(lab, from + i + 1);
	BitReader r;
	r.init(lab, pos);
	for (int64_t j = a; j < b; j++) {
		const uint64_t v = KIND == 0 ? (width ? r.bits((unsigned)width) : 0) : r.gamma();
		out[j] = (int64_t)(int32_t)(uint32_t)v;  // the Java field is an int
	}
	if (r.tell() != end) raise(lab, ERR_DEGREE);
}

// Fixed-width labels need no per-node work: the labels of consecutive nodes are consecutive in the stream, so label j of
// the range sits at (offset of node `from`) + j * width.  One thread per label, coalesced 8-byte stores; the offsets are
// only used to check, one thread per node, that they say the same (k_labels_check).
__device__ __forceinline__ int64_t fixed_label(const uint32_t *__restrict__ words, uint64_t nw, uint64_t bit, int width) {
	const uint64_t w = bit >> 5;
	const uint32_t a = w < nw ? __ldg(words + w) : 0u, b = w + 1 < nw ? __ldg(words + w + 1) : 0u;
	const uint64_t two = (uint64_t)a << 32 | b;  // bits [32w, 32w + 64) of the stream, first bit on top
	const uint32_t v = width ? (uint32_t)((two << (bit & 31u)) >> (64 - width)) : 0u;
	return (int64_t)(int32_t)v;  // the Java field is an int
}

__global__ void __launch_bounds__(TPB) k_labels_fixed(GraphDev lab, int64_t from, int64_t count, const int64_t *__restrict__ rp, int width,
                                                      int64_t *__restrict__ out, uint64_t capacity) {
	__shared__ uint64_t s_base;
	if (threadIdx.x == 0) s_base = ef_get(lab, from);
	__syncthreads();
	const uint64_t arcs = (uint64_t)(rp[count] - rp[0]);
	const uint64_t j = ((uint64_t)blockIdx.x * TPB + threadIdx.x) * 2;  // two labels per thread, one 16-byte store
	if (j >= arcs) return;
	if (arcs > capacity) { if (j == 0) raise(lab, ERR_CAPACITY); return; }
	const uint32_t *words = (const uint32_t *)lab.words;
	const uint64_t nw = lab.nwords * 2, bit = s_base + j * (uint64_t)width;
	const int64_t v0 = fixed_label(words, nw, bit, width);
	if (j + 1 < arcs) {
		const int64_t v1 = fixed_label(words, nw, bit + (uint64_t)width, width);
		if ((reinterpret_cast<uintptr_t>(out) & 15u) == 0) *reinterpret_cast<longlong2 *>(out + j) = make_longlong2(v0, v1);
		else { out[j] = v0; out[j + 1] = v1; }
	} else out[j] = v0;
}

__global__ void __launch_bounds__(TPB) k_labels_check(GraphDev lab, int64_t from, int64_t count, const int64_t *__restrict__ rp, int width) {
	const int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
	if (i >= count) return;
	if (ef_get(lab, from + i + 1) - ef_get(lab, from + i) != (uint64_t)(rp[i + 1] - rp[i]) * (uint64_t)width) raise(lab, ERR_DEGREE);
}

// Gamma-coded labels, load-balanced: the labels of a node can only be read one after the other, and outdegrees are heavy-
// tailed, so a warp whose 32 lanes take 32 consecutive nodes runs at the pace of its longest node with most lanes idle.
// As in the successor kernels a block first looks at LT nodes, sorts them by outdegree class in shared memory and then
// decodes 32 nodes of one class per warp in lock step.
static constexpr int LRPT = 4;
static constexpr int LT = TPB * LRPT;

__global__ void __launch_bounds__(TPB) k_labels_gamma(GraphDev lab, int64_t from, int64_t count, const int64_t *__restrict__ rp,
                                                      int64_t *__restrict__ out, uint64_t capacity) {
	__shared__ uint32_t hist[NBUCKETS + 1];
	__shared__ uint16_t order[LT];
	const int64_t base = (int64_t)blockIdx.x * LT;
	const int64_t rp0 = rp[0];
	int bucket[LRPT];
#pragma unroll
	for (int q = 0; q < LRPT; q++) {
		const int64_t i = base + q * TPB + threadIdx.x;
		bucket[q] = -1;
		if (i < count) {
			const int64_t a = rp[i] - rp0, b = rp[i + 1] - rp0;
			if ((uint64_t)b > capacity) raise(lab, ERR_CAPACITY);
			else if (b > a) bucket[q] = work_bucket((uint32_t)(b - a > 0x7fffffff ? 0x7fffffff : b - a));
			else if (ef_get(lab, from + i + 1) != ef_get(lab, from + i)) raise(lab, ERR_DEGREE);  // no arcs: no label bits
		}
	}
	uint32_t total;
	block_bin_sort<LRPT>(bucket, hist, order, &total);
#pragma unroll 1
	for (uint32_t tb = (threadIdx.x & ~31u); tb < total; tb += TPB) {
		const uint32_t t = tb + (threadIdx.x & 31u);
		const bool active = t < total;
		int64_t j = 0, b = 0;
		uint64_t end = 0;
		BitReader r;
		if (active) {
			const int64_t i = base + order[t];
			j = rp[i] - rp0;
			b = rp[i + 1] - rp0;
			end = ef_get(lab, from + i + 1);
			r.init(lab, ef_get(lab, from + i));
		}
		while (__any_sync(0xffffffffu, j < b)) {  // (lanes without a node have j == b)
			if (j < b) {
				out[j] = (int64_t)(int32_t)(uint32_t)r.gamma();  // the Java field is an int
				j++;
			}
		}
		if (active && r.tell() != end) raise(lab, ERR_DEGREE);
	}
}

void launch_labels(const Launch &L, const GraphDev &lab, int kind, int width, int64_t from, int64_t count, const int64_t *rowptr,
                   int64_t *out, uint64_t capacity, uint64_t arcs_bound) {
	if (count <= 0) return;
	if (kind == 0) {
		k_labels_check<<<grid_for(count, TPB), TPB, 0, L.stream>>>(lab, from, count, rowptr, width);
		WGBV_LAUNCHED(L);
		if (arcs_bound) {
			k_labels_fixed<<<grid_for((int64_t)((arcs_bound + 1) / 2), TPB), TPB, 0, L.stream>>>(lab, from, count, rowptr, width, out, capacity);
			WGBV_LAUNCHED(L);
		}
	} else {
		k_labels_gamma<<<grid_for(count, LT), TPB, 0, L.stream>>>(lab, from, count, rowptr, out, capacity);
		WGBV_LAUNCHED(L);
	}
}

// List-valued labels.  FILL = false: the number of elements of every arc's list goes to lens[arc]; FILL = true: lens holds
// the exclusive prefix sums and the elements go to values.  One lane per node (the lists of a node's arcs follow one
// another in the stream and each starts with its own length).
template <bool FILL>
__global__ void __launch_bounds__(TPB) k_label_lists(GraphDev lab, int64_t from, int64_t count, const int64_t *__restrict__ rp, int width, int is_long,
                                                     int64_t *lens, int64_t *__restrict__ values, uint64_t capacity) {
	const int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
	if (i >= count) return;
	const int64_t a = rp[i] - rp[0], b = rp[i + 1] - rp[0];
	const uint64_t pos = ef_get(lab, from + i), end = ef_get(lab, from + i + 1);
	BitReader r;
	r.init(lab, pos);
	for (int64_t j = a; j < b; j++) {
		const uint64_t n = r.gamma();
		if (n > 0x7fffffffull || r.tell() + n * (uint64_t)width > end) { raise(lab, ERR_DEGREE); return; }
		if (!FILL) {
			lens[j] = (int64_t)n;
			r.seek(r.tell() + n * (uint64_t)width);
		} else {
			const int64_t o = lens[j];
			if ((uint64_t)o + n > capacity) { raise(lab, ERR_CAPACITY); return; }
			for (uint64_t t = 0; t < n; t++) {
				const uint64_t v = width ? r.bits((unsigned)width) : 0;
				values[o + (int64_t)t] = is_long ? (int64_t)v : (int64_t)(int32_t)(uint32_t)v;  // (the Java fields are long[] / int[])
			}
		}
	}
	if (r.tell() != end) raise(lab, ERR_DEGREE);
}

void launch_label_lists(const Launch &L, const GraphDev &lab, bool fill, int width, int is_long, int64_t from, int64_t count, const int64_t *rowptr,
                        int64_t *lens, int64_t *values, uint64_t capacity) {
	if (count <= 0) return;
	if (fill) k_label_lists<true><<<grid_for(count, TPB), TPB, 0, L.stream>>>(lab, from, count, rowptr, width, is_long, lens, values, capacity);
	else k_label_lists<false><<<grid_for(count, TPB), TPB, 0, L.stream>>>(lab, from, count, rowptr, width, is_long, lens, values, capacity);
	WGBV_LAUNCHED(L);
}

}  // namespace wgbv
