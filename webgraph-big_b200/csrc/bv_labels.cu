// bv_labels.cu -- arc labels stored as a bit stream, decoded into an array parallel to the CSR successors.
//
// The reference keeps the labels of a BitStreamArcLabelledImmutableGraph in a `.labels` bit stream -- the labels of the
// arcs of node 0 in successor order, then those of node 1, ... -- and the bit position of every node's first label in
// `.labeloffsets`, gamma-coded differences exactly like `.offsets` (labelling/BitStreamArcLabelledImmutableGraph.java:
// 65-100, 339-363, 654-681).  Its node iterator positions the stream at the node's offset and calls
// label.fromBitStream() once per successor (:487-560).  Here the stream is held like the graph's (32-bit MSB-first
// words + Elias-Fano offsets, both built by the loader the graph uses) and one lane reads one node's labels:
//   kind WGBV_LABEL_FIXED_INT  FixedWidthIntLabel: readInt(width)   (FixedWidthIntLabel.java:70-73)
//   kind WGBV_LABEL_GAMMA_INT  GammaCodedIntLabel: readGamma()      (GammaCodedIntLabel.java:60-64)
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

void launch_labels(const Launch &L, const GraphDev &lab, int kind, int width, int64_t from, int64_t count, const int64_t *rowptr,
                   int64_t *out, uint64_t capacity) {
	if (count <= 0) return;
	if (kind == 0) k_labels<0><<<grid_for(count, TPB), TPB, 0, L.stream>>>(lab, from, count, rowptr, width, out, capacity);
	else k_labels<1><<<grid_for(count, TPB), TPB, 0, L.stream>>>(lab, from, count, rowptr, width, out, capacity);
	WGBV_LAUNCHED(L);
}

}  // namespace wgbv
