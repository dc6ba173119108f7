// wgbv.hpp -- header-only C++ host mirror of webgraph-big's ImmutableGraph API over the C ABI of libwgbv.
//
// Same method names, argument meaning and error behaviour as the reference's Java classes
// (ImmutableGraph.java:201-447, NodeIterator.java:34-133, LazyLongIterator.java:28-43, BVGraph.java:821-1265),
// with the reference's exceptions as C++ exception types.  Link with -lwgbv; no CUDA headers are needed.
#pragma once
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../../include/wgbv.h"

namespace wgbv {

struct IllegalArgumentException : std::invalid_argument { using std::invalid_argument::invalid_argument; };
struct IllegalStateException : std::logic_error { using std::logic_error::logic_error; };
struct IOException : std::runtime_error { using std::runtime_error::runtime_error; };
struct UnsupportedOperationException : std::logic_error { using std::logic_error::logic_error; };
struct NoSuchElementException : std::out_of_range { using std::out_of_range::out_of_range; };
struct CudaError : std::runtime_error { using std::runtime_error::runtime_error; };

inline void check(int rc) {
	if (rc == WGBV_OK) return;
	const std::string msg = wgbv_last_error();
	switch (rc) {
	case WGBV_EINDEX:
	case WGBV_EARG: throw IllegalArgumentException(msg);
	case WGBV_ESTATE: throw IllegalStateException(msg);
	case WGBV_EIO: throw IOException(msg);
	case WGBV_EUNSUPPORTED: throw UnsupportedOperationException(msg);
	default: throw CudaError(msg);
	}
}

// LazyLongIterator.java:28-43: nextLong() returns -1 when the list is exhausted.
class LazyLongIterator {
	const int64_t *p_, *end_;
	std::shared_ptr<const std::vector<int64_t>> keep_;

public:
	LazyLongIterator(const int64_t *b, const int64_t *e, std::shared_ptr<const std::vector<int64_t>> keep = nullptr)
	    : p_(b), end_(e), keep_(std::move(keep)) {}
	int64_t nextLong() { return p_ == end_ ? -1 : *p_++; }
	int64_t skip(int64_t n) {
		const int64_t k = n < end_ - p_ ? n : end_ - p_;
		p_ += k;
		return k;
	}
};

class Graph;

// BVGraphNodeIterator (BVGraph.java:1100-1245) served from chunked bulk decodes on the GPU.
class NodeIterator {
	friend class Graph;
	std::shared_ptr<Graph> g_;
	int64_t from_, curr_, limit_, c0_ = 0, c1_ = 0;
	std::vector<int64_t> rowptr_, col_;
	NodeIterator(std::shared_ptr<Graph> g, int64_t from, int64_t upper);

public:
	static constexpr int64_t CHUNK_NODES = 1 << 18;
	bool hasNext() const { return curr_ < limit_; }
	int64_t nextLong();
	int64_t outdegree() const {
		if (curr_ == from_ - 1) throw IllegalStateException("nextLong() has not been called");  // BVGraph.java:1207
		return rowptr_[curr_ - c0_ + 1] - rowptr_[curr_ - c0_];
	}
	// valid until the next nextLong(), like the reference's window slot (BVGraph.java:1171-1172)
	std::pair<const int64_t *, int64_t> successorArray() const {
		const int64_t d = outdegree();
		return {col_.data() + rowptr_[curr_ - c0_], d};
	}
	LazyLongIterator successors() const {
		auto a = successorArray();
		return LazyLongIterator(a.first, a.first + a.second);
	}
	int64_t skip(int64_t n) {
		if (n < 0) throw IllegalArgumentException("Argument must be nonnegative");
		int64_t k = 0;
		while (k < n && hasNext()) { nextLong(); k++; }
		return k;
	}
	NodeIterator copy(int64_t upperBound) const;  // BVGraph.java:1223-1229
};

// A BVGraph decoded on the GPU: the sibling of it.unimi.dsi.big.webgraph.BVGraph.
class Graph : public std::enable_shared_from_this<Graph> {
	struct Shared {
		wgbv_graph *h = nullptr;
		std::string basename;
		~Shared() { if (h) wgbv_close(h); }
	};
	std::shared_ptr<Shared> sh_;
	wgbv_stream *st_ = nullptr;
	explicit Graph(std::shared_ptr<Shared> sh) : sh_(std::move(sh)) { check(wgbv_stream_create(sh_->h, nullptr, &st_)); }

public:
	Graph(const Graph &) = delete;
	Graph &operator=(const Graph &) = delete;
	~Graph() { if (st_) wgbv_stream_destroy(st_); }

	// BVGraph.load(basename), BVGraph.java:1345-1383
	static std::shared_ptr<Graph> load(const std::string &basename, int device = 0, int shard_index = 0, int shard_count = 1) {
		auto sh = std::make_shared<Shared>();
		sh->basename = basename;
		check(wgbv_open(basename.c_str(), device, shard_index, shard_count, &sh->h));
		return std::shared_ptr<Graph>(new Graph(sh));
	}
	int64_t numNodes() const { return wgbv_num_nodes(sh_->h); }
	int64_t numArcs() const { return wgbv_num_arcs(sh_->h); }
	bool randomAccess() const { return true; }
	bool hasCopiableIterators() const { return true; }
	const std::string &basename() const { return sh_->basename; }
	// flyweight copy: shares the device-resident graph, owns its stream (ImmutableGraph.java:187-197, :447)
	std::shared_ptr<Graph> copy() const { return std::shared_ptr<Graph>(new Graph(sh_)); }

	int64_t outdegree(int64_t x) {  // BVGraph.java:821-842
		int64_t d = 0;
		check(wgbv_outdegrees(st_, &x, 1, &d));
		return d;
	}
	// batched random access: (rowptr[k+1], col)
	std::pair<std::vector<int64_t>, std::vector<int64_t>> successorsBatch(const std::vector<int64_t> &ids) {
		std::vector<int64_t> rp(ids.size() + 1), col;
		int64_t arcs = 0;
		check(wgbv_successors_batch_host(st_, ids.data(), ids.size(), rp.data(), nullptr, 0, &arcs));
		col.resize((size_t)arcs);
		check(wgbv_successors_batch_host(st_, ids.data(), ids.size(), rp.data(), col.data(), col.size(), &arcs));
		return {std::move(rp), std::move(col)};
	}
	std::vector<int64_t> successorArray(int64_t x) { return successorsBatch({x}).second; }  // ImmutableGraph.java:329-333
	LazyLongIterator successors(int64_t x) {                                                 // BVGraph.java:860-867
		auto v = std::make_shared<const std::vector<int64_t>>(successorArray(x));
		return LazyLongIterator(v->data(), v->data() + v->size(), v);
	}
	// the bulk call: CSR of [from, to) in host vectors
	std::pair<std::vector<int64_t>, std::vector<int64_t>> decodeToCSR(int64_t from, int64_t to) {
		std::vector<int64_t> rp((size_t)(to > from ? to - from : 0) + 1), col;
		int64_t arcs = 0;
		check(wgbv_decode_csr_host(st_, from, to, rp.data(), nullptr, 0, &arcs));
		col.resize((size_t)arcs);
		check(wgbv_decode_csr_host(st_, from, to, rp.data(), col.data(), col.size(), &arcs));
		return {std::move(rp), std::move(col)};
	}
	NodeIterator nodeIterator(int64_t from = 0) {  // BVGraph.java:1257
		if (from < 0 || from > numNodes()) throw IllegalArgumentException("Node index out of range: " + std::to_string(from));
		return NodeIterator(shared_from_this(), from, numNodes());
	}
	// ImmutableGraph.java:405-436: contiguous ranges of ceil(n / howMany) nodes
	std::vector<NodeIterator> splitNodeIterators(int howMany) {
		if (howMany < 1) throw IllegalArgumentException("howMany < 1");
		const int64_t n = numNodes(), m = (n + howMany - 1) / howMany;
		std::vector<NodeIterator> res;
		for (int64_t f = 0; f < n; f += m) res.push_back(NodeIterator(copy(), f, f + m < n ? f + m : n));
		while ((int)res.size() < howMany) res.push_back(NodeIterator(shared_from_this(), n, n));
		return res;
	}
	// Transform.transpose on a CSR that already lives in device memory (Transform.java:1098-1144); all four are device pointers
	void transposeDevice(int64_t n, const int64_t *d_rowptr, const int64_t *d_col, int64_t *d_t_rowptr, int64_t *d_t_col) {
		check(wgbv_transpose_csr(st_, n, d_rowptr, d_col, d_t_rowptr, d_t_col));
	}
	// Transform.symmetrizeOffline = union(g, transpose(g)) (Transform.java:573-574) of a device CSR; returns the arcs of the result
	// (d_s_col == nullptr: only the row pointers and the count)
	int64_t symmetrizeDevice(int64_t n, const int64_t *d_rowptr, const int64_t *d_col, int64_t *d_s_rowptr, int64_t *d_s_col, uint64_t capacity) {
		int64_t arcs = 0;
		check(wgbv_symmetrize_csr(st_, n, d_rowptr, d_col, d_s_rowptr, d_s_col, capacity, &arcs));
		return arcs;
	}
	// Transform.union (Transform.java:1659-1663) of two device CSRs over the same nodes
	int64_t unionDevice(int64_t n, const int64_t *a_rowptr, const int64_t *a_col, const int64_t *b_rowptr, const int64_t *b_col, int64_t *d_u_rowptr,
	                    int64_t *d_u_col, uint64_t capacity) {
		int64_t arcs = 0;
		check(wgbv_union_csr(st_, n, a_rowptr, a_col, b_rowptr, b_col, d_u_rowptr, d_u_col, capacity, &arcs));
		return arcs;
	}
	// host-side scans of this copy deliver successors as uint32 (graphs of at most 2^32 nodes): half the PCIe bytes
	void setNarrowScan(bool on) { check(wgbv_scan_set_narrow(st_, on ? 1 : 0)); }
	wgbv_stream *stream() const { return st_; }
};

// BVGraph.store (BVGraph.java:2329-2590) for a CSR in device memory: the bytes of the .graph file and the record lengths
// (whose gamma-coded differences are the .offsets file); split_nodes = 0: one run, as a single-threaded store.
inline std::vector<uint8_t> storeDevice(int device, const wgbv_props &props, int64_t first_node, int64_t n, const int64_t *d_rowptr, const int64_t *d_col,
                                        int64_t split_nodes, std::vector<uint64_t> *node_bits = nullptr) {
	std::vector<uint8_t> out((size_t)(props.arcs > 0 ? props.arcs : 0) * 10 + (size_t)n * 4 + 64);
	if (node_bits) node_bits->assign((size_t)(n > 0 ? n : 1), 0);
	uint64_t bits = 0;
	int rc = wgbv_encode_csr(device, &props, first_node, n, d_rowptr, d_col, split_nodes, out.data(), out.size(), &bits, node_bits ? node_bits->data() : nullptr);
	if (rc == WGBV_EARG && bits / 8 + 1 > out.size()) {  // (the first guess was too small)
		out.assign(bits / 8 + 64, 0);
		rc = wgbv_encode_csr(device, &props, first_node, n, d_rowptr, d_col, split_nodes, out.data(), out.size(), &bits, node_bits ? node_bits->data() : nullptr);
	}
	check(rc);
	out.resize((bits + 7) / 8);
	return out;
}

// BVGraph.writeOffsets (BVG:2595-2609): the .offsets stream of a graph shipped without one (host, one sequential pass).
inline std::vector<uint8_t> writeOffsets(const wgbv_props &props, const std::vector<uint8_t> &graph) {
	uint64_t size = 0;
	check(wgbv_write_offsets(&props, graph.data(), graph.size(), nullptr, 0, &size));
	std::vector<uint8_t> out((size_t)size);
	check(wgbv_write_offsets(&props, graph.data(), graph.size(), out.data(), out.size(), &size));
	return out;
}

// The labels of a BitStreamArcLabelledImmutableGraph (labelling/BitStreamArcLabelledImmutableGraph.java:65-100) on the
// device: decode() returns them parallel to the successors that Graph::decodeToCSR returns for the same range.
class ArcLabels {
	wgbv_labels *h_ = nullptr;

public:
	ArcLabels(int device, int64_t nodes, int kind, int width, const std::vector<uint8_t> &labels, const std::vector<uint8_t> &labeloffsets) {
		check(wgbv_labels_open_mem(device, nodes, kind, width, labels.data(), labels.size(), labeloffsets.data(), labeloffsets.size(), &h_));
	}
	ArcLabels(const ArcLabels &) = delete;
	ArcLabels &operator=(const ArcLabels &) = delete;
	~ArcLabels() { wgbv_labels_close(h_); }
	std::vector<int64_t> decode(Graph &g, int64_t from, int64_t to, const std::vector<int64_t> &rowptr) {
		if ((int64_t)rowptr.size() != to - from + 1) throw IllegalArgumentException("rowptr must have to - from + 1 entries");
		std::vector<int64_t> out((size_t)(rowptr.back() - rowptr.front()));
		check(wgbv_labels_decode_host(g.stream(), h_, from, to, rowptr.data(), out.data(), out.size()));
		return out;
	}
	// list-valued labels (FixedWidthIntListLabel / FixedWidthLongListLabel): arc j's list is values[list_ptr[j] .. list_ptr[j + 1])
	void decodeLists(Graph &g, int64_t from, int64_t to, const std::vector<int64_t> &rowptr, std::vector<int64_t> &list_ptr, std::vector<int64_t> &values) {
		if ((int64_t)rowptr.size() != to - from + 1) throw IllegalArgumentException("rowptr must have to - from + 1 entries");
		list_ptr.resize((size_t)(rowptr.back() - rowptr.front()) + 1);
		int64_t total = 0;
		check(wgbv_labels_decode_lists_host(g.stream(), h_, from, to, rowptr.data(), list_ptr.data(), nullptr, 0, &total));
		values.resize((size_t)total);
		check(wgbv_labels_decode_lists_host(g.stream(), h_, from, to, rowptr.data(), list_ptr.data(), values.data(), values.size(), &total));
	}
};

inline NodeIterator::NodeIterator(std::shared_ptr<Graph> g, int64_t from, int64_t upper)
    : g_(std::move(g)), from_(from), curr_(from - 1), limit_((upper < g_->numNodes() ? upper : g_->numNodes()) - 1), c0_(from), c1_(from) {}

inline int64_t NodeIterator::nextLong() {
	if (!hasNext()) throw NoSuchElementException("no more nodes");  // BVGraph.java:1165
	if (++curr_ >= c1_) {
		c0_ = curr_;
		c1_ = c0_ + CHUNK_NODES < limit_ + 1 ? c0_ + CHUNK_NODES : limit_ + 1;
		auto csr = g_->decodeToCSR(c0_, c1_);
		rowptr_ = std::move(csr.first);
		col_ = std::move(csr.second);
	}
	return curr_;
}

inline NodeIterator NodeIterator::copy(int64_t upperBound) const { return NodeIterator(g_->copy(), curr_ + 1, upperBound); }

}  // namespace wgbv
