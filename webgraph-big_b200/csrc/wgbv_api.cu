// wgbv_api.cu -- the C ABI of libwgbv (include/wgbv.h): handles, device memory, and the orchestration of the
// decode pipeline.  Host code only; the kernels live in bv_kernels.cu.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <new>
#include <string>
#include <vector>

#include "bv_kernels.cuh"
#include "bv_loader.h"

using namespace wgbv;

#define CU(call)                                                                                                  \
	do {                                                                                                          \
		const cudaError_t e_ = (call);                                                                            \
		if (e_ != cudaSuccess) return fail(WGBV_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
	} while (0)

// ------------------------------------------------------------------------------------------------ handles
struct wgbv_graph {
	wgbv_props props{};
	int device = 0;
	int def_codes = 2;  // 0: codings resolved at run time, 1: default codings, 2: default codings with zeta_k = 3
	int64_t shard_lo = 0, shard_hi = 0;  // nodes this handle decodes
	int64_t avail_lo = 0;                // first node whose record is held (shard_lo minus the halo)
	int64_t halo_nodes = 0;              // window_size * max_ref_count, or -1 if unbounded
	GraphDev dev{};
	uint64_t *d_words = nullptr, *d_lower = nullptr, *d_upper = nullptr, *d_hints = nullptr, *d_sparse = nullptr;
	uint64_t device_bytes = 0;
	uint64_t byte0 = 0, byte1 = 0, data_words = 0;  // the byte range of the caller's buffer that is resident
	EliasFano ef;  // host copy, for accounting
	int streams = 0;
	int sm_count = 148;
	int fused_rows = 0;  // rows per tile of the one-pass kernel for this graph (0: not used)
};

// A grow-only device buffer.
struct DevBuf {
	void *p = nullptr;
	size_t cap = 0;
	int ensure(size_t bytes) {
		if (bytes <= cap) return WGBV_OK;
		if (p) cudaFree(p);
		p = nullptr;
		cap = 0;
		size_t want = bytes + bytes / 8 + 256;
		cudaError_t e = cudaMalloc(&p, want);
		if (e != cudaSuccess) {
			want = bytes;
			e = cudaMalloc(&p, want);
		}
		if (e != cudaSuccess) { p = nullptr; return fail(WGBV_ECUDA, "cudaMalloc of %zu bytes failed: %s", want, cudaGetErrorString(e)); }
		cap = want;
		return WGBV_OK;
	}
	void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
	template <typename T> T *as() const { return (T *)p; }
};

struct wgbv_stream {
	wgbv_graph *g = nullptr;
	cudaStream_t stream = nullptr;
	bool own_stream = false;
	GraphDev dev{};            // the graph's view with this stream's error word
	uint32_t *d_err = nullptr;  // [0] sticky error bits, [1..3] spare
	uint32_t *d_stats = nullptr;  // [0] unresolved, [1] max depth, [2..3] spare
	uint64_t *h_pin = nullptr;  // pinned, mapped host scratch for read-backs (32 words)
	uint64_t *d_pin = nullptr;  // the device's address of h_pin: kernels write scalars for the host there (launch_peek)
	uint64_t launches = 0;
	DevBuf deg, ref, depth, rp, scan_state, ticket, halo, levels, list, list2, col, rowptr, ids, batch;
	DevBuf t_pairs[2], t_indeg, t_counts, t_pos;  // wgbv_transpose_csr
	DevBuf sym_rowptr, sym_col;                   // wgbv_symmetrize_csr: the transpose
	DevBuf heavy_rows, heavy_sorted, heavy_cls, heavy_counters, giants, gchunks, level_pos, row_info, tile_done;
	bool front_in_extras = true;  // false: always run header and scan as their own launches (WGBV_SPLIT_FRONT=1)
	bool use_fused = false;       // true: bulk decodes take the one-pass tile kernel where it applies (WGBV_FUSED=1)
	// The decodes of the fused kernel whose outcome has not been read yet: if one of them reported ERR_FUSED_FALLBACK
	// (something only the general path handles), the call that collects the errors reruns them on the general path
	// (the error word does not say which one it was; rerunning a good one gives the same result).
	struct FusedCall {
		int64_t from, to;
		int64_t *d_rowptr, *d_col;
		uint64_t col_capacity;
	};
	std::vector<FusedCall> fused_pending;
	uint64_t fused_fallbacks = 0;  // decodes that were rerun
	uint32_t last_slow_tiles = 0, last_tiles = 0;
	DevBuf c_node, c_dst, c_parent, c_deg, c_isq, c_level, c_len, c_arcs, c_row0, c_soff, c_store, c_info;
	cudaEvent_t ev[6]{};
	bool ev_valid = false;
	Launch launch() { return Launch{stream, &launches}; }

	// streaming scan (wgbv_scan_*): a ring of chunk buffers, decoded on `stream`, copied back on `copy_stream`
	struct ScanSlot {
		DevBuf d_rowptr, d_col;  // with narrow delivery d_col holds the uint32 copy (the decode goes to the stream's col)
		int64_t *h_rowptr = nullptr, *h_col = nullptr;
		size_t h_rowptr_cap = 0, h_col_cap = 0;  // elements
		cudaEvent_t decoded = nullptr, ready = nullptr;
		int64_t first = 0, last = 0, arcs = 0;
		uint32_t *h_err = nullptr;  // pinned copy of the error word taken after this chunk
	};
	static constexpr int SCAN_SLOTS = 3;
	ScanSlot slot[SCAN_SLOTS];
	cudaStream_t copy_stream = nullptr, upload_stream = nullptr;
	cudaEvent_t uploaded[2]{};          // the words a chunk needs are on the device (alternating)
	const uint8_t *up_graph = nullptr;  // host copy of the bit stream to upload while scanning (wgbv_scan_begin_upload)
	uint64_t up_next_word = 0;          // first 64-bit unit of the device buffer not uploaded yet
	int64_t up_chunks = 0;              // chunks whose words have been queued
	bool scanning = false;
	bool narrow = false;                // deliver successors as uint32 (wgbv_scan_set_narrow)
	int64_t scan_to = 0, scan_chunk = 0, scan_issue = 0, scan_take = 0;  // next node to issue / to hand out
	int64_t issued = 0, taken = 0;                                       // chunks issued / handed out
};

// ------------------------------------------------------------------------------------------------ misc
extern "C" const char *wgbv_last_error(void) { return last_error_message(); }

extern "C" int wgbv_parse_properties(const char *text, wgbv_props *out) {
	clear_error();
	return parse_properties_text(text, out);
}

static int check_codings(const wgbv_props &p) {
	auto gd = [](int c) { return c == WGBV_GAMMA || c == WGBV_DELTA; };
	auto gdu = [](int c) { return c == WGBV_GAMMA || c == WGBV_DELTA || c == WGBV_UNARY; };
	if (!gd(p.outdegree_coding)) return fail(WGBV_EUNSUPPORTED, "The required outdegree coding (%d) is not supported.", p.outdegree_coding);
	if (!gdu(p.reference_coding)) return fail(WGBV_EUNSUPPORTED, "The required reference coding (%d) is not supported.", p.reference_coding);
	if (!gdu(p.block_count_coding)) return fail(WGBV_EUNSUPPORTED, "The required block count coding (%d) is not supported.", p.block_count_coding);
	if (!gdu(p.block_coding)) return fail(WGBV_EUNSUPPORTED, "The required block coding (%d) is not supported.", p.block_coding);
	const int r = p.residual_coding;
	if (r != WGBV_GAMMA && r != WGBV_ZETA && r != WGBV_DELTA && r != WGBV_GOLOMB && r != WGBV_NIBBLE)
		return fail(WGBV_EUNSUPPORTED, "The required residuals coding (%d) is not supported.", r);
	if (!gd(p.offset_coding)) return fail(WGBV_EUNSUPPORTED, "The required offset coding (%d) is not supported.", p.offset_coding);
	if (p.window_size < 0 || p.window_size > 255) return fail(WGBV_EUNSUPPORTED, "window sizes above 255 are not supported");
	if (p.zeta_k < 1 || p.zeta_k > 32) return fail(WGBV_EUNSUPPORTED, "zeta_k out of range");
	return WGBV_OK;
}

static int select_device(int device) {
	int n = 0;
	const cudaError_t e = cudaGetDeviceCount(&n);
	if (e != cudaSuccess || n == 0) return fail(WGBV_ECUDA, "no usable CUDA device (%s); libwgbv has no CPU fallback", e != cudaSuccess ? cudaGetErrorString(e) : "0 devices");
	if (device < 0 || device >= n) return fail(WGBV_ECUDA, "device %d out of range (%d devices)", device, n);
	CU(cudaSetDevice(device));
	return WGBV_OK;
}

static int upload(uint64_t **dst, const std::vector<uint64_t> &v, uint64_t *total) {
	const size_t bytes = std::max<size_t>(v.size(), 1) * 8;
	CU(cudaMalloc((void **)dst, bytes));
	if (!v.empty()) CU(cudaMemcpy(*dst, v.data(), v.size() * 8, cudaMemcpyHostToDevice));
	*total += bytes;
	return WGBV_OK;
}

extern "C" void wgbv_close(wgbv_graph *g) {
	if (!g) return;
	cudaSetDevice(g->device);
	cudaFree(g->d_words);
	cudaFree(g->d_lower);
	cudaFree(g->d_upper);
	cudaFree(g->d_hints);
	cudaFree(g->d_sparse);
	delete g;
}

// Builds the device-resident graph for nodes [avail_lo, shard_hi) from `off`, the absolute bit offsets (inside
// `graph`) of nodes [off_first, off_first + off.size() - 1].
// WGBV_TIMING=1: the phases of an open, on stderr
struct LoadTimer {
	bool on;
	std::chrono::steady_clock::time_point t;
	LoadTimer() : on(getenv("WGBV_TIMING") != nullptr), t(std::chrono::steady_clock::now()) {}
	void lap(const char *what) {
		if (!on) return;
		const auto now = std::chrono::steady_clock::now();
		fprintf(stderr, "[wgbv] %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(now - t).count());
		t = now;
	}
};

// The offsets come either as a host vector (`off`) or, decoded on the device, as `d_off` (then `off` is empty).
static int build_graph(const wgbv_props *props, const uint8_t *graph, uint64_t graph_bytes, const std::vector<uint64_t> &off,
                       const uint64_t *d_off, int64_t off_first, int64_t avail_lo, int64_t shard_lo, int64_t shard_hi, int64_t halo_nodes,
                       int device, wgbv_graph **out) {
	wgbv_graph *g = new (std::nothrow) wgbv_graph();
	if (!g) return fail(WGBV_EARG, "out of host memory");
	g->props = *props;
	g->device = device;
	const bool defaults = props->outdegree_coding == WGBV_GAMMA && props->block_coding == WGBV_GAMMA && props->residual_coding == WGBV_ZETA &&
	                      props->reference_coding == WGBV_UNARY && props->block_count_coding == WGBV_GAMMA;
	g->def_codes = defaults ? (props->zeta_k == 3 ? 2 : 1) : 0;
	g->halo_nodes = halo_nodes;
	cudaDeviceGetAttribute(&g->sm_count, cudaDevAttrMultiProcessorCount, device);
	g->shard_lo = shard_lo;
	g->shard_hi = shard_hi;
	g->avail_lo = avail_lo;
	int rc;

	// the slice of the bit stream that holds [avail_lo, shard_hi), starting on a 64-bit word of the file
	uint64_t first_off, last_off;
	if (d_off) {
		if (cudaMemcpy(&first_off, d_off + (avail_lo - off_first), 8, cudaMemcpyDeviceToHost) != cudaSuccess ||
		    cudaMemcpy(&last_off, d_off + (shard_hi - off_first), 8, cudaMemcpyDeviceToHost) != cudaSuccess) {
			delete g;
			return fail(WGBV_ECUDA, "reading the decoded offsets failed: %s", cudaGetErrorString(cudaGetLastError()));
		}
	} else {
		first_off = off[(size_t)(avail_lo - off_first)];
		last_off = off[(size_t)(shard_hi - off_first)];
	}
	const uint64_t bit0 = first_off & ~63ull;
	const uint64_t bit1 = last_off;
	const uint64_t byte0 = bit0 >> 3;
	const uint64_t byte1 = std::min<uint64_t>((bit1 + 7) >> 3, graph_bytes);
	const uint64_t data_words = (byte1 > byte0 ? byte1 - byte0 + 7 : 0) / 8;
	const uint64_t nwords = data_words + 2;
	g->byte0 = byte0;
	g->byte1 = byte1;
	g->data_words = data_words;
	auto cleanup = [&](int code) { wgbv_close(g); return code; };
	LoadTimer lt;
	if (cudaMalloc((void **)&g->d_words, nwords * 8) != cudaSuccess) return cleanup(fail(WGBV_ECUDA, "cudaMalloc of the bit stream (%llu bytes) failed", (unsigned long long)nwords * 8));
	g->device_bytes += nwords * 8;
	if (cudaMemset(g->d_words, 0, nwords * 8) != cudaSuccess) return cleanup(fail(WGBV_ECUDA, "cudaMemset failed"));
	if (byte1 > byte0 && cudaMemcpy(g->d_words, graph + byte0, byte1 - byte0, cudaMemcpyHostToDevice) != cudaSuccess)
		return cleanup(fail(WGBV_ECUDA, "upload of the bit stream failed"));
	Launch L{nullptr, nullptr};
	launch_bswap(L, g->d_words, data_words);
	lt.lap("upload of the bit stream");

	// Elias-Fano list of the offsets of [avail_lo, shard_hi] relative to bit0
	if (d_off) {
		if ((rc = device_ef_build(d_off + (avail_lo - off_first), (uint64_t)(shard_hi - avail_lo + 1), bit0, last_off, g->ef, &g->d_lower,
		                          &g->d_upper, &g->d_hints, &g->d_sparse, &g->device_bytes)))
			return cleanup(rc);
		lt.lap("Elias-Fano on the device");
	} else {
		g->ef.build(off.data() + (avail_lo - off_first), (uint64_t)(shard_hi - avail_lo + 1), bit0);
		lt.lap("Elias-Fano construction");
		if ((rc = upload(&g->d_lower, g->ef.lower, &g->device_bytes))) return cleanup(rc);
		if ((rc = upload(&g->d_upper, g->ef.upper, &g->device_bytes))) return cleanup(rc);
		if ((rc = upload(&g->d_hints, g->ef.hints, &g->device_bytes))) return cleanup(rc);
		if ((rc = upload(&g->d_sparse, g->ef.sparse, &g->device_bytes))) return cleanup(rc);
	}
	if (cudaDeviceSynchronize() != cudaSuccess) return cleanup(fail(WGBV_ECUDA, "device error while loading: %s", cudaGetErrorString(cudaGetLastError())));
	lt.lap("upload of the offsets list");

	GraphDev &d = g->dev;
	d.words = g->d_words;
	d.nwords = nwords;
	d.bit_base = bit0;
	d.bit_end = bit1;
	d.ef_lower = g->d_lower;
	d.ef_upper = g->d_upper;
	d.ef_hints = g->d_hints;
	d.ef_sparse = g->d_sparse;
	d.ef_first = avail_lo;
	d.ef_count = shard_hi - avail_lo + 1;
	d.ef_l = g->ef.l;
	d.nodes = props->nodes;
	d.window = props->window_size;
	d.min_interval = props->min_interval_length;
	d.zeta_k = props->zeta_k;
	d.outdegree_coding = props->outdegree_coding;
	d.block_coding = props->block_coding;
	d.residual_coding = props->residual_coding;
	d.reference_coding = props->reference_coding;
	d.block_count_coding = props->block_count_coding;
	d.err = nullptr;
	{
		// the one-pass tile kernel: bounded chains, node ids that fit 32 bits, and a density its tiles can hold
		const int64_t held = shard_hi - avail_lo;
		const double rec_bits = held > 0 ? (double)(bit1 - bit0) / (double)held : 0.0;
		const double mean_deg = props->nodes > 0 && props->arcs > 0 ? (double)props->arcs / (double)props->nodes : rec_bits / 12.0;
		if (halo_nodes >= 0 && halo_nodes <= TILE_MAX_HALO && props->max_ref_count <= 8 && (uint64_t)props->nodes < 0xffffffffull)
			g->fused_rows = fused_tile_rows((int)halo_nodes, mean_deg, rec_bits);
	}
	*out = g;
	return WGBV_OK;
}

// halo: a reference chain has at most max_ref_count hops of at most window_size nodes each (BVG:701, 2258); -1 = unbounded
static int64_t halo_of(const wgbv_props *props) {
	if (props->window_size == 0 || props->max_ref_count <= 0) return 0;
	if (props->max_ref_count > (1 << 20)) return -1;
	return (int64_t)props->window_size * (int64_t)props->max_ref_count;
}

// .offsets on the device (bv_offsets.cu): gamma-coded streams of at least 2^20 entries; WGBV_DEVICE_OFFSETS=1 forces it
// for any size, =0 keeps the host decoder.
static bool offsets_on_device(const wgbv_props *props, int64_t count) {
	if (props->offset_coding != WGBV_GAMMA || count < 1) return false;
	const char *e = getenv("WGBV_DEVICE_OFFSETS");
	if (e && e[0] == '0') return false;
	if (e && e[0] == '1') return true;
	return count >= (1 << 20);
}

// Decodes the stream on the device and opens [shard_lo, shard_hi) of it (no sharding decision is taken from the offsets).
static int open_with_device_offsets(const wgbv_props *props, const uint8_t *graph, uint64_t graph_bytes, const uint8_t *offsets,
                                    uint64_t offsets_bytes, int64_t off_first, int64_t count, int64_t avail_lo, int64_t shard_lo,
                                    int64_t shard_hi, int device, wgbv_graph **out) {
	uint64_t *d_off = nullptr;
	LoadTimer lt;
	int rc = device_decode_offsets(offsets, offsets_bytes, count, &d_off);
	if (rc) return rc;
	lt.lap(".offsets on the device");
	uint64_t last = 0;
	if (cudaMemcpy(&last, d_off + (count - 1), 8, cudaMemcpyDeviceToHost) != cudaSuccess) {
		cudaFree(d_off);
		return fail(WGBV_ECUDA, "reading the decoded offsets failed: %s", cudaGetErrorString(cudaGetLastError()));
	}
	if (last > graph_bytes * 8) {
		cudaFree(d_off);
		return fail(WGBV_EIO, "the last offset (%llu) lies beyond the end of the bit stream (%llu bits)", (unsigned long long)last, (unsigned long long)graph_bytes * 8);
	}
	static const std::vector<uint64_t> none;
	rc = build_graph(props, graph, graph_bytes, none, d_off, off_first, avail_lo, shard_lo, shard_hi, halo_of(props), device, out);
	cudaFree(d_off);
	return rc;
}

static int check_offsets(const std::vector<uint64_t> &off, uint64_t graph_bytes) {
	if (off.back() > graph_bytes * 8) return fail(WGBV_EIO, "the last offset (%llu) lies beyond the end of the bit stream (%llu bits)", (unsigned long long)off.back(), (unsigned long long)graph_bytes * 8);
	for (size_t i = 0; i + 1 < off.size(); i++)
		if (off[i + 1] < off[i]) return fail(WGBV_EIO, "offsets are not monotone at entry %zu", i);
	return WGBV_OK;
}

extern "C" int wgbv_open_mem(const wgbv_props *props, const uint8_t *graph, uint64_t graph_bytes, const uint8_t *offsets,
                             uint64_t offsets_bytes, int device, int shard_index, int shard_count, wgbv_graph **out) {
	clear_error();
	if (!props || !out || (!graph && graph_bytes)) return fail(WGBV_EARG, "null argument");
	*out = nullptr;
	if (shard_count < 1 || shard_index < 0 || shard_index >= shard_count) return fail(WGBV_EARG, "bad shard %d of %d", shard_index, shard_count);
	int rc = check_codings(*props);
	if (rc) return rc;
	if ((rc = select_device(device))) return rc;
	if (offsets && shard_count == 1 && offsets_on_device(props, props->nodes + 1))
		return open_with_device_offsets(props, graph, graph_bytes, offsets, offsets_bytes, 0, props->nodes + 1, 0, 0, props->nodes, device, out);
	std::vector<uint64_t> off;
	// (no .offsets: the records are found by the sequential pass of BVGraph.writeOffsets, see wgbv_write_offsets)
	if ((rc = offsets ? decode_offsets(*props, offsets, offsets_bytes, props->nodes + 1, off) : build_offsets(*props, graph, graph_bytes, off))) return rc;
	if ((rc = check_offsets(off, graph_bytes))) return rc;
	const int64_t halo = halo_of(props);
	if (halo < 0 && shard_count > 1)
		return fail(WGBV_EUNSUPPORTED, "maxrefcount=%d makes reference chains unbounded: the graph cannot be sharded", props->max_ref_count);
	std::vector<int64_t> bounds;
	plan_shards(off, shard_count, bounds);
	const int64_t lo = bounds[(size_t)shard_index], hi = bounds[(size_t)shard_index + 1];
	const int64_t avail_lo = halo < 0 ? 0 : std::max<int64_t>(0, lo - halo);
	return build_graph(props, graph, graph_bytes, off, nullptr, 0, avail_lo, lo, hi, halo, device, out);
}

extern "C" int wgbv_write_offsets(const wgbv_props *props, const uint8_t *graph, uint64_t graph_bytes, uint8_t *out, uint64_t out_capacity,
                                  uint64_t *out_bytes) {
	clear_error();
	if (!props || (!graph && graph_bytes) || !out_bytes) return fail(WGBV_EARG, "null argument");
	int rc = check_codings(*props);
	if (rc) return rc;
	std::vector<uint64_t> off;
	if ((rc = build_offsets(*props, graph, graph_bytes, off))) return rc;
	std::vector<uint8_t> stream;
	encode_offsets(*props, off, stream);
	*out_bytes = stream.size();
	if (!out) return WGBV_OK;
	if (out_capacity < stream.size()) return fail(WGBV_EARG, "out_capacity (%llu) is smaller than the %llu bytes of the offsets stream", (unsigned long long)out_capacity, (unsigned long long)stream.size());
	if (!stream.empty()) memcpy(out, stream.data(), stream.size());
	return WGBV_OK;
}

extern "C" int wgbv_plan_shards(const wgbv_props *props, const uint8_t *offsets, uint64_t offsets_bytes, int shard_count,
                                int64_t *bounds_out, int64_t *halo_out) {
	clear_error();
	if (!props || !offsets || !bounds_out) return fail(WGBV_EARG, "null argument");
	if (shard_count < 1) return fail(WGBV_EARG, "bad shard count %d", shard_count);
	std::vector<uint64_t> off;
	int rc;
	if ((rc = decode_offsets(*props, offsets, offsets_bytes, props->nodes + 1, off))) return rc;
	const int64_t halo = halo_of(props);
	if (halo < 0 && shard_count > 1)
		return fail(WGBV_EUNSUPPORTED, "maxrefcount=%d makes reference chains unbounded: the graph cannot be sharded", props->max_ref_count);
	std::vector<int64_t> bounds;
	plan_shards(off, shard_count, bounds);
	for (int k = 0; k <= shard_count; k++) bounds_out[k] = bounds[(size_t)k];
	if (halo_out) *halo_out = halo;
	return WGBV_OK;
}

extern "C" int wgbv_open_slice(const wgbv_props *props, int64_t first_node, int64_t node_count, int64_t halo_nodes,
                               const uint8_t *bits, uint64_t nbytes, const uint8_t *offsets, uint64_t offsets_bytes, int device,
                               wgbv_graph **out) {
	clear_error();
	if (!props || !out || (!bits && nbytes) || !offsets) return fail(WGBV_EARG, "null argument");
	*out = nullptr;
	if (first_node < 0 || node_count < 0 || halo_nodes < 0 || halo_nodes > node_count || first_node + node_count > props->nodes)
		return fail(WGBV_EINDEX, "Node index out of range: slice [%lld, %lld) of %lld nodes", (long long)first_node, (long long)(first_node + node_count), (long long)props->nodes);
	int rc = check_codings(*props);
	if (rc) return rc;
	if ((rc = select_device(device))) return rc;
	if (offsets_on_device(props, node_count + 1))
		return open_with_device_offsets(props, bits, nbytes, offsets, offsets_bytes, first_node, node_count + 1, first_node,
		                                first_node + halo_nodes, first_node + node_count, device, out);
	std::vector<uint64_t> off;
	LoadTimer lt;
	if ((rc = decode_offsets(*props, offsets, offsets_bytes, node_count + 1, off))) return rc;
	lt.lap("decoding of .offsets");
	if ((rc = check_offsets(off, nbytes))) return rc;
	return build_graph(props, bits, nbytes, off, nullptr, first_node, first_node, first_node + halo_nodes, first_node + node_count, halo_of(props), device, out);
}

extern "C" int wgbv_upload_stream(wgbv_stream *s, const uint8_t *graph, uint64_t graph_bytes) {
	clear_error();
	if (!s || (!graph && graph_bytes)) return fail(WGBV_EARG, "null argument");
	wgbv_graph *g = s->g;
	CU(cudaSetDevice(g->device));
	if (graph_bytes < g->byte1) return fail(WGBV_EARG, "the buffer (%llu bytes) is shorter than the stream that was opened (%llu bytes)", (unsigned long long)graph_bytes, (unsigned long long)g->byte1);
	// the words are rewritten in place (raw bytes first, byte-swapped afterwards): another stream decoding the same graph
	// meanwhile would read garbage without any error
	if (g->streams > 1) return fail(WGBV_ESTATE, "the bit stream can only be uploaded again while this is the graph's only stream (%d exist)", g->streams);
	if (g->data_words) {
		CU(cudaMemsetAsync(g->d_words + g->data_words - 1, 0, 8, s->stream));
		CU(cudaMemcpyAsync(g->d_words, graph + g->byte0, g->byte1 - g->byte0, cudaMemcpyHostToDevice, s->stream));
		launch_bswap(s->launch(), g->d_words, g->data_words);
	}
	return WGBV_OK;
}

extern "C" int wgbv_open(const char *basename, int device, int shard_index, int shard_count, wgbv_graph **out) {
	clear_error();
	if (!basename || !out) return fail(WGBV_EARG, "null argument");
	std::vector<uint8_t> ptxt, graph, offs;
	int rc;
	const std::string base(basename);
	if ((rc = read_file(base + ".properties", ptxt))) return rc;
	ptxt.push_back(0);
	wgbv_props p;
	if ((rc = parse_properties_text((const char *)ptxt.data(), &p))) return rc;
	if ((rc = read_file(base + ".graph", graph))) return rc;
	if ((rc = read_file(base + ".offsets", offs))) return rc;
	return wgbv_open_mem(&p, graph.data(), graph.size(), offs.data(), offs.size(), device, shard_index, shard_count, out);
}

extern "C" int64_t wgbv_num_nodes(const wgbv_graph *g) { return g ? g->props.nodes : 0; }
extern "C" int64_t wgbv_num_arcs(const wgbv_graph *g) { return g ? g->props.arcs : 0; }
extern "C" int wgbv_get_props(const wgbv_graph *g, wgbv_props *out) {
	if (!g || !out) return fail(WGBV_EARG, "null argument");
	*out = g->props;
	return WGBV_OK;
}
extern "C" int wgbv_shard_range(const wgbv_graph *g, int64_t *first, int64_t *last) {
	if (!g || !first || !last) return fail(WGBV_EARG, "null argument");
	*first = g->shard_lo;
	*last = g->shard_hi;
	return WGBV_OK;
}
extern "C" uint64_t wgbv_device_bytes(const wgbv_graph *g) { return g ? g->device_bytes : 0; }
extern "C" uint64_t wgbv_offsets_bytes(const wgbv_graph *g) { return g ? g->ef.bytes() : 0; }

extern "C" uint64_t wgbv_algorithmic_bytes(const wgbv_graph *g, int64_t from, int64_t to, int64_t arcs) {
	if (!g || from < g->avail_lo || to > g->shard_hi || to < from) return 0;
	const uint64_t bits = g->ef.get((uint64_t)(to - g->avail_lo)) - g->ef.get((uint64_t)(from - g->avail_lo));
	const uint64_t nodes = (uint64_t)(to - from);
	const uint64_t ef_bits = nodes * (2 + (uint64_t)g->ef.l);
	return (bits + 7) / 8 + (ef_bits + 7) / 8 + 8 * (nodes + 1) + 8 * (uint64_t)arcs;
}

// ------------------------------------------------------------------------------------------------ streams
extern "C" int wgbv_stream_create(wgbv_graph *g, void *cuda_stream, wgbv_stream **out) {
	clear_error();
	if (!g || !out) return fail(WGBV_EARG, "null argument");
	*out = nullptr;
	CU(cudaSetDevice(g->device));
	wgbv_stream *s = new (std::nothrow) wgbv_stream();
	if (!s) return fail(WGBV_EARG, "out of host memory");
	s->g = g;
	if (cuda_stream) s->stream = (cudaStream_t)cuda_stream;
	else {
		if (cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking) != cudaSuccess) { delete s; return fail(WGBV_ECUDA, "cudaStreamCreate failed"); }
		s->own_stream = true;
	}
	if (cudaMalloc((void **)&s->d_err, 32) != cudaSuccess || cudaMemset(s->d_err, 0, 32) != cudaSuccess ||
	    cudaHostAlloc((void **)&s->h_pin, 32 * 8, cudaHostAllocMapped) != cudaSuccess ||
	    cudaHostGetDevicePointer((void **)&s->d_pin, s->h_pin, 0) != cudaSuccess) {
		wgbv_stream_destroy(s);
		return fail(WGBV_ECUDA, "allocation of the stream state failed");
	}
	s->d_stats = s->d_err + 4;
	for (auto &e : s->ev) cudaEventCreate(&e);
	s->dev = g->dev;
	s->dev.err = s->d_err;
	// WGBV_FUSED=1 sends bulk decodes of graphs with bounded chains through the one-pass tile kernel (bv_fused.cu).  It
	// moves the algorithmic minimum of bytes and needs no scratch arrays, but as measured on B200 (profiles/r02_notes.md)
	// it is still slower than the level-synchronous kernels (8.9 vs 7.5 ms per 2e7 nodes), so those stay the default.
	const char *fu = getenv("WGBV_FUSED");
	s->use_fused = fu && fu[0] == '1';
	const char *sf = getenv("WGBV_SPLIT_FRONT");
	s->front_in_extras = !(sf && sf[0] == '1');
	g->streams++;
	*out = s;
	return WGBV_OK;
}

extern "C" void wgbv_stream_destroy(wgbv_stream *s) {
	if (!s) return;
	cudaSetDevice(s->g->device);
	if (s->stream) cudaStreamSynchronize(s->stream);
	for (DevBuf *b : {&s->t_pairs[0], &s->t_pairs[1], &s->t_indeg, &s->t_counts, &s->t_pos, &s->sym_rowptr, &s->sym_col}) b->release();
	for (DevBuf *b : {&s->deg, &s->ref, &s->depth, &s->rp, &s->scan_state, &s->ticket, &s->halo, &s->levels, &s->list, &s->list2, &s->col,
	                  &s->rowptr, &s->ids, &s->batch, &s->heavy_rows, &s->heavy_sorted, &s->heavy_cls, &s->heavy_counters, &s->giants, &s->gchunks, &s->tile_done, &s->level_pos, &s->row_info, &s->c_node, &s->c_dst, &s->c_parent, &s->c_deg, &s->c_isq, &s->c_level,
	                  &s->c_len, &s->c_arcs, &s->c_row0, &s->c_soff, &s->c_store, &s->c_info})
		b->release();
	for (auto &sl : s->slot) {
		sl.d_rowptr.release();
		sl.d_col.release();
		if (sl.h_rowptr) cudaFreeHost(sl.h_rowptr);
		if (sl.h_col) cudaFreeHost(sl.h_col);
		if (sl.h_err) cudaFreeHost(sl.h_err);
		if (sl.decoded) cudaEventDestroy(sl.decoded);
		if (sl.ready) cudaEventDestroy(sl.ready);
	}
	if (s->copy_stream) cudaStreamDestroy(s->copy_stream);
	if (s->upload_stream) cudaStreamDestroy(s->upload_stream);
	for (auto &e : s->uploaded) if (e) cudaEventDestroy(e);
	if (s->d_err) cudaFree(s->d_err);
	if (s->h_pin) cudaFreeHost(s->h_pin);
	for (auto &e : s->ev) if (e) cudaEventDestroy(e);
	if (s->own_stream && s->stream) cudaStreamDestroy(s->stream);
	s->g->streams--;
	delete s;
}

static int error_from_bits(wgbv_stream *s, uint32_t e);

static int decode_general(wgbv_stream *s, int64_t from, int64_t to, int64_t *d_rowptr, int64_t *d_col, uint64_t col_capacity, bool trust_maxref);

// Reads and clears the sticky error word; maps it onto a status code.  A decode of the fused kernel that met something
// only the general path handles (ERR_FUSED_FALLBACK: a reference chain longer than the properties promise, a slow
// tile at the start of a range that needs halo rows) is rerun here, on the general path, before anything is reported.
static int collect_errors(wgbv_stream *s) {
	uint32_t *h = (uint32_t *)(s->h_pin + 8);
	for (int attempt = 0;; attempt++) {
		CU(cudaMemcpyAsync(h, s->d_err, 12, cudaMemcpyDeviceToHost, s->stream));
		CU(cudaStreamSynchronize(s->stream));
		const uint32_t e = h[0];
		if (!s->fused_pending.empty()) s->last_slow_tiles = h[2];
		if (!e) { s->fused_pending.clear(); return WGBV_OK; }
		CU(cudaMemsetAsync(s->d_err, 0, 4, s->stream));
		if ((e & (ERR_FUSED_FALLBACK | ERR_FUSED_DEEP)) && !s->fused_pending.empty() && attempt == 0) {
			const std::vector<wgbv_stream::FusedCall> calls = s->fused_pending;
			s->fused_pending.clear();
			for (const auto &c : calls) {
				s->fused_fallbacks++;
				const int rc = decode_general(s, c.from, c.to, c.d_rowptr, c.d_col, c.col_capacity, !(e & ERR_FUSED_DEEP));
				if (rc) return rc;
			}
			continue;
		}
		s->fused_pending.clear();
		return error_from_bits(s, e & ~(ERR_FUSED_FALLBACK | ERR_FUSED_DEEP));
	}
}

static int error_from_bits(wgbv_stream *s, uint32_t e) {
	if (e & ERR_INDEX) return fail(WGBV_EINDEX, "Node index out of range");
	if (e & ERR_REF_WINDOW) return fail(WGBV_ESTATE, "The required reference is incompatible with the window size (%d)", s->g->props.window_size);
	if (e & ERR_OVERRUN) return fail(WGBV_EIO, "a code runs past the end of the bit stream");
	if (e & ERR_REF_UNRESOLVED) return fail(WGBV_ESTATE, "a reference chain is longer than maxrefcount allows and leaves the loaded range");
	if (e & ERR_NOT_SORTED) return fail(WGBV_ESTATE, "malformed successor list: copied successors and extras overlap");
	if (e & ERR_CAPACITY) return fail(WGBV_EARG, "output capacity exceeded");
	return fail(WGBV_ESTATE, "malformed bit stream (inconsistent counts, error bits 0x%x)", e);
}

extern "C" int wgbv_stream_sync(wgbv_stream *s) {
	clear_error();
	if (!s) return fail(WGBV_EARG, "null argument");
	CU(cudaSetDevice(s->g->device));
	CU(cudaStreamSynchronize(s->stream));
	const cudaError_t e = cudaGetLastError();
	if (e != cudaSuccess) return fail(WGBV_ECUDA, "CUDA error: %s", cudaGetErrorString(e));
	return collect_errors(s);
}

extern "C" uint64_t wgbv_kernel_launches(const wgbv_stream *s) { return s ? s->launches : 0; }

// ------------------------------------------------------------------------------------------------ bulk decode
static int check_range(const wgbv_graph *g, int64_t from, int64_t to) {
	if (from < 0 || from > g->props.nodes || to < from || to > g->props.nodes) return fail(WGBV_EINDEX, "Node index out of range: [%lld, %lld)", (long long)from, (long long)to);
	if (from < g->shard_lo || to > g->shard_hi) return fail(WGBV_EINDEX, "Node index out of range for this shard [%lld, %lld): [%lld, %lld)", (long long)g->shard_lo, (long long)g->shard_hi, (long long)from, (long long)to);
	return WGBV_OK;
}

// First node that has to be decoded so that every row of [from, ...) finds its reference chain.
static int64_t halo_start(const wgbv_graph *g, int64_t from) {
	if (g->halo_nodes < 0) return g->avail_lo;
	return std::max<int64_t>(g->avail_lo, from - g->halo_nodes);
}

struct Prepared {
	int64_t lo = 0, cnt = 0, halo_n = 0;
	int64_t halo_arcs = 0, arcs = 0;
	uint32_t max_depth = 0;
	bool have_depth = false;
	const int64_t *rp = nullptr;
	bool no_walk = false;         // true: the depths come from k_depth whatever the properties say
	int64_t *front_rp = nullptr;  // not null: stages 1 and 2 have NOT run; k_extras does them itself and writes the row pointers here
};

// Stages 1, 2 and 2b for [from, to): degrees, row pointers and depth levels.  Leaves the row pointers of
// [lo, to) in s->rp and writes those of [from, to) to d_rowptr.  Synchronises once to learn the arc counts.
// Stage 2b: reference-chain depth of every row of the prepared range (only the level-synchronous path needs it).
static int compute_depth(wgbv_stream *s, int64_t from, Prepared &P) {
	const Launch L = s->launch();
	int rc;
	if ((rc = s->depth.ensure((size_t)std::max<int64_t>(P.cnt, 1) * 4))) return rc;
	uint32_t *hs = (uint32_t *)(s->h_pin + 4);
	for (int pass = 0;; pass++) {
		CU(cudaMemsetAsync(s->d_stats, 0, 8, s->stream));
		launch_depth(L, s->dev, s->ref.as<uint8_t>(), s->depth.as<uint32_t>(), P.lo, from, P.cnt, pass == 0, s->d_stats);
		CU(cudaMemcpyAsync(hs, s->d_stats, 8, cudaMemcpyDeviceToHost, s->stream));
		CU(cudaStreamSynchronize(s->stream));
		if (hs[0] == 0) break;
		if (pass > (1 << 26)) return fail(WGBV_ESTATE, "reference chains do not resolve");
	}
	P.max_depth = hs[1];
	P.have_depth = true;
	return WGBV_OK;
}

static int prepare(wgbv_stream *s, int64_t from, int64_t to, int64_t *d_rowptr, Prepared &P, bool trust_maxref = true) {
	wgbv_graph *g = s->g;
	const Launch L = s->launch();
	// (a file whose chains are longer than its properties say: every row the handle holds may be needed)
	P.lo = trust_maxref ? halo_start(g, from) : g->avail_lo;
	P.no_walk = !trust_maxref;
	P.cnt = to - P.lo;
	P.halo_n = from - P.lo;
	if (P.cnt >= (int64_t)0xfffffff0ll) return fail(WGBV_EARG, "a single decode call is limited to 2^32-16 nodes; split the range");
	int rc;
	const size_t c = (size_t)std::max<int64_t>(P.cnt, 1);
	if ((rc = s->deg.ensure(c * 4)) || (rc = s->ref.ensure(c)) || (rc = s->rp.ensure((c + 1) * 8)) ||
	    (rc = s->scan_state.ensure(((c + 1) / 2048 + 2) * 8)) || (rc = s->ticket.ensure(64)))
		return rc;
	cudaEventRecord(s->ev[0], s->stream);
	launch_header(L, s->dev, g->def_codes, P.lo, P.cnt, s->deg.as<uint32_t>(), s->ref.as<uint8_t>());
	cudaEventRecord(s->ev[1], s->stream);
	int64_t *rp = (P.halo_n == 0 && d_rowptr) ? d_rowptr : s->rp.as<int64_t>();
	launch_scan(L, s->deg.as<uint32_t>(), rp, P.cnt, s->scan_state.as<uint64_t>(), s->ticket.as<uint32_t>());
	if (rp != d_rowptr && d_rowptr) launch_emit_rowptr(L, rp, P.halo_n, to - from, d_rowptr);
	P.rp = rp;
	cudaEventRecord(s->ev[2], s->stream);
	PeekArgs pk;
	pk.add(rp + P.halo_n, 8);
	pk.add(rp + P.cnt, 8);
	launch_peek(L, pk, s->d_pin);
	CU(cudaStreamSynchronize(s->stream));
	P.halo_arcs = (int64_t)s->h_pin[0];
	P.arcs = (int64_t)s->h_pin[1] - P.halo_arcs;
	cudaEventRecord(s->ev[3], s->stream);
	return WGBV_OK;
}

extern "C" int wgbv_decode_degrees(wgbv_stream *s, int64_t from, int64_t to, int64_t *d_rowptr, int64_t *arcs_out) {
	clear_error();
	if (!s || !d_rowptr) return fail(WGBV_EARG, "null argument");
	CU(cudaSetDevice(s->g->device));
	int rc = check_range(s->g, from, to);
	if (rc) return rc;
	// no halo is needed for degrees alone
	wgbv_graph *g = s->g;
	const Launch L = s->launch();
	const int64_t cnt = to - from;
	const size_t c = (size_t)std::max<int64_t>(cnt, 1);
	if (cnt >= (int64_t)0xfffffff0ll) return fail(WGBV_EARG, "a single decode call is limited to 2^32-16 nodes; split the range");
	if ((rc = s->deg.ensure(c * 4)) || (rc = s->ref.ensure(c)) || (rc = s->scan_state.ensure(((c + 1) / 2048 + 2) * 8)) || (rc = s->ticket.ensure(64))) return rc;
	launch_header(L, s->dev, g->def_codes, from, cnt, s->deg.as<uint32_t>(), s->ref.as<uint8_t>());
	launch_scan(L, s->deg.as<uint32_t>(), d_rowptr, cnt, s->scan_state.as<uint64_t>(), s->ticket.as<uint32_t>());
	CU(cudaMemcpyAsync(s->h_pin, d_rowptr + cnt, 8, cudaMemcpyDeviceToHost, s->stream));
	if ((rc = collect_errors(s))) return rc;
	if (arcs_out) *arcs_out = (int64_t)s->h_pin[0];
	return WGBV_OK;
}

// Stage 3 on a prepared range.
static int decode_rows(wgbv_stream *s, int64_t from, int64_t to, Prepared &P, int64_t *d_col, uint64_t col_capacity) {
	wgbv_graph *g = s->g;
	const Launch L = s->launch();
	int rc;
	if ((uint64_t)P.arcs > col_capacity) return fail(WGBV_EARG, "col_capacity (%llu) is smaller than the %lld arcs of [%lld, %lld)", (unsigned long long)col_capacity, (long long)P.arcs, (long long)from, (long long)to);
	if (P.cnt <= 0 || to <= from) return WGBV_OK;  // nothing to decode (the row pointers have been written)
	RangeDev R;
	R.lo = P.lo;
	R.from = from;
	R.hi = to;
	R.deg = s->deg.as<uint32_t>();
	R.ref = s->ref.as<uint8_t>();
	R.depth = s->depth.as<uint32_t>();
	R.rp = P.rp;
	R.col = d_col;
	R.halo = s->halo.as<int64_t>();
	R.col_capacity = col_capacity;
	R.info = nullptr;
	// Shallow chains (max_ref_count <= 8, the usual case): no separate depth pass -- the level-list kernels walk the
	// chains themselves.  Otherwise (or if a chain turns out longer than the properties promise) k_depth runs first.
	bool walk = !P.have_depth && !P.no_walk && g->halo_nodes >= 0 && g->props.max_ref_count <= 8;
	// (max_ref_count <= 0 promises that no row has a reference; with a window the stream may still hold some: check)
	if (walk && g->props.window_size > 0 && g->props.max_ref_count <= 0) walk = false;
	if (!walk && !P.have_depth && (rc = compute_depth(s, from, P))) return rc;
	if ((rc = s->halo.ensure((size_t)std::max<int64_t>(P.halo_arcs, 1) * 8)) || (rc = s->depth.ensure((size_t)std::max<int64_t>(P.cnt, 1) * 4))) return rc;
	R.depth = walk ? nullptr : s->depth.as<uint32_t>();
	R.halo = s->halo.as<int64_t>();
	if ((rc = s->row_info.ensure((size_t)std::max<int64_t>(P.cnt, 1) * sizeof(RowInfo)))) return rc;
	R.info = s->row_info.as<RowInfo>();
	// rows with many extras go through a global, class-sorted list: at most total_arcs / threshold of them
	HeavyDev H;
	const size_t hcap = (size_t)((P.arcs + P.halo_arcs) / heavy_threshold() + 1);
	if ((rc = s->heavy_rows.ensure(hcap * 4)) || (rc = s->heavy_sorted.ensure(hcap * 4)) || (rc = s->heavy_cls.ensure(hcap)) ||
	    (rc = s->heavy_counters.ensure(130 * 8)) || (rc = s->giants.ensure(sizeof(GiantRow) * GIANT_MAX)) ||
	    (rc = s->gchunks.ensure(sizeof(GiantChunk) * (size_t)GIANT_MAX * GIANT_CHUNKS)))
		return rc;
	H.rows = s->heavy_rows.as<uint32_t>();
	H.capacity = hcap;
	H.sorted = s->heavy_sorted.as<uint32_t>();
	H.cls = s->heavy_cls.as<uint8_t>();
	H.counters = s->heavy_counters.as<unsigned long long>();
	H.giants = s->giants.as<GiantRow>();
	H.gchunks = s->gchunks.as<GiantChunk>();
	uint32_t nlev = walk ? (uint32_t)std::max(g->props.max_ref_count, 0) : P.max_depth;
	uint32_t *deep = s->d_err + 3;
	if (P.front_rp) {
		R.deg_w = s->deg.as<uint32_t>();
		R.ref_w = s->ref.as<uint8_t>();
		R.rp_w = P.front_rp;
		R.rp_from_w = P.halo_arcs;
		R.tile_state = s->scan_state.as<uint64_t>();
		R.tile_ticket = s->ticket.as<uint32_t>();
	}
	launch_extras(L, s->dev, g->def_codes, R, H);
	R.depth = s->depth.as<uint32_t>();
	R.rp_w = nullptr;
	for (int attempt = 0; attempt < 2 && nlev > 0 && nlev <= 8; attempt++) {
		// shallow chains: ordered lists of the rows of every level
		const size_t nblocks = (size_t)((P.cnt + 2047) / 2048);
		if ((rc = s->levels.ensure(nlev * nblocks * 4)) || (rc = s->level_pos.ensure((nlev * nblocks + 1) * 8)) ||
		    (rc = s->list.ensure((size_t)std::max<int64_t>(P.cnt, 1) * 4)) || (rc = s->scan_state.ensure(((nlev * nblocks + 1) / 2048 + 2) * 8)))
			return rc;
		CU(cudaMemsetAsync(deep, 0, 4, s->stream));
		launch_level_lists(L, s->dev, walk ? s->ref.as<uint8_t>() : nullptr, s->depth.as<uint32_t>(), P.lo, from, P.cnt, 0, nlev, s->levels.as<uint32_t>(),
		                   s->level_pos.as<int64_t>(), s->list.as<uint32_t>(), s->scan_state.as<uint64_t>(), s->ticket.as<uint32_t>(), deep);
		const int64_t *bounds = (const int64_t *)(s->h_pin + 16);
		PeekArgs pk;
		for (uint32_t l = 0; l <= nlev; l++) pk.add(s->level_pos.as<int64_t>() + (size_t)l * nblocks, 8);
		launch_peek(L, pk, s->d_pin + 16);
		PeekArgs pd;
		pd.add(deep, 4);
		launch_peek(L, pd, s->d_pin + 10);
		CU(cudaStreamSynchronize(s->stream));
		if (walk && s->h_pin[10]) {
			// a chain longer than max_ref_count: the properties lied (or the file is malformed); take the general route
			walk = false;
			if ((rc = compute_depth(s, from, P))) return rc;
			nlev = P.max_depth;
			continue;
		}
		for (uint32_t lev = 1; lev <= nlev; lev++)
			if (bounds[lev] > bounds[lev - 1])
				launch_copy_level(L, s->dev, g->def_codes, R, lev, s->list.as<uint32_t>() + bounds[lev - 1], (uint64_t)(bounds[lev] - bounds[lev - 1]));
		nlev = 0;  // done
	}
	if (nlev == 0) {
	} else {
		// deep chains (depths from k_depth): the same ordered lists, eight levels per pass; every pass also lists the rows
		// that lie deeper, and the next pass works on that list only (two list buffers take turns)
		const uint32_t *ids = nullptr;
		int64_t nids = P.cnt;
		DevBuf *bufs[2] = {&s->list, &s->list2};
		int cur = 0;
		for (uint32_t level0 = 0; level0 < P.max_depth && nids > 0; level0 += 8) {
			const uint32_t n8 = std::min<uint32_t>(8, P.max_depth - level0);
			const size_t nblocks = (size_t)deep_list_blocks(nids);
			if ((rc = s->levels.ensure(9 * nblocks * 4)) || (rc = s->level_pos.ensure((9 * nblocks + 1) * 8)) ||
			    (rc = bufs[cur]->ensure((size_t)std::max<int64_t>(nids, 1) * 4)) || (rc = s->scan_state.ensure(((9 * nblocks + 1) / 2048 + 2) * 8)))
				return rc;
			uint32_t *list = bufs[cur]->as<uint32_t>();
			launch_deep_lists(L, s->depth.as<uint32_t>(), ids, nids, level0, n8, s->levels.as<uint32_t>(), s->level_pos.as<int64_t>(), list,
			                  s->scan_state.as<uint64_t>(), s->ticket.as<uint32_t>());
			const int64_t *bounds = (const int64_t *)(s->h_pin + 16);
			PeekArgs pk;
			for (uint32_t l = 0; l <= n8 + 1; l++) pk.add(s->level_pos.as<int64_t>() + (size_t)l * nblocks, 8);
			launch_peek(L, pk, s->d_pin + 16);
			CU(cudaStreamSynchronize(s->stream));
			for (uint32_t l = 1; l <= n8; l++)
				if (bounds[l] > bounds[l - 1])
					launch_copy_level(L, s->dev, g->def_codes, R, level0 + l, list + bounds[l - 1], (uint64_t)(bounds[l] - bounds[l - 1]));
			ids = list + bounds[n8];
			nids = bounds[n8 + 1] - bounds[n8];
			cur ^= 1;
		}
	}
	cudaEventRecord(s->ev[4], s->stream);
	s->ev_valid = true;
	const cudaError_t e = cudaGetLastError();
	if (e != cudaSuccess) return fail(WGBV_ECUDA, "kernel launch failed: %s", cudaGetErrorString(e));
	return WGBV_OK;
}

// The level-synchronous path for [from, to): valid for every graph the library opens.
static int decode_general(wgbv_stream *s, int64_t from, int64_t to, int64_t *d_rowptr, int64_t *d_col, uint64_t col_capacity, bool trust_maxref) {
	int rc;
	Prepared P;
	wgbv_graph *g = s->g;
	// Bounded, shallow reference chains and a range that needs no halo run without the separate header and scan
	// launches and without the read-back of the arc count: k_extras does stages 1 and 2 for its own tiles
	// (RangeDev::rp_w) and the capacity is checked on the device.  WGBV_SPLIT_FRONT=1 keeps the separate stages.
	if (trust_maxref && s->front_in_extras && to > from && g->halo_nodes >= 0 && g->props.max_ref_count <= 8 &&
	    !(g->props.window_size > 0 && g->props.max_ref_count <= 0) && to - halo_start(g, from) < (int64_t)0xfffffff0ll) {
		const Launch L = s->launch();
		P.lo = halo_start(g, from);
		P.cnt = to - P.lo;
		P.halo_n = from - P.lo;
		P.halo_arcs = 0;
		const size_t c = (size_t)P.cnt;
		if ((rc = s->deg.ensure(c * 4)) || (rc = s->ref.ensure(c)) || (rc = s->scan_state.ensure(((c + 1) / 2048 + (size_t)extras_tiles(P.cnt) + 2) * 8)) ||
		    (rc = s->ticket.ensure(64)))
			return rc;
		for (int i = 0; i < 2; i++) cudaEventRecord(s->ev[i], s->stream);
		if (P.halo_n > 0) {
			// halo rows (references of the first rows of the range reach back to them): their successors go to a scratch
			// array whose size must be known before the launch, so their outdegrees are read and added up first -- a few
			// rows, one read-back.  The row pointers then cover [lo, to) as in the general path and those of [from, to)
			// are copied out at the end.
			if ((rc = s->rp.ensure((c + 1) * 8))) return rc;
			launch_header(L, s->dev, g->def_codes, P.lo, P.halo_n, s->deg.as<uint32_t>(), s->ref.as<uint8_t>());
			launch_scan(L, s->deg.as<uint32_t>(), s->rp.as<int64_t>(), P.halo_n, s->scan_state.as<uint64_t>(), s->ticket.as<uint32_t>());
			PeekArgs pk;
			pk.add(s->rp.as<int64_t>() + P.halo_n, 8);
			launch_peek(L, pk, s->d_pin);
			CU(cudaStreamSynchronize(s->stream));
			P.halo_arcs = (int64_t)s->h_pin[0];
		}
		for (int i = 2; i < 4; i++) cudaEventRecord(s->ev[i], s->stream);  // (stages 1-2 have no time of their own here)
		// an upper bound on the arcs of the range: sizes the scratch of the heavy rows (never more than the whole graph has)
		P.arcs = (int64_t)std::min<uint64_t>(col_capacity, (uint64_t)std::max<int64_t>(g->props.arcs, 0));
		P.front_rp = P.halo_n ? s->rp.as<int64_t>() : d_rowptr;
		P.rp = P.front_rp;
		if ((rc = decode_rows(s, from, to, P, d_col, col_capacity))) return rc;
		if (P.halo_n) launch_emit_rowptr(L, P.rp, P.halo_n, to - from, d_rowptr);
		return WGBV_OK;
	}
	if ((rc = prepare(s, from, to, d_rowptr, P, trust_maxref))) return rc;
	return decode_rows(s, from, to, P, d_col, col_capacity);
}

// One pass of the fused tile kernel over [from, to) (bv_fused.cu); asynchronous, no read-back.
static bool fused_applies(const wgbv_stream *s) {
	const wgbv_graph *g = s->g;
	return s->use_fused && g->fused_rows > 0;
}

static int decode_fused(wgbv_stream *s, int64_t from, int64_t to, int64_t *d_rowptr, int64_t *d_col, uint64_t col_capacity) {
	wgbv_graph *g = s->g;
	int rc;
	FusedDev F;
	F.from = from;
	F.to = to;
	F.halo = (int32_t)g->halo_nodes;
	F.tile_rows = g->fused_rows;
	const int64_t ntiles = (to - from + F.tile_rows - 1) / F.tile_rows;
	if (ntiles >= (int64_t)0x7fffffff) return fail(WGBV_EARG, "a single decode call is limited to 2^31 tiles; split the range");
	F.ntiles = (uint32_t)ntiles;
	if ((rc = s->scan_state.ensure((size_t)ntiles * 8)) || (rc = s->ticket.ensure(64)) || (rc = s->tile_done.ensure((size_t)ntiles * 4))) return rc;
	F.rowptr = d_rowptr;
	F.col = d_col;
	F.col_capacity = col_capacity;
	F.tile_state = s->scan_state.as<uint64_t>();
	F.ticket = s->ticket.as<uint32_t>();
	F.slow_count = s->d_err + 2;
	F.tile_done = s->tile_done.as<uint32_t>();
	for (int i = 0; i < 4; i++) cudaEventRecord(s->ev[i], s->stream);
	CU(cudaMemsetAsync(F.tile_state, 0, (size_t)ntiles * 8, s->stream));
	CU(cudaMemsetAsync(F.ticket, 0, 64, s->stream));
	CU(cudaMemsetAsync(F.slow_count, 0, 4, s->stream));
	const int e = launch_fused(s->launch(), s->dev, g->def_codes, F, g->sm_count);
	if (e) return fail(WGBV_ECUDA, "launch of the fused tile kernel failed: %s", cudaGetErrorString((cudaError_t)e));
	cudaEventRecord(s->ev[4], s->stream);
	s->ev_valid = true;
	s->fused_pending.push_back({from, to, d_rowptr, d_col, col_capacity});
	s->last_tiles = F.ntiles;
	const cudaError_t le = cudaGetLastError();
	if (le != cudaSuccess) return fail(WGBV_ECUDA, "kernel launch failed: %s", cudaGetErrorString(le));
	return WGBV_OK;
}

extern "C" int wgbv_decode_csr(wgbv_stream *s, int64_t from, int64_t to, int64_t *d_rowptr, int64_t *d_col, uint64_t col_capacity) {
	clear_error();
	if (!s || !d_rowptr || (!d_col && col_capacity)) return fail(WGBV_EARG, "null argument");
	CU(cudaSetDevice(s->g->device));
	int rc = check_range(s->g, from, to);
	if (rc) return rc;
	if (s->fused_pending.size() >= 16) {
		// many decodes in flight whose outcome nobody has asked for: read it before the list of possible reruns grows
		if ((rc = collect_errors(s))) return rc;
	}
	if (to == from) {
		CU(cudaMemsetAsync(d_rowptr, 0, 8, s->stream));  // rowptr[0] = 0: an empty range has no other output
		return WGBV_OK;
	}
	if (fused_applies(s)) return decode_fused(s, from, to, d_rowptr, d_col, col_capacity);
	return decode_general(s, from, to, d_rowptr, d_col, col_capacity, true);
}

extern "C" int wgbv_fused_stats(wgbv_stream *s, uint64_t out[4]) {
	clear_error();
	if (!s || !out) return fail(WGBV_EARG, "null argument");
	out[0] = s->last_tiles;
	out[1] = s->last_slow_tiles;
	out[2] = s->fused_fallbacks;
	out[3] = fused_applies(s) ? 1 : 0;
	return WGBV_OK;
}

extern "C" int wgbv_last_stage_ms(wgbv_stream *s, float out[5]) {
	clear_error();
	if (!s || !out) return fail(WGBV_EARG, "null argument");
	if (!s->ev_valid) return fail(WGBV_ESTATE, "no decode has run on this stream");
	CU(cudaStreamSynchronize(s->stream));
	for (int i = 0; i < 4; i++) CU(cudaEventElapsedTime(&out[i], s->ev[i], s->ev[i + 1]));
	CU(cudaEventElapsedTime(&out[4], s->ev[0], s->ev[4]));
	return WGBV_OK;
}

extern "C" int wgbv_decode_csr_host(wgbv_stream *s, int64_t from, int64_t to, int64_t *h_rowptr, int64_t *h_col, uint64_t col_capacity,
                                    int64_t *arcs_out) {
	clear_error();
	if (!s || !h_rowptr) return fail(WGBV_EARG, "null argument");
	CU(cudaSetDevice(s->g->device));
	int rc = check_range(s->g, from, to);
	if (rc) return rc;
	const int64_t cnt = to - from;
	if ((rc = s->rowptr.ensure((size_t)(cnt + 1) * 8))) return rc;
	if (h_col && cnt > 0 && fused_applies(s)) {
		// one pass of the tile kernel into a device buffer of the caller's capacity; the arc count comes back with the row pointers
		if ((rc = s->col.ensure((size_t)std::max<uint64_t>(col_capacity, 1) * 8))) return rc;
		if ((rc = decode_fused(s, from, to, s->rowptr.as<int64_t>(), s->col.as<int64_t>(), col_capacity))) return rc;
		if ((rc = collect_errors(s))) return rc;
		CU(cudaMemcpyAsync(h_rowptr, s->rowptr.as<int64_t>(), (size_t)(cnt + 1) * 8, cudaMemcpyDeviceToHost, s->stream));
		CU(cudaStreamSynchronize(s->stream));
		const int64_t arcs = h_rowptr[cnt];
		if (arcs_out) *arcs_out = arcs;
		if (arcs > 0) {
			CU(cudaMemcpyAsync(h_col, s->col.as<int64_t>(), (size_t)arcs * 8, cudaMemcpyDeviceToHost, s->stream));
			CU(cudaStreamSynchronize(s->stream));
		}
		return WGBV_OK;
	}
	Prepared P;
	if ((rc = prepare(s, from, to, s->rowptr.as<int64_t>(), P))) return rc;
	if (arcs_out) *arcs_out = P.arcs;
	CU(cudaMemcpyAsync(h_rowptr, s->rowptr.as<int64_t>(), (size_t)(cnt + 1) * 8, cudaMemcpyDeviceToHost, s->stream));
	if (h_col) {
		if ((uint64_t)P.arcs > col_capacity) return fail(WGBV_EARG, "col_capacity (%llu) is smaller than the %lld arcs of the range", (unsigned long long)col_capacity, (long long)P.arcs);
		if ((rc = s->col.ensure((size_t)std::max<int64_t>(P.arcs, 1) * 8))) return rc;
		if ((rc = decode_rows(s, from, to, P, s->col.as<int64_t>(), (uint64_t)P.arcs))) return rc;
		if (P.arcs) CU(cudaMemcpyAsync(h_col, s->col.as<int64_t>(), (size_t)P.arcs * 8, cudaMemcpyDeviceToHost, s->stream));
	}
	return collect_errors(s);
}

// ------------------------------------------------------------------------------------------------ streaming scan
static int error_from_bits(wgbv_stream *s, uint32_t e);

static int pinned_ensure(int64_t **p, size_t *cap, size_t elems) {
	if (elems <= *cap) return WGBV_OK;
	if (*p) cudaFreeHost(*p);
	*p = nullptr;
	*cap = 0;
	const size_t want = elems + elems / 4 + 1024;
	if (cudaMallocHost((void **)p, want * 8) != cudaSuccess) return fail(WGBV_ECUDA, "cudaMallocHost of %zu bytes failed", want * 8);
	*cap = want;
	return WGBV_OK;
}

// Streaming upload: queues, on the upload stream, the part of the bit stream that chunk number `c` (nodes up to `b`)
// needs and that has not been sent yet.  Pieces are disjoint 64-bit-unit ranges in stream order, each ending with the
// unit that holds the last bit of its chunk, so a piece never rewrites words an earlier chunk may still be reading.
static int scan_upload_chunk(wgbv_stream *s, int64_t c, int64_t b) {
	wgbv_graph *g = s->g;
	if (c < s->up_chunks) return WGBV_OK;
	const uint64_t end_bit = g->ef.get((uint64_t)(b - g->avail_lo));  // relative to the first resident word
	uint64_t end_word = b >= s->scan_to ? g->data_words : std::min<uint64_t>(g->data_words, (end_bit >> 6) + 1);
	if (end_word > s->up_next_word) {
		const uint64_t w0 = s->up_next_word, n = end_word - w0;
		const uint64_t src0 = g->byte0 + w0 * 8, avail = g->byte1 > src0 ? g->byte1 - src0 : 0;
		const uint64_t bytes = std::min<uint64_t>(n * 8, avail);
		if (bytes < n * 8) CU(cudaMemsetAsync(g->d_words + w0 + bytes / 8, 0, (size_t)(n * 8 - (bytes / 8) * 8), s->upload_stream));
		if (bytes) CU(cudaMemcpyAsync(g->d_words + w0, s->up_graph + src0, (size_t)bytes, cudaMemcpyHostToDevice, s->upload_stream));
		Launch L{s->upload_stream, &s->launches};
		launch_bswap(L, g->d_words + w0, n);
		s->up_next_word = end_word;
	}
	CU(cudaEventRecord(s->uploaded[c & 1], s->upload_stream));
	s->up_chunks = c + 1;
	return WGBV_OK;
}

// Queues the decode of the next chunk and its copy back to pinned host memory.
static int scan_issue_one(wgbv_stream *s) {
	wgbv_stream::ScanSlot &sl = s->slot[s->issued % wgbv_stream::SCAN_SLOTS];
	const int64_t a = s->scan_issue, b = std::min(a + s->scan_chunk, s->scan_to);
	int rc;
	if (s->up_graph) {
		// this chunk's words, then (so that they travel while this chunk is decoded) the next chunk's
		if ((rc = scan_upload_chunk(s, s->issued, b))) return rc;
		CU(cudaStreamWaitEvent(s->stream, s->uploaded[s->issued & 1], 0));
		if (b < s->scan_to && (rc = scan_upload_chunk(s, s->issued + 1, std::min(b + s->scan_chunk, s->scan_to)))) return rc;
	}
	if ((rc = sl.d_rowptr.ensure((size_t)(b - a + 1) * 8))) return rc;
	Prepared P;
	if ((rc = prepare(s, a, b, sl.d_rowptr.as<int64_t>(), P))) return rc;  // synchronises `stream`: the previous chunk is decoded
	const size_t esz = s->narrow ? 4 : 8;  // bytes per delivered successor
	if ((rc = sl.d_col.ensure((size_t)std::max<int64_t>(P.arcs, 1) * esz))) return rc;
	if (s->narrow && (rc = s->col.ensure((size_t)std::max<int64_t>(P.arcs, 1) * 8))) return rc;
	// (the pinned buffers are sized in 8-byte units)
	if ((rc = pinned_ensure(&sl.h_rowptr, &sl.h_rowptr_cap, (size_t)(b - a + 1))) ||
	    (rc = pinned_ensure(&sl.h_col, &sl.h_col_cap, (size_t)(std::max<int64_t>(P.arcs, 1) * esz + 7) / 8)))
		return rc;
	if (s->narrow) {
		// decode into the stream's own buffer (the next chunk's decode follows the narrowing kernel in stream order)
		if ((rc = decode_rows(s, a, b, P, s->col.as<int64_t>(), (uint64_t)P.arcs))) return rc;
		launch_narrow(s->launch(), s->col.as<int64_t>(), sl.d_col.as<uint32_t>(), (uint64_t)P.arcs);
	} else if ((rc = decode_rows(s, a, b, P, sl.d_col.as<int64_t>(), (uint64_t)P.arcs))) return rc;
	CU(cudaEventRecord(sl.decoded, s->stream));
	CU(cudaStreamWaitEvent(s->copy_stream, sl.decoded, 0));
	CU(cudaMemcpyAsync(sl.h_rowptr, sl.d_rowptr.as<int64_t>(), (size_t)(b - a + 1) * 8, cudaMemcpyDeviceToHost, s->copy_stream));
	if (P.arcs) CU(cudaMemcpyAsync(sl.h_col, sl.d_col.as<char>(), (size_t)P.arcs * esz, cudaMemcpyDeviceToHost, s->copy_stream));
	CU(cudaMemcpyAsync(sl.h_err, s->d_err, 4, cudaMemcpyDeviceToHost, s->copy_stream));
	CU(cudaEventRecord(sl.ready, s->copy_stream));
	sl.first = a;
	sl.last = b;
	sl.arcs = P.arcs;
	s->scan_issue = b;
	s->issued++;
	return WGBV_OK;
}

extern "C" int wgbv_scan_begin(wgbv_stream *s, int64_t from, int64_t to, int64_t chunk_nodes) {
	clear_error();
	if (!s) return fail(WGBV_EARG, "null argument");
	if (chunk_nodes < 1) return fail(WGBV_EARG, "chunk_nodes must be positive");
	CU(cudaSetDevice(s->g->device));
	int rc = check_range(s->g, from, to);
	if (rc) return rc;
	if (!s->copy_stream) {
		CU(cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking));
		for (auto &sl : s->slot) {
			CU(cudaEventCreateWithFlags(&sl.decoded, cudaEventDisableTiming));
			CU(cudaEventCreateWithFlags(&sl.ready, cudaEventDisableTiming));
			CU(cudaMallocHost((void **)&sl.h_err, 64));
		}
	}
	CU(cudaStreamSynchronize(s->copy_stream));
	s->up_graph = nullptr;
	s->scanning = true;
	s->scan_to = to;
	s->scan_chunk = chunk_nodes;
	s->scan_issue = s->scan_take = from;
	s->issued = s->taken = 0;
	return WGBV_OK;
}

extern "C" int wgbv_scan_begin_upload(wgbv_stream *s, const uint8_t *graph, uint64_t graph_bytes, int64_t from, int64_t to, int64_t chunk_nodes) {
	clear_error();
	if (!s || !graph) return fail(WGBV_EARG, "null argument");
	wgbv_graph *g = s->g;
	if (graph_bytes < g->byte1) return fail(WGBV_EARG, "the buffer (%llu bytes) is shorter than the stream that was opened (%llu bytes)", (unsigned long long)graph_bytes, (unsigned long long)g->byte1);
	if (from != g->shard_lo) return fail(WGBV_EARG, "a scan that uploads the stream has to start at the first node of the handle (%lld)", (long long)g->shard_lo);
	if (g->streams > 1) return fail(WGBV_ESTATE, "the bit stream can only be uploaded again while this is the graph's only stream (%d exist)", g->streams);
	int rc = wgbv_scan_begin(s, from, to, chunk_nodes);
	if (rc) return rc;
	if (!s->upload_stream) {
		CU(cudaStreamCreateWithFlags(&s->upload_stream, cudaStreamNonBlocking));
		for (auto &e : s->uploaded) CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
	}
	CU(cudaStreamSynchronize(s->stream));
	CU(cudaStreamSynchronize(s->upload_stream));
	s->up_graph = graph;
	s->up_next_word = 0;
	s->up_chunks = 0;
	return WGBV_OK;
}

extern "C" int wgbv_scan_next(wgbv_stream *s, int64_t *first, int64_t *last, const int64_t **h_rowptr, const int64_t **h_col, int64_t *arcs) {
	clear_error();
	if (!s || !first || !last || !h_rowptr || !h_col) return fail(WGBV_EARG, "null argument");
	if (!s->scanning) return fail(WGBV_ESTATE, "wgbv_scan_begin has not been called");
	CU(cudaSetDevice(s->g->device));
	if (s->scan_take >= s->scan_to && s->taken == s->issued) return 0;
	int rc;
	// keep the pipeline full: the chunk handed out now plus the ones decoding / copying behind it; the slot of the chunk
	// handed out by the PREVIOUS call becomes reusable here (its buffers were valid until this call)
	while (s->scan_issue < s->scan_to && s->issued - s->taken < wgbv_stream::SCAN_SLOTS - 1)
		if ((rc = scan_issue_one(s))) return rc;
	if (s->issued == s->taken) return 0;
	wgbv_stream::ScanSlot &sl = s->slot[s->taken % wgbv_stream::SCAN_SLOTS];
	CU(cudaEventSynchronize(sl.ready));
	if (sl.h_err[0]) {
		cudaMemsetAsync(s->d_err, 0, 4, s->stream);
		return error_from_bits(s, sl.h_err[0]);
	}
	*first = sl.first;
	*last = sl.last;
	*h_rowptr = sl.h_rowptr;
	*h_col = sl.h_col;
	if (arcs) *arcs = sl.arcs;
	s->scan_take = sl.last;
	s->taken++;
	return 1;
}

extern "C" int wgbv_scan_set_narrow(wgbv_stream *s, int on) {
	clear_error();
	if (!s) return fail(WGBV_EARG, "null argument");
	if (s->scanning) return fail(WGBV_ESTATE, "a scan is running on this stream");
	if (on && (uint64_t)s->g->props.nodes > (1ull << 32))
		return fail(WGBV_EUNSUPPORTED, "node ids of a graph with %lld nodes do not fit 32 bits", (long long)s->g->props.nodes);
	s->narrow = on != 0;
	return WGBV_OK;
}

extern "C" void wgbv_scan_end(wgbv_stream *s) {
	if (!s || !s->scanning) return;
	cudaSetDevice(s->g->device);
	cudaStreamSynchronize(s->stream);
	if (s->copy_stream) cudaStreamSynchronize(s->copy_stream);
	if (s->upload_stream) cudaStreamSynchronize(s->upload_stream);
	if (s->up_graph && s->up_next_word < s->g->data_words) {
		// the scan was abandoned: send the rest so that the handle holds the whole stream again
		s->scan_to = s->g->shard_hi;
		scan_upload_chunk(s, s->up_chunks, s->scan_to);
		cudaStreamSynchronize(s->upload_stream);
	}
	s->up_graph = nullptr;
	s->scanning = false;
}

// ------------------------------------------------------------------------------------------------ union / symmetrize
extern "C" int wgbv_union_csr(wgbv_stream *s, int64_t n, const int64_t *d_a_rowptr, const int64_t *d_a_col, const int64_t *d_b_rowptr,
                              const int64_t *d_b_col, int64_t *d_u_rowptr, int64_t *d_u_col, uint64_t u_capacity, int64_t *arcs_out) {
	clear_error();
	if (!s || !d_a_rowptr || !d_b_rowptr || !d_u_rowptr || n < 0) return fail(WGBV_EARG, "null argument or negative node count");
	CU(cudaSetDevice(s->g->device));
	const Launch L = s->launch();
	int rc;
	const size_t nn = (size_t)std::max<int64_t>(n, 1);
	if ((rc = s->t_indeg.ensure(nn * 4)) || (rc = s->scan_state.ensure(((nn + 1) / 2048 + 2) * 8)) || (rc = s->ticket.ensure(64))) return rc;
	launch_union_count(L, n, d_a_rowptr, d_a_col, d_b_rowptr, d_b_col, s->t_indeg.as<uint32_t>(), s->d_err);
	launch_scan(L, s->t_indeg.as<uint32_t>(), d_u_rowptr, n, s->scan_state.as<uint64_t>(), s->ticket.as<uint32_t>());
	PeekArgs pk;
	pk.add(d_u_rowptr + n, 8);
	launch_peek(L, pk, s->d_pin);
	CU(cudaStreamSynchronize(s->stream));
	const int64_t arcs = (int64_t)s->h_pin[0];
	if (arcs_out) *arcs_out = arcs;
	if (d_u_col) {
		if ((uint64_t)arcs > u_capacity) return fail(WGBV_EARG, "u_capacity (%llu) is smaller than the %lld arcs of the union", (unsigned long long)u_capacity, (long long)arcs);
		launch_union_fill(L, n, d_a_rowptr, d_a_col, d_b_rowptr, d_b_col, d_u_rowptr, d_u_col, u_capacity, s->d_err);
	}
	const cudaError_t e = cudaGetLastError();
	if (e != cudaSuccess) return fail(WGBV_ECUDA, "kernel launch failed: %s", cudaGetErrorString(e));
	return collect_errors(s);
}

extern "C" int wgbv_symmetrize_csr(wgbv_stream *s, int64_t n, const int64_t *d_rowptr, const int64_t *d_col, int64_t *d_s_rowptr, int64_t *d_s_col,
                                   uint64_t s_capacity, int64_t *arcs_out) {
	clear_error();
	if (!s || !d_rowptr || !d_s_rowptr || n < 0) return fail(WGBV_EARG, "null argument or negative node count");
	CU(cudaSetDevice(s->g->device));
	int64_t h[2];
	CU(cudaMemcpyAsync(&h[0], d_rowptr, 8, cudaMemcpyDeviceToHost, s->stream));
	CU(cudaMemcpyAsync(&h[1], d_rowptr + n, 8, cudaMemcpyDeviceToHost, s->stream));
	CU(cudaStreamSynchronize(s->stream));
	const int64_t arcs = h[1] - h[0];
	if (arcs < 0) return fail(WGBV_EARG, "row pointers are not increasing");
	int rc;
	if ((rc = s->sym_rowptr.ensure((size_t)(n + 1) * 8)) || (rc = s->sym_col.ensure((size_t)std::max<int64_t>(arcs, 1) * 8))) return rc;
	if ((rc = wgbv_transpose_csr(s, n, d_rowptr, d_col, s->sym_rowptr.as<int64_t>(), s->sym_col.as<int64_t>()))) return rc;
	return wgbv_union_csr(s, n, d_rowptr, d_col, s->sym_rowptr.as<int64_t>(), s->sym_col.as<int64_t>(), d_s_rowptr, d_s_col, s_capacity, arcs_out);
}

// ------------------------------------------------------------------------------------------------ compression
extern "C" int wgbv_encode_csr(int device, const wgbv_props *props, int64_t first_node, int64_t n, const int64_t *d_rowptr, const int64_t *d_col,
                               int64_t split_nodes, uint8_t *graph_out, uint64_t graph_capacity, uint64_t *graph_bits, uint64_t *node_bits_out) {
	clear_error();
	if (!props || !d_rowptr || !graph_bits || n < 0 || split_nodes < 0) return fail(WGBV_EARG, "null argument or negative size");
	int rc = check_codings(*props);
	if (rc) return rc;
	if ((rc = select_device(device))) return rc;
	*graph_bits = 0;
	if (n == 0) return WGBV_OK;
	EncDev E;
	E.rowptr = d_rowptr;
	E.col = d_col;
	E.n = n;
	E.first_node = first_node;
	E.split = split_nodes > 0 ? split_nodes : n;
	E.window = props->window_size;
	E.maxref = props->max_ref_count;
	E.minint = props->min_interval_length;
	E.zetak = props->zeta_k;
	E.outdegree_coding = props->outdegree_coding;
	E.block_coding = props->block_coding;
	E.residual_coding = props->residual_coding;
	E.reference_coding = props->reference_coding;
	E.blockcount_coding = props->block_count_coding;
	cudaStream_t st = nullptr;
	DevBuf cost, chosen, bits, off, state, ticket, words;
	auto done = [&](int code) {
		for (DevBuf *b : {&cost, &chosen, &bits, &off, &state, &ticket, &words}) b->release();
		if (st) cudaStreamDestroy(st);
		return code;
	};
	if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess) return done(fail(WGBV_ECUDA, "cudaStreamCreate failed"));
	int64_t h_rp[2];
	if (cudaMemcpy(&h_rp[0], d_rowptr, 8, cudaMemcpyDeviceToHost) != cudaSuccess || cudaMemcpy(&h_rp[1], d_rowptr + n, 8, cudaMemcpyDeviceToHost) != cudaSuccess)
		return done(fail(WGBV_ECUDA, "reading the row pointers failed: %s", cudaGetErrorString(cudaGetLastError())));
	const int64_t arcs = h_rp[1] - h_rp[0];
	if (arcs < 0 || (arcs && !d_col)) return done(fail(WGBV_EARG, "row pointers are not increasing, or no successors given"));
	const size_t cyc = (size_t)props->window_size + 1;
	// room for the stream: the caller's capacity bounds what is worth producing
	const uint64_t words_cap = (graph_capacity / 8 + 2) * 2;
	if ((rc = cost.ensure((size_t)n * cyc * 4)) || (rc = chosen.ensure((size_t)n)) || (rc = bits.ensure((size_t)n * 8)) || (rc = off.ensure((size_t)(n + 1) * 8)) ||
	    (rc = state.ensure(((size_t)(n + 1) / 2048 + 2) * 8)) || (rc = ticket.ensure(64)) || (rc = words.ensure((size_t)words_cap * 4)))
		return done(rc);
	uint64_t total = 0;
	const int e = encode_csr_device(st, E, cost.as<uint32_t>(), chosen.as<uint8_t>(), bits.as<unsigned long long>(), off.as<int64_t>(),
	                                state.as<uint64_t>(), ticket.as<uint32_t>(), words.as<uint32_t>(), words_cap, &total);
	*graph_bits = total;
	if (e == -2) return done(fail(WGBV_EARG, "graph_capacity (%llu bytes) is too small for the %llu bits of the stream", (unsigned long long)graph_capacity, (unsigned long long)total));
	if (e) return done(fail(WGBV_ECUDA, "compression failed on the device: %s", cudaGetErrorString(cudaGetLastError())));
	const uint64_t nbytes = (total + 7) / 8;
	if (nbytes > graph_capacity || (!graph_out && nbytes)) return done(fail(WGBV_EARG, "graph_capacity (%llu bytes) is too small for the %llu bits of the stream", (unsigned long long)graph_capacity, (unsigned long long)total));
	if (nbytes && cudaMemcpyAsync(graph_out, words.p, (size_t)nbytes, cudaMemcpyDeviceToHost, st) != cudaSuccess) return done(fail(WGBV_ECUDA, "copy of the stream failed"));
	if (node_bits_out && cudaMemcpyAsync(node_bits_out, bits.p, (size_t)n * 8, cudaMemcpyDeviceToHost, st) != cudaSuccess) return done(fail(WGBV_ECUDA, "copy of the record lengths failed"));
	if (cudaStreamSynchronize(st) != cudaSuccess) return done(fail(WGBV_ECUDA, "device error while compressing: %s", cudaGetErrorString(cudaGetLastError())));
	return done(WGBV_OK);
}

// ------------------------------------------------------------------------------------------------ transpose
extern "C" int wgbv_transpose_csr(wgbv_stream *s, int64_t n, const int64_t *d_rowptr, const int64_t *d_col, int64_t *d_t_rowptr, int64_t *d_t_col) {
	clear_error();
	if (!s || !d_rowptr || !d_t_rowptr || n < 0) return fail(WGBV_EARG, "null argument or negative node count");
	if ((uint64_t)n >= (1ull << 32)) return fail(WGBV_EUNSUPPORTED, "wgbv_transpose_csr packs node ids in 32 bits: %lld nodes are too many", (long long)n);
	CU(cudaSetDevice(s->g->device));
	const Launch L = s->launch();
	int rc;
	PeekArgs pk;
	pk.add(d_rowptr, 8);
	pk.add(d_rowptr + n, 8);
	launch_peek(L, pk, s->d_pin);
	CU(cudaStreamSynchronize(s->stream));
	const int64_t arcs = (int64_t)s->h_pin[1] - (int64_t)s->h_pin[0];
	if (arcs < 0) return fail(WGBV_EARG, "row pointers are not increasing");
	if (arcs && (!d_col || !d_t_col)) return fail(WGBV_EARG, "null argument");
	const size_t nn = (size_t)std::max<int64_t>(n, 1), na = (size_t)std::max<int64_t>(arcs, 1);
	const uint64_t tiles = transpose_tiles((uint64_t)arcs);
	const size_t ncounts = (size_t)std::max<uint64_t>(tiles * (uint64_t)transpose_digits(), 1);
	const size_t scan_items = std::max(ncounts, nn);
	if ((rc = s->t_indeg.ensure(nn * 4)) || (rc = s->t_pairs[0].ensure(na * 8)) || (rc = s->t_counts.ensure(ncounts * 4)) ||
	    (rc = s->t_pos.ensure((ncounts + 1) * 8)) || (rc = s->scan_state.ensure(((scan_items + 1) / 2048 + 2) * 8)) || (rc = s->ticket.ensure(64)))
		return rc;
	const int passes = arcs ? transpose_passes((uint64_t)n) : 0;
	if (passes > 1 && (rc = s->t_pairs[1].ensure(na * 8))) return rc;
	CU(cudaMemsetAsync(s->t_indeg.as<uint32_t>(), 0, nn * 4, s->stream));
	launch_transpose_pack(L, s->dev, n, d_rowptr, d_col, (uint64_t)arcs, s->t_pairs[0].as<uint64_t>(), s->t_indeg.as<uint32_t>(), s->t_counts.as<uint32_t>());
	// in-degrees -> row pointers of the transpose
	launch_scan(L, s->t_indeg.as<uint32_t>(), d_t_rowptr, n, s->scan_state.as<uint64_t>(), s->ticket.as<uint32_t>());
	int cur = 0;
	for (int p = 0; p < passes; p++) {
		const uint64_t *in = s->t_pairs[cur].as<uint64_t>();
		if (p > 0) launch_radix_hist(L, in, (uint64_t)arcs, p, s->t_counts.as<uint32_t>());  // (pass 0: counted by the pack kernel)
		launch_scan(L, s->t_counts.as<uint32_t>(), s->t_pos.as<int64_t>(), (int64_t)ncounts, s->scan_state.as<uint64_t>(), s->ticket.as<uint32_t>());
		const bool last = p == passes - 1;
		launch_radix_scatter(L, in, (uint64_t)arcs, p, s->t_pos.as<int64_t>(), last ? nullptr : s->t_pairs[cur ^ 1].as<uint64_t>(), last ? d_t_col : nullptr);
		cur ^= 1;
	}
	const cudaError_t e = cudaGetLastError();
	if (e != cudaSuccess) return fail(WGBV_ECUDA, "kernel launch failed: %s", cudaGetErrorString(e));
	return collect_errors(s);
}

// ------------------------------------------------------------------------------------------------ arc labels
// A label stream with its offsets is held exactly like a graph's bit stream with its offsets: the handle wraps a
// wgbv_graph opened over (.labels, .labeloffsets) with neutral format parameters.
struct wgbv_labels {
	wgbv_graph *g = nullptr;
	int kind = 0, width = 0;
};

extern "C" int wgbv_labels_open_mem(int device, int64_t nodes, int kind, int width, const uint8_t *labels, uint64_t labels_bytes,
                                    const uint8_t *labeloffsets, uint64_t labeloffsets_bytes, wgbv_labels **out) {
	clear_error();
	if (!out || (!labels && labels_bytes) || !labeloffsets || nodes < 0) return fail(WGBV_EARG, "null argument or negative node count");
	*out = nullptr;
	if (kind < WGBV_LABEL_FIXED_INT || kind > WGBV_LABEL_FIXED_LONG_LIST) return fail(WGBV_EUNSUPPORTED, "unknown label kind %d", kind);
	if ((kind == WGBV_LABEL_FIXED_INT || kind == WGBV_LABEL_FIXED_INT_LIST) && (width < 0 || width > 32)) return fail(WGBV_EARG, "a fixed-width int label has 0 to 32 bits, not %d", width);
	if (kind == WGBV_LABEL_FIXED_LONG_LIST && (width < 0 || width > 64)) return fail(WGBV_EARG, "a fixed-width long label has 0 to 64 bits, not %d", width);
	wgbv_props p{};
	p.nodes = nodes;
	p.window_size = 0;
	p.max_ref_count = 0;
	p.zeta_k = 3;
	p.outdegree_coding = p.block_coding = p.block_count_coding = p.offset_coding = WGBV_GAMMA;
	p.residual_coding = WGBV_ZETA;
	p.reference_coding = WGBV_UNARY;
	wgbv_graph *g = nullptr;
	const int rc = wgbv_open_mem(&p, labels, labels_bytes, labeloffsets, labeloffsets_bytes, device, 0, 1, &g);
	if (rc) return rc;
	wgbv_labels *l = new (std::nothrow) wgbv_labels();
	if (!l) { wgbv_close(g); return fail(WGBV_EARG, "out of host memory"); }
	l->g = g;
	l->kind = kind;
	l->width = width;
	*out = l;
	return WGBV_OK;
}

extern "C" void wgbv_labels_close(wgbv_labels *l) {
	if (!l) return;
	wgbv_close(l->g);
	delete l;
}

extern "C" int wgbv_labels_decode(wgbv_stream *s, const wgbv_labels *l, int64_t from, int64_t to, const int64_t *d_rowptr,
                                  int64_t *d_labels, uint64_t capacity) {
	clear_error();
	if (!s || !l || !d_rowptr || (!d_labels && capacity)) return fail(WGBV_EARG, "null argument");
	if (from < 0 || to < from || to > l->g->props.nodes) return fail(WGBV_EINDEX, "Node index out of range: [%lld, %lld)", (long long)from, (long long)to);
	if (l->g->device != s->g->device) return fail(WGBV_EARG, "the labels and the stream live on different devices");
	if (l->kind >= WGBV_LABEL_FIXED_INT_LIST) return fail(WGBV_EUNSUPPORTED, "list-valued labels are decoded by wgbv_labels_decode_lists");
	CU(cudaSetDevice(s->g->device));
	GraphDev lab = l->g->dev;
	lab.err = s->d_err;
	// (the device knows the exact number of labels; `capacity` bounds it, and so does the length of the stream)
	const uint64_t stream_bits = l->g->dev.bit_end;
	const uint64_t bound = l->kind == WGBV_LABEL_FIXED_INT && l->width > 0 ? std::min<uint64_t>(capacity, stream_bits / (uint64_t)l->width) : capacity;
	launch_labels(s->launch(), lab, l->kind, l->width, from, to - from, d_rowptr, d_labels, capacity, bound);
	const cudaError_t e = cudaGetLastError();
	if (e != cudaSuccess) return fail(WGBV_ECUDA, "kernel launch failed: %s", cudaGetErrorString(e));
	return WGBV_OK;
}

extern "C" int wgbv_labels_decode_lists(wgbv_stream *s, const wgbv_labels *l, int64_t from, int64_t to, const int64_t *d_rowptr,
                                        int64_t *d_list_ptr, int64_t *d_values, uint64_t capacity, int64_t *values_out) {
	clear_error();
	if (!s || !l || !d_rowptr || !d_list_ptr) return fail(WGBV_EARG, "null argument");
	if (from < 0 || to < from || to > l->g->props.nodes) return fail(WGBV_EINDEX, "Node index out of range: [%lld, %lld)", (long long)from, (long long)to);
	if (l->g->device != s->g->device) return fail(WGBV_EARG, "the labels and the stream live on different devices");
	if (l->kind < WGBV_LABEL_FIXED_INT_LIST) return fail(WGBV_EUNSUPPORTED, "scalar labels are decoded by wgbv_labels_decode");
	CU(cudaSetDevice(s->g->device));
	GraphDev lab = l->g->dev;
	lab.err = s->d_err;
	const Launch L = s->launch();
	const int is_long = l->kind == WGBV_LABEL_FIXED_LONG_LIST;
	int rc;
	PeekArgs pk;
	pk.add(d_rowptr, 8);
	pk.add(d_rowptr + (to - from), 8);
	launch_peek(L, pk, s->d_pin);
	CU(cudaStreamSynchronize(s->stream));
	const int64_t arcs = (int64_t)s->h_pin[1] - (int64_t)s->h_pin[0];
	if (arcs < 0) return fail(WGBV_EARG, "row pointers are not increasing");
	if ((rc = s->scan_state.ensure((((size_t)arcs + 1) / 2048 + 2) * 8)) || (rc = s->ticket.ensure(64)) || (rc = s->c_arcs.ensure(((size_t)arcs + 1) * 8))) return rc;
	// lengths of all lists, their prefix sums, then the elements
	launch_label_lists(L, lab, false, l->width, is_long, from, to - from, d_rowptr, s->c_arcs.as<int64_t>(), nullptr, 0);
	launch_scan64(L, s->c_arcs.as<unsigned long long>(), d_list_ptr, arcs, s->scan_state.as<uint64_t>(), s->ticket.as<uint32_t>());
	PeekArgs pt;
	pt.add(d_list_ptr + arcs, 8);
	launch_peek(L, pt, s->d_pin);
	if ((rc = collect_errors(s))) return rc;
	const int64_t total = (int64_t)s->h_pin[0];
	if (values_out) *values_out = total;
	if (d_values) {
		if ((uint64_t)total > capacity) return fail(WGBV_EARG, "capacity (%llu) is smaller than the %lld label elements of the range", (unsigned long long)capacity, (long long)total);
		launch_label_lists(L, lab, true, l->width, is_long, from, to - from, d_rowptr, d_list_ptr, d_values, capacity);
	}
	const cudaError_t e = cudaGetLastError();
	if (e != cudaSuccess) return fail(WGBV_ECUDA, "kernel launch failed: %s", cudaGetErrorString(e));
	return WGBV_OK;
}

extern "C" int wgbv_labels_decode_host(wgbv_stream *s, const wgbv_labels *l, int64_t from, int64_t to, const int64_t *h_rowptr,
                                       int64_t *h_labels, uint64_t capacity) {
	clear_error();
	if (!s || !l || !h_rowptr) return fail(WGBV_EARG, "null argument");
	if (from < 0 || to < from || to > l->g->props.nodes) return fail(WGBV_EINDEX, "Node index out of range: [%lld, %lld)", (long long)from, (long long)to);
	const int64_t cnt = to - from, arcs = h_rowptr[cnt] - h_rowptr[0];
	if (arcs < 0) return fail(WGBV_EARG, "row pointers are not increasing");
	if ((uint64_t)arcs > capacity) return fail(WGBV_EARG, "capacity (%llu) is smaller than the %lld arcs of the range", (unsigned long long)capacity, (long long)arcs);
	if (arcs && !h_labels) return fail(WGBV_EARG, "null argument");
	CU(cudaSetDevice(s->g->device));
	int rc;
	if ((rc = s->rowptr.ensure((size_t)(cnt + 1) * 8)) || (rc = s->col.ensure((size_t)std::max<int64_t>(arcs, 1) * 8))) return rc;
	CU(cudaMemcpyAsync(s->rowptr.as<int64_t>(), h_rowptr, (size_t)(cnt + 1) * 8, cudaMemcpyHostToDevice, s->stream));
	if ((rc = wgbv_labels_decode(s, l, from, to, s->rowptr.as<int64_t>(), s->col.as<int64_t>(), (uint64_t)arcs))) return rc;
	if (arcs) CU(cudaMemcpyAsync(h_labels, s->col.as<int64_t>(), (size_t)arcs * 8, cudaMemcpyDeviceToHost, s->stream));
	return collect_errors(s);
}

extern "C" int wgbv_labels_decode_lists_host(wgbv_stream *s, const wgbv_labels *l, int64_t from, int64_t to, const int64_t *h_rowptr,
                                             int64_t *h_list_ptr, int64_t *h_values, uint64_t capacity, int64_t *values_out) {
	clear_error();
	if (!s || !l || !h_rowptr || !h_list_ptr) return fail(WGBV_EARG, "null argument");
	if (from < 0 || to < from || to > l->g->props.nodes) return fail(WGBV_EINDEX, "Node index out of range: [%lld, %lld)", (long long)from, (long long)to);
	const int64_t cnt = to - from, arcs = h_rowptr[cnt] - h_rowptr[0];
	if (arcs < 0) return fail(WGBV_EARG, "row pointers are not increasing");
	CU(cudaSetDevice(s->g->device));
	int rc;
	int64_t total = 0;
	if ((rc = s->rowptr.ensure((size_t)(cnt + 1) * 8)) || (rc = s->col.ensure((size_t)(arcs + 1) * 8))) return rc;
	CU(cudaMemcpyAsync(s->rowptr.as<int64_t>(), h_rowptr, (size_t)(cnt + 1) * 8, cudaMemcpyHostToDevice, s->stream));
	if ((rc = wgbv_labels_decode_lists(s, l, from, to, s->rowptr.as<int64_t>(), s->col.as<int64_t>(), nullptr, 0, &total))) return rc;
	if (values_out) *values_out = total;
	CU(cudaMemcpyAsync(h_list_ptr, s->col.as<int64_t>(), (size_t)(arcs + 1) * 8, cudaMemcpyDeviceToHost, s->stream));
	if (h_values) {
		if ((uint64_t)total > capacity) return fail(WGBV_EARG, "capacity (%llu) is smaller than the %lld label elements of the range", (unsigned long long)capacity, (long long)total);
		if ((rc = s->c_soff.ensure((size_t)std::max<int64_t>(total, 1) * 8))) return rc;
		GraphDev lab = l->g->dev;
		lab.err = s->d_err;
		launch_label_lists(s->launch(), lab, true, l->width, l->kind == WGBV_LABEL_FIXED_LONG_LIST, from, cnt, s->rowptr.as<int64_t>(), s->col.as<int64_t>(),
		                   s->c_soff.as<int64_t>(), (uint64_t)total);
		if (total) CU(cudaMemcpyAsync(h_values, s->c_soff.as<int64_t>(), (size_t)total * 8, cudaMemcpyDeviceToHost, s->stream));
	}
	return collect_errors(s);
}

// ------------------------------------------------------------------------------------------------ random access
static int batch_degrees(wgbv_stream *s, const int64_t *d_ids, uint64_t k, int64_t *d_rowptr, int64_t *arcs_out) {
	wgbv_graph *g = s->g;
	const Launch L = s->launch();
	int rc;
	const size_t c = (size_t)std::max<uint64_t>(k, 1);
	if (k >= 0x7ffffff0ull) return fail(WGBV_EARG, "a batch is limited to 2^31-16 queries");
	if ((rc = s->deg.ensure(c * 4)) || (rc = s->scan_state.ensure(((c + 1) / 2048 + 2) * 8)) || (rc = s->ticket.ensure(64))) return rc;
	launch_batch_degrees(L, s->dev, g->def_codes, d_ids, k, g->shard_lo, g->shard_hi, s->deg.as<uint32_t>());
	launch_scan(L, s->deg.as<uint32_t>(), d_rowptr, (int64_t)k, s->scan_state.as<uint64_t>(), s->ticket.as<uint32_t>());
	CU(cudaMemcpyAsync(s->h_pin, d_rowptr + k, 8, cudaMemcpyDeviceToHost, s->stream));
	if ((rc = collect_errors(s))) return rc;
	if (arcs_out) *arcs_out = (int64_t)s->h_pin[0];
	return WGBV_OK;
}

extern "C" int wgbv_successors_batch_degrees(wgbv_stream *s, const int64_t *d_ids, uint64_t k, int64_t *d_rowptr, int64_t *arcs_out) {
	clear_error();
	if (!s || (!d_ids && k) || !d_rowptr) return fail(WGBV_EARG, "null argument");
	CU(cudaSetDevice(s->g->device));
	return batch_degrees(s, d_ids, k, d_rowptr, arcs_out);
}

extern "C" int wgbv_successors_batch(wgbv_stream *s, const int64_t *d_ids, uint64_t k, int64_t *d_rowptr, int64_t *d_col,
                                     uint64_t col_capacity) {
	clear_error();
	if (!s || (!d_ids && k) || !d_rowptr || (!d_col && col_capacity)) return fail(WGBV_EARG, "null argument");
	CU(cudaSetDevice(s->g->device));
	wgbv_graph *g = s->g;
	const Launch L = s->launch();
	int rc;
	int64_t arcs = 0;
	if ((rc = batch_degrees(s, d_ids, k, d_rowptr, &arcs))) return rc;
	if ((uint64_t)arcs > col_capacity) return fail(WGBV_EARG, "col_capacity (%llu) is smaller than the %lld arcs of the batch", (unsigned long long)col_capacity, (long long)arcs);
	if (k == 0) return WGBV_OK;
	// chains
	const size_t c = (size_t)k;
	if ((rc = s->c_len.ensure(c * 4)) || (rc = s->c_arcs.ensure(c * 8)) || (rc = s->c_row0.ensure((c + 1) * 8)) || (rc = s->c_soff.ensure((c + 1) * 8))) return rc;
	CU(cudaMemsetAsync(s->d_stats, 0, 8, s->stream));
	launch_chain_walk(L, s->dev, g->def_codes, d_ids, k, g->avail_lo, s->c_len.as<uint32_t>(), s->c_arcs.as<unsigned long long>(), s->d_stats);
	launch_scan(L, s->c_len.as<uint32_t>(), s->c_row0.as<int64_t>(), (int64_t)k, s->scan_state.as<uint64_t>(), s->ticket.as<uint32_t>());
	launch_scan64(L, s->c_arcs.as<unsigned long long>(), s->c_soff.as<int64_t>(), (int64_t)k, s->scan_state.as<uint64_t>(), s->ticket.as<uint32_t>());
	uint32_t *hs = (uint32_t *)(s->h_pin + 4);
	CU(cudaMemcpyAsync(hs, s->d_stats, 8, cudaMemcpyDeviceToHost, s->stream));
	CU(cudaMemcpyAsync(s->h_pin, s->c_row0.as<int64_t>() + k, 8, cudaMemcpyDeviceToHost, s->stream));
	CU(cudaMemcpyAsync(s->h_pin + 1, s->c_soff.as<int64_t>() + k, 8, cudaMemcpyDeviceToHost, s->stream));
	if ((rc = collect_errors(s))) return rc;
	const uint64_t rows = s->h_pin[0], store = s->h_pin[1];
	const uint32_t max_len = hs[1];
	if (rows >= 0x7ffffff0ull) return fail(WGBV_EARG, "the batch expands to too many chain rows; split it");
	const size_t r = (size_t)std::max<uint64_t>(rows, 1);
	if ((rc = s->c_node.ensure(r * 8)) || (rc = s->c_dst.ensure(r * 8)) || (rc = s->c_parent.ensure(r * 4)) || (rc = s->c_deg.ensure(r * 4)) ||
	    (rc = s->c_isq.ensure(r)) || (rc = s->c_level.ensure(r * 4)) || (rc = s->c_info.ensure(r * sizeof(RowInfo))) || (rc = s->c_store.ensure((size_t)std::max<uint64_t>(store, 1) * 8)))
		return rc;
	launch_chain_fill(L, s->dev, g->def_codes, d_ids, k, s->c_row0.as<int64_t>(), s->c_soff.as<int64_t>(), d_rowptr, s->c_node.as<int64_t>(),
	                  s->c_dst.as<int64_t>(), s->c_parent.as<int32_t>(), s->c_deg.as<uint32_t>(), s->c_isq.as<uint8_t>(), s->c_level.as<uint32_t>());
	ChainDev C;
	C.node = s->c_node.as<int64_t>();
	C.dst_off = s->c_dst.as<int64_t>();
	C.parent = s->c_parent.as<int32_t>();
	C.deg = s->c_deg.as<uint32_t>();
	C.is_query = s->c_isq.as<uint8_t>();
	C.col = d_col;
	C.store = s->c_store.as<int64_t>();
	C.info = s->c_info.as<RowInfo>();
	launch_chain_extras(L, s->dev, g->def_codes, C, rows);
	for (uint32_t lev = 1; lev < max_len; lev++) launch_chain_copy(L, s->dev, g->def_codes, C, rows, s->c_level.as<uint32_t>(), lev);
	const cudaError_t e = cudaGetLastError();
	if (e != cudaSuccess) return fail(WGBV_ECUDA, "kernel launch failed: %s", cudaGetErrorString(e));
	return WGBV_OK;
}

extern "C" int wgbv_successors_batch_host(wgbv_stream *s, const int64_t *h_ids, uint64_t k, int64_t *h_rowptr, int64_t *h_col,
                                          uint64_t col_capacity, int64_t *arcs_out) {
	clear_error();
	if (!s || (!h_ids && k) || !h_rowptr) return fail(WGBV_EARG, "null argument");
	CU(cudaSetDevice(s->g->device));
	// BVG:863: an id outside [0, n) is an error before anything is decoded
	for (uint64_t i = 0; i < k; i++)
		if (h_ids[i] < 0 || h_ids[i] >= s->g->props.nodes) return fail(WGBV_EINDEX, "Node index out of range: %lld", (long long)h_ids[i]);
	int rc;
	if ((rc = s->ids.ensure((size_t)std::max<uint64_t>(k, 1) * 8)) || (rc = s->rowptr.ensure((size_t)(k + 1) * 8))) return rc;
	if (k) CU(cudaMemcpyAsync(s->ids.as<int64_t>(), h_ids, (size_t)k * 8, cudaMemcpyHostToDevice, s->stream));
	int64_t arcs = 0;
	if ((rc = batch_degrees(s, s->ids.as<int64_t>(), k, s->rowptr.as<int64_t>(), &arcs))) return rc;
	if (arcs_out) *arcs_out = arcs;
	CU(cudaMemcpyAsync(h_rowptr, s->rowptr.as<int64_t>(), (size_t)(k + 1) * 8, cudaMemcpyDeviceToHost, s->stream));
	if (h_col) {
		if ((uint64_t)arcs > col_capacity) return fail(WGBV_EARG, "col_capacity (%llu) is smaller than the %lld arcs of the batch", (unsigned long long)col_capacity, (long long)arcs);
		if ((rc = s->col.ensure((size_t)std::max<int64_t>(arcs, 1) * 8))) return rc;
		if ((rc = wgbv_successors_batch(s, s->ids.as<int64_t>(), k, s->rowptr.as<int64_t>(), s->col.as<int64_t>(), (uint64_t)arcs))) return rc;
		if (arcs) CU(cudaMemcpyAsync(h_col, s->col.as<int64_t>(), (size_t)arcs * 8, cudaMemcpyDeviceToHost, s->stream));
	}
	return collect_errors(s);
}

extern "C" int wgbv_outdegrees(wgbv_stream *s, const int64_t *h_ids, uint64_t k, int64_t *h_out) {
	clear_error();
	if (!s || (!h_ids && k) || (!h_out && k)) return fail(WGBV_EARG, "null argument");
	CU(cudaSetDevice(s->g->device));
	for (uint64_t i = 0; i < k; i++)
		if (h_ids[i] < 0 || h_ids[i] >= s->g->props.nodes) return fail(WGBV_EINDEX, "Node index out of range: %lld", (long long)h_ids[i]);
	if (k == 0) return WGBV_OK;
	int rc;
	if ((rc = s->ids.ensure((size_t)k * 8)) || (rc = s->deg.ensure((size_t)k * 4))) return rc;
	CU(cudaMemcpyAsync(s->ids.as<int64_t>(), h_ids, (size_t)k * 8, cudaMemcpyHostToDevice, s->stream));
	launch_batch_degrees(s->launch(), s->dev, s->g->def_codes, s->ids.as<int64_t>(), k, s->g->shard_lo, s->g->shard_hi, s->deg.as<uint32_t>());
	std::vector<uint32_t> tmp((size_t)k);
	CU(cudaMemcpyAsync(tmp.data(), s->deg.as<uint32_t>(), (size_t)k * 4, cudaMemcpyDeviceToHost, s->stream));
	if ((rc = collect_errors(s))) return rc;
	for (uint64_t i = 0; i < k; i++) h_out[i] = tmp[(size_t)i];
	return WGBV_OK;
}

// ------------------------------------------------------------------------------------------------ checksum
extern "C" int wgbv_csr_checksum(wgbv_stream *s, int64_t first_node, int64_t rows, const int64_t *d_rowptr, const int64_t *d_col,
                                 uint64_t *checksum_out) {
	clear_error();
	if (!s || !d_rowptr || !checksum_out) return fail(WGBV_EARG, "null argument");
	CU(cudaSetDevice(s->g->device));
	int rc;
	if ((rc = s->ticket.ensure(64))) return rc;
	unsigned long long *acc = s->ticket.as<unsigned long long>() + 1;
	CU(cudaMemsetAsync(acc, 0, 8, s->stream));
	launch_checksum(s->launch(), first_node, rows, d_rowptr, d_col, acc);
	CU(cudaMemcpyAsync(s->h_pin + 2, acc, 8, cudaMemcpyDeviceToHost, s->stream));
	CU(cudaStreamSynchronize(s->stream));
	const cudaError_t e = cudaGetLastError();
	if (e != cudaSuccess) return fail(WGBV_ECUDA, "CUDA error: %s", cudaGetErrorString(e));
	*checksum_out = s->h_pin[2];
	return WGBV_OK;
}
