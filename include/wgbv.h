/*
 * wgbv.h -- C ABI of libwgbv, a B200 (sm_100a) BVGraph decompression engine.
 *
 * This is the drop-in boundary behind webgraph-big's ImmutableGraph API for the BVGraph -> CSR decode path:
 * every entry point names the reference interface it stands in for.  "BVG" is
 * src/it/unimi/dsi/big/webgraph/BVGraph.java and "IG" is src/it/unimi/dsi/big/webgraph/ImmutableGraph.java of
 * webgraph-big 3.7.1.  Plain C: opaque handles, plain pointers and sizes, integer status codes -- callable from
 * Java (Panama FFM or JNI), Python ctypes, C and C++ (see INTEGRATION.md for the Java binding).
 *
 * There is no CPU fallback: every decode entry point runs hand-written CUDA kernels and fails with WGBV_ECUDA
 * when no usable device is present.
 *
 * Conventions
 *   - Every function that can fail returns 0 or one of the negative codes below; wgbv_last_error() then gives a
 *     thread-local, human-readable message.  The codes map one to one onto the exceptions of the reference.
 *   - A wgbv_graph is immutable after opening and may be shared by any number of threads (it is what a
 *     BVGraph and all its copy()s share: the bit stream and the Elias-Fano offsets, BVG:553-578).
 *   - A wgbv_stream holds the mutable per-user state (one CUDA stream plus scratch memory): the counterpart of
 *     one flyweight copy() (IG:187-197).  One stream must not be used by two threads at once.
 *   - Node ids are 64-bit and global (also when the handle only holds a shard of the graph).
 *   - "device" pointers are CUDA device pointers valid on the graph's device; "host" pointers are ordinary
 *     (preferably pinned) host memory.
 */
#ifndef WGBV_H
#define WGBV_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define WGBV_API __attribute__((visibility("default")))
#else
#define WGBV_API
#endif

/* Status codes (reference exception in brackets). */
#define WGBV_OK 0
#define WGBV_EINDEX (-1)       /* IllegalArgumentException "Node index out of range" (BVG:823, 863, 1000, 1128) */
#define WGBV_ESTATE (-2)       /* IllegalStateException: reference > window size (BVG:701), malformed stream */
#define WGBV_EIO (-3)          /* IOException: missing file, wrong graphclass, missing/too-new version (BVG:1491-1497) */
#define WGBV_EUNSUPPORTED (-4) /* UnsupportedOperationException: unknown coding id (BVG:631, 658, 699, 733, 763, 794) */
#define WGBV_ECUDA (-5)        /* no usable device, CUDA failure, out of device memory */
#define WGBV_EARG (-6)         /* IllegalArgumentException for any other bad argument (null pointer, capacity too small) */

/* Coding ids of the compression flags (CompressionFlags.java:26-46). */
#define WGBV_DELTA 1
#define WGBV_GAMMA 2
#define WGBV_GOLOMB 3
#define WGBV_SKEWED_GOLOMB 4
#define WGBV_UNARY 5
#define WGBV_ZETA 6
#define WGBV_NIBBLE 7

/* The decoded content of a .properties file (the fields BVGraph.loadInternal reads, BVG:1495-1503, with the
 * compression flags already unpacked as setFlags does, BVG:1281-1289). */
typedef struct wgbv_props {
	int64_t nodes;
	int64_t arcs;
	int32_t window_size;
	int32_t max_ref_count;
	int32_t min_interval_length;
	int32_t zeta_k;
	int32_t outdegree_coding;   /* GAMMA (default) | DELTA */
	int32_t block_coding;       /* GAMMA (default) | DELTA | UNARY */
	int32_t residual_coding;    /* ZETA (default) | GAMMA | DELTA | GOLOMB | NIBBLE */
	int32_t reference_coding;   /* UNARY (default) | GAMMA | DELTA */
	int32_t block_count_coding; /* GAMMA (default) | DELTA | UNARY */
	int32_t offset_coding;      /* GAMMA (default) | DELTA */
} wgbv_props;

typedef struct wgbv_graph wgbv_graph;
typedef struct wgbv_stream wgbv_stream;

/* Thread-local message of the last failure on the calling thread ("" if none). */
WGBV_API const char *wgbv_last_error(void);

/* Parses the text of a .properties file into *out, applying the checks of BVGraph.loadInternal (BVG:1482-1503:
 * graphclass must name BVGraph, standard or big; version present and <= 0) and string2Flags/setFlags
 * (BVG:1316-1331, 1281-1289).  Host only. */
WGBV_API int wgbv_parse_properties(const char *text, wgbv_props *out);

/*
 * Opening.  Replaces BVGraph.load(basename) (BVG:1345-1383 -> loadInternal BVG:1479-1574): reads
 * <basename>.properties, <basename>.graph and <basename>.offsets, turns the gamma/delta coded offsets
 * (OffsetsLongIterator, BVG:870-898) into an Elias-Fano list in device memory (the role of
 * EliasFanoMonotoneLongBigList at BVG:1558) and uploads the bit stream.
 *
 * shard_index / shard_count select a contiguous, bit-offset-balanced shard of the node range (shard_count = 1:
 * the whole graph).  A shard also keeps the window_size * max_ref_count predecessors of its first node (the
 * GPU analogue of the window pre-fill of nodeIterator(from), BVG:1139-1143) so that decoding never needs
 * another device.  Graphs whose max_ref_count makes that halo unbounded can only be opened with shard_count 1.
 */
WGBV_API int wgbv_open(const char *basename, int device, int shard_index, int shard_count, wgbv_graph **out);

/* Same, from memory: `graph` and `offsets` are the bytes of the .graph and .offsets files. */
WGBV_API int wgbv_open_mem(const wgbv_props *props, const uint8_t *graph, uint64_t graph_bytes, const uint8_t *offsets,
                  uint64_t offsets_bytes, int device, int shard_index, int shard_count, wgbv_graph **out);

/*
 * BVGraph.writeOffsets (BVG:2595-2609; the `-o` option of the reference's BVGraph main, BVG:2654ff): the .offsets stream of
 * a graph that was shipped without one -- nodes + 1 gamma (or delta, per the properties) codes, the first one 0, then the
 * length in bits of every record.  Host code: a record can only be found by parsing every record before it (how many
 * codes a record has depends on the outdegree of the row it refers to), which is why the reference, too, walks the file
 * with a single node iterator.  One pass, lengths only.  *out_bytes receives the size of the stream; nothing is copied
 * (WGBV_EARG) if out_capacity is smaller -- call with out == NULL to learn the size.  wgbv_open_mem accepts
 * offsets == NULL and then does the same pass itself.
 */
WGBV_API int wgbv_write_offsets(const wgbv_props *props, const uint8_t *graph, uint64_t graph_bytes, uint8_t *out, uint64_t out_capacity,
                                uint64_t *out_bytes);

/* The shard plan wgbv_open uses, without touching a device (host only): bounds_out[k] is the first node of shard k,
 * bounds_out[shard_count] = nodes; shards are contiguous and balanced by compressed bits (the arc-balancing idea of
 * algo/EliasFanoCumulativeOutdegreeList.java:139-151 applied to bit offsets).  *halo_out = window_size *
 * max_ref_count, the number of predecessor nodes every shard keeps (cf. the window pre-fill, BVG:1139-1143). */
WGBV_API int wgbv_plan_shards(const wgbv_props *props, const uint8_t *offsets, uint64_t offsets_bytes, int shard_count,
                              int64_t *bounds_out, int64_t *halo_out);

/*
 * Opening from a pre-cut slice, for callers that read only their own byte range of a large file (one process per
 * GPU, each holding one shard).  `bits` is a piece of the .graph bit stream and `offsets` a gamma/delta coded stream
 * of node_count + 1 differences in the .offsets format (BVG:870-898) whose running sums are the bit positions, inside
 * `bits`, of the records of nodes first_node .. first_node + node_count (the last one being the end of the slice).
 * The first halo_nodes nodes of the slice are only kept as reference material: the handle decodes
 * [first_node + halo_nodes, first_node + node_count).  props describes the whole graph (props->nodes is global).
 * The loader counterpart is the same loadInternal (BVG:1479-1574) restricted to a node range.
 */
WGBV_API int wgbv_open_slice(const wgbv_props *props, int64_t first_node, int64_t node_count, int64_t halo_nodes,
                             const uint8_t *bits, uint64_t nbytes, const uint8_t *offsets, uint64_t offsets_bytes, int device,
                             wgbv_graph **out);

/* Releases the device memory of a graph.  All its streams must have been destroyed. */
WGBV_API void wgbv_close(wgbv_graph *g);

/* numNodes() IG:245, numArcs() IG:263 (the value stored in .properties), and the properties themselves. */
WGBV_API int64_t wgbv_num_nodes(const wgbv_graph *g);
WGBV_API int64_t wgbv_num_arcs(const wgbv_graph *g);
WGBV_API int wgbv_get_props(const wgbv_graph *g, wgbv_props *out);

/* The node range [*first, *last) this handle can decode (the shard; the whole graph if unsharded), the bytes of
 * device memory the handle holds, and the bytes the Elias-Fano offsets take (for bits-per-pointer reports, cf.
 * BVG:1565-1566). */
WGBV_API int wgbv_shard_range(const wgbv_graph *g, int64_t *first, int64_t *last);
WGBV_API uint64_t wgbv_device_bytes(const wgbv_graph *g);
WGBV_API uint64_t wgbv_offsets_bytes(const wgbv_graph *g);

/* Per-user state: the counterpart of ImmutableGraph.copy() (IG:447, BVG:553-578).  `cuda_stream` is a
 * cudaStream_t to run on (e.g. torch's current stream) or NULL for a stream owned by the library. */
WGBV_API int wgbv_stream_create(wgbv_graph *g, void *cuda_stream, wgbv_stream **out);
WGBV_API void wgbv_stream_destroy(wgbv_stream *s);
/* Blocks until everything queued on the stream has finished; reports errors found by the kernels. */
WGBV_API int wgbv_stream_sync(wgbv_stream *s);

/* Copies the bit stream from host memory to the device again, asynchronously on the stream (the upload half of
 * loadInternal's "read the whole graph into memory", BVG:1520-1528, for callers that keep the compressed graph in
 * pinned host memory and stream it in per pass).  `graph` must be the buffer (or an identical copy) the handle was
 * opened from.  The words are rewritten in place, so this (and wgbv_scan_begin_upload) is only allowed while the
 * stream is the only one of its graph: WGBV_ESTATE otherwise. */
WGBV_API int wgbv_upload_stream(wgbv_stream *s, const uint8_t *graph, uint64_t graph_bytes);

/*
 * Bulk decode of nodes [from, to) to CSR -- the new bulk call of the GPU-backed graph.  It produces what a full
 * sequential scan produces (nodeIterator(from) + nextLong()/outdegree()/successorBigArray() until `to`,
 * BVG:1100-1245; per node BVGraph.successors(x, ibs, window, outd), BVG:995-1097): for every node its
 * outdegree and its successors in increasing order.
 *
 * Two steps, so that the caller can size the column array:
 *   wgbv_decode_degrees  header pass + device prefix scan: d_rowptr[i] = number of arcs of nodes [from, from+i),
 *                        i = 0..to-from (device, int64), *arcs_out = d_rowptr[to-from] (host; forces a sync).
 *                        Also what outdegree(x) (BVG:821-842) returns for every x in the range:
 *                        outdegree(from+i) = d_rowptr[i+1] - d_rowptr[i].
 *   wgbv_decode_csr      the complete decode; d_col must hold at least the arcs of [from, to) (col_capacity in
 *                        elements); writes d_rowptr (to-from+1 int64) and d_col (int64 successors).  Asynchronous
 *                        on the stream: call wgbv_stream_sync before reading and to collect decode errors.
 */
WGBV_API int wgbv_decode_degrees(wgbv_stream *s, int64_t from, int64_t to, int64_t *d_rowptr, int64_t *arcs_out);
WGBV_API int wgbv_decode_csr(wgbv_stream *s, int64_t from, int64_t to, int64_t *d_rowptr, int64_t *d_col, uint64_t col_capacity);

/* The same with HOST buffers (pinned memory recommended): uploads nothing but the request, decodes on the device and
 * copies row pointers and successors back.  h_col may be NULL to only count (then *arcs_out is still set).
 * Synchronous.  This is the call a host-side ImmutableGraph implementation sits on. */
WGBV_API int wgbv_decode_csr_host(wgbv_stream *s, int64_t from, int64_t to, int64_t *h_rowptr, int64_t *h_col,
                         uint64_t col_capacity, int64_t *arcs_out);

/*
 * Streaming scan of [from, to): the bulk form of nodeIterator(from) (BVG:1100-1245) for host-side consumers.  The
 * library decodes chunk_nodes nodes at a time on the device and copies each chunk back into pinned host buffers it
 * owns, keeping the next chunks decoding and copying while the caller works on the current one.
 *   wgbv_scan_begin   starts a scan (one scan per wgbv_stream at a time)
 *   wgbv_scan_next    returns 1 and the next chunk -- nodes [*first, *last), *h_rowptr[0 .. last-first] (relative to the
 *                     chunk: h_rowptr[0] = 0) and *h_col[0 .. *arcs) -- 0 when the range is exhausted, or a negative
 *                     status.  The buffers stay valid until the next wgbv_scan_next / wgbv_scan_end on this stream:
 *                     the ownership rule of the reference's iterator, whose arrays are valid until the next
 *                     nextLong() (BVG:1171-1172).
 *   wgbv_scan_end     waits for outstanding work and ends the scan.
 */
WGBV_API int wgbv_scan_begin(wgbv_stream *s, int64_t from, int64_t to, int64_t chunk_nodes);
/* The same scan for callers that keep the compressed graph in (pinned) host memory and stream it in for every pass:
 * the bit stream is uploaded piece by piece while the scan runs -- the words of chunk k+1 travel to the device while
 * chunk k is decoded and chunk k-1 is copied back -- instead of in one wgbv_upload_stream before it.  `graph` must be
 * the buffer (or an identical copy) the handle was opened from; `from` must be the first node of the handle. */
WGBV_API int wgbv_scan_begin_upload(wgbv_stream *s, const uint8_t *graph, uint64_t graph_bytes, int64_t from, int64_t to, int64_t chunk_nodes);
WGBV_API int wgbv_scan_next(wgbv_stream *s, int64_t *first, int64_t *last, const int64_t **h_rowptr, const int64_t **h_col, int64_t *arcs);
WGBV_API void wgbv_scan_end(wgbv_stream *s);
/* Narrow delivery for the scans of this stream (no counterpart in the reference, whose successor lists are long[]):
 * with on != 0 the chunks' successors reach the host as uint32 -- *h_col of wgbv_scan_next then points to
 * uint32_t[*arcs] -- which halves the bytes that cross PCIe, the bound of a host-side scan.  Only for graphs whose
 * node ids fit (wgbv_num_nodes <= 2^32: WGBV_EUNSUPPORTED otherwise); row pointers stay int64.  The values are the
 * same; a Java binding widens them when it fills its long arrays.  Not while a scan is running (WGBV_ESTATE). */
WGBV_API int wgbv_scan_set_narrow(wgbv_stream *s, int on);

/*
 * Transpose of a CSR in device memory: the device form of what Transform.transposeOffline
 * (Transform.java:1098-1144) computes from a sequential scan -- for every node y the increasing list of the x with
 * y in successors(x).  d_rowptr[0 .. n] / d_col describe a graph with n nodes whose successors are node ids of the same
 * graph (the output of wgbv_decode_csr over [0, n)); d_t_rowptr gets n + 1 entries and d_t_col as many as d_col has
 * (d_rowptr[n] - d_rowptr[0]).  All four are device pointers; the call is asynchronous on the stream except for one
 * read-back of the arc count.  n < 2^32 (WGBV_EUNSUPPORTED beyond); a successor outside [0, n) gives WGBV_EINDEX.
 */
WGBV_API int wgbv_transpose_csr(wgbv_stream *s, int64_t n, const int64_t *d_rowptr, const int64_t *d_col, int64_t *d_t_rowptr, int64_t *d_t_col);

/*
 * Union of two CSRs over the same n nodes, in device memory: for every node the increasing union of its two successor
 * lists, duplicates dropped -- Transform.union (Transform.java:1659-1663, UnionImmutableGraph) of the reference.  d_u_rowptr gets n + 1 entries,
 * d_u_col the successors (u_capacity entries available; pass d_u_col = NULL to learn the size only); *arcs_out, if not
 * NULL, the number of arcs of the union.  WGBV_EARG if u_capacity is too small (*arcs_out says how much is needed).
 */
WGBV_API int wgbv_union_csr(wgbv_stream *s, int64_t n, const int64_t *d_a_rowptr, const int64_t *d_a_col, const int64_t *d_b_rowptr,
                            const int64_t *d_b_col, int64_t *d_u_rowptr, int64_t *d_u_col, uint64_t u_capacity, int64_t *arcs_out);

/*
 * Symmetrization of a CSR in device memory: the union of the graph and its transpose, Transform.symmetrizeOffline
 * (Transform.java:573-574: "return union(g, transposeOffline(g, ...))").  The transpose is built in scratch memory of the
 * stream (as many entries as the graph has arcs).  Arguments as for wgbv_union_csr; n < 2^32 as for wgbv_transpose_csr.
 */
WGBV_API int wgbv_symmetrize_csr(wgbv_stream *s, int64_t n, const int64_t *d_rowptr, const int64_t *d_col, int64_t *d_s_rowptr, int64_t *d_s_col,
                                 uint64_t s_capacity, int64_t *arcs_out);

/*
 * BVGraph compression of a CSR in device memory: what BVGraph.store / storeInternal write to the .graph
 * file (BVGraph.java:2163-2327; per node: the candidate loop over the window with refCount < maxRefCount, BVG:2249-2270,
 * diffComp BVG:1977-2159, intervalize BVG:1595-1618).  props gives windowsize, maxrefcount, minintervallength, zetak and
 * the codings (nodes / arcs are not read); d_rowptr[0 .. n] / d_col are device pointers to the successor lists (sorted,
 * without duplicates) of nodes first_node .. first_node + n - 1.  split_nodes = 0 compresses in one run, exactly as a
 * single-threaded store; split_nodes = k > 0 starts every k nodes with an empty window, as the threads of the reference's
 * parallel compressor do (BVG:2408-2457) -- the streams of the splits are already concatenated bit-exactly.
 * Outputs (host memory): graph_out[0 .. ceil(*graph_bits / 8)) the bytes of the .graph file (capacity in bytes:
 * WGBV_EARG with the size needed in *graph_bits if it is too small); node_bits_out, if not NULL, the n record lengths
 * in bits, whose gamma-coded differences are the .offsets file (BVG:2228, 2311).
 */
WGBV_API int wgbv_encode_csr(int device, const wgbv_props *props, int64_t first_node, int64_t n, const int64_t *d_rowptr, const int64_t *d_col,
                             int64_t split_nodes, uint8_t *graph_out, uint64_t graph_capacity, uint64_t *graph_bits, uint64_t *node_bits_out);

/*
 * Arc labels: the `.labels` / `.labeloffsets` pair of a BitStreamArcLabelledImmutableGraph
 * (labelling/BitStreamArcLabelledImmutableGraph.java:65-100; offsets :339-363; what its node iterator reads per arc
 * :487-560) decoded into an int64 array parallel to the CSR successors of the underlying graph.  kind selects the
 * label prototype: WGBV_LABEL_FIXED_INT = FixedWidthIntLabel(key, width), width 0..32 (FixedWidthIntLabel.java:70-73);
 * WGBV_LABEL_GAMMA_INT = GammaCodedIntLabel(key) (GammaCodedIntLabel.java:60-64).  `nodes` is the node count of the
 * underlying graph; labeloffsets holds nodes + 1 gamma-coded differences.
 *   wgbv_labels_decode       d_rowptr[0 .. to-from] (device; what wgbv_decode_csr / wgbv_decode_degrees wrote for the
 *                            same range) gives every node's outdegree and the place of its labels in d_labels
 *                            (device, `capacity` elements).  Asynchronous on the stream; errors at wgbv_stream_sync.
 *   wgbv_labels_decode_host  the same with host arrays (synchronous).
 * A stream, offsets and outdegrees that do not belong together give WGBV_ESTATE (the reference would return garbage).
 * List-valued prototypes -- WGBV_LABEL_FIXED_INT_LIST = FixedWidthIntListLabel(key, width) (width 0..32; per arc
 * readGamma() elements of readInt(width) each, FixedWidthIntListLabel.java:73-78) and WGBV_LABEL_FIXED_LONG_LIST =
 * FixedWidthLongListLabel(key, width) (width 0..64, FixedWidthLongListLabel.java:78-84) -- have a ragged result:
 *   wgbv_labels_decode_lists d_list_ptr (device, arcs + 1 entries) receives the start of every arc's list inside d_values
 *                            (device, `capacity` elements; NULL: only the starts and *values_out, the total, are produced).
 *                            Synchronises the stream once to learn the total.  wgbv_labels_decode refuses these kinds.
 *   wgbv_labels_decode_lists_host  the same with host arrays; call it with h_values NULL to learn *values_out, then again
 *                            with room for that many elements (or once with a capacity known to suffice).
 */
typedef struct wgbv_labels wgbv_labels;
enum { WGBV_LABEL_FIXED_INT = 0, WGBV_LABEL_GAMMA_INT = 1, WGBV_LABEL_FIXED_INT_LIST = 2, WGBV_LABEL_FIXED_LONG_LIST = 3 };
WGBV_API int wgbv_labels_open_mem(int device, int64_t nodes, int kind, int width, const uint8_t *labels, uint64_t labels_bytes,
                                  const uint8_t *labeloffsets, uint64_t labeloffsets_bytes, wgbv_labels **out);
WGBV_API void wgbv_labels_close(wgbv_labels *l);
WGBV_API int wgbv_labels_decode(wgbv_stream *s, const wgbv_labels *l, int64_t from, int64_t to, const int64_t *d_rowptr,
                                int64_t *d_labels, uint64_t capacity);
WGBV_API int wgbv_labels_decode_host(wgbv_stream *s, const wgbv_labels *l, int64_t from, int64_t to, const int64_t *h_rowptr,
                                     int64_t *h_labels, uint64_t capacity);
WGBV_API int wgbv_labels_decode_lists(wgbv_stream *s, const wgbv_labels *l, int64_t from, int64_t to, const int64_t *d_rowptr,
                                      int64_t *d_list_ptr, int64_t *d_values, uint64_t capacity, int64_t *values_out);
WGBV_API int wgbv_labels_decode_lists_host(wgbv_stream *s, const wgbv_labels *l, int64_t from, int64_t to, const int64_t *h_rowptr,
                                           int64_t *h_list_ptr, int64_t *h_values, uint64_t capacity, int64_t *values_out);

/* outdegree(x) for a batch of arbitrary nodes (BVG:821-851): h_out[i] = outdegree(h_ids[i]).  Host buffers. */
WGBV_API int wgbv_outdegrees(wgbv_stream *s, const int64_t *h_ids, uint64_t k, int64_t *h_out);

/*
 * Batched random access: successors(x) (BVG:860-867, with the recursive reference resolution of BVG:1084) for k
 * arbitrary nodes.  Two steps like the bulk call:
 *   wgbv_successors_batch_degrees  d_rowptr[i] = arcs of queries 0..i-1 (k+1 int64, device), *arcs_out on host
 *   wgbv_successors_batch          fills d_col (capacity in elements) with the successor lists, query after query
 * d_ids is a device array of k node ids.  Any id outside [0, numNodes) fails with WGBV_EINDEX (BVG:863) before
 * anything is decoded; ids outside a shard's range likewise.
 */
WGBV_API int wgbv_successors_batch_degrees(wgbv_stream *s, const int64_t *d_ids, uint64_t k, int64_t *d_rowptr, int64_t *arcs_out);
WGBV_API int wgbv_successors_batch(wgbv_stream *s, const int64_t *d_ids, uint64_t k, int64_t *d_rowptr, int64_t *d_col,
                          uint64_t col_capacity);
/* Host-buffer form (ids, row pointers and successors in host memory).  h_col may be NULL to only count. */
WGBV_API int wgbv_successors_batch_host(wgbv_stream *s, const int64_t *h_ids, uint64_t k, int64_t *h_rowptr, int64_t *h_col,
                               uint64_t col_capacity, int64_t *arcs_out);

/* Order-sensitive checksum of a CSR held in device memory: sum over the rows of a 64-bit hash of
 * (node id, outdegree, successors in order).  Equal CSRs give equal checksums whatever the sharding; used to
 * verify decodes that are too large to compare element by element (the role of ImmutableGraph.equals, IG:759-779,
 * at scale).  first_node is the id of row 0.  Synchronous. */
WGBV_API int wgbv_csr_checksum(wgbv_stream *s, int64_t first_node, int64_t rows, const int64_t *d_rowptr, const int64_t *d_col,
                      uint64_t *checksum_out);

/*
 * Row pointers from outdegrees with NO allocation.  (This entry point has no counterpart in the reference: the
 * Java code never has to think about where a prefix sum keeps its scratch.)
 *
 * d_rowptr[i] = d_degrees[0] + ... + d_degrees[i-1] for i = 0..count (count + 1 int64 values), computed on the
 * CURRENT device by the same single-pass prefix-sum kernel design as wgbv_decode_degrees, except that its
 * tile-status table lives in a file-scope __device__ array inside the library instead of cudaMalloc'ed memory.
 * It is meant for callers that own their input and output buffers but cannot call the allocator when they need the
 * row pointers: processes whose device memory is entirely held by another component's pool, and latency-critical
 * threads that must not trigger the device-wide synchronisation cudaMalloc / cudaFree may cause.  The function
 * calls neither of them, nor any library routine that allocates.
 *
 *   - Limit: count <= WGBV_STATIC_SCAN_MAX_COUNT (about 134 million degrees); larger inputs fail with WGBV_EARG.
 *     The static array is 8 * (WGBV_STATIC_SCAN_TILES + 32) bytes (about 513 KiB); every GPU on which the library
 *     is loaded keeps it until the process exits.
 *   - The first call on each device looks up the device address of that array (cudaGetSymbolAddress); make this
 *     first call before starting any stream capture on that device.
 *   - All calls on a device share the array, which keeps the contents of the previous call; every call clears the
 *     words it uses first.  Calls whose kernels could overlap on the GPU (different streams without ordering) race
 *     on it: order them.
 *   - d_degrees and d_rowptr are device pointers on the current device; the work is queued on cuda_stream
 *     (a cudaStream_t, NULL for the default stream) and is asynchronous.
 */
#define WGBV_STATIC_SCAN_TILES 65536
#define WGBV_STATIC_SCAN_MAX_COUNT ((int64_t)WGBV_STATIC_SCAN_TILES * 2048 - 1)
WGBV_API int wgbv_scan_degrees_static(const uint32_t *d_degrees, int64_t count, int64_t *d_rowptr, void *cuda_stream);

/* Timing of the stages of the last wgbv_decode_csr on this stream, in milliseconds (CUDA events on the stream):
 * out[0] header pass, [1] scan, [2] depth levels, [3] successor decode, [4] total.  Synchronises the stream. */
WGBV_API int wgbv_last_stage_ms(wgbv_stream *s, float out[5]);

/* Number of kernels the library has launched on this stream since it was created (for launch accounting). */
WGBV_API uint64_t wgbv_kernel_launches(const wgbv_stream *s);

/* Accounting of the one-pass tile kernel (csrc/bv_fused.cu; no counterpart in the reference): out[0] = tiles of the last
 * wgbv_decode_csr that took it, out[1] = how many of them did not fit shared memory and were decoded by the slow-tile
 * kernel (valid after wgbv_stream_sync), out[2] = decodes of this stream that had to be rerun on the level-synchronous
 * path, out[3] = 1 if decodes of this handle take the tile kernel (WGBV_FUSED=1 in the environment, bounded reference
 * chains, node ids below 2^32 - 1). */
WGBV_API int wgbv_fused_stats(wgbv_stream *s, uint64_t out[4]);

/* Algorithmic bytes of decoding [from, to) (compressed bits in + Elias-Fano bits touched + int64 row pointers
 * out + int64 successors out), given the arc count of the range. */
WGBV_API uint64_t wgbv_algorithmic_bytes(const wgbv_graph *g, int64_t from, int64_t to, int64_t arcs);

#ifdef __cplusplus
}
#endif
#endif /* WGBV_H */
