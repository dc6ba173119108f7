// bv_kernels.cuh -- launch interface between the C ABI (wgbv_api.cu) and the kernels (bv_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "bv_device.cuh"

namespace wgbv {

enum : uint32_t {
	DEPTH_UNRESOLVED = 0xffffffffu,  // chain not followed to its root yet
	DEPTH_INVALID = 0xfffffffeu,     // chain leaves the decodable range (halo rows only)
};

// What stage 3a learns about a row with a reference and stage 3b needs again: kept so that the copy kernels do not
// have to look the row's offset up and parse its head a second time.  Two shapes, told apart by the outdegrees of the
// row and of its reference row (row_is_masked), which every kernel has at hand:
//   masked rows: a = bits 0..63, b = bits 64..127 of the mask of copied positions of the reference row
//   other rows:  a = bit position of the first block, b = ncopied | block count << 32
struct RowInfo {
	uint64_t a, b;
};
// Rows of fewer than WARP_ROW successors whose reference row has at most 128 are merged by one lane from the mask of
// copied positions (row_copy_masked); all the others by a whole warp from the block codes (row_copy_warp).
static constexpr uint32_t WARP_ROW = 128;
__host__ __device__ inline bool row_is_masked(uint32_t d, uint32_t dref) { return d < WARP_ROW && dref <= 128; }

// One bulk decode job: nodes [lo, hi) are decoded, of which [from, hi) were requested; [lo, from) is the halo.
struct RangeDev {
	int64_t lo, from, hi;
	const uint32_t *deg;    // outdegrees of [lo, hi)
	const uint8_t *ref;     // references of [lo, hi) (0: none)
	const uint32_t *depth;  // reference-chain depth of [lo, hi)
	const int64_t *rp;      // exclusive prefix sum of deg, hi-lo+1 entries, rp[0] = 0
	int64_t *col;           // successors of [from, hi): row x starts at col + rp[x-lo] - rp[from-lo]
	int64_t *halo;          // successors of [lo, from): row x starts at halo + rp[x-lo]
	uint64_t col_capacity;  // elements available at col
	RowInfo *info;          // written by k_extras for every row with a reference, read by k_copy_level (may be nullptr)
	// One-launch front end: k_extras itself reads outdegree and reference of its rows (stage 1),
	// scans the outdegrees across tiles by decoupled look-back (stage 2) and writes deg_w / ref_w / rp_w -- the very
	// arrays deg / ref / rp point to -- before it parses the rest.  tile_state: one zeroed word per tile; tile_ticket:
	// one zeroed word.  All unused when rp_w is nullptr.
	uint32_t *deg_w = nullptr;
	uint8_t *ref_w = nullptr;
	int64_t *rp_w = nullptr;
	uint64_t *tile_state = nullptr;
	uint32_t *tile_ticket = nullptr;
	int64_t rp_from_w = 0;  // what rp[from - lo] is going to be (the arcs of the halo rows, which the host has counted)
};

// One random-access job list: `rows` chain rows, each decoded into its own slot.
struct ChainDev {
	const int64_t *node;     // node id of each chain row
	const int64_t *dst_off;  // element offset of the row's slot (in `store` for ancestors, in `col` for the query row)
	const int32_t *parent;   // index of the row holding the reference list (-1: none)
	const uint32_t *deg;     // outdegree of each chain row
	const uint8_t *is_query; // 1: the row is a query's own list (goes to col), 0: ancestor (goes to store)
	int64_t *col;
	int64_t *store;
	RowInfo *info;  // per chain row with a parent: what k_chain_extras parsed (mask of copied positions, or where the blocks start), for k_chain_copy
};

// The global list of rows whose extras are too long to be decoded inside their tile (see bv_kernels.cu).
// A giant row (tens of thousands of residuals and more: BVGraphSlowTest.java:30-96 has a row of 2^30 successors) is not
// decoded by one warp but by GIANT_CHUNKS warps side by side: the bit range of its residuals is cut into that many
// chunks, every warp finds a code start at the head of its chunk (a parse started GIANT_LEAD bits earlier has fallen in
// step by then), counts its codes and sums their gaps (k_giant_scan); one warp per row checks that every chunk starts
// where the one before it ended -- rescanning it if not -- and turns counts and sums into starting indices and values
// (k_giant_fix); then the warps write (k_giant_write).
struct GiantRow {
	int64_t x;              // node id
	uint64_t p0, p1;        // bit range of the residuals (relative to GraphDev::bit_base)
	uint64_t cnt_pos;       // position of the interval count (only used when ic != 0)
	int64_t *ext;           // the extras part of the row's slot (ne entries; the residuals go to its tail)
	uint32_t nres, ne, ic, pad;
};
struct GiantChunk {
	uint64_t entry, exit;   // first code starting in the chunk; first code starting after it
	int64_t total, base;    // sum of (gap + 1) over the chunk's codes; value its first code builds on
	uint32_t count, idx0;   // codes in the chunk; index of the first one
};
static constexpr int GIANT_MAX = 1024;     // giant rows per decode handled this way (more: decoded by one warp each)
static constexpr int GIANT_CHUNKS = 256;
static constexpr uint32_t GIANT_LEAD = 4096;

struct HeavyDev {
	uint32_t *rows;    // row indices (relative to RangeDev.lo) in arrival order; nullptr disables the heavy path
	unsigned long long capacity;  // entries the three arrays can hold (sized from the arc count the properties promise)
	uint8_t *cls;      // class of each entry
	uint32_t *sorted;  // row indices sorted by class, heaviest first
	unsigned long long *counters;  // [0] entries, [1..64] per-class counts, [65..128] per-class cursors, [129] giant rows
	GiantRow *giants;      // GIANT_MAX entries (nullptr: giant rows are decoded by one warp each)
	GiantChunk *gchunks;   // GIANT_MAX * GIANT_CHUNKS entries
};

// The fused tile kernel (bv_fused.cu): one pass over the bit stream, a tile of rows per block, everything on chip.
static constexpr int TILE_MAX_HALO = 64;  // largest window_size * max_ref_count the fused kernel handles
struct FusedDev {
	int64_t from, to;        // nodes to decode; rows before `from` (down to the first row the handle holds) serve as halo
	int32_t halo;            // window_size * max_ref_count of the graph
	int32_t tile_rows;       // rows per tile: 256 - halo
	uint32_t ntiles;
	int64_t *rowptr;         // to - from + 1 entries, rowptr[0] = 0
	int64_t *col;
	uint64_t col_capacity;
	uint64_t *tile_state;    // ntiles zeroed words: look-back state of the arc-count scan
	uint32_t *ticket;        // zeroed: [0] tiles handed out, [1] chunks of k_slow handed out
	uint32_t *slow_count;    // zeroed: tiles left to k_slow
	uint32_t *tile_done;     // ntiles words, written by the kernels: 1 = the tile's successors are complete
};

// BVGraph compression of a device CSR (bv_encode.cu)
struct EncDev {
	const int64_t *rowptr, *col;
	int64_t n;           // rows
	int64_t first_node;  // node id of row 0 (successors are coded relative to the node id)
	int64_t split;       // nodes per compression split (>= 1)
	int32_t window, maxref, minint, zetak;
	int32_t outdegree_coding, block_coding, residual_coding, reference_coding, blockcount_coding;
};
// prices, chooses and writes; returns 0, -1 (CUDA error) or -2 (words_capacity too small: *total_bits_out says how much is needed)
int encode_csr_device(cudaStream_t stream, const EncDev &E, uint32_t *d_cost, uint8_t *d_chosen, unsigned long long *d_bits, int64_t *d_off,
                      uint64_t *scan_state, uint32_t *ticket, uint32_t *d_words, uint64_t words_capacity, uint64_t *total_bits_out);

struct Launch {
	cudaStream_t stream;
	uint64_t *launches;  // incremented once per kernel launched
};

// stage 1: outdegree and reference of every node of [lo, lo+count)
void launch_header(const Launch &L, const GraphDev &g, int def, int64_t lo, int64_t count, uint32_t *deg, uint8_t *ref);
// stage 2: exclusive prefix sum (single pass, decoupled look-back).  state: count/2048+2 words, ticket: 1 word
void launch_scan(const Launch &L, const uint32_t *deg, int64_t *rp, int64_t count, uint64_t *state, uint32_t *ticket);
// rowptr_out[i] = rp[skip + i] - rp[skip] (+ base), i = 0..count
void launch_emit_rowptr(const Launch &L, const int64_t *rp, int64_t skip, int64_t count, int64_t *rowptr_out);
// stage 2b: reference-chain depths.  stats[0] = nodes still unresolved after this pass, stats[1] = max depth seen
void launch_depth(const Launch &L, const GraphDev &g, const uint8_t *ref, uint32_t *depth, int64_t lo, int64_t from,
                  int64_t count, int first_pass, uint32_t *stats);
// ordered level lists, at most 8 levels per pass (nlevels <= 8; deeper chains take several passes): counts and pos hold nlevels * ceil(count / 2048) entries (+1 for pos);
// the pass lists levels level0+1 .. level0+nlevels; level level0+l (l 1-based) occupies list[pos[(l-1) * nblocks] .. pos[l * nblocks]).  With ref != nullptr the depths are computed
// on the way (bounded walk, *deep set if a chain is longer than nlevels) and stored to depth[]; with ref == nullptr depth[] is input.
uint32_t deep_list_blocks(int64_t count);
uint32_t heavy_threshold();  // rows with at least this many extras go through the global class-sorted list
void launch_deep_lists(const Launch &L, const uint32_t *depth, const uint32_t *ids, int64_t count, uint32_t level0, uint32_t nlevels, uint32_t *counts,
                       int64_t *pos, uint32_t *list, uint64_t *scan_state, uint32_t *ticket);
void launch_level_lists(const Launch &L, const GraphDev &g, const uint8_t *ref, uint32_t *depth, int64_t lo, int64_t from, int64_t count,
                        uint32_t level0, uint32_t nlevels, uint32_t *counts, int64_t *pos, uint32_t *list, uint64_t *scan_state, uint32_t *ticket,
                        uint32_t *deep);
// stage 3a: intervals + residuals of every row (whole row when it has no reference)
void launch_extras(const Launch &L, const GraphDev &g, int def, const RangeDev &r, const HeavyDev &h);
uint32_t extras_tiles(int64_t count);
// stage 3b: copy-block resolution of the rows at one depth level; list == nullptr: every node is inspected
void launch_copy_level(const Launch &L, const GraphDev &g, int def, const RangeDev &r, uint32_t level,
                       const uint32_t *list, uint64_t list_count);
// checksum of a CSR
void launch_checksum(const Launch &L, int64_t first_node, int64_t rows, const int64_t *rowptr, const int64_t *col,
                     unsigned long long *out);

// random access
void launch_batch_degrees(const Launch &L, const GraphDev &g, int def, const int64_t *ids, uint64_t k, int64_t lo,
                          int64_t hi, uint32_t *deg);
// walks the reference chain of every query: chain_len[q] rows, chain_arcs[q] successors in ancestor rows
void launch_chain_walk(const Launch &L, const GraphDev &g, int def, const int64_t *ids, uint64_t k, int64_t lo,
                       uint32_t *chain_len, unsigned long long *chain_arcs, uint32_t *stats);
void launch_chain_fill(const Launch &L, const GraphDev &g, int def, const int64_t *ids, uint64_t k,
                       const int64_t *chain_row0, const int64_t *store_off0, const int64_t *rowptr, int64_t *node,
                       int64_t *dst_off, int32_t *parent, uint32_t *deg, uint8_t *is_query, uint32_t *level);
void launch_chain_extras(const Launch &L, const GraphDev &g, int def, const ChainDev &c, uint64_t rows);
void launch_chain_copy(const Launch &L, const GraphDev &g, int def, const ChainDev &c, uint64_t rows,
                       const uint32_t *level, uint32_t this_level);
void launch_scan64(const Launch &L, const unsigned long long *in, int64_t *out, int64_t count, uint64_t *state, uint32_t *ticket);
// the whole decode of [F.from, F.to) in one pass (k_fused) plus the tiles that do not fit shared memory (k_slow);
// returns a cudaError_t as int
int launch_fused(const Launch &L, const GraphDev &g, int def, const FusedDev &F, int sm_count);
int fused_tile_rows(int halo, double mean_degree, double mean_record_bits);
// transpose of a device CSR (bv_transpose.cu): pack arcs into target << 32 | source words and count in-degrees, then
// LSD radix passes over the target half (hist -> scan of counts[digit * tiles + tile] -> stable scatter; the last
// scatter writes the sources to t_col)
uint64_t transpose_tiles(uint64_t arcs);
int transpose_digits();
int transpose_passes(uint64_t nodes);
// (counts: the tile histograms of radix pass 0, which the pack kernel produces on the way)
void launch_transpose_pack(const Launch &L, const GraphDev &g, int64_t n, const int64_t *rowptr, const int64_t *col, uint64_t arcs,
                           uint64_t *pairs, uint32_t *indeg, uint32_t *counts);
// union of two CSRs over the same nodes (bv_transpose.cu): per-row counts, then (after a scan of the counts) the rows
void launch_union_count(const Launch &L, int64_t n, const int64_t *arp, const int64_t *acol, const int64_t *brp, const int64_t *bcol, uint32_t *cnt,
                        uint32_t *err);
void launch_union_fill(const Launch &L, int64_t n, const int64_t *arp, const int64_t *acol, const int64_t *brp, const int64_t *bcol,
                       const int64_t *urp, int64_t *ucol, uint64_t capacity, uint32_t *err);
void launch_radix_hist(const Launch &L, const uint64_t *in, uint64_t count, int pass, uint32_t *counts);
void launch_radix_scatter(const Launch &L, const uint64_t *in, uint64_t count, int pass, const int64_t *pos, uint64_t *out, int64_t *t_col);
// .offsets on the device (bv_offsets.cu): h_stream holds `count` gamma-coded differences; *d_off_out (cudaMalloc'ed, the
// caller frees it) receives their running sums.  device_ef_build turns off[0 .. n) - base into the Elias-Fano arrays
// ef_get reads, leaves them on the device and copies them into `ef`.
struct EliasFano;
int device_decode_offsets(const uint8_t *h_stream, uint64_t nbytes, int64_t count, uint64_t **d_off_out);
int device_ef_build(const uint64_t *d_off, uint64_t n, uint64_t base, uint64_t last_value, EliasFano &ef, uint64_t **d_lower,
                    uint64_t **d_upper, uint64_t **d_hints, uint64_t **d_sparse, uint64_t *device_bytes);
// arc labels (bv_labels.cu): `lab` holds the label stream and its offsets the way a graph handle holds the graph's;
// rowptr[0 .. count] are the row pointers of nodes [from, from + count); out is parallel to their successors
// arcs_bound: an upper bound on the labels of the range (sizes the grid of the fixed-width kernel; the kernel reads the
// exact number from rowptr)
void launch_labels(const Launch &L, const GraphDev &lab, int kind, int width, int64_t from, int64_t count, const int64_t *rowptr,
                   int64_t *out, uint64_t capacity, uint64_t arcs_bound);
// list-valued labels: fill = false writes every arc's list length to lens[arc]; fill = true reads the exclusive prefix sums from
// lens and writes the elements to values
void launch_label_lists(const Launch &L, const GraphDev &lab, bool fill, int width, int is_long, int64_t from, int64_t count, const int64_t *rowptr,
                        int64_t *lens, int64_t *values, uint64_t capacity);
// byte-swap the uploaded stream into big-endian 64-bit words, in place
void launch_bswap(const Launch &L, uint64_t *words, uint64_t nwords);
// A handful of device scalars for the host, written by a kernel straight into mapped pinned host memory: unlike a small
// cudaMemcpyAsync they do not queue behind a large device-to-host copy that another stream has on the copy engine.
struct PeekArgs {
	static constexpr int MAX = 12;
	const void *src[MAX];
	uint8_t bytes[MAX];  // 4 or 8
	uint32_t n = 0;
	void add(const void *p, int size) { src[n] = p; bytes[n] = (uint8_t)size; n++; }
};
void launch_peek(const Launch &L, const PeekArgs &a, uint64_t *mapped_dst);
// out[i] = (uint32_t)in[i]: successors for narrow delivery (wgbv_scan_set_narrow)
void launch_narrow(const Launch &L, const int64_t *in, uint32_t *out, uint64_t count);

}  // namespace wgbv
