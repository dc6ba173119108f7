// bv_noalloc.cu -- wgbv_scan_degrees_static: outdegrees -> CSR row pointers without allocating anything.
//
// Who needs this.  Every other entry point of libwgbv takes its scratch memory from cudaMalloc (through the
// grow-only buffers of a wgbv_stream).  Some hosts cannot afford that call at the moment they need row pointers:
//   * a process whose device memory has already been handed, all of it, to another component's pool (a JVM-side
//     arena, a framework's caching allocator): a cudaMalloc there fails, or succeeds only after that pool is
//     trimmed;
//   * a latency-sensitive thread (one that serves queries while other streams are busy): cudaMalloc and cudaFree
//     can synchronise the whole device, so one allocation stalls every stream of the process.
// Such callers already own the input (outdegrees) and the output (row pointers) buffers; the only thing the
// single-pass prefix sum needs beyond them is its tile-status table and its ticket.  Here those live in a
// file-scope __device__ array that is part of the loaded library image: no allocation call is ever made.
//
// Cost and limits of that choice.  The array is WGBV_STATIC_SCAN_TILES + 32 eight-byte words (about 513 KiB) and
// every GPU that loads the library holds it until the process exits.  It bounds the input to
// WGBV_STATIC_SCAN_MAX_COUNT degrees per call; longer inputs are refused, not truncated.  All calls on a device
// share the one array, and it keeps whatever the previous call left in it, so each call first clears the words it
// is going to use; two calls whose kernels overlap on the GPU (different streams, no ordering between them) would
// race on it and must not be issued.
//
// The kernel is a variant of k_scan (bv_kernels.cu), same single-pass decoupled look-back, with the status words
// and the ticket taken from the static array.
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

#include "bv_loader.h"

namespace {

constexpr int SCAN_TPB = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_TPB * SCAN_ITEMS;
constexpr int MAX_DEVICES = 64;

// [0] ticket (with padding up to 32 words), then one status word per tile
static __device__ __align__(256) unsigned long long g_scan_scratch[32 + WGBV_STATIC_SCAN_TILES];

enum : uint64_t { ST_AGG = 1ull << 62, ST_PFX = 2ull << 62, ST_MASK = 3ull << 62 };

__global__ void __launch_bounds__(SCAN_TPB) k_scan_static(const uint32_t *__restrict__ in, int64_t *__restrict__ out, int64_t count,
                                                          unsigned long long *scratch) {
	volatile unsigned long long *state = scratch + 32;
	__shared__ uint32_t s_tile;
	__shared__ uint64_t s_warp[SCAN_TPB / 32];
	__shared__ uint64_t s_base;
	if (threadIdx.x == 0) s_tile = (uint32_t)atomicAdd(scratch, 1ull);  // tiles are taken in order: a tile only waits for running ones
	__syncthreads();
	const uint32_t tile = s_tile;
	const int64_t base = (int64_t)tile * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
	uint64_t v[SCAN_ITEMS], sum = 0;
#pragma unroll
	for (int j = 0; j < SCAN_ITEMS; j++) {
		v[j] = base + j < count ? (uint64_t)in[base + j] : 0;
		sum += v[j];
	}
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint64_t inc = sum;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const uint64_t t = __shfl_up_sync(0xffffffffu, inc, o);
		if (lane >= o) inc += t;
	}
	if (lane == 31) s_warp[warp] = inc;
	__syncthreads();
	if (warp == 0) {
		uint64_t w = lane < SCAN_TPB / 32 ? s_warp[lane] : 0;
#pragma unroll
		for (int o = 1; o < SCAN_TPB / 32; o <<= 1) {
			const uint64_t t = __shfl_up_sync(0xffffffffu, w, o);
			if (lane >= o) w += t;
		}
		const uint64_t tile_total = __shfl_sync(0xffffffffu, w, SCAN_TPB / 32 - 1);
		if (lane < SCAN_TPB / 32) s_warp[lane] = w;
		uint64_t exclusive = 0;
		if (tile == 0) {
			if (lane == 0) state[0] = ST_PFX | tile_total;
		} else {
			if (lane == 0) state[tile] = ST_AGG | tile_total;
			int64_t look = (int64_t)tile - 1;
			for (;;) {
				const int64_t idx = look - lane;
				uint64_t s = idx >= 0 ? state[idx] : ST_PFX;
				while (__any_sync(0xffffffffu, (s & ST_MASK) == 0))
					if ((s & ST_MASK) == 0) s = state[idx];
				const unsigned pfx = __ballot_sync(0xffffffffu, (s & ST_MASK) == ST_PFX);
				const int first = pfx ? __ffs(pfx) - 1 : 32;
				uint64_t contrib = (lane <= first && idx >= 0) ? (s & ~ST_MASK) : 0;
#pragma unroll
				for (int o = 16; o; o >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, o);
				exclusive += contrib;
				if (pfx) break;
				look -= 32;
			}
			if (lane == 0) state[tile] = ST_PFX | (exclusive + tile_total);
		}
		if (lane == 0) s_base = exclusive;
	}
	__syncthreads();
	uint64_t run = s_base + (warp ? s_warp[warp - 1] : 0) + (inc - sum);
#pragma unroll
	for (int j = 0; j < SCAN_ITEMS; j++) {
		if (base + j <= count) out[base + j] = (int64_t)run;
		run += v[j];
	}
}

std::mutex g_setup_lock;
unsigned long long *g_scratch_ptr[MAX_DEVICES] = {};  // device address of g_scan_scratch on each device, once looked up

}  // namespace

extern "C" int wgbv_scan_degrees_static(const uint32_t *d_degrees, int64_t count, int64_t *d_rowptr, void *cuda_stream) {
	using wgbv::fail;
	wgbv::clear_error();
	if (count < 0 || (!d_degrees && count) || !d_rowptr) return fail(WGBV_EARG, "null argument or negative count");
	if (count > WGBV_STATIC_SCAN_MAX_COUNT)
		return fail(WGBV_EARG, "wgbv_scan_degrees_static handles at most %lld degrees per call (got %lld): its scratch is a fixed-size static array",
		            (long long)WGBV_STATIC_SCAN_MAX_COUNT, (long long)count);
	int dev = -1;
	cudaError_t e = cudaGetDevice(&dev);
	if (e != cudaSuccess) return fail(WGBV_ECUDA, "cudaGetDevice failed: %s", cudaGetErrorString(e));
	if (dev < 0 || dev >= MAX_DEVICES) return fail(WGBV_ECUDA, "device index %d not supported", dev);
	// Set-up, once per device, with that device current.  In host code the name g_scan_scratch is only a handle for
	// the symbol, not a device address: the address has to be asked for.
	unsigned long long *scratch;
	{
		std::lock_guard<std::mutex> hold(g_setup_lock);
		if (!g_scratch_ptr[dev]) {
			void *p = nullptr;
			e = cudaGetSymbolAddress(&p, g_scan_scratch);
			if (e != cudaSuccess || !p) return fail(WGBV_ECUDA, "cudaGetSymbolAddress(g_scan_scratch) failed on device %d: %s", dev, cudaGetErrorString(e));
			g_scratch_ptr[dev] = (unsigned long long *)p;
		}
		scratch = g_scratch_ptr[dev];
	}
	cudaStream_t stream = (cudaStream_t)cuda_stream;
	const int64_t tiles = (count + 1 + SCAN_TILE - 1) / SCAN_TILE;  // count + 1 outputs
	// the array still holds the previous call's ticket and status words: clear the part this call uses
	e = cudaMemsetAsync(scratch, 0, (size_t)(32 + tiles) * sizeof(unsigned long long), stream);
	if (e != cudaSuccess) return fail(WGBV_ECUDA, "cudaMemsetAsync failed: %s", cudaGetErrorString(e));
	k_scan_static<<<(unsigned)tiles, SCAN_TPB, 0, stream>>>(d_degrees, d_rowptr, count, scratch);
	e = cudaGetLastError();
	if (e != cudaSuccess) return fail(WGBV_ECUDA, "launch of k_scan_static failed: %s", cudaGetErrorString(e));
	return WGBV_OK;
}
