// bv_offsets.cu -- the .offsets stream decoded on the device and turned into the Elias-Fano list there.
//
// The reference reads the n+1 gamma-coded differences one after the other (OffsetsLongIterator, BVG:870-898) and feeds
// the running sum to sux4j's Elias-Fano builder (BVG:1558): strictly serial, 0.77 s for 10^8 nodes on the GPU box's
// host even with a fast reader.  On the device the codes are found WITHOUT knowing where they start:
//
//   * For every bit position p of a block of OB_BITS bits the length a gamma code would have if one started at p is a
//     count of leading zeros; J[p] = where the chain of codes that starts at p leaves p's 64-bit chunk (a backward
//     sweep over the chunk, one thread per chunk, in shared memory).
//   * Following J from each of the possible entry positions e of the block (the previous block's last code may reach
//     up to 128 bits into it) gives the block's exit as a function F_b(e).  Gamma streams resynchronise within a few
//     codes, so F_b is almost always constant and the entry of block b+1 is known at once; the rare other blocks walk
//     back to the nearest block with a constant exit and forward again (k_off_resolve).
//   * With its entry known a block walks its codes chunk by chunk (thread t starts where the chain enters chunk t),
//     counts them and sums their values; two scans over the blocks give every thread the index and the running sum of
//     its first code, and a second walk writes the offsets.
//
// The Elias-Fano arrays (layout of EliasFano::build in bv_loader.cpp, what ef_get reads) are then built from the
// offsets by three data-parallel kernels, and only the finished list -- about 1.1 bytes per node -- travels to the host.
#include "bv_block.cuh"
#include "bv_kernels.cuh"
#include "bv_loader.h"

namespace wgbv {

static constexpr int OB_WORDS = 512;            // 32-bit words of the stream per block
static constexpr int OB_BITS = OB_WORDS * 32;   // 16384 positions
static constexpr int OB_CHUNK = OB_BITS / TPB;  // positions per thread: 64
static constexpr int OB_ENTRIES = 130;          // possible entry positions of a block: 0 .. 129
static constexpr int OB_NOCONST = 255;

struct OffSmem {
	uint32_t words[OB_WORDS + 8];  // the block's bits, MSB first, plus what a code starting near the end may need
	uint16_t J[OB_BITS];           // chunk exits, position p of chunk t at (p % 64) * TPB + t
	uint16_t centry[TPB];          // where the chain of codes enters chunk t (0xffff: it does not)
	uint64_t wsum[TPB / 32];
	uint32_t wcnt[TPB / 32];
	uint32_t exits[OB_ENTRIES];
};

__device__ __forceinline__ uint16_t &jref(OffSmem &S, uint32_t p) { return S.J[(p & (OB_CHUNK - 1)) * TPB + (p >> 6)]; }

// the 64 bits of the block that start at position p (first bit in the most significant place)
__device__ __forceinline__ uint64_t window64(const OffSmem &S, uint32_t p) {
	const uint32_t w = p >> 5, s = p & 31u;
	const uint32_t a = S.words[w], b = S.words[w + 1], c = S.words[w + 2];
	return (uint64_t)__funnelshift_l(b, a, s) << 32 | __funnelshift_l(c, b, s);
}

// length of the gamma code that starts at p (129: no code can start there -- more than 63 leading zeros)
__device__ __forceinline__ uint32_t gamma_len(const OffSmem &S, uint32_t p, uint32_t *zeros) {
	const uint64_t w = window64(S, p);
	const uint32_t z = w ? (uint32_t)__clzll((long long)w) : 64u;
	*zeros = z;
	return z < 64 ? 2 * z + 1 : 129u;
}

// value of the gamma code at p with z < 64 leading zeros
__device__ __forceinline__ uint64_t gamma_value(const OffSmem &S, uint32_t p, uint32_t z) {
	const uint64_t w = window64(S, p + z);  // starts with the code's leading one
	return (w >> (63u - z)) - 1;
}

__device__ __forceinline__ void load_block(OffSmem &S, const uint32_t *__restrict__ stream, uint64_t stream_words, uint64_t b) {
	const uint64_t w0 = b * OB_WORDS;
	for (int i = threadIdx.x; i < OB_WORDS + 8; i += TPB) {
		const uint64_t w = w0 + i;
		S.words[i] = w < stream_words ? __byte_perm(stream[w], 0, 0x0123) : 0u;
	}
	__syncthreads();
}

// J for every position of the block; thread t owns chunk t
__device__ __forceinline__ void chunk_exits(OffSmem &S) {
	const uint32_t c0 = threadIdx.x * OB_CHUNK, c1 = c0 + OB_CHUNK;
	for (int p = (int)c1 - 1; p >= (int)c0; p--) {
		uint32_t z;
		const uint32_t nx = (uint32_t)p + gamma_len(S, (uint32_t)p, &z);
		jref(S, (uint32_t)p) = (uint16_t)(nx >= c1 ? nx : jref(S, nx));
	}
	__syncthreads();
}

// pass 1: the exit of every block as a function of its entry; cflag[b] = the exit if it is the same for every entry
__global__ void __launch_bounds__(TPB) k_off_tables(const uint32_t *__restrict__ stream, uint64_t stream_words, uint8_t *__restrict__ F,
                                                    uint8_t *__restrict__ cflag) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	OffSmem &S = *reinterpret_cast<OffSmem *>(smem_raw);
	load_block(S, stream, stream_words, blockIdx.x);
	chunk_exits(S);
	uint32_t ex = 0;
	if (threadIdx.x < OB_ENTRIES) {
		uint32_t pos = threadIdx.x;
		while (pos < OB_BITS) pos = jref(S, pos);
		ex = pos - OB_BITS;
		if (ex > OB_ENTRIES - 1) ex = OB_ENTRIES - 1;
		S.exits[threadIdx.x] = ex;
		F[(size_t)blockIdx.x * OB_ENTRIES + threadIdx.x] = (uint8_t)ex;
	}
	__syncthreads();
	const bool same = threadIdx.x >= OB_ENTRIES || ex == S.exits[0];
	const int all = __syncthreads_and(same);
	if (threadIdx.x == 0) cflag[blockIdx.x] = (uint8_t)(all ? S.exits[0] : OB_NOCONST);
}

// pass 2: the entry of every block
__global__ void __launch_bounds__(TPB) k_off_resolve(const uint8_t *__restrict__ F, const uint8_t *__restrict__ cflag, uint64_t nblocks,
                                                     uint8_t *__restrict__ entry) {
	const uint64_t b = (uint64_t)blockIdx.x * TPB + threadIdx.x;
	if (b >= nblocks) return;
	// the nearest block at or before b whose entry does not depend on anything: block 0, or one that follows a block with
	// a constant exit; from there the exit functions are applied one after the other
	uint64_t k = b;
	while (k > 0 && cflag[k - 1] == OB_NOCONST) k--;
	uint32_t e = k == 0 ? 0u : cflag[k - 1];
	for (uint64_t j = k; j < b; j++) e = F[j * OB_ENTRIES + e];
	entry[b] = (uint8_t)e;
}

// passes 3 and 4: count / sum of the codes of every block, then the offsets themselves
template <bool WRITE>
__global__ void __launch_bounds__(TPB) k_off_walk(const uint32_t *__restrict__ stream, uint64_t stream_words, const uint8_t *__restrict__ entry,
                                                  uint32_t *__restrict__ bcnt, unsigned long long *__restrict__ bsum,
                                                  const int64_t *__restrict__ cbase, const int64_t *__restrict__ sbase, uint64_t count,
                                                  uint64_t *__restrict__ out) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	OffSmem &S = *reinterpret_cast<OffSmem *>(smem_raw);
	load_block(S, stream, stream_words, blockIdx.x);
	chunk_exits(S);
	S.centry[threadIdx.x] = 0xffff;
	__syncthreads();
	if (threadIdx.x == 0) {
		uint32_t pos = entry[blockIdx.x];
		while (pos < OB_BITS) {
			S.centry[pos >> 6] = (uint16_t)pos;
			pos = jref(S, pos);
		}
	}
	__syncthreads();
	const uint32_t c1 = (threadIdx.x + 1) * OB_CHUNK;
	const uint32_t start = S.centry[threadIdx.x];
	uint32_t cnt = 0;
	uint64_t sum = 0;
	if (start != 0xffff) {
		for (uint32_t p = start; p < c1;) {
			uint32_t z;
			const uint32_t len = gamma_len(S, p, &z);
			if (z < 64) { sum += gamma_value(S, p, z); cnt++; }
			p += len;
		}
	}
	// block-wide exclusive scan of (cnt, sum)
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint32_t ic = cnt;
	uint64_t is = sum;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const uint32_t tc = __shfl_up_sync(0xffffffffu, ic, o);
		const uint64_t ts = __shfl_up_sync(0xffffffffu, is, o);
		if (lane >= o) { ic += tc; is += ts; }
	}
	if (lane == 31) { S.wcnt[warp] = ic; S.wsum[warp] = is; }
	__syncthreads();
	uint32_t bc = 0;
	uint64_t bs = 0;
	for (int w = 0; w < warp; w++) { bc += S.wcnt[w]; bs += S.wsum[w]; }
	if (!WRITE) {
		if (threadIdx.x == TPB - 1) {
			bcnt[blockIdx.x] = bc + ic;
			bsum[blockIdx.x] = bs + is;
		}
		return;
	}
	uint64_t idx = (uint64_t)cbase[blockIdx.x] + bc + (ic - cnt);
	uint64_t run = (uint64_t)sbase[blockIdx.x] + bs + (is - sum);
	if (start != 0xffff) {
		for (uint32_t p = start; p < c1;) {
			uint32_t z;
			const uint32_t len = gamma_len(S, p, &z);
			if (z < 64) {
				run += gamma_value(S, p, z);
				if (idx < count) out[idx] = run;
				idx++;
			}
			p += len;
		}
	}
}

// ------------------------------------------------------------------------------------------------ Elias-Fano on the device
// lower bits: one thread per 64-bit word of the array (the elements that overlap it)
__global__ void __launch_bounds__(TPB) k_ef_lower(const uint64_t *__restrict__ off, uint64_t n, uint64_t base, int l, uint64_t nwords,
                                                  uint64_t *__restrict__ lower) {
	const uint64_t w = (uint64_t)blockIdx.x * TPB + threadIdx.x;
	if (w >= nwords) return;
	uint64_t word = 0;
	if (l) {
		const uint64_t b0 = w * 64, b1 = b0 + 64;
		const uint64_t lmask = (1ull << l) - 1;
		for (uint64_t i = b0 / (uint64_t)l; i < n && i * (uint64_t)l < b1; i++) {
			const uint64_t v = (off[i] - base) & lmask, lb = i * (uint64_t)l;
			if (lb >= b0) word |= v << (lb - b0);
			else word |= v >> (b0 - lb);
		}
	}
	lower[w] = word;
}

// upper bits: element i sets bit (v_i >> l) + i
__global__ void __launch_bounds__(TPB) k_ef_upper(const uint64_t *__restrict__ off, uint64_t n, uint64_t base, int l,
                                                  unsigned long long *__restrict__ upper) {
	const uint64_t i = (uint64_t)blockIdx.x * TPB + threadIdx.x;
	if (i >= n) return;
	const uint64_t pos = ((off[i] - base) >> l) + i;
	atomicOr(&upper[pos >> 6], 1ull << (pos & 63));
}

// hints: bit position of element 64k; blocks spread over more than 4096 bits keep their 64 positions explicitly
__global__ void __launch_bounds__(TPB) k_ef_hints(const uint64_t *__restrict__ off, uint64_t n, uint64_t base, int l, uint64_t nblocks,
                                                  uint64_t *__restrict__ hints, uint64_t *__restrict__ sparse, unsigned long long *sparse_blocks,
                                                  uint64_t sparse_capacity) {
	const uint64_t k = (uint64_t)blockIdx.x * TPB + threadIdx.x;
	if (k >= nblocks) return;
	const uint64_t i0 = k * 64, i1 = i0 + 64 < n ? i0 + 64 : n;
	const uint64_t first = ((off[i0] - base) >> l) + i0, last = ((off[i1 - 1] - base) >> l) + (i1 - 1);
	if (last - first > 4096) {
		const unsigned long long idx = atomicAdd(sparse_blocks, 1ull);
		if (idx < sparse_capacity) {
			for (uint64_t j = 0; j < 64; j++) sparse[idx * 64 + j] = i0 + j < i1 ? ((off[i0 + j] - base) >> l) + i0 + j : last;
			hints[k] = (1ull << 63) | idx;
		}
	} else hints[k] = first;
}

// ------------------------------------------------------------------------------------------------ host drivers
#define OFF_CU(x)                                                                                            \
	do {                                                                                                     \
		const cudaError_t e_ = (x);                                                                          \
		if (e_ != cudaSuccess) { rc = fail(WGBV_ECUDA, "%s: %s", #x, cudaGetErrorString(e_)); goto done; }   \
	} while (0)

int device_decode_offsets(const uint8_t *h_stream, uint64_t nbytes, int64_t count, uint64_t **d_off_out) {
	*d_off_out = nullptr;
	if (count <= 0) return fail(WGBV_EARG, "no offsets to decode");
	int rc = WGBV_OK;
	const uint64_t stream_words = (nbytes + 3) / 4;
	const uint64_t nblocks = (nbytes * 8 + OB_BITS - 1) / OB_BITS;
	if (nblocks == 0) return fail(WGBV_EIO, "the offsets stream ends before %lld offsets were read", (long long)count);
	uint32_t *d_stream = nullptr, *d_cnt = nullptr, *d_ticket = nullptr;
	uint8_t *d_F = nullptr, *d_cflag = nullptr, *d_entry = nullptr;
	unsigned long long *d_sum = nullptr;
	int64_t *d_cbase = nullptr, *d_sbase = nullptr;
	uint64_t *d_state = nullptr, *d_off = nullptr;
	int64_t total = 0;
	const Launch L{nullptr, nullptr};
	const size_t smem = sizeof(OffSmem);
	OFF_CU(cudaMalloc((void **)&d_stream, stream_words * 4 + 64));
	OFF_CU(cudaMemset(d_stream + (stream_words ? stream_words - 1 : 0), 0, 4 + 64));
	OFF_CU(cudaMemcpy(d_stream, h_stream, nbytes, cudaMemcpyHostToDevice));
	OFF_CU(cudaMalloc((void **)&d_F, nblocks * OB_ENTRIES));
	OFF_CU(cudaMalloc((void **)&d_cflag, nblocks));
	OFF_CU(cudaMalloc((void **)&d_entry, nblocks));
	OFF_CU(cudaMalloc((void **)&d_cnt, nblocks * 4));
	OFF_CU(cudaMalloc((void **)&d_sum, nblocks * 8));
	OFF_CU(cudaMalloc((void **)&d_cbase, (nblocks + 1) * 8));
	OFF_CU(cudaMalloc((void **)&d_sbase, (nblocks + 1) * 8));
	OFF_CU(cudaMalloc((void **)&d_state, ((nblocks + 1) / 2048 + 2) * 8));
	OFF_CU(cudaMalloc((void **)&d_ticket, 64));
	OFF_CU(cudaMalloc((void **)&d_off, (size_t)count * 8));
	OFF_CU(cudaFuncSetAttribute(k_off_tables, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
	OFF_CU(cudaFuncSetAttribute(k_off_walk<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
	OFF_CU(cudaFuncSetAttribute(k_off_walk<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
	k_off_tables<<<(unsigned)nblocks, TPB, smem>>>(d_stream, stream_words, d_F, d_cflag);
	k_off_resolve<<<grid_for((int64_t)nblocks, TPB), TPB>>>(d_F, d_cflag, nblocks, d_entry);
	k_off_walk<false><<<(unsigned)nblocks, TPB, smem>>>(d_stream, stream_words, d_entry, d_cnt, d_sum, nullptr, nullptr, (uint64_t)count, nullptr);
	launch_scan(L, d_cnt, d_cbase, (int64_t)nblocks, d_state, d_ticket);
	launch_scan64(L, d_sum, d_sbase, (int64_t)nblocks, d_state, d_ticket);
	k_off_walk<true><<<(unsigned)nblocks, TPB, smem>>>(d_stream, stream_words, d_entry, nullptr, nullptr, d_cbase, d_sbase, (uint64_t)count, d_off);
	OFF_CU(cudaMemcpy(&total, d_cbase + nblocks, 8, cudaMemcpyDeviceToHost));
	OFF_CU(cudaGetLastError());
	if (total < count) {
		rc = fail(WGBV_EIO, "the offsets stream ends before %lld offsets were read", (long long)count);
		goto done;
	}
	*d_off_out = d_off;
	d_off = nullptr;
done:
	cudaFree(d_stream); cudaFree(d_F); cudaFree(d_cflag); cudaFree(d_entry); cudaFree(d_cnt); cudaFree(d_sum);
	cudaFree(d_cbase); cudaFree(d_sbase); cudaFree(d_state); cudaFree(d_ticket); cudaFree(d_off);
	return rc;
}

// Builds the list of off[0 .. n) - base on the device; the arrays stay there (out pointers) and a copy goes to `ef`.
int device_ef_build(const uint64_t *d_off, uint64_t n, uint64_t base, uint64_t last_value, EliasFano &ef, uint64_t **d_lower,
                    uint64_t **d_upper, uint64_t **d_hints, uint64_t **d_sparse, uint64_t *device_bytes) {
	int rc = WGBV_OK;
	ef.count = n;
	const uint64_t u = last_value - base + 1;
	ef.l = 0;
	if (u / n >= 2) ef.l = 63 - __builtin_clzll(u / n);
	const int l = ef.l;
	const uint64_t upper_bits = n + ((last_value - base) >> l) + 1;
	const uint64_t nupper = (upper_bits + 63) / 64 + 2, nlower = (n * (uint64_t)l + 63) / 64 + 2, nhints = (n + 63) / 64 + 1;
	const uint64_t nblocks = (n + 63) / 64;
	const uint64_t sparse_cap = upper_bits / 4096 + 1;  // a sparse block covers more than 4096 upper bits
	unsigned long long *d_nsparse = nullptr;
	uint64_t *sp = nullptr;
	unsigned long long nsparse = 0;
	*d_lower = *d_upper = *d_hints = *d_sparse = nullptr;
	OFF_CU(cudaMalloc((void **)d_lower, nlower * 8));
	OFF_CU(cudaMalloc((void **)d_upper, nupper * 8));
	OFF_CU(cudaMalloc((void **)d_hints, nhints * 8));
	OFF_CU(cudaMalloc((void **)&sp, sparse_cap * 64 * 8));
	OFF_CU(cudaMalloc((void **)&d_nsparse, 8));
	OFF_CU(cudaMemset(*d_upper, 0, nupper * 8));
	OFF_CU(cudaMemset(*d_hints, 0, nhints * 8));
	OFF_CU(cudaMemset(d_nsparse, 0, 8));
	k_ef_lower<<<grid_for((int64_t)nlower, TPB), TPB>>>(d_off, n, base, l, nlower, *d_lower);
	k_ef_upper<<<grid_for((int64_t)n, TPB), TPB>>>(d_off, n, base, l, (unsigned long long *)*d_upper);
	k_ef_hints<<<grid_for((int64_t)nblocks, TPB), TPB>>>(d_off, n, base, l, nblocks, *d_hints, sp, d_nsparse, sparse_cap);
	OFF_CU(cudaMemcpy(&nsparse, d_nsparse, 8, cudaMemcpyDeviceToHost));
	OFF_CU(cudaGetLastError());
	if (nsparse > sparse_cap) { rc = fail(WGBV_ECUDA, "internal: more sparse blocks than the bound"); goto done; }
	{
		// the explicit positions, trimmed to what is used (at least one word: ef_get never sees a null array)
		const uint64_t nsp = nsparse ? nsparse * 64 : 1;
		OFF_CU(cudaMalloc((void **)d_sparse, nsp * 8));
		if (nsparse) OFF_CU(cudaMemcpy(*d_sparse, sp, nsp * 8, cudaMemcpyDeviceToDevice));
		else OFF_CU(cudaMemset(*d_sparse, 0, 8));
		ef.lower.resize((size_t)nlower);
		ef.upper.resize((size_t)nupper);
		ef.hints.resize((size_t)nhints);
		ef.sparse.resize((size_t)nsp);
		OFF_CU(cudaMemcpy(ef.lower.data(), *d_lower, nlower * 8, cudaMemcpyDeviceToHost));
		OFF_CU(cudaMemcpy(ef.upper.data(), *d_upper, nupper * 8, cudaMemcpyDeviceToHost));
		OFF_CU(cudaMemcpy(ef.hints.data(), *d_hints, nhints * 8, cudaMemcpyDeviceToHost));
		OFF_CU(cudaMemcpy(ef.sparse.data(), *d_sparse, nsp * 8, cudaMemcpyDeviceToHost));
		if (device_bytes) *device_bytes += (nlower + nupper + nhints + nsp) * 8;
	}
done:
	cudaFree(sp);
	cudaFree(d_nsparse);
	if (rc) {
		cudaFree(*d_lower); cudaFree(*d_upper); cudaFree(*d_hints); cudaFree(*d_sparse);
		*d_lower = *d_upper = *d_hints = *d_sparse = nullptr;
	}
	return rc;
}

}  // namespace wgbv
