// K2: device-wide exclusive scan (reduce / scan-of-sums / apply), used for the CSR offsets
// (create_csr_edge's prefix sum) and for the radix-pass bucket bases.
// HBM-bound streaming: 4 B read (x2) + 4 or 8 B write per element.
#include "common.cuh"

namespace pgq {

static constexpr int SCAN_THREADS = 512;
static constexpr int SCAN_ITEMS = 8;
static constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t x) {
	const unsigned lane = threadIdx.x & 31;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		uint32_t y = __shfl_up_sync(0xffffffffu, x, d);
		if (lane >= (unsigned)d) x += y;
	}
	return x;
}

// Block-wide exclusive scan of one value per thread; returns exclusive prefix, *total = block sum.
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t x, uint32_t *smem /*[32]*/, uint32_t *total) {
	const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint32_t incl = warp_incl_scan(x);
	if (lane == 31) smem[warp] = incl;
	__syncthreads();
	if (warp == 0) {
		uint32_t w = (lane < (blockDim.x >> 5)) ? smem[lane] : 0;
		uint32_t wi = warp_incl_scan(w);
		smem[lane] = wi - w; // exclusive warp base
		if (lane == 31) smem[32] = wi;
	}
	__syncthreads();
	uint32_t res = smem[warp] + incl - x;
	*total = smem[32];
	__syncthreads();
	return res;
}

__device__ __forceinline__ void load_items(const uint32_t *in, uint64_t n, uint64_t base, uint32_t (&it)[SCAN_ITEMS]) {
	uint64_t i0 = base + (uint64_t)threadIdx.x * SCAN_ITEMS;
	if (i0 + SCAN_ITEMS <= n) {
		const uint4 *p = reinterpret_cast<const uint4 *>(in + i0);
		uint4 a = p[0], b = p[1];
		it[0] = a.x; it[1] = a.y; it[2] = a.z; it[3] = a.w;
		it[4] = b.x; it[5] = b.y; it[6] = b.z; it[7] = b.w;
	} else {
#pragma unroll
		for (int k = 0; k < SCAN_ITEMS; k++) it[k] = (i0 + k < n) ? in[i0 + k] : 0u;
	}
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_reduce_kernel(const uint32_t *__restrict__ in, uint64_t n,
                                                                   uint32_t *__restrict__ sums) {
	__shared__ uint32_t sm[33];
	uint32_t it[SCAN_ITEMS];
	load_items(in, n, (uint64_t)blockIdx.x * SCAN_TILE, it);
	uint32_t s = 0;
#pragma unroll
	for (int k = 0; k < SCAN_ITEMS; k++) s += it[k];
	uint32_t total;
	block_excl_scan(s, sm, &total);
	if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

// single block: exclusive scan of the per-tile sums in place; writes grand total
__global__ void __launch_bounds__(1024) scan_sums_kernel(uint32_t *__restrict__ sums, uint32_t nsums,
                                                         uint64_t *__restrict__ d_total) {
	__shared__ uint32_t sm[33];
	uint32_t carry = 0;
	for (uint32_t base = 0; base < nsums; base += blockDim.x) {
		uint32_t i = base + threadIdx.x;
		uint32_t x = i < nsums ? sums[i] : 0u;
		uint32_t total;
		uint32_t ex = block_excl_scan(x, sm, &total);
		if (i < nsums) sums[i] = carry + ex;
		carry += total;
	}
	if (threadIdx.x == 0 && d_total) *d_total = carry;
}

template <typename OutT>
__global__ void __launch_bounds__(SCAN_THREADS) scan_apply_kernel(const uint32_t *__restrict__ in, uint64_t n,
                                                                  const uint32_t *__restrict__ sums,
                                                                  OutT *__restrict__ out) {
	__shared__ uint32_t sm[33];
	uint32_t it[SCAN_ITEMS];
	const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE;
	load_items(in, n, base, it);
	uint32_t s = 0;
#pragma unroll
	for (int k = 0; k < SCAN_ITEMS; k++) s += it[k];
	uint32_t total;
	uint32_t ex = block_excl_scan(s, sm, &total) + sums[blockIdx.x];
	uint64_t i0 = base + (uint64_t)threadIdx.x * SCAN_ITEMS;
#pragma unroll
	for (int k = 0; k < SCAN_ITEMS; k++) {
		if (i0 + k < n) out[i0 + k] = (OutT)ex;
		ex += it[k];
	}
}

template <typename OutT>
static int scan_impl(Ctx *c, const uint32_t *d_in, OutT *d_out, uint64_t n, uint64_t *d_total) {
	if (n == 0) {
		if (d_total) PGQ_CUDA(cudaMemsetAsync(d_total, 0, sizeof(uint64_t), c->stream));
		return PGQ_OK;
	}
	uint32_t ntiles = div_up(n, SCAN_TILE);
	PGQ_TRY(c->scan_tmp.reserve((size_t)ntiles * sizeof(uint32_t) + 64));
	uint32_t *sums = c->scan_tmp.as<uint32_t>();
	scan_reduce_kernel<<<ntiles, SCAN_THREADS, 0, c->stream>>>(d_in, n, sums);
	scan_sums_kernel<<<1, 1024, 0, c->stream>>>(sums, ntiles, d_total);
	scan_apply_kernel<OutT><<<ntiles, SCAN_THREADS, 0, c->stream>>>(d_in, n, sums, d_out);
	count_launch(c, 3);
	PGQ_CUDA(cudaGetLastError());
	return PGQ_OK;
}

int exclusive_scan_u32(Ctx *c, const uint32_t *d_in, uint32_t *d_out, uint64_t n, uint64_t *d_total) {
	return scan_impl<uint32_t>(c, d_in, d_out, n, d_total);
}

int exclusive_scan_u32_to_i64(Ctx *c, const uint32_t *d_in, int64_t *d_out, uint64_t n) {
	return scan_impl<int64_t>(c, d_in, d_out, n, nullptr);
}

} // namespace pgq
