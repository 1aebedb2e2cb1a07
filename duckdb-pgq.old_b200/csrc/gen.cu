// Synthetic edge tables written straight into HBM (bench/test helper, not the product path).
// Shares include/pgq_gen.h verbatim with the CPU generator so both sides emit identical rows.
#include "common.cuh"
#include "../../include/pgq_gen.h"

namespace pgq {

__global__ void __launch_bounds__(256) gen_rmat_kernel(uint64_t seed, int scale, uint64_t n, int64_t *src,
                                                       int64_t *dst) {
	uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
		int64_t s, d;
		pgq_rmat_edge(seed, scale, i, &s, &d);
		src[i] = s;
		dst[i] = d;
	}
}

// rows [row0, row0+n) of the same stream, written to src[0..n), dst[0..n)
__global__ void __launch_bounds__(256) gen_rmat_rows_kernel(uint64_t seed, int scale, uint64_t row0, uint64_t n,
                                                            int64_t *src, int64_t *dst) {
	uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
		int64_t s, d;
		pgq_rmat_edge(seed, scale, row0 + i, &s, &d);
		src[i] = s;
		dst[i] = d;
	}
}

__global__ void __launch_bounds__(256) gen_knows_kernel(uint64_t seed, uint64_t V, uint64_t E, uint64_t mult,
                                                        uint64_t add, int64_t *src, int64_t *dst) {
	uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < E; i += stride) {
		int64_t s, d;
		pgq_knows_edge(seed, V, E, mult, add, i, &s, &d);
		src[i] = s;
		dst[i] = d;
	}
}

int gen_rmat_device(Ctx *c, uint64_t seed, int scale, int64_t E, int64_t *d_src, int64_t *d_dst) {
	if (E <= 0) return PGQ_OK;
	gen_rmat_kernel<<<c->sm_count * 16, 256, 0, c->stream>>>(seed, scale, (uint64_t)E, d_src, d_dst);
	count_launch(c);
	PGQ_CUDA(cudaGetLastError());
	return PGQ_OK;
}

int gen_rmat_rows_device(Ctx *c, uint64_t seed, int scale, int64_t row0, int64_t n, int64_t *d_src,
                         int64_t *d_dst) {
	if (n <= 0) return PGQ_OK;
	gen_rmat_rows_kernel<<<c->sm_count * 16, 256, 0, c->stream>>>(seed, scale, (uint64_t)row0, (uint64_t)n, d_src,
	                                                              d_dst);
	count_launch(c);
	PGQ_CUDA(cudaGetLastError());
	return PGQ_OK;
}

int gen_knows_device(Ctx *c, uint64_t seed, int64_t V, int64_t E, uint64_t mult, uint64_t add, int64_t *d_src,
                     int64_t *d_dst) {
	if (E <= 0) return PGQ_OK;
	gen_knows_kernel<<<c->sm_count * 16, 256, 0, c->stream>>>(seed, (uint64_t)V, (uint64_t)E, mult, add, d_src,
	                                                          d_dst);
	count_launch(c);
	PGQ_CUDA(cudaGetLastError());
	return PGQ_OK;
}

} // namespace pgq
