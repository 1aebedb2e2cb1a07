// 1-D vertex partition (BASELINE.json configs[4]): every rank keeps the in-edges
// of the targets it owns and builds that slice of the reverse CSR with the same stable sort as the
// single-GPU build.  The in-list order is the arrival order of the rows (enough for lengths and
// reachability; parents are not offered on this path).
#include "common.cuh"
#include "nccl_dl.cuh"

namespace pgq {

// keep[i] = 1 when dst[i] is owned by this rank; err |= 1 for ids outside [0, gV)
__global__ void __launch_bounds__(256) part_flag_kernel(const int64_t *__restrict__ src,
                                                        const int64_t *__restrict__ dst, uint32_t n, uint64_t gV,
                                                        PartMap pm, uint32_t rank, uint32_t *__restrict__ keep,
                                                        uint32_t *__restrict__ err) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const uint64_t s = (uint64_t)src[i], d = (uint64_t)dst[i];
	uint32_t k = 0;
	if (s >= gV || d >= gV) atomicOr(err, 1u);
	else k = part_owner(d, pm) == rank ? 1u : 0u;
	keep[i] = k;
}

// stable append of the kept rows: local target index + renumbered (rank-major) source
__global__ void __launch_bounds__(256) part_scatter_kernel(const int64_t *__restrict__ src,
                                                           const int64_t *__restrict__ dst, uint32_t n, uint64_t gV,
                                                           PartMap pm, uint32_t rank,
                                                           const uint32_t *__restrict__ pos,
                                                           uint32_t *__restrict__ ltgt,
                                                           uint32_t *__restrict__ lsrc) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const uint64_t s = (uint64_t)src[i], d = (uint64_t)dst[i];
	if (s < gV && d < gV && part_owner(d, pm) == rank) {
		const uint32_t p = pos[i];
		ltgt[p] = part_local(d, pm);
		lsrc[p] = part_newid(s, pm);
	}
}

// DevBuf::reserve drops the contents; this one keeps the first `used` bytes
static int grow_keep(DevBuf *b, size_t used, size_t need, cudaStream_t st) {
	if (need <= b->cap) return PGQ_OK;
	size_t cap = b->cap ? b->cap : (size_t)1 << 20;
	while (cap < need) cap += cap / 2 + ((size_t)1 << 20);
	DevBuf nb;
	PGQ_TRY(nb.reserve(cap));
	if (used) PGQ_CUDA(cudaMemcpyAsync(nb.p, b->p, used, cudaMemcpyDeviceToDevice, st));
	PGQ_CUDA(cudaStreamSynchronize(st));
	b->release();
	*b = nb;
	return PGQ_OK;
}

int part_add_edges_device(Ctx *c, Csr *g, const int64_t *d_src, const int64_t *d_dst, int64_t n) {
	if (n <= 0) return PGQ_OK;
	if (n >= ((int64_t)1 << 32) - 1) {
		set_error("part_add_edges: at most 2^32-2 rows per call");
		return PGQ_ERR_INVALID;
	}
	cudaStream_t st = c->stream;
	TmpBuf keep, tot;
	PGQ_TRY(keep.reserve((size_t)n * sizeof(uint32_t)));
	PGQ_TRY(tot.reserve(32));
	uint32_t *d_err = tot.as<uint32_t>() + 4;
	PGQ_CUDA(cudaMemsetAsync(tot.p, 0, 32, st));
	const PartMap pm = part_map(g->gV, c->part_world);
	const uint32_t rank = (uint32_t)c->part_rank;
	part_flag_kernel<<<div_up(n, 256), 256, 0, st>>>(d_src, d_dst, (uint32_t)n, (uint64_t)g->gV, pm, rank,
	                                                 keep.as<uint32_t>(), d_err);
	count_launch(c);
	PGQ_TRY(exclusive_scan_u32(c, keep.as<uint32_t>(), keep.as<uint32_t>(), (uint64_t)n, tot.as<uint64_t>()));
	uint64_t h[4] = {0, 0, 0, 0};
	PGQ_CUDA(cudaMemcpyAsync(h, tot.p, 32, cudaMemcpyDeviceToHost, st));
	PGQ_CUDA(cudaStreamSynchronize(st));
	if (reinterpret_cast<uint32_t *>(h)[4] & 1u) {
		set_error("create_csr_edge: src/dst id outside [0, v_size)");
		return PGQ_ERR_INVALID;
	}
	const uint64_t add = h[0];
	if (g->n_local + add >= ((uint64_t)1 << 32) - 1) {
		set_error("partition slice holds too many edges for 32-bit positions");
		return PGQ_ERR_INVALID;
	}
	if (add) {
		PGQ_TRY(grow_keep(&g->d_ltgt, g->n_local * 4, (g->n_local + add) * 4, st));
		PGQ_TRY(grow_keep(&g->d_lsrc, g->n_local * 4, (g->n_local + add) * 4, st));
		part_scatter_kernel<<<div_up(n, 256), 256, 0, st>>>(d_src, d_dst, (uint32_t)n, (uint64_t)g->gV, pm, rank,
		                                                    keep.as<uint32_t>(),
		                                                    g->d_ltgt.as<uint32_t>() + g->n_local,
		                                                    g->d_lsrc.as<uint32_t>() + g->n_local);
		count_launch(c);
		g->n_local += add;
	}
	PGQ_CUDA(cudaStreamSynchronize(st));
	return PGQ_OK;
}


int part_finalize(Ctx *c, Csr *g) {
	cudaStream_t st = c->stream;
	const int64_t Vp = g->V;
	const uint64_t E = g->n_local;
	g->E = (int64_t)E;
	PGQ_TRY(g->d_ioff.reserve((size_t)(Vp + 1) * sizeof(uint32_t)));
	PGQ_TRY(g->d_isrc.reserve((size_t)(E ? E : 1) * sizeof(uint32_t)));
	const uint32_t *tkeys = nullptr;
	if (E) {
		int bits = 1;
		while (bits < 32 && ((int64_t)1 << bits) < Vp) bits++;
		SortIO io;
		io.k32 = g->d_ltgt.as<uint32_t>();
		io.p32 = g->d_lsrc.as<uint32_t>();
		io.n = (uint32_t)E;
		io.key_bits = bits;
		io.out_p0 = g->d_isrc.as<uint32_t>();
		io.keys_scratch = &tkeys;
		PGQ_TRY(radix_sort_stable(c, io));
	}
	PGQ_TRY(part_offsets_u32(c, tkeys, (uint32_t)E, (uint32_t)Vp, g->d_ioff.as<uint32_t>()));
	PGQ_CUDA(cudaStreamSynchronize(st));
	g->d_ltgt.release();
	g->d_lsrc.release();
	g->prepared = true; // the slice IS the reverse CSR; csr_prepare only adds the in-degree classes
	g->state = Csr::READY;
	return csr_prepare(c, g);
}

} // namespace pgq
