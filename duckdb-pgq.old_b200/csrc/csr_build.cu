// CSR construction on the device (SURVEY.md §8(a) a4 = restatement of create_csr_vertex +
// create_csr_edge, rules R1-R3):
//   K1 degree histogram over `src`            (8 B/edge read, 4 B atomic per edge)
//   K2 exclusive scan -> v[0..V+1] (int64)    (scan.cu)
//   K3 stable scatter = stable radix sort by src carrying dst (+ input row)   (radix_sort.cu)
// and the reverse CSR used by the pull-direction MS-BFS (same kernels, key = target).
#include "common.cuh"

namespace pgq {

// ---- K1 ----------------------------------------------------------------------------------
// deg[src]++ ; also validates 0 <= src,dst < V (the reference would write out of bounds).
__global__ void __launch_bounds__(256) degree_hist_kernel(const int64_t *__restrict__ src,
                                                          const int64_t *__restrict__ dst, uint32_t n,
                                                          uint64_t V, uint32_t *__restrict__ deg,
                                                          uint32_t *__restrict__ err) {
	uint32_t stride = gridDim.x * blockDim.x;
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
		uint64_t s = (uint64_t)src[i], d = (uint64_t)dst[i];
		if (s >= V || d >= V) {
			atomicOr(err, 1u);
			continue;
		}
		atomicAdd(&deg[s], 1u);
	}
}

__global__ void __launch_bounds__(256) hist_u32_kernel(const uint32_t *__restrict__ key, uint32_t n,
                                                       uint32_t *__restrict__ deg) {
	uint32_t stride = gridDim.x * blockDim.x;
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) atomicAdd(&deg[key[i]], 1u);
}

// create_csr_vertex counts (host-provided, uploaded) must equal the histogram.
__global__ void __launch_bounds__(256) check_cnt_kernel(const int64_t *__restrict__ cnt,
                                                        const uint32_t *__restrict__ deg, uint32_t V,
                                                        uint32_t *__restrict__ err) {
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < V && cnt[i] != (int64_t)deg[i]) atomicOr(err, 2u);
}

// eid_csr[p] = eid_in[row[p]]
__global__ void __launch_bounds__(256) gather_eid_kernel(const int64_t *__restrict__ eid_in,
                                                         const uint32_t *__restrict__ row, uint32_t n,
                                                         int64_t *__restrict__ out) {
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) out[i] = eid_in[row[i]];
}

// src_of[p] = vertex whose forward range contains slot p  (upper_bound(off, p) - 1)
__global__ void __launch_bounds__(256) expand_src_kernel(const int64_t *__restrict__ off, uint32_t V, uint32_t n,
                                                         uint32_t *__restrict__ out) {
	uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= n) return;
	uint32_t lo = 0, hi = V; // find first index in [0,V] with off[idx] > p
	while (lo < hi) {
		uint32_t mid = lo + ((hi - lo) >> 1);
		if (off[mid] > (int64_t)p) hi = mid;
		else lo = mid + 1;
	}
	out[p] = lo - 1;
}

static int key_bits_for(int64_t V) {
	int b = 1;
	while (b < 32 && ((int64_t)1 << b) < V) b++;
	return b;
}

int csr_device_build(Ctx *c, Csr *g) {
	const int64_t V = g->V, E = g->E;
	if (V >= ((int64_t)1 << 32) - 2 || E >= ((int64_t)1 << 32) - 1) {
		set_error("csr too large for one GPU slice (V=%lld E=%lld; 32-bit device ids)", (long long)V,
		          (long long)E);
		return PGQ_ERR_INVALID;
	}
	cudaStream_t st = c->stream;
	// outputs first, so the event-timed region below holds kernels only
	PGQ_TRY(g->d_off.reserve((size_t)(V + 2) * sizeof(int64_t)));
	PGQ_TRY(g->d_tgt.reserve((size_t)(E > 0 ? E : 1) * sizeof(uint32_t)));
	if (g->keep_rows) PGQ_TRY(g->d_row.reserve((size_t)(E > 0 ? E : 1) * sizeof(uint32_t)));
	if (g->have_eid && E > 0) PGQ_TRY(g->d_eid.reserve((size_t)E * sizeof(int64_t)));
	// deg[V+2] (two trailing zeros so the exclusive scan yields v[V] = v[V+1] = E)
	TmpBuf deg;
	PGQ_TRY(deg.reserve((size_t)(V + 2) * sizeof(uint32_t) + 16));
	uint32_t *d_deg = deg.as<uint32_t>();
	uint32_t *d_err = d_deg + (V + 2);
	PGQ_CUDA(cudaEventRecord(c->ev_a, st));
	PGQ_CUDA(cudaMemsetAsync(d_deg, 0, (size_t)(V + 2) * sizeof(uint32_t) + 16, st));
	const int64_t *s64 = g->d_src64.as<int64_t>(), *d64 = g->d_dst64.as<int64_t>();
	if (E > 0) {
		uint32_t grid = div_up(E, 256);
		uint32_t cap = (uint32_t)c->sm_count * 32;
		if (grid > cap) grid = cap;
		degree_hist_kernel<<<grid, 256, 0, st>>>(s64, d64, (uint32_t)E, (uint64_t)V, d_deg, d_err);
		count_launch(c);
	}
	TmpBuf d_cnt;
	if (g->have_cnt && V > 0) {
		PGQ_TRY(d_cnt.reserve((size_t)V * sizeof(int64_t)));
		PGQ_CUDA(cudaMemcpyAsync(d_cnt.p, g->cnt_host.data(), (size_t)V * sizeof(int64_t),
		                         cudaMemcpyHostToDevice, st));
		c->stats.h2d_bytes += (uint64_t)V * sizeof(int64_t);
		check_cnt_kernel<<<div_up(V, 256), 256, 0, st>>>(d_cnt.as<int64_t>(), d_deg, (uint32_t)V, d_err);
		count_launch(c);
	}
	// K2
	PGQ_TRY(exclusive_scan_u32_to_i64(c, d_deg, g->d_off.as<int64_t>(), (uint64_t)(V + 2)));
	// K3
	if (E > 0) {
		SortIO io;
		io.k64 = s64;
		io.p64 = d64;
		io.n = (uint32_t)E;
		io.key_bits = key_bits_for(V);
		io.out_p0 = g->d_tgt.as<uint32_t>();
		io.out_p1 = g->keep_rows ? g->d_row.as<uint32_t>() : nullptr;
		PGQ_TRY(radix_sort_stable(c, io));
		if (g->have_eid) {
			gather_eid_kernel<<<div_up(E, 256), 256, 0, st>>>(g->d_eid64.as<int64_t>(), g->d_row.as<uint32_t>(),
			                                                  (uint32_t)E, g->d_eid.as<int64_t>());
			count_launch(c);
		}
	}
	PGQ_CUDA(cudaEventRecord(c->ev_b, st));
	uint32_t h_err = 0;
	PGQ_CUDA(cudaMemcpyAsync(&h_err, d_err, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
	PGQ_CUDA(cudaStreamSynchronize(st));
	float ms = 0;
	PGQ_CUDA(cudaEventElapsedTime(&ms, c->ev_a, c->ev_b));
	c->stats.csr_build_ms = ms;
	deg.release();
	d_cnt.release();
	if (h_err & 1u) {
		set_error("create_csr_edge: src/dst id outside [0, v_size)");
		return PGQ_ERR_INVALID;
	}
	if (h_err & 2u) {
		set_error("create_csr_vertex counts disagree with the edge rows");
		return PGQ_ERR_MISMATCH;
	}
	return PGQ_OK;
}

// ---- reverse CSR + in-degree bins ------------------------------------------------------------

struct BinCounters {
	uint32_t n_mid, n_hub, n_chunks, pad;
};

__global__ void __launch_bounds__(256) bin_kernel(const uint32_t *__restrict__ ioff, uint32_t V,
                                                  uint32_t small_max, uint32_t hub_min, uint32_t chunk,
                                                  uint32_t *__restrict__ mid, uint32_t *__restrict__ hub,
                                                  HubChunk *__restrict__ chunks, BinCounters *__restrict__ cnt) {
	uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
	if (v >= V) return;
	uint32_t lo = ioff[v], hi = ioff[v + 1], deg = hi - lo;
	if (deg > hub_min) {
		uint32_t h = atomicAdd(&cnt->n_hub, 1u);
		hub[h] = v;
		uint32_t k = (deg + chunk - 1) / chunk;
		uint32_t c0 = atomicAdd(&cnt->n_chunks, k);
		for (uint32_t j = 0; j < k; j++) {
			HubChunk hc;
			hc.hub = h;
			hc.begin = lo + j * chunk;
			hc.end = min(hi, lo + (j + 1) * chunk);
			chunks[c0 + j] = hc;
		}
	} else if (deg > small_max) {
		uint32_t m = atomicAdd(&cnt->n_mid, 1u);
		mid[m] = v;
	}
}

int csr_prepare(Ctx *c, Csr *g) {
	const int64_t V = g->V, E = g->E;
	cudaStream_t st = c->stream;
	const bool need_rev = !g->prepared;
	const bool need_bins = need_rev || g->bins_small_max != c->opt.small_max || g->bins_hub_min != c->opt.hub_min;
	if (!need_bins) return PGQ_OK;
	PGQ_CUDA(cudaEventRecord(c->ev_a, st));
	if (need_rev) {
		TmpBuf ideg, srcof;
		PGQ_TRY(ideg.reserve((size_t)(V + 1) * sizeof(uint32_t)));
		PGQ_CUDA(cudaMemsetAsync(ideg.p, 0, (size_t)(V + 1) * sizeof(uint32_t), st));
		PGQ_TRY(g->d_ioff.reserve((size_t)(V + 1) * sizeof(uint32_t)));
		PGQ_TRY(g->d_isrc.reserve((size_t)(E > 0 ? E : 1) * sizeof(uint32_t)));
		PGQ_TRY(g->d_ipos.reserve((size_t)(E > 0 ? E : 1) * sizeof(uint32_t)));
		if (E > 0) {
			uint32_t grid = div_up(E, 256);
			uint32_t cap = (uint32_t)c->sm_count * 32;
			hist_u32_kernel<<<grid > cap ? cap : grid, 256, 0, st>>>(g->d_tgt.as<uint32_t>(), (uint32_t)E,
			                                                         ideg.as<uint32_t>());
			count_launch(c);
		}
		PGQ_TRY(exclusive_scan_u32(c, ideg.as<uint32_t>(), g->d_ioff.as<uint32_t>(), (uint64_t)(V + 1), nullptr));
		if (E > 0) {
			PGQ_TRY(srcof.reserve((size_t)E * sizeof(uint32_t)));
			expand_src_kernel<<<div_up(E, 256), 256, 0, st>>>(g->d_off.as<int64_t>(), (uint32_t)V, (uint32_t)E,
			                                                  srcof.as<uint32_t>());
			count_launch(c);
			SortIO io;
			io.k32 = g->d_tgt.as<uint32_t>();
			io.p32 = srcof.as<uint32_t>();
			io.n = (uint32_t)E;
			io.key_bits = key_bits_for(V);
			io.out_p0 = g->d_isrc.as<uint32_t>();
			io.out_p1 = g->d_ipos.as<uint32_t>();
			PGQ_TRY(radix_sort_stable(c, io));
		}
		PGQ_CUDA(cudaStreamSynchronize(st));
		ideg.release();
		srcof.release();
		g->prepared = true;
	}
	// bins
	{
		uint32_t small_max = (uint32_t)c->opt.small_max, hub_min = (uint32_t)c->opt.hub_min;
		uint32_t chunk = (uint32_t)c->opt.hub_chunk;
		if (hub_min < small_max) hub_min = small_max;
		size_t max_mid = (size_t)(E / ((int64_t)small_max + 1)) + 1;
		size_t max_hub = (size_t)(E / ((int64_t)hub_min + 1)) + 1;
		size_t max_chunks = (size_t)(E / chunk) + max_hub + 1;
		PGQ_TRY(g->d_mid.reserve(max_mid * sizeof(uint32_t)));
		PGQ_TRY(g->d_hub.reserve(max_hub * sizeof(uint32_t)));
		PGQ_TRY(g->d_hubchunks.reserve(max_chunks * sizeof(HubChunk)));
		TmpBuf cnt;
		PGQ_TRY(cnt.reserve(sizeof(BinCounters)));
		PGQ_CUDA(cudaMemsetAsync(cnt.p, 0, sizeof(BinCounters), st));
		if (V > 0) {
			bin_kernel<<<div_up(V, 256), 256, 0, st>>>(g->d_ioff.as<uint32_t>(), (uint32_t)V, small_max, hub_min,
			                                           chunk, g->d_mid.as<uint32_t>(), g->d_hub.as<uint32_t>(),
			                                           g->d_hubchunks.as<HubChunk>(), cnt.as<BinCounters>());
			count_launch(c);
		}
		BinCounters h{};
		PGQ_CUDA(cudaMemcpyAsync(&h, cnt.p, sizeof(h), cudaMemcpyDeviceToHost, st));
		PGQ_CUDA(cudaEventRecord(c->ev_b, st));
		PGQ_CUDA(cudaStreamSynchronize(st));
		cnt.release();
		g->n_mid = h.n_mid;
		g->n_hub = h.n_hub;
		g->n_hubchunks = h.n_chunks;
		g->bins_small_max = c->opt.small_max;
		g->bins_hub_min = c->opt.hub_min;
	}
	if (need_rev) {
		float ms = 0;
		PGQ_CUDA(cudaEventElapsedTime(&ms, c->ev_a, c->ev_b));
		c->stats.transpose_ms = ms;
	}
	PGQ_CUDA(cudaGetLastError());
	return PGQ_OK;
}

void Csr::free_all() {
	if (h_src) cudaFreeHost(h_src);
	if (h_dst) cudaFreeHost(h_dst);
	if (h_eid) cudaFreeHost(h_eid);
	h_src = h_dst = h_eid = nullptr;
	d_src64.release(); d_dst64.release(); d_eid64.release();
	d_off.release(); d_tgt.release(); d_row.release(); d_eid.release();
	d_ioff.release(); d_isrc.release(); d_ipos.release();
	d_mid.release(); d_hub.release(); d_hubchunks.release();
}

} // namespace pgq
