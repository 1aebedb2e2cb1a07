// CSR construction on the device (SURVEY.md §8(a) a4 = restatement of create_csr_vertex +
// create_csr_edge, rules R1-R3):
//   K1 degree histogram over `src`            (8 B/edge read, 4 B atomic per edge)
//   K2 exclusive scan -> v[0..V+1] (int64)    (scan.cu)
//   K3 stable scatter = stable radix sort by src carrying dst (+ input row)   (radix_sort.cu)
// and the reverse CSR used by the pull-direction MS-BFS (same kernels, key = target).
#include "common.cuh"

#include <atomic>

namespace pgq {

// ---- K1+K2 ----------------------------------------------------------------------------------
// Offsets from the SORTED key array (no atomics, no hot hub counters): position p starts the range
// of key[p] when key[p-1] differs; the vertices in between are empty and start at p too.
//   off[s] = first position whose key >= s,  off[nv .. nv+tail) = n.
template <typename OutT>
__global__ void __launch_bounds__(256) offsets_from_sorted_kernel(const uint32_t *__restrict__ key, uint32_t n,
                                                                  uint32_t nv, uint32_t tail,
                                                                  OutT *__restrict__ off) {
	// four consecutive keys per thread (one 16-byte load) + the key before them
	const uint64_t p0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
	if (p0 > n) return;
	uint32_t kk[4];
	if (p0 + 4 <= n) {
		const uint4 q = *reinterpret_cast<const uint4 *>(key + p0);
		kk[0] = q.x; kk[1] = q.y; kk[2] = q.z; kk[3] = q.w;
	} else {
#pragma unroll
		for (int j = 0; j < 4; j++) kk[j] = p0 + j < n ? key[p0 + j] : 0u;
	}
	uint32_t first = p0 ? key[p0 - 1] + 1 : 0; // keys (first .. k] start at this position
#pragma unroll
	for (int j = 0; j < 4; j++) {
		const uint64_t p = p0 + j;
		if (p > n) break;
		if (p == n) { // closes the array
			for (uint32_t s = first; s < nv + tail; s++) off[s] = (OutT)n;
			break;
		}
		const uint32_t k = kk[j];
		for (uint32_t s = first; s <= k && s < nv; s++) off[s] = (OutT)p;
		first = k + 1;
	}
}

// create_csr_vertex counts (host-provided, uploaded) must equal the histogram.
__global__ void __launch_bounds__(256) check_cnt_kernel(const int64_t *__restrict__ cnt,
                                                        const int64_t *__restrict__ off, uint32_t V,
                                                        uint32_t *__restrict__ err) {
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < V && cnt[i] != off[i + 1] - off[i]) atomicOr(err, 2u);
}

// eid_csr[p] = eid_in[row[p]]
__global__ void __launch_bounds__(256) gather_eid_kernel(const int64_t *__restrict__ eid_in,
                                                         const uint32_t *__restrict__ row, uint32_t n,
                                                         int64_t *__restrict__ out) {
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) out[i] = eid_in[row[i]];
}

int part_offsets_u32(Ctx *c, const uint32_t *keys, uint32_t n, uint32_t nv, uint32_t *off) {
	offsets_from_sorted_kernel<uint32_t><<<div_up(div_up((uint64_t)n + 1, 4), 256), 256, 0, c->stream>>>(keys, n, nv, 1u, off);
	count_launch(c);
	PGQ_CUDA(cudaGetLastError());
	return PGQ_OK;
}

static int key_bits_for(int64_t V) {
	int b = 1;
	while (b < 32 && ((int64_t)1 << b) < V) b++;
	return b;
}

static std::atomic<uint64_t> g_csr_gen{1};

int csr_device_build(Ctx *c, Csr *g) {
	const int64_t V = g->V, E = g->E;
	g->gen = g_csr_gen.fetch_add(1);
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
	TmpBuf errbuf;
	PGQ_TRY(errbuf.reserve(16));
	uint32_t *d_err = errbuf.as<uint32_t>();
	PGQ_TRY(g->d_fsrc.reserve((size_t)(E > 0 ? E : 1) * sizeof(uint32_t)));
	PGQ_CUDA(cudaEventRecord(c->ev_a, st));
	PGQ_CUDA(cudaMemsetAsync(d_err, 0, 16, st));
	const int64_t *s64 = g->d_src64.as<int64_t>(), *d64 = g->d_dst64.as<int64_t>();
	// K3 first: stable sort by src; its sorted keys give K1+K2 (offsets) without a histogram
	if (E > 0) {
		SortIO io;
		io.k64 = s64;
		io.p64 = d64;
		io.n = (uint32_t)E;
		io.key_bits = key_bits_for(V);
		io.out_p0 = g->d_tgt.as<uint32_t>();
		io.out_p1 = g->keep_rows ? g->d_row.as<uint32_t>() : nullptr;
		io.out_keys = g->d_fsrc.as<uint32_t>();
		io.vmax = (uint64_t)V;
		io.err = d_err;
		PGQ_TRY(radix_sort_stable(c, io));
	}
	offsets_from_sorted_kernel<int64_t><<<div_up(div_up((uint64_t)E + 1, 4), 256), 256, 0, st>>>(
	    g->d_fsrc.as<uint32_t>(), (uint32_t)E, (uint32_t)V, 2u, g->d_off.as<int64_t>());
	count_launch(c);
	TmpBuf d_cnt;
	if (g->have_cnt && V > 0) {
		PGQ_TRY(d_cnt.reserve((size_t)V * sizeof(int64_t)));
		PGQ_CUDA(cudaMemcpyAsync(d_cnt.p, g->cnt_host.data(), (size_t)V * sizeof(int64_t),
		                         cudaMemcpyHostToDevice, st));
		c->stats.h2d_bytes += (uint64_t)V * sizeof(int64_t);
		check_cnt_kernel<<<div_up(V, 256), 256, 0, st>>>(d_cnt.as<int64_t>(), g->d_off.as<int64_t>(), (uint32_t)V,
		                                                 d_err);
		count_launch(c);
	}
	if (E > 0) {
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
	d_cnt.release();
	if (h_err & 1u) {
		set_error("create_csr_edge: src/dst id outside [0, v_size)");
		return PGQ_ERR_INVALID;
	}
	if (h_err & 2u) {
		set_error("create_csr_vertex counts disagree with the edge rows");
		return PGQ_ERR_MISMATCH;
	}
	if (h_err & 4u) {
		set_error("csr build: radix sort look-back watchdog fired (a tile never published its bucket totals)");
		return PGQ_ERR_CUDA;
	}
	return PGQ_OK;
}

// ---- reverse CSR + in-degree bins ------------------------------------------------------------

// In-degree classes of the pull kernels: each class is a vertex list in ascending id order
// (flag -> exclusive scan -> scatter), so neighbouring list entries touch neighbouring words.
__global__ void __launch_bounds__(256) class_flag_kernel(const uint32_t *__restrict__ ioff, uint32_t V, uint32_t dlo,
                                                         uint32_t dhi, uint32_t *__restrict__ flag) {
	uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
	if (v >= V) return;
	uint32_t deg = ioff[v + 1] - ioff[v];
	flag[v] = (deg >= dlo && deg <= dhi) ? 1u : 0u;
}

__global__ void __launch_bounds__(256) class_scatter_kernel(const uint32_t *__restrict__ ioff, uint32_t V,
                                                            uint32_t dlo, uint32_t dhi,
                                                            const uint32_t *__restrict__ pos,
                                                            uint32_t *__restrict__ out) {
	uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
	if (v >= V) return;
	uint32_t deg = ioff[v + 1] - ioff[v];
	if (deg >= dlo && deg <= dhi) out[pos[v]] = v;
}

// chunk descriptors of the hub class: in-edges [begin,end) of hub list entry h
__global__ void __launch_bounds__(256) hub_chunk_kernel(const uint32_t *__restrict__ ioff,
                                                        const uint32_t *__restrict__ hub, uint32_t nhub,
                                                        uint32_t chunk, HubChunk *__restrict__ chunks,
                                                        uint32_t *__restrict__ nchunks) {
	uint32_t h = blockIdx.x * blockDim.x + threadIdx.x;
	if (h >= nhub) return;
	uint32_t v = hub[h];
	uint32_t lo = ioff[v], hi = ioff[v + 1];
	uint32_t k = (hi - lo + chunk - 1) / chunk;
	uint32_t c0 = atomicAdd(nchunks, k);
	for (uint32_t j = 0; j < k; j++) {
		HubChunk hc;
		hc.hub = h;
		hc.begin = lo + j * chunk;
		hc.end = min(hi, lo + (j + 1) * chunk);
		chunks[c0 + j] = hc;
	}
}

static int build_class(Ctx *c, Csr *g, uint32_t dlo, uint32_t dhi, uint32_t *flag, uint64_t *d_total, DevBuf *out,
                       uint32_t *n_out) {
	const uint32_t V = (uint32_t)g->V;
	cudaStream_t st = c->stream;
	*n_out = 0;
	if (V == 0 || dlo > dhi) return PGQ_OK;
	class_flag_kernel<<<div_up(V, 256), 256, 0, st>>>(g->d_ioff.as<uint32_t>(), V, dlo, dhi, flag);
	count_launch(c);
	PGQ_TRY(exclusive_scan_u32(c, flag, flag, V, d_total));
	uint64_t total = 0;
	PGQ_CUDA(cudaMemcpyAsync(&total, d_total, sizeof(total), cudaMemcpyDeviceToHost, st));
	PGQ_CUDA(cudaStreamSynchronize(st));
	PGQ_TRY(out->reserve((size_t)(total ? total : 1) * sizeof(uint32_t)));
	if (total) {
		class_scatter_kernel<<<div_up(V, 256), 256, 0, st>>>(g->d_ioff.as<uint32_t>(), V, dlo, dhi, flag,
		                                                    out->as<uint32_t>());
		count_launch(c);
	}
	*n_out = (uint32_t)total;
	return PGQ_OK;
}

int csr_prepare(Ctx *c, Csr *g) {
	const int64_t V = g->V, E = g->E;
	cudaStream_t st = c->stream;
	const bool need_rev = !g->prepared;
	const bool need_bins = need_rev || g->bins_tiny_max != c->opt.tiny_max || g->bins_small_max != c->opt.small_max ||
	                       g->bins_med_max != c->opt.med_max ||
	                       g->bins_hub_min != c->opt.hub_min;
	if (!need_bins) return PGQ_OK;
	g->gen = g_csr_gen.fetch_add(1); // cached loop graphs and replicas of the old arrays are stale
	PGQ_CUDA(cudaEventRecord(c->ev_a, st));
	if (need_rev) {
		const uint32_t *tkeys = nullptr; // sorted targets, left in the sort's scratch
		PGQ_TRY(g->d_ioff.reserve((size_t)(V + 1) * sizeof(uint32_t)));
		PGQ_TRY(g->d_isrc.reserve((size_t)(E > 0 ? E : 1) * sizeof(uint32_t)));
		PGQ_TRY(g->d_ipos.reserve((size_t)(E > 0 ? E : 1) * sizeof(uint32_t)));
		if (E > 0) {
			// key = target, payload = source of each forward slot (= the forward build's sorted keys)
			SortIO io;
			io.k32 = g->d_tgt.as<uint32_t>();
			io.p32 = g->d_fsrc.as<uint32_t>();
			io.n = (uint32_t)E;
			io.key_bits = key_bits_for(V);
			io.out_p0 = g->d_isrc.as<uint32_t>();
			io.out_p1 = g->d_ipos.as<uint32_t>();
			io.keys_scratch = &tkeys;
			PGQ_TRY(radix_sort_stable(c, io));
		}
		offsets_from_sorted_kernel<uint32_t><<<div_up(div_up((uint64_t)E + 1, 4), 256), 256, 0, st>>>(
		    tkeys, (uint32_t)E, (uint32_t)V, 1u, g->d_ioff.as<uint32_t>());
		count_launch(c);
		PGQ_CUDA(cudaStreamSynchronize(st));
		g->d_fsrc.release(); // only needed to seed the reverse build
		g->prepared = true;
	}
	// in-degree classes
	uint32_t med_max_eff = 0;
	{
		uint32_t tiny_max = (uint32_t)c->opt.tiny_max, small_max = (uint32_t)c->opt.small_max;
		uint32_t hub_min = (uint32_t)c->opt.hub_min, chunk = (uint32_t)c->opt.hub_chunk;
		uint32_t med_max = (uint32_t)c->opt.med_max;
		if (small_max < tiny_max) tiny_max = small_max;
		if (med_max < small_max) med_max = small_max;
		med_max_eff = med_max;
		if (hub_min < med_max) hub_min = med_max;
		TmpBuf flag, tot;
		PGQ_TRY(flag.reserve((size_t)(V + 1) * sizeof(uint32_t)));
		PGQ_TRY(tot.reserve(16));
		uint32_t *f = flag.as<uint32_t>();
		uint64_t *dt = tot.as<uint64_t>();
		PGQ_TRY(build_class(c, g, 1, tiny_max, f, dt, &g->d_tiny, &g->n_tiny));
		PGQ_TRY(build_class(c, g, tiny_max + 1, small_max, f, dt, &g->d_small, &g->n_small));
		PGQ_TRY(build_class(c, g, small_max + 1, med_max, f, dt, &g->d_med, &g->n_med));
		PGQ_TRY(build_class(c, g, med_max + 1, hub_min, f, dt, &g->d_mid, &g->n_mid));
		PGQ_TRY(build_class(c, g, hub_min + 1, 0xFFFFFFFFu, f, dt, &g->d_hub, &g->n_hub));
		g->n_hubchunks = 0;
		if (g->n_hub) {
			size_t max_chunks = (size_t)(E / chunk) + g->n_hub + 1;
			PGQ_TRY(g->d_hubchunks.reserve(max_chunks * sizeof(HubChunk)));
			uint32_t *dn = reinterpret_cast<uint32_t *>(dt);
			PGQ_CUDA(cudaMemsetAsync(dn, 0, sizeof(uint32_t), st));
			hub_chunk_kernel<<<div_up(g->n_hub, 256), 256, 0, st>>>(g->d_ioff.as<uint32_t>(), g->d_hub.as<uint32_t>(),
			                                                        g->n_hub, chunk, g->d_hubchunks.as<HubChunk>(), dn);
			count_launch(c);
			PGQ_CUDA(cudaMemcpyAsync(&g->n_hubchunks, dn, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
		}
		PGQ_CUDA(cudaEventRecord(c->ev_b, st));
		PGQ_CUDA(cudaStreamSynchronize(st));
		g->bins_tiny_max = c->opt.tiny_max;
		g->bins_small_max = c->opt.small_max;
		g->bins_med_max = c->opt.med_max;
		g->eff_med_max = med_max_eff; // clamped: d_mid and d_hub hold exactly the vertices above it
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
	if (!replicas.empty()) {
		int cur = 0;
		cudaGetDevice(&cur);
		for (auto &kv : replicas) {
			cudaSetDevice(kv.first);
			kv.second->free_all();
			delete kv.second;
		}
		replicas.clear();
		cudaSetDevice(cur);
	}
	pinned_release(h_src, h_bytes);
	pinned_release(h_dst, h_bytes);
	pinned_release(h_eid, h_bytes);
	h_src = h_dst = h_eid = nullptr;
	d_src64.release(); d_dst64.release(); d_eid64.release();
	d_off.release(); d_tgt.release(); d_row.release(); d_eid.release(); d_fsrc.release();
	d_w_in.release(); d_w.release(); d_iw.release(); d_iw16.release(); d_iw32n.release(); d_wstat.release();
	d_ioff.release(); d_isrc.release(); d_ipos.release();
	d_tiny.release(); d_small.release(); d_med.release(); d_mid.release(); d_hub.release(); d_hubchunks.release();
	d_ltgt.release(); d_lsrc.release();
}

} // namespace pgq
