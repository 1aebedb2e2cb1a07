// cheapest_path_length (rule R10 in DESIGN.md section 5b): lane-batched
// Bellman-Ford over a CSR that carries one 8-byte weight (int64 or double) per edge.
//
// Layout: dist[V][64] of T (512 B per vertex: one coalesced row per warp access).  Each round is a
// bottom-up relaxation over the reverse CSR: one warp per vertex n, 32 in-edges per step — every
// lane loads one (source, weight) pair and tests the source's bit in the "changed last round"
// bitmap; for the sources that pass, the warp reads their 512-byte rows (each thread owns lanes
// `lane` and `lane + 32`) and keeps the minimum of dist[v] + w.  Rows are updated in place
// (8-byte loads/stores are single transactions and distances only decrease, so a relaxation that
// sees a newer value only converges sooner); the fixed point min_v(dist[v] + w) does not depend on
// the order, which is what makes the result bit-identical to the oracle's push formulation for
// int64 and for double weights alike.  A round that changes nothing ends the batch; more than V
// rounds means a negative cycle.
#include "common.cuh"

#include <cmath>
#include <limits>

namespace pgq {

static constexpr int SSSP_LANES = 64;

template <typename T>
struct WInf;
template <>
struct WInf<long long> {
	__host__ __device__ static long long value() { return 0x3FFFFFFFFFFFFFFFll; } // INT64_MAX / 2 (R10)
	__host__ __device__ static bool is_inf(long long x) { return x >= 0x3FFFFFFFFFFFFFFFll; }
};
template <>
struct WInf<double> {
	__host__ __device__ static double value() { return std::numeric_limits<double>::infinity(); }
	__host__ __device__ static bool is_inf(double x) { return x == std::numeric_limits<double>::infinity(); }
};

// w_csr[p] = w_rows[row[p]]  (raw 8-byte values)
__global__ void __launch_bounds__(256) gather_w_kernel(const unsigned long long *__restrict__ in,
                                                       const uint32_t *__restrict__ idx, uint32_t n,
                                                       unsigned long long *__restrict__ out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) out[i] = in[idx[i]];
}

template <typename T>
__global__ void __launch_bounds__(256) sssp_fill_kernel(T *__restrict__ dist, size_t n) {
	const T inf = WInf<T>::value();
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
		dist[i] = inf;
}

template <typename T>
__global__ void sssp_seed_kernel(const uint32_t *__restrict__ lane_src, T *__restrict__ dist,
                                 uint32_t *__restrict__ chg) {
	const int l = threadIdx.x;
	const uint32_t s = lane_src[l];
	if (s == 0xFFFFFFFFu) return;
	dist[(size_t)s * SSSP_LANES + l] = (T)0;
	atomicOr(&chg[s >> 5], 1u << (s & 31));
}

// min over the in-edges base0, base0+step, ... (32 consecutive edges per visit) of dist[src] + w, for
// lanes `lane` and `lane + 32`; only sources whose bit is set in chg_prev are read.
template <typename T>
__device__ __forceinline__ bool sssp_scan(const uint32_t *__restrict__ isrc, const T *__restrict__ iw, const T *dist,
                                          const uint32_t *__restrict__ chg_prev, uint32_t base0, uint32_t hi,
                                          uint32_t step, unsigned lane, T &best0, T &best1) {
	bool touched = false;
	for (uint32_t base = base0; base < hi; base += step) {
		const uint32_t j = base + lane;
		uint32_t v = 0;
		T w = (T)0;
		bool hot = false;
		if (j < hi) {
			v = isrc[j];
			hot = (__ldg(chg_prev + (v >> 5)) >> (v & 31)) & 1u;
			if (hot) w = iw[j];
		}
		unsigned m = __ballot_sync(0xffffffffu, hot);
		touched |= m != 0;
		while (m) { // four source rows (8 loads per thread) in flight per step
			uint32_t sv[4];
			T sw[4], d0[4], d1[4];
			bool on[4];
#pragma unroll
			for (int u = 0; u < 4; u++) {
				on[u] = m != 0;
				const int k = on[u] ? __ffs(m) - 1 : 0;
				m &= m - 1; // 0 stays 0
				sv[u] = __shfl_sync(0xffffffffu, v, k);
				sw[u] = __shfl_sync(0xffffffffu, w, k);
			}
#pragma unroll
			for (int u = 0; u < 4; u++) {
				// volatile: rows are updated in place by other warps; never cache them in registers
				const T *srow = dist + (size_t)sv[u] * SSSP_LANES;
				d0[u] = on[u] ? *reinterpret_cast<const volatile T *>(srow + lane) : WInf<T>::value();
				d1[u] = on[u] ? *reinterpret_cast<const volatile T *>(srow + lane + 32) : WInf<T>::value();
			}
#pragma unroll
			for (int u = 0; u < 4; u++) {
				if (!WInf<T>::is_inf(d0[u])) {
					const T c0 = d0[u] + sw[u];
					if (c0 < best0) best0 = c0;
				}
				if (!WInf<T>::is_inf(d1[u])) {
					const T c1 = d1[u] + sw[u];
					if (c1 < best1) best1 = c1;
				}
			}
		}
	}
	return touched;
}

// one warp per vertex with at most warp_max in-edges
template <typename T>
__global__ void __launch_bounds__(256) sssp_relax_kernel(const uint32_t *__restrict__ ioff,
                                                         const uint32_t *__restrict__ isrc, const T *__restrict__ iw,
                                                         T *dist, const uint32_t *__restrict__ chg_prev,
                                                         uint32_t *__restrict__ chg_next, uint32_t *__restrict__ flag,
                                                         uint32_t V, uint32_t warp_max) {
	const uint32_t n = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const unsigned lane = threadIdx.x & 31;
	if (n >= V) return;
	const uint32_t lo = ioff[n], hi = ioff[n + 1];
	if (lo == hi || hi - lo > warp_max) return;
	T *row = dist + (size_t)n * SSSP_LANES;
	T best0 = WInf<T>::value(), best1 = WInf<T>::value();
	// the vertex's own row is only read when some in-neighbour changed
	if (!sssp_scan<T>(isrc, iw, dist, chg_prev, lo, hi, 32, lane, best0, best1)) return; // warp-uniform
	const T cur0 = row[lane], cur1 = row[lane + 32];
	bool ch = false;
	if (best0 < cur0) {
		row[lane] = best0;
		ch = true;
	}
	if (best1 < cur1) {
		row[lane + 32] = best1;
		ch = true;
	}
	if (__any_sync(0xffffffffu, ch) && lane == 0) {
		atomicOr(&chg_next[n >> 5], 1u << (n & 31));
		*flag = 1u;
	}
}

// one block (8 warps) per vertex of a high in-degree class list; warps take alternate 32-edge groups
template <typename T>
__global__ void __launch_bounds__(256) sssp_relax_big_kernel(const uint32_t *__restrict__ list,
                                                             const uint32_t *__restrict__ ioff,
                                                             const uint32_t *__restrict__ isrc,
                                                             const T *__restrict__ iw, T *dist,
                                                             const uint32_t *__restrict__ chg_prev,
                                                             uint32_t *__restrict__ chg_next,
                                                             uint32_t *__restrict__ flag) {
	__shared__ T part[8][SSSP_LANES];
	__shared__ int s_touched;
	const uint32_t n = list[blockIdx.x];
	const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint32_t lo = ioff[n], hi = ioff[n + 1];
	if (threadIdx.x == 0) s_touched = 0;
	__syncthreads();
	T best0 = WInf<T>::value(), best1 = WInf<T>::value();
	const bool touched = sssp_scan<T>(isrc, iw, dist, chg_prev, lo + warp * 32, hi, 8 * 32, lane, best0, best1);
	part[warp][lane] = best0;
	part[warp][lane + 32] = best1;
	if (touched && lane == 0) s_touched = 1;
	__syncthreads();
	if (!s_touched || threadIdx.x >= SSSP_LANES) return;
	T best = part[0][threadIdx.x];
#pragma unroll
	for (int k = 1; k < 8; k++) {
		const T x = part[k][threadIdx.x];
		if (x < best) best = x;
	}
	T *row = dist + (size_t)n * SSSP_LANES;
	const bool ch = best < row[threadIdx.x];
	if (ch) row[threadIdx.x] = best;
	if (__any_sync(0xffffffffu, ch) && lane == 0) {
		atomicOr(&chg_next[n >> 5], 1u << (n & 31));
		*flag = 1u;
	}
}

template <typename T>
__global__ void sssp_result_kernel(const uint32_t *__restrict__ lane_src, const uint32_t *__restrict__ lane_dst,
                                   const T *__restrict__ dist, T *__restrict__ out) {
	const int l = threadIdx.x;
	out[l] = lane_src[l] == 0xFFFFFFFFu ? WInf<T>::value() : dist[(size_t)lane_dst[l] * SSSP_LANES + l];
}

// ---- weights ---------------------------------------------------------------------------------

int csr_set_weights(Ctx *c, Csr *g, int wtype, const void *w, int64_t row0, int64_t n) {
	if (wtype != PGQ_WEIGHT_I64 && wtype != PGQ_WEIGHT_F64) {
		set_error("weight type must be PGQ_WEIGHT_I64 or PGQ_WEIGHT_F64");
		return PGQ_ERR_INVALID;
	}
	if (g->is_part || !g->keep_rows) {
		set_error("weights need a CSR built with edge ids or PGQ_CSR_KEEP_ROWS (row tracking)");
		return PGQ_ERR_INVALID;
	}
	if (g->w_rows == 0) g->wtype = wtype;
	if (wtype != g->wtype || row0 != g->w_rows || n < 0 || row0 + n > g->E) {
		set_error("weights must arrive in input-row order, one type, %lld rows expected next", (long long)g->w_rows);
		return PGQ_ERR_INVALID;
	}
	cudaStream_t st = c->stream;
	if (g->w_rows == 0) {
		g->w_ready = false;
		g->d_iw.release();
		PGQ_TRY(g->d_w_in.reserve((size_t)(g->E > 0 ? g->E : 1) * 8));
	}
	if (n) {
		PGQ_CUDA(cudaMemcpyAsync(static_cast<char *>(g->d_w_in.p) + (size_t)row0 * 8, w, (size_t)n * 8,
		                         cudaMemcpyHostToDevice, st));
		PGQ_CUDA(cudaStreamSynchronize(st)); // w is the caller's buffer
		c->stats.h2d_bytes += (uint64_t)n * 8;
	}
	g->w_rows += n;
	if (g->w_rows == g->E) { // all rows here: bring them into CSR slot order (same slot rule as the targets)
		PGQ_TRY(g->d_w.reserve((size_t)(g->E > 0 ? g->E : 1) * 8));
		if (g->E > 0) {
			gather_w_kernel<<<div_up(g->E, 256), 256, 0, st>>>(g->d_w_in.as<unsigned long long>(),
			                                                   g->d_row.as<uint32_t>(), (uint32_t)g->E,
			                                                   g->d_w.as<unsigned long long>());
			count_launch(c);
			PGQ_CUDA(cudaStreamSynchronize(st));
		}
		g->d_w_in.release();
		g->w_ready = true;
		g->w_rows = 0; // a later call starts a fresh set of weights
	}
	return PGQ_OK;
}

int csr_download_weights(Ctx *c, Csr *g, void *out) {
	if (!g->w_ready) {
		set_error("csr has no weights");
		return PGQ_ERR_STATE;
	}
	if (g->E > 0) {
		PGQ_CUDA(cudaMemcpyAsync(out, g->d_w.p, (size_t)g->E * 8, cudaMemcpyDeviceToHost, c->stream));
		PGQ_CUDA(cudaStreamSynchronize(c->stream));
		c->stats.d2h_bytes += (uint64_t)g->E * 8;
	}
	return PGQ_OK;
}

// ---- batches -----------------------------------------------------------------------------------

template <typename T>
static int sssp_batches(Ctx *c, Csr *g, const LaneTable &lt, T *h_out) {
	cudaStream_t st = c->stream;
	const int L = SSSP_LANES;
	const uint32_t V = (uint32_t)g->V;
	const size_t cells = (size_t)V * L;
	const uint32_t cw = div_up((uint64_t)V, 32) + 1;
	PGQ_TRY(c->sssp_dist.reserve(cells * sizeof(T) + 16));
	PGQ_TRY(c->sssp_chg.reserve((size_t)2 * cw * sizeof(uint32_t) + 16));
	const size_t nslots = (size_t)lt.nbatch * L;
	PGQ_TRY(c->bfs_lane.reserve(nslots * (2 * sizeof(uint32_t) + sizeof(T)) + 64));
	T *d_out = c->bfs_lane.as<T>(); // 8-byte aligned first
	uint32_t *d_src = reinterpret_cast<uint32_t *>(d_out + nslots);
	uint32_t *d_dst = d_src + nslots;
	uint32_t *d_flag = c->sssp_chg.as<uint32_t>() + (size_t)2 * cw;
	PGQ_CUDA(cudaMemcpyAsync(d_src, lt.src.data(), nslots * 4, cudaMemcpyHostToDevice, st));
	PGQ_CUDA(cudaMemcpyAsync(d_dst, lt.dst.data(), nslots * 4, cudaMemcpyHostToDevice, st));
	c->stats.h2d_bytes += (uint64_t)nslots * 8;
	T *dist = c->sssp_dist.as<T>();
	uint32_t *chg[2] = {c->sssp_chg.as<uint32_t>(), c->sssp_chg.as<uint32_t>() + cw};
	const T *iw = g->d_iw.as<T>();
	uint32_t *h_flag = reinterpret_cast<uint32_t *>(c->h_ctrl);
	for (int64_t b = 0; b < lt.nbatch; b++) {
		sssp_fill_kernel<T><<<c->sm_count * 8, 256, 0, st>>>(dist, cells);
		PGQ_CUDA(cudaMemsetAsync(chg[0], 0, (size_t)cw * 4, st));
		sssp_seed_kernel<T><<<1, L, 0, st>>>(d_src + b * L, dist, chg[0]);
		count_launch(c, 2);
		c->stats.batches++;
		int cur = 0;
		for (int64_t round = 1;; round++) {
			if (round > (int64_t)V + 1) {
				set_error("cheapest_path_length: negative cycle reachable from a source");
				return PGQ_ERR_INVALID;
			}
			PGQ_CUDA(cudaMemsetAsync(chg[cur ^ 1], 0, (size_t)cw * 4, st));
			PGQ_CUDA(cudaMemsetAsync(d_flag, 0, 4, st));
			sssp_relax_kernel<T><<<div_up((uint64_t)V * 32, 256), 256, 0, st>>>(
			    g->d_ioff.as<uint32_t>(), g->d_isrc.as<uint32_t>(), iw, dist, chg[cur], chg[cur ^ 1], d_flag, V,
			    g->eff_med_max);
			count_launch(c);
			// in-degree above med_max: exactly the `mid` and `hub` class lists of csr_prepare
			if (g->n_mid) {
				sssp_relax_big_kernel<T><<<g->n_mid, 256, 0, st>>>(g->d_mid.as<uint32_t>(), g->d_ioff.as<uint32_t>(),
				                                                 g->d_isrc.as<uint32_t>(), iw, dist, chg[cur],
				                                                 chg[cur ^ 1], d_flag);
				count_launch(c);
			}
			if (g->n_hub) {
				sssp_relax_big_kernel<T><<<g->n_hub, 256, 0, st>>>(g->d_hub.as<uint32_t>(), g->d_ioff.as<uint32_t>(),
				                                                 g->d_isrc.as<uint32_t>(), iw, dist, chg[cur],
				                                                 chg[cur ^ 1], d_flag);
				count_launch(c);
			}
			PGQ_CUDA(cudaMemcpyAsync(h_flag, d_flag, 4, cudaMemcpyDeviceToHost, st));
			PGQ_CUDA(cudaStreamSynchronize(st));
			c->stats.levels++;
			if (!*h_flag) break;
			cur ^= 1;
		}
		sssp_result_kernel<T><<<1, L, 0, st>>>(d_src + b * L, d_dst + b * L, dist, d_out + b * L);
		count_launch(c);
	}
	PGQ_CUDA(cudaMemcpyAsync(h_out, d_out, nslots * sizeof(T), cudaMemcpyDeviceToHost, st));
	PGQ_CUDA(cudaStreamSynchronize(st));
	c->stats.d2h_bytes += (uint64_t)nslots * sizeof(T);
	PGQ_CUDA(cudaGetLastError());
	return PGQ_OK;
}

// out8[slot] = distance of the lane slot (raw 8 bytes of T), +inf when the destination was not reached
int sssp_run(Ctx *c, Csr *g, const LaneTable &lt, void *h_out8) {
	if (lt.nbatch == 0) return PGQ_OK;
	if (lt.W != SSSP_LANES) {
		set_error("internal: cheapest_path_length lane tables are %d wide", SSSP_LANES);
		return PGQ_ERR_INVALID;
	}
	PGQ_TRY(csr_prepare(c, g));
	if (!g->d_iw.p && g->E > 0) { // weights in reverse-CSR order: iw[j] = w[ipos[j]]
		PGQ_TRY(g->d_iw.reserve((size_t)g->E * 8));
		gather_w_kernel<<<div_up(g->E, 256), 256, 0, c->stream>>>(g->d_w.as<unsigned long long>(),
		                                                          g->d_ipos.as<uint32_t>(), (uint32_t)g->E,
		                                                          g->d_iw.as<unsigned long long>());
		count_launch(c);
	}
	if (g->wtype == PGQ_WEIGHT_I64) return sssp_batches<long long>(c, g, lt, static_cast<long long *>(h_out8));
	return sssp_batches<double>(c, g, lt, static_cast<double *>(h_out8));
}

} // namespace pgq
