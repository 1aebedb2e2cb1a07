// cheapest_path_length on narrow distance rows: 16-bit and 32-bit lanes (included by sssp.cu only).
//
// For non-negative int64 weights the relaxation rounds run on rows of 64 narrow unsigned lanes:
//   N16: uint16 lanes, one 128-byte line per vertex; finite distances 0 .. 0x3FFF, "not reached" 0x8000, weights
//        clamped to 0x4000 (uint16 per edge);
//   N32: uint32 lanes, 256 bytes per vertex; finite 0 .. 0x3FFFFFFF, "not reached" 0x80000000, weights clamped to
//        0x40000000 (uint32 per edge).
//   A sum of a finite distance and a weight stays below "not reached"; "not reached" plus a weight stays below the
//   wrap: nothing wraps, and unsigned minima keep every real candidate below every unreached one.  A sum with the
//   second-highest bit set is a real distance that does not fit (or went through a clamped weight): if such a value is
//   ever ACCEPTED into a row -- it beats the row AND the lane's bound, the destination's current distance -- the batch
//   is flagged, stops, and is redone on the next wider rows (16 -> 32 -> 8-byte).  Distances are exact integers at
//   every width: same answers.
// A thread owns 16 bytes of a row (8 or 4 lanes), so 8 (N16) or 16 (N32) threads own a row and a warp relaxes four or
// two source rows per load instruction; the relaxation itself is one DPX instruction per 32-bit word:
// min(best, dist + w) = VIADDMNMX.U16x2 (__viaddmin_u16x2) or VIADDMNMX.U32 (__viaddmin_u32), native on sm_100a.
// Hot in-edges of a 32-edge step are compacted through a per-warp shared-memory stage so that each thread group
// picks its own source row.
#pragma once

namespace pgq {

// Source rows may come from this SM's L1: a row that another SM updated during this round can then be read stale,
// which is harmless -- whoever updates a row also sets its bit in the next round's changed-bitmap, so the fresh value
// is read next round (L1 is invalidated at every kernel boundary), and the round that changes nothing still reads only
// final values.  High out-degree sources are read from many in-lists per round, which is what L1 catches.
// Measured (gpurun check 17c, C4 graph, 1024 pairs): 86.6 ms through L1 against 84.6 ms from L2 only -- L2 only it is.
#ifdef PGQ_SSSP_ROWS_L1
#define SSSPN_ROW_LOAD(p) __ldca(p)
#else
#define SSSPN_ROW_LOAD(p) __ldcg(p)
#endif

struct N16 {
	using S = uint16_t;                        // a lane in memory, also a weight in memory
	static constexpr int TPR = 8;              // threads per row (16 bytes each)
	static constexpr int LPW = 2;              // lanes per 32-bit word
	static constexpr uint32_t INF = 0x8000u;   // one lane: not reached
	static constexpr uint32_t INFW = 0x80008000u; // a word of them
	static constexpr uint32_t LOSTW = 0x40004000u; // per lane: a distance that does not fit
	static constexpr long long WCLAMP = 0x4000;
	__device__ static uint32_t wword(S w) { return (uint32_t)w * 0x10001u; }
	__device__ static uint32_t relax(uint32_t d, uint32_t w, uint32_t best) { return __viaddmin_u16x2(d, w, best); }
	__device__ static uint32_t vmin(uint32_t a, uint32_t b) { return __vminu2(a, b); }
	__device__ static uint32_t ltmask(uint32_t a, uint32_t b) { return __vcmpltu2(a, b); } // all-ones per lane a < b
	__device__ static uint32_t lane_of(uint32_t word, int i) { return i ? word >> 16 : word & 0xFFFFu; }
	__device__ static uint32_t pack(const uint32_t *lanes) { return lanes[0] | (lanes[1] << 16); }
};

struct N32 {
	using S = uint32_t;
	static constexpr int TPR = 16;
	static constexpr int LPW = 1;
	static constexpr uint32_t INF = 0x80000000u;
	static constexpr uint32_t INFW = 0x80000000u;
	static constexpr uint32_t LOSTW = 0x40000000u;
	static constexpr long long WCLAMP = 0x40000000ll;
	__device__ static uint32_t wword(S w) { return w; }
	__device__ static uint32_t relax(uint32_t d, uint32_t w, uint32_t best) { return __viaddmin_u32(d, w, best); }
	__device__ static uint32_t vmin(uint32_t a, uint32_t b) { return a < b ? a : b; }
	__device__ static uint32_t ltmask(uint32_t a, uint32_t b) { return a < b ? 0xFFFFFFFFu : 0u; }
	__device__ static uint32_t lane_of(uint32_t word, int) { return word; }
	__device__ static uint32_t pack(const uint32_t *lanes) { return lanes[0]; }
};

struct StageN {
	uint32_t v[32];
	uint32_t w[32];
};

// int64 weights (known non-negative) clamped for the narrow rows
template <class N>
__global__ void __launch_bounds__(256) narrow_w_kernel(const long long *__restrict__ in, uint32_t n,
                                                       typename N::S *__restrict__ out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) out[i] = in[i] > N::WCLAMP ? (typename N::S)N::WCLAMP : (typename N::S)in[i];
}

template <class N>
__global__ void __launch_bounds__(256) ssspn_fill_kernel(uint4 *__restrict__ rows, size_t n16) {
	const uint4 inf = make_uint4(N::INFW, N::INFW, N::INFW, N::INFW);
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x)
		rows[i] = inf;
}

template <class N>
__global__ void ssspn_seed_kernel(const uint32_t *__restrict__ lane_src, typename N::S *__restrict__ dist,
                                  uint32_t *__restrict__ chg, SsspCtl *__restrict__ ctl, uint32_t mark_max) {
	using S = typename N::S;
	const int l = threadIdx.x;
	if (l == 0) {
		ctl->done = 0;
		ctl->use_mark = SSSP_LANES <= mark_max ? 1u : 0u; // at most 64 sources changed
		ctl->n_changed = 0;
		ctl->rounds = 0;
		ctl->overflow_seen = 0;
		ctl->hot_seen = ctl->scanned_seen = 0;
		ctl->overflow = 0;
		ctl->hot = 0;
		ctl->scanned = 0;
	}
	reinterpret_cast<S *>(ctl->bound)[l] = (S)N::INF;
	const uint32_t s = lane_src[l];
	if (s == 0xFFFFFFFFu) return;
	dist[(size_t)s * SSSP_LANES + l] = 0;
	atomicOr(&chg[s >> 5], 1u << (s & 31));
}

template <class N>
__global__ void __launch_bounds__(1024) ssspn_round_end_kernel(const uint32_t *__restrict__ lane_src,
                                                               const uint32_t *__restrict__ lane_dst,
                                                               const typename N::S *__restrict__ dist,
                                                               const uint32_t *__restrict__ chg_next, uint32_t nwords,
                                                               SsspCtl *__restrict__ ctl, int prune, uint32_t mark_max) {
	using S = typename N::S;
	if (ctl->done) return;
	const uint32_t any = sssp_count_changed(chg_next, nwords, ctl);
	const int l = threadIdx.x;
	if (l < SSSP_LANES && prune && lane_src[l] != 0xFFFFFFFFu)
		reinterpret_cast<S *>(ctl->bound)[l] = dist[(size_t)lane_dst[l] * SSSP_LANES + l];
	if (l == 0) sssp_close_round(ctl, any, true, mark_max); // an overflowed batch is redone on wider rows anyway
}

template <class N>
__global__ void ssspn_result_kernel(const uint32_t *__restrict__ lane_src, const uint32_t *__restrict__ lane_dst,
                                    const typename N::S *__restrict__ dist, long long *__restrict__ out) {
	const int l = threadIdx.x;
	long long r = WInf<long long>::value();
	if (lane_src[l] != 0xFFFFFFFFu) {
		const uint32_t d = dist[(size_t)lane_dst[l] * SSSP_LANES + l];
		if (d < N::INF) r = (long long)d;
	}
	out[l] = r;
}

// min over the in-edges base0, base0+step, ... (32 consecutive edges per visit) of dist[src] + w.  On return thread
// (q = lane % TPR, g = lane / TPR) holds, in best[0..3], the minima of its 16 bytes of the row over the source rows
// its group g took; ssspn_fold merges the groups.
template <class N, bool STATS>
__device__ __forceinline__ bool ssspn_scan(const uint32_t *__restrict__ isrc, const typename N::S *__restrict__ iw,
                                           const typename N::S *dist, const uint32_t *__restrict__ chg_prev,
                                           uint32_t base0, uint32_t hi, uint32_t step, unsigned lane, StageN &sg,
                                           uint32_t (&best)[4], SsspCtl *ctl) {
	constexpr int TPR = N::TPR, G = 32 / TPR;
	const unsigned q = lane % TPR, g = lane / TPR;
	const uint4 *rows = reinterpret_cast<const uint4 *>(dist);
	bool touched = false;
	uint32_t n_hot = 0, n_scan = 0;
	for (uint32_t base = base0; base < hi; base += step) {
		const uint32_t j = base + lane;
		uint32_t v = 0;
		bool hot = false;
		if (j < hi) {
			v = isrc[j];
			hot = (__ldg(chg_prev + (v >> 5)) >> (v & 31)) & 1u;
		}
		const unsigned m = __ballot_sync(0xffffffffu, hot);
		if (STATS) n_scan += min(32u, hi - base);
		if (!m) continue; // warp-uniform
		touched = true;
		if (hot) {
			const int r = __popc(m & ((1u << lane) - 1u));
			sg.v[r] = v;
			sg.w[r] = N::wword(iw[j]);
		}
		__syncwarp();
		const int cnt = __popc(m);
		if (STATS) n_hot += cnt;
		for (int i = (int)g; i < cnt; i += 4 * G) { // up to four source rows per thread in flight
			uint4 d[4];
			uint32_t w[4];
#pragma unroll
			for (int u = 0; u < 4; u++) {
				const int k = i + G * u;
				d[u] = make_uint4(N::INFW, N::INFW, N::INFW, N::INFW);
				w[u] = 0;
				if (k < cnt) {
					w[u] = sg.w[k];
					d[u] = SSSPN_ROW_LOAD(rows + (size_t)sg.v[k] * TPR + q);
				}
			}
#pragma unroll
			for (int u = 0; u < 4; u++) {
				best[0] = N::relax(d[u].x, w[u], best[0]);
				best[1] = N::relax(d[u].y, w[u], best[1]);
				best[2] = N::relax(d[u].z, w[u], best[2]);
				best[3] = N::relax(d[u].w, w[u], best[3]);
			}
		}
		__syncwarp();
	}
	if (STATS && lane == 0) {
		atomicAdd(&ctl->hot, (unsigned long long)n_hot);
		atomicAdd(&ctl->scanned, (unsigned long long)n_scan);
	}
	return touched;
}

// the thread groups' minima -> every thread of the warp (threads 0 .. TPR-1 are the ones that use them)
template <class N>
__device__ __forceinline__ void ssspn_fold(uint32_t (&best)[4]) {
#pragma unroll
	for (int off = N::TPR; off < 32; off <<= 1) {
#pragma unroll
		for (int k = 0; k < 4; k++) best[k] = N::vmin(best[k], __shfl_xor_sync(0xffffffffu, best[k], off));
	}
}

// Whole warp calls; threads 0 .. TPR-1 hold the candidates of their 16 bytes of row n.  The row takes the candidates
// that beat it and the lane's bound; marks n changed.
template <class N>
__device__ __forceinline__ void ssspn_commit(typename N::S *dist, uint32_t n, unsigned lane, const uint32_t (&best)[4],
                                             uint32_t *__restrict__ chg_next, SsspCtl *__restrict__ ctl) {
	bool ch = false, lost = false;
	if (lane < N::TPR) {
		uint4 *rowp = reinterpret_cast<uint4 *>(dist) + (size_t)n * N::TPR + lane;
		const uint4 cur = __ldcg(rowp);
		const uint4 bd = *(reinterpret_cast<const uint4 *>(ctl->bound) + lane);
		uint32_t c[4] = {cur.x, cur.y, cur.z, cur.w};
		const uint32_t b[4] = {bd.x, bd.y, bd.z, bd.w};
#pragma unroll
		for (int k = 0; k < 4; k++) {
			const uint32_t mk = N::ltmask(best[k], N::vmin(c[k], b[k])); // all-ones per lane that is taken
			ch |= mk != 0;
			lost |= (best[k] & mk & N::LOSTW) != 0;
			c[k] = (best[k] & mk) | (c[k] & ~mk);
		}
		if (ch) __stcg(rowp, make_uint4(c[0], c[1], c[2], c[3]));
		if (lost) ctl->overflow = 1u;
	}
	if (__any_sync(0xffffffffu, ch) && lane == 0) atomicOr(&chg_next[n >> 5], 1u << (n & 31));
}

// one warp per vertex with at most warp_max in-edges
template <class N, bool STATS>
__global__ void __launch_bounds__(256) ssspn_relax_kernel(const uint32_t *__restrict__ ioff,
                                                          const uint32_t *__restrict__ isrc,
                                                          const typename N::S *__restrict__ iw, typename N::S *dist,
                                                          const uint32_t *__restrict__ chg_prev,
                                                          uint32_t *__restrict__ chg_next, SsspCtl *__restrict__ ctl,
                                                          const uint32_t *__restrict__ mark, uint32_t V,
                                                          uint32_t warp_max) {
	__shared__ StageN stage[8];
	if (ctl->done) return;
	const uint32_t n = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const unsigned lane = threadIdx.x & 31;
	if (n >= V) return;
	if (ctl->use_mark && !((mark[n >> 5] >> (n & 31)) & 1u)) return; // sparse round: no in-neighbour changed
	const uint32_t lo = ioff[n], hi = ioff[n + 1];
	if (lo == hi || hi - lo > warp_max) return;
	uint32_t best[4] = {N::INFW, N::INFW, N::INFW, N::INFW};
	// the vertex's own row is only read when some in-neighbour changed
	if (!ssspn_scan<N, STATS>(isrc, iw, dist, chg_prev, lo, hi, 32, lane, stage[threadIdx.x >> 5], best, ctl)) return;
	ssspn_fold<N>(best);
	ssspn_commit<N>(dist, n, lane, best, chg_next, ctl);
}

// 8 warps over the in-edges [lo, hi); on return warp 0 holds the block's minima (threads 0 .. TPR-1), and the return
// value (block-uniform) says whether any source row was read
template <class N, bool STATS>
__device__ __forceinline__ bool ssspn_block_scan(const uint32_t *__restrict__ isrc, const typename N::S *__restrict__ iw,
                                                 const typename N::S *dist, const uint32_t *__restrict__ chg_prev,
                                                 uint32_t lo, uint32_t hi, uint32_t (&best)[4], SsspCtl *ctl) {
	__shared__ StageN stage[8];
	__shared__ uint32_t part[8][N::TPR * 4];
	__shared__ int s_touched;
	const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	if (threadIdx.x == 0) s_touched = 0;
	__syncthreads();
	const bool touched = ssspn_scan<N, STATS>(isrc, iw, dist, chg_prev, lo + warp * 32, hi, 8 * 32, lane, stage[warp],
	                                          best, ctl);
	ssspn_fold<N>(best);
	if (lane < N::TPR) {
#pragma unroll
		for (int k = 0; k < 4; k++) part[warp][lane * 4 + k] = best[k];
	}
	if (touched && lane == 0) s_touched = 1;
	__syncthreads();
	if (!s_touched) return false;
	if (warp == 0 && lane < N::TPR) {
#pragma unroll
		for (int k = 0; k < 4; k++) {
			uint32_t x = part[0][lane * 4 + k];
#pragma unroll
			for (int w = 1; w < 8; w++) x = N::vmin(x, part[w][lane * 4 + k]);
			best[k] = x;
		}
	}
	return true;
}

// one block per vertex of the `mid` class list
template <class N, bool STATS>
__global__ void __launch_bounds__(256) ssspn_relax_big_kernel(const uint32_t *__restrict__ list,
                                                              const uint32_t *__restrict__ ioff,
                                                              const uint32_t *__restrict__ isrc,
                                                              const typename N::S *__restrict__ iw,
                                                              typename N::S *dist,
                                                              const uint32_t *__restrict__ chg_prev,
                                                              uint32_t *__restrict__ chg_next,
                                                              SsspCtl *__restrict__ ctl,
                                                              const uint32_t *__restrict__ mark) {
	if (ctl->done) return;
	const uint32_t n = list[blockIdx.x];
	if (ctl->use_mark && !((mark[n >> 5] >> (n & 31)) & 1u)) return;
	uint32_t best[4] = {N::INFW, N::INFW, N::INFW, N::INFW};
	if (!ssspn_block_scan<N, STATS>(isrc, iw, dist, chg_prev, ioff[n], ioff[n + 1], best, ctl)) return;
	if (threadIdx.x >= 32) return;
	ssspn_commit<N>(dist, n, threadIdx.x, best, chg_next, ctl);
}

// one block per chunk of a hub vertex's in-list; minima meet in acc[hub][64] (uint32 per lane at either width)
template <class N, bool STATS>
__global__ void __launch_bounds__(256) ssspn_hub_kernel(const HubChunk *__restrict__ chunks,
                                                        const uint32_t *__restrict__ isrc,
                                                        const typename N::S *__restrict__ iw,
                                                        const typename N::S *dist,
                                                        const uint32_t *__restrict__ chg_prev, uint32_t *acc,
                                                        uint32_t *acc_touched, SsspCtl *__restrict__ ctl) {
	if (ctl->done) return;
	const HubChunk hc = chunks[blockIdx.x];
	uint32_t best[4] = {N::INFW, N::INFW, N::INFW, N::INFW};
	if (!ssspn_block_scan<N, STATS>(isrc, iw, dist, chg_prev, hc.begin, hc.end, best, ctl)) return;
	if (threadIdx.x >= N::TPR) return;
	uint32_t *row = acc + (size_t)hc.hub * SSSP_LANES + threadIdx.x * (4 * N::LPW);
#pragma unroll
	for (int k = 0; k < 4; k++) {
#pragma unroll
		for (int i = 0; i < N::LPW; i++) {
			const uint32_t x = N::lane_of(best[k], i);
			if (x < N::INF) atomicMin(row + N::LPW * k + i, x);
		}
	}
	if (threadIdx.x == 0) acc_touched[hc.hub] = 1u;
}

template <class N>
__global__ void __launch_bounds__(256) ssspn_acc_fill_kernel(uint32_t *__restrict__ acc, uint32_t n) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) acc[i] = N::INF;
}

// one warp per hub vertex
template <class N>
__global__ void __launch_bounds__(256) ssspn_hub_commit_kernel(const uint32_t *__restrict__ hub, uint32_t nhub,
                                                               uint32_t *acc, uint32_t *acc_touched,
                                                               typename N::S *dist, uint32_t *__restrict__ chg_next,
                                                               SsspCtl *__restrict__ ctl) {
	if (ctl->done) return;
	const uint32_t h = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const unsigned lane = threadIdx.x & 31;
	if (h >= nhub || !acc_touched[h]) return; // warp-uniform
	uint32_t best[4] = {N::INFW, N::INFW, N::INFW, N::INFW};
	if (lane < N::TPR) {
		uint32_t *row = acc + (size_t)h * SSSP_LANES + lane * (4 * N::LPW);
#pragma unroll
		for (int k = 0; k < 4; k++) {
			uint32_t lanes[2] = {0, 0};
#pragma unroll
			for (int i = 0; i < N::LPW; i++) {
				lanes[i] = row[N::LPW * k + i];
				row[N::LPW * k + i] = N::INF;
			}
			best[k] = N::pack(lanes);
		}
	}
	__syncwarp();
	if (lane == 0) acc_touched[h] = 0u;
	ssspn_commit<N>(dist, hub[h], lane, best, chg_next, ctl);
}

} // namespace pgq
