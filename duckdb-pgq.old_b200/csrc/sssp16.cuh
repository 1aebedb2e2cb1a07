// cheapest_path_length on 16-bit distance rows (included by sssp.cu only).
//
// For non-negative int64 weights the relaxation rounds run on rows of 64 x uint16 (one 128-byte line per vertex):
//   finite distances are 0 .. 0x3FFF, "not reached" is 0x8000, weights are clamped to 0x4000 (uint16 per edge).
//   A sum of a finite distance and a weight is then at most 0x7FFF, below "not reached"; "not reached" plus a weight
//   is 0x8000 .. 0xC000: nothing wraps, and unsigned minima keep every real candidate below every unreached one.
//   A sum in 0x4000 .. 0x7FFF is a real distance that does not fit (or went through a clamped weight): if such a
//   value is ever accepted into a row -- it beats the row AND the lane's bound, the destination's current distance --
//   the batch is flagged and redone on 8-byte rows.  Distances are exact integers either way: same answers.
// Eight threads own a row (16 bytes = 8 lanes each), so a warp relaxes four source rows per load instruction, and the
// relaxation itself is one VIADDMNMX.U16x2 per two lanes: min(best, dist + w) (__viaddmin_u16x2, native on sm_100a).
// Hot in-edges of a 32-edge step are compacted through a per-warp shared-memory stage so that each 8-thread group
// picks its own source row.
#pragma once

namespace pgq {

static constexpr uint32_t N16_INF = 0x8000u;
static constexpr uint32_t N16_INF2 = 0x80008000u;
static constexpr uint32_t N16_LOST2 = 0x40004000u; // bit 14 of either half: a distance that does not fit
static constexpr long long N16_WCLAMP = 0x4000;

// Source rows may come from this SM's L1: a row that another SM updated during this round can then be read stale,
// which is harmless -- whoever updates a row also sets its bit in the next round's changed-bitmap, so the fresh value
// is read next round (L1 is invalidated at every kernel boundary), and the round that changes nothing still reads only
// final values.  High out-degree sources are read from many in-lists per round, which is what L1 catches.
// Measured (gpurun check 17c, C4 graph, 1024 pairs): 86.6 ms through L1 against 84.6 ms from L2 only -- L2 only it is.
#ifdef PGQ_SSSP_ROWS_L1
#define SSSP16_ROW_LOAD(p) __ldca(p)
#else
#define SSSP16_ROW_LOAD(p) __ldcg(p)
#endif

struct Stage16 {
	uint32_t v[32];
	uint32_t w2[32];
};

// int64 weights (known non-negative) clamped for the 16-bit rows
__global__ void __launch_bounds__(256) narrow_w_kernel(const long long *__restrict__ in, uint32_t n,
                                                       uint16_t *__restrict__ out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) out[i] = in[i] > N16_WCLAMP ? (uint16_t)N16_WCLAMP : (uint16_t)in[i];
}

__global__ void __launch_bounds__(256) sssp16_fill_kernel(uint4 *__restrict__ rows, size_t n16) {
	const uint4 inf = make_uint4(N16_INF2, N16_INF2, N16_INF2, N16_INF2);
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x)
		rows[i] = inf;
}

__global__ void sssp16_seed_kernel(const uint32_t *__restrict__ lane_src, uint16_t *__restrict__ dist,
                                   uint32_t *__restrict__ chg, SsspCtl *__restrict__ ctl, uint32_t mark_max) {
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
	reinterpret_cast<uint16_t *>(ctl->bound)[l] = (uint16_t)N16_INF;
	const uint32_t s = lane_src[l];
	if (s == 0xFFFFFFFFu) return;
	dist[(size_t)s * SSSP_LANES + l] = 0;
	atomicOr(&chg[s >> 5], 1u << (s & 31));
}

__global__ void __launch_bounds__(1024) sssp16_round_end_kernel(const uint32_t *__restrict__ lane_src,
                                                               const uint32_t *__restrict__ lane_dst,
                                                               const uint16_t *__restrict__ dist,
                                                               const uint32_t *__restrict__ chg_next, uint32_t nwords,
                                                               SsspCtl *__restrict__ ctl, int prune, uint32_t mark_max) {
	if (ctl->done) return;
	const uint32_t any = sssp_count_changed(chg_next, nwords, ctl);
	const int l = threadIdx.x;
	if (l < SSSP_LANES && prune && lane_src[l] != 0xFFFFFFFFu)
		reinterpret_cast<uint16_t *>(ctl->bound)[l] = dist[(size_t)lane_dst[l] * SSSP_LANES + l];
	if (l == 0) sssp_close_round(ctl, any, true, mark_max); // an overflowed batch is redone on 8-byte rows anyway
}

__global__ void sssp16_result_kernel(const uint32_t *__restrict__ lane_src, const uint32_t *__restrict__ lane_dst,
                                     const uint16_t *__restrict__ dist, long long *__restrict__ out) {
	const int l = threadIdx.x;
	long long r = WInf<long long>::value();
	if (lane_src[l] != 0xFFFFFFFFu) {
		const uint32_t d = dist[(size_t)lane_dst[l] * SSSP_LANES + l];
		if (d < N16_INF) r = (long long)d;
	}
	out[l] = r;
}

// min over the in-edges base0, base0+step, ... (32 consecutive edges per visit) of dist[src] + w.  On return thread
// (q = lane & 7, g = lane >> 3) holds, in best[0..3], the minima of row lanes 8q .. 8q+7 over the source rows its
// group g took; sssp16_fold merges the four groups.
template <bool STATS>
__device__ __forceinline__ bool sssp16_scan(const uint32_t *__restrict__ isrc, const uint16_t *__restrict__ iw,
                                            const uint16_t *dist, const uint32_t *__restrict__ chg_prev,
                                            uint32_t base0, uint32_t hi, uint32_t step, unsigned lane, Stage16 &sg,
                                            uint32_t (&best)[4], SsspCtl *ctl) {
	const unsigned q = lane & 7, g = lane >> 3;
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
			sg.w2[r] = (uint32_t)iw[j] * 0x10001u;
		}
		__syncwarp();
		const int cnt = __popc(m);
		if (STATS) n_hot += cnt;
		for (int i = (int)g; i < cnt; i += 16) { // up to four source rows per thread (16 per warp) in flight
			uint4 d[4];
			uint32_t w[4];
#pragma unroll
			for (int u = 0; u < 4; u++) {
				const int k = i + 4 * u;
				d[u] = make_uint4(N16_INF2, N16_INF2, N16_INF2, N16_INF2);
				w[u] = 0;
				if (k < cnt) {
					w[u] = sg.w2[k];
					d[u] = SSSP16_ROW_LOAD(rows + (size_t)sg.v[k] * 8 + q);
				}
			}
#pragma unroll
			for (int u = 0; u < 4; u++) {
				best[0] = __viaddmin_u16x2(d[u].x, w[u], best[0]);
				best[1] = __viaddmin_u16x2(d[u].y, w[u], best[1]);
				best[2] = __viaddmin_u16x2(d[u].z, w[u], best[2]);
				best[3] = __viaddmin_u16x2(d[u].w, w[u], best[3]);
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

// the four 8-thread groups' minima -> every thread of the warp (threads 0..7 are the ones that use them)
__device__ __forceinline__ void sssp16_fold(uint32_t (&best)[4]) {
#pragma unroll
	for (int k = 0; k < 4; k++) {
		best[k] = __vminu2(best[k], __shfl_xor_sync(0xffffffffu, best[k], 8));
		best[k] = __vminu2(best[k], __shfl_xor_sync(0xffffffffu, best[k], 16));
	}
}

// Whole warp calls; threads 0..7 hold the candidates of row lanes 8*lane .. 8*lane+7.  The row takes the candidates
// that beat it and the lane's bound; marks n changed.
__device__ __forceinline__ void sssp16_commit(uint16_t *dist, uint32_t n, unsigned lane, const uint32_t (&best)[4],
                                              uint32_t *__restrict__ chg_next, SsspCtl *__restrict__ ctl) {
	bool ch = false, lost = false;
	if (lane < 8) {
		uint4 *rowp = reinterpret_cast<uint4 *>(dist) + (size_t)n * 8 + lane;
		const uint4 cur = __ldcg(rowp);
		const uint4 bd = *(reinterpret_cast<const uint4 *>(ctl->bound) + lane);
		uint32_t c[4] = {cur.x, cur.y, cur.z, cur.w};
		const uint32_t b[4] = {bd.x, bd.y, bd.z, bd.w};
#pragma unroll
		for (int k = 0; k < 4; k++) {
			const uint32_t mk = __vcmpltu2(best[k], __vminu2(c[k], b[k])); // 0xFFFF per half that is taken
			ch |= mk != 0;
			lost |= (best[k] & mk & N16_LOST2) != 0;
			c[k] = (best[k] & mk) | (c[k] & ~mk);
		}
		if (ch) __stcg(rowp, make_uint4(c[0], c[1], c[2], c[3]));
		if (lost) ctl->overflow = 1u;
	}
	if (__any_sync(0xffffffffu, ch) && lane == 0) atomicOr(&chg_next[n >> 5], 1u << (n & 31));
}

// one warp per vertex with at most warp_max in-edges
template <bool STATS>
__global__ void __launch_bounds__(256) sssp16_relax_kernel(const uint32_t *__restrict__ ioff,
                                                           const uint32_t *__restrict__ isrc,
                                                           const uint16_t *__restrict__ iw, uint16_t *dist,
                                                           const uint32_t *__restrict__ chg_prev,
                                                           uint32_t *__restrict__ chg_next, SsspCtl *__restrict__ ctl,
                                                           const uint32_t *__restrict__ mark, uint32_t V,
                                                           uint32_t warp_max) {
	__shared__ Stage16 stage[8];
	if (ctl->done) return;
	const uint32_t n = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const unsigned lane = threadIdx.x & 31;
	if (n >= V) return;
	if (ctl->use_mark && !((mark[n >> 5] >> (n & 31)) & 1u)) return; // sparse round: no in-neighbour changed
	const uint32_t lo = ioff[n], hi = ioff[n + 1];
	if (lo == hi || hi - lo > warp_max) return;
	uint32_t best[4] = {N16_INF2, N16_INF2, N16_INF2, N16_INF2};
	// the vertex's own row is only read when some in-neighbour changed
	if (!sssp16_scan<STATS>(isrc, iw, dist, chg_prev, lo, hi, 32, lane, stage[threadIdx.x >> 5], best, ctl)) return;
	sssp16_fold(best);
	sssp16_commit(dist, n, lane, best, chg_next, ctl);
}

// 8 warps over the in-edges [lo, hi); on return warp 0 holds the block's minima (threads 0..7), and the return value
// (block-uniform) says whether any source row was read
template <bool STATS>
__device__ __forceinline__ bool sssp16_block_scan(const uint32_t *__restrict__ isrc, const uint16_t *__restrict__ iw,
                                                  const uint16_t *dist, const uint32_t *__restrict__ chg_prev,
                                                  uint32_t lo, uint32_t hi, uint32_t (&best)[4], SsspCtl *ctl) {
	__shared__ Stage16 stage[8];
	__shared__ uint32_t part[8][32];
	__shared__ int s_touched;
	const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	if (threadIdx.x == 0) s_touched = 0;
	__syncthreads();
	const bool touched = sssp16_scan<STATS>(isrc, iw, dist, chg_prev, lo + warp * 32, hi, 8 * 32, lane, stage[warp],
	                                        best, ctl);
	sssp16_fold(best);
	if (lane < 8) {
#pragma unroll
		for (int k = 0; k < 4; k++) part[warp][lane * 4 + k] = best[k];
	}
	if (touched && lane == 0) s_touched = 1;
	__syncthreads();
	if (!s_touched) return false;
	if (warp == 0 && lane < 8) {
#pragma unroll
		for (int k = 0; k < 4; k++) {
			uint32_t x = part[0][lane * 4 + k];
#pragma unroll
			for (int w = 1; w < 8; w++) x = __vminu2(x, part[w][lane * 4 + k]);
			best[k] = x;
		}
	}
	return true;
}

// one block per vertex of the `mid` class list
template <bool STATS>
__global__ void __launch_bounds__(256) sssp16_relax_big_kernel(const uint32_t *__restrict__ list,
                                                               const uint32_t *__restrict__ ioff,
                                                               const uint32_t *__restrict__ isrc,
                                                               const uint16_t *__restrict__ iw, uint16_t *dist,
                                                               const uint32_t *__restrict__ chg_prev,
                                                               uint32_t *__restrict__ chg_next,
                                                               SsspCtl *__restrict__ ctl,
                                                               const uint32_t *__restrict__ mark) {
	if (ctl->done) return;
	const uint32_t n = list[blockIdx.x];
	if (ctl->use_mark && !((mark[n >> 5] >> (n & 31)) & 1u)) return;
	uint32_t best[4] = {N16_INF2, N16_INF2, N16_INF2, N16_INF2};
	if (!sssp16_block_scan<STATS>(isrc, iw, dist, chg_prev, ioff[n], ioff[n + 1], best, ctl)) return;
	if (threadIdx.x >= 32) return;
	sssp16_commit(dist, n, threadIdx.x, best, chg_next, ctl);
}

// one block per chunk of a hub vertex's in-list; minima meet in acc[hub][64] (uint32 per lane)
template <bool STATS>
__global__ void __launch_bounds__(256) sssp16_hub_kernel(const HubChunk *__restrict__ chunks,
                                                         const uint32_t *__restrict__ isrc,
                                                         const uint16_t *__restrict__ iw, const uint16_t *dist,
                                                         const uint32_t *__restrict__ chg_prev, uint32_t *acc,
                                                         uint32_t *acc_touched, SsspCtl *__restrict__ ctl) {
	if (ctl->done) return;
	const HubChunk hc = chunks[blockIdx.x];
	uint32_t best[4] = {N16_INF2, N16_INF2, N16_INF2, N16_INF2};
	if (!sssp16_block_scan<STATS>(isrc, iw, dist, chg_prev, hc.begin, hc.end, best, ctl)) return;
	if (threadIdx.x >= 8) return;
	uint32_t *row = acc + (size_t)hc.hub * SSSP_LANES + threadIdx.x * 8;
#pragma unroll
	for (int k = 0; k < 4; k++) {
		const uint32_t a = best[k] & 0xFFFFu, b = best[k] >> 16;
		if (a < N16_INF) atomicMin(row + 2 * k, a);
		if (b < N16_INF) atomicMin(row + 2 * k + 1, b);
	}
	if (threadIdx.x == 0) acc_touched[hc.hub] = 1u;
}

__global__ void __launch_bounds__(256) sssp16_acc_fill_kernel(uint32_t *__restrict__ acc, uint32_t n) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) acc[i] = N16_INF;
}

// one warp per hub vertex
__global__ void __launch_bounds__(256) sssp16_hub_commit_kernel(const uint32_t *__restrict__ hub, uint32_t nhub,
                                                                uint32_t *acc, uint32_t *acc_touched, uint16_t *dist,
                                                                uint32_t *__restrict__ chg_next,
                                                                SsspCtl *__restrict__ ctl) {
	if (ctl->done) return;
	const uint32_t h = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const unsigned lane = threadIdx.x & 31;
	if (h >= nhub || !acc_touched[h]) return; // warp-uniform
	uint32_t best[4] = {N16_INF2, N16_INF2, N16_INF2, N16_INF2};
	if (lane < 8) {
		uint32_t *row = acc + (size_t)h * SSSP_LANES + lane * 8;
#pragma unroll
		for (int k = 0; k < 4; k++) {
			best[k] = row[2 * k] | (row[2 * k + 1] << 16);
			row[2 * k] = N16_INF;
			row[2 * k + 1] = N16_INF;
		}
	}
	__syncwarp();
	if (lane == 0) acc_touched[h] = 0u;
	sssp16_commit(dist, hub[h], lane, best, chg_next, ctl);
}

} // namespace pgq
