// cheapest_path_length (rule R10 in DESIGN.md section 5b): lane-batched
// Bellman-Ford over a CSR that carries one 8-byte weight (int64 or double) per edge.
//
// Layout: one distance row per vertex, 64 lanes (64 source rows of the query) wide.  Each round is a bottom-up
// relaxation over the reverse CSR: a warp per vertex n (a block per vertex / per 4096-edge chunk for the long
// in-lists), 32 in-edges per step -- every thread loads one source id and tests the source's bit in the "changed
// last round" bitmap; for the sources that pass, the warp reads their rows and keeps the minimum of dist[v] + w.
// Rows are updated in place (distances only decrease, so a relaxation that sees a newer value only converges
// sooner); the fixed point min_v(dist[v] + w) does not depend on the order, which is what makes the result
// bit-identical to the oracle's push formulation for int64 and for double weights alike.  A round that changes
// nothing ends the batch; more than V rounds means a negative cycle.
//
// Round 2 (all measured on the C4 graph, 1024 pairs, int64 weights 1..100: 559 ms -> 80 ms):
//  - the control block no longer has the "changed" flag next to `done` (every warp's first load queued behind
//    450 k stores to that line: 3.1 ms -> 0.55 ms per dense round); the round-end kernel counts the changed-bitmap;
//  - non-negative int64 weights run on 16-bit rows (sssp_narrow.cuh: one 128-byte line per vertex, VIADDMNMX.U16x2,
//    exact; batches whose distances do not fit are redone on 32-bit rows (VIADDMNMX.U32), then on 8-byte rows);
//  - with non-negative weights the destination's current distance bounds what a row accepts;
//  - hub in-lists are relaxed in 4096-edge chunks (one 59 k-edge in-list used to set the length of every round);
//  - sparse rounds relax only the out-neighbours of the changed vertices (sssp_mark_kernel);
//  - up to four lane batches in flight on their own streams, and the host looks at a batch every 4 rounds instead
//    of waiting after each one.
#include "common.cuh"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <type_traits>
#include <vector>

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

// Per-batch control block, device side; the host reads a copy of its first 32 bytes every few rounds instead of
// waiting once per round.  Three groups of 128-byte lines, so that what every warp of a relax kernel READS (done,
// the bounds) never shares a line with what the relax kernels WRITE: with the "changed" flag next to `done`, the
// first load of every warp queued behind 450 k stores to the same line (23 % of all stall samples, ncu, check 17).
// Whether a round changed anything is read off the changed-bitmap by the round-end kernel instead.
struct SsspCtl {
	// line 0: written by the seed / round-end kernels only
	uint32_t done;     // a round changed nothing (or 16-bit rows overflowed): later launches of this batch return at once
	uint32_t rounds;   // rounds run
	uint32_t overflow_seen;
	uint32_t use_mark; // the coming round is sparse: only vertices marked by sssp_mark_kernel are relaxed
	unsigned long long hot_seen, scanned_seen;
	uint32_t n_changed; // vertices the last round changed
	unsigned char pad1[92];
	// lines 1-4: raw value of dist[dst_l][l] at the start of the round (8-byte rows: R per lane; 16-bit rows: the
	// first 128 bytes, uint16 per lane); written by the round-end kernel only
	unsigned long long bound[SSSP_LANES];
	// line 5: written by the relax kernels
	uint32_t overflow; // 16-bit rows only: a distance that may still matter did not fit
	uint32_t pad2;
	unsigned long long hot;     // stats mode: source rows read
	unsigned long long scanned; // stats mode: in-edges scanned
	unsigned char pad3[104];
};
static_assert(sizeof(SsspCtl) == 128 * 6, "SsspCtl: three groups of whole lines");

// Whole block (the round-end kernel): how many vertices did the round mark in the changed-bitmap?
__device__ __forceinline__ uint32_t sssp_count_changed(const uint32_t *__restrict__ chg, uint32_t nwords,
                                                       SsspCtl *ctl) {
	__shared__ uint32_t s_total;
	if (threadIdx.x == 0) s_total = 0;
	__syncthreads();
	uint32_t acc = 0;
	const uint4 *c4 = reinterpret_cast<const uint4 *>(chg); // the bitmap is 16-byte aligned, nwords a multiple of 4
	for (uint32_t i = threadIdx.x; i < nwords / 4; i += blockDim.x) {
		const uint4 x = c4[i];
		acc += __popc(x.x) + __popc(x.y) + __popc(x.z) + __popc(x.w);
	}
	acc = __reduce_add_sync(0xffffffffu, acc);
	if ((threadIdx.x & 31) == 0 && acc) atomicAdd(&s_total, acc);
	__syncthreads();
	return s_total;
}

// thread 0 of the round-end kernel.  A round after which few vertices changed is followed by a sparse round: the
// changed vertices' out-neighbours are marked (sssp_mark_kernel) and only those are relaxed, so the tail rounds stop
// paying for a scan of every in-list (130 us per round on the C4 graph, whatever the number of changes).
__device__ __forceinline__ void sssp_close_round(SsspCtl *ctl, uint32_t any, bool stop_on_overflow,
                                                 uint32_t mark_max) {
	ctl->n_changed = any;
	ctl->use_mark = any <= mark_max ? 1u : 0u;
	ctl->rounds++;
	ctl->overflow_seen = ctl->overflow;
	ctl->hot_seen = ctl->hot;
	ctl->scanned_seen = ctl->scanned;
	if (!any || (stop_on_overflow && ctl->overflow)) ctl->done = 1;
}

// Sparse rounds: one warp per word of the changed-bitmap; the out-neighbours of every changed vertex get their bit
// in `mark` (forward CSR: off is the reference's int64 v array, tgt the narrowed e array).
__global__ void __launch_bounds__(256) sssp_mark_kernel(const uint32_t *__restrict__ chg_prev, uint32_t nwords,
                                                        const int64_t *__restrict__ off,
                                                        const uint32_t *__restrict__ tgt, uint32_t *__restrict__ mark,
                                                        const SsspCtl *__restrict__ ctl) {
	if (ctl->done || !ctl->use_mark) return;
	const uint32_t wi = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const unsigned lane = threadIdx.x & 31;
	if (wi >= nwords) return;
	uint32_t bits = chg_prev[wi];
	while (bits) {
		const uint32_t v = wi * 32 + (__ffs(bits) - 1);
		bits &= bits - 1;
		const int64_t lo = off[v], hi = off[v + 1];
		for (int64_t j = lo + lane; j < hi; j += 32) {
			const uint32_t t = tgt[j];
			atomicOr(&mark[t >> 5], 1u << (t & 31));
		}
	}
}

// Row types.  R = a lane's value in registers, S = storage element, W = weight in registers.  A warp owns a row;
// thread t of the warp owns two lanes of it, lane_of(t, 0) and lane_of(t, 1).
template <typename T>
struct Row64 { // 64 lanes x 8 bytes, T = long long | double; thread t owns lanes t and t + 32
	using R = T;
	using S = T;
	using W = T;
	static constexpr bool narrow = false;
	__device__ static R inf() { return WInf<T>::value(); }
	__device__ static int lane_of(unsigned t, int k) { return (int)t + 32 * k; }
	__device__ static void load2(const S *dist, uint32_t v, unsigned t, R &a, R &b) {
		// volatile: rows are updated in place by other warps; never cache them in registers
		const S *row = dist + (size_t)v * SSSP_LANES;
		a = *reinterpret_cast<const volatile S *>(row + t);
		b = *reinterpret_cast<const volatile S *>(row + t + 32);
	}
	__device__ static R add(R d, W w) { return WInf<T>::is_inf(d) ? d : d + w; }
	__device__ static void store2(S *dist, uint32_t v, unsigned t, R a, R b, bool sa, bool sb) {
		S *row = dist + (size_t)v * SSSP_LANES;
		if (sa) row[t] = a;
		if (sb) row[t + 32] = b;
	}
	__device__ static R load1(const S *dist, uint32_t v, int l) { return dist[(size_t)v * SSSP_LANES + l]; }
	__device__ static void store1(S *dist, uint32_t v, int l, R x) { dist[(size_t)v * SSSP_LANES + l] = x; }
	__device__ static unsigned long long raw(R x) {
		unsigned long long u;
		memcpy(&u, &x, 8);
		return u;
	}
	__device__ static R unraw(unsigned long long u) {
		R x;
		memcpy(&x, &u, 8);
		return x;
	}
	__device__ static T result(R x) { return x; }
	__device__ static void atomic_min(R *addr, R x) {
		if constexpr (std::is_same<T, long long>::value) atomicMin(addr, x);
		else { // double: compare-and-swap until x is no longer smaller (NaN never is)
			unsigned long long *a = reinterpret_cast<unsigned long long *>(addr);
			unsigned long long old = *a;
			while (x < __longlong_as_double((long long)old)) {
				const unsigned long long seen = atomicCAS(a, old, (unsigned long long)__double_as_longlong(x));
				if (seen == old) break;
				old = seen;
			}
		}
	}
};

// Narrow rows for non-negative int64 weights (sssp_narrow.cuh has their kernels): 64 lanes x 2 bytes (one 128-byte
// line per vertex), and 64 lanes x 4 bytes for the batches whose distances do not fit 14 bits
struct N16;
struct N32;
struct Row16 {
	using R = uint32_t; // hub accumulator element
	using S = uint16_t;
	using W = uint16_t;
	using N = N16;
	static constexpr bool narrow = true;
};
struct Row32 {
	using R = uint32_t;
	using S = uint32_t;
	using W = uint32_t;
	using N = N32;
	static constexpr bool narrow = true;
};

// w_csr[p] = w_rows[row[p]]  (raw 8-byte values)
__global__ void __launch_bounds__(256) gather_w_kernel(const unsigned long long *__restrict__ in,
                                                       const uint32_t *__restrict__ idx, uint32_t n,
                                                       unsigned long long *__restrict__ out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) out[i] = in[idx[i]];
}

// Sums of whole-numbered doubles are exact while they stay below 2^53: with every weight a whole number of at most
// 2^20 and fewer than 2^32 edges on a path, double weights can take the integer rows and give the same bits.
static constexpr double SSSP_WHOLE_MAX = 1048576.0;

// stat[0] bit 0: some weight is negative (or NaN); bit 1 (double weights): some weight is not a whole number <= 2^20
__global__ void __launch_bounds__(256) wstat_kernel(const unsigned long long *__restrict__ w, uint32_t n, int f64,
                                                    uint32_t *__restrict__ stat) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	bool bad = false, frac = false;
	if (i < n) {
		if (f64) {
			const double x = __longlong_as_double((long long)w[i]);
			bad = !(x >= 0.0);
			frac = !(x == floor(x) && x <= SSSP_WHOLE_MAX);
		} else
			bad = (long long)w[i] < 0;
	}
	const uint32_t bits = (__any_sync(0xffffffffu, bad) ? 1u : 0u) | (__any_sync(0xffffffffu, frac) ? 2u : 0u);
	if (bits && (threadIdx.x & 31) == 0) atomicOr(stat, bits);
}

// whole-numbered double weights as int64, and the int64 answers back as doubles (+inf = not reached)
__global__ void __launch_bounds__(256) whole_w_kernel(const double *__restrict__ in, uint32_t n, long long *__restrict__ out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) out[i] = (long long)in[i];
}

template <class TR>
__global__ void __launch_bounds__(256) sssp_fill_kernel(typename TR::S *__restrict__ dist, size_t n) {
	const typename TR::S inf = TR::inf();
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
		dist[i] = inf;
}

// one block of 64 threads: sources at distance 0, control block reset, bounds = not reached
template <class TR>
__global__ void sssp_seed_kernel(const uint32_t *__restrict__ lane_src, typename TR::S *__restrict__ dist,
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
	ctl->bound[l] = TR::raw(TR::inf());
	const uint32_t s = lane_src[l];
	if (s == 0xFFFFFFFFu) return;
	TR::store1(dist, s, l, (typename TR::R)0);
	atomicOr(&chg[s >> 5], 1u << (s & 31));
}

// after the relax kernels of a round (one block of 256): nothing changed -> done; else the lanes' bounds for the next
template <class TR>
__global__ void __launch_bounds__(1024) sssp_round_end_kernel(const uint32_t *__restrict__ lane_src,
                                                             const uint32_t *__restrict__ lane_dst,
                                                             const typename TR::S *__restrict__ dist,
                                                             const uint32_t *__restrict__ chg_next, uint32_t nwords,
                                                             SsspCtl *__restrict__ ctl, int prune, uint32_t mark_max) {
	if (ctl->done) return;
	const uint32_t any = sssp_count_changed(chg_next, nwords, ctl);
	const int l = threadIdx.x;
	if (l < SSSP_LANES && prune && lane_src[l] != 0xFFFFFFFFu)
		ctl->bound[l] = TR::raw(TR::load1(dist, lane_dst[l], l));
	if (l == 0) sssp_close_round(ctl, any, false, mark_max);
}

// min over the in-edges base0, base0+step, ... (32 consecutive edges per visit) of dist[src] + w, for the
// thread's two lanes; only sources whose bit is set in chg_prev are read.
template <class TR, bool STATS>
__device__ __forceinline__ bool sssp_scan(const uint32_t *__restrict__ isrc, const typename TR::W *__restrict__ iw,
                                          const typename TR::S *dist, const uint32_t *__restrict__ chg_prev,
                                          uint32_t base0, uint32_t hi, uint32_t step, unsigned lane,
                                          typename TR::R &best0, typename TR::R &best1, SsspCtl *ctl) {
	using R = typename TR::R;
	using W = typename TR::W;
	bool touched = false;
	uint32_t n_hot = 0, n_scan = 0;
	for (uint32_t base = base0; base < hi; base += step) {
		const uint32_t j = base + lane;
		uint32_t v = 0;
		W w = (W)0;
		bool hot = false;
		if (j < hi) {
			v = isrc[j];
			hot = (__ldg(chg_prev + (v >> 5)) >> (v & 31)) & 1u;
			if (hot) w = iw[j];
		}
		unsigned m = __ballot_sync(0xffffffffu, hot);
		touched |= m != 0;
		if (STATS) {
			n_hot += __popc(m);
			n_scan += min(32u, hi - base);
		}
		while (m) { // four source rows in flight per step
			uint32_t sv[4];
			W sw[4];
			R d0[4], d1[4];
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
				d0[u] = d1[u] = TR::inf();
				if (on[u]) TR::load2(dist, sv[u], lane, d0[u], d1[u]);
			}
#pragma unroll
			for (int u = 0; u < 4; u++) {
				const R c0 = TR::add(d0[u], sw[u]), c1 = TR::add(d1[u], sw[u]);
				if (c0 < best0) best0 = c0;
				if (c1 < best1) best1 = c1;
			}
		}
	}
	if (STATS && lane == 0) {
		atomicAdd(&ctl->hot, (unsigned long long)n_hot);
		atomicAdd(&ctl->scanned, (unsigned long long)n_scan);
	}
	return touched;
}

// The row of vertex n takes the candidates that beat it and the lane's bound; marks n changed.
template <class TR>
__device__ __forceinline__ void sssp_commit(typename TR::S *dist, uint32_t n, unsigned lane, typename TR::R best0,
                                            typename TR::R best1, uint32_t *__restrict__ chg_next,
                                            SsspCtl *__restrict__ ctl) {
	using R = typename TR::R;
	R cur0, cur1;
	TR::load2(dist, n, lane, cur0, cur1);
	const R b0 = TR::unraw(ctl->bound[TR::lane_of(lane, 0)]), b1 = TR::unraw(ctl->bound[TR::lane_of(lane, 1)]);
	bool s0 = best0 < cur0 && best0 < b0, s1 = best1 < cur1 && best1 < b1;
	TR::store2(dist, n, lane, best0, best1, s0, s1);
	if (__any_sync(0xffffffffu, s0 || s1) && lane == 0) atomicOr(&chg_next[n >> 5], 1u << (n & 31));
}

// one warp per vertex with at most warp_max in-edges
template <class TR, bool STATS>
__global__ void __launch_bounds__(256) sssp_relax_kernel(const uint32_t *__restrict__ ioff,
                                                         const uint32_t *__restrict__ isrc,
                                                         const typename TR::W *__restrict__ iw, typename TR::S *dist,
                                                         const uint32_t *__restrict__ chg_prev,
                                                         uint32_t *__restrict__ chg_next, SsspCtl *__restrict__ ctl,
                                                         const uint32_t *__restrict__ mark, uint32_t V,
                                                         uint32_t warp_max) {
	if (ctl->done) return;
	const uint32_t n = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const unsigned lane = threadIdx.x & 31;
	if (n >= V) return;
	if (ctl->use_mark && !((mark[n >> 5] >> (n & 31)) & 1u)) return; // sparse round: no in-neighbour changed
	const uint32_t lo = ioff[n], hi = ioff[n + 1];
	if (lo == hi || hi - lo > warp_max) return;
	typename TR::R best0 = TR::inf(), best1 = TR::inf();
	// the vertex's own row is only read when some in-neighbour changed
	if (!sssp_scan<TR, STATS>(isrc, iw, dist, chg_prev, lo, hi, 32, lane, best0, best1, ctl)) return; // warp-uniform
	sssp_commit<TR>(dist, n, lane, best0, best1, chg_next, ctl);
}

// one block (8 warps) per vertex of a high in-degree class list; warps take alternate 32-edge groups
template <class TR, bool STATS>
__global__ void __launch_bounds__(256) sssp_relax_big_kernel(const uint32_t *__restrict__ list,
                                                             const uint32_t *__restrict__ ioff,
                                                             const uint32_t *__restrict__ isrc,
                                                             const typename TR::W *__restrict__ iw,
                                                             typename TR::S *dist,
                                                             const uint32_t *__restrict__ chg_prev,
                                                             uint32_t *__restrict__ chg_next,
                                                             SsspCtl *__restrict__ ctl,
                                                             const uint32_t *__restrict__ mark) {
	using R = typename TR::R;
	__shared__ R part[8][SSSP_LANES];
	__shared__ int s_touched;
	if (ctl->done) return;
	const uint32_t n = list[blockIdx.x];
	if (ctl->use_mark && !((mark[n >> 5] >> (n & 31)) & 1u)) return;
	const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint32_t lo = ioff[n], hi = ioff[n + 1];
	if (threadIdx.x == 0) s_touched = 0;
	__syncthreads();
	R best0 = TR::inf(), best1 = TR::inf();
	const bool touched =
	    sssp_scan<TR, STATS>(isrc, iw, dist, chg_prev, lo + warp * 32, hi, 8 * 32, lane, best0, best1, ctl);
	part[warp][TR::lane_of(lane, 0)] = best0;
	part[warp][TR::lane_of(lane, 1)] = best1;
	if (touched && lane == 0) s_touched = 1;
	__syncthreads();
	if (!s_touched || warp != 0) return;
	best0 = part[0][TR::lane_of(lane, 0)];
	best1 = part[0][TR::lane_of(lane, 1)];
#pragma unroll
	for (int k = 1; k < 8; k++) {
		const R x0 = part[k][TR::lane_of(lane, 0)], x1 = part[k][TR::lane_of(lane, 1)];
		if (x0 < best0) best0 = x0;
		if (x1 < best1) best1 = x1;
	}
	sssp_commit<TR>(dist, n, lane, best0, best1, chg_next, ctl);
}

// Hub vertices (in-degree above hub_min): one block per 4096-edge chunk of the in-list (csr_prepare's HubChunk list),
// so the longest in-list no longer sets the length of a round.  The chunks' minima meet in acc[hub][64] through
// atomic minima; sssp_hub_commit_kernel then treats the row like any other and re-arms acc.
template <class TR, bool STATS>
__global__ void __launch_bounds__(256) sssp_hub_kernel(const HubChunk *__restrict__ chunks,
                                                       const uint32_t *__restrict__ isrc,
                                                       const typename TR::W *__restrict__ iw,
                                                       const typename TR::S *dist,
                                                       const uint32_t *__restrict__ chg_prev, typename TR::R *acc,
                                                       uint32_t *acc_touched, SsspCtl *__restrict__ ctl) {
	using R = typename TR::R;
	__shared__ R part[8][SSSP_LANES];
	__shared__ int s_touched;
	if (ctl->done) return;
	const HubChunk hc = chunks[blockIdx.x];
	const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	if (threadIdx.x == 0) s_touched = 0;
	__syncthreads();
	R best0 = TR::inf(), best1 = TR::inf();
	const bool touched = sssp_scan<TR, STATS>(isrc, iw, dist, chg_prev, hc.begin + warp * 32, hc.end, 8 * 32, lane,
	                                          best0, best1, ctl);
	part[warp][TR::lane_of(lane, 0)] = best0;
	part[warp][TR::lane_of(lane, 1)] = best1;
	if (touched && lane == 0) s_touched = 1;
	__syncthreads();
	if (!s_touched || warp != 0) return;
	best0 = part[0][TR::lane_of(lane, 0)];
	best1 = part[0][TR::lane_of(lane, 1)];
#pragma unroll
	for (int k = 1; k < 8; k++) {
		const R x0 = part[k][TR::lane_of(lane, 0)], x1 = part[k][TR::lane_of(lane, 1)];
		if (x0 < best0) best0 = x0;
		if (x1 < best1) best1 = x1;
	}
	R *row = acc + (size_t)hc.hub * SSSP_LANES;
	if (best0 < TR::inf()) TR::atomic_min(row + TR::lane_of(lane, 0), best0);
	if (best1 < TR::inf()) TR::atomic_min(row + TR::lane_of(lane, 1), best1);
	if (lane == 0) acc_touched[hc.hub] = 1u;
}

template <class TR>
__global__ void __launch_bounds__(256) sssp_acc_fill_kernel(typename TR::R *__restrict__ acc, uint32_t n) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) acc[i] = TR::inf();
}

// one warp per hub vertex
template <class TR>
__global__ void __launch_bounds__(256) sssp_hub_commit_kernel(const uint32_t *__restrict__ hub, uint32_t nhub,
                                                              typename TR::R *acc, uint32_t *acc_touched,
                                                              typename TR::S *dist, uint32_t *__restrict__ chg_next,
                                                              SsspCtl *__restrict__ ctl) {
	using R = typename TR::R;
	if (ctl->done) return;
	const uint32_t h = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const unsigned lane = threadIdx.x & 31;
	if (h >= nhub || !acc_touched[h]) return; // warp-uniform
	R *row = acc + (size_t)h * SSSP_LANES;
	const R best0 = row[TR::lane_of(lane, 0)], best1 = row[TR::lane_of(lane, 1)];
	row[TR::lane_of(lane, 0)] = TR::inf();
	row[TR::lane_of(lane, 1)] = TR::inf();
	__syncwarp();
	if (lane == 0) acc_touched[h] = 0u;
	sssp_commit<TR>(dist, hub[h], lane, best0, best1, chg_next, ctl);
}

// out[l] = the lane's answer as T (int64 or double); a narrow batch that overflowed leaves its lanes to the redo
template <class TR, typename T>
__global__ void sssp_result_kernel(const uint32_t *__restrict__ lane_src, const uint32_t *__restrict__ lane_dst,
                                   const typename TR::S *__restrict__ dist, T *__restrict__ out) {
	const int l = threadIdx.x;
	out[l] = lane_src[l] == 0xFFFFFFFFu ? WInf<T>::value() : (T)TR::result(TR::load1(dist, lane_dst[l], l));
}

} // namespace pgq
#include "sssp_narrow.cuh"
namespace pgq {

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
		g->d_iw16.release();
		g->d_iw32n.release();
		g->d_iwi.release();
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
		// non-negative weights (no NaN) let the destination's distance bound the search; int64 ones also take 16-bit rows
		PGQ_TRY(g->d_wstat.reserve(16));
		PGQ_CUDA(cudaMemsetAsync(g->d_wstat.p, 0, 16, st));
		if (g->E > 0) {
			wstat_kernel<<<div_up(g->E, 256), 256, 0, st>>>(g->d_w.as<unsigned long long>(), (uint32_t)g->E,
			                                                wtype == PGQ_WEIGHT_F64 ? 1 : 0, g->d_wstat.as<uint32_t>());
			count_launch(c);
		}
		uint32_t bad = 1;
		PGQ_CUDA(cudaMemcpyAsync(&bad, g->d_wstat.p, 4, cudaMemcpyDeviceToHost, st));
		PGQ_CUDA(cudaStreamSynchronize(st));
		g->w_nonneg = (bad & 1u) == 0;
		g->w_whole = wtype == PGQ_WEIGHT_F64 && bad == 0; // non-negative whole numbers stored as DOUBLE
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

static constexpr int SSSP_CHUNK = 4; // rounds launched between two looks at the control block

// Runs the lane batches listed in `todo` with rows of type TR, up to opt.sssp_slots of them in flight, each on its
// own stream.  The host never waits per round: it launches SSSP_CHUNK rounds, copies the control block and moves on
// to the next slot; launches that follow a round which changed nothing return at once (ctl->done).  Batches whose
// 16-bit rows overflowed come back in `redo`.
template <class TR, typename T>
static int sssp_batches(Ctx *c, Csr *g, const LaneTable &lt, const std::vector<int64_t> &todo,
                        const typename TR::W *iw, bool prune, T *d_out, const uint32_t *d_src,
                        const uint32_t *d_dst, std::vector<int64_t> *redo) {
	using S = typename TR::S;
	using R = typename TR::R;
	const int L = SSSP_LANES;
	const uint32_t V = (uint32_t)g->V;
	const size_t cells = (size_t)V * L;
	const uint32_t cw = (uint32_t)((div_up((uint64_t)V, 32) + 4) & ~3ull); // words per bitmap, a multiple of 4
	// a round is sparse when the vertices changed before it have, at the average out-degree, fewer than V / 4
	// out-neighbours between them: marking those beats scanning every in-list (measured on the C4 graph: a round with
	// 5.7 k changed vertices of 450 k marks most of the graph and saves nothing; one with 1 k saves 2/3 of the scan)
	const uint32_t mark_max =
	    (uint32_t)std::min<uint64_t>((uint64_t)V * V / (4 * (uint64_t)std::max<int64_t>(g->E, 1)), 0x7FFFFFFFu);
	const size_t acc_cells = (size_t)g->n_hub * L; // R is 4 or 8 bytes: sized for 8
	if (!c->h_sssp) PGQ_CUDA(cudaHostAlloc(&c->h_sssp, 4 * sizeof(SsspCtl), cudaHostAllocDefault));
	SsspCtl *h_ctl = static_cast<SsspCtl *>(c->h_sssp);
	int nslot = (int)std::min<int64_t>(c->opt.sssp_slots, (int64_t)todo.size());
	if (nslot < 1) nslot = 1;
	int got = 0;
	for (; got < nslot; got++) {
		SsspSlot &sl = c->sssp_slot[got];
		if (!sl.st) PGQ_CUDA(cudaStreamCreateWithFlags(&sl.st, cudaStreamNonBlocking));
		if (!sl.ev) PGQ_CUDA(cudaEventCreateWithFlags(&sl.ev, cudaEventDisableTiming));
		int rc = sl.dist.reserve(cells * sizeof(S) + 16);
		if (rc == PGQ_OK) rc = sl.chg.reserve((size_t)3 * cw * sizeof(uint32_t) + 16);
		if (rc == PGQ_OK) rc = sl.ctl.reserve(sizeof(SsspCtl));
		if (rc == PGQ_OK) rc = sl.acc.reserve(acc_cells * 8 + (size_t)g->n_hub * 4 + 16);
		if (rc == PGQ_ERR_OOM && got > 0) { // fewer batches in flight
			(void)cudaGetLastError();
			break;
		}
		if (rc != PGQ_OK) return rc;
	}
	nslot = got;
	PGQ_CUDA(cudaEventRecord(c->ev_fork, c->stream)); // lane tables and weights are in place after this point
	struct Run {
		int64_t batch = -1;
		int64_t rounds = 0;
		int cur = 0;
	};
	Run run[4];
	size_t next = 0;
	int active = 0;
	auto launch_chunk = [&](int si) -> int {
		SsspSlot &sl = c->sssp_slot[si];
		Run &r = run[si];
		uint32_t *chg[2] = {sl.chg.as<uint32_t>(), sl.chg.as<uint32_t>() + cw};
		S *dist = sl.dist.as<S>();
		SsspCtl *ctl = sl.ctl.as<SsspCtl>();
		const uint32_t *ls = d_src + r.batch * L, *ld = d_dst + r.batch * L;
		const bool stats = c->opt.stats != 0;
		R *acc = sl.acc.as<R>();
		uint32_t *acc_touched = reinterpret_cast<uint32_t *>(sl.acc.as<char>() + acc_cells * 8);
		const uint32_t *ioff = g->d_ioff.as<uint32_t>(), *isrc = g->d_isrc.as<uint32_t>();
		const uint32_t *mid = g->d_mid.as<uint32_t>(), *hub = g->d_hub.as<uint32_t>();
		const HubChunk *chunks = g->d_hubchunks.as<HubChunk>();
		const unsigned g_warp = div_up((uint64_t)V * 32, 256), g_hubc = div_up((uint64_t)g->n_hub * 32, 256);
		for (int k = 0; k < SSSP_CHUNK; k++) {
			uint32_t *prev = chg[r.cur], *nxt = chg[r.cur ^ 1];
			uint32_t *mark = sl.chg.as<uint32_t>() + (size_t)2 * cw;
			PGQ_CUDA(cudaMemsetAsync(nxt, 0, (size_t)cw * 4, sl.st));
			PGQ_CUDA(cudaMemsetAsync(mark, 0, (size_t)cw * 4, sl.st));
			sssp_mark_kernel<<<div_up((uint64_t)cw * 32, 256), 256, 0, sl.st>>>(prev, cw, g->d_off.as<int64_t>(),
			                                                                  g->d_tgt.as<uint32_t>(), mark, ctl);
			// in-degree up to med_max: a warp per vertex; above: exactly the `mid` and `hub` class lists of csr_prepare
			if constexpr (TR::narrow) {
				using NW = typename TR::N;
				(stats ? ssspn_relax_kernel<NW, true> : ssspn_relax_kernel<NW, false>)<<<g_warp, 256, 0, sl.st>>>(
				    ioff, isrc, iw, dist, prev, nxt, ctl, mark, V, g->eff_med_max);
				if (g->n_mid)
					(stats ? ssspn_relax_big_kernel<NW, true> : ssspn_relax_big_kernel<NW, false>)<<<g->n_mid, 256, 0, sl.st>>>(
					    mid, ioff, isrc, iw, dist, prev, nxt, ctl, mark);
				if (g->n_hubchunks) {
					(stats ? ssspn_hub_kernel<NW, true> : ssspn_hub_kernel<NW, false>)<<<g->n_hubchunks, 256, 0, sl.st>>>(
					    chunks, isrc, iw, dist, prev, acc, acc_touched, ctl);
					ssspn_hub_commit_kernel<NW><<<g_hubc, 256, 0, sl.st>>>(hub, g->n_hub, acc, acc_touched, dist, nxt, ctl);
				}
				ssspn_round_end_kernel<NW><<<1, 1024, 0, sl.st>>>(ls, ld, dist, nxt, cw, ctl, prune ? 1 : 0, mark_max);
			} else {
				(stats ? sssp_relax_kernel<TR, true> : sssp_relax_kernel<TR, false>)<<<g_warp, 256, 0, sl.st>>>(
				    ioff, isrc, iw, dist, prev, nxt, ctl, mark, V, g->eff_med_max);
				if (g->n_mid)
					(stats ? sssp_relax_big_kernel<TR, true> : sssp_relax_big_kernel<TR, false>)<<<g->n_mid, 256, 0, sl.st>>>(
					    mid, ioff, isrc, iw, dist, prev, nxt, ctl, mark);
				if (g->n_hubchunks) {
					(stats ? sssp_hub_kernel<TR, true> : sssp_hub_kernel<TR, false>)<<<g->n_hubchunks, 256, 0, sl.st>>>(
					    chunks, isrc, iw, dist, prev, acc, acc_touched, ctl);
					sssp_hub_commit_kernel<TR><<<g_hubc, 256, 0, sl.st>>>(hub, g->n_hub, acc, acc_touched, dist, nxt, ctl);
				}
				sssp_round_end_kernel<TR><<<1, 1024, 0, sl.st>>>(ls, ld, dist, nxt, cw, ctl, prune ? 1 : 0, mark_max);
			}
			count_launch(c, 3 + (g->n_mid ? 1 : 0) + (g->n_hubchunks ? 2 : 0));
			r.cur ^= 1;
			r.rounds++;
		}
		PGQ_CUDA(cudaMemcpyAsync(h_ctl + si, ctl, 32, cudaMemcpyDeviceToHost, sl.st));
		PGQ_CUDA(cudaEventRecord(sl.ev, sl.st));
		return PGQ_OK;
	};
	auto start = [&](int si) -> int {
		SsspSlot &sl = c->sssp_slot[si];
		Run &r = run[si];
		r.batch = todo[next++];
		r.rounds = 0;
		r.cur = 0;
		active++;
		PGQ_CUDA(cudaStreamWaitEvent(sl.st, c->ev_fork, 0));
		PGQ_CUDA(cudaMemsetAsync(sl.chg.p, 0, (size_t)cw * 4, sl.st));
		if constexpr (TR::narrow) {
			ssspn_fill_kernel<typename TR::N><<<c->sm_count * 8, 256, 0, sl.st>>>(sl.dist.as<uint4>(),
			                                                                     cells * sizeof(S) / 16);
			ssspn_seed_kernel<typename TR::N><<<1, L, 0, sl.st>>>(d_src + r.batch * L, sl.dist.as<S>(), sl.chg.as<uint32_t>(),
			                                       sl.ctl.as<SsspCtl>(), mark_max);
		} else {
			sssp_fill_kernel<TR><<<c->sm_count * 8, 256, 0, sl.st>>>(sl.dist.as<S>(), cells);
			sssp_seed_kernel<TR><<<1, L, 0, sl.st>>>(d_src + r.batch * L, sl.dist.as<S>(), sl.chg.as<uint32_t>(),
			                                         sl.ctl.as<SsspCtl>(), mark_max);
		}
		count_launch(c, 2);
		if (g->n_hub) { // the hub accumulators start at "not reached"
			if constexpr (TR::narrow)
				ssspn_acc_fill_kernel<typename TR::N><<<div_up((uint64_t)acc_cells, 256), 256, 0, sl.st>>>(sl.acc.as<R>(),
				                                                                        (uint32_t)acc_cells);
			else
				sssp_acc_fill_kernel<TR><<<div_up((uint64_t)acc_cells, 256), 256, 0, sl.st>>>(sl.acc.as<R>(),
				                                                                          (uint32_t)acc_cells);
			PGQ_CUDA(cudaMemsetAsync(sl.acc.as<char>() + acc_cells * 8, 0, (size_t)g->n_hub * 4, sl.st));
			count_launch(c);
		}
		c->stats.batches++;
		return launch_chunk(si);
	};
	int rc_all = PGQ_OK;
	for (int si = 0; si < nslot && next < todo.size() && rc_all == PGQ_OK; si++) rc_all = start(si);
	int turn = 0;
	while (active > 0 && rc_all == PGQ_OK) {
		// oldest first: slots are served round-robin, so the one at `turn` has waited longest
		int si = -1;
		for (int k = 0; k < nslot; k++) {
			const int s = (turn + k) % nslot;
			if (run[s].batch >= 0) {
				si = s;
				break;
			}
		}
		turn = (si + 1) % nslot;
		SsspSlot &sl = c->sssp_slot[si];
		Run &r = run[si];
		if (cudaEventSynchronize(sl.ev) != cudaSuccess) {
			set_error("cheapest_path_length: batch failed: %s", cudaGetErrorString(cudaGetLastError()));
			rc_all = PGQ_ERR_CUDA;
			break;
		}
		const SsspCtl &hc = h_ctl[si];
		if (!hc.done) {
			if (r.rounds > (int64_t)V + 1) {
				set_error("cheapest_path_length: negative cycle reachable from a source");
				rc_all = PGQ_ERR_INVALID;
				break;
			}
			rc_all = launch_chunk(si);
			continue;
		}
		c->stats.levels += hc.rounds;
		c->stats.k_gathered += hc.hot_seen;    // stats mode: source rows read
		c->stats.k_scanned += hc.scanned_seen; // stats mode: in-edges scanned
		if (TR::narrow) {
			if (hc.overflow_seen) c->stats.sssp_redo_batches++;
			else if (sizeof(S) == 2) c->stats.sssp_narrow_batches++;
			else c->stats.sssp_narrow32_batches++;
		}
		if (TR::narrow && hc.overflow_seen) redo->push_back(r.batch);
		else {
			if constexpr (TR::narrow)
				ssspn_result_kernel<typename TR::N><<<1, L, 0, sl.st>>>(d_src + r.batch * L, d_dst + r.batch * L, sl.dist.as<S>(),
				                                         d_out + r.batch * L);
			else
				sssp_result_kernel<TR, T><<<1, L, 0, sl.st>>>(d_src + r.batch * L, d_dst + r.batch * L, sl.dist.as<S>(),
				                                            d_out + r.batch * L);
			count_launch(c);
		}
		r.batch = -1;
		active--;
		if (next < todo.size()) rc_all = start(si);
	}
	// the caller's stream continues after every slot
	for (int si = 0; si < nslot; si++) {
		if (rc_all == PGQ_OK) {
			PGQ_CUDA(cudaEventRecord(c->sssp_slot[si].ev, c->sssp_slot[si].st));
			PGQ_CUDA(cudaStreamWaitEvent(c->stream, c->sssp_slot[si].ev, 0));
		} else cudaStreamSynchronize(c->sssp_slot[si].st);
	}
	return rc_all;
}

// out8[slot] = distance of the lane slot (raw 8 bytes of T), +inf when the destination was not reached
template <typename T>
static int sssp_typed(Ctx *c, Csr *g, const LaneTable &lt, T *h_out, const T *iw64) {
	cudaStream_t st = c->stream;
	const int L = SSSP_LANES;
	const size_t nslots = (size_t)lt.nbatch * L;
	PGQ_TRY(c->bfs_lane.reserve(nslots * (2 * sizeof(uint32_t) + sizeof(T)) + 64));
	T *d_out = c->bfs_lane.as<T>(); // 8-byte aligned first
	uint32_t *d_src = reinterpret_cast<uint32_t *>(d_out + nslots);
	uint32_t *d_dst = d_src + nslots;
	PGQ_CUDA(cudaMemcpyAsync(d_src, lt.src.data(), nslots * 4, cudaMemcpyHostToDevice, st));
	PGQ_CUDA(cudaMemcpyAsync(d_dst, lt.dst.data(), nslots * 4, cudaMemcpyHostToDevice, st));
	c->stats.h2d_bytes += (uint64_t)nslots * 8;
	const bool prune = g->w_nonneg && c->opt.sssp_prune;
	std::vector<int64_t> todo((size_t)lt.nbatch), redo;
	for (int64_t b = 0; b < lt.nbatch; b++) todo[(size_t)b] = b;
	if constexpr (std::is_same<T, long long>::value) {
		// option sssp_narrow: 1 = 16-bit rows, then 32-bit rows for what did not fit; 2 = 32-bit rows only; 0 = neither
		if (g->w_nonneg && c->opt.sssp_narrow == 1) {
			if (!g->d_iw16.p && g->E > 0) {
				PGQ_TRY(g->d_iw16.reserve((size_t)g->E * 2));
				narrow_w_kernel<N16><<<div_up(g->E, 256), 256, 0, st>>>(iw64, (uint32_t)g->E,
				                                                        g->d_iw16.as<uint16_t>());
				count_launch(c);
			}
			PGQ_TRY((sssp_batches<Row16, T>(c, g, lt, todo, g->d_iw16.as<uint16_t>(), prune, d_out, d_src, d_dst,
			                                &redo)));
			todo.swap(redo);
			redo.clear();
		}
		if (g->w_nonneg && c->opt.sssp_narrow && !todo.empty()) {
			if (!g->d_iw32n.p && g->E > 0) {
				PGQ_TRY(g->d_iw32n.reserve((size_t)g->E * 4));
				narrow_w_kernel<N32><<<div_up(g->E, 256), 256, 0, st>>>(iw64, (uint32_t)g->E,
				                                                        g->d_iw32n.as<uint32_t>());
				count_launch(c);
			}
			PGQ_TRY((sssp_batches<Row32, T>(c, g, lt, todo, g->d_iw32n.as<uint32_t>(), prune, d_out, d_src, d_dst,
			                                &redo)));
			todo.swap(redo);
			redo.clear();
		}
	}
	if (!todo.empty())
		PGQ_TRY((sssp_batches<Row64<T>, T>(c, g, lt, todo, iw64, prune, d_out, d_src, d_dst, &redo)));
	PGQ_CUDA(cudaMemcpyAsync(h_out, d_out, nslots * sizeof(T), cudaMemcpyDeviceToHost, st));
	PGQ_CUDA(cudaStreamSynchronize(st));
	c->stats.d2h_bytes += (uint64_t)nslots * sizeof(T);
	PGQ_CUDA(cudaGetLastError());
	return PGQ_OK;
}

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
	if (g->wtype == PGQ_WEIGHT_I64)
		return sssp_typed<long long>(c, g, lt, static_cast<long long *>(h_out8), g->d_iw.as<long long>());
	if (g->w_whole && c->opt.sssp_narrow && g->E > 0) {
		// DOUBLE weights that are all non-negative whole numbers: the integer rows give the same bits (see SSSP_WHOLE_MAX)
		if (!g->d_iwi.p) {
			PGQ_TRY(g->d_iwi.reserve((size_t)g->E * 8));
			whole_w_kernel<<<div_up(g->E, 256), 256, 0, c->stream>>>(g->d_iw.as<double>(), (uint32_t)g->E,
			                                                         g->d_iwi.as<long long>());
			count_launch(c);
		}
		long long *hi = static_cast<long long *>(h_out8);
		PGQ_TRY(sssp_typed<long long>(c, g, lt, hi, g->d_iwi.as<long long>()));
		double *hd = static_cast<double *>(h_out8);
		const size_t nslots = (size_t)lt.nbatch * SSSP_LANES;
		for (size_t k = 0; k < nslots; k++) {
			const long long v = hi[k]; // read before the same 8 bytes are rewritten
			hd[k] = WInf<long long>::is_inf(v) ? WInf<double>::value() : (double)v;
		}
		return PGQ_OK;
	}
	return sssp_typed<double>(c, g, lt, static_cast<double *>(h_out8), g->d_iw.as<double>());
}

} // namespace pgq
