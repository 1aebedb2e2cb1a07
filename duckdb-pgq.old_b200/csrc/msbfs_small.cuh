// Whole-batch MS-BFS kernel for small graphs (BASELINE configs[0]: 10 k vertices / 180 k edges).
//
// On a graph whose bitmaps and edges live in L2, a level of the general path costs ~100 us of
// launches plus one host round trip while its kernels run for a few microseconds.  Here one
// cooperative launch runs the complete batch: zero the bitmaps, plant the sources, then per level a
// bottom-up sweep (one warp per vertex, its 32/TPV thread groups striding over the in-list; teams of
// 8 warps for the vertices of the higher in-degree classes), ONE grid barrier, and the destination
// check / lane retirement / stop decision evaluated redundantly by every block from the same data
// (level results travel through three rotating slots of the control block, so no second barrier is
// needed).  No host involvement until the batch is done.  Measured floor 4.4 us per level (a chain
// graph) against ~110 us per level for the general path on the same graph.  Semantics are those of the general path (R4/R5 of SURVEY.md §8(a)):
//   next[n] = OR_{v->n} visit[v] & ~seen[n] & active;  seen[n] |= next[n];
//   lanes whose destination is in `seen` get found = iter and retire, exhausted lanes retire.
// Every level writes next[n] for every n, so neither the ping-pong pair nor the per-level history
// buffers (shortestpath) need clearing.
#pragma once
#include "msbfs_kernels.cuh"

namespace pgq {

static constexpr int SMALL_THREADS = 1024;

struct SmallArgs {
	const uint32_t *ioff; // reverse CSR
	const uint32_t *isrc;
	uint32_t V;
	uint32_t team_threads;  // pass B: threads per vertex team (256 or 1024)
	uint32_t warp_max;      // in-degree up to which a warp handles the vertex; above: the class lists
	const uint32_t *big[3]; // vertices with more in-edges (med / mid / hub lists of csr_prepare)
	uint32_t nbig[3];
	uint32_t *seen;
	uint32_t *buf[2];      // frontier ping-pong (hist == nullptr)
	uint32_t *const *hist; // level buffers 0..max_levels (shortestpath keeps every frontier)
	int max_levels;
	int level_cap; // stop after this many levels whatever is still open (iterativelength upper bound, R11)
	const uint32_t *lane_src;
	const uint32_t *lane_dst;
	int32_t *found;
	BfsCtrl *ctrl;
};

// L2 loads for words other blocks wrote before the last grid barrier
template <int SW>
__device__ __forceinline__ Slice<SW> ldcg_slice(const uint32_t *p) {
	Slice<SW> s;
	if (SW == 4) {
		uint4 v = __ldcg(reinterpret_cast<const uint4 *>(p));
		s.w[0] = v.x; s.w[1] = v.y; s.w[2] = v.z; s.w[3] = v.w;
	} else {
		uint2 v = __ldcg(reinterpret_cast<const uint2 *>(p));
		s.w[0] = v.x; s.w[1] = v.y;
	}
	return s;
}

// OR of the frontier words of in-edges j0, j0+stride, ... < hi (slice t), two loads in flight
template <int W>
__device__ __forceinline__ Slice<LaneCfg<W>::SW> small_gather(const SmallArgs &a, const uint32_t *cur, uint32_t j0,
                                                              uint32_t hi, uint32_t stride, unsigned t) {
	constexpr int SW = LaneCfg<W>::SW, WPV = LaneCfg<W>::WPV;
	Slice<SW> acc = slice_zero<SW>();
	uint32_t j = j0;
	for (; j + stride < hi; j += 2 * stride) {
		const uint32_t s0 = __ldg(a.isrc + j), s1 = __ldg(a.isrc + j + stride);
		const Slice<SW> w0 = ldcg_slice<SW>(cur + (size_t)s0 * WPV + t * SW);
		const Slice<SW> w1 = ldcg_slice<SW>(cur + (size_t)s1 * WPV + t * SW);
#pragma unroll
		for (int i = 0; i < SW; i++) acc.w[i] |= w0.w[i] | w1.w[i];
	}
	if (j < hi) {
		const Slice<SW> w0 = ldcg_slice<SW>(cur + (size_t)__ldg(a.isrc + j) * WPV + t * SW);
#pragma unroll
		for (int i = 0; i < SW; i++) acc.w[i] |= w0.w[i];
	}
	return acc;
}

// next[v] = news (always written); new bits go to seen[v], the block's alive mask and change flag
template <int W>
__device__ __forceinline__ void small_finish(const SmallArgs &a, uint32_t *next, size_t wo,
                                             const Slice<LaneCfg<W>::SW> &news, Slice<LaneCfg<W>::SW> sn,
                                             uint32_t *s_alive, uint32_t *s_change, unsigned t) {
	constexpr int SW = LaneCfg<W>::SW;
	st_slice<SW>(next + wo, news);
	if (slice_any<SW>(news)) {
#pragma unroll
		for (int i = 0; i < SW; i++) {
			sn.w[i] |= news.w[i];
			if (news.w[i] & ~s_alive[t * SW + i]) atomicOr(&s_alive[t * SW + i], news.w[i]);
		}
		st_slice<SW>(a.seen + wo, sn);
		*s_change = 1;
	}
}

// Grid barrier on a monotonically increasing counter (zeroed by the host before the launch; the
// cooperative launch guarantees every block is resident).  Barrier number k is complete when the
// counter reaches k * gridDim.x.  About 1 us against ~3 us for cooperative_groups' grid.sync().
__device__ __forceinline__ void small_barrier(unsigned *counter, unsigned &target) {
	__syncthreads();
	if (threadIdx.x == 0) {
		target += gridDim.x;
		__threadfence();
		atomicAdd(counter, 1u);
		while (*reinterpret_cast<volatile unsigned *>(counter) < target) {
		}
		__threadfence();
	}
	__syncthreads();
}

template <int W>
__global__ void __launch_bounds__(SMALL_THREADS) msbfs_small_kernel(SmallArgs a) {
	using L = LaneCfg<W>;
	constexpr int SW = L::SW, TPV = L::TPV, WPV = L::WPV, G = L::G;
	constexpr int NW = W / 32;
	__shared__ __align__(16) uint32_t s_active[NW]; // every block keeps its own (identical) lane mask
	__shared__ uint32_t s_alive[NW];
	__shared__ uint32_t s_acc[(SMALL_THREADS / 256) * NW];
	__shared__ uint32_t s_change, s_cnt;
	BfsCtrl *ctrl = a.ctrl;
	unsigned bar_target = 0;
	const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
	const unsigned lane = threadIdx.x & 31, t = lane % TPV, grp = lane / TPV;
	const uint32_t gwarp = tid >> 5, nwarps = nth >> 5;
	const size_t words = (size_t)a.V * WPV;
	uint32_t *cur = a.hist ? a.hist[0] : a.buf[0];

	for (size_t i = tid; i < words; i += nth) {
		a.seen[i] = 0;
		cur[i] = 0;
	}
	if (threadIdx.x < (SMALL_THREADS / 256) * NW) s_acc[threadIdx.x] = 0;
	if (threadIdx.x == 0) s_cnt = 0;
	__syncthreads();
	if (threadIdx.x < NW) { // lanes in use
		uint32_t m = 0;
		for (int b = 0; b < 32; b++)
			if (a.lane_src[threadIdx.x * 32 + b] != 0xFFFFFFFFu) m |= 1u << b;
		s_active[threadIdx.x] = m;
		atomicAdd(&s_cnt, (uint32_t)__popc(m));
	}
	if (blockIdx.x == 0)
		for (int l = threadIdx.x; l < W; l += SMALL_THREADS) a.found[l] = -1;
	small_barrier(&ctrl->small_bar, bar_target);
	if (s_cnt == 0) return;
	if (blockIdx.x == 0) { // plant the sources
		for (int l = threadIdx.x; l < W; l += SMALL_THREADS) {
			const uint32_t s = a.lane_src[l];
			if (s != 0xFFFFFFFFu) atomicOr(&cur[(size_t)s * WPV + (l >> 5)], 1u << (l & 31));
		}
	}
	small_barrier(&ctrl->small_bar, bar_target);

	for (int iter = 1;; iter++) {
		uint32_t *next = a.hist ? a.hist[iter] : a.buf[iter & 1];
		uint32_t *slot = ctrl->small_slots[iter % 3]; // [0..NW) alive lanes of this level, [16] change
		if (threadIdx.x < NW) s_alive[threadIdx.x] = 0;
		if (threadIdx.x == 0) {
			s_change = 0;
			s_cnt = 0;
		}
		__syncthreads();
		const Slice<SW> act = ld_slice<SW>(s_active + t * SW);
		// pass A: one warp per vertex with at most warp_max in-edges
		for (uint32_t v = gwarp; v < a.V; v += nwarps) {
			const uint32_t lo = __ldg(a.ioff + v), hi = __ldg(a.ioff + v + 1);
			if (hi - lo > a.warp_max) continue; // pass B
			const size_t wo = (size_t)v * WPV + t * SW;
			Slice<SW> news = slice_zero<SW>();
			Slice<SW> sn = slice_zero<SW>();
			if (lo < hi) {
				sn = ld_slice<SW>(a.seen + wo); // only this warp ever writes seen[v]
				Slice<SW> unseen;
#pragma unroll
				for (int i = 0; i < SW; i++) unseen.w[i] = ~sn.w[i] & act.w[i];
				if (__any_sync(0xffffffffu, slice_any<SW>(unseen))) {
					Slice<SW> acc = small_gather<W>(a, cur, lo + grp, hi, G, t);
					acc = warp_or_groups<SW, TPV>(acc);
#pragma unroll
					for (int i = 0; i < SW; i++) news.w[i] = acc.w[i] & unseen.w[i];
				}
			}
			if (grp == 0) small_finish<W>(a, next, wo, news, sn, s_alive, &s_change, t);
		}
		// pass B: a team of 8 warps per vertex with more in-edges (the in-degree classes above
		// `small`); the block's teams work on different vertices and meet on their own named barrier
		{
			const unsigned TEAM = a.team_threads, NTEAM = SMALL_THREADS / TEAM;
			const unsigned team = threadIdx.x / TEAM, tt = threadIdx.x % TEAM;
			uint32_t *acc_t = s_acc + team * NW;
			for (int cls = 0; cls < 3; cls++) {
				const uint32_t *list = a.big[cls];
				const uint32_t nlist = a.nbig[cls];
				for (uint32_t idx = blockIdx.x * NTEAM + team; idx < nlist; idx += gridDim.x * NTEAM) {
					const uint32_t v = __ldg(list + idx);
					const uint32_t lo = __ldg(a.ioff + v), hi = __ldg(a.ioff + v + 1);
					const size_t wo = (size_t)v * WPV + t * SW;
					const Slice<SW> sn = ldcg_slice<SW>(a.seen + wo);
					Slice<SW> unseen;
#pragma unroll
					for (int i = 0; i < SW; i++) unseen.w[i] = ~sn.w[i] & act.w[i];
					const bool open = __any_sync(0xffffffffu, slice_any<SW>(unseen)); // same in every warp
					if (open) {
						Slice<SW> acc = small_gather<W>(a, cur, lo + (tt >> 5) * G + grp, hi, (TEAM / 32) * G, t);
						acc = warp_or_groups<SW, TPV>(acc);
						if (grp == 0) {
#pragma unroll
							for (int i = 0; i < SW; i++)
								if (acc.w[i]) atomicOr(&acc_t[t * SW + i], acc.w[i]);
						}
					}
					asm volatile("bar.sync %0, %1;" ::"r"(team + 1), "r"(TEAM) : "memory");
					if (tt < TPV) { // the team's first warp, group 0: lane == t
						Slice<SW> news;
#pragma unroll
						for (int i = 0; i < SW; i++) {
							news.w[i] = acc_t[t * SW + i] & unseen.w[i];
							acc_t[t * SW + i] = 0;
						}
						small_finish<W>(a, next, wo, news, sn, s_alive, &s_change, t);
					}
					asm volatile("bar.sync %0, %1;" ::"r"(team + 1), "r"(TEAM) : "memory");
				}
			}
		}
		__syncthreads();
		if (threadIdx.x < NW && s_alive[threadIdx.x]) atomicOr(&slot[threadIdx.x], s_alive[threadIdx.x]);
		if (threadIdx.x == 0 && s_change) slot[16] = 1u;
		small_barrier(&ctrl->small_bar, bar_target);
		// Every block evaluates the level's outcome for itself (identical inputs, identical result):
		// destination check, lane retirement, continue/stop.  Block 0 records it.
		// The slot of level iter-1 is free again: nobody reads it after this barrier, and it is next
		// written in level iter+2, i.e. after block 0 has passed the next barrier.
		if (blockIdx.x == 0 && threadIdx.x < 20) ctrl->small_slots[(iter + 2) % 3][threadIdx.x] = 0;
		if (threadIdx.x < W) {
			const int l = threadIdx.x;
			const uint32_t word = s_active[l >> 5];
			const bool on = (word >> (l & 31)) & 1u;
			bool hit = false;
			if (on) {
				const uint32_t d = a.lane_dst[l];
				hit = (__ldcg(&a.seen[(size_t)d * WPV + (l >> 5)]) >> (l & 31)) & 1u;
				if (hit && blockIdx.x == 0) a.found[l] = iter;
			}
			const unsigned hm = __ballot_sync(0xffffffffu, hit);
			if (lane == 0) {
				const uint32_t nw = word & ~hm & __ldcg(&slot[l >> 5]); // found or exhausted lanes retire
				s_active[l >> 5] = nw;
				atomicAdd(&s_cnt, (uint32_t)__popc(nw));
			}
		}
		__syncthreads();
		bool cont = __ldcg(&slot[16]) != 0 && s_cnt > 0 && iter < a.level_cap;
		bool overflow = false;
		if (cont && a.hist && iter >= a.max_levels) { // no buffer for another level
			cont = false;
			overflow = true;
		}
		if (!cont) {
			if (tid == 0) {
				ctrl->n_active = s_cnt;
				ctrl->small_levels = (uint32_t)iter;
				ctrl->small_overflow = overflow ? 1u : 0u;
			}
			break;
		}
		cur = next;
	}
}

} // namespace pgq
