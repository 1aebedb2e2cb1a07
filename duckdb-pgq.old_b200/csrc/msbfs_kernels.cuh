// MS-BFS level kernels (K4-K7), pull direction over the reverse CSR.
//
// Semantics restated from SURVEY.md §8(a) R4/R5/R7: one level computes, for every vertex n,
//   next[n] = (OR over in-neighbours v of visit[v]) & ~seen[n];  seen[n] |= next[n]
// which is exactly the reference's push loop (next[e[off]] |= visit[v]; next &= ~seen;
// seen |= next) evaluated from the target side — no atomics, deterministic.  Lanes retired by
// the destination check are masked out through `active` (per-lane early exit).
//
// Parent rule R7 falls out of the reverse CSR's order: the in-list of n is sorted by
// (source id ascending, forward CSR offset ascending), so the FIRST in-edge whose source has
// lane l in the frontier, at the level n is first reached by l, is the reference's parent.
// shortestpath therefore runs the same level kernels as iterativelength and only KEEPS every
// level's frontier; bfs_paths_hist_kernel finds parents by that rule while walking back from the
// destination (no V x W parent array, no per-bit parent stores in the hot loop).
//
// Layout: bitmaps are uint32[V][W/32]; a thread owns a 128-bit slice (64-bit when W=64) of one
// vertex word and moves it with one vector load/store.  HBM-bound: per level
//   4*V*(W/8) + E_act*(4 + W/8) bytes.
#pragma once
#include "common.cuh"

namespace pgq {

// Device-side level loop.  After every level the check kernel rotates the frontier
// views and decides the direction of the next level; the level kernels read their frontier pointers
// from here, so the same kernel nodes serve every level of a CUDA-graph WHILE loop and the host is
// not needed between the init kernel and the result copy.  The host-driven loop (options "stats",
// "timing", "trace", "loop"=0) reads `mode` back after every level and launches the same bodies.
enum LoopMode : uint32_t {
	MODE_STOP = 0,
	MODE_PUSH = 1,         // top-down over the work queue
	MODE_PULL = 2,         // bottom-up, frontier-bitmap filter off (the frontier owns most edges)
	MODE_PULL_F_EARLY = 3, // bottom-up, filter on, few vertices discovered so far (edge-parallel medium class)
	MODE_PULL_F_LATE = 4,  // bottom-up, filter on, tail levels
	MODE_COUNT = 5
};

struct LoopState {
	// this worker's buffers (constant over the batch)
	uint32_t *visit[2];
	uint32_t *fbit[3];
	uint2 *queue[2];
	uint32_t *const *hist; // shortestpath: table of per-level frontier buffers, else nullptr
	uint32_t hist_cap;     // entries of hist
	uint32_t pad0;
	// views of the level about to run
	const uint32_t *cur_visit;
	uint32_t *cur_next;
	const uint32_t *cur_fbit; // bitmap of cur_visit
	uint32_t *nxt_fbit;       // built for cur_next
	uint32_t *old_fbit;       // describes what cur_next still holds (frontier of two levels ago)
	const uint2 *cur_queue;
	uint2 *nxt_queue;
	uint32_t cur, fb[3], q, have_queue;
	uint32_t iter;        // level about to run (1-based)
	uint32_t mode;        // LoopMode of that level
	uint32_t stop_reason; // 1 finished (no lane active / nothing changed / level cap), 2 hist table full
	uint32_t level_cap;
	unsigned long long cum_verts; // frontier vertices of all levels so far
	// direction heuristic (Options)
	uint32_t direction, push_beta, push_vbeta, filter_gamma;
	unsigned long long E, V, n_listed;
	uint32_t levels_run, push_levels;
	// this batch's lanes (destination check)
	const uint32_t *lane_dst;
	int32_t *found;
	uint32_t nk[MODE_COUNT]; // kernels one level of each mode launches (launch statistics without a host in the loop)
	uint32_t pad1;
};

// Launch bounds of the gather kernels.  A/B builds (tools/r02_ab_build.sh) add a minimum number of resident blocks
// per SM; note that asking for a minimum of 1 is NOT neutral: ptxas then spends 62 registers on pull_list instead of
// 48 and the level gets slower (gpurun check 20).
#ifdef PGQ_LIST_MINB
#define PGQ_LIST_BOUNDS __launch_bounds__(128, PGQ_LIST_MINB)
#else
#define PGQ_LIST_BOUNDS __launch_bounds__(128)
#endif
#ifdef PGQ_WARP_MINB
#define PGQ_WARP_BOUNDS __launch_bounds__(128, PGQ_WARP_MINB)
#else
#define PGQ_WARP_BOUNDS __launch_bounds__(128)
#endif
#ifdef PGQ_SEG_MINB
#define PGQ_SEG_BOUNDS __launch_bounds__(128, PGQ_SEG_MINB)
#else
#define PGQ_SEG_BOUNDS __launch_bounds__(128)
#endif

struct BfsCtrl {
	uint32_t active[32]; // lane mask, W/32 words used (W <= 1024); first so slices are 16 B aligned
	uint32_t alive[32];  // lanes that gained a frontier bit in the running level (others are exhausted)
	uint32_t change_out; // snapshot of `change` taken by the check kernel
	uint32_t n_active;   // lanes still searching
	uint32_t pad[2];
	unsigned long long edges_lane;
	unsigned long long edges_unique;
	unsigned long long verts_active;
	unsigned long long pad2;
	// snapshot of the frontier the running level consumes (written by the init / check kernels only)
	unsigned long long out_edges;
	uint32_t out_verts;
	uint32_t out_items;
	// The frontier being produced by the running level: the ONLY words of this block the level kernels write (one
	// atomicAdd per warp or block), on a 128-byte line of their own -- every warp of a level kernel reads `active`,
	// `out_items` and the views in `L`, and a load queues behind the atomics aimed at its line (the effect measured on
	// cheapest_path_length's control block, profiles/r02_sssp_relax_before_ctl_fix.txt).
	alignas(128) unsigned long long next_edges; // sum of out-degrees of vertices with a non-empty next word
	uint32_t next_verts;
	uint32_t next_items; // work items written to the next push queue
	uint32_t change;     // set by level kernels when any new bit appeared
	alignas(128) uint32_t line_end; // (the next field starts a new line)
	// whole-batch kernel for small graphs (msbfs_small_kernel)
	uint32_t small_bar;      // grid barrier counter (zeroed by the host before the launch)
	uint32_t small_levels;   // levels run
	uint32_t small_overflow; // ran out of level buffers (shortestpath history): host falls back
	uint32_t pad3;
	uint32_t small_slots[3][20]; // per level (mod 3): alive lanes [0..15], change flag [16]
	// what the level kernels actually touched (instrumented instantiations only, option "stats"=1):
	// the roofline model of DESIGN.md section 4 is computed from these
	unsigned long long k_listed;   // class-list entries visited (list entry + two offsets + done bit)
	unsigned long long k_open;     // vertices whose `seen` word was read
	unsigned long long k_new;      // vertices that stored a `next` word and rewrote `seen`
	unsigned long long k_scanned;  // in-edges (pull) / out-edges (push) whose endpoint id was read
	unsigned long long k_gathered; // frontier words gathered
	unsigned long long kp_items;   // push: work items (queue entry + frontier word + two offsets)
	unsigned long long kp_edges;   // push: out-edges whose target id was read
	unsigned long long kp_seen;    // push: 4-byte atomicOr on `seen`
	unsigned long long kp_next;    // push: 4-byte atomicOr on `next`
	// running totals of this worker (never reset on the device; the host folds the deltas into pgq_stats_t)
	unsigned long long tot_levels, tot_push_levels, tot_launches;
	LoopState L;
};

// Instrumented kernels: per-thread counts -> one atomicAdd per warp and counter (untimed pass only).
__device__ __forceinline__ void stat_add(unsigned long long *dst, uint32_t mine) {
	const uint32_t tot = __reduce_add_sync(0xffffffffu, mine);
	if ((threadIdx.x & 31) == 0 && tot) atomicAdd(dst, (unsigned long long)tot);
}

// Lanes that gained a frontier bit in the running level ("alive").  A lane with an empty frontier
// can never reach anything new; the check kernel retires it so it stops blocking the pull
// kernels' early exit.  Producers OR their new words into a block-shared mask and flush it once
// per block into slot blockIdx % ALIVE_SLOTS (spread addresses: no hot atomics); the check kernel
// folds the slots into BfsCtrl::alive.
static constexpr uint32_t ALIVE_SLOTS = 128;

// Producers keep a block-shared copy of their slot (blockIdx % ALIVE_SLOTS), preloaded from the
// slot so knowledge carries over between blocks, and touch the global slot only for bits the copy
// does not have yet: after the first few blocks of a level this is shared-memory reads only.
__device__ __forceinline__ void alive_init(uint32_t *salive, const uint32_t *alive_slots) {
	if (threadIdx.x < 16) salive[threadIdx.x] = alive_slots[(blockIdx.x % ALIVE_SLOTS) * 16 + threadIdx.x];
	__syncthreads();
}

template <int SW>
__device__ __forceinline__ void alive_note(uint32_t *salive, uint32_t *alive_slots, uint32_t word0, const uint32_t *w) {
#pragma unroll
	for (int i = 0; i < SW; i++) {
		const uint32_t nb = w[i] & ~salive[word0 + i];
		if (nb) {
			atomicOr(&salive[word0 + i], nb);
			atomicOr(&alive_slots[(blockIdx.x % ALIVE_SLOTS) * 16 + word0 + i], nb);
		}
	}
}

// Push work item: out-edges [off[v] + chunk*PUSH_CHUNK, +PUSH_CHUNK) of frontier vertex v.
static constexpr uint32_t PUSH_CHUNK = 128; // a warp walks its chunk 32 edges per step with dependent atomics: 1024-edge chunks left level 1 waiting 0.15 ms on the longest one

struct PullArgs {
	const uint32_t *ioff;
	const uint32_t *isrc;
	const uint32_t *visit;
	uint32_t *next;
	uint32_t *seen;
	BfsCtrl *ctrl;
	uint32_t V;
	uint32_t small_max;
	const int64_t *off;   // forward offsets (out-degree of new frontier vertices, direction heuristic)
	const uint32_t *list; // mid or hub vertex list
	uint32_t nlist;
	const HubChunk *chunks;
	uint32_t nchunks;
	uint32_t *hubacc;
	const uint32_t *fcur; // 1 bit per vertex: visit word non-empty (may over-approximate, never under)
	uint32_t *fnext;      // same for `next`, produced by this level
	uint32_t *alive_slots;   // [ALIVE_SLOTS][16] per-block-slot OR of the words produced this level
	// 1 bit per vertex: seen[n] already holds every lane that is still active.  `active` only ever shrinks,
	// so a done vertex stays done for the rest of the batch and costs neither its 64-byte `seen` read nor
	// anything else.  nullptr = not tracked (partitioned path).
	uint32_t *done;
	// 1 = visit / next / fcur / fnext come from ctrl->L (the level loop's rotating views), 0 = from the
	// fields above (partitioned path)
	uint32_t dyn;
	// dense = 1: every listed vertex stores its `next` word, zero or not (unfiltered levels, where most
	// vertices gain bits anyway and one streaming store beats a scrub pass plus scattered stores);
	// dense = 0: `next` is all-zero on entry and only words that gained bits are written.
	uint32_t dense;
	uint32_t use_done; // test the done bitmap (tail levels); it is maintained on every level
	// 1-D partition, fused exchange: the other ranks' `next` replicas (peer memory over NVLink), each
	// already offset to THIS rank's slice.  A vertex that gained bits is stored to every replica by
	// the thread that produced it, so the frontier words never go through a separate collective.
	uint32_t *const *peer_next;
	uint32_t npeers;
};


template <int W>
struct LaneCfg {
	static constexpr int SW = (W >= 128) ? 4 : 2;         // uint32 words per thread slice
	static constexpr int TPV = (W >= 128) ? (W / 128) : 1; // threads per vertex word
	static constexpr int WPV = W / 32;                     // uint32 words per vertex
	static constexpr int G = 32 / TPV;                     // vertex-slices (edge groups) per warp
};

template <int SW>
struct Slice {
	uint32_t w[SW];
};

template <int SW>
__device__ __forceinline__ Slice<SW> slice_zero() {
	Slice<SW> s;
#pragma unroll
	for (int i = 0; i < SW; i++) s.w[i] = 0;
	return s;
}

template <int SW>
__device__ __forceinline__ Slice<SW> ld_slice(const uint32_t *p) {
	Slice<SW> s;
	if (SW == 4) {
		uint4 v = *reinterpret_cast<const uint4 *>(p);
		s.w[0] = v.x; s.w[1] = v.y; s.w[2] = v.z; s.w[3] = v.w;
	} else {
		uint2 v = *reinterpret_cast<const uint2 *>(p);
		s.w[0] = v.x; s.w[1] = v.y;
	}
	return s;
}

// read-only (non-coherent) gather of a frontier slice written by the previous launch
template <int SW>
__device__ __forceinline__ Slice<SW> ldg_slice(const uint32_t *p) {
	Slice<SW> s;
	if (SW == 4) {
		uint4 v = __ldg(reinterpret_cast<const uint4 *>(p));
		s.w[0] = v.x; s.w[1] = v.y; s.w[2] = v.z; s.w[3] = v.w;
	} else {
		uint2 v = __ldg(reinterpret_cast<const uint2 *>(p));
		s.w[0] = v.x; s.w[1] = v.y;
	}
	return s;
}

template <int SW>
__device__ __forceinline__ void st_slice(uint32_t *p, const Slice<SW> &s) {
	if (SW == 4) *reinterpret_cast<uint4 *>(p) = make_uint4(s.w[0], s.w[1], s.w[2], s.w[3]);
	else *reinterpret_cast<uint2 *>(p) = make_uint2(s.w[0], s.w[1]);
}

template <int SW>
__device__ __forceinline__ void st_peers(const PullArgs &a, size_t wo, const Slice<SW> &news) {
	for (uint32_t p = 0; p < a.npeers; p++) st_slice<SW>(a.peer_next[p] + wo, news);
}

template <int SW>
__device__ __forceinline__ bool slice_any(const Slice<SW> &s) {
	uint32_t o = 0;
#pragma unroll
	for (int i = 0; i < SW; i++) o |= s.w[i];
	return o != 0;
}

template <int SW>
__device__ __forceinline__ bool slice_covers(const Slice<SW> &acc, const Slice<SW> &need) {
	uint32_t miss = 0;
#pragma unroll
	for (int i = 0; i < SW; i++) miss |= need.w[i] & ~acc.w[i];
	return miss == 0;
}

// OR-reduce across the G groups of a warp (lanes with equal lane % TPV); result in every lane.
template <int SW, int TPV>
__device__ __forceinline__ Slice<SW> warp_or_groups(Slice<SW> s) {
#pragma unroll
	for (int d = TPV; d < 32; d <<= 1) {
#pragma unroll
		for (int i = 0; i < SW; i++) s.w[i] |= __shfl_xor_sync(0xffffffffu, s.w[i], d);
	}
	return s;
}

__device__ __forceinline__ void snapshot_frontier(BfsCtrl *ctrl) {
	ctrl->out_edges = ctrl->next_edges;
	ctrl->out_verts = ctrl->next_verts;
	ctrl->out_items = ctrl->next_items;
	ctrl->next_edges = 0;
	ctrl->next_verts = 0;
	ctrl->next_items = 0;
}

__device__ __forceinline__ bool fbit_test(const uint32_t *f, uint32_t v) {
	return (__ldg(f + (v >> 5)) >> (v & 31)) & 1u;
}

// A/B switches for bisecting kernel cost (tools/r02_ab.sh builds library variants with -DPGQ_AB_...)
#ifdef PGQ_AB_NODONE
static constexpr bool kDone = false;
#else
static constexpr bool kDone = true;
#endif
#ifdef PGQ_AB_NODYN
static constexpr bool kDyn = false;
#else
static constexpr bool kDyn = true;
#endif
#ifdef PGQ_AB_DENSE
static constexpr bool kSparse = false;
#else
static constexpr bool kSparse = true;
#endif

// The level's frontier views.  With the device-side level loop they rotate from level to level and are read
// from the control block.  Only `visit` (and `fcur` on filtered levels) is looked up before the gather loop;
// `next` / `fnext` are looked up where the results are stored, AFTER the loop: keeping all four pointers live
// across the loop cost 10 % of a batch (measured: 8.84 ms -> 7.9 ms per s24 batch, profiles/r02_ab_views.md).
__device__ __forceinline__ const uint32_t *view_visit(const PullArgs &a) {
	return (kDyn && a.dyn) ? a.ctrl->L.cur_visit : a.visit;
}
__device__ __forceinline__ const uint32_t *view_fcur(const PullArgs &a) {
	return (kDyn && a.dyn) ? a.ctrl->L.cur_fbit : a.fcur;
}
__device__ __forceinline__ uint32_t *view_next(const PullArgs &a) {
	return (kDyn && a.dyn) ? a.ctrl->L.cur_next : a.next;
}
__device__ __forceinline__ uint32_t *view_fnext(const PullArgs &a) {
	return (kDyn && a.dyn) ? a.ctrl->L.nxt_fbit : a.fnext;
}

// done bit of vertex n (plain load: other vertices' bits of the same word may be set by this launch)
__device__ __forceinline__ bool done_test(const uint32_t *done, uint32_t n) {
	return done && ((done[n >> 5] >> (n & 31)) & 1u);
}

template <int TPV>
__device__ __forceinline__ bool group_any(bool x) {
	unsigned b = x ? 1u : 0u;
#pragma unroll
	for (int d = 1; d < TPV; d <<= 1) b |= __shfl_xor_sync(0xffffffffu, b, d);
	return b != 0;
}

// ---- K4/K7: pull over one in-degree class --------------------------------------------------------
// A vertex is served by GPV thread groups (a group = the TPV threads that together hold one
// W-bit word); a warp therefore serves VPW = G/GPV list entries.  Round r of a vertex covers
// in-edges [lo + r*GPV*U, +GPV*U): sub-group sg takes edges base + u*GPV + sg.  The sources of
// the NEXT round are fetched before the gathers of the current one are consumed, so a round costs
// one gather latency instead of index + bitmap + gather.  FILTER skips in-neighbours whose
// frontier bit is clear before their word is touched (pays off unless the frontier holds most
// edges).  Loops are warp-uniform (trip count = the slowest vertex of the warp) because the
// reductions use full-mask shuffles.
template <int W, int GPV, bool FILTER, bool STATS = false>
__global__ void PGQ_LIST_BOUNDS pull_list_kernel(PullArgs a) {
	// unfiltered (dense) levels: every listed vertex stores its word and no done-bitmap code at all (the code shape
	// of round 1: measured 3.11 vs 3.40 ms on the dense level of an s24 batch); filtered levels: sparse stores + done
	constexpr bool SP = FILTER;
	const uint32_t *const visit_p = view_visit(a);
	const uint32_t *const fcur_p = view_fcur(a);
	using C = LaneCfg<W>;
	constexpr int SW = C::SW, TPV = C::TPV, WPV = C::WPV, G = C::G, U = 4;
	static_assert(GPV >= 1 && GPV <= G, "groups per vertex");
	constexpr int VPW = G / GPV;
	constexpr uint32_t INV = 0xFFFFFFFFu, FULL = 0xffffffffu;
	__shared__ uint32_t salive[16];
	alive_init(salive, a.alive_slots);
	const uint32_t wid = (uint32_t)(((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
	if ((uint64_t)wid * VPW >= a.nlist) return; // warp-uniform (an exited warp counts as arrived at barriers)
	const uint32_t lane = threadIdx.x & 31, g = lane / TPV, t = lane % TPV;
	const uint32_t vslot = g / GPV, sg = g % GPV;
	const uint64_t idx = (uint64_t)wid * VPW + vslot;
	const bool listed = idx < a.nlist;
	const uint32_t n = listed ? a.list[idx] : 0u;
	const bool valid = listed && !(SP && kDone && a.use_done && done_test(a.done, n)); // a done vertex is skipped altogether
	uint32_t lo = 0, hi = 0;
	if (valid) {
		lo = a.ioff[n];
		hi = a.ioff[n + 1];
	}
	const size_t wo = (size_t)n * WPV + (size_t)t * SW;
	Slice<SW> s = slice_zero<SW>(), unseen = slice_zero<SW>();
	uint32_t c_scan = 0, c_gath = 0; // STATS
	// Software pipeline, three rounds deep: sources of round r+2 are being loaded and the frontier
	// bits of round r+1 tested while the gathers of round r are in flight.  The first two rounds'
	// sources do not depend on `seen`: issue them before it.
	uint32_t j = lo + sg;
	uint32_t vs[U], v1[U];
#pragma unroll
	for (int u = 0; u < U; u++) {
		const uint32_t jj = j + u * GPV;
		vs[u] = jj < hi ? a.isrc[jj] : INV;
		if (STATS && t == 0 && jj < hi) c_scan++;
	}
#pragma unroll
	for (int u = 0; u < U; u++) {
		const uint32_t jj = j + GPV * U + u * GPV;
		v1[u] = jj < hi ? a.isrc[jj] : INV;
		if (STATS && t == 0 && jj < hi) c_scan++;
	}
	if (valid) {
		s = ld_slice<SW>(a.seen + wo);
		const Slice<SW> act = ld_slice<SW>(a.ctrl->active + t * SW);
#pragma unroll
		for (int i = 0; i < SW; i++) unseen.w[i] = ~s.w[i] & act.w[i];
	}
	bool open = slice_any<SW>(unseen); // identical in all sub-groups of the vertex (same t)
	Slice<SW> acc = slice_zero<SW>();  // this sub-group's OR
	Slice<SW> tot = slice_zero<SW>();
	uint32_t fm = 0; // bit u: edge u of the current round has its source in the frontier
#pragma unroll
	for (int u = 0; u < U; u++) {
		bool f = vs[u] != INV;
		if (FILTER && f) f = fbit_test(fcur_p, vs[u]);
		fm |= (f ? 1u : 0u) << u;
	}
	while (__any_sync(FULL, open && j < hi)) {
		Slice<SW> w[U];
#pragma unroll
		for (int u = 0; u < U; u++) {
			w[u] = (open && ((fm >> u) & 1u)) ? ldg_slice<SW>(visit_p + (size_t)vs[u] * WPV + t * SW) : slice_zero<SW>();
			if (STATS && t == 0 && open && ((fm >> u) & 1u)) c_gath++;
		}
		const uint32_t jn = j + GPV * U;
		uint32_t v2[U];
#pragma unroll
		for (int u = 0; u < U; u++) {
			const uint32_t jj = jn + GPV * U + u * GPV;
			v2[u] = (open && jj < hi) ? a.isrc[jj] : INV;
			if (STATS && t == 0 && open && jj < hi) c_scan++;
		}
		uint32_t fn = 0;
#pragma unroll
		for (int u = 0; u < U; u++) {
			bool f = v1[u] != INV;
			if (FILTER && f) f = fbit_test(fcur_p, v1[u]);
			fn |= (f ? 1u : 0u) << u;
		}
#pragma unroll
		for (int u = 0; u < U; u++) {
#pragma unroll
			for (int i = 0; i < SW; i++) acc.w[i] |= w[u].w[i];
		}
		tot = acc;
#pragma unroll
		for (int d = TPV; d < TPV * GPV; d <<= 1) {
#pragma unroll
			for (int i = 0; i < SW; i++) tot.w[i] |= __shfl_xor_sync(FULL, tot.w[i], d);
		}
		open = open && !slice_covers<SW>(tot, unseen);
		j = jn;
		fm = fn;
#pragma unroll
		for (int u = 0; u < U; u++) {
			vs[u] = v1[u];
			v1[u] = v2[u];
		}
	}
	Slice<SW> news;
#pragma unroll
	for (int i = 0; i < SW; i++) news.w[i] = tot.w[i] & unseen.w[i];
	// `next` is all-zero on entry (the level's clear pass scrubbed what the frontier of two levels ago
	// left there), so only words that gained bits are written: a vertex without news costs no store.
	const bool writer = valid && sg == 0;
	bool rest = false; // some active lane of my slice is still unseen after this level
#pragma unroll
	for (int i = 0; i < SW; i++) rest |= (unseen.w[i] & ~news.w[i]) != 0u;
	const unsigned bal = __ballot_sync(FULL, writer && slice_any<SW>(news));
	const unsigned rbal = __ballot_sync(FULL, writer && rest);
	const unsigned gmask = ((1u << TPV) - 1u) << (vslot * (GPV * TPV));
	const bool vnew = (bal & gmask) != 0;
	if (writer && (!SP || !kSparse || vnew || a.dense)) {
		// the WHOLE word is written when any slice gained bits: a reader that passed the frontier-bit test
		// reads all of it
		st_slice<SW>(view_next(a) + wo, news);
		if (slice_any<SW>(news)) {
#pragma unroll
			for (int i = 0; i < SW; i++) s.w[i] |= news.w[i];
			st_slice<SW>(a.seen + wo, s);
			alive_note<SW>(salive, a.alive_slots, t * SW, news.w);
		}
	}
	const bool leader = writer && t == 0;
	if (leader && vnew) atomicOr(&view_fnext(a)[n >> 5], 1u << (n & 31));
	if (SP && kDone && leader && a.done && !(rbal & gmask)) atomicOr(&a.done[n >> 5], 1u << (n & 31));
	if (a.npeers && writer && vnew) st_peers<SW>(a, wo, news); // whole word, zero slices included
	// no global counters here: frontier_stats_kernel derives them from fnext afterwards
	if (STATS) {
		stat_add(&a.ctrl->k_listed, (listed && sg == 0 && t == 0) ? 1u : 0u);
		stat_add(&a.ctrl->k_open, (leader) ? 1u : 0u);
		stat_add(&a.ctrl->k_new, (leader && vnew) ? 1u : 0u);
		stat_add(&a.ctrl->k_scanned, c_scan);
		stat_add(&a.ctrl->k_gathered, c_gath);
	}
}

// ---- K4 tiny: in-degree 1..4, straight-line ------------------------------------------------------
// Same result as pull_list_kernel<W, 1, FILTER> for vertices with at most 4 in-edges, without the
// round loop, the prefetch pipeline and the cross-group shuffles: this class is 25 % of an R-MAT's
// vertices and 3 % of its edges, so its cost is instructions and dependent loads per VERTEX.
template <int W, bool FILTER, bool STATS = false>
__global__ void __launch_bounds__(128) pull_tiny_kernel(PullArgs a) {
	// unfiltered (dense) levels: every listed vertex stores its word and no done-bitmap code at all (the code shape
	// of round 1: measured 3.11 vs 3.40 ms on the dense level of an s24 batch); filtered levels: sparse stores + done
	constexpr bool SP = FILTER;
	const uint32_t *const visit_p = view_visit(a);
	const uint32_t *const fcur_p = view_fcur(a);
	using C = LaneCfg<W>;
	constexpr int SW = C::SW, TPV = C::TPV, WPV = C::WPV, G = C::G, U = 4;
	constexpr uint32_t INV = 0xFFFFFFFFu, FULL = 0xffffffffu;
	__shared__ uint32_t salive[16];
	alive_init(salive, a.alive_slots);
	const uint32_t wid = (uint32_t)(((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
	if ((uint64_t)wid * G >= a.nlist) return; // warp-uniform
	const uint32_t lane = threadIdx.x & 31, g = lane / TPV, t = lane % TPV;
	const uint64_t idx = (uint64_t)wid * G + g;
	const bool inlist = idx < a.nlist;
	const uint32_t n = inlist ? a.list[idx] : 0u;
	const bool listed = inlist && !(SP && kDone && a.use_done && done_test(a.done, n)); // a done vertex is skipped altogether
	uint32_t lo = 0, hi = 0;
	if (listed) {
		lo = a.ioff[n];
		hi = a.ioff[n + 1];
	}
	uint32_t vs[U];
	uint32_t c_scan = 0, c_gath = 0; // STATS
#pragma unroll
	for (int u = 0; u < U; u++) {
		vs[u] = lo + u < hi ? a.isrc[lo + u] : INV;
		if (STATS && t == 0 && lo + u < hi) c_scan++;
	}
	const size_t wo = (size_t)n * WPV + (size_t)t * SW;
	Slice<SW> s = slice_zero<SW>(), unseen = slice_zero<SW>();
	if (listed) {
		s = ld_slice<SW>(a.seen + wo);
		const Slice<SW> act = ld_slice<SW>(a.ctrl->active + t * SW);
#pragma unroll
		for (int i = 0; i < SW; i++) unseen.w[i] = ~s.w[i] & act.w[i];
	}
	Slice<SW> acc = slice_zero<SW>();
	if (slice_any<SW>(unseen)) {
		Slice<SW> w[U];
#pragma unroll
		for (int u = 0; u < U; u++) {
			bool f = vs[u] != INV;
			if (FILTER && f) f = fbit_test(fcur_p, vs[u]);
			w[u] = f ? ldg_slice<SW>(visit_p + (size_t)vs[u] * WPV + t * SW) : slice_zero<SW>();
			if (STATS && t == 0 && f) c_gath++;
		}
#pragma unroll
		for (int u = 0; u < U; u++) {
#pragma unroll
			for (int i = 0; i < SW; i++) acc.w[i] |= w[u].w[i];
		}
	}
	Slice<SW> news;
#pragma unroll
	for (int i = 0; i < SW; i++) news.w[i] = acc.w[i] & unseen.w[i];
	const bool mine = slice_any<SW>(news);
	bool rest = false; // some active lane of my slice is still unseen after this level
#pragma unroll
	for (int i = 0; i < SW; i++) rest |= (unseen.w[i] & ~news.w[i]) != 0u;
	const unsigned bal = __ballot_sync(FULL, mine);
	const unsigned rbal = __ballot_sync(FULL, listed && rest);
	const unsigned gmask = ((1u << TPV) - 1u) << (g * TPV);
	const bool vnew = (bal & gmask) != 0;
	if (listed && (!SP || !kSparse || vnew || a.dense)) { // sparse: `next` is all-zero on entry, only words that gained bits are written (whole word)
		st_slice<SW>(view_next(a) + wo, news);
		if (mine) {
#pragma unroll
			for (int i = 0; i < SW; i++) s.w[i] |= news.w[i];
			st_slice<SW>(a.seen + wo, s);
			alive_note<SW>(salive, a.alive_slots, t * SW, news.w);
		}
	}
	if (listed && vnew) {
		if (t == 0) atomicOr(&view_fnext(a)[n >> 5], 1u << (n & 31));
		if (a.npeers) st_peers<SW>(a, wo, news);
	}
	if (SP && kDone && listed && t == 0 && a.done && !(rbal & gmask)) atomicOr(&a.done[n >> 5], 1u << (n & 31));
	if (STATS) {
		stat_add(&a.ctrl->k_listed, (inlist && t == 0) ? 1u : 0u);
		stat_add(&a.ctrl->k_open, (listed && t == 0) ? 1u : 0u);
		stat_add(&a.ctrl->k_new, (listed && t == 0 && vnew) ? 1u : 0u);
		stat_add(&a.ctrl->k_scanned, c_scan);
		stat_add(&a.ctrl->k_gathered, c_gath);
	}
}

// ---- K4 mid (lengths): one warp per vertex, edge-parallel scan ------------------------------------
// Each LANE owns one in-edge of a 32-edge round: one coalesced source load and one frontier-bit
// test per edge (instead of once per thread of the group), a ballot, and then the G groups gather
// only the words whose test passed (source ids handed over by shuffle).  Rounds are pipelined two
// deep like pull_list_kernel.  Early exit once every unseen lane of the vertex is covered.
template <int W, bool FILTER, bool STATS = false>
__global__ void PGQ_WARP_BOUNDS pull_warp_kernel(PullArgs a) {
	// unfiltered (dense) levels: every listed vertex stores its word and no done-bitmap code at all (the code shape
	// of round 1: measured 3.11 vs 3.40 ms on the dense level of an s24 batch); filtered levels: sparse stores + done
	constexpr bool SP = FILTER;
	const uint32_t *const visit_p = view_visit(a);
	const uint32_t *const fcur_p = view_fcur(a);
	using C = LaneCfg<W>;
	constexpr int SW = C::SW, TPV = C::TPV, WPV = C::WPV, G = C::G;
	constexpr int STEPS = 32 / G;
	constexpr uint32_t INV = 0xFFFFFFFFu, FULL = 0xffffffffu;
	__shared__ uint32_t salive[16];
	alive_init(salive, a.alive_slots);
	const uint32_t wid = (uint32_t)(((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
	if (wid >= a.nlist) return; // warp-uniform
	const uint32_t lane = threadIdx.x & 31, g = lane / TPV, t = lane % TPV;
	const uint32_t n = a.list[wid];
	if (SP && kDone && a.use_done && done_test(a.done, n)) { // warp-uniform: the vertex is skipped altogether
		if (STATS && lane == 0) atomicAdd(&a.ctrl->k_listed, 1ull);
		return;
	}
	const uint32_t lo = a.ioff[n], hi = a.ioff[n + 1];
	uint32_t v0 = lo + lane < hi ? a.isrc[lo + lane] : INV;
	uint32_t v1 = lo + 32 + lane < hi ? a.isrc[lo + 32 + lane] : INV;
	uint32_t c_scan = 0, c_gath = 0; // STATS
	if (STATS) c_scan = (v0 != INV ? 1u : 0u) + (v1 != INV ? 1u : 0u);
	const size_t wo = (size_t)n * WPV + (size_t)t * SW;
	Slice<SW> s = ld_slice<SW>(a.seen + wo);
	const Slice<SW> act = ld_slice<SW>(a.ctrl->active + t * SW);
	Slice<SW> unseen;
#pragma unroll
	for (int i = 0; i < SW; i++) unseen.w[i] = ~s.w[i] & act.w[i];
	Slice<SW> acc = slice_zero<SW>();
	if (__any_sync(FULL, slice_any<SW>(unseen))) {
		bool p0 = v0 != INV;
		if (FILTER && p0) p0 = fbit_test(fcur_p, v0);
		for (uint32_t base = lo; base < hi; base += 32) {
			const unsigned mask = __ballot_sync(FULL, p0);
			Slice<SW> w[STEPS];
#pragma unroll
			for (int u = 0; u < STEPS; u++) {
				const int sl = u * G + g;
				const uint32_t src = STEPS == 1 ? v0 : __shfl_sync(FULL, v0, sl);
				w[u] = ((mask >> sl) & 1u) ? ldg_slice<SW>(visit_p + (size_t)src * WPV + t * SW) : slice_zero<SW>();
				if (STATS && t == 0 && ((mask >> sl) & 1u)) c_gath++;
			}
			const uint32_t j2 = base + 64 + lane;
			const uint32_t v2 = j2 < hi ? a.isrc[j2] : INV;
			if (STATS && j2 < hi) c_scan++;
			bool p1 = v1 != INV;
			if (FILTER && p1) p1 = fbit_test(fcur_p, v1);
#pragma unroll
			for (int u = 0; u < STEPS; u++) {
#pragma unroll
				for (int i = 0; i < SW; i++) acc.w[i] |= w[u].w[i];
			}
			if (mask) {
				const Slice<SW> tot = warp_or_groups<SW, TPV>(acc);
				if (__all_sync(FULL, slice_covers<SW>(tot, unseen))) break; // every unseen lane reached
			}
			v0 = v1;
			p0 = p1;
			v1 = v2;
		}
	}
	const Slice<SW> tot = warp_or_groups<SW, TPV>(acc);
	Slice<SW> news;
#pragma unroll
	for (int i = 0; i < SW; i++) news.w[i] = tot.w[i] & unseen.w[i];
	const bool wany = __any_sync(FULL, slice_any<SW>(news));
	bool rest = false; // some active lane of my slice is still unseen after this level
#pragma unroll
	for (int i = 0; i < SW; i++) rest |= (unseen.w[i] & ~news.w[i]) != 0u;
	const bool wrest = __any_sync(FULL, rest);
	if (g == 0 && (!SP || !kSparse || wany || a.dense)) { // sparse: `next` is all-zero on entry, only words that gained bits are written (whole word)
		st_slice<SW>(view_next(a) + wo, news);
		if (slice_any<SW>(news)) {
#pragma unroll
			for (int i = 0; i < SW; i++) s.w[i] |= news.w[i];
			st_slice<SW>(a.seen + wo, s);
			alive_note<SW>(salive, a.alive_slots, t * SW, news.w);
		}
	}
	if (wany && lane == 0) atomicOr(&view_fnext(a)[n >> 5], 1u << (n & 31));
	if (SP && kDone && !wrest && lane == 0 && a.done) atomicOr(&a.done[n >> 5], 1u << (n & 31));
	if (a.npeers && wany && g == 0) st_peers<SW>(a, wo, news);
	if (STATS) {
		if (lane == 0) {
			atomicAdd(&a.ctrl->k_listed, 1ull);
			atomicAdd(&a.ctrl->k_open, 1ull);
			if (wany) atomicAdd(&a.ctrl->k_new, 1ull);
		}
		stat_add(&a.ctrl->k_scanned, c_scan);
		stat_add(&a.ctrl->k_gathered, c_gath);
	}
}

// ---- K4 small class, filtered levels: SEG lanes per vertex, edge-parallel frontier test ----------------
// On a filtered level most in-edges fail the frontier-bit test and nothing is gathered for them, so the
// cost of pull_list_kernel<W, GPV, true> is the test itself, which every one of the TPV threads of a
// word repeats for every edge (issue-bound: profiles/r01_ncu_iter2_filtered.txt).  Here the SEG lanes that
// serve a vertex each own ONE in-edge of a SEG-edge round (one source load + one bit test per edge), a
// ballot collects the hits of every segment, and the segment's SEG/TPV word groups gather only those,
// source ids handed over by shuffle.  pull_warp_kernel is the SEG = 32 case of the same idea; this one
// packs 32/SEG vertices into a warp for in-degrees far below 32.  Rounds are pipelined two deep.
template <int W, int SEG, bool STATS = false>
__global__ void PGQ_SEG_BOUNDS pull_seg_kernel(PullArgs a) {
	const uint32_t *const visit_p = view_visit(a);
	const uint32_t *const fcur_p = view_fcur(a);
	using C = LaneCfg<W>;
	constexpr int SW = C::SW, TPV = C::TPV, WPV = C::WPV;
	constexpr int VPW = 32 / SEG;  // vertices per warp
	constexpr int GPS = SEG / TPV; // word groups per segment
	static_assert(SEG >= TPV && SEG <= 32 && (SEG & (SEG - 1)) == 0, "segment shape");
	constexpr uint32_t INV = 0xFFFFFFFFu, FULL = 0xffffffffu;
	constexpr uint32_t SEGMASK = SEG == 32 ? 0xffffffffu : ((1u << SEG) - 1u);
	__shared__ uint32_t salive[16];
	alive_init(salive, a.alive_slots);
	const uint32_t wid = (uint32_t)(((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
	if ((uint64_t)wid * VPW >= a.nlist) return; // warp-uniform
	const uint32_t lane = threadIdx.x & 31, seg = lane / SEG, sl = lane % SEG;
	const uint32_t g = sl / TPV, t = sl % TPV; // word group inside the segment, slice inside the word
	const uint64_t idx = (uint64_t)wid * VPW + seg;
	const bool listed = idx < a.nlist;
	const uint32_t n = listed ? a.list[idx] : 0u;
	const bool valid = listed && !(kDone && a.use_done && done_test(a.done, n)); // a done vertex is skipped altogether
	uint32_t lo = 0, hi = 0;
	if (valid) {
		lo = a.ioff[n];
		hi = a.ioff[n + 1];
	}
	uint32_t v0 = lo + sl < hi ? a.isrc[lo + sl] : INV;
	uint32_t v1 = lo + SEG + sl < hi ? a.isrc[lo + SEG + sl] : INV;
	uint32_t c_scan = 0, c_gath = 0; // STATS
	if (STATS) c_scan = (v0 != INV ? 1u : 0u) + (v1 != INV ? 1u : 0u);
	const size_t wo = (size_t)n * WPV + (size_t)t * SW;
	Slice<SW> s = slice_zero<SW>(), unseen = slice_zero<SW>();
	if (valid) {
		s = ld_slice<SW>(a.seen + wo);
		const Slice<SW> act = ld_slice<SW>(a.ctrl->active + t * SW);
#pragma unroll
		for (int i = 0; i < SW; i++) unseen.w[i] = ~s.w[i] & act.w[i];
	}
	// a vertex is scanned while ANY of its slices still misses lanes (the edge tests are shared by the
	// whole segment); a covered slice just stops gathering
	bool mine_open = slice_any<SW>(unseen);
	bool vopen = (__ballot_sync(FULL, mine_open) >> (seg * SEG)) & SEGMASK;
	Slice<SW> acc = slice_zero<SW>();
	bool p0 = vopen && v0 != INV && fbit_test(fcur_p, v0);
	uint32_t base = lo;
	while (__any_sync(FULL, vopen && base < hi)) {
		const unsigned bal = __ballot_sync(FULL, p0);
		// sources of the round after next, frontier bits of the next round: in flight during the gathers
		const uint32_t j2 = base + 2 * SEG + sl;
		const uint32_t v2 = (vopen && j2 < hi) ? a.isrc[j2] : INV;
		if (STATS && vopen && j2 < hi) c_scan++;
		const bool p1 = vopen && v1 != INV && fbit_test(fcur_p, v1);
		unsigned m = (bal >> (seg * SEG)) & SEGMASK; // hits of my vertex in this round
		while (__any_sync(FULL, m != 0u)) {
			unsigned mm = m; // my group takes the g-th lowest hit
#pragma unroll
			for (int k = 0; k < GPS - 1; k++)
				if (k < (int)g) mm &= mm - 1u;
			const uint32_t src = __shfl_sync(FULL, v0, (int)(seg * SEG) + (mm ? __ffs((int)mm) - 1 : 0));
			if (mm && mine_open) {
				if (STATS) c_gath++; // per 16-byte slice here (a covered slice stops gathering): / TPV below
				const Slice<SW> w = ldg_slice<SW>(visit_p + (size_t)src * WPV + t * SW);
#pragma unroll
				for (int i = 0; i < SW; i++) acc.w[i] |= w.w[i];
			}
#pragma unroll
			for (int k = 0; k < GPS; k++) m &= m - 1u;
		}
		if (bal) { // something was gathered somewhere in the warp: early-exit test per vertex
			Slice<SW> tot = acc;
#pragma unroll
			for (int d = TPV; d < SEG; d <<= 1) {
#pragma unroll
				for (int i = 0; i < SW; i++) tot.w[i] |= __shfl_xor_sync(FULL, tot.w[i], d);
			}
			mine_open = mine_open && !slice_covers<SW>(tot, unseen);
			vopen = (__ballot_sync(FULL, mine_open) >> (seg * SEG)) & SEGMASK;
		}
		v0 = v1;
		v1 = v2;
		p0 = p1 && vopen;
		base += SEG;
	}
	Slice<SW> tot = acc;
#pragma unroll
	for (int d = TPV; d < SEG; d <<= 1) {
#pragma unroll
		for (int i = 0; i < SW; i++) tot.w[i] |= __shfl_xor_sync(FULL, tot.w[i], d);
	}
	Slice<SW> news;
#pragma unroll
	for (int i = 0; i < SW; i++) news.w[i] = tot.w[i] & unseen.w[i];
	const bool writer = valid && g == 0;
	bool rest = false; // some active lane of my slice is still unseen after this level
#pragma unroll
	for (int i = 0; i < SW; i++) rest |= (unseen.w[i] & ~news.w[i]) != 0u;
	const unsigned nb = __ballot_sync(FULL, writer && slice_any<SW>(news));
	const unsigned rb = __ballot_sync(FULL, writer && rest);
	const bool vnew = ((nb >> (seg * SEG)) & SEGMASK) != 0u;
	const bool vrest = ((rb >> (seg * SEG)) & SEGMASK) != 0u;
	if (writer && (!kSparse || vnew || a.dense)) { // sparse: `next` is all-zero on entry, only words that gained bits are written (whole word)
		st_slice<SW>(view_next(a) + wo, news);
		if (slice_any<SW>(news)) {
#pragma unroll
			for (int i = 0; i < SW; i++) s.w[i] |= news.w[i];
			st_slice<SW>(a.seen + wo, s);
			alive_note<SW>(salive, a.alive_slots, t * SW, news.w);
		}
	}
	if (writer && t == 0 && vnew) atomicOr(&view_fnext(a)[n >> 5], 1u << (n & 31));
	if (kDone && writer && t == 0 && a.done && !vrest) atomicOr(&a.done[n >> 5], 1u << (n & 31));
	if (a.npeers && writer && vnew) st_peers<SW>(a, wo, news); // whole word, zero slices included
	if (STATS) {
		stat_add(&a.ctrl->k_listed, (listed && sl == 0) ? 1u : 0u);
		stat_add(&a.ctrl->k_open, (valid && sl == 0) ? 1u : 0u);
		stat_add(&a.ctrl->k_new, (valid && sl == 0 && vnew) ? 1u : 0u);
		stat_add(&a.ctrl->k_scanned, c_scan);
		const uint32_t gsl = __reduce_add_sync(FULL, c_gath); // 16-byte slices -> words (rounded up)
		if (lane == 0 && gsl) atomicAdd(&a.ctrl->k_gathered, (unsigned long long)((gsl + TPV - 1) / TPV));
	}
}

// ---- K4 hubs: one block per chunk of a hub's in-list, partial ORs merged with atomicOr ---------
template <int W, bool STATS = false>
__global__ void __launch_bounds__(256) pull_hub_kernel(PullArgs a) {
	const uint32_t *const visit_p = view_visit(a);
	const uint32_t *const fcur_p = view_fcur(a);
	using C = LaneCfg<W>;
	constexpr int SW = C::SW, TPV = C::TPV, WPV = C::WPV, G = C::G, U = 4;
	const HubChunk hc = a.chunks[blockIdx.x];
	const uint32_t n = a.list[hc.hub];
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
	const uint32_t g = lane / TPV, t = lane % TPV;
	const size_t wo = (size_t)n * WPV + (size_t)t * SW;
	Slice<SW> s = ld_slice<SW>(a.seen + wo);
	Slice<SW> act = ld_slice<SW>(a.ctrl->active + t * SW);
	Slice<SW> unseen;
#pragma unroll
	for (int i = 0; i < SW; i++) unseen.w[i] = ~s.w[i] & act.w[i];
	if (STATS && threadIdx.x == 0) { // every chunk block reads the hub's `seen` word
		atomicAdd(&a.ctrl->k_listed, 1ull);
		atomicAdd(&a.ctrl->k_open, 1ull);
	}
	if (!__any_sync(0xffffffffu, slice_any<SW>(unseen))) return; // same decision in every warp
	Slice<SW> acc = slice_zero<SW>();
	Slice<SW> tot = slice_zero<SW>();
	uint32_t c_scan = 0, c_gath = 0; // STATS
	for (uint32_t j0 = hc.begin + warp * (G * U); j0 < hc.end; j0 += nwarps * G * U) {
		uint32_t vs[U];
#pragma unroll
		for (int u = 0; u < U; u++) {
			uint32_t j = j0 + u * G + g;
			vs[u] = j < hc.end ? a.isrc[j] : 0xFFFFFFFFu;
			if (STATS && t == 0 && j < hc.end) c_scan++;
		}
#pragma unroll
		for (int u = 0; u < U; u++)
			if (vs[u] != 0xFFFFFFFFu && !fbit_test(fcur_p, vs[u])) vs[u] = 0xFFFFFFFFu;
#pragma unroll
		for (int u = 0; u < U; u++) {
			if (vs[u] != 0xFFFFFFFFu) {
				if (STATS && t == 0) c_gath++;
				Slice<SW> w = ldg_slice<SW>(visit_p + (size_t)vs[u] * WPV + t * SW);
#pragma unroll
				for (int i = 0; i < SW; i++) acc.w[i] |= w.w[i];
			}
		}
		tot = warp_or_groups<SW, TPV>(acc);
		if (__all_sync(0xffffffffu, slice_covers<SW>(tot, unseen))) break;
	}
	tot = warp_or_groups<SW, TPV>(acc);
	if (g == 0) {
		uint32_t *dst = a.hubacc + (size_t)hc.hub * WPV + (size_t)t * SW;
#pragma unroll
		for (int i = 0; i < SW; i++) {
			uint32_t x = tot.w[i] & unseen.w[i];
			if (x) atomicOr(dst + i, x);
		}
	}
	if (STATS) {
		stat_add(&a.ctrl->k_scanned, c_scan);
		stat_add(&a.ctrl->k_gathered, c_gath);
	}
}

// one TPV-thread group per hub: next = hubacc & ~seen & active; seen |= next; hubacc = 0
template <int W>
__global__ void __launch_bounds__(256) hub_finalize_kernel(PullArgs a) {
	const uint32_t *const visit_p = view_visit(a);
	const uint32_t *const fcur_p = view_fcur(a);
	using C = LaneCfg<W>;
	constexpr int SW = C::SW, TPV = C::TPV, WPV = C::WPV;
	__shared__ uint32_t salive[16];
	alive_init(salive, a.alive_slots);
	const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t h = gid / TPV, t = gid % TPV;
	Slice<SW> news = slice_zero<SW>();
	if (h < a.nlist) {
		const uint32_t n = a.list[h];
		const size_t wo = (size_t)n * WPV + (size_t)t * SW;
		uint32_t *ha = a.hubacc + (size_t)h * WPV + (size_t)t * SW;
		Slice<SW> acc = ld_slice<SW>(ha);
		st_slice<SW>(ha, slice_zero<SW>());
		Slice<SW> s = ld_slice<SW>(a.seen + wo);
		Slice<SW> act = ld_slice<SW>(a.ctrl->active + t * SW);
#pragma unroll
		for (int i = 0; i < SW; i++) news.w[i] = acc.w[i] & ~s.w[i] & act.w[i];
	}
	const bool anyn = slice_any<SW>(news);
	const bool gnew = group_any<TPV>(anyn);
	if (h < a.nlist && (!kSparse || gnew || a.dense)) { // sparse: `next` is all-zero on entry, only words that gained bits are written (whole word)
		const uint32_t n = a.list[h];
		const size_t wo = (size_t)n * WPV + (size_t)t * SW;
		st_slice<SW>(view_next(a) + wo, news);
		if (anyn) {
			Slice<SW> s = ld_slice<SW>(a.seen + wo);
#pragma unroll
			for (int i = 0; i < SW; i++) s.w[i] |= news.w[i];
			st_slice<SW>(a.seen + wo, s);
		}
	}
	const bool leader = gnew && t == 0 && h < a.nlist;
	const uint32_t hn = h < a.nlist ? a.list[h] : 0u;
	if (leader) atomicOr(&view_fnext(a)[hn >> 5], 1u << (hn & 31));
	if (anyn) alive_note<SW>(salive, a.alive_slots, t * SW, news.w);
	if (a.npeers && gnew && h < a.nlist) st_peers<SW>(a, (size_t)hn * WPV + (size_t)t * SW, news);
}

// After a pull level: vertices in the new frontier and the sum of their out-degrees (the inputs of
// the direction switch), plus the `change` flag, from the frontier bitmap.  One atomic per block.
static __global__ void __launch_bounds__(256) frontier_stats_kernel(uint32_t nwords, uint32_t V,
                                                             const int64_t *__restrict__ off, BfsCtrl *ctrl) {
	constexpr uint32_t FULL = 0xffffffffu;
	const uint32_t *const fbit = ctrl->L.nxt_fbit; // the frontier this level produced
	__shared__ unsigned long long sdeg[8];
	__shared__ uint32_t scnt[8];
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint32_t m = i < nwords ? fbit[i] : 0u;
	uint32_t cnt = __popc(m);
	unsigned long long deg = 0;
	while (m) {
		const int k = __ffs(m) - 1;
		// runs of set bits: one subtraction per run instead of per vertex
		const uint32_t inv = ~(m >> k);
		const uint32_t run = inv ? (uint32_t)(__ffs(inv) - 1) : 32u; // consecutive ones starting at k (1..32)
		const uint32_t n0 = i * 32 + k;
		uint32_t n1 = n0 + run;
		if (n1 > V) n1 = V;
		if (n0 < V) deg += (unsigned long long)(off[n1] - off[n0]);
		if (k + run >= 32) break;
		m &= ~((1u << (k + run)) - 1u);
	}
#pragma unroll
	for (int d = 16; d > 0; d >>= 1) {
		deg += __shfl_xor_sync(FULL, deg, d);
		cnt += __shfl_xor_sync(FULL, cnt, d);
	}
	if (lane == 0) {
		sdeg[warp] = deg;
		scnt[warp] = cnt;
	}
	__syncthreads();
	if (threadIdx.x == 0) {
#pragma unroll
		for (int w = 1; w < 8; w++) {
			deg += sdeg[w];
			cnt += scnt[w];
		}
		if (cnt) {
			atomicAdd(&ctrl->next_edges, deg);
			atomicAdd(&ctrl->next_verts, cnt);
			ctrl->change = 1u;
		}
	}
}

// ---- batch init / K6 destination check / stats / K8 path reconstruction -----------------------

// Appends vertex n to the push queue being built: one work item per PUSH_CHUNK out-edges (at
// least one, so zero-out-degree vertices still get their frontier word cleaned up).
__device__ __forceinline__ void enqueue_vertex(uint32_t n, const int64_t *off, uint2 *q, BfsCtrl *ctrl) {
	const unsigned long long deg = (unsigned long long)(off[n + 1] - off[n]);
	uint32_t k = (uint32_t)((deg + PUSH_CHUNK - 1) / PUSH_CHUNK);
	if (k == 0) k = 1;
	const uint32_t base = atomicAdd(&ctrl->next_items, k);
	for (uint32_t c = 0; c < k; c++) q[base + c] = make_uint2(n, c);
	atomicAdd(&ctrl->next_edges, deg);
	atomicAdd(&ctrl->next_verts, 1u);
}

// Direction of the level about to run from the statistics of its input frontier (out_edges, out_verts)
// -- the host loop of round 1, evaluated on the device by the init and check kernels.
__device__ __forceinline__ uint32_t loop_pick_mode(const BfsCtrl *ctrl, const LoopState &L) {
	const unsigned long long oe = ctrl->out_edges, ov = ctrl->out_verts;
	// top-down while the frontier's out-edges are few
	const bool push = L.direction == 0 && oe * L.push_beta <= L.E && ov * L.push_vbeta <= L.V + (unsigned long long)L.push_vbeta * 1024ull;
	if (push) return MODE_PUSH;
	// the bitmap test is dead weight once the frontier holds most of the edges
	const bool filter = oe * L.filter_gamma <= L.E;
	if (!filter) return MODE_PULL;
	// "early": fewer vertex discoveries so far than vertices that can be discovered at all
	return L.cum_verts < L.n_listed ? MODE_PULL_F_EARLY : MODE_PULL_F_LATE;
}

// rotating views of the level L.iter from cur / fb / q (and the level-buffer table of shortestpath)
__device__ __forceinline__ void loop_set_views(LoopState &L) {
	if (L.hist) {
		L.cur_visit = L.hist[L.iter - 1];
		L.cur_next = L.hist[L.iter];
	} else {
		L.cur_visit = L.visit[L.cur];
		L.cur_next = L.visit[L.cur ^ 1];
	}
	L.cur_fbit = L.fbit[L.fb[0]];
	L.nxt_fbit = L.fbit[L.fb[1]];
	L.old_fbit = L.fbit[L.fb[2]];
	L.cur_queue = L.queue[L.q];
	L.nxt_queue = L.queue[L.q ^ 1];
}

// Per-batch parameters of the level loop (host -> init kernel, by value).
struct LoopInit {
	uint32_t *visit[2];
	uint32_t *fbit[3];
	uint2 *queue[2];
	uint32_t *const *hist;
	uint32_t hist_cap;
	uint32_t level_cap;
	uint32_t direction, push_beta, push_vbeta, filter_gamma;
	unsigned long long E, V, n_listed;
	const uint32_t *lane_dst;
	uint32_t nk[MODE_COUNT];
};

// One block of W threads.  Sets the source bits (R4: visit only, seen untouched), the active mask,
// found = -1, the level-1 push queue (the distinct sources) and the loop state of level 1.
// lane_src == 0xFFFFFFFF marks an unused lane slot.  The frontier bitmaps must be zero on entry.
static __global__ void bfs_init_kernel(const uint32_t *__restrict__ lane_src, int W, int32_t *__restrict__ found,
                                BfsCtrl *ctrl, const int64_t *__restrict__ off, LoopInit li) {
	__shared__ uint32_t cnt;
	const int lane = threadIdx.x;
	uint32_t *const visit = li.hist ? li.hist[0] : li.visit[0];
	uint2 *const queue = li.queue[0];
	uint32_t *const inq = li.fbit[0];
	if (lane == 0) {
		cnt = 0;
		ctrl->change = 0; ctrl->change_out = 0;
		ctrl->edges_lane = 0; ctrl->edges_unique = 0; ctrl->verts_active = 0;
		ctrl->next_edges = 0; ctrl->next_verts = 0; ctrl->next_items = 0;
		ctrl->k_listed = 0; ctrl->k_open = 0; ctrl->k_new = 0; ctrl->k_scanned = 0; ctrl->k_gathered = 0;
		ctrl->kp_items = 0; ctrl->kp_edges = 0; ctrl->kp_seen = 0; ctrl->kp_next = 0;
	}
	__syncthreads();
	const uint32_t s = lane_src[lane];
	const bool used = s != 0xFFFFFFFFu;
	found[lane] = -1;
	if (used) {
		atomicOr(&visit[(size_t)s * (W / 32) + (lane >> 5)], 1u << (lane & 31));
		const uint32_t bit = 1u << (s & 31);
		if (!(atomicOr(&inq[s >> 5], bit) & bit)) enqueue_vertex(s, off, queue, ctrl);
	}
	const unsigned m = __ballot_sync(0xffffffffu, used);
	if ((lane & 31) == 0) {
		ctrl->active[lane >> 5] = m;
		ctrl->alive[lane >> 5] = 0;
		atomicAdd(&cnt, (uint32_t)__popc(m));
	}
	__syncthreads();
	if (lane == 0) {
		ctrl->n_active = cnt;
		__threadfence();
		snapshot_frontier(ctrl);
		LoopState &L = ctrl->L;
		for (int i = 0; i < 2; i++) { L.visit[i] = li.visit[i]; L.queue[i] = li.queue[i]; }
		for (int i = 0; i < 3; i++) { L.fbit[i] = li.fbit[i]; L.fb[i] = (uint32_t)i; }
		L.hist = li.hist;
		L.hist_cap = li.hist_cap;
		L.cur = 0; L.q = 0; L.have_queue = 1;
		L.iter = 1;
		L.level_cap = li.level_cap;
		L.cum_verts = ctrl->out_verts;
		L.direction = li.direction; L.push_beta = li.push_beta; L.push_vbeta = li.push_vbeta;
		L.filter_gamma = li.filter_gamma;
		L.E = li.E; L.V = li.V; L.n_listed = li.n_listed;
		L.levels_run = 0; L.push_levels = 0;
		L.stop_reason = 0;
		L.lane_dst = li.lane_dst;
		L.found = found;
		for (int i = 0; i < (int)MODE_COUNT; i++) L.nk[i] = li.nk[i];
		loop_set_views(L);
		if (cnt == 0 || L.level_cap == 0) {
			L.mode = MODE_STOP;
			L.stop_reason = 1;
		} else if (L.hist && L.hist_cap < 2) {
			L.mode = MODE_STOP;
			L.stop_reason = 2;
		} else {
			L.mode = loop_pick_mode(ctrl, L);
		}
	}
}

// After a level: lanes whose destination has been seen get found=iter and leave the active mask;
// the frontier statistics of the level are snapshotted for the host's direction decision.
static __global__ void bfs_check_kernel(const uint32_t *__restrict__ seen, int W, BfsCtrl *ctrl,
                                 uint32_t *__restrict__ alive_slots, int use_handles,
                                 cudaGraphConditionalHandle h_while, cudaGraphConditionalHandle h_switch) {
	const int iter = (int)ctrl->L.iter; // the level that just ran
	const uint32_t *const lane_dst = ctrl->L.lane_dst;
	int32_t *const found = ctrl->L.found;
	__shared__ uint32_t cnt;
	__shared__ uint32_t salive[16];
	const int lane = threadIdx.x;
	if (lane == 0) cnt = 0;
	if (lane < 16) salive[lane] = 0;
	__syncthreads();
	// fold (and re-zero) the per-slot alive masks of the level that just ran
	for (uint32_t idx = lane; idx < ALIVE_SLOTS * 16; idx += blockDim.x) {
		const uint32_t v = alive_slots[idx];
		if (v) {
			alive_slots[idx] = 0;
			atomicOr(&salive[idx & 15], v);
		}
	}
	__syncthreads();
	const uint32_t word = ctrl->active[lane >> 5];
	const bool act = (word >> (lane & 31)) & 1u;
	bool hit = false;
	if (act) {
		const uint32_t d = lane_dst[lane];
		hit = (seen[(size_t)d * (W / 32) + (lane >> 5)] >> (lane & 31)) & 1u;
		if (hit) found[lane] = iter;
	}
	const unsigned hm = __ballot_sync(0xffffffffu, hit);
	__syncthreads(); // all reads of active[] done before any warp rewrites its word
	if ((lane & 31) == 0) {
		const uint32_t nw = word & ~hm & salive[lane >> 5]; // found or exhausted lanes retire
		ctrl->alive[lane >> 5] = salive[lane >> 5];
		ctrl->active[lane >> 5] = nw;
		atomicAdd(&cnt, (uint32_t)__popc(nw));
	}
	__syncthreads();
	if (lane == 0) {
		ctrl->n_active = cnt;
		const uint32_t changed = ctrl->change;
		ctrl->change_out = changed;
		ctrl->change = 0;
		snapshot_frontier(ctrl);
		// ---- the level loop's bookkeeping: the output of the level becomes the current frontier ----
		LoopState &L = ctrl->L;
		L.levels_run++;
		ctrl->tot_levels++;
		ctrl->tot_launches += L.nk[L.mode] + 1u; // + this kernel
		if (L.mode == MODE_PUSH) { // the queue it built describes the new frontier
			L.push_levels++;
			ctrl->tot_push_levels++;
			L.q ^= 1u;
			L.have_queue = 1;
		} else {
			L.have_queue = 0;
		}
		L.cur ^= 1u;
		const uint32_t t = L.fb[0];
		L.fb[0] = L.fb[1];
		L.fb[1] = L.fb[2];
		L.fb[2] = t;
		L.cum_verts += ctrl->out_verts;
		L.iter = (uint32_t)iter + 1u;
		if (!changed || cnt == 0 || (uint32_t)iter >= L.level_cap) {
			L.mode = MODE_STOP;
			L.stop_reason = 1;
		} else if (L.hist && L.iter >= L.hist_cap) { // no buffer for the next level's output: the host extends the table
			L.mode = MODE_STOP;
			L.stop_reason = 2;
		} else {
			L.mode = loop_pick_mode(ctrl, L);
		}
		if (L.mode != MODE_STOP) loop_set_views(L); // a stopped loop keeps the views of its last level
		if (use_handles) {
			cudaGraphSetConditional(h_while, L.mode != MODE_STOP ? 1u : 0u);
			cudaGraphSetConditional(h_switch, L.mode - 1u); // body index; MODE_STOP selects none
		}
	}
}

// First node of the loop graph: the conditional handles from the mode the init kernel (or the check
// kernel of an interrupted loop) left behind.
static __global__ void loop_begin_kernel(BfsCtrl *ctrl, cudaGraphConditionalHandle h_while,
                                  cudaGraphConditionalHandle h_switch) {
	cudaGraphSetConditional(h_while, ctrl->L.mode != MODE_STOP ? 1u : 0u);
	cudaGraphSetConditional(h_switch, ctrl->L.mode - 1u);
}

// The host extended the level-buffer table of an interrupted shortestpath batch: resume.
static __global__ void loop_resume_kernel(BfsCtrl *ctrl, uint32_t *const *hist, uint32_t hist_cap) {
	LoopState &L = ctrl->L;
	L.hist = hist;
	L.hist_cap = hist_cap;
	if (L.stop_reason == 2 && L.iter < hist_cap) {
		L.stop_reason = 0;
		L.mode = loop_pick_mode(ctrl, L);
		loop_set_views(L);
	}
}

// ---- K4 top-down (push) for sparse frontiers ---------------------------------------------------
// One warp per work item (frontier vertex v, chunk of its out-edges).  For every out-edge v->n and
// every non-empty 32-lane word of visit[v]:  old = atomicOr(seen[n].w, bits); the bits that were
// not in `old` are exactly next[n] = visit[v] & ~seen[n] of the reference (R4/R5), so they are
// ORed into next[n] and n joins the next queue the first time that happens this level.
struct PushArgs {
	const int64_t *off;
	const uint32_t *tgt;
	uint32_t *seen;
	BfsCtrl *ctrl; // frontier words, work queues and the next frontier's bitmap (= dedupe bitmap of the next
	               // queue) are the level loop's rotating views in ctrl->L; item count = ctrl->out_items
	uint32_t *alive_slots;
};

template <int W, bool STATS = false>
__global__ void __launch_bounds__(256) push_kernel(PushArgs a) {
	constexpr int WPV = W / 32;
	uint32_t c_scan = 0, c_gath = 0, c_items = 0, c_new = 0; // STATS
	__shared__ uint32_t salive[16];
	alive_init(salive, a.alive_slots);
	const LoopState &L = a.ctrl->L;
	const uint32_t *const visit_p = L.cur_visit;
	uint32_t *const next_p = L.cur_next;
	uint32_t *const fnext_p = L.nxt_fbit;
	const uint2 *const queue_p = L.cur_queue;
	uint2 *const queue_next_p = L.nxt_queue;
	const uint32_t nitems = a.ctrl->out_items;
	const uint32_t lane = threadIdx.x & 31;
	const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
	bool anynew = false;
	for (uint32_t it = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; it < nitems; it += nwarps) {
		const uint2 item = queue_p[it];
		const uint32_t v = item.x;
		uint32_t word = 0;
		if (lane < WPV) word = visit_p[(size_t)v * WPV + lane] & a.ctrl->active[lane];
		const unsigned nz = __ballot_sync(0xffffffffu, word != 0);
		if (STATS && lane == 0) c_items++;
		if (nz == 0) continue;
		const int64_t e0 = a.off[v] + (int64_t)item.y * PUSH_CHUNK;
		int64_t e1 = a.off[v + 1];
		if (e1 > e0 + PUSH_CHUNK) e1 = e0 + PUSH_CHUNK;
		for (int64_t j0 = e0; j0 < e1; j0 += 32) {
			const int64_t j = j0 + lane;
			const bool have = j < e1;
			const uint32_t n = have ? a.tgt[j] : 0u;
			if (STATS && have) c_scan++;
			bool mynew = false;
			unsigned rem = nz;
			while (rem) { // warp-uniform loop over the non-empty words of visit[v]
				const int w = __ffs(rem) - 1;
				rem &= rem - 1;
				const uint32_t bits = __shfl_sync(0xffffffffu, word, w);
				if (have) {
					const uint32_t old = atomicOr(&a.seen[(size_t)n * WPV + w], bits);
					const uint32_t nb = bits & ~old;
					if (STATS) c_gath++;
					if (nb) {
						if (STATS) c_new++;
						atomicOr(&next_p[(size_t)n * WPV + w], nb);
						alive_note<1>(salive, a.alive_slots, (uint32_t)w, &nb);
						mynew = true;
					}
				}
			}
			bool enq = false; // first discovery of n on this level: it joins the next queue
			if (mynew) {
				anynew = true;
				const uint32_t bit = 1u << (n & 31);
				enq = !(atomicOr(&fnext_p[n >> 5], bit) & bit);
			}
			// Warp-aggregated enqueue: the three queue counters are single addresses, and a level-1 push
			// discovers tens of thousands of vertices -- one atomic per warp step instead of three per vertex.
			const unsigned em = __ballot_sync(0xffffffffu, enq);
			if (em) {
				unsigned long long deg = 0;
				uint32_t k = 0;
				if (enq) {
					deg = (unsigned long long)(a.off[n + 1] - a.off[n]);
					k = (uint32_t)((deg + PUSH_CHUNK - 1) / PUSH_CHUNK);
					if (k == 0) k = 1;
				}
				uint32_t inc = k;
#pragma unroll
				for (int o = 1; o < 32; o <<= 1) {
					const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
					if ((int)lane >= o) inc += y;
				}
				const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
				unsigned long long dsum = deg;
#pragma unroll
				for (int o = 16; o > 0; o >>= 1) dsum += __shfl_xor_sync(0xffffffffu, dsum, o);
				uint32_t base = 0;
				if (lane == 0) {
					base = atomicAdd(&a.ctrl->next_items, total);
					atomicAdd(&a.ctrl->next_edges, dsum);
					atomicAdd(&a.ctrl->next_verts, (uint32_t)__popc(em));
				}
				base = __shfl_sync(0xffffffffu, base, 0) + inc - k;
				for (uint32_t ch = 0; ch < k; ch++) queue_next_p[base + ch] = make_uint2(n, ch);
			}
		}
	}
	if (__any_sync(0xffffffffu, anynew) && lane == 0) a.ctrl->change = 1u;
	if (STATS) {
		stat_add(&a.ctrl->kp_items, c_items);
		stat_add(&a.ctrl->kp_edges, c_scan);
		stat_add(&a.ctrl->kp_seen, c_gath);
		stat_add(&a.ctrl->kp_next, c_new);
	}
}

// Before a push level: the `next` buffer still holds the frontier of two levels ago, described by
// its bitmap.  Zero exactly those words, then the bitmap (cost ~ that frontier, not V).
template <int W>
__device__ __forceinline__ void clear_by_bitmap_body(uint32_t *__restrict__ fbit, uint32_t nwords,
                                                     uint32_t *__restrict__ visit) {
	// One warp per 32 bitmap words (1024 vertices).  The words a stale frontier leaves behind are 16-64 bytes
	// each and often neighbours: the warp walks the set bits of one bitmap word at a time and zeroes 32 / Q
	// vertices per store instruction, Q = 16-byte quarters per vertex word handled by Q adjacent lanes
	// (one thread per bitmap word writing its up to 32 x 64 bytes serially ran at 1.5 TB/s).
	constexpr int WPV = W / 32;
	constexpr int Q = WPV >= 4 ? WPV / 4 : 1; // lanes per vertex
	constexpr int VPS = 32 / Q;               // vertices per store instruction
	const uint32_t lane = threadIdx.x & 31;
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; // my bitmap word
	uint32_t mine = i < nwords ? fbit[i] : 0u;
	if (!__any_sync(0xffffffffu, mine != 0u)) return;
	if (mine) fbit[i] = 0;
	const uint32_t wbase = i - lane; // first bitmap word of the warp
	const uint32_t g = lane / Q, t = lane % Q;
	unsigned todo = __ballot_sync(0xffffffffu, mine != 0u);
	while (todo) {
		const int w = __ffs((int)todo) - 1;
		todo &= todo - 1u;
		uint32_t m = __shfl_sync(0xffffffffu, mine, w);
		while (m) { // warp-uniform
			uint32_t mm = m; // group g takes the g-th lowest set bit
#pragma unroll
			for (int k = 0; k < VPS - 1; k++)
				if (k < (int)g) mm &= mm - 1u;
			if (mm) {
				const uint32_t v = (wbase + (uint32_t)w) * 32u + (uint32_t)(__ffs((int)mm) - 1);
				uint32_t *p = visit + (size_t)v * WPV;
				if (WPV >= 4) *reinterpret_cast<uint4 *>(p + t * 4) = make_uint4(0u, 0u, 0u, 0u);
				else *reinterpret_cast<uint2 *>(p) = make_uint2(0u, 0u);
			}
#pragma unroll
			for (int k = 0; k < VPS; k++) m &= m - 1u;
		}
	}
}

template <int W>
__global__ void __launch_bounds__(256) clear_by_bitmap_kernel(uint32_t *__restrict__ fbit, uint32_t nwords,
                                                              uint32_t *__restrict__ visit) {
	clear_by_bitmap_body<W>(fbit, nwords, visit);
}

// Start of every level of the loop: the buffer that becomes `next` still holds the frontier of two
// levels ago, named by old_fbit.  Zero exactly those words and that bitmap (which is the one the level
// after this builds); afterwards `next` is all-zero, so the level kernels store only words that gained
// bits.  words == 0 (shortestpath: `next` is a fresh level buffer): only the bitmap is zeroed.
template <int W>
__global__ void __launch_bounds__(256) level_clear_kernel(BfsCtrl *ctrl, uint32_t nwords, int words) {
	const LoopState &L = ctrl->L;
	// words == 2 (dense-store level: every listed vertex rewrites its word): only level 2 scrubs, because
	// `next` then still holds the source words, some at in-degree-0 vertices that no class list names
	if (words == 1 || (words == 2 && L.iter == 2)) {
		clear_by_bitmap_body<W>(L.old_fbit, nwords, L.cur_next);
	} else {
		const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i < nwords) L.old_fbit[i] = 0;
	}
}

// Start of a batch that reuses the frontier buffers of the previous batch of this worker: their only
// non-zero words are the ones the bitmaps of its last two frontiers name (ctrl->L as that batch left
// it).  which = 0: the last `visit` buffer / its bitmap, 1: the last `next` buffer / the bitmap before.
template <int W>
__global__ void __launch_bounds__(256) scrub_kernel(BfsCtrl *ctrl, uint32_t nwords, int which) {
	const LoopState &L = ctrl->L;
	if (which == 0) clear_by_bitmap_body<W>(L.fbit[L.fb[0]], nwords, L.visit[L.cur]);
	else clear_by_bitmap_body<W>(L.fbit[L.fb[2]], nwords, L.visit[L.cur ^ 1]);
}

// shortestpath: zero the level buffer the level about to run writes (16 bytes per thread and step)
static __global__ void __launch_bounds__(256) zero_next_kernel(BfsCtrl *ctrl, size_t n16) {
	uint4 *p = reinterpret_cast<uint4 *>(ctrl->L.cur_next);
	const size_t stride = (size_t)gridDim.x * blockDim.x;
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) p[i] = make_uint4(0u, 0u, 0u, 0u);
}

// pull -> push transition: rebuild the work queue from the frontier bitmap.  One thread per bitmap
// word, two passes (count items, warp-scan, one atomic per warp, write).  The vertex / edge totals
// of this frontier were already published by frontier_stats_kernel; only the item count is new.
static __global__ void __launch_bounds__(256) compact_bitmap_kernel(uint32_t nwords, uint32_t V,
                                                             const int64_t *__restrict__ off, BfsCtrl *ctrl) {
	if (ctrl->L.have_queue) return; // the previous level was top-down too: its output queue is the input
	const uint32_t *const fbit = ctrl->L.cur_fbit;
	uint2 *const queue = const_cast<uint2 *>(ctrl->L.cur_queue);
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t lane = threadIdx.x & 31;
	const uint32_t m = i < nwords ? fbit[i] : 0u;
	uint32_t items = 0;
	for (uint32_t mm = m; mm; mm &= mm - 1) {
		const uint32_t n = i * 32 + (__ffs(mm) - 1);
		if (n >= V) break;
		const unsigned long long deg = (unsigned long long)(off[n + 1] - off[n]);
		const uint32_t k = (uint32_t)((deg + PUSH_CHUNK - 1) / PUSH_CHUNK);
		items += k ? k : 1u;
	}
	uint32_t incl = items;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		const uint32_t y = __shfl_up_sync(0xffffffffu, incl, d);
		if (lane >= (uint32_t)d) incl += y;
	}
	const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
	if (total == 0) return; // warp-uniform
	uint32_t base = 0;
	if (lane == 0) base = atomicAdd(&ctrl->next_items, total);
	base = __shfl_sync(0xffffffffu, base, 0);
	uint32_t pos = base + incl - items;
	for (uint32_t mm = m; mm; mm &= mm - 1) {
		const uint32_t n = i * 32 + (__ffs(mm) - 1);
		if (n >= V) break;
		const unsigned long long deg = (unsigned long long)(off[n + 1] - off[n]);
		uint32_t k = (uint32_t)((deg + PUSH_CHUNK - 1) / PUSH_CHUNK);
		if (k == 0) k = 1;
		for (uint32_t c = 0; c < k; c++) queue[pos++] = make_uint2(n, c);
	}
}

// publish the rebuilt queue's item count
static __global__ void bfs_snapshot_items_kernel(BfsCtrl *ctrl) {
	if (ctrl->L.have_queue) return;
	ctrl->out_items = ctrl->next_items;
	ctrl->next_items = 0;
}

// TEPS numerator of the level about to run (stats mode only; not part of the timed path).
template <int W>
__global__ void __launch_bounds__(256) bfs_count_kernel(const uint32_t *__restrict__ visit,
                                                        const int64_t *__restrict__ off, uint32_t V,
                                                        BfsCtrl *ctrl) {
	constexpr int WPV = W / 32;
	const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
	unsigned long long el = 0, eu = 0, va = 0;
	if (n < V) {
		uint32_t pc = 0;
#pragma unroll
		for (int i = 0; i < WPV; i++) pc += __popc(visit[(size_t)n * WPV + i] & ctrl->active[i]);
		if (pc) {
			unsigned long long deg = (unsigned long long)(off[n + 1] - off[n]);
			el = deg * pc; eu = deg; va = 1;
		}
	}
#pragma unroll
	for (int d = 16; d > 0; d >>= 1) {
		el += __shfl_xor_sync(0xffffffffu, el, d);
		eu += __shfl_xor_sync(0xffffffffu, eu, d);
		va += __shfl_xor_sync(0xffffffffu, va, d);
	}
	if ((threadIdx.x & 31) == 0 && va) {
		atomicAdd(&ctrl->edges_lane, el);
		atomicAdd(&ctrl->edges_unique, eu);
		atomicAdd(&ctrl->verts_active, va);
	}
}

// ---- 1-D partitioned variants (msbfs_part_run) -----------------------------------------------------
// Every rank runs these on its own replica of the full frontier structures, with identical inputs.

// source bits + frontier bitmap + active mask (no push queue: this path is bottom-up only)
static __global__ void bfs_init_part_kernel(const uint32_t *__restrict__ lane_src, int W, uint32_t *__restrict__ visit,
                                     int32_t *__restrict__ found, BfsCtrl *ctrl, uint32_t *__restrict__ fbit0) {
	__shared__ uint32_t cnt;
	const int lane = threadIdx.x;
	if (lane == 0) {
		cnt = 0;
		ctrl->change = 0; ctrl->change_out = 0;
		ctrl->edges_lane = 0; ctrl->edges_unique = 0; ctrl->verts_active = 0;
		ctrl->next_edges = 0; ctrl->next_verts = 0; ctrl->next_items = 0;
		ctrl->out_edges = 0; ctrl->out_verts = 0; ctrl->out_items = 0;
	}
	__syncthreads();
	const uint32_t s = lane_src[lane];
	const bool used = s != 0xFFFFFFFFu;
	found[lane] = -1;
	// words outside the bitmap are never read, so only the source words need a defined value
	if (used)
		for (int i = 0; i < W / 32; i++) visit[(size_t)s * (W / 32) + i] = 0;
	__syncthreads();
	if (used) {
		atomicOr(&visit[(size_t)s * (W / 32) + (lane >> 5)], 1u << (lane & 31));
		atomicOr(&fbit0[s >> 5], 1u << (s & 31));
	}
	const unsigned m = __ballot_sync(0xffffffffu, used);
	if ((lane & 31) == 0) {
		ctrl->active[lane >> 5] = m;
		ctrl->alive[lane >> 5] = 0;
		atomicAdd(&cnt, (uint32_t)__popc(m));
	}
	__syncthreads();
	if (lane == 0) ctrl->n_active = cnt;
}

// Instrumented pass of the partitioned path: TEPS numerator of the level about to run, same definition as
// bfs_count_kernel (sum over lanes of the out-degrees of the lane's frontier vertices).  A rank holds no
// out-degrees, but every edge v -> n is an in-edge of exactly one rank's slice: each rank counts, over its own
// in-edges, the lanes that have the edge's source in their frontier; the ranks' sums add up to the numerator.
template <int W>
__global__ void __launch_bounds__(256) part_count_kernel(const uint32_t *__restrict__ isrc, uint32_t nedges,
                                                         const uint32_t *__restrict__ visit,
                                                         const uint32_t *__restrict__ fbit, BfsCtrl *ctrl) {
	constexpr int WPV = W / 32;
	const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
	unsigned long long el = 0;
	if (j < nedges) {
		const uint32_t v = isrc[j];
		if (fbit_test(fbit, v)) {
			uint32_t pc = 0;
#pragma unroll
			for (int i = 0; i < WPV; i++) pc += __popc(visit[(size_t)v * WPV + i] & ctrl->active[i]);
			el = pc;
		}
	}
#pragma unroll
	for (int d = 16; d > 0; d >>= 1) el += __shfl_xor_sync(0xffffffffu, el, d);
	if ((threadIdx.x & 31) == 0 && el) atomicAdd(&ctrl->edges_lane, el);
}

// fold (and re-zero) this rank's alive slots into 16 words for the allgather
static __global__ void alive_fold_kernel(uint32_t *__restrict__ alive_slots, uint32_t *__restrict__ out16) {
	__shared__ uint32_t salive[16];
	if (threadIdx.x < 16) salive[threadIdx.x] = 0;
	__syncthreads();
	for (uint32_t idx = threadIdx.x; idx < ALIVE_SLOTS * 16; idx += blockDim.x) {
		const uint32_t v = alive_slots[idx];
		if (v) {
			alive_slots[idx] = 0;
			atomicOr(&salive[idx & 15], v);
		}
	}
	__syncthreads();
	if (threadIdx.x < 16) out16[threadIdx.x] = salive[threadIdx.x];
}

// After the allgather every rank holds the complete new frontier: a lane is found at this level iff
// next[dst] carries its bit (seen[dst] gains a bit exactly when next[dst] does), and a lane is
// alive iff some rank produced a bit for it.  Identical on all ranks, no further exchange.
static __global__ void bfs_check_part_kernel(const uint32_t *__restrict__ next, const uint32_t *__restrict__ lane_dst, int W,
                                      int32_t *__restrict__ found, BfsCtrl *ctrl, int iter,
                                      const uint32_t *__restrict__ alive_all, int world,
                                      const uint32_t *__restrict__ fnext) {
	__shared__ uint32_t cnt;
	__shared__ uint32_t anyalive;
	const int lane = threadIdx.x;
	if (lane == 0) {
		cnt = 0;
		anyalive = 0;
	}
	__syncthreads();
	const uint32_t word = ctrl->active[lane >> 5];
	const bool act = (word >> (lane & 31)) & 1u;
	bool hit = false;
	if (act) {
		const uint32_t d = lane_dst[lane];
		// the word is meaningful only when the vertex is in the new frontier
		hit = ((fnext[d >> 5] >> (d & 31)) & 1u) && ((next[(size_t)d * (W / 32) + (lane >> 5)] >> (lane & 31)) & 1u);
		if (hit) found[lane] = iter;
	}
	const unsigned hm = __ballot_sync(0xffffffffu, hit);
	__syncthreads();
	if ((lane & 31) == 0) {
		uint32_t al = 0;
		for (int r = 0; r < world; r++) al |= alive_all[r * 16 + (lane >> 5)];
		if (al) atomicOr(&anyalive, 1u);
		const uint32_t nw = word & ~hm & al;
		ctrl->alive[lane >> 5] = al;
		ctrl->active[lane >> 5] = nw;
		atomicAdd(&cnt, (uint32_t)__popc(nw));
	}
	__syncthreads();
	if (lane == 0) {
		ctrl->n_active = cnt;
		ctrl->change_out = anyalive;
		ctrl->change = 0;
	}
}

// K8: path reconstruction from the per-level frontier history (R7/R8).  One BLOCK per lane walks
// dst -> src.  The parent of `cur` (discovered at level k+1) is the FIRST in-edge, in reverse-CSR
// order (source id, then forward offset), whose source holds this lane's bit in the frontier of
// level k: the block tests 256 in-edges per step and takes the lowest hit (a hub's in-list can be
// 10^5 entries long before the first hit).  Level 0's frontier is the source itself, found by
// binary search in the id-sorted in-list.
// mode 0: [src,..,dst] (len L+1); mode 1: [src,e0,v1,..,dst] (len 2L+1).
static __global__ void __launch_bounds__(256) bfs_paths_hist_kernel(const uint32_t *const *__restrict__ hist,
                                                             const uint32_t *__restrict__ ioff,
                                                             const uint32_t *__restrict__ isrc,
                                                             const uint32_t *__restrict__ ipos,
                                                             const int64_t *__restrict__ eid,
                                                             const uint32_t *__restrict__ row,
                                                             const uint32_t *__restrict__ lane_src,
                                                             const uint32_t *__restrict__ lane_dst,
                                                             const int32_t *__restrict__ found,
                                                             const unsigned long long *__restrict__ lane_off, int W,
                                                             int mode, int64_t *__restrict__ out) {
	__shared__ uint32_t s_first;
	const int lane = blockIdx.x; // BFS lane served by this block
	const uint32_t tid = threadIdx.x;
	if (lane >= W) return;
	const int L = found[lane];
	if (L < 0) return;
	const int WPV = W / 32, wsel = lane >> 5;
	const uint32_t bit = 1u << (lane & 31);
	uint32_t cur = lane_dst[lane];
	int64_t *o = out + lane_off[lane];
	if (tid == 0) o[mode == 0 ? L : 2 * L] = (int64_t)cur;
	for (int k = L - 1; k >= 0; k--) {
		const uint32_t lo = ioff[cur], hi = ioff[cur + 1];
		uint32_t pq = 0xFFFFFFFFu;
		if (k == 0) {
			// frontier 0 = {src}: first in-edge from src = lower bound in the id-sorted in-list
			const uint32_t s = lane_src[lane];
			uint32_t a = lo, b = hi;
			while (a < b) {
				const uint32_t mid = a + ((b - a) >> 1);
				if (isrc[mid] < s) a = mid + 1;
				else b = mid;
			}
			if (a < hi && isrc[a] == s) pq = a;
		} else {
			const uint32_t *fr = hist[k];
			for (uint32_t base = lo; base < hi; base += 256) {
				if (tid == 0) s_first = 0xFFFFFFFFu;
				__syncthreads();
				const uint32_t j = base + tid;
				bool ok = false;
				if (j < hi) ok = (fr[(size_t)isrc[j] * WPV + wsel] & bit) != 0;
				const unsigned m = __ballot_sync(0xffffffffu, ok);
				if (m && (tid & 31) == 0) atomicMin(&s_first, base + (tid & ~31u) + (uint32_t)(__ffs(m) - 1));
				__syncthreads();
				pq = s_first;
				__syncthreads(); // everyone has read s_first before the next step resets it
				if (pq != 0xFFFFFFFFu) break;
			}
		}
		if (pq == 0xFFFFFFFFu) { // cannot happen for a found lane; leave a visible marker
			if (tid == 0) o[mode == 0 ? k : 2 * k] = -1;
			return;
		}
		const uint32_t pv = isrc[pq];
		if (tid == 0) {
			if (mode == 0) {
				o[k] = (int64_t)pv;
			} else {
				const uint32_t p = ipos[pq];
				o[2 * k + 1] = eid ? eid[p] : (int64_t)row[p];
				o[2 * k] = (int64_t)pv;
			}
		}
		cur = pv;
	}
}

} // namespace pgq
