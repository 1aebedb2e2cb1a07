// Host driver of the bitset MS-BFS (iterativelength / reachability / shortestpath batches).
// Mirrors the reference's batch loop: clear seen/visit, seed the lanes,
// run level steps while a lane is active and something changed, check destinations after each
// level, retire found lanes.
//
// The level loop itself lives on the device (msbfs_kernels.cuh LoopState): the check kernel that ends
// a level rotates the frontier views and picks the direction of the next one, and the whole loop is
// one CUDA graph -- loop_begin -> WHILE { SWITCH(mode) { push | pull | pull filtered early | pull
// filtered late } -> check } -- launched once per batch.  Between the init kernel and the result
// copy the host does nothing: batches of a call are enqueued back to back on several streams (and,
// in a multi-GPU context, several devices) by ONE host thread.  The host-driven loop of round 1 is
// kept for the instrumented passes (options "stats", "timing", "trace") and as "loop"=0; it launches
// the same level bodies from the mode it reads back after every level.
#include "msbfs_kernels.cuh"
#include "msbfs_small.cuh"
#include "nccl_dl.cuh"

#include <atomic>
#include <functional>
#include <thread>

namespace pgq {

namespace {

struct BatchBufs {
	uint32_t *seen, *visit[2], *hubacc;
	uint2 *queue[2];
	uint32_t *fbit[3]; // rotating frontier bitmaps (1 bit per vertex, padded): current / next / two levels ago
	size_t fbit_bytes;
	uint32_t fbit_words;
	uint32_t *alive_slots; // [ALIVE_SLOTS][16], all-zero between levels
	uint32_t *done;        // 1 bit per vertex (same geometry as fbit), or nullptr
	BfsCtrl *ctrl;
};

// ---- launchers.  A kernel of a level body is either launched on a stream (host-driven loop) or added as an
// explicit kernel node, chained after the previous one, to a body graph of the loop graph (device-side loop).
// (Explicit nodes, not stream capture: graphs whose conditional bodies were filled by capture crashed inside
// cuGraphDestroy on driver 580.159.)  nk counts the kernels issued. ----
template <class T>
struct ident {
	using type = T;
};

struct Launch {
	cudaStream_t st = nullptr;
	cudaGraph_t graph = nullptr; // graph mode when set
	cudaGraphNode_t last = nullptr;
	uint32_t nk = 0;
	cudaError_t err = cudaSuccess;
	template <class... P>
	void run(void (*k)(P...), dim3 grid, dim3 block, typename ident<P>::type... args) {
		nk++;
		if (!graph) {
			k<<<grid, block, 0, st>>>(args...);
			return;
		}
		void *ptrs[] = {(void *)&args...};
		cudaKernelNodeParams kp{};
		kp.func = (void *)k;
		kp.gridDim = grid;
		kp.blockDim = block;
		kp.sharedMemBytes = 0;
		kp.kernelParams = ptrs;
		cudaGraphNode_t node = nullptr;
		const cudaError_t e = cudaGraphAddKernelNode(&node, graph, last ? &last : nullptr, last ? 1 : 0, &kp);
		if (e != cudaSuccess && err == cudaSuccess) err = e;
		last = node;
	}
};

// STATS picks the instrumented instantiation of the untimed pass.
template <int W, int GPV, bool STATS = false>
void launch_class(PullArgs &a, const DevBuf &list, uint32_t nlist, bool filter, Launch &L) {
	if (!nlist) return;
	using C = LaneCfg<W>;
	constexpr int VPW = C::G / GPV;
	a.list = list.as<uint32_t>();
	a.nlist = nlist;
	const uint32_t warps = div_up(nlist, VPW);
	const uint32_t grid = div_up((uint64_t)warps * 32, 128);
	if (filter) L.run(pull_list_kernel<W, GPV, true, STATS>, grid, 128, a);
	else L.run(pull_list_kernel<W, GPV, false, STATS>, grid, 128, a);
}

// filtered levels, small in-degrees: SEG lanes per vertex, edge-parallel frontier test
template <int W, int SEG, bool STATS = false>
void launch_seg(PullArgs &a, const DevBuf &list, uint32_t nlist, Launch &L) {
	if (!nlist) return;
	a.list = list.as<uint32_t>();
	a.nlist = nlist;
	const uint32_t warps = div_up(nlist, 32 / SEG);
	L.run(pull_seg_kernel<W, SEG, STATS>, div_up((uint64_t)warps * 32, 128), 128, a);
}

template <int W, bool STATS = false>
void launch_warp(PullArgs &a, const DevBuf &list, uint32_t nlist, bool filter, Launch &L) {
	if (!nlist) return;
	a.list = list.as<uint32_t>();
	a.nlist = nlist;
	const uint32_t grid = div_up((uint64_t)nlist * 32, 128);
	if (filter) L.run(pull_warp_kernel<W, true, STATS>, grid, 128, a);
	else L.run(pull_warp_kernel<W, false, STATS>, grid, 128, a);
}

// tiny class: straight-line kernel when the class really is <= 4 in-edges, else the generic one
template <int W, bool STATS = false>
void launch_tiny(PullArgs &a, Csr *g, bool filter, Launch &L) {
	if (!g->n_tiny) return;
	if (g->bins_tiny_max > 4) {
		launch_class<W, 1, STATS>(a, g->d_tiny, g->n_tiny, filter, L);
		return;
	}
	using C = LaneCfg<W>;
	a.list = g->d_tiny.as<uint32_t>();
	a.nlist = g->n_tiny;
	const uint32_t grid = div_up((uint64_t)div_up(g->n_tiny, C::G) * 32, 128);
	if (filter) L.run(pull_tiny_kernel<W, true, STATS>, grid, 128, a);
	else L.run(pull_tiny_kernel<W, false, STATS>, grid, 128, a);
}

template <int W, bool STATS = false>
void launch_hubs(PullArgs &a, Csr *g, Launch &L) {
	if (!g->n_hub) return;
	using C = LaneCfg<W>;
	a.list = g->d_hub.as<uint32_t>();
	a.nlist = g->n_hub;
	a.chunks = g->d_hubchunks.as<HubChunk>();
	a.nchunks = g->n_hubchunks;
	L.run(pull_hub_kernel<W, STATS>, g->n_hubchunks, 256, a);
	L.run(hub_finalize_kernel<W>, div_up((uint64_t)g->n_hub * C::TPV, 256), 256, a);
}

// One level of `mode`: the body of one SWITCH case of the loop graph and, identically, what the host loop
// launches.  Everything that changes from level to level (frontier buffers, bitmaps, queues) is read from
// ctrl->L by the kernels.
template <int W, bool STATS>
int launch_body(Ctx *c, Csr *g, const BatchBufs &b, uint32_t mode, bool keep_hist, Launch &L,
                const LoopState *hl = nullptr) {
	const uint32_t cgrid = div_up(b.fbit_words, 256);
	const size_t bm = (size_t)g->V * (W / 8);
	// `next` becomes all-zero: scrub the words its bitmap names (ping-pong pair) or zero the fresh level
	// buffer (shortestpath); either way the bitmap two levels old is zeroed -- it is the one the level after
	// this builds.  Unfiltered bottom-up levels store densely (most vertices gain bits there), every other level
	// sparsely after the scrub.
	// Stores of a bottom-up level: dense = every listed vertex writes its `next` word (round 1's form).  The
	// unfiltered instantiations always store densely (compile-time); on filtered levels it is option dense_store
	// (default 1: measured 2.09 / 0.56 ms dense against 2.16 / 0.74 ms sparse on levels 2 / 6 of an s24 batch,
	// profiles/r02_ab_views.md).  Top-down levels scrub `next` by bitmap first, as before.
	const bool dense = mode == MODE_PULL || (mode != MODE_PUSH && c->opt.dense_store);
	L.run(level_clear_kernel<W>, cgrid, 256, b.ctrl, b.fbit_words, keep_hist ? 0 : dense ? 2 : 1);
	if (keep_hist && bm) L.run(zero_next_kernel, (uint32_t)c->sm_count * 4, 256, b.ctrl, bm / 16);
	if (mode == MODE_PUSH) {
		// pull -> push transition: the queue is rebuilt from the frontier bitmap (both kernels return at once
		// when the previous level left a queue)
		L.run(compact_bitmap_kernel, cgrid, 256, b.fbit_words, (uint32_t)g->V, g->d_off.as<int64_t>(), b.ctrl);
		L.run(bfs_snapshot_items_kernel, 1, 1, b.ctrl);
		PushArgs pa{};
		pa.off = g->d_off.as<int64_t>();
		pa.tgt = g->d_tgt.as<uint32_t>();
		pa.seen = b.seen;
		pa.ctrl = b.ctrl;
		pa.alive_slots = b.alive_slots;
		L.run(push_kernel<W, STATS>, (uint32_t)c->sm_count * 8, 256, pa);
	} else {
		const bool filter = mode != MODE_PULL, early = mode == MODE_PULL_F_EARLY;
		PullArgs a{};
		a.ioff = g->d_ioff.as<uint32_t>();
		a.isrc = g->d_isrc.as<uint32_t>();
		a.seen = b.seen;
		a.ctrl = b.ctrl;
		a.V = (uint32_t)g->V;
		a.hubacc = b.hubacc;
		a.off = g->d_off.as<int64_t>();
		a.alive_slots = b.alive_slots;
		a.done = b.done;
		a.dyn = 1;
		if (hl && !c->opt.dyn) { // host-driven loop, A/B switch: the level's views as kernel arguments
			a.dyn = 0;
			a.visit = hl->cur_visit;
			a.next = hl->cur_next;
			a.fcur = hl->cur_fbit;
			a.fnext = hl->nxt_fbit;
		}
		a.dense = dense ? 1u : 0u;
		// the done test only on sparse-store tail levels: a skipped vertex stores nothing, and a dense level relies
		// on every listed vertex overwriting what `next` still holds (testing it everywhere left stale frontier
		// words behind: measured wrong frontiers from level 2 of the NEXT batch on)
		a.use_done = (b.done && mode == MODE_PULL_F_LATE && !dense) ? 1u : 0u;
		launch_tiny<W, STATS>(a, g, filter, L);
		// measured on the filtered levels of R-MAT s24 (levels 2 / 6): pull_list 2.23 / 0.62 ms, 8 lanes per
		// vertex 2.08 / 0.56 ms, 16 lanes 2.31 / 0.73 ms; the same kernel for the tiny class: no change
		if (filter && c->opt.seg_small) launch_seg<W, 8, STATS>(a, g->d_small, g->n_small, L);
		else launch_class<W, 2, STATS>(a, g->d_small, g->n_small, filter, L);
		// medium in-degrees: the edge-parallel warp scan wins while most vertices still scan their whole
		// in-list against a sparse frontier (early levels); two thread groups per vertex win when every
		// word is gathered anyway (dense levels) or most vertices have nothing left to find (tail)
		if (filter && early) launch_warp<W, STATS>(a, g->d_med, g->n_med, true, L);
		else launch_class<W, 2, STATS>(a, g->d_med, g->n_med, filter, L);
		launch_warp<W, STATS>(a, g->d_mid, g->n_mid, filter, L);
		launch_hubs<W, STATS>(a, g, L);
		L.run(frontier_stats_kernel, cgrid, 256, b.fbit_words, (uint32_t)g->V, g->d_off.as<int64_t>(), b.ctrl);
	}
	if (L.graph) {
		if (L.err != cudaSuccess) {
			set_error("loop graph: cudaGraphAddKernelNode -> %s", cudaGetErrorString(L.err));
			return PGQ_ERR_CUDA;
		}
	} else {
		PGQ_CUDA(cudaGetLastError());
	}
	return PGQ_OK;
}

int launch_body_w(Ctx *c, Csr *g, const BatchBufs &b, int W, bool stats, uint32_t mode, bool keep_hist, Launch &L,
                  const LoopState *hl = nullptr) {
	if (stats) {
		switch (W) {
		case 64: return launch_body<64, true>(c, g, b, mode, keep_hist, L, hl);
		case 128: return launch_body<128, true>(c, g, b, mode, keep_hist, L, hl);
		case 256: return launch_body<256, true>(c, g, b, mode, keep_hist, L, hl);
		case 512: return launch_body<512, true>(c, g, b, mode, keep_hist, L, hl);
		}
	} else {
		switch (W) {
		case 64: return launch_body<64, false>(c, g, b, mode, keep_hist, L, hl);
		case 128: return launch_body<128, false>(c, g, b, mode, keep_hist, L, hl);
		case 256: return launch_body<256, false>(c, g, b, mode, keep_hist, L, hl);
		case 512: return launch_body<512, false>(c, g, b, mode, keep_hist, L, hl);
		}
	}
	set_error("unsupported lane width %d", W);
	return PGQ_ERR_INVALID;
}

void launch_check(const BatchBufs &b, int W, cudaStream_t st, int use_handles, cudaGraphConditionalHandle hw,
                  cudaGraphConditionalHandle hs) {
	bfs_check_kernel<<<1, W, 0, st>>>(b.seen, W, b.ctrl, b.alive_slots, use_handles, hw, hs);
}

// zero the words of the previous batch's last two frontiers (see scrub_kernel)
int launch_scrub_w(Ctx *c, const BatchBufs &b, int W, int which) {
	const uint32_t grid = div_up(b.fbit_words, 256);
	switch (W) {
	case 64: scrub_kernel<64><<<grid, 256, 0, c->stream>>>(b.ctrl, b.fbit_words, which); break;
	case 128: scrub_kernel<128><<<grid, 256, 0, c->stream>>>(b.ctrl, b.fbit_words, which); break;
	case 256: scrub_kernel<256><<<grid, 256, 0, c->stream>>>(b.ctrl, b.fbit_words, which); break;
	case 512: scrub_kernel<512><<<grid, 256, 0, c->stream>>>(b.ctrl, b.fbit_words, which); break;
	default: return PGQ_ERR_INVALID;
	}
	count_launch(c);
	return PGQ_OK;
}

void launch_count(Ctx *c, Csr *g, const BatchBufs &b, int W, const uint32_t *visit) {
	uint32_t grid = div_up(g->V, 256);
	if (!grid) return;
	const int64_t *off = g->d_off.as<int64_t>();
	switch (W) {
	case 64: bfs_count_kernel<64><<<grid, 256, 0, c->stream>>>(visit, off, (uint32_t)g->V, b.ctrl); break;
	case 128: bfs_count_kernel<128><<<grid, 256, 0, c->stream>>>(visit, off, (uint32_t)g->V, b.ctrl); break;
	case 256: bfs_count_kernel<256><<<grid, 256, 0, c->stream>>>(visit, off, (uint32_t)g->V, b.ctrl); break;
	case 512: bfs_count_kernel<512><<<grid, 256, 0, c->stream>>>(visit, off, (uint32_t)g->V, b.ctrl); break;
	}
	count_launch(c);
}

int read_ctrl(Ctx *c, const BatchBufs &b, BfsCtrl *h) {
	PGQ_CUDA(cudaMemcpyAsync(c->h_ctrl, b.ctrl, sizeof(BfsCtrl), cudaMemcpyDeviceToHost, c->stream));
	PGQ_CUDA(cudaStreamSynchronize(c->stream));
	c->stats.host_syncs++;
	*h = *reinterpret_cast<BfsCtrl *>(c->h_ctrl);
	return PGQ_OK;
}

// levels / launches the device counted since the last fold (the loop runs without the host)
void fold_totals(Ctx *c, const BfsCtrl &h) {
	c->stats.levels += h.tot_levels - c->tot_seen[0];
	c->stats.push_levels += h.tot_push_levels - c->tot_seen[1];
	c->stats.kernel_launches += h.tot_launches - c->tot_seen[2];
	c->tot_seen[0] = h.tot_levels;
	c->tot_seen[1] = h.tot_push_levels;
	c->tot_seen[2] = h.tot_launches;
}

// Bytes THIS implementation must move for the levels one batch ran, from what its kernels touched
// (instrumented pass).  No credit for cache hits: a gathered word counts W/8 bytes
// wherever it was served from.
uint64_t model_bytes_of(const BfsCtrl &h, const Csr *g, int W, size_t fbit_bytes, bool keep_hist) {
	const uint64_t wb = (uint64_t)(W / 8), bm = (uint64_t)g->V * wb;
	uint64_t m = 0;
	// bottom-up: list entry + two in-offsets per listed vertex, `seen` per open vertex, `next` + `seen` stores
	// per vertex that gained bits, a source id per scanned in-edge, a frontier word per gather
	m += h.k_listed * 12ull + h.k_open * wb + h.k_new * 2ull * wb + h.k_scanned * 4ull + h.k_gathered * wb;
	// top-down: queue item + frontier word + two out-offsets per item, a target id per edge, a 4-byte
	// read-modify-write per atomic
	m += h.kp_items * (8ull + wb + 16ull) + h.kp_edges * 4ull + (h.kp_seen + h.kp_next) * 8ull;
	// per level: the old bitmap read + zeroed, the new one read for the statistics (+ a fresh level buffer)
	m += (uint64_t)h.L.levels_run * (3ull * fbit_bytes + (keep_hist ? bm : 0ull));
	// per batch: `seen` zeroed, the bitmaps zeroed; every frontier word is scrubbed once before its buffer is reused
	m += bm + 4ull * fbit_bytes + (keep_hist ? 0ull : h.L.cum_verts * wb);
	return m;
}

int alloc_bufs(Ctx *c, Csr *g, int W, BatchBufs *b) {
	const size_t bm = (size_t)(g->V > 0 ? g->V : 1) * (W / 8);
	PGQ_TRY(c->bfs_seen.reserve(bm));
	PGQ_TRY(c->bfs_visit[0].reserve(bm));
	PGQ_TRY(c->bfs_visit[1].reserve(bm));
	if (!c->bfs_ctrl.p) c->ctrl_fresh = true;
	PGQ_TRY(c->bfs_ctrl.reserve(sizeof(BfsCtrl)));
	if (c->ctrl_fresh) { // running totals start at zero
		PGQ_CUDA(cudaMemsetAsync(c->bfs_ctrl.p, 0, sizeof(BfsCtrl), c->stream));
		c->tot_seen[0] = c->tot_seen[1] = c->tot_seen[2] = 0;
		c->ctrl_fresh = false;
		c->bfs_dirty.valid = false;
	}
	const size_t hb = (size_t)(g->n_hub > 0 ? g->n_hub : 1) * (W / 8);
	if (hb > c->bfs_hubacc.cap) {
		PGQ_TRY(c->bfs_hubacc.reserve(hb));
	}
	// hubacc must be zero on entry of every level (finalize re-zeroes it)
	PGQ_CUDA(cudaMemsetAsync(c->bfs_hubacc.p, 0, c->bfs_hubacc.cap, c->stream));
	b->seen = c->bfs_seen.as<uint32_t>();
	b->visit[0] = c->bfs_visit[0].as<uint32_t>();
	b->visit[1] = c->bfs_visit[1].as<uint32_t>();
	b->hubacc = c->bfs_hubacc.as<uint32_t>();
	b->ctrl = c->bfs_ctrl.as<BfsCtrl>();
	PGQ_TRY(c->bfs_alive.reserve((size_t)ALIVE_SLOTS * 16 * sizeof(uint32_t)));
	PGQ_CUDA(cudaMemsetAsync(c->bfs_alive.p, 0, (size_t)ALIVE_SLOTS * 16 * sizeof(uint32_t), c->stream));
	b->alive_slots = c->bfs_alive.as<uint32_t>();
	const size_t qcap = (size_t)g->V + (size_t)(g->E / PUSH_CHUNK) + 2;
	// padded: pull_small stores whole 8/16/32-vertex units, its grid may overhang V by < 256 vertices
	b->fbit_words = (uint32_t)(((size_t)g->V + 256 + 31) / 32);
	b->fbit_bytes = (size_t)b->fbit_words * sizeof(uint32_t);
	for (int i = 0; i < 2; i++) {
		PGQ_TRY(c->bfs_queue[i].reserve(qcap * sizeof(uint2)));
		b->queue[i] = c->bfs_queue[i].as<uint2>();
	}
	// three frontier bitmaps in one allocation
	PGQ_TRY(c->bfs_inq[0].reserve(3 * b->fbit_bytes));
	for (int i = 0; i < 3; i++) b->fbit[i] = c->bfs_inq[0].as<uint32_t>() + (size_t)i * b->fbit_words;
	b->done = nullptr;
	if (c->opt.done_bitmap) {
		PGQ_TRY(c->bfs_done.reserve(b->fbit_bytes));
		b->done = c->bfs_done.as<uint32_t>();
	}
	return PGQ_OK;
}

// ---- shortestpath: per-level frontier history -----------------------------------------------------------
static constexpr int PGQ_HIST_OVER = 0x7002; // internal: the level history would exceed its budget

size_t hist_budget(Ctx *c) {
	if (c->opt.hist_budget_mb > 0) return (size_t)c->opt.hist_budget_mb << 20;
	size_t fr = 0, tot = 0;
	if (cudaMemGetInfo(&fr, &tot) != cudaSuccess) {
		(void)cudaGetLastError();
		return (size_t)4 << 30;
	}
	return (fr + c->hist_bytes) / 4; // a quarter of what is free once the history itself is not counted
}

// Level buffers 0..levels-1 of bm bytes each (grown, never zeroed here) and their device pointer table.
int hist_reserve(Ctx *c, size_t levels, size_t bm) {
	if (levels * (bm ? bm : 16) > hist_budget(c)) return PGQ_HIST_OVER;
	if (c->bfs_hist.size() < levels) c->bfs_hist.resize(levels);
	bool changed = false;
	for (size_t k = 0; k < levels; k++) {
		DevBuf &hb = c->bfs_hist[k];
		if (hb.cap < (bm ? bm : 16)) {
			c->hist_bytes -= hb.actual;
			PGQ_TRY(hb.reserve(bm ? bm : 16));
			c->hist_bytes += hb.actual;
			changed = true;
		}
	}
	if (changed || c->bfs_hist_tab.cap < c->bfs_hist.size() * sizeof(uint32_t *)) {
		std::vector<uint32_t *> hp(c->bfs_hist.size());
		for (size_t k = 0; k < hp.size(); k++) hp[k] = c->bfs_hist[k].as<uint32_t>();
		PGQ_CUDA(cudaStreamSynchronize(c->stream)); // nothing in flight reads the old table
		PGQ_TRY(c->bfs_hist_tab.reserve(hp.size() * sizeof(uint32_t *)));
		PGQ_CUDA(cudaMemcpyAsync(c->bfs_hist_tab.p, hp.data(), hp.size() * sizeof(uint32_t *), cudaMemcpyHostToDevice,
		                         c->stream));
		PGQ_CUDA(cudaStreamSynchronize(c->stream)); // hp is a host temporary
	}
	return PGQ_OK;
}

void hist_trim(Ctx *c) {
	for (DevBuf &hb : c->bfs_hist) hb.release();
	c->bfs_hist.clear();
	c->bfs_hist_tab.release();
	c->hist_bytes = 0;
}

// ---- small graphs: the whole batch in one cooperative launch (msbfs_small.cuh) ------------------
static constexpr int SMALL_HIST_LEVELS = 48;      // level buffers kept for shortestpath
static constexpr int PGQ_SMALL_OVERFLOW = 0x7001; // internal: history too short, use the general path

static constexpr int SMALL_PENDING_MAX = 512;

// Level counts of small-graph batches are copied back asynchronously; fold them into the stats.
int small_flush(Ctx *c) {
	if (!c->small_pending) return PGQ_OK;
	PGQ_CUDA(cudaStreamSynchronize(c->stream));
	const uint32_t *lv = reinterpret_cast<const uint32_t *>(static_cast<const char *>(c->h_ctrl) + 2048);
	for (int i = 0; i < c->small_pending; i++) c->stats.levels += lv[i];
	c->small_pending = 0;
	return PGQ_OK;
}

bool small_eligible(const Ctx *c, const Csr *g, int W, bool keep_hist) {
	if (!c->opt.small_path || c->opt.stats || g->V <= 0 || g->is_part) return false;
	const uint64_t bm = (uint64_t)g->V * (W / 8), eb = (uint64_t)g->E * (W / 8);
	if (keep_hist && bm > (2ull << 20)) return false; // 49 level buffers
	return bm <= (8ull << 20) && eb <= ((uint64_t)c->opt.small_edge_mb << 20);
}

template <int W>
int small_grid(Ctx *c, int *out) {
	static int per_sm = -1; // per instantiation
	if (per_sm < 0) {
		int nb = 0;
		PGQ_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, msbfs_small_kernel<W>, SMALL_THREADS, 0));
		per_sm = nb;
	}
	*out = per_sm;
	return PGQ_OK;
}

int run_batch_small(Ctx *c, Csr *g, const BatchBufs &b, int W, bool keep_hist, const uint32_t *d_lane_src,
                    const uint32_t *d_lane_dst, int32_t *d_found, int *out_levels) {
	cudaStream_t st = c->stream;
	const size_t bm = (size_t)g->V * (W / 8);
	SmallArgs a{};
	a.ioff = g->d_ioff.as<uint32_t>();
	a.isrc = g->d_isrc.as<uint32_t>();
	a.V = (uint32_t)g->V;
	a.team_threads = 256;
	a.warp_max = (uint32_t)g->bins_small_max; // the lists below hold exactly the vertices above it
	a.big[0] = g->d_med.as<uint32_t>();
	a.nbig[0] = g->n_med;
	a.big[1] = g->d_mid.as<uint32_t>();
	a.nbig[1] = g->n_mid;
	a.big[2] = g->d_hub.as<uint32_t>();
	a.nbig[2] = g->n_hub;
	a.seen = b.seen;
	a.buf[0] = b.visit[0];
	a.buf[1] = b.visit[1];
	a.lane_src = d_lane_src;
	a.lane_dst = d_lane_dst;
	a.found = d_found;
	a.ctrl = b.ctrl;
	a.max_levels = 0x7fffffff;
	a.level_cap = (int)c->opt.level_cap;
	if (keep_hist) {
		const int rc = hist_reserve(c, (size_t)SMALL_HIST_LEVELS + 1, bm);
		if (rc == PGQ_HIST_OVER) return PGQ_SMALL_OVERFLOW;
		PGQ_TRY(rc);
		a.hist = c->bfs_hist_tab.as<uint32_t *const>();
		a.max_levels = SMALL_HIST_LEVELS;
	}
	c->bfs_dirty.valid = false; // the frontier pair is left in a state the scrub does not describe
	int per_sm = 0;
	switch (W) {
	case 64: PGQ_TRY(small_grid<64>(c, &per_sm)); break;
	case 128: PGQ_TRY(small_grid<128>(c, &per_sm)); break;
	case 256: PGQ_TRY(small_grid<256>(c, &per_sm)); break;
	default: PGQ_TRY(small_grid<512>(c, &per_sm)); break;
	}
	if (per_sm < 1) return PGQ_SMALL_OVERFLOW;
	int grid = c->sm_count; // one 1024-thread block per SM: the grid barrier scales with the block count
	const int want = (int)div_up((uint64_t)g->V, SMALL_THREADS / 32);
	if (grid > want) grid = want;
	if (grid < 1) grid = 1;
	void *args[] = {&a};
	const void *fn = nullptr;
	switch (W) {
	case 64: fn = (const void *)msbfs_small_kernel<64>; break;
	case 128: fn = (const void *)msbfs_small_kernel<128>; break;
	case 256: fn = (const void *)msbfs_small_kernel<256>; break;
	default: fn = (const void *)msbfs_small_kernel<512>; break;
	}
	if (c->opt.timing) PGQ_CUDA(cudaEventRecord(c->ev_a, st));
	PGQ_CUDA(cudaMemsetAsync(&b.ctrl->small_bar, 0, offsetof(BfsCtrl, k_listed) - offsetof(BfsCtrl, small_bar), st));
	PGQ_CUDA(cudaLaunchCooperativeKernel(fn, dim3((unsigned)grid), dim3(SMALL_THREADS), args, 0, st));
	count_launch(c);
	if (c->opt.timing) PGQ_CUDA(cudaEventRecord(c->ev_b, st));
	if (!keep_hist && !c->opt.timing) {
		// nothing the host needs before the results: the level count is fetched with them
		if (c->small_pending == SMALL_PENDING_MAX) PGQ_TRY(small_flush(c));
		uint32_t *slot = reinterpret_cast<uint32_t *>(static_cast<char *>(c->h_ctrl) + 2048) + c->small_pending++;
		PGQ_CUDA(cudaMemcpyAsync(slot, &b.ctrl->small_levels, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
		c->stats.batches++;
		if (out_levels) *out_levels = 0;
		return PGQ_OK;
	}
	BfsCtrl h;
	PGQ_TRY(read_ctrl(c, b, &h));
	if (h.small_overflow) return PGQ_SMALL_OVERFLOW;
	c->stats.batches++;
	c->stats.levels += h.small_levels;
	if (c->opt.timing) {
		float ms = 0;
		PGQ_CUDA(cudaEventElapsedTime(&ms, c->ev_a, c->ev_b));
		c->stats.level_kernel_ms += ms;
		c->stats.level_kernel_launches += 1;
		if (c->opt.trace)
			fprintf(stderr, "[pgq] small-graph kernel: %u levels %.3f ms, grid %d x %d\n", h.small_levels, ms, grid,
			        SMALL_THREADS);
	}
	if (out_levels) *out_levels = (int)h.small_levels;
	return PGQ_OK;
}

// ---- the level loop as a CUDA graph ---------------------------------------------------------------------
// loop_begin -> WHILE(h_while) { SWITCH(h_switch) { body(mode 1..4) } -> check }.  The bodies are recorded
// by stream capture of launch_body; the conditional handles are set by the check kernel
// (cudaGraphSetConditional).  One graph per worker and (CSR, W, keep_hist): it embeds the worker's buffers.
int loop_graph_build(Ctx *c, Csr *g, const BatchBufs &b, int W, bool keep_hist, LoopGraph *lg, uint32_t *nk_out) {
	lg->destroy();
	cudaGraph_t G = nullptr;
	PGQ_CUDA(cudaGraphCreate(&G, 0));
	lg->graph = G;
	cudaGraphConditionalHandle hw, hs;
	PGQ_CUDA(cudaGraphConditionalHandleCreate(&hw, G, 0, 0));
	PGQ_CUDA(cudaGraphConditionalHandleCreate(&hs, G, 0, 0));
	BfsCtrl *ctrl = b.ctrl;
	cudaGraphNode_t n_begin, n_while, n_switch, n_check;
	{
		void *args[] = {&ctrl, &hw, &hs};
		cudaKernelNodeParams kp{};
		kp.func = (void *)loop_begin_kernel;
		kp.gridDim = dim3(1);
		kp.blockDim = dim3(1);
		kp.kernelParams = args;
		PGQ_CUDA(cudaGraphAddKernelNode(&n_begin, G, nullptr, 0, &kp));
	}
	cudaGraphNodeParams wp{};
	wp.type = cudaGraphNodeTypeConditional;
	wp.conditional.handle = hw;
	wp.conditional.type = cudaGraphCondTypeWhile;
	wp.conditional.size = 1;
	PGQ_CUDA(cudaGraphAddNode(&n_while, G, &n_begin, 1, &wp));
	cudaGraph_t body = wp.conditional.phGraph_out[0];
	cudaGraphNodeParams sp{};
	sp.type = cudaGraphNodeTypeConditional;
	sp.conditional.handle = hs;
	sp.conditional.type = cudaGraphCondTypeSwitch;
	sp.conditional.size = MODE_COUNT - 1;
	PGQ_CUDA(cudaGraphAddNode(&n_switch, body, nullptr, 0, &sp));
	for (uint32_t m = MODE_PUSH; m < MODE_COUNT; m++) {
		Launch L;
		L.graph = sp.conditional.phGraph_out[m - 1];
		PGQ_TRY(launch_body_w(c, g, b, W, false, m, keep_hist, L));
		nk_out[m] = L.nk;
	}
	nk_out[MODE_STOP] = 0;
	{
		const uint32_t *seen = b.seen;
		uint32_t *alive = b.alive_slots;
		int use = 1;
		int w = W;
		void *args[] = {&seen, &w, &ctrl, &alive, &use, &hw, &hs};
		cudaKernelNodeParams kp{};
		kp.func = (void *)bfs_check_kernel;
		kp.gridDim = dim3(1);
		kp.blockDim = dim3((unsigned)W);
		kp.kernelParams = args;
		PGQ_CUDA(cudaGraphAddKernelNode(&n_check, body, &n_switch, 1, &kp));
	}
	PGQ_CUDA(cudaGraphInstantiate(&lg->exec, G, 0));
	// The template is not needed once the executable graph exists.  It is destroyed HERE, before the first
	// launch: destroying it later (at rebuild time, after launches) crashed inside cuGraphDestroy on driver
	// 580.159 while it walked the nested WHILE -> SWITCH body graphs.  PGQ_B200_GRAPH_TEMPLATE=late restores that
	// order, =leak never destroys templates (diagnostics).
	{
		static const char *mode = getenv("PGQ_B200_GRAPH_TEMPLATE");
		if (!mode || !strcmp(mode, "early")) {
			PGQ_CUDA(cudaGraphDestroy(G));
			lg->graph = nullptr;
		} else if (!strcmp(mode, "leak")) {
			lg->graph = nullptr;
		}
	}
	lg->g = g;
	lg->gen = g->gen;
	lg->W = W;
	lg->keep_hist = keep_hist;
	lg->seen = b.seen;
	lg->ctrl = b.ctrl;
	lg->done = b.done;
	lg->seg_small = c->opt.seg_small * 16 + c->opt.dense_store * 4 + c->opt.done_bitmap;
	return PGQ_OK;
}

// the worker's loop graph for this batch shape (built on first use, reused by every later batch)
int loop_graph_get(Ctx *c, Csr *g, const BatchBufs &b, int W, bool keep_hist, LoopGraph **out, uint32_t *nk) {
	LoopGraph *lg = &c->loop_graph[keep_hist ? 1 : 0];
	const bool hit = lg->exec && lg->g == g && lg->gen == g->gen && lg->W == W && lg->keep_hist == keep_hist &&
	                 lg->seen == b.seen && lg->ctrl == b.ctrl && lg->done == b.done &&
	                 lg->seg_small == c->opt.seg_small * 16 + c->opt.dense_store * 4 + c->opt.done_bitmap;
	if (!hit) {
		uint32_t fresh[MODE_COUNT] = {0, 0, 0, 0, 0};
		static std::mutex build_mu; // graph construction + capture: one worker thread at a time
		std::lock_guard<std::mutex> lk(build_mu);
		const int rc = loop_graph_build(c, g, b, W, keep_hist, lg, fresh);
		if (rc != PGQ_OK) {
			lg->destroy();
			return rc;
		}
		for (int i = 0; i < (int)MODE_COUNT; i++) c->loop_nk[keep_hist ? 1 : 0][i] = fresh[i];
	}
	for (int i = 0; i < (int)MODE_COUNT; i++) nk[i] = c->loop_nk[keep_hist ? 1 : 0][i];
	*out = lg;
	return PGQ_OK;
}

// persisting-L2 window on `seen` for the worker's stream (north-star item; measured)
void apply_l2_window(Ctx *c, const BatchBufs &b, size_t bm) {
	const int want = c->opt.l2_persist ? 1 : 0;
	if (c->l2_window_on == want && (!want || c->l2_window_ptr == (const void *)b.seen)) return;
	cudaStreamAttrValue v{};
	if (want) {
		int maxw = 0;
		cudaDeviceGetAttribute(&maxw, cudaDevAttrMaxAccessPolicyWindowSize, c->device);
		v.accessPolicyWindow.base_ptr = b.seen;
		v.accessPolicyWindow.num_bytes = bm < (size_t)maxw ? bm : (size_t)maxw;
		v.accessPolicyWindow.hitRatio = 1.0f;
		v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
		v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
	} else {
		v.accessPolicyWindow.num_bytes = 0;
	}
	if (cudaStreamSetAttribute(c->stream, cudaStreamAttributeAccessPolicyWindow, &v) != cudaSuccess)
		(void)cudaGetLastError();
	c->l2_window_on = want;
	c->l2_window_ptr = b.seen;
}

// One batch of W lanes.  d_lane_src/dst/found point at this batch's W slots.
// keep_hist (shortestpath): level k's frontier stays in c->bfs_hist[k] instead of a ping-pong pair, so
// the path kernel can apply the parent rule afterwards; the batch is then complete when this returns and
// *out_levels = levels run.  Otherwise (device loop) the batch is only ENQUEUED: nothing is waited for.
int run_batch(Ctx *c, Csr *g, const BatchBufs &b_in, int W, bool keep_hist, const uint32_t *d_lane_src,
              const uint32_t *d_lane_dst, int32_t *d_found, int *out_levels = nullptr) {
	cudaStream_t st = c->stream;
	BatchBufs b = b_in;
	const size_t bm = (size_t)g->V * (W / 8);
	if (small_eligible(c, g, W, keep_hist)) {
		const int rc = run_batch_small(c, g, b, W, keep_hist, d_lane_src, d_lane_dst, d_found, out_levels);
		if (rc != PGQ_SMALL_OVERFLOW) return rc;
	}
	const bool host_loop = !c->opt.loop_now || c->opt.stats || c->opt.timing || c->opt.trace || c->loop_unsupported;
	apply_l2_window(c, b, bm);
	size_t hist_cap = 0;
	if (keep_hist) {
		hist_cap = c->bfs_hist.size() >= 16 ? c->bfs_hist.size() : 16;
		int rc = hist_reserve(c, hist_cap, bm);
		if (rc == PGQ_HIST_OVER && hist_cap > 2) {
			// what earlier calls left allocated no longer fits the budget: give it back, start from two levels
			PGQ_CUDA(cudaStreamSynchronize(st));
			hist_trim(c);
			hist_cap = 2;
			rc = hist_reserve(c, hist_cap, bm);
		}
		if (rc != PGQ_OK) return rc; // PGQ_HIST_OVER: the caller narrows the lane words
		if (bm) {
			PGQ_CUDA(cudaMemsetAsync(b.seen, 0, bm, st));
			PGQ_CUDA(cudaMemsetAsync(c->bfs_hist[0].p, 0, bm, st));
		}
		c->bfs_dirty.valid = false;
	} else if (bm) {
		PGQ_CUDA(cudaMemsetAsync(b.seen, 0, bm, st));
		// The frontier buffers must start all-zero.  After a clean batch on the same buffers their only
		// non-zero words are the ones its last two frontier bitmaps name (ctrl->L still describes them):
		// scrub those instead of streaming 2 x V x W/8 bytes of zeros.
		BfsDirty &d = c->bfs_dirty;
		if (d.valid && d.W == W && d.V == g->V && d.visit0 == b.visit[0] && d.fbit0 == b.fbit[0]) {
			PGQ_TRY(launch_scrub_w(c, b, W, 0));
			PGQ_TRY(launch_scrub_w(c, b, W, 1));
		} else {
			PGQ_CUDA(cudaMemsetAsync(b.visit[0], 0, bm, st));
			PGQ_CUDA(cudaMemsetAsync(b.visit[1], 0, bm, st));
		}
		d.valid = false; // until this batch is enqueued completely
	}
	PGQ_CUDA(cudaMemsetAsync(b.fbit[0], 0, 3 * b.fbit_bytes, st)); // all three frontier bitmaps
	if (b.done) PGQ_CUDA(cudaMemsetAsync(b.done, 0, b.fbit_bytes, st));
	LoopInit li{};
	li.visit[0] = b.visit[0];
	li.visit[1] = b.visit[1];
	for (int i = 0; i < 3; i++) li.fbit[i] = b.fbit[i];
	li.queue[0] = b.queue[0];
	li.queue[1] = b.queue[1];
	li.hist = keep_hist ? c->bfs_hist_tab.as<uint32_t *const>() : nullptr;
	li.hist_cap = (uint32_t)hist_cap;
	li.level_cap = (uint32_t)(c->opt.level_cap > 0x7fffffff ? 0x7fffffff : c->opt.level_cap);
	li.direction = (uint32_t)c->opt.direction;
	li.push_beta = (uint32_t)c->opt.push_beta;
	li.push_vbeta = (uint32_t)c->opt.push_vbeta;
	li.filter_gamma = (uint32_t)c->opt.filter_gamma;
	li.E = (unsigned long long)g->E;
	li.V = (unsigned long long)g->V;
	li.n_listed = (unsigned long long)g->n_tiny + g->n_small + g->n_med + g->n_mid + g->n_hub;
	li.lane_dst = d_lane_dst;
	LoopGraph *lg = nullptr;
	if (!host_loop) {
		const int rc = loop_graph_get(c, g, b, W, keep_hist, &lg, li.nk);
		if (rc != PGQ_OK) {
			// a driver that refuses conditional nodes: remember it and fall through to the host loop
			c->loop_unsupported = true;
			(void)cudaGetLastError();
			fprintf(stderr, "[pgq_b200] level-loop graph unavailable (%s): falling back to the host-driven loop\n",
			        last_error_str());
			lg = nullptr;
		}
	}
	bfs_init_kernel<<<1, W, 0, st>>>(d_lane_src, W, d_found, b.ctrl, g->d_off.as<int64_t>(), li);
	count_launch(c);
	c->stats.batches++;
	BfsCtrl h;
	if (lg) {
		// ---- device loop: one graph launch per batch, re-launched only when the level history ran out ----
		for (;;) {
			PGQ_CUDA(cudaGraphLaunch(lg->exec, st));
			if (!keep_hist) break;
			PGQ_TRY(read_ctrl(c, b, &h)); // shortestpath sizes its output on the host anyway
			if (h.L.stop_reason != 2) break;
			size_t want = hist_cap * 2;
			int rc = hist_reserve(c, want, bm);
			if (rc == PGQ_HIST_OVER) {
				want = hist_cap + 1;
				rc = hist_reserve(c, want, bm);
			}
			if (rc != PGQ_OK) return rc;
			hist_cap = want;
			loop_resume_kernel<<<1, 1, 0, st>>>(b.ctrl, c->bfs_hist_tab.as<uint32_t *const>(), (uint32_t)hist_cap);
			count_launch(c);
		}
		if (keep_hist) {
			fold_totals(c, h);
			if (out_levels) *out_levels = (int)h.L.levels_run;
		}
	} else {
		// ---- host loop: the mode the device picked is read back after every level ----
		PGQ_TRY(read_ctrl(c, b, &h));
		while (h.L.mode != MODE_STOP || h.L.stop_reason == 2) {
			if (h.L.stop_reason == 2) { // level history full
				size_t want = hist_cap * 2;
				int rc = hist_reserve(c, want, bm);
				if (rc == PGQ_HIST_OVER) {
					want = hist_cap + 1;
					rc = hist_reserve(c, want, bm);
				}
				if (rc != PGQ_OK) return rc;
				hist_cap = want;
				loop_resume_kernel<<<1, 1, 0, st>>>(b.ctrl, c->bfs_hist_tab.as<uint32_t *const>(), (uint32_t)hist_cap);
				count_launch(c);
				PGQ_TRY(read_ctrl(c, b, &h));
				continue;
			}
			if (c->opt.stats) launch_count(c, g, b, W, h.L.cur_visit);
			if (c->opt.timing) PGQ_CUDA(cudaEventRecord(c->ev_a, st));
			Launch L;
			L.st = st;
			const uint32_t mode = h.L.mode;
			PGQ_TRY(launch_body_w(c, g, b, W, c->opt.stats != 0, mode, keep_hist, L, &h.L));
			const uint32_t nk = L.nk;
			count_launch(c, nk);
			if (c->opt.timing) PGQ_CUDA(cudaEventRecord(c->ev_b, st));
			launch_check(b, W, st, 0, 0, 0);
			const uint32_t iter = h.L.iter;
			PGQ_TRY(read_ctrl(c, b, &h));
			if (c->opt.timing) {
				float ms = 0;
				PGQ_CUDA(cudaEventElapsedTime(&ms, c->ev_a, c->ev_b));
				c->stats.level_kernel_ms += ms;
				c->stats.level_kernel_launches += nk;
				if (c->opt.trace)
					fprintf(stderr, "[pgq] iter %u %s %.3f ms  next: verts %u edges %llu items %u  active %u change %u\n",
					        iter, mode == MODE_PUSH ? "push" : mode == MODE_PULL ? "pull" : mode == MODE_PULL_F_EARLY ? "pull-fe" : "pull-fl",
					        ms, h.out_verts, (unsigned long long)h.out_edges, h.out_items, h.n_active, h.change_out);
			}
		}
		fold_totals(c, h);
		if (out_levels) *out_levels = (int)h.L.levels_run;
		if (c->opt.stats) {
			// counters accumulate in the control block over the batch (reset by the init kernel)
			c->stats.edges_lane += h.edges_lane;
			c->stats.edges_unique += h.edges_unique;
			// SURVEY.md §8(d), the REFERENCE algorithm's bytes: 4*V*(W/8) per level + E_act*(4 + W/8) + 8*V_act
			c->stats.algo_bytes += h.edges_unique * (uint64_t)(4 + W / 8) + 8ull * h.verts_active +
			                       (uint64_t)h.L.levels_run * 4ull * (uint64_t)g->V * (uint64_t)(W / 8);
			// what THIS implementation touched
			const uint64_t wb = (uint64_t)(W / 8);
			c->stats.k_listed += h.k_listed;
			c->stats.k_open += h.k_open;
			c->stats.k_new += h.k_new;
			c->stats.k_scanned += h.k_scanned;
			c->stats.k_gathered += h.k_gathered;
			c->stats.kp_items += h.kp_items;
			c->stats.kp_edges += h.kp_edges;
			c->stats.kp_seen += h.kp_seen;
			c->stats.kp_next += h.kp_next;
			c->stats.model_bytes += model_bytes_of(h, g, W, b.fbit_bytes, keep_hist);
			(void)wb;
		}
	}
	if (bm && !keep_hist) { // what the frontier buffers hold once this batch is through, for the next batch's scrub
		BfsDirty &d = c->bfs_dirty;
		d.valid = true;
		d.W = W;
		d.V = g->V;
		d.visit0 = b.visit[0];
		d.fbit0 = b.fbit[0];
	}
	PGQ_CUDA(cudaGetLastError());
	return PGQ_OK;
}

// ---- workers: one (device, stream, buffer set) each ----------------------------------------------------
struct Worker {
	Ctx *w = nullptr;
	Csr *g = nullptr; // the CSR (replica) on the worker's device
	BatchBufs b{};
	uint32_t *d_src = nullptr, *d_dst = nullptr; // this device's copy of the call's lane tables
	int32_t *d_found = nullptr;
};

// Read-only copy of the query-side arrays of g on device `dev` (multi-GPU contexts), made on first use.
#define PGQ_REPL(field)                                                                                   \
	do {                                                                                                  \
		if (g->field.p && g->field.cap) {                                                                 \
			PGQ_TRY(r->field.reserve(g->field.cap));                                                      \
			PGQ_CUDA(cudaMemcpyPeerAsync(r->field.p, dev, g->field.p, c->device, g->field.cap, c->stream)); \
		}                                                                                                 \
	} while (0)

int csr_replica(Ctx *c, Csr *g, int dev, Csr **out) {
	auto it = g->replicas.find(dev);
	if (it != g->replicas.end() && it->second->gen == g->gen) {
		*out = it->second;
		return PGQ_OK;
	}
	if (it != g->replicas.end()) {
		PGQ_CUDA(cudaSetDevice(dev));
		it->second->free_all();
		delete it->second;
		g->replicas.erase(it);
	}
	PGQ_CUDA(cudaSetDevice(c->device));
	int can = 0;
	cudaDeviceCanAccessPeer(&can, c->device, dev);
	if (can && cudaDeviceEnablePeerAccess(dev, 0) != cudaSuccess) (void)cudaGetLastError(); // already enabled
	PGQ_CUDA(cudaSetDevice(dev));
	cudaDeviceCanAccessPeer(&can, dev, c->device);
	if (can && cudaDeviceEnablePeerAccess(c->device, 0) != cudaSuccess) (void)cudaGetLastError();
	Csr *r = new Csr();
	r->device = dev;
	r->V = g->V;
	r->E = g->E;
	r->state = g->state;
	r->have_eid = g->have_eid;
	r->keep_rows = g->keep_rows;
	r->prepared = g->prepared;
	r->n_tiny = g->n_tiny; r->n_small = g->n_small; r->n_med = g->n_med; r->n_mid = g->n_mid; r->n_hub = g->n_hub;
	r->n_hubchunks = g->n_hubchunks;
	r->bins_tiny_max = g->bins_tiny_max; r->bins_small_max = g->bins_small_max; r->bins_med_max = g->bins_med_max;
	r->bins_hub_min = g->bins_hub_min; r->eff_med_max = g->eff_med_max;
	r->gen = g->gen;
	g->replicas[dev] = r;
	PGQ_REPL(d_off);
	PGQ_REPL(d_tgt);
	PGQ_REPL(d_row);
	PGQ_REPL(d_eid);
	PGQ_REPL(d_ioff);
	PGQ_REPL(d_isrc);
	PGQ_REPL(d_ipos);
	PGQ_REPL(d_tiny);
	PGQ_REPL(d_small);
	PGQ_REPL(d_med);
	PGQ_REPL(d_mid);
	PGQ_REPL(d_hub);
	PGQ_REPL(d_hubchunks);
	PGQ_CUDA(cudaSetDevice(c->device));
	PGQ_CUDA(cudaStreamSynchronize(c->stream));
	*out = r;
	return PGQ_OK;
}
#undef PGQ_REPL

// helper context number `idx` of device `dev` (created on first use; lives until pgq_shutdown)
int get_helper(Ctx *c, int dev, int idx, Ctx **out) {
	int seen = 0;
	for (Ctx *h : c->helpers)
		if (h->device == dev && seen++ == idx) {
			*out = h;
			return PGQ_OK;
		}
	PGQ_CUDA(cudaSetDevice(dev));
	Ctx *h = new Ctx();
	h->device = dev;
	h->sm_count = c->sm_count;
	const int rc = ctx_init_resources(h);
	if (rc != PGQ_OK) {
		ctx_release_resources(h);
		delete h;
		cudaSetDevice(c->device);
		return rc;
	}
	c->helpers.push_back(h);
	*out = h;
	return PGQ_OK;
}

// The workers of one call: `streams` per device, over as many devices of the context as the call has
// batches to keep busy (a 1024-row call = 2 batches stays on the primary).  Lane tables (already on the
// primary, written on c->stream) are copied to every other device; all worker streams wait for them.
int make_workers(Ctx *c, Csr *g, const BatchBufs &b0, int W, int64_t nbatch, bool keep_hist, uint32_t *d_src,
                 uint32_t *d_dst, int32_t *d_found, int64_t nslots, std::vector<Worker> *out) {
	out->clear();
	Worker w0;
	w0.w = c;
	w0.g = g;
	w0.b = b0;
	w0.d_src = d_src;
	w0.d_dst = d_dst;
	w0.d_found = d_found;
	out->push_back(w0);
	const bool single = c->opt.timing || c->opt.stats || c->opt.trace || g->is_part || small_eligible(c, g, W, keep_hist);
	if (single || nbatch < 2) return PGQ_OK;
	int ndev = (int)c->devices.size();
	if (ndev < 1) ndev = 1;
	if ((int64_t)ndev > nbatch / 2) ndev = (int)(nbatch / 2 > 0 ? nbatch / 2 : 1);
	const int streams = (int)c->opt.streams;
	int64_t want = (int64_t)ndev * streams;
	if (want > nbatch) want = nbatch;
	PGQ_CUDA(cudaEventRecord(c->ev_fork, c->stream)); // the lane tables are in place after this point
	// stream-major order: worker k sits on device k % ndev, so consecutive batches go to different GPUs
	for (int64_t k = 1; k < want; k++) {
		const int di = (int)(k % ndev), si = (int)(k / ndev);
		const int dev = ndev > 1 ? c->devices[(size_t)di] : c->device;
		Ctx *h = nullptr;
		// the primary's own stream is its worker 0
		int rc = get_helper(c, dev, dev == c->device ? si - 1 : si, &h);
		if (rc != PGQ_OK) return rc;
		Worker wk;
		wk.w = h;
		wk.g = g;
		h->opt = c->opt;
		PGQ_CUDA(cudaSetDevice(dev));
		if (dev != c->device) {
			rc = csr_replica(c, g, dev, &wk.g);
			if (rc != PGQ_OK) {
				cudaSetDevice(c->device);
				return rc;
			}
			PGQ_CUDA(cudaSetDevice(dev));
		}
		rc = alloc_bufs(h, wk.g, W, &wk.b);
		if (rc == PGQ_ERR_OOM) { // no room for another set of bitmaps: the other workers take every batch
			(void)cudaGetLastError();
			continue;
		}
		if (rc != PGQ_OK) {
			cudaSetDevice(c->device);
			return rc;
		}
		if (dev == c->device) {
			wk.d_src = d_src;
			wk.d_dst = d_dst;
			wk.d_found = d_found;
			PGQ_CUDA(cudaStreamWaitEvent(h->stream, c->ev_fork, 0));
		} else {
			// the first worker of a device holds the device's lane tables; later ones share them
			Worker *first = nullptr;
			for (Worker &o : *out)
				if (o.w->device == dev) {
					first = &o;
					break;
				}
			if (first) {
				wk.d_src = first->d_src;
				wk.d_dst = first->d_dst;
				wk.d_found = first->d_found;
				PGQ_CUDA(cudaStreamWaitEvent(h->stream, first->w->ev_fork, 0));
			} else {
				PGQ_TRY(h->bfs_lane.reserve((size_t)nslots * 12 + 64));
				wk.d_src = h->bfs_lane.as<uint32_t>();
				wk.d_dst = wk.d_src + nslots;
				wk.d_found = reinterpret_cast<int32_t *>(wk.d_dst + nslots);
				PGQ_CUDA(cudaStreamWaitEvent(h->stream, c->ev_fork, 0));
				PGQ_CUDA(cudaMemcpyPeerAsync(wk.d_src, dev, d_src, c->device, (size_t)nslots * 4, h->stream));
				PGQ_CUDA(cudaMemcpyPeerAsync(wk.d_dst, dev, d_dst, c->device, (size_t)nslots * 4, h->stream));
				PGQ_CUDA(cudaEventRecord(h->ev_fork, h->stream));
			}
		}
		out->push_back(wk);
	}
	PGQ_CUDA(cudaSetDevice(c->device));
	return PGQ_OK;
}

using BatchFn = std::function<int(Worker &, int64_t)>; // (worker, batch index)

// All batches of one call over the workers.  With the device-side level loop a lengths batch is a handful
// of asynchronous launches, so ONE host thread deals them round-robin and nothing waits until the end.
// Batches that meet the host (shortestpath, host loop) get a thread per worker and are handed out dynamically.
int run_batches(Ctx *c, std::vector<Worker> &ws, int64_t nbatch, bool threaded, const BatchFn &fn) {
	const int nw = (int)ws.size();
	int rc_all = PGQ_OK;
	std::string err_all;
	if (nw == 1 || !threaded) {
		for (int64_t bi = 0; bi < nbatch && rc_all == PGQ_OK; bi++) {
			Worker &wk = ws[(size_t)(bi % nw)];
			if (nw > 1 && cudaSetDevice(wk.w->device) != cudaSuccess) {
				set_error("cudaSetDevice(%d) failed", wk.w->device);
				rc_all = PGQ_ERR_CUDA;
				break;
			}
			rc_all = fn(wk, bi);
		}
		for (int i = 1; i < nw; i++) {
			cudaSetDevice(ws[(size_t)i].w->device);
			if (cudaStreamSynchronize(ws[(size_t)i].w->stream) != cudaSuccess && rc_all == PGQ_OK) {
				set_error("worker stream failed: %s", cudaGetErrorString(cudaGetLastError()));
				rc_all = PGQ_ERR_CUDA;
			}
		}
		cudaSetDevice(c->device);
		if (rc_all != PGQ_OK) err_all = last_error_str();
	} else {
		struct Res {
			int rc = PGQ_OK;
			std::string err;
		};
		std::vector<Res> res((size_t)nw);
		std::atomic<int64_t> next_batch{0};
		std::atomic<int> failed{0};
		auto drain = [&](Worker &wk, Res &r) {
			while (!failed.load(std::memory_order_relaxed)) {
				const int64_t bi = next_batch.fetch_add(1);
				if (bi >= nbatch) break;
				const int rc = fn(wk, bi);
				if (rc != PGQ_OK) {
					r.rc = rc;
					r.err = last_error_str();
					failed.store(1);
					break;
				}
			}
		};
		std::vector<std::thread> threads;
		for (int i = 1; i < nw; i++) {
			threads.emplace_back([&, i] {
				Worker &wk = ws[(size_t)i];
				Res &r = res[(size_t)i];
				if (cudaSetDevice(wk.w->device) != cudaSuccess) {
					set_error("helper thread: cudaSetDevice(%d) failed", wk.w->device);
					r.rc = PGQ_ERR_CUDA;
					r.err = last_error_str();
					failed.store(1);
					return;
				}
				drain(wk, r);
				if (cudaStreamSynchronize(wk.w->stream) != cudaSuccess && r.rc == PGQ_OK) {
					set_error("helper stream failed");
					r.rc = PGQ_ERR_CUDA;
					r.err = last_error_str();
					failed.store(1);
				}
			});
		}
		drain(ws[0], res[0]);
		for (std::thread &t : threads) t.join();
		for (const Res &r : res)
			if (r.rc != PGQ_OK && rc_all == PGQ_OK) {
				rc_all = r.rc;
				err_all = r.err;
			}
	}
	// the device-side running totals of every worker, then the helpers' host-side counters
	for (int i = 0; i < nw && rc_all == PGQ_OK; i++) {
		Worker &wk = ws[(size_t)i];
		cudaSetDevice(wk.w->device);
		BfsCtrl h;
		const int rc = read_ctrl(wk.w, wk.b, &h);
		wk.w->stats.host_syncs--; // end-of-call bookkeeping, not part of a batch
		if (rc != PGQ_OK) {
			rc_all = rc;
			err_all = last_error_str();
			break;
		}
		fold_totals(wk.w, h);
	}
	cudaSetDevice(c->device);
	for (int i = 1; i < nw; i++) {
		Ctx *h = ws[(size_t)i].w;
		c->stats.kernel_launches += h->stats.kernel_launches;
		c->stats.batches += h->stats.batches;
		c->stats.levels += h->stats.levels;
		c->stats.push_levels += h->stats.push_levels;
		c->stats.h2d_bytes += h->stats.h2d_bytes;
		c->stats.d2h_bytes += h->stats.d2h_bytes;
		c->stats.host_syncs += h->stats.host_syncs;
		h->stats = pgq_stats_t{};
	}
	if (rc_all != PGQ_OK) {
		set_error("%s", err_all.c_str());
		return rc_all;
	}
	return PGQ_OK;
}

// Device-side level loop for this call?  Building a worker's loop graph costs ~0.8 ms and pays back ~0.05 ms per
// batch (gpurun check 14, tools/graph_build_cost.py: a fresh CSR's first 1024-row call 3.2 ms with the graph against
// 1.65 ms with the host loop, later calls 1.50 against 1.59 ms).  DuckPGQ builds a CSR per query, so a small call on a
// fresh CSR takes the host-driven loop; calls with many batches, and CSRs that keep being queried, take the graph.
bool use_device_loop(Ctx *c, Csr *g, int64_t nbatch) {
	const bool yes = c->opt.loop && !c->loop_unsupported &&
	                 (nbatch >= c->opt.loop_min_batches || (int64_t)g->query_calls >= c->opt.loop_min_calls);
	g->query_calls++;
	c->opt.loop_now = yes ? 1 : 0; // the helper contexts copy the options per call (make_workers)
	return yes;
}

int pinned_found(Ctx *c, size_t bytes, int32_t **out) {
	if (bytes > c->h_found_cap) {
		if (c->h_found) cudaFreeHost(c->h_found);
		c->h_found = nullptr;
		c->h_found_cap = 0;
		size_t cap = 4096;
		while (cap < bytes) cap <<= 1;
		PGQ_CUDA(cudaHostAlloc(&c->h_found, cap, cudaHostAllocPortable));
		c->h_found_cap = cap;
	}
	*out = static_cast<int32_t *>(c->h_found);
	return PGQ_OK;
}

// ---- K8: one shortestpath batch: BFS with level history, then the paths of its rows in ROW order (R9) --
struct PathSeg {
	int64_t r0 = 0, r1 = 0;     // rows [r0, r1)
	std::vector<int64_t> vals;  // their list values, offsets in paths->entries are segment-local until stitched
};

// lanes [0, W) at d_src/d_dst/d_found (device) serve the rows lane_row[0..W) (host, -1 = unused) within
// [r0, r1).  When the level history of a W-lane batch would not fit its budget the batch is redone as
// W/64 batches of 64 lanes over slices of the same lane arrays (lanes are independent: same results).
int path_batch(Worker &wk, int W, const uint32_t *d_src, const uint32_t *d_dst, int32_t *d_found,
               const int64_t *lane_row, int64_t r0, int64_t r1, const uint8_t *kind, PathOut *paths,
               const int64_t *row_src, std::vector<PathSeg> *segs) {
	Ctx *w = wk.w;
	Csr *g = wk.g;
	cudaStream_t ws = w->stream;
	int nlev = 0;
	BatchBufs bw{};
	PGQ_TRY(alloc_bufs(w, g, W, &bw));
	int rc = run_batch(w, g, bw, W, true, d_src, d_dst, d_found, &nlev);
	if (rc == PGQ_HIST_OVER) {
		if (W <= 64) {
			set_error("shortestpath: the level history of one 64-lane batch (%d+ levels x %zu bytes) exceeds its budget "
			          "of %zu MiB; raise option hist_budget_mb", nlev, (size_t)g->V * 8, hist_budget(w) >> 20);
			return PGQ_ERR_OOM;
		}
		for (int k = 0; k < W / 64; k++) {
			// rows of this 64-lane group: from its first lane's row (group 0: from r0) up to the next group's
			int64_t a = r0, b = r1;
			if (k > 0) {
				a = -1;
				for (int l = k * 64; l < W && a < 0; l++) a = lane_row[l];
				if (a < 0) break; // no lanes left: the rows before were taken by the previous group
			}
			{
				int64_t nb = -1;
				for (int l = (k + 1) * 64; l < W && nb < 0; l++) nb = lane_row[l];
				if (nb >= 0) b = nb;
			}
			PGQ_TRY(path_batch(wk, 64, d_src + k * 64, d_dst + k * 64, d_found + k * 64, lane_row + k * 64, a, b, kind,
			                   paths, row_src, segs));
			if (b == r1) break;
		}
		return PGQ_OK;
	}
	PGQ_TRY(rc);
	std::vector<int32_t> hf((size_t)W);
	PGQ_CUDA(cudaMemcpyAsync(hf.data(), d_found, (size_t)W * 4, cudaMemcpyDeviceToHost, ws));
	PGQ_CUDA(cudaStreamSynchronize(ws));
	w->stats.host_syncs++;
	w->stats.d2h_bytes += (uint64_t)W * 4;
	std::vector<uint64_t> lane_off((size_t)W, 0);
	uint64_t bt = 0; // values of this segment so far
	int lane = 0;
	for (int64_t r = r0; r < r1; r++) {
		const uint8_t kd = kind[r];
		uint64_t len = 0;
		if (kd == 1) {
			len = 1;
		} else if (kd == 2) {
			while (lane < W && lane_row[lane] != r) lane++;
			if (lane < W && hf[(size_t)lane] >= 0) {
				len = paths->mode == 0 ? (uint64_t)hf[(size_t)lane] + 1 : 2ull * (uint64_t)hf[(size_t)lane] + 1;
				lane_off[(size_t)lane] = bt;
			}
		}
		paths->entries[r].offset = bt; // segment-local until stitched
		paths->entries[r].length = len;
		paths->valid[r] = len ? 1 : 0;
		bt += len;
	}
	segs->emplace_back();
	PathSeg &seg = segs->back();
	seg.r0 = r0;
	seg.r1 = r1;
	seg.vals.resize(bt);
	if (!bt) return PGQ_OK;
	// per-worker output buffer: [W lane offsets][values]; persistent, a per-call buffer would cudaMalloc every batch
	DevBuf &d_out = w->bfs_pathout;
	const size_t need = (size_t)W * 8 + bt * sizeof(int64_t);
	if (need > d_out.cap) PGQ_TRY(d_out.reserve(need * 2));
	unsigned long long *d_lo = d_out.as<unsigned long long>();
	int64_t *d_vals = reinterpret_cast<int64_t *>(d_lo + W);
	PGQ_CUDA(cudaMemcpyAsync(d_lo, lane_off.data(), (size_t)W * 8, cudaMemcpyHostToDevice, ws));
	bfs_paths_hist_kernel<<<W, 256, 0, ws>>>(w->bfs_hist_tab.as<const uint32_t *const>(), g->d_ioff.as<uint32_t>(),
	                                         g->d_isrc.as<uint32_t>(), g->d_ipos.as<uint32_t>(),
	                                         g->have_eid ? g->d_eid.as<int64_t>() : nullptr,
	                                         g->keep_rows ? g->d_row.as<uint32_t>() : nullptr, d_src, d_dst, d_found, d_lo,
	                                         W, paths->mode, d_vals);
	count_launch(w);
	PGQ_CUDA(cudaMemcpyAsync(seg.vals.data(), d_vals, bt * sizeof(int64_t), cudaMemcpyDeviceToHost, ws));
	PGQ_CUDA(cudaStreamSynchronize(ws));
	w->stats.host_syncs++;
	w->stats.d2h_bytes += bt * 8;
	// trivial rows: [src]
	for (int64_t r = r0; r < r1; r++)
		if (kind[r] == 1) seg.vals[paths->entries[r].offset] = row_src[r];
	return PGQ_OK;
}

} // namespace

// found[slot] for all lane slots of all batches (host).  Path output per batch when `paths`.
int msbfs_run(Ctx *c, Csr *g, const LaneTable &lt, int32_t *h_found, PathOut *paths, const int64_t *row_src,
              int64_t nrows) {
	(void)nrows;
	const int W = lt.W;
	const int64_t nslots = lt.nbatch * W;
	if (nslots == 0) return PGQ_OK;
	PGQ_TRY(csr_prepare(c, g));
	const bool want_paths = paths != nullptr;
	BatchBufs b{};
	PGQ_TRY(alloc_bufs(c, g, W, &b));
	cudaStream_t st = c->stream;
	// lane tables: src | dst | found
	PGQ_TRY(c->bfs_lane.reserve((size_t)nslots * 12 + 64));
	uint32_t *d_src = c->bfs_lane.as<uint32_t>();
	uint32_t *d_dst = d_src + nslots;
	int32_t *d_found = reinterpret_cast<int32_t *>(d_dst + nslots);
	PGQ_CUDA(cudaMemcpyAsync(d_src, lt.src.data(), (size_t)nslots * 4, cudaMemcpyHostToDevice, st));
	PGQ_CUDA(cudaMemcpyAsync(d_dst, lt.dst.data(), (size_t)nslots * 4, cudaMemcpyHostToDevice, st));
	c->stats.h2d_bytes += (uint64_t)nslots * 8;
	const bool dev_loop = use_device_loop(c, g, lt.nbatch);
	std::vector<Worker> ws;
	PGQ_TRY(make_workers(c, g, b, W, lt.nbatch, want_paths, d_src, d_dst, d_found, nslots, &ws));
	const bool host_loop = !dev_loop || c->opt.stats || c->opt.timing || c->opt.trace || c->loop_unsupported;
	if (!want_paths) {
		int32_t *stage = nullptr;
		PGQ_TRY(pinned_found(c, (size_t)nslots * 4, &stage));
		// one host thread deals all batches when they stay on one device; with several devices a thread per worker:
		// a graph launch blocks while its stream's queue is full, and a single dealer stuck on one device's stream
		// starved the other devices (measured: 0.90x on 2 GPUs with one dealer)
		bool multi_dev = false;
		for (const Worker &wk : ws) multi_dev |= wk.w->device != c->device;
		PGQ_TRY(run_batches(c, ws, lt.nbatch, (host_loop || multi_dev) && ws.size() > 1, [&](Worker &wk, int64_t bi) {
			PGQ_TRY(run_batch(wk.w, wk.g, wk.b, W, false, wk.d_src + bi * W, wk.d_dst + bi * W, wk.d_found + bi * W));
			PGQ_CUDA(cudaMemcpyAsync(stage + bi * W, wk.d_found + bi * W, (size_t)W * 4, cudaMemcpyDeviceToHost,
			                         wk.w->stream));
			// worker threads take batches dynamically; a batch is only ENQUEUED here, so without this wait the first
			// thread to start would queue nearly every batch on its own stream (measured: 0.87x on 2 GPUs)
			if (multi_dev && !host_loop) PGQ_CUDA(cudaStreamSynchronize(wk.w->stream));
			return (int)PGQ_OK;
		}));
		PGQ_CUDA(cudaStreamSynchronize(st));
		memcpy(h_found, stage, (size_t)nslots * 4);
		c->stats.d2h_bytes += (uint64_t)nslots * 4;
	} else {
		// each batch reconstructs its paths in row order with segment-local list offsets; the segments
		// (possibly finished out of order on different streams / devices) are stitched afterwards
		std::vector<std::vector<PathSeg>> segs((size_t)lt.nbatch);
		const int rc = run_batches(c, ws, lt.nbatch, ws.size() > 1, [&](Worker &wk, int64_t bi) {
			return path_batch(wk, W, wk.d_src + bi * W, wk.d_dst + bi * W, wk.d_found + bi * W, lt.row.data() + bi * W,
			                  lt.batch_row_begin[bi], lt.batch_row_end[bi], lt.kind.data(), paths, row_src,
			                  &segs[(size_t)bi]);
		});
		// level buffers beyond a working set are given back (a long-chain query must not pin HBM for good)
		for (Worker &wk : ws)
			if (wk.w->hist_bytes > ((size_t)2 << 30)) {
				cudaSetDevice(wk.w->device);
				cudaStreamSynchronize(wk.w->stream);
				hist_trim(wk.w);
			}
		cudaSetDevice(c->device);
		PGQ_TRY(rc);
		uint64_t total = 0; // running list offset over all rows
		for (auto &bs : segs)
			for (PathSeg &s : bs) total += s.vals.size();
		int64_t *dstp = nullptr;
		if (paths->sink) { // straight into the caller's buffer (the DuckDB shim: the LIST child vector)
			dstp = paths->sink(paths->sink_user, total);
			if (!dstp && total) {
				set_error("shortestpath: the caller's reserve callback returned no buffer for %llu values",
				          (unsigned long long)total);
				return PGQ_ERR_OOM;
			}
		} else {
			paths->vals.resize(total);
			dstp = paths->vals.data();
		}
		paths->total = total;
		total = 0;
		for (auto &bs : segs)
			for (PathSeg &s : bs) {
				for (int64_t r = s.r0; r < s.r1; r++) paths->entries[r].offset += total;
				if (!s.vals.empty()) memcpy(dstp + total, s.vals.data(), s.vals.size() * sizeof(int64_t));
				total += s.vals.size();
			}
	}
	PGQ_TRY(small_flush(c));
	return PGQ_OK;
}

// ---- 1-D vertex-partitioned variant (BASELINE.json configs[4]) ------------------------------------
// g is this rank's slice of the reverse CSR (local index space, global sources).  Frontier words
// and bitmap are replicated at G*Vp vertices; each level every rank fills its slice and one
// grouped ncclAllGather (words, bitmap, alive masks) completes all replicas.  Bottom-up only, the
// bitmap filter always on (so words outside the bitmap are never read).

namespace {

struct PartBufs {
	uint32_t *visit[2], *fbit[2], *seen, *hubacc, *alive_slots, *alive_local, *alive_all;
	uint32_t *const *peer_tab; // fused exchange: [2][G-1] peers' replicas offset to my slice, or nullptr
	BfsCtrl *ctrl;
	size_t words_per_slice, fwords_per_slice; // uint32 counts
};

template <int W>
int launch_part_level(Ctx *c, Csr *g, const PartBufs &b, int cur, int iter) {
	cudaStream_t st = c->stream;
	const int G = c->part_world, r = c->part_rank;
	const size_t fwords = b.fwords_per_slice * G;
	if (iter == 2 && !b.peer_tab) { // scrub the source words of level 0 out of the buffer that becomes `next` (zeroes the bitmap too)
		clear_by_bitmap_kernel<W><<<div_up(fwords, 256), 256, 0, st>>>(b.fbit[cur ^ 1], (uint32_t)fwords,
		                                                              b.visit[cur ^ 1]);
		count_launch(c);
	} else {
		PGQ_CUDA(cudaMemsetAsync(b.fbit[cur ^ 1], 0, fwords * sizeof(uint32_t), st));
	}
	PullArgs a{};
	a.ioff = g->d_ioff.as<uint32_t>();
	a.isrc = g->d_isrc.as<uint32_t>();
	a.visit = b.visit[cur];
	a.next = b.visit[cur ^ 1] + b.words_per_slice * r; // local vertex 0 = global vertex part_base
	a.seen = b.seen;
	a.ctrl = b.ctrl;
	a.V = (uint32_t)g->V;
	a.hubacc = b.hubacc;
	a.fcur = b.fbit[cur];
	a.fnext = b.fbit[cur ^ 1] + b.fwords_per_slice * r;
	a.alive_slots = b.alive_slots;
	a.dense = 1; // every owned vertex stores its word (the allgather ships the whole slice)
	if (b.peer_tab) { // fused exchange: kernels store new words into every peer's replica
		a.peer_next = b.peer_tab + (size_t)(cur ^ 1) * (G - 1);
		a.npeers = (uint32_t)(G - 1);
	}
	Launch L;
	L.st = st;
	launch_tiny<W>(a, g, true, L);
	launch_class<W, 2>(a, g->d_small, g->n_small, true, L);
	launch_warp<W>(a, g->d_med, g->n_med, true, L);
	launch_warp<W>(a, g->d_mid, g->n_mid, true, L);
	launch_hubs<W>(a, g, L);
	count_launch(c, L.nk);
	alive_fold_kernel<<<1, 1024, 0, st>>>(b.alive_slots, b.alive_local);
	count_launch(c);
	return PGQ_OK;
}

int launch_part_level_w(Ctx *c, Csr *g, const PartBufs &b, int W, int cur, int iter) {
	switch (W) {
	case 64: return launch_part_level<64>(c, g, b, cur, iter);
	case 128: return launch_part_level<128>(c, g, b, cur, iter);
	case 256: return launch_part_level<256>(c, g, b, cur, iter);
	case 512: return launch_part_level<512>(c, g, b, cur, iter);
	}
	return PGQ_ERR_INVALID;
}

// exchange of one level: next words, frontier bitmap, alive masks (in place, one NCCL group)
int part_exchange(Ctx *c, const PartBufs &b, int nxt) {
	const int G = c->part_world, r = c->part_rank;
	cudaStream_t st = c->stream;
	if (G == 1) {
		PGQ_CUDA(cudaMemcpyAsync(b.alive_all, b.alive_local, 16 * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
		return PGQ_OK;
	}
	const NcclApi *nc = nccl_api();
	if (!nc) return PGQ_ERR_CUDA;
	ncclComm_t comm = (ncclComm_t)c->nccl_comm;
	PGQ_NCCL(nc, nc->GroupStart());
	if (!b.peer_tab) {
		PGQ_NCCL(nc, nc->AllGather(b.visit[nxt] + b.words_per_slice * r, b.visit[nxt], b.words_per_slice, ncclUint32, comm, st));
		c->stats.coll_bytes += b.words_per_slice * sizeof(uint32_t);
	}
	PGQ_NCCL(nc, nc->AllGather(b.fbit[nxt] + b.fwords_per_slice * r, b.fbit[nxt], b.fwords_per_slice, ncclUint32, comm, st));
	PGQ_NCCL(nc, nc->AllGather(b.alive_local, b.alive_all, 16, ncclUint32, comm, st));
	PGQ_NCCL(nc, nc->GroupEnd());
	c->stats.coll_bytes += (b.fwords_per_slice + 16) * sizeof(uint32_t);
	return PGQ_OK;
}

} // namespace

int msbfs_part_run(Ctx *c, Csr *g, const LaneTable &lt, int32_t *h_found) {
	const int W = lt.W, G = c->part_world;
	const int64_t nslots = lt.nbatch * W;
	if (nslots == 0) return PGQ_OK;
	cudaStream_t st = c->stream;
	const size_t Vp = (size_t)g->V, WPV = (size_t)W / 32;
	PartBufs b{};
	b.words_per_slice = Vp * WPV;
	b.fwords_per_slice = Vp / 32; // Vp is a multiple of 256
	const size_t bm = b.words_per_slice * G * sizeof(uint32_t), fb = b.fwords_per_slice * G * sizeof(uint32_t);
	const bool p2p = G > 1 && c->part_p2p && c->part_p2p_W == W && c->part_p2p_Vp == g->V;
	if (!p2p) {
		PGQ_TRY(c->bfs_visit[0].reserve(bm));
		PGQ_TRY(c->bfs_visit[1].reserve(bm));
	}
	PGQ_TRY(c->bfs_seen.reserve(b.words_per_slice * sizeof(uint32_t)));
	PGQ_TRY(c->bfs_inq[0].reserve(2 * fb));
	PGQ_TRY(c->bfs_ctrl.reserve(sizeof(BfsCtrl)));
	PGQ_TRY(c->bfs_alive.reserve((size_t)ALIVE_SLOTS * 16 * sizeof(uint32_t)));
	PGQ_TRY(c->part_alive.reserve((size_t)(1 + G) * 16 * sizeof(uint32_t)));
	const size_t hb = (size_t)(g->n_hub > 0 ? g->n_hub : 1) * (W / 8);
	PGQ_TRY(c->bfs_hubacc.reserve(hb));
	PGQ_CUDA(cudaMemsetAsync(c->bfs_hubacc.p, 0, c->bfs_hubacc.cap, st));
	PGQ_CUDA(cudaMemsetAsync(c->bfs_alive.p, 0, (size_t)ALIVE_SLOTS * 16 * sizeof(uint32_t), st));
	c->bfs_dirty.valid = false; // the shared frontier buffers are used with another geometry here
	b.visit[0] = p2p ? c->part_visit[0].as<uint32_t>() : c->bfs_visit[0].as<uint32_t>();
	b.visit[1] = p2p ? c->part_visit[1].as<uint32_t>() : c->bfs_visit[1].as<uint32_t>();
	b.peer_tab = p2p ? c->part_peer_tab.as<uint32_t *const>() : nullptr;
	b.seen = c->bfs_seen.as<uint32_t>();
	b.fbit[0] = c->bfs_inq[0].as<uint32_t>();
	b.fbit[1] = b.fbit[0] + b.fwords_per_slice * G;
	b.hubacc = c->bfs_hubacc.as<uint32_t>();
	b.alive_slots = c->bfs_alive.as<uint32_t>();
	b.alive_local = c->part_alive.as<uint32_t>();
	b.alive_all = b.alive_local + 16;
	b.ctrl = c->bfs_ctrl.as<BfsCtrl>();
	PGQ_TRY(c->bfs_lane.reserve((size_t)nslots * 12 + 64));
	uint32_t *d_src = c->bfs_lane.as<uint32_t>();
	uint32_t *d_dst = d_src + nslots;
	int32_t *d_found = reinterpret_cast<int32_t *>(d_dst + nslots);
	PGQ_CUDA(cudaMemcpyAsync(d_src, lt.src.data(), (size_t)nslots * 4, cudaMemcpyHostToDevice, st));
	PGQ_CUDA(cudaMemcpyAsync(d_dst, lt.dst.data(), (size_t)nslots * 4, cudaMemcpyHostToDevice, st));
	c->stats.h2d_bytes += (uint64_t)nslots * 8;
	for (int64_t bi = 0; bi < lt.nbatch; bi++) {
		PGQ_CUDA(cudaMemsetAsync(b.seen, 0, b.words_per_slice * sizeof(uint32_t), st));
		if (!p2p) { // with the fused exchange no word outside the bitmap is ever read: nothing to clear
			PGQ_CUDA(cudaMemsetAsync(b.visit[0], 0, bm, st));
			PGQ_CUDA(cudaMemsetAsync(b.visit[1], 0, bm, st));
		}
		PGQ_CUDA(cudaMemsetAsync(b.fbit[0], 0, 2 * fb, st));
		bfs_init_part_kernel<<<1, W, 0, st>>>(d_src + bi * W, W, b.visit[0], d_found + bi * W, b.ctrl, b.fbit[0]);
		count_launch(c);
		BfsCtrl h;
		PGQ_CUDA(cudaMemcpyAsync(c->h_ctrl, b.ctrl, sizeof(BfsCtrl), cudaMemcpyDeviceToHost, st));
		PGQ_CUDA(cudaStreamSynchronize(st));
		h = *reinterpret_cast<BfsCtrl *>(c->h_ctrl);
		c->stats.batches++;
		uint32_t active = h.n_active;
		int cur = 0;
		for (int iter = 1; active; iter++) {
			if (c->opt.stats && g->E > 0) { // TEPS numerator of this level (untimed instrumented pass)
				const uint32_t ne = (uint32_t)g->E, grid = div_up(ne, 256);
				const uint32_t *isrc = g->d_isrc.as<uint32_t>();
				switch (W) {
				case 64: part_count_kernel<64><<<grid, 256, 0, st>>>(isrc, ne, b.visit[cur], b.fbit[cur], b.ctrl); break;
				case 128: part_count_kernel<128><<<grid, 256, 0, st>>>(isrc, ne, b.visit[cur], b.fbit[cur], b.ctrl); break;
				case 256: part_count_kernel<256><<<grid, 256, 0, st>>>(isrc, ne, b.visit[cur], b.fbit[cur], b.ctrl); break;
				default: part_count_kernel<512><<<grid, 256, 0, st>>>(isrc, ne, b.visit[cur], b.fbit[cur], b.ctrl); break;
				}
				count_launch(c);
			}
			if (c->opt.timing) PGQ_CUDA(cudaEventRecord(c->ev_a, st));
			uint64_t before = c->stats.kernel_launches;
			PGQ_TRY(launch_part_level_w(c, g, b, W, cur, iter));
			if (c->opt.timing) PGQ_CUDA(cudaEventRecord(c->ev_b, st));
			PGQ_TRY(part_exchange(c, b, cur ^ 1));
			if (c->opt.timing) PGQ_CUDA(cudaEventRecord(c->ev_c, st));
			bfs_check_part_kernel<<<1, W, 0, st>>>(b.visit[cur ^ 1], d_dst + bi * W, W, d_found + bi * W, b.ctrl, iter,
			                                       b.alive_all, G, b.fbit[cur ^ 1]);
			count_launch(c);
			PGQ_CUDA(cudaMemcpyAsync(c->h_ctrl, b.ctrl, sizeof(BfsCtrl), cudaMemcpyDeviceToHost, st));
			PGQ_CUDA(cudaStreamSynchronize(st));
			h = *reinterpret_cast<BfsCtrl *>(c->h_ctrl);
			c->stats.levels++;
			if (c->opt.timing) {
				float ms = 0, cms = 0;
				PGQ_CUDA(cudaEventElapsedTime(&ms, c->ev_a, c->ev_b));
				PGQ_CUDA(cudaEventElapsedTime(&cms, c->ev_b, c->ev_c));
				c->stats.level_kernel_ms += ms;
				c->stats.coll_ms += cms;
				c->stats.level_kernel_launches += c->stats.kernel_launches - before;
				if (c->opt.trace)
					fprintf(stderr, "[pgq part r%d] iter %d pull %.3f ms  exchange %.3f ms  active %u alive %u\n",
					        c->part_rank, iter, ms, cms, h.n_active, h.change_out);
			}
			cur ^= 1;
			if (!h.change_out) break;
			active = h.n_active;
		}
		if (c->opt.stats) c->stats.edges_lane += h.edges_lane; // accumulated in the control block over the batch
	}
	PGQ_CUDA(cudaMemcpyAsync(h_found, d_found, (size_t)nslots * 4, cudaMemcpyDeviceToHost, st));
	PGQ_CUDA(cudaStreamSynchronize(st));
	c->stats.d2h_bytes += (uint64_t)nslots * 4;
	PGQ_CUDA(cudaGetLastError());
	return PGQ_OK;
}

// ---- device-resident variant: slot i <-> row i -------------------------------------------------

__global__ void __launch_bounds__(256) lanes_from_rows_kernel(const int64_t *__restrict__ src,
                                                              const int64_t *__restrict__ dst, int64_t n,
                                                              int64_t nslots, uint64_t V,
                                                              uint32_t *__restrict__ lane_src,
                                                              uint32_t *__restrict__ lane_dst, uint32_t *err) {
	int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= nslots) return;
	uint32_t s = 0xFFFFFFFFu, d = 0;
	if (i < n) {
		uint64_t a = (uint64_t)src[i], b = (uint64_t)dst[i];
		if (a >= V || b >= V) atomicOr(err, 1u);
		else if (a != b) {
			s = (uint32_t)a;
			d = (uint32_t)b;
		}
	}
	lane_src[i] = s;
	lane_dst[i] = d;
}

__global__ void __launch_bounds__(256) rows_from_lanes_kernel(const int64_t *__restrict__ src,
                                                              const int64_t *__restrict__ dst, int64_t n,
                                                              const int32_t *__restrict__ found,
                                                              int64_t *__restrict__ out, uint8_t *__restrict__ valid) {
	int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	if (src[i] == dst[i]) {
		out[i] = 0;
		valid[i] = 1;
	} else {
		int32_t f = found[i];
		out[i] = f >= 0 ? (int64_t)f : -1;
		valid[i] = f >= 0;
	}
}

int msbfs_run_device(Ctx *c, Csr *g, int64_t n, const int64_t *d_srcs, const int64_t *d_dsts, int64_t *d_out,
                     uint8_t *d_out_valid) {
	if (n == 0) return PGQ_OK;
	const int W = c->opt.lanes ? (int)c->opt.lanes : pick_lane_width(n);
	const int64_t nbatch = (n + W - 1) / W, nslots = nbatch * W;
	PGQ_TRY(csr_prepare(c, g));
	BatchBufs b{};
	PGQ_TRY(alloc_bufs(c, g, W, &b));
	cudaStream_t st = c->stream;
	PGQ_TRY(c->bfs_lane.reserve((size_t)nslots * 12 + 64));
	uint32_t *d_src = c->bfs_lane.as<uint32_t>() + 4; // [0] = error flag
	uint32_t *d_err = c->bfs_lane.as<uint32_t>();
	uint32_t *d_dst = d_src + nslots;
	int32_t *d_found = reinterpret_cast<int32_t *>(d_dst + nslots);
	PGQ_CUDA(cudaMemsetAsync(d_err, 0, 16, st));
	lanes_from_rows_kernel<<<div_up(nslots, 256), 256, 0, st>>>(d_srcs, d_dsts, n, nslots, (uint64_t)g->V, d_src,
	                                                            d_dst, d_err);
	count_launch(c);
	const bool dev_loop = use_device_loop(c, g, nbatch);
	std::vector<Worker> ws;
	PGQ_TRY(make_workers(c, g, b, W, nbatch, false, d_src, d_dst, d_found, nslots, &ws));
	const bool host_loop = !dev_loop || c->opt.stats || c->opt.timing || c->opt.trace || c->loop_unsupported;
	bool multi_dev = false;
	for (const Worker &wk : ws) multi_dev |= wk.w->device != c->device;
	int32_t *stage = nullptr;
	if (multi_dev) PGQ_TRY(pinned_found(c, (size_t)nslots * 4, &stage));
	PGQ_TRY(run_batches(c, ws, nbatch, (host_loop || multi_dev) && ws.size() > 1, [&](Worker &wk, int64_t bi) {
		PGQ_TRY(run_batch(wk.w, wk.g, wk.b, W, false, wk.d_src + bi * W, wk.d_dst + bi * W, wk.d_found + bi * W));
		if (multi_dev) { // results of every device meet in pinned host memory, then return to the primary
			PGQ_CUDA(cudaMemcpyAsync(stage + bi * W, wk.d_found + bi * W, (size_t)W * 4, cudaMemcpyDeviceToHost,
			                         wk.w->stream));
			if (!host_loop) PGQ_CUDA(cudaStreamSynchronize(wk.w->stream)); // dynamic hand-out: see msbfs_run
		}
		return (int)PGQ_OK;
	}));
	if (multi_dev) PGQ_CUDA(cudaMemcpyAsync(d_found, stage, (size_t)nslots * 4, cudaMemcpyHostToDevice, st));
	// helper streams of this device are complete (run_batches waited); order the result kernel after them
	rows_from_lanes_kernel<<<div_up(n, 256), 256, 0, st>>>(d_srcs, d_dsts, n, d_found, d_out, d_out_valid);
	count_launch(c);
	uint32_t h_err = 0;
	PGQ_CUDA(cudaMemcpyAsync(&h_err, d_err, 4, cudaMemcpyDeviceToHost, st));
	PGQ_CUDA(cudaStreamSynchronize(st));
	PGQ_TRY(small_flush(c));
	if (h_err) {
		set_error("iterativelength: src/dst id outside [0, v_size)");
		return PGQ_ERR_INVALID;
	}
	return PGQ_OK;
}

} // namespace pgq
