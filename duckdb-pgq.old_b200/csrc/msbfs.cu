// Host driver of the bitset MS-BFS (iterativelength / reachability / shortestpath batches).
// Mirrors the reference's batch loop: clear seen/visit, seed the lanes,
// run level steps while a lane is active and something changed, check destinations after each
// level, retire found lanes.  One stream, one host sync per level (reads 16 bytes of control).
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
	BfsCtrl *ctrl;
};

template <int W, int GPV>
void launch_class(Ctx *c, PullArgs &a, const DevBuf &list, uint32_t nlist, bool filter) {
	if (!nlist) return;
	using C = LaneCfg<W>;
	constexpr int VPW = C::G / GPV;
	a.list = list.as<uint32_t>();
	a.nlist = nlist;
	const uint32_t warps = div_up(nlist, VPW);
	const uint32_t grid = div_up((uint64_t)warps * 32, 128);
	if (filter) pull_list_kernel<W, GPV, true><<<grid, 128, 0, c->stream>>>(a);
	else pull_list_kernel<W, GPV, false><<<grid, 128, 0, c->stream>>>(a);
	count_launch(c);
}

// filtered levels, small in-degrees: SEG lanes per vertex, edge-parallel frontier test
template <int W, int SEG>
void launch_seg(Ctx *c, PullArgs &a, const DevBuf &list, uint32_t nlist) {
	if (!nlist) return;
	a.list = list.as<uint32_t>();
	a.nlist = nlist;
	const uint32_t warps = div_up(nlist, 32 / SEG);
	pull_seg_kernel<W, SEG><<<div_up((uint64_t)warps * 32, 128), 128, 0, c->stream>>>(a);
	count_launch(c);
}

template <int W>
void launch_warp(Ctx *c, PullArgs &a, const DevBuf &list, uint32_t nlist, bool filter) {
	if (!nlist) return;
	a.list = list.as<uint32_t>();
	a.nlist = nlist;
	const uint32_t grid = div_up((uint64_t)nlist * 32, 128);
	if (filter) pull_warp_kernel<W, true><<<grid, 128, 0, c->stream>>>(a);
	else pull_warp_kernel<W, false><<<grid, 128, 0, c->stream>>>(a);
	count_launch(c);
}

// tiny class: straight-line kernel when the class really is <= 4 in-edges, else the generic one
template <int W>
void launch_tiny(Ctx *c, PullArgs &a, Csr *g, bool filter) {
	if (!g->n_tiny) return;
	if (g->bins_tiny_max > 4) {
		launch_class<W, 1>(c, a, g->d_tiny, g->n_tiny, filter);
		return;
	}
	using C = LaneCfg<W>;
	a.list = g->d_tiny.as<uint32_t>();
	a.nlist = g->n_tiny;
	const uint32_t grid = div_up((uint64_t)div_up(g->n_tiny, C::G) * 32, 128);
	if (filter) pull_tiny_kernel<W, true><<<grid, 128, 0, c->stream>>>(a);
	else pull_tiny_kernel<W, false><<<grid, 128, 0, c->stream>>>(a);
	count_launch(c);
}

// One bottom-up level: frontier visit[cur]/fbit[cur] -> next visit[cur^1]/fbit[cur^1].
template <int W>
int launch_level(Ctx *c, Csr *g, const BatchBufs &b, int cur, const int *fb, int iter, bool filter, bool early) {
	using C = LaneCfg<W>;
	cudaStream_t st = c->stream;
	// fb[0] describes visit[cur], fb[1] is built for visit[cur^1], fb[2] describes what visit[cur^1]
	// still holds (the frontier of two levels ago).
	// fnext is built with atomicOr: start from zero.  At level 2 `next` still holds the source words
	// (level-0 frontier), some at in-degree-0 vertices that no class rewrites: clear them by bitmap.
	PGQ_CUDA(cudaMemsetAsync(b.fbit[fb[1]], 0, b.fbit_bytes, st));
	if (iter == 2) {
		clear_by_bitmap_kernel<W><<<div_up(b.fbit_words, 256), 256, 0, st>>>(b.fbit[fb[2]], b.fbit_words,
		                                                                    b.visit[cur ^ 1]);
		count_launch(c);
	}
	PullArgs a{};
	a.ioff = g->d_ioff.as<uint32_t>();
	a.isrc = g->d_isrc.as<uint32_t>();
	a.visit = b.visit[cur];
	a.next = b.visit[cur ^ 1];
	a.seen = b.seen;
	a.ctrl = b.ctrl;
	a.V = (uint32_t)g->V;
	a.hubacc = b.hubacc;
	a.off = g->d_off.as<int64_t>();
	a.fcur = b.fbit[fb[0]];
	a.fnext = b.fbit[fb[1]];
	a.alive_slots = b.alive_slots;
	launch_tiny<W>(c, a, g, filter);
	// measured on the filtered levels of R-MAT s24 (levels 2 / 6): pull_list 2.23 / 0.62 ms, 8 lanes per
	// vertex 2.08 / 0.56 ms, 16 lanes 2.31 / 0.73 ms; the same kernel for the tiny class: no change
	if (filter && c->opt.seg_small) launch_seg<W, 8>(c, a, g->d_small, g->n_small);
	else launch_class<W, 2>(c, a, g->d_small, g->n_small, filter);
	// medium in-degrees: the edge-parallel warp scan wins while most vertices still scan their whole
	// in-list against a sparse frontier (early levels); two thread groups per vertex win when every
	// word is gathered anyway (dense levels) or most vertices have nothing left to find (tail)
	if (filter && early) launch_warp<W>(c, a, g->d_med, g->n_med, true);
	else launch_class<W, 2>(c, a, g->d_med, g->n_med, filter);
	launch_warp<W>(c, a, g->d_mid, g->n_mid, filter);
	if (g->n_hub) {
		a.list = g->d_hub.as<uint32_t>();
		a.nlist = g->n_hub;
		a.chunks = g->d_hubchunks.as<HubChunk>();
		a.nchunks = g->n_hubchunks;
		pull_hub_kernel<W><<<g->n_hubchunks, 256, 0, st>>>(a);
		hub_finalize_kernel<W><<<div_up((uint64_t)g->n_hub * C::TPV, 256), 256, 0, st>>>(a);
		count_launch(c, 2);
	}
	frontier_stats_kernel<<<div_up(b.fbit_words, 256), 256, 0, st>>>(b.fbit[fb[1]], b.fbit_words, (uint32_t)g->V,
	                                                               g->d_off.as<int64_t>(), b.ctrl);
	count_launch(c);
	return PGQ_OK;
}

int launch_level_w(Ctx *c, Csr *g, const BatchBufs &b, int W, int cur, const int *fb, int iter, bool filter,
                   bool early) {
	switch (W) {
	case 64: return launch_level<64>(c, g, b, cur, fb, iter, filter, early);
	case 128: return launch_level<128>(c, g, b, cur, fb, iter, filter, early);
	case 256: return launch_level<256>(c, g, b, cur, fb, iter, filter, early);
	case 512: return launch_level<512>(c, g, b, cur, fb, iter, filter, early);
	}
	set_error("unsupported lane width %d", W);
	return PGQ_ERR_INVALID;
}

// One top-down level over queue[q] (frontier in visit[cur]); builds queue[q^1] and fbit[cur^1].
template <int W>
int launch_push(Ctx *c, Csr *g, const BatchBufs &b, int cur, const int *fb, int q) {
	PushArgs a{};
	a.off = g->d_off.as<int64_t>();
	a.tgt = g->d_tgt.as<uint32_t>();
	a.visit = b.visit[cur];
	a.next = b.visit[cur ^ 1];
	a.seen = b.seen;
	a.ctrl = b.ctrl;
	a.queue = b.queue[q];
	a.queue_next = b.queue[q ^ 1];
	a.fnext = b.fbit[fb[1]];
	// `next` still holds the frontier of two levels ago, described by fb[2]: zero exactly those words
	PGQ_CUDA(cudaMemsetAsync(b.fbit[fb[1]], 0, b.fbit_bytes, c->stream));
	clear_by_bitmap_kernel<W><<<div_up(b.fbit_words, 256), 256, 0, c->stream>>>(b.fbit[fb[2]], b.fbit_words,
	                                                                           b.visit[cur ^ 1]);
	const uint32_t grid = (uint32_t)c->sm_count * 8;
	a.alive_slots = b.alive_slots;
	push_kernel<W><<<grid, 256, 0, c->stream>>>(a);
	count_launch(c, 2);
	return PGQ_OK;
}

int launch_push_w(Ctx *c, Csr *g, const BatchBufs &b, int W, int cur, const int *fb, int q) {
	switch (W) {
	case 64: return launch_push<64>(c, g, b, cur, fb, q);
	case 128: return launch_push<128>(c, g, b, cur, fb, q);
	case 256: return launch_push<256>(c, g, b, cur, fb, q);
	case 512: return launch_push<512>(c, g, b, cur, fb, q);
	}
	return PGQ_ERR_INVALID;
}

// zero the words of visit[which] named by bitmap fbit[bi] (and the bitmap)
int launch_clear_w(Ctx *c, const BatchBufs &b, int W, int bi, int which) {
	const uint32_t grid = div_up(b.fbit_words, 256);
	switch (W) {
	case 64: clear_by_bitmap_kernel<64><<<grid, 256, 0, c->stream>>>(b.fbit[bi], b.fbit_words, b.visit[which]); break;
	case 128: clear_by_bitmap_kernel<128><<<grid, 256, 0, c->stream>>>(b.fbit[bi], b.fbit_words, b.visit[which]); break;
	case 256: clear_by_bitmap_kernel<256><<<grid, 256, 0, c->stream>>>(b.fbit[bi], b.fbit_words, b.visit[which]); break;
	case 512: clear_by_bitmap_kernel<512><<<grid, 256, 0, c->stream>>>(b.fbit[bi], b.fbit_words, b.visit[which]); break;
	default: return PGQ_ERR_INVALID;
	}
	count_launch(c);
	return PGQ_OK;
}

// pull -> push transition: queue[q] from the frontier bitmap fbit[cur].
int launch_compact(Ctx *c, Csr *g, const BatchBufs &b, const int *fb, int q) {
	cudaStream_t st = c->stream;
	compact_bitmap_kernel<<<div_up(b.fbit_words, 256), 256, 0, st>>>(b.fbit[fb[0]], b.fbit_words, (uint32_t)g->V,
	                                                                 g->d_off.as<int64_t>(), b.queue[q], b.ctrl);
	bfs_snapshot_items_kernel<<<1, 1, 0, st>>>(b.ctrl);
	count_launch(c, 2);
	return PGQ_OK;
}

void launch_count(Ctx *c, Csr *g, const BatchBufs &b, int W, int cur) {
	uint32_t grid = div_up(g->V, 256);
	if (!grid) return;
	const int64_t *off = g->d_off.as<int64_t>();
	switch (W) {
	case 64: bfs_count_kernel<64><<<grid, 256, 0, c->stream>>>(b.visit[cur], off, (uint32_t)g->V, b.ctrl); break;
	case 128: bfs_count_kernel<128><<<grid, 256, 0, c->stream>>>(b.visit[cur], off, (uint32_t)g->V, b.ctrl); break;
	case 256: bfs_count_kernel<256><<<grid, 256, 0, c->stream>>>(b.visit[cur], off, (uint32_t)g->V, b.ctrl); break;
	case 512: bfs_count_kernel<512><<<grid, 256, 0, c->stream>>>(b.visit[cur], off, (uint32_t)g->V, b.ctrl); break;
	}
	count_launch(c);
}

int read_ctrl(Ctx *c, const BatchBufs &b, BfsCtrl *h) {
	PGQ_CUDA(cudaMemcpyAsync(c->h_ctrl, b.ctrl, sizeof(BfsCtrl), cudaMemcpyDeviceToHost, c->stream));
	PGQ_CUDA(cudaStreamSynchronize(c->stream));
	*h = *reinterpret_cast<BfsCtrl *>(c->h_ctrl);
	return PGQ_OK;
}

int alloc_bufs(Ctx *c, Csr *g, int W, BatchBufs *b) {
	const size_t bm = (size_t)(g->V > 0 ? g->V : 1) * (W / 8);
	PGQ_TRY(c->bfs_seen.reserve(bm));
	PGQ_TRY(c->bfs_visit[0].reserve(bm));
	PGQ_TRY(c->bfs_visit[1].reserve(bm));
	PGQ_TRY(c->bfs_ctrl.reserve(sizeof(BfsCtrl)));
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
	return PGQ_OK;
}

// A zeroed frontier buffer for level `k` of a batch that keeps its history (shortestpath).
int hist_level(Ctx *c, size_t k, size_t bm, uint32_t **out) {
	if (c->bfs_hist.size() <= k) c->bfs_hist.resize(k + 1);
	PGQ_TRY(c->bfs_hist[k].reserve(bm ? bm : 16));
	if (bm) PGQ_CUDA(cudaMemsetAsync(c->bfs_hist[k].p, 0, bm, c->stream));
	*out = c->bfs_hist[k].as<uint32_t>();
	return PGQ_OK;
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
		if (c->bfs_hist.size() < (size_t)SMALL_HIST_LEVELS + 1) c->bfs_hist.resize((size_t)SMALL_HIST_LEVELS + 1);
		std::vector<uint32_t *> hp((size_t)SMALL_HIST_LEVELS + 1);
		for (int k = 0; k <= SMALL_HIST_LEVELS; k++) {
			PGQ_TRY(c->bfs_hist[(size_t)k].reserve(bm));
			hp[(size_t)k] = c->bfs_hist[(size_t)k].as<uint32_t>();
		}
		PGQ_TRY(c->bfs_hist_tab.reserve(hp.size() * sizeof(uint32_t *)));
		PGQ_CUDA(cudaMemcpyAsync(c->bfs_hist_tab.p, hp.data(), hp.size() * sizeof(uint32_t *), cudaMemcpyHostToDevice,
		                         st));
		PGQ_CUDA(cudaStreamSynchronize(st)); // hp is a host temporary
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
	PGQ_CUDA(cudaMemsetAsync(&b.ctrl->small_bar, 0, sizeof(BfsCtrl) - offsetof(BfsCtrl, small_bar), st));
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

// One batch of W lanes.  d_lane_src/dst/found point at this batch's W slots.
// keep_hist (shortestpath): level k's frontier stays in c->bfs_hist[k] instead of a ping-pong pair, so
// the path kernel can apply the parent rule afterwards; *out_levels = levels run.
int run_batch(Ctx *c, Csr *g, const BatchBufs &b_in, int W, bool keep_hist, const uint32_t *d_lane_src,
              const uint32_t *d_lane_dst, int32_t *d_found, int *out_levels = nullptr) {
	cudaStream_t st = c->stream;
	BatchBufs b = b_in;
	const size_t bm = (size_t)g->V * (W / 8);
	if (small_eligible(c, g, W, keep_hist)) {
		const int rc = run_batch_small(c, g, b, W, keep_hist, d_lane_src, d_lane_dst, d_found, out_levels);
		if (rc != PGQ_SMALL_OVERFLOW) return rc;
	}
	if (keep_hist) {
		if (bm) PGQ_CUDA(cudaMemsetAsync(b.seen, 0, bm, st));
		PGQ_TRY(hist_level(c, 0, bm, &b.visit[0]));
		b.visit[1] = nullptr; // set per level
		c->bfs_dirty.valid = false;
	} else if (bm) {
		PGQ_CUDA(cudaMemsetAsync(b.seen, 0, bm, st));
		// The frontier buffers must start all-zero (a top-down level ORs into them).  After a clean
		// batch on the same buffers their only non-zero words are the ones the last two frontier
		// bitmaps name: scrub those instead of streaming 2 x V x W/8 bytes of zeros.
		BfsDirty &d = c->bfs_dirty;
		if (d.valid && d.W == W && d.V == g->V && d.visit0 == b.visit[0] && d.fbit0 == b.fbit[0]) {
			PGQ_TRY(launch_clear_w(c, b, W, d.map[0], 0));
			PGQ_TRY(launch_clear_w(c, b, W, d.map[1], 1));
		} else {
			PGQ_CUDA(cudaMemsetAsync(b.visit[0], 0, bm, st));
			PGQ_CUDA(cudaMemsetAsync(b.visit[1], 0, bm, st));
		}
		d.valid = false; // until this batch completes
	}
	PGQ_CUDA(cudaMemsetAsync(b.fbit[0], 0, 3 * b.fbit_bytes, st)); // all three frontier bitmaps
	bfs_init_kernel<<<1, W, 0, st>>>(d_lane_src, W, b.visit[0], d_found, b.ctrl, g->d_off.as<int64_t>(), b.queue[0],
	                                 b.fbit[0]);
	count_launch(c);
	BfsCtrl h;
	PGQ_TRY(read_ctrl(c, b, &h));
	c->stats.batches++;
	const uint64_t levels_at_start = c->stats.levels;
	uint32_t active = h.n_active;
	int cur = 0, q = 0;
	unsigned long long cum_verts = h.out_verts; // frontier vertices of all levels so far
	int fb[3] = {0, 1, 2}; // bitmap of visit[cur] / being built for visit[cur^1] / of what visit[cur^1] holds
	bool have_queue = true; // queue[q] describes visit[cur] and visit[cur^1] is all-zero
	for (int iter = 1; active && iter <= (int)c->opt.level_cap; iter++) {
		// direction switch on frontier density: top-down while the frontier's out-edges are few
		if (keep_hist) { // fresh zeroed buffer for this level's output; the input stays where it is
			b.visit[cur] = c->bfs_hist[(size_t)iter - 1].as<uint32_t>();
			PGQ_TRY(hist_level(c, (size_t)iter, bm, &b.visit[cur ^ 1]));
		}
		const bool push = c->opt.direction == 0 &&
		                  (unsigned long long)h.out_edges * (unsigned long long)c->opt.push_beta <=
		                      (unsigned long long)g->E &&
		                  (unsigned long long)h.out_verts * (unsigned long long)c->opt.push_vbeta <=
		                      (unsigned long long)g->V + (unsigned long long)c->opt.push_vbeta * 1024ull;
		if (c->opt.stats) launch_count(c, g, b, W, cur);
		if (c->opt.timing) PGQ_CUDA(cudaEventRecord(c->ev_a, st));
		uint64_t before = c->stats.kernel_launches;
		if (push) {
			if (!have_queue) PGQ_TRY(launch_compact(c, g, b, fb, q));
			PGQ_TRY(launch_push_w(c, g, b, W, cur, fb, q));
			have_queue = true;
			q ^= 1;
			c->stats.push_levels++;
		} else {
			// the bitmap test is dead weight once the frontier holds most of the edges
			const bool filter = (unsigned long long)h.out_edges * (unsigned long long)c->opt.filter_gamma <=
			                    (unsigned long long)g->E;
			// "early": fewer vertex discoveries so far than vertices that can be discovered at all
			const bool early = cum_verts < (unsigned long long)g->n_tiny + g->n_small + g->n_med + g->n_mid + g->n_hub;
			PGQ_TRY(launch_level_w(c, g, b, W, cur, fb, iter, filter, early));
			have_queue = false;
		}
		if (c->opt.timing) PGQ_CUDA(cudaEventRecord(c->ev_b, st));
		bfs_check_kernel<<<1, W, 0, st>>>(b.seen, d_lane_dst, W, d_found, b.ctrl, iter, b.alive_slots);
		count_launch(c);
		PGQ_TRY(read_ctrl(c, b, &h));
		c->stats.levels++;
		if (c->opt.timing) {
			float ms = 0;
			PGQ_CUDA(cudaEventElapsedTime(&ms, c->ev_a, c->ev_b));
			c->stats.level_kernel_ms += ms;
			c->stats.level_kernel_launches += c->stats.kernel_launches - 1 - before;
			if (c->opt.trace)
				fprintf(stderr, "[pgq] iter %d %s %.3f ms  next: verts %u edges %llu items %u  active %u change %u\n",
				        iter, push ? "push" : "pull", ms, h.out_verts, (unsigned long long)h.out_edges, h.out_items,
				        h.n_active, h.change_out);
		}
		// the level's output becomes the current frontier
		cur ^= 1;
		{
			const int t = fb[0];
			fb[0] = fb[1];
			fb[1] = fb[2];
			fb[2] = t;
		}
		if (!h.change_out) break;
		active = h.n_active;
		cum_verts += h.out_verts;
	}
	if (out_levels) *out_levels = (int)(c->stats.levels - levels_at_start);
	if (bm && !keep_hist) { // what the frontier buffers hold now, for the next batch's scrub
		BfsDirty &d = c->bfs_dirty;
		d.valid = true;
		d.W = W;
		d.V = g->V;
		d.visit0 = b.visit[0];
		d.fbit0 = b.fbit[0];
		d.map[cur] = fb[0];
		d.map[cur ^ 1] = fb[2];
	}
	if (c->opt.stats) {
		// counters accumulate in the control block over the batch (reset by the init kernel)
		PGQ_TRY(read_ctrl(c, b, &h));
		c->stats.edges_lane += h.edges_lane;
		c->stats.edges_unique += h.edges_unique;
		// 4*V*(W/8) per level + E_act*(4 + W/8) + 8*V_act
		c->stats.algo_bytes += h.edges_unique * (uint64_t)(4 + W / 8) + 8ull * h.verts_active;
	}
	PGQ_CUDA(cudaGetLastError());
	return PGQ_OK;
}

// All batches of one call (lengths / reachability).  A batch on the general path meets the host once
// per level and leaves the GPU idle across launch gaps, memsets and kernel tails (about 12 % of a
// batch on R-MAT s24, most of it on mid-sized graphs), and batches are independent: so up to
// `streams` of them are in flight at once, each on a helper context with its own stream, control
// block and frontier buffers, driven by its own host thread.  The CSR is shared read-only.
using BatchFn = std::function<int(Ctx *, const BatchBufs &, int64_t)>; // (worker context, its buffers, batch index)

int run_batches(Ctx *c, Csr *g, const BatchBufs &b0, int W, int64_t nbatch, bool keep_hist, const BatchFn &fn) {
	int nw = (int)std::min<int64_t>(c->opt.streams, nbatch);
	if (nw < 2 || c->opt.timing || c->opt.stats || c->opt.trace || g->is_part || small_eligible(c, g, W, keep_hist)) {
		for (int64_t bi = 0; bi < nbatch; bi++) PGQ_TRY(fn(c, b0, bi));
		return PGQ_OK;
	}
	while ((int)c->helpers.size() < nw - 1) {
		Ctx *h = new Ctx();
		h->device = c->device;
		h->sm_count = c->sm_count;
		const int rc = ctx_init_resources(h);
		if (rc != PGQ_OK) {
			ctx_release_resources(h);
			delete h;
			return rc;
		}
		c->helpers.push_back(h);
	}
	PGQ_CUDA(cudaEventRecord(c->ev_fork, c->stream)); // the lane tables are in place after this point
	struct Res {
		int rc = PGQ_OK;
		std::string err;
	};
	std::vector<Res> res((size_t)nw);
	std::atomic<int64_t> next_batch{0};
	std::atomic<int> failed{0};
	auto drain = [&](Ctx *w, const BatchBufs &bw, Res &r) {
		while (!failed.load(std::memory_order_relaxed)) {
			const int64_t bi = next_batch.fetch_add(1);
			if (bi >= nbatch) break;
			const int rc = fn(w, bw, bi);
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
			Ctx *w = c->helpers[(size_t)i - 1];
			Res &r = res[(size_t)i];
			auto fail = [&](int rc) {
				r.rc = rc;
				r.err = last_error_str();
				failed.store(1);
			};
			if (cudaSetDevice(w->device) != cudaSuccess || cudaStreamWaitEvent(w->stream, c->ev_fork, 0) != cudaSuccess) {
				set_error("helper stream setup failed");
				return fail(PGQ_ERR_CUDA);
			}
			w->opt = c->opt;
			BatchBufs bw{};
			const int rc = alloc_bufs(w, g, W, &bw);
			if (rc == PGQ_ERR_OOM) return; // no room for a second set of bitmaps: the other workers take every batch
			if (rc != PGQ_OK) return fail(rc);
			drain(w, bw, r);
			if (cudaStreamSynchronize(w->stream) != cudaSuccess) {
				set_error("helper stream failed");
				fail(PGQ_ERR_CUDA);
			}
		});
	}
	drain(c, b0, res[0]);
	for (std::thread &t : threads) t.join();
	for (Ctx *h : c->helpers) {
		c->stats.kernel_launches += h->stats.kernel_launches;
		c->stats.batches += h->stats.batches;
		c->stats.levels += h->stats.levels;
		c->stats.push_levels += h->stats.push_levels;
		c->stats.h2d_bytes += h->stats.h2d_bytes;
		c->stats.d2h_bytes += h->stats.d2h_bytes;
		h->stats = pgq_stats_t{};
	}
	for (const Res &r : res)
		if (r.rc != PGQ_OK) {
			set_error("%s", r.err.c_str());
			return r.rc;
		}
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
	// lane tables: src | dst | found | lane_off(u64, paths only)
	size_t lane_bytes = (size_t)nslots * (2 * sizeof(uint32_t) + sizeof(int32_t)) + (size_t)W * sizeof(uint64_t) + 64;
	PGQ_TRY(c->bfs_lane.reserve(lane_bytes));
	uint64_t *d_lane_off = c->bfs_lane.as<uint64_t>(); // first, 8 B aligned
	uint32_t *d_src = reinterpret_cast<uint32_t *>(d_lane_off + W);
	uint32_t *d_dst = d_src + nslots;
	int32_t *d_found = reinterpret_cast<int32_t *>(d_dst + nslots);
	PGQ_CUDA(cudaMemcpyAsync(d_src, lt.src.data(), (size_t)nslots * 4, cudaMemcpyHostToDevice, st));
	PGQ_CUDA(cudaMemcpyAsync(d_dst, lt.dst.data(), (size_t)nslots * 4, cudaMemcpyHostToDevice, st));
	c->stats.h2d_bytes += (uint64_t)nslots * 8;
	const uint64_t levels_before = c->stats.levels;
	if (!want_paths) {
		PGQ_TRY(run_batches(c, g, b, W, lt.nbatch, false, [&](Ctx *w, const BatchBufs &bw, int64_t bi) {
			return run_batch(w, g, bw, W, false, d_src + bi * W, d_dst + bi * W, d_found + bi * W);
		}));
	} else {
		// ---- K8: each batch reconstructs its paths in ROW order (R9) with batch-local list offsets;
		// the batches (possibly finished out of order on different streams) are stitched afterwards.
		std::vector<std::vector<int64_t>> batch_vals((size_t)lt.nbatch);
		PGQ_TRY(run_batches(c, g, b, W, lt.nbatch, true, [&](Ctx *w, const BatchBufs &bw, int64_t bi) {
			cudaStream_t ws = w->stream;
			int nlev = 0;
			PGQ_TRY(run_batch(w, g, bw, W, true, d_src + bi * W, d_dst + bi * W, d_found + bi * W, &nlev));
			int32_t *hf = h_found + bi * W;
			PGQ_CUDA(cudaMemcpyAsync(hf, d_found + bi * W, (size_t)W * 4, cudaMemcpyDeviceToHost, ws));
			PGQ_CUDA(cudaStreamSynchronize(ws));
			w->stats.d2h_bytes += (uint64_t)W * 4;
			const int64_t r0 = lt.batch_row_begin[bi], r1 = lt.batch_row_end[bi];
			std::vector<uint64_t> lane_off((size_t)W);
			uint64_t bt = 0; // values of this batch so far
			int lane = 0;
			for (int64_t r = r0; r < r1; r++) {
				const uint8_t kind = lt.kind[r];
				uint64_t len = 0;
				if (kind == 1) {
					len = 1;
				} else if (kind == 2) {
					while (lane < W && lt.row[bi * W + lane] != r) lane++;
					if (lane < W && hf[lane] >= 0) {
						len = paths->mode == 0 ? (uint64_t)hf[lane] + 1 : 2ull * (uint64_t)hf[lane] + 1;
						lane_off[lane] = bt;
					}
				}
				paths->entries[r].offset = bt; // batch-local until stitched
				paths->entries[r].length = len;
				paths->valid[r] = len ? 1 : 0;
				bt += len;
			}
			std::vector<int64_t> &vals = batch_vals[(size_t)bi];
			vals.resize(bt);
			if (!bt) return (int)PGQ_OK;
			DevBuf &d_out = w->bfs_pathout; // persistent: a per-call buffer would cudaMalloc/cudaFree every batch
			if (bt * sizeof(int64_t) > d_out.cap) PGQ_TRY(d_out.reserve(bt * sizeof(int64_t) * 2));
			uint64_t *d_lo = d_lane_off;
			if (w != c) { // helpers keep their own lane-offset table
				PGQ_TRY(w->bfs_lane.reserve((size_t)W * sizeof(uint64_t) + 64));
				d_lo = w->bfs_lane.as<uint64_t>();
			}
			PGQ_CUDA(cudaMemcpyAsync(d_lo, lane_off.data(), (size_t)W * 8, cudaMemcpyHostToDevice, ws));
			// device table of the level buffers (they may have been reallocated by this batch)
			std::vector<const uint32_t *> hp((size_t)nlev + 1);
			for (int k = 0; k <= nlev; k++) hp[(size_t)k] = w->bfs_hist[(size_t)k].as<uint32_t>();
			PGQ_TRY(w->bfs_hist_tab.reserve(hp.size() * sizeof(uint32_t *)));
			PGQ_CUDA(cudaMemcpyAsync(w->bfs_hist_tab.p, hp.data(), hp.size() * sizeof(uint32_t *),
			                         cudaMemcpyHostToDevice, ws));
			bfs_paths_hist_kernel<<<W, 256, 0, ws>>>(
			    w->bfs_hist_tab.as<const uint32_t *const>(), g->d_ioff.as<uint32_t>(), g->d_isrc.as<uint32_t>(),
			    g->d_ipos.as<uint32_t>(), g->have_eid ? g->d_eid.as<int64_t>() : nullptr,
			    g->keep_rows ? g->d_row.as<uint32_t>() : nullptr, d_src + bi * W, d_dst + bi * W, d_found + bi * W,
			    (const unsigned long long *)d_lo, W, paths->mode, d_out.as<int64_t>());
			count_launch(w);
			PGQ_CUDA(cudaMemcpyAsync(vals.data(), d_out.p, bt * sizeof(int64_t), cudaMemcpyDeviceToHost, ws));
			PGQ_CUDA(cudaStreamSynchronize(ws));
			w->stats.d2h_bytes += bt * 8;
			// trivial rows: [src]
			for (int64_t r = r0; r < r1; r++)
				if (lt.kind[r] == 1) vals[paths->entries[r].offset] = row_src[r];
			return (int)PGQ_OK;
		}));
		uint64_t total = 0; // running list offset over all rows
		for (int64_t bi = 0; bi < lt.nbatch; bi++) total += batch_vals[(size_t)bi].size();
		paths->vals.resize(total);
		total = 0;
		for (int64_t bi = 0; bi < lt.nbatch; bi++) {
			const std::vector<int64_t> &vals = batch_vals[(size_t)bi];
			for (int64_t r = lt.batch_row_begin[bi]; r < lt.batch_row_end[bi]; r++) paths->entries[r].offset += total;
			if (!vals.empty()) memcpy(paths->vals.data() + total, vals.data(), vals.size() * sizeof(int64_t));
			total += vals.size();
		}
	}
	if (!want_paths) {
		PGQ_CUDA(cudaMemcpyAsync(h_found, d_found, (size_t)nslots * 4, cudaMemcpyDeviceToHost, st));
		PGQ_CUDA(cudaStreamSynchronize(st));
		c->stats.d2h_bytes += (uint64_t)nslots * 4;
	}
	PGQ_TRY(small_flush(c));
	if (c->opt.stats)
		c->stats.algo_bytes += (c->stats.levels - levels_before) * 4ull * (uint64_t)g->V * (uint64_t)(W / 8);
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
	using C = LaneCfg<W>;
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
	if (b.peer_tab) { // fused exchange: kernels store new words into every peer's replica
		a.peer_next = b.peer_tab + (size_t)(cur ^ 1) * (G - 1);
		a.npeers = (uint32_t)(G - 1);
	}
	launch_tiny<W>(c, a, g, true);
	launch_class<W, 2>(c, a, g->d_small, g->n_small, true);
	launch_warp<W>(c, a, g->d_med, g->n_med, true);
	launch_warp<W>(c, a, g->d_mid, g->n_mid, true);
	if (g->n_hub) {
		a.list = g->d_hub.as<uint32_t>();
		a.nlist = g->n_hub;
		a.chunks = g->d_hubchunks.as<HubChunk>();
		a.nchunks = g->n_hubchunks;
		pull_hub_kernel<W><<<g->n_hubchunks, 256, 0, st>>>(a);
		hub_finalize_kernel<W><<<div_up((uint64_t)g->n_hub * C::TPV, 256), 256, 0, st>>>(a);
		count_launch(c, 2);
	}
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
	const uint64_t levels_before = c->stats.levels;
	PGQ_TRY(run_batches(c, g, b, W, nbatch, false, [&](Ctx *w, const BatchBufs &bw, int64_t bi) {
		return run_batch(w, g, bw, W, false, d_src + bi * W, d_dst + bi * W, d_found + bi * W);
	}));
	rows_from_lanes_kernel<<<div_up(n, 256), 256, 0, st>>>(d_srcs, d_dsts, n, d_found, d_out, d_out_valid);
	count_launch(c);
	uint32_t h_err = 0;
	PGQ_CUDA(cudaMemcpyAsync(&h_err, d_err, 4, cudaMemcpyDeviceToHost, st));
	PGQ_CUDA(cudaStreamSynchronize(st));
	PGQ_TRY(small_flush(c));
	if (c->opt.stats)
		c->stats.algo_bytes += (c->stats.levels - levels_before) * 4ull * (uint64_t)g->V * (uint64_t)(W / 8);
	if (h_err) {
		set_error("iterativelength: src/dst id outside [0, v_size)");
		return PGQ_ERR_INVALID;
	}
	return PGQ_OK;
}

} // namespace pgq
