// C ABI of pgq_b200 (include/pgq_b200.h): context, CSR lifecycle, queries, helpers.
// Thread-safety: every entry point takes the context mutex (DuckDB may enter a scalar function
// from several worker threads at once: /root/reference/src/parallel/pipeline.cpp:79-123).
#include <execinfo.h>
#include <fcntl.h>
#include <signal.h>
#include <stdarg.h>
#include <stdlib.h>
#include <unistd.h>
#include <string.h>

#include "common.cuh"

#include <map>
#include <mutex>
#include <utility>
#include <vector>

#include <cmath>
#include "nccl_dl.cuh"

namespace pgq {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(g_err, sizeof(g_err), fmt, ap);
	va_end(ap);
}

static Csr *find_csr(Ctx *c, int32_t id) {
	auto it = c->csrs.find(id);
	if (it == c->csrs.end()) {
		set_error("unknown csr id %d", id);
		return nullptr;
	}
	return it->second;
}

static constexpr int64_t STAGE_ROWS = 4 << 20; // rows per pinned staging block (32 MiB per column)

static int flush_staged(Ctx *c, Csr *g) {
	if (g->staged == 0) return PGQ_OK;
	size_t bytes = (size_t)g->staged * sizeof(int64_t);
	PGQ_CUDA(cudaMemcpyAsync(g->d_src64.as<int64_t>() + g->uploaded, g->h_src, bytes, cudaMemcpyHostToDevice,
	                         c->stream));
	PGQ_CUDA(cudaMemcpyAsync(g->d_dst64.as<int64_t>() + g->uploaded, g->h_dst, bytes, cudaMemcpyHostToDevice,
	                         c->stream));
	c->stats.h2d_bytes += 2 * bytes;
	if (g->have_eid) {
		PGQ_CUDA(cudaMemcpyAsync(g->d_eid64.as<int64_t>() + g->uploaded, g->h_eid, bytes, cudaMemcpyHostToDevice,
		                         c->stream));
		c->stats.h2d_bytes += bytes;
	}
	PGQ_CUDA(cudaStreamSynchronize(c->stream)); // staging block is reused
	g->uploaded += g->staged;
	g->staged = 0;
	return PGQ_OK;
}

static int begin_edges(Ctx *c, Csr *g, int64_t E, bool with_eid) {
	if (g->state != Csr::CREATED) {
		set_error("create_csr_edge: csr already has its edge array");
		return PGQ_ERR_STATE;
	}
	if (E < 0) {
		set_error("negative edge count");
		return PGQ_ERR_INVALID;
	}
	g->E = E;
	g->have_eid = with_eid;
	if (with_eid) g->keep_rows = true;
	size_t bytes = (size_t)(E > 0 ? E : 1) * sizeof(int64_t);
	PGQ_TRY(g->d_src64.reserve(bytes));
	PGQ_TRY(g->d_dst64.reserve(bytes));
	if (with_eid) PGQ_TRY(g->d_eid64.reserve(bytes));
	g->uploaded = 0;
	g->staged = 0;
	g->state = Csr::EDGES_OPEN;
	return PGQ_OK;
}

static int finalize(Ctx *c, Csr *g) {
	if (g->state != Csr::EDGES_OPEN) {
		set_error("csr finalize: no open edge array");
		return PGQ_ERR_STATE;
	}
	PGQ_TRY(flush_staged(c, g));
	if (g->uploaded != g->E) {
		set_error("create_csr_edge: got %lld rows, num_edges said %lld", (long long)g->uploaded, (long long)g->E);
		return PGQ_ERR_MISMATCH;
	}
	const int rc = csr_device_build(c, g);
	// the 16-24 B/edge staging copies go whether the build worked or not
	g->d_src64.release();
	g->d_dst64.release();
	g->d_eid64.release();
	pinned_release(g->h_src, g->h_bytes);
	pinned_release(g->h_dst, g->h_bytes);
	pinned_release(g->h_eid, g->h_bytes);
	g->h_src = g->h_dst = g->h_eid = nullptr;
	g->staged_cap = 0;
	if (rc != PGQ_OK) { // back to "vertices only": the edge rows can be sent again after pgq_csr_begin_edges
		g->state = Csr::CREATED;
		g->uploaded = g->staged = 0;
		return rc;
	}
	g->state = Csr::READY;
	return PGQ_OK;
}

} // namespace pgq


namespace pgq {

// ---- small-block caches (see common.cuh) ---------------------------------------------------------------
namespace {
constexpr size_t CACHE_MAX_BLOCK = (size_t)16 << 20;
constexpr size_t CACHE_MAX_TOTAL = (size_t)512 << 20;
constexpr size_t PINNED_MAX_BLOCK = (size_t)32 << 20;
constexpr size_t PINNED_MAX_TOTAL = (size_t)256 << 20;
std::mutex g_cache_mu;
struct DevCache {
	std::vector<void *> free_list[32]; // by log2(size)
	size_t total = 0;
};
std::map<int, DevCache> g_devcache; // per device
std::vector<std::pair<void *, size_t>> g_pinned;
size_t g_pinned_total = 0;

inline int size_class(size_t bytes, size_t *rounded) {
	int k = 8; // 256 B
	while (((size_t)1 << k) < bytes) k++;
	*rounded = (size_t)1 << k;
	return k;
}
} // namespace

void *devcache_alloc(size_t bytes, size_t *actual) {
	void *p = nullptr;
	if (bytes <= CACHE_MAX_BLOCK) {
		size_t rounded;
		const int k = size_class(bytes, &rounded);
		int dev = 0;
		cudaGetDevice(&dev);
		{
			std::lock_guard<std::mutex> lk(g_cache_mu);
			DevCache &dc = g_devcache[dev];
			if (!dc.free_list[k].empty()) {
				p = dc.free_list[k].back();
				dc.free_list[k].pop_back();
				dc.total -= rounded;
			}
		}
		bytes = rounded;
	}
	if (!p) {
		cudaError_t e = cudaMalloc(&p, bytes);
		if (e != cudaSuccess) { // give the cached blocks back and try once more
			(void)cudaGetLastError();
			devcache_trim();
			e = cudaMalloc(&p, bytes);
		}
		if (e != cudaSuccess) {
			set_error("cudaMalloc(%zu) -> %s", bytes, cudaGetErrorString(e));
			(void)cudaGetLastError();
			*actual = 0;
			return nullptr;
		}
	}
	*actual = bytes;
	return p;
}

void devcache_free(void *p, size_t actual) {
	if (!p) return;
	if (actual <= CACHE_MAX_BLOCK && actual >= 256 && (actual & (actual - 1)) == 0) {
		int dev = 0;
		cudaGetDevice(&dev);
		cudaDeviceSynchronize(); // what cudaFree guaranteed: nothing in flight still uses the block
		size_t rounded;
		const int k = size_class(actual, &rounded);
		std::lock_guard<std::mutex> lk(g_cache_mu);
		DevCache &dc = g_devcache[dev];
		if (dc.total + rounded <= CACHE_MAX_TOTAL) {
			dc.free_list[k].push_back(p);
			dc.total += rounded;
			return;
		}
	}
	cudaFree(p);
}

void devcache_trim() {
	int dev = 0;
	cudaGetDevice(&dev);
	std::vector<void *> victims;
	{
		std::lock_guard<std::mutex> lk(g_cache_mu);
		DevCache &dc = g_devcache[dev];
		for (auto &fl : dc.free_list) {
			victims.insert(victims.end(), fl.begin(), fl.end());
			fl.clear();
		}
		dc.total = 0;
	}
	for (void *p : victims) cudaFree(p);
}

void *pinned_acquire(size_t bytes, size_t *actual) {
	if (bytes <= PINNED_MAX_BLOCK) {
		size_t rounded;
		size_class(bytes, &rounded);
		bytes = rounded;
		std::lock_guard<std::mutex> lk(g_cache_mu);
		for (size_t i = 0; i < g_pinned.size(); i++) {
			if (g_pinned[i].second == bytes) {
				void *p = g_pinned[i].first;
				g_pinned.erase(g_pinned.begin() + (long)i);
				g_pinned_total -= bytes;
				*actual = bytes;
				return p;
			}
		}
	}
	void *p = nullptr;
	if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) {
		set_error("cudaHostAlloc(%zu) failed", bytes);
		(void)cudaGetLastError();
		*actual = 0;
		return nullptr;
	}
	*actual = bytes;
	return p;
}

void pinned_release(void *p, size_t actual) {
	if (!p) return;
	if (actual <= PINNED_MAX_BLOCK) {
		std::lock_guard<std::mutex> lk(g_cache_mu);
		if (g_pinned_total + actual <= PINNED_MAX_TOTAL) {
			g_pinned.emplace_back(p, actual);
			g_pinned_total += actual;
			return;
		}
	}
	cudaFreeHost(p);
}

const char *last_error_str() {
	return g_err;
}

int ctx_init_resources(Ctx *c) {
	PGQ_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
	PGQ_CUDA(cudaEventCreate(&c->ev_a));
	PGQ_CUDA(cudaEventCreate(&c->ev_b));
	PGQ_CUDA(cudaEventCreate(&c->ev_c));
	PGQ_CUDA(cudaEventCreate(&c->ev_timer));
	PGQ_CUDA(cudaEventCreate(&c->ev_timer2));
	PGQ_CUDA(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
	PGQ_CUDA(cudaHostAlloc(&c->h_ctrl, 4096, cudaHostAllocDefault));
	if (const char *fg = getenv("PGQ_B200_L2_FETCH")) { // A/B knob: cudaLimitMaxL2FetchGranularity (32, 64 or 128 bytes)
		const int bytes = atoi(fg);
		if (bytes == 32 || bytes == 64 || bytes == 128) (void)cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)bytes);
		size_t got = 0;
		if (cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity) == cudaSuccess && getenv("PGQ_B200_VERBOSE"))
			fprintf(stderr, "[pgq] L2 fetch granularity %zu bytes\n", got);
		(void)cudaGetLastError();
	}
	return PGQ_OK;
}

void ctx_release_resources(Ctx *c) {
	for (Ctx *h : c->helpers) {
		cudaSetDevice(h->device);
		if (h->stream) cudaStreamSynchronize(h->stream);
		ctx_release_resources(h);
		if (h->device != c->device) devcache_trim();
		delete h;
	}
	c->helpers.clear();
	cudaSetDevice(c->device);
	for (LoopGraph &lg : c->loop_graph) lg.destroy();
	if (c->h_found) cudaFreeHost(c->h_found);
	c->h_found = nullptr;
	c->h_found_cap = 0;
	c->bfs_done.release();
	c->hist_bytes = 0;
	DevBuf *bufs[] = {&c->scan_tmp, &c->sort_keys[0], &c->sort_keys[1], &c->sort_p0[0], &c->sort_p0[1],
	                  &c->sort_p1[0], &c->sort_p1[1], &c->sort_hist, &c->sort_pack, &c->bfs_seen, &c->bfs_visit[0],
	                  &c->bfs_visit[1], &c->bfs_hist_tab, &c->bfs_pathout, &c->bfs_hubacc, &c->bfs_queue[0], &c->bfs_queue[1], &c->bfs_inq[0],
	                  &c->bfs_inq[1], &c->bfs_lane, &c->bfs_ctrl, &c->bfs_alive, &c->part_alive, &c->part_visit[0], &c->part_visit[1], &c->part_peer_tab, &c->l2_flush};
	for (DevBuf *b : bufs) b->release();
	for (SsspSlot &sl : c->sssp_slot) sl.destroy();
	if (c->h_sssp) cudaFreeHost(c->h_sssp);
	c->h_sssp = nullptr;
	for (DevBuf &hb : c->bfs_hist) hb.release();
	c->bfs_hist.clear();
	if (c->h_ctrl) cudaFreeHost(c->h_ctrl);
	c->h_ctrl = nullptr;
	if (c->ev_a) cudaEventDestroy(c->ev_a);
	if (c->ev_b) cudaEventDestroy(c->ev_b);
	if (c->ev_c) cudaEventDestroy(c->ev_c);
	if (c->ev_timer) cudaEventDestroy(c->ev_timer);
	if (c->ev_timer2) cudaEventDestroy(c->ev_timer2);
	if (c->ev_fork) cudaEventDestroy(c->ev_fork);
	if (c->stream) cudaStreamDestroy(c->stream);
	c->stream = nullptr;
}

} // namespace pgq

using namespace pgq;

#define LOCK(ctx)                                   \
	if (!(ctx)) {                                   \
		pgq::set_error("null context");             \
		return PGQ_ERR_INVALID;                     \
	}                                               \
	Ctx *c = &(ctx)->c;                             \
	std::lock_guard<std::mutex> _guard(c->mu);      \
	if (cudaSetDevice(c->device) != cudaSuccess) {  \
		pgq::set_error("cudaSetDevice failed");     \
		return PGQ_ERR_CUDA;                        \
	}

extern "C" {

const char *pgq_last_error(void) {
	return g_err;
}

// PGQ_B200_BACKTRACE=1: frames of a crashing thread on stderr (addresses resolve with addr2line -e libpgq_b200.so)
static void crash_handler(int sig) {
	const char msg[] = "[pgq_b200] fatal signal, backtrace:\n";
	const int fd = open("/tmp/pgq_b200_crash.txt", O_WRONLY | O_CREAT | O_APPEND, 0644);
	if (fd >= 0 && write(fd, msg, sizeof(msg) - 1) < 0) {
	}
	if (write(2, msg, sizeof(msg) - 1) < 0) {
	}
	void *frames[64];
	const int n = backtrace(frames, 64);
	if (fd >= 0) {
		backtrace_symbols_fd(frames, n, fd);
		close(fd);
	}
	backtrace_symbols_fd(frames, n, 2);
	signal(sig, SIG_DFL);
	raise(sig);
}

static void install_crash_handler() {
	void *warm[4];
	(void)backtrace(warm, 4); // the first call loads libgcc: not something to do inside a signal handler
	static char altstack[1 << 16];
	stack_t ss;
	ss.ss_sp = altstack;
	ss.ss_size = sizeof(altstack);
	ss.ss_flags = 0;
	sigaltstack(&ss, nullptr); // (the installing thread only: enough to tell a stack overflow apart)
	struct sigaction sa;
	memset(&sa, 0, sizeof(sa));
	sa.sa_handler = crash_handler;
	sa.sa_flags = SA_ONSTACK;
	sigaction(SIGSEGV, &sa, nullptr);
	sigaction(SIGBUS, &sa, nullptr);
	sigaction(SIGABRT, &sa, nullptr);
}

static int init_one(int device, pgq_ctx_t **out_ctx) {
	if (getenv("PGQ_B200_BACKTRACE")) install_crash_handler();
	if (!out_ctx) return PGQ_ERR_INVALID;
	*out_ctx = nullptr;
	int ndev = 0;
	cudaError_t e = cudaGetDeviceCount(&ndev);
	if (e != cudaSuccess || ndev == 0) {
		set_error("no CUDA device: %s (pgq_b200 has no CPU fallback)", cudaGetErrorString(e));
		(void)cudaGetLastError();
		return PGQ_ERR_CUDA;
	}
	if (device < 0 || device >= ndev) {
		set_error("device %d out of range (%d devices)", device, ndev);
		return PGQ_ERR_INVALID;
	}
	PGQ_CUDA(cudaSetDevice(device));
	cudaDeviceProp prop;
	PGQ_CUDA(cudaGetDeviceProperties(&prop, device));
	if (prop.major < 10) {
		set_error("device %d is sm_%d%d; pgq_b200 is built for sm_100a only", device, prop.major, prop.minor);
		return PGQ_ERR_CUDA;
	}
	pgq_ctx_t *ctx = new pgq_ctx_t();
	Ctx *c = &ctx->c;
	c->device = device;
	c->sm_count = prop.multiProcessorCount;
	int rc = ctx_init_resources(c);
	if (rc != PGQ_OK) {
		ctx_release_resources(c);
		delete ctx;
		return rc;
	}
	c->devices.assign(1, device);
	c->devices_all = c->devices;
	*out_ctx = ctx;
	return PGQ_OK;
}

int pgq_init(int device, pgq_ctx_t **out_ctx) {
	return init_one(device, out_ctx);
}

int pgq_init_devices(const int *devices, int n, pgq_ctx_t **out_ctx) {
	if (!out_ctx || !devices || n < 1) {
		set_error("pgq_init_devices: need at least one device");
		return PGQ_ERR_INVALID;
	}
	int ndev = 0;
	cudaError_t e = cudaGetDeviceCount(&ndev);
	if (e != cudaSuccess || ndev == 0) {
		set_error("no CUDA device: %s (pgq_b200 has no CPU fallback)", cudaGetErrorString(e));
		(void)cudaGetLastError();
		*out_ctx = nullptr;
		return PGQ_ERR_CUDA;
	}
	for (int i = 0; i < n; i++) {
		if (devices[i] < 0 || devices[i] >= ndev) {
			set_error("device %d out of range (%d devices)", devices[i], ndev);
			return PGQ_ERR_INVALID;
		}
		for (int j = 0; j < i; j++)
			if (devices[j] == devices[i]) {
				set_error("device %d listed twice", devices[i]);
				return PGQ_ERR_INVALID;
			}
		cudaDeviceProp prop;
		PGQ_CUDA(cudaGetDeviceProperties(&prop, devices[i]));
		if (prop.major < 10) {
			set_error("device %d is sm_%d%d; pgq_b200 is built for sm_100a only", devices[i], prop.major, prop.minor);
			return PGQ_ERR_CUDA;
		}
	}
	PGQ_TRY(init_one(devices[0], out_ctx));
	Ctx *c = &(*out_ctx)->c;
	c->devices.assign(devices, devices + n);
	c->devices_all = c->devices;
	return PGQ_OK;
}

int pgq_device_count(pgq_ctx_t *ctx, int *out_n) {
	if (!ctx || !out_n) return PGQ_ERR_INVALID;
	*out_n = (int)ctx->c.devices.size();
	return PGQ_OK;
}

int pgq_shutdown(pgq_ctx_t *ctx) {
	if (!ctx) return PGQ_OK;
	Ctx *c = &ctx->c;
	cudaSetDevice(c->device);
	cudaStreamSynchronize(c->stream);
	for (auto &kv : c->csrs) {
		kv.second->free_all();
		delete kv.second;
	}
	c->csrs.clear();
	for (void *m : c->part_peer_open) cudaIpcCloseMemHandle(m);
	c->part_peer_open.clear();
	if (c->nccl_comm) {
		const NcclApi *nc = nccl_api();
		if (nc) nc->CommDestroy((ncclComm_t)c->nccl_comm);
		c->nccl_comm = nullptr;
	}
	ctx_release_resources(c);
	devcache_trim(); // cached blocks of this device go back to the driver with the context
	delete ctx;
	return PGQ_OK;
}

int pgq_set_option(pgq_ctx_t *ctx, const char *key, int64_t value) {
	LOCK(ctx);
	if (!key) return PGQ_ERR_INVALID;
	if (!strcmp(key, "lanes")) {
		if (value != 0 && value != 64 && value != 128 && value != 256 && value != 512) {
			set_error("lanes must be 0 (auto), 64, 128, 256 or 512");
			return PGQ_ERR_INVALID;
		}
		c->opt.lanes = value;
	} else if (!strcmp(key, "stats")) c->opt.stats = value;
	else if (!strcmp(key, "timing")) c->opt.timing = value;
	else if (!strcmp(key, "tiny_max")) c->opt.tiny_max = value < 0 ? 0 : value;
	else if (!strcmp(key, "small_max")) c->opt.small_max = value < 0 ? 0 : value;
	else if (!strcmp(key, "med_max")) c->opt.med_max = value < 0 ? 0 : value;
	else if (!strcmp(key, "hub_min")) c->opt.hub_min = value < 1 ? 1 : value;
	else if (!strcmp(key, "hub_chunk")) c->opt.hub_chunk = value < 256 ? 256 : value;
	else if (!strcmp(key, "direction")) c->opt.direction = value;
	else if (!strcmp(key, "trace")) c->opt.trace = value;
	else if (!strcmp(key, "push_vbeta")) c->opt.push_vbeta = value < 1 ? 1 : value;
	else if (!strcmp(key, "filter_gamma")) c->opt.filter_gamma = value < 0 ? 0 : value;
	else if (!strcmp(key, "push_beta")) c->opt.push_beta = value < 1 ? 1 : value;
	else if (!strcmp(key, "small_path")) c->opt.small_path = value;
	else if (!strcmp(key, "small_edge_mb")) c->opt.small_edge_mb = value < 0 ? 0 : value;
	else if (!strcmp(key, "trim_cache")) devcache_trim(); // give cached small blocks back to the driver now
	else if (!strcmp(key, "seg_small")) c->opt.seg_small = value;
	else if (!strcmp(key, "sort_wide")) c->opt.sort_wide = value;
	else if (!strcmp(key, "sort_lookback")) c->opt.sort_lookback = value;
	else if (!strcmp(key, "sort_wide_min")) c->opt.sort_wide_min = value < 0 ? 0 : value;
	else if (!strcmp(key, "streams")) c->opt.streams = value < 1 ? 1 : value > 16 ? 16 : value;
	else if (!strcmp(key, "loop")) c->opt.loop = value;
	else if (!strcmp(key, "sssp_narrow")) c->opt.sssp_narrow = value;
	else if (!strcmp(key, "sssp_prune")) c->opt.sssp_prune = value;
	else if (!strcmp(key, "sssp_slots")) c->opt.sssp_slots = value < 1 ? 1 : (value > 4 ? 4 : value);
	else if (!strcmp(key, "loop_min_batches")) c->opt.loop_min_batches = value < 0 ? 0 : value;
	else if (!strcmp(key, "loop_min_calls")) c->opt.loop_min_calls = value < 0 ? 0 : value;
	else if (!strcmp(key, "done_bitmap")) c->opt.done_bitmap = value ? 1 : 0;
	else if (!strcmp(key, "dense_store")) c->opt.dense_store = value ? 1 : 0;
	else if (!strcmp(key, "dyn")) c->opt.dyn = value ? 1 : 0;
	else if (!strcmp(key, "l2_persist")) c->opt.l2_persist = value;
	else if (!strcmp(key, "hist_budget_mb")) c->opt.hist_budget_mb = value < 0 ? 0 : value;
	else if (!strcmp(key, "devices_used")) { // at most this many devices of a multi-GPU context take batches
		if (value < 1 || value > (int64_t)c->devices_all.size()) {
			set_error("devices_used must be in [1, %zu]", c->devices_all.size());
			return PGQ_ERR_INVALID;
		}
		c->devices.assign(c->devices_all.begin(), c->devices_all.begin() + value);
	}
	else {
		set_error("unknown option '%s'", key);
		return PGQ_ERR_INVALID;
	}
	return PGQ_OK;
}

int pgq_get_stats(pgq_ctx_t *ctx, pgq_stats_t *out) {
	LOCK(ctx);
	if (!out) return PGQ_ERR_INVALID;
	*out = c->stats;
	return PGQ_OK;
}

int pgq_reset_stats(pgq_ctx_t *ctx) {
	LOCK(ctx);
	double b = c->stats.csr_build_ms, t = c->stats.transpose_ms;
	memset(&c->stats, 0, sizeof(c->stats));
	c->stats.csr_build_ms = b;
	c->stats.transpose_ms = t;
	return PGQ_OK;
}

// ---- CSR lifecycle ------------------------------------------------------------------------

int pgq_csr_create(pgq_ctx_t *ctx, int32_t id, int64_t v_size) {
	LOCK(ctx);
	if (v_size < 0) {
		set_error("negative v_size");
		return PGQ_ERR_INVALID;
	}
	auto it = c->csrs.find(id);
	if (it != c->csrs.end()) { // re-creating an id replaces it (the reference deletes after the query)
		it->second->free_all();
		delete it->second;
		c->csrs.erase(it);
	}
	Csr *g = new Csr();
	g->V = v_size;
	g->device = c->device;
	c->csrs[id] = g;
	return PGQ_OK;
}

int pgq_csr_set_degrees(pgq_ctx_t *ctx, int32_t id, const int64_t *vid, const int64_t *cnt, int64_t n,
                        int64_t *out_sum) {
	LOCK(ctx);
	Csr *g = find_csr(c, id);
	if (!g) return PGQ_ERR_INVALID;
	if (g->state != Csr::CREATED) {
		set_error("create_csr_vertex after create_csr_edge started");
		return PGQ_ERR_STATE;
	}
	if (!g->have_cnt) {
		g->cnt_host.assign((size_t)g->V, 0);
		g->have_cnt = true;
	}
	int64_t sum = 0;
	for (int64_t i = 0; i < n; i++) {
		if (vid[i] < 0 || vid[i] >= g->V) {
			set_error("create_csr_vertex: vertex id %lld outside [0,%lld)", (long long)vid[i], (long long)g->V);
			return PGQ_ERR_INVALID;
		}
		g->cnt_host[(size_t)vid[i]] = cnt[i];
		sum += cnt[i];
	}
	if (out_sum) *out_sum += sum;
	return PGQ_OK;
}

int pgq_csr_begin_edges(pgq_ctx_t *ctx, int32_t id, int64_t e_size) {
	LOCK(ctx);
	Csr *g = find_csr(c, id);
	if (!g) return PGQ_ERR_INVALID;
	if (g->state != Csr::CREATED || g->pending_begin) { // a built (or half-built) graph is never reopened silently
		set_error("create_csr_edge: csr %d already has its edge array (delete it or re-create it first)", id);
		return PGQ_ERR_STATE;
	}
	if (e_size < 0) {
		set_error("negative edge count");
		return PGQ_ERR_INVALID;
	}
	g->E = e_size;
	g->pending_begin = true;
	return PGQ_OK;
}

int pgq_csr_add_edges(pgq_ctx_t *ctx, int32_t id, const int64_t *src, const int64_t *dst, const int64_t *eid,
                      int64_t n) {
	LOCK(ctx);
	Csr *g = find_csr(c, id);
	if (!g) return PGQ_ERR_INVALID;
	if (g->pending_begin) { // whether edge ids exist is only known at the first row batch
		g->pending_begin = false;
		PGQ_TRY(begin_edges(c, g, g->E, eid != nullptr));
	}
	if (g->state != Csr::EDGES_OPEN) {
		set_error("create_csr_edge before CsrInitializeEdge (pgq_csr_begin_edges)");
		return PGQ_ERR_STATE;
	}
	if ((eid != nullptr) != g->have_eid) {
		set_error("create_csr_edge: edge ids given for some rows only");
		return PGQ_ERR_INVALID;
	}
	if (g->uploaded + g->staged + n > g->E) {
		set_error("create_csr_edge: more rows than num_edges=%lld", (long long)g->E);
		return PGQ_ERR_MISMATCH;
	}
	if (!g->h_src) {
		g->staged_cap = g->E < STAGE_ROWS ? (g->E > 0 ? g->E : 1) : STAGE_ROWS;
		size_t bytes = (size_t)g->staged_cap * sizeof(int64_t);
		g->h_src = static_cast<int64_t *>(pinned_acquire(bytes, &g->h_bytes));
		g->h_dst = static_cast<int64_t *>(pinned_acquire(bytes, &g->h_bytes));
		if (g->have_eid) g->h_eid = static_cast<int64_t *>(pinned_acquire(bytes, &g->h_bytes));
		if (!g->h_src || !g->h_dst || (g->have_eid && !g->h_eid)) { // all or nothing: a retry allocates again
			pinned_release(g->h_src, g->h_bytes);
			pinned_release(g->h_dst, g->h_bytes);
			pinned_release(g->h_eid, g->h_bytes);
			g->h_src = g->h_dst = g->h_eid = nullptr;
			g->staged_cap = 0;
			return PGQ_ERR_OOM;
		}
	}
	int64_t done = 0;
	while (done < n) {
		int64_t room = g->staged_cap - g->staged;
		int64_t k = n - done < room ? n - done : room;
		memcpy(g->h_src + g->staged, src + done, (size_t)k * sizeof(int64_t));
		memcpy(g->h_dst + g->staged, dst + done, (size_t)k * sizeof(int64_t));
		if (g->have_eid) memcpy(g->h_eid + g->staged, eid + done, (size_t)k * sizeof(int64_t));
		g->staged += k;
		done += k;
		if (g->staged == g->staged_cap) PGQ_TRY(flush_staged(c, g));
	}
	return PGQ_OK;
}

int pgq_csr_finalize(pgq_ctx_t *ctx, int32_t id) {
	LOCK(ctx);
	Csr *g = find_csr(c, id);
	if (!g) return PGQ_ERR_INVALID;
	if (g->pending_begin) { // zero edge rows arrived
		g->pending_begin = false;
		PGQ_TRY(begin_edges(c, g, g->E, false));
	}
	return finalize(c, g);
}

static int build_common(Ctx *c, int32_t id, int64_t v_size, int64_t e_size, const int64_t *src, const int64_t *dst,
                        const int64_t *eid, int flags, bool device_ptrs) {
	if (v_size < 0 || e_size < 0) {
		set_error("negative size");
		return PGQ_ERR_INVALID;
	}
	auto it = c->csrs.find(id);
	if (it != c->csrs.end()) {
		it->second->free_all();
		delete it->second;
		c->csrs.erase(it);
	}
	Csr *g = new Csr();
	g->V = v_size;
	g->device = c->device;
	c->csrs[id] = g;
	g->keep_rows = (flags & PGQ_CSR_ROWIDS) != 0;
	if (device_ptrs) { // inputs already resident: no staging copy of the edge table
		g->E = e_size;
		g->have_eid = eid != nullptr;
		if (g->have_eid) g->keep_rows = true;
		g->state = Csr::EDGES_OPEN;
	} else {
		PGQ_TRY(begin_edges(c, g, e_size, eid != nullptr));
	}
	size_t bytes = (size_t)e_size * sizeof(int64_t);
	cudaMemcpyKind kind = device_ptrs ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
	if (e_size > 0) {
		if (device_ptrs) {
			// inputs already resident: build straight from the caller's arrays, no copy
			g->d_src64.p = const_cast<int64_t *>(src);
			g->d_dst64.p = const_cast<int64_t *>(dst);
			g->borrowed_edges = true;
			if (eid) g->d_eid64.p = const_cast<int64_t *>(eid);
		} else {
			PGQ_CUDA(cudaMemcpyAsync(g->d_src64.p, src, bytes, kind, c->stream));
			PGQ_CUDA(cudaMemcpyAsync(g->d_dst64.p, dst, bytes, kind, c->stream));
			c->stats.h2d_bytes += 2 * bytes;
			if (eid) {
				PGQ_CUDA(cudaMemcpyAsync(g->d_eid64.p, eid, bytes, kind, c->stream));
				c->stats.h2d_bytes += bytes;
			}
		}
	}
	g->uploaded = e_size;
	int rc = csr_device_build(c, g);
	if (g->borrowed_edges) { // never free the caller's arrays
		g->d_src64.p = nullptr; g->d_src64.cap = 0;
		g->d_dst64.p = nullptr; g->d_dst64.cap = 0;
		g->d_eid64.p = nullptr; g->d_eid64.cap = 0;
		g->borrowed_edges = false;
	}
	g->d_src64.release();
	g->d_dst64.release();
	g->d_eid64.release();
	if (rc != PGQ_OK) return rc;
	g->state = Csr::READY;
	return PGQ_OK;
}

int pgq_csr_build(pgq_ctx_t *ctx, int32_t id, int64_t v_size, int64_t e_size, const int64_t *src,
                  const int64_t *dst, const int64_t *eid, int flags) {
	LOCK(ctx);
	return build_common(c, id, v_size, e_size, src, dst, eid, flags, false);
}

int pgq_csr_build_device(pgq_ctx_t *ctx, int32_t id, int64_t v_size, int64_t e_size, const int64_t *d_src,
                         const int64_t *d_dst, const int64_t *d_eid, int flags) {
	LOCK(ctx);
	return build_common(c, id, v_size, e_size, d_src, d_dst, d_eid, flags, true);
}

__global__ void __launch_bounds__(256) widen_u32_kernel(const uint32_t *__restrict__ in, uint32_t n,
                                                        int64_t *__restrict__ out) {
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) out[i] = (int64_t)in[i];
}

int pgq_csr_download(pgq_ctx_t *ctx, int32_t id, int64_t *v, int64_t *e, int64_t *eid) {
	LOCK(ctx);
	Csr *g = find_csr(c, id);
	if (!g) return PGQ_ERR_INVALID;
	if (g->state != Csr::READY) {
		set_error("csr %d not finalized", id);
		return PGQ_ERR_STATE;
	}
	if (g->is_part) {
		set_error("csr %d is vertex-partitioned: only pgq_part_iterativelength / pgq_part_reachability apply", id);
		return PGQ_ERR_STATE;
	}
	cudaStream_t st = c->stream;
	if (v) {
		PGQ_CUDA(cudaMemcpyAsync(v, g->d_off.p, (size_t)(g->V + 2) * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
		c->stats.d2h_bytes += (uint64_t)(g->V + 2) * 8;
	}
	TmpBuf wide;
	if ((e || (eid && !g->have_eid)) && g->E > 0) PGQ_TRY(wide.reserve((size_t)g->E * sizeof(int64_t)));
	if (e && g->E > 0) {
		widen_u32_kernel<<<div_up(g->E, 256), 256, 0, st>>>(g->d_tgt.as<uint32_t>(), (uint32_t)g->E,
		                                                    wide.as<int64_t>());
		count_launch(c);
		PGQ_CUDA(cudaMemcpyAsync(e, wide.p, (size_t)g->E * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
		PGQ_CUDA(cudaStreamSynchronize(st));
		c->stats.d2h_bytes += (uint64_t)g->E * 8;
	}
	if (eid && g->E > 0) {
		if (g->have_eid) {
			PGQ_CUDA(cudaMemcpyAsync(eid, g->d_eid.p, (size_t)g->E * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
		} else if (g->keep_rows) {
			widen_u32_kernel<<<div_up(g->E, 256), 256, 0, st>>>(g->d_row.as<uint32_t>(), (uint32_t)g->E,
			                                                    wide.as<int64_t>());
			count_launch(c);
			PGQ_CUDA(cudaMemcpyAsync(eid, wide.p, (size_t)g->E * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
		} else {
			set_error("csr %d was built without edge ids", id);
			return PGQ_ERR_STATE;
		}
		c->stats.d2h_bytes += (uint64_t)g->E * 8;
	}
	PGQ_CUDA(cudaStreamSynchronize(st));
	return PGQ_OK;
}

int pgq_csr_info(pgq_ctx_t *ctx, int32_t id, int64_t *v_size, int64_t *e_size) {
	LOCK(ctx);
	Csr *g = find_csr(c, id);
	if (!g) return PGQ_ERR_INVALID;
	if (v_size) *v_size = g->V;
	if (e_size) *e_size = g->E;
	return PGQ_OK;
}

int pgq_csr_delete(pgq_ctx_t *ctx, int32_t id) {
	LOCK(ctx);
	auto it = c->csrs.find(id);
	if (it == c->csrs.end()) return PGQ_OK;
	cudaStreamSynchronize(c->stream);
	it->second->free_all();
	delete it->second;
	c->csrs.erase(it);
	return PGQ_OK;
}

int pgq_csr_prepare(pgq_ctx_t *ctx, int32_t id) {
	LOCK(ctx);
	Csr *g = find_csr(c, id);
	if (!g) return PGQ_ERR_INVALID;
	if (g->state != Csr::READY) {
		set_error("csr %d not finalized", id);
		return PGQ_ERR_STATE;
	}
	return csr_prepare(c, g);
}

// ---- queries ---------------------------------------------------------------------------------

// Lane filling exactly like the reference's loop: rows are consumed in order;
// NULL rows and src==dst rows are answered without a lane; every other row takes the next lane.
static int build_lane_table(Ctx *c, Csr *g, int64_t n, const int64_t *src, const int64_t *dst,
                            const uint8_t *sv, const uint8_t *dv, LaneTable *lt, int force_w = 0) {
	lt->kind.assign((size_t)n, 0);
	int64_t need = 0;
	for (int64_t i = 0; i < n; i++) {
		if ((sv && !sv[i]) || (dv && !dv[i])) continue;
		const int64_t vmax = g->is_part ? g->gV : g->V;
		if (src[i] < 0 || src[i] >= vmax || dst[i] < 0 || dst[i] >= vmax) {
			set_error("vertex id outside [0, v_size) in row %lld", (long long)i);
			return PGQ_ERR_INVALID;
		}
		if (src[i] == dst[i]) lt->kind[i] = 1;
		else {
			lt->kind[i] = 2;
			need++;
		}
	}
	const int W = force_w ? force_w : c->opt.lanes ? (int)c->opt.lanes : pick_lane_width(need);
	lt->W = W;
	int64_t started = 0;
	while (started < n) {
		const int64_t b = lt->nbatch;
		lt->src.resize((size_t)(b + 1) * W, 0xFFFFFFFFu);
		lt->dst.resize((size_t)(b + 1) * W, 0u);
		lt->row.resize((size_t)(b + 1) * W, -1);
		lt->batch_row_begin.push_back(started);
		for (int lane = 0; lane < W; lane++) {
			while (started < n) {
				int64_t r = started++;
				if (lt->kind[r] == 2) {
					lt->src[(size_t)b * W + lane] = (uint32_t)src[r];
					lt->dst[(size_t)b * W + lane] = (uint32_t)dst[r];
					lt->row[(size_t)b * W + lane] = r;
					break;
				}
			}
		}
		lt->batch_row_end.push_back(started);
		lt->nbatch++;
	}
	return PGQ_OK;
}

static int query_lengths(Ctx *c, int32_t id, int64_t n, const int64_t *src, const int64_t *dst, const uint8_t *sv,
                         const uint8_t *dv, int64_t *out_len, uint8_t *out_bool, uint8_t *out_valid,
                         int64_t lower = 0, int64_t upper = -1) {
	Csr *g = find_csr(c, id);
	if (!g) return PGQ_ERR_INVALID;
	if (g->state != Csr::READY) {
		set_error("csr %d not finalized", id);
		return PGQ_ERR_STATE;
	}
	if (n <= 0) return PGQ_OK;
	LaneTable lt;
	PGQ_TRY(build_lane_table(c, g, n, src, dst, sv, dv, &lt));
	std::vector<int32_t> found((size_t)lt.nbatch * lt.W, -1);
	if (g->is_part) {
		// lanes address the replicated frontier structures by the internal rank-major numbering
		const PartMap pm = part_map(g->gV, c->part_world);
		for (size_t k = 0; k < lt.src.size(); k++) {
			if (lt.src[k] == 0xFFFFFFFFu) continue;
			lt.src[k] = part_newid(lt.src[k], pm);
			lt.dst[k] = part_newid(lt.dst[k], pm);
		}
		if (upper >= 0) {
			set_error("csr %d is vertex-partitioned: bounded iterativelength is not available on that path", id);
			return PGQ_ERR_STATE;
		}
		PGQ_TRY(msbfs_part_run(c, g, lt, found.data()));
	} else if (upper != 0) { // upper == 0: no level runs, every lane stays unfound
		// the level loops (host loop and whole-batch kernel) read the cap from the options the helper
		// contexts copy; it is reset before anything else can run (the context lock is held)
		c->opt.level_cap = upper < 0 || upper > 0x7fffffff ? 0x7fffffff : upper;
		const int rc = msbfs_run(c, g, lt, found.data(), nullptr, src, n);
		c->opt.level_cap = 0x7fffffff;
		PGQ_TRY(rc);
	}
	for (int64_t i = 0; i < n; i++) {
		if (lt.kind[i] == 0) {
			out_valid[i] = 0;
			if (out_len) out_len[i] = -1;
			if (out_bool) out_bool[i] = 0;
		} else if (lt.kind[i] == 1) { // src == dst: a length-0 path, in bounds only when lower <= 0 (R11)
			out_valid[i] = lower <= 0;
			if (out_len) out_len[i] = lower <= 0 ? 0 : -1;
			if (out_bool) out_bool[i] = 1;
		}
	}
	for (size_t s = 0; s < found.size(); s++) {
		int64_t r = lt.row[s];
		if (r < 0) continue;
		if (out_len) { // iterativelength: NULL when no path (or a shortest length below the lower bound, R11)
			const bool ok = found[s] >= 0 && found[s] >= lower;
			out_valid[r] = ok;
			out_len[r] = ok ? found[s] : -1;
		} else { // reachability: false, not NULL (R6)
			out_valid[r] = 1;
			out_bool[r] = found[s] >= 0;
		}
	}
	return PGQ_OK;
}

// ---- weighted next row (f4) ---------------------------------------------------------------------

int pgq_csr_set_weights(pgq_ctx_t *ctx, int32_t id, int wtype, const void *w, int64_t row0, int64_t n) {
	LOCK(ctx);
	Csr *g = find_csr(c, id);
	if (!g) return PGQ_ERR_INVALID;
	if (g->state != Csr::READY) {
		set_error("csr %d not finalized", id);
		return PGQ_ERR_STATE;
	}
	if (n > 0 && !w) {
		set_error("null weights");
		return PGQ_ERR_INVALID;
	}
	return csr_set_weights(c, g, wtype, w, row0, n);
}

int pgq_csr_download_weights(pgq_ctx_t *ctx, int32_t id, void *out) {
	LOCK(ctx);
	Csr *g = find_csr(c, id);
	if (!g) return PGQ_ERR_INVALID;
	return csr_download_weights(c, g, out);
}

int pgq_csr_weight_type(pgq_ctx_t *ctx, int32_t id, int *wtype) {
	LOCK(ctx);
	Csr *g = find_csr(c, id);
	if (!g || !wtype) return PGQ_ERR_INVALID;
	*wtype = g->w_ready ? g->wtype : 0;
	return PGQ_OK;
}

int pgq_cheapest_path_length(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *src, const int64_t *dst,
                             const uint8_t *src_valid, const uint8_t *dst_valid, void *out, uint8_t *out_valid) {
	LOCK(ctx);
	Csr *g = find_csr(c, id);
	if (!g) return PGQ_ERR_INVALID;
	if (g->state != Csr::READY || !g->w_ready) {
		set_error("csr %d has no weights (pgq_csr_set_weights after finalize)", id);
		return PGQ_ERR_STATE;
	}
	if (n <= 0) return PGQ_OK;
	LaneTable lt;
	PGQ_TRY(build_lane_table(c, g, n, src, dst, src_valid, dst_valid, &lt, 64));
	std::vector<uint64_t> d8((size_t)lt.nbatch * lt.W, 0);
	PGQ_TRY(sssp_run(c, g, lt, d8.data()));
	const bool f64 = g->wtype == PGQ_WEIGHT_F64;
	int64_t *oi = static_cast<int64_t *>(out);
	double *od = static_cast<double *>(out);
	for (int64_t i = 0; i < n; i++) { // NULL rows and src == dst rows took no lane
		out_valid[i] = lt.kind[i] == 1;
		if (f64) od[i] = 0.0;
		else oi[i] = 0;
	}
	for (size_t s = 0; s < d8.size(); s++) {
		const int64_t r = lt.row[s];
		if (r < 0) continue;
		if (f64) {
			double d;
			memcpy(&d, &d8[s], 8);
			out_valid[r] = !std::isinf(d);
			od[r] = std::isinf(d) ? 0.0 : d;
		} else {
			const int64_t d = (int64_t)d8[s];
			out_valid[r] = d < INT64_MAX / 2;
			oi[r] = d < INT64_MAX / 2 ? d : 0;
		}
	}
	return PGQ_OK;
}

int pgq_iterativelength(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *src, const int64_t *dst,
                        const uint8_t *src_valid, const uint8_t *dst_valid, int64_t *out, uint8_t *out_valid) {
	LOCK(ctx);
	return query_lengths(c, id, n, src, dst, src_valid, dst_valid, out, nullptr, out_valid);
}

int pgq_iterativelength_bounded(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *src, const int64_t *dst,
                                const uint8_t *src_valid, const uint8_t *dst_valid, int64_t lower, int64_t upper,
                                int64_t *out, uint8_t *out_valid) {
	LOCK(ctx);
	return query_lengths(c, id, n, src, dst, src_valid, dst_valid, out, nullptr, out_valid, lower, upper);
}

int pgq_reachability(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *src, const int64_t *dst,
                     const uint8_t *src_valid, const uint8_t *dst_valid, uint8_t *out, uint8_t *out_valid) {
	LOCK(ctx);
	return query_lengths(c, id, n, src, dst, src_valid, dst_valid, nullptr, out, out_valid);
}

static int shortestpath_common(Ctx *c, int32_t id, int64_t n, const int64_t *src, const int64_t *dst,
                               const uint8_t *src_valid, const uint8_t *dst_valid, int mode,
                               pgq_list_entry_t *entries, uint8_t *out_valid, int64_t **out_vals,
                               pgq_reserve_fn reserve, void *user, uint64_t *out_total) {
	Csr *g = find_csr(c, id);
	if (!g) return PGQ_ERR_INVALID;
	if (g->state != Csr::READY) {
		set_error("csr %d not finalized", id);
		return PGQ_ERR_STATE;
	}
	if (g->is_part) {
		set_error("csr %d is vertex-partitioned: only pgq_part_iterativelength / pgq_part_reachability apply", id);
		return PGQ_ERR_STATE;
	}
	if (mode != 0 && mode != 1) {
		set_error("shortestpath mode must be 0 or 1");
		return PGQ_ERR_INVALID;
	}
	if (mode == 1 && !g->have_eid && !g->keep_rows) {
		set_error("shortestpath mode 1 needs a csr built with edge ids");
		return PGQ_ERR_STATE;
	}
	if (out_vals) *out_vals = nullptr;
	*out_total = 0;
	if (n <= 0) {
		if (reserve) (void)reserve(user, 0);
		return PGQ_OK;
	}
	LaneTable lt;
	PGQ_TRY(build_lane_table(c, g, n, src, dst, src_valid, dst_valid, &lt));
	std::vector<int32_t> found((size_t)lt.nbatch * lt.W, -1);
	PathOut po;
	po.mode = mode;
	po.entries = entries;
	po.valid = out_valid;
	po.sink = reserve;
	po.sink_user = user;
	PGQ_TRY(msbfs_run(c, g, lt, found.data(), &po, src, n));
	if (reserve) {
		*out_total = po.total;
		return PGQ_OK;
	}
	int64_t *vals = (int64_t *)malloc((po.vals.size() ? po.vals.size() : 1) * sizeof(int64_t));
	if (!vals) {
		set_error("host allocation failed");
		return PGQ_ERR_OOM;
	}
	if (!po.vals.empty()) memcpy(vals, po.vals.data(), po.vals.size() * sizeof(int64_t));
	*out_vals = vals;
	*out_total = po.vals.size();
	return PGQ_OK;
}

int pgq_shortestpath(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *src, const int64_t *dst,
                     const uint8_t *src_valid, const uint8_t *dst_valid, int mode, pgq_list_entry_t *entries,
                     uint8_t *out_valid, int64_t **out_vals, uint64_t *out_total) {
	LOCK(ctx);
	if (!out_vals || !out_total) return PGQ_ERR_INVALID;
	return shortestpath_common(c, id, n, src, dst, src_valid, dst_valid, mode, entries, out_valid, out_vals, nullptr,
	                           nullptr, out_total);
}

int pgq_shortestpath_sink(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *src, const int64_t *dst,
                          const uint8_t *src_valid, const uint8_t *dst_valid, int mode, pgq_list_entry_t *entries,
                          uint8_t *out_valid, pgq_reserve_fn reserve, void *user, uint64_t *out_total) {
	LOCK(ctx);
	if (!reserve || !out_total) return PGQ_ERR_INVALID;
	return shortestpath_common(c, id, n, src, dst, src_valid, dst_valid, mode, entries, out_valid, nullptr, reserve,
	                           user, out_total);
}

void pgq_free(void *p) {
	free(p);
}

int pgq_iterativelength_device(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *d_src, const int64_t *d_dst,
                               int64_t *d_out, uint8_t *d_out_valid) {
	LOCK(ctx);
	Csr *g = find_csr(c, id);
	if (!g) return PGQ_ERR_INVALID;
	if (g->state != Csr::READY) {
		set_error("csr %d not finalized", id);
		return PGQ_ERR_STATE;
	}
	if (g->is_part) {
		set_error("csr %d is vertex-partitioned: only pgq_part_iterativelength / pgq_part_reachability apply", id);
		return PGQ_ERR_STATE;
	}
	return msbfs_run_device(c, g, n, d_src, d_dst, d_out, d_out_valid);
}

// ---- bench / test helpers -----------------------------------------------------------------------

// ---- 1-D vertex partition (collective entry points) ------------------------------------------------
int pgq_nccl_unique_id(void *out128) {
	const NcclApi *nc = nccl_api();
	if (!nc) return PGQ_ERR_CUDA;
	ncclUniqueId id;
	PGQ_NCCL(nc, nc->GetUniqueId(&id));
	static_assert(sizeof(id) == 128, "ncclUniqueId is 128 bytes");
	memcpy(out128, &id, sizeof(id));
	return PGQ_OK;
}

int pgq_part_init(pgq_ctx_t *ctx, int rank, int world, const void *unique_id128) {
	LOCK(ctx);
	if (world < 1 || rank < 0 || rank >= world) {
		set_error("pgq_part_init: bad rank/world");
		return PGQ_ERR_INVALID;
	}
	PGQ_CUDA(cudaSetDevice(c->device));
	c->part_rank = rank;
	c->part_world = world;
	if (world > 1) {
		const NcclApi *nc = nccl_api();
		if (!nc) return PGQ_ERR_CUDA;
		if (!unique_id128) {
			set_error("pgq_part_init: unique id required for world > 1");
			return PGQ_ERR_INVALID;
		}
		ncclUniqueId id;
		memcpy(&id, unique_id128, sizeof(id));
		ncclComm_t comm;
		PGQ_NCCL(nc, nc->CommInitRank(&comm, world, id, rank));
		c->nccl_comm = comm;
	}
	return PGQ_OK;
}

int pgq_part_range(pgq_ctx_t *ctx, int64_t v_size, int64_t *out_base, int64_t *out_count) {
	LOCK(ctx);
	// scrambled block-cyclic ownership (common.cuh PartMap): *out_count = slice size of the replicas
	const PartMap pm = part_map(v_size, c->part_world);
	if (out_base) *out_base = (int64_t)pm.Vp * c->part_rank;
	if (out_count) *out_count = (int64_t)pm.Vp;
	return PGQ_OK;
}

int pgq_part_csr_create(pgq_ctx_t *ctx, int32_t id, int64_t v_size) {
	LOCK(ctx);
	if (v_size < 0 || v_size > ((int64_t)1 << 30)) {
		set_error("pgq_part_csr_create: v_size out of range (32-bit vertex ids)");
		return PGQ_ERR_INVALID;
	}
	auto it = c->csrs.find(id);
	if (it != c->csrs.end()) {
		it->second->free_all();
		delete it->second;
		c->csrs.erase(it);
	}
	Csr *g = new Csr();
	g->device = c->device;
	g->is_part = true;
	g->gV = v_size;
	g->V = (int64_t)part_map(v_size, c->part_world).Vp;
	g->part_base = g->V * c->part_rank;
	g->state = Csr::EDGES_OPEN;
	c->csrs[id] = g;
	return PGQ_OK;
}

int pgq_part_csr_add_edges_device(pgq_ctx_t *ctx, int32_t id, const int64_t *d_src, const int64_t *d_dst,
                                  int64_t n) {
	LOCK(ctx);
	Csr *g = find_csr(c, id);
	if (!g) return PGQ_ERR_INVALID;
	if (!g->is_part || g->state != Csr::EDGES_OPEN) {
		set_error("csr %d is not an open partitioned csr", id);
		return PGQ_ERR_STATE;
	}
	return part_add_edges_device(c, g, d_src, d_dst, n);
}

int pgq_part_csr_finalize(pgq_ctx_t *ctx, int32_t id) {
	LOCK(ctx);
	Csr *g = find_csr(c, id);
	if (!g) return PGQ_ERR_INVALID;
	if (!g->is_part || g->state != Csr::EDGES_OPEN) {
		set_error("csr %d is not an open partitioned csr", id);
		return PGQ_ERR_STATE;
	}
	return part_finalize(c, g);
}

int pgq_part_p2p_close(pgq_ctx_t *ctx) {
	LOCK(ctx);
	PGQ_CUDA(cudaStreamSynchronize(c->stream));
	for (void *m : c->part_peer_open) cudaIpcCloseMemHandle(m);
	c->part_peer_open.clear();
	c->part_p2p = false;
	return PGQ_OK;
}

int pgq_part_p2p_export(pgq_ctx_t *ctx, int32_t id, int lanes, void *out_handles128) {
	LOCK(ctx);
	Csr *g = find_csr(c, id);
	if (!g) return PGQ_ERR_INVALID;
	if (!g->is_part || (lanes != 64 && lanes != 128 && lanes != 256 && lanes != 512)) {
		set_error("pgq_part_p2p_export: partitioned csr and lanes in {64,128,256,512} required");
		return PGQ_ERR_INVALID;
	}
	PGQ_CUDA(cudaSetDevice(c->device));
	c->part_p2p = false;
	const size_t bytes = (size_t)g->V * c->part_world * (lanes / 8);
	cudaIpcMemHandle_t h[2];
	static_assert(sizeof(h) == PGQ_P2P_HANDLE_BYTES, "two 64-byte IPC handles");
	for (int j = 0; j < 2; j++) {
		c->part_visit[j].release();
		PGQ_TRY(c->part_visit[j].reserve(bytes));
		PGQ_CUDA(cudaIpcGetMemHandle(&h[j], c->part_visit[j].p));
	}
	memcpy(out_handles128, h, sizeof(h));
	c->part_p2p_W = lanes;
	c->part_p2p_Vp = g->V;
	return PGQ_OK;
}

int pgq_part_p2p_import(pgq_ctx_t *ctx, const void *all_handles) {
	LOCK(ctx);
	PGQ_CUDA(cudaSetDevice(c->device));
	const int G = c->part_world, r = c->part_rank;
	if (G < 2 || !c->part_visit[0].p) {
		set_error("pgq_part_p2p_import: export first (world > 1)");
		return PGQ_ERR_STATE;
	}
	const size_t slice_words = (size_t)c->part_p2p_Vp * (c->part_p2p_W / 32);
	std::vector<uint32_t *> tab((size_t)2 * (G - 1));
	const cudaIpcMemHandle_t *hs = reinterpret_cast<const cudaIpcMemHandle_t *>(all_handles);
	for (int p = 0, k = 0; p < G; p++) {
		if (p == r) continue;
		for (int j = 0; j < 2; j++) {
			void *base = nullptr;
			PGQ_CUDA(cudaIpcOpenMemHandle(&base, hs[(size_t)p * 2 + j], cudaIpcMemLazyEnablePeerAccess));
			c->part_peer_open.push_back(base);
			tab[(size_t)j * (G - 1) + k] = reinterpret_cast<uint32_t *>(base) + slice_words * r;
		}
		k++;
	}
	PGQ_TRY(c->part_peer_tab.reserve(tab.size() * sizeof(uint32_t *)));
	PGQ_CUDA(cudaMemcpy(c->part_peer_tab.p, tab.data(), tab.size() * sizeof(uint32_t *), cudaMemcpyHostToDevice));
	c->part_p2p = true;
	return PGQ_OK;
}

int pgq_part_iterativelength(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *src, const int64_t *dst,
                             const uint8_t *src_valid, const uint8_t *dst_valid, int64_t *out,
                             uint8_t *out_valid) {
	LOCK(ctx);
	Csr *g = find_csr(c, id);
	if (!g) return PGQ_ERR_INVALID;
	if (!g->is_part) {
		set_error("csr %d is not partitioned", id);
		return PGQ_ERR_STATE;
	}
	return query_lengths(c, id, n, src, dst, src_valid, dst_valid, out, nullptr, out_valid);
}

int pgq_part_reachability(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *src, const int64_t *dst,
                          const uint8_t *src_valid, const uint8_t *dst_valid, uint8_t *out, uint8_t *out_valid) {
	LOCK(ctx);
	Csr *g = find_csr(c, id);
	if (!g) return PGQ_ERR_INVALID;
	if (!g->is_part) {
		set_error("csr %d is not partitioned", id);
		return PGQ_ERR_STATE;
	}
	return query_lengths(c, id, n, src, dst, src_valid, dst_valid, nullptr, out, out_valid);
}

int pgq_dev_alloc(pgq_ctx_t *ctx, uint64_t bytes, void **out_dptr) {
	LOCK(ctx);
	if (!out_dptr) return PGQ_ERR_INVALID;
	*out_dptr = nullptr;
	cudaError_t e = cudaMalloc(out_dptr, bytes ? bytes : 16);
	if (e != cudaSuccess) {
		set_error("cudaMalloc(%llu) -> %s", (unsigned long long)bytes, cudaGetErrorString(e));
		(void)cudaGetLastError();
		return PGQ_ERR_OOM;
	}
	return PGQ_OK;
}

int pgq_dev_free(pgq_ctx_t *ctx, void *dptr) {
	LOCK(ctx);
	if (dptr) PGQ_CUDA(cudaFree(dptr));
	return PGQ_OK;
}

int pgq_dev_h2d(pgq_ctx_t *ctx, void *dptr, const void *hptr, uint64_t bytes) {
	LOCK(ctx);
	PGQ_CUDA(cudaMemcpyAsync(dptr, hptr, bytes, cudaMemcpyHostToDevice, c->stream));
	PGQ_CUDA(cudaStreamSynchronize(c->stream));
	return PGQ_OK;
}

int pgq_dev_d2h(pgq_ctx_t *ctx, void *hptr, const void *dptr, uint64_t bytes) {
	LOCK(ctx);
	PGQ_CUDA(cudaMemcpyAsync(hptr, dptr, bytes, cudaMemcpyDeviceToHost, c->stream));
	PGQ_CUDA(cudaStreamSynchronize(c->stream));
	return PGQ_OK;
}

int pgq_host_alloc_pinned(uint64_t bytes, void **out_hptr) {
	if (!out_hptr) return PGQ_ERR_INVALID;
	*out_hptr = nullptr;
	cudaError_t e = cudaHostAlloc(out_hptr, bytes ? bytes : 16, cudaHostAllocDefault);
	if (e != cudaSuccess) {
		set_error("cudaHostAlloc(%llu) -> %s", (unsigned long long)bytes, cudaGetErrorString(e));
		(void)cudaGetLastError();
		return PGQ_ERR_OOM;
	}
	return PGQ_OK;
}

int pgq_host_free_pinned(void *hptr) {
	if (hptr) cudaFreeHost(hptr);
	return PGQ_OK;
}

int pgq_sync(pgq_ctx_t *ctx) {
	LOCK(ctx);
	PGQ_CUDA(cudaStreamSynchronize(c->stream));
	return PGQ_OK;
}

int pgq_gen_rmat_device(pgq_ctx_t *ctx, uint64_t seed, int scale, int64_t e_size, int64_t *d_src, int64_t *d_dst) {
	LOCK(ctx);
	PGQ_TRY(gen_rmat_device(c, seed, scale, e_size, d_src, d_dst));
	PGQ_CUDA(cudaStreamSynchronize(c->stream));
	return PGQ_OK;
}

int pgq_gen_rmat_rows_device(pgq_ctx_t *ctx, uint64_t seed, int scale, int64_t row0, int64_t n, int64_t *d_src,
                             int64_t *d_dst) {
	LOCK(ctx);
	PGQ_TRY(gen_rmat_rows_device(c, seed, scale, row0, n, d_src, d_dst));
	return PGQ_OK;
}

int pgq_gen_knows_device(pgq_ctx_t *ctx, uint64_t seed, int64_t v_size, int64_t e_size, uint64_t mult, uint64_t add,
                         int64_t *d_src, int64_t *d_dst) {
	LOCK(ctx);
	PGQ_TRY(gen_knows_device(c, seed, v_size, e_size, mult, add, d_src, d_dst));
	PGQ_CUDA(cudaStreamSynchronize(c->stream));
	return PGQ_OK;
}

int pgq_probe_gather(pgq_ctx_t *ctx, int64_t nwords, int64_t ngathers, int unroll, double *out_gbs, double *out_ms) {
	LOCK(ctx);
	return probe_gather(c, nwords, ngathers, unroll, out_gbs, out_ms);
}

int pgq_flush_l2(pgq_ctx_t *ctx) {
	LOCK(ctx);
	const size_t bytes = (size_t)256 << 20; // > 126 MB L2
	PGQ_TRY(c->l2_flush.reserve(bytes));
	PGQ_CUDA(cudaMemsetAsync(c->l2_flush.p, 0x5a, bytes, c->stream));
	return PGQ_OK;
}

int pgq_timer_start(pgq_ctx_t *ctx) {
	LOCK(ctx);
	PGQ_CUDA(cudaEventRecord(c->ev_timer, c->stream));
	return PGQ_OK;
}

int pgq_timer_stop_ms(pgq_ctx_t *ctx, double *out_ms) {
	LOCK(ctx);
	PGQ_CUDA(cudaEventRecord(c->ev_timer2, c->stream));
	PGQ_CUDA(cudaEventSynchronize(c->ev_timer2));
	float ms = 0;
	PGQ_CUDA(cudaEventElapsedTime(&ms, c->ev_timer, c->ev_timer2));
	if (out_ms) *out_ms = ms;
	return PGQ_OK;
}

} // extern "C"
