// Shared declarations of the pgq_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/pgq_b200.h"

namespace pgq {

void set_error(const char *fmt, ...);

#define PGQ_CUDA(call)                                                                              \
	do {                                                                                            \
		cudaError_t _e = (call);                                                                    \
		if (_e != cudaSuccess) {                                                                    \
			pgq::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e));    \
			return (_e == cudaErrorMemoryAllocation) ? PGQ_ERR_OOM : PGQ_ERR_CUDA;                   \
		}                                                                                           \
	} while (0)

#define PGQ_TRY(call)                 \
	do {                              \
		int _rc = (call);             \
		if (_rc != PGQ_OK) return _rc; \
	} while (0)

// Small-block cache behind DevBuf (api.cu).  A config-1 sized CSR build (180 k edges) spends 0.1 ms in
// kernels and, without this, 3.5 ms in a dozen cudaMalloc / cudaFree / cudaHostAlloc calls.  Blocks of at
// most 16 MiB are rounded up to a power of two and kept per device when released (up to 512 MiB);
// releasing synchronises the device first, which is what cudaFree did for the callers before.
void *devcache_alloc(size_t bytes, size_t *actual); // nullptr: out of memory (error text set)
void devcache_free(void *p, size_t actual);
void devcache_trim();                               // cudaFree everything cached (context shutdown)
void *pinned_acquire(size_t bytes, size_t *actual); // pinned host staging blocks, same idea
void pinned_release(void *p, size_t actual);

// Owning device buffer (grows, never shrinks unless released).
struct DevBuf {
	void *p = nullptr;
	size_t cap = 0;
	size_t actual = 0; // bytes behind p (>= cap: cached blocks are rounded up)
	int reserve(size_t bytes) {
		if (bytes <= cap) return PGQ_OK;
		release();
		p = devcache_alloc(bytes ? bytes : 16, &actual);
		if (!p) return PGQ_ERR_OOM;
		cap = bytes;
		return PGQ_OK;
	}
	void release() {
		if (p) devcache_free(p, actual);
		p = nullptr;
		cap = 0;
		actual = 0;
	}
	template <class T>
	T *as() const {
		return reinterpret_cast<T *>(p);
	}
};

// One in-flight cheapest_path_length batch (sssp.cu): its own stream, distance rows, changed bitmaps and control block.
struct SsspSlot {
	cudaStream_t st = nullptr;
	cudaEvent_t ev = nullptr;
	DevBuf dist, chg, ctl, acc; // acc: per hub vertex, the minima its in-list chunks found this round + a touched flag
	void destroy() {
		acc.release();
		dist.release();
		chg.release();
		ctl.release();
		if (ev) cudaEventDestroy(ev);
		if (st) cudaStreamDestroy(st);
		ev = nullptr;
		st = nullptr;
	}
};

// Scoped temporary device buffer.
struct TmpBuf : DevBuf {
	~TmpBuf() {
		release();
	}
};

// Hub chunk descriptor: in-edges [begin,end) of hub `hub` (index into hub_list).
struct HubChunk {
	uint32_t hub;
	uint32_t begin;
	uint32_t end;
};

// One CSR (R1): forward arrays in the reference layout plus the reverse CSR used by the
// pull-direction MS-BFS.  Vertex ids and edge positions are 32-bit on the device
// (V < 2^32, E < 2^32 per GPU); download widens to the reference's int64.
struct Csr {
	int64_t V = 0, E = 0;
	enum State { CREATED, EDGES_OPEN, READY } state = CREATED;
	// create_csr_vertex rows (host staging, validated against the device histogram)
	std::vector<int64_t> cnt_host; // size V when any row was given
	bool have_cnt = false;
	// create_csr_edge rows staged in pinned host memory until finalize
	int64_t *h_src = nullptr, *h_dst = nullptr, *h_eid = nullptr;
	size_t h_bytes = 0; // size of each pinned staging block (pinned_acquire)
	int64_t staged = 0, staged_cap = 0;
	bool have_eid = false;
	bool pending_begin = false; // begin_edges deferred until the first row batch tells whether eids exist
	// device edge table (int64, input order) while EDGES_OPEN
	DevBuf d_src64, d_dst64, d_eid64;
	int64_t uploaded = 0;
	bool borrowed_edges = false; // d_src64/d_dst64 point at caller-owned device arrays
	// forward CSR
	DevBuf d_off;  // int64[V+2], reference `v`
	DevBuf d_tgt;  // uint32[E], reference `e` narrowed
	DevBuf d_row;  // uint32[E] input row of each CSR slot (only when edge ids are kept)
	DevBuf d_eid;  // int64[E] edge ids in CSR order (only when caller gave eids)
	DevBuf d_fsrc; // uint32[E] source of each forward slot (sorted keys of the build; consumed by csr_prepare)
	bool keep_rows = false;
	// edge weights (sssp.cu): raw 8-byte values, int64 or double
	int wtype = 0;          // 0 none, PGQ_WEIGHT_I64, PGQ_WEIGHT_F64
	bool w_ready = false;
	int64_t w_rows = 0;     // input rows received so far
	DevBuf d_w_in;          // [E] input-row order while rows arrive
	DevBuf d_w;             // [E] CSR slot order
	DevBuf d_iw;            // [E] reverse-CSR order (lazy)
	DevBuf d_iw16;          // uint16[E] reverse-CSR order, int64 weights clamped to 0x4000 (16-bit distance rows, lazy)
	DevBuf d_iw32n;         // uint32[E] reverse-CSR order, int64 weights clamped to 0x40000000 (32-bit distance rows, lazy)
	DevBuf d_wstat;         // device: weight statistics (sssp.cu WStat)
	TmpBuf d_iwi;           // int64[E] reverse-CSR order: DOUBLE weights that are all whole numbers, as integers (lazy;
	                        // released with the object: every owner deletes the Csr right after free_all)
	bool w_whole = false;   // DOUBLE weights, every one a non-negative whole number <= 2^20: integer rows give the same bits
	bool w_nonneg = false;  // every weight >= 0 (and no NaN): the destination's distance bounds the search
	// reverse CSR (lazy)
	bool prepared = false;
	DevBuf d_ioff;  // uint32[V+1]
	DevBuf d_isrc;  // uint32[E] source of each in-edge, ascending (src, fwd pos) per target
	DevBuf d_ipos;  // uint32[E] forward CSR position of each in-edge
	// in-degree classes, each an ascending vertex list (in-degree 0 is in no list: never a target)
	DevBuf d_tiny;  // 1 <= indeg <= tiny_max          one thread group per vertex, one round
	DevBuf d_small; // tiny_max < indeg <= small_max   two thread groups per vertex
	DevBuf d_med;   // small_max < indeg <= med_max    two groups per vertex on dense levels, a warp on filtered ones
	DevBuf d_mid;   // med_max < indeg <= hub_min      one warp per vertex
	DevBuf d_hub;   // indeg > hub_min                 one block per chunk of the in-list
	DevBuf d_hubchunks; // HubChunk[]
	uint32_t n_tiny = 0, n_small = 0, n_med = 0, n_mid = 0, n_hub = 0, n_hubchunks = 0;
	int64_t bins_tiny_max = -1, bins_small_max = -1, bins_med_max = -1, bins_hub_min = -1;
	uint32_t eff_med_max = 0; // med_max after clamping to >= small_max
	// 1-D partition (part.cu): this object then describes the slice of the reverse CSR owned by the
	// rank — V = owned vertex count (local index space), E = local in-edges, d_isrc = GLOBAL sources
	// identity of the device arrays: bumped whenever they are (re)built, so cached level-loop graphs
	// (msbfs.cu) notice a CSR that was rebuilt under the same id / address
	uint64_t gen = 0;
	uint32_t query_calls = 0; // query calls served so far (device-loop heuristic)
	int device = 0; // the GPU the arrays live on
	// multi-GPU contexts (pgq_init_devices): read-only copies of the query-side arrays on the other
	// devices, made on first use by a multi-device query (msbfs.cu csr_replica)
	std::map<int, Csr *> replicas;
	bool is_part = false;
	int64_t gV = 0;         // global vertex count
	int64_t part_base = 0;  // first owned vertex
	DevBuf d_ltgt, d_lsrc;  // staged local in-edges: local target index, global source
	uint64_t n_local = 0;
	void free_all();
};

struct Options {
	int64_t lanes = 0;
	int64_t stats = 0;
	int64_t timing = 0;
	int64_t tiny_max = 4;
	int64_t small_max = 64;
	int64_t med_max = 512;
	int64_t hub_min = 4096;
	int64_t hub_chunk = 4096;
	int64_t direction = 0; // 0 auto (push when the frontier is sparse), 1 pull only
	int64_t trace = 0;     // with timing=1: per-level line on stderr
	int64_t push_beta = 32; // push when frontier out-edges * push_beta <= E
	int64_t push_vbeta = 16; // ... and frontier vertices * push_vbeta <= V (+ slack for tiny graphs)
	int64_t filter_gamma = 2; // pull tests the frontier bitmap while frontier out-edges * gamma <= E
	int64_t small_edge_mb = 256; // ... while E * W/8 stays below this many MiB (one level's gather volume)
	int64_t small_path = 1;   // whole-batch cooperative kernel for graphs that live in L2 (msbfs_small.cuh)
	int64_t streams = 4;      // general path: batches of one call in flight at once (own stream + buffers each)
	int64_t seg_small = 1;    // filtered levels: small in-degree class through pull_seg_kernel (0 = pull_list_kernel)
	int64_t sort_wide = 1;    // keys of <= 22 bits: two 11-bit passes (radix_sort_wide.cu) instead of 8-bit ones
	int64_t sort_lookback = 1; // two-pass sort: decoupled look-back instead of histogram pre-pass + scan
	int64_t sort_wide_min = 32768; // ... from this many keys on
	int64_t level_cap = 0x7fffffff; // per call, not user-settable: iterativelength's upper bound (R11)
	int64_t loop = 1;          // 1: device-side level loop (CUDA graph WHILE/SWITCH nodes), 0: host loop, one sync per level
	int64_t sssp_narrow = 1;   // cheapest_path_length, non-negative int64 weights: 1 = 16-bit then 32-bit distance rows, 2 = 32-bit only, 0 = 8-byte (all exact)
	int64_t sssp_prune = 1;    // ... drop candidates that cannot beat the destination's distance (non-negative weights)
	int64_t sssp_slots = 4;    // ... lane batches in flight
	int64_t loop_now = 1;      // per call, not user-settable: this call uses the device loop (msbfs.cu use_device_loop)
	int64_t loop_min_batches = 8; // calls with fewer batches on a CSR that has served < loop_min_calls calls take the host loop
	int64_t loop_min_calls = 4;
	int64_t done_bitmap = 0;   // with dense_store=0: tail levels skip vertices whose `seen` already holds every active lane
	int64_t dyn = 1;           // host-driven loop only: 0 = frontier views as kernel arguments instead of from the control block
	int64_t dense_store = 1;   // filtered bottom-up levels write every listed vertex's `next` word too (0: sparse stores after a scrub)
	int64_t l2_persist = 0;    // access-policy window (persisting L2) on `seen` for the batch streams
	int64_t hist_budget_mb = 0; // shortestpath level history per worker: 0 = a quarter of the free HBM
};

// What the two frontier buffers hold after the last completed batch (msbfs.cu run_batch).
struct BfsDirty {
	bool valid = false;
	int W = 0;
	int64_t V = 0;
	const void *visit0 = nullptr, *fbit0 = nullptr;
	// which bitmaps name the non-zero words is in the worker's device control block (BfsCtrl::L)
};

// The level loop of one worker as an instantiated CUDA graph (msbfs.cu): loop_begin -> WHILE { SWITCH(mode)
// { push | pull | pull filtered early | pull filtered late } -> check }.  Valid for one (CSR, lane width,
// buffers) combination; rebuilt when any of them changes.
struct LoopGraph {
	cudaGraph_t graph = nullptr;
	cudaGraphExec_t exec = nullptr;
	const void *g = nullptr;
	uint64_t gen = 0;
	int W = 0;
	bool keep_hist = false;
	const void *seen = nullptr, *ctrl = nullptr, *done = nullptr;
	int64_t seg_small = -1;
	void destroy() {
		if (exec) cudaGraphExecDestroy(exec);
		if (graph) cudaGraphDestroy(graph);
		exec = nullptr;
		graph = nullptr;
		g = nullptr;
	}
};

struct Ctx {
	int device = 0;
	int sm_count = 148;
	cudaStream_t stream = nullptr;
	std::mutex mu;
	std::map<int32_t, Csr *> csrs;
	Options opt;
	pgq_stats_t stats{};
	// scratch reused across calls
	DevBuf scan_tmp;
	DevBuf sort_keys[2], sort_p0[2], sort_p1[2], sort_hist, sort_pack;
	DevBuf bfs_seen, bfs_visit[2], bfs_hubacc;
	std::vector<DevBuf> bfs_hist; // shortestpath: one frontier buffer per level of the running batch
	DevBuf bfs_hist_tab;          // device pointer table of bfs_hist
	DevBuf bfs_pathout;           // path values of one batch before the D2H copy
	DevBuf bfs_queue[2], bfs_inq[2]; // push work queues + dedupe bitmaps
	BfsDirty bfs_dirty;
	// 1-D partition
	int part_rank = 0, part_world = 1;
	void *nccl_comm = nullptr;
	DevBuf part_alive; // [1 + world][16] local / gathered alive masks
	// fused exchange over peer memory (pgq_part_p2p_export / _import)
	DevBuf part_visit[2];     // this rank's replicas of the frontier words, IPC-exported
	DevBuf part_peer_tab;     // device: uint32_t* [2][world-1] peers' replicas, offset to this rank's slice
	std::vector<void *> part_peer_open; // IPC mappings to close at shutdown
	bool part_p2p = false;
	int part_p2p_W = 0;
	int64_t part_p2p_Vp = 0;
	DevBuf bfs_alive; // per-block-slot alive lane masks (msbfs_kernels.cuh)
	DevBuf bfs_lane; // per-call lane tables
	SsspSlot sssp_slot[4]; // cheapest_path_length: up to four lane batches in flight (sssp.cu)
	void *h_sssp = nullptr; // pinned: one SsspCtl mirror per slot
	DevBuf bfs_ctrl; // device control block
	DevBuf l2_flush;
	void *h_ctrl = nullptr; // pinned host mirror of the control block (4 KiB; upper half: small-path level counts)
	int small_pending = 0;  // small-graph batches whose level count has not been folded into stats yet
	cudaEvent_t ev_a = nullptr, ev_b = nullptr, ev_c = nullptr, ev_timer = nullptr, ev_timer2 = nullptr;
	std::vector<cudaEvent_t> ev_pool; // level-kernel timing events
	size_t ev_used = 0;
	// batch-level concurrency (msbfs.cu run_batches): helper contexts with their own stream, control
	// block and frontier buffers; they share the caller's CSRs read-only and hold none of their own
	std::vector<Ctx *> helpers;
	cudaEvent_t ev_fork = nullptr;
	// multi-GPU context (pgq_init_devices): devices[0] == device; queries spread their batches over all
	// of them (CSR replicated), everything else runs on the primary
	std::vector<int> devices;
	std::vector<int> devices_all; // as given to pgq_init_devices (option "devices_used" narrows `devices`)
	DevBuf bfs_done;        // 1 bit per vertex: seen holds every active lane (msbfs_kernels.cuh)
	LoopGraph loop_graph[2]; // [keep_hist]
	bool loop_unsupported = false; // conditional graph nodes refused by the driver: host loop from then on
	uint64_t tot_seen[3] = {0, 0, 0}; // device-side running totals already folded into stats (levels, push, launches)
	bool ctrl_fresh = true;  // the control block has not been zeroed since it was allocated
	void *h_found = nullptr; // pinned staging of the per-call found[] table
	size_t h_found_cap = 0;
	size_t hist_bytes = 0;   // bytes behind bfs_hist
	uint32_t loop_nk[2][8] = {}; // kernels per level of each mode in loop_graph[i]
	int l2_window_on = -1;   // access-policy window currently set on the stream (-1 unknown)
	const void *l2_window_ptr = nullptr;
};

int ctx_init_resources(Ctx *c);    // stream, events, pinned control mirror (api.cu)
void ctx_release_resources(Ctx *c); // everything a Ctx owns except its CSRs (api.cu)
const char *last_error_str();

inline void count_launch(Ctx *c, uint64_t n = 1) {
	c->stats.kernel_launches += n;
}

static inline uint32_t div_up(uint64_t a, uint64_t b) {
	return (uint32_t)((a + b - 1) / b);
}

// ---- scan.cu ----
// exclusive scan of n uint32 values into uint32 (out may alias in); total written to *d_total (u64) if non-null
int exclusive_scan_u32(Ctx *c, const uint32_t *d_in, uint32_t *d_out, uint64_t n, uint64_t *d_total);
// exclusive scan of n uint32 values into int64 out with a +0 base
int exclusive_scan_u32_to_i64(Ctx *c, const uint32_t *d_in, int64_t *d_out, uint64_t n);

// ---- radix_sort.cu ----
// Sort description (host).
struct SortIO {
	const int64_t *k64 = nullptr, *p64 = nullptr; // 64-bit first-pass inputs, or
	const uint32_t *k32 = nullptr, *p32 = nullptr; // 32-bit first-pass inputs
	uint32_t n = 0;
	int key_bits = 1;
	uint32_t *out_p0 = nullptr; // payload0 in sorted order
	uint32_t *out_p1 = nullptr; // original index in sorted order (nullptr = not wanted)
	uint32_t *out_keys = nullptr; // sorted keys (nullptr = not wanted)
	const uint32_t **keys_scratch = nullptr; // or: sorted keys left in the context's free ping-pong buffer
	// first-pass validation of 64-bit inputs: sets *err |= 1 when a key or payload is outside [0, vmax)
	uint64_t vmax = 0;
	uint32_t *err = nullptr;
};

int radix_sort_stable(Ctx *c, const SortIO &io);
// radix_sort_wide.cu: the same sort in two passes of <= 11-bit digits (keys of at most 22 bits)
bool radix_sort_wide_eligible(const Ctx *c, const SortIO &io);
int radix_sort_wide(Ctx *c, const SortIO &io);

// ---- csr_build.cu ----
int csr_set_weights(Ctx *c, Csr *g, int wtype, const void *w, int64_t row0, int64_t n); // sssp.cu
int csr_download_weights(Ctx *c, Csr *g, void *out);
int csr_device_build(Ctx *c, Csr *g);   // K1-K3 from d_src64/d_dst64/d_eid64
int csr_prepare(Ctx *c, Csr *g);        // reverse CSR + degree bins

// ---- msbfs.cu ----
struct LaneTable {
	// host-side compact description of the lanes of all batches of one call
	std::vector<uint32_t> src, dst;   // per lane slot (batch*W + lane); 0xFFFFFFFF = unused
	std::vector<int64_t> row;         // row index of the lane slot, -1 unused
	std::vector<int64_t> batch_row_begin, batch_row_end; // rows consumed by each batch [begin,end)
	std::vector<uint8_t> kind;        // per row: 0 NULL, 1 trivial (src==dst), 2 searched in a lane
	int W = 0;
	int64_t nbatch = 0;
};
static inline int pick_lane_width(int64_t lanes_needed) {
	return lanes_needed <= 64 ? 64 : lanes_needed <= 128 ? 128 : lanes_needed <= 256 ? 256 : 512;
}
// Runs all batches; found[slot] = level at which dst was seen (or -1).  If want_paths, path values
// are produced per batch through the callback-free interface in paths.cu (see msbfs.cu).
struct PathOut {
	int mode = 0;
	pgq_list_entry_t *entries = nullptr; // per row
	uint8_t *valid = nullptr;            // per row
	std::vector<int64_t> vals;
	// when set: asked once for the final buffer of `total` values (the caller's list child buffer); vals stays empty
	int64_t *(*sink)(void *user, uint64_t total) = nullptr;
	void *sink_user = nullptr;
	uint64_t total = 0;
};
int msbfs_run(Ctx *c, Csr *g, const LaneTable &lt, int32_t *h_found, PathOut *paths,
              const int64_t *row_src /*host, for trivial rows in path mode*/, int64_t nrows);
int msbfs_run_device(Ctx *c, Csr *g, int64_t n, const int64_t *d_src, const int64_t *d_dst, int64_t *d_out,
                     uint8_t *d_out_valid);
// partitioned variant (collective): g is the rank's slice
int msbfs_part_run(Ctx *c, Csr *g, const LaneTable &lt, int32_t *h_found);
// cheapest_path_length batches (sssp.cu): h_out8[slot] = distance (int64 or double by g->wtype), +inf = unreached
int sssp_run(Ctx *c, Csr *g, const LaneTable &lt, void *h_out8);

// ---- part.cu ----
int part_add_edges_device(Ctx *c, Csr *g, const int64_t *d_src, const int64_t *d_dst, int64_t n);
int part_finalize(Ctx *c, Csr *g);
int part_offsets_u32(Ctx *c, const uint32_t *keys, uint32_t n, uint32_t nv, uint32_t *off); // csr_build.cu
// Ownership of the 1-D partition.  Vertices are grouped in 32-vertex blocks (one frontier-bitmap
// word); the block index is scrambled by a bijection on k-bit integers (k = ceil(log2(#blocks)))
// and dealt round-robin: owner = mix(blk) % G, local block = mix(blk) / G.  Plain ranges or plain
// block-cyclic dealing leave R-MAT 3:1 unbalanced between ranks because in-degree correlates with
// every id bit.  Internally vertices are renumbered rank-major, new(v) = owner*Vp + local(v), so a
// rank's slice of the replicated frontier structures is contiguous (in-place allgather) and the
// stored sources index those structures directly.  Non-power-of-two block counts pad to 2^k.
struct PartMap {
	uint32_t G = 1;
	uint32_t kbits = 0;
	uint64_t Vp = 0; // vertices per rank in the replicated structures (multiple of 256)
};
static inline PartMap part_map(int64_t V, int world) {
	PartMap m;
	m.G = (uint32_t)world;
	const uint64_t nblk = (uint64_t)(V + 31) / 32;
	while (((uint64_t)1 << m.kbits) < nblk) m.kbits++;
	const uint64_t lblk = (((uint64_t)1 << m.kbits) + m.G - 1) / m.G;
	m.Vp = (lblk * 32 + 255) / 256 * 256;
	return m;
}
__host__ __device__ static inline uint64_t part_mix(uint64_t blk, uint32_t k) {
	if (k < 2) return blk;
	const uint64_t mask = ((uint64_t)1 << k) - 1;
	const uint32_t h = (k + 1) / 2;
	uint64_t x = blk;
	x ^= x >> h;
	x = (x * 0x9E3779B97F4A7C15ull) & mask;
	x ^= x >> h;
	x = (x * 0xC2B2AE3D27D4EB4Full) & mask;
	x ^= x >> h;
	return x;
}
__host__ __device__ static inline uint32_t part_owner(uint64_t v, const PartMap &m) {
	return (uint32_t)(part_mix(v >> 5, m.kbits) % m.G);
}
__host__ __device__ static inline uint32_t part_local(uint64_t v, const PartMap &m) {
	return (uint32_t)(((part_mix(v >> 5, m.kbits) / m.G) << 5) | (v & 31u));
}
__host__ __device__ static inline uint32_t part_newid(uint64_t v, const PartMap &m) {
	return (uint32_t)(part_owner(v, m) * m.Vp + part_local(v, m));
}
int part_add_edges_device(Ctx *c, Csr *g, const int64_t *d_src, const int64_t *d_dst, int64_t n);
int part_finalize(Ctx *c, Csr *g);
int part_offsets_u32(Ctx *c, const uint32_t *keys, uint32_t n, uint32_t nv, uint32_t *off); // csr_build.cu
// ---- gen.cu ----
int gen_rmat_device(Ctx *c, uint64_t seed, int scale, int64_t E, int64_t *d_src, int64_t *d_dst);
int gen_rmat_rows_device(Ctx *c, uint64_t seed, int scale, int64_t row0, int64_t n, int64_t *d_src, int64_t *d_dst);
int probe_gather(Ctx *c, int64_t nwords, int64_t ngathers, int unroll, double *out_gbs, double *out_ms);
int gen_knows_device(Ctx *c, uint64_t seed, int64_t V, int64_t E, uint64_t mult, uint64_t add, int64_t *d_src,
                     int64_t *d_dst);

} // namespace pgq

struct pgq_ctx {
	pgq::Ctx c;
};
