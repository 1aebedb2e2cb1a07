/*
 * pgq_b200 — C ABI of the B200-native DuckPGQ graph hot path.
 *
 * This is the drop-in boundary: plain C types, caller-owned buffers, int return codes
 * (0 = ok, <0 = error; text via pgq_last_error()).  No torch / C++ types cross it.
 * The DuckDB extension (duckdb-pgq.old_b200/ext/pgq_b200_extension.cpp) and the Python
 * host (duckdb-pgq.old_b200/__init__.py, ctypes) are both thin callers of these entry
 * points.  There is NO CPU fallback: every compute entry point fails with PGQ_ERR_CUDA
 * when no sm_100 device / driver is present.
 *
 * What each entry point replaces.  The PGQ scalar functions themselves are not in the
 * mounted reference; their semantics are rules R1-R9 of SURVEY.md §8(a).
 * The DuckDB-side contracts they plug into ARE in the tree and are cited per function:
 *   scalar_function_t     /root/reference/src/include/duckdb/function/scalar_function.hpp:24
 *   one call = one chunk  /root/reference/src/execution/expression_executor/execute_function.cpp:19-40
 *   <=1024 rows per call  /root/reference/src/include/duckdb/common/vector_size.hpp:16-18
 *   list_entry_t layout   /root/reference/src/include/duckdb/common/types.hpp:188-195
 *   validity = 1 bit/row  /root/reference/src/include/duckdb/common/types/validity_mask.hpp:151-156,190-196
 *                         (the shim expands/packs bits to the one-byte-per-row arrays used here)
 */
#ifndef PGQ_B200_H
#define PGQ_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PGQ_OK 0
#define PGQ_ERR_INVALID (-1)  /* bad argument / unknown csr id / id out of range */
#define PGQ_ERR_CUDA (-2)     /* CUDA runtime error or no device */
#define PGQ_ERR_OOM (-3)      /* device or host allocation failed */
#define PGQ_ERR_STATE (-4)    /* call order violated (e.g. query before finalize) */
#define PGQ_ERR_MISMATCH (-5) /* create_csr_vertex counts disagree with the edge rows */

typedef struct pgq_ctx pgq_ctx_t;

/* list_entry_t {uint64 offset; uint64 length} — types.hpp:188-195 */
typedef struct {
	uint64_t offset;
	uint64_t length;
} pgq_list_entry_t;

typedef struct {
	uint64_t kernel_launches; /* kernels of this library launched since the last reset */
	uint64_t batches;         /* MS-BFS batches */
	uint64_t levels;          /* level steps */
	uint64_t push_levels;     /* of which top-down (sparse frontier) */
	uint64_t edges_lane;      /* TEPS numerator (only when option "stats"=1), SURVEY.md §8(d) */
	uint64_t edges_unique;    /* unique edge visits (only when option "stats"=1) */
	uint64_t algo_bytes;      /* algorithmic bytes of the level kernels, §8(d) formula ("stats"=1) */
	double level_kernel_ms;   /* CUDA-event time of the level (pull/push) kernels ("timing"=1) */
	uint64_t level_kernel_launches;
	double csr_build_ms;      /* CUDA-event time of the last device CSR build (K1-K3) */
	double transpose_ms;      /* CUDA-event time of the last reverse-CSR build */
	uint64_t h2d_bytes;       /* bytes copied host->device by this library since reset */
	uint64_t d2h_bytes;
	uint64_t coll_bytes;      /* bytes this rank contributed to per-level collectives (partitioned path) */
	double coll_ms;           /* CUDA-event time of those collectives ("timing"=1) */
	/* what the level kernels of THIS implementation touched ("stats"=1, instrumented instantiations;
	 * DESIGN.md section 4 turns these into model_bytes, the numerator of bench.py's roofline.frac) */
	uint64_t model_bytes;     /* bytes this algorithm must move for the levels it ran (no cache credit) */
	uint64_t k_listed;        /* pull: class-list entries visited */
	uint64_t k_open;          /* pull: vertices whose `seen` word was read (not yet done) */
	uint64_t k_new;           /* pull: vertices that stored a `next` word and rewrote `seen` */
	uint64_t k_scanned;       /* pull: in-edges whose source id was read */
	uint64_t k_gathered;      /* pull: frontier words gathered */
	uint64_t kp_items;        /* push: work items */
	uint64_t kp_edges;        /* push: out-edges whose target id was read */
	uint64_t kp_seen;         /* push: 4-byte atomicOr on `seen` */
	uint64_t kp_next;         /* push: 4-byte atomicOr on `next` */
	uint64_t host_syncs;      /* stream synchronisations issued inside MS-BFS batches (level loop) */
	/* cheapest_path_length (levels = relaxation rounds; "stats"=1 also fills k_scanned / k_gathered with the
	 * in-edges scanned and the source rows read) */
	uint64_t sssp_narrow_batches; /* lane batches answered on 16-bit distance rows */
	uint64_t sssp_redo_batches;   /* batches that did not fit their narrow rows and were redone on wider ones */
	uint64_t sssp_narrow32_batches; /* lane batches answered on 32-bit distance rows */
} pgq_stats_t;

/* ---- context ------------------------------------------------------------------------- */
int pgq_init(int device, pgq_ctx_t **out_ctx);
/* Multi-GPU context for a single-process host (DuckDB is one process with worker threads:
 * /root/reference/src/parallel/task_scheduler.cpp:163-195).  devices[0] is the primary: CSRs are built
 * there; the first query that has enough batches copies the query-side arrays to the other devices over
 * NVLink (peer copies) and from then on every iterativelength / reachability / shortestpath call deals
 * its 512-lane batches round-robin over all devices (source-batch sharding, no data-path collective).
 * Same entry points, same results as a one-device context. */
int pgq_init_devices(const int *devices, int n, pgq_ctx_t **out_ctx);
int pgq_device_count(pgq_ctx_t *ctx, int *out_n);
int pgq_shutdown(pgq_ctx_t *ctx);
const char *pgq_last_error(void); /* thread-local message of the last failing call */

/* options (cheapest_path_length): "sssp_narrow" (non-negative int64 weights; 1: 16-bit distance rows, batches that do
 *          not fit are redone on 32-bit rows, then on 8-byte rows; 2: 32-bit rows first; 0: 8-byte rows; all exact), "sssp_prune" (1: with non-negative weights, drop
 *          candidates that cannot beat the destination's distance), "sssp_slots" (1..4 lane batches in flight).
 * options: "lanes" (0=auto, 64,128,256,512), "stats" (0/1), "timing" (0/1),
 *          "small_max", "hub_min" (in-degree bin thresholds), "direction" (0 auto, 1 pull only),
 *          "push_beta" (top-down when frontier out-edges * beta <= E),
 *          "loop" (1 = level loop on the device as a CUDA graph with WHILE/SWITCH nodes, 0 = host loop),
 *          "streams" (batches of one call in flight per device), "devices_used" (multi-GPU contexts),
 *          "done_bitmap", "l2_persist" (measurement switches), "hist_budget_mb" (shortestpath level history) */
int pgq_set_option(pgq_ctx_t *ctx, const char *key, int64_t value);
int pgq_get_stats(pgq_ctx_t *ctx, pgq_stats_t *out);
int pgq_reset_stats(pgq_ctx_t *ctx);

/* ---- CSR lifecycle: mirrors create_csr_vertex / create_csr_edge (R1-R3) ------------------ */
/* CsrInitializeVertex: allocate v[V+2] zero-filled for csr `id` (first create_csr_vertex call). */
int pgq_csr_create(pgq_ctx_t *ctx, int32_t id, int64_t v_size);
/* create_csr_vertex rows: v[vid+2] = cnt; *out_sum += sum(cnt) (what the SQL caller SUMs). */
int pgq_csr_set_degrees(pgq_ctx_t *ctx, int32_t id, const int64_t *vid, const int64_t *cnt, int64_t n,
                        int64_t *out_sum);
/* CsrInitializeEdge: e.resize(E) + prefix sum (first create_csr_edge call). */
int pgq_csr_begin_edges(pgq_ctx_t *ctx, int32_t id, int64_t e_size);
/* create_csr_edge rows, appended in ARRIVAL order (= the stable key; R3).  eid may be NULL. */
int pgq_csr_add_edges(pgq_ctx_t *ctx, int32_t id, const int64_t *src, const int64_t *dst,
                      const int64_t *eid, int64_t n);
/* Runs K1 (degree histogram) + K2 (exclusive scan) + K3 (stable scatter) on the device. */
int pgq_csr_finalize(pgq_ctx_t *ctx, int32_t id);

#define PGQ_CSR_ROWIDS 1 /* eid==NULL: edge ids = input row numbers (kept for mode-1 paths/download) */
/* One-shot build from HOST arrays (create + begin + add + finalize). */
int pgq_csr_build(pgq_ctx_t *ctx, int32_t id, int64_t v_size, int64_t e_size, const int64_t *src,
                  const int64_t *dst, const int64_t *eid, int flags);
/* One-shot build from DEVICE int64 arrays (inputs already resident in HBM). */
int pgq_csr_build_device(pgq_ctx_t *ctx, int32_t id, int64_t v_size, int64_t e_size, const int64_t *d_src,
                         const int64_t *d_dst, const int64_t *d_eid, int flags);
/* Reference-layout arrays: v[V+2], e[E], eid[E] (any pointer may be NULL to skip). */
int pgq_csr_download(pgq_ctx_t *ctx, int32_t id, int64_t *v, int64_t *e, int64_t *eid);
int pgq_csr_info(pgq_ctx_t *ctx, int32_t id, int64_t *v_size, int64_t *e_size);
int pgq_csr_delete(pgq_ctx_t *ctx, int32_t id);
/* Build the reverse CSR + degree bins now (otherwise done lazily by the first query). */
int pgq_csr_prepare(pgq_ctx_t *ctx, int32_t id);

/* ---- queries: iterativelength / reachability / shortestpath (R4-R9) ------------------------ */
/* HOST buffers, one byte per row validity (NULL = all valid).  out_valid[i]==0 -> SQL NULL. */
int pgq_iterativelength(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *src, const int64_t *dst,
                        const uint8_t *src_valid, const uint8_t *dst_valid, int64_t *out,
                        uint8_t *out_valid);
/* 6-argument iterativelength(csr_id, v_size, src, dst, lower, upper) of the later DuckPGQ code (SURVEY.md
 * §8(f) f1; rule R11 in DESIGN.md §4, [RECALL] -- not present in /root/reference): the search stops after
 * `upper` levels (negative = unbounded); a pair whose shortest length is below `lower` gives NULL;
 * src == dst gives 0 when lower <= 0, else NULL.  Not available on a vertex-partitioned csr. */
int pgq_iterativelength_bounded(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *src, const int64_t *dst,
                                const uint8_t *src_valid, const uint8_t *dst_valid, int64_t lower, int64_t upper,
                                int64_t *out, uint8_t *out_valid);
int pgq_reachability(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *src, const int64_t *dst,
                     const uint8_t *src_valid, const uint8_t *dst_valid, uint8_t *out, uint8_t *out_valid);
/* mode 0: [src,...,dst]; mode 1: [src,e0,v1,e1,...,dst].  *out_vals is allocated by the
 * library (pgq_free), *out_total = number of int64 values; entries[i] = list_entry_t. */
int pgq_shortestpath(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *src, const int64_t *dst,
                     const uint8_t *src_valid, const uint8_t *dst_valid, int mode,
                     pgq_list_entry_t *entries, uint8_t *out_valid, int64_t **out_vals,
                     uint64_t *out_total);
void pgq_free(void *p);
/* The same, with the values written straight into a buffer of the caller: once the total number of values is
 * known `reserve(user, total)` is called exactly once (also for total == 0) and returns where the `total`
 * int64 values go -- the DuckDB shim answers with ListVector::Reserve + the child vector's data pointer
 * (/root/reference/src/function/scalar/list/range.cpp:118-149), so no intermediate malloc / memcpy. */
typedef int64_t *(*pgq_reserve_fn)(void *user, uint64_t total);
int pgq_shortestpath_sink(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *src, const int64_t *dst,
                          const uint8_t *src_valid, const uint8_t *dst_valid, int mode,
                          pgq_list_entry_t *entries, uint8_t *out_valid, pgq_reserve_fn reserve, void *user,
                          uint64_t *out_total);

/* ---- weighted next row ------------------------------------------------------
 * Replaces the 7-argument create_csr_edge(csr_id, v_size, num_edges, src, dst, edge_rowid, weight)
 * and cheapest_path_length(csr_id, v_size, src, dst) of DuckPGQ ([RECALL], rule R10 in
 * DESIGN.md section 5b; not present in /root/reference).  Weights are given per INPUT ROW
 * (the order the rows went through pgq_csr_add_edges), in one or more consecutive chunks after
 * pgq_csr_finalize; the CSR must track rows (edge ids given, or PGQ_CSR_KEEP_ROWS).  The library
 * moves them into the edge's CSR slot.  wtype: BIGINT weights give BIGINT lengths, DOUBLE gives DOUBLE. */
#define PGQ_WEIGHT_I64 1
#define PGQ_WEIGHT_F64 2
int pgq_csr_set_weights(pgq_ctx_t *ctx, int32_t id, int wtype, const void *w, int64_t row0, int64_t n);
/* weights in CSR slot order (8 bytes each), and their type (0 = none) */
int pgq_csr_download_weights(pgq_ctx_t *ctx, int32_t id, void *out);
int pgq_csr_weight_type(pgq_ctx_t *ctx, int32_t id, int *wtype);
/* Lane-batched Bellman-Ford (64 lanes per batch).  out: int64[n] or double[n] by the weight type;
 * out_valid[i] = 0 (SQL NULL) for NULL inputs and unreachable destinations; src == dst gives 0.
 * A negative cycle reachable from a source is an error (the reference would not terminate). */
int pgq_cheapest_path_length(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *src, const int64_t *dst,
                             const uint8_t *src_valid, const uint8_t *dst_valid, void *out, uint8_t *out_valid);

/* DEVICE-resident variant (all rows valid): d_out int64[n] (-1 = no path), d_out_valid uint8[n]. */
int pgq_iterativelength_device(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *d_src,
                               const int64_t *d_dst, int64_t *d_out, uint8_t *d_out_valid);

/* Allocation-free variant of pgq_iterativelength_device.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same arguments and results as pgq_iterativelength_device, but every byte of
 * scratch (control block, lane tables, `seen`, both frontier buffers) lives in one file-scope static __device__
 * array of 8 MiB (csrc/noalloc.cu) that every GPU loading the library holds until exit; the call performs no
 * cudaMalloc / cudaFree / cudaHostAlloc.  Limits, rejected with an error naming them: n <= 4096 rows;
 * 64 KiB + 3 * roundup256(V * lanes/8) <= 8 MiB (lanes = 64..512 by n, or option "lanes"): V <= ~43 000 at 512
 * lanes, ~348 000 at 64; the csr must have been prepared before (pgq_csr_prepare or any earlier query -- that
 * step allocates).  The array is shared by all calls on a device and keeps the last call's contents (what must
 * be zero is cleared per batch); calls that overlap on the GPU (two contexts on one device) race on it. */
int pgq_iterativelength_device_noalloc(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *d_src,
                                       const int64_t *d_dst, int64_t *d_out, uint8_t *d_out_valid);

/* ---- 1-D vertex partition across GPUs (BASELINE.json configs[4]) -------------------------------
 * One process per GPU.  32-vertex blocks are dealt to the ranks by a scrambling bijection of the
 * block index (plain ranges or plain block-cyclic dealing leave R-MAT ranks 3:1 unbalanced); a rank
 * owns its blocks' vertices as targets, together with
 * their in-edges (a slice of the reverse CSR).  The W-bit
 * frontier words and the frontier bitmap are replicated; after every bottom-up level each rank's
 * slice is exchanged with ncclAllGather over NVLink.  `seen` exists only for owned vertices.
 * Every function below is COLLECTIVE: all ranks call it with identical arguments; query results
 * are produced on every rank.  Lengths / reachability only (no parents). */
int pgq_nccl_unique_id(void *out128); /* rank 0; the caller broadcasts the 128 bytes */
int pgq_part_init(pgq_ctx_t *ctx, int rank, int world, const void *unique_id128);
int pgq_part_range(pgq_ctx_t *ctx, int64_t v_size, int64_t *out_base, int64_t *out_count);
int pgq_part_csr_create(pgq_ctx_t *ctx, int32_t id, int64_t v_size);
/* Rows of the GLOBAL edge table in arrival order (device int64 arrays).  Every rank is shown every
 * row (or at least the rows whose dst it owns); rows owned by other ranks are dropped. */
int pgq_part_csr_add_edges_device(pgq_ctx_t *ctx, int32_t id, const int64_t *d_src, const int64_t *d_dst,
                                  int64_t n);
int pgq_part_csr_finalize(pgq_ctx_t *ctx, int32_t id);
/* Fused exchange (optional, world > 1): the level kernels store new frontier words straight into
 * every rank's replica through NVLink peer memory instead of a per-level ncclAllGather of the word
 * array; only the 1-bit frontier bitmap and the 64-byte alive masks still go through NCCL (which
 * doubles as the level barrier).  export: allocates this rank's replicas for (csr id, lanes) and
 * writes 2 x 64 bytes of CUDA IPC handles; the caller all-gathers them (rank-major) and every rank
 * calls import with the world x 2 x 64 bytes. */
#define PGQ_P2P_HANDLE_BYTES 128
int pgq_part_p2p_export(pgq_ctx_t *ctx, int32_t id, int lanes, void *out_handles128);
/* Before a SECOND export (another lane width / csr): every rank closes its mappings of the peers' old replicas, then
 * the caller barriers, then everybody exports again (an exporter must not free memory a peer still maps). */
int pgq_part_p2p_close(pgq_ctx_t *ctx);
int pgq_part_p2p_import(pgq_ctx_t *ctx, const void *all_handles);
int pgq_part_iterativelength(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *src, const int64_t *dst,
                             const uint8_t *src_valid, const uint8_t *dst_valid, int64_t *out,
                             uint8_t *out_valid);
int pgq_part_reachability(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *src, const int64_t *dst,
                          const uint8_t *src_valid, const uint8_t *dst_valid, uint8_t *out, uint8_t *out_valid);

/* ---- bench / test helpers (not part of the reference surface) ------------------------------ */
int pgq_dev_alloc(pgq_ctx_t *ctx, uint64_t bytes, void **out_dptr);
int pgq_dev_free(pgq_ctx_t *ctx, void *dptr);
int pgq_dev_h2d(pgq_ctx_t *ctx, void *dptr, const void *hptr, uint64_t bytes);
int pgq_dev_d2h(pgq_ctx_t *ctx, void *hptr, const void *dptr, uint64_t bytes);
int pgq_host_alloc_pinned(uint64_t bytes, void **out_hptr);
int pgq_host_free_pinned(void *hptr);
int pgq_sync(pgq_ctx_t *ctx);
/* synthetic edge tables written straight into HBM (same streams as oracle/pgq_gen.h) */
int pgq_gen_rmat_device(pgq_ctx_t *ctx, uint64_t seed, int scale, int64_t e_size, int64_t *d_src,
                        int64_t *d_dst);
/* rows [row0, row0+n) of the same R-MAT stream (chunked generation for graphs larger than a buffer) */
int pgq_gen_rmat_rows_device(pgq_ctx_t *ctx, uint64_t seed, int scale, int64_t row0, int64_t n, int64_t *d_src,
                             int64_t *d_dst);
int pgq_gen_knows_device(pgq_ctx_t *ctx, uint64_t seed, int64_t v_size, int64_t e_size, uint64_t mult,
                         uint64_t add, int64_t *d_src, int64_t *d_dst);
/* write >L2-sized scratch so the next timed step starts with a cold L2 */
/* Measurement probe (bench only): `ngathers` random 64-byte gathers (4 threads x 16 bytes, `unroll` in flight
 * per thread, indices streamed) from a table of `nwords` words -- the dense bottom-up level's access pattern
 * without its arithmetic.  *out_gbs counts 68 bytes per gather (word + index). */
int pgq_probe_gather(pgq_ctx_t *ctx, int64_t nwords, int64_t ngathers, int unroll, double *out_gbs, double *out_ms);
int pgq_flush_l2(pgq_ctx_t *ctx);
/* CUDA-event stopwatch on the library's stream */
int pgq_timer_start(pgq_ctx_t *ctx);
int pgq_timer_stop_ms(pgq_ctx_t *ctx, double *out_ms);

#ifdef __cplusplus
}
#endif
#endif
