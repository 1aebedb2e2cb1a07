/*
 * TEST INFRASTRUCTURE ONLY — CPU restatement ("oracle") of the DuckPGQ graph hot path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this.  The product path (duckdb-pgq.old_b200/) never links or calls it.
 *
 * PARITY UNPINNED BY THE REFERENCE: /root/reference is a plain DuckDB snapshot that does
 * not contain the PGQ scalar functions.  This file therefore restates the
 * public DuckPGQ algorithm from the rules R1-R9 frozen in SURVEY.md §8(a); each function
 * cites the rule it follows, plus the DuckDB type contracts that *are* in the tree
 * (list_entry_t: src/include/duckdb/common/types.hpp:188-195; 1024-row chunks:
 * src/include/duckdb/common/vector_size.hpp:16-18).  Independent pins that do exist:
 * hand-checked golden graphs (tests/golden/), a textbook queue BFS (orc_bfs_simple), and
 * hop distances computed by the reference DuckDB itself through the recursive-CTE
 * formulation of benchmark/ldbc/queries/bi-10-shortestpath.sql:8-31
 * (tests/golden/make_cte_golden.py).
 */
#ifndef PGQ_ORACLE_H
#define PGQ_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_LANE_LIMIT 512 /* R5: LANE_LIMIT of the reference's std::bitset<512> */
#define ORC_CHUNK 1024     /* STANDARD_VECTOR_SIZE, vector_size.hpp:16-18 */

typedef struct {
	uint64_t batches;      /* MS-BFS batches started */
	uint64_t levels;       /* level steps executed over all batches */
	uint64_t edges_lane;   /* TEPS numerator: sum over levels, over lanes still active at the
	                          start of the level, of out-degrees of that lane's frontier vertices */
	uint64_t edges_unique; /* sum over levels of out-degrees of vertices with >=1 active lane bit */
} orc_stats_t;

/* ---- CSR (R1-R3) ---------------------------------------------------------------- */

/* SQL side of the reference query template: per-vertex out-degree (GROUP BY count). */
void orc_count_degrees(int64_t V, int64_t E, const int64_t *src, int64_t *cnt /*[V]*/);

/* create_csr_vertex rows (R2): v[vid+2] = cnt; returns sum of cnt over the rows. v has V+2 slots. */
int64_t orc_create_csr_vertex(int64_t *v, int64_t n, const int64_t *vid, const int64_t *cnt);

/* First create_csr_edge call (R3): in-place inclusive prefix sum over v[1..V+1]. */
void orc_csr_initialize_edge(int64_t *v, int64_t V);

/* create_csr_edge rows (R3): pos = ++v[src+1]; e[pos-1]=dst; eids[pos-1]=eid (eid==NULL -> row_base+i). */
void orc_create_csr_edge(int64_t *v, int64_t *e, int64_t *eids, int64_t n, const int64_t *src,
                         const int64_t *dst, const int64_t *eid, int64_t row_base);

/* All of the above in input-row order: v[V+2], e[E], eids[E] (eids may be NULL). */
void orc_csr_build(int64_t V, int64_t E, const int64_t *src, const int64_t *dst, const int64_t *eid,
                   int64_t *v, int64_t *e, int64_t *eids);

/* ---- MS-BFS (R4-R9) ------------------------------------------------------------- */

/* iterativelength (R4,R5): one "call" over n rows (the reference sees n<=1024 per call; any n
 * works here, batches of 512 lanes are formed in row order).  valid arrays may be NULL (=all
 * valid), one byte per row.  out_valid[i]=0 -> SQL NULL (out[i] = -1). */
void orc_iterativelength(int64_t V, const int64_t *v, const int64_t *e, int64_t n, const int64_t *src,
                         const int64_t *dst, const uint8_t *src_valid, const uint8_t *dst_valid,
                         int64_t *out, uint8_t *out_valid, orc_stats_t *stats);

/* iterativelength with lower/upper bounds (R11): at most `upper` levels (negative =
 * unbounded); a shortest length below `lower` is NULL. */
void orc_iterativelength_bounded(int64_t V, const int64_t *v, const int64_t *e, int64_t n, const int64_t *src,
                                 const int64_t *dst, const uint8_t *src_valid, const uint8_t *dst_valid,
                                 int64_t lower, int64_t upper, int64_t *out, uint8_t *out_valid,
                                 orc_stats_t *stats);

/* reachability (R6): out[i] in {0,1}; NULL only for NULL inputs. */
void orc_reachability(int64_t V, const int64_t *v, const int64_t *e, int64_t n, const int64_t *src,
                      const int64_t *dst, const uint8_t *src_valid, const uint8_t *dst_valid, uint8_t *out,
                      uint8_t *out_valid, orc_stats_t *stats);

/* shortestpath (R7-R9).  mode 0: [src,...,dst] vertex ids; mode 1: [src,e0,v1,e1,...,dst]
 * interleaved with edge ids (eids==NULL -> CSR edge ids are the input row numbers are
 * unavailable, so mode 1 requires eids).  faithful!=0 runs the reference's literal
 * per-edge 512-lane parent loop (slow; the timed baseline), faithful==0 an equivalent
 * set-bit formulation.  Results: out_entries[i] = {offset,length} (list_entry_t layout),
 * *out_vals malloc'd (free with orc_free), *out_total = number of int64 values. */
typedef struct {
	uint64_t offset;
	uint64_t length;
} orc_list_entry_t;

int orc_shortestpath(int64_t V, const int64_t *v, const int64_t *e, const int64_t *eids, int64_t n,
                     const int64_t *src, const int64_t *dst, const uint8_t *src_valid,
                     const uint8_t *dst_valid, int mode, int faithful, orc_list_entry_t *out_entries,
                     uint8_t *out_valid, int64_t **out_vals, uint64_t *out_total, orc_stats_t *stats);

void orc_free(void *p);

/* Textbook single-source queue BFS: dist[V] (-1 unreachable).  Independent of the bitset code. */
void orc_bfs_simple(int64_t V, const int64_t *v, const int64_t *e, int64_t source, int32_t *dist);

/* ---- weighted next row (R10; pgq_oracle_weighted.c) ------------------- */
/* weights of the input rows -> CSR slot order (raw 8-byte values: int64 or double) */
void orc_csr_weights(int64_t V, int64_t E, const int64_t *v, const int64_t *src, const void *w_in, void *w_out);
/* cheapest_path_length: lane-batched Bellman-Ford; returns 1 if a negative cycle was met */
int orc_cheapest_path_length_i64(int64_t V, const int64_t *v, const int64_t *e, const int64_t *w, int64_t n,
                                 const int64_t *src, const int64_t *dst, const uint8_t *src_valid,
                                 const uint8_t *dst_valid, int64_t *out, uint8_t *out_valid, uint64_t *rounds);
int orc_cheapest_path_length_f64(int64_t V, const int64_t *v, const int64_t *e, const double *w, int64_t n,
                                 const int64_t *src, const int64_t *dst, const uint8_t *src_valid,
                                 const uint8_t *dst_valid, double *out, uint8_t *out_valid, uint64_t *rounds);
/* independent check: binary-heap Dijkstra, non-negative int64 weights; INT64_MAX/2 = unreachable */
void orc_dijkstra_i64(int64_t V, const int64_t *v, const int64_t *e, const int64_t *w, int64_t source,
                      int64_t *dist);

/* ---- multi-threaded driver = the reference's execution shape ----------------------- */
/* Splits the n rows into ORC_CHUNK-row calls handed to nthreads workers (dynamic assignment),
 * like DuckDB's PhysicalProjection under PRAGMA threads (src/parallel/pipeline.cpp:104-123).
 * kind 0 = iterativelength (out = int64[n]), 1 = reachability (out = uint8[n]).
 * Returns wall seconds. */
double orc_run_chunks_mt(int kind, int nthreads, int64_t V, const int64_t *v, const int64_t *e, int64_t n,
                         const int64_t *src, const int64_t *dst, void *out, uint8_t *out_valid,
                         orc_stats_t *stats);

/* ---- generators (pgq_gen.h) ---------------------------------------------------------- */
void orc_gen_rmat(uint64_t seed, int scale, int64_t E, int64_t row0, int64_t nrows, int64_t *src,
                  int64_t *dst, int nthreads);
void orc_gen_knows(uint64_t seed, int64_t V, int64_t E, int64_t row0, int64_t nrows, int64_t *src,
                   int64_t *dst, int nthreads);
/* affine-bijection parameters used by the knows generator for a given V */
void orc_knows_params(int64_t V, uint64_t *mult, uint64_t *add);
/* pairs: uniform over vertices with out-degree > 0 (rejection on v[]); v = CSR offsets (V+2) */
void orc_gen_pairs(uint64_t seed, int64_t V, const int64_t *v, int64_t n, int64_t *src, int64_t *dst);

#ifdef __cplusplus
}
#endif
#endif
