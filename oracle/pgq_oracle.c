/*
 * TEST INFRASTRUCTURE ONLY — see pgq_oracle.h.  CPU restatement of the DuckPGQ hot path
 * following rules R1-R9 of SURVEY.md §8(a).  PARITY UNPINNED BY THE REFERENCE (the PGQ
 * functions are not in /root/reference).
 *
 * Deliberately written the way the reference runs: dense O(V) sweeps per level over
 * std::bitset<512>-shaped words, int64 CSR arrays, one <=1024-row call at a time.
 */
#define _GNU_SOURCE
#include "pgq_oracle.h"
#include "../include/pgq_gen.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "pgq_oracle_internal.h"

/* ---------------------------------------------------------------------------------- */
/* CSR: R1-R3                                                                          */
/* ---------------------------------------------------------------------------------- */

void orc_count_degrees(int64_t V, int64_t E, const int64_t *src, int64_t *cnt) {
	memset(cnt, 0, (size_t)V * sizeof(int64_t));
	for (int64_t i = 0; i < E; i++) cnt[src[i]]++;
}

/* R2: create_csr_vertex — per row v[src+2] = cnt; the SQL caller SUMs the returned cnt. */
int64_t orc_create_csr_vertex(int64_t *v, int64_t n, const int64_t *vid, const int64_t *cnt) {
	int64_t sum = 0;
	for (int64_t i = 0; i < n; i++) {
		v[vid[i] + 2] = cnt[i];
		sum += cnt[i];
	}
	return sum;
}

/* R3 (first create_csr_edge call): for i in 1..V+1: v[i] += v[i-1]  =>  v[i+1] = start(i). */
void orc_csr_initialize_edge(int64_t *v, int64_t V) {
	for (int64_t i = 1; i < V + 2; i++) v[i] += v[i - 1];
}

/* R3 per row: pos = ++v[src+1]; e[pos-1] = dst; edge_ids[pos-1] = edge_rowid. */
void orc_create_csr_edge(int64_t *v, int64_t *e, int64_t *eids, int64_t n, const int64_t *src,
                         const int64_t *dst, const int64_t *eid, int64_t row_base) {
	for (int64_t i = 0; i < n; i++) {
		int64_t pos = ++v[src[i] + 1];
		e[pos - 1] = dst[i];
		if (eids) eids[pos - 1] = eid ? eid[i] : row_base + i;
	}
}

void orc_csr_build(int64_t V, int64_t E, const int64_t *src, const int64_t *dst, const int64_t *eid,
                   int64_t *v, int64_t *e, int64_t *eids) {
	int64_t *cnt = (int64_t *)malloc((size_t)(V > 0 ? V : 1) * sizeof(int64_t));
	int64_t *vid = (int64_t *)malloc((size_t)(V > 0 ? V : 1) * sizeof(int64_t));
	orc_count_degrees(V, E, src, cnt);
	for (int64_t i = 0; i < V; i++) vid[i] = i;
	memset(v, 0, (size_t)(V + 2) * sizeof(int64_t)); /* R1: v[V+2] zero-filled */
	orc_create_csr_vertex(v, V, vid, cnt);
	orc_csr_initialize_edge(v, V);
	orc_create_csr_edge(v, e, eids, E, src, dst, eid, 0);
	free(cnt);
	free(vid);
}

/* ---------------------------------------------------------------------------------- */
/* MS-BFS level step: R4/R5                                                            */
/* ---------------------------------------------------------------------------------- */

/* One level: next=0; for every v with visit[v].any(): next[e[off]] |= visit[v];
 * then next &= ~seen; seen |= next; change |= next.any().
 * `active` (lane mask of not-yet-retired lanes) only feeds the stats counters. */
static int level_step(int64_t V, const int64_t *v, const int64_t *e, bs512 *seen, const bs512 *visit,
                      bs512 *next, const bs512 *active, orc_stats_t *st) {
	int change = 0;
	memset(next, 0, (size_t)V * sizeof(bs512));
	for (int64_t i = 0; i < V; i++) {
		if (!bs_any(&visit[i])) continue;
		const bs512 vi = visit[i];
		if (st) {
			uint64_t pc = 0;
			for (int k = 0; k < WORDS; k++) pc += (uint64_t)__builtin_popcountll(vi.w[k] & active->w[k]);
			uint64_t deg = (uint64_t)(v[i + 1] - v[i]);
			st->edges_lane += deg * pc;
			if (pc) st->edges_unique += deg;
		}
		for (int64_t off = v[i]; off < v[i + 1]; off++) {
			bs512 *nx = &next[e[off]];
			for (int k = 0; k < WORDS; k++) nx->w[k] |= vi.w[k];
		}
	}
	for (int64_t i = 0; i < V; i++) {
		uint64_t any = 0;
		for (int k = 0; k < WORDS; k++) {
			uint64_t nw = next[i].w[k] & ~seen[i].w[k];
			next[i].w[k] = nw;
			seen[i].w[k] |= nw;
			any |= nw;
		}
		change |= (any != 0);
	}
	if (st) st->levels++;
	return change;
}

/* Shared driver for iterativelength (kind 0) and reachability (kind 1). */
static void msbfs_lengths(int kind, int64_t V, const int64_t *v, const int64_t *e, int64_t n,
                          const int64_t *src, const int64_t *dst, const uint8_t *src_valid,
                          const uint8_t *dst_valid, int64_t *out_len, uint8_t *out_bool, uint8_t *out_valid,
                          orc_stats_t *st, int64_t lower, int64_t upper) {
	size_t vv = (size_t)(V > 0 ? V : 1);
	bs512 *seen = (bs512 *)malloc(vv * sizeof(bs512));
	bs512 *visit1 = (bs512 *)malloc(vv * sizeof(bs512));
	bs512 *visit2 = (bs512 *)malloc(vv * sizeof(bs512));
	int64_t lane_to_num[ORC_LANE_LIMIT];
	int64_t started = 0;

	while (started < n) {
		memset(seen, 0, vv * sizeof(bs512)); /* R4: seen does NOT get the source bit */
		memset(visit1, 0, vv * sizeof(bs512));
		bs512 active_mask;
		memset(&active_mask, 0, sizeof(active_mask));
		int64_t active = 0;
		/* fill lanes 0..511 in row order */
		for (int lane = 0; lane < ORC_LANE_LIMIT; lane++) {
			lane_to_num[lane] = -1;
			while (started < n) {
				int64_t row = started++;
				/* NULL src -> NULL (reference).  NULL dst: the reference reads the slot anyway;
				 * frozen here as NULL result (documented in DESIGN.md). */
				if (!row_valid(src_valid, row) || !row_valid(dst_valid, row)) {
					out_valid[row] = 0;
					if (out_len) out_len[row] = -1;
					if (out_bool) out_bool[row] = 0;
				} else if (src[row] == dst[row]) {
					out_valid[row] = lower <= 0; /* R11: a length-0 path is in bounds only when lower <= 0 */
					if (out_len) out_len[row] = lower <= 0 ? 0 : -1; /* length-0 path needs no search */
					if (out_bool) out_bool[row] = 1;
				} else {
					bs_set(&visit1[src[row]], lane);
					bs_set(&active_mask, lane);
					lane_to_num[lane] = row;
					active++;
					break;
				}
			}
		}
		if (st) st->batches++;
		for (int64_t iter = 1; active && iter <= upper; iter++) { /* R11: at most `upper` levels */
			bs512 *vis = (iter & 1) ? visit1 : visit2;
			bs512 *nxt = (iter & 1) ? visit2 : visit1;
			if (!level_step(V, v, e, seen, vis, nxt, &active_mask, st)) break;
			for (int lane = 0; lane < ORC_LANE_LIMIT; lane++) {
				int64_t row = lane_to_num[lane];
				if (row >= 0 && bs_test(&seen[dst[row]], lane)) {
					out_valid[row] = iter >= lower; /* R11: shortest length below the lower bound -> NULL */
					if (out_len) out_len[row] = iter >= lower ? iter : -1; /* found at iter => path length */
					if (out_bool) out_bool[row] = 1;
					lane_to_num[lane] = -1; /* retired, NOT refilled until the batch ends (R5) */
					active_mask.w[lane >> 6] &= ~(1ull << (lane & 63));
					active--;
				}
			}
		}
		/* no change any more: still-active lanes have no path */
		for (int lane = 0; lane < ORC_LANE_LIMIT; lane++) {
			int64_t row = lane_to_num[lane];
			if (row >= 0) {
				if (kind == 0) {
					out_valid[row] = 0;
					out_len[row] = -1;
				} else {
					out_valid[row] = 1; /* R6: reachability returns false, not NULL */
					out_bool[row] = 0;
				}
				lane_to_num[lane] = -1;
			}
		}
	}
	free(seen);
	free(visit1);
	free(visit2);
}

void orc_iterativelength(int64_t V, const int64_t *v, const int64_t *e, int64_t n, const int64_t *src,
                         const int64_t *dst, const uint8_t *src_valid, const uint8_t *dst_valid,
                         int64_t *out, uint8_t *out_valid, orc_stats_t *stats) {
	msbfs_lengths(0, V, v, e, n, src, dst, src_valid, dst_valid, out, NULL, out_valid, stats, 0, INT64_MAX);
}

/* R11 [RECALL, low confidence; frozen here]: iterativelength(csr_id, v_size, src, dst, lower, upper) of the
 * later DuckPGQ code.  The level loop stops after `upper` levels; a destination first seen at level L
 * gives L when lower <= L (<= upper by construction), otherwise NULL -- the search is a BFS with a `seen`
 * set, so L is the only length it can report for the pair.  src == dst gives 0 when lower <= 0, else NULL.
 * A negative upper bound means "unbounded" (what the 4-argument form does). */
void orc_iterativelength_bounded(int64_t V, const int64_t *v, const int64_t *e, int64_t n, const int64_t *src,
                                 const int64_t *dst, const uint8_t *src_valid, const uint8_t *dst_valid,
                                 int64_t lower, int64_t upper, int64_t *out, uint8_t *out_valid,
                                 orc_stats_t *stats) {
	msbfs_lengths(0, V, v, e, n, src, dst, src_valid, dst_valid, out, NULL, out_valid, stats, lower,
	              upper < 0 ? INT64_MAX : upper);
}

void orc_reachability(int64_t V, const int64_t *v, const int64_t *e, int64_t n, const int64_t *src,
                      const int64_t *dst, const uint8_t *src_valid, const uint8_t *dst_valid, uint8_t *out,
                      uint8_t *out_valid, orc_stats_t *stats) {
	msbfs_lengths(1, V, v, e, n, src, dst, src_valid, dst_valid, NULL, out, out_valid, stats, 0, INT64_MAX);
}
