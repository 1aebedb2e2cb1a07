/*
 * TEST INFRASTRUCTURE ONLY — see pgq_oracle.h.  shortestpath restatement, rules R7-R9 of
 * SURVEY.md §8(a).  PARITY UNPINNED BY THE REFERENCE.
 *
 * R7 tie-break: in the expand loop, for each frontier vertex v ASCENDING, each CSR offset
 * ASCENDING, target n, every lane l:  if parents_v[n][l]==-1 && visit[v][l] then
 * parents_v[n][l]=v, parents_e[n][l]=edge_ids[off].  parents_v[src][lane]=src up front.
 * R8 output: vertex-only [src,..,dst] (mode 0) or interleaved [src,e0,v1,..,dst] (mode 1).
 * R9 list layout: result_data[row] = {offset=total_len, length}, values appended in row
 * order of reconstruction (list_entry_t: src/include/duckdb/common/types.hpp:188-195).
 */
#define _GNU_SOURCE
#include "pgq_oracle_internal.h"

#include <stdlib.h>
#include <string.h>

typedef struct {
	int64_t *vals;
	uint64_t n, cap;
} vec64;

static void vec_push(vec64 *a, int64_t x) {
	if (a->n == a->cap) {
		a->cap = a->cap ? a->cap * 2 : 1024;
		a->vals = (int64_t *)realloc(a->vals, a->cap * sizeof(int64_t));
	}
	a->vals[a->n++] = x;
}

/* level step with parent recording.  has_parent mirrors (parents_v != -1) per lane and is
 * only consulted when faithful==0. */
static int level_step_parents(int64_t V, const int64_t *v, const int64_t *e, const int64_t *eids,
                              int64_t *parents_v, int64_t *parents_e, bs512 *has_parent, bs512 *seen,
                              const bs512 *visit, bs512 *next, const bs512 *active, int faithful,
                              orc_stats_t *st) {
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
			int64_t n = e[off];
			int64_t edge_id = eids ? eids[off] : -1;
			bs512 *nx = &next[n];
			for (int k = 0; k < WORDS; k++) nx->w[k] |= vi.w[k];
			int64_t *pv = parents_v + (size_t)n * ORC_LANE_LIMIT;
			int64_t *pe = parents_e ? parents_e + (size_t)n * ORC_LANE_LIMIT : NULL;
			if (faithful) {
				/* the reference's literal 512-iteration inner loop */
				for (int l = 0; l < ORC_LANE_LIMIT; l++) {
					int take = (pv[l] == -1) && bs_test(&vi, l);
					pv[l] = take ? i : pv[l];
					if (pe) pe[l] = take ? edge_id : pe[l];
				}
			} else {
				bs512 *hp = &has_parent[n];
				for (int k = 0; k < WORDS; k++) {
					uint64_t nb = vi.w[k] & ~hp->w[k];
					hp->w[k] |= nb;
					while (nb) {
						int b = __builtin_ctzll(nb);
						nb &= nb - 1;
						pv[k * 64 + b] = i;
						if (pe) pe[k * 64 + b] = edge_id;
					}
				}
			}
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

int orc_shortestpath(int64_t V, const int64_t *v, const int64_t *e, const int64_t *eids, int64_t n,
                     const int64_t *src, const int64_t *dst, const uint8_t *src_valid,
                     const uint8_t *dst_valid, int mode, int faithful, orc_list_entry_t *out_entries,
                     uint8_t *out_valid, int64_t **out_vals, uint64_t *out_total, orc_stats_t *st) {
	if (mode == 1 && !eids) return -1;
	size_t vv = (size_t)(V > 0 ? V : 1);
	bs512 *seen = (bs512 *)malloc(vv * sizeof(bs512));
	bs512 *visit1 = (bs512 *)malloc(vv * sizeof(bs512));
	bs512 *visit2 = (bs512 *)malloc(vv * sizeof(bs512));
	bs512 *has_parent = (bs512 *)malloc(vv * sizeof(bs512));
	int64_t *parents_v = (int64_t *)malloc(vv * ORC_LANE_LIMIT * sizeof(int64_t));
	int64_t *parents_e = mode == 1 ? (int64_t *)malloc(vv * ORC_LANE_LIMIT * sizeof(int64_t)) : NULL;
	if (!seen || !visit1 || !visit2 || !has_parent || !parents_v || (mode == 1 && !parents_e)) return -2;
	int64_t lane_to_num[ORC_LANE_LIMIT];
	int64_t started = 0;
	vec64 vals = {NULL, 0, 0};
	vec64 tmp = {NULL, 0, 0};
	uint64_t total_len = 0;

	while (started < n) {
		memset(seen, 0, vv * sizeof(bs512));
		memset(visit1, 0, vv * sizeof(bs512));
		memset(has_parent, 0, vv * sizeof(bs512));
		for (size_t i = 0; i < vv * ORC_LANE_LIMIT; i++) parents_v[i] = -1;
		if (parents_e)
			for (size_t i = 0; i < vv * ORC_LANE_LIMIT; i++) parents_e[i] = -1;
		bs512 active_mask;
		memset(&active_mask, 0, sizeof(active_mask));
		int64_t active = 0;
		int64_t batch_first_row = started;
		for (int lane = 0; lane < ORC_LANE_LIMIT; lane++) {
			lane_to_num[lane] = -1;
			while (started < n) {
				int64_t row = started++;
				if (!row_valid(src_valid, row) || !row_valid(dst_valid, row)) {
					out_valid[row] = 0;
				} else if (src[row] == dst[row]) {
					out_valid[row] = 2; /* marker: trivial [src] list, emitted below in row order */
				} else {
					bs_set(&visit1[src[row]], lane);
					bs_set(&active_mask, lane);
					parents_v[(size_t)src[row] * ORC_LANE_LIMIT + lane] = src[row]; /* mark source with itself */
					if (parents_e) parents_e[(size_t)src[row] * ORC_LANE_LIMIT + lane] = -2; /* no incoming edge */
					bs_set(&has_parent[src[row]], lane);
					lane_to_num[lane] = row;
					active++;
					break;
				}
			}
		}
		int64_t batch_end_row = started;
		if (st) st->batches++;
		int64_t remaining = active;
		bs512 found;
		memset(&found, 0, sizeof(found));
		for (int64_t iter = 1; remaining; iter++) {
			bs512 *vis = (iter & 1) ? visit1 : visit2;
			bs512 *nxt = (iter & 1) ? visit2 : visit1;
			if (!level_step_parents(V, v, e, eids, parents_v, parents_e, has_parent, seen, vis, nxt,
			                        &active_mask, faithful, st))
				break;
			for (int lane = 0; lane < ORC_LANE_LIMIT; lane++) {
				int64_t row = lane_to_num[lane];
				if (row >= 0 && !bs_test(&found, lane) && bs_test(&seen[dst[row]], lane)) {
					bs_set(&found, lane);
					active_mask.w[lane >> 6] &= ~(1ull << (lane & 63));
					remaining--;
				}
			}
		}
		/* Reconstruct in ROW order over the rows this batch consumed (trivial rows included),
		 * so offsets are monotone in the row index (R9). */
		int lane_of_row_cursor = 0;
		for (int64_t row = batch_first_row; row < batch_end_row; row++) {
			if (out_valid[row] == 0 && (!row_valid(src_valid, row) || !row_valid(dst_valid, row))) {
				out_entries[row].offset = total_len;
				out_entries[row].length = 0;
				continue;
			}
			if (out_valid[row] == 2) {
				out_valid[row] = 1;
				vec_push(&vals, src[row]);
				out_entries[row].offset = total_len;
				out_entries[row].length = 1;
				total_len += 1;
				continue;
			}
			/* find this row's lane (lanes were filled in row order) */
			while (lane_of_row_cursor < ORC_LANE_LIMIT && lane_to_num[lane_of_row_cursor] != row)
				lane_of_row_cursor++;
			int lane = lane_of_row_cursor;
			if (lane >= ORC_LANE_LIMIT || !bs_test(&found, lane)) {
				out_valid[row] = 0; /* no path */
				out_entries[row].offset = total_len;
				out_entries[row].length = 0;
				continue;
			}
			/* walk dst -> parents -> src, then reverse */
			tmp.n = 0;
			int64_t source_v = src[row];
			int64_t cur = dst[row];
			int ok = 1;
			vec_push(&tmp, cur);
			while (cur != source_v) {
				int64_t pv = parents_v[(size_t)cur * ORC_LANE_LIMIT + lane];
				if (pv == -1 || (pv == cur && cur != source_v)) {
					ok = 0;
					break;
				}
				if (mode == 1) vec_push(&tmp, parents_e[(size_t)cur * ORC_LANE_LIMIT + lane]);
				vec_push(&tmp, pv);
				cur = pv;
			}
			if (!ok) {
				out_valid[row] = 0;
				out_entries[row].offset = total_len;
				out_entries[row].length = 0;
				continue;
			}
			out_valid[row] = 1;
			out_entries[row].offset = total_len;
			out_entries[row].length = tmp.n;
			for (uint64_t k = tmp.n; k > 0; k--) vec_push(&vals, tmp.vals[k - 1]);
			total_len += tmp.n;
		}
	}
	free(seen);
	free(visit1);
	free(visit2);
	free(has_parent);
	free(parents_v);
	free(parents_e);
	free(tmp.vals);
	if (!vals.vals) vals.vals = (int64_t *)malloc(sizeof(int64_t));
	*out_vals = vals.vals;
	*out_total = total_len;
	return 0;
}

void orc_free(void *p) {
	free(p);
}
