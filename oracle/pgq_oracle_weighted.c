/*
 * TEST INFRASTRUCTURE ONLY — oracle for the weighted next-row of SURVEY.md §8(f) f4:
 * cheapest_path_length over a CSR that carries one weight per edge.
 *
 * PARITY UNPINNED: like the rest of the PGQ functions this is absent from /root/reference.
 * Rule R10 (frozen here, [RECALL] of DuckPGQ's cheapest_path_length): the 7-argument
 * create_csr_edge(csr_id, v_size, num_edges, src, dst, edge_rowid, weight) stores `weight` in the
 * edge's CSR slot (same slot rule as R3); cheapest_path_length(csr_id, v_size, src, dst) runs a
 * lane-batched Bellman-Ford: dists[v][lane] = +inf (int64: INT64_MAX/2), dists[src][lane] = 0,
 * "modified" vertices relax their out-edges (new = dists[v][lane] + w; if new < dists[n][lane]
 * store and mark n) until nothing changes; result = dists[dst][lane], NULL when still +inf,
 * NULL src/dst -> NULL, src == dst -> 0.  BIGINT weights give BIGINT, DOUBLE weights DOUBLE.
 * Two independent checks: orc_dijkstra below (binary heap, non-negative weights) and, for
 * doubles, the fact that Bellman-Ford's fixed point min_v fl(d[v] + w) does not depend on the
 * relaxation order.
 */
#include "pgq_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define ORC_WLANES 64
#define ORC_I64_INF (INT64_MAX / 2)

/* weights into CSR slot order, R3's slot rule: slot = v[src] + (earlier rows with the same src) */
void orc_csr_weights(int64_t V, int64_t E, const int64_t *v, const int64_t *src, const void *w_in, void *w_out) {
	int64_t *next = (int64_t *)malloc((size_t)(V > 0 ? V : 1) * sizeof(int64_t));
	const uint64_t *wi = (const uint64_t *)w_in; /* raw 8-byte values: int64 or double */
	uint64_t *wo = (uint64_t *)w_out;
	for (int64_t i = 0; i < V; i++) next[i] = v[i];
	for (int64_t r = 0; r < E; r++) wo[next[src[r]]++] = wi[r];
	free(next);
}

#define DEFINE_BF(NAME, T, INF, IS_INF)                                                                          \
	int NAME(int64_t V, const int64_t *v, const int64_t *e, const T *w, int64_t n, const int64_t *src,           \
	         const int64_t *dst, const uint8_t *src_valid, const uint8_t *dst_valid, T *out, uint8_t *out_valid,  \
	         uint64_t *rounds_out) {                                                                              \
		T *dist = (T *)malloc((size_t)(V > 0 ? V : 1) * ORC_WLANES * sizeof(T));                                  \
		uint8_t *mod = (uint8_t *)calloc((size_t)(V > 0 ? V : 1), 1), *nmod = (uint8_t *)calloc((size_t)(V > 0 ? V : 1), 1); \
		int64_t lane_row[ORC_WLANES];                                                                             \
		uint64_t rounds = 0;                                                                                      \
		int rc = 0;                                                                                               \
		int64_t row = 0;                                                                                          \
		while (row < n) {                                                                                         \
			int lanes = 0;                                                                                        \
			for (int64_t i = 0; i < V * ORC_WLANES; i++) dist[i] = INF;                                           \
			memset(mod, 0, (size_t)V);                                                                            \
			while (row < n && lanes < ORC_WLANES) {                                                               \
				const int64_t r = row++;                                                                          \
				if ((src_valid && !src_valid[r]) || (dst_valid && !dst_valid[r])) {                               \
					out[r] = 0;                                                                                   \
					out_valid[r] = 0;                                                                             \
				} else if (src[r] == dst[r]) {                                                                    \
					out[r] = 0;                                                                                   \
					out_valid[r] = 1;                                                                             \
				} else {                                                                                          \
					dist[src[r] * ORC_WLANES + lanes] = 0;                                                        \
					mod[src[r]] = 1;                                                                              \
					lane_row[lanes++] = r;                                                                        \
				}                                                                                                 \
			}                                                                                                     \
			int changed = lanes > 0;                                                                              \
			int64_t it = 0;                                                                                       \
			while (changed) {                                                                                     \
				if (++it > V + 1) {                                                                               \
					rc = 1; /* negative cycle reachable from a source */                                          \
					break;                                                                                        \
				}                                                                                                 \
				changed = 0;                                                                                      \
				rounds++;                                                                                         \
				memset(nmod, 0, (size_t)V);                                                                       \
				for (int64_t a = 0; a < V; a++) {                                                                 \
					if (!mod[a]) continue;                                                                        \
					for (int64_t off = v[a]; off < v[a + 1]; off++) {                                             \
						const int64_t b = e[off];                                                                 \
						const T ww = w[off];                                                                      \
						for (int l = 0; l < lanes; l++) {                                                         \
							const T da = dist[a * ORC_WLANES + l];                                                \
							if (IS_INF(da)) continue;                                                             \
							const T nd = da + ww;                                                                 \
							if (nd < dist[b * ORC_WLANES + l]) {                                                  \
								dist[b * ORC_WLANES + l] = nd;                                                    \
								nmod[b] = 1;                                                                      \
								changed = 1;                                                                      \
							}                                                                                     \
						}                                                                                         \
					}                                                                                             \
				}                                                                                                 \
				uint8_t *t = mod;                                                                                 \
				mod = nmod;                                                                                       \
				nmod = t;                                                                                         \
			}                                                                                                     \
			for (int l = 0; l < lanes; l++) {                                                                     \
				const int64_t r = lane_row[l];                                                                    \
				const T d = dist[dst[r] * ORC_WLANES + l];                                                        \
				out_valid[r] = IS_INF(d) ? 0 : 1;                                                                 \
				out[r] = IS_INF(d) ? 0 : d;                                                                       \
			}                                                                                                     \
			if (rc) break;                                                                                        \
		}                                                                                                         \
		if (rounds_out) *rounds_out = rounds;                                                                     \
		free(dist);                                                                                               \
		free(mod);                                                                                                \
		free(nmod);                                                                                               \
		return rc;                                                                                                \
	}

#define I64_IS_INF(x) ((x) >= ORC_I64_INF)
#define F64_IS_INF(x) (isinf(x))
DEFINE_BF(orc_cheapest_path_length_i64, int64_t, ORC_I64_INF, I64_IS_INF)
DEFINE_BF(orc_cheapest_path_length_f64, double, INFINITY, F64_IS_INF)

/* Independent check: single-source Dijkstra with a binary heap (non-negative int64 weights).
 * dist[V], ORC_I64_INF for unreachable. */
typedef struct {
	int64_t d;
	int64_t v;
} heap_item;

static void heap_push(heap_item *h, int64_t *n, heap_item x) {
	int64_t i = (*n)++;
	while (i > 0) {
		const int64_t p = (i - 1) / 2;
		if (h[p].d <= x.d) break;
		h[i] = h[p];
		i = p;
	}
	h[i] = x;
}

static heap_item heap_pop(heap_item *h, int64_t *n) {
	const heap_item top = h[0];
	const heap_item x = h[--(*n)];
	int64_t i = 0;
	for (;;) {
		int64_t c = 2 * i + 1;
		if (c >= *n) break;
		if (c + 1 < *n && h[c + 1].d < h[c].d) c++;
		if (x.d <= h[c].d) break;
		h[i] = h[c];
		i = c;
	}
	h[i] = x;
	return top;
}

void orc_dijkstra_i64(int64_t V, const int64_t *v, const int64_t *e, const int64_t *w, int64_t source,
                      int64_t *dist) {
	const int64_t E = v[V];
	heap_item *h = (heap_item *)malloc((size_t)(E + V + 1) * sizeof(heap_item));
	int64_t hn = 0;
	for (int64_t i = 0; i < V; i++) dist[i] = ORC_I64_INF;
	dist[source] = 0;
	heap_push(h, &hn, (heap_item) {0, source});
	while (hn > 0) {
		const heap_item it = heap_pop(h, &hn);
		if (it.d != dist[it.v]) continue;
		for (int64_t off = v[it.v]; off < v[it.v + 1]; off++) {
			const int64_t nd = it.d + w[off];
			if (nd < dist[e[off]]) {
				dist[e[off]] = nd;
				heap_push(h, &hn, (heap_item) {nd, e[off]});
			}
		}
	}
	free(h);
}
