/*
 * TEST INFRASTRUCTURE ONLY — see pgq_oracle.h.  Textbook BFS cross-check, the multi-threaded
 * chunk driver (the reference's execution shape) and the synthetic input generators.
 */
#define _GNU_SOURCE
#include "pgq_oracle.h"
#include "../include/pgq_gen.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

void orc_bfs_simple(int64_t V, const int64_t *v, const int64_t *e, int64_t source, int32_t *dist) {
	for (int64_t i = 0; i < V; i++) dist[i] = -1;
	if (source < 0 || source >= V) return;
	int64_t *q = (int64_t *)malloc((size_t)(V > 0 ? V : 1) * sizeof(int64_t));
	int64_t head = 0, tail = 0;
	dist[source] = 0;
	q[tail++] = source;
	while (head < tail) {
		int64_t u = q[head++];
		for (int64_t off = v[u]; off < v[u + 1]; off++) {
			int64_t n = e[off];
			if (dist[n] < 0) {
				dist[n] = dist[u] + 1;
				q[tail++] = n;
			}
		}
	}
	free(q);
}

/* ---- multi-threaded chunk driver ------------------------------------------------------ */

typedef struct {
	int kind;
	int64_t V;
	const int64_t *v, *e;
	int64_t n;
	const int64_t *src, *dst;
	void *out;
	uint8_t *out_valid;
	int64_t next_chunk; /* guarded by mu */
	pthread_mutex_t mu;
	orc_stats_t stats; /* guarded by mu */
	int want_stats;
} mt_job_t;

static void *mt_worker(void *arg) {
	mt_job_t *j = (mt_job_t *)arg;
	for (;;) {
		pthread_mutex_lock(&j->mu);
		int64_t c = j->next_chunk++;
		pthread_mutex_unlock(&j->mu);
		int64_t lo = c * ORC_CHUNK;
		if (lo >= j->n) break;
		int64_t cnt = j->n - lo < ORC_CHUNK ? j->n - lo : ORC_CHUNK;
		orc_stats_t st;
		memset(&st, 0, sizeof(st));
		if (j->kind == 0) {
			orc_iterativelength(j->V, j->v, j->e, cnt, j->src + lo, j->dst + lo, NULL, NULL,
			                    (int64_t *)j->out + lo, j->out_valid + lo, j->want_stats ? &st : NULL);
		} else {
			orc_reachability(j->V, j->v, j->e, cnt, j->src + lo, j->dst + lo, NULL, NULL,
			                 (uint8_t *)j->out + lo, j->out_valid + lo, j->want_stats ? &st : NULL);
		}
		if (j->want_stats) {
			pthread_mutex_lock(&j->mu);
			j->stats.batches += st.batches;
			j->stats.levels += st.levels;
			j->stats.edges_lane += st.edges_lane;
			j->stats.edges_unique += st.edges_unique;
			pthread_mutex_unlock(&j->mu);
		}
	}
	return NULL;
}

double orc_run_chunks_mt(int kind, int nthreads, int64_t V, const int64_t *v, const int64_t *e, int64_t n,
                         const int64_t *src, const int64_t *dst, void *out, uint8_t *out_valid,
                         orc_stats_t *stats) {
	mt_job_t j;
	memset(&j, 0, sizeof(j));
	j.kind = kind; j.V = V; j.v = v; j.e = e; j.n = n; j.src = src; j.dst = dst;
	j.out = out; j.out_valid = out_valid; j.want_stats = stats != NULL;
	pthread_mutex_init(&j.mu, NULL);
	if (nthreads < 1) nthreads = 1;
	pthread_t *th = (pthread_t *)malloc((size_t)nthreads * sizeof(pthread_t));
	struct timespec t0, t1;
	clock_gettime(CLOCK_MONOTONIC, &t0);
	for (int t = 0; t < nthreads; t++) pthread_create(&th[t], NULL, mt_worker, &j);
	for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
	clock_gettime(CLOCK_MONOTONIC, &t1);
	free(th);
	pthread_mutex_destroy(&j.mu);
	if (stats) *stats = j.stats;
	return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}

/* ---- generators ------------------------------------------------------------------------- */

typedef struct {
	int which; /* 0 rmat, 1 knows */
	uint64_t seed;
	int scale;
	uint64_t V, E, mult, add;
	int64_t row0, lo, hi;
	int64_t *src, *dst;
} gen_job_t;

static void *gen_worker(void *arg) {
	gen_job_t *g = (gen_job_t *)arg;
	for (int64_t i = g->lo; i < g->hi; i++) {
		uint64_t row = (uint64_t)(g->row0 + i);
		if (g->which == 0)
			pgq_rmat_edge(g->seed, g->scale, row, &g->src[i], &g->dst[i]);
		else
			pgq_knows_edge(g->seed, g->V, g->E, g->mult, g->add, row, &g->src[i], &g->dst[i]);
	}
	return NULL;
}

static void gen_run(gen_job_t proto, int64_t nrows, int nthreads) {
	if (nthreads < 1) nthreads = 1;
	if (nthreads > 256) nthreads = 256;
	pthread_t th[256];
	gen_job_t jobs[256];
	int64_t per = (nrows + nthreads - 1) / nthreads;
	int started = 0;
	for (int t = 0; t < nthreads; t++) {
		jobs[t] = proto;
		jobs[t].lo = t * per;
		jobs[t].hi = (t + 1) * per < nrows ? (t + 1) * per : nrows;
		if (jobs[t].lo >= jobs[t].hi) break;
		pthread_create(&th[t], NULL, gen_worker, &jobs[t]);
		started++;
	}
	for (int t = 0; t < started; t++) pthread_join(th[t], NULL);
}

void orc_gen_rmat(uint64_t seed, int scale, int64_t E, int64_t row0, int64_t nrows, int64_t *src,
                  int64_t *dst, int nthreads) {
	gen_job_t p;
	memset(&p, 0, sizeof(p));
	p.which = 0; p.seed = seed; p.scale = scale; p.V = 1ull << scale; p.E = (uint64_t)E;
	p.row0 = row0; p.src = src; p.dst = dst;
	gen_run(p, nrows, nthreads);
}

static uint64_t gcd_u64(uint64_t a, uint64_t b) {
	while (b) {
		uint64_t t = a % b;
		a = b;
		b = t;
	}
	return a;
}

void orc_knows_params(int64_t V, uint64_t *mult, uint64_t *add) {
	uint64_t v = (uint64_t)(V > 0 ? V : 1);
	uint64_t m = 2654435761ull % v;
	if (m < 2) m = 1;
	while (gcd_u64(m, v) != 1) m++;
	*mult = m;
	*add = 0x9E3779B9ull % v;
}

void orc_gen_knows(uint64_t seed, int64_t V, int64_t E, int64_t row0, int64_t nrows, int64_t *src,
                   int64_t *dst, int nthreads) {
	gen_job_t p;
	memset(&p, 0, sizeof(p));
	p.which = 1; p.seed = seed; p.V = (uint64_t)V; p.E = (uint64_t)E;
	orc_knows_params(V, &p.mult, &p.add);
	p.row0 = row0; p.src = src; p.dst = dst;
	gen_run(p, nrows, nthreads);
}

void orc_gen_pairs(uint64_t seed, int64_t V, const int64_t *v, int64_t n, int64_t *src, int64_t *dst) {
	for (int64_t i = 0; i < n; i++) {
		for (int which = 0; which < 2; which++) {
			uint64_t c = 0;
			for (uint64_t t = 0; t < 4096; t++) {
				c = pgq_pair_candidate(seed, (uint64_t)V, (uint64_t)i, (uint64_t)which, t);
				if (v[c + 1] > v[c]) break; /* out-degree > 0 */
			}
			if (which == 0) src[i] = (int64_t)c; else dst[i] = (int64_t)c;
		}
	}
}
