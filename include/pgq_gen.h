/*
 * Synthetic input generators (bench / test inputs only; no query algorithm lives here).
 *
 * Counter-based (stateless) generators so that the CPU side here and the CUDA
 * side (duckdb-pgq.old_b200/csrc/gen.cu) and the CPU oracle (oracle/) produce bit-identical edge tables from
 * (seed, row index) alone.  All arithmetic is integer: no float thresholds, so
 * host and device cannot disagree on rounding.
 *
 * Shapes follow SURVEY.md §8(d):
 *   - R-MAT (Graph500 Kronecker) a=.57 b=.19 c=.19 d=.05, V=2^scale, E=ef*V,
 *     no dedup, no self-loop removal, ids not permuted; edge table order =
 *     generation order (that order *is* the stable key of the CSR build).
 *   - LDBC-SNB "knows"-shaped: V persons, E/2 undirected friendships with a
 *     power-law-ish degree, materialised in both directions exactly like the
 *     reference's loader does (first all (p1,p2) rows, then all (p2,p1) rows:
 *     /root/reference/benchmark/ldbc/snb-load.sql:23-24).
 *
 * This header is shared verbatim (by #include) with the CUDA generator so the
 * two cannot drift; PGQ_HD expands to __host__ __device__ under nvcc.
 */
#ifndef PGQ_GEN_H
#define PGQ_GEN_H

#include <stdint.h>

#ifdef __CUDACC__
#define PGQ_HD __host__ __device__ __forceinline__
#else
#define PGQ_HD static inline
#endif

PGQ_HD uint64_t pgq_mix64(uint64_t x) {
	/* splitmix64 finaliser */
	x += 0x9E3779B97F4A7C15ull;
	x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
	x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
	return x ^ (x >> 31);
}

PGQ_HD uint64_t pgq_hash3(uint64_t seed, uint64_t a, uint64_t b) {
	return pgq_mix64(pgq_mix64(seed ^ (a * 0xD1B54A32D192ED03ull)) + b * 0x8CB92BA72F3D8DD7ull);
}

/* R-MAT thresholds on a uniform u32: a, a+b, a+b+c with a=.57 b=.19 c=.19 */
#define PGQ_RMAT_A 2448131359u   /* floor(0.57 * 2^32) */
#define PGQ_RMAT_AB 3264175145u  /* floor(0.76 * 2^32) */
#define PGQ_RMAT_ABC 4080218931u /* floor(0.95 * 2^32) */

/* Edge `i` of the R-MAT(scale) stream with the given seed. */
PGQ_HD void pgq_rmat_edge(uint64_t seed, int scale, uint64_t i, int64_t *src, int64_t *dst) {
	uint64_t s = 0, d = 0;
	for (int lvl = 0; lvl < scale; lvl += 2) {
		uint64_t r = pgq_hash3(seed, i, (uint64_t)(lvl >> 1));
		for (int k = 0; k < 2 && lvl + k < scale; k++) {
			uint32_t u = (uint32_t)(r >> (32 * k));
			uint64_t sb, db;
			if (u < PGQ_RMAT_A) {
				sb = 0; db = 0;
			} else if (u < PGQ_RMAT_AB) {
				sb = 0; db = 1;
			} else if (u < PGQ_RMAT_ABC) {
				sb = 1; db = 0;
			} else {
				sb = 1; db = 1;
			}
			s = (s << 1) | sb;
			d = (d << 1) | db;
		}
	}
	*src = (int64_t)s;
	*dst = (int64_t)d;
}

/* Skewed person pick for the knows-shaped generator: density ~ x^(-1/2) over
 * [0,V), then an affine bijection (mult coprime to V) to spread hubs over ids. */
PGQ_HD uint64_t pgq_knows_pick(uint64_t r, uint64_t V, uint64_t mult, uint64_t add) {
	uint64_t u = r >> 32;               /* u32 uniform */
	uint64_t t = (u * u) >> 32;         /* square: skew towards 0 */
	uint64_t idx = (t * V) >> 32;       /* [0,V) */
	return (idx * mult + add) % V;
}

/* Undirected friendship k (0 <= k < E/2) -> endpoints (a,b), a != b when V>1. */
PGQ_HD void pgq_knows_pair(uint64_t seed, uint64_t V, uint64_t mult, uint64_t add, uint64_t k, int64_t *a,
                           int64_t *b) {
	uint64_t pa = pgq_knows_pick(pgq_hash3(seed, k, 0), V, mult, add);
	uint64_t pb = pgq_knows_pick(pgq_hash3(seed, k, 1), V, mult, add);
	if (pa == pb) {
		pb = (pb + 1) % V;
	}
	*a = (int64_t)pa;
	*b = (int64_t)pb;
}

/* Row `row` of the directed knows table with E rows (E even): rows [0,E/2) are
 * (a,b), rows [E/2,E) are the reversed (b,a). */
PGQ_HD void pgq_knows_edge(uint64_t seed, uint64_t V, uint64_t E, uint64_t mult, uint64_t add, uint64_t row,
                           int64_t *src, int64_t *dst) {
	uint64_t half = E / 2;
	int64_t a, b;
	if (row < half) {
		pgq_knows_pair(seed, V, mult, add, row, &a, &b);
		*src = a; *dst = b;
	} else {
		pgq_knows_pair(seed, V, mult, add, row - half, &a, &b);
		*src = b; *dst = a;
	}
}

/* Candidate vertex for pair `i`, attempt `t` (rejection sampling on out-degree
 * > 0 is done by the caller, which owns the offsets array). */
PGQ_HD uint64_t pgq_pair_candidate(uint64_t seed, uint64_t V, uint64_t i, uint64_t which, uint64_t t) {
	return pgq_hash3(seed ^ 0xA5A5A5A55A5A5A5Aull, i * 2 + which, t) % V;
}

#endif
