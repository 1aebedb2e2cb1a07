/* TEST INFRASTRUCTURE ONLY — shared 512-lane bitset helpers of the oracle (see pgq_oracle.h). */
#ifndef PGQ_ORACLE_INTERNAL_H
#define PGQ_ORACLE_INTERNAL_H
#include "pgq_oracle.h"
#include <stddef.h>

#define WORDS (ORC_LANE_LIMIT / 64)

typedef struct {
	uint64_t w[WORDS];
} bs512;

static inline int bs_any(const bs512 *b) {
	uint64_t o = 0;
	for (int i = 0; i < WORDS; i++) o |= b->w[i];
	return o != 0;
}
static inline int bs_test(const bs512 *b, int lane) {
	return (int)((b->w[lane >> 6] >> (lane & 63)) & 1);
}
static inline void bs_set(bs512 *b, int lane) {
	b->w[lane >> 6] |= (1ull << (lane & 63));
}

static inline int row_valid(const uint8_t *valid, int64_t i) {
	return valid == NULL || valid[i] != 0;
}

#endif
