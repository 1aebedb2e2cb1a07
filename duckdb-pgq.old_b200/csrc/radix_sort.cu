// K3 core: stable LSD radix sort (8-bit digits) of (key, payload0[, payload1]) triples.
//
// Why a sort: create_csr_edge's neighbour order inside a vertex is the ARRIVAL order of the
// rows.  A stable sort of the edge table by `src` puts every edge at
// e[v[src] + (number of earlier rows with the same src)] — exactly the reference's slot —
// without atomics, so the CSR is bit-identical and deterministic.
//
// Per pass: tile digit histogram -> device-wide exclusive scan (bin-major) -> stable scatter.
// Stability inside a tile: each warp owns a contiguous run of the tile and walks it 32 keys at
// a time; __match_any_sync ranks equal digits by lane, per-warp counters carry across rounds,
// a block-wide scan over warps orders the warps.  No shared/global atomics anywhere.
// HBM-bound: per pass 2 key reads + 1 payload read + writes.
#include "common.cuh"

namespace pgq {

static constexpr int RS_THREADS = 256;
static constexpr int RS_WARPS = RS_THREADS / 32;
static constexpr int RS_ITEMS = 16;
static constexpr int RS_TILE = RS_THREADS * RS_ITEMS;

struct RsArgs {
	const int64_t *k64;   // first pass, 64-bit inputs: key = (u32)k64[i], p0 = (u32)p64[i]
	const int64_t *p64;
	const uint32_t *kin;  // otherwise
	const uint32_t *p0in;
	const uint32_t *p1in; // nullptr -> p1 = i (iota)
	uint32_t *kout;       // nullptr -> keys not written (last pass)
	uint32_t *p0out;
	uint32_t *p1out;      // nullptr -> payload1 not carried
	uint32_t *hist;       // [256][ntiles] bin-major (hist kernel writes counts, scatter reads bases)
	uint32_t n;
	uint32_t ntiles;
	int shift;
	uint64_t vmax;        // FIRST64 only: ids must be < vmax
	uint32_t *err;
};

template <bool FIRST64>
__device__ __forceinline__ uint32_t rs_load_key(const RsArgs &a, uint32_t i) {
	if (FIRST64) {
		const uint64_t k = (uint64_t)a.k64[i];
		if (a.err && k >= a.vmax) atomicOr(a.err, 1u); // the reference would write out of bounds here
		return (uint32_t)k;
	}
	return a.kin[i];
}

// Lanes of the warp holding the same 8-bit digit, from 8 ballots (one per digit bit).  The
// hardware MATCH.ANY costs one iteration per distinct value in the warp (measured: the digit pass
// with 32 distinct digits per warp ran 1.3x slower than the skewed top-digit pass); the ballots
// cost the same whatever the distribution.
__device__ __forceinline__ unsigned rs_match8(uint32_t d, bool valid) {
	unsigned peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
	for (int b = 0; b < 8; b++) {
		const bool bit = (d >> b) & 1u;
		const unsigned m = __ballot_sync(0xffffffffu, bit);
		peers &= bit ? m : ~m;
	}
	return peers;
}

// Ranks the tile's keys. keys[r], rank[r] per round; leaves per-warp exclusive bases in cnt and
// tile totals in tot.
template <bool FIRST64, bool WANT_RANK>
__device__ __forceinline__ void rs_rank_tile(const RsArgs &a, uint32_t (*cnt)[256], uint32_t *tot,
                                             uint32_t (&keys)[RS_ITEMS], uint32_t (&rank)[RS_ITEMS]) {
	const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const unsigned lt_mask = (1u << lane) - 1u;
	for (int i = threadIdx.x; i < RS_WARPS * 256; i += RS_THREADS) (&cnt[0][0])[i] = 0;
	__syncthreads();
	const uint64_t wbase = (uint64_t)blockIdx.x * RS_TILE + (uint64_t)warp * (RS_ITEMS * 32);
	// all loads first: 16 independent coalesced requests in flight per thread before any ranking
#pragma unroll
	for (int r = 0; r < RS_ITEMS; r++) {
		const uint64_t idx = wbase + (uint64_t)r * 32 + lane;
		keys[r] = idx < a.n ? rs_load_key<FIRST64>(a, (uint32_t)idx) : 0u;
	}
#pragma unroll
	for (int r = 0; r < RS_ITEMS; r++) {
		const bool valid = wbase + (uint64_t)r * 32 + lane < a.n;
		const uint32_t d = (keys[r] >> a.shift) & 0xFFu;
		if (WANT_RANK) {
			const unsigned peers = rs_match8(d, valid);
			const unsigned lower = peers & lt_mask;
			uint32_t base = 0;
			if (valid) base = cnt[warp][d];
			__syncwarp();
			if (valid && lower == 0) cnt[warp][d] = base + __popc(peers);
			__syncwarp();
			rank[r] = base + __popc(lower);
		} else if (valid) {
			atomicAdd(&cnt[warp][d], 1u); // counting only: order-free shared-memory reduction
		}
	}
	__syncthreads();
	// exclusive scan over warps, one digit per thread (RS_THREADS == 256)
	{
		const int d = threadIdx.x;
		uint32_t run = 0;
#pragma unroll
		for (int w = 0; w < RS_WARPS; w++) {
			uint32_t x = cnt[w][d];
			cnt[w][d] = run;
			run += x;
		}
		tot[d] = run;
	}
	__syncthreads();
}

template <bool FIRST64>
__global__ void __launch_bounds__(RS_THREADS) rs_hist_kernel(RsArgs a) {
	__shared__ uint32_t cnt[RS_WARPS][256];
	__shared__ uint32_t tot[256];
	uint32_t keys[RS_ITEMS], rank[RS_ITEMS];
	rs_rank_tile<FIRST64, false>(a, cnt, tot, keys, rank);
	a.hist[(uint64_t)threadIdx.x * a.ntiles + blockIdx.x] = tot[threadIdx.x];
}

// Scatter with a tile-local reorder: the ranked items are first placed at their position inside
// the tile's own sorted order in shared memory, then written out by consecutive threads, so each
// digit bucket receives contiguous runs (avg RS_TILE/256 items) instead of 32 unrelated 4-byte
// stores per warp instruction.  Keys stay in shared memory (they give the digit, hence the global
// position, of every sorted slot); payloads go through one 16 KB staging array, one at a time.
template <bool FIRST64>
__global__ void __launch_bounds__(RS_THREADS, 4) rs_scatter_kernel(RsArgs a) {
	__shared__ uint32_t cnt[RS_WARPS][256];
	__shared__ uint32_t tot[256];    // tile totals, then (global bucket base - local bucket start)
	__shared__ uint32_t lstart[256]; // local bucket start inside the tile's sorted order
	__shared__ uint32_t wsum[RS_WARPS];
	__shared__ uint32_t skey[RS_TILE];
	__shared__ uint32_t sval[RS_TILE];
	uint32_t keys[RS_ITEMS], lp[RS_ITEMS];
	rs_rank_tile<FIRST64, true>(a, cnt, tot, keys, lp);
	const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	{ // exclusive scan of the tile totals over the 256 digits, one digit per thread
		const uint32_t t = tot[threadIdx.x];
		uint32_t inc = t;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
			if ((int)lane >= o) inc += y;
		}
		if (lane == 31) wsum[warp] = inc;
		__syncthreads();
		uint32_t wb = 0;
#pragma unroll
		for (int w = 0; w < RS_WARPS; w++) wb += (w < (int)warp) ? wsum[w] : 0u;
		const uint32_t ls = wb + inc - t;
		lstart[threadIdx.x] = ls;
		tot[threadIdx.x] = a.hist[(uint64_t)threadIdx.x * a.ntiles + blockIdx.x] - ls;
	}
	__syncthreads();
	const uint64_t tbase = (uint64_t)blockIdx.x * RS_TILE;
	const uint64_t wbase = tbase + (uint64_t)warp * (RS_ITEMS * 32);
	const uint32_t ntile = (uint32_t)((a.n - tbase) < (uint64_t)RS_TILE ? (a.n - tbase) : (uint64_t)RS_TILE);
#pragma unroll
	for (int r = 0; r < RS_ITEMS; r++) {
		const uint64_t idx = wbase + (uint64_t)r * 32 + lane;
		if (idx < a.n) {
			const uint32_t d = (keys[r] >> a.shift) & 0xFFu;
			lp[r] = lstart[d] + cnt[warp][d] + lp[r];
			skey[lp[r]] = keys[r];
		}
	}
	__syncthreads();
#define RS_GPOS(j) (tot[(skey[j] >> a.shift) & 0xFFu] + (j))
	if (a.kout) {
#pragma unroll
		for (int k = 0; k < RS_ITEMS; k++) {
			const uint32_t j = threadIdx.x + k * RS_THREADS;
			if (j < ntile) a.kout[RS_GPOS(j)] = skey[j];
		}
	}
	// payload 0
#pragma unroll
	for (int r = 0; r < RS_ITEMS; r++) {
		const uint64_t idx = wbase + (uint64_t)r * 32 + lane;
		if (idx < a.n) {
			if (FIRST64) {
				const uint64_t pv = (uint64_t)a.p64[idx];
				if (a.err && pv >= a.vmax) atomicOr(a.err, 1u);
				sval[lp[r]] = (uint32_t)pv;
			} else {
				sval[lp[r]] = a.p0in[idx];
			}
		}
	}
	__syncthreads();
#pragma unroll
	for (int k = 0; k < RS_ITEMS; k++) {
		const uint32_t j = threadIdx.x + k * RS_THREADS;
		if (j < ntile) a.p0out[RS_GPOS(j)] = sval[j];
	}
	if (a.p1out) {
		__syncthreads();
#pragma unroll
		for (int r = 0; r < RS_ITEMS; r++) {
			const uint64_t idx = wbase + (uint64_t)r * 32 + lane;
			if (idx < a.n) sval[lp[r]] = a.p1in ? a.p1in[idx] : (uint32_t)idx;
		}
		__syncthreads();
#pragma unroll
		for (int k = 0; k < RS_ITEMS; k++) {
			const uint32_t j = threadIdx.x + k * RS_THREADS;
			if (j < ntile) a.p1out[RS_GPOS(j)] = sval[j];
		}
	}
}

int radix_sort_stable(Ctx *c, const SortIO &io) {
	if (io.n == 0) return PGQ_OK;
	if (radix_sort_wide_eligible(c, io)) return radix_sort_wide(c, io);
	int passes = (io.key_bits + 7) / 8;
	if (passes < 1) passes = 1;
	const uint32_t ntiles = div_up(io.n, RS_TILE);
	const bool want_p1 = io.out_p1 != nullptr;
	const size_t bytes = (size_t)io.n * sizeof(uint32_t);
	if (io.keys_scratch) PGQ_TRY(c->sort_keys[(passes - 1) & 1].reserve(bytes));
	if (passes > 1) {
		for (int b = 0; b < (passes > 2 ? 2 : 1); b++) {
			PGQ_TRY(c->sort_keys[b].reserve(bytes));
			PGQ_TRY(c->sort_p0[b].reserve(bytes));
			if (want_p1) PGQ_TRY(c->sort_p1[b].reserve(bytes));
		}
	}
	PGQ_TRY(c->sort_hist.reserve((size_t)256 * ntiles * sizeof(uint32_t)));
	for (int p = 0; p < passes; p++) {
		RsArgs a{};
		const bool first = p == 0, last = p == passes - 1;
		const bool first64 = first && io.k64 != nullptr;
		if (first) {
			a.k64 = io.k64; a.p64 = io.p64; a.kin = io.k32; a.p0in = io.p32; a.p1in = nullptr;
		} else {
			int in = (p - 1) & 1;
			a.kin = c->sort_keys[in].as<uint32_t>();
			a.p0in = c->sort_p0[in].as<uint32_t>();
			a.p1in = want_p1 ? c->sort_p1[in].as<uint32_t>() : nullptr;
		}
		if (last) {
			a.kout = io.out_keys; a.p0out = io.out_p0; a.p1out = io.out_p1;
			if (io.keys_scratch) { // the ping-pong buffer the last pass does not read is free
				a.kout = c->sort_keys[p & 1].as<uint32_t>();
				*io.keys_scratch = a.kout;
			}
		} else {
			int out = p & 1;
			a.kout = c->sort_keys[out].as<uint32_t>();
			a.p0out = c->sort_p0[out].as<uint32_t>();
			a.p1out = want_p1 ? c->sort_p1[out].as<uint32_t>() : nullptr;
		}
		a.hist = c->sort_hist.as<uint32_t>();
		a.n = io.n; a.ntiles = ntiles; a.shift = 8 * p;
		a.vmax = io.vmax; a.err = first64 ? io.err : nullptr;
		if (first64) rs_hist_kernel<true><<<ntiles, RS_THREADS, 0, c->stream>>>(a);
		else rs_hist_kernel<false><<<ntiles, RS_THREADS, 0, c->stream>>>(a);
		count_launch(c);
		PGQ_TRY(exclusive_scan_u32(c, a.hist, a.hist, (uint64_t)256 * ntiles, nullptr));
		if (first64) rs_scatter_kernel<true><<<ntiles, RS_THREADS, 0, c->stream>>>(a);
		else rs_scatter_kernel<false><<<ntiles, RS_THREADS, 0, c->stream>>>(a);
		count_launch(c);
		PGQ_CUDA(cudaGetLastError());
	}
	return PGQ_OK;
}

} // namespace pgq
