// K3, keys of at most 22 bits: the same stable LSD radix sort as radix_sort.cu in TWO passes of up to
// 11-bit digits instead of three 8-bit ones.
//
// Why: the 8-bit scatter is instruction-bound (profiles/r01_csr_build_v2.txt: 54 % issue-active, ~160 SASS
// instructions per key and pass, 1/3 of them the ballots that find equal digits inside a warp), so
// the cost of a sort is (ballots per key = key bits, the same however they are split) + (per-pass work:
// loads, staging, position arithmetic, stores, the histogram pre-pass) x passes.  Two passes instead of
// three remove a third of the second term and a third of the memory traffic.  What makes 2048 buckets
// per pass affordable on this GPU is shared memory: a 8192-key tile is ranked with per-warp 16-bit
// counters (8 warps x 2048 x 2 B = 32 KB) and reordered in a 64 KB staging area, two blocks per SM.
//
// Between the passes the (key, payload) pairs travel PACKED as one 8-byte item: one 8-byte load and
// one 8-byte store per key, and a bucket's run inside a tile (8192 / 2048 = 4 items on uniform digits)
// is a whole 32-byte sector.
//
// Per pass, as in radix_sort.cu: tile digit histogram -> device-wide exclusive scan (bin-major) ->
// stable scatter (rank inside the tile, place in the tile's sorted order in shared memory, write out
// with consecutive threads).  No global atomics, deterministic, bit-identical to the 8-bit path.
#include "common.cuh"

namespace pgq {

static constexpr int RW_MAXBITS = 11;          // keys of at most 22 bits: two passes of <= 11-bit digits
static constexpr int RW_NB = 1 << RW_MAXBITS; // 2048 buckets
// keys of 23 / 24 bits (R-MAT s24: C3's own graph) CAN take two passes of 12-bit digits: 4096 buckets, 512-thread
// blocks (16 warps x 4096 x 2 B of counters = 128 KB + 64 KB staging + 16 KB bases = 208 KB, one block per SM, the
// same 16 warps per SM as two 256-thread blocks of the 11-bit shape) instead of three 8-bit passes.  Measured slower
// (10.4 vs 9.1 ms at s24), so it is opt-in (option sort_wide=3); the tests run it for bit-identity.
static constexpr int RW_MAXBITS_BIG = 12;
static constexpr int RW_NB_BIG = 1 << RW_MAXBITS_BIG;
static constexpr int RW_THREADS = 256;
static constexpr int RW_WARPS = RW_THREADS / 32;
static constexpr int RW_ITEMS = 32;
static constexpr int RW_TILE = RW_THREADS * RW_ITEMS; // 8192 keys

struct RwArgs {
	const int64_t *k64;   // input form A (first pass): 64-bit columns, validated against vmax
	const int64_t *p64;
	const uint32_t *k32;  // input form B (first pass): 32-bit columns
	const uint32_t *p32;
	const uint2 *pin;     // input form C (second pass): packed (key, payload0)
	const uint32_t *p1in; // payload1 of the previous pass; nullptr on the first pass (payload1 = index)
	uint2 *pout;          // output, not the last pass: packed
	uint32_t *kout;       // output, last pass: sorted keys (may be nullptr)
	uint32_t *p0out;      //                    payload0
	uint32_t *p1out;      // payload1 (nullptr = not carried)
	uint32_t *hist;       // [1 << nbits][ntiles] bin-major: tile counts, then (after the scan) tile bases
	// decoupled look-back instead of the histogram pre-pass (status != nullptr):
	uint32_t *status;      // [ntiles][RW_NB] tile-major, zeroed: (flag << 30) | count, flag 1 = tile total, 2 = inclusive prefix
	uint32_t *ticket;      // zeroed: tiles are numbered in the order their blocks start
	const uint32_t *gdig;  // [RW_NB] global start of every bucket of this pass
	uint32_t n;
	uint32_t ntiles;
	uint32_t nb;           // bucket stride of status / gdig: RW_NB or RW_NB_BIG
	int shift;
	int nbits;
	uint64_t vmax;
	uint32_t *err;
};

// Input forms (template parameter, so a tile's loads are straight-line code the compiler can batch).
enum { RW_IN_64 = 0, RW_IN_32 = 1, RW_IN_PACKED = 2 };

template <int IN>
__device__ __forceinline__ uint32_t rw_load_key(const RwArgs &a, uint32_t i) {
	if (IN == RW_IN_64) return (uint32_t)(uint64_t)a.k64[i];
	if (IN == RW_IN_PACKED) return a.pin[i].x;
	return a.k32[i];
}

// bad |= "id outside [0, vmax)" for the 64-bit form (reported once per warp after the loads)
template <int IN>
__device__ __forceinline__ uint2 rw_load_item(const RwArgs &a, uint32_t i, bool &bad) {
	if (IN == RW_IN_64) {
		const uint64_t k = (uint64_t)a.k64[i], p = (uint64_t)a.p64[i];
		bad |= (k >= a.vmax) | (p >= a.vmax);
		return make_uint2((uint32_t)k, (uint32_t)p);
	}
	if (IN == RW_IN_PACKED) return a.pin[i];
	return make_uint2(a.k32[i], a.p32[i]);
}

// Tile digit counts (order-free: shared-memory atomics).
template <int IN>
__global__ void __launch_bounds__(RW_THREADS) rw_hist_kernel(RwArgs a) {
	__shared__ uint32_t h[RW_NB_BIG];
	for (int i = threadIdx.x; i < RW_NB_BIG; i += RW_THREADS) h[i] = 0;
	__syncthreads();
	const uint32_t tbase = blockIdx.x * (uint32_t)RW_TILE;
	const uint32_t mask = (1u << a.nbits) - 1u;
	uint32_t keys[RW_ITEMS];
	if (a.n - tbase >= (uint32_t)RW_TILE) { // full tile: unguarded loads, all in flight together
#pragma unroll
		for (int r = 0; r < RW_ITEMS; r++) keys[r] = rw_load_key<IN>(a, tbase + r * RW_THREADS + threadIdx.x);
#pragma unroll
		for (int r = 0; r < RW_ITEMS; r++) atomicAdd(&h[(keys[r] >> a.shift) & mask], 1u);
	} else {
		for (int r = 0; r < RW_ITEMS; r++) {
			const uint32_t idx = tbase + r * RW_THREADS + threadIdx.x;
			if (idx < a.n) atomicAdd(&h[(rw_load_key<IN>(a, idx) >> a.shift) & mask], 1u);
		}
	}
	__syncthreads();
	const uint32_t nb = 1u << a.nbits;
	for (uint32_t d = threadIdx.x; d < nb; d += RW_THREADS) a.hist[(uint64_t)d * a.ntiles + blockIdx.x] = h[d];
}

// ---- look-back variant: one read of the first-pass keys gives the GLOBAL bucket sizes of both passes ----
template <int IN>
__global__ void __launch_bounds__(RW_THREADS) rw_global_hist_kernel(RwArgs a, int bits0, int bits1, uint32_t *gh) {
	__shared__ uint32_t h[2 * RW_NB_BIG];
	for (int i = threadIdx.x; i < 2 * RW_NB_BIG; i += RW_THREADS) h[i] = 0;
	const uint32_t NBS = a.nb;
	__syncthreads();
	const uint32_t m0 = (1u << bits0) - 1u, m1 = (1u << bits1) - 1u; // bits1 == 0: m1 = 0, everything in bucket 0
	for (uint32_t tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
		const uint32_t tbase = tile * (uint32_t)RW_TILE;
		if (a.n - tbase >= (uint32_t)RW_TILE) {
			uint32_t keys[RW_ITEMS];
#pragma unroll
			for (int r = 0; r < RW_ITEMS; r++) keys[r] = rw_load_key<IN>(a, tbase + r * RW_THREADS + threadIdx.x);
#pragma unroll
			for (int r = 0; r < RW_ITEMS; r++) {
				atomicAdd(&h[keys[r] & m0], 1u);
				atomicAdd(&h[NBS + ((keys[r] >> bits0) & m1)], 1u);
			}
		} else {
			for (uint32_t idx = tbase + threadIdx.x; idx < a.n; idx += RW_THREADS) {
				const uint32_t k = rw_load_key<IN>(a, idx);
				atomicAdd(&h[k & m0], 1u);
				atomicAdd(&h[NBS + ((k >> bits0) & m1)], 1u);
			}
		}
	}
	__syncthreads();
	for (uint32_t i = threadIdx.x; i < 2 * NBS; i += RW_THREADS)
		if (h[i]) atomicAdd(&gh[i], h[i]);
}

// exclusive scan of each pass's NB bucket sizes, in place (one 1024-thread block per pass)
template <int NB>
__global__ void __launch_bounds__(1024) rw_gdig_scan_kernel(uint32_t *gh) {
	constexpr int PT = NB / 1024; // 2 or 4 consecutive buckets per thread
	__shared__ uint32_t wsum[32];
	uint32_t *g = gh + blockIdx.x * NB;
	const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint32_t x[PT];
	uint32_t inc = 0;
#pragma unroll
	for (int k = 0; k < PT; k++) {
		x[k] = g[PT * threadIdx.x + k];
		inc += x[k];
	}
	const uint32_t mine = inc;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
		if ((int)lane >= o) inc += y;
	}
	if (lane == 31) wsum[warp] = inc;
	__syncthreads();
	uint32_t ex = inc - mine;
	for (int w = 0; w < (int)warp; w++) ex += wsum[w];
#pragma unroll
	for (int k = 0; k < PT; k++) {
		g[PT * threadIdx.x + k] = ex;
		ex += x[k];
	}
}

static constexpr uint32_t RW_FLAG_A = 1u << 30, RW_FLAG_P = 2u << 30, RW_VMASK = (1u << 30) - 1u;

// DPT consecutive status words of one tile, read / written past L1 (they are polled)
template <int DPT>
__device__ __forceinline__ void rw_status_load(const uint32_t *p, uint32_t (&w)[DPT]) {
	if (DPT == 8) {
		asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
		             : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(p) : "memory");
		asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
		             : "=r"(w[4 % DPT]), "=r"(w[5 % DPT]), "=r"(w[6 % DPT]), "=r"(w[7 % DPT]) : "l"(p + 4) : "memory");
	} else {
		asm volatile("ld.volatile.global.v2.u32 {%0,%1}, [%2];" : "=r"(w[0]), "=r"(w[1]) : "l"(p) : "memory");
	}
}
template <int DPT>
__device__ __forceinline__ void rw_status_store(uint32_t *p, const uint32_t (&w)[DPT]) {
	if (DPT == 8) {
		asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};"
		             :: "l"(p), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
		asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};"
		             :: "l"(p + 4), "r"(w[4 % DPT]), "r"(w[5 % DPT]), "r"(w[6 % DPT]), "r"(w[7 % DPT]) : "memory");
	} else {
		asm volatile("st.volatile.global.v2.u32 [%0], {%1,%2};" :: "l"(p), "r"(w[0]), "r"(w[1]) : "memory");
	}
}

// Scatter kernel configuration: THREADS x ITEMS = RW_TILE keys per block.
//   <256, 32>: two blocks per SM (16 warps), long per-warp rank chains (default: 2.11 ms at R-MAT s22);
//   <1024, 8>: one block per SM (32 warps), per-warp counters take 128 KB of shared memory (2.74 ms; option sort_wide=2).
template <int THREADS, int NBITS>
struct RwCfg {
	static constexpr int NB = 1 << NBITS;
	static constexpr int WARPS = THREADS / 32;
	static constexpr int ITEMS = RW_TILE / THREADS;
	static constexpr int DPT = NB / THREADS; // digits per thread in the block scans (8 or 2)
	static constexpr size_t OFF_STAGE = 0;                                     // uint2[RW_TILE]
	static constexpr size_t OFF_GBASE = OFF_STAGE + (size_t)RW_TILE * 8;       // uint32[NB]
	static constexpr size_t OFF_CNT = OFF_GBASE + (size_t)NB * 4;              // uint16[WARPS][NB]
	static constexpr size_t OFF_WSUM = OFF_CNT + (size_t)WARPS * NB * 2;       // uint32[WARPS]
	static constexpr size_t SMEM = OFF_WSUM + (size_t)WARPS * 4 + 32;
	static_assert((size_t)WARPS * NB * 2 >= (size_t)RW_TILE * 4, "payload1 staging reuses the counter area");
	static_assert(DPT == 8 || DPT == 2, "one 16-byte or one 4-byte row of 16-bit counters per thread");
	static_assert(ITEMS % 4 == 0, "ballots are batched four rounds at a time");
	static_assert(SMEM <= 227 * 1024, "shared memory per block");
};

template <int IN, int THREADS, int NBITS>
__global__ void __launch_bounds__(THREADS, THREADS == 256 ? 2 : 1) rw_scatter_kernel(RwArgs a) {
	using C = RwCfg<THREADS, NBITS>;
	constexpr int WARPS = C::WARPS, ITEMS = C::ITEMS, DPT = C::DPT, NB = C::NB;
	extern __shared__ __align__(16) unsigned char rw_smem[];
	uint2 *stage = reinterpret_cast<uint2 *>(rw_smem + C::OFF_STAGE);
	uint32_t *gbase = reinterpret_cast<uint32_t *>(rw_smem + C::OFF_GBASE); // global bucket base - local start
	uint16_t *cnt = reinterpret_cast<uint16_t *>(rw_smem + C::OFF_CNT);
	uint32_t *wsum = reinterpret_cast<uint32_t *>(rw_smem + C::OFF_WSUM);
	const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const unsigned lt_mask = (1u << lane) - 1u;
	const uint32_t mask = (1u << a.nbits) - 1u;
	{
		uint4 *z = reinterpret_cast<uint4 *>(cnt);
#pragma unroll
		for (int i = 0; i < WARPS * NB * 2 / 16 / THREADS; i++) z[i * THREADS + threadIdx.x] = make_uint4(0u, 0u, 0u, 0u);
	}
	// Look-back variant: tiles are numbered in the order their blocks START, so every predecessor a tile
	// waits for is already running (forward progress whatever order the hardware schedules blocks in).
	__shared__ uint32_t s_tile;
	if (a.status && threadIdx.x == 0) s_tile = atomicAdd(a.ticket, 1u);
	if (a.status) __syncthreads();
	const uint32_t tile = a.status ? s_tile : blockIdx.x;
	// Each warp owns a contiguous run of the tile and walks it 32 keys at a time (stability).
	const uint32_t tbase = tile * (uint32_t)RW_TILE;
	const uint32_t wbase = tbase + warp * (uint32_t)(ITEMS * 32) + lane;
	uint32_t keys[ITEMS], vals[ITEMS];
	bool bad = false;
	if (a.n - tbase >= (uint32_t)RW_TILE) { // full tile: unguarded loads, all in flight together
#pragma unroll
		for (int r = 0; r < ITEMS; r++) {
			const uint2 it = rw_load_item<IN>(a, wbase + r * 32, bad);
			keys[r] = it.x;
			vals[r] = it.y;
		}
	} else {
#pragma unroll
		for (int r = 0; r < ITEMS; r++) {
			const uint32_t idx = wbase + r * 32;
			uint2 it = make_uint2(0u, 0u);
			if (idx < a.n) it = rw_load_item<IN>(a, idx, bad);
			keys[r] = it.x;
			vals[r] = it.y;
		}
	}
	// the reference would write out of bounds for such ids; the digits are masked, so nothing does here
	if (IN == RW_IN_64 && a.err && bad) atomicOr(a.err, 1u);
	__syncthreads();
	uint16_t *mycnt = cnt + warp * NB;
	uint32_t lp[ITEMS / 2]; // two 16-bit values per register: rank inside (warp, digit), later the tile position
#pragma unroll
	for (int r0 = 0; r0 < ITEMS; r0 += 4) {
		// Lanes of the warp holding the same digit, one ballot per digit bit (bits above nbits are zero
		// in every lane: peers &= ~0).  Four rounds' ballots are independent of the counters: issue them
		// together, then the four dependent counter updates.
		unsigned peers[4];
#pragma unroll
		for (int q = 0; q < 4; q++) {
			const uint32_t d = (keys[r0 + q] >> a.shift) & mask;
			unsigned pm = __ballot_sync(0xffffffffu, wbase + (r0 + q) * 32 < a.n);
#pragma unroll
			for (int b = 0; b < NBITS; b++) {
				const bool bit = (d >> b) & 1u;
				const unsigned m = __ballot_sync(0xffffffffu, bit);
				pm &= bit ? m : ~m;
			}
			peers[q] = pm;
		}
#pragma unroll
		for (int q = 0; q < 4; q++) {
			const int r = r0 + q;
			const bool valid = wbase + r * 32 < a.n;
			const uint32_t d = (keys[r] >> a.shift) & mask;
			const unsigned lower = peers[q] & lt_mask;
			uint32_t base = 0;
			if (valid) base = mycnt[d];
			__syncwarp();
			if (valid && lower == 0) mycnt[d] = (uint16_t)(base + __popc(peers[q]));
			__syncwarp();
			const uint32_t rank = base + __popc(lower);
			if (r & 1) lp[r >> 1] |= rank << 16;
			else lp[r >> 1] = rank;
		}
	}
	__syncthreads();
	// Exclusive scan over the warps, digit by digit; thread t owns digits DPT*t .. DPT*t+DPT-1 (one row
	// of DPT 16-bit counters per warp).
	uint32_t tot[DPT];
#pragma unroll
	for (int k = 0; k < DPT; k++) tot[k] = 0;
#pragma unroll 8
	for (int w = 0; w < WARPS; w++) {
		if (DPT == 8) {
			uint4 *row = reinterpret_cast<uint4 *>(cnt + w * NB) + threadIdx.x;
			const uint4 q = *row;
			const uint32_t x[4] = {q.x, q.y, q.z, q.w};
			uint32_t o[4];
#pragma unroll
			for (int k = 0; k < 4; k++) {
				o[k] = tot[(2 * k) % DPT] | (tot[(2 * k + 1) % DPT] << 16); // tile counts stay below 65536
				tot[(2 * k) % DPT] += x[k] & 0xFFFFu;
				tot[(2 * k + 1) % DPT] += x[k] >> 16;
			}
			*row = make_uint4(o[0], o[1], o[2], o[3]);
		} else {
			uint32_t *row = reinterpret_cast<uint32_t *>(cnt + w * NB) + threadIdx.x;
			const uint32_t x = *row;
			*row = tot[0] | (tot[1 % DPT] << 16);
			tot[0] += x & 0xFFFFu;
			tot[1 % DPT] += x >> 16;
		}
	}
	// Look-back: publish this tile's bucket totals, then add up the predecessors' until one of them
	// already knows its inclusive prefix (per bucket).
	uint32_t pre[DPT];
#pragma unroll
	for (int k = 0; k < DPT; k++) pre[k] = 0;
	if (a.status) {
		uint32_t *mine = a.status + (size_t)tile * NB + threadIdx.x * DPT;
		uint32_t w[DPT];
		if (tile != 0) {
#pragma unroll
			for (int k = 0; k < DPT; k++) w[k] = tot[k] | RW_FLAG_A;
			rw_status_store<DPT>(mine, w);
			uint32_t open = (1u << DPT) - 1u; // buckets still looking back
			uint32_t spins = 0;
			for (uint32_t t = tile - 1; open;) {
				rw_status_load<DPT>(a.status + (size_t)t * NB + threadIdx.x * DPT, w);
				bool ready = true;
#pragma unroll
				for (int k = 0; k < DPT; k++) ready &= (w[k] >> 30) != 0u;
				if (!ready) {
					if (++spins > (1u << 22)) { // never expected; do not hang the GPU
						if (a.err) atomicOr(a.err, 4u);
						break;
					}
					__nanosleep(20);
					continue;
				}
#pragma unroll
				for (int k = 0; k < DPT; k++) {
					if ((open >> k) & 1u) {
						pre[k] += w[k] & RW_VMASK;
						if ((w[k] >> 30) == 2u) open &= ~(1u << k);
					}
				}
				t--; // tile 0 always publishes inclusive prefixes, so t never passes it with `open` set
			}
		}
#pragma unroll
		for (int k = 0; k < DPT; k++) w[k] = ((pre[k] + tot[k]) & RW_VMASK) | RW_FLAG_P;
		rw_status_store<DPT>(mine, w);
	}
	// Exclusive scan of the tile totals over the digits.
	{
		uint32_t tsum = 0;
#pragma unroll
		for (int k = 0; k < DPT; k++) tsum += tot[k];
		uint32_t inc = tsum;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
			if ((int)lane >= o) inc += y;
		}
		if (lane == 31) wsum[warp] = inc;
		__syncthreads();
		uint32_t ex = inc - tsum;
#pragma unroll 8
		for (int w = 0; w < WARPS; w++) ex += (w < (int)warp) ? wsum[w] : 0u;
		const uint32_t nb = 1u << a.nbits;
		uint32_t ls[DPT]; // local start of each of my buckets inside the tile's sorted order
#pragma unroll
		for (int k = 0; k < DPT; k++) {
			const uint32_t d = threadIdx.x * DPT + k;
			ls[k] = ex;
			if (a.status) gbase[d] = a.gdig[d] + pre[k] - ex;
			else gbase[d] = (d < nb ? a.hist[(uint64_t)d * a.ntiles + tile] : 0u) - ex;
			ex += tot[k];
		}
		// fold the bucket start into every warp's exclusive base: one random shared-memory read per key
		// (instead of two) when the items are placed; all values stay below RW_TILE <= 65535
#pragma unroll 8
		for (int w = 0; w < WARPS; w++) {
			if (DPT == 8) {
				uint4 *row = reinterpret_cast<uint4 *>(cnt + w * NB) + threadIdx.x;
				uint4 q = *row;
				q.x += ls[0] | (ls[1 % DPT] << 16);
				q.y += ls[2 % DPT] | (ls[3 % DPT] << 16);
				q.z += ls[4 % DPT] | (ls[5 % DPT] << 16);
				q.w += ls[6 % DPT] | (ls[7 % DPT] << 16);
				*row = q;
			} else {
				uint32_t *row = reinterpret_cast<uint32_t *>(cnt + w * NB) + threadIdx.x;
				*row += ls[0] | (ls[1 % DPT] << 16);
			}
		}
	}
	__syncthreads();
	// Tile position of every item; items go to the staging area in the tile's sorted order.
#pragma unroll
	for (int r = 0; r < ITEMS; r++) {
		const bool valid = wbase + r * 32 < a.n;
		const uint32_t d = (keys[r] >> a.shift) & mask;
		const uint32_t rank = (r & 1) ? (lp[r >> 1] >> 16) : (lp[r >> 1] & 0xFFFFu);
		uint32_t p = 0;
		if (valid) {
			p = (uint32_t)mycnt[d] + rank;
			stage[p] = make_uint2(keys[r], vals[r]);
		}
		if (r & 1) lp[r >> 1] = (lp[r >> 1] & 0xFFFFu) | (p << 16);
		else lp[r >> 1] = (lp[r >> 1] & 0xFFFF0000u) | p;
	}
	__syncthreads();
	const uint32_t ntile = a.n - tbase < (uint32_t)RW_TILE ? a.n - tbase : (uint32_t)RW_TILE;
	// Consecutive threads write consecutive sorted slots: every bucket receives one contiguous run.
	if (a.pout) {
#pragma unroll 8
		for (int k = 0; k < ITEMS; k++) {
			const uint32_t j = threadIdx.x + k * THREADS;
			if (j < ntile) {
				const uint2 it = stage[j];
				a.pout[gbase[(it.x >> a.shift) & mask] + j] = it;
			}
		}
	} else {
#pragma unroll 8
		for (int k = 0; k < ITEMS; k++) {
			const uint32_t j = threadIdx.x + k * THREADS;
			if (j < ntile) {
				const uint2 it = stage[j];
				const uint32_t g = gbase[(it.x >> a.shift) & mask] + j;
				if (a.kout) a.kout[g] = it.x;
				a.p0out[g] = it.y;
			}
		}
	}
	if (a.p1out) {
		// payload1 goes through the counter area, which nobody reads any more (barrier above)
		uint32_t *stage1 = reinterpret_cast<uint32_t *>(cnt);
#pragma unroll
		for (int r = 0; r < ITEMS; r++) {
			const uint32_t idx = wbase + r * 32;
			const uint32_t p = (r & 1) ? (lp[r >> 1] >> 16) : (lp[r >> 1] & 0xFFFFu);
			if (idx < a.n) stage1[p] = a.p1in ? a.p1in[idx] : idx;
		}
		__syncthreads();
#pragma unroll 8
		for (int k = 0; k < ITEMS; k++) {
			const uint32_t j = threadIdx.x + k * THREADS;
			if (j < ntile) a.p1out[gbase[(stage[j].x >> a.shift) & mask] + j] = stage1[j];
		}
	}
}

template <int THREADS, int NBITS>
static int rw_launch_scatter(Ctx *c, int form, const RwArgs &a) {
	using C = RwCfg<THREADS, NBITS>;
	// per device and cheap: set on every call rather than cached per process
#define RW_LAUNCH(IN)                                                                                                  \
	do {                                                                                                               \
		PGQ_CUDA(cudaFuncSetAttribute(rw_scatter_kernel<IN, THREADS, NBITS>,                                           \
		                              cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM));                     \
		rw_scatter_kernel<IN, THREADS, NBITS><<<a.ntiles, THREADS, C::SMEM, c->stream>>>(a);                           \
	} while (0)
	if (form == RW_IN_64) RW_LAUNCH(RW_IN_64);
	else if (form == RW_IN_32) RW_LAUNCH(RW_IN_32);
	else RW_LAUNCH(RW_IN_PACKED);
#undef RW_LAUNCH
	count_launch(c);
	return PGQ_OK;
}

bool radix_sort_wide_eligible(const Ctx *c, const SortIO &io) {
	// 23 / 24-bit keys: the 12-bit shape only on request (sort_wide=3): measured 10.4 ms against 9.1 ms for the three
	// 8-bit passes at R-MAT s24 (268 M keys) -- 16 warps of one 208 KB block rank slower than two 110 KB blocks
	const int maxbits = c->opt.sort_wide == 3 ? 2 * RW_MAXBITS_BIG : 2 * RW_MAXBITS;
	return c->opt.sort_wide != 0 && io.key_bits >= 1 && io.key_bits <= maxbits &&
	       (int64_t)io.n >= c->opt.sort_wide_min && io.n < 0xFFFFFFFFu - 2u * (uint32_t)RW_TILE;
}

int radix_sort_wide(Ctx *c, const SortIO &io) {
	if (io.n == 0) return PGQ_OK;
	const bool big = io.key_bits > 2 * RW_MAXBITS; // 23 / 24-bit keys: 12-bit digits, 4096 buckets, 512-thread blocks
	const uint32_t NBS = big ? RW_NB_BIG : RW_NB;
	const int passes = io.key_bits <= RW_MAXBITS ? 1 : 2;
	const int bits0 = passes == 1 ? io.key_bits : (io.key_bits + 1) / 2;
	const int bits1 = io.key_bits - bits0;
	const uint32_t ntiles = div_up(io.n, RW_TILE);
	const bool want_p1 = io.out_p1 != nullptr;
	const size_t n4 = (size_t)io.n * sizeof(uint32_t);
	if (passes == 2) {
		PGQ_TRY(c->sort_pack.reserve(2 * n4));
		if (want_p1) PGQ_TRY(c->sort_p1[0].reserve(n4));
	}
	uint32_t *keys_out = io.out_keys;
	if (io.keys_scratch) {
		PGQ_TRY(c->sort_keys[0].reserve(n4));
		keys_out = c->sort_keys[0].as<uint32_t>();
		*io.keys_scratch = keys_out;
	}
	// look-back variant (option sort_lookback, counts must fit 30 bits): no per-tile histogram pre-pass and no
	// device-wide scan; one extra read of the first-pass keys gives the global bucket sizes of both passes
	const bool lookback = c->opt.sort_lookback != 0 && io.n < (1u << 30);
	const size_t status_bytes = (size_t)ntiles * NBS * sizeof(uint32_t);
	uint32_t *d_gdig = nullptr, *d_ticket = nullptr;
	if (lookback) {
		PGQ_TRY(c->sort_hist.reserve(status_bytes + 2 * NBS * sizeof(uint32_t) + 256));
		d_gdig = reinterpret_cast<uint32_t *>(static_cast<char *>(c->sort_hist.p) + status_bytes);
		d_ticket = d_gdig + 2 * NBS;
		PGQ_CUDA(cudaMemsetAsync(d_gdig, 0, 2 * NBS * sizeof(uint32_t) + 256, c->stream));
		RwArgs a{};
		a.k64 = io.k64;
		a.k32 = io.k32;
		a.n = io.n;
		a.ntiles = ntiles;
		a.nb = NBS;
		const unsigned grid = ntiles < (unsigned)c->sm_count * 8u ? ntiles : (unsigned)c->sm_count * 8u;
		if (io.k64) rw_global_hist_kernel<RW_IN_64><<<grid, RW_THREADS, 0, c->stream>>>(a, bits0, bits1, d_gdig);
		else rw_global_hist_kernel<RW_IN_32><<<grid, RW_THREADS, 0, c->stream>>>(a, bits0, bits1, d_gdig);
		if (big) rw_gdig_scan_kernel<RW_NB_BIG><<<2, 1024, 0, c->stream>>>(d_gdig);
		else rw_gdig_scan_kernel<RW_NB><<<2, 1024, 0, c->stream>>>(d_gdig);
		count_launch(c, 2);
	} else {
		PGQ_TRY(c->sort_hist.reserve(((size_t)ntiles << bits0) * sizeof(uint32_t)));
	}
	for (int p = 0; p < passes; p++) {
		RwArgs a{};
		const bool first = p == 0, last = p == passes - 1;
		if (first) {
			a.k64 = io.k64;
			a.p64 = io.p64;
			a.k32 = io.k32;
			a.p32 = io.p32;
			a.vmax = io.vmax;
		} else {
			a.pin = c->sort_pack.as<uint2>();
			a.p1in = want_p1 ? c->sort_p1[0].as<uint32_t>() : nullptr;
		}
		// out-of-range ids (64-bit form) and the look-back watchdog report here; without a caller's flag the
		// watchdog gets a private one, checked below
		a.err = io.err ? io.err : (lookback ? d_ticket + 8 : nullptr);
		if (last) {
			a.kout = keys_out;
			a.p0out = io.out_p0;
			a.p1out = io.out_p1;
		} else {
			a.pout = c->sort_pack.as<uint2>();
			a.p1out = want_p1 ? c->sort_p1[0].as<uint32_t>() : nullptr;
		}
		a.hist = c->sort_hist.as<uint32_t>();
		a.n = io.n;
		a.ntiles = ntiles;
		a.nb = NBS;
		a.shift = first ? 0 : bits0;
		a.nbits = first ? bits0 : bits1;
		const int form = !first ? RW_IN_PACKED : io.k64 ? RW_IN_64 : RW_IN_32;
		if (lookback) {
			a.status = c->sort_hist.as<uint32_t>();
			a.ticket = d_ticket + p;
			a.gdig = d_gdig + p * NBS;
			PGQ_CUDA(cudaMemsetAsync(a.status, 0, status_bytes, c->stream));
		} else {
			if (form == RW_IN_64) rw_hist_kernel<RW_IN_64><<<ntiles, RW_THREADS, 0, c->stream>>>(a);
			else if (form == RW_IN_32) rw_hist_kernel<RW_IN_32><<<ntiles, RW_THREADS, 0, c->stream>>>(a);
			else rw_hist_kernel<RW_IN_PACKED><<<ntiles, RW_THREADS, 0, c->stream>>>(a);
			count_launch(c);
			PGQ_TRY(exclusive_scan_u32(c, a.hist, a.hist, (uint64_t)ntiles << a.nbits, nullptr));
		}
		if (big) PGQ_TRY((rw_launch_scatter<512, RW_MAXBITS_BIG>(c, form, a)));
		else if (c->opt.sort_wide == 2) PGQ_TRY((rw_launch_scatter<1024, RW_MAXBITS>(c, form, a)));
		else PGQ_TRY((rw_launch_scatter<256, RW_MAXBITS>(c, form, a)));
		PGQ_CUDA(cudaGetLastError());
	}
	if (lookback && !io.err) {
		uint32_t h = 0;
		PGQ_CUDA(cudaMemcpyAsync(&h, d_ticket + 8, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
		PGQ_CUDA(cudaStreamSynchronize(c->stream));
		if (h & 4u) {
			set_error("radix sort look-back watchdog fired (a tile never published its bucket totals)");
			return PGQ_ERR_CUDA;
		}
	}
	return PGQ_OK;
}

} // namespace pgq
