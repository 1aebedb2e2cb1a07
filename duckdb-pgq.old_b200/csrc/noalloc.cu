// pgq_iterativelength_device_noalloc — an allocation-free variant of pgq_iterativelength_device.
//
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It runs the
// whole-batch small-graph kernel (msbfs_small.cuh: one cooperative launch per 64..512-lane batch) with ALL of
// its scratch -- control block, lane tables, `seen` and the two frontier buffers -- inside one file-scope
// static __device__ array of NOALLOC_SCRATCH_BYTES = 8 MiB.  Every GPU that loads the library holds that array
// until the process exits.  The caller passes inputs and outputs (device pointers) exactly as for
// pgq_iterativelength_device; this entry point calls no cudaMalloc / cudaFree / cudaHostAlloc and no allocating
// library routine.
//
// Limits (rejected with PGQ_ERR_INVALID / PGQ_ERR_STATE and a message naming them):
//   n <= 4096 rows per call;  64 KiB + 3 * roundup256(V * W/8) <= 8 MiB  (W = lane width of the call: 64..512,
//   chosen like the existing entry point does), e.g. V <= 43 000 at W = 512, V <= 348 000 at W = 64;
//   the reverse CSR and in-degree class lists must exist already (pgq_csr_prepare, or any earlier query).
//
// The array is shared by every call on the device and keeps the last call's contents.  What must start at zero
// is cleared per batch (the kernel's barrier counter and level slots, the error flag; the kernel itself zeroes
// `seen` and the first frontier, plants `found = -1`, and every level rewrites the whole next frontier).  Calls
// that overlap on the GPU -- two contexts on one device, or the same context from two host threads without the
// library's per-context lock -- would race on it; calls on one context are serialised by that lock and by
// stream order.
//
// In host code the array's NAME is a host-side stand-in, not a device address; the address used by the launches
// comes from cudaGetSymbolAddress, once per device, with that device current (noalloc_scratch()).
#include "msbfs_small.cuh"

#include <map>
#include <mutex>

namespace pgq {

namespace {

constexpr size_t NOALLOC_SCRATCH_BYTES = (size_t)8 << 20;
constexpr int64_t NOALLOC_MAX_ROWS = 4096;
constexpr size_t NA_CTRL_OFF = 0;       // BfsCtrl (< 4 KiB)
constexpr size_t NA_ERR_OFF = 4096;     // 16-byte error flag
constexpr size_t NA_LANE_OFF = 4352;    // lane src | dst | found, NA_MAX_SLOTS entries each
constexpr int64_t NA_MAX_SLOTS = NOALLOC_MAX_ROWS + 512;
constexpr size_t NA_BITMAP_OFF = 65536; // seen | frontier 0 | frontier 1
static_assert(sizeof(BfsCtrl) <= NA_ERR_OFF, "control block outgrew its slot");
static_assert(NA_LANE_OFF + 3 * NA_MAX_SLOTS * 4 <= NA_BITMAP_OFF, "lane tables outgrew their slot");

static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

// rows -> lane slots (slot i <-> row i; unused / trivial rows take no lane), same rules as the existing entry point
__global__ void __launch_bounds__(256) na_lanes_kernel(const int64_t *__restrict__ src, const int64_t *__restrict__ dst,
                                                       int64_t n, int64_t nslots, uint64_t V,
                                                       uint32_t *__restrict__ lane_src, uint32_t *__restrict__ lane_dst,
                                                       uint32_t *err) {
	const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= nslots) return;
	uint32_t s = 0xFFFFFFFFu, d = 0;
	if (i < n) {
		const uint64_t a = (uint64_t)src[i], b = (uint64_t)dst[i];
		if (a >= V || b >= V) atomicOr(err, 1u);
		else if (a != b) {
			s = (uint32_t)a;
			d = (uint32_t)b;
		}
	}
	lane_src[i] = s;
	lane_dst[i] = d;
}

__global__ void __launch_bounds__(256) na_rows_kernel(const int64_t *__restrict__ src, const int64_t *__restrict__ dst,
                                                      int64_t n, const int32_t *__restrict__ found,
                                                      int64_t *__restrict__ out, uint8_t *__restrict__ valid) {
	const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	if (src[i] == dst[i]) {
		out[i] = 0;
		valid[i] = 1;
	} else {
		const int32_t f = found[i];
		out[i] = f >= 0 ? (int64_t)f : -1;
		valid[i] = f >= 0;
	}
}

// Device address of the static scratch on the CURRENT device (looked up once per device).
int noalloc_scratch(int device, unsigned char **out) {
	static std::mutex mu;
	static std::map<int, unsigned char *> addr;
	std::lock_guard<std::mutex> lk(mu);
	auto it = addr.find(device);
	if (it == addr.end()) {
		unsigned char *p = nullptr;
		const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
		if (e != cudaSuccess || !p) {
			set_error("cudaGetSymbolAddress(noalloc scratch) on device %d -> %s", device, cudaGetErrorString(e));
			(void)cudaGetLastError();
			return PGQ_ERR_CUDA;
		}
		it = addr.insert(std::make_pair(device, p)).first;
	}
	*out = it->second;
	return PGQ_OK;
}

template <int W>
int na_per_sm(int *out) {
	static int per_sm = -1; // per instantiation
	if (per_sm < 0) {
		int nb = 0;
		PGQ_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, msbfs_small_kernel<W>, SMALL_THREADS, 0));
		per_sm = nb;
	}
	*out = per_sm;
	return PGQ_OK;
}

} // namespace

} // namespace pgq

using namespace pgq;

extern "C" int pgq_iterativelength_device_noalloc(pgq_ctx_t *ctx, int32_t id, int64_t n, const int64_t *d_src,
                                                  const int64_t *d_dst, int64_t *d_out, uint8_t *d_out_valid) {
	if (!ctx) {
		set_error("null context");
		return PGQ_ERR_INVALID;
	}
	Ctx *c = &ctx->c;
	std::lock_guard<std::mutex> guard(c->mu);
	PGQ_CUDA(cudaSetDevice(c->device));
	auto it = c->csrs.find(id);
	if (it == c->csrs.end()) {
		set_error("unknown csr id %d", id);
		return PGQ_ERR_INVALID;
	}
	Csr *g = it->second;
	if (g->state != Csr::READY || g->is_part) {
		set_error("csr %d is not a finalized single-GPU csr", id);
		return PGQ_ERR_STATE;
	}
	if (!g->prepared || g->bins_small_max < 0) {
		set_error("pgq_iterativelength_device_noalloc: csr %d has no reverse CSR yet; call pgq_csr_prepare once "
		          "(it allocates) before the allocation-free calls", id);
		return PGQ_ERR_STATE;
	}
	if (n < 0 || n > NOALLOC_MAX_ROWS) {
		set_error("pgq_iterativelength_device_noalloc: %lld rows, the limit is %lld per call", (long long)n,
		          (long long)NOALLOC_MAX_ROWS);
		return PGQ_ERR_INVALID;
	}
	if (n == 0) return PGQ_OK;
	if (!d_src || !d_dst || !d_out || !d_out_valid) {
		set_error("pgq_iterativelength_device_noalloc: null device pointer");
		return PGQ_ERR_INVALID;
	}
	const int W = c->opt.lanes ? (int)c->opt.lanes : pick_lane_width(n);
	const int64_t nbatch = (n + W - 1) / W, nslots = nbatch * W;
	const size_t bm = (((size_t)g->V * (size_t)(W / 8)) + 255) & ~(size_t)255;
	if (g->V <= 0 || NA_BITMAP_OFF + 3 * bm > NOALLOC_SCRATCH_BYTES) {
		set_error("pgq_iterativelength_device_noalloc: V=%lld at %d lanes needs %zu bytes of scratch, the static "
		          "array holds %zu (64 KiB + 3 * V * lanes/8 <= 8 MiB)", (long long)g->V, W, NA_BITMAP_OFF + 3 * bm,
		          NOALLOC_SCRATCH_BYTES);
		return PGQ_ERR_INVALID;
	}
	unsigned char *base = nullptr;
	PGQ_TRY(noalloc_scratch(c->device, &base));
	cudaStream_t st = c->stream;
	BfsCtrl *ctrl = reinterpret_cast<BfsCtrl *>(base + NA_CTRL_OFF);
	uint32_t *d_err = reinterpret_cast<uint32_t *>(base + NA_ERR_OFF);
	uint32_t *lane_src = reinterpret_cast<uint32_t *>(base + NA_LANE_OFF);
	uint32_t *lane_dst = lane_src + NA_MAX_SLOTS;
	int32_t *found = reinterpret_cast<int32_t *>(lane_dst + NA_MAX_SLOTS);
	uint32_t *seen = reinterpret_cast<uint32_t *>(base + NA_BITMAP_OFF);
	uint32_t *buf0 = reinterpret_cast<uint32_t *>(base + NA_BITMAP_OFF + bm);
	uint32_t *buf1 = reinterpret_cast<uint32_t *>(base + NA_BITMAP_OFF + 2 * bm);
	int per_sm = 0;
	const void *fn = nullptr;
	switch (W) {
	case 64: PGQ_TRY(na_per_sm<64>(&per_sm)); fn = (const void *)msbfs_small_kernel<64>; break;
	case 128: PGQ_TRY(na_per_sm<128>(&per_sm)); fn = (const void *)msbfs_small_kernel<128>; break;
	case 256: PGQ_TRY(na_per_sm<256>(&per_sm)); fn = (const void *)msbfs_small_kernel<256>; break;
	default: PGQ_TRY(na_per_sm<512>(&per_sm)); fn = (const void *)msbfs_small_kernel<512>; break;
	}
	if (per_sm < 1) {
		set_error("pgq_iterativelength_device_noalloc: the whole-batch kernel does not fit an SM on this device");
		return PGQ_ERR_CUDA;
	}
	int grid = c->sm_count; // one 1024-thread block per SM (cooperative launch: all blocks resident)
	const int want = (int)div_up((uint64_t)g->V, SMALL_THREADS / 32);
	if (grid > want) grid = want;
	if (grid < 1) grid = 1;
	PGQ_CUDA(cudaMemsetAsync(d_err, 0, 16, st));
	na_lanes_kernel<<<div_up(nslots, 256), 256, 0, st>>>(d_src, d_dst, n, nslots, (uint64_t)g->V, lane_src, lane_dst,
	                                                     d_err);
	count_launch(c);
	for (int64_t bi = 0; bi < nbatch; bi++) {
		SmallArgs a{};
		a.ioff = g->d_ioff.as<uint32_t>();
		a.isrc = g->d_isrc.as<uint32_t>();
		a.V = (uint32_t)g->V;
		a.team_threads = 256;
		a.warp_max = (uint32_t)g->bins_small_max;
		a.big[0] = g->d_med.as<uint32_t>();
		a.nbig[0] = g->n_med;
		a.big[1] = g->d_mid.as<uint32_t>();
		a.nbig[1] = g->n_mid;
		a.big[2] = g->d_hub.as<uint32_t>();
		a.nbig[2] = g->n_hub;
		a.seen = seen;
		a.buf[0] = buf0;
		a.buf[1] = buf1;
		a.hist = nullptr;
		a.max_levels = 0x7fffffff;
		a.level_cap = 0x7fffffff;
		a.lane_src = lane_src + bi * W;
		a.lane_dst = lane_dst + bi * W;
		a.found = found + bi * W;
		a.ctrl = ctrl;
		// the kernel's grid-barrier counter and level slots must start at zero (they keep the last batch's values)
		PGQ_CUDA(cudaMemsetAsync(&ctrl->small_bar, 0, offsetof(BfsCtrl, k_listed) - offsetof(BfsCtrl, small_bar), st));
		void *args[] = {&a};
		PGQ_CUDA(cudaLaunchCooperativeKernel(fn, dim3((unsigned)grid), dim3(SMALL_THREADS), args, 0, st));
		count_launch(c);
		c->stats.batches++;
	}
	na_rows_kernel<<<div_up(n, 256), 256, 0, st>>>(d_src, d_dst, n, found, d_out, d_out_valid);
	count_launch(c);
	// the error flag comes back through the context's pinned control mirror (allocated at pgq_init)
	// (bytes [1536, 1540): past the BfsCtrl mirror, before the small-path level counts at 2048)
	static_assert(sizeof(BfsCtrl) <= 1536, "pinned mirror layout");
	uint32_t *h_err = reinterpret_cast<uint32_t *>(static_cast<char *>(c->h_ctrl) + 1536);
	PGQ_CUDA(cudaMemcpyAsync(h_err, d_err, 4, cudaMemcpyDeviceToHost, st));
	PGQ_CUDA(cudaStreamSynchronize(st));
	PGQ_CUDA(cudaGetLastError());
	if (*h_err) {
		set_error("iterativelength: src/dst id outside [0, v_size)");
		return PGQ_ERR_INVALID;
	}
	return PGQ_OK;
}
