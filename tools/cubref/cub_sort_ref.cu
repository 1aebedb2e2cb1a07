// Reference point for the CSR build's sort (not part of the product, not linked into it): cub::DeviceRadixSort::SortPairs
// (CUB's onesweep, the library baseline) on the same (source, target) rows, keys narrowed to uint32 beforehand.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -shared -Xcompiler -fPIC -o libcubref.so cub_sort_ref.cu
#include <cub/device/device_radix_sort.cuh>
#include <cstdint>
#include <cstdio>

__global__ void narrow_kernel(const long long *a, const long long *b, uint32_t *ka, uint32_t *kb, size_t n) {
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) {
		ka[i] = (uint32_t)a[i];
		kb[i] = (uint32_t)b[i];
	}
}

// d_src / d_dst: int64 device columns; returns the milliseconds of ONE SortPairs over key bits [0, bits) (CUDA events,
// median of reps), keys/values already narrowed (the narrowing pass is timed separately into *narrow_ms)
extern "C" int cubref_sort_ms(const long long *d_src, const long long *d_dst, long long n, int bits, int reps, float *sort_ms,
                              float *narrow_ms) {
	uint32_t *k0, *k1, *v0, *v1;
	cudaMalloc(&k0, n * 4); cudaMalloc(&k1, n * 4); cudaMalloc(&v0, n * 4); cudaMalloc(&v1, n * 4);
	cudaEvent_t a, b;
	cudaEventCreate(&a); cudaEventCreate(&b);
	cudaEventRecord(a);
	narrow_kernel<<<(unsigned)((n + 255) / 256), 256>>>(d_src, d_dst, k0, v0, (size_t)n);
	cudaEventRecord(b); cudaEventSynchronize(b);
	cudaEventElapsedTime(narrow_ms, a, b);
	size_t tmp_bytes = 0;
	cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, k0, k1, v0, v1, (int)n, 0, bits);
	void *tmp; cudaMalloc(&tmp, tmp_bytes);
	float best = 1e30f, all[32];
	if (reps > 32) reps = 32;
	for (int r = 0; r < reps; r++) {
		cudaEventRecord(a);
		cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, k0, k1, v0, v1, (int)n, 0, bits);
		cudaEventRecord(b); cudaEventSynchronize(b);
		cudaEventElapsedTime(&all[r], a, b);
		if (all[r] < best) best = all[r];
	}
	// median
	for (int i = 0; i < reps; i++) for (int j = i + 1; j < reps; j++) if (all[j] < all[i]) { float t = all[i]; all[i] = all[j]; all[j] = t; }
	*sort_ms = all[reps / 2];
	int rc = cudaGetLastError() == cudaSuccess ? 0 : 1;
	cudaFree(k0); cudaFree(k1); cudaFree(v0); cudaFree(v1); cudaFree(tmp);
	cudaEventDestroy(a); cudaEventDestroy(b);
	return rc;
}
