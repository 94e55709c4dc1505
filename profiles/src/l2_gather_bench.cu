// Gather-bandwidth microbenchmark for the roofline of the hash-grid encode ("L2 gather ceiling to be taken from a
// measured L2 bandwidth microbenchmark on the box"). Random 4-byte and 8-byte reads (one 32-byte sector each) from a table that fits
// L2 (24 MB: the 2^19 model's hash table) and from one that does not (268 MB: the 2^22 model's). Every thread issues 16 independent
// loads per trip, like a lane of the march does for two levels. Prints one JSON object. Not part of the product.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o profiles/_bin/l2_gather_bench profiles/src/l2_gather_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <typename T>
__global__ void k_gather(const T* __restrict__ table, uint32_t mask, uint32_t trips, unsigned long long* __restrict__ sink) {
	uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
	unsigned long long acc = 0;
	for (uint32_t t = 0; t < trips; ++t) {
		uint32_t idx[16];
#pragma unroll
		for (int i = 0; i < 16; ++i) { s = s * 1664525u + 1013904223u; idx[i] = (s >> 4) & mask; }
		T v[16];
#pragma unroll
		for (int i = 0; i < 16; ++i) v[i] = __ldg(table + idx[i]);
#pragma unroll
		for (int i = 0; i < 16; ++i) acc += (unsigned long long)(*reinterpret_cast<const uint32_t*>(&v[i]));
	}
	if (acc == 0xdeadbeefULL) *sink = acc;
}

template <typename T>
static double run(size_t bytes, int blocks, int threads, uint32_t trips) {
	T* table; unsigned long long* sink;
	const size_t n = bytes / sizeof(T);
	cudaMalloc(&table, bytes); cudaMalloc(&sink, 8);
	cudaMemset(table, 1, bytes);
	uint32_t mask = 1; while ((size_t)mask * 2 <= n) mask *= 2; mask -= 1;
	cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
	k_gather<T><<<blocks, threads>>>(table, mask, trips / 4, sink);      // warm-up: pulls the table into L2 where it fits
	double best = 0.0;
	for (int rep = 0; rep < 5; ++rep) {
		cudaEventRecord(a);
		k_gather<T><<<blocks, threads>>>(table, mask, trips, sink);
		cudaEventRecord(b); cudaEventSynchronize(b);
		float ms; cudaEventElapsedTime(&ms, a, b);
		const double loads = (double)blocks * threads * trips * 16.0;
		const double gloads = loads / (ms * 1e-3) / 1e9;
		if (gloads > best) best = gloads;
	}
	cudaFree(table); cudaFree(sink);
	return best;   // G loads / s
}

int main() {
	cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
	const int blocks = p.multiProcessorCount * 8, threads = 256;
	const uint32_t trips = 256;
	const double l2_4 = run<uint32_t>(24u << 20, blocks, threads, trips), l2_8 = run<uint2>(24u << 20, blocks, threads, trips);
	const double hbm_4 = run<uint32_t>(268u << 20, blocks, threads, trips), hbm_8 = run<uint2>(268u << 20, blocks, threads, trips);
	printf("{\"gpu\": \"%s\", \"sms\": %d, \"l2_gather_gloads_4B\": %.2f, \"l2_gather_gbs\": %.1f, \"l2_gather_sector_gbs\": %.1f, \"l2_gather_gloads_8B\": %.2f, \"l2_gather_gbs_8B\": %.1f, "
	       "\"hbm_gather_gloads_4B\": %.2f, \"hbm_gather_gbs\": %.1f, \"hbm_gather_sector_gbs\": %.1f, \"hbm_gather_gloads_8B\": %.2f, \"hbm_gather_gbs_8B\": %.1f, "
	       "\"how\": \"random reads, 16 independent loads per thread per trip, %d blocks x %d threads, best of 5; gbs = useful bytes, sector_gbs = 32 B per load; tables 24 MB (L2 resident) and 268 MB\"}\n",
	       p.name, p.multiProcessorCount, l2_4, l2_4 * 4, l2_4 * 32, l2_8, l2_8 * 8, hbm_4, hbm_4 * 4, hbm_4 * 32, hbm_8, hbm_8 * 8, blocks, threads);
	return 0;
}
