// synerfgine_b200 — device-side BVH construction and refit.
//
// The reference builds every object's BVH on the host at load time (TriangleBvhWithBranchingFactor<2>::build, src/triangle_bvh.cu:615-718:
// recursive median split with std::nth_element, single-threaded) and never touches it again: VirtualObject::next_frame
// (include/synerfgine/virtual_object.cuh:53-64) animates the rigid transform only, and rays are taken into object space.
// snb_bvh_build (snb_scene.cpp) restates that builder and stays the parity mode (triangle order and ids bit-exact).
//
// This file is the B200 path for meshes that are large or that deform:
//   * build: linear BVH (Karras 2012) — centroid Morton codes, one radix sort, the radix tree over clusters of `prims_per_leaf`
//     consecutive triangles, bottom-up bounds — entirely on the device, in the SAME node format the traversal kernels read
//     (8 floats: lo, hi, left, right; the children of a node are adjacent; a leaf holds the negative [begin, end) triangle range), so
//     bvh_closest / the shadow any-hit walk it unchanged;
//   * refit: bounds of every node recomputed bottom-up from the current triangles (vertex animation) for ANY tree in that format,
//     the host-built one included.
// The only library piece is cub::DeviceRadixSort for the 64-bit key sort (a load-time step, not the render path).
#pragma once
#include <algorithm>
#include <cub/device/device_radix_sort.cuh>
#include "snb_device.cuh"   // included from snb_context.cuh after DevBuf

namespace snb {

struct LbvhScratch {
	DevBuf keys, keys_sorted, sort_tmp, islot, parent, flags, bounds;
};

// order-preserving float <-> uint map for atomicMin / atomicMax
SNB_D unsigned int f2ord(float f) { const unsigned int u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
SNB_D float ord2f(unsigned int u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u); }

SNB_D v3 tri_centroid(const float* __restrict__ t) {
	return mk3((t[0] + t[3] + t[6]) * (1.0f / 3.0f), (t[1] + t[4] + t[7]) * (1.0f / 3.0f), (t[2] + t[5] + t[8]) * (1.0f / 3.0f));
}

__global__ void k_lbvh_init_bounds(unsigned int* __restrict__ b) {
	if (threadIdx.x < 3) b[threadIdx.x] = 0xFFFFFFFFu; else if (threadIdx.x < 6) b[threadIdx.x] = 0u;
}

// centroid bounds: warp reduction, then 6 atomics per warp
__global__ void __launch_bounds__(256) k_lbvh_bounds(const float* __restrict__ tris, int n, unsigned int* __restrict__ b) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	v3 lo = mk3(3.4e38f, 3.4e38f, 3.4e38f), hi = mk3(-3.4e38f, -3.4e38f, -3.4e38f);
	if (i < n) { const v3 c = tri_centroid(tris + (size_t)i * 9); lo = c; hi = c; }
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		lo.x = fminf(lo.x, __shfl_xor_sync(0xFFFFFFFFu, lo.x, o)); lo.y = fminf(lo.y, __shfl_xor_sync(0xFFFFFFFFu, lo.y, o)); lo.z = fminf(lo.z, __shfl_xor_sync(0xFFFFFFFFu, lo.z, o));
		hi.x = fmaxf(hi.x, __shfl_xor_sync(0xFFFFFFFFu, hi.x, o)); hi.y = fmaxf(hi.y, __shfl_xor_sync(0xFFFFFFFFu, hi.y, o)); hi.z = fmaxf(hi.z, __shfl_xor_sync(0xFFFFFFFFu, hi.z, o));
	}
	if ((threadIdx.x & 31) == 0) {
		atomicMin(b + 0, f2ord(lo.x)); atomicMin(b + 1, f2ord(lo.y)); atomicMin(b + 2, f2ord(lo.z));
		atomicMax(b + 3, f2ord(hi.x)); atomicMax(b + 4, f2ord(hi.y)); atomicMax(b + 5, f2ord(hi.z));
	}
}

SNB_D unsigned int expand_bits_10(unsigned int v) {   // 10 bits -> every third bit
	v = (v * 0x00010001u) & 0xFF0000FFu;
	v = (v * 0x00000101u) & 0x0F00F00Fu;
	v = (v * 0x00000011u) & 0xC30C30C3u;
	v = (v * 0x00000005u) & 0x49249249u;
	return v;
}

// key = 30-bit Morton code of the centroid << 32 | triangle index: unique, so the radix tree needs no tie handling
__global__ void __launch_bounds__(256) k_lbvh_keys(const float* __restrict__ tris, int n, const unsigned int* __restrict__ b, unsigned long long* __restrict__ keys) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const v3 lo = mk3(ord2f(b[0]), ord2f(b[1]), ord2f(b[2])), hi = mk3(ord2f(b[3]), ord2f(b[4]), ord2f(b[5]));
	const v3 c = tri_centroid(tris + (size_t)i * 9);
	const float ex = fmaxf(hi.x - lo.x, 1e-30f), ey = fmaxf(hi.y - lo.y, 1e-30f), ez = fmaxf(hi.z - lo.z, 1e-30f);
	const unsigned int qx = (unsigned int)fminf(fmaxf((c.x - lo.x) / ex * 1024.0f, 0.0f), 1023.0f);
	const unsigned int qy = (unsigned int)fminf(fmaxf((c.y - lo.y) / ey * 1024.0f, 0.0f), 1023.0f);
	const unsigned int qz = (unsigned int)fminf(fmaxf((c.z - lo.z) / ez * 1024.0f, 0.0f), 1023.0f);
	const unsigned int morton = (expand_bits_10(qx) << 2) | (expand_bits_10(qy) << 1) | expand_bits_10(qz);
	keys[i] = ((unsigned long long)morton << 32) | (unsigned int)i;
}

// triangles into Morton order (one thread per float so the stores coalesce)
__global__ void __launch_bounds__(256) k_lbvh_gather(const float* __restrict__ tris, const unsigned long long* __restrict__ keys_sorted, int n, float* __restrict__ out, int* __restrict__ order_out) {
	const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= (long long)n * 9) return;
	const int j = (int)(g / 9), k = (int)(g % 9);
	const unsigned int src = (unsigned int)(keys_sorted[j] & 0xFFFFFFFFull);
	out[g] = tris[(size_t)src * 9 + k];
	if (order_out && k == 0) order_out[j] = (int)src;
}

// new vertex positions of a device-built object: src in the caller's order -> dst in the tree's order
__global__ void __launch_bounds__(256) k_gather_tris(const float* __restrict__ src, const int* __restrict__ order, int n, float* __restrict__ dst) {
	const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= (long long)n * 9) return;
	dst[g] = src[(size_t)order[g / 9] * 9 + (g % 9)];
}

// delta(i, j) of Karras 2012 over the cluster keys (cluster c = triangles [c * ppl, (c + 1) * ppl) in sorted order, key of its first triangle)
SNB_D int lbvh_delta(const unsigned long long* __restrict__ keys, int ppl, int L, int i, int j) {
	if (j < 0 || j >= L) return -1;
	return __clzll((long long)(keys[(size_t)i * ppl] ^ keys[(size_t)j * ppl]));
}

// One thread per internal node i of the radix tree (L - 1 of them): its children go to slots 1 + 2i and 2 + 2i of the node array; a leaf
// child is written right away (range only; bounds come from the fit), an internal child j learns its slot (islot[j]).
__global__ void __launch_bounds__(256) k_lbvh_topology(const unsigned long long* __restrict__ keys, int n, int ppl, int L, float* __restrict__ nodes, int* __restrict__ islot) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= L - 1) return;
	const int d = (lbvh_delta(keys, ppl, L, i, i + 1) - lbvh_delta(keys, ppl, L, i, i - 1)) >= 0 ? 1 : -1;
	const int dmin = lbvh_delta(keys, ppl, L, i, i - d);
	int lmax = 2;
	while (lbvh_delta(keys, ppl, L, i, i + lmax * d) > dmin) lmax <<= 1;
	int l = 0;
	for (int t = lmax >> 1; t >= 1; t >>= 1) if (lbvh_delta(keys, ppl, L, i, i + (l + t) * d) > dmin) l += t;
	const int j = i + l * d;
	const int dnode = lbvh_delta(keys, ppl, L, i, j);
	int s = 0, t = l;
	do {
		t = (t + 1) >> 1;
		if (lbvh_delta(keys, ppl, L, i, i + (s + t) * d) > dnode) s += t;
	} while (t > 1);
	const int gamma = i + s * d + min(d, 0);
	const int lo = min(i, j), hi = max(i, j);
	const int child[2] = {gamma, gamma + 1};
	const bool leaf[2] = {lo == gamma, hi == gamma + 1};
#pragma unroll
	for (int c = 0; c < 2; ++c) {
		const int slot = 1 + 2 * i + c;
		if (leaf[c]) {
			const int begin = child[c] * ppl, end = min((child[c] + 1) * ppl, n);
			int* p = reinterpret_cast<int*>(nodes + (size_t)slot * 8);
			p[6] = -begin - 1; p[7] = -end - 1;
		} else islot[child[c]] = slot;
	}
	if (i == 0) islot[0] = 0;
}

// internal node j -> its slot gets the child pointers (left = 1 + 2j, right = left + 2 like the host builder writes them)
__global__ void __launch_bounds__(256) k_lbvh_link(int L, const int* __restrict__ islot, float* __restrict__ nodes) {
	const int j = blockIdx.x * blockDim.x + threadIdx.x;
	if (j >= L - 1) return;
	int* p = reinterpret_cast<int*>(nodes + (size_t)islot[j] * 8);
	p[6] = 1 + 2 * j; p[7] = 3 + 2 * j;
}

__global__ void k_lbvh_single_leaf(int n, float* __restrict__ nodes) {
	int* p = reinterpret_cast<int*>(nodes);
	p[6] = -0 - 1; p[7] = -n - 1;
}

// ---- bounds: generic bottom-up fit over the node format (used by the build and by refit) ----------------------------------------------
// parent[] of every node from the child pointers
__global__ void __launch_bounds__(256) k_bvh_parents(const float* __restrict__ nodes, int n_nodes, int* __restrict__ parent, unsigned int* __restrict__ flags) {
	const int s = blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= n_nodes) return;
	flags[s] = 0u;
	if (s == 0) parent[0] = -1;
	const int left = reinterpret_cast<const int*>(nodes + (size_t)s * 8)[6];
	if (left >= 0) { parent[left] = s; parent[left + 1] = s; }
}

// One thread per node; leaves compute their box from their triangles and climb: the second child to arrive at a parent (atomic counter)
// merges the two child boxes and carries on, so every internal node is written exactly once, after both children. No thread waits on
// another one: the first arrival just leaves.
__global__ void __launch_bounds__(256) k_bvh_fit(float* nodes, int n_nodes, const float* __restrict__ tris, const int* __restrict__ parent, unsigned int* flags) {
	const int s = blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= n_nodes) return;
	const int* pi = reinterpret_cast<const int*>(nodes + (size_t)s * 8);
	const int left = pi[6], right = pi[7];
	if (left >= 0) return;
	const int begin = -left - 1, end = -right - 1;
	v3 lo = mk3(0.0f, 0.0f, 0.0f), hi = lo;
	if (end > begin) {
		lo = mk3(3.4e38f, 3.4e38f, 3.4e38f); hi = mk3(-3.4e38f, -3.4e38f, -3.4e38f);
		for (int t = begin; t < end; ++t) {
			const float* p = tris + (size_t)t * 9;
#pragma unroll
			for (int k = 0; k < 3; ++k) {
				lo.x = fminf(lo.x, p[k * 3]); lo.y = fminf(lo.y, p[k * 3 + 1]); lo.z = fminf(lo.z, p[k * 3 + 2]);
				hi.x = fmaxf(hi.x, p[k * 3]); hi.y = fmaxf(hi.y, p[k * 3 + 1]); hi.z = fmaxf(hi.z, p[k * 3 + 2]);
			}
		}
	}
	volatile float* o = nodes + (size_t)s * 8;
	o[0] = lo.x; o[1] = lo.y; o[2] = lo.z; o[3] = hi.x; o[4] = hi.y; o[5] = hi.z;
	int cur = parent[s];
	while (cur >= 0) {
		__threadfence();
		if (atomicAdd(flags + cur, 1u) == 0u) return;      // first child: the sibling will finish this node
		__threadfence();
		const int cl = reinterpret_cast<const int*>(nodes + (size_t)cur * 8)[6];
		const volatile float* a = nodes + (size_t)cl * 8;
		const volatile float* b = a + 8;
		volatile float* w = nodes + (size_t)cur * 8;
		w[0] = fminf(a[0], b[0]); w[1] = fminf(a[1], b[1]); w[2] = fminf(a[2], b[2]);
		w[3] = fmaxf(a[3], b[3]); w[4] = fmaxf(a[4], b[4]); w[5] = fmaxf(a[5], b[5]);
		cur = parent[cur];
	}
}

// depth of the deepest leaf (validation: the traversal stack holds 32 entries)
__global__ void __launch_bounds__(256) k_bvh_max_depth(const float* __restrict__ nodes, int n_nodes, const int* __restrict__ parent, int* __restrict__ out) {
	const int s = blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= n_nodes) return;
	if (reinterpret_cast<const int*>(nodes + (size_t)s * 8)[6] >= 0) return;
	int d = 0;
	for (int cur = parent[s]; cur >= 0; cur = parent[cur]) ++d;
	atomicMax(out, d);
}

inline int lbvh_node_count(int n_tris, int ppl) { const int L = (n_tris + ppl - 1) / ppl; return 2 * L - 1; }

// Refit: nodes_dev (any tree in the node format) gets the bounds of the triangles as they are now. 2 launches (3 with max_depth_dev).
inline cudaError_t bvh_refit_device(LbvhScratch& sc, cudaStream_t st, const float* tris_dev, float* nodes_dev, int n_nodes, int* max_depth_dev = nullptr) {
	cudaError_t e;
	if ((e = sc.parent.reserve((size_t)n_nodes * 4)) != cudaSuccess) return e;
	if ((e = sc.flags.reserve((size_t)n_nodes * 4)) != cudaSuccess) return e;
	const int blocks = (n_nodes + 255) / 256;
	k_bvh_parents<<<blocks, 256, 0, st>>>(nodes_dev, n_nodes, sc.parent.as<int>(), sc.flags.as<unsigned int>());
	k_bvh_fit<<<blocks, 256, 0, st>>>(nodes_dev, n_nodes, tris_dev, sc.parent.as<int>(), sc.flags.as<unsigned int>());
	if (max_depth_dev) k_bvh_max_depth<<<blocks, 256, 0, st>>>(nodes_dev, n_nodes, sc.parent.as<int>(), max_depth_dev);
	return cudaGetLastError();
}

// Build: tris_dev (n x 9, any order) -> tris_out_dev (n x 9 in Morton order; may not alias tris_dev), nodes_dev (lbvh_node_count x 8),
// order_out_dev (optional: source index of every output triangle). Asynchronous on `st`; *launches = kernels launched (the sort counted as one).
inline cudaError_t lbvh_build_device(LbvhScratch& sc, cudaStream_t st, const float* tris_dev, int n, int ppl, float* tris_out_dev, float* nodes_dev, int* order_out_dev,
                                     int* max_depth_dev, int* launches) {
	cudaError_t e;
	const int L = (n + ppl - 1) / ppl, n_nodes = 2 * L - 1;
	if ((e = sc.keys.reserve((size_t)n * 8)) != cudaSuccess) return e;
	if ((e = sc.keys_sorted.reserve((size_t)n * 8)) != cudaSuccess) return e;
	if ((e = sc.bounds.reserve(32)) != cudaSuccess) return e;
	if ((e = sc.islot.reserve((size_t)std::max(1, L) * 4)) != cudaSuccess) return e;
	size_t tmp_bytes = 0;
	cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, sc.keys.as<unsigned long long>(), sc.keys_sorted.as<unsigned long long>(), n, 0, 64, st);
	if ((e = sc.sort_tmp.reserve(std::max<size_t>(tmp_bytes, 16))) != cudaSuccess) return e;
	const int blocks = (n + 255) / 256;
	k_lbvh_init_bounds<<<1, 32, 0, st>>>(sc.bounds.as<unsigned int>());
	k_lbvh_bounds<<<blocks, 256, 0, st>>>(tris_dev, n, sc.bounds.as<unsigned int>());
	k_lbvh_keys<<<blocks, 256, 0, st>>>(tris_dev, n, sc.bounds.as<unsigned int>(), sc.keys.as<unsigned long long>());
	if ((e = cub::DeviceRadixSort::SortKeys(sc.sort_tmp.p, tmp_bytes, sc.keys.as<unsigned long long>(), sc.keys_sorted.as<unsigned long long>(), n, 0, 64, st)) != cudaSuccess) return e;
	k_lbvh_gather<<<(int)(((long long)n * 9 + 255) / 256), 256, 0, st>>>(tris_dev, sc.keys_sorted.as<unsigned long long>(), n, tris_out_dev, order_out_dev);
	if ((e = cudaMemsetAsync(nodes_dev, 0, (size_t)n_nodes * 32, st)) != cudaSuccess) return e;
	int count = 5;
	if (L == 1) { k_lbvh_single_leaf<<<1, 1, 0, st>>>(n, nodes_dev); ++count; }
	else {
		const int ib = (L - 1 + 255) / 256;
		k_lbvh_topology<<<ib, 256, 0, st>>>(sc.keys_sorted.as<unsigned long long>(), n, ppl, L, nodes_dev, sc.islot.as<int>());
		k_lbvh_link<<<ib, 256, 0, st>>>(L, sc.islot.as<int>(), nodes_dev);
		count += 2;
	}
	if ((e = bvh_refit_device(sc, st, tris_out_dev, nodes_dev, n_nodes, max_depth_dev)) != cudaSuccess) return e;
	count += max_depth_dev ? 3 : 2;
	if (launches) *launches = count;
	return cudaGetLastError();
}

} // namespace snb
