// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
// NerfNetwork inference: hash-grid encoding -> density MLP -> SH -> rgb MLP.
// Device-only code in the reference (tcnn). Pinned since round 2 against the reference's own kernel_grid / kernel_mlp_fused run on the GPU
// (oracle/_ref/libref_dev.so, tests/test_gpu_refdev.py): features bit for bit, network outputs within fp16-accumulation tolerance.
#include "network.h"

namespace orc {

static float g_scale_override[32];
static uint32_t g_scale_override_n = 0;

// tcnn/common_device.h:841-850 grid_scale / grid_resolution; grid.h:692-723 offset table.
GridLayout make_grid_layout(uint32_t n_levels, uint32_t F, uint32_t log2_hashmap_size, uint32_t base_resolution, float per_level_scale) {
	GridLayout g;
	g.n_levels = n_levels; g.F = F;
	float log2_pls = std::log2(per_level_scale);
	uint32_t offset = 0;
	for (uint32_t i = 0; i < n_levels; ++i) {
		float scale = std::exp2(i * log2_pls) * base_resolution - 1.0f;
		uint32_t resolution = (uint32_t)std::ceil(scale) + 1;
		uint32_t max_params = std::numeric_limits<uint32_t>::max() / 2;
		uint32_t params_in_level = std::pow((float)resolution, 3) > (float)max_params ? max_params : resolution * resolution * resolution;
		params_in_level = (params_in_level + 7u) / 8u * 8u;
		params_in_level = std::min(params_in_level, 1u << log2_hashmap_size);
		g.scale[i] = scale; g.resolution[i] = resolution; g.offset[i] = offset;
		offset += params_in_level;
	}
	g.offset[n_levels] = offset;
	return g;
}

// tcnn/common_device.h:631-642,651-655 (coherent_prime_hash) and :697-713 (grid_index, GridType::Hash).
uint32_t grid_index(uint32_t hashmap_size, uint32_t resolution, uint32_t x, uint32_t y, uint32_t z) {
	uint32_t pos[3] = {x, y, z};
	uint32_t stride = 1, index = 0;
	for (uint32_t dim = 0; dim < 3 && stride <= hashmap_size; ++dim) {
		index += pos[dim] * stride;
		stride *= resolution;
	}
	if (hashmap_size < stride) index = (x * 1u) ^ (y * 2654435761u) ^ (z * 805459861u);
	return index % hashmap_size;
}

// grid.h:48-169 kernel_grid forward, Linear interpolation, accumulation in half with half FMA (grid.h:162).
void grid_encode(const GridLayout& g, const half_t* grid, vec3 x, half_t* out /* n_levels*F */, uint32_t* out_idx /* n_levels*8 or null */) {
	for (uint32_t level = 0; level < g.n_levels; ++level) {
		const half_t* lg = grid + (size_t)g.offset[level] * g.F;
		uint32_t hashmap_size = g.offset[level + 1] - g.offset[level];
		float scale = g.scale[level];
		uint32_t resolution = g.resolution[level];
		float pos[3]; uint32_t pos_grid[3];
		for (int d = 0; d < 3; ++d) { // pos_fract, common_device.h:853-857 (identity interpolation)
			float p = std::fmaf(scale, x[d], 0.5f);
			float tmp = std::floor(p);
			pos_grid[d] = (uint32_t)(int)tmp;
			pos[d] = p - tmp;
		}
		half_t result[8] = {};
		for (uint32_t idx = 0; idx < 8; ++idx) {
			float weight = 1;
			uint32_t pl[3];
			for (uint32_t d = 0; d < 3; ++d) {
				if ((idx & (1u << d)) == 0) { weight *= 1 - pos[d]; pl[d] = pos_grid[d]; }
				else { weight *= pos[d]; pl[d] = pos_grid[d] + 1; }
			}
			uint32_t index = grid_index(hashmap_size, resolution, pl[0], pl[1], pl[2]);
			if (out_idx) out_idx[level * 8 + idx] = index;
			half_t w = f2h(weight);
			for (uint32_t f = 0; f < g.F; ++f) result[f] = hfma(w, lg[(size_t)index * g.F + f], result[f]);
		}
		for (uint32_t f = 0; f < g.F; ++f) out[level * g.F + f] = result[f];
	}
}

// tcnn/common_device.h:340-365 sh_enc degree 4 fed with dir*2-1 (spherical_harmonics.h:66-70).
void sh4_encode(vec3 dir_warped, half_t* out16) {
	float x = dir_warped.x * 2.f - 1.f, y = dir_warped.y * 2.f - 1.f, z = dir_warped.z * 2.f - 1.f;
	float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
	out16[0] = f2h(0.28209479177387814f);
	out16[1] = f2h(-0.48860251190291987f * y);
	out16[2] = f2h(0.48860251190291987f * z);
	out16[3] = f2h(-0.48860251190291987f * x);
	out16[4] = f2h(1.0925484305920792f * xy);
	out16[5] = f2h(-1.0925484305920792f * yz);
	out16[6] = f2h(0.94617469575755997f * z2 - 0.31539156525251999f);
	out16[7] = f2h(-1.0925484305920792f * xz);
	out16[8] = f2h(0.54627421529603959f * x2 - 0.54627421529603959f * y2);
	out16[9] = f2h(0.59004358992664352f * y * (-3.0f * x2 + y2));
	out16[10] = f2h(2.8906114426405538f * xy * z);
	out16[11] = f2h(0.45704579946446572f * y * (1.0f - 5.0f * z2));
	out16[12] = f2h(0.3731763325901154f * z * (5.0f * z2 - 3.0f));
	out16[13] = f2h(0.45704579946446572f * x * (1.0f - 5.0f * z2));
	out16[14] = f2h(1.4453057213202769f * z * (x2 - y2));
	out16[15] = f2h(0.59004358992664352f * x * (-x2 + 3.0f * y2));
}

// One dense layer y = W x, W row-major [n_out x n_in] (fully_fused_mlp.cu:659-671), fp16 operands,
// fp32 accumulation, output rounded to fp16. (The reference accumulates in fp16 wmma fragments —
// hardware-specific rounding that no CPU can restate; BASELINE.json tolerates 1e-2 on the framebuffer.)
static void dense(const half_t* W, const half_t* x, uint32_t n_in, uint32_t n_out, bool relu, half_t* y) {
	for (uint32_t o = 0; o < n_out; ++o) {
		float acc = 0.0f;
		for (uint32_t i = 0; i < n_in; ++i) acc += h2f(W[(size_t)o * n_in + i]) * h2f(x[i]);
		if (relu && acc < 0.0f) acc = 0.0f;
		y[o] = f2h(acc);
	}
}

// nerf_network.h:105-139 inference_mixed_precision_impl; returns rgb (network outputs 0..2) and log-density.
void network_query(const NetworkView& net, vec3 pos_warped, vec3 dir_warped, half_t out4[4], half_t* feat32_out) {
	half_t feat[32], h1[64], h2[64], rgb_in[32], out16[16];
	grid_encode(net.grid_layout, net.grid, pos_warped, feat, nullptr);
	if (feat32_out) std::memcpy(feat32_out, feat, sizeof(feat));
	dense(net.density_w1, feat, 32, 64, true, h1);
	dense(net.density_w2, h1, 64, 16, false, rgb_in);      // density output = rows 0..15 of the rgb input
	sh4_encode(dir_warped, rgb_in + 16);
	dense(net.rgb_w1, rgb_in, 32, 64, true, h1);
	dense(net.rgb_w2, h1, 64, 64, true, h2);
	dense(net.rgb_w3, h2, 64, 16, false, out16);
	out4[0] = out16[0]; out4[1] = out16[1]; out4[2] = out16[2];
	out4[3] = rgb_in[0];                                   // extract_density, nerf_network.h:31-43,132-138
}

NetworkView make_network_view(const orc_model& m) {
	NetworkView v;
	const half_t* p = (const half_t*)m.params;
	v.density_w1 = p; p += 64 * 32;
	v.density_w2 = p; p += 16 * 64;
	v.rgb_w1 = p; p += 64 * 32;
	v.rgb_w2 = p; p += 64 * 64;
	v.rgb_w3 = p; p += 16 * 64;
	v.grid = p;
	v.grid_layout = make_grid_layout(m.n_levels, m.n_features_per_level, m.log2_hashmap_size, m.base_resolution, m.per_level_scale);
	// kernel_grid evaluates grid_scale ON THE DEVICE: exp2f there is MUFU.EX2 and the multiply-subtract is one FFMA (SASS of the
	// reference build, profiles/r02_refdev_notes.md), which differs from the host's libm value in the last place for most levels. The
	// hardware function has no host restatement, so GPU parity tests hand its values in (orc_set_grid_scales); the offset table stays the
	// host's, as in tcnn (grid.h:692-723 runs on the host).
	if (g_scale_override_n == m.n_levels)
		for (uint32_t i = 0; i < m.n_levels; ++i) { v.grid_layout.scale[i] = g_scale_override[i]; v.grid_layout.resolution[i] = (uint32_t)std::ceil(g_scale_override[i]) + 1; }
	return v;
}

} // namespace orc

using namespace orc;

extern "C" {

// Size (in fp16 elements) of the params blob for a model description.
// Per-level scales as the DEVICE computes them (n = 0 returns to the host formula). See make_network_view.
void orc_set_grid_scales(const float* scales, uint32_t n) {
	g_scale_override_n = n <= 32 ? n : 0;
	for (uint32_t i = 0; i < g_scale_override_n; ++i) g_scale_override[i] = scales[i];
}

uint64_t orc_n_params(uint32_t n_levels, uint32_t F, uint32_t log2_hashmap_size, uint32_t base_resolution, float per_level_scale) {
	GridLayout g = make_grid_layout(n_levels, F, log2_hashmap_size, base_resolution, per_level_scale);
	return (uint64_t)(64 * 32 + 16 * 64 + 64 * 32 + 64 * 64 + 16 * 64) + (uint64_t)g.offset[n_levels] * F;
}

void orc_grid_layout(uint32_t n_levels, uint32_t F, uint32_t log2_hashmap_size, uint32_t base_resolution, float per_level_scale,
                     float* scale, uint32_t* resolution, uint32_t* offset) {
	GridLayout g = make_grid_layout(n_levels, F, log2_hashmap_size, base_resolution, per_level_scale);
	for (uint32_t i = 0; i < n_levels; ++i) { scale[i] = g.scale[i]; resolution[i] = g.resolution[i]; offset[i] = g.offset[i]; }
	offset[n_levels] = g.offset[n_levels];
}

// coords: n x 7 floats (NerfCoordinate: pos.xyz, dt, dir.xyz). out: n x 4 halves (r,g,b,sigma raw network outputs).
// Optional taps: features n x 32 halves, hash indices n x (n_levels*8) u32.
void orc_network_inference(const orc_model* model, const float* coords, uint64_t n, uint16_t* out, uint16_t* features, uint32_t* hash_idx) {
	NetworkView net = make_network_view(*model);
	for (uint64_t i = 0; i < n; ++i) {
		const float* c = coords + i * 7;
		half_t o4[4]; half_t feat[32];
		network_query(net, {c[0], c[1], c[2]}, {c[4], c[5], c[6]}, o4, feat);
		std::memcpy(out + i * 4, o4, 8);
		if (features) std::memcpy(features + i * 32, feat, 64);
		if (hash_idx) { half_t tmp[32]; grid_encode(net.grid_layout, net.grid, {c[0], c[1], c[2]}, tmp, hash_idx + i * net.grid_layout.n_levels * 8); }
	}
}

}
