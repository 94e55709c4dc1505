// synerfgine_b200 — occupancy-cascade DDA march, start jitter and pinhole ray generation.
//
// Replaces (reference, include/neural-graphics-primitives/nerf_device.cuh): cascaded_grid_idx_at :317-329,
// density_grid_occupied_at :331-341, distance_to_next_voxel :360-368, to/from_stepping_space :379-421,
// advance_n_steps / calc_dt :423-429, advance_to_next_voxel :431-441, mip_from_pos :443-448,
// if_unoccupied_advance_to_next_occupied_voxel :464-495; random_val.cuh:161-325 (Sobol + Owen scramble);
// common_device.cuh:361-368,403-470 (pinhole uv_to_ray, rolling-shutter transform);
// src/testbed_nerf.cu:334-363 (advance_pos_nerf), :1855-1970 (init_rays_with_payload_kernel_nerf).
//
// B200 notes: the stepping-space constants (log1p(c), a, b, exp(a·l), exp(b·l)) depend only on the cone
// angle, so they are evaluated once per frame on the host with the same det_* formulas and carried in
// MarchEnv instead of being re-derived at every probe. The 2 MiB bitfield stays L2/L1 resident; probes
// use the read-only path.
#pragma once
#include "snb_math.cuh"

#ifndef SNB_GATHER_HINT
#define SNB_GATHER_HINT 0   // L1 eviction-priority hints on the gathers (tuning hook)
#endif

namespace snb {

constexpr uint32_t kGridSize = 128;
constexpr uint32_t kGridCells = 128u * 128u * 128u;
constexpr uint32_t kCascades = 8;
constexpr float kSqrt3 = 1.73205080757f;
constexpr float kMinStep = kSqrt3 / 1024;                       // MIN_CONE_STEPSIZE
constexpr float kMaxStep = kMinStep * (1 << (kCascades - 1)) * 1024 / kGridSize; // MAX_CONE_STEPSIZE
constexpr uint32_t kMarchIter = 10000;                          // MARCH_ITER, src/testbed_nerf.cu:47

struct Stepping {      // constants of the exponential stepping function for one cone angle
	float cone;        // cone_angle_constant
	float l;           // log(1 + cone)
	float a, b;        // stepping-space bounds of the exponential segment
	float at, bt;      // exp(a*l), exp(b*l)
};

SNB_HD Stepping make_stepping(float cone) {
	Stepping s; s.cone = cone; s.l = s.a = s.b = s.at = s.bt = 0.0f;
	if (cone <= 1e-5f) return s;
	s.l = det_logf(1.0f + cone);
	s.a = (det_logf(kMinStep) - det_logf(s.l)) / s.l;
	s.b = (det_logf(kMaxStep) - det_logf(s.l)) / s.l;
	s.at = det_expf(s.a * s.l);
	s.bt = det_expf(s.b * s.l);
	return s;
}

// Coarse look-ahead grid: kCoarseRes^3 bits over the extent of cascade max_mip, a dilated superset of every occupied
// cell of cascades 0..max_mip (built once per model by k_build_coarse). It never changes a march result: it only proves
// "no occupied cell ahead", in which case the DDA would have walked to the edge of the render box and returned kMaxDepth.
constexpr int kCoarseRes = 32;
constexpr int kCoarseWords = kCoarseRes * kCoarseRes * kCoarseRes / 32;   // 1024 words = 4 KB (staged in shared memory)

struct MarchEnv {
	const uint8_t* bitfield;   // kCascades * 128^3 / 8 bytes, Morton order per cascade (the reference's layout); nullptr = all occupied
	const uint8_t* bits_xyz;   // device-side copy of the same bits in x-fastest linear order per cascade (k_bitfield_to_linear):
	                           // the DDA probes cells ~2.5 times per step and a Morton index costs ~28 instructions, a linear one 3
	uint32_t min_mip, max_mip;
	Aabb render_aabb;
	m3 to_local;               // render_aabb_to_local
	uint32_t local_identity;   // to_local is exactly the identity (set by snb_set_model): in_render_box skips the 3x3 product
	Stepping st;
	const uint32_t* coarse;    // kCoarseWords words or nullptr (no look-ahead)
	v3 coarse_lo;              // world position of coarse cell (0,0,0)
	float coarse_inv_cell;     // coarse cells per world unit
};

SNB_HD uint32_t spread3(uint32_t v) { // tcnn/common_device.h:761-770
	v = (v * 0x00010001u) & 0xFF0000FFu;
	v = (v * 0x00000101u) & 0x0F00F00Fu;
	v = (v * 0x00000011u) & 0xC30C30C3u;
	v = (v * 0x00000005u) & 0x49249249u;
	return v;
}
SNB_HD uint32_t morton3(uint32_t x, uint32_t y, uint32_t z) { return spread3(x) | (spread3(y) << 1) | (spread3(z) << 2); }

// 2^e for small |e| as an exponent-field constant: x * pow2i(e) == scalbnf(x, e) whenever neither under- nor overflows,
// which holds for every use below (|e| <= 8, operands of order 1).
SNB_HD float pow2i(int e) { return u2f((uint32_t)(127 + e) << 23); }

SNB_HD uint32_t cascade_cell(v3 pos, uint32_t mip) {
	const float mip_scale = pow2i(-(int)mip);
	pos = pos - 0.5f;
	pos = pos * mip_scale;
	pos = pos + 0.5f;
	const int ix = (int)(pos.x * (float)kGridSize), iy = (int)(pos.y * (float)kGridSize), iz = (int)(pos.z * (float)kGridSize);
	if (ix < 0 || ix >= (int)kGridSize || iy < 0 || iy >= (int)kGridSize || iz < 0 || iz >= (int)kGridSize) return 0xFFFFFFFFu;
	return morton3((uint32_t)ix, (uint32_t)iy, (uint32_t)iz);
}

#ifdef __CUDACC__
// Same predicate as cell_occupied below, read from the linear-order copy. The cell coordinates are computed exactly as in
// cascade_cell (cascaded_grid_idx_at); only the bit's address differs.
SNB_D bool cell_occupied_xyz(v3 pos, const uint8_t* __restrict__ bits_xyz, uint32_t mip) {
	const float mip_scale = pow2i(-(int)mip);
	pos = pos - 0.5f;
	pos = pos * mip_scale;
	pos = pos + 0.5f;
	const int ix = (int)(pos.x * (float)kGridSize), iy = (int)(pos.y * (float)kGridSize), iz = (int)(pos.z * (float)kGridSize);
	if ((uint32_t)(ix | iy | iz) >= kGridSize) return false;      // any coordinate negative or >= 128
	const uint32_t idx = (uint32_t)ix | ((uint32_t)iy << 7) | ((uint32_t)iz << 14) | (mip << 21);
	return (__ldg(bits_xyz + (idx >> 3)) >> (idx & 7u)) & 1u;
}
#endif

SNB_HD bool cell_occupied(v3 pos, const uint8_t* bitfield, uint32_t mip) {
	const uint32_t idx = cascade_cell(pos, mip);
	if (idx == 0xFFFFFFFFu) return false;
#ifdef __CUDA_ARCH__
#if SNB_GATHER_HINT
	uint32_t byte;   // keep the bitfield in L1 against the hash-grid gathers streaming through it
	asm("ld.global.nc.L1::evict_last.u8 %0, [%1];" : "=r"(byte) : "l"(bitfield + (idx / 8 + (kGridCells * mip) / 8)));
#else
	const uint8_t byte = __ldg(bitfield + (idx / 8 + (kGridCells * mip) / 8));
#endif
#else
	const uint8_t byte = bitfield[idx / 8 + (kGridCells * mip) / 8];
#endif
	return byte & (1u << (idx % 8));
}

SNB_HD float voxel_exit_distance(v3 pos, v3 dir, v3 idir, float res, float inv_res) {
	const v3 p = res * (pos - 0.5f);
	const float tx = (floorf(p.x + 0.5f + 0.5f * copysignf(1.0f, dir.x)) - p.x) * idir.x;
	const float ty = (floorf(p.y + 0.5f + 0.5f * copysignf(1.0f, dir.y)) - p.y) * idir.y;
	const float tz = (floorf(p.z + 0.5f + 0.5f * copysignf(1.0f, dir.z)) - p.z) * idir.z;
	const float t = fminf(fminf(tx, ty), tz);
	return fmaxf(t * inv_res, 0.0f);   // res is a power of two: t / res == t * (1 / res) exactly
}

SNB_HD float to_step_space(float t, const Stepping& s) {
	if (s.cone <= 1e-5f) return t / kMinStep;
	if (t <= s.at) return (t - s.at) / kMinStep + s.a;
	else if (t <= s.bt) return det_logf(t) / s.l;
	else return (t - s.bt) / kMaxStep + s.b;
}
SNB_HD float from_step_space(float n, const Stepping& s) {
	if (s.cone <= 1e-5f) return n * kMinStep;
	if (n <= s.a) return (n - s.a) * kMinStep + s.at;
	else if (n <= s.b) return det_expf(n * s.l);
	else return (n - s.b) * kMaxStep + s.bt;
}
SNB_HD float advance_steps(float t, const Stepping& s, float n) { return from_step_space(to_step_space(t, s) + n, s); }
SNB_HD float step_size(float t, const Stepping& s) { return advance_steps(t, s, 1.0f) - t; }

SNB_HD float advance_past_voxel(float t, const Stepping& s, v3 pos, v3 dir, v3 idir, uint32_t mip) {
	const float res = (float)kGridSize * pow2i(-(int)mip), inv_res = pow2i((int)mip) * (1.0f / (float)kGridSize);
	float t_target = t + voxel_exit_distance(pos, dir, idir, res, inv_res);
	t = to_step_space(t, s);
	t_target = to_step_space(t_target, s);
	return from_step_space(t + ceilf(fmaxf(t_target - t, 0.5f)), s);
}

#ifdef __CUDACC__
// Out-of-line twins for the persistent march kernel: its loop inlined to_step_space three times and from_step_space
// twice (~500 SASS instructions) and no longer fitted the instruction cache (profiles/r01_ncu_fused_v3.md). Same
// arithmetic, one copy of the code.
__device__ __noinline__ float to_step_space_call(float t, Stepping s) { return to_step_space(t, s); }
__device__ __noinline__ float from_step_space_call(float n, Stepping s) { return from_step_space(n, s); }
__device__ __forceinline__ float step_size_call(float t, const Stepping& s) { return from_step_space_call(to_step_space_call(t, s) + 1.0f, s) - t; }
__device__ __forceinline__ float advance_past_voxel_call(float t, const Stepping& s, v3 pos, v3 dir, v3 idir, uint32_t mip) {
	const float res = (float)kGridSize * pow2i(-(int)mip), inv_res = pow2i((int)mip) * (1.0f / (float)kGridSize);
	float t_target = t + voxel_exit_distance(pos, dir, idir, res, inv_res);
	t = to_step_space_call(t, s);
	t_target = to_step_space_call(t_target, s);
	return from_step_space_call(t + ceilf(fmaxf(t_target - t, 0.5f)), s);
}
#endif

SNB_HD uint32_t mip_of(v3 pos, uint32_t max_cascade = kCascades - 1) {
	// frexpf exponent from the exponent field; zero / subnormal inputs give a negative m either way and clamp to 0
	const float maxval = vmax(vabs(pos - 0.5f));
	const int exponent = (int)((f2u(maxval) >> 23) & 0xFFu) - 126;
	int m = exponent + 1;
	if (m < 0) m = 0;
	if (m > (int)max_cascade) m = (int)max_cascade;
	return (uint32_t)m;
}
SNB_HD uint32_t clamp_mip(uint32_t mip, const MarchEnv& env) {
	if (mip < env.min_mip) mip = env.min_mip;
	if (mip > env.max_mip) mip = env.max_mip;
	return mip;
}

// Conservative "is anything occupied ahead of pos along d?" — a 3-D DDA through the coarse grid. Returns true when
// every coarse cell the ray touches from pos to the edge of the coarse domain is empty. Positions outside the domain
// (or a missing grid) answer false, i.e. "cannot tell".
SNB_HD bool coarse_clear_ahead(v3 pos, v3 d, const MarchEnv& env) {
	if (!env.coarse) return false;
	const float R = (float)kCoarseRes;
	const v3 g = (pos - env.coarse_lo) * env.coarse_inv_cell;
	if (!(g.x >= 0.0f && g.x < R && g.y >= 0.0f && g.y < R && g.z >= 0.0f && g.z < R)) return false;
	int ix = (int)g.x, iy = (int)g.y, iz = (int)g.z;
	const int sx = d.x > 0.0f ? 1 : -1, sy = d.y > 0.0f ? 1 : -1, sz = d.z > 0.0f ? 1 : -1;
	const float big = 3.0e38f;
	const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
	const float dx = ax > 1e-12f ? 1.0f / ax : big, dy = ay > 1e-12f ? 1.0f / ay : big, dz = az > 1e-12f ? 1.0f / az : big;   // grid units of t per cell
	float tx = ax > 1e-12f ? (sx > 0 ? ((float)(ix + 1) - g.x) : (g.x - (float)ix)) * dx : big;
	float ty = ay > 1e-12f ? (sy > 0 ? ((float)(iy + 1) - g.y) : (g.y - (float)iy)) * dy : big;
	float tz = az > 1e-12f ? (sz > 0 ? ((float)(iz + 1) - g.z) : (g.z - (float)iz)) * dz : big;
	for (int guard = 0; guard < 3 * kCoarseRes + 3; ++guard) {
		const uint32_t bit = (uint32_t)ix + (uint32_t)kCoarseRes * ((uint32_t)iy + (uint32_t)kCoarseRes * (uint32_t)iz);
		if (env.coarse[bit >> 5] & (1u << (bit & 31))) return false;
		if (tx <= ty && tx <= tz) { ix += sx; tx += dx; if ((unsigned)ix >= (unsigned)kCoarseRes) return true; }
		else if (ty <= tz) { iy += sy; ty += dy; if ((unsigned)iy >= (unsigned)kCoarseRes) return true; }
		else { iz += sz; tz += dz; if ((unsigned)iz >= (unsigned)kCoarseRes) return true; }
	}
	return false;
}

#ifdef __CUDACC__
// Out-of-line twin of coarse_clear_ahead for the persistent march kernel (instruction-cache budget): the walk runs once per
// kLookAheadEvery empty cells per ray, its ~150 instructions need not sit inside the hot loop. `coarse` may point to shared memory.
__device__ __noinline__ bool coarse_clear_ahead_call(float px, float py, float pz, float dx, float dy, float dz, const uint32_t* coarse, float lox, float loy, float loz,
                                                     float inv_cell) {
	MarchEnv e;
	e.coarse = coarse; e.coarse_lo = mk3(lox, loy, loz); e.coarse_inv_cell = inv_cell;
	return coarse_clear_ahead(mk3(px, py, pz), mk3(dx, dy, dz), e);
}
#endif

// m_render_aabb.contains(render_aabb_to_local * pos). With the identity the product is pos itself bit for bit:
// (1*x + 0*y) + 0*z == x for finite coordinates (a -0 may become +0, which no comparison sees), so it is skipped.
SNB_HD bool in_render_box(v3 pos, const MarchEnv& env) {
	return env.render_aabb.contains(env.local_identity ? pos : mul(env.to_local, pos));
}

SNB_HD bool env_cell_occupied(v3 pos, const MarchEnv& env, uint32_t mip) {
#ifdef __CUDA_ARCH__
	return cell_occupied_xyz(pos, env.bits_xyz, mip);
#else
	return cell_occupied(pos, env.bitfield, mip);
#endif
}

// Skip empty space: returns the first t >= t_in whose enclosing cascade cell is occupied, or kMaxDepth
// once the ray leaves the render box. Every kLookAheadEvery-th empty cell the coarse grid is consulted: if nothing
// can be occupied ahead, the remaining DDA steps (which could only end in kMaxDepth) are not walked.
#ifndef SNB_LOOKAHEAD_EVERY
#define SNB_LOOKAHEAD_EVERY 8   // empty cells between two look-ahead walks (power of two; tuning hook)
#endif
constexpr int kLookAheadEvery = SNB_LOOKAHEAD_EVERY;
SNB_HD float skip_to_occupied(float t, v3 o, v3 d, v3 idir, const MarchEnv& env) {
	int n_empty = 0;
	while (true) {
		const v3 pos = o + d * t;
		if (t >= kMaxDepth || !in_render_box(pos, env)) return kMaxDepth;
		uint32_t mip = clamp_mip(mip_of(pos), env);
		if (!env.bitfield || env_cell_occupied(pos, env, mip)) return t;
		if ((n_empty++ % kLookAheadEvery) == 0 && coarse_clear_ahead(pos, d, env)) return kMaxDepth;
		while (mip < env.max_mip && !env_cell_occupied(pos, env, mip + 1)) ++mip;
		t = advance_past_voxel(t, env.st, pos, d, idir, mip);
	}
}

// ---- low-discrepancy sampler (Burley 2020 Owen-scrambled Sobol, reference random_val.cuh:161-325) ----
SNB_HD uint32_t sobol_dim(uint32_t index, uint32_t dim) {
	// dim 0: van der Corput (v_i = 2^(31-i)) == bit reversal; dim 1: v_i = v_{i-1} ^ (v_{i-1} >> 1).
	uint32_t X = 0, v = 0x80000000u;
	for (uint32_t bit = 0; bit < 32; ++bit) {
		if ((index >> bit) & 1u) X ^= v;
		v = dim == 0 ? (v >> 1) : (v ^ (v >> 1));
	}
	return X;
}
SNB_HD uint32_t bitrev32(uint32_t x) {
#ifdef __CUDA_ARCH__
	return __brev(x);
#else
	x = ((x & 0xaaaaaaaau) >> 1) | ((x & 0x55555555u) << 1);
	x = ((x & 0xccccccccu) >> 2) | ((x & 0x33333333u) << 2);
	x = ((x & 0xf0f0f0f0u) >> 4) | ((x & 0x0f0f0f0fu) << 4);
	x = ((x & 0xff00ff00u) >> 8) | ((x & 0x00ff00ffu) << 8);
	return (x >> 16) | (x << 16);
#endif
}
SNB_HD uint32_t lk_permute(uint32_t x, uint32_t seed) {
	x += seed;
	x ^= x * 0x6c50b47cu;
	x ^= x * 0xb82f1e52u;
	x ^= x * 0xc7afe638u;
	x ^= x * 0x8d22f6e6u;
	return x;
}
SNB_HD uint32_t owen_scramble(uint32_t x, uint32_t seed) { return bitrev32(lk_permute(bitrev32(x), seed)); }
SNB_HD uint32_t seed_mix(uint32_t seed, uint32_t v) { return seed ^ (v + (seed << 6) + (seed >> 2)); }
SNB_HD float ld_value(uint32_t index, uint32_t seed, uint32_t dim = 0) {
	const float S = 2.3283064365386963e-10f; // 2^-32
	index = owen_scramble(index, seed);
	return (float)owen_scramble(sobol_dim(index, dim), seed_mix(seed, dim)) * S;
}
SNB_HD v2 ld_pixel_offset(uint32_t spp) {
	const uint32_t seed = 0xdeadbeefu;
	v2 a, b, r;
	a.x = ld_value(0, seed, 0); a.y = ld_value(0, seed, 1);
	b.x = ld_value(spp, seed, 0); b.y = ld_value(spp, seed, 1);
	const float ox = 0.5f - a.x + b.x, oy = 0.5f - a.y + b.y;
	r.x = ox - floorf(ox); r.y = oy - floorf(oy);
	return r;
}

} // namespace snb
