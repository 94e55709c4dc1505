// synerfgine_b200 — NerfNetwork inference building blocks: multiresolution hash-grid encoding, SH-4 direction
// encoding and the two width-64 MLPs, evaluated for the 32 samples a warp owns (one sample per lane).
//
// Replaces (reference): tcnn kernel_grid forward (dependencies/tiny-cuda-nn/include/tiny-cuda-nn/encodings/grid.h:48-169),
// grid_index / coherent_prime_hash / pos_fract (tiny-cuda-nn/common_device.h:631-718,841-868), kernel_sh degree 4
// (encodings/spherical_harmonics.h:45-72, common_device.h:340-360), kernel_mlp_fused (src/fully_fused_mlp.cu:499-557)
// and the wiring of NerfNetwork::inference_mixed_precision_impl + extract_density
// (include/neural-graphics-primitives/nerf_network.h:31-43,105-139).
//
// This file is the warp-level (mma.sync m16n8k16, fp32 accumulate) evaluator used by the fused march kernel and
// by the parity taps. The CTA-level tcgen05/TMEM evaluator for the same weights lives in snb_mlp_tc.cuh.
//
// Numerics: features accumulate in fp16 with single-rounding half FMAs in corner order 0..7 (grid.h:144,162);
// every MLP layer output is rounded to fp16 (ReLU applied before rounding on hidden layers); accumulation is fp32.
#pragma once
#include "snb_device.cuh"

#ifndef SNB_ENC_GROUP
#define SNB_ENC_GROUP 2   // hash-grid levels gathered together per loop trip (ILP vs code size; tuning hook)
#endif

namespace snb {

// ---- weight fragment order for mma.m16n8k16 (B operand = W[n][k], "col" layout) --------------------------------
// For a layer with N outputs and K inputs: word index ((nt * (K/16) + kt) * 32 + lane) * 2 + {0,1} holds
//   {W[nt*8+g][kt*16+2t], W[nt*8+g][kt*16+2t+1]} and {W[nt*8+g][kt*16+8+2t], W[..][kt*16+8+2t+1]},  g = lane>>2, t = lane&3.
constexpr int kFragDW1 = 0;                       // 64x32 -> 1024 words
constexpr int kFragDW2 = kFragDW1 + 64 * 32 / 2;  // 16x64 ->  512
constexpr int kFragRW1 = kFragDW2 + 16 * 64 / 2;  // 64x32 -> 1024
constexpr int kFragRW2 = kFragRW1 + 64 * 32 / 2;  // 64x64 -> 2048
constexpr int kFragRW3 = kFragRW2 + 64 * 64 / 2;  // 16x64 ->  512
constexpr int kFragWords = kFragRW3 + 16 * 64 / 2; // 5120 words = 20 KB

inline void build_mma_fragments(const uint16_t* mlp_rowmajor, uint32_t* frag_out) {
	struct L { int off_h, off_w, N, K; };
	const L layers[5] = {{kOffDW1, kFragDW1, 64, 32}, {kOffDW2, kFragDW2, 16, 64}, {kOffRW1, kFragRW1, 64, 32}, {kOffRW2, kFragRW2, 64, 64}, {kOffRW3, kFragRW3, 16, 64}};
	for (const L& l : layers) {
		const uint16_t* W = mlp_rowmajor + l.off_h;
		const int KT = l.K / 16;
		for (int nt = 0; nt < l.N / 8; ++nt) for (int kt = 0; kt < KT; ++kt) for (int lane = 0; lane < 32; ++lane) {
			const int g = lane >> 2, t = lane & 3;
			const uint16_t* row = W + (size_t)(nt * 8 + g) * l.K + kt * 16;
			uint32_t* dst = frag_out + l.off_w + ((nt * KT + kt) * 32 + lane) * 2;
			dst[0] = (uint32_t)row[2 * t] | ((uint32_t)row[2 * t + 1] << 16);
			dst[1] = (uint32_t)row[8 + 2 * t] | ((uint32_t)row[8 + 2 * t + 1] << 16);
		}
	}
}

// ---- hash grid -------------------------------------------------------------------------------------------------
SNB_HD uint32_t grid_cell_index(uint32_t hashmap_size, uint32_t res, uint32_t x, uint32_t y, uint32_t z) {
	// Dense addressing while the running stride still fits the level, else the coherent prime hash.
	uint32_t stride = 1, index = 0;
	index += x * stride; stride *= res;
	if (stride <= hashmap_size) { index += y * stride; stride *= res; }
	if (stride <= hashmap_size) { index += z * stride; stride *= res; }
	if (hashmap_size < stride) index = x ^ (y * 2654435761u) ^ (z * 805459861u);
	return index % hashmap_size;
}

// Per-level constants of the lookup, 32 B, staged in shared memory once per CTA. `mode` is decided once per level by
// running the stride rule of grid_cell_index: 0 = dense through all three axes, 1 = hashed, 2 = anything else (generic).
struct __align__(16) LevelDesc {
	float scale;
	uint32_t s1, s2;      // dense strides of y and z (uint32 wrap-around arithmetic, as the reference's)
	uint32_t size;        // entries of this level
	const void* base;     // first entry of this level in the concatenated table (one 64-bit add per level, not per corner)
	uint32_t mode;
	uint32_t mask;        // size - 1 when size is a power of two, else 0
};
enum { kLevelDense = 0, kLevelHashed = 1, kLevelGeneric = 2 };

SNB_HD LevelDesc make_level_desc(const GridDesc& g, int level, const __half* table) {
	LevelDesc d;
	const uint32_t res = g.res[level];
	d.scale = g.scale[level];
	d.base = table + (size_t)g.offset[level] * g.n_feat;
	d.size = g.offset[level + 1] - g.offset[level];
	d.s1 = res; d.s2 = res * res;
	d.mask = (d.size & (d.size - 1u)) == 0u ? d.size - 1u : 0u;
	uint32_t stride = res;
	bool y_in = false, z_in = false;
	if (stride <= d.size) { y_in = true; stride *= res; }
	if (stride <= d.size) { z_in = true; stride *= res; }
	if (d.size < stride) d.mode = d.mask ? kLevelHashed : kLevelGeneric;
	else d.mode = (y_in && z_in) ? kLevelDense : kLevelGeneric;
	return d;
}

template <int F> struct FeatVec;
template <> struct FeatVec<1> { typedef unsigned short T; };
template <> struct FeatVec<2> { typedef uint32_t T; };
template <> struct FeatVec<4> { typedef uint2 T; };
template <> struct FeatVec<8> { typedef uint4 T; };

SNB_D __half2 u2h2(uint32_t u) { return *reinterpret_cast<__half2*>(&u); }
SNB_D uint32_t h22u(__half2 h) { return *reinterpret_cast<uint32_t*>(&h); }

__device__ __noinline__ uint32_t grid_cell_index_slow(uint32_t hashmap_size, uint32_t res, uint32_t x, uint32_t y, uint32_t z) { return grid_cell_index(hashmap_size, res, x, y, z); }

// ---- one level, split in two halves so that the gathers of several levels can be in flight together ---------------
// The 8 indices share their per-axis terms (dense: base + {0,1} + {0,s1} + {0,s2}; hashed: (x|x+1) ^ (y*P1 | +P1) ^
// (z*P2 | +P2)), which is the reference's index for each corner in uint32 arithmetic, evaluated once instead of 8 times.
struct LevelFrac { float rx, ry, rz; };

// Gathers of the hashed (fine) levels have next to no reuse beyond the x-neighbour that shares their 32 B sector: they are
// allocated in L1 as first-to-evict, so that the occupancy bitfield and the dense coarse levels keep their lines.
#if SNB_GATHER_HINT
SNB_D unsigned short ldg_stream(const unsigned short* p) { unsigned short v; asm("ld.global.nc.L1::evict_first.u16 %0, [%1];" : "=h"(v) : "l"(p)); return v; }
SNB_D uint32_t ldg_stream(const uint32_t* p) { uint32_t v; asm("ld.global.nc.L1::evict_first.b32 %0, [%1];" : "=r"(v) : "l"(p)); return v; }
SNB_D uint2 ldg_stream(const uint2* p) { uint2 v; asm("ld.global.nc.L1::evict_first.v2.b32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p)); return v; }
SNB_D uint4 ldg_stream(const uint4* p) { uint4 v; asm("ld.global.nc.L1::evict_first.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p)); return v; }
#endif

// Part 1: position in the level, corner indices (idx_out for the parity tap), issue the 8 gathers.
template <int F, bool TAP>
SNB_D void level_gather(const LevelDesc& L, v3 x, typename FeatVec<F>::T (&vals)[8], LevelFrac& fr, uint32_t* idx_out) {
	const float px = fmaf(L.scale, x.x, 0.5f), py = fmaf(L.scale, x.y, 0.5f), pz = fmaf(L.scale, x.z, 0.5f);
	const float flx = floorf(px), fly = floorf(py), flz = floorf(pz);
	const uint32_t gx = (uint32_t)(int)flx, gy = (uint32_t)(int)fly, gz = (uint32_t)(int)flz;
	fr.rx = px - flx; fr.ry = py - fly; fr.rz = pz - flz;
	uint32_t idx[8];
	if (L.mode == kLevelHashed) {
		// (a ^ b ^ c) & m == (a & m) ^ (b & m) ^ (c & m): mask the six per-axis terms, then one 3-input XOR per corner
		const uint32_t hy = gy * 2654435761u, hz = gz * 805459861u, m = L.mask;
		const uint32_t x0 = gx & m, x1 = (gx + 1u) & m, y0 = hy & m, y1 = (hy + 2654435761u) & m, z0 = hz & m, z1 = (hz + 805459861u) & m;
		idx[0] = x0 ^ y0 ^ z0; idx[1] = x1 ^ y0 ^ z0; idx[2] = x0 ^ y1 ^ z0; idx[3] = x1 ^ y1 ^ z0;
		idx[4] = x0 ^ y0 ^ z1; idx[5] = x1 ^ y0 ^ z1; idx[6] = x0 ^ y1 ^ z1; idx[7] = x1 ^ y1 ^ z1;
	} else {
		const uint32_t b0 = gx + gy * L.s1 + gz * L.s2;
		const uint32_t b1 = b0 + L.s1, b2 = b0 + L.s2, b3 = b1 + L.s2;
		idx[0] = b0; idx[1] = b0 + 1u; idx[2] = b1; idx[3] = b1 + 1u; idx[4] = b2; idx[5] = b2 + 1u; idx[6] = b3; idx[7] = b3 + 1u;
		// dense level and the largest corner index in range (and no uint32 wrap): nothing to reduce. Otherwise (last cell of a
		// dense level, positions outside [0,1], or a level the stride rule treats specially) take the reference's formula per corner.
		if (L.mode != kLevelDense || b0 > idx[7] || idx[7] >= L.size) {
#pragma unroll
			for (int c = 0; c < 8; ++c) idx[c] = grid_cell_index_slow(L.size, L.s1, gx + (c & 1), gy + ((c >> 1) & 1), gz + ((c >> 2) & 1));
		}
	}
	typedef typename FeatVec<F>::T V;
	const V* base = reinterpret_cast<const V*>(L.base);
#if SNB_GATHER_HINT
	if (L.mode == kLevelHashed) {
#pragma unroll
		for (int c = 0; c < 8; ++c) vals[c] = ldg_stream(base + idx[c]);
	} else
#endif
	{
#pragma unroll
		for (int c = 0; c < 8; ++c) vals[c] = __ldg(base + idx[c]);
	}
	if (TAP) {
#pragma unroll
		for (int c = 0; c < 8; ++c) idx_out[c] = idx[c];
	}
}

// Part 2: trilinear weights ((1 * wx) * wy) * wz rounded to half once (grid.h:144), fp16 FMA accumulate in corner order.
// `out` receives F halves packed in F/2 words (F == 1: low half of out[0]).
template <int F>
SNB_D void level_blend(const typename FeatVec<F>::T (&vals)[8], const LevelFrac& fr, uint32_t* out) {
	const float rx = fr.rx, ry = fr.ry, rz = fr.rz;
	const float ax = 1.0f - rx, ay = 1.0f - ry, az = 1.0f - rz;
	const float wxy0 = ax * ay, wxy1 = rx * ay, wxy2 = ax * ry, wxy3 = rx * ry;
	__half w[8];
	w[0] = __float2half_rn(wxy0 * az); w[1] = __float2half_rn(wxy1 * az); w[2] = __float2half_rn(wxy2 * az); w[3] = __float2half_rn(wxy3 * az);
	w[4] = __float2half_rn(wxy0 * rz); w[5] = __float2half_rn(wxy1 * rz); w[6] = __float2half_rn(wxy2 * rz); w[7] = __float2half_rn(wxy3 * rz);
	if constexpr (F == 1) {
		__half r = __ushort_as_half((unsigned short)0);
#pragma unroll
		for (int c = 0; c < 8; ++c) r = __hfma(w[c], __ushort_as_half(vals[c]), r);
		out[0] = (uint32_t)__half_as_ushort(r);
	} else if constexpr (F == 2) {
		__half2 r = u2h2(0u);
#pragma unroll
		for (int c = 0; c < 8; ++c) r = __hfma2(__half2half2(w[c]), u2h2(vals[c]), r);
		out[0] = h22u(r);
	} else if constexpr (F == 4) {
		__half2 r0 = u2h2(0u), r1 = u2h2(0u);
#pragma unroll
		for (int c = 0; c < 8; ++c) { const __half2 ww = __half2half2(w[c]); r0 = __hfma2(ww, u2h2(vals[c].x), r0); r1 = __hfma2(ww, u2h2(vals[c].y), r1); }
		out[0] = h22u(r0); out[1] = h22u(r1);
	} else {
		__half2 r0 = u2h2(0u), r1 = u2h2(0u), r2 = u2h2(0u), r3 = u2h2(0u);
#pragma unroll
		for (int c = 0; c < 8; ++c) {
			const __half2 ww = __half2half2(w[c]);
			r0 = __hfma2(ww, u2h2(vals[c].x), r0); r1 = __hfma2(ww, u2h2(vals[c].y), r1);
			r2 = __hfma2(ww, u2h2(vals[c].z), r2); r3 = __hfma2(ww, u2h2(vals[c].w), r3);
		}
		out[0] = h22u(r0); out[1] = h22u(r1); out[2] = h22u(r2); out[3] = h22u(r3);
	}
}

// All levels of one sample -> the 32 feature halves of this lane's row of the staging tile (level-major,
// feature-minor, as the reference's AoS row). Levels are processed in groups of G: all 8*G gathers of a group are issued
// before the first one is consumed, so a sample pays L/G dependent L2 round trips instead of L (with in-order issue a plain
// unrolled loop serialises them: each level's HFMA chain blocks the next level's loads). The group loop itself is not
// unrolled: the march kernel's hot loop has to stay inside the instruction cache (profiles/r01_ncu_fused_v2.md, r01_ncu_fused_v3.md).
// CS = byte stride between consecutive 16-byte chunks (8 halves) of the row: 16 for a plain row-major tile, 128 for the
// tcgen05 K-major core-matrix layout (snb_mlp_tc.cuh).
template <int F, bool TAP, int CS = 16>
SNB_D void encode_position_to_row(const LevelDesc* __restrict__ levels, v3 x, __half* row, uint32_t* idx_out) {
	constexpr int L = kFeat / F;
	constexpr int G = F <= 2 ? SNB_ENC_GROUP : 1;
	typedef typename FeatVec<F>::T V;
#pragma unroll 1
	for (int l0 = 0; l0 < L; l0 += G) {
		V vals[G][8];
		LevelFrac fr[G];
#pragma unroll
		for (int g = 0; g < G; ++g) level_gather<F, TAP>(levels[l0 + g], x, vals[g], fr[g], TAP ? idx_out + (l0 + g) * 8 : nullptr);
#pragma unroll
		for (int g = 0; g < G; ++g) {
			uint32_t o[F >= 2 ? F / 2 : 1];
			level_blend<F>(vals[g], fr[g], o);
			const int l = l0 + g;
			unsigned char* rb = reinterpret_cast<unsigned char*>(row);
			if constexpr (F == 1) *reinterpret_cast<__half*>(rb + (l / 8) * CS + (l % 8) * 2) = __ushort_as_half((unsigned short)o[0]);
			else if constexpr (F == 2) *reinterpret_cast<uint32_t*>(rb + (l / 4) * CS + (l % 4) * 4) = o[0];
			else if constexpr (F == 4) *reinterpret_cast<uint2*>(rb + (l / 2) * CS + (l % 2) * 8) = make_uint2(o[0], o[1]);
			else *reinterpret_cast<uint4*>(rb + l * CS) = make_uint4(o[0], o[1], o[2], o[3]);
		}
	}
}

// SH degree 4 of dir*2-1 (dir is the warped direction in [0,1]^3), rounded to fp16; 16 halves in 8 words.
SNB_D void encode_direction_sh4(v3 dir_warped, uint32_t (&out)[8]) {
	const float x = dir_warped.x * 2.f - 1.f, y = dir_warped.y * 2.f - 1.f, z = dir_warped.z * 2.f - 1.f;
	const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
	float s[16];
	s[0] = 0.28209479177387814f;
	s[1] = -0.48860251190291987f * y;
	s[2] = 0.48860251190291987f * z;
	s[3] = -0.48860251190291987f * x;
	s[4] = 1.0925484305920792f * xy;
	s[5] = -1.0925484305920792f * yz;
	s[6] = 0.94617469575755997f * z2 - 0.31539156525251999f;
	s[7] = -1.0925484305920792f * xz;
	s[8] = 0.54627421529603959f * x2 - 0.54627421529603959f * y2;
	s[9] = 0.59004358992664352f * y * (-3.0f * x2 + y2);
	s[10] = 2.8906114426405538f * xy * z;
	s[11] = 0.45704579946446572f * y * (1.0f - 5.0f * z2);
	s[12] = 0.3731763325901154f * z * (5.0f * z2 - 3.0f);
	s[13] = 0.45704579946446572f * x * (1.0f - 5.0f * z2);
	s[14] = 1.4453057213202769f * z * (x2 - y2);
	s[15] = 0.59004358992664352f * x * (-x2 + 3.0f * y2);
#pragma unroll
	for (int i = 0; i < 8; ++i) out[i] = h22u(__floats2half2_rn(s[2 * i], s[2 * i + 1]));
}

// ---- warp-level MLP (mma.sync) -----------------------------------------------------------------------------------
constexpr int kTileStrideH = 56;                    // halves per staged row: [32 features | 16 SH | 8 pad] = 112 B -> conflict-free ldmatrix
constexpr int kTileShCol = 32;                      // SH block of the row (rewritten only when a slot gets a new ray)
constexpr int kWarpTileBytes = 32 * kTileStrideH * 2; // 3584
constexpr int kWarpOutBytes = 32 * 8;               // rgbσ halves per row
constexpr int kWarpSmemBytes = kWarpTileBytes + kWarpOutBytes;

SNB_D void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
	asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
	             : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
SNB_D void ldmatrix_x4(uint32_t (&r)[4], const void* smem_row_ptr) {
	const uint32_t addr = (uint32_t)__cvta_generic_to_shared(smem_row_ptr);
	asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
SNB_D uint32_t pack_relu_h2(float a, float b, bool relu) {
	// round first, then clamp both halves with one HMNMX2: max(round(x), 0) == round(max(x, 0)) (rounding is monotonic)
	const __half2 h = __floats2half2_rn(a, b);
	return relu ? h22u(__hmax2(h, u2h2(0u))) : h22u(h);
}

// Load the A fragments (2 m-tiles x KT k-tiles) of a staged [32 x 16*KT] half tile.
template <int KT>
SNB_D void load_a_frags(const __half* tile, int lane, uint32_t (&a)[2][KT][4]) {
#pragma unroll
	for (int m = 0; m < 2; ++m)
#pragma unroll
		for (int kt = 0; kt < KT; ++kt) {
			const int row = 16 * m + (lane & 7) + ((lane >> 3) & 1) * 8;
			const int col = 16 * kt + (lane >> 4) * 8;
			ldmatrix_x4(a[m][kt], tile + row * kTileStrideH + col);
		}
}

// One dense layer for the warp's 32 rows. Output n-tiles [NT0, NT0 + 2*KT_OUT) are produced two at a time and
// folded straight into the next layer's A fragments (C fragment of n-tiles 2k,2k+1 == A fragment of k-tile k).
template <bool GLOBAL_W> SNB_D uint2 load_wfrag(const uint32_t* p) {
	if (GLOBAL_W) return __ldg(reinterpret_cast<const uint2*>(p));   // weights read through L1 (read-only path)
	return *reinterpret_cast<const uint2*>(p);                       // weights staged in shared memory
}

template <int KT_IN, int KT_OUT, bool RELU, bool GLOBAL_W>
SNB_D void dense_to_frags(const uint32_t* __restrict__ wfrag, int lane, const uint32_t (&a)[2][KT_IN][4], uint32_t (&out)[2][KT_OUT][4]) {
#pragma unroll
	for (int ko = 0; ko < KT_OUT; ++ko) {
		float acc[2][2][4];
#pragma unroll
		for (int m = 0; m < 2; ++m)
#pragma unroll
			for (int h = 0; h < 2; ++h)
#pragma unroll
				for (int i = 0; i < 4; ++i) acc[m][h][i] = 0.0f;
#pragma unroll
		for (int h = 0; h < 2; ++h) {
			const int nt = 2 * ko + h;
#pragma unroll
			for (int kt = 0; kt < KT_IN; ++kt) {
				const uint2 b = load_wfrag<GLOBAL_W>(wfrag + ((nt * KT_IN + kt) * 32 + lane) * 2);
				mma16816(acc[0][h], a[0][kt], b.x, b.y);
				mma16816(acc[1][h], a[1][kt], b.x, b.y);
			}
		}
#pragma unroll
		for (int m = 0; m < 2; ++m) {
			out[m][ko][0] = pack_relu_h2(acc[m][0][0], acc[m][0][1], RELU);
			out[m][ko][1] = pack_relu_h2(acc[m][0][2], acc[m][0][3], RELU);
			out[m][ko][2] = pack_relu_h2(acc[m][1][0], acc[m][1][1], RELU);
			out[m][ko][3] = pack_relu_h2(acc[m][1][2], acc[m][1][3], RELU);
		}
	}
}

// Full network for the warp's 32 samples.
//   the tile holds each lane's 32 encoded halves in its row; sh_frag: the rays' SH tile already in A-fragment form (constant per ray);
//   returns this lane's {r, g, b, sigma} as raw fp16 network outputs (activations are applied by the compositor).
// smem_w: fragment-ordered weights (kFragWords words); tile/out_tile: this warp's staging areas.
template <bool GLOBAL_W = false>
SNB_D uint2 warp_network_eval(const uint32_t* __restrict__ smem_w, __half* tile, __half* out_tile, int lane, const uint32_t (&sh_frag)[2][4]) {
	// the features were staged row-major by encode_position_to_row; pull them back as A fragments
	__syncwarp();
	uint32_t a_feat[2][2][4];
	load_a_frags<2>(tile, lane, a_feat);
	uint32_t h1[2][4][4];
	dense_to_frags<2, 4, true, GLOBAL_W>(smem_w + kFragDW1, lane, a_feat, h1);          // 32 -> 64, ReLU
	uint32_t rgb_in[2][2][4];
	{
		uint32_t dens[2][1][4];
		dense_to_frags<4, 1, false, GLOBAL_W>(smem_w + kFragDW2, lane, h1, dens);        // 64 -> 16
#pragma unroll
		for (int m = 0; m < 2; ++m)
#pragma unroll
			for (int i = 0; i < 4; ++i) { rgb_in[m][0][i] = dens[m][0][i]; rgb_in[m][1][i] = sh_frag[m][i]; }
	}
	uint32_t g1[2][4][4];
	dense_to_frags<2, 4, true, GLOBAL_W>(smem_w + kFragRW1, lane, rgb_in, g1);          // 32 -> 64, ReLU
	uint32_t g2[2][4][4];
	dense_to_frags<4, 4, true, GLOBAL_W>(smem_w + kFragRW2, lane, g1, g2);              // 64 -> 64, ReLU
	// last layer: only n-tile 0 (outputs 0..7; rgb = 0..2) is needed
	float acc[2][4];
#pragma unroll
	for (int m = 0; m < 2; ++m)
#pragma unroll
		for (int i = 0; i < 4; ++i) acc[m][i] = 0.0f;
#pragma unroll
	for (int kt = 0; kt < 4; ++kt) {
		const uint2 b = load_wfrag<GLOBAL_W>(smem_w + kFragRW3 + (kt * 32 + lane) * 2);
		mma16816(acc[0], g2[0][kt], b.x, b.y);
		mma16816(acc[1], g2[1][kt], b.x, b.y);
	}
	// transpose {r,g,b,sigma} back to the owning lane through the warp's out tile
	const int g = lane >> 2, t = lane & 3;
	uint32_t* o32 = reinterpret_cast<uint32_t*>(out_tile);
	if (t == 0) {
#pragma unroll
		for (int m = 0; m < 2; ++m) {
			o32[(16 * m + g) * 2] = h22u(__floats2half2_rn(acc[m][0], acc[m][1]));
			o32[(16 * m + g + 8) * 2] = h22u(__floats2half2_rn(acc[m][2], acc[m][3]));
		}
	} else if (t == 1) {
		// b in the low half; sigma (density output 0, already fp16 in rgb_in k-tile 0, lane t==0) is filled below
#pragma unroll
		for (int m = 0; m < 2; ++m) {
			reinterpret_cast<__half*>(o32)[(16 * m + g) * 4 + 2] = __float2half_rn(acc[m][0]);
			reinterpret_cast<__half*>(o32)[(16 * m + g + 8) * 4 + 2] = __float2half_rn(acc[m][2]);
		}
	}
	if (t == 0) {
		// rgb_in[m][0][0] = {row g: cols 0,1}, [1] = {row g+8: cols 0,1}; column 0 is the log-density
#pragma unroll
		for (int m = 0; m < 2; ++m) {
			reinterpret_cast<unsigned short*>(o32)[(16 * m + g) * 4 + 3] = (unsigned short)(rgb_in[m][0][0] & 0xFFFFu);
			reinterpret_cast<unsigned short*>(o32)[(16 * m + g + 8) * 4 + 3] = (unsigned short)(rgb_in[m][0][1] & 0xFFFFu);
		}
	}
	__syncwarp();
	const uint2 r = *reinterpret_cast<const uint2*>(o32 + lane * 2);
	__syncwarp();
	return r;
}

// The same network with the two 16-row m-tiles evaluated one after the other in a rolled loop: half the code of
// warp_network_eval (the persistent march kernel has to stay inside the 32 KB instruction cache, see
// profiles/r01_ncu_fused_v3.md) and half the live fragments; weight fragments are fetched twice instead. The SH block is read
// from the staged row (k-tile 2) instead of being carried in registers. Bit-identical results (same mma order per row).
template <int KT_IN, int KT_OUT, bool RELU, bool GLOBAL_W>
SNB_D void dense_to_frags_1m(const uint32_t* __restrict__ wfrag, int lane, const uint32_t (&a)[KT_IN][4], uint32_t (&out)[KT_OUT][4]) {
#pragma unroll
	for (int ko = 0; ko < KT_OUT; ++ko) {
		float acc[2][4];
#pragma unroll
		for (int h = 0; h < 2; ++h)
#pragma unroll
			for (int i = 0; i < 4; ++i) acc[h][i] = 0.0f;
#pragma unroll
		for (int h = 0; h < 2; ++h) {
			const int nt = 2 * ko + h;
#pragma unroll
			for (int kt = 0; kt < KT_IN; ++kt) {
				const uint2 b = load_wfrag<GLOBAL_W>(wfrag + ((nt * KT_IN + kt) * 32 + lane) * 2);
				mma16816(acc[h], a[kt], b.x, b.y);
			}
		}
		out[ko][0] = pack_relu_h2(acc[0][0], acc[0][1], RELU);
		out[ko][1] = pack_relu_h2(acc[0][2], acc[0][3], RELU);
		out[ko][2] = pack_relu_h2(acc[1][0], acc[1][1], RELU);
		out[ko][3] = pack_relu_h2(acc[1][2], acc[1][3], RELU);
	}
}

template <bool GLOBAL_W = false>
SNB_D uint2 warp_network_eval_rolled(const uint32_t* __restrict__ smem_w, __half* tile, __half* out_tile, int lane) {
	__syncwarp();
	const int g = lane >> 2, t = lane & 3;
	uint32_t* o32 = reinterpret_cast<uint32_t*>(out_tile);
#pragma unroll 1
	for (int m = 0; m < 2; ++m) {
		const __half* rows = tile + (16 * m + (lane & 7) + ((lane >> 3) & 1) * 8) * kTileStrideH + (lane >> 4) * 8;
		uint32_t a_feat[2][4];
		ldmatrix_x4(a_feat[0], rows);
		ldmatrix_x4(a_feat[1], rows + 16);
		uint32_t h1[4][4];
		dense_to_frags_1m<2, 4, true, GLOBAL_W>(smem_w + kFragDW1, lane, a_feat, h1);       // 32 -> 64, ReLU
		uint32_t rgb_in[2][4];
		{
			uint32_t dens[1][4];
			dense_to_frags_1m<4, 1, false, GLOBAL_W>(smem_w + kFragDW2, lane, h1, dens);     // 64 -> 16
#pragma unroll
			for (int i = 0; i < 4; ++i) rgb_in[0][i] = dens[0][i];
		}
		ldmatrix_x4(rgb_in[1], rows + kTileShCol);                                        // SH block of the rows
		uint32_t g1[4][4];
		dense_to_frags_1m<2, 4, true, GLOBAL_W>(smem_w + kFragRW1, lane, rgb_in, g1);       // 32 -> 64, ReLU
		uint32_t g2[4][4];
		dense_to_frags_1m<4, 4, true, GLOBAL_W>(smem_w + kFragRW2, lane, g1, g2);           // 64 -> 64, ReLU
		float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
		for (int kt = 0; kt < 4; ++kt) {
			const uint2 b = load_wfrag<GLOBAL_W>(smem_w + kFragRW3 + (kt * 32 + lane) * 2);
			mma16816(acc, g2[kt], b.x, b.y);
		}
		if (t == 0) {
			o32[(16 * m + g) * 2] = h22u(__floats2half2_rn(acc[0], acc[1]));
			o32[(16 * m + g + 8) * 2] = h22u(__floats2half2_rn(acc[2], acc[3]));
			reinterpret_cast<unsigned short*>(o32)[(16 * m + g) * 4 + 3] = (unsigned short)(rgb_in[0][0] & 0xFFFFu);
			reinterpret_cast<unsigned short*>(o32)[(16 * m + g + 8) * 4 + 3] = (unsigned short)(rgb_in[0][1] & 0xFFFFu);
		} else if (t == 1) {
			reinterpret_cast<__half*>(o32)[(16 * m + g) * 4 + 2] = __float2half_rn(acc[0]);
			reinterpret_cast<__half*>(o32)[(16 * m + g + 8) * 4 + 2] = __float2half_rn(acc[2]);
		}
	}
	__syncwarp();
	const uint2 r = *reinterpret_cast<const uint2*>(o32 + lane * 2);
	__syncwarp();
	return r;
}

// SH block of the staging tile: a lane (re)writes its row when its slot receives a new ray ...
SNB_D void stage_sh_row(__half* tile, int lane, const uint32_t (&sh)[8]) {
	uint4* dst = reinterpret_cast<uint4*>(tile + lane * kTileStrideH + kTileShCol);
	dst[0] = make_uint4(sh[0], sh[1], sh[2], sh[3]);
	dst[1] = make_uint4(sh[4], sh[5], sh[6], sh[7]);
}
// ... and the warp reloads the A fragments of that block (k-tile 2 of the staged row).
SNB_D void load_sh_frags(const __half* tile, int lane, uint32_t (&sh_frag)[2][4]) {
#pragma unroll
	for (int m = 0; m < 2; ++m) {
		const int row = 16 * m + (lane & 7) + ((lane >> 3) & 1) * 8;
		const int col = kTileShCol + (lane >> 4) * 8;
		ldmatrix_x4(sh_frag[m], tile + row * kTileStrideH + col);
	}
}

} // namespace snb
