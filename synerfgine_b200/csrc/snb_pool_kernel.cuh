// synerfgine_b200 — the persistent NeRF march kernel, ray-pool form.
//
// Replaces (reference, src/testbed_nerf.cu): compact_kernel_nerf :1830-1853, generate_next_nerf_network_inputs :790-837,
// the network kernels it feeds (tcnn kernel_grid, kernel_sh, kernel_mlp_fused x2, extract_density), composite_kernel_nerf
// :577-788 / composite_kernel_nerf_alt :476-575, extract_from_payload :1578-1612, shade_kernel_nerf :1788-1828 and the host
// loops NerfTracer::trace / trace_alt :2128-2401.
//
// Why a pool (profiles/r01_ncu_fused_v3.md): with one ray per lane the empty-space DDA ran with 2-5 of 32
// lanes active and was ~35 % of all issued instructions although it is only ~0.5 DDA steps per network sample. A CTA-wide
// pool with compacted lists fixed the packing but spent 44 % of its time in the two CTA barriers per iteration. This form
// keeps warps autonomous (no barrier, no atomics on shared memory): every LANE owns kSlotsPerLane ray slots in shared memory
// (slot j of lane l lives at [j][l]: conflict-free), each in one of three states,
//   SEARCH (in empty space) -> one search pass = up to kSearchSteps DDA steps of a lane's searching slot,
//   READY  (holds a sample position) -> network phase: hash-grid gather, both MLPs, composite, probe of the next cell,
//   EMPTY  -> refilled from the global ray queue with one warp-aggregated atomicAdd (ballot + popc),
// and a lane works on whichever of its slots needs the phase the warp is running. With two slots per lane ~94 % of the
// lanes have a READY ray in the network phase and ~45 % a SEARCH ray in a search pass (one slot per lane: 94 % / 16 %).
// No per-sample data ever touches HBM; there is no host round trip.
#pragma once
#include "snb_nerf_kernels.cuh"

namespace snb {

constexpr int kPoolThreads = 128;          // 4 autonomous warps per CTA, 4 CTAs per SM
constexpr int kPoolWarps = kPoolThreads / 32;
#ifndef SNB_SLOTS_PER_LANE
#define SNB_SLOTS_PER_LANE 2
#endif
#ifndef SNB_SEARCH_STEPS
#define SNB_SEARCH_STEPS 2
#endif
#ifndef SNB_MIN_FILL
#define SNB_MIN_FILL 28
#endif
constexpr int kSlotsPerLane = SNB_SLOTS_PER_LANE;
constexpr int kPoolSlots = kSlotsPerLane * kPoolThreads;
constexpr int kSearchSteps = SNB_SEARCH_STEPS;            // DDA steps a searching slot may take per pass
constexpr int kPoolMinFill = SNB_MIN_FILL;           // run the network phase once this many lanes hold a READY slot (or nobody searches)
constexpr int kMaxSearchPasses = 8;        // ... or after this many passes

enum : uint32_t { kSlotEmpty = 0u, kSlotSearch = 1u, kSlotReady = 2u };
// meta word of a slot: [1:0] state, [5:2] steps composited in the current batch of 8, [8:6] empty cells since the last
// look-ahead (mod 8), [31:16] march iteration counter of the reference's host loop (1, 9, 17, ...)
SNB_D uint32_t meta_pack(uint32_t state, uint32_t k, uint32_t n_empty, uint32_t step_i) { return state | (k << 2) | ((n_empty & 7u) << 6) | (step_i << 16); }

struct WarpPool {                          // per warp: kSlotsPerLane x 32 rays, structure of arrays, [slot][lane]
	float ox[kSlotsPerLane][32], oy[kSlotsPerLane][32], oz[kSlotsPerLane][32];
	float dx[kSlotsPerLane][32], dy[kSlotsPerLane][32], dz[kSlotsPerLane][32], t[kSlotsPerLane][32];
	float ix[kSlotsPerLane][32], iy[kSlotsPerLane][32], iz[kSlotsPerLane][32];   // 1 / d, evaluated once per ray
	float4 rgba[kSlotsPerLane][32];
	float depth[kSlotsPerLane][32], maxw[kSlotsPerLane][32];
	uint32_t idx[kSlotsPerLane][32], meta[kSlotsPerLane][32];
	uint4 sh[kSlotsPerLane][2][32];        // SH-4 of the ray direction, 16 halves, computed at refill
};

struct PoolSmem {
	unsigned char warp[kPoolWarps][kWarpSmemBytes];           // per-warp [32 x (32 feat | 16 SH | pad)] staging tile + rgbσ tile
	LevelDesc levels[kMaxLevels];
	WarpPool pool[kPoolWarps];
};

template <int F, bool ALT>
__global__ void __launch_bounds__(kPoolThreads, 4)
k_nerf_march_pool(DevModel model, DevView view, NerfOut out, int render_mode, const RayRec* __restrict__ queue, const unsigned int* __restrict__ queue_tail,
                  unsigned int* __restrict__ queue_head, NerfCounters* __restrict__ counters) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	PoolSmem& S = *reinterpret_cast<PoolSmem*>(smem_raw);
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	WarpPool& P = S.pool[warp];

	if (tid < kFeat / F) S.levels[tid] = make_level_desc(model.grid, tid, model.grid_params);
#pragma unroll
	for (int j = 0; j < kSlotsPerLane; ++j) P.meta[j][lane] = kSlotEmpty;
	__half* tile = reinterpret_cast<__half*>(S.warp[warp]);
	__half* out_tile = reinterpret_cast<__half*>(S.warp[warp] + kWarpTileBytes);
	{ // zero this lane's tile row once so never-filled rows hold finite values
		uint4* row = reinterpret_cast<uint4*>(tile + lane * kTileStrideH);
#pragma unroll
		for (int i = 0; i < kTileStrideH * 2 / 16; ++i) row[i] = make_uint4(0, 0, 0, 0);
	}
	__syncthreads();

	const unsigned int n_rays = *queue_tail;
	const MarchEnv env = model.env;   // look-ahead grid (4 KB) and MLP weights (20 KB) are read through L1, not staged: 40 KB smem / CTA
	const v3 cam_fwd = view.cam.c2;
	const v3 aabb_diag = model.train_aabb.diag();
	bool drained = false;
	uint32_t n_samples = 0, n_hits = 0, n_launched = 0, n_dda = 0, n_iter = 0, n_pass = 0;

	// ray is done: write its pixel (if it gathered any opacity) and free the slot
	auto finish = [&](int j, v3 d, float t, uint32_t step_i) {
		P.meta[j][lane] = kSlotEmpty;
		if (step_i + kBatchSteps >= kMarchIter) return;   // died in the last possible iteration: never compacted into the hit list
		const float4 rgba = P.rgba[j][lane];
		const float depth = P.depth[j][lane];
		if (ALT) t = depth / dot(cam_fwd, d);             // payload.t rewrite at the end of composite_kernel_nerf_alt
		if (rgba.w > 0.001f) { write_pixel(view, out, ALT, render_mode, P.idx[j][lane], d, t, rgba, depth); ++n_hits; }
	};

	for (uint32_t iter = 0;; ++iter) {
		if (lane == 0) ++n_iter;
		// -- refill empty slots: one atomic per warp and slot row ------------------------------------------------------
		bool any = false;
#pragma unroll 1
		for (int j = 0; j < kSlotsPerLane; ++j) {
			uint32_t st = P.meta[j][lane] & 3u;
			if (drained) { any |= st != kSlotEmpty; continue; }
			const unsigned int need = __ballot_sync(0xFFFFFFFFu, st == kSlotEmpty);
			if (!need) { any = true; continue; }
			unsigned int base = 0;
			if (lane == 0) base = atomicAdd(queue_head, __popc(need));
			base = __shfl_sync(0xFFFFFFFFu, base, 0);
			const unsigned int mine = base + __popc(need & ((1u << lane) - 1u));
			if (st == kSlotEmpty && mine < n_rays) {
				const float4 q0 = __ldg(reinterpret_cast<const float4*>(queue + mine));
				const float4 q1 = __ldg(reinterpret_cast<const float4*>(queue + mine) + 1);
				P.ox[j][lane] = q0.x; P.oy[j][lane] = q0.y; P.oz[j][lane] = q0.z; P.t[j][lane] = q0.w;
				P.dx[j][lane] = q1.x; P.dy[j][lane] = q1.y; P.dz[j][lane] = q1.z; P.idx[j][lane] = __float_as_uint(q1.w);
				P.rgba[j][lane] = make_float4(0, 0, 0, 0); P.depth[j][lane] = 0.0f; P.maxw[j][lane] = 0.0f;
				{ const v3 idir = 1.0f / mk3(q1.x, q1.y, q1.z); P.ix[j][lane] = idir.x; P.iy[j][lane] = idir.y; P.iz[j][lane] = idir.z; }
				uint32_t sh[8];
				encode_direction_sh4((mk3(q1.x, q1.y, q1.z) + 1.0f) * 0.5f, sh);   // warp_direction, nerf_device.cuh:290-292
				P.sh[j][0][lane] = make_uint4(sh[0], sh[1], sh[2], sh[3]); P.sh[j][1][lane] = make_uint4(sh[4], sh[5], sh[6], sh[7]);
				st = kSlotReady;          // ray generation already advanced t to the first occupied cell
				P.meta[j][lane] = meta_pack(kSlotReady, 0u, 0u, 1u);
				++n_launched;
			}
			if (base + __popc(need) >= n_rays) drained = true;
			any |= st != kSlotEmpty;
		}
		if (__ballot_sync(0xFFFFFFFFu, any) == 0u) break;   // every slot of the warp is empty and the queue is drained

		// -- search passes: if_unoccupied_advance_to_next_occupied_voxel, kSearchSteps DDA steps of one searching slot per lane.
		// Deferred until fewer than kPoolMinFill lanes hold a READY slot, so that a pass finds many lanes with work. -----------
#pragma unroll 1
		for (int pass = 0; pass < kMaxSearchPasses; ++pass) {
			int js = -1;
			bool has_ready = false;
#pragma unroll
			for (int j = 0; j < kSlotsPerLane; ++j) {
				const uint32_t s = P.meta[j][lane] & 3u;
				if (s == kSlotSearch && js < 0) js = j;
				has_ready |= s == kSlotReady;
			}
			if (__ballot_sync(0xFFFFFFFFu, js >= 0) == 0u) break;
			if (__popc(__ballot_sync(0xFFFFFFFFu, has_ready)) >= kPoolMinFill) break;
			if (lane == 0) ++n_pass;
			if (js >= 0) {
				const v3 o = mk3(P.ox[js][lane], P.oy[js][lane], P.oz[js][lane]), d = mk3(P.dx[js][lane], P.dy[js][lane], P.dz[js][lane]);
				float t = P.t[js][lane];
				const uint32_t m = P.meta[js][lane];
				uint32_t n_empty = (m >> 6) & 7u;
				const v3 idir = mk3(P.ix[js][lane], P.iy[js][lane], P.iz[js][lane]);
				uint32_t s = kSlotSearch;
#pragma unroll 1
				for (int step = 0; step < kSearchSteps; ++step) {
					const v3 pos = o + d * t;
					if (t >= kMaxDepth || !in_render_box(pos, env)) { s = kSlotEmpty; break; }   // left the volume
					uint32_t mip = clamp_mip(mip_of(pos), env);
					if (!env.bitfield || cell_occupied_xyz(pos, env.bits_xyz, mip)) { s = kSlotReady; n_empty = 0; break; }
					if (((n_empty++) & (uint32_t)(kLookAheadEvery - 1)) == 0u && coarse_clear_ahead(pos, d, env)) { s = kSlotEmpty; break; }
					while (mip < env.max_mip && !cell_occupied_xyz(pos, env.bits_xyz, mip + 1)) ++mip;
					t = advance_past_voxel(t, env.st, pos, d, idir, mip);
					++n_dda;
				}
				if (s == kSlotEmpty) finish(js, d, t, m >> 16);
				else { P.t[js][lane] = t; P.meta[js][lane] = (m & ~(3u | (7u << 6))) | s | ((n_empty & 7u) << 6); }
			}
		}

		// -- network phase: one sample of one READY slot per lane (slots served round-robin) ---------------------------------------
		int jr = -1;
		{
			const int rot = (int)(iter % (uint32_t)kSlotsPerLane);
#pragma unroll
			for (int jj = 0; jj < kSlotsPerLane; ++jj) {
				int j = jj + rot; if (j >= kSlotsPerLane) j -= kSlotsPerLane;
				if (jr < 0 && (P.meta[j][lane] & 3u) == kSlotReady) jr = j;
			}
		}
		const bool valid = jr >= 0;
		if (__ballot_sync(0xFFFFFFFFu, valid) == 0u) continue;
		v3 o = mk3(0, 0, 0), d = mk3(1, 0, 0), wpos = mk3(0, 0, 0);
		float t = 0.0f, wdt = 0.0f;
		if (valid) {
			o = mk3(P.ox[jr][lane], P.oy[jr][lane], P.oz[jr][lane]); d = mk3(P.dx[jr][lane], P.dy[jr][lane], P.dz[jr][lane]); t = P.t[jr][lane];
			const float dt = step_size(t, env.st);
			wpos = (o + d * t - model.train_aabb.lo) / aabb_diag;   // warp_position
			wdt = dt_warp(dt);
			t += dt;
			__half* row = tile + lane * kTileStrideH;
			encode_position_to_row<F, false>(S.levels, wpos, row, nullptr);
			reinterpret_cast<uint4*>(row + kTileShCol)[0] = P.sh[jr][0][lane];
			reinterpret_cast<uint4*>(row + kTileShCol)[1] = P.sh[jr][1][lane];
		}
		__syncwarp();
		uint32_t sh_frag[2][4];
		load_sh_frags(tile, lane, sh_frag);
		const uint2 net = warp_network_eval<true>(model.mlp_frag, tile, out_tile, lane, sh_frag);
		if (valid) {
			++n_samples;
			const uint32_t m = P.meta[jr][lane];
			uint32_t k = (m >> 2) & 15u, step_i = m >> 16;
			float4 rgba = P.rgba[jr][lane];
			float depth = P.depth[jr][lane], max_weight = P.maxw[jr][lane];
			const __half2 rg = u2h2(net.x), bs = u2h2(net.y);
			const v3 pos = model.train_aabb.lo + wpos * aabb_diag;   // unwarp_position
			const float T = 1.f - rgba.w;
			const float dt = dt_unwarp(wdt);
			const float alpha = 1.f - __expf(-act_density(__high2float(bs), model.density_act) * dt);
			const float weight = alpha * T;
			v3 rgb = mk3(act_rgb(__low2float(rg), model.rgb_act), act_rgb(__high2float(rg), model.rgb_act), act_rgb(__low2float(bs), model.rgb_act));
			if (!ALT) {
				if (render_mode == 3) rgb = (pos - 0.5f) / 2.0f + 0.5f;
				else if (render_mode == 4) rgb = splat3(dot(cam_fwd, pos - o) * 1.0f);
			}
			rgba.x += rgb.x * weight; rgba.y += rgb.y * weight; rgba.z += rgb.z * weight; rgba.w += weight;
			if (ALT) {
				depth = dot(cam_fwd, pos - view.cam.c3);
				if (weight > max_weight) max_weight = weight;
			} else if (weight > max_weight) {
				max_weight = weight;
				depth = dot(cam_fwd, pos - view.cam.c3);
			}
			bool alive = true, dropped = false;
			if (rgba.w > (1.0f - model.min_transmittance)) {
				const float inv_a = 1.0f / rgba.w;
				rgba.x *= inv_a; rgba.y *= inv_a; rgba.z *= inv_a; rgba.w *= inv_a;
				alive = false;
			} else if (++k == (uint32_t)kBatchSteps) {
				if (ALT) t = depth / dot(cam_fwd, d);            // batch boundary: the sng compositor rewinds t to the last composited sample
				k = 0;
				step_i += kBatchSteps;
				if (step_i >= kMarchIter) { alive = false; dropped = true; }   // iteration budget exhausted: dropped, not written
			}
			P.rgba[jr][lane] = rgba; P.depth[jr][lane] = depth; P.maxw[jr][lane] = max_weight;
			if (dropped) P.meta[jr][lane] = kSlotEmpty;
			else if (!alive) finish(jr, d, t, step_i);
			else {
				// probe the next cell here (the ray is in registers): READY again if it is occupied, else SEARCH
				const v3 npos = o + d * t;
				uint32_t s = kSlotSearch;
				if (t < kMaxDepth && in_render_box(npos, env)) {
					const uint32_t mip = clamp_mip(mip_of(npos), env);
					if (!env.bitfield || cell_occupied_xyz(npos, env.bits_xyz, mip)) s = kSlotReady;
				}
				P.t[jr][lane] = t;
				P.meta[jr][lane] = meta_pack(s, k, 0u, step_i);
			}
		}
	}

	// -- counters: one atomic per warp ----------------------------------------------------------------------------------
#pragma unroll
	for (int s = 16; s > 0; s >>= 1) {
		n_samples += __shfl_xor_sync(0xFFFFFFFFu, n_samples, s);
		n_hits += __shfl_xor_sync(0xFFFFFFFFu, n_hits, s);
		n_launched += __shfl_xor_sync(0xFFFFFFFFu, n_launched, s);
		n_dda += __shfl_xor_sync(0xFFFFFFFFu, n_dda, s);
	}
	if (lane == 0 && counters) {
		atomicAdd(&counters->iterations, (unsigned long long)n_iter);
		atomicAdd(&counters->net_phases, (unsigned long long)n_pass);
		atomicAdd(&counters->dda_steps, (unsigned long long)n_dda);
		atomicAdd(&counters->samples, (unsigned long long)n_samples);
		atomicAdd(&counters->hits, (unsigned long long)n_hits);
		atomicAdd(&counters->rays, (unsigned long long)n_launched);
	}
}

} // namespace snb
