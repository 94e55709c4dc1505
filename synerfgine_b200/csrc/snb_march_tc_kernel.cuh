// synerfgine_b200 — the persistent NeRF march kernel with the MLPs on tcgen05 tensor cores ("tc" form).
//
// Same replacement set as k_nerf_march_fused (snb_nerf_kernels.cuh): generate_next_nerf_network_inputs, compact_kernel_nerf,
// kernel_grid, kernel_sh, kernel_mlp_fused x2, extract_density, composite_kernel_nerf(_alt), extract_from_payload /
// shade_kernel_nerf and the host loops of NerfTracer::trace(_alt) (reference src/testbed_nerf.cu:476-837,1578-1853,2128-2401).
//
// One CTA = 128 threads = 128 ray slots = one 128-row UMMA tile. Per iteration every thread (1) refills its slot from the
// global queue (warp-aggregated atomicAdd), (2) takes up to kTcSearchSteps empty-space DDA steps, (3) if it holds a sample,
// gathers the hash-grid features straight into its row of the K-major operand buffer, (4) all 128 threads run both MLPs
// through tc::cta_network_eval_compact (tcgen05.mma into TMEM, tcgen05.ld epilogues), (5) composites in registers.
// Ray state and the MLP epilogue need few registers (accumulators live in TMEM), so 5 CTAs = 20 warps fit an SM
// (37 KB shared memory and 64 TMEM columns each) against 16 warps for the mma.sync form.
#pragma once
#include "snb_nerf_kernels.cuh"
#include "snb_mlp_tc.cuh"

namespace snb {

constexpr int kTcSearchSteps = 2;
constexpr int kTcSmemBytes = tc::kCompactBytes + 1024 /* alignment slack */ + (int)sizeof(LevelDesc) * kMaxLevels + kCoarseWords * 4;

template <int F, bool ALT>
__global__ void __launch_bounds__(tc::kTileRows, 5)
k_nerf_march_tc(DevModel model, DevView view, NerfOut out, int render_mode, const RayRec* __restrict__ queue, const unsigned int* __restrict__ queue_tail,
                unsigned int* __restrict__ queue_head, NerfCounters* __restrict__ counters) {
	extern __shared__ unsigned char smem_dyn[];
	unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
	LevelDesc* levels = reinterpret_cast<LevelDesc*>(smem + tc::kCompactBytes);
	const int tid = threadIdx.x, lane = tid & 31;
	uint32_t* s_coarse = reinterpret_cast<uint32_t*>(smem + tc::kCompactBytes + sizeof(LevelDesc) * kMaxLevels);
	if (tid < kFeat / F) levels[tid] = make_level_desc(model.grid, tid, model.grid_params);
	if (model.env.coarse) for (int i = tid; i < kCoarseWords; i += tc::kTileRows) s_coarse[i] = __ldg(model.env.coarse + i);
	tc::CtaMlp mlp;
	tc::cta_mlp_init_compact(mlp, smem, model.mlp_umma);
	unsigned char* frow = smem + tc::kCompactBuf + tc::chunk_offset<32>(tid, 0);   // chunk kc of this thread's feature row: frow + kc * 128
	__syncthreads();

	const unsigned int n_rays = *queue_tail;
	MarchEnv env = model.env;
	if (env.coarse) env.coarse = s_coarse;
	const v3 cam_fwd = view.cam.c2;
	const v3 aabb_diag = model.train_aabb.diag();

	// per-slot ray state
	v3 o = mk3(0, 0, 0), d = mk3(1, 0, 0), idir = mk3(1, 1, 1);
	float t = 0.0f, depth = 0.0f, max_weight = 0.0f;
	float4 rgba = make_float4(0, 0, 0, 0);
	uint32_t idx = 0, step_i = 1;
	int k = 0, n_empty = 0;
	bool alive = false, drained = false, searching = false;
	uint32_t sh[8];
#pragma unroll
	for (int i = 0; i < 8; ++i) sh[i] = 0u;
	uint32_t n_samples = 0, n_hits = 0, n_launched = 0, n_dda = 0, n_iter = 0;

	auto finish = [&]() {
		alive = false; searching = false;
		if (step_i + kBatchSteps >= kMarchIter) return;   // died in the last possible iteration: never compacted into the hit list
		if (ALT) t = depth / dot(cam_fwd, d);             // payload.t rewrite at the end of composite_kernel_nerf_alt
		if (rgba.w > 0.001f) { write_pixel(view, out, ALT, render_mode, idx, d, t, rgba, depth); ++n_hits; }
	};

	for (;;) {
		// -- refill empty slots: one atomic per warp -----------------------------------------------------------------
		const unsigned int need = __ballot_sync(0xFFFFFFFFu, !alive);
		if (need && !drained) {
			unsigned int base = 0;
			if (lane == 0) base = atomicAdd(queue_head, __popc(need));
			base = __shfl_sync(0xFFFFFFFFu, base, 0);
			const unsigned int mine = base + __popc(need & ((1u << lane) - 1u));
			if (!alive && mine < n_rays) {
				const float4 q0 = __ldg(reinterpret_cast<const float4*>(queue + mine));
				const float4 q1 = __ldg(reinterpret_cast<const float4*>(queue + mine) + 1);
				o = mk3(q0.x, q0.y, q0.z); t = q0.w; d = mk3(q1.x, q1.y, q1.z); idx = __float_as_uint(q1.w);
				idir = 1.0f / d;
				rgba = make_float4(0, 0, 0, 0); depth = 0.0f; max_weight = 0.0f; step_i = 1; k = 0; n_empty = 0; alive = true;
				searching = false;                               // ray generation already advanced t to the first occupied cell
				++n_launched;
				encode_direction_sh4((d + 1.0f) * 0.5f, sh);   // warp_direction, nerf_device.cuh:290-292
			}
			if (base + __popc(need) >= n_rays) drained = true;
		}
		if (__syncthreads_or(alive ? 1 : 0) == 0) break;        // CTA-uniform exit: the evaluator below is CTA-collective
		if (tid == 0) ++n_iter;

		// -- search: bounded number of DDA steps of if_unoccupied_advance_to_next_occupied_voxel ---------------------------------
		if (alive && searching) {
#pragma unroll 1
			for (int step = 0; step < kTcSearchSteps; ++step) {
				const v3 pos = o + d * t;
				if (t >= kMaxDepth || !in_render_box(pos, env)) { finish(); break; }   // left the volume
				uint32_t mip = clamp_mip(mip_of(pos), env);
				if (!env.bitfield || cell_occupied_xyz(pos, env.bits_xyz, mip)) { searching = false; n_empty = 0; break; }
				if (((n_empty++) & (kLookAheadEvery - 1)) == 0 && coarse_clear_ahead(pos, d, env)) { finish(); break; }
				while (mip < env.max_mip && !cell_occupied_xyz(pos, env.bits_xyz, mip + 1)) ++mip;
				t = advance_past_voxel(t, env.st, pos, d, idir, mip);
				++n_dda;
			}
		}
		const bool valid = alive && !searching;

		// -- sample position + hash-grid gather into this thread's operand row -------------------------------------------------------
		v3 wpos = mk3(0, 0, 0);
		float wdt = 0.0f;
		if (valid) {
			const float dt = step_size(t, env.st);
			wpos = (o + d * t - model.train_aabb.lo) / aabb_diag;   // warp_position
			wdt = dt_warp(dt);
			t += dt;
			encode_position_to_row<F, false, 128>(levels, wpos, reinterpret_cast<__half*>(frow), nullptr);
		} else {
#pragma unroll
			for (int kc = 0; kc < 4; ++kc) *reinterpret_cast<uint4*>(frow + kc * 128) = make_uint4(0, 0, 0, 0);
		}
		// -- both MLPs on the tensor cores (CTA-collective) ------------------------------------------------------------------------
		const uint2 net = tc::cta_network_eval_compact(mlp, sh);

		// -- composite -------------------------------------------------------------------------------------------------
		if (valid) {
			++n_samples;
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
			if (rgba.w > (1.0f - model.min_transmittance)) {
				const float inv_a = 1.0f / rgba.w;
				rgba.x *= inv_a; rgba.y *= inv_a; rgba.z *= inv_a; rgba.w *= inv_a;
				finish();
			} else {
				if (++k == kBatchSteps) {
					if (ALT) t = depth / dot(cam_fwd, d);            // batch boundary: the sng compositor rewinds t to the last composited sample
					k = 0;
					step_i += kBatchSteps;
					if (step_i >= kMarchIter) alive = false;         // iteration budget exhausted: ray is dropped, not written
				}
				searching = true;                                    // the next cell is probed by the search phase
			}
		}
	}
	tc::cta_mlp_free_compact(mlp);

	// -- counters: one atomic per warp ----------------------------------------------------------------------------------
#pragma unroll
	for (int s = 16; s > 0; s >>= 1) {
		n_samples += __shfl_xor_sync(0xFFFFFFFFu, n_samples, s);
		n_hits += __shfl_xor_sync(0xFFFFFFFFu, n_hits, s);
		n_launched += __shfl_xor_sync(0xFFFFFFFFu, n_launched, s);
		n_dda += __shfl_xor_sync(0xFFFFFFFFu, n_dda, s);
	}
	if (tid == 0 && counters) atomicAdd(&counters->iterations, (unsigned long long)n_iter);
	if (lane == 0 && counters) {
		atomicAdd(&counters->dda_steps, (unsigned long long)n_dda);
		atomicAdd(&counters->samples, (unsigned long long)n_samples);
		atomicAdd(&counters->hits, (unsigned long long)n_hits);
		atomicAdd(&counters->rays, (unsigned long long)n_launched);
	}
}

} // namespace snb
