// synerfgine_b200 — the persistent NeRF march as a warp-specialised pipeline with both MLPs on the 5th-generation tensor cores
// (tcgen05.mma, accumulators in TMEM, weights by one bulk-TMA copy): march option "ws".
//
// Same replacement set as k_nerf_march_fused (snb_nerf_kernels.cuh): generate_next_nerf_network_inputs, compact_kernel_nerf,
// kernel_grid, kernel_sh, kernel_mlp_fused x2, extract_density, composite_kernel_nerf(_alt), extract_from_payload /
// shade_kernel_nerf and the host loops of NerfTracer::trace(_alt) (reference src/testbed_nerf.cu:476-837,1578-1853,2128-2401;
// tcnn src/fully_fused_mlp.cu:499-557).
//
// What makes a pipeline possible: along a ray the sample POSITIONS never depend on the network (t advances by calc_dt, and even the
// sng compositor's rewind, src/testbed_nerf.cu:574, uses the last sample's position only); the network decides where compositing
// stops. So one CTA per SM runs three roles that meet only at mbarriers:
//   * 16 PRODUCER warps = 4 groups of 128 ray slots. Each thread owns a ray: queue refill (warp-aggregated atomicAdd), occupancy
//     DDA, hash-grid gather straight into its row of the group's [128 x 32] operand tile (K-major core-matrix layout, 2 stages per
//     group) + a 16-byte row record {dt, depth, ray, flags}. Producers run ahead of the network by up to two tiles.
//   * 1 MMA warp: one thread issues the five-layer tcgen05.mma chain of a tile (M = 128, N = 64 | 16, K = 16 per instruction) into
//     the TMEM columns of one of two CHAINS (chain c serves groups 2c, 2c+1; 128 columns each), commits to an mbarrier, and
//     polls both chains with mbarrier.test_wait, so the tensor pipe works on one chain while the other's epilogue runs.
//   * 2 x 4 EPILOGUE warps (one warpgroup per chain, thread = tile row = TMEM lane): tcgen05.ld -> ReLU -> fp16 -> next A operand
//     in shared memory (hidden activations never leave the SM), the SH block of the ray joins the density outputs for the rgb
//     MLP, and after the last layer the sample is composited in registers; a ray that saturates is written out and its id is
//     posted so that its producer drops it (the one or two samples already in flight are ignored).
// No __syncthreads in steady state. Per-sample buffers never touch HBM.
//
// Batch schedule of the sng compositor (AltState, snb_nerf_kernels.cuh): the producer applies batch lengths and rewinds; the
// epilogue counts a ray into batch b when it composites that batch's first sample, i.e. exactly when the ray survived batch b-1.
#pragma once
#include "snb_nerf_kernels.cuh"
#include "snb_mlp_tc.cuh"

namespace snb {
namespace ws {

constexpr int kGroups = 4;
constexpr int kChains = 2;
constexpr int kStages = 2;
constexpr int kRows = 128;
constexpr int kProducerWarps = kGroups * 4;
constexpr int kEpilogueWarp0 = kProducerWarps;
constexpr int kMmaWarp = kProducerWarps + kChains * 4;
constexpr int kThreads = (kMmaWarp + 1) * 32;              // 800 threads: 80 registers each
constexpr uint32_t kTmemCols = 256;                        // per chain 128: hidden accumulator @0..63, 16-wide accumulator @64..79
constexpr uint32_t kTmemChain = 128;
constexpr uint32_t kColSmall = 64;

struct __align__(16) RowMeta { float dt; float depth; uint32_t ray; uint32_t flags; };
enum : uint32_t { kValid = 1u, kEnd = 2u, kDrop = 4u, kNoWrite = 8u };   // bits 8..15: batch index + 1 on the first sample of a batch
constexpr uint32_t kTerminate = 0x7E57D0E5u;               // tile header: the group has no more work
constexpr uint32_t kExit = 0xFFFFFFFFu;                    // tile_info: the chain is finished

// shared-memory carve-up (bytes) from a 1024-aligned base
constexpr int kTileF = kRows * 32 * 2;                     // 8192
constexpr int kOffW = 0;
constexpr int kOffF = kOffW + kMlpHalves * 2;
constexpr int kOffMeta = kOffF + kGroups * kStages * kTileF;
constexpr int kOffH = kOffMeta + kGroups * kStages * kRows * (int)sizeof(RowMeta);
constexpr int kOffR = kOffH + kChains * kRows * 64 * 2;
constexpr int kOffDone = kOffR + kChains * kRows * 32 * 2;
constexpr int kOffLevels = kOffDone + kGroups * kRows * 4;
constexpr int kOffCoarse = kOffLevels + (int)sizeof(LevelDesc) * kMaxLevels;
constexpr int kOffBar = kOffCoarse + kCoarseWords * 4;
// barriers (8 B each): full[g][s], empty[g][s], mma_done[c], epi_done[c], weights
constexpr int kBarFull = 0, kBarEmpty = kGroups * kStages, kBarMma = 2 * kGroups * kStages, kBarEpi = kBarMma + kChains, kBarW = kBarEpi + kChains, kNumBars = kBarW + 1;
constexpr int kOffHdr = kOffBar + kNumBars * 8;            // uint32 hdr[g][s]
constexpr int kOffInfo = kOffHdr + kGroups * kStages * 4;  // uint32 tile_info[c]
constexpr int kOffTmem = kOffInfo + kChains * 4;           // uint32 TMEM base
constexpr int kSmemBytes = kOffTmem + 16;
constexpr int kSmemAlloc = kSmemBytes + 1024;              // alignment slack

SNB_D void mbar_arrive(uint32_t mbar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" :: "r"(mbar) : "memory"); }
SNB_D bool mbar_test(uint32_t mbar, uint32_t parity) {
	uint32_t done;
	asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(done) : "r"(mbar), "r"(parity) : "memory");
	return done != 0;
}
// Bounded wait with back-off: a lost arrival must trap instead of hanging the GPU.
SNB_D void mbar_wait_long(uint32_t mbar, uint32_t parity) {
	uint32_t done = 0;
	for (uint32_t spins = 0; !done; ++spins) {
		asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(done) : "r"(mbar), "r"(parity) : "memory");
		if (spins > (1u << 26)) __trap();
	}
}
SNB_D bool group_any(int group, bool pred) {   // named barrier 1 + group over the group's 128 threads, OR-reduction of pred
	uint32_t r;
	asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.u32 p, %1, 0;\n\tbarrier.red.or.pred q, %2, %3, p;\n\tselp.u32 %0, 1, 0, q;\n\t}\n"
	             : "=r"(r) : "r"((uint32_t)pred), "r"(1 + group), "r"(kRows) : "memory");
	return r != 0;
}

struct Acc { uint32_t ray; uint32_t fin; float r, g, b, a, depth, maxw; };

} // namespace ws

template <int F, bool ALT>
__global__ void __launch_bounds__(ws::kThreads, 1)
k_nerf_march_ws(DevModel model, DevView view, NerfOut out, int render_mode, const RayRec* __restrict__ queue, const unsigned int* __restrict__ queue_tail,
                unsigned int* __restrict__ queue_head, NerfCounters* __restrict__ counters, AltCtl alt, unsigned long long* __restrict__ prof) {
	using namespace ws;
	if (alt.conditional && !alt.st->redo) return;   // later pass of the batch-schedule iteration: only after a mismatch (uniform)
	extern __shared__ unsigned char smem_dyn[];
	unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
	__shared__ uint8_t s_sched[kAltSched];
	__shared__ uint32_t s_hist[kAltSched];
	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	const uint32_t bar0 = tc::smem_u32(smem + kOffBar);
	auto bar = [&](int i) { return bar0 + 8u * (uint32_t)i; };
	volatile uint32_t* s_hdr = reinterpret_cast<volatile uint32_t*>(smem + kOffHdr);
	volatile uint32_t* s_info = reinterpret_cast<volatile uint32_t*>(smem + kOffInfo);
	volatile uint32_t* s_done = reinterpret_cast<volatile uint32_t*>(smem + kOffDone);
	LevelDesc* s_levels = reinterpret_cast<LevelDesc*>(smem + kOffLevels);
	uint32_t* s_coarse = reinterpret_cast<uint32_t*>(smem + kOffCoarse);
	const unsigned int n_rays = *queue_tail;

	// ---- one-time set-up -------------------------------------------------------------------------------------------------------
	if (tid == 0) {
		for (int i = 0; i < kGroups * kStages; ++i) { tc::mbar_init(bar(kBarFull + i), kRows); tc::mbar_init(bar(kBarEmpty + i), kRows); }
		for (int c = 0; c < kChains; ++c) { tc::mbar_init(bar(kBarMma + c), 1); tc::mbar_init(bar(kBarEpi + c), kRows); }
		tc::mbar_init(bar(kBarW), 1);
		tc::mbar_fence_init();
	}
	if (warp == kMmaWarp) tc::tmem_alloc(tc::smem_u32(smem + kOffTmem), kTmemCols);
	if (tid < kFeat / F) s_levels[tid] = make_level_desc(model.grid, tid, model.grid_params);
	if (model.env.coarse) for (int i = tid; i < kCoarseWords; i += kThreads) s_coarse[i] = __ldg(model.env.coarse + i);
	for (int i = tid; i < kGroups * kRows; i += kThreads) s_done[i] = 0xFFFFFFFFu;
	if (tid < kAltSched) {
		uint32_t steps = kBatchSteps;
		if (ALT && alt.fixed_steps) steps = alt.fixed_steps;
		else if (ALT && alt.st) {
			const uint32_t s0 = alt_steps_for(alt.target_queries, n_rays);
			steps = (tid == 0 && !alt.external) ? s0 : alt.st->sched[tid];
			if (steps == 0) steps = s0;
		}
		s_sched[tid] = (uint8_t)steps;
		s_hist[tid] = 0u;
	}
	tc::fence_before_sync();
	__syncthreads();
	tc::fence_after_sync();
	const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(smem + kOffTmem);
	const bool count_batches = ALT && alt.st && !alt.fixed_steps;
	const v3 cam_fwd = view.cam.c2;
	uint32_t n_samples = 0, n_hits = 0, n_launched = 0;

	if (warp < kProducerWarps) {
		// ===== PRODUCER ==========================================================================================================
		const int g = warp >> 2, row = tid & (kRows - 1);
		MarchEnv env = model.env;
		if (env.coarse) env.coarse = s_coarse;
		const v3 aabb_diag = model.train_aabb.diag();
		v3 o = mk3(0, 0, 0), d = mk3(1, 0, 0), idir = mk3(1, 1, 1);
		float t = 0.0f;
		uint32_t ray = 0xFFFFFFFFu, step_i = 1, batch = 0, batch_steps = s_sched[0], pending = 0;
		int k = 0, n_empty = 0;
		bool alive = false, drained = false;
		uint32_t use = 0;                                           // tiles this group has published
		long long t_wait = 0, t_bar = 0, t_search = 0, t_enc = 0;   // pipeline balance (prof): where a producer warp's time goes
		const long long t_begin = clock64();
		for (;;) {
			if (alive && s_done[g * kRows + row] == ray) alive = false;   // the epilogue saturated this ray: drop it (samples in flight are ignored there)
			// -- refill: one atomic per warp ------------------------------------------------------------------------------------------
			const unsigned int need = __ballot_sync(0xFFFFFFFFu, !alive);
			if (need && !drained) {
				unsigned int base = 0;
				if (lane == 0) base = atomicAdd(queue_head, __popc(need));
				base = __shfl_sync(0xFFFFFFFFu, base, 0);
				const unsigned int mine = base + __popc(need & ((1u << lane) - 1u));
				if (!alive && mine < n_rays) {
					const float4 q0 = __ldg(reinterpret_cast<const float4*>(queue + mine));
					const float4 q1 = __ldg(reinterpret_cast<const float4*>(queue + mine) + 1);
					o = mk3(q0.x, q0.y, q0.z); t = q0.w; d = mk3(q1.x, q1.y, q1.z);
					idir = 1.0f / d;
					ray = mine; step_i = 1; k = 0; n_empty = 0; batch = 0; batch_steps = s_sched[0]; pending = 0; alive = true;
					++n_launched;
				}
				if (base + __popc(need) >= n_rays) drained = true;
			}
			const uint32_t stage = use & 1u;
			unsigned char* f_tile = smem + kOffF + (g * kStages + stage) * kTileF;
			RowMeta* meta = reinterpret_cast<RowMeta*>(smem + kOffMeta) + (g * kStages + stage) * kRows + row;
			long long c0 = clock64();
			const bool any_alive = group_any(g, alive);
			long long c1 = clock64(); t_bar += c1 - c0;
			if (!any_alive) {
				// the whole group is out of work: hand the consumer a terminate tile and leave
				tc::mbar_wait(bar(kBarEmpty + g * kStages + stage), ((use >> 1) & 1u) ^ 1u);
				if (row == 0) s_hdr[g * kStages + stage] = kTerminate;
				mbar_arrive(bar(kBarFull + g * kStages + stage));
				break;
			}
			// -- search phase: empty-space DDA, one step per pass, until enough lanes of the warp hold a sample ---------------------
			bool valid = false;
			uint32_t end_flags = 0;
			if (alive && pending) { end_flags = pending; pending = 0; alive = false; }
			{
				bool searching = alive;
#pragma unroll 1
				for (;;) {
					if (searching) {
						const v3 pos = o + d * t;
						if (t >= kMaxDepth || !in_render_box(pos, env)) { end_flags = kEnd; searching = false; alive = false; }
						else {
							uint32_t mip = clamp_mip(mip_of(pos), env);
							if (!env.bitfield || cell_occupied_xyz(pos, env.bits_xyz, mip)) { valid = true; searching = false; n_empty = 0; }
							else if (((n_empty++) & (kLookAheadEvery - 1)) == 0 && coarse_clear_ahead_call(pos.x, pos.y, pos.z, d.x, d.y, d.z, env.coarse, env.coarse_lo.x, env.coarse_lo.y, env.coarse_lo.z, env.coarse_inv_cell)) { end_flags = kEnd; searching = false; alive = false; }
							else {
								while (mip < env.max_mip && !cell_occupied_xyz(pos, env.bits_xyz, mip + 1)) ++mip;
								t = advance_past_voxel_call(t, env.st, pos, d, idir, mip);
							}
						}
					}
					if (__ballot_sync(0xFFFFFFFFu, searching) == 0) break;
					if (__popc(__ballot_sync(0xFFFFFFFFu, valid)) >= kMinFill) break;
				}
			}
			c0 = clock64(); t_search += c0 - c1;
			// a ray that dies in the last possible lock-step iteration is never compacted into the hit list (src/testbed_nerf.cu:2165-2271)
			const uint32_t nowrite = (step_i + batch_steps >= kMarchIter) ? (uint32_t)kNoWrite : 0u;
			if (end_flags == kEnd && nowrite) end_flags = kDrop;
			// -- publish this lane's row of the tile --------------------------------------------------------------------------------
			tc::mbar_wait(bar(kBarEmpty + g * kStages + stage), ((use >> 1) & 1u) ^ 1u);
			c1 = clock64(); t_wait += c1 - c0;
			RowMeta rec; rec.dt = 0.0f; rec.depth = 0.0f; rec.ray = ray; rec.flags = end_flags;
			if (valid) {
				const float dt = step_size_call(t, env.st);
				const v3 wpos = (o + d * t - model.train_aabb.lo) / aabb_diag;          // warp_position
				const v3 pos = model.train_aabb.lo + wpos * aabb_diag;                   // unwarp_position, as the compositor sees it
				rec.dt = dt_unwarp(dt_warp(dt));
				rec.depth = dot(cam_fwd, pos - view.cam.c3);
				rec.flags = kValid | nowrite | (k == 0 ? ((batch < (uint32_t)kAltSched ? batch : (uint32_t)kAltSched - 1u) + 1u) << 8 : 0u);
				t += dt;
				encode_position_to_row<F, false, 128>(s_levels, wpos, reinterpret_cast<__half*>(f_tile + tc::chunk_offset<32>(row, 0)), nullptr);
				if (++k == (int)batch_steps) {
					// batch boundary: the sng compositor rewinds t to the last composited sample (which is this one if the ray is still alive;
					// if it is not, the epilogue says so and everything produced from here on is ignored)
					if (ALT) t = rec.depth / dot(cam_fwd, d);
					k = 0;
					step_i += batch_steps;
					if (step_i >= kMarchIter) pending = kDrop;                            // iteration budget exhausted: dropped, not written
					else { ++batch; batch_steps = batch < (uint32_t)kAltSched ? s_sched[batch] : ((ALT && alt.fixed_steps) ? alt.fixed_steps : (uint32_t)kBatchSteps); }
				}
			}
			*meta = rec;
			if (row == 0) s_hdr[g * kStages + stage] = 0u;
			tc::fence_async_smem();                                                       // generic-proxy writes -> visible to the tensor core's async proxy
			mbar_arrive(bar(kBarFull + g * kStages + stage));
			++use;
			t_enc += clock64() - c1;
		}
		if (prof && lane == 0) {
			atomicAdd(prof + 0, (unsigned long long)(clock64() - t_begin)); atomicAdd(prof + 1, (unsigned long long)t_wait); atomicAdd(prof + 2, (unsigned long long)t_bar);
			atomicAdd(prof + 3, (unsigned long long)t_search); atomicAdd(prof + 4, (unsigned long long)t_enc); atomicAdd(prof + 5, (unsigned long long)use);
		}
	} else if (warp < kMmaWarp) {
		// ===== EPILOGUE ==========================================================================================================
		const int c = (warp - kEpilogueWarp0) >> 2, row = (warp & 3) * 32 + lane;
		const uint32_t lane_base = tmem + (uint32_t)c * kTmemChain + ((uint32_t)(row & ~31) << 16);
		unsigned char* h_buf = smem + kOffH + c * kRows * 64 * 2;
		unsigned char* r_buf = smem + kOffR + c * kRows * 32 * 2;
		Acc acc0 = {0xFFFFFFFFu, 0u, 0, 0, 0, 0, 0, 0}, acc1 = acc0;
		uint32_t par_mma = 0;
		long long t_idle = 0;
		const long long t_begin = clock64();
		auto layer_done = [&]() { tc::fence_async_smem(); tc::fence_before_sync(); mbar_arrive(bar(kBarEpi + c)); };
		auto await_mma = [&]() { const long long w0 = clock64(); mbar_wait_long(bar(kBarMma + c), par_mma); par_mma ^= 1u; tc::fence_after_sync(); t_idle += clock64() - w0; };
		for (;;) {
			await_mma();                                                                  // layer 1 of a new tile, or the exit signal
			const uint32_t info = s_info[c];
			if (info == kExit) break;
			const int g = (int)(info >> 1), stage = (int)(info & 1u);
			const RowMeta rec = *(reinterpret_cast<const RowMeta*>(smem + kOffMeta) + (g * kStages + stage) * kRows + row);
			uint4 sh0 = make_uint4(0, 0, 0, 0), sh1 = sh0;
			if (rec.flags & kValid) {                                                     // SH-4 block of the ray (encoded once, by ray generation)
				sh0 = __ldg(reinterpret_cast<const uint4*>(queue + rec.ray) + 2); sh1 = __ldg(reinterpret_cast<const uint4*>(queue + rec.ray) + 3);
			}
			tc::hidden_to_smem(lane_base, h_buf, row);                                    // density 32 -> 64, ReLU
			mbar_arrive(bar(kBarEmpty + g * kStages + stage));                            // features and row record consumed: the producers may refill the stage
			layer_done();
			await_mma();                                                                  // density 64 -> 16
			uint32_t sigma_bits;
			{
				uint32_t v[16];
				tc::tmem_ld16(lane_base + kColSmall, v);
				tc::tmem_wait_ld();
				uint32_t p[8];
#pragma unroll
				for (int i = 0; i < 8; ++i) p[i] = pack_relu_h2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]), false);
				sigma_bits = p[0] & 0xFFFFu;                                              // density output 0 = log-density (extract_density)
				*reinterpret_cast<uint4*>(r_buf + tc::chunk_offset<32>(row, 0)) = make_uint4(p[0], p[1], p[2], p[3]);
				*reinterpret_cast<uint4*>(r_buf + tc::chunk_offset<32>(row, 1)) = make_uint4(p[4], p[5], p[6], p[7]);
				*reinterpret_cast<uint4*>(r_buf + tc::chunk_offset<32>(row, 2)) = sh0;
				*reinterpret_cast<uint4*>(r_buf + tc::chunk_offset<32>(row, 3)) = sh1;
			}
			layer_done();
			await_mma();                                                                  // rgb [16 | 16 SH] -> 64, ReLU
			tc::hidden_to_smem(lane_base, h_buf, row);
			layer_done();
			await_mma();                                                                  // rgb 64 -> 64, ReLU
			tc::hidden_to_smem(lane_base, h_buf, row);
			layer_done();
			await_mma();                                                                  // rgb 64 -> 16 (3 live)
			uint32_t v[4];
			tc::tmem_ld4(lane_base + kColSmall, v);
			tc::tmem_wait_ld();
			tc::fence_before_sync();
			mbar_arrive(bar(kBarEpi + c));                                                // TMEM of this chain is free for the next tile
			// -- composite (composite_kernel_nerf(_alt)) ------------------------------------------------------------------------------
			if (rec.flags & (kValid | kEnd | kDrop)) {
				Acc& acc = (g & 1) ? acc1 : acc0;
				if (acc.ray != rec.ray) { acc.ray = rec.ray; acc.fin = 0u; acc.r = acc.g = acc.b = acc.a = 0.0f; acc.depth = 0.0f; acc.maxw = 0.0f; }
				if (!acc.fin) {
					bool finish = false, write = true;
					if (rec.flags & kValid) {
						const uint32_t bidx = (rec.flags >> 8) & 0xFFu;
						if (bidx && count_batches) atomicAdd(&s_hist[bidx - 1u], 1u);
						++n_samples;
						const float T = 1.f - acc.a;
						const float sigma = __half2float(__ushort_as_half((unsigned short)sigma_bits));
						const float alpha = 1.f - __expf(-act_density(sigma, model.density_act) * rec.dt);
						const float weight = alpha * T;
						const float cr = act_rgb(__half2float(__float2half_rn(__uint_as_float(v[0]))), model.rgb_act);
						const float cg = act_rgb(__half2float(__float2half_rn(__uint_as_float(v[1]))), model.rgb_act);
						const float cb = act_rgb(__half2float(__float2half_rn(__uint_as_float(v[2]))), model.rgb_act);
						acc.r += cr * weight; acc.g += cg * weight; acc.b += cb * weight; acc.a += weight;
						if (ALT) { acc.depth = rec.depth; if (weight > acc.maxw) acc.maxw = weight; }
						else if (weight > acc.maxw) { acc.maxw = weight; acc.depth = rec.depth; }
						if (acc.a > (1.0f - model.min_transmittance)) {
							const float inv_a = 1.0f / acc.a;
							acc.r *= inv_a; acc.g *= inv_a; acc.b *= inv_a; acc.a *= inv_a;
							finish = true; write = !(rec.flags & kNoWrite);
							s_done[g * kRows + row] = rec.ray;                                // tell the producer
						}
					} else { finish = true; write = (rec.flags & kEnd) != 0u; }
					if (finish) {
						acc.fin = 1u;
						if (write && acc.a > 0.001f) {
							const float4 q1 = __ldg(reinterpret_cast<const float4*>(queue + rec.ray) + 1);
							const v3 dir = mk3(q1.x, q1.y, q1.z);
							const float tt = ALT ? acc.depth / dot(cam_fwd, dir) : 0.0f;        // payload.t rewrite at the end of composite_kernel_nerf_alt
							write_pixel(view, out, ALT, render_mode, __float_as_uint(q1.w), dir, tt, make_float4(acc.r, acc.g, acc.b, acc.a), acc.depth);
							++n_hits;
						}
					}
				}
			}
		}
		if (prof && lane == 0) { atomicAdd(prof + 6, (unsigned long long)(clock64() - t_begin)); atomicAdd(prof + 7, (unsigned long long)t_idle); }
	} else if (lane == 0) {
		// ===== MMA ISSUER (one thread) ============================================================================================
		const uint32_t sW = tc::smem_u32(smem + kOffW);
		tc::mbar_expect_tx(bar(kBarW), kMlpHalves * 2);
		tc::tma_bulk_g2s(sW, model.mlp_umma, kMlpHalves * 2, bar(kBarW));
		tc::mbar_wait(bar(kBarW), 0);
		uint32_t used[kGroups] = {0, 0, 0, 0};
		bool gdone[kGroups] = {false, false, false, false};
		int layer[kChains] = {0, 0};                                                      // 0 = idle, l = layer l issued and waiting for its epilogue
		uint32_t par_epi[kChains] = {0, 0};
		int prefer[kChains] = {0, 0};
		for (;;) {
			bool finished = true;
#pragma unroll
			for (int c = 0; c < kChains; ++c) {
				const uint32_t d_hidden = tmem + (uint32_t)c * kTmemChain, d_small = d_hidden + kColSmall;
				const uint32_t sH = tc::smem_u32(smem + kOffH + c * kRows * 64 * 2), sR = tc::smem_u32(smem + kOffR + c * kRows * 32 * 2);
				if (layer[c] == 0) {
#pragma unroll
					for (int i = 0; i < 2; ++i) {
						const int g = 2 * c + ((prefer[c] + i) & 1);
						if (gdone[g] || layer[c] != 0) continue;
						const uint32_t stage = used[g] & 1u;
						if (!mbar_test(bar(kBarFull + g * kStages + stage), (used[g] >> 1) & 1u)) continue;
						if (s_hdr[g * kStages + stage] == kTerminate) { gdone[g] = true; continue; }
						tc::fence_after_sync();
						s_info[c] = (uint32_t)g * 2u + stage;
						__threadfence_block();
						tc::issue_layer<32, 64>(d_hidden, tc::smem_u32(smem + kOffF + (g * kStages + (int)stage) * kTileF), sW + kOffDW1 * 2, bar(kBarMma + c));
						++used[g]; layer[c] = 1; prefer[c] = (g & 1) ^ 1;
					}
					if (layer[c] != 0 || !(gdone[2 * c] && gdone[2 * c + 1])) finished = false;
				} else {
					finished = false;
					if (mbar_test(bar(kBarEpi + c), par_epi[c])) {
						par_epi[c] ^= 1u;
						tc::fence_after_sync();
						switch (layer[c]) {
							case 1: tc::issue_layer<64, 16>(d_small, sH, sW + kOffDW2 * 2, bar(kBarMma + c)); layer[c] = 2; break;
							case 2: tc::issue_layer<32, 64>(d_hidden, sR, sW + kOffRW1 * 2, bar(kBarMma + c)); layer[c] = 3; break;
							case 3: tc::issue_layer<64, 64>(d_hidden, sH, sW + kOffRW2 * 2, bar(kBarMma + c)); layer[c] = 4; break;
							case 4: tc::issue_layer<64, 16>(d_small, sH, sW + kOffRW3 * 2, bar(kBarMma + c)); layer[c] = 5; break;
							default: layer[c] = 0; break;                                   // the last epilogue has drained TMEM: chain idle
						}
					}
				}
			}
			if (finished) break;
		}
		for (int c = 0; c < kChains; ++c) { s_info[c] = kExit; __threadfence_block(); mbar_arrive(bar(kBarMma + c)); }
	}

	// ---- teardown ------------------------------------------------------------------------------------------------------------------
	tc::fence_before_sync();
	__syncthreads();
	if (warp == kMmaWarp) tc::tmem_dealloc(tmem, kTmemCols);
#pragma unroll
	for (int s = 16; s > 0; s >>= 1) {
		n_samples += __shfl_xor_sync(0xFFFFFFFFu, n_samples, s);
		n_hits += __shfl_xor_sync(0xFFFFFFFFu, n_hits, s);
		n_launched += __shfl_xor_sync(0xFFFFFFFFu, n_launched, s);
	}
	if (lane == 0 && counters) {
		if (n_samples) atomicAdd(&counters->samples, (unsigned long long)n_samples);
		if (n_hits) atomicAdd(&counters->hits, (unsigned long long)n_hits);
		if (n_launched) atomicAdd(&counters->rays, (unsigned long long)n_launched);
	}
	if (count_batches && tid < kAltSched && s_hist[tid]) atomicAdd(&alt.st->hist[tid], s_hist[tid]);
}

} // namespace snb
