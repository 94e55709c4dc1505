// synerfgine_b200 — NeRF layer kernels: ray generation + start jitter, the persistent fused
// march -> encode -> MLP -> composite kernel, and the parity taps.
//
// Replaces (reference, src/testbed_nerf.cu): init_rays_with_payload_kernel_nerf :1855-1970, advance_pos_nerf_kernel
// :334-384, compact_kernel_nerf :1830-1853, generate_next_nerf_network_inputs :790-837, composite_kernel_nerf
// :577-788 / composite_kernel_nerf_alt :476-575, extract_from_payload :1578-1612, shade_kernel_nerf :1788-1828 and
// the host loops NerfTracer::trace / trace_alt :2128-2401 (one D2H + stream sync per iteration there; none here).
//
// Design (DESIGN.md §"NeRF march"): rays are independent, so instead of a global wavefront with a compaction
// kernel and a host round trip per iteration, a persistent grid of warps each owns 32 ray slots. A slot whose ray
// finished is refilled from a global queue with one warp-aggregated atomicAdd (ballot + popc); every loop
// iteration evaluates one sample per live slot: DDA skip -> hash-grid gather -> both MLPs on tensor cores ->
// front-to-back composite in registers. Samples, features and network outputs never touch HBM; a finished ray
// writes its pixel once. The per-ray batch counter (8 steps, MAX_STEPS_INBETWEEN_COMPACTION) is kept because
// composite_kernel_nerf_alt rewinds payload.t at every batch boundary (src/testbed_nerf.cu:574).
#pragma once
#include "snb_camera.cuh"
#include "snb_network.cuh"

namespace snb {

constexpr int kBatchSteps = 8;       // n_steps_between_compaction for <= 262144 alive rays (src/testbed_nerf.cu:2189-2190)
#ifndef SNB_LANE_WARPS
#define SNB_LANE_WARPS 12      // warps per CTA of the lane march (tuning hook): 2 CTAs x 12 warps = 24 warps / SM at 80 registers, one 20 KB weight copy per CTA
#endif
constexpr int kFusedWarps = SNB_LANE_WARPS;
constexpr int kFusedThreads = kFusedWarps * 32;
constexpr int kTapWarps = 4, kTapThreads = kTapWarps * 32;   // parity tap kernel (k_network_inference): fixed, static shared memory

struct NerfCounters { unsigned long long samples, rays, hits, batches, iterations, dda_steps, net_phases, reserved; };

struct NerfOut {                     // tile-sized output buffers
	float4* fb;                      // rgba
	float* depth;
	float* positions;                // xyz per pixel (alt path) or nullptr
	// parity tap of the fused march (snb_set_sample_tap; all nullptr normally): per tile pixel the samples the kernel itself evaluates
	float* tap_t; uint32_t* tap_cell; uint32_t* tap_n; uint32_t tap_cap;
	// sng path, lane kernel: finished rays store the compositor's colour as it is and k_resolve_srgb converts the tile once the march is over
	// (extract_from_payload's srgb_to_linear: three powf per ray, 12 KB of code, otherwise run by one or two lanes of a warp inside the march loop)
	uint32_t raw_rgb;
};

// ---- batch schedule of the sng compositor ------------------------------------------------------------------------------------
// NerfTracer::trace_alt marches all alive rays in lock step, S = clamp(2^21 / n_alive, 1, 8) samples per iteration
// (src/testbed_nerf.cu:2189-2190), and composite_kernel_nerf_alt rewinds payload.t to the last composited sample after every
// iteration (:574), so the last sample of a batch is evaluated again as the first sample of the next: the image depends on
// the sequence S_0, S_1, ... Here rays run independently, so the sequence is an INPUT (kAltSched entries, 8 afterwards; entry 0
// is always derived from the queued-ray count, which is exact) and the kernel counts the rays that start each batch; k_alt_verify
// derives the sequence those counts imply and flags a mismatch, in which case the march is repeated with the corrected sequence
// (conditional launches, no host round trip). n_alive_j depends only on S_0..S_{j-1}, so every pass fixes at least one more
// entry; from one frame to the next the sequence rarely changes and one pass suffices.
constexpr int kAltSched = 64;
struct AltState {                     // device-resident, one per context
	uint32_t hist[kAltSched];         // rays that started batch j in the last pass (n_alive_j of the reference)
	uint8_t sched[kAltSched];         // batch lengths used by the next pass; 0 = not known yet (use S_0)
	uint32_t redo;                    // set by k_alt_verify: the pass just made used a sequence its own counts contradict
	uint32_t passes, converged, last_mismatch;
};
// Checkpointed prefix (option alt_schedule=lockstep; chosen by "auto" for frames whose camera or world moved). While the batch length is
// below 8 the reference's n_alive genuinely couples all rays. Instead of guessing the whole sequence and re-marching the frame when the
// guess was wrong, the prefix is marched in CHUNKS of kLockChunk batches per launch: every ray parks its compositor state at the end of
// the chunk, a small kernel checks the chunk's lengths against the counts it produced and either COMMITS the chunk (the parked states
// become the checkpoint) or corrects the first wrong length and has the next launch redo the chunk FROM THE CHECKPOINT — a wrong guess
// costs a few batches, not a frame. As soon as a committed length is 8 (n_alive <= target / 8, and it only falls) the rest of the march
// is one free-running launch from the checkpoint. Per ray the arithmetic is that of the speculative path, so frames are bit-identical.
constexpr int kLockChunk = 1;
struct RayPark { float4 rgba; float t, depth, max_weight; uint32_t step_i; uint32_t k, dead, pad0, pad1; };   // 48 B per queued ray, indexed by queue position
struct LockState {                     // device-resident, behind AltState
	uint32_t cur_batch;                // checkpoint: batches < cur_batch are committed
	uint32_t stop_batch;               // the attempt in flight covers [cur_batch, stop_batch)
	uint32_t done;                     // the committed length of batch cur_batch is 8 (or no ray is left): prefix finished
	uint32_t sel;                      // which list / park buffer holds the checkpoint
	uint32_t n_in, n_out;              // rays alive at the checkpoint / parked by the attempt in flight
	uint32_t attempts, redos;          // statistics
	unsigned long long pend_samples, pend_hits, pend_reused;   // counters of the attempt in flight (added to NerfCounters on commit)
};
struct AltCtl {
	AltState* st;                     // nullptr: stock path / fixed batch length
	uint32_t target_queries;          // 2 * 1024 * 1024 in the reference
	uint32_t fixed_steps;             // != 0: constant batch length (option alt_batch), no counting
	uint32_t conditional;             // run only if st->redo is set (second and third pass)
	uint32_t external;                // the caller owns the sequence (a tile of a larger frame): entry 0 is taken from it too
	// lock-step prefix (PH instances of the lane kernel only)
	LockState* lk;
	RayPark* park[2];                 // ping-pong: the checkpoint's states / the states the attempt in flight parks
	uint32_t* lists[2];               // ping-pong lists of queue positions of the rays still alive
	uint32_t phase;                   // 0 = not phased, 1 = prefix attempt (one chunk, then park), 2 = final launch (run to completion)
	uint32_t reuse_dup;               // re-use the network output for the sample the rewind makes the compositor evaluate twice (option dup_reuse)
};
SNB_HD uint32_t alt_steps_for(uint32_t target, uint32_t n_alive) {
	uint32_t s = n_alive ? target / n_alive : 8u;
	return s < 1u ? 1u : (s > 8u ? 8u : s);
}
__global__ void k_alt_verify(AltState* st, uint32_t target, const unsigned int* __restrict__ queue_tail, uint32_t last_pass, uint32_t external) {
	__shared__ uint32_t s_bad, s_used[kAltSched], s_n[kAltSched], s_want[kAltSched];
	const int j = threadIdx.x;
	if (j == 0) s_bad = 0xFFFFFFFFu;
	__syncthreads();
	if (st->passes != 0 && !st->redo) return;         // a conditional pass that did not run: nothing new to verify (uniform)
	const uint32_t n = st->hist[j];
	uint32_t used = st->sched[j];
	if ((j == 0 && !external) || used == 0) used = alt_steps_for(target, *queue_tail);
	const uint32_t want = n ? alt_steps_for(target, n) : 8u;
	s_used[j] = used; s_n[j] = n; s_want[j] = want;
	if (n && want != used) atomicMin(&s_bad, (uint32_t)j);
	if (j == kAltSched - 1 && n > target / 8u) atomicMin(&s_bad, (uint32_t)kAltSched);   // still below 8 after the last tracked batch: inexact beyond
	__syncthreads();
	if (j != 0) return;
	const uint32_t bad = s_bad;
	const bool ok = bad == 0xFFFFFFFFu;
	if (!external) {   // external: the counts of one tile say nothing about the frame; the caller sums them
		// Entries up to the first mismatch are exact (n_alive_j depends on S_0..S_{j-1} only). Behind it the counts were taken under a
		// sequence that is now known to be wrong, so the guess for the next pass is not want[j] but a replay: n_alive as a function of
		// the samples composited so far (piecewise linear through the counts of this pass) is close to independent of how the samples
		// are cut into batches, and walking it with the corrected lengths predicts where the later thresholds fall. The next
		// pass is verified like this one; the replay only decides how many passes that takes.
		uint32_t c = 0, k = 0, c_k = 0;               // samples composited at the start of batch i (new sequence) / batch k (this pass)
		for (uint32_t i = 0; i < (uint32_t)kAltSched; ++i) {
			uint32_t steps = s_want[i];
			if (!ok && i > bad) {
				while (k + 1 < (uint32_t)kAltSched && c_k + s_used[k] <= c) { c_k += s_used[k]; ++k; }
				const uint32_t n0 = s_n[k], n1 = (k + 1 < (uint32_t)kAltSched) ? s_n[k + 1] : s_n[k];
				const uint32_t frac_num = c > c_k ? c - c_k : 0u, frac_den = s_used[k] ? s_used[k] : 1u;
				const uint32_t a = n0 - (uint32_t)(((unsigned long long)(n0 > n1 ? n0 - n1 : 0u) * (frac_num < frac_den ? frac_num : frac_den)) / frac_den);
				steps = a ? alt_steps_for(target, a) : 8u;
			}
			st->sched[i] = (uint8_t)steps;
			c += steps;
		}
	}
	st->passes += 1; st->converged = ok ? 1u : 0u; st->last_mismatch = bad;
	st->redo = (!ok && !last_pass && bad < (uint32_t)kAltSched) ? 1u : 0u;
}
// Checkpointed prefix, start of a frame (after ray generation). Batch 0's length is exact (it follows from the ray count); the other
// entries keep the last frame's values as guesses (8 where there is none). When batch 0 is already 8 there is no prefix.
__global__ void k_lock_begin(AltState* st, LockState* lk, uint32_t target, const unsigned int* __restrict__ queue_tail, unsigned int* __restrict__ queue_head,
                             NerfCounters* __restrict__ counters) {
	const uint32_t j = threadIdx.x;
	const uint32_t n = *queue_tail;
	const uint32_t s0 = alt_steps_for(target, n);
	if (j < (uint32_t)kAltSched) {
		st->hist[j] = j == 0 ? n : 0u;
		if (j == 0) st->sched[0] = (uint8_t)s0;
		else if (st->sched[j] < s0) st->sched[j] = (uint8_t)s0;   // no guess yet (0): assume the length stays; lengths never fall along a march
	}
	if (j == 0) {
		lk->cur_batch = 0; lk->stop_batch = (uint32_t)kLockChunk; lk->sel = 0; lk->n_in = n; lk->n_out = 0; lk->attempts = 0; lk->redos = 0;
		lk->pend_samples = 0ull; lk->pend_hits = 0ull; lk->pend_reused = 0ull;
		lk->done = (s0 >= 8u || n == 0u) ? 1u : 0u;
		*queue_head = 0u; st->passes = 0; st->redo = 0; st->converged = 0;
		if (counters) { counters->samples = 0; counters->rays = n; counters->hits = 0; counters->batches = 0; }
	}
}
// After a prefix attempt over [b, stop): hist[j], b < j < stop, are the rays that started batch j in this attempt and n_out the rays
// parked at `stop`; each is the reference's n_alive_j provided every earlier length of the attempt was right.
__global__ void k_lock_next(AltState* st, LockState* lk, uint32_t target, unsigned int* __restrict__ queue_head, NerfCounters* __restrict__ counters) {
	if (lk->done) return;
	const uint32_t b = lk->cur_batch, stop = lk->stop_batch;
	lk->attempts += 1;
	uint32_t bad = 0xFFFFFFFFu;
	for (uint32_t j = b + 1; j < stop; ++j) {
		const uint32_t n = st->hist[j];
		const uint32_t want = n ? alt_steps_for(target, n) : 8u;
		if (want != st->sched[j]) { bad = j; st->sched[j] = (uint8_t)want; break; }
	}
	if (bad == 0xFFFFFFFFu) {
		// commit: the parked states are the new checkpoint
		const uint32_t n = lk->n_out;
		const uint32_t steps = n ? alt_steps_for(target, n) : 8u;
		if (stop < (uint32_t)kAltSched) { st->hist[stop] = n; st->sched[stop] = (uint8_t)steps; }
		for (uint32_t j = stop + 1; j < (uint32_t)kAltSched; ++j) if (st->sched[j] < steps) st->sched[j] = (uint8_t)steps;   // guesses stay monotone
		lk->cur_batch = stop; lk->sel ^= 1u; lk->n_in = n;
		lk->stop_batch = stop + (uint32_t)kLockChunk < (uint32_t)kAltSched ? stop + (uint32_t)kLockChunk : (uint32_t)kAltSched;
		if (counters) { counters->samples += lk->pend_samples; counters->hits += lk->pend_hits; counters->batches += lk->pend_reused; }
		if (steps >= 8u || n == 0u || stop >= (uint32_t)kAltSched) lk->done = 1u;
	} else {
		// redo the chunk from the checkpoint with the corrected length (every entry before `bad` is exact now: progress is guaranteed)
		for (uint32_t j = bad + 1; j < (uint32_t)kAltSched; ++j) if (st->sched[j] < st->sched[bad]) st->sched[j] = st->sched[bad];
		for (uint32_t j = b + 1; j < (uint32_t)kAltSched; ++j) st->hist[j] = 0u;
		lk->redos += 1;
	}
	lk->n_out = 0; lk->pend_samples = 0ull; lk->pend_hits = 0ull; lk->pend_reused = 0ull;
	*queue_head = 0u;
}

// Start of a march pass: the first pass always runs and resets the per-frame state; later passes run only after a mismatch and
// first undo what the previous pass wrote (tile buffers are cleared like ray generation clears them).
__global__ void __launch_bounds__(256) k_alt_begin_pass(AltState* st, uint32_t first_pass, NerfOut out, uint32_t n_px, unsigned int* __restrict__ queue_head,
                                                        NerfCounters* __restrict__ counters) {
	if (!first_pass && !st->redo) return;
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i == 0) { *queue_head = 0u; if (first_pass) { st->passes = 0; st->redo = 0; st->converged = 0; } if (counters) { counters->samples = 0; counters->rays = 0; counters->hits = 0; counters->batches = 0; } }
	if (i < (uint32_t)kAltSched) st->hist[i] = 0u;
	if (first_pass) return;
	for (uint32_t p = i; p < n_px; p += gridDim.x * blockDim.x) {
		out.fb[p] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
		out.depth[p] = kMaxDepth;
		if (out.positions) { out.positions[p * 3 + 0] = 0.0f; out.positions[p * 3 + 1] = 0.0f; out.positions[p * 3 + 2] = 0.0f; }
	}
}

SNB_D float act_logistic(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }
SNB_D float act_rgb(float v, int act) {
	switch (act) {
		case 0: return v;
		case 1: return v > 0.0f ? v : 0.0f;
		case 2: return act_logistic(v);
		default: return __expf(fminf(fmaxf(v, -10.0f), 10.0f));
	}
}
SNB_D float act_density(float v, int act) {
	switch (act) {
		case 0: return v;
		case 1: return v > 0.0f ? v : 0.0f;
		case 2: return act_logistic(v);
		default: return __expf(v);
	}
}
SNB_HD float dt_warp(float dt) { const float mx = kMinStep * (1 << 7); return (dt - kMinStep) / (mx - kMinStep); }
SNB_HD float dt_unwarp(float dt) { const float mx = kMinStep * (1 << 7); return dt * (mx - kMinStep) + kMinStep; }

SNB_D int tile_local_index(const DevView& v, uint32_t gidx) {
	const int x = (int)(gidx % (uint32_t)v.width) - v.tx0, y = (int)(gidx / (uint32_t)v.width) - v.ty0;
	return x + v.tw * y;
}

// ---- per-level grid scale (once per model) -------------------------------------------------------------------------------------
// tcnn's kernel_grid evaluates grid_scale(level, log2_per_level_scale, base_resolution) = exp2f(level * log2_pls) * base - 1 on the
// device in every thread (common_device.h:841-845): exp2f is MUFU.EX2 there and nvcc contracts the multiply-subtract into one FFMA.
// The host's libm value differs in the last place on most levels, which moves ~17 % of the fp16 features by an ulp, so the scales
// are evaluated here once, by the same two operations, and kept in the level table.
__global__ void k_grid_scales(uint32_t n_levels, float log2_per_level_scale, uint32_t base_resolution, float* __restrict__ out) {
	const uint32_t level = threadIdx.x;
	if (level < n_levels) out[level] = fmaf(exp2f((float)level * log2_per_level_scale), (float)base_resolution, -1.0f);
}

// ---- coarse look-ahead grid (once per model) ------------------------------------------------------------------------
// One thread per bitfield byte of cascades 0..max_mip; every set bit marks the coarse cells its world box touches,
// dilated by one coarse cell so that a DDA grazing a cell corner can never miss it.
__global__ void __launch_bounds__(128) k_build_coarse(const uint8_t* __restrict__ bitfield, uint32_t n_bytes, v3 coarse_lo, float inv_cell, uint32_t* __restrict__ coarse) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_bytes) return;
	const uint8_t byte = bitfield[i];
	if (!byte) return;
	const uint32_t mip = i / (kGridCells / 8);
	const float cell = scalbnf(1.0f / (float)kGridSize, (int)mip), half = scalbnf(0.5f, (int)mip);
	for (int j = 0; j < 8; ++j) {
		if (!(byte & (1u << j))) continue;
		const uint32_t m = (i % (kGridCells / 8)) * 8 + j;
		uint32_t xyz[3];
		for (int a = 0; a < 3; ++a) {
			uint32_t x = (m >> a) & 0x49249249u;
			x = (x | (x >> 2)) & 0xc30c30c3u; x = (x | (x >> 4)) & 0x0f00f00fu; x = (x | (x >> 8)) & 0xff0000ffu; x = (x | (x >> 16)) & 0x0000ffffu;
			xyz[a] = x;
		}
		int lo[3], hi[3];
		const float base[3] = {coarse_lo.x, coarse_lo.y, coarse_lo.z};
		for (int a = 0; a < 3; ++a) {
			const float wlo = (float)xyz[a] * cell - half + 0.5f, whi = wlo + cell;
			lo[a] = max(0, (int)floorf((wlo - base[a]) * inv_cell) - 1);
			hi[a] = min(kCoarseRes - 1, (int)floorf((whi - base[a]) * inv_cell) + 1);
		}
		for (int z = lo[2]; z <= hi[2]; ++z) for (int y = lo[1]; y <= hi[1]; ++y) for (int x = lo[0]; x <= hi[0]; ++x) {
			const uint32_t bit = (uint32_t)x + (uint32_t)kCoarseRes * ((uint32_t)y + (uint32_t)kCoarseRes * (uint32_t)z);
			atomicOr(&coarse[bit >> 5], 1u << (bit & 31));
		}
	}
}

// ---- linear-order copy of the occupancy bitfield (once per model) ---------------------------------------------------------
// One thread per output byte = 8 x-consecutive cells of one cascade; bit (x, y, z) of cascade m comes from Morton bit
// morton3(x, y, z) of the reference-layout bitfield. Read by cell_occupied_xyz (snb_march.cuh).
__global__ void __launch_bounds__(128) k_bitfield_to_linear(const uint8_t* __restrict__ morton_bits, uint32_t n_bytes, uint8_t* __restrict__ bits_xyz) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_bytes) return;
	const uint32_t mip = i / (kGridCells / 8), cell0 = (i % (kGridCells / 8)) * 8;
	const uint32_t x0 = cell0 & 127u, y = (cell0 >> 7) & 127u, z = cell0 >> 14;
	uint32_t out = 0;
	for (uint32_t j = 0; j < 8; ++j) {
		const uint32_t m = morton3(x0 + j, y, z);
		out |= ((uint32_t)(morton_bits[mip * (kGridCells / 8) + (m >> 3)] >> (m & 7u)) & 1u) << j;
	}
	bits_xyz[i] = (uint8_t)out;
}

// ---- ray generation: one thread per tile pixel, 8x4 pixel patches per warp so queue neighbours are image neighbours ----
// Patches [patch_begin, patch_end) of the tile (row-major patch order): one launch per band of the frame.
template <bool EX>
__global__ void __launch_bounds__(128) k_nerf_raygen(DevView view, MarchEnv env, NerfOut out, RayRec* __restrict__ queue, unsigned int* __restrict__ queue_tail,
                                                     int patch_begin, int patch_end) {
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int patches_x = (view.tw + 7) / 8;
	const int patch = patch_begin + blockIdx.x * 4 + warp;
	if (patch >= patch_end) return;                            // whole warp
	const int px = (patch % patches_x) * 8 + (lane & 7), py = (patch / patches_x) * 4 + (lane >> 3);
	bool alive = false;
	RayRec rec;
	if (px < view.tw && py < view.th) {
		const int x = view.tx0 + px, y = view.ty0 + py;
		const int lidx = px + view.tw * py;
		out.fb[lidx] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
		out.depth[lidx] = kMaxDepth;
		if (out.positions) { out.positions[lidx * 3 + 0] = 0.0f; out.positions[lidx * 3 + 1] = 0.0f; out.positions[lidx * 3 + 2] = 0.0f; }
		v3 o, d; float t;
		bool valid = true;
		alive = nerf_primary_ray<EX>(x, y, view, env, o, d, t, &valid);
		const uint32_t idx = (uint32_t)x + (uint32_t)view.width * (uint32_t)y;
		if (EX && view.ex.enabled) {
			// init_rays_with_payload_kernel_nerf, src/testbed_nerf.cu:1918-1950: a negative plane_z (slice plane) kills the ray and writes the
			// plane depth; the environment map is the background of every valid ray, whether it meets the render box or not.
			if (valid && view.ex.plane_z < 0.0f) { alive = false; out.depth[lidx] = -view.ex.plane_z; }
			else if (valid && view.ex.envmap) out.fb[lidx] = envmap_lookup(view.ex, d);
		}
		if (alive) {
			// start jitter + skip to the first occupied cell (advance_pos_nerf)
			const v3 idir = 1.0f / d;
			t = advance_steps(t, env.st, ld_value(view.spp, idx * 786433u));
			t = skip_to_occupied(t, o, d, idir, env);
			alive = t < kMaxDepth;
		}
		rec.ox = o.x; rec.oy = o.y; rec.oz = o.z; rec.t = t; rec.dx = d.x; rec.dy = d.y; rec.dz = d.z; rec.idx = idx;
	}
	const unsigned int mask = __ballot_sync(0xFFFFFFFFu, alive);
	if (mask == 0) return;
	unsigned int base = 0;
	if (lane == 0) base = atomicAdd(queue_tail, __popc(mask));
	base = __shfl_sync(0xFFFFFFFFu, base, 0);
	if (alive) {
		float4* q = reinterpret_cast<float4*>(queue + base + __popc(mask & ((1u << lane) - 1u)));
		q[0] = make_float4(rec.ox, rec.oy, rec.oz, rec.t);
		q[1] = make_float4(rec.dx, rec.dy, rec.dz, __uint_as_float(rec.idx));
		uint32_t sh[8];
		encode_direction_sh4((mk3(rec.dx, rec.dy, rec.dz) + 1.0f) * 0.5f, sh);   // warp_direction, nerf_device.cuh:290-292
		reinterpret_cast<uint4*>(q)[2] = make_uint4(sh[0], sh[1], sh[2], sh[3]);
		reinterpret_cast<uint4*>(q)[3] = make_uint4(sh[4], sh[5], sh[6], sh[7]);
	}
}

// Finished ray -> pixel (extract_from_payload for the sng path, shade_kernel_nerf for the stock path). Runs once per
// ray, so it is kept out of line: the march loop has to stay small (instruction cache).
// `rewrite`: the pixel may hold the result of an abandoned prefix attempt (checkpointed prefix): every field is written, those a clean
// frame would leave untouched get the values ray generation cleared them to.
__device__ __noinline__ void write_pixel(const DevView& view, const NerfOut& out, bool alt, int render_mode, uint32_t idx, v3 dir, float t, float4 rgba, float depth,
                                         bool rewrite = false) {
	const int l = tile_local_index(view, idx);
	float4 tmp = rgba;
	if (alt) {
		const v3 orig_pos = view.cam.c3 + dir * t;
		if (render_mode == 4) { const float v = (float)fmax(0.0, 1.0 - (double)(depth / 3.0f)); tmp.x = tmp.y = tmp.z = v; }
		else if (render_mode == 3) { tmp.x = srgb_to_linear(orig_pos.x * 0.5f + 0.5f); tmp.y = srgb_to_linear(orig_pos.y * 0.5f + 0.5f); tmp.z = srgb_to_linear(orig_pos.z * 0.5f + 0.5f); }
		else { tmp.x = srgb_to_linear(tmp.x); tmp.y = srgb_to_linear(tmp.y); tmp.z = srgb_to_linear(tmp.z); }
		if (out.positions) { out.positions[l * 3 + 0] = orig_pos.x; out.positions[l * 3 + 1] = orig_pos.y; out.positions[l * 3 + 2] = orig_pos.z; }
	} else if (render_mode == 1) {
		tmp.x = srgb_to_linear(tmp.x); tmp.y = srgb_to_linear(tmp.y); tmp.z = srgb_to_linear(tmp.z);
	}
	if (!alt && view.ex.enabled && view.ex.envmap) {
		// shade_kernel_nerf, src/testbed_nerf.cu:1824: frame_buffer = tmp + frame_buffer * (1 - tmp.a) over the environment map
		const float4 bg = out.fb[l]; const float k = 1.0f - tmp.w;
		out.fb[l] = make_float4(tmp.x + bg.x * k, tmp.y + bg.y * k, tmp.z + bg.z * k, tmp.w + bg.w * k);
	} else out.fb[l] = tmp;   // frame buffer was cleared to 0 by ray generation: tmp + 0 * (1 - a)
	if (tmp.w > 0.2f) out.depth[l] = depth;
	else if (rewrite) out.depth[l] = kMaxDepth;
}
// write_pixel of the sng path without the colour conversion (NerfOut::raw_rgb): the same stores, rgb left in sRGB for k_resolve_srgb.
SNB_D void write_pixel_raw(const DevView& view, const NerfOut& out, uint32_t idx, v3 dir, float t, float4 rgba, float depth, bool rewrite) {
	const int l = tile_local_index(view, idx);
	const v3 orig_pos = view.cam.c3 + dir * t;
	if (out.positions) { out.positions[l * 3 + 0] = orig_pos.x; out.positions[l * 3 + 1] = orig_pos.y; out.positions[l * 3 + 2] = orig_pos.z; }
	out.fb[l] = rgba;
	if (rgba.w > 0.2f) out.depth[l] = depth;
	else if (rewrite) out.depth[l] = kMaxDepth;
}
// The conversion write_pixel_raw left out, over the whole tile: pixels no ray wrote hold 0, which srgb_to_linear maps to 0.
__global__ void __launch_bounds__(256) k_resolve_srgb(float4* __restrict__ fb, uint32_t n_px) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_px) return;
	float4 v = fb[i];
	if (v.w == 0.0f) return;                                // never written (or cleared): rgb is 0 too
	v.x = srgb_to_linear(v.x); v.y = srgb_to_linear(v.y); v.z = srgb_to_linear(v.z);
	fb[i] = v;
}
__device__ __noinline__ void clear_pixel(const DevView& view, const NerfOut& out, uint32_t idx) {
	const int l = tile_local_index(view, idx);
	out.fb[l] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
	out.depth[l] = kMaxDepth;
	if (out.positions) { out.positions[l * 3 + 0] = 0.0f; out.positions[l * 3 + 1] = 0.0f; out.positions[l * 3 + 2] = 0.0f; }
}

// Parity tap (snb_set_sample_tap): sample number `ord` of pixel idx as the march kernel evaluates it. Out of line: the hot loop keeps its registers.
__device__ __noinline__ void tap_sample(const NerfOut& out, const DevView& view, uint32_t idx, uint32_t ord, v3 sp, float t, uint32_t min_mip, uint32_t max_mip) {
	const int l = tile_local_index(view, idx);
	if (ord < out.tap_cap) {
		uint32_t mip = mip_of(sp);
		if (mip < min_mip) mip = min_mip;
		if (mip > max_mip) mip = max_mip;
		out.tap_t[(size_t)l * out.tap_cap + ord] = t;
		out.tap_cell[(size_t)l * out.tap_cap + ord] = (mip << 24) | (cascade_cell(sp, mip) & 0xFFFFFFu);
	}
	out.tap_n[l] = ord + 1u;
}

// ---- the persistent fused kernel -----------------------------------------------------------------------------------
// grid = SMs x SNB_LANE_CTAS (2), block = SNB_LANE_WARPS (12) warps; dynamic smem (kLaneSmemBytes, 70 KB) = fragment-ordered
// MLP weights (20 KB, shared by the CTA's warps) + per-warp staging tile (3.5 KB) + per-warp output tile (256 B) + level
// table (1 KB) + look-ahead grid (4 KB). Warps never synchronise after the prologue; the CTA shape only sets how many warps
// share one weight copy and thereby how much of the SM's 256 KB is left as L1 for the gathers (profiles/r01_ncu_fused_v3.md).
//
// One loop iteration of a warp = (1) refill finished slots from the queue, (2) search phase: lanes whose ray sits in
// empty space take DDA steps until at least kMinFill lanes hold a sample position (lanes that already have one
// wait), (3) encode + both MLPs for the lanes that have a sample, (4) composite. A lane still searching after (2)
// simply carries on in the next iteration, so one ray crossing a long empty gap no longer stalls the other 31.
#ifndef SNB_LANE_CTAS
#define SNB_LANE_CTAS 2     // resident CTAs per SM the register budget is set for (tuning hook)
#endif
#ifndef SNB_LANE_UNROLLED_MLP
#define SNB_LANE_UNROLLED_MLP 0
#endif
#ifndef SNB_LANE_MIN_FILL
#define SNB_LANE_MIN_FILL 20   // lanes holding a sample before the network runs (tuning hook)
#endif
constexpr int kMinFill = SNB_LANE_MIN_FILL;
// W = warps per CTA. Two shapes are built: kFusedWarps (12 warps, 80 registers: 24 warps / SM — the faster one when every lane
// has several rays to work through, i.e. a full 1080p tile on one GPU) and kFusedWarpsFew (10 warps, 96 registers, no spills —
// the faster one when a tile holds about one ray per lane and the march lasts as long as its longest rays: the N >= 2 strips).
constexpr int kFusedWarpsFew = 10;
constexpr int lane_smem_bytes(int w) { return kFragWords * 4 + w * kWarpSmemBytes + (int)sizeof(LevelDesc) * kMaxLevels + kCoarseWords * 4; }
constexpr int kLaneSmemBytes = lane_smem_bytes(kFusedWarps);
// PH = lock-step prefix instance (see LockState): rays resume from / park into RayPark records at batch boundaries.
template <int F, bool ALT, int W, bool PH = false>
__global__ void __launch_bounds__(W * 32, SNB_LANE_CTAS)
k_nerf_march_fused(DevModel model, DevView view, NerfOut out, int render_mode, const RayRec* __restrict__ queue, const unsigned int* __restrict__ queue_tail,
                   unsigned int* __restrict__ queue_head, NerfCounters* __restrict__ counters, AltCtl alt) {
	if (alt.conditional && !alt.st->redo) return;   // second / third pass of the batch-schedule iteration: only after a mismatch (uniform)
	if (PH && alt.phase == 1 && alt.lk->done) return;   // prefix launch enqueued for a batch the frame did not need (uniform)
	// dynamic shared memory (kLaneSmemBytes): [fragment-ordered weights | per-warp staging + output tiles | level table | look-ahead grid]
	extern __shared__ __align__(16) unsigned char s_lane[];
	uint32_t* s_w = reinterpret_cast<uint32_t*>(s_lane);
	unsigned char (*s_warp)[kWarpSmemBytes] = reinterpret_cast<unsigned char (*)[kWarpSmemBytes]>(s_lane + kFragWords * 4);
	LevelDesc* s_levels = reinterpret_cast<LevelDesc*>(s_lane + kFragWords * 4 + W * kWarpSmemBytes);
	uint32_t* s_coarse = reinterpret_cast<uint32_t*>(s_lane + kFragWords * 4 + W * kWarpSmemBytes + sizeof(LevelDesc) * kMaxLevels);   // look-ahead occupancy (4 KB)
	if (model.env.coarse) for (int i = threadIdx.x; i < kCoarseWords; i += W * 32) s_coarse[i] = __ldg(model.env.coarse + i);
	for (int i = threadIdx.x; i < kFragWords / 4; i += W * 32)
		reinterpret_cast<uint4*>(s_w)[i] = __ldg(reinterpret_cast<const uint4*>(model.mlp_frag) + i);
	if (threadIdx.x < kFeat / F) s_levels[threadIdx.x] = make_level_desc(model.grid, threadIdx.x, model.grid_params);
	// batch schedule of the sng compositor (see AltState): lengths of batches 0..kAltSched-1 (8 afterwards) and this CTA's share of the counts
	__shared__ uint8_t s_sched[kAltSched];
	__shared__ uint32_t s_hist[kAltSched];
	const unsigned int n_rays = PH ? alt.lk->n_in : *queue_tail;
	const uint32_t batch0 = PH ? alt.lk->cur_batch : 0u;                     // PH: the batch every ray of this launch starts with
	const bool resume = PH && batch0 > 0u;
	const uint32_t stop_batch = (PH && alt.phase == 1) ? alt.lk->stop_batch : 0xFFFFFFFFu;
	const uint32_t sel = PH ? alt.lk->sel : 0u;
	const uint32_t* __restrict__ list_in = PH ? alt.lists[sel] : nullptr;
	uint32_t* __restrict__ list_out = PH ? alt.lists[sel ^ 1u] : nullptr;
	const RayPark* __restrict__ park_in = PH ? alt.park[sel] : nullptr;
	RayPark* __restrict__ park_out = PH ? alt.park[sel ^ 1u] : nullptr;
	if (threadIdx.x < kAltSched) {
		uint32_t steps = kBatchSteps;
		if (ALT && alt.fixed_steps) steps = alt.fixed_steps;
		else if (PH) steps = alt.st->sched[threadIdx.x];                      // exact up to batch0, guesses beyond (k_lock_begin / k_lock_next)
		else if (ALT && alt.st) {
			const uint32_t s0 = alt_steps_for(alt.target_queries, n_rays);
			steps = (threadIdx.x == 0 && !alt.external) ? s0 : alt.st->sched[threadIdx.x];
			if (steps == 0) steps = s0;
		}
		s_sched[threadIdx.x] = (uint8_t)steps;
		s_hist[threadIdx.x] = 0u;
	}
	__syncthreads();
	const bool count_batches = ALT && alt.st && !alt.fixed_steps;

	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	__half* tile = reinterpret_cast<__half*>(s_warp[warp]);
	__half* out_tile = reinterpret_cast<__half*>(s_warp[warp] + kWarpTileBytes);
	MarchEnv env = model.env;
	if (env.coarse) env.coarse = s_coarse;
	const v3 cam_fwd = view.cam.c2;
	const v3 aabb_diag = model.train_aabb.diag();

	// per-slot ray state
	v3 o = mk3(0, 0, 0), d = mk3(1, 0, 0), idir = mk3(1, 1, 1);
	float t = 0.0f, depth = 0.0f;
	float4 rgba = make_float4(0, 0, 0, 0);
	uint32_t idx = 0, step_i = 1;
	uint32_t qpos = 0;   // PH: queue position of the slot's ray (its RayPark record)
	int k = 0, n_empty = 0;
	uint32_t batch = 0, batch_steps = s_sched[0];   // index and length of the ray's current batch
	bool alive = false, drained = false;
	float max_weight = 0.0f;
	{ // zero the whole row once so never-filled rows hold finite values
		uint4* row = reinterpret_cast<uint4*>(tile + lane * kTileStrideH);
#pragma unroll
		for (int i = 0; i < kTileStrideH * 2 / 16; ++i) row[i] = make_uint4(0, 0, 0, 0);
	}
	uint32_t n_samples = 0, n_hits = 0, n_launched = 0, n_reused = 0;

	auto finish = [&]() {
		alive = false;
		if (step_i + batch_steps >= kMarchIter) return;   // died in the last possible iteration: never compacted into the hit list
		if (ALT) t = depth / dot(cam_fwd, d);             // payload.t rewrite at the end of composite_kernel_nerf_alt
		if (rgba.w > 0.001f) {
			if (ALT && out.raw_rgb) write_pixel_raw(view, out, idx, d, t, rgba, depth, PH);
			else write_pixel(view, out, ALT, render_mode, idx, d, t, rgba, depth, PH);
			++n_hits;
		}
		else if (PH) clear_pixel(view, out, idx);   // an abandoned attempt may have written this pixel: back to the cleared state
	};

	bool fin = false;   // ray ended in the previous iteration: one finish() site keeps the loop small
	for (;;) {
		if (fin) { finish(); fin = false; }
		// -- refill empty slots: one atomic per warp -----------------------------------------------------------------
		const unsigned int need = __ballot_sync(0xFFFFFFFFu, !alive);
		if (need && !drained) {
			unsigned int base = 0;
			if (lane == 0) base = atomicAdd(queue_head, __popc(need));
			base = __shfl_sync(0xFFFFFFFFu, base, 0);
			const unsigned int mine = base + __popc(need & ((1u << lane) - 1u));
			if (!alive && mine < n_rays) {
				const unsigned int qp = resume ? list_in[mine] : mine;
				const float4 q0 = __ldg(reinterpret_cast<const float4*>(queue + qp));
				const float4 q1 = __ldg(reinterpret_cast<const float4*>(queue + qp) + 1);
				o = mk3(q0.x, q0.y, q0.z); t = q0.w; d = mk3(q1.x, q1.y, q1.z); idx = __float_as_uint(q1.w);
				idir = 1.0f / d;
				rgba = make_float4(0, 0, 0, 0); depth = 0.0f; max_weight = 0.0f; step_i = 1; k = 0; n_empty = 0; alive = true;
				batch = 0; batch_steps = s_sched[0];
				if (PH) {
					qpos = qp;
					if (resume) {   // compositor state as it was at the end of batch batch0 - 1
						const RayPark pk = park_in[qp];
						rgba = pk.rgba; t = pk.t; depth = pk.depth; max_weight = pk.max_weight; step_i = pk.step_i; k = (int)pk.k;
						batch = batch0; batch_steps = batch0 < (uint32_t)kAltSched ? s_sched[batch0] : (uint32_t)kBatchSteps;
						if (pk.dead) alive = false;   // counted alive at the start of this batch, ended while its re-composited sample was folded in
					}
				} else {
					if (count_batches) atomicAdd(&s_hist[0], 1u);
					++n_launched;
				}
				// SH block of the staged row (k-tile 2): encoded by ray generation, copied as is
				const uint4 s0 = __ldg(reinterpret_cast<const uint4*>(queue + qp) + 2), s1 = __ldg(reinterpret_cast<const uint4*>(queue + qp) + 3);
				uint4* dst = reinterpret_cast<uint4*>(tile + lane * kTileStrideH + kTileShCol);
				dst[0] = s0; dst[1] = s1;
			}
			if (base + __popc(need) >= n_rays) drained = true;
		}
		if (__ballot_sync(0xFFFFFFFFu, alive) == 0) break;

		// -- search phase: the skip loop of if_unoccupied_advance_to_next_occupied_voxel, one DDA step per pass -----------
		bool valid = false;
		{
			bool searching = alive;
#pragma unroll 1
			for (;;) {
				if (searching) {
					const v3 pos = o + d * t;
					if (t >= kMaxDepth || !in_render_box(pos, env)) { fin = true; searching = false; }   // left the volume
					else {
						uint32_t mip = clamp_mip(mip_of(pos), env);
						if (!env.bitfield || cell_occupied_xyz(pos, env.bits_xyz, mip)) { valid = true; searching = false; n_empty = 0; }
						else if (((n_empty++) & (kLookAheadEvery - 1)) == 0 && coarse_clear_ahead_call(pos.x, pos.y, pos.z, d.x, d.y, d.z, env.coarse, env.coarse_lo.x, env.coarse_lo.y, env.coarse_lo.z, env.coarse_inv_cell)) { fin = true; searching = false; }
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
		if (__ballot_sync(0xFFFFFFFFu, valid) == 0) continue;

		// -- sample position -------------------------------------------------------------------------------------------
		v3 wpos = mk3(0, 0, 0);
		float wdt = 0.0f;
		if (valid) {
			const float dt = step_size_call(t, env.st);
			wpos = (o + d * t - model.train_aabb.lo) / aabb_diag;   // warp_position
			wdt = dt_warp(dt);
			if constexpr (!ALT) { if (out.tap_t) tap_sample(out, view, idx, step_i - 1u + (uint32_t)k, o + d * t, t, env.min_mip, env.max_mip); }   // uniform; off outside the parity test
			t += dt;
			// -- hash-grid gather straight into this lane's row of the staging tile ---------------------------------------
			encode_position_to_row<F, false>(s_levels, wpos, tile + lane * kTileStrideH, nullptr);
		}
		// -- both MLPs ---------------------------------------------------------------------------------------------------
#if SNB_LANE_UNROLLED_MLP   // tuning hook: both 16-row m-tiles per weight fetch (half the shared-memory wavefronts, twice the code)
		uint32_t sh_frag[2][4];
		__syncwarp();
		load_sh_frags(tile, lane, sh_frag);
		const uint2 net = warp_network_eval(s_w, tile, out_tile, lane, sh_frag);
#else
		const uint2 net = warp_network_eval_rolled(s_w, tile, out_tile, lane);
#endif

		// -- composite -------------------------------------------------------------------------------------------------
		if (valid) {
			++n_samples;
			const __half2 rg = u2h2(net.x), bs = u2h2(net.y);
			const float sigma = act_density(__high2float(bs), model.density_act);
			const v3 rgb_net = mk3(act_rgb(__low2float(rg), model.rgb_act), act_rgb(__high2float(rg), model.rgb_act), act_rgb(__low2float(bs), model.rgb_act));
			// One trip per composited sample. A second trip happens only with alt.reuse_dup (sng path): the compositor's rewind makes the first
			// sample of the next batch the last sample of this one again (same position up to the rounding of the rewind), so its network
			// output is re-used instead of being gathered and evaluated a second time; position, dt, depth and all bookkeeping of that
			// sample are computed as always.
			bool parking = false;
#pragma unroll 1
			for (int rep = 0; ; ++rep) {
				const v3 pos = model.train_aabb.lo + wpos * aabb_diag;   // unwarp_position
				const float T = 1.f - rgba.w;
				const float dt = dt_unwarp(wdt);
				const float alpha = 1.f - __expf(-sigma * dt);
				const float weight = alpha * T;
				v3 rgb = rgb_net;
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
					fin = true;
					break;
				}
				if (parking) { ++k; break; }                          // the folded sample of a ray that is about to park (its batch length is not known yet)
				if (++k != (int)batch_steps) break;
				// batch boundary: the sng compositor rewinds t to the last composited sample
				if (ALT) t = depth / dot(cam_fwd, d);
				k = 0;
				step_i += batch_steps;
				if (step_i >= kMarchIter) { alive = false; break; }         // iteration budget exhausted: ray is dropped, not written
				++batch;
				const uint32_t prev_steps = batch_steps;
				if (PH && batch == stop_batch) {
					// lock-step prefix: the ray is alive at the start of batch `stop`: count it and list it now; its state is parked below,
					// after the sample the rewind repeats has been folded in (exactly what the free-running path does at this point)
					list_out[atomicAdd(&alt.lk->n_out, 1u)] = qpos;
					parking = true;
				}
				else if (batch < (uint32_t)kAltSched) { batch_steps = s_sched[batch]; if (count_batches) atomicAdd(&s_hist[batch], 1u); }
				else batch_steps = ALT && alt.fixed_steps ? alt.fixed_steps : kBatchSteps;
				// lengths never fall along a march, so a batch that follows one of >= 2 samples cannot end with the folded sample
				if (!ALT || !alt.reuse_dup || rep != 0 || prev_steps < 2u) break;
				// the re-composited sample of the new batch, at the rewound t (what the search phase + sample position would find next iteration)
				const v3 p2 = o + d * t;
				if (t >= kMaxDepth || !in_render_box(p2, env)) break;
				if (env.bitfield && !cell_occupied_xyz(p2, env.bits_xyz, clamp_mip(mip_of(p2), env))) break;   // the rewind left the cell: a new sample after all
				const float dt2 = step_size_call(t, env.st);
				wpos = (p2 - model.train_aabb.lo) / aabb_diag;
				wdt = dt_warp(dt2);
				t += dt2;
				++n_reused;
			}
			if (PH && parking) {
				RayPark pk; pk.rgba = rgba; pk.t = t; pk.depth = depth; pk.max_weight = max_weight; pk.step_i = step_i; pk.k = (uint32_t)k; pk.dead = fin ? 1u : 0u; pk.pad0 = 0; pk.pad1 = 0;
				park_out[qpos] = pk;
				if (!fin) alive = false;       // fin: finish() writes the pixel at the top of the next iteration
			}
		}
	}

	// -- counters: one atomic per warp ----------------------------------------------------------------------------------
#pragma unroll
	for (int s = 16; s > 0; s >>= 1) {
		n_samples += __shfl_xor_sync(0xFFFFFFFFu, n_samples, s);
		n_hits += __shfl_xor_sync(0xFFFFFFFFu, n_hits, s);
		n_launched += __shfl_xor_sync(0xFFFFFFFFu, n_launched, s);
		n_reused += __shfl_xor_sync(0xFFFFFFFFu, n_reused, s);
	}
	if (PH && alt.phase == 1) {   // a prefix attempt: its counts become final only when k_lock_next commits the chunk
		if (lane == 0) { atomicAdd(&alt.lk->pend_samples, (unsigned long long)n_samples); atomicAdd(&alt.lk->pend_hits, (unsigned long long)n_hits); if (n_reused) atomicAdd(&alt.lk->pend_reused, (unsigned long long)n_reused); }
	} else if (lane == 0 && counters) {
		atomicAdd(&counters->samples, (unsigned long long)n_samples);
		atomicAdd(&counters->hits, (unsigned long long)n_hits);
		atomicAdd(&counters->rays, (unsigned long long)n_launched);
		if (n_reused) atomicAdd(&counters->batches, (unsigned long long)n_reused);   // samples composited from a re-used network output
	}
	if (count_batches) {
		__syncthreads();
		if (threadIdx.x < kAltSched && s_hist[threadIdx.x]) atomicAdd(&alt.st->hist[threadIdx.x], s_hist[threadIdx.x]);
	}
}

// ---- parity taps ---------------------------------------------------------------------------------------------------
// Per-pixel march sample sequence with no early termination (what generate_next_nerf_network_inputs would emit).
__global__ void __launch_bounds__(128) k_march_trace(DevView view, MarchEnv env, uint32_t max_samples, float* __restrict__ out_t, float* __restrict__ out_dt,
                                                     uint32_t* __restrict__ out_cell, uint32_t* __restrict__ n_out, float* __restrict__ origin_dir) {
	const uint32_t ray = blockIdx.x * blockDim.x + threadIdx.x;
	if (ray >= (uint32_t)(view.width * view.height)) return;
	const int x = (int)(ray % (uint32_t)view.width), y = (int)(ray / (uint32_t)view.width);
	v3 o, d; float t;
	bool alive = nerf_primary_ray<true>(x, y, view, env, o, d, t);
	if (alive) {
		const v3 idir0 = 1.0f / d;
		t = advance_steps(t, env.st, ld_value(view.spp, ray * 786433u));
		t = skip_to_occupied(t, o, d, idir0, env);
		alive = t < kMaxDepth;
	}
	if (origin_dir) { float* p = origin_dir + (size_t)ray * 6; p[0] = o.x; p[1] = o.y; p[2] = o.z; p[3] = d.x; p[4] = d.y; p[5] = d.z; }
	uint32_t n = 0;
	if (alive) {
		const v3 idir = 1.0f / d;
		for (; n < max_samples; ++n) {
			t = skip_to_occupied(t, o, d, idir, env);
			if (t >= kMaxDepth) break;
			const float dt = step_size(t, env.st);
			const v3 pos = o + d * t;
			const uint32_t mip = clamp_mip(mip_of(pos), env);
			const size_t at = (size_t)ray * max_samples + n;
			out_t[at] = t; out_dt[at] = dt; out_cell[at] = (mip << 24) | (cascade_cell(pos, mip) & 0xFFFFFFu);
			t += dt;
		}
	}
	for (uint32_t j = n; j < max_samples; ++j) { const size_t at = (size_t)ray * max_samples + j; out_t[at] = 0.0f; out_dt[at] = 0.0f; out_cell[at] = 0xFFFFFFFFu; }
	n_out[ray] = n;
}

// NerfNetwork inference on explicit NerfCoordinates (n x 7 floats) through the same encode + warp evaluator as the fused kernel.
template <int F>
__global__ void __launch_bounds__(kTapThreads) k_network_inference(DevModel model, const float* __restrict__ coords, uint64_t n, uint16_t* __restrict__ out,
                                                                     uint16_t* __restrict__ features, uint32_t* __restrict__ hash_idx) {
	__shared__ __align__(16) uint32_t s_w[kFragWords];
	__shared__ __align__(16) unsigned char s_warp[kTapWarps][kWarpSmemBytes];
	__shared__ LevelDesc s_levels[kMaxLevels];
	for (int i = threadIdx.x; i < kFragWords / 4; i += kTapThreads)
		reinterpret_cast<uint4*>(s_w)[i] = __ldg(reinterpret_cast<const uint4*>(model.mlp_frag) + i);
	if (threadIdx.x < kFeat / F) s_levels[threadIdx.x] = make_level_desc(model.grid, threadIdx.x, model.grid_params);
	__syncthreads();
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	__half* tile = reinterpret_cast<__half*>(s_warp[warp]);
	__half* out_tile = reinterpret_cast<__half*>(s_warp[warp] + kWarpTileBytes);
	__half* row = tile + lane * kTileStrideH;
	const uint64_t n_tiles = (n + 31) / 32;
	for (uint64_t tile_i = (uint64_t)blockIdx.x * kTapWarps + warp; tile_i < n_tiles; tile_i += (uint64_t)gridDim.x * kTapWarps) {
		const uint64_t i = tile_i * 32 + lane;
		const bool valid = i < n;
		uint32_t sh[8];
#pragma unroll
		for (int j = 0; j < 8; ++j) sh[j] = 0u;
#pragma unroll
		for (int j = 0; j < 4; ++j) reinterpret_cast<uint4*>(row)[j] = make_uint4(0, 0, 0, 0);
		if (valid) {
			const float* c = coords + i * 7;
			const v3 p = mk3(c[0], c[1], c[2]), dir = mk3(c[4], c[5], c[6]);
			if (hash_idx) encode_position_to_row<F, true>(s_levels, p, row, hash_idx + i * (uint64_t)(kFeat / F) * 8);
			else encode_position_to_row<F, false>(s_levels, p, row, nullptr);
			encode_direction_sh4(dir, sh);
			if (features) {
#pragma unroll
				for (int j = 0; j < 16; ++j) reinterpret_cast<uint32_t*>(features)[i * 16 + j] = reinterpret_cast<const uint32_t*>(row)[j];
			}
		}
		stage_sh_row(tile, lane, sh);
		__syncwarp();
		uint32_t sh_frag[2][4];
		load_sh_frags(tile, lane, sh_frag);
		const uint2 r = warp_network_eval(s_w, tile, out_tile, lane, sh_frag);
		if (valid) reinterpret_cast<uint2*>(out)[i] = r;
	}
}

// ---- screen-space normals from the position buffer (write_normals_to_buffer, src/testbed_nerf.cu:1523-1576) -------------
__global__ void __launch_bounds__(128) k_write_normals(int width, int height, const float* __restrict__ positions, float* __restrict__ normals) {
	const int x = blockIdx.x * 16 + (threadIdx.x & 15), y = blockIdx.y * 8 + (threadIdx.x >> 4);
	if (x >= width || y >= height) return;
	const int offs[9][2] = {{1, 0}, {0, 1}, {-1, 0}, {0, -1}, {2, 0}, {0, 2}, {-2, 0}, {0, -2}, {1, 0}};
	const int idx = x + width * y;
	const v3 pos = mk3(positions[idx * 3], positions[idx * 3 + 1], positions[idx * 3 + 2]);
	float factor = 0.0f;
	v3 N = mk3(0, 0, 0);
#pragma unroll
	for (int k = 0; k < 8; ++k) {
		const int tx = x + offs[k + 1][0], ty = y + offs[k + 1][1], bx = x + offs[k][0], by = y + offs[k][1];
		if (tx >= width || tx < 0 || ty >= height || ty < 0 || bx >= width || bx < 0 || by >= height || by < 0) continue;
		const float* tp = positions + (size_t)(ty * width + tx) * 3;
		const float* bp = positions + (size_t)(by * width + bx) * 3;
		const v3 T = mk3(tp[0], tp[1], tp[2]) - pos;
		const v3 B = mk3(bp[0], bp[1], bp[2]) - pos;
		N = N + normalize(cross(normalize(T), B));
		factor += 1.0f;
	}
	N = factor == 0.0f ? N : N / factor;
	N = normalize(N);
	normals[idx * 3] = N.x; normals[idx * 3 + 1] = N.y; normals[idx * 3 + 2] = N.z;
}

} // namespace snb
