// synerfgine_b200 — parity taps, counters and the occupancy bitfield builder (part of the snb_api.cu translation unit).

namespace snb {

SNB_HD uint32_t compact3(uint32_t x) { // inverse of spread3 (tcnn morton3D_invert)
	x &= 0x49249249u;
	x = (x | (x >> 2)) & 0xc30c30c3u;
	x = (x | (x >> 4)) & 0x0f00f00fu;
	x = (x | (x >> 8)) & 0xff0000ffu;
	x = (x | (x >> 16)) & 0x0000ffffu;
	return x;
}

// density > min(0.01, mean) -> one bit per cell, 8 cells per byte; cascades above max_cascade are zeroed.
__global__ void __launch_bounds__(128) k_density_to_bits(uint32_t n_bytes, uint32_t n_nonzero_bytes, const float* __restrict__ density, float thresh, uint8_t* __restrict__ bits_out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_bytes) return;
	if (i >= n_nonzero_bytes) { bits_out[i] = 0; return; }
	uint8_t bits = 0;
#pragma unroll
	for (int j = 0; j < 8; ++j) bits |= density[(size_t)i * 8 + j] > thresh ? (uint8_t)(1u << j) : (uint8_t)0;
	bits_out[i] = bits;
}

// 2x2x2 OR-pool of one cascade into the centre octant of the next (Morton order on both sides).
__global__ void __launch_bounds__(128) k_bits_max_pool(uint32_t n_bytes_next_octant, const uint8_t* __restrict__ prev, uint8_t* __restrict__ next) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_bytes_next_octant) return;
	uint8_t bits = 0;
#pragma unroll
	for (int j = 0; j < 8; ++j) bits |= prev[(size_t)i * 8 + j] > 0 ? (uint8_t)(1u << j) : (uint8_t)0;
	const uint32_t x = compact3(i >> 0) + kGridSize / 8, y = compact3(i >> 1) + kGridSize / 8, z = compact3(i >> 2) + kGridSize / 8;
	next[morton3(x, y, z)] |= bits;
}

// Parity tap of the generator: state words of subsequence i after curand_init(seed, i, 0) as k_init_rng leaves them, and the first
// n_uniform draws of rng_uniform from that state (what light_sample / scatter / the lens jitter consume).
__global__ void __launch_bounds__(128) k_rng_tap(uint32_t n, const Rng* __restrict__ states, uint32_t* __restrict__ out6, float* __restrict__ uniforms, uint32_t n_uniform) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	Rng r = states[i];
	out6[i * 6 + 0] = r.v0; out6[i * 6 + 1] = r.v1; out6[i * 6 + 2] = r.v2; out6[i * 6 + 3] = r.v3; out6[i * 6 + 4] = r.v4; out6[i * 6 + 5] = r.d;
	for (uint32_t k = 0; k < n_uniform; ++k) uniforms[(size_t)i * n_uniform + k] = rng_uniform(r);
}

} // namespace snb

extern "C" {

int snb_rng_tap(snb_ctx* c, uint64_t seed, uint32_t n, uint32_t n_uniform, uint32_t* states6_dev, float* uniforms_dev) {
	if (!c || !states6_dev || (n_uniform && !uniforms_dev) || n == 0) return SNB_ERR_ARG;
	SNB_CUDA(c, cudaSetDevice(c->device));
	DevBuf tmp;
	SNB_CUDA(c, tmp.reserve((size_t)n * sizeof(Rng)));
	k_init_rng<<<(n + 127) / 128, 128, 0, c->stream>>>(n, (unsigned long long)seed, (int)n, 0, 0, (int)n, tmp.as<Rng>());
	++c->launches;
	k_rng_tap<<<(n + 127) / 128, 128, 0, c->stream>>>(n, tmp.as<Rng>(), states6_dev, uniforms_dev, n_uniform);
	++c->launches;
	const cudaError_t e = cudaStreamSynchronize(c->stream);
	tmp.release();
	if (e != cudaSuccess) return c->cuda_fail(e, "snb_rng_tap");
	return SNB_OK;
}

int snb_march_trace(snb_ctx* c, const snb_view* view, uint32_t max_samples, float* t_dev, float* dt_dev, uint32_t* cell_dev, uint32_t* n_dev, float* origin_dir_dev) {
	if (!c || !t_dev || !dt_dev || !cell_dev || !n_dev || max_samples == 0) return SNB_ERR_ARG;
	int rc = check_view(c, view, nullptr); if (rc) return rc;
	if (!c->has_model) return c->fail(SNB_ERR_STATE, "no model");
	SNB_CUDA(c, cudaSetDevice(c->device));
	const DevView dv = make_stock_view(c, view, nullptr);
	const uint32_t n = (uint32_t)(dv.width * dv.height);
	k_march_trace<<<(n + 127) / 128, 128, 0, c->stream>>>(dv, c->model.env, max_samples, t_dev, dt_dev, cell_dev, n_dev, origin_dir_dev);
	SNB_LAUNCHED(c);
	return SNB_OK;
}

int snb_network_inference(snb_ctx* c, const float* coords_dev, uint64_t n, uint16_t* out_dev, uint16_t* features_dev, uint32_t* hash_idx_dev) {
	if (!c || !coords_dev || !out_dev) return SNB_ERR_ARG;
	if (!c->has_model) return c->fail(SNB_ERR_STATE, "no model");
	if (n == 0) return SNB_OK;
	SNB_CUDA(c, cudaSetDevice(c->device));
	if (c->mlp_variant == 1) {
		const uint64_t tiles = (n + tc::kTileRows - 1) / tc::kTileRows;
		const int blocks = (int)std::min<uint64_t>(tiles, (uint64_t)c->sm_count * 2);
		const int smem = tc::kSmemBytes + 1024 + (int)sizeof(LevelDesc) * kMaxLevels;
#define SNB_TC_TAP(FF) do { \
		SNB_CUDA(c, cudaFuncSetAttribute(k_network_inference_tc<FF>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); \
		k_network_inference_tc<FF><<<blocks, tc::kTileRows, smem, c->stream>>>(c->model, coords_dev, n, out_dev, features_dev, hash_idx_dev); } while (0)
		switch (c->feat_per_level) {
			case 1: SNB_TC_TAP(1); break;
			case 2: SNB_TC_TAP(2); break;
			case 4: SNB_TC_TAP(4); break;
			default: SNB_TC_TAP(8); break;
		}
#undef SNB_TC_TAP
		SNB_LAUNCHED(c);
		return SNB_OK;
	}
	const uint64_t tiles = (n + 31) / 32;
	const int blocks = (int)std::min<uint64_t>((tiles + kTapWarps - 1) / kTapWarps, (uint64_t)c->sm_count * 8);
	switch (c->feat_per_level) {
		case 1: k_network_inference<1><<<blocks, kTapThreads, 0, c->stream>>>(c->model, coords_dev, n, out_dev, features_dev, hash_idx_dev); break;
		case 2: k_network_inference<2><<<blocks, kTapThreads, 0, c->stream>>>(c->model, coords_dev, n, out_dev, features_dev, hash_idx_dev); break;
		case 4: k_network_inference<4><<<blocks, kTapThreads, 0, c->stream>>>(c->model, coords_dev, n, out_dev, features_dev, hash_idx_dev); break;
		default: k_network_inference<8><<<blocks, kTapThreads, 0, c->stream>>>(c->model, coords_dev, n, out_dev, features_dev, hash_idx_dev); break;
	}
	SNB_LAUNCHED(c);
	return SNB_OK;
}

int snb_trace_rays(snb_ctx* c, const float* rays_dev, uint32_t n, int32_t* obj_id_dev, int32_t* tri_id_dev, float* t_dev) {
	if (!c || !rays_dev || !obj_id_dev || !tri_id_dev || !t_dev) return SNB_ERR_ARG;
	if (n == 0) return SNB_OK;
	SNB_CUDA(c, cudaSetDevice(c->device));
	k_trace_rays<<<(n + 127) / 128, 128, 0, c->stream>>>(n, rays_dev, c->scene, obj_id_dev, tri_id_dev, t_dev);
	SNB_LAUNCHED(c);
	return SNB_OK;
}

int snb_get_counters(snb_ctx* c, uint64_t* out4) {
	if (!c || !out4) return SNB_ERR_ARG;
	SNB_CUDA(c, cudaSetDevice(c->device));
	NerfCounters h;
	SNB_CUDA(c, cudaMemcpyAsync(&h, c->d_counters.p, sizeof h, cudaMemcpyDeviceToHost, c->stream));
	SNB_CUDA(c, cudaStreamSynchronize(c->stream));
	out4[0] = h.samples; out4[1] = h.rays; out4[2] = h.hits; out4[3] = h.batches;
	return SNB_OK;
}

int snb_get_march_stats(snb_ctx* c, uint64_t* out4) {
	if (!c || !out4) return SNB_ERR_ARG;
	SNB_CUDA(c, cudaSetDevice(c->device));
	NerfCounters h;
	SNB_CUDA(c, cudaMemcpyAsync(&h, c->d_counters.p, sizeof h, cudaMemcpyDeviceToHost, c->stream));
	SNB_CUDA(c, cudaStreamSynchronize(c->stream));
	out4[0] = h.iterations; out4[1] = h.dda_steps; out4[2] = h.net_phases; out4[3] = (uint64_t)c->fused_blocks_per_sm * (uint64_t)c->sm_count;
	return SNB_OK;
}

int snb_get_alt_status(snb_ctx* c, uint32_t* out132) {
	if (!c || !out132) return SNB_ERR_ARG;
	SNB_CUDA(c, cudaSetDevice(c->device));
	AltState h;
	SNB_CUDA(c, cudaMemcpyAsync(&h, c->d_alt.p, sizeof h, cudaMemcpyDeviceToHost, c->stream));
	SNB_CUDA(c, cudaStreamSynchronize(c->stream));
	out132[0] = h.passes; out132[1] = h.converged; out132[2] = h.last_mismatch; out132[3] = c->alt_target_queries;
	for (int j = 0; j < kAltSched; ++j) { out132[4 + j] = h.hist[j]; out132[4 + kAltSched + j] = h.sched[j]; }
	return SNB_OK;
}

int snb_set_alt_schedule(snb_ctx* c, const uint8_t* steps, uint32_t n) {
	if (!c || (n && !steps) || n > (uint32_t)kAltSched) return SNB_ERR_ARG;
	for (uint32_t j = 0; j < n; ++j) if (steps[j] > 8) return c->fail(SNB_ERR_ARG, "batch lengths are 0 (unknown) or 1..8");
	SNB_CUDA(c, cudaSetDevice(c->device));
	uint8_t sched[kAltSched];
	for (int j = 0; j < kAltSched; ++j) sched[j] = (uint32_t)j < n ? steps[j] : (n ? (uint8_t)8 : (uint8_t)0);
	SNB_CUDA(c, cudaMemcpyAsync(c->d_alt.as<unsigned char>() + offsetof(AltState, sched), sched, sizeof sched, cudaMemcpyHostToDevice, c->stream));
	SNB_CUDA(c, cudaStreamSynchronize(c->stream));   // `sched` is a stack array
	return SNB_OK;
}

// The march kernel's own samples: the following snb_render_nerf calls (stock path, lane march) record, per tile pixel, t and
// cell = mip << 24 | Morton index of every sample k_nerf_march_fused evaluates (first `cap` per ray) and the count. NULL clears the tap.
int snb_set_sample_tap(snb_ctx* c, float* t_dev, uint32_t* cell_dev, uint32_t* n_dev, uint32_t cap) {
	if (!c) return SNB_ERR_ARG;
	if (!t_dev || !cell_dev || !n_dev || cap == 0) { c->tap_t = nullptr; c->tap_cell = nullptr; c->tap_n = nullptr; c->tap_cap = 0; return SNB_OK; }
	c->tap_t = t_dev; c->tap_cell = cell_dev; c->tap_n = n_dev; c->tap_cap = cap;
	return SNB_OK;
}

int snb_set_camera_ex(snb_ctx* c, const snb_camera_ex* ex) {
	if (!c) return SNB_ERR_ARG;
	if (!ex) { c->has_cam_ex = false; c->cam_ex = CameraEx{}; return SNB_OK; }
	if (ex->lens_mode < 0 || ex->lens_mode > SNB_LENS_EQUIRECTANGULAR) return c->fail(SNB_ERR_ARG, "unknown lens mode");
	if (ex->aperture_size != 0.0f && ex->plane_z == 0.0f) return c->fail(SNB_ERR_ARG, "an open aperture needs a focus distance (plane_z != 0)");
	CameraEx e = {};
	e.enabled = 1;
	e.cam1.c0 = ld3(ex->camera1); e.cam1.c1 = ld3(ex->camera1 + 3); e.cam1.c2 = ld3(ex->camera1 + 6); e.cam1.c3 = ld3(ex->camera1 + 9);
	for (int i = 0; i < 4; ++i) e.rolling[i] = ex->rolling_shutter[i];
	e.parallax = ld3(ex->parallax_shift);
	e.near_distance = ex->near_distance; e.plane_z = ex->plane_z; e.aperture = ex->aperture_size;
	e.lens_mode = ex->lens_mode;
	for (int i = 0; i < 7; ++i) e.lens[i] = ex->lens_params[i];
	c->cam_ex = e; c->has_cam_ex = true;
	return SNB_OK;
}

int snb_set_envmap(snb_ctx* c, const float* rgba_host, int32_t w, int32_t h) {
	if (!c) return SNB_ERR_ARG;
	if (!rgba_host || w <= 0 || h <= 0) { c->env_w = c->env_h = 0; return SNB_OK; }
	SNB_CUDA(c, cudaSetDevice(c->device));
	SNB_CUDA(c, c->d_envmap.reserve((size_t)w * h * 16));
	SNB_CUDA(c, cudaMemcpyAsync(c->d_envmap.p, rgba_host, (size_t)w * h * 16, cudaMemcpyHostToDevice, c->stream));
	SNB_CUDA(c, cudaStreamSynchronize(c->stream));   // the caller's buffer may go away
	c->env_w = w; c->env_h = h;
	return SNB_OK;
}

int snb_get_grid_scales(snb_ctx* c, float* out_host, uint32_t cap) {
	if (!c || !out_host) return SNB_ERR_ARG;
	if (!c->has_model) return c->fail(SNB_ERR_STATE, "no model");
	const uint32_t n = c->model.grid.n_levels;
	if (cap < n) return c->fail(SNB_ERR_ARG, "snb_get_grid_scales: buffer too small");
	for (uint32_t i = 0; i < n; ++i) out_host[i] = c->model.grid.scale[i];
	return (int)n;
}

int snb_get_ws_profile(snb_ctx* c, uint64_t* out8) {
	if (!c || !out8) return SNB_ERR_ARG;
	SNB_CUDA(c, cudaSetDevice(c->device));
	SNB_CUDA(c, cudaMemcpyAsync(out8, &c->d_counters.as<NerfCounters>()[1], 8 * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->stream));
	SNB_CUDA(c, cudaStreamSynchronize(c->stream));
	return SNB_OK;
}

int snb_get_stage_ms(snb_ctx* c, float* out5) {
	if (!c || !out5) return SNB_ERR_ARG;
	SNB_CUDA(c, cudaSetDevice(c->device));
	SNB_CUDA(c, cudaStreamSynchronize(c->stream));
	for (int s = 0; s < snb_ctx::kStages; ++s) {
		out5[s] = 0.0f;
		if (c->ev_set[s]) cudaEventElapsedTime(&out5[s], c->ev[s][0], c->ev[s][1]);
	}
	return SNB_OK;
}

int snb_density_to_bitfield(snb_ctx* c, const float* density_host, uint32_t max_cascade, uint8_t* bitfield_out_host) {
	if (!c || !density_host || !bitfield_out_host || max_cascade >= kCascades) return SNB_ERR_ARG;
	SNB_CUDA(c, cudaSetDevice(c->device));
	const size_t n_cells = (size_t)kGridCells * (max_cascade + 1);
	// mean of max(d, 0) over cascade 0 (update_density_grid_mean_and_bitfield reduces NERF_GRID_N_CELLS elements)
	double acc = 0.0;
	for (uint32_t i = 0; i < kGridCells; ++i) acc += (double)std::max(density_host[i], 0.0f);
	const float mean = (float)(acc / (double)kGridCells);
	const float thresh = std::min(0.01f, mean);   // NERF_MIN_OPTICAL_THICKNESS
	DevBuf d_density, d_bits;
	const uint32_t n_bytes = kGridCells / 8 * kCascades;
	cudaError_t e = d_density.reserve(n_cells * 4);
	if (e == cudaSuccess) e = d_bits.reserve(n_bytes);
	if (e != cudaSuccess) { d_density.release(); d_bits.release(); return c->cuda_fail(e, "bitfield scratch"); }
	cudaMemcpyAsync(d_density.p, density_host, n_cells * 4, cudaMemcpyHostToDevice, c->stream);
	k_density_to_bits<<<(n_bytes + 127) / 128, 128, 0, c->stream>>>(n_bytes, kGridCells / 8 * (max_cascade + 1), d_density.as<float>(), thresh, d_bits.as<uint8_t>());
	++c->launches;
	for (uint32_t level = 1; level < kCascades; ++level) {
		const uint32_t n = kGridCells / 64;
		k_bits_max_pool<<<(n + 127) / 128, 128, 0, c->stream>>>(n, d_bits.as<uint8_t>() + (size_t)(level - 1) * kGridCells / 8, d_bits.as<uint8_t>() + (size_t)level * kGridCells / 8);
		++c->launches;
	}
	cudaMemcpyAsync(bitfield_out_host, d_bits.p, n_bytes, cudaMemcpyDeviceToHost, c->stream);
	e = cudaStreamSynchronize(c->stream);
	if (e == cudaSuccess) e = cudaGetLastError();
	d_density.release(); d_bits.release();
	if (e != cudaSuccess) return c->cuda_fail(e, "snb_density_to_bitfield");
	return SNB_OK;
}

} // extern "C"
