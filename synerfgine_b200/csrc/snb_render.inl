// synerfgine_b200 — render entry points (part of the snb_api.cu translation unit).

static DevView make_view(const snb_view* v, const snb_tile* tile) {
	DevView d;
	d.cam.c0 = ld3(v->camera); d.cam.c1 = ld3(v->camera + 3); d.cam.c2 = ld3(v->camera + 6); d.cam.c3 = ld3(v->camera + 9);
	d.focal.x = v->focal_length[0]; d.focal.y = v->focal_length[1];
	d.center.x = v->screen_center[0]; d.center.y = v->screen_center[1];
	d.width = v->width; d.height = v->height;
	if (tile && tile->w > 0 && tile->h > 0) { d.tx0 = tile->x0; d.ty0 = tile->y0; d.tw = tile->w; d.th = tile->h; }
	else { d.tx0 = 0; d.ty0 = 0; d.tw = v->width; d.th = v->height; }
	d.spp = v->spp; d.snap = v->snap_to_pixel_centers;
	d.pix_off = ld_pixel_offset(d.snap ? 0u : d.spp);
	d.ex = CameraEx{};
	return d;
}

// The stock path's view: the context's camera extras and environment map on top of the pinhole view.
static DevView make_stock_view(snb_ctx* c, const snb_view* v, const snb_tile* tile) {
	DevView d = make_view(v, tile);
	const bool env = c->env_w > 0 && c->env_h > 0;
	if (!c->has_cam_ex && !env) return d;
	if (c->has_cam_ex) d.ex = c->cam_ex;
	else { d.ex.cam1 = d.cam; d.ex.rolling[3] = 1.0f; d.ex.plane_z = 1.0f; }   // envmap alone: camera1 = camera0, nothing else
	d.ex.enabled = 1;
	d.ex.envmap = env ? c->d_envmap.as<float4>() : nullptr; d.ex.env_w = c->env_w; d.ex.env_h = c->env_h;
	return d;
}

static int check_view(snb_ctx* c, const snb_view* v, const snb_tile* t) {
	if (!v || v->width <= 0 || v->height <= 0) return c->fail(SNB_ERR_ARG, "bad view");
	if (t && t->w > 0 && t->h > 0 && (t->x0 < 0 || t->y0 < 0 || t->x0 + t->w > v->width || t->y0 + t->h > v->height)) return c->fail(SNB_ERR_ARG, "tile outside the view");
	return SNB_OK;
}

// Per-context, per-kernel-variant launch set-up (the dynamic shared-memory opt-in is a per-device attribute and a context belongs to
// one device; nothing process-wide is cached).
template <int F, bool ALT, int W, bool PH = false>
static int launch_lane(snb_ctx* c, const DevView& dv, const NerfOut& out, int render_mode, int n_px, const RayRec* queue, unsigned int* qctl, const AltCtl& alt) {
	const int slot = (F == 1 ? 0 : F == 2 ? 1 : F == 4 ? 2 : 3) * 4 + (ALT ? 2 : 0) + (W == kFusedWarps ? 1 : 0) + (PH ? 16 : 0);
	int& per_sm = c->lane_blocks_per_sm[slot];
	if (per_sm == 0) {
		SNB_CUDA(c, cudaFuncSetAttribute(k_nerf_march_fused<F, ALT, W, PH>, cudaFuncAttributeMaxDynamicSharedMemorySize, lane_smem_bytes(W)));
		SNB_CUDA(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_nerf_march_fused<F, ALT, W, PH>, W * 32, lane_smem_bytes(W)));
		per_sm = std::max(1, per_sm);
	}
	c->fused_blocks_per_sm = per_sm;
	const int blocks = std::max(1, std::min(c->sm_count * per_sm, (n_px + W * 32 - 1) / (W * 32)));
	k_nerf_march_fused<F, ALT, W, PH><<<blocks, W * 32, lane_smem_bytes(W), c->stream>>>(c->model, dv, out, render_mode, queue, qctl, qctl + 1, c->d_counters.as<NerfCounters>(), alt);
	SNB_LAUNCHED(c);
	return SNB_OK;
}

template <int F, bool ALT>
static int launch_ws(snb_ctx* c, const DevView& dv, const NerfOut& out, int render_mode, int n_px, const RayRec* queue, unsigned int* qctl, const AltCtl& alt) {
	const int slot = (F == 1 ? 0 : F == 2 ? 1 : F == 4 ? 2 : 3) * 2 + (ALT ? 1 : 0);
	if (!c->ws_ready[slot]) {
		SNB_CUDA(c, cudaFuncSetAttribute(k_nerf_march_ws<F, ALT>, cudaFuncAttributeMaxDynamicSharedMemorySize, ws::kSmemAlloc));
		c->ws_ready[slot] = true;
	}
	c->fused_blocks_per_sm = 1;
	const int slots = ws::kGroups * ws::kRows;   // ray slots per CTA
	const int blocks = std::max(1, std::min(c->sm_count, (n_px + slots - 1) / slots));
	k_nerf_march_ws<F, ALT><<<blocks, ws::kThreads, ws::kSmemAlloc, c->stream>>>(c->model, dv, out, render_mode, queue, qctl, qctl + 1, c->d_counters.as<NerfCounters>(), alt,
	                                                                                 reinterpret_cast<unsigned long long*>(&c->d_counters.as<NerfCounters>()[1]));
	SNB_LAUNCHED(c);
	return SNB_OK;
}

template <int F, bool ALT>
static int launch_fused(snb_ctx* c, const DevView& dv, const NerfOut& out, int render_mode, int n_px, const RayRec* queue, unsigned int* qctl, const AltCtl& alt) {
	// "ws": warp-specialised tcgen05 pipeline (default). The stock path's Positions / Depth visualisations need the sample position in
	// the compositor, which only the lane kernel carries.
	if (c->march_variant == 6 && !alt.phase && (ALT || (render_mode != SNB_MODE_POSITIONS && render_mode != SNB_MODE_DEPTH))) return launch_ws<F, ALT>(c, dv, out, render_mode, n_px, queue, qctl, alt);
	// CTA shape by the rays a lane will see (option "lane_warps": 0 = by tile size, else 10 | 12): results do not depend on it
	const bool few = c->lane_warps ? c->lane_warps == kFusedWarpsFew : (size_t)n_px < (size_t)c->sm_count * SNB_LANE_CTAS * kFusedWarps * 32 * 16;
	if constexpr (ALT) {
		if (alt.phase)   // lock-step prefix instance (lane kernel only)
			return few ? launch_lane<F, true, kFusedWarpsFew, true>(c, dv, out, render_mode, n_px, queue, qctl, alt) : launch_lane<F, true, kFusedWarps, true>(c, dv, out, render_mode, n_px, queue, qctl, alt);
	}
	return few ? launch_lane<F, ALT, kFusedWarpsFew>(c, dv, out, render_mode, n_px, queue, qctl, alt) : launch_lane<F, ALT, kFusedWarps>(c, dv, out, render_mode, n_px, queue, qctl, alt);
}

template <bool ALT>
static int launch_fused_f(snb_ctx* c, const DevView& dv, const NerfOut& out, int render_mode, int n_px, const RayRec* queue, unsigned int* qctl, const AltCtl& alt) {
	switch (c->feat_per_level) {
		case 1: return launch_fused<1, ALT>(c, dv, out, render_mode, n_px, queue, qctl, alt);
		case 2: return launch_fused<2, ALT>(c, dv, out, render_mode, n_px, queue, qctl, alt);
		case 4: return launch_fused<4, ALT>(c, dv, out, render_mode, n_px, queue, qctl, alt);
		default: return launch_fused<8, ALT>(c, dv, out, render_mode, n_px, queue, qctl, alt);
	}
}

// NeRF layer into tile-sized buffers. alt = sng compositor (trace_alt) + positions.
//
// sng path: the batch lengths of the reference's lock-step march (AltState, snb_nerf_kernels.cuh) are carried on the device from
// frame to frame; a pass whose own ray counts contradict the lengths it used is repeated (at most kAltPasses passes, each later
// one a conditional launch that returns at once when the previous pass verified) — all enqueued without a host round trip.
static const int kAltPasses = 4;
static const int kLockAttempts = 16;  // prefix chunk attempts enqueued per moving frame (kLockChunk batches each; C2 needs 1)
static int render_nerf_layer(snb_ctx* c, const DevView& dv, bool alt, int render_mode, float* fb, float* depth, float* positions) {
	if (!c->has_model) return c->fail(SNB_ERR_STATE, "no model: call snb_set_model first");
	const int n_px = dv.tw * dv.th;
	SNB_CUDA(c, c->d_queue.reserve(((size_t)n_px + 64) * sizeof(RayRec)));
	SNB_CUDA(c, cudaMemsetAsync(c->d_qctl.p, 0, 16, c->stream));
	SNB_CUDA(c, cudaMemsetAsync(c->d_counters.p, 0, 2 * sizeof(NerfCounters), c->stream));   // counters + pipeline-balance cycles of the ws march
	NerfOut out; out.fb = reinterpret_cast<float4*>(fb); out.depth = depth; out.positions = positions;
	out.tap_t = nullptr; out.tap_cell = nullptr; out.tap_n = nullptr; out.tap_cap = 0;
	// lane kernel on the sng path: colours stay in sRGB until k_resolve_srgb below (the Positions / Depth visualisations write other values: not deferred)
	out.raw_rgb = (alt && !dv.ex.enabled && c->march_variant == 3 && c->defer_srgb && render_mode != SNB_MODE_POSITIONS && render_mode != SNB_MODE_DEPTH) ? 1u : 0u;   // (an environment map pre-fills pixels the march never writes: stock path only today)
	if (!alt && c->tap_t && c->march_variant == 3) { out.tap_t = c->tap_t; out.tap_cell = c->tap_cell; out.tap_n = c->tap_n; out.tap_cap = c->tap_cap; }
	unsigned int* qctl = c->d_qctl.as<unsigned int>();
	const int patches_x = (dv.tw + 7) / 8, patches_y = (dv.th + 3) / 4;
	const int patches = patches_x * patches_y;
	RayRec* q0 = c->d_queue.as<RayRec>();
	c->stage_begin(0);
	if (dv.ex.enabled) k_nerf_raygen<true><<<(patches + 3) / 4, 128, 0, c->stream>>>(dv, c->model.env, out, q0, qctl, 0, patches);
	else k_nerf_raygen<false><<<(patches + 3) / 4, 128, 0, c->stream>>>(dv, c->model.env, out, q0, qctl, 0, patches);
	SNB_LAUNCHED(c);
	c->stage_end(0);
	c->stage_begin(1);
	int rc = SNB_OK;
	AltCtl ctl; ctl.st = nullptr; ctl.target_queries = c->alt_target_queries; ctl.fixed_steps = 0; ctl.conditional = 0; ctl.external = c->alt_mode == 1;
	ctl.lk = nullptr; ctl.park[0] = ctl.park[1] = nullptr; ctl.lists[0] = ctl.lists[1] = nullptr; ctl.phase = 0; ctl.reuse_dup = c->dup_reuse ? 1u : 0u;
	if (!alt) {
		// Testbed::render_nerf hands camera_matrix1 to the compositor (src/testbed_nerf.cu:2761): depth is measured along the END-of-shutter camera
		DevView mv = dv;
		if (dv.ex.enabled) mv.cam = dv.ex.cam1;
		rc = launch_fused_f<false>(c, mv, out, render_mode, patches * 32, q0, qctl, ctl);
	}
	else if (c->alt_mode == 2) { ctl.fixed_steps = (uint32_t)c->alt_batch; rc = launch_fused_f<true>(c, dv, out, render_mode, patches * 32, q0, qctl, ctl); }
	else {
		AltState* st = c->d_alt.as<AltState>();
		ctl.st = st;
		const int passes = c->alt_mode == 1 ? 1 : kAltPasses;   // "external": the caller owns the sequence (multi-GPU: counts are summed over ranks)
		// "auto": a frame whose camera (or tile) changed since the last one would need its batch lengths guessed (and re-marched when the
		// guess is wrong): it marches the lock-step prefix instead. A still frame keeps the verified sequence of the last one: one speculative
		// pass. The march reads the model and the view only — moving OBJECTS do not change n_alive — so the scene does not enter the test.
		const bool moved = std::memcmp(&dv.cam, c->alt_last_cam, sizeof c->alt_last_cam) != 0 || dv.tw != c->alt_last_tw || dv.th != c->alt_last_th || dv.tx0 != c->alt_last_tx0 || dv.ty0 != c->alt_last_ty0 ||
		                   dv.focal.x != c->alt_last_focal;
		std::memcpy(c->alt_last_cam, &dv.cam, sizeof c->alt_last_cam); c->alt_last_tw = dv.tw; c->alt_last_th = dv.th; c->alt_last_tx0 = dv.tx0; c->alt_last_ty0 = dv.ty0; c->alt_last_focal = dv.focal.x;
		const bool lockstep = c->march_variant == 3 && (c->alt_mode == 3 || (c->alt_mode == 0 && c->alt_auto_lockstep && moved));
		int first_pass = 0;
		if (lockstep) {
			LockState* lk = reinterpret_cast<LockState*>(st + 1);
			const size_t cap = (size_t)patches * 32 + 64;
			SNB_CUDA(c, c->d_park.reserve(cap * 2 * sizeof(RayPark)));
			SNB_CUDA(c, c->d_lock_lists.reserve(cap * 2 * sizeof(uint32_t)));
			ctl.lk = lk; ctl.park[0] = c->d_park.as<RayPark>(); ctl.park[1] = ctl.park[0] + cap;
			ctl.lists[0] = c->d_lock_lists.as<uint32_t>(); ctl.lists[1] = ctl.lists[0] + cap;
			k_lock_begin<<<1, kAltSched, 0, c->stream>>>(st, lk, c->alt_target_queries, qctl, qctl + 1, c->d_counters.as<NerfCounters>());
			SNB_LAUNCHED(c);
			ctl.phase = 1;
			for (int p = 0; p < kLockAttempts && rc == SNB_OK; ++p) {   // chunk attempts; those beyond the frame's prefix return at once (lk->done)
				rc = launch_fused_f<true>(c, dv, out, render_mode, patches * 32, q0, qctl, ctl);
				if (rc != SNB_OK) break;
				k_lock_next<<<1, 1, 0, c->stream>>>(st, lk, c->alt_target_queries, qctl + 1, c->d_counters.as<NerfCounters>());
				SNB_LAUNCHED(c);
			}
			ctl.phase = 2;
			if (rc == SNB_OK) rc = launch_fused_f<true>(c, dv, out, render_mode, patches * 32, q0, qctl, ctl);
			if (rc == SNB_OK) {
				k_alt_verify<<<1, kAltSched, 0, c->stream>>>(st, c->alt_target_queries, qctl, 0u, 0u);   // exact unless the prefix outlasted the attempts: then the passes below repair it
				SNB_LAUNCHED(c);
			}
			ctl.phase = 0; ctl.lk = nullptr;
			first_pass = 1;
		}
		for (int pass = first_pass; pass < passes && rc == SNB_OK; ++pass) {
			k_alt_begin_pass<<<pass == 0 ? 1 : std::min(2 * c->sm_count, (n_px + 255) / 256), 256, 0, c->stream>>>(st, pass == 0, out, (uint32_t)n_px, qctl + 1, c->d_counters.as<NerfCounters>());
			SNB_LAUNCHED(c);
			ctl.conditional = pass > 0;
			rc = launch_fused_f<true>(c, dv, out, render_mode, patches * 32, q0, qctl, ctl);
			if (rc != SNB_OK) break;
			k_alt_verify<<<1, kAltSched, 0, c->stream>>>(st, c->alt_target_queries, qctl, (pass == passes - 1 || c->alt_mode == 1) ? 1u : 0u, c->alt_mode == 1 ? 1u : 0u);
			SNB_LAUNCHED(c);
		}
	}
	if (rc == SNB_OK && out.raw_rgb) {
		k_resolve_srgb<<<(n_px + 255) / 256, 256, 0, c->stream>>>(out.fb, (uint32_t)n_px);
		SNB_LAUNCHED(c);
	}
	c->stage_end(1);
	return rc;
}

// Per-pixel generator states live in FULL-FRAME arrays indexed by the global pixel (the state belongs to the pixel, not to the
// tile that happens to render it): seeded once per (seed, resolution) like Engine::resize does (engine.cu:236-255), and a tile
// that moves between frames (tiles.StripBalancer) keeps using the same states without any re-seeding.
static int ensure_rng(snb_ctx* c, const DevView& syn, const DevView& nerf) {
	const bool same_tile = c->rng_tx0 == syn.tx0 && c->rng_ty0 == syn.ty0 && c->rng_tw == syn.tw && c->rng_th == syn.th;
	const bool same_frame = !c->rng_dirty && c->rng_w == syn.width && c->rng_h == syn.height && c->rng_scale == syn.width / nerf.width;
	if (!same_tile) { c->syn_reset = true; c->rng_tx0 = syn.tx0; c->rng_ty0 = syn.ty0; c->rng_tw = syn.tw; c->rng_th = syn.th; }   // the mesh layer's temporal blend is tile-local
	if (same_frame) return SNB_OK;
	const uint32_t n_syn = (uint32_t)(syn.width * syn.height), n_nerf = (uint32_t)(nerf.width * nerf.height);
	SNB_CUDA(c, c->d_rng_syn.reserve((size_t)n_syn * sizeof(Rng)));
	SNB_CUDA(c, c->d_rng_nerf.reserve((size_t)n_nerf * sizeof(Rng)));
	k_init_rng<<<(n_syn + 127) / 128, 128, 0, c->stream>>>(n_syn, (unsigned long long)c->rng_seed, syn.width, 0, 0, syn.width, c->d_rng_syn.as<Rng>());
	SNB_LAUNCHED(c);
	k_init_rng<<<(n_nerf + 127) / 128, 128, 0, c->stream>>>(n_nerf, (unsigned long long)c->rng_seed, nerf.width, 0, 0, nerf.width, c->d_rng_nerf.as<Rng>());
	SNB_LAUNCHED(c);
	c->rng_dirty = false; c->rng_w = syn.width; c->rng_h = syn.height;
	c->rng_scale = syn.width / nerf.width;
	c->syn_reset = true;
	return SNB_OK;
}

static RtParams make_rt(const snb_rt_params* rt) {
	RtParams r;
	r.n_steps = rt->n_steps; r.samples = rt->samples; r.bounces = rt->bounces; r.shadow_iters = rt->shadow_iters;
	r.lens = rt->lens_angle_constant; r.depth_offset = rt->depth_offset; r.syn_shadow_factor = rt->syn_shadow_factor;
	r.view_nerf_shadow = rt->view_nerf_shadow; r.buffer_type = rt->buffer_type;
	return r;
}

// Mesh layer into the context's syn buffers (tile-sized).
static int render_syn_layer(snb_ctx* c, const DevView& dv, const snb_rt_params* rt, cudaStream_t st) {
	if (!st) st = c->stream;
	if (!c->has_model) return c->fail(SNB_ERR_STATE, "no model: call snb_set_model first");
	if (rt->samples == 0 || rt->bounces == 0 || rt->shadow_iters == 0) return c->fail(SNB_ERR_ARG, "samples, bounces and shadow_iters must be >= 1");
	const uint32_t n = (uint32_t)(dv.tw * dv.th);
	if (c->syn_w != dv.tw || c->syn_h != dv.th) { c->syn_reset = true; c->syn_w = dv.tw; c->syn_h = dv.th; }
	SNB_CUDA(c, c->d_syn_origins.reserve((size_t)n * 12));
	SNB_CUDA(c, c->d_syn_dirs.reserve((size_t)n * 12));
	SNB_CUDA(c, c->d_syn_rgba.reserve((size_t)n * 16));
	SNB_CUDA(c, c->d_syn_depth.reserve((size_t)n * 4));
	const dim3 grid2((dv.tw + 15) / 16, (dv.th + 7) / 8);
	c->stage_begin(3, st);
	if (c->syn_reset) {
		k_syn_raygen<<<grid2, 128, 0, st>>>(dv, c->d_syn_origins.as<float>(), c->d_syn_dirs.as<float>(), nullptr, nullptr, c->d_syn_rgba.as<float4>(), c->d_syn_depth.as<float>());
		SNB_LAUNCHED(c);
		c->syn_reset = false;
	}
	const int sb = c->sng_block;   // 0 = row-linear 128-thread blocks; else threads per block of the 8 x 4-patch-per-warp mapping
	const dim3 grid_px = sb ? dim3(((dv.tw + 7) / 8 + sb / 32 - 1) / (sb / 32), (dv.th + 3) / 4) : dim3((n + 127) / 128);
	k_raytrace<<<grid_px, sb ? sb : 128, 0, st>>>(n, c->d_syn_origins.as<float>(), c->d_syn_dirs.as<float>(), dv.cam.c0, c->model.env, c->scene, make_rt(rt),
	                                              c->d_rng_syn.as<Rng>(), c->d_syn_rgba.as<float4>(), c->d_syn_depth.as<float>(), dv.width, dv.tx0, dv.ty0, dv.tw, sb);
	SNB_LAUNCHED(c);
	c->stage_end(3, st);
	return SNB_OK;
}

// Normals + mesh shadows on the NeRF layer (tile-sized buffers; window reads stop at the tile edge like at the image edge).
static int nerf_post(snb_ctx* c, const DevView& dv, const snb_rt_params* rt, int render_mode, float* fb, float* positions, float* normals) {
	const dim3 grid2((dv.tw + 15) / 16, (dv.th + 7) / 8);
	c->stage_begin(2);
	k_write_normals<<<grid2, 128, 0, c->stream>>>(dv.tw, dv.th, positions, normals);
	SNB_LAUNCHED(c);
	if (rt->show_shadow_on_nerf) {
		const uint32_t n = (uint32_t)(dv.tw * dv.th);
		SNB_CUDA(c, c->d_terms.reserve((size_t)n * 4));
		MarchEnv env = c->model.env; env.min_mip = 0;
		const int sb = c->terms_block;
		const dim3 grid_px = sb ? dim3(((dv.tw + 7) / 8 + sb / 32 - 1) / (sb / 32), (dv.th + 3) / 4) : dim3((n + 127) / 128);
		const float4* layer = (render_mode != 8 && c->skip_dead_terms) ? reinterpret_cast<const float4*>(fb) : nullptr;
		k_nerf_shadow_terms<<<grid_px, sb ? sb : 128, 0, c->stream>>>(n, positions, normals, env, c->scene, rt->nerf_shadow_intensity, rt->nerf_on_nerf_shadow_threshold,
		                                                            c->d_rng_nerf.as<Rng>(), c->d_terms.as<float>(), layer, dv.tw, dv.th,
		                                                            rt->nerf_shadow_kernel / 2, dv.width, dv.tx0, dv.ty0, sb);
		SNB_LAUNCHED(c);
		k_nerf_shadow_apply<<<grid2, 128, 0, c->stream>>>(dv.tw, dv.th, rt->nerf_shadow_kernel / 2, rt->nerf_shadow_intensity, render_mode, c->d_terms.as<float>(),
		                                                  reinterpret_cast<float4*>(fb), nullptr);
		SNB_LAUNCHED(c);
	}
	c->stage_end(2);
	return SNB_OK;
}

static DevView nerf_view_of(const DevView& syn, int scale) {
	DevView n = syn;
	if (scale > 1) {
		n.width = syn.width / scale; n.height = syn.height / scale;
		n.tx0 = syn.tx0 / scale; n.ty0 = syn.ty0 / scale; n.tw = syn.tw / scale; n.th = syn.th / scale;
		n.focal.x = syn.focal.x / (float)scale; n.focal.y = syn.focal.y / (float)scale;
	}
	return n;
}

extern "C" {

int snb_render_nerf(snb_ctx* c, const snb_view* view, const snb_tile* tile, int render_mode, float* fb, float* depth) {
	if (!c || !fb || !depth) return SNB_ERR_ARG;
	int rc = check_view(c, view, tile); if (rc) return rc;
	SNB_CUDA(c, cudaSetDevice(c->device));
	return render_nerf_layer(c, make_stock_view(c, view, tile), false, render_mode, fb, depth, nullptr);
}

int snb_render_syn(snb_ctx* c, const snb_view* view, const snb_tile* tile, const snb_rt_params* rt, float* syn_rgba, float* syn_depth) {
	if (!c || !rt) return SNB_ERR_ARG;
	int rc = check_view(c, view, tile); if (rc) return rc;
	SNB_CUDA(c, cudaSetDevice(c->device));
	const DevView dv = make_view(view, tile);
	const int scale = std::max(1, rt->syn_px_scale);
	rc = ensure_rng(c, dv, nerf_view_of(dv, scale)); if (rc) return rc;
	rc = render_syn_layer(c, dv, rt, nullptr); if (rc) return rc;
	const size_t n = (size_t)dv.tw * dv.th;
	if (syn_rgba) SNB_CUDA(c, cudaMemcpyAsync(syn_rgba, c->d_syn_rgba.p, n * 16, cudaMemcpyDeviceToDevice, c->stream));
	if (syn_depth) SNB_CUDA(c, cudaMemcpyAsync(syn_depth, c->d_syn_depth.p, n * 4, cudaMemcpyDeviceToDevice, c->stream));
	return SNB_OK;
}

int snb_render_nerf_with_buffers(snb_ctx* c, const snb_view* view, const snb_tile* tile, const snb_rt_params* rt, int render_mode,
                                 float* fb, float* depth, float* positions, float* normals) {
	if (!c || !rt || !fb || !depth) return SNB_ERR_ARG;
	int rc = check_view(c, view, tile); if (rc) return rc;
	SNB_CUDA(c, cudaSetDevice(c->device));
	const DevView dv = make_view(view, tile);
	const size_t n = (size_t)dv.tw * dv.th;
	if (!positions) { SNB_CUDA(c, c->d_positions.reserve(n * 12)); positions = c->d_positions.as<float>(); }
	if (!normals) { SNB_CUDA(c, c->d_normals.reserve(n * 12)); normals = c->d_normals.as<float>(); }
	// RNG states of the NeRF layer are per NeRF pixel; (re)seeded on resize like Engine::resize (engine.cu:236-255).
	DevView syn = dv; const int scale = std::max(1, rt->syn_px_scale);
	if (scale > 1) { syn.width *= scale; syn.height *= scale; syn.tx0 *= scale; syn.ty0 *= scale; syn.tw *= scale; syn.th *= scale; }
	rc = ensure_rng(c, syn, dv); if (rc) return rc;
	rc = render_nerf_layer(c, dv, true, render_mode, fb, depth, positions); if (rc) return rc;
	return nerf_post(c, dv, rt, render_mode, fb, positions, normals);
}

int snb_overlay(snb_ctx* c, int syn_w, int syn_h, const snb_rt_params* rt, const float* syn_rgba, const float* syn_depth, const float* nerf_rgba,
                const float* nerf_depth, float* out_rgba, float* out_depth) {
	if (!c || !rt || !syn_rgba || !syn_depth || !nerf_rgba || !nerf_depth || !out_rgba || syn_w <= 0 || syn_h <= 0) return SNB_ERR_ARG;
	SNB_CUDA(c, cudaSetDevice(c->device));
	const dim3 grid2((syn_w + 15) / 16, (syn_h + 7) / 8);
	c->stage_begin(4);
	k_overlay<<<grid2, 128, 0, c->stream>>>(syn_w, syn_h, std::max(1, rt->syn_px_scale), reinterpret_cast<const float4*>(syn_rgba), syn_depth,
	                                        reinterpret_cast<const float4*>(nerf_rgba), nerf_depth, rt->show_nerf, 1, rt->tonemap_curve, rt->exposure, rt->depth_offset,
	                                        reinterpret_cast<float4*>(out_rgba), out_depth);
	SNB_LAUNCHED(c);
	c->stage_end(4);
	return SNB_OK;
}

int snb_accumulate_tonemap(snb_ctx* c, int n_pixels, const float* frame, float* accum, uint32_t spp, int cs_srgb, int out_srgb, float exposure,
                           const float* bg, int curve, float* out) {
	if (!c || !frame || !accum || !bg || !out || n_pixels <= 0) return SNB_ERR_ARG;
	SNB_CUDA(c, cudaSetDevice(c->device));
	k_accumulate_tonemap<<<(n_pixels + 127) / 128, 128, 0, c->stream>>>(n_pixels, reinterpret_cast<const float4*>(frame), reinterpret_cast<float4*>(accum), spp, cs_srgb, out_srgb,
	                                                                   exposure, make_float4(bg[0], bg[1], bg[2], bg[3]), curve, reinterpret_cast<float4*>(out));
	SNB_LAUNCHED(c);
	return SNB_OK;
}

int snb_frame(snb_ctx* c, const snb_view* view, const snb_tile* tile, const snb_rt_params* rt, float* out_rgba, float* out_depth) {
	if (!c || !rt || !out_rgba) return SNB_ERR_ARG;
	int rc = check_view(c, view, tile); if (rc) return rc;
	SNB_CUDA(c, cudaSetDevice(c->device));
	const int scale = std::max(1, rt->syn_px_scale);
	const DevView syn = make_view(view, tile);
	if (scale > 1 && (syn.width % scale || syn.height % scale || syn.tx0 % scale || syn.ty0 % scale || syn.tw % scale || syn.th % scale))
		return c->fail(SNB_ERR_ARG, "view and tile must be multiples of syn_px_scale");
	const DevView nerf = nerf_view_of(syn, scale);
	const size_t n_nerf = (size_t)nerf.tw * nerf.th;
	rc = ensure_rng(c, syn, nerf); if (rc) return rc;
	// 1. mesh layer (RayTracer::render) — forked onto the second stream, joined before the overlay
	const bool fork = c->overlap_layers && rt->show_nerf;
	if (fork) {
		if (!c->stream2) {
			int prio_lo = 0, prio_hi = 0;   // numerically lower = higher priority
			SNB_CUDA(c, cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
			const int prio = c->mesh_priority > 0 ? prio_hi : c->mesh_priority < 0 ? prio_lo : 0;
			SNB_CUDA(c, cudaStreamCreateWithPriority(&c->stream2, cudaStreamNonBlocking, prio));
			SNB_CUDA(c, cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
			SNB_CUDA(c, cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
		}
	}
	if (fork) {
		SNB_CUDA(c, cudaEventRecord(c->ev_fork, c->stream));
		SNB_CUDA(c, cudaStreamWaitEvent(c->stream2, c->ev_fork, 0));
	}
	rc = render_syn_layer(c, syn, rt, fork ? c->stream2 : c->stream); if (rc) return rc;
	if (fork) SNB_CUDA(c, cudaEventRecord(c->ev_join, c->stream2));
	// 2. NeRF layer with mesh shadows (Testbed::render sng overload)
	SNB_CUDA(c, c->d_nerf_fb.reserve(n_nerf * 16));
	SNB_CUDA(c, c->d_nerf_depth.reserve(n_nerf * 4));
	if (rt->show_nerf) {
		SNB_CUDA(c, c->d_positions.reserve(n_nerf * 12));
		SNB_CUDA(c, c->d_normals.reserve(n_nerf * 12));
		rc = render_nerf_layer(c, nerf, true, SNB_MODE_SHADE, c->d_nerf_fb.as<float>(), c->d_nerf_depth.as<float>(), c->d_positions.as<float>());
		if (rc) return rc;
		rc = nerf_post(c, nerf, rt, SNB_MODE_SHADE, c->d_nerf_fb.as<float>(), c->d_positions.as<float>(), c->d_normals.as<float>()); if (rc) return rc;
	}
	if (fork) SNB_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_join, 0));
	// 3. overlay (RayTracer::overlay), output colour space sRGB (engine.cu:406)
	return snb_overlay(c, syn.tw, syn.th, rt, c->d_syn_rgba.as<float>(), c->d_syn_depth.as<float>(), c->d_nerf_fb.as<float>(), c->d_nerf_depth.as<float>(), out_rgba, out_depth);
}

int snb_frame_host(snb_ctx* c, const snb_view* view, const snb_tile* tile, const snb_rt_params* rt, float* out_rgba_host, float* out_depth_host) {
	if (!c || !out_rgba_host) return SNB_ERR_ARG;
	int rc = check_view(c, view, tile); if (rc) return rc;
	const DevView dv = make_view(view, tile);
	const size_t n = (size_t)dv.tw * dv.th;
	SNB_CUDA(c, cudaSetDevice(c->device));
	SNB_CUDA(c, c->d_out_rgba.reserve(n * 16));
	SNB_CUDA(c, c->d_out_depth.reserve(n * 4));
	rc = snb_frame(c, view, tile, rt, c->d_out_rgba.as<float>(), c->d_out_depth.as<float>()); if (rc) return rc;
	SNB_CUDA(c, cudaMemcpyAsync(out_rgba_host, c->d_out_rgba.p, n * 16, cudaMemcpyDeviceToHost, c->stream));
	if (out_depth_host) SNB_CUDA(c, cudaMemcpyAsync(out_depth_host, c->d_out_depth.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
	SNB_CUDA(c, cudaStreamSynchronize(c->stream));
	return SNB_OK;
}

// Pipelined form of snb_frame_host for callers that stream frames: the device-to-host copy of frame i runs on a copy stream while
// frame i + 1 renders. `slot` (0 | 1) names the double buffer; a slot's host buffers (pinned memory, or the copy is not asynchronous)
// may be read after snb_frame_host_wait(ctx, slot) and must not be reused before it.
int snb_frame_host_async(snb_ctx* c, const snb_view* view, const snb_tile* tile, const snb_rt_params* rt, float* out_rgba_host, float* out_depth_host, int32_t slot) {
	if (!c || !out_rgba_host || slot < 0 || slot > 1) return SNB_ERR_ARG;
	int rc = check_view(c, view, tile); if (rc) return rc;
	const DevView dv = make_view(view, tile);
	const size_t n = (size_t)dv.tw * dv.th;
	SNB_CUDA(c, cudaSetDevice(c->device));
	if (!c->copy_stream) {
		SNB_CUDA(c, cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
		for (int k = 0; k < 2; ++k) { SNB_CUDA(c, cudaEventCreateWithFlags(&c->ev_frame_done[k], cudaEventDisableTiming)); SNB_CUDA(c, cudaEventCreateWithFlags(&c->ev_copy_done[k], cudaEventDisableTiming)); }
	}
	if (c->d_pipe_rgba[slot].cap < n * 16 || c->d_pipe_depth[slot].cap < n * 4) {
		SNB_CUDA(c, cudaStreamSynchronize(c->copy_stream));       // growing a buffer a copy may still read
		SNB_CUDA(c, c->d_pipe_rgba[slot].reserve(n * 16));
		SNB_CUDA(c, c->d_pipe_depth[slot].reserve(n * 4));
	}
	if (c->pipe_used[slot]) SNB_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_copy_done[slot], 0));   // the slot's last copy has left the device buffer
	rc = snb_frame(c, view, tile, rt, c->d_pipe_rgba[slot].as<float>(), c->d_pipe_depth[slot].as<float>()); if (rc) return rc;
	SNB_CUDA(c, cudaEventRecord(c->ev_frame_done[slot], c->stream));
	SNB_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev_frame_done[slot], 0));
	SNB_CUDA(c, cudaMemcpyAsync(out_rgba_host, c->d_pipe_rgba[slot].p, n * 16, cudaMemcpyDeviceToHost, c->copy_stream));
	if (out_depth_host) SNB_CUDA(c, cudaMemcpyAsync(out_depth_host, c->d_pipe_depth[slot].p, n * 4, cudaMemcpyDeviceToHost, c->copy_stream));
	SNB_CUDA(c, cudaEventRecord(c->ev_copy_done[slot], c->copy_stream));
	c->pipe_used[slot] = true;
	return SNB_OK;
}

int snb_frame_host_wait(snb_ctx* c, int32_t slot) {
	if (!c || slot < 0 || slot > 1) return SNB_ERR_ARG;
	if (!c->pipe_used[slot]) return SNB_OK;
	SNB_CUDA(c, cudaSetDevice(c->device));
	SNB_CUDA(c, cudaEventSynchronize(c->ev_copy_done[slot]));
	return SNB_OK;
}

} // extern "C"
