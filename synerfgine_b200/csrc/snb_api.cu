// synerfgine_b200 — C ABI implementation (include/snb.h): lifetime, inputs, the render entry points and taps.
// Host orchestration follows Engine::frame (reference src/synerfgine/engine.cu:352-433), Testbed::render sng
// overload (src/testbed.cu:4353-4404), render_nerf_with_buffers / shade_nerf_shadows / render_nerf
// (src/testbed_nerf.cu:2467-2837) and RayTracer::render / overlay (src/synerfgine/raytracer.cu:312-392) — minus
// every host<->device sync inside the frame: the march loop is device-resident, so a frame is a fixed sequence of
// launches on one stream.
#include <cstring>
#include <cstdlib>
#include <cmath>
#include <algorithm>
#include <cstdio>
#include <cctype>
#include <string>
#include <vector>
#include "snb_context.cuh"

using namespace snb;

static v3 ld3(const float* p) { return mk3(p[0], p[1], p[2]); }
static m3 ldm3(const float* p) { m3 m; m.c0 = ld3(p); m.c1 = ld3(p + 3); m.c2 = ld3(p + 6); return m; }

// tcnn grid_scale / grid_resolution (common_device.h:841-850) and the level offset table (encodings/grid.h:692-723).
static GridDesc make_grid_desc(uint32_t n_levels, uint32_t F, uint32_t log2_T, uint32_t base_res, float per_level_scale) {
	GridDesc g; memset(&g, 0, sizeof g);
	g.n_levels = n_levels; g.n_feat = F;
	const float log2_pls = std::log2(per_level_scale);
	uint32_t offset = 0;
	for (uint32_t i = 0; i < n_levels; ++i) {
		const float scale = std::exp2(i * log2_pls) * base_res - 1.0f;
		const uint32_t res = (uint32_t)std::ceil(scale) + 1;
		const uint32_t max_params = 0xFFFFFFFFu / 2;
		uint32_t n = std::pow((float)res, 3) > (float)max_params ? max_params : res * res * res;
		n = (n + 7u) / 8u * 8u;
		n = std::min(n, 1u << log2_T);
		g.scale[i] = scale; g.res[i] = res; g.offset[i] = offset;
		offset += n;
	}
	g.offset[n_levels] = offset;
	return g;
}

extern "C" {

int snb_set_option(snb_ctx* c, const char* key, const char* value) {
	if (!c || !key || !value) return SNB_ERR_ARG;
	const std::string k(key), v(value);
	if (k == "march") {
		if (v == "lane") c->march_variant = 3;
		else if (v == "ws") c->march_variant = 6;
		else return c->fail(SNB_ERR_ARG, "march: expected ws | lane");
	} else if (k == "bvh") {
		if (v == "reference" || v == "host") c->bvh_builder = 0;
		else if (v == "lbvh") c->bvh_builder = 1;
		else return c->fail(SNB_ERR_ARG, "bvh: expected reference | lbvh");
	} else if (k == "bvh_leaf") {
		const int b = atoi(v.c_str());
		if (b < 1 || b > 64) return c->fail(SNB_ERR_ARG, "bvh_leaf: expected 1..64");
		c->bvh_leaf = b;
	} else if (k == "alt_schedule") {
		if (v == "auto") c->alt_mode = 0;
		else if (v == "external") c->alt_mode = 1;
		else if (v == "fixed") c->alt_mode = 2;
		else if (v == "lockstep") c->alt_mode = 3;
		else return c->fail(SNB_ERR_ARG, "alt_schedule: expected auto | external | fixed | lockstep");
	} else if (k == "alt_auto_lockstep") {
		c->alt_auto_lockstep = v != "0";
	} else if (k == "dup_reuse") {
		c->dup_reuse = v != "0";
	} else if (k == "defer_srgb") {
		c->defer_srgb = v != "0";
	} else if (k == "alt_batch") {
		const int b = atoi(v.c_str());
		if (b < 1 || b > 8) return c->fail(SNB_ERR_ARG, "alt_batch: expected 1..8");
		c->alt_batch = b;
	} else if (k == "alt_target_queries") {
		const long long t = atoll(v.c_str());
		if (t < 1 || t > 0x7FFFFFFFll) return c->fail(SNB_ERR_ARG, "alt_target_queries: expected a positive 31-bit count");
		c->alt_target_queries = (uint32_t)t;
	} else if (k == "mlp") {
		if (v == "mma") c->mlp_variant = 0;
		else if (v == "tc") c->mlp_variant = 1;
		else return c->fail(SNB_ERR_ARG, "mlp: expected mma | tc");
	} else if (k == "overlap") {
		c->overlap_layers = v != "0";
	} else if (k == "mesh_priority") {
		c->mesh_priority = v == "high" ? 1 : v == "low" ? -1 : 0;
	} else if (k == "lane_warps") {
		const int w = atoi(v.c_str());
		if (w != 0 && w != kFusedWarpsFew && w != kFusedWarps) return c->fail(SNB_ERR_ARG, "lane_warps: expected 0 | 10 | 12");
		c->lane_warps = w;
	} else if (k == "sng_block" || k == "terms_block") {
		const int b = atoi(v.c_str());
		if ((b != 0 && b != 32 && b != 64 && b != 128 && b != 256) || (k == "sng_block" ? b > SNB_RT_THREADS : b > 128)) return c->fail(SNB_ERR_ARG, "sng_block / terms_block: expected 0 | 32 | 64 | 128");
		(k == "sng_block" ? c->sng_block : c->terms_block) = b;
	} else if (k == "dead_terms") {
		c->skip_dead_terms = v != "0";
	} else return c->fail(SNB_ERR_ARG, "unknown option");
	return SNB_OK;
}

const char* snb_version(void) { return "synerfgine_b200 0.1 (sm_100a)"; }

int snb_create(int device, void* cuda_stream, snb_ctx** out) {
	if (!out) return SNB_ERR_ARG;
	*out = nullptr;
	int n_dev = 0;
	if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) return SNB_ERR_CUDA;   // no CPU fallback
	if (device < 0 || device >= n_dev) return SNB_ERR_ARG;
	if (cudaSetDevice(device) != cudaSuccess) return SNB_ERR_CUDA;
	cudaDeviceProp prop;
	if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return SNB_ERR_CUDA;
	if (prop.major < 10) return SNB_ERR_UNSUPPORTED;   // sm_100a binary only
	snb_ctx* c = new snb_ctx();
	c->device = device;
	c->sm_count = prop.multiProcessorCount;
	// SNB_<OPTION> in the environment sets the default of snb_set_option(ctx, "<option>", ...) (A/B runs without touching the caller)
	static const char* const kEnvOptions[] = {"march", "mlp", "overlap", "dead_terms", "lane_warps", "sng_block", "terms_block", "mesh_priority", "alt_schedule", "alt_batch", "alt_auto_lockstep", "dup_reuse", "defer_srgb", "bvh", "bvh_leaf"};
	for (const char* opt : kEnvOptions) {
		std::string env = std::string("SNB_") + opt;
		for (char& ch : env) ch = (char)toupper((unsigned char)ch);
		if (const char* mv = getenv(env.c_str())) snb_set_option(c, opt, mv);
	}
	if (cuda_stream) c->stream = (cudaStream_t)cuda_stream;
	else { if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { delete c; return SNB_ERR_CUDA; } c->own_stream = true; }
	if (c->d_qctl.reserve(64) != cudaSuccess || c->d_counters.reserve(2 * sizeof(NerfCounters)) != cudaSuccess || c->d_alt.reserve(sizeof(AltState) + sizeof(LockState)) != cudaSuccess) { delete c; return SNB_ERR_CUDA; }
	cudaMemsetAsync(c->d_counters.p, 0, 2 * sizeof(NerfCounters), c->stream);
	cudaMemsetAsync(c->d_alt.p, 0, sizeof(AltState) + sizeof(LockState), c->stream);
	memset(&c->model, 0, sizeof c->model);
	*out = c;
	return SNB_OK;
}

void snb_destroy(snb_ctx* c) {
	if (!c) return;
	cudaSetDevice(c->device);
	cudaStreamSynchronize(c->stream);
	DevBuf* bufs[] = {&c->d_grid, &c->d_mlp, &c->d_mlp_frag, &c->d_mlp_umma, &c->d_bitfield, &c->d_bits_xyz, &c->d_coarse, &c->d_present, &c->d_objects, &c->d_materials, &c->d_lights, &c->d_queue, &c->d_qctl,
	                  &c->d_counters, &c->d_alt, &c->d_rng_nerf, &c->d_rng_syn, &c->d_syn_origins, &c->d_syn_dirs, &c->d_syn_rgba, &c->d_syn_depth, &c->d_nerf_fb, &c->d_nerf_depth,
	                  &c->d_positions, &c->d_normals, &c->d_terms, &c->d_out_rgba, &c->d_out_depth, &c->d_envmap, &c->d_park, &c->d_lock_lists, &c->d_tris_in, &c->d_bvh_depth, &c->lbvh.keys, &c->lbvh.keys_sorted, &c->lbvh.sort_tmp,
	                  &c->lbvh.islot, &c->lbvh.parent, &c->lbvh.flags, &c->lbvh.bounds};
	for (DevBuf* b : bufs) b->release();
	for (auto& o : c->obj_store) { o.nodes.release(); o.tris.release(); o.order.release(); }
	if (c->copy_stream) {
		cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream);
		for (int k = 0; k < 2; ++k) { cudaEventDestroy(c->ev_frame_done[k]); cudaEventDestroy(c->ev_copy_done[k]); c->d_pipe_rgba[k].release(); c->d_pipe_depth[k].release(); }
	}
	if (c->stream2) { cudaStreamSynchronize(c->stream2); cudaStreamDestroy(c->stream2); cudaEventDestroy(c->ev_fork); cudaEventDestroy(c->ev_join); }
	if (c->own_stream) cudaStreamDestroy(c->stream);
	delete c;
}

const char* snb_last_error(const snb_ctx* c) { return c ? c->err.c_str() : "null context"; }
uint64_t snb_launch_count(const snb_ctx* c) { return c ? c->launches : 0; }

int snb_sync(snb_ctx* c) {
	if (!c) return SNB_ERR_ARG;
	SNB_CUDA(c, cudaStreamSynchronize(c->stream));
	return SNB_OK;
}

static int set_model_impl(snb_ctx* c, const snb_model_desc* m, bool on_device);
int snb_set_model(snb_ctx* c, const snb_model_desc* m) { return set_model_impl(c, m, false); }
// The same with `params` and `bitfield` pointing to DEVICE memory of the context's GPU (a model that was just trained or loaded there):
// the hash table and the bitfield are copied device-to-device; only the 20 KB of MLP weights visit the host, to be laid out in fragment order.
int snb_set_model_dev(snb_ctx* c, const snb_model_desc* m) { return set_model_impl(c, m, true); }

static int set_model_impl(snb_ctx* c, const snb_model_desc* m, bool on_device) {
	if (!c || !m || !m->params) return SNB_ERR_ARG;
	SNB_CUDA(c, cudaSetDevice(c->device));
	const cudaMemcpyKind kind = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
	std::vector<uint16_t> mlp_host;
	const uint16_t* mlp_params = m->params;
	if (on_device) {
		mlp_host.resize(kMlpHalves);
		SNB_CUDA(c, cudaMemcpyAsync(mlp_host.data(), m->params, kMlpHalves * 2, cudaMemcpyDeviceToHost, c->stream));
		SNB_CUDA(c, cudaStreamSynchronize(c->stream));
		mlp_params = mlp_host.data();
	}
	const uint32_t F = m->n_features_per_level, L = m->n_levels;
	if (!(F == 1 || F == 2 || F == 4 || F == 8) || L * F != (uint32_t)kFeat || L > (uint32_t)kMaxLevels)
		return c->fail(SNB_ERR_UNSUPPORTED, "encoding must satisfy n_levels * n_features_per_level == 32 with F in {1,2,4,8}");
	if (m->max_cascade >= kCascades) return c->fail(SNB_ERR_ARG, "max_cascade must be < 8");
	GridDesc g = make_grid_desc(L, F, m->log2_hashmap_size, m->base_resolution, m->per_level_scale);
	const uint64_t n_grid = (uint64_t)g.offset[L] * F;
	if (m->n_params != (uint64_t)kMlpHalves + n_grid) return c->fail(SNB_ERR_ARG, "n_params does not match the network architecture");
	SNB_CUDA(c, c->d_mlp.reserve(kMlpHalves * 2));
	SNB_CUDA(c, c->d_mlp_frag.reserve(kFragWords * 4));
	SNB_CUDA(c, c->d_grid.reserve(n_grid * 2));
	SNB_CUDA(c, cudaMemcpyAsync(c->d_mlp.p, m->params, kMlpHalves * 2, kind, c->stream));
	SNB_CUDA(c, cudaMemcpyAsync(c->d_grid.p, m->params + kMlpHalves, n_grid * 2, kind, c->stream));
	std::vector<uint32_t> frag(kFragWords);
	build_mma_fragments(mlp_params, frag.data());
	SNB_CUDA(c, cudaMemcpyAsync(c->d_mlp_frag.p, frag.data(), kFragWords * 4, cudaMemcpyHostToDevice, c->stream));
	std::vector<uint16_t> umma(kMlpHalves);
	tc::build_umma_weights(mlp_params, umma.data());
	SNB_CUDA(c, c->d_mlp_umma.reserve(kMlpHalves * 2));
	SNB_CUDA(c, cudaMemcpyAsync(c->d_mlp_umma.p, umma.data(), kMlpHalves * 2, cudaMemcpyHostToDevice, c->stream));
	const size_t bf_bytes = (size_t)kCascades * kGridCells / 8;
	if (m->bitfield) {
		SNB_CUDA(c, c->d_bitfield.reserve(bf_bytes));
		SNB_CUDA(c, cudaMemcpyAsync(c->d_bitfield.p, m->bitfield, bf_bytes, kind, c->stream));
	}
	{   // level scales as the device evaluates them (k_grid_scales); resolutions follow, the offset table stays the host's like tcnn's
		DevBuf tmp;
		SNB_CUDA(c, tmp.reserve(kMaxLevels * sizeof(float)));
		float* d_scales = tmp.as<float>();
		k_grid_scales<<<1, 32, 0, c->stream>>>(L, std::log2(m->per_level_scale), m->base_resolution, d_scales);
		SNB_LAUNCHED(c);
		float h_scales[kMaxLevels];
		SNB_CUDA(c, cudaMemcpyAsync(h_scales, d_scales, L * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
		SNB_CUDA(c, cudaStreamSynchronize(c->stream));   // host staging vectors go out of scope
		tmp.release();
		for (uint32_t i = 0; i < L; ++i) { g.scale[i] = h_scales[i]; g.res[i] = (uint32_t)std::ceil(h_scales[i]) + 1; }
	}

	DevModel& d = c->model;
	d.grid = g;
	d.grid_params = c->d_grid.as<__half>();
	d.mlp_rowmajor = c->d_mlp.as<__half>();
	d.mlp_frag = c->d_mlp_frag.as<uint32_t>();
	d.mlp_umma = c->d_mlp_umma.as<__half>();
	d.train_aabb.lo = ld3(m->aabb_min); d.train_aabb.hi = ld3(m->aabb_max);
	d.rgb_act = m->rgb_activation; d.density_act = m->density_activation;
	d.min_transmittance = m->min_transmittance;
	d.env.bitfield = m->bitfield ? c->d_bitfield.as<uint8_t>() : nullptr;
	d.env.bits_xyz = nullptr;
	if (m->bitfield) {
		SNB_CUDA(c, c->d_bits_xyz.reserve(bf_bytes));
		k_bitfield_to_linear<<<((uint32_t)bf_bytes + 127) / 128, 128, 0, c->stream>>>(c->d_bitfield.as<uint8_t>(), (uint32_t)bf_bytes, c->d_bits_xyz.as<uint8_t>());
		SNB_LAUNCHED(c);
		d.env.bits_xyz = c->d_bits_xyz.as<uint8_t>();
	}
	d.env.min_mip = 0; d.env.max_mip = m->max_cascade;
	d.env.render_aabb.lo = ld3(m->render_aabb_min); d.env.render_aabb.hi = ld3(m->render_aabb_max);
	d.env.to_local = ldm3(m->render_aabb_to_local);
	{
		static const float ident[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
		d.env.local_identity = std::memcmp(m->render_aabb_to_local, ident, sizeof ident) == 0 ? 1u : 0u;   // bitwise: -0 entries do not qualify
	}
	d.env.st = make_stepping(m->cone_angle_constant);
	// coarse look-ahead grid over the extent of cascade max_cascade (snb_march.cuh)
	d.env.coarse = nullptr;
	const float extent = std::ldexp(1.0f, (int)m->max_cascade);
	d.env.coarse_lo = mk3(0.5f - 0.5f * extent, 0.5f - 0.5f * extent, 0.5f - 0.5f * extent);
	d.env.coarse_inv_cell = (float)kCoarseRes / extent;
	if (m->bitfield) {
		SNB_CUDA(c, c->d_coarse.reserve(kCoarseWords * 4));
		SNB_CUDA(c, cudaMemsetAsync(c->d_coarse.p, 0, kCoarseWords * 4, c->stream));
		const uint32_t n_bytes = kGridCells / 8 * (m->max_cascade + 1);
		k_build_coarse<<<(n_bytes + 127) / 128, 128, 0, c->stream>>>(c->d_bitfield.as<uint8_t>(), n_bytes, d.env.coarse_lo, d.env.coarse_inv_cell, c->d_coarse.as<uint32_t>());
		SNB_LAUNCHED(c);
		SNB_CUDA(c, cudaStreamSynchronize(c->stream));
		d.env.coarse = c->d_coarse.as<uint32_t>();
	}
	c->feat_per_level = F;
	c->has_model = true;
	return SNB_OK;
}

static int upload_scene_pods(snb_ctx* c, const snb_object* objects, uint32_t n_objects, const snb_material* materials, uint32_t n_materials,
                             const snb_light* lights, uint32_t n_lights) {
	std::vector<DevObject> ho(n_objects);
	for (uint32_t i = 0; i < n_objects; ++i) {
		DevObject& o = ho[i];
		o.nodes = c->obj_store[i].nodes.as<float>(); o.tris = c->obj_store[i].tris.as<float>();
		o.n_nodes = c->obj_store[i].n_nodes; o.n_tris = c->obj_store[i].n_tris;
		o.rot = ldm3(objects[i].rot);
		o.inv_sr = mul(identity3() / objects[i].scale, inverse(o.rot));   // ngp::ray_intersect_nodes, src/triangle_bvh.cu:313-319
		o.pos = ld3(objects[i].pos); o.scale = objects[i].scale; o.mat_id = objects[i].mat_id;
	}
	std::vector<DevMaterial> hm(n_materials);
	for (uint32_t i = 0; i < n_materials; ++i) {
		hm[i].ka = ld3(materials[i].ka); hm[i].kd = ld3(materials[i].kd); hm[i].ks = ld3(materials[i].ks);
		hm[i].n = materials[i].n; hm[i].rg = materials[i].rg; hm[i].type = materials[i].type; hm[i].spec_angle = materials[i].spec_angle;
	}
	std::vector<DevLight> hl(n_lights);
	for (uint32_t i = 0; i < n_lights; ++i) { hl[i].pos = ld3(lights[i].pos); hl[i].type = lights[i].type; hl[i].intensity = lights[i].intensity; hl[i].size = lights[i].size; }
	SNB_CUDA(c, c->d_objects.reserve(std::max<size_t>(1, n_objects) * sizeof(DevObject)));
	SNB_CUDA(c, c->d_materials.reserve(std::max<size_t>(1, n_materials) * sizeof(DevMaterial)));
	SNB_CUDA(c, c->d_lights.reserve(std::max<size_t>(1, n_lights) * sizeof(DevLight)));
	if (n_objects) SNB_CUDA(c, cudaMemcpyAsync(c->d_objects.p, ho.data(), n_objects * sizeof(DevObject), cudaMemcpyHostToDevice, c->stream));
	if (n_materials) SNB_CUDA(c, cudaMemcpyAsync(c->d_materials.p, hm.data(), n_materials * sizeof(DevMaterial), cudaMemcpyHostToDevice, c->stream));
	if (n_lights) SNB_CUDA(c, cudaMemcpyAsync(c->d_lights.p, hl.data(), n_lights * sizeof(DevLight), cudaMemcpyHostToDevice, c->stream));
	// No stream synchronisation: copies from pageable host memory return once the source has been staged, and they are ordered on the
	// context's stream behind the previous frame (whose mesh layer on the second stream was joined before its overlay), so a per-frame
	// snb_update_scene (animated objects, Engine::update_world_objects) does not drain the GPU.
	c->scene.objects = c->d_objects.as<DevObject>(); c->scene.n_objects = n_objects;
	c->scene.materials = c->d_materials.as<DevMaterial>(); c->scene.n_materials = n_materials;
	c->scene.lights = c->d_lights.as<DevLight>(); c->scene.n_lights = n_lights;
	c->syn_reset = true;   // anything moved -> accumulation restarts (engine.cu:80-127)
	return SNB_OK;
}

static int check_scene_pods(snb_ctx* c, const snb_object* objects, uint32_t n_objects, const snb_material* materials, uint32_t n_materials, const snb_light* lights, uint32_t n_lights) {
	if ((n_objects && !objects) || (n_materials && !materials) || (n_lights && !lights)) return c->fail(SNB_ERR_ARG, "scene arrays missing");
	for (uint32_t i = 0; i < n_objects; ++i) {
		if (objects[i].mat_id < 0 || (uint32_t)objects[i].mat_id >= n_materials) return c->fail(SNB_ERR_ARG, "object material index out of range");
		if (!(objects[i].scale != 0.0f)) return c->fail(SNB_ERR_ARG, "object scale must not be 0");
	}
	return SNB_OK;
}

int snb_set_scene(snb_ctx* c, const snb_object* objects, uint32_t n_objects, const snb_material* materials, uint32_t n_materials,
                  const snb_light* lights, uint32_t n_lights) {
	if (!c) return SNB_ERR_ARG;
	int rc = check_scene_pods(c, objects, n_objects, materials, n_materials, lights, n_lights); if (rc) return rc;
	SNB_CUDA(c, cudaSetDevice(c->device));
	const bool device_build = c->bvh_builder == 1;
	for (uint32_t i = 0; i < n_objects; ++i) {
		if (!objects[i].tris || objects[i].n_tris <= 0) return c->fail(SNB_ERR_ARG, "object without triangles");
		if (!device_build && (!objects[i].nodes || objects[i].n_nodes <= 0)) return c->fail(SNB_ERR_ARG, "object without BVH nodes (pass them, or set option bvh=lbvh to build on the device)");
	}
	for (auto& o : c->obj_store) { o.nodes.release(); o.tris.release(); o.order.release(); }
	c->obj_store.assign(n_objects, ObjectStore());
	if (device_build) { SNB_CUDA(c, c->d_bvh_depth.reserve(4)); SNB_CUDA(c, cudaMemsetAsync(c->d_bvh_depth.p, 0, 4, c->stream)); }
	for (uint32_t i = 0; i < n_objects; ++i) {
		ObjectStore& s = c->obj_store[i];
		s.n_tris = objects[i].n_tris;
		SNB_CUDA(c, s.tris.reserve((size_t)s.n_tris * 36));
		if (!device_build) {
			s.n_nodes = objects[i].n_nodes;
			SNB_CUDA(c, s.nodes.reserve((size_t)s.n_nodes * 32));
			SNB_CUDA(c, cudaMemcpyAsync(s.nodes.p, objects[i].nodes, (size_t)s.n_nodes * 32, cudaMemcpyHostToDevice, c->stream));
			SNB_CUDA(c, cudaMemcpyAsync(s.tris.p, objects[i].tris, (size_t)s.n_tris * 36, cudaMemcpyHostToDevice, c->stream));
		} else {
			// linear BVH on the device (snb_bvh_dev.cuh): triangles go up in the caller's order, come out in Morton order with `order` kept
			s.n_nodes = lbvh_node_count(s.n_tris, c->bvh_leaf);
			SNB_CUDA(c, s.nodes.reserve((size_t)s.n_nodes * 32));
			SNB_CUDA(c, s.order.reserve((size_t)s.n_tris * 4));
			SNB_CUDA(c, c->d_tris_in.reserve((size_t)s.n_tris * 36));
			SNB_CUDA(c, cudaMemcpyAsync(c->d_tris_in.p, objects[i].tris, (size_t)s.n_tris * 36, cudaMemcpyHostToDevice, c->stream));
			int launches = 0;
			SNB_CUDA(c, lbvh_build_device(c->lbvh, c->stream, c->d_tris_in.as<float>(), s.n_tris, c->bvh_leaf, s.tris.as<float>(), s.nodes.as<float>(), s.order.as<int>(),
			                              c->d_bvh_depth.as<int>(), &launches));
			c->launches += (uint64_t)launches;
			s.permuted = true;
		}
	}
	if (device_build && n_objects) {
		SNB_CUDA(c, cudaMemcpyAsync(&c->bvh_max_depth, c->d_bvh_depth.p, 4, cudaMemcpyDeviceToHost, c->stream));
		SNB_CUDA(c, cudaStreamSynchronize(c->stream));
		if (c->bvh_max_depth > 31) return c->fail(SNB_ERR_STATE, "device-built BVH deeper than the traversal stack (32): raise bvh_leaf");
	}
	return upload_scene_pods(c, objects, n_objects, materials, n_materials, lights, n_lights);
}

int snb_update_scene(snb_ctx* c, const snb_object* objects, uint32_t n_objects, const snb_material* materials, uint32_t n_materials,
                     const snb_light* lights, uint32_t n_lights) {
	if (!c) return SNB_ERR_ARG;
	if (n_objects != c->obj_store.size()) return c->fail(SNB_ERR_STATE, "snb_update_scene: object count differs from snb_set_scene");
	int rc = check_scene_pods(c, objects, n_objects, materials, n_materials, lights, n_lights); if (rc) return rc;
	SNB_CUDA(c, cudaSetDevice(c->device));
	return upload_scene_pods(c, objects, n_objects, materials, n_materials, lights, n_lights);
}

// Deforming mesh: new vertex positions of object `index` (n_tris x 9 floats in the order snb_set_scene received them), then a bottom-up
// refit of its BVH bounds on the device. Topology (and with it every triangle id) stays.
int snb_update_object_triangles(snb_ctx* c, uint32_t index, const float* tris_host) {
	if (!c || !tris_host) return SNB_ERR_ARG;
	if (index >= c->obj_store.size()) return c->fail(SNB_ERR_ARG, "object index out of range");
	SNB_CUDA(c, cudaSetDevice(c->device));
	ObjectStore& s = c->obj_store[index];
	if (s.permuted) {
		SNB_CUDA(c, c->d_tris_in.reserve((size_t)s.n_tris * 36));
		SNB_CUDA(c, cudaMemcpyAsync(c->d_tris_in.p, tris_host, (size_t)s.n_tris * 36, cudaMemcpyHostToDevice, c->stream));
		k_gather_tris<<<(int)(((long long)s.n_tris * 9 + 255) / 256), 256, 0, c->stream>>>(c->d_tris_in.as<float>(), s.order.as<int>(), s.n_tris, s.tris.as<float>());
		SNB_LAUNCHED(c);
	} else SNB_CUDA(c, cudaMemcpyAsync(s.tris.p, tris_host, (size_t)s.n_tris * 36, cudaMemcpyHostToDevice, c->stream));
	SNB_CUDA(c, bvh_refit_device(c->lbvh, c->stream, s.tris.as<float>(), s.nodes.as<float>(), s.n_nodes));
	c->launches += 2;
	c->syn_reset = true;
	return SNB_OK;
}

// ≙ TriangleBvh::build on the device, for arrays the caller owns: tris_dev (n x 9) -> tris_out_dev (Morton order), nodes_out_dev
// ((2 * ceil(n / prims_per_leaf) - 1) x 8 floats), order_out_dev (n ints or NULL). Synchronises when max_depth_out is given.
int snb_bvh_build_dev(snb_ctx* c, const float* tris_dev, int32_t n_tris, int32_t prims_per_leaf, float* tris_out_dev, float* nodes_out_dev, int32_t nodes_cap,
                      int32_t* order_out_dev, int32_t* n_nodes_out, int32_t* max_depth_out) {
	if (!c || !tris_dev || !tris_out_dev || !nodes_out_dev || n_tris <= 0 || prims_per_leaf <= 0 || tris_dev == tris_out_dev) return SNB_ERR_ARG;
	const int n_nodes = lbvh_node_count(n_tris, prims_per_leaf);
	if (n_nodes_out) *n_nodes_out = n_nodes;
	if (nodes_cap < n_nodes) return c->fail(SNB_ERR_ARG, "node buffer too small");
	SNB_CUDA(c, cudaSetDevice(c->device));
	if (max_depth_out) { SNB_CUDA(c, c->d_bvh_depth.reserve(4)); SNB_CUDA(c, cudaMemsetAsync(c->d_bvh_depth.p, 0, 4, c->stream)); }
	int launches = 0;
	SNB_CUDA(c, lbvh_build_device(c->lbvh, c->stream, tris_dev, n_tris, prims_per_leaf, tris_out_dev, nodes_out_dev, order_out_dev, max_depth_out ? c->d_bvh_depth.as<int>() : nullptr, &launches));
	c->launches += (uint64_t)launches;
	if (max_depth_out) {
		SNB_CUDA(c, cudaMemcpyAsync(max_depth_out, c->d_bvh_depth.p, 4, cudaMemcpyDeviceToHost, c->stream));
		SNB_CUDA(c, cudaStreamSynchronize(c->stream));
	}
	return SNB_OK;
}

int snb_bvh_refit_dev(snb_ctx* c, const float* tris_dev, float* nodes_dev, int32_t n_nodes) {
	if (!c || !tris_dev || !nodes_dev || n_nodes <= 0) return SNB_ERR_ARG;
	SNB_CUDA(c, cudaSetDevice(c->device));
	SNB_CUDA(c, bvh_refit_device(c->lbvh, c->stream, tris_dev, nodes_dev, n_nodes));
	c->launches += 2;
	return SNB_OK;
}

int snb_seed_rng(snb_ctx* c, uint64_t seed) {
	if (!c) return SNB_ERR_ARG;
	c->rng_seed = seed; c->rng_dirty = true;
	return SNB_OK;
}

int snb_reset_accumulation(snb_ctx* c) {
	if (!c) return SNB_ERR_ARG;
	c->syn_reset = true;
	return SNB_OK;
}

} // extern "C"

// ---- allocation-free refit (no counterpart in the reference) -----------------------------
// Some callers cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). snb_bvh_refit_dev keeps its
// scratch (parent index + arrival counter per node) in grow-only context buffers; this variant keeps it in a fixed array in device memory
// instead. Every GPU that loads the library holds the array (SNB_REFIT_NOALLOC_MAX_NODES x 8 B = 2 MiB) until the process exits.
static __device__ __align__(256) unsigned int g_refit_noalloc_scratch[2 * SNB_REFIT_NOALLOC_MAX_NODES];   // [0, MAX): parent, [MAX, 2 MAX): arrival flags

#include <mutex>
static const int kNoAllocMaxDevices = 64;
static std::mutex g_refit_noalloc_mutex;
static unsigned int* g_refit_noalloc_ptr[kNoAllocMaxDevices] = {};   // device address of the array, looked up once per device

extern "C" int snb_bvh_refit_dev_noalloc(snb_ctx* c, const float* tris_dev, float* nodes_dev, int32_t n_nodes) {
	if (!c || !tris_dev || !nodes_dev || n_nodes <= 0) return SNB_ERR_ARG;
	if (n_nodes > SNB_REFIT_NOALLOC_MAX_NODES) {
		char msg[160];
		snprintf(msg, sizeof msg, "snb_bvh_refit_dev_noalloc: %d nodes exceed the fixed scratch of %d nodes (use snb_bvh_refit_dev)", (int)n_nodes, (int)SNB_REFIT_NOALLOC_MAX_NODES);
		return c->fail(SNB_ERR_ARG, msg);
	}
	if (c->device < 0 || c->device >= kNoAllocMaxDevices) return c->fail(SNB_ERR_ARG, "snb_bvh_refit_dev_noalloc: device ordinal out of range");
	SNB_CUDA(c, cudaSetDevice(c->device));
	unsigned int* scratch = nullptr;
	{
		// In host code `g_refit_noalloc_scratch` is only a handle: the device address comes from the runtime, once per device, with that device current.
		std::lock_guard<std::mutex> lock(g_refit_noalloc_mutex);
		if (!g_refit_noalloc_ptr[c->device]) {
			void* p = nullptr;
			SNB_CUDA(c, cudaGetSymbolAddress(&p, g_refit_noalloc_scratch));
			g_refit_noalloc_ptr[c->device] = static_cast<unsigned int*>(p);
		}
		scratch = g_refit_noalloc_ptr[c->device];
	}
	int* parent = reinterpret_cast<int*>(scratch);
	unsigned int* flags = scratch + SNB_REFIT_NOALLOC_MAX_NODES;
	const int blocks = (n_nodes + 255) / 256;
	k_bvh_parents<<<blocks, 256, 0, c->stream>>>(nodes_dev, n_nodes, parent, flags);   // also clears the arrival flags of [0, n_nodes): the array keeps the last call's contents
	SNB_LAUNCHED(c);
	k_bvh_fit<<<blocks, 256, 0, c->stream>>>(nodes_dev, n_nodes, tris_dev, parent, flags);
	SNB_LAUNCHED(c);
	return SNB_OK;
}

#include "snb_render.inl"
#include "snb_taps.inl"
#include "snb_present.inl"
