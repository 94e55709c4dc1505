// ORACLE — TEST INFRASTRUCTURE ONLY. The REFERENCE's own CUDA code, compiled for sm_100a where it lies under
// /root/reference (nothing copied; oracle/Makefile.refdev), behind a thin extern "C" surface:
//   * tiny-cuda-nn (GridEncoding kernel_grid, FullyFusedMLP kernel_mlp_fused, SphericalHarmonics kernel_sh) through the
//     reference's NerfNetwork<__half> (include/neural-graphics-primitives/nerf_network.h),
//   * Testbed::NerfTracer::init_rays_from_camera / trace_alt / trace and the NeRF kernels of src/testbed_nerf.cu
//     (init_rays_with_payload_kernel_nerf, advance_pos_nerf_kernel, compact_kernel_nerf, generate_next_nerf_network_inputs,
//     composite_kernel_nerf(_alt), extract_from_payload, shade_kernel_nerf, write_normals_to_buffer, shade_with_shadow,
//     grid_to_bitfield, bitfield_max_pool),
//   * sng::raytrace / init_rays_with_payload_kernel_nerf / overlay_nerf (src/synerfgine/raytracer.cu), sng::init_rand_state,
//     depth_test_world / depth_test_nerf (src/synerfgine/common.cu), ngp::ray_intersect_nodes (src/triangle_bvh.cu).
// Two uses: (1) tests/ pin the device-only bodies of the CPU oracle (and the product) on real reference output;
// (2) bench.py's `cuda_baseline` leg times the reference's kernels on the same B200 on the same inputs.
// The product (libsnb.so) never links, loads or calls this library.
//
// What is NOT the reference's here, and why: the host glue of Testbed::render_nerf_with_buffers / shade_nerf_shadows
// (src/testbed_nerf.cu:2467-2677), Testbed::render (src/testbed.cu:4353-4404), RayTracer::render / overlay
// (src/synerfgine/raytracer.cu:312-392) and Engine::frame (src/synerfgine/engine.cu:352-433) lives in classes that cannot be
// constructed headless (GLFW window, ImGui, GL textures), so the ~60 lines of launch sequence are restated below with the
// members turned into arguments; every kernel they launch and the whole march loop (NerfTracer) are the reference's.
// GLTexture exists only under NGP_GUI; RayTracer's declaration needs the type to parse, so an empty stand-in type is
// declared (never instantiated).
#include <cstdio>
#include <cstring>
#include <memory>
#include <vector>

#include <neural-graphics-primitives/common.h>
#include <neural-graphics-primitives/render_buffer.h>
#include <imgui/imgui.h>                   // declarations only: light.cuh names ImGui:: without including it (testbed.h does so under NGP_GUI)

namespace ngp {
class GLTexture : public SurfaceProvider {   // parse-only stand-in (see header comment)
public:
	cudaSurfaceObject_t surface() override { return 0; }
	cudaArray_t array() override { return nullptr; }
	ivec2 resolution() const override { return ivec2(0); }
	void resize(const ivec2&, int) override {}
	void load_gpu(const float*, ivec2, int, std::vector<float>&) {}
	void load_gpu(const vec4*, ivec2, std::vector<vec4>&) {}
};
}

#include <testbed_nerf.cu>                 // -I$(REF)/src
#include <synerfgine/common.cu>
#include <synerfgine/raytracer.cu>

// src/triangle_bvh.cu:309 defines `ngp::ray_intersect_nodes` with an explicit qualification inside namespace ngp, which g++
// rejects. Giving the name `ngp` a second meaning INSIDE namespace ngp (a nested namespace that re-exports ::ngp and
// declares that one overload) makes the line a legal out-of-namespace definition; nothing of the file is changed.
namespace ngp { namespace ngp {
using namespace ::ngp;
__host__ __device__ std::pair<int, float> ray_intersect_nodes(const vec3& ro, const vec3& rd, TriangleBvhNode* __restrict__ bvhnodes, Triangle* __restrict__ triangles);
} }
#include <triangle_bvh.cu>

#define EXPORT extern "C" __attribute__((visibility("default")))

// No using-directive for ngp at global scope: together with the nested ngp::ngp above it would make the bare name `ngp` ambiguous
// (nvcc's generated host stubs spell types as ngp::X at global scope).
namespace refdev {
using namespace ::ngp;
using sng::Light; using sng::Material; using sng::ObjectTransform;

namespace {

thread_local char g_err[512] = "";
template <typename F> int guarded(F&& f) {
	try { f(); return 0; }
	catch (const std::exception& e) { snprintf(g_err, sizeof g_err, "%s", e.what()); return -1; }
	catch (...) { snprintf(g_err, sizeof g_err, "unknown exception"); return -1; }
}

vec3 V3(const float* p) { return vec3(p[0], p[1], p[2]); }
mat3 M3(const float* p) { return mat3(V3(p), V3(p + 3), V3(p + 6)); }
mat4x3 M43(const float* p) { return mat4x3(V3(p), V3(p + 3), V3(p + 6), V3(p + 9)); }

struct RefModel {
	std::shared_ptr<NerfNetwork<network_precision_t>> network;
	GPUMemory<network_precision_t> params;
	GPUMemory<uint8_t> bitfield;
	BoundingBox aabb, render_aabb;
	mat3 render_aabb_to_local = mat3::identity();
	uint32_t max_cascade = 0;
	float cone_angle_constant = 0.0f;
	ENerfActivation rgb_activation = ENerfActivation::Logistic, density_activation = ENerfActivation::Exponential;
	float min_transmittance = 0.01f;
};

struct RefScene {
	std::vector<GPUMemory<TriangleBvhNode>> nodes;
	std::vector<GPUMemory<Triangle>> tris;
	GPUMemory<ObjectTransform> world;
	GPUMemory<Light> lights;
	GPUMemory<Material> materials;
};

// Engine + RayTracer state that survives between frames (engine.cuh:79-120, raytracer.cuh:150-182).
struct RefEngine {
	ivec2 nerf_res{0}, syn_res{0};
	int syn_px_scale = 1;
	GPUMemory<vec4> nerf_fb; GPUMemory<float> nerf_depth;          // CudaRenderBufferView of the NeRF view
	GPUMemory<vec3> nerf_normals, nerf_positions;
	GPUMemory<curandState_t> nerf_rand, syn_rand;
	GPUMemory<vec3> syn_origin, syn_dir;
	GPUMemory<vec4> syn_acc_rgba, syn_fb; GPUMemory<float> syn_acc_depth, syn_fb_depth;
	bool reset_color_buffer = true;
	uint32_t last_n_hit = 0;
	// stock-path camera extras (refdev_engine_set_camera_ex): the arguments Testbed::render_nerf forwards beyond the pinhole camera
	bool has_ex = false;
	mat4x3 camera1; vec4 rolling_shutter{0.0f, 0.0f, 0.0f, 1.0f}; vec3 parallax_shift{0.0f};
	float near_distance = 0.0f, plane_z = 0.0f, aperture_size = 0.0f;
	Lens lens{};
	GPUMemory<vec4> envmap; ivec2 envmap_res{0};
};
struct RefCameraEx { float camera1[12]; float rolling_shutter[4]; float parallax_shift[3]; float near_distance, plane_z, aperture_size; int lens_mode; float lens_params[7]; };

__global__ void k_dump_rand(uint32_t n, const curandState_t* __restrict__ st, uint32_t* __restrict__ out6, float* __restrict__ uniforms, uint32_t n_uniform) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	curandState_t s = st[i];
	for (int k = 0; k < 5; ++k) out6[i * 6 + k] = s.v[k];
	out6[i * 6 + 5] = s.d;
	for (uint32_t k = 0; k < n_uniform; ++k) uniforms[(size_t)i * n_uniform + k] = curand_uniform(&s);
}

__global__ void k_trace_rays(uint32_t n, const float* __restrict__ rays, const ObjectTransform* __restrict__ world, uint32_t n_objects, int* __restrict__ obj_id,
                             int* __restrict__ tri_id, float* __restrict__ t_out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const vec3 o(rays[i * 6], rays[i * 6 + 1], rays[i * 6 + 2]), d(rays[i * 6 + 3], rays[i * 6 + 4], rays[i * 6 + 5]);
	int32_t oid = -1;
	// the HitRecord overload is the one raytrace() uses; the triangle id comes from the plain overload's traversal
	float best = MAX_DEPTH(); int best_tri = -1;
	const vec3 off = o + d * MIN_DEPTH();
	for (uint32_t c = 0; c < n_objects; ++c) {
		const ObjectTransform& obj = world[c];
		auto [ti, hd] = ::ngp::ray_intersect_nodes(off, d, obj.scale, obj.pos, obj.rot, obj.g_node, obj.g_tris);
		if (hd < best && hd > MIN_DEPTH()) { best = hd; best_tri = ti; oid = (int32_t)c; }
	}
	const float t = sng::depth_test_world(o, d, world, n_objects, oid);
	obj_id[i] = oid; tri_id[i] = oid >= 0 ? best_tri : -1; t_out[i] = t;
}

__global__ void k_shadow_dda(uint32_t n, const float* __restrict__ src_dst, uint32_t n_steps, float cone, const uint8_t* __restrict__ grid, uint32_t min_mip, uint32_t max_mip,
                             BoundingBox aabb, mat3 to_local, float* __restrict__ out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const vec3 src(src_dst[i * 6], src_dst[i * 6 + 1], src_dst[i * 6 + 2]), dst(src_dst[i * 6 + 3], src_dst[i * 6 + 4], src_dst[i * 6 + 5]);
	out[i] = sng::depth_test_nerf(n_steps, cone, src, dst, grid, min_mip, max_mip, aabb, to_local);
}

} // namespace

EXPORT const char* refdev_last_error() { return g_err; }

EXPORT int refdev_init(int device) {
	return guarded([&] { CUDA_CHECK_THROW(cudaSetDevice(device)); });
}

// ---- model: NerfNetwork<__half> built as Testbed::reset_network does (src/testbed.cu:3651-3895) ---------------------------------
EXPORT void* refdev_model_create(uint32_t n_levels, uint32_t n_features, uint32_t log2_hashmap_size, uint32_t base_resolution, float per_level_scale,
                                 const uint16_t* params_host, uint64_t n_params, const uint8_t* bitfield_host, const float* aabb6, const float* render_aabb6,
                                 const float* render_aabb_to_local9, uint32_t max_cascade, float cone_angle_constant, int rgb_activation, int density_activation,
                                 float min_transmittance) {
	RefModel* m = nullptr;
	const int rc = guarded([&] {
		using json = nlohmann::json;
		json enc = {{"otype", "HashGrid"}, {"n_levels", n_levels}, {"n_features_per_level", n_features}, {"log2_hashmap_size", log2_hashmap_size},
		            {"base_resolution", base_resolution}, {"per_level_scale", per_level_scale}, {"n_pos_dims", 3}};
		json dir_enc = {{"otype", "Composite"}, {"nested", json::array({json{{"n_dims_to_encode", 3}, {"otype", "SphericalHarmonics"}, {"degree", 4}}, json{{"otype", "Identity"}}})}};
		json net = {{"otype", "FullyFusedMLP"}, {"activation", "ReLU"}, {"output_activation", "None"}, {"n_neurons", 64}, {"n_hidden_layers", 1}};
		json rgb_net = {{"otype", "FullyFusedMLP"}, {"activation", "ReLU"}, {"output_activation", "None"}, {"n_neurons", 64}, {"n_hidden_layers", 2}};
		m = new RefModel();
		m->network = std::make_shared<NerfNetwork<network_precision_t>>(3u, 3u, 0u, 4u, enc, dir_enc, net, rgb_net);
		if (m->network->n_params() != n_params) throw std::runtime_error(fmt::format("n_params mismatch: reference network has {}, caller passed {}", m->network->n_params(), n_params));
		m->params.resize(n_params);
		m->params.copy_from_host((const network_precision_t*)params_host, n_params);
		m->network->set_params((network_precision_t*)m->params.data(), (network_precision_t*)m->params.data(), nullptr);
		if (bitfield_host) { m->bitfield.resize(NERF_GRID_N_CELLS() * NERF_CASCADES() / 8); m->bitfield.copy_from_host(bitfield_host); }
		m->aabb = BoundingBox(V3(aabb6), V3(aabb6 + 3));
		m->render_aabb = BoundingBox(V3(render_aabb6), V3(render_aabb6 + 3));
		m->render_aabb_to_local = M3(render_aabb_to_local9);
		m->max_cascade = max_cascade; m->cone_angle_constant = cone_angle_constant;
		m->rgb_activation = (ENerfActivation)rgb_activation; m->density_activation = (ENerfActivation)density_activation;
		m->min_transmittance = min_transmittance;
	});
	if (rc) { delete m; return nullptr; }
	return m;
}
EXPORT void refdev_model_free(void* p) { delete (RefModel*)p; }
EXPORT uint32_t refdev_padded_output_width(void* p) { return ((RefModel*)p)->network->padded_output_width(); }

// NerfNetwork::inference_mixed_precision (nerf_network.h:105-139): coords n x 7 floats (device, NerfCoordinate AoS), n a multiple of 256;
// out = [padded_output_width x n] halves, row-major as trace() allocates it (row r of sample i at r * n + i).
EXPORT int refdev_network_inference(void* p, const float* coords_dev, uint32_t n, uint16_t* out_dev) {
	RefModel* m = (RefModel*)p;
	return guarded([&] {
		if (n % BATCH_SIZE_GRANULARITY) throw std::runtime_error("n must be a multiple of 256");
		GPUMatrix<float> in((float*)coords_dev, 7, n);
		GPUMatrix<network_precision_t, RM> out((network_precision_t*)out_dev, m->network->padded_output_width(), n);
		m->network->inference_mixed_precision(nullptr, in, out);
		CUDA_CHECK_THROW(cudaStreamSynchronize(nullptr));
	});
}
// GridEncoding forward alone (kernel_grid, grid.h:48-212,765-793): positions taken from the same coords; out = [32 x n] halves SoA
// (feature f of sample i at f * n + i), the layout NerfNetwork feeds the density MLP with.
EXPORT int refdev_grid_encode(void* p, const float* coords_dev, uint32_t n, uint16_t* out_dev) {
	RefModel* m = (RefModel*)p;
	return guarded([&] {
		if (n % BATCH_SIZE_GRANULARITY) throw std::runtime_error("n must be a multiple of 256");
		auto enc = m->network->pos_encoding();
		GPUMatrix<float> in((float*)coords_dev, 7, n);
		GPUMatrixDynamic<network_precision_t> out((network_precision_t*)out_dev, enc->padded_output_width(), n, enc->preferred_output_layout());
		if (enc->preferred_output_layout() != SoA) throw std::runtime_error("unexpected grid output layout");
		enc->inference_mixed_precision(nullptr, in.slice_rows(0, enc->input_width()), out);
		CUDA_CHECK_THROW(cudaStreamSynchronize(nullptr));
	});
}
// Timed network pieces for the CUDA baseline: ms per call of the full network on n samples (average over reps, CUDA events).
EXPORT float refdev_time_network(void* p, const float* coords_dev, uint32_t n, uint16_t* out_dev, int reps, int grid_only) {
	RefModel* m = (RefModel*)p;
	float ms = -1.0f;
	guarded([&] {
		cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
		GPUMatrix<float> in((float*)coords_dev, 7, n);
		GPUMatrix<network_precision_t, RM> out((network_precision_t*)out_dev, m->network->padded_output_width(), n);
		auto enc = m->network->pos_encoding();
		GPUMatrixDynamic<network_precision_t> eout((network_precision_t*)out_dev, enc->padded_output_width(), n, enc->preferred_output_layout());
		for (int i = 0; i < 2; ++i) { if (grid_only) enc->inference_mixed_precision(nullptr, in.slice_rows(0, 3), eout); else m->network->inference_mixed_precision(nullptr, in, out); }
		cudaEventRecord(a, nullptr);
		for (int i = 0; i < reps; ++i) { if (grid_only) enc->inference_mixed_precision(nullptr, in.slice_rows(0, 3), eout); else m->network->inference_mixed_precision(nullptr, in, out); }
		cudaEventRecord(b, nullptr);
		CUDA_CHECK_THROW(cudaEventSynchronize(b));
		cudaEventElapsedTime(&ms, a, b); ms /= reps;
		cudaEventDestroy(a); cudaEventDestroy(b);
	});
	return ms;
}

// grid_to_bitfield + bitfield_max_pool as update_density_grid_mean_and_bitfield drives them (src/testbed_nerf.cu:285-332,3212-3229).
EXPORT int refdev_density_to_bitfield(const float* density_host, uint32_t max_cascade, uint8_t* bitfield_out_host) {
	return guarded([&] {
		const uint32_t n_elements = NERF_GRID_N_CELLS() * (max_cascade + 1);
		GPUMemory<float> grid(NERF_GRID_N_CELLS() * NERF_CASCADES()), mean(1);
		CUDA_CHECK_THROW(cudaMemset(grid.data(), 0, grid.size() * sizeof(float)));
		CUDA_CHECK_THROW(cudaMemcpy(grid.data(), density_host, (size_t)n_elements * sizeof(float), cudaMemcpyHostToDevice));
		GPUMemory<uint8_t> bits(NERF_GRID_N_CELLS() * NERF_CASCADES() / 8);
		CUDA_CHECK_THROW(cudaMemset(mean.data(), 0, sizeof(float)));
		reduce_sum(grid.data(), [n_elements] __device__ (float val) { return fmaxf(val, 0.f) / (n_elements); }, mean.data(), n_elements, nullptr);
		linear_kernel(grid_to_bitfield, 0, nullptr, NERF_GRID_N_CELLS() * NERF_CASCADES() / 8, NERF_GRID_N_CELLS() * (max_cascade + 1) / 8, grid.data(), bits.data(), mean.data());
		for (uint32_t level = 1; level < NERF_CASCADES(); ++level)
			linear_kernel(bitfield_max_pool, 0, nullptr, NERF_GRID_N_CELLS() / 64, bits.data() + grid_mip_offset(level - 1) / 8, bits.data() + grid_mip_offset(level) / 8);
		CUDA_CHECK_THROW(cudaMemcpy(bitfield_out_host, bits.data(), bits.size(), cudaMemcpyDeviceToHost));
	});
}

// ---- scene: what Engine::set_virtual_world / update_world_objects upload (engine.cu:21-127) -------------------------------------
// objects: per object {nodes (n_nodes x 8 floats: bb.min, bb.max, left, right), tris (n_tris x 9), rot9 (column-major), pos3, scale, mat_id}
EXPORT void* refdev_scene_create(uint32_t n_objects, const float* const* nodes, const int32_t* n_nodes, const float* const* tris, const int32_t* n_tris, const float* rot9,
                                 const float* pos3, const float* scale, const int32_t* mat_id, uint32_t n_materials, const float* mats /* n x 13: ka kd ks n rg type spec */,
                                 uint32_t n_lights, const float* lights /* n x 6: pos type intensity size */) {
	RefScene* s = nullptr;
	const int rc = guarded([&] {
		static_assert(sizeof(TriangleBvhNode) == 32 && sizeof(Triangle) == 36, "layouts the product's ABI also assumes");
		s = new RefScene();
		std::vector<ObjectTransform> world;
		s->nodes.resize(n_objects); s->tris.resize(n_objects);
		for (uint32_t i = 0; i < n_objects; ++i) {
			s->nodes[i].resize(n_nodes[i]); s->tris[i].resize(n_tris[i]);
			CUDA_CHECK_THROW(cudaMemcpy(s->nodes[i].data(), nodes[i], (size_t)n_nodes[i] * 32, cudaMemcpyHostToDevice));
			CUDA_CHECK_THROW(cudaMemcpy(s->tris[i].data(), tris[i], (size_t)n_tris[i] * 36, cudaMemcpyHostToDevice));
			world.emplace_back(s->nodes[i].data(), s->tris[i].data(), M3(rot9 + 9 * i), V3(pos3 + 3 * i), scale[i], mat_id[i]);
		}
		std::vector<Material> hm;
		for (uint32_t i = 0; i < n_materials; ++i) {
			const float* p = mats + 13 * i;
			Material mt(i, V3(p + 3), p[10], p[9], (sng::MaterialType)(int)p[11]);
			mt.ka = V3(p); mt.kd = V3(p + 3); mt.ks = V3(p + 6); mt.n = p[9]; mt.rg = p[10]; mt.spec_angle = p[12];
			hm.push_back(mt);
		}
		std::vector<Light> hl;
		for (uint32_t i = 0; i < n_lights; ++i) {
			const float* p = lights + 6 * i;
			nlohmann::json cfg = {{"pos", {p[0], p[1], p[2]}}, {"intensity", p[4]}, {"size", p[5]}, {"type", (int)p[3] == 0 ? "point" : "directional"}};
			hl.emplace_back(i, cfg);
		}
		if (n_objects) s->world.resize_and_copy_from_host(world);
		if (n_materials) s->materials.resize_and_copy_from_host(hm);
		if (n_lights) s->lights.resize_and_copy_from_host(hl);
	});
	if (rc) { delete s; return nullptr; }
	return s;
}
EXPORT void refdev_scene_free(void* p) { delete (RefScene*)p; }

// ---- engine state: Engine::resize (engine.cu:236-255) + RayTracer::enlarge (raytracer.cu:259-281) ---------------------------------
EXPORT void* refdev_engine_create(int nerf_w, int nerf_h, int syn_px_scale) {
	RefEngine* e = nullptr;
	const int rc = guarded([&] {
		e = new RefEngine();
		e->nerf_res = ivec2(nerf_w, nerf_h); e->syn_px_scale = syn_px_scale; e->syn_res = e->nerf_res * syn_px_scale;
		const size_t n = (size_t)nerf_w * nerf_h, ns = next_multiple((size_t)product(e->syn_res), size_t(BATCH_SIZE_GRANULARITY));
		e->nerf_fb.resize(n); e->nerf_depth.resize(n); e->nerf_normals.resize(n); e->nerf_positions.resize(n); e->nerf_rand.resize(n);
		linear_kernel(sng::init_rand_state, 0, nullptr, (uint32_t)n, e->nerf_rand.data());
		e->syn_origin.resize(ns); e->syn_dir.resize(ns); e->syn_acc_rgba.resize(ns); e->syn_acc_depth.resize(ns); e->syn_fb.resize(ns); e->syn_fb_depth.resize(ns); e->syn_rand.resize(ns);
		linear_kernel(sng::init_rand_state, 0, nullptr, (uint32_t)ns, e->syn_rand.data());
		CUDA_CHECK_THROW(cudaStreamSynchronize(nullptr));
	});
	if (rc) { delete e; return nullptr; }
	return e;
}
EXPORT void refdev_engine_free(void* p) { delete (RefEngine*)p; }
EXPORT void refdev_engine_reset_accumulation(void* p) { ((RefEngine*)p)->reset_color_buffer = true; }
EXPORT int refdev_engine_reseed(void* p) {
	RefEngine* e = (RefEngine*)p;
	return guarded([&] {
		linear_kernel(sng::init_rand_state, 0, nullptr, (uint32_t)e->nerf_rand.size(), e->nerf_rand.data());
		linear_kernel(sng::init_rand_state, 0, nullptr, (uint32_t)e->syn_rand.size(), e->syn_rand.data());
		CUDA_CHECK_THROW(cudaStreamSynchronize(nullptr));
	});
}

struct RefView { float camera[12]; float focal_length[2]; float screen_center[2]; uint32_t spp; int snap_to_pixel_centers; };
struct RefRt {
	uint32_t n_steps, samples, bounces, shadow_iters; float lens_angle_constant, depth_offset, syn_shadow_factor; int view_nerf_shadow, buffer_type;
	int nerf_shadow_kernel; float nerf_shadow_intensity, nerf_on_nerf_shadow_threshold; int show_shadow_on_nerf, show_nerf; float exposure; int tonemap_curve;
};

namespace {

// RayTracer::render (raytracer.cu:312-370): arguments as Engine::frame passes them (engine.cu:386-398).
void ref_syn_layer(RefEngine* e, RefModel* m, RefScene* s, const RefView& v, const RefRt& rt, cudaStream_t stream) {
	const ivec2 res = e->syn_res;
	const uint32_t n_elements = product(res);
	const vec2 focal(v.focal_length[0] * e->syn_px_scale, v.focal_length[1] * e->syn_px_scale);
	const mat4x3 cam = M43(v.camera);
	if (e->reset_color_buffer) {
		const dim3 threads = {16, 8, 1};
		const dim3 blocks = {div_round_up((uint32_t)res.x, threads.x), div_round_up((uint32_t)res.y, threads.y), 1};
		sng::init_rays_with_payload_kernel_nerf<<<blocks, threads, 0, stream>>>(v.spp, res, focal, cam, vec2(v.screen_center[0], v.screen_center[1]), e->syn_fb.data(),
		                                                                     e->syn_fb_depth.data(), e->syn_origin.data(), e->syn_dir.data(), e->syn_acc_rgba.data(), e->syn_acc_depth.data());
		e->reset_color_buffer = false;
	}
	linear_kernel(sng::raytrace, 0, stream, n_elements, e->syn_origin.data(), e->syn_dir.data(), cam, s->lights.data(), s->lights.size(), s->materials.data(), s->materials.size(),
	              s->world.data(), s->world.size(), m->cone_angle_constant, rt.lens_angle_constant, (uint32_t)rt.n_steps, (bool)rt.view_nerf_shadow, m->render_aabb,
	              m->render_aabb_to_local, ivec2(focal), m->bitfield.data(), 0u, m->max_cascade, (sng::ImgBufferType)rt.buffer_type, (size_t)rt.samples, (size_t)rt.bounces,
	              (size_t)rt.shadow_iters, 1.0f, rt.depth_offset, e->syn_rand.data(), e->syn_acc_rgba.data(), e->syn_acc_depth.data(), rt.syn_shadow_factor);
}

// Testbed::render sng overload -> render_nerf_with_buffers + shade_nerf_shadows (testbed.cu:4353-4404, testbed_nerf.cu:2467-2677).
// alt = true: the sng path (trace_alt + extract_from_payload + normals + shade_with_shadow); false: the stock path of
// Testbed::render_nerf (trace + shade_kernel_nerf, testbed_nerf.cu:2679-2837).
void ref_nerf_layer(RefEngine* e, RefModel* m, RefScene* s, const RefView& v, const RefRt* rt, bool alt, ERenderMode render_mode, cudaStream_t stream) {
	const ivec2 res = e->nerf_res;
	const vec2 focal(v.focal_length[0], v.focal_length[1]);
	const mat4x3 cam = M43(v.camera);
	Testbed::NerfTracer tracer;
	// Testbed::render clears the frame whenever accumulation restarts (view.render_buffer->clear_frame, testbed.cu:4375-4380); every call
	// here is such a frame. init_rays_with_payload_kernel_nerf only rewrites rgb, so without this the alpha of a recycled allocation leaks in.
	CUDA_CHECK_THROW(cudaMemsetAsync(e->nerf_fb.data(), 0, e->nerf_fb.size() * sizeof(vec4), stream));
	const bool ex = !alt && e->has_ex;
	const mat4x3 cam1 = ex ? e->camera1 : cam;
	const Buffer2DView<const vec4> envmap = (!alt && e->envmap_res.x > 0) ? Buffer2DView<const vec4>{e->envmap.data(), e->envmap_res} : Buffer2DView<const vec4>{};
	tracer.init_rays_from_camera(v.spp, m->network->padded_output_width(), m->network->n_extra_dims(), res, focal, cam, cam1, ex ? e->rolling_shutter : vec4(0.0f, 0.0f, 0.0f, 1.0f),
	                             vec2(v.screen_center[0], v.screen_center[1]), ex ? e->parallax_shift : vec3(0.0f), (bool)v.snap_to_pixel_centers, m->render_aabb, m->render_aabb_to_local,
	                             ex ? e->near_distance : 0.0f, ex ? e->plane_z : 0.0f /* unused without aperture / Slice mode */, ex ? e->aperture_size : 0.0f, Foveation{}, ex ? e->lens : Lens{},
	                             envmap, Buffer2DView<const vec2>{},
	                             e->nerf_fb.data(), e->nerf_depth.data(), Buffer2DView<const uint8_t>{}, m->bitfield.data(), -1, m->max_cascade, m->cone_angle_constant, render_mode, stream);
	const float depth_scale = 1.0f;
	if (alt) {
		CUDA_CHECK_THROW(cudaMemsetAsync(e->nerf_normals.data(), 0, e->nerf_normals.size() * sizeof(vec3), stream));
		CUDA_CHECK_THROW(cudaMemsetAsync(e->nerf_positions.data(), 0, e->nerf_positions.size() * sizeof(vec3), stream));
		const uint32_t n_hit = tracer.trace_alt(m->network, m->render_aabb, m->render_aabb_to_local, m->aabb, focal, m->cone_angle_constant, m->bitfield.data(), render_mode, cam,
		                                        depth_scale, 0, -1, m->rgb_activation, m->density_activation, -1, m->max_cascade, m->min_transmittance, 0.0f, 0, nullptr,
		                                        e->syn_acc_rgba.data(), e->syn_acc_depth.data(), res, (size_t)e->syn_px_scale, stream);
		e->last_n_hit = n_hit;
		RaysNerfSoa& rays_hit = tracer.rays_hit();
		linear_kernel(extract_from_payload, 0, stream, n_hit, cam, rays_hit.rgba, rays_hit.depth, rays_hit.payload, render_mode, e->nerf_fb.data(), e->nerf_depth.data(), e->nerf_positions.data());
		const dim3 threads = {16, 8, 1};
		const dim3 blocks = {div_round_up((uint32_t)res.x, threads.x), div_round_up((uint32_t)res.y, threads.y), 1};
		CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
		write_normals_to_buffer<<<blocks, threads, 0, stream>>>(res, e->nerf_positions.data(), e->nerf_normals.data(), e->nerf_fb.data(), cam[3], render_mode);
		CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
		if (rt && rt->show_shadow_on_nerf) {
			linear_kernel(shade_with_shadow, 0, stream, (uint32_t)product(res), res, cam, e->nerf_fb.data(), e->nerf_depth.data(), e->nerf_positions.data(), e->nerf_normals.data(),
			              s->lights.data(), (uint32_t)s->lights.size(), s->world.data(), (uint32_t)s->world.size(), s->materials.data(), (uint32_t)s->materials.size(), e->nerf_rand.data(),
			              m->bitfield.data(), MAX_STEPS_INBETWEEN_COMPACTION, m->cone_angle_constant, m->max_cascade, m->render_aabb, mat4x3(m->render_aabb_to_local), render_mode,
			              e->nerf_fb.data(), e->nerf_depth.data(), rt->nerf_shadow_intensity, 0.0f, rt->nerf_on_nerf_shadow_threshold, (int)rt->nerf_shadow_kernel, 0.0f, 0, 0.0f);
			CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
		}
	} else {
		// Testbed::render_nerf passes camera_matrix1 to trace() and camera_matrix0 to shade_kernel_nerf (testbed_nerf.cu:2761,2815)
		const uint32_t n_hit = tracer.trace(m->network, m->render_aabb, m->render_aabb_to_local, m->aabb, focal, m->cone_angle_constant, m->bitfield.data(), render_mode, cam1, depth_scale,
		                                    0, -1, m->rgb_activation, m->density_activation, -1, m->max_cascade, m->min_transmittance, 0.0f, 0, nullptr, stream);
		e->last_n_hit = n_hit;
		RaysNerfSoa& rays_hit = tracer.rays_hit();
		linear_kernel(shade_kernel_nerf, 0, stream, n_hit, false /* gbuffer_hard_edges */, cam, depth_scale, rays_hit.rgba, rays_hit.depth, rays_hit.payload, render_mode, false /* train_in_linear_colors */,
		              e->nerf_fb.data(), e->nerf_depth.data());
	}
}

void ref_overlay(RefEngine* e, const RefRt& rt, cudaStream_t stream) {
	const ivec2 res = e->syn_res;
	const dim3 threads = {16, 8, 1};
	const dim3 blocks = {div_round_up((uint32_t)res.x, threads.x), div_round_up((uint32_t)res.y, threads.y), 1};
	sng::overlay_nerf<<<blocks, threads, 0, stream>>>((bool)rt.show_nerf, res, e->syn_px_scale, e->syn_fb.data(), e->syn_fb_depth.data(), e->syn_acc_rgba.data(), e->syn_acc_depth.data(),
	                                               e->nerf_fb.data(), e->nerf_depth.data(), EColorSpace::SRGB, (ETonemapCurve)rt.tonemap_curve, rt.exposure, rt.depth_offset);
}

} // namespace

// Engine::frame minus GL (engine.cu:352-433): RayTracer::render -> Testbed::render (sng) -> RayTracer::overlay, with the syncs the
// reference has between them (RayTracer::sync after each of its launches, Testbed::render syncs at entry and exit).
// stage_ms (may be NULL): [0] mesh layer, [1] NeRF layer incl. shadows, [2] overlay, [3] whole frame (CUDA events on the one stream).
EXPORT int refdev_frame(void* engine, void* model, void* scene, const RefView* view, const RefRt* rt, float* stage_ms) {
	RefEngine* e = (RefEngine*)engine; RefModel* m = (RefModel*)model; RefScene* s = (RefScene*)scene;
	return guarded([&] {
		cudaStream_t st = nullptr;
		cudaEvent_t ev[4];
		for (auto& x : ev) cudaEventCreate(&x);
		cudaEventRecord(ev[0], st);
		ref_syn_layer(e, m, s, *view, *rt, st);
		CUDA_CHECK_THROW(cudaStreamSynchronize(st));
		cudaEventRecord(ev[1], st);
		if (rt->show_nerf) ref_nerf_layer(e, m, s, *view, rt, true, ERenderMode::Shade, st);
		CUDA_CHECK_THROW(cudaStreamSynchronize(st));
		cudaEventRecord(ev[2], st);
		ref_overlay(e, *rt, st);
		cudaEventRecord(ev[3], st);
		CUDA_CHECK_THROW(cudaStreamSynchronize(st));
		if (stage_ms) {
			cudaEventElapsedTime(&stage_ms[0], ev[0], ev[1]); cudaEventElapsedTime(&stage_ms[1], ev[1], ev[2]);
			cudaEventElapsedTime(&stage_ms[2], ev[2], ev[3]); cudaEventElapsedTime(&stage_ms[3], ev[0], ev[3]);
		}
		for (auto& x : ev) cudaEventDestroy(x);
	});
}

// The NeRF layer alone. alt = 1: sng path with positions / normals (and mesh shadows when rt && scene are given); 0: stock render_nerf path.
EXPORT int refdev_render_nerf(void* engine, void* model, void* scene, const RefView* view, const RefRt* rt, int alt, int render_mode, float* ms_out) {
	RefEngine* e = (RefEngine*)engine; RefModel* m = (RefModel*)model; RefScene* s = (RefScene*)scene;
	return guarded([&] {
		cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
		cudaEventRecord(a, nullptr);
		ref_nerf_layer(e, m, s, *view, (rt && s) ? rt : nullptr, alt != 0, (ERenderMode)render_mode, nullptr);
		cudaEventRecord(b, nullptr);
		CUDA_CHECK_THROW(cudaEventSynchronize(b));
		if (ms_out) cudaEventElapsedTime(ms_out, a, b);
		cudaEventDestroy(a); cudaEventDestroy(b);
	});
}
EXPORT int refdev_render_syn(void* engine, void* model, void* scene, const RefView* view, const RefRt* rt) {
	RefEngine* e = (RefEngine*)engine;
	return guarded([&] { ref_syn_layer(e, (RefModel*)model, (RefScene*)scene, *view, *rt, nullptr); CUDA_CHECK_THROW(cudaStreamSynchronize(nullptr)); });
}

// Buffers of the last render, copied to host: which = 0 final rgba (syn res x 4), 1 final depth, 2 nerf rgba, 3 nerf depth, 4 positions (x3), 5 normals (x3),
// 6 syn acc rgba, 7 syn acc depth.
EXPORT int refdev_engine_read(void* engine, int which, float* out_host) {
	RefEngine* e = (RefEngine*)engine;
	return guarded([&] {
		const size_t n = (size_t)product(e->nerf_res), ns = (size_t)product(e->syn_res);
		const void* src = nullptr; size_t bytes = 0;
		switch (which) {
			case 0: src = e->syn_fb.data(); bytes = ns * 16; break;
			case 1: src = e->syn_fb_depth.data(); bytes = ns * 4; break;
			case 2: src = e->nerf_fb.data(); bytes = n * 16; break;
			case 3: src = e->nerf_depth.data(); bytes = n * 4; break;
			case 4: src = e->nerf_positions.data(); bytes = n * 12; break;
			case 5: src = e->nerf_normals.data(); bytes = n * 12; break;
			case 6: src = e->syn_acc_rgba.data(); bytes = ns * 16; break;
			case 7: src = e->syn_acc_depth.data(); bytes = ns * 4; break;
			default: throw std::runtime_error("bad buffer id");
		}
		CUDA_CHECK_THROW(cudaMemcpy(out_host, src, bytes, cudaMemcpyDeviceToHost));
	});
}
EXPORT uint32_t refdev_engine_last_n_hit(void* engine) { return ((RefEngine*)engine)->last_n_hit; }

// Camera extras + environment map of the following stock-path renders (refdev_render_nerf with alt = 0). ex = NULL: pinhole; envmap = NULL: none.
EXPORT int refdev_engine_set_camera_ex(void* engine, const RefCameraEx* ex, const float* envmap_rgba_host, int env_w, int env_h) {
	RefEngine* e = (RefEngine*)engine;
	return guarded([&] {
		e->has_ex = ex != nullptr;
		if (ex) {
			e->camera1 = M43(ex->camera1);
			e->rolling_shutter = vec4(ex->rolling_shutter[0], ex->rolling_shutter[1], ex->rolling_shutter[2], ex->rolling_shutter[3]);
			e->parallax_shift = V3(ex->parallax_shift);
			e->near_distance = ex->near_distance; e->plane_z = ex->plane_z; e->aperture_size = ex->aperture_size;
			e->lens = Lens{}; e->lens.mode = (ELensMode)ex->lens_mode;
			for (int i = 0; i < 7; ++i) e->lens.params[i] = ex->lens_params[i];
		}
		if (envmap_rgba_host && env_w > 0 && env_h > 0) {
			e->envmap.resize((size_t)env_w * env_h);
			CUDA_CHECK_THROW(cudaMemcpy(e->envmap.data(), envmap_rgba_host, (size_t)env_w * env_h * sizeof(vec4), cudaMemcpyHostToDevice));
			e->envmap_res = ivec2(env_w, env_h);
		} else e->envmap_res = ivec2(0);
	});
}

// ---- small taps -----------------------------------------------------------------------------------------------------------------------
// Ray set-up + start jitter exactly as NerfTracer::init_rays_from_camera leaves the payloads (testbed_nerf.cu:2037-2120): out n x 8 floats
// {origin3, dir3, t, alive}.
static int init_rays_impl(RefModel* m, const RefView* v, int w, int h, const RefCameraEx* ex, float* out_host);
EXPORT int refdev_init_rays(void* model, const RefView* v, int w, int h, float* out_host) { return init_rays_impl((RefModel*)model, v, w, h, nullptr, out_host); }
// The same with the camera extras of uv_to_ray (lens, aperture, parallax, near plane, end-of-shutter camera).
EXPORT int refdev_init_rays_ex(void* model, const RefView* v, int w, int h, const RefCameraEx* ex, float* out_host) { return init_rays_impl((RefModel*)model, v, w, h, ex, out_host); }
static int init_rays_impl(RefModel* m, const RefView* v, int w, int h, const RefCameraEx* ex, float* out_host) {
	return guarded([&] {
		const size_t n = (size_t)w * h;
		GPUMemory<vec4> fb(n); GPUMemory<float> depth(n);
		Testbed::NerfTracer tracer;
		const mat4x3 cam = M43(v->camera);
		Lens lens{};
		if (ex) { lens.mode = (ELensMode)ex->lens_mode; for (int i = 0; i < 7; ++i) lens.params[i] = ex->lens_params[i]; }
		tracer.init_rays_from_camera(v->spp, m->network->padded_output_width(), 0, ivec2(w, h), vec2(v->focal_length[0], v->focal_length[1]), cam, ex ? M43(ex->camera1) : cam,
		                             ex ? vec4(ex->rolling_shutter[0], ex->rolling_shutter[1], ex->rolling_shutter[2], ex->rolling_shutter[3]) : vec4(0.0f, 0.0f, 0.0f, 1.0f),
		                             vec2(v->screen_center[0], v->screen_center[1]), ex ? V3(ex->parallax_shift) : vec3(0.0f), (bool)v->snap_to_pixel_centers, m->render_aabb, m->render_aabb_to_local,
		                             ex ? ex->near_distance : 0.0f, ex ? ex->plane_z : 0.0f, ex ? ex->aperture_size : 0.0f,
		                             Foveation{}, lens, Buffer2DView<const vec4>{}, Buffer2DView<const vec2>{}, fb.data(), depth.data(), Buffer2DView<const uint8_t>{}, m->bitfield.data(), -1,
		                             m->max_cascade, m->cone_angle_constant, ERenderMode::Shade, nullptr);
		std::vector<NerfPayload> pl(n);
		CUDA_CHECK_THROW(cudaMemcpy(pl.data(), tracer.rays_init().payload, n * sizeof(NerfPayload), cudaMemcpyDeviceToHost));
		for (size_t i = 0; i < n; ++i) {
			float* o = out_host + i * 8;
			o[0] = pl[i].origin.x; o[1] = pl[i].origin.y; o[2] = pl[i].origin.z; o[3] = pl[i].dir.x; o[4] = pl[i].dir.y; o[5] = pl[i].dir.z; o[6] = pl[i].t; o[7] = pl[i].alive ? 1.0f : 0.0f;
		}
	});
}

// generate_next_nerf_network_inputs (testbed_nerf.cu:790-837) for explicit rays: in n x 7 {origin3, dir3, t}; out n_steps NerfCoordinates per ray
// (n x n_steps x 7 floats: warped pos3, warped dt, warped dir3; zeros past the ray's end), steps_out n, t_out n (payload.t afterwards).
EXPORT int refdev_march_samples(void* model, const float* rays_host, uint32_t n, uint32_t n_steps, float* coords_out_host, uint32_t* steps_out_host, float* t_out_host) {
	RefModel* m = (RefModel*)model;
	return guarded([&] {
		std::vector<NerfPayload> pl(n);
		for (uint32_t i = 0; i < n; ++i) {
			const float* r = rays_host + (size_t)i * 7;
			pl[i].origin = V3(r); pl[i].dir = V3(r + 3); pl[i].t = r[6]; pl[i].max_weight = 0.0f; pl[i].idx = i; pl[i].n_steps = 0; pl[i].alive = true;
		}
		GPUMemory<NerfPayload> d_pl(n); d_pl.copy_from_host(pl);
		GPUMemory<float> coords((size_t)n * n_steps * 7);
		CUDA_CHECK_THROW(cudaMemset(coords.data(), 0, coords.size() * sizeof(float)));
		PitchedPtr<NerfCoordinate> input_data((NerfCoordinate*)coords.data(), 1, 0, 0);
		linear_kernel(generate_next_nerf_network_inputs, 0, nullptr, n, m->render_aabb, m->render_aabb_to_local, m->aabb, vec2(1.0f), vec3(0.0f, 0.0f, 1.0f), d_pl.data(), input_data,
		              n_steps, m->bitfield.data(), 0u, m->max_cascade, m->cone_angle_constant, (const float*)nullptr);
		CUDA_CHECK_THROW(cudaDeviceSynchronize());
		d_pl.copy_to_host(pl);
		std::vector<float> c(coords.size()); coords.copy_to_host(c);
		// the kernel writes step j of ray i at column i + j * n: regroup per ray
		for (uint32_t i = 0; i < n; ++i) {
			for (uint32_t j = 0; j < n_steps; ++j) memcpy(coords_out_host + ((size_t)i * n_steps + j) * 7, c.data() + ((size_t)i + (size_t)j * n) * 7, 28);
			steps_out_host[i] = pl[i].n_steps; t_out_host[i] = pl[i].t;
		}
	});
}

EXPORT int refdev_rand_states(uint32_t n, uint32_t n_uniform, uint32_t* states6_host, float* uniforms_host) {
	return guarded([&] {
		GPUMemory<curandState_t> st(n); GPUMemory<uint32_t> out(n * 6); GPUMemory<float> un((size_t)n * std::max(1u, n_uniform));
		linear_kernel(sng::init_rand_state, 0, nullptr, n, st.data());
		k_dump_rand<<<div_round_up(n, 128u), 128>>>(n, st.data(), out.data(), un.data(), n_uniform);
		CUDA_CHECK_THROW(cudaDeviceSynchronize());
		out.copy_to_host(states6_host, n * 6);
		if (n_uniform) un.copy_to_host(uniforms_host, (size_t)n * n_uniform);
	});
}

EXPORT int refdev_trace_rays(void* scene, const float* rays_host, uint32_t n, int32_t* obj_id_host, int32_t* tri_id_host, float* t_host) {
	RefScene* s = (RefScene*)scene;
	return guarded([&] {
		GPUMemory<float> rays((size_t)n * 6), t(n); GPUMemory<int> oid(n), tid(n);
		rays.copy_from_host(rays_host, (size_t)n * 6);
		k_trace_rays<<<div_round_up(n, 128u), 128>>>(n, rays.data(), s->world.data(), (uint32_t)s->world.size(), oid.data(), tid.data(), t.data());
		CUDA_CHECK_THROW(cudaDeviceSynchronize());
		oid.copy_to_host(obj_id_host, n); tid.copy_to_host(tri_id_host, n); t.copy_to_host(t_host, n);
	});
}

EXPORT int refdev_shadow_dda(void* model, const float* src_dst_host, uint32_t n, uint32_t n_steps, float* out_host) {
	RefModel* m = (RefModel*)model;
	return guarded([&] {
		GPUMemory<float> sd((size_t)n * 6), out(n);
		sd.copy_from_host(src_dst_host, (size_t)n * 6);
		k_shadow_dda<<<div_round_up(n, 128u), 128>>>(n, sd.data(), n_steps, m->cone_angle_constant, m->bitfield.data(), 0u, m->max_cascade, m->render_aabb, m->render_aabb_to_local, out.data());
		CUDA_CHECK_THROW(cudaDeviceSynchronize());
		out.copy_to_host(out_host, n);
	});
}

} // namespace refdev
