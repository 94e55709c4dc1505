// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
// NeRF layer of the frame: marching loop, compositing, payload
// extraction, screen-space normals, accumulate + tonemap.
//
// Batch schedule. The reference marches all alive rays in lock-step batches of
// S = clamp(2^21 / n_alive, 1, 8) samples (testbed_nerf.cu:2189-2190). In the stock path
// (trace + composite_kernel_nerf) the per-ray result does not depend on S. In the sng path
// (trace_alt + composite_kernel_nerf_alt) it does: payload.t is rewound to the last composited sample
// after every batch (testbed_nerf.cu:574), so the last sample of a batch is evaluated again as the
// first sample of the next. S == 8 whenever n_alive <= 262144, i.e. always for frames or tiles of at
// most 512x512 rays. This oracle takes the schedule as a parameter: steps_per_batch = 8 reproduces the
// reference for <= 262144 rays per render call and defines the tile-wise semantics for larger frames;
// orc_render_nerf_wavefront() replays the reference's global schedule for cross-checks.
#include "march.h"
#include "network.h"
#include <vector>

namespace orc {

constexpr uint32_t MARCH_ITER = 10000; // testbed_nerf.cu:47

static float logistic(float x) { return 1.0f / (1.0f + std::exp(-x)); } // tcnn/common_device.h:42-44
// nerf_device.cuh:204-213
static float network_to_rgb(float val, int act) {
	switch (act) {
		case ORC_ACT_NONE: return val;
		case ORC_ACT_RELU: return val > 0.0f ? val : 0.0f;
		case ORC_ACT_LOGISTIC: return logistic(val);
		case ORC_ACT_EXPONENTIAL: return std::exp(std::min(std::max(val, -10.0f), 10.0f));
	}
	return 0.0f;
}
// nerf_device.cuh:234-243
static float network_to_density(float val, int act) {
	switch (act) {
		case ORC_ACT_NONE: return val;
		case ORC_ACT_RELU: return val > 0.0f ? val : 0.0f;
		case ORC_ACT_LOGISTIC: return logistic(val);
		case ORC_ACT_EXPONENTIAL: return std::exp(val);
	}
	return 0.0f;
}
// nerf_device.cuh:306-314
static float warp_dt(float dt) { float max_stepsize = (1.73205080757f / 1024) * (1 << 7); return (dt - (1.73205080757f / 1024)) / (max_stepsize - (1.73205080757f / 1024)); }
static float unwarp_dt(float dt) { float max_stepsize = (1.73205080757f / 1024) * (1 << 7); return dt * (max_stepsize - (1.73205080757f / 1024)) + (1.73205080757f / 1024); }

// common_device.cuh:35-41, 59-65
float srgb_to_linear(float s) { return s <= 0.04045f ? s / 12.92f : std::pow((s + 0.055f) / 1.055f, 2.4f); }
float linear_to_srgb(float l) { return l < 0.0031308f ? 12.92f * l : 1.055f * std::pow(l, 0.41666f) - 0.055f; }

struct RayState { NerfPayload p; vec4 rgba; float depth; };

struct Sample { vec3 warped_pos; float warped_dt; };

// generate_next_nerf_network_inputs (testbed_nerf.cu:790-837) for one ray; returns samples produced.
static uint32_t generate_samples(RayState& r, const MarchEnv& env, const BoundingBox& train_aabb, float cone, uint32_t n_steps, Sample* out) {
	NerfPayload& payload = r.p;
	vec3 origin = payload.origin, dir = payload.dir;
	vec3 idir = 1.0f / dir;
	float t = payload.t;
	for (uint32_t j = 0; j < n_steps; ++j) {
		t = if_unoccupied_advance_to_next_occupied_voxel(t, cone, {origin, dir}, idir, env);
		if (t >= MAX_DEPTH) { payload.n_steps = j; return j; }
		float dt = calc_dt(t, cone);
		out[j].warped_pos = train_aabb.relative_pos(origin + dir * t); // warp_position, nerf_device.cuh:266-271
		out[j].warped_dt = warp_dt(dt);
		t += dt;
	}
	payload.t = t;
	payload.n_steps = n_steps;
	return n_steps;
}

// composite_kernel_nerf (testbed_nerf.cu:577-788) / composite_kernel_nerf_alt (:476-575) for one ray,
// render modes Shade / Depth / Positions, no glow, no show_accel.
static void composite(RayState& r, const Sample* samples, const half_t* net_out /* n x 4 */, uint32_t n_steps, uint32_t current_step,
                      const BoundingBox& train_aabb, const mat4x3& cam, const orc_model& m, int render_mode, bool alt, float depth_scale) {
	NerfPayload& payload = r.p;
	vec3 cam_fwd = cam.c[2];
	vec4 local_rgba = r.rgba;
	float local_depth = r.depth;
	vec3 origin = payload.origin;
	uint32_t actual_n_steps = payload.n_steps;
	uint32_t j = 0;
	for (; j < actual_n_steps; ++j) {
		const half_t* o = net_out + j * 4;
		vec3 warped_pos = samples[j].warped_pos;
		vec3 pos = train_aabb.min + warped_pos * train_aabb.diag(); // unwarp_position, nerf_device.cuh:273-278
		float T = 1.f - local_rgba.w;
		float dt = unwarp_dt(samples[j].warped_dt);
		float alpha = 1.f - std::exp(-network_to_density(h2f(o[3]), m.density_activation) * dt);
		float weight = alpha * T;
		vec3 rgb = {network_to_rgb(h2f(o[0]), m.rgb_activation), network_to_rgb(h2f(o[1]), m.rgb_activation), network_to_rgb(h2f(o[2]), m.rgb_activation)};
		if (!alt) {
			if (render_mode == ORC_MODE_POSITIONS) rgb = (pos - 0.5f) / 2.0f + 0.5f;
			else if (render_mode == ORC_MODE_DEPTH) rgb = V3(dot(cam_fwd, pos - origin) * depth_scale);
		}
		local_rgba.x += rgb.x * weight; local_rgba.y += rgb.y * weight; local_rgba.z += rgb.z * weight; local_rgba.w += weight;
		if (alt) {
			local_depth = dot(cam_fwd, pos - cam.c[3]);
			if (weight > payload.max_weight) payload.max_weight = weight;
		} else if (weight > payload.max_weight) {
			payload.max_weight = weight;
			local_depth = dot(cam_fwd, pos - cam.c[3]);
		}
		if (local_rgba.w > (1.0f - m.min_transmittance)) {
			float a = local_rgba.w;
			local_rgba.x /= a; local_rgba.y /= a; local_rgba.z /= a; local_rgba.w /= a;
			break;
		}
	}
	if (j < n_steps) {
		payload.alive = false;
		payload.n_steps = j + current_step;
	}
	r.rgba = local_rgba;
	r.depth = local_depth;
	if (alt) payload.t = local_depth / dot(cam_fwd, payload.dir); // testbed_nerf.cu:574
}

static void run_batch(RayState& r, const MarchEnv& env, const NetworkView& net, const BoundingBox& train_aabb, const mat4x3& cam,
                      const orc_model& m, uint32_t S, uint32_t current_step, int render_mode, bool alt, uint64_t* sample_counter) {
	Sample samples[8];
	half_t net_out[8 * 4];
	uint32_t n = generate_samples(r, env, train_aabb, m.cone_angle_constant, S, samples);
	vec3 wdir = (r.p.dir + 1.0f) * 0.5f; // warp_direction, nerf_device.cuh:290-292
	for (uint32_t j = 0; j < n; ++j) network_query(net, samples[j].warped_pos, wdir, net_out + j * 4, nullptr);
	if (sample_counter) *sample_counter += n;
	composite(r, samples, net_out, S, current_step, train_aabb, cam, m, render_mode, alt, 1.0f);
}

// extract_from_payload (testbed_nerf.cu:1578-1612, alt) / shade_kernel_nerf (:1788-1828, stock; linear_colors = false).
static void write_hit(const RayState& r, const mat4x3& cam, int render_mode, bool alt, float* fb, float* depth_buffer, float* positions) {
	const NerfPayload& payload = r.p;
	vec4 tmp = r.rgba;
	uint32_t px = payload.idx;
	if (alt) {
		vec3 orig_pos = cam.c[3] + payload.dir * payload.t;
		if (render_mode == ORC_MODE_DEPTH) { float v = (float)std::max(0.0, 1.0 - (double)(r.depth / 3.0f)); tmp.x = tmp.y = tmp.z = v; }
		else if (render_mode == ORC_MODE_POSITIONS) {
			// sng::vec3_to_col (common.cu:287-289): v * 0.5 + 0.5
			tmp.x = srgb_to_linear(orig_pos.x * 0.5f + 0.5f); tmp.y = srgb_to_linear(orig_pos.y * 0.5f + 0.5f); tmp.z = srgb_to_linear(orig_pos.z * 0.5f + 0.5f);
		} else { tmp.x = srgb_to_linear(tmp.x); tmp.y = srgb_to_linear(tmp.y); tmp.z = srgb_to_linear(tmp.z); }
		if (positions) { positions[px * 3 + 0] = orig_pos.x; positions[px * 3 + 1] = orig_pos.y; positions[px * 3 + 2] = orig_pos.z; }
	} else {
		if (render_mode == ORC_MODE_SHADE) { tmp.x = srgb_to_linear(tmp.x); tmp.y = srgb_to_linear(tmp.y); tmp.z = srgb_to_linear(tmp.z); }
	}
	float* f = fb + (size_t)px * 4;
	float k = 1.0f - tmp.w;
	f[0] = tmp.x + f[0] * k; f[1] = tmp.y + f[1] * k; f[2] = tmp.z + f[2] * k; f[3] = tmp.w + f[3] * k;
	if (tmp.w > 0.2f) depth_buffer[px] = r.depth;
}

} // namespace orc

using namespace orc;

extern "C" {

// Per-ray schedule with a fixed batch size (see the header comment). fb (RGBA, in/out) and depth are
// width*height; positions (alt only, may be NULL) is width*height*3 and is zero-filled first like
// render_nerf_with_buffers does (testbed_nerf.cu:2538-2539). Returns the number of network samples.
static uint64_t render_nerf_per_ray(const orc_model* model, const orc_view* view, int alt, int render_mode, const uint32_t* sched, uint32_t n_sched, uint32_t tail_steps,
                                    float* fb, float* depth, float* positions, uint16_t* n_steps_out);

uint64_t orc_render_nerf(const orc_model* model, const orc_view* view, int alt, int render_mode, uint32_t steps_per_batch,
                         float* fb, float* depth, float* positions, uint16_t* n_steps_out) {
	return render_nerf_per_ray(model, view, alt, render_mode, nullptr, 0, steps_per_batch, fb, depth, positions, n_steps_out);
}

// The same per-ray loop with an explicit batch-length sequence (batch b uses sched[b], 8 past the end): what the product's march
// evaluates once its sequence is the one the reference's lock-step loop would have produced (orc_render_nerf_wavefront's schedule_out).
uint64_t orc_render_nerf_sched(const orc_model* model, const orc_view* view, int alt, int render_mode, const uint32_t* sched, uint32_t n_sched,
                               float* fb, float* depth, float* positions) {
	return render_nerf_per_ray(model, view, alt, render_mode, sched, n_sched, 8, fb, depth, positions, nullptr);
}

}

static uint64_t render_nerf_per_ray(const orc_model* model, const orc_view* view, int alt, int render_mode, const uint32_t* sched, uint32_t n_sched, uint32_t steps_per_batch,
                                    float* fb, float* depth, float* positions, uint16_t* n_steps_out) {
	MarchEnv env = make_env(*model);
	NetworkView net = make_network_view(*model);
	BoundingBox train_aabb = load_aabb(model->aabb_min, model->aabb_max);
	mat4x3 cam = alt ? load_camera(view->camera) : compositor_camera(*view);   // trace() gets camera_matrix1 (testbed_nerf.cu:2761)
	const uint32_t N = (uint32_t)view->width * view->height;
	if (positions) std::memset(positions, 0, sizeof(float) * 3 * N);
	uint64_t n_samples = 0;
	const uint32_t S = steps_per_batch;
	for (int y = 0; y < view->height; ++y) for (int x = 0; x < view->width; ++x) {
		RayState r;
		init_ray_payload(x, y, *view, env, r.p);
		uint32_t idx = r.p.idx;
		init_ray_background(x, y, *view, env, r.p, fb + (size_t)idx * 4, depth + idx); // rgb clear or envmap texel, testbed_nerf.cu:1946-1952
		r.rgba = {0, 0, 0, 0}; r.depth = 0.0f;
		advance_pos_nerf(r.p, env, view->spp, model->cone_angle_constant);
		bool finished = !r.p.alive; // rays dead before the first iteration are compacted in iteration 1
		uint32_t b = 0;
		for (uint32_t i = 1; r.p.alive && i < MARCH_ITER; ++b) {
			const uint32_t Sb = b < n_sched ? sched[b] : S;
			run_batch(r, env, net, train_aabb, cam, *model, Sb, i, render_mode, alt != 0, &n_samples);
			// A ray that dies in the last possible iteration is never compacted into the hit list (:2165-2271).
			if (!r.p.alive && i + Sb < MARCH_ITER) finished = true;
			i += Sb;
		}
		if (n_steps_out) n_steps_out[idx] = r.p.n_steps;
		if (finished && r.rgba.w > 0.001f) write_hit(r, cam, render_mode, alt != 0, fb, depth, positions);
	}
	return n_samples;
}

extern "C" {

// Replay of the reference's global wavefront schedule (trace / trace_alt, testbed_nerf.cu:2128-2401):
// all rays advance batch by batch and S is recomputed from the global alive count.
// target_queries = 2 * 1024 * 1024 in the reference (:2188); alive_out (may be NULL) receives n_alive of every iteration.
uint64_t orc_render_nerf_wavefront(const orc_model* model, const orc_view* view, int alt, int render_mode, uint32_t target_queries,
                                   float* fb, float* depth, float* positions, uint32_t* schedule_out, uint32_t* alive_out, uint32_t schedule_cap) {
	MarchEnv env = make_env(*model);
	NetworkView net = make_network_view(*model);
	BoundingBox train_aabb = load_aabb(model->aabb_min, model->aabb_max);
	mat4x3 cam = alt ? load_camera(view->camera) : compositor_camera(*view);   // trace() gets camera_matrix1 (testbed_nerf.cu:2761)
	const uint32_t N = (uint32_t)view->width * view->height;
	if (positions) std::memset(positions, 0, sizeof(float) * 3 * N);
	std::vector<RayState> rays(N);
	for (int y = 0; y < view->height; ++y) for (int x = 0; x < view->width; ++x) {
		RayState& r = rays[x + view->width * y];
		init_ray_payload(x, y, *view, env, r.p);
		init_ray_background(x, y, *view, env, r.p, fb + (size_t)r.p.idx * 4, depth + r.p.idx);
		r.rgba = {0, 0, 0, 0}; r.depth = 0.0f;
		advance_pos_nerf(r.p, env, view->spp, model->cone_angle_constant);
	}
	std::vector<uint32_t> alive, next;
	std::vector<uint8_t> hit(N, 0);
	for (uint32_t k = 0; k < N; ++k) alive.push_back(k);
	uint64_t n_samples = 0;
	uint32_t it = 0;
	uint32_t i = 1;
	while (i < MARCH_ITER) {
		next.clear();
		for (uint32_t k : alive) { // compact_kernel_nerf, testbed_nerf.cu:1830-1853
			if (rays[k].p.alive) next.push_back(k);
			else if (rays[k].rgba.w > 0.001f) hit[k] = 1;
		}
		alive.swap(next);
		uint32_t n_alive = (uint32_t)alive.size();
		if (n_alive == 0) break;
		uint32_t S = std::min(std::max(target_queries / n_alive, 1u), 8u);
		if (schedule_out && it < schedule_cap) schedule_out[it] = S;
		if (alive_out && it < schedule_cap) alive_out[it] = n_alive;
		++it;
		for (uint32_t k : alive) run_batch(rays[k], env, net, train_aabb, cam, *model, S, i, render_mode, alt != 0, &n_samples);
		i += S;
	}
	for (uint32_t k = 0; k < N; ++k) if (hit[k]) write_hit(rays[k], cam, render_mode, alt != 0, fb, depth, positions);
	return n_samples;
}

// write_normals_to_buffer (testbed_nerf.cu:1523-1576) with sng::get_normal (common.cuh:55-57).
void orc_write_normals(int width, int height, const float* positions, float* normals) {
	static const int offs[9][2] = {{1, 0}, {0, 1}, {-1, 0}, {0, -1}, {2, 0}, {0, 2}, {-2, 0}, {0, -2}, {1, 0}};
	for (int y = 0; y < height; ++y) for (int x = 0; x < width; ++x) {
		uint32_t idx = x + width * y;
		vec3 pos = {positions[idx * 3], positions[idx * 3 + 1], positions[idx * 3 + 2]};
		float factor = 0.0f;
		vec3 N = {0, 0, 0};
		for (int t = 0; t < 8; ++t) {
			int tx = x + offs[t + 1][0], ty = y + offs[t + 1][1], bx = x + offs[t][0], by = y + offs[t][1];
			if (tx >= width || tx < 0 || ty >= height || ty < 0 || bx >= width || bx < 0 || by >= height || by < 0) continue;
			const float* tp = positions + (size_t)(ty * width + tx) * 3;
			const float* bp = positions + (size_t)(by * width + bx) * 3;
			vec3 T = vec3{tp[0], tp[1], tp[2]} - pos;
			vec3 B = vec3{bp[0], bp[1], bp[2]} - pos;
			N += normalize(cross(normalize(T), B));
			factor += 1.0f;
		}
		N = factor == 0.0f ? N : N / factor;
		N = normalize(N);
		normals[idx * 3] = N.x; normals[idx * 3 + 1] = N.y; normals[idx * 3 + 2] = N.z;
	}
}

// CudaRenderBuffer::accumulate + tonemap (render_buffer.cu:268-302, 304-379, 569-601, 671-735) for one
// sample: color_space Linear/SRGB accumulate, background blend, exposure, tonemap curve, output space.
void orc_tonemap_curve(float* rgb, int curve); // sng.cpp (same polynomial as render_buffer.cu:304-361)

void orc_accumulate_tonemap(int n_pixels, const float* frame_buffer, float* accumulate_buffer, uint32_t spp,
                            int color_space_srgb, int output_srgb, float exposure, const float* background_rgba, int tonemap_curve, float* out_rgba) {
	float sample_count = (float)spp;
	for (int i = 0; i < n_pixels; ++i) {
		const float* c = frame_buffer + (size_t)i * 4;
		float* a = accumulate_buffer + (size_t)i * 4;
		if (spp == 0) a[0] = a[1] = a[2] = a[3] = 0.0f;
		float col[3] = {c[0], c[1], c[2]};
		if (color_space_srgb) for (int k = 0; k < 3; ++k) col[k] = linear_to_srgb(col[k]);
		for (int k = 0; k < 3; ++k) a[k] = (a[k] * sample_count + col[k]) / (sample_count + 1);
		a[3] = (a[3] * sample_count + c[3]) / (sample_count + 1);
		float bg[4] = {background_rgba[0], background_rgba[1], background_rgba[2], background_rgba[3]};
		if (!color_space_srgb) for (int k = 0; k < 3; ++k) bg[k] = srgb_to_linear(bg[k]);
		float color[4] = {a[0], a[1], a[2], a[3]};
		float weight = (1 - color[3]) * bg[3];
		for (int k = 0; k < 3; ++k) color[k] += bg[k] * weight;
		color[3] += weight;
		if (color_space_srgb) for (int k = 0; k < 3; ++k) color[k] = srgb_to_linear(color[k]);
		float e = std::pow(2.0f, exposure);
		for (int k = 0; k < 3; ++k) color[k] *= e;
		orc_tonemap_curve(color, tonemap_curve);
		if (output_srgb) for (int k = 0; k < 3; ++k) color[k] = linear_to_srgb(color[k]);
		for (int k = 0; k < 4; ++k) out_rgba[(size_t)i * 4 + k] = color[k];
	}
}

}
