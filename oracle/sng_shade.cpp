// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
// The synerfgine shading layer: sng::raytrace + shade_object, mesh shadows
// on the NeRF (shade_with_shadow), and the depth-tested overlay. All device-only code in the reference.
#include "sng_common.h"
#include <vector>

namespace orc {

constexpr float PI_F = 3.14159265358979323846f; // tcnn/common_device.h:39

// include/synerfgine/common.cuh:30-32
static vec3 reflect(vec3 incident, vec3 normal) { return 2.0f * dot(incident, normal) * normal - incident; }
// common.cuh:33-36 (frame given as a matrix)
static vec3 cone_random_frame(vec3 orig, const mat3& frame, float longi, float latid) {
	vec3 offset = {std::cos(longi) * std::sin(latid), std::sin(longi) * std::sin(latid), std::cos(longi)};
	return orig + frame * offset;
}
// common.cuh:37-48 (frame built from an up vector)
static vec3 cone_random_up(vec3 orig, vec3 up, float longi, float latid) {
	vec3 N = normalize(orig);
	vec3 B = normalize(cross(N, up));
	vec3 T = cross(B, N);
	mat3 frame = {{T, B, N}};
	vec3 offset = {std::sin(longi) * std::cos(latid), std::sin(longi) * std::sin(latid), std::cos(longi)};
	return orig + frame * offset;
}
static float fractf(float x) { return x - std::floor(x); }
static float smoothstep1(float v) { return v * v * (3.0f - 2.0f * v); } // tcnn/common_device.h:801-803

// light.cuh:71-77
static vec3 light_sample(const orc_light& l, Xorwow& rng) {
	float a = fractf(xorwow_uniform(rng));
	float b = fractf(xorwow_uniform(rng));
	float c = fractf(xorwow_uniform(rng));
	vec3 offset = {a, b, c};
	return vec3{l.pos[0], l.pos[1], l.pos[2]} + offset * l.size * 1.0f;
}

// material.cuh:96-98
static vec3 local_color(const orc_material& m, vec3 L, vec3 N, vec3 R, vec3 V, const orc_light& light) {
	vec3 kd = {m.kd[0], m.kd[1], m.kd[2]}, ks = {m.ks[0], m.ks[1], m.ks[2]};
	return std::max(0.0f, dot(L, N)) * kd * light.intensity + std::pow(std::max(0.0f, dot(R, V)), m.n) * ks;
}

struct SampledRay { vec3 pos{0, 0, 0}, dir{0, 0, 0}; float pdf = 0.0f, attenuation = 1.0f; }; // common.cuh:130-135

// material.cuh:111-123
static void scatter(const orc_material& m, const HitRecord& hit, vec3 src_dir, SampledRay& ray, Xorwow& rng) {
	ray.pos = hit.pos;
	if (m.type != 0 && m.type != 1) return;
	float specular_angle = m.type == 0 ? PI_F / 2 : m.spec_angle;
	ray.dir = reflect(-src_dir, hit.normal);
	float longi = xorwow_uniform(rng) * specular_angle;
	double latid = xorwow_uniform(rng) * 2.0 * PI_F;
	ray.dir = cone_random_frame(hit.normal, hit.perturb_matrix, longi, (float)latid);
	ray.pdf = 1.0f / std::max(1.0f, specular_angle * 2.0f);
	ray.attenuation *= m.rg;
}

struct SceneView {
	const orc_object* objs; uint32_t n_objs;
	const orc_material* mats; uint32_t n_mats;
	const orc_light* lights; uint32_t n_lights;
};

// raytracer.cu:6-57
static vec3 shade_object(vec3 wi, SampledRay& ray, uint32_t shadow_count, HitRecord& hit, const SceneView& sc, uint32_t n_steps,
                         float cone, const MarchEnv& env, Xorwow& rng, float& out_nerf_shadow, bool no_shadow, float syn_shadow_factor) {
	if (hit.material_idx < 0) return {0, 0, 0};
	const orc_material& material = sc.mats[hit.material_idx];
	vec3 color = {0, 0, 0};
	for (uint32_t l = 0; l < sc.n_lights; ++l) {
		const orc_light& light = sc.lights[l];
		for (uint32_t s = 0; s < shadow_count; ++s) {
			if (light.type == 0) {
				vec3 lpos = light_sample(light, rng);
				vec3 L = lpos - hit.pos;
				float full_dist = length(L);
				L = normalize(L);
				vec3 invL = 1.0f / L;
				int32_t obj_hit = -1;
				float syn_shadow = no_shadow ? 1.0f : depth_test_world(hit.pos, L, sc.objs, sc.n_objs, obj_hit);
				float nerf_shadow = no_shadow ? 1.0f : depth_test_nerf((float)(syn_shadow + 1.0), n_steps, cone, hit.pos, L, invL, env);
				out_nerf_shadow = std::min(nerf_shadow / full_dist, out_nerf_shadow);
				float shadow = std::min(std::min(nerf_shadow, syn_shadow), full_dist);
				float shadow_mask = smoothstep1(shadow / full_dist);
				shadow_mask = std::pow(shadow_mask, syn_shadow_factor);
				vec3 R = reflect(L, hit.normal);
				vec3 V = normalize(-wi);
				color += local_color(material, L, hit.normal, R, V, light) * shadow_mask;
			} else if (light.type == 1) {
				vec3 lpos = light_sample(light, rng);
				vec3 L = normalize(lpos - hit.pos);
				vec3 R = reflect(L, hit.normal);
				vec3 V = normalize(-wi);
				color += local_color(material, L, hit.normal, R, V, light);
			}
		}
	}
	color /= (float)shadow_count;
	color += vec3{material.ka[0], material.ka[1], material.ka[2]};
	scatter(material, hit, wi, ray, rng);
	return color;
}

// raytracer.cu:304-361 tonemap curves == src/synerfgine/common.cu:186-243 sng_tonemap
static void tonemap_curve(float* x, int curve) {
	if (curve == 0) return;
	for (int k = 0; k < 3; ++k) x[k] = std::max(x[k], 0.0f);
	float k0, k1, k2, k3, k4, k5;
	if (curve == 1) {
		k0 = 0.6f * 0.6f * 2.51f; k1 = 0.6f * 0.03f; k2 = 0.0f; k3 = 0.6f * 0.6f * 2.43f; k4 = 0.6f * 0.59f; k5 = 0.14f;
	} else if (curve == 2) {
		const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
		k0 = A * F - A * E; k1 = C * B * F - B * E; k2 = 0.0f; k3 = A * F; k4 = B * F; k5 = D * F * F;
		const float W = 11.2f;
		const float nom = k0 * (W * W) + k1 * W + k2;
		const float denom = k3 * (W * W) + k4 * W + k5;
		const float white_scale = denom / nom;
		k0 = 4.0f * k0 * white_scale; k1 = 2.0f * k1 * white_scale; k2 = k2 * white_scale; k3 = 4.0f * k3; k4 = 2.0f * k4;
	} else {
		float Y = dot(vec3{0.2126f, 0.7152f, 0.0722f}, vec3{x[0], x[1], x[2]});
		float s = 1.f / (Y + 1.0f);
		for (int k = 0; k < 3; ++k) x[k] = x[k] * s;
		return;
	}
	for (int k = 0; k < 3; ++k) {
		float sq = x[k] * x[k];
		float nom = sq * k0 + k1 * x[k] + k2;
		float denom = k3 * sq + k4 * x[k] + k5;
		x[k] = nom / denom;
	}
}

} // namespace orc

using namespace orc;

extern "C" {

void orc_tonemap_curve(float* rgb, int curve) { tonemap_curve(rgb, curve); }

// sng::init_rays_with_payload_kernel_nerf (raytracer.cu:59-99): uv = pixel / resolution (no half-pixel
// offset), pinhole uv_to_ray, normalised direction; clears the four per-pixel buffers.
void orc_syn_init_rays(const orc_view* view, float* origins, float* dirs, float* frame_rgba, float* frame_depth, float* acc_rgba, float* acc_depth) {
	mat4x3 cam = load_camera(view->camera);
	ivec2 res = {view->width, view->height};
	for (int y = 0; y < res.y; ++y) for (int x = 0; x < res.x; ++x) {
		uint32_t idx = x + res.x * y;
		vec2 uv = {(float)x / (float)res.x, (float)y / (float)res.y};
		Ray ray = uv_to_ray_pinhole(uv, res, {view->focal_length[0], view->focal_length[1]}, cam, {view->screen_center[0], view->screen_center[1]});
		ray.d = normalize(ray.d);
		for (int k = 0; k < 3; ++k) { origins[idx * 3 + k] = ray.o[k]; dirs[idx * 3 + k] = ray.d[k]; }
		float* bufs[2] = {frame_rgba, acc_rgba};
		for (float* b : bufs) if (b) { b[idx * 4] = b[idx * 4 + 1] = b[idx * 4 + 2] = 0.0f; b[idx * 4 + 3] = 1.0f; }
		if (frame_depth) frame_depth[idx] = MAX_DEPTH;
		if (acc_depth) acc_depth[idx] = MAX_DEPTH;
	}
}

// sng::raytrace (raytracer.cu:101-218), one call per frame. rng6: n x 6 XORWOW state words, updated in place.
void orc_syn_raytrace(uint32_t n, const float* origins, const float* dirs, const orc_view* view, const orc_model* model,
                      const orc_object* objs, uint32_t n_objs, const orc_material* mats, uint32_t n_mats, const orc_light* lights, uint32_t n_lights,
                      const orc_rt_params* rt, uint32_t* rng6, float* acc_rgba, float* acc_depth) {
	MarchEnv env = make_env(*model);
	mat4x3 cam = load_camera(view->camera);
	SceneView sc = {objs, n_objs, mats, n_mats, lights, n_lights};
	const vec3 up_vec = cam.c[0];
	const float cone = model->cone_angle_constant;
	for (uint32_t i = 0; i < n; ++i) {
		Xorwow rng; for (int k = 0; k < 5; ++k) rng.v[k] = rng6[i * 6 + k]; rng.d = rng6[i * 6 + 5];
		vec3 src_position = {origins[i * 3], origins[i * 3 + 1], origins[i * 3 + 2]};
		vec3 src_direction = {dirs[i * 3], dirs[i * 3 + 1], dirs[i * 3 + 2]};
		vec3 shade = {0, 0, 0}, normal = {0, 0, 0}, view_pos = {0, 0, 0}, next_pos = {0, 0, 0}, view_dir = {0, 0, 0}, next_dir = {0, 0, 0};
		float nerf_shadow = 1.0f;
		for (uint32_t spp = 0; spp < rt->samples; ++spp) {
			SampledRay ray;
			float longi = xorwow_uniform(rng) * rt->lens_angle_constant;
			double latid = rt->lens_angle_constant ? 0.0 : xorwow_uniform(rng) * 2.0 * PI_F; // inverted ternary kept (raytracer.cu:149)
			ray.pos = src_position;
			ray.dir = cone_random_up(src_direction, up_vec, longi, (float)latid);
			ray.pdf = 1.0f / (float)rt->bounces;
			ray.attenuation = 1.0f;
			vec3 shade_s = {0, 0, 0};
			for (uint32_t bounce = 0; bounce < rt->bounces; ++bounce) {
				vec3 src_pos = ray.pos, src_dir = ray.dir;
				HitRecord hit;
				int32_t hit_obj_id = -1;
				depth_test_world(src_pos, src_dir, objs, n_objs, hit_obj_id, hit);
				if (!bounce) { normal += hit.normal; view_pos += src_pos; view_dir += src_dir; next_pos += hit.pos; }
				if (hit_obj_id < 0) break;
				SampledRay next_ray;
				vec3 color = shade_object(src_dir, next_ray, rt->shadow_iters, hit, sc, rt->n_steps, cone, env, rng, nerf_shadow, !rt->view_nerf_shadow, rt->syn_shadow_factor);
				shade_s += color * ray.pdf * ray.attenuation;
				if (!bounce) next_dir += next_ray.dir;
				ray = next_ray;
			}
			shade += shade_s;
		}
		float weight = (float)rt->samples;
		view_pos /= weight; view_dir /= weight; next_pos /= weight; next_dir /= weight; normal /= weight; shade /= weight;
		float depth = dot(src_direction, next_pos - src_position);
		acc_depth[i] = depth;
		float* a = acc_rgba + (size_t)i * 4;
		vec3 curr = {a[0], a[1], a[2]};
		vec3 outc;
		switch (rt->buffer_type) {
			case 5: outc = normal * 0.5f + V3(0.5f); break;                 // Normal
			case 3: outc = next_dir * 0.5f + V3(0.5f); break;               // NextDirection
			case 4: outc = view_dir * 0.5f + V3(0.5f); break;               // SrcDirection
			case 1: outc = length(next_pos - view_pos) + MIN_DEPTH > MAX_DEPTH ? V3(0.0f) : normalize(next_pos) * 0.5f + V3(0.5f); break; // NextOrigin
			case 2: outc = normalize(view_pos) * 0.5f + V3(0.5f); break;    // SrcOrigin
			case 6: outc = V3(depth); break;                                // Depth
			case 7: outc = V3(nerf_shadow); break;                          // NerfShadow
			default:
				if (dot(curr, curr) > 0.001f) shade = shade * 0.5f + curr * 0.5f; // temporal blend, raytracer.cu:212-215
				outc = shade;
		}
		a[0] = outc.x; a[1] = outc.y; a[2] = outc.z;
		for (int k = 0; k < 5; ++k) rng6[i * 6 + k] = rng.v[k]; rng6[i * 6 + 5] = rng.d;
	}
}

// shadow_for_px (testbed_nerf.cu:1614-1700) per pixel, then shade_with_shadow (:1702-1786).
// The reference evaluates shadow_for_px(q) once per (pixel p, neighbour q) pair and mutates
// rand_state[q] from every thread whose window contains q — a data race, so its output is not
// defined for kernel sizes > 1. This restatement resolves it the only deterministic way that agrees
// with the reference when the window is 1x1: shadow_for_px(q) is evaluated once per pixel q from
// rand_state[q] (advancing it once), and every window reuses that value. Summation order over the
// window is the reference's (x offset outer, y offset inner).
void orc_shade_nerf_shadows(int width, int height, const orc_model* model, const float* positions, const float* normals,
                            const orc_object* objs, uint32_t n_objs, const orc_light* lights, uint32_t n_lights,
                            const orc_rt_params* rt, uint32_t* rng6, int render_mode, float* rgba, float* shadow_out) {
	MarchEnv env = make_env(*model);
	env.min_mip = 0;
	const float cone = model->cone_angle_constant;
	const uint32_t N = (uint32_t)width * height;
	std::vector<float> S(N);
	const float intensity = rt->nerf_shadow_intensity, thr = rt->nerf_on_nerf_shadow_threshold;
	for (uint32_t q = 0; q < N; ++q) {
		Xorwow rng; for (int k = 0; k < 5; ++k) rng.v[k] = rng6[q * 6 + k]; rng.d = rng6[q * 6 + 5];
		vec3 orig_pos = {positions[q * 3], positions[q * 3 + 1], positions[q * 3 + 2]};
		vec3 normal = {normals[q * 3], normals[q * 3 + 1], normals[q * 3 + 2]};
		float overall = 1.0f;
		for (uint32_t i = 0; i < n_lights; ++i) {
			const orc_light& light = lights[i];
			if (light.type == 0) {
				vec3 lpos = light_sample(light, rng);
				vec3 l = normalize(lpos - orig_pos);
				float full_d = length(lpos - orig_pos);
				int32_t hit_obj_id = -1;
				float syn_depth = depth_test_world(orig_pos, l, objs, n_objs, hit_obj_id);
				float syn_shadow_mask = syn_depth / full_d;
				overall = std::min(overall, std::pow(syn_shadow_mask, intensity));
				vec3 fract_offset = full_d * thr * lpos;
				float nerf_depth = std::min(full_d, depth_test_nerf_to(8, cone, orig_pos + fract_offset, lpos, env)); // n_steps = MAX_STEPS_INBETWEEN_COMPACTION (:2660)
				float nerf_shadow_mask = (float)((double)(nerf_depth * (1.0f - std::min(light.intensity, 0.0f))) / ((double)full_d * (1.0 - (double)thr)));
				overall = std::min(overall, nerf_shadow_mask);
			} else if (light.type == 1) {
				vec3 l = normalize(vec3{light.pos[0], light.pos[1], light.pos[2]} - orig_pos);
				overall = (float)std::min(1.0, (double)overall + std::min(0.0, (double)dot(l, normal)) * (double)light.intensity);
			}
		}
		S[q] = overall;
		for (int k = 0; k < 5; ++k) rng6[q * 6 + k] = rng.v[k]; rng6[q * 6 + 5] = rng.d;
	}
	int ks = rt->nerf_shadow_kernel / 2;
	for (int y = 0; y < height; ++y) for (int x = 0; x < width; ++x) {
		float sum = 0.0f; int blend = 0;
		for (int i = -ks; i <= ks; ++i) for (int j = -ks; j <= ks; ++j) {
			int fx = x + i, fy = y + j;
			if (fx < 0 || fy < 0 || fx >= width || fy >= height) continue;
			sum += S[fy * width + fx]; ++blend;
		}
		sum /= (float)blend;
		sum = std::pow(sum, intensity);
		uint32_t idx = x + width * y;
		if (shadow_out) shadow_out[idx] = sum;
		float* c = rgba + (size_t)idx * 4;
		if (render_mode == ORC_MODE_SHADOWDEPTH) { c[0] = c[1] = c[2] = sum; }
		else { c[0] = srgb_to_linear(c[0]) * sum; c[1] = srgb_to_linear(c[1]) * sum; c[2] = srgb_to_linear(c[2]) * sum; } // second linearisation kept (:1783)
	}
}

// sng::overlay_nerf (raytracer.cu:220-258), output colour space sRGB as Engine::frame passes (engine.cu:406).
void orc_overlay(int syn_w, int syn_h, int syn_px_scale, const float* syn_rgba, const float* syn_depth, const float* nerf_rgba, const float* nerf_depth,
                 int show_nerf, int srgb_out, int tonemap, float exposure, float depth_offset, float* final_rgba, float* final_depth) {
	int nerf_w = syn_w / syn_px_scale;
	for (int y = 0; y < syn_h; ++y) for (int x = 0; x < syn_w; ++x) {
		int sid = x + y * syn_w;
		int nid = (x / syn_px_scale) + (y / syn_px_scale) * nerf_w;
		const float* s = syn_rgba + (size_t)sid * 4; const float* nn = nerf_rgba + (size_t)nid * 4;
		float sdepth = syn_depth[sid], ndepth = nerf_depth[nid];
		const float* use = (!show_nerf || sdepth - depth_offset < ndepth) ? s : nn;
		float c[4] = {use[0], use[1], use[2], use[3]};
		float e = std::pow(2.0f, exposure);
		for (int k = 0; k < 3; ++k) c[k] *= e;
		tonemap_curve(c, tonemap);
		if (srgb_out) for (int k = 0; k < 3; ++k) c[k] = linear_to_srgb(c[k]);
		for (int k = 0; k < 4; ++k) final_rgba[(size_t)sid * 4 + k] = c[k];
		final_depth[sid] = sdepth;
	}
}

}
