// synerfgine_b200 — the synerfgine mesh layer: BVH closest-hit traversal, Lambertian / glossy distribution-ray
// shading with shadow rays through the meshes and through the NeRF occupancy cascade, mesh shadows cast onto the
// NeRF, and the depth-tested overlay.
//
// Replaces (reference): ngp::ray_intersect_nodes + ray_intersect_nodes_f<2> (src/triangle_bvh.cu:263-319),
// Triangle::ray_intersect / normal / get_perturb_matrix (include/neural-graphics-primitives/triangle.cuh:40-59,164-169),
// sng::depth_test_world x2 and depth_test_nerf x2 (src/synerfgine/common.cu:36-102), sng::init_rand_state (:22-26),
// shade_object / init_rays_with_payload_kernel_nerf / raytrace / overlay_nerf (src/synerfgine/raytracer.cu:6-258),
// Material::local_color / scatter (include/synerfgine/material.cuh:96-123), Light::sample (light.cuh:71-77),
// shadow_for_px / shade_with_shadow (src/testbed_nerf.cu:1614-1786), sng_tonemap (common.cu:186-243),
// accumulate_kernel / tonemap_kernel (src/render_buffer.cu:268-302,569-601).
//
// B200 notes: the world->object matrix is evaluated once per scene update instead of per ray; the RNG state is the
// 24-byte XORWOW core (no Box-Muller cache) seeded by the toolkit's curand_init so sequences equal curand's.
#pragma once
#include <curand_kernel.h>
#include "snb_device.cuh"

namespace snb {

// ---- XORWOW (curand's default generator) ----------------------------------------------------------------------------
struct Rng { uint32_t v0, v1, v2, v3, v4, d; };
SNB_D uint32_t rng_next(Rng& s) {
	const uint32_t t = s.v0 ^ (s.v0 >> 2);
	s.v0 = s.v1; s.v1 = s.v2; s.v2 = s.v3; s.v3 = s.v4;
	s.v4 = (s.v4 ^ (s.v4 << 4)) ^ (t ^ (t << 1));
	s.d += 362437u;
	return s.v4 + s.d;
}
SNB_D float rng_uniform(Rng& s) { return rng_next(s) * 2.3283064e-10f + (2.3283064e-10f / 2.0f); }

__global__ void __launch_bounds__(128) k_init_rng(uint32_t n, unsigned long long seed, int width, int tx0, int ty0, int tw, Rng* __restrict__ states) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	// subsequence = global pixel index so a tile draws the same numbers as the full frame
	const uint32_t gx = tx0 + (int)(i % (uint32_t)tw), gy = ty0 + (int)(i / (uint32_t)tw);
	curandStateXORWOW_t st;
	curand_init(seed, (unsigned long long)gx + (unsigned long long)width * gy, 0, &st);
	Rng r; r.v0 = st.v[0]; r.v1 = st.v[1]; r.v2 = st.v[2]; r.v3 = st.v[3]; r.v4 = st.v[4]; r.d = st.d;
	states[i] = r;
}

// ---- triangle + BVH ---------------------------------------------------------------------------------------------------
struct Tri { v3 a, b, c; };
SNB_D Tri load_tri(const float* __restrict__ tris, int i) {
	const float* p = tris + (size_t)i * 9;
	Tri t;
	t.a = mk3(__ldg(p + 0), __ldg(p + 1), __ldg(p + 2));
	t.b = mk3(__ldg(p + 3), __ldg(p + 4), __ldg(p + 5));
	t.c = mk3(__ldg(p + 6), __ldg(p + 7), __ldg(p + 8));
	return t;
}
SNB_D float tri_intersect(const Tri& tri, v3 ro, v3 rd) {
	const v3 v1v0 = tri.b - tri.a, v2v0 = tri.c - tri.a, rov0 = ro - tri.a;
	const v3 n = cross(v1v0, v2v0);
	const v3 q = cross(rov0, rd);
	const float d = 1.0f / dot(rd, n);
	const float u = d * -dot(q, v2v0);
	const float v = d * dot(q, v1v0);
	float t = d * -dot(n, rov0);
	if (u < 0.0f || u > 1.0f || v < 0.0f || (u + v) > 1.0f || t < 0.0f) t = 3.402823466e+38f;
	return t;
}
SNB_D v3 tri_normal(const Tri& t) { return normalize(cross(t.b - t.a, t.c - t.a)); }

// Closest hit in one object's BVH (object space). Depth-first, near child first, 32-entry stack.
SNB_D void bvh_closest(v3 ro, v3 rd, const float* __restrict__ nodes, const float* __restrict__ tris, int& out_idx, float& out_t) {
	int stack[32]; int sp = 0;
	stack[sp++] = 0;
	float mint = kMaxDepth;
	int best = -1;
	while (sp > 0) {
		const int ni = stack[--sp];
		const float4 n1 = __ldg(reinterpret_cast<const float4*>(nodes + (size_t)ni * 8) + 1);
		const int left = __float_as_int(n1.z), right = __float_as_int(n1.w);
		if (left < 0) {
			const int end = -right - 1;
			for (int i = -left - 1; i < end; ++i) {
				const float t = tri_intersect(load_tri(tris, i), ro, rd);
				if (t < mint) { mint = t; best = i; }
			}
		} else {
			float dist[2];
#pragma unroll
			for (int c = 0; c < 2; ++c) {
				const float4 a = __ldg(reinterpret_cast<const float4*>(nodes + (size_t)(left + c) * 8));
				const float4 b = __ldg(reinterpret_cast<const float4*>(nodes + (size_t)(left + c) * 8) + 1);
				Aabb bb; bb.lo = mk3(a.x, a.y, a.z); bb.hi = mk3(a.w, b.x, b.y);
				dist[c] = bb.ray_intersect(ro, rd).x;
			}
			int far_i = left, near_i = left + 1; float far_d = dist[0], near_d = dist[1];
			if (dist[0] < dist[1]) { far_i = left + 1; near_i = left; far_d = dist[1]; near_d = dist[0]; }
			if (far_d < mint && sp < 32) stack[sp++] = far_i;
			if (near_d < mint && sp < 32) stack[sp++] = near_i;
		}
	}
	out_idx = best; out_t = mint;
}

SNB_D void object_closest(v3 ro, v3 rd, const DevObject& obj, int& out_idx, float& out_t) {
	const v3 oro = mul(obj.inv_sr, ro - obj.pos);
	const v3 ord = mul(obj.inv_sr, rd);
	bvh_closest(oro, ord, obj.nodes, obj.tris, out_idx, out_t);
}

// Closest hit over all objects: distance, object id, triangle id.
SNB_D float world_closest(v3 origin, v3 dir, const DevScene& sc, int& obj_id, int& tri_id) {
	float depth = kMaxDepth;
	const v3 off_origin = origin + dir * kMinDepth;
	for (uint32_t c = 0; c < sc.n_objects; ++c) {
		int ti; float hd;
		object_closest(off_origin, dir, sc.objects[c], ti, hd);
		if (hd < depth && hd > kMinDepth) { obj_id = (int)c; tri_id = ti; depth = hd; }
	}
	return depth;
}

struct Hit {
	v3 pos, normal;
	m3 frame;            // perturb_matrix {T, B, N}
	float t;
	int material_idx, object_idx;
};

SNB_D float world_closest_hit(v3 origin, v3 dir, const DevScene& sc, int& obj_id, Hit& hit) {
	int tri_id = -1;
	hit.t = world_closest(origin, dir, sc, obj_id, tri_id);
	hit.material_idx = -1; hit.object_idx = -1;
	hit.normal = mk3(0, 0, 0); hit.frame = identity3();
	if (obj_id >= 0) {
		const DevObject& ob = sc.objects[obj_id];
		const Tri tri = load_tri(ob.tris, tri_id);
		const v3 N = tri_normal(tri);
		hit.material_idx = ob.mat_id;
		hit.object_idx = obj_id;
		hit.normal = mul(ob.rot, N);
		const v3 T = normalize((tri.a + tri.b + tri.c) / 3.0f - tri.a);
		hit.frame.c0 = T; hit.frame.c1 = cross(T, N); hit.frame.c2 = N;
	}
	hit.pos = origin + hit.t * dir;
	return hit.t;
}

// Shadow ray through the occupancy cascade only: distance reached within n_steps DDA steps, capped at full_d.
SNB_D float nerf_shadow_distance(float full_d, uint32_t n_steps, v3 src, v3 L, v3 invL, const MarchEnv& env) {
	float s = 0.0f;
	for (uint32_t j = 0; j < n_steps; ++j) {
		s = skip_to_occupied(s, src, L, invL, env);
		if (s >= full_d) { s = full_d; break; }
		s += step_size(s, env.st);
	}
	return s;
}
SNB_D float nerf_shadow_distance_to(uint32_t n_steps, v3 src, v3 dst, const MarchEnv& env) {
	const float full_d = length(dst - src);
	const v3 L = normalize(dst - src);
	return nerf_shadow_distance(full_d, n_steps, src, L, 1.0f / L, env);
}

// ---- shading ------------------------------------------------------------------------------------------------------------
SNB_D v3 reflect_about(v3 incident, v3 normal) { return 2.0f * dot(incident, normal) * normal - incident; }
SNB_D float fract1(float x) { return x - floorf(x); }
SNB_D v3 light_sample(const DevLight& l, Rng& rng) {
	const float a = fract1(rng_uniform(rng)), b = fract1(rng_uniform(rng)), c = fract1(rng_uniform(rng));
	return l.pos + mk3(a, b, c) * l.size * 1.0f;
}
SNB_D v3 phong(const DevMaterial& m, v3 L, v3 N, v3 R, v3 V, const DevLight& light) {
	return fmaxf(0.0f, dot(L, N)) * m.kd * light.intensity + powf(fmaxf(0.0f, dot(R, V)), m.n) * m.ks;
}
struct NextRay { v3 pos, dir; float pdf, attenuation; };

SNB_D void scatter(const DevMaterial& m, const Hit& hit, v3 src_dir, NextRay& ray, Rng& rng) {
	ray.pos = hit.pos;
	if (m.type != 0 && m.type != 1) return;
	const float spec = m.type == 0 ? kPi / 2 : m.spec_angle;
	const float longi = rng_uniform(rng) * spec;
	const double latid_d = rng_uniform(rng) * 2.0 * kPi;
	const float latid = (float)latid_d;
	const v3 offset = mk3(cosf(longi) * sinf(latid), sinf(longi) * sinf(latid), cosf(longi));
	ray.dir = hit.normal + mul(hit.frame, offset);          // cone_random(normal, perturb_matrix, ...) overwrites reflect()
	ray.pdf = 1.0f / fmaxf(1.0f, spec * 2.0f);
	ray.attenuation *= m.rg;
}

struct RtParams {
	uint32_t n_steps, samples, bounces, shadow_iters;
	float lens, depth_offset, syn_shadow_factor;
	int view_nerf_shadow, buffer_type;
};

SNB_D v3 shade_hit(v3 wi, NextRay& ray, const Hit& hit, const DevScene& sc, const RtParams& rt, const MarchEnv& env, Rng& rng, float& out_nerf_shadow) {
	if (hit.material_idx < 0) return mk3(0, 0, 0);
	const DevMaterial& mat = sc.materials[hit.material_idx];
	const bool no_shadow = !rt.view_nerf_shadow;
	v3 color = mk3(0, 0, 0);
	for (uint32_t l = 0; l < sc.n_lights; ++l) {
		const DevLight& light = sc.lights[l];
		for (uint32_t s = 0; s < rt.shadow_iters; ++s) {
			if (light.type == 0) {
				const v3 lpos = light_sample(light, rng);
				v3 L = lpos - hit.pos;
				const float full_dist = length(L);
				L = normalize(L);
				const v3 invL = 1.0f / L;
				int oid = -1, tid = -1;
				const float syn_shadow = no_shadow ? 1.0f : world_closest(hit.pos, L, sc, oid, tid);
				const float nerf_shadow = no_shadow ? 1.0f : nerf_shadow_distance((float)(syn_shadow + 1.0), rt.n_steps, hit.pos, L, invL, env);
				out_nerf_shadow = fminf(nerf_shadow / full_dist, out_nerf_shadow);
				const float shadow = fminf(fminf(nerf_shadow, syn_shadow), full_dist);
				const float v = shadow / full_dist;
				float mask = v * v * (3.0f - 2.0f * v);         // smoothstep
				mask = powf(mask, rt.syn_shadow_factor);
				const v3 R = reflect_about(L, hit.normal);
				const v3 V = normalize(-wi);
				color = color + phong(mat, L, hit.normal, R, V, light) * mask;
			} else if (light.type == 1) {
				const v3 lpos = light_sample(light, rng);
				const v3 L = normalize(lpos - hit.pos);
				const v3 R = reflect_about(L, hit.normal);
				const v3 V = normalize(-wi);
				color = color + phong(mat, L, hit.normal, R, V, light);
			}
		}
	}
	color = color / (float)rt.shadow_iters;
	color = color + mat.ka;
	scatter(mat, hit, wi, ray, rng);
	return color;
}

// ---- kernels --------------------------------------------------------------------------------------------------------------
// Primary rays of the mesh layer: uv = pixel / resolution (no half-pixel offset), pinhole, normalised; clears the
// per-pixel buffers. Runs only when accumulation is reset.
__global__ void __launch_bounds__(128) k_syn_raygen(DevView view, float* __restrict__ origins, float* __restrict__ dirs, float4* __restrict__ frame_rgba,
                                                    float* __restrict__ frame_depth, float4* __restrict__ acc_rgba, float* __restrict__ acc_depth) {
	const int px = blockIdx.x * 16 + (threadIdx.x & 15), py = blockIdx.y * 8 + (threadIdx.x >> 4);
	if (px >= view.tw || py >= view.th) return;
	const int x = view.tx0 + px, y = view.ty0 + py, l = px + view.tw * py;
	v2 uv; uv.x = (float)x / (float)view.width; uv.y = (float)y / (float)view.height;
	v3 o, d;
	{
		v3 dir = mk3((uv.x - view.center.x) * (float)view.width / view.focal.x, (uv.y - view.center.y) * (float)view.height / view.focal.y, 1.0f);
		const v3 head = mk3(0, 0, 0);
		dir = dir - head * 0.0f;
		dir = mul(rot_of(view.cam), dir);
		o = mul(rot_of(view.cam), head) + view.cam.c3;
		o = o + dir * 0.0f;
		d = normalize(dir);
	}
	origins[l * 3] = o.x; origins[l * 3 + 1] = o.y; origins[l * 3 + 2] = o.z;
	dirs[l * 3] = d.x; dirs[l * 3 + 1] = d.y; dirs[l * 3 + 2] = d.z;
	if (frame_rgba) frame_rgba[l] = make_float4(0, 0, 0, 1);
	if (acc_rgba) acc_rgba[l] = make_float4(0, 0, 0, 1);
	if (frame_depth) frame_depth[l] = kMaxDepth;
	if (acc_depth) acc_depth[l] = kMaxDepth;
}

// Pixel of this thread for the two divergent per-pixel kernels (k_raytrace, k_nerf_shadow_terms). tiled: a warp covers an
// 8 x 4 pixel patch (grid.x * warps-per-block patches across, grid.y patches down) instead of 32 pixels of one row, so the
// lanes that hit the mesh / sit on written NeRF pixels cluster into fewer warps; results are per pixel and do not depend
// on the mapping. Not tiled: the row-linear index (option "sng_block" = "0").
SNB_D bool tile_pixel(int tiled, uint32_t n, int width, uint32_t& idx) {
	if (!tiled) { idx = blockIdx.x * blockDim.x + threadIdx.x; return idx < n; }
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	const uint32_t x = (blockIdx.x * (blockDim.x >> 5) + warp) * 8u + (lane & 7u), y = blockIdx.y * 4u + (lane >> 3);
	idx = y * (uint32_t)width + x;
	return x < (uint32_t)width && idx < n;
}

#ifndef SNB_RT_THREADS
#define SNB_RT_THREADS 128      // largest k_raytrace block (tuning hook; option "sng_block" picks 32..SNB_RT_THREADS at run time)
#endif
#ifdef SNB_RT_MIN_BLOCKS         // resident blocks per SM the register budget of k_raytrace is set for (tuning hook; unset = ptxas' own choice, 96 registers)
#define SNB_RT_BOUNDS __launch_bounds__(SNB_RT_THREADS, SNB_RT_MIN_BLOCKS)
#else
#define SNB_RT_BOUNDS __launch_bounds__(SNB_RT_THREADS)
#endif
// One thread per mesh-layer pixel: samples x bounces closest-hit + shading; 50/50 temporal blend into acc_rgba.
__global__ void SNB_RT_BOUNDS k_raytrace(uint32_t n, const float* __restrict__ origins, const float* __restrict__ dirs, v3 up_vec, MarchEnv env,
                                                  DevScene sc, RtParams rt, Rng* __restrict__ rng_states, float4* __restrict__ acc_rgba, float* __restrict__ acc_depth,
                                                  int full_w, int tx0, int ty0, int tw, int tiled) {
	uint32_t i;
	if (!tile_pixel(tiled, n, tw, i)) return;
	const size_t gi = (size_t)(ty0 + (int)(i / (uint32_t)tw)) * full_w + tx0 + (int)(i % (uint32_t)tw);   // generator states are per GLOBAL pixel
	Rng rng = rng_states[gi];
	const v3 src_position = mk3(origins[i * 3], origins[i * 3 + 1], origins[i * 3 + 2]);
	const v3 src_direction = mk3(dirs[i * 3], dirs[i * 3 + 1], dirs[i * 3 + 2]);
	v3 shade = mk3(0, 0, 0), normal = mk3(0, 0, 0), view_pos = mk3(0, 0, 0), next_pos = mk3(0, 0, 0), view_dir = mk3(0, 0, 0), next_dir = mk3(0, 0, 0);
	float nerf_shadow = 1.0f;
	for (uint32_t spp = 0; spp < rt.samples; ++spp) {
		NextRay ray;
		const float longi = rng_uniform(rng) * rt.lens;
		const double latid_d = rt.lens ? 0.0 : rng_uniform(rng) * 2.0 * kPi;   // the reference's inverted ternary (raytracer.cu:149)
		const float latid = (float)latid_d;
		{ // cone_random(src_direction, up_vec, longi, latid)
			const v3 N = normalize(src_direction);
			const v3 B = normalize(cross(N, up_vec));
			const v3 T = cross(B, N);
			m3 frame; frame.c0 = T; frame.c1 = B; frame.c2 = N;
			const v3 offset = mk3(sinf(longi) * cosf(latid), sinf(longi) * sinf(latid), cosf(longi));
			ray.dir = src_direction + mul(frame, offset);
		}
		ray.pos = src_position;
		ray.pdf = 1.0f / (float)rt.bounces;
		ray.attenuation = 1.0f;
		v3 shade_s = mk3(0, 0, 0);
		for (uint32_t bounce = 0; bounce < rt.bounces; ++bounce) {
			const v3 src_pos = ray.pos, src_dir = ray.dir;
			Hit hit;
			int hit_obj = -1;
			world_closest_hit(src_pos, src_dir, sc, hit_obj, hit);
			if (!bounce) { normal = normal + hit.normal; view_pos = view_pos + src_pos; view_dir = view_dir + src_dir; next_pos = next_pos + hit.pos; }
			if (hit_obj < 0) break;
			NextRay next; next.pos = mk3(0, 0, 0); next.dir = mk3(0, 0, 0); next.pdf = 0.0f; next.attenuation = 1.0f;
			const v3 color = shade_hit(src_dir, next, hit, sc, rt, env, rng, nerf_shadow);
			shade_s = shade_s + color * ray.pdf * ray.attenuation;
			if (!bounce) next_dir = next_dir + next.dir;
			ray = next;
		}
		shade = shade + shade_s;
	}
	const float weight = (float)rt.samples;
	view_pos = view_pos / weight; view_dir = view_dir / weight; next_pos = next_pos / weight; next_dir = next_dir / weight; normal = normal / weight; shade = shade / weight;
	const float depth = dot(src_direction, next_pos - src_position);
	acc_depth[i] = depth;
	const float4 prev = acc_rgba[i];
	const v3 curr = mk3(prev.x, prev.y, prev.z);
	v3 outc;
	switch (rt.buffer_type) {
		case 5: outc = normal * 0.5f + splat3(0.5f); break;
		case 3: outc = next_dir * 0.5f + splat3(0.5f); break;
		case 4: outc = view_dir * 0.5f + splat3(0.5f); break;
		case 1: outc = length(next_pos - view_pos) + kMinDepth > kMaxDepth ? splat3(0.0f) : normalize(next_pos) * 0.5f + splat3(0.5f); break;
		case 2: outc = normalize(view_pos) * 0.5f + splat3(0.5f); break;
		case 6: outc = splat3(depth); break;
		case 7: outc = splat3(nerf_shadow); break;
		default:
			if (dot(curr, curr) > 0.001f) shade = shade * 0.5f + curr * 0.5f;
			outc = shade;
	}
	acc_rgba[i] = make_float4(outc.x, outc.y, outc.z, prev.w);
	rng_states[gi] = rng;
}

// Mesh shadows on the NeRF, pass 1: one shadow term per NeRF pixel from its world position / normal.
//
// Dead-term elimination (`layer` != nullptr): pass 2 multiplies srgb_to_linear(rgb(p)) by the window mean around p, and a pixel
// the march never wrote has rgb == 0 (frame buffer cleared by ray generation), so its output is 0 whatever the mean is. term(q)
// is therefore only ever observable through a pixel p of q's window that the march wrote (alpha(p) > 0). When no such pixel
// exists the BVH traversal and the shadow DDA are skipped; the RNG still draws what light_sample would have drawn, so the
// generator states — and with them every later frame — stay identical. Not used in ShadowDepth mode (which shows the mean itself).
__global__ void __launch_bounds__(128) k_nerf_shadow_terms(uint32_t n, const float* __restrict__ positions, const float* __restrict__ normals, MarchEnv env, DevScene sc,
                                                          float intensity, float thr, Rng* __restrict__ rng_states, float* __restrict__ terms,
                                                          const float4* __restrict__ layer, int width, int height, int half_k, int full_w, int tx0, int ty0, int tiled) {
	uint32_t q;
	if (!tile_pixel(tiled, n, width, q)) return;
	const size_t gq = (size_t)(ty0 + (int)(q / (uint32_t)width)) * full_w + tx0 + (int)(q % (uint32_t)width);   // generator states are per GLOBAL pixel
	Rng rng = rng_states[gq];
	bool observable = true;
	if (layer) {
		observable = layer[q].w > 0.0f;
		if (!observable) {
			const int x = (int)(q % (uint32_t)width), y = (int)(q / (uint32_t)width);
			for (int j = max(0, y - half_k); j <= min(height - 1, y + half_k) && !observable; ++j)
				for (int i = max(0, x - half_k); i <= min(width - 1, x + half_k); ++i)
					if (layer[j * width + i].w > 0.0f) { observable = true; break; }
		}
	}
	if (!observable) {
		for (uint32_t i = 0; i < sc.n_lights; ++i)
			if (sc.lights[i].type == 0) (void)light_sample(sc.lights[i], rng);
		terms[q] = 1.0f;
		rng_states[gq] = rng;
		return;
	}
	const v3 orig_pos = mk3(positions[q * 3], positions[q * 3 + 1], positions[q * 3 + 2]);
	const v3 normal = mk3(normals[q * 3], normals[q * 3 + 1], normals[q * 3 + 2]);
	float overall = 1.0f;
	for (uint32_t i = 0; i < sc.n_lights; ++i) {
		const DevLight& light = sc.lights[i];
		if (light.type == 0) {
			const v3 lpos = light_sample(light, rng);
			const v3 l = normalize(lpos - orig_pos);
			const float full_d = length(lpos - orig_pos);
			int oid = -1, tid = -1;
			const float syn_depth = world_closest(orig_pos, l, sc, oid, tid);
			overall = fminf(overall, powf(syn_depth / full_d, intensity));
			const v3 fract_offset = full_d * thr * lpos;       // along the light position vector, as the reference does (:1662)
			const float nerf_depth = fminf(full_d, nerf_shadow_distance_to(8, orig_pos + fract_offset, lpos, env));
			const float mask = (float)((double)(nerf_depth * (1.0f - fminf(light.intensity, 0.0f))) / ((double)full_d * (1.0 - (double)thr)));
			overall = fminf(overall, mask);
		} else if (light.type == 1) {
			const v3 l = normalize(light.pos - orig_pos);
			overall = (float)fmin(1.0, (double)overall + fmin(0.0, (double)dot(l, normal)) * (double)light.intensity);
		}
	}
	terms[q] = overall;
	rng_states[gq] = rng;
}

// Pass 2: (2*(K/2)+1)^2 window mean of the terms, pow, darken (with the reference's second srgb_to_linear).
__global__ void __launch_bounds__(128) k_nerf_shadow_apply(int width, int height, int half_k, float intensity, int render_mode, const float* __restrict__ terms,
                                                          float4* __restrict__ rgba, float* __restrict__ shadow_out) {
	const int x = blockIdx.x * 16 + (threadIdx.x & 15), y = blockIdx.y * 8 + (threadIdx.x >> 4);
	if (x >= width || y >= height) return;
	float sum = 0.0f; int blend = 0;
	for (int i = -half_k; i <= half_k; ++i) for (int j = -half_k; j <= half_k; ++j) {
		const int fx = x + i, fy = y + j;
		if (fx < 0 || fy < 0 || fx >= width || fy >= height) continue;
		sum += terms[fy * width + fx]; ++blend;
	}
	sum /= (float)blend;
	sum = powf(sum, intensity);
	const int idx = x + width * y;
	if (shadow_out) shadow_out[idx] = sum;
	float4 c = rgba[idx];
	if (render_mode == 8) { c.x = c.y = c.z = sum; }
	else { c.x = srgb_to_linear(c.x) * sum; c.y = srgb_to_linear(c.y) * sum; c.z = srgb_to_linear(c.z) * sum; }
	rgba[idx] = c;
}

SNB_D void tonemap_rgb(float (&x)[3], int curve) {
	if (curve == 0) return;
	for (int k = 0; k < 3; ++k) x[k] = fmaxf(x[k], 0.0f);
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
		const float Y = dot(mk3(0.2126f, 0.7152f, 0.0722f), mk3(x[0], x[1], x[2]));
		const float s = 1.f / (Y + 1.0f);
		for (int k = 0; k < 3; ++k) x[k] = x[k] * s;
		return;
	}
	for (int k = 0; k < 3; ++k) {
		const float sq = x[k] * x[k];
		const float nom = sq * k0 + k1 * x[k] + k2;
		const float denom = k3 * sq + k4 * x[k] + k5;
		x[k] = nom / denom;
	}
}

// Depth-tested composite of the mesh layer over the NeRF layer + exposure, tonemap, sRGB.
__global__ void __launch_bounds__(128) k_overlay(int syn_w, int syn_h, int scale, const float4* __restrict__ syn_rgba, const float* __restrict__ syn_depth,
                                                const float4* __restrict__ nerf_rgba, const float* __restrict__ nerf_depth, int show_nerf, int srgb_out, int curve,
                                                float exposure, float depth_offset, float4* __restrict__ out_rgba, float* __restrict__ out_depth) {
	const int x = blockIdx.x * 16 + (threadIdx.x & 15), y = blockIdx.y * 8 + (threadIdx.x >> 4);
	if (x >= syn_w || y >= syn_h) return;
	const int nerf_w = syn_w / scale;
	const int sid = x + y * syn_w, nid = (x / scale) + (y / scale) * nerf_w;
	const float sdepth = syn_depth[sid];
	const float4 use = (!show_nerf || sdepth - depth_offset < nerf_depth[nid]) ? syn_rgba[sid] : nerf_rgba[nid];
	float c[3] = {use.x, use.y, use.z};
	const float e = powf(2.0f, exposure);
	for (int k = 0; k < 3; ++k) c[k] *= e;
	tonemap_rgb(c, curve);
	if (srgb_out) for (int k = 0; k < 3; ++k) c[k] = linear_to_srgb(c[k]);
	out_rgba[sid] = make_float4(c[0], c[1], c[2], use.w);
	if (out_depth) out_depth[sid] = sdepth;
}

// Running mean over spp + background blend + exposure + tonemap (render_frame_epilogue).
__global__ void __launch_bounds__(128) k_accumulate_tonemap(int n, const float4* __restrict__ frame, float4* __restrict__ accum, uint32_t spp, int cs_srgb, int out_srgb,
                                                           float exposure, float4 background, int curve, float4* __restrict__ out) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const float sample_count = (float)spp;
	const float4 c = frame[i];
	float4 a = spp == 0 ? make_float4(0, 0, 0, 0) : accum[i];
	float col[3] = {c.x, c.y, c.z};
	if (cs_srgb) for (int k = 0; k < 3; ++k) col[k] = linear_to_srgb(col[k]);
	a.x = (a.x * sample_count + col[0]) / (sample_count + 1);
	a.y = (a.y * sample_count + col[1]) / (sample_count + 1);
	a.z = (a.z * sample_count + col[2]) / (sample_count + 1);
	a.w = (a.w * sample_count + c.w) / (sample_count + 1);
	accum[i] = a;
	float bg[3] = {background.x, background.y, background.z};
	if (!cs_srgb) for (int k = 0; k < 3; ++k) bg[k] = srgb_to_linear(bg[k]);
	float color[3] = {a.x, a.y, a.z};
	float alpha = a.w;
	const float weight = (1 - alpha) * background.w;
	for (int k = 0; k < 3; ++k) color[k] += bg[k] * weight;
	alpha += weight;
	if (cs_srgb) for (int k = 0; k < 3; ++k) color[k] = srgb_to_linear(color[k]);
	const float e = powf(2.0f, exposure);
	for (int k = 0; k < 3; ++k) color[k] *= e;
	tonemap_rgb(color, curve);
	if (out_srgb) for (int k = 0; k < 3; ++k) color[k] = linear_to_srgb(color[k]);
	out[i] = make_float4(color[0], color[1], color[2], alpha);
}

// Parity tap: closest hit of arbitrary world-space rays.
__global__ void __launch_bounds__(128) k_trace_rays(uint32_t n, const float* __restrict__ rays, DevScene sc, int* __restrict__ obj_id, int* __restrict__ tri_id, float* __restrict__ t_out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const v3 o = mk3(rays[i * 6], rays[i * 6 + 1], rays[i * 6 + 2]), d = mk3(rays[i * 6 + 3], rays[i * 6 + 4], rays[i * 6 + 5]);
	int oid = -1, tid = -1;
	const float t = world_closest(o, d, sc, oid, tid);
	obj_id[i] = oid; tri_id[i] = oid >= 0 ? tid : -1; t_out[i] = t;
}

} // namespace snb
