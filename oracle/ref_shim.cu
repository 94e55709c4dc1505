// ORACLE — TEST INFRASTRUCTURE ONLY. Thin extern "C" shim over the REFERENCE's own NGP_HOST_DEVICE headers, compiled where
// they lie under /root/reference (never copied) and executed on the host. It pins the oracle's restatement of the
// host-callable subset: march helpers, LD sampler, camera/ray generation, AABB and triangle intersection, sRGB.
// Build: oracle/Makefile target `ref` -> oracle/_ref/libref_shim.so (git-ignored). Goldens made from it:
// tests/golden/make_ref_golden.py -> tests/golden/ref_host.npz.
#include <cstdio>
#include <neural-graphics-primitives/common.h>
#include <neural-graphics-primitives/common_device.cuh>
#include <neural-graphics-primitives/random_val.cuh>
#include <neural-graphics-primitives/bounding_box.cuh>
#include <neural-graphics-primitives/triangle.cuh>
#include <neural-graphics-primitives/nerf_device.cuh>

using namespace ngp;

static vec3 V(const float* p) { return vec3(p[0], p[1], p[2]); }

extern "C" {

float ref_calc_dt(float t, float cone) { return calc_dt(t, cone); }
float ref_advance_n_steps(float t, float cone, float n) { return advance_n_steps(t, cone, n); }
float ref_to_stepping_space(float t, float cone) { return to_stepping_space(t, cone); }
float ref_from_stepping_space(float n, float cone) { return from_stepping_space(n, cone); }
uint32_t ref_cascaded_grid_idx_at(const float* pos, uint32_t mip) { return cascaded_grid_idx_at(V(pos), mip); }
uint32_t ref_mip_from_pos(const float* pos) { return mip_from_pos(V(pos)); }
uint32_t ref_morton3D(uint32_t x, uint32_t y, uint32_t z) { return tcnn::morton3D(x, y, z); }
float ref_ld_random_val(uint32_t index, uint32_t seed) { return ld_random_val(index, seed); }
void ref_ld_random_pixel_offset(uint32_t spp, float* out) { vec2 o = ld_random_pixel_offset(spp); out[0] = o.x; out[1] = o.y; }
float ref_srgb_to_linear(float x) { return srgb_to_linear(x); }
float ref_linear_to_srgb(float x) { return linear_to_srgb(x); }

float ref_march_to_occupied(const uint8_t* bitfield, uint32_t max_mip, const float* aabb_min, const float* aabb_max, float cone, const float* o, const float* d, float t) {
	Ray ray{V(o), V(d)};
	return if_unoccupied_advance_to_next_occupied_voxel(t, cone, ray, vec3(1.0f) / ray.d, bitfield, 0, max_mip, BoundingBox(V(aabb_min), V(aabb_max)), mat3::identity());
}

// init_rays_with_payload_kernel_nerf's ray set-up (src/testbed_nerf.cu:1884-1949) composed from the reference's own
// host-callable pieces: pixel offset, rolling-shutter camera, uv_to_ray, normalisation, AABB entry.
void ref_init_ray(const float* cam12, const float* focal2, const float* center2, int w, int h, uint32_t spp, int snap, const float* aabb_min, const float* aabb_max,
                  uint32_t x, uint32_t y, float* out8) {
	mat4x3 cam(V(cam12), V(cam12 + 3), V(cam12 + 6), V(cam12 + 9));
	ivec2 res(w, h);
	uint32_t idx = x + res.x * y;
	vec2 pixel_offset = ld_random_pixel_offset(snap ? 0 : spp);
	vec2 uv = vec2{(float)x + pixel_offset.x, (float)y + pixel_offset.y} / vec2(res);
	mat4x3 camera = get_xform_given_rolling_shutter({cam, cam}, vec4(0.0f, 0.0f, 0.0f, 1.0f), uv, ld_random_val(spp, idx * 72239731));
	Ray ray = uv_to_ray(spp, uv, res, vec2(focal2[0], focal2[1]), camera, vec2(center2[0], center2[1]));
	ray.d = normalize(ray.d);
	BoundingBox bb(V(aabb_min), V(aabb_max));
	float t = fmaxf(bb.ray_intersect(ray.o, ray.d).x, 0.0f) + 1e-6f;
	out8[0] = ray.o.x; out8[1] = ray.o.y; out8[2] = ray.o.z; out8[3] = ray.d.x; out8[4] = ray.d.y; out8[5] = ray.d.z;
	out8[6] = t; out8[7] = bb.contains(ray(t)) ? 1.0f : 0.0f;
}

float ref_triangle_ray_intersect(const float* tri9, const float* ro, const float* rd) {
	Triangle tri{V(tri9), V(tri9 + 3), V(tri9 + 6)};
	return tri.ray_intersect(V(ro), V(rd));
}
void ref_triangle_normal(const float* tri9, float* out3) {
	Triangle tri{V(tri9), V(tri9 + 3), V(tri9 + 6)};
	vec3 n = tri.normal(); out3[0] = n.x; out3[1] = n.y; out3[2] = n.z;
}
void ref_aabb_ray_intersect(const float* mn, const float* mx, const float* ro, const float* rd, float* out2) {
	vec2 r = BoundingBox(V(mn), V(mx)).ray_intersect(V(ro), V(rd)); out2[0] = r.x; out2[1] = r.y;
}

}
