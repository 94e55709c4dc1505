// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
#pragma once
#include "oracle_common.h"
#include "oracle.h"

namespace orc {

struct Ray { vec3 o, d; vec3 at(float t) const { return o + d * t; } };

struct NerfPayload { // nerf_device.cuh:145-153
	vec3 origin, dir;
	float t, max_weight;
	uint32_t idx;
	uint16_t n_steps;
	bool alive;
};

struct MarchEnv {
	const uint8_t* bitfield;
	uint32_t min_mip, max_mip;
	BoundingBox render_aabb;
	mat3 render_aabb_to_local;
};

uint32_t morton3D(uint32_t x, uint32_t y, uint32_t z);
uint32_t cascaded_grid_idx_at(vec3 pos, uint32_t mip);
bool density_grid_occupied_at(vec3 pos, const uint8_t* bitfield, uint32_t mip);
float to_stepping_space(float t, float cone_angle);
float from_stepping_space(float n, float cone_angle);
float advance_n_steps(float t, float cone_angle, float n);
float calc_dt(float t, float cone_angle);
uint32_t mip_from_pos(vec3 pos, uint32_t max_cascade = 7);
float advance_to_next_voxel(float t, float cone_angle, vec3 pos, vec3 dir, vec3 idir, uint32_t mip);
float if_unoccupied_advance_to_next_occupied_voxel(float t, float cone_angle, const Ray& ray, vec3 idir, const MarchEnv& env);
float ld_random_val(uint32_t index, uint32_t seed, uint32_t dim = 0);
vec2 ld_random_pixel_offset(uint32_t spp);
Ray uv_to_ray_pinhole(vec2 uv, ivec2 resolution, vec2 focal_length, const mat4x3& camera, vec2 screen_center);
void init_ray_payload(uint32_t x, uint32_t y, const orc_view& view, const MarchEnv& env, NerfPayload& payload);
// Stock-path extras (orc_set_camera_ex / orc_set_envmap). init_ray_background: what init_rays_with_payload_kernel_nerf leaves in the
// frame / depth buffers for pixel (x, y) besides the payload; compositor_camera: the matrix Testbed::render_nerf hands to trace().
extern bool g_has_camera_ex;
extern orc_camera_ex g_camera_ex;
void init_ray_background(uint32_t x, uint32_t y, const orc_view& view, const MarchEnv& env, const NerfPayload& payload, float* fb4, float* depth1);
mat4x3 compositor_camera(const orc_view& view);
void advance_pos_nerf(NerfPayload& payload, const MarchEnv& env, uint32_t sample_index, float cone_angle_constant);

MarchEnv make_env(const orc_model& m);
mat4x3 load_camera(const float* c);
BoundingBox load_aabb(const float* mn, const float* mx);
mat3 load_mat3(const float* m);

} // namespace orc
