// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
#pragma once
#include "march.h"

namespace orc {

// curand XORWOW state (curand_kernel.h curandStateXORWOW: d, v[5]); Box-Muller fields are unused here.
struct Xorwow { uint32_t v[5]; uint32_t d; };
void xorwow_init_sequence(uint64_t seed, uint32_t n, Xorwow* out); // states of subsequences 0..n-1, offset 0
uint32_t xorwow_next(Xorwow& s);
float xorwow_uniform(Xorwow& s);

struct Triangle { vec3 a, b, c; };
struct BvhNode { vec3 bb_min, bb_max; int left_idx, right_idx; };

struct HitRecord { // include/synerfgine/common.cuh:118-128
	vec3 pos{0, 0, 0}, normal{0, 0, 0};
	mat3 perturb_matrix = mat3::identity();
	float t = MAX_DEPTH;
	int32_t material_idx = -1, object_idx = -1;
	bool front_face = true;
};

float triangle_ray_intersect(const Triangle& tri, vec3 ro, vec3 rd);
vec3 triangle_normal(const Triangle& tri);
void bvh_ray_intersect(vec3 ro, vec3 rd, const BvhNode* nodes, const Triangle* tris, int& out_idx, float& out_t);
void object_ray_intersect(vec3 ro, vec3 rd, const orc_object& obj, int& out_idx, float& out_t);
float depth_test_world(vec3 origin, vec3 dir, const orc_object* objs, uint32_t n_objs, int32_t& out_obj_id);
float depth_test_world(vec3 origin, vec3 dir, const orc_object* objs, uint32_t n_objs, int32_t& out_obj_id, HitRecord& hit);
float depth_test_nerf(float full_d, uint32_t n_steps, float cone, vec3 src, vec3 L, vec3 invL, const MarchEnv& env);
float depth_test_nerf_to(uint32_t n_steps, float cone, vec3 src, vec3 dst, const MarchEnv& env);

float srgb_to_linear(float s);
float linear_to_srgb(float l);

} // namespace orc
