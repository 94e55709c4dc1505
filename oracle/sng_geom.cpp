// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
// Geometry + RNG substrate of the synerfgine layer: TriangleBvh traversal,
// depth_test_world / depth_test_nerf, and the curand XORWOW generator the reference seeds per pixel.
#include "sng_common.h"
#include <vector>

namespace orc {

// ---- XORWOW (Marsaglia 2003, "Xorshift RNGs"; seeding + sequence spacing as curand_kernel.h:770-800) ----
// curand_init(seed, subsequence, 0): scramble the seed into v[0..4], d; then jump ahead by
// subsequence * 2^67 steps of the xorshift part (the Weyl counter d is not advanced by sequence jumps).
// The jump is a GF(2) linear map on the 160 state bits; we build it by squaring the one-step matrix 67
// times instead of reading the toolkit's precalculated table (tests compare both).
static inline void xorshift_step(uint32_t v[5]) { // curand_kernel.h curand(curandStateXORWOW_t*)
	uint32_t t = v[0] ^ (v[0] >> 2);
	v[0] = v[1]; v[1] = v[2]; v[2] = v[3]; v[3] = v[4];
	v[4] = (v[4] ^ (v[4] << 4)) ^ (t ^ (t << 1));
}

struct BitMatrix { uint32_t row[160][5]; };

static void mat_apply(const BitMatrix& M, const uint32_t in[5], uint32_t out[5]) {
	uint32_t r[5] = {0, 0, 0, 0, 0};
	for (int i = 0; i < 5; ++i) for (int j = 0; j < 32; ++j) if (in[i] & (1u << j)) {
		const uint32_t* row = M.row[i * 32 + j];
		for (int k = 0; k < 5; ++k) r[k] ^= row[k];
	}
	for (int k = 0; k < 5; ++k) out[k] = r[k];
}

static const BitMatrix& sequence_jump_matrix() {
	static BitMatrix S;
	static bool built = false;
	if (built) return S;
	BitMatrix M;
	for (int b = 0; b < 160; ++b) {
		uint32_t v[5] = {0, 0, 0, 0, 0};
		v[b / 32] = 1u << (b & 31);
		xorshift_step(v);
		for (int k = 0; k < 5; ++k) M.row[b][k] = v[k];
	}
	for (int s = 0; s < 67; ++s) { // XORWOW_SEQUENCE_SPACING = 67 (curand_precalc.h:54)
		BitMatrix Q;
		for (int b = 0; b < 160; ++b) mat_apply(M, M.row[b], Q.row[b]);
		M = Q;
	}
	S = M; built = true;
	return S;
}

void xorwow_init_sequence(uint64_t seed, uint32_t n, Xorwow* out) {
	uint32_t s0 = ((uint32_t)seed) ^ 0xaad26b49u;
	uint32_t s1 = (uint32_t)(seed >> 32) ^ 0xf7dcefddu;
	uint32_t t0 = 1099087573u * s0;
	uint32_t t1 = 2591861531u * s1;
	Xorwow st;
	st.d = 6615241u + t1 + t0;
	st.v[0] = 123456789u + t0;
	st.v[1] = 362436069u ^ t0;
	st.v[2] = 521288629u + t1;
	st.v[3] = 88675123u ^ t1;
	st.v[4] = 5783321u + t0;
	const BitMatrix& S = sequence_jump_matrix();
	for (uint32_t i = 0; i < n; ++i) {
		out[i] = st;
		mat_apply(S, st.v, st.v);
	}
}

uint32_t xorwow_next(Xorwow& s) { xorshift_step(s.v); s.d += 362437u; return s.v[4] + s.d; }
// curand_uniform: x * 2^-32 + 2^-33 in fp32 (curand_uniform.h _curand_uniform)
float xorwow_uniform(Xorwow& s) { return xorwow_next(s) * 2.3283064e-10f + (2.3283064e-10f / 2.0f); }

// ---- triangle / BVH -------------------------------------------------------------------------------
// triangle.cuh:45-59
float triangle_ray_intersect(const Triangle& tri, vec3 ro, vec3 rd) {
	vec3 v1v0 = tri.b - tri.a;
	vec3 v2v0 = tri.c - tri.a;
	vec3 rov0 = ro - tri.a;
	vec3 n = cross(v1v0, v2v0);
	vec3 q = cross(rov0, rd);
	float d = 1.0f / dot(rd, n);
	float u = d * -dot(q, v2v0);
	float v = d * dot(q, v1v0);
	float t = d * -dot(n, rov0);
	if (u < 0.0f || u > 1.0f || v < 0.0f || (u + v) > 1.0f || t < 0.0f) t = std::numeric_limits<float>::max();
	return t;
}
vec3 triangle_normal(const Triangle& tri) { return normalize(cross(tri.b - tri.a, tri.c - tri.a)); } // triangle.cuh:40-42

// triangle_bvh.cu:263-307 ray_intersect_nodes_f<2>: DFS, 32-entry stack, far child pushed first.
void bvh_ray_intersect(vec3 ro, vec3 rd, const BvhNode* nodes, const Triangle* tris, int& out_idx, float& out_t) {
	int stack[64]; int sp = 0;
	stack[sp++] = 0;
	float mint = MAX_DEPTH;
	int shortest_idx = -1;
	while (sp > 0) {
		int idx = stack[--sp];
		const BvhNode& node = nodes[idx];
		if (node.left_idx < 0) {
			int end = -node.right_idx - 1;
			for (int i = -node.left_idx - 1; i < end; ++i) {
				float t = triangle_ray_intersect(tris[i], ro, rd);
				if (t < mint) { mint = t; shortest_idx = i; }
			}
		} else {
			float dist[2]; uint32_t cidx[2];
			for (uint32_t i = 0; i < 2; ++i) {
				const BvhNode& c = nodes[i + node.left_idx];
				BoundingBox bb = {c.bb_min, c.bb_max};
				dist[i] = bb.ray_intersect(ro, rd).x;
				cidx[i] = i + node.left_idx;
			}
			if (dist[0] < dist[1]) { std::swap(dist[0], dist[1]); std::swap(cidx[0], cidx[1]); } // descending sort (triangle_bvh.cu:146-161)
			for (uint32_t i = 0; i < 2; ++i) if (dist[i] < mint) stack[sp++] = (int)cidx[i];
		}
	}
	out_idx = shortest_idx; out_t = mint;
}

// triangle_bvh.cu:313-319
void object_ray_intersect(vec3 ro, vec3 rd, const orc_object& obj, int& out_idx, float& out_t) {
	mat3 m_scale = mat3::identity() / obj.scale;
	mat3 m_rotate = inverse(load_mat3(obj.rot));
	vec3 p = {obj.pos[0], obj.pos[1], obj.pos[2]};
	mat3 sr = m_scale * m_rotate;
	vec3 oro = sr * (ro - p);
	vec3 ord = sr * rd;
	bvh_ray_intersect(oro, ord, (const BvhNode*)obj.nodes, (const Triangle*)obj.tris, out_idx, out_t);
}

// src/synerfgine/common.cu:36-49
float depth_test_world(vec3 origin, vec3 dir, const orc_object* objs, uint32_t n_objs, int32_t& out_obj_id) {
	float depth = MAX_DEPTH;
	vec3 offset_origin = origin + dir * MIN_DEPTH;
	for (uint32_t c = 0; c < n_objs; ++c) {
		int tri_id; float hit_d;
		object_ray_intersect(offset_origin, dir, objs[c], tri_id, hit_d);
		if (hit_d < depth && hit_d > MIN_DEPTH) { out_obj_id = (int32_t)c; depth = hit_d; }
	}
	return depth;
}

// src/synerfgine/common.cu:51-67
float depth_test_world(vec3 origin, vec3 dir, const orc_object* objs, uint32_t n_objs, int32_t& out_obj_id, HitRecord& hit) {
	vec3 offset_origin = origin + dir * MIN_DEPTH;
	for (uint32_t c = 0; c < n_objs; ++c) {
		int tri_id; float hit_d;
		object_ray_intersect(offset_origin, dir, objs[c], tri_id, hit_d);
		if (hit_d < hit.t && hit_d > MIN_DEPTH) {
			const Triangle& tri = ((const Triangle*)objs[c].tris)[tri_id];
			out_obj_id = (int32_t)c;
			hit.t = hit_d;
			hit.material_idx = objs[c].mat_id;
			hit.normal = load_mat3(objs[c].rot) * triangle_normal(tri);
			// Triangle::get_perturb_matrix, triangle.cuh:164-169
			vec3 N = triangle_normal(tri);
			vec3 T = normalize((tri.a + tri.b + tri.c) / 3.0f - tri.a);
			vec3 B = cross(T, N);
			hit.perturb_matrix = {{T, B, N}};
			hit.object_idx = (int32_t)c;
		}
	}
	hit.pos = origin + hit.t * dir;
	hit.front_face = dot(dir, hit.normal) < 0.0f;
	return hit.t;
}

// src/synerfgine/common.cu:69-83
float depth_test_nerf(float full_d, uint32_t n_steps, float cone, vec3 src, vec3 L, vec3 invL, const MarchEnv& env) {
	float nerf_shadow = 0.0f;
	for (uint32_t j = 0; j < n_steps; ++j) {
		nerf_shadow = if_unoccupied_advance_to_next_occupied_voxel(nerf_shadow, cone, {src, L}, invL, env);
		if (nerf_shadow >= full_d) { nerf_shadow = full_d; break; }
		float dt = calc_dt(nerf_shadow, cone);
		nerf_shadow += dt;
	}
	return nerf_shadow;
}

// src/synerfgine/common.cu:85-102
float depth_test_nerf_to(uint32_t n_steps, float cone, vec3 src, vec3 dst, const MarchEnv& env) {
	float full_d = length(dst - src);
	vec3 L = normalize(dst - src);
	vec3 invL = 1.0f / L;
	return depth_test_nerf(full_d, n_steps, cone, src, L, invL, env);
}

} // namespace orc

using namespace orc;

extern "C" {

void orc_xorwow_states(uint64_t seed, uint32_t n, uint32_t* out6 /* n x 6: v0..v4, d */) {
	std::vector<Xorwow> st(n);
	xorwow_init_sequence(seed, n, st.data());
	for (uint32_t i = 0; i < n; ++i) { for (int k = 0; k < 5; ++k) out6[i * 6 + k] = st[i].v[k]; out6[i * 6 + 5] = st[i].d; }
}
void orc_xorwow_jump_matrix(uint32_t* out800) { // same layout as curand_precalc.h precalc_xorwow_matrix[0]
	const BitMatrix& S = sequence_jump_matrix();
	for (int b = 0; b < 160; ++b) for (int k = 0; k < 5; ++k) out800[b * 5 + k] = S.row[b][k];
}
void orc_xorwow_uniforms(uint64_t seed, uint32_t subsequence, uint32_t n, float* out) {
	std::vector<Xorwow> st(subsequence + 1);
	xorwow_init_sequence(seed, subsequence + 1, st.data());
	Xorwow s = st[subsequence];
	for (uint32_t i = 0; i < n; ++i) out[i] = xorwow_uniform(s);
}

// rays: n x 6 (origin, dir). Outputs: closest object id, triangle id (in that object's triangle array) and t.
void orc_trace_world(const orc_object* objs, uint32_t n_objs, const float* rays, uint32_t n, int32_t* obj_id, int32_t* tri_id, float* t_out) {
	for (uint32_t i = 0; i < n; ++i) {
		vec3 o = {rays[i * 6], rays[i * 6 + 1], rays[i * 6 + 2]}, d = {rays[i * 6 + 3], rays[i * 6 + 4], rays[i * 6 + 5]};
		vec3 oo = o + d * MIN_DEPTH;
		float best = MAX_DEPTH; int32_t bo = -1, bt = -1;
		for (uint32_t c = 0; c < n_objs; ++c) {
			int ti; float hd;
			object_ray_intersect(oo, d, objs[c], ti, hd);
			if (hd < best && hd > MIN_DEPTH) { best = hd; bo = (int32_t)c; bt = ti; }
		}
		obj_id[i] = bo; tri_id[i] = bt; t_out[i] = best;
	}
}

}
