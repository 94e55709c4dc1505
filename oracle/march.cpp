// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
// Ray generation, start jitter and the occupancy-cascade DDA march.
#include "march.h"
#include <vector>

namespace orc {

// ---- small linear algebra -------------------------------------------------------------------
// tcnn/vec.h determinant/adjoint/inverse for 3x3 (cofactor expansion, inverse = adjoint / det).
float determinant(const mat3& m) {
	return
		 m.c[0][0] * (m.c[1][1] * m.c[2][2] - m.c[1][2] * m.c[2][1])
		-m.c[0][1] * (m.c[1][0] * m.c[2][2] - m.c[1][2] * m.c[2][0])
		+m.c[0][2] * (m.c[1][0] * m.c[2][1] - m.c[1][1] * m.c[2][0]);
}

static inline float det2(float a, float b, float c, float d) { return a * d - b * c; }

mat3 inverse(const mat3& m) {
	// adjoint (tcnn/vec.h adjoint<3>): signed 2x2 minors, transposed.
	const float m00 = det2(m.c[1][1], m.c[1][2], m.c[2][1], m.c[2][2]);
	const float m01 = det2(m.c[0][1], m.c[0][2], m.c[2][1], m.c[2][2]);
	const float m02 = det2(m.c[0][1], m.c[0][2], m.c[1][1], m.c[1][2]);
	const float m10 = det2(m.c[1][0], m.c[1][2], m.c[2][0], m.c[2][2]);
	const float m11 = det2(m.c[0][0], m.c[0][2], m.c[2][0], m.c[2][2]);
	const float m12 = det2(m.c[0][0], m.c[0][2], m.c[1][0], m.c[1][2]);
	const float m20 = det2(m.c[1][0], m.c[1][1], m.c[2][0], m.c[2][1]);
	const float m21 = det2(m.c[0][0], m.c[0][1], m.c[2][0], m.c[2][1]);
	const float m22 = det2(m.c[0][0], m.c[0][1], m.c[1][0], m.c[1][1]);
	mat3 adj = {{{m00, -m01, m02}, {-m10, m11, -m12}, {m20, -m21, m22}}};
	return adj / determinant(m);
}

// bounding_box.cuh:163-211
vec2 BoundingBox::ray_intersect(vec3 pos, vec3 dir) const {
	const float fmaxv = std::numeric_limits<float>::max();
	float tmin = (min.x - pos.x) / dir.x;
	float tmax = (max.x - pos.x) / dir.x;
	if (tmin > tmax) std::swap(tmin, tmax);
	float tymin = (min.y - pos.y) / dir.y;
	float tymax = (max.y - pos.y) / dir.y;
	if (tymin > tymax) std::swap(tymin, tymax);
	if (tmin > tymax || tymin > tmax) return {fmaxv, fmaxv};
	if (tymin > tmin) tmin = tymin;
	if (tymax < tmax) tmax = tymax;
	float tzmin = (min.z - pos.z) / dir.z;
	float tzmax = (max.z - pos.z) / dir.z;
	if (tzmin > tzmax) std::swap(tzmin, tzmax);
	if (tmin > tzmax || tzmin > tmax) return {fmaxv, fmaxv};
	if (tzmin > tmin) tmin = tzmin;
	if (tzmax < tmax) tmax = tzmax;
	return {tmin, tmax};
}

mat4x3 load_camera(const float* c) {
	mat4x3 m;
	for (int i = 0; i < 4; ++i) m.c[i] = {c[i * 3 + 0], c[i * 3 + 1], c[i * 3 + 2]};
	return m;
}
BoundingBox load_aabb(const float* mn, const float* mx) { return {{mn[0], mn[1], mn[2]}, {mx[0], mx[1], mx[2]}}; }
mat3 load_mat3(const float* m) { return {{{m[0], m[1], m[2]}, {m[3], m[4], m[5]}, {m[6], m[7], m[8]}}}; }

MarchEnv make_env(const orc_model& m) {
	MarchEnv e;
	e.bitfield = m.bitfield;
	e.min_mip = 0;
	e.max_mip = m.max_cascade;
	e.render_aabb = load_aabb(m.render_aabb_min, m.render_aabb_max);
	e.render_aabb_to_local = load_mat3(m.render_aabb_to_local);
	return e;
}

// ---- occupancy cascade ---------------------------------------------------------------------------
// tcnn/common_device.h:761-776
static inline uint32_t expand_bits(uint32_t v) {
	v = (v * 0x00010001u) & 0xFF0000FFu;
	v = (v * 0x00000101u) & 0x0F00F00Fu;
	v = (v * 0x00000011u) & 0xC30C30C3u;
	v = (v * 0x00000005u) & 0x49249249u;
	return v;
}
uint32_t morton3D(uint32_t x, uint32_t y, uint32_t z) { return expand_bits(x) | (expand_bits(y) << 1) | (expand_bits(z) << 2); }

constexpr uint32_t GRIDSIZE = 128;                      // nerf_device.cuh:25
constexpr uint32_t GRID_N_CELLS = 128u * 128u * 128u;   // :26
constexpr uint32_t N_CASCADES = 8;                      // :30
constexpr float SQRT3 = 1.73205080757f;                 // :32
constexpr float STEPSIZE = SQRT3 / 1024;                // :33 (NERF_STEPS 1024)
constexpr float MIN_CONE_STEPSIZE = STEPSIZE;           // :34
constexpr float MAX_CONE_STEPSIZE = STEPSIZE * (1 << (N_CASCADES - 1)) * 1024 / GRIDSIZE; // :36

// nerf_device.cuh:317-329
uint32_t cascaded_grid_idx_at(vec3 pos, uint32_t mip) {
	float mip_scale = std::scalbn(1.0f, -(int)mip);
	pos -= V3(0.5f);
	pos *= mip_scale;
	pos += V3(0.5f);
	int ix = (int)(pos.x * (float)GRIDSIZE), iy = (int)(pos.y * (float)GRIDSIZE), iz = (int)(pos.z * (float)GRIDSIZE);
	if (ix < 0 || ix >= (int)GRIDSIZE || iy < 0 || iy >= (int)GRIDSIZE || iz < 0 || iz >= (int)GRIDSIZE) return 0xFFFFFFFFu;
	return morton3D(ix, iy, iz);
}

// nerf_device.cuh:331-341
bool density_grid_occupied_at(vec3 pos, const uint8_t* bitfield, uint32_t mip) {
	uint32_t idx = cascaded_grid_idx_at(pos, mip);
	if (idx == 0xFFFFFFFFu) return false;
	return bitfield[idx / 8 + (GRID_N_CELLS * mip) / 8] & (1 << (idx % 8));
}

static inline float fsign(float a) { return std::copysign(1.0f, a); }

// nerf_device.cuh:360-368
static float distance_to_next_voxel(vec3 pos, vec3 dir, vec3 idir, float res) {
	vec3 p = res * (pos - 0.5f);
	float tx = (std::floor(p.x + 0.5f + 0.5f * fsign(dir.x)) - p.x) * idir.x;
	float ty = (std::floor(p.y + 0.5f + 0.5f * fsign(dir.y)) - p.y) * idir.y;
	float tz = (std::floor(p.z + 0.5f + 0.5f * fsign(dir.z)) - p.z) * idir.z;
	float t = std::min(std::min(tx, ty), tz);
	return std::fmax(t / res, 0.0f);
}

// nerf_device.cuh:379-399
float to_stepping_space(float t, float cone_angle) {
	if (cone_angle <= 1e-5f) return t / MIN_CONE_STEPSIZE;
	float log1p_c = m_logf(1.0f + cone_angle);
	float a = (m_logf(MIN_CONE_STEPSIZE) - m_logf(log1p_c)) / log1p_c;
	float b = (m_logf(MAX_CONE_STEPSIZE) - m_logf(log1p_c)) / log1p_c;
	float at = m_expf(a * log1p_c);
	float bt = m_expf(b * log1p_c);
	if (t <= at) return (t - at) / MIN_CONE_STEPSIZE + a;
	else if (t <= bt) return m_logf(t) / log1p_c;
	else return (t - bt) / MAX_CONE_STEPSIZE + b;
}

// nerf_device.cuh:401-421
float from_stepping_space(float n, float cone_angle) {
	if (cone_angle <= 1e-5f) return n * MIN_CONE_STEPSIZE;
	float log1p_c = m_logf(1.0f + cone_angle);
	float a = (m_logf(MIN_CONE_STEPSIZE) - m_logf(log1p_c)) / log1p_c;
	float b = (m_logf(MAX_CONE_STEPSIZE) - m_logf(log1p_c)) / log1p_c;
	float at = m_expf(a * log1p_c);
	float bt = m_expf(b * log1p_c);
	if (n <= a) return (n - a) * MIN_CONE_STEPSIZE + at;
	else if (n <= b) return m_expf(n * log1p_c);
	else return (n - b) * MAX_CONE_STEPSIZE + bt;
}

// nerf_device.cuh:423-429
float advance_n_steps(float t, float cone_angle, float n) { return from_stepping_space(to_stepping_space(t, cone_angle) + n, cone_angle); }
float calc_dt(float t, float cone_angle) { return advance_n_steps(t, cone_angle, 1.0f) - t; }

// nerf_device.cuh:431-441
float advance_to_next_voxel(float t, float cone_angle, vec3 pos, vec3 dir, vec3 idir, uint32_t mip) {
	float res = std::scalbn((float)GRIDSIZE, -(int)mip);
	float t_target = t + distance_to_next_voxel(pos, dir, idir, res);
	t = to_stepping_space(t, cone_angle);
	t_target = to_stepping_space(t_target, cone_angle);
	return from_stepping_space(t + std::ceil(std::fmax(t_target - t, 0.5f)), cone_angle);
}

// nerf_device.cuh:443-448
uint32_t mip_from_pos(vec3 pos, uint32_t max_cascade) {
	int exponent;
	float maxval = vmax(vabs(pos - 0.5f));
	std::frexp(maxval, &exponent);
	return (uint32_t)std::min(std::max(exponent + 1, 0), (int)max_cascade);
}

// nerf_device.cuh:464-495 (MIP_FROM_DT = false)
float if_unoccupied_advance_to_next_occupied_voxel(float t, float cone_angle, const Ray& ray, vec3 idir, const MarchEnv& env) {
	while (true) {
		vec3 pos = ray.at(t);
		if (t >= MAX_DEPTH || !env.render_aabb.contains(env.render_aabb_to_local * pos)) return MAX_DEPTH;
		uint32_t mip = std::min(std::max(mip_from_pos(pos), env.min_mip), env.max_mip);
		if (!env.bitfield || density_grid_occupied_at(pos, env.bitfield, mip)) return t;
		while (mip < env.max_mip && !density_grid_occupied_at(pos, env.bitfield, mip + 1)) ++mip;
		t = advance_to_next_voxel(t, cone_angle, pos, ray.d, idir, mip);
	}
}

// ---- low-discrepancy sampler (random_val.cuh:161-291) ----------------------------------------
// Sobol direction numbers: dimension 0 is the van der Corput sequence (v_i = 2^(31-i)); dimension 1
// obeys v_i = v_{i-1} ^ (v_{i-1} >> 1) with v_0 = 2^31 (Joe-Kuo / Burley 2020 tables reproduce this).
static uint32_t sobol(uint32_t index, uint32_t dim) {
	uint32_t X = 0;
	uint32_t v = 0x80000000u;
	for (uint32_t bit = 0; bit < 32; ++bit) {
		uint32_t mask = (index >> bit) & 1;
		X ^= mask * v;
		v = dim == 0 ? (v >> 1) : (v ^ (v >> 1));
	}
	return X;
}
static uint32_t hash_combine(uint32_t seed, uint32_t v) { return seed ^ (v + (seed << 6) + (seed >> 2)); }
static uint32_t reverse_bits(uint32_t x) {
	x = (((x & 0xaaaaaaaau) >> 1) | ((x & 0x55555555u) << 1));
	x = (((x & 0xccccccccu) >> 2) | ((x & 0x33333333u) << 2));
	x = (((x & 0xf0f0f0f0u) >> 4) | ((x & 0x0f0f0f0fu) << 4));
	x = (((x & 0xff00ff00u) >> 8) | ((x & 0x00ff00ffu) << 8));
	return ((x >> 16) | (x << 16));
}
static uint32_t laine_karras_permutation(uint32_t x, uint32_t seed) {
	x += seed;
	x ^= x * 0x6c50b47cu;
	x ^= x * 0xb82f1e52u;
	x ^= x * 0xc7afe638u;
	x ^= x * 0x8d22f6e6u;
	return x;
}
static uint32_t nested_uniform_scramble_base2(uint32_t x, uint32_t seed) {
	x = reverse_bits(x);
	x = laine_karras_permutation(x, seed);
	return reverse_bits(x);
}
float ld_random_val(uint32_t index, uint32_t seed, uint32_t dim) {
	const float S = float(1.0 / (1ull << 32));
	index = nested_uniform_scramble_base2(index, seed);
	return (float)nested_uniform_scramble_base2(sobol(index, dim), hash_combine(seed, dim)) * S;
}
static vec2 ld_random_val_2d(uint32_t index, uint32_t seed) {
	const float S = float(1.0 / (1ull << 32));
	index = nested_uniform_scramble_base2(index, seed);
	uint32_t X0 = nested_uniform_scramble_base2(sobol(index, 0), hash_combine(seed, 0));
	uint32_t X1 = nested_uniform_scramble_base2(sobol(index, 1), hash_combine(seed, 1));
	return {(float)X0 * S, (float)X1 * S};
}
static float fractf(float x) { return x - std::floor(x); }
// random_val.cuh:320-325
vec2 ld_random_pixel_offset(uint32_t spp) {
	vec2 a = ld_random_val_2d(0, 0xdeadbeefu), b = ld_random_val_2d(spp, 0xdeadbeefu);
	vec2 offset = {0.5f - a.x + b.x, 0.5f - a.y + b.y};
	return {fractf(offset.x), fractf(offset.y)};
}

// ---- camera ---------------------------------------------------------------------------------------
struct quat { float w, x, y, z; };
// tcnn/vec.h:1056-1086 (matrix -> quaternion)
static quat quat_from_mat3(const mat3& m) {
	quat q;
	float tr = m.c[0][0] + m.c[1][1] + m.c[2][2];
	if (tr > 0.0f) {
		float S = std::sqrt(tr + 1.0f) * 2.0f;
		q.w = 0.25f * S;
		q.x = (m.c[1][2] - m.c[2][1]) / S;
		q.y = (m.c[2][0] - m.c[0][2]) / S;
		q.z = (m.c[0][1] - m.c[1][0]) / S;
	} else if (m.c[0][0] > m.c[1][1] && m.c[0][0] > m.c[2][2]) {
		float S = std::sqrt(1.0f + m.c[0][0] - m.c[1][1] - m.c[2][2]) * 2.0f;
		q.w = (m.c[1][2] - m.c[2][1]) / S;
		q.x = 0.25f * S;
		q.y = (m.c[1][0] + m.c[0][1]) / S;
		q.z = (m.c[2][0] + m.c[0][2]) / S;
	} else if (m.c[1][1] > m.c[2][2]) {
		float S = std::sqrt(1.0f + m.c[1][1] - m.c[0][0] - m.c[2][2]) * 2.0f;
		q.w = (m.c[2][0] - m.c[0][2]) / S;
		q.x = (m.c[1][0] + m.c[0][1]) / S;
		q.y = 0.25f * S;
		q.z = (m.c[2][1] + m.c[1][2]) / S;
	} else {
		float S = std::sqrt(1.0f + m.c[2][2] - m.c[0][0] - m.c[1][1]) * 2.0f;
		q.w = (m.c[0][1] - m.c[1][0]) / S;
		q.x = (m.c[2][0] + m.c[0][2]) / S;
		q.y = (m.c[2][1] + m.c[1][2]) / S;
		q.z = 0.25f * S;
	}
	return q;
}
static float qdot(quat a, quat b) { return (a.w * b.w + a.x * b.x) + (a.y * b.y + a.z * b.z); } // vec.h:1098
static quat qscale(quat a, float s) { return {a.w * s, a.x * s, a.y * s, a.z * s}; }
static quat qscale_l(float s, quat a) { return {s * a.w, s * a.x, s * a.y, s * a.z}; }
static quat qadd(quat a, quat b) { return {a.w + b.w, a.x + b.x, a.y + b.y, a.z + b.z}; }
static quat qdiv(quat a, float s) { return {a.w / s, a.x / s, a.y / s, a.z / s}; }
// vec.h:1121-1142
static quat qslerp(quat x, quat y, float t) {
	quat z = y;
	float cos_theta = qdot(x, y);
	if (cos_theta < 0.0f) { z = {-y.w, -y.x, -y.y, -y.z}; cos_theta = -cos_theta; }
	if (cos_theta > 1.0f - std::numeric_limits<float>::epsilon()) {
		return qadd(qscale(x, 1.0f - t), qscale(z, t));  // mix(), vec.h:1102
	}
	float angle = m_acosf(cos_theta);
	return qdiv(qadd(qscale_l(m_sinf((1.0f - t) * angle), x), qscale_l(m_sinf(t * angle), z)), m_sinf(angle));
}
// vec.h:1099-1110: note length2() already takes a sqrt, so "length" is the 4th root of dot(a,a).
static quat qnormalize(quat a) {
	float len = std::sqrt(std::sqrt(qdot(a, a)));
	if (len <= 0.0f) return {1.0f, 0.0f, 0.0f, 0.0f};
	return qdiv(a, len);
}
// vec.h:1163-1173
static mat3 quat_to_mat3(quat q) {
	float qxx = q.x * q.x, qyy = q.y * q.y, qzz = q.z * q.z;
	float qxz = q.x * q.z, qxy = q.x * q.y, qyz = q.y * q.z;
	float qwx = q.w * q.x, qwy = q.w * q.y, qwz = q.w * q.z;
	return {{
		{1.0f - 2.0f * (qyy + qzz), 2.0f * (qxy + qwz), 2.0f * (qxz - qwy)},
		{2.0f * (qxy - qwz), 1.0f - 2.0f * (qxx + qzz), 2.0f * (qyz + qwx)},
		{2.0f * (qxz + qwy), 2.0f * (qyz - qwx), 1.0f - 2.0f * (qxx + qyy)},
	}};
}

// common_device.cuh:361-368 with start == end and rolling_shutter = (0,0,0,1) (testbed.h:1042).
static mat4x3 xform_given_rolling_shutter(const mat4x3& cam, float motionblur_time, vec2 uv) {
	float pixel_t = 0.0f + 0.0f * uv.x + 0.0f * uv.y + 1.0f * motionblur_time;
	vec3 pos = cam.c[3] + (cam.c[3] - cam.c[3]) * pixel_t;
	quat q = quat_from_mat3(to_mat3(cam));
	mat3 rot = quat_to_mat3(qnormalize(qslerp(q, q, pixel_t)));
	return {{rot.c[0], rot.c[1], rot.c[2], pos}};
}

// common_device.cuh:403-470 restricted to the pinhole lens, no foveation / parallax / DoF / near plane.
Ray uv_to_ray_pinhole(vec2 uv, ivec2 resolution, vec2 focal_length, const mat4x3& camera, vec2 screen_center) {
	vec3 dir = {
		(uv.x - screen_center.x) * (float)resolution.x / focal_length.x,
		(uv.y - screen_center.y) * (float)resolution.y / focal_length.y,
		1.0f,
	};
	vec3 head_pos = {0.0f, 0.0f, 0.0f};
	dir -= head_pos * 0.0f;
	dir = to_mat3(camera) * dir;
	vec3 origin = to_mat3(camera) * head_pos + camera.c[3];
	origin += dir * 0.0f;
	return {origin, dir};
}

// ---- the general camera model (installed by orc_set_camera_ex / orc_set_envmap) -------------------------------------------------------
bool g_has_camera_ex = false;
orc_camera_ex g_camera_ex;
static std::vector<float> g_envmap; static int g_env_w = 0, g_env_h = 0;

// common_device.cuh:361-368
static mat4x3 xform_given_rolling_shutter2(const mat4x3& start, const mat4x3& end, const float* rs, vec2 uv, float motionblur_time) {
	float pixel_t = rs[0] + rs[1] * uv.x + rs[2] * uv.y + rs[3] * motionblur_time;
	vec3 pos = start.c[3] + (end.c[3] - start.c[3]) * pixel_t;
	mat3 rot = quat_to_mat3(qnormalize(qslerp(quat_from_mat3(to_mat3(start)), quat_from_mat3(to_mat3(end)), pixel_t)));
	return {{rot.c[0], rot.c[1], rot.c[2], pos}};
}
// common_device.cuh:249-288
static void opencv_delta(const float* p, float u, float v, float* du, float* dv) {
	const float k1 = p[0], k2 = p[1], p1 = p[2], p2 = p[3];
	const float u2 = u * u, uv = u * v, v2 = v * v, r2 = u2 + v2;
	const float radial = k1 * r2 + k2 * r2 * r2;
	*du = u * radial + 2.0f * p1 * uv + p2 * (r2 + 2.0f * u2);
	*dv = v * radial + 2.0f * p2 * uv + p1 * (r2 + 2.0f * v2);
}
static void fisheye_delta(const float* p, float u, float v, float* du, float* dv) {
	const float r = std::sqrt(u * u + v * v);
	if (r > (float)std::numeric_limits<double>::epsilon()) {
		const float theta = std::atan(r), theta2 = theta * theta, theta4 = theta2 * theta2, theta6 = theta4 * theta2, theta8 = theta4 * theta4;
		const float thetad = theta * (1.0f + p[0] * theta2 + p[1] * theta4 + p[2] * theta6 + p[3] * theta8);
		*du = u * thetad / r - u; *dv = v * thetad / r - v;
	} else { *du = 0.0f; *dv = 0.0f; }
}
// common_device.cuh:290-329
template <typename F> static void iterative_undistortion(const float* params, float* u, float* v, F fun) {
	const float eps = std::numeric_limits<float>::epsilon();
	const float x0[2] = {*u, *v};
	float x[2] = {*u, *v};
	for (uint32_t i = 0; i < 100; ++i) {
		const float step0 = std::max(eps, std::fabs(1e-6f * x[0])), step1 = std::max(eps, std::fabs(1e-6f * x[1]));
		float dx[2], b0[2], f0[2], b1[2], f1[2];
		fun(params, x[0], x[1], &dx[0], &dx[1]);
		fun(params, x[0] - step0, x[1], &b0[0], &b0[1]); fun(params, x[0] + step0, x[1], &f0[0], &f0[1]);
		fun(params, x[0], x[1] - step1, &b1[0], &b1[1]); fun(params, x[0], x[1] + step1, &f1[0], &f1[1]);
		// J[col][row]
		const float J00 = 1 + (f0[0] - b0[0]) / (2 * step0), J10 = (f1[0] - b1[0]) / (2 * step1), J01 = (f0[1] - b0[1]) / (2 * step0), J11 = 1 + (f1[1] - b1[1]) / (2 * step1);
		const float rx = x[0] + dx[0] - x0[0], ry = x[1] + dx[1] - x0[1];
		const float det = J00 * J11 - J10 * J01;
		const float sx = (J11 * rx - J10 * ry) / det, sy = (-J01 * rx + J00 * ry) / det;
		x[0] -= sx; x[1] -= sy;
		if (sx * sx + sy * sy < 1e-10f) break;
	}
	*u = x[0]; *v = x[1];
}
// random_val.cuh:112-128
static vec2 square2disk_shirley(vec2 sq) {
	const float pi = 3.14159265358979323846f;
	float phi, r;
	const float a = sq.x, b = sq.y;
	if (a * a > b * b) { r = a; phi = (pi / 4.0f) * (b / a); } else { r = b; phi = (pi / 2.0f) - (pi / 4.0f) * (a / b); }
	return {r * std::cos(phi), r * std::sin(phi)};
}
// common_device.cuh:403-470 without foveation / hidden-area mask / distortion grid. Returns false for Ray::invalid().
static bool uv_to_ray_general(uint32_t spp, vec2 uv, ivec2 resolution, vec2 focal_length, const mat4x3& camera, vec2 screen_center,
                              const orc_camera_ex& ex, float aperture_size, Ray& out) {
	const float pi = 3.14159265358979323846f;
	vec3 dir;
	if (ex.lens_mode == 2) {          // f_theta_undistortion, :370-383
		const float xpix = (uv.x - screen_center.x) * ex.lens_params[5], ypix = (uv.y - screen_center.y) * ex.lens_params[6];
		const float norm = std::sqrt(xpix * xpix + ypix * ypix);
		const float* q = ex.lens_params;
		const float alpha = q[0] + norm * (q[1] + norm * (q[2] + norm * (q[3] + norm * q[4])));
		float sin_alpha = std::sin(alpha); const float cos_alpha = std::cos(alpha);
		if (cos_alpha <= std::numeric_limits<float>::min() || norm == 0.0f) return false;
		sin_alpha *= 1.0f / norm;
		dir = {sin_alpha * xpix, sin_alpha * ypix, cos_alpha};
		if (dir.x == 0.0f && dir.y == 0.0f && dir.z == 0.0f) return false;
	} else if (ex.lens_mode == 3) {   // latlong_to_dir, :385-392
		const float theta = (uv.y - 0.5f) * pi, phi = (uv.x - 0.5f) * pi * 2.0f;
		dir = {std::sin(phi) * std::cos(theta), std::sin(theta), std::cos(phi) * std::cos(theta)};
	} else if (ex.lens_mode == 5) {   // equirectangular_to_dir, :394-401
		const float ct = (uv.y - 0.5f) * 2.0f, st = std::sqrt(std::max(1.0f - ct * ct, 0.0f)), phi = (uv.x - 0.5f) * pi * 2.0f;
		dir = {std::sin(phi) * st, ct, std::cos(phi) * st};
	} else {
		dir = {(uv.x - screen_center.x) * (float)resolution.x / focal_length.x, (uv.y - screen_center.y) * (float)resolution.y / focal_length.y, 1.0f};
		if (ex.lens_mode == 1) iterative_undistortion(ex.lens_params, &dir.x, &dir.y, opencv_delta);
		else if (ex.lens_mode == 4) iterative_undistortion(ex.lens_params, &dir.x, &dir.y, fisheye_delta);
	}
	vec3 head_pos = {ex.parallax_shift[0], ex.parallax_shift[1], 0.0f};
	dir -= head_pos * ex.parallax_shift[2];
	dir = to_mat3(camera) * dir;
	vec3 origin = to_mat3(camera) * head_pos + camera.c[3];
	if (aperture_size != 0.0f) {
		vec3 lookat = origin + dir * ex.plane_z;
		const int px = (int)(uv.x * (float)resolution.x), py = (int)(uv.y * (float)resolution.y);
		vec2 r2 = ld_random_val_2d(spp, (uint32_t)px * 19349663u + (uint32_t)py * 96925573u);
		vec2 disk = square2disk_shirley({r2.x * 2.0f - 1.0f, r2.y * 2.0f - 1.0f});
		vec2 blur = {aperture_size * disk.x, aperture_size * disk.y};
		origin += camera.c[0] * blur.x + camera.c[1] * blur.y;
		dir = (lookat - origin) / ex.plane_z;
	}
	origin += dir * ex.near_distance;
	out = {origin, dir};
	return true;
}
// envmap.cuh:24-50
static vec4 read_envmap(vec3 dir) {
	const float pi = 3.14159265358979323846f;
	vec3 d = {dir.z, -dir.x, dir.y};
	const float theta = std::acos(std::fmin(std::fmax(d.z, -1.0f), 1.0f)), phi = std::atan2(d.y, d.x);
	const float cx = theta / pi, cy = phi / (2.0f * pi) + 0.5f;
	const float fx = cy * (float)(g_env_w - 1), fy = cx * (float)(g_env_h - 1);
	const int tx = (int)fx, ty = (int)fy;
	const float wx = fx - (float)tx, wy = fy - (float)ty;
	auto rd = [&](int x, int y) {
		if (x < 0) x += g_env_w; else if (x >= g_env_w) x -= g_env_w;
		y = std::max(std::min(y, g_env_h - 1), 0);
		const float* t = g_envmap.data() + ((size_t)y * g_env_w + x) * 4;
		return vec4{t[0], t[1], t[2], t[3]};
	};
	const vec4 a = rd(tx, ty), b = rd(tx + 1, ty), c = rd(tx, ty + 1), e = rd(tx + 1, ty + 1);
	const float w00 = (1 - wx) * (1 - wy), w10 = wx * (1 - wy), w01 = (1 - wx) * wy, w11 = wx * wy;
	return {((w00 * a.x + w10 * b.x) + w01 * c.x) + w11 * e.x, ((w00 * a.y + w10 * b.y) + w01 * c.y) + w11 * e.y,
	        ((w00 * a.z + w10 * b.z) + w01 * c.z) + w11 * e.z, ((w00 * a.w + w10 * b.w) + w01 * c.w) + w11 * e.w};
}

// testbed_nerf.cu:1855-1970 (init_rays_with_payload_kernel_nerf). payload.n_steps carries 0; payload.t = 0 marks an invalid ray.
void init_ray_payload(uint32_t x, uint32_t y, const orc_view& view, const MarchEnv& env, NerfPayload& payload) {
	ivec2 res = {view.width, view.height};
	uint32_t idx = x + res.x * y;
	vec2 off = ld_random_pixel_offset(view.snap_to_pixel_centers ? 0 : view.spp);
	vec2 uv = {((float)x + off.x) / (float)res.x, ((float)y + off.y) / (float)res.y};
	mat4x3 cam0 = load_camera(view.camera);
	Ray ray;
	bool valid = true;
	if (!g_has_camera_ex) {
		mat4x3 camera = xform_given_rolling_shutter(cam0, ld_random_val(view.spp, idx * 72239731u), uv);
		ray = uv_to_ray_pinhole(uv, res, {view.focal_length[0], view.focal_length[1]}, camera, {view.screen_center[0], view.screen_center[1]});
	} else {
		const orc_camera_ex& ex = g_camera_ex;
		float aperture = ex.plane_z < 0.0f ? 0.0f : ex.aperture_size;
		mat4x3 camera = xform_given_rolling_shutter2(cam0, load_camera(ex.camera1), ex.rolling_shutter, uv, ld_random_val(view.spp, idx * 72239731u));
		valid = uv_to_ray_general(view.spp, uv, res, {view.focal_length[0], view.focal_length[1]}, camera, {view.screen_center[0], view.screen_center[1]}, ex, aperture, ray);
	}
	payload.max_weight = 0.0f;
	payload.idx = idx; // (the reference leaves idx stale for dead rays; they are never read)
	payload.n_steps = 0;
	if (!valid) { payload.origin = {0, 0, 0}; payload.dir = {0, 0, 0}; payload.t = 0.0f; payload.alive = false; return; }
	ray.d = normalize(ray.d);
	float t = std::fmax(env.render_aabb.ray_intersect(env.render_aabb_to_local * ray.o, env.render_aabb_to_local * ray.d).x, 0.0f) + 1e-6f;
	payload.origin = ray.o;
	payload.dir = ray.d;
	payload.t = t;
	if (g_has_camera_ex && g_camera_ex.plane_z < 0.0f) { payload.alive = false; return; }   // slice plane, :1918-1928
	if (!env.render_aabb.contains(env.render_aabb_to_local * ray.at(t))) { payload.alive = false; return; }
	payload.alive = true;
}

// What the kernel leaves in frame_buffer[idx] / depth_buffer[idx] (:1912,1926,1946-1950): rgb cleared (alpha kept) or the envmap texel.
void init_ray_background(uint32_t, uint32_t, const orc_view&, const MarchEnv&, const NerfPayload& payload, float* fb4, float* depth1) {
	*depth1 = MAX_DEPTH;
	const bool valid = !(payload.dir.x == 0.0f && payload.dir.y == 0.0f && payload.dir.z == 0.0f);
	if (g_has_camera_ex && !valid) return;
	if (g_has_camera_ex && g_camera_ex.plane_z < 0.0f) { *depth1 = -g_camera_ex.plane_z; return; }
	if (g_env_w > 0) { vec4 e = read_envmap(payload.dir); fb4[0] = e.x; fb4[1] = e.y; fb4[2] = e.z; fb4[3] = e.w; }
	else { fb4[0] = fb4[1] = fb4[2] = 0.0f; }
}

mat4x3 compositor_camera(const orc_view& view) { return g_has_camera_ex ? load_camera(g_camera_ex.camera1) : load_camera(view.camera); }

// testbed_nerf.cu:334-363
void advance_pos_nerf(NerfPayload& payload, const MarchEnv& env, uint32_t sample_index, float cone_angle_constant) {
	if (!payload.alive) return;
	vec3 origin = payload.origin, dir = payload.dir;
	vec3 idir = 1.0f / dir;
	float cone_angle = cone_angle_constant; // calc_cone_angle, nerf_device.cuh:370-377
	float t = advance_n_steps(payload.t, cone_angle, ld_random_val(sample_index, payload.idx * 786433u));
	t = if_unoccupied_advance_to_next_occupied_voxel(t, cone_angle, {origin, dir}, idir, env);
	if (t >= MAX_DEPTH) payload.alive = false; else payload.t = t;
}

} // namespace orc

using namespace orc;

extern "C" {

// Per-ray sample sequence (the "occupancy traversal sample indices" of BASELINE.json): for every pixel
// of the view, init + jitter + up to max_samples march steps exactly as generate_next_nerf_network_inputs
// (testbed_nerf.cu:790-837) would emit them with no early termination. Outputs per sample:
// t, dt, and cell = mip << 24 | morton index (0xFFFFFFFF padding). n_out[ray] = samples produced.
void orc_march_trace(const orc_model* model, const orc_view* view, uint32_t max_samples,
                     float* out_t, float* out_dt, uint32_t* out_cell, uint32_t* n_out, float* out_origin_dir) {
	MarchEnv env = make_env(*model);
	const float cone = model->cone_angle_constant;
	for (int y = 0; y < view->height; ++y) for (int x = 0; x < view->width; ++x) {
		NerfPayload p;
		init_ray_payload(x, y, *view, env, p);
		advance_pos_nerf(p, env, view->spp, cone);
		uint32_t ray = x + view->width * y;
		if (out_origin_dir) { for (int k = 0; k < 3; ++k) { out_origin_dir[ray * 6 + k] = p.origin[k]; out_origin_dir[ray * 6 + 3 + k] = p.dir[k]; } }
		uint32_t n = 0;
		if (p.alive) {
			vec3 idir = 1.0f / p.dir;
			float t = p.t;
			for (; n < max_samples; ++n) {
				t = if_unoccupied_advance_to_next_occupied_voxel(t, cone, {p.origin, p.dir}, idir, env);
				if (t >= MAX_DEPTH) break;
				float dt = calc_dt(t, cone);
				vec3 pos = p.origin + p.dir * t;
				uint32_t mip = std::min(std::max(mip_from_pos(pos), env.min_mip), env.max_mip);
				size_t o = (size_t)ray * max_samples + n;
				out_t[o] = t; out_dt[o] = dt; out_cell[o] = (mip << 24) | (cascaded_grid_idx_at(pos, mip) & 0xFFFFFFu);
				t += dt;
			}
		}
		for (uint32_t k = n; k < max_samples; ++k) { size_t o = (size_t)ray * max_samples + k; out_t[o] = 0; out_dt[o] = 0; out_cell[o] = 0xFFFFFFFFu; }
		n_out[ray] = n;
	}
}

float orc_ld_random_val(uint32_t index, uint32_t seed) { return ld_random_val(index, seed, 0); }
void orc_ld_random_pixel_offset(uint32_t spp, float* out) { vec2 o = ld_random_pixel_offset(spp); out[0] = o.x; out[1] = o.y; }
float orc_calc_dt(float t, float cone) { return calc_dt(t, cone); }
float orc_advance_n_steps(float t, float cone, float n) { return advance_n_steps(t, cone, n); }
uint32_t orc_cascaded_grid_idx_at(const float* pos, uint32_t mip) { return cascaded_grid_idx_at({pos[0], pos[1], pos[2]}, mip); }
uint32_t orc_mip_from_pos(const float* pos) { return mip_from_pos({pos[0], pos[1], pos[2]}); }
float orc_march_to_occupied(const orc_model* model, const float* o, const float* d, float t) {
	MarchEnv env = make_env(*model);
	vec3 dir = {d[0], d[1], d[2]};
	return if_unoccupied_advance_to_next_occupied_voxel(t, model->cone_angle_constant, {{o[0], o[1], o[2]}, dir}, 1.0f / dir, env);
}
void orc_set_camera_ex(const orc_camera_ex* ex) { g_has_camera_ex = ex != nullptr; if (ex) g_camera_ex = *ex; }
void orc_set_envmap(const float* rgba, int32_t w, int32_t h) {
	if (!rgba || w <= 0 || h <= 0) { g_envmap.clear(); g_env_w = g_env_h = 0; return; }
	g_envmap.assign(rgba, rgba + (size_t)w * h * 4); g_env_w = w; g_env_h = h;
}
void orc_init_ray(const orc_view* view, const orc_model* model, uint32_t x, uint32_t y, float* out8) {
	MarchEnv env = make_env(*model);
	NerfPayload p; init_ray_payload(x, y, *view, env, p);
	for (int k = 0; k < 3; ++k) { out8[k] = p.origin[k]; out8[3 + k] = p.dir[k]; }
	out8[6] = p.t; out8[7] = p.alive ? 1.0f : 0.0f;
}

}
