// ORACLE — TEST INFRASTRUCTURE ONLY.
// Scalar CPU restatement of the SyNeRFgine render hot path. Nothing under oracle/ is ever linked
// into, imported by, or called from the product (synerfgine_b200/). Only tests/, bench.py's
// cpu_baseline / --impl reference leg and __graft_entry__.smoke() may load it, as the checker.
//
// Every function cites the reference file:line it restates (paths relative to /root/reference).
// Arithmetic is IEEE fp32, operation order as written in the reference, no FMA contraction
// (build with -ffp-contract=off); fused multiply-adds appear only where the reference calls
// fmaf()/fma() explicitly.
//
// Parity status: the host-callable subset (march helpers, camera, AABB, triangle, BVH traversal,
// LD sampler) is pinned bit-exactly against the reference's own headers compiled on the host
// (oracle/_ref, see oracle/Makefile + tests/test_oracle_pinned.py). Device-only bodies (kernels,
// curand, tcnn grid/MLP) are pinned against the reference's own device code compiled for sm_100a from the
// sources where they lie (oracle/Makefile.refdev -> oracle/_ref/libref_dev.so, tests/test_gpu_refdev.py).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <algorithm>
#include <limits>

namespace orc {

// ---- transcendental mode -------------------------------------------------------------------
// LIBM: logf/expf from glibc — what the reference's NGP_HOST_DEVICE headers do on the host.
// DET : a deterministic polynomial (oracle/det_math.h) shared formula with the CUDA product so
//       that the exponential-stepping march is bit-identical on CPU and GPU.
enum MathMode { MATH_LIBM = 0, MATH_DET = 1 };
extern int g_math_mode;
float det_logf(float x);
float det_expf(float x);
float det_sinf(float x);
float det_acosf(float x);
inline float m_sinf(float x) { return g_math_mode == MATH_DET ? det_sinf(x) : std::sin(x); }
inline float m_acosf(float x) { return g_math_mode == MATH_DET ? det_acosf(x) : std::acos(x); }
inline float m_logf(float x) { return g_math_mode == MATH_DET ? det_logf(x) : std::log(x); }
inline float m_expf(float x) { return g_math_mode == MATH_DET ? det_expf(x) : std::exp(x); }

// ---- tiny vector algebra with the reference's evaluation order (tcnn/vec.h) --------------------
struct vec2 { float x, y; };
struct vec3 {
	float x, y, z;
	float& operator[](int i) { return (&x)[i]; }
	const float& operator[](int i) const { return (&x)[i]; }
};
struct vec4 { float x, y, z, w; };
struct ivec2 { int x, y; };

inline vec3 V3(float a) { return {a, a, a}; }
inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline vec3 operator*(vec3 a, vec3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline vec3 operator/(vec3 a, vec3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline vec3 operator*(float s, vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline vec3 operator/(vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline vec3 operator/(float s, vec3 a) { return {s / a.x, s / a.y, s / a.z}; }
inline vec3 operator+(vec3 a, float s) { return {a.x + s, a.y + s, a.z + s}; }
inline vec3 operator-(vec3 a, float s) { return {a.x - s, a.y - s, a.z - s}; }
inline vec3 operator-(vec3 a) { return {-a.x, -a.y, -a.z}; }
inline vec3& operator+=(vec3& a, vec3 b) { a = a + b; return a; }
inline vec3& operator-=(vec3& a, vec3 b) { a = a - b; return a; }
inline vec3& operator*=(vec3& a, float s) { a = a * s; return a; }
inline vec3& operator/=(vec3& a, float s) { a = a / s; return a; }

// tcnn/vec.h:426 REDUCTION_OP(dot): result = 0; result += a[i]*b[i] in index order.
inline float dot(vec3 a, vec3 b) { float r = 0.0f; r += a.x * b.x; r += a.y * b.y; r += a.z * b.z; return r; }
inline float length2(vec3 a) { return dot(a, a); }
inline float length(vec3 a) { return std::sqrt(length2(a)); }
// tcnn/vec.h:465-474
inline vec3 normalize(vec3 v) { float len = length(v); if (len <= 0.0f) return {1.0f, 0.0f, 0.0f}; return v / len; }
// tcnn/vec.h:476-483
inline vec3 cross(vec3 a, vec3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline float vmax(vec3 a) { float r = -std::numeric_limits<float>::infinity(); r = std::max(r, a.x); r = std::max(r, a.y); r = std::max(r, a.z); return r; }
inline vec3 vabs(vec3 a) { return {std::fabs(a.x), std::fabs(a.y), std::fabs(a.z)}; }
inline vec3 vmaxv(vec3 a, vec3 b) { return {std::max(a.x, b.x), std::max(a.y, b.y), std::max(a.z, b.z)}; }

// Column-major 3x3 (tcnn tmat<float,3,3>): c[i] is column i.
struct mat3 {
	vec3 c[3];
	static mat3 identity() { return {{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}}; }
};
// tcnn/vec.h:575-585: result[j] = 0; result[j] += m[i][j]*v[i] for i in columns.
inline vec3 operator*(const mat3& m, vec3 v) {
	vec3 r = {0.0f, 0.0f, 0.0f};
	for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) r[j] += m.c[i][j] * v[i];
	return r;
}
inline mat3 operator*(const mat3& a, const mat3& b) { return {{a * b.c[0], a * b.c[1], a * b.c[2]}}; }
inline mat3 operator/(const mat3& a, float s) { return {{a.c[0] / s, a.c[1] / s, a.c[2] / s}}; }
// mat4x3: 4 columns of vec3; column 3 is the translation (camera origin).
struct mat4x3 { vec3 c[4]; };
inline mat3 to_mat3(const mat4x3& m) { return {{m.c[0], m.c[1], m.c[2]}}; }

float determinant(const mat3& m);
mat3 inverse(const mat3& m);

struct BoundingBox {
	vec3 min, max;
	// bounding_box.cuh:163-211
	vec2 ray_intersect(vec3 pos, vec3 dir) const;
	// bounding_box.cuh:217-222
	bool contains(vec3 p) const {
		return p.x >= min.x && p.x <= max.x && p.y >= min.y && p.y <= max.y && p.z >= min.z && p.z <= max.z;
	}
	vec3 diag() const { return max - min; }
	vec3 relative_pos(vec3 p) const { return (p - min) / diag(); }
};

constexpr float MAX_DEPTH = 16384.0f; // common.h MAX_DEPTH()
constexpr float MIN_DEPTH = 1e-5f;   // common.h MIN_DEPTH() (sng)

// ---- half precision -----------------------------------------------------------------------------
typedef _Float16 half_t;
inline half_t f2h(float f) { return (half_t)f; }              // RNE like __float2half_rn
inline float h2f(half_t h) { return (float)h; }
inline half_t hfma(half_t a, half_t b, half_t c) { return (half_t)((double)a * (double)b + (double)c); } // single rounding like HFMA

} // namespace orc
