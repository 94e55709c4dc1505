// synerfgine_b200 — small fp32 vector algebra + deterministic transcendentals shared by every kernel.
//
// Arithmetic contract (DESIGN.md "Numerics"): this translation unit is compiled with
//   -fmad=false -prec-div=true -prec-sqrt=true -ftz=false
// so every +,-,*,/ and sqrt below is a single IEEE-754 fp32 operation evaluated in the order
// written; fused multiply-adds appear only where fmaf() is spelled out. That is what makes the
// occupancy-cascade march reproducible bit for bit on any device and against the CPU oracle.
//
// Evaluation orders follow tcnn's vec.h (reference: dependencies/tiny-cuda-nn/include/tiny-cuda-nn/vec.h):
// dot() accumulates x,y,z in index order starting from 0 (vec.h:426), mat*vec accumulates column by
// column (vec.h:575-585), normalize() returns (1,0,0) for zero length (vec.h:465-474).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <math.h>

#define SNB_HD __host__ __device__ __forceinline__
#define SNB_D __device__ __forceinline__

namespace snb {

struct v2 { float x, y; };
struct v3 { float x, y, z; };
struct v4 { float x, y, z, w; };
struct m3 { v3 c0, c1, c2; };      // column-major 3x3
struct m4x3 { v3 c0, c1, c2, c3; }; // camera: right, up(down), forward, origin

SNB_HD v3 mk3(float a, float b, float c) { v3 r; r.x = a; r.y = b; r.z = c; return r; }
SNB_HD v3 splat3(float a) { return mk3(a, a, a); }
SNB_HD v3 operator+(v3 a, v3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
SNB_HD v3 operator-(v3 a, v3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
SNB_HD v3 operator*(v3 a, v3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
SNB_HD v3 operator/(v3 a, v3 b) { return mk3(a.x / b.x, a.y / b.y, a.z / b.z); }
SNB_HD v3 operator*(v3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
SNB_HD v3 operator*(float s, v3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
SNB_HD v3 operator/(v3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
SNB_HD v3 operator/(float s, v3 a) { return mk3(s / a.x, s / a.y, s / a.z); }
SNB_HD v3 operator+(v3 a, float s) { return mk3(a.x + s, a.y + s, a.z + s); }
SNB_HD v3 operator-(v3 a, float s) { return mk3(a.x - s, a.y - s, a.z - s); }
SNB_HD v3 operator-(v3 a) { return mk3(-a.x, -a.y, -a.z); }
SNB_HD float comp(v3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

SNB_HD float dot(v3 a, v3 b) { float r = 0.0f; r += a.x * b.x; r += a.y * b.y; r += a.z * b.z; return r; }
SNB_HD float length(v3 a) { return sqrtf(dot(a, a)); }
SNB_HD v3 normalize(v3 v) { float len = length(v); if (len <= 0.0f) return mk3(1.0f, 0.0f, 0.0f); return v / len; }
SNB_HD v3 cross(v3 a, v3 b) { return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
SNB_HD v3 vabs(v3 a) { return mk3(fabsf(a.x), fabsf(a.y), fabsf(a.z)); }
SNB_HD float vmax(v3 a) { return fmaxf(fmaxf(a.x, a.y), a.z); }

SNB_HD v3 mul(const m3& m, v3 v) {
	v3 r = mk3(0.0f, 0.0f, 0.0f);
	r.x += m.c0.x * v.x; r.y += m.c0.y * v.x; r.z += m.c0.z * v.x;
	r.x += m.c1.x * v.y; r.y += m.c1.y * v.y; r.z += m.c1.z * v.y;
	r.x += m.c2.x * v.z; r.y += m.c2.y * v.z; r.z += m.c2.z * v.z;
	return r;
}
SNB_HD m3 mul(const m3& a, const m3& b) { m3 r; r.c0 = mul(a, b.c0); r.c1 = mul(a, b.c1); r.c2 = mul(a, b.c2); return r; }
SNB_HD m3 operator/(const m3& a, float s) { m3 r; r.c0 = a.c0 / s; r.c1 = a.c1 / s; r.c2 = a.c2 / s; return r; }
SNB_HD m3 identity3() { m3 r; r.c0 = mk3(1, 0, 0); r.c1 = mk3(0, 1, 0); r.c2 = mk3(0, 0, 1); return r; }
SNB_HD m3 rot_of(const m4x3& m) { m3 r; r.c0 = m.c0; r.c1 = m.c1; r.c2 = m.c2; return r; }

// 3x3 inverse as adjugate / determinant, cofactor order of tcnn vec.h (adjoint<3>, determinant<3>).
SNB_HD float det2(float a, float b, float c, float d) { return a * d - b * c; }
SNB_HD float determinant(const m3& m) {
	return m.c0.x * (m.c1.y * m.c2.z - m.c1.z * m.c2.y)
	     - m.c0.y * (m.c1.x * m.c2.z - m.c1.z * m.c2.x)
	     + m.c0.z * (m.c1.x * m.c2.y - m.c1.y * m.c2.x);
}
SNB_HD m3 inverse(const m3& m) {
	const float m00 = det2(m.c1.y, m.c1.z, m.c2.y, m.c2.z);
	const float m01 = det2(m.c0.y, m.c0.z, m.c2.y, m.c2.z);
	const float m02 = det2(m.c0.y, m.c0.z, m.c1.y, m.c1.z);
	const float m10 = det2(m.c1.x, m.c1.z, m.c2.x, m.c2.z);
	const float m11 = det2(m.c0.x, m.c0.z, m.c2.x, m.c2.z);
	const float m12 = det2(m.c0.x, m.c0.z, m.c1.x, m.c1.z);
	const float m20 = det2(m.c1.x, m.c1.y, m.c2.x, m.c2.y);
	const float m21 = det2(m.c0.x, m.c0.y, m.c2.x, m.c2.y);
	const float m22 = det2(m.c0.x, m.c0.y, m.c1.x, m.c1.y);
	m3 adj; adj.c0 = mk3(m00, -m01, m02); adj.c1 = mk3(-m10, m11, -m12); adj.c2 = mk3(m20, -m21, m22);
	return adj / determinant(m);
}

struct Aabb {
	v3 lo, hi;
	SNB_HD bool contains(v3 p) const { return p.x >= lo.x && p.x <= hi.x && p.y >= lo.y && p.y <= hi.y && p.z >= lo.z && p.z <= hi.z; }
	SNB_HD v3 diag() const { return hi - lo; }
	// Slab test with the reference's early-outs (include/neural-graphics-primitives/bounding_box.cuh:163-211).
	SNB_HD v2 ray_intersect(v3 pos, v3 dir) const {
		const float big = 3.402823466e+38f;
		v2 miss; miss.x = big; miss.y = big;
		float tmin = (lo.x - pos.x) / dir.x, tmax = (hi.x - pos.x) / dir.x;
		if (tmin > tmax) { float s = tmin; tmin = tmax; tmax = s; }
		float tymin = (lo.y - pos.y) / dir.y, tymax = (hi.y - pos.y) / dir.y;
		if (tymin > tymax) { float s = tymin; tymin = tymax; tymax = s; }
		if (tmin > tymax || tymin > tmax) return miss;
		if (tymin > tmin) tmin = tymin;
		if (tymax < tmax) tmax = tymax;
		float tzmin = (lo.z - pos.z) / dir.z, tzmax = (hi.z - pos.z) / dir.z;
		if (tzmin > tzmax) { float s = tzmin; tzmin = tzmax; tzmax = s; }
		if (tmin > tzmax || tzmin > tmax) return miss;
		if (tzmin > tmin) tmin = tzmin;
		if (tzmax < tmax) tmax = tzmax;
		v2 r; r.x = tmin; r.y = tmax; return r;
	}
};

constexpr float kMaxDepth = 16384.0f; // MAX_DEPTH(), include/neural-graphics-primitives/common.h
constexpr float kMinDepth = 1e-5f;    // MIN_DEPTH()
constexpr float kPi = 3.14159265358979323846f;

SNB_HD uint32_t f2u(float f) {
#ifdef __CUDA_ARCH__
	return __float_as_uint(f);
#else
	union { float f; uint32_t u; } c; c.f = f; return c.u;
#endif
}
SNB_HD float u2f(uint32_t u) {
#ifdef __CUDA_ARCH__
	return __uint_as_float(u);
#else
	union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}

// ---- deterministic transcendentals (spec: oracle/det_math.cpp header; written independently here) ----
// log: split x = m * 2^e with m in [2/3, 4/3), s = (m-1)/(m+1), odd atanh series to s^11.
SNB_HD float det_logf(float x) {
	const uint32_t ux = f2u(x);
	const int32_t e = (int32_t)(ux - 0x3f2aaaabu) >> 23;
	const float m = u2f(ux - ((uint32_t)e << 23));
	const float s = (m - 1.0f) / (m + 1.0f);
	const float z = s * s;
	float p = 0.181818181818181818f;
	p = fmaf(p, z, 0.222222222222222222f);
	p = fmaf(p, z, 0.285714285714285714f);
	p = fmaf(p, z, 0.4f);
	p = fmaf(p, z, 0.666666666666666667f);
	p = p * z;
	const float lm = fmaf(s, p, 2.0f * s);
	const float fe = (float)e;
	return fmaf(fe, 0.693145751953125f, fmaf(fe, 1.42860682030941723212e-6f, lm));
}
// exp: n = floor(x*log2(e) + 1/2) clamped, Cody-Waite reduction, Taylor to r^8, exponent-field add.
SNB_HD float det_expf(float x) {
	float fn = floorf(fmaf(x, 1.44269504088896341f, 0.5f));
	if (fn < -125.0f) fn = -125.0f;
	if (fn > 126.0f) fn = 126.0f;
	float r = fmaf(-fn, 0.693145751953125f, x);
	r = fmaf(-fn, 1.42860682030941723212e-6f, r);
	float q = 2.48015873015873016e-5f;
	q = fmaf(q, r, 1.98412698412698413e-4f);
	q = fmaf(q, r, 1.38888888888888889e-3f);
	q = fmaf(q, r, 8.33333333333333333e-3f);
	q = fmaf(q, r, 4.16666666666666667e-2f);
	q = fmaf(q, r, 1.66666666666666667e-1f);
	q = fmaf(q, r, 0.5f);
	q = fmaf(q, r, 1.0f);
	q = fmaf(q, r, 1.0f);
	const int32_t n = (int32_t)fn;
	return u2f(f2u(q) + ((uint32_t)n << 23));
}
SNB_HD float det_sinf(float x) { // x in [0, pi/2]
	const float z = x * x;
	float p = 1.0f / 6227020800.0f;
	p = fmaf(p, z, -1.0f / 39916800.0f);
	p = fmaf(p, z, 1.0f / 362880.0f);
	p = fmaf(p, z, -1.0f / 5040.0f);
	p = fmaf(p, z, 1.0f / 120.0f);
	p = fmaf(p, z, -1.0f / 6.0f);
	p = fmaf(p, z, 1.0f);
	return x * p;
}
SNB_HD float det_asin_small(float s) { // s in [0, 0.5]
	const float z = s * s;
	float p = 2027025.0f / 175472640.0f;
	p = fmaf(p, z, 135135.0f / 9676800.0f);
	p = fmaf(p, z, 10395.0f / 599040.0f);
	p = fmaf(p, z, 945.0f / 42240.0f);
	p = fmaf(p, z, 105.0f / 3456.0f);
	p = fmaf(p, z, 15.0f / 336.0f);
	p = fmaf(p, z, 3.0f / 40.0f);
	p = fmaf(p, z, 1.0f / 6.0f);
	p = fmaf(p, z, 1.0f);
	return s * p;
}
SNB_HD float det_acosf(float c) { // c in [0, 1]
	if (c > 0.5f) return 2.0f * det_asin_small(sqrtf((1.0f - c) * 0.5f));
	return 1.57079632679489662f - det_asin_small(c);
}

// sRGB transfer (include/neural-graphics-primitives/common_device.cuh:35-41,59-65).
SNB_HD float srgb_to_linear(float s) { return s <= 0.04045f ? s / 12.92f : powf((s + 0.055f) / 1.055f, 2.4f); }
SNB_HD float linear_to_srgb(float l) { return l < 0.0031308f ? 12.92f * l : 1.055f * powf(l, 0.41666f) - 0.055f; }

} // namespace snb
