// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
// Deterministic logf / expf built from IEEE +,-,*,/ and fmaf only, so the same formula evaluated by
// the CUDA product (synerfgine_b200/csrc/det_math.cuh, written independently against this spec)
// yields bit-identical results. Spec:
//   det_logf(x), x > 0 normal:  x = m * 2^e with m in [2/3, 4/3);  s = (m-1)/(m+1);  z = s*s
//       p = z*(2/3 + z*(2/5 + z*(2/7 + z*(2/9 + z*(2/11)))))  (Horner with fmaf)
//       log(m) = fmaf(s, p, 2*s);  result = fmaf(e, LN2_HI, fmaf(e, LN2_LO, log(m)))
//   det_expf(x): n = floor(fmaf(x, LOG2E, 0.5)), clamped to [-125, 126];
//       r = fmaf(-n, LN2_HI, x); r = fmaf(-n, LN2_LO, r)
//       q = 1 + r + r^2/2 + ... + r^8/40320 (Horner with fmaf); result = q * 2^n (exponent-field add)
//   det_sinf(x), x in [0, pi/2]: odd Taylor polynomial through x^13, Horner with fmaf in z = x*x:
//       s = x * (1 + z*(-1/6 + z*(1/120 + z*(-1/5040 + z*(1/362880 + z*(-1/39916800 + z/6227020800))))))
//   det_acosf(c), c in [0, 1]:  asin_small(s) = s * (1 + z*(1/6 + z*(3/40 + z*(15/336 + z*(105/3456
//       + z*(945/42240 + z*(10395/599040 + z*(135135/9676800 + z*(2027025/175472640))))))))), z = s*s
//       c > 0.5:  2 * asin_small(sqrt((1 - c) * 0.5));   else:  PI_2 - asin_small(c)
//       (used only by the camera quaternion slerp, tcnn/vec.h:1121-1142, when its trig branch is taken)
// They replace logf/expf only inside the exponential stepping functions
// (nerf_device.cuh:379-431 to_stepping_space / from_stepping_space) when math mode is MATH_DET.
#include "oracle_common.h"

namespace orc {

int g_math_mode = MATH_LIBM;

static inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

static const float LN2_HI = 0.693145751953125f;      // 0x1.62e400p-1, 16 significant bits
static const float LN2_LO = 1.42860682030941723212e-6f; // ln2 - LN2_HI
static const float LOG2E = 1.44269504088896341f;

float det_logf(float x) {
	uint32_t ux = f2u(x);
	// exponent such that the mantissa lands in [2/3, 4/3): 0x3f2aaaab is 2/3 rounded up.
	int32_t e = (int32_t)(ux - 0x3f2aaaabu) >> 23;
	float m = u2f(ux - ((uint32_t)e << 23));
	float s = (m - 1.0f) / (m + 1.0f);
	float z = s * s;
	float p = 0.181818181818181818f;           // 2/11
	p = std::fmaf(p, z, 0.222222222222222222f); // 2/9
	p = std::fmaf(p, z, 0.285714285714285714f); // 2/7
	p = std::fmaf(p, z, 0.4f);                  // 2/5
	p = std::fmaf(p, z, 0.666666666666666667f); // 2/3
	p = p * z;
	float lm = std::fmaf(s, p, 2.0f * s);
	float fe = (float)e;
	return std::fmaf(fe, LN2_HI, std::fmaf(fe, LN2_LO, lm));
}

float det_expf(float x) {
	float fn = std::floor(std::fmaf(x, LOG2E, 0.5f));
	if (fn < -125.0f) fn = -125.0f;
	if (fn > 126.0f) fn = 126.0f;
	float r = std::fmaf(-fn, LN2_HI, x);
	r = std::fmaf(-fn, LN2_LO, r);
	float q = 2.48015873015873016e-5f;            // 1/40320
	q = std::fmaf(q, r, 1.98412698412698413e-4f); // 1/5040
	q = std::fmaf(q, r, 1.38888888888888889e-3f); // 1/720
	q = std::fmaf(q, r, 8.33333333333333333e-3f); // 1/120
	q = std::fmaf(q, r, 4.16666666666666667e-2f); // 1/24
	q = std::fmaf(q, r, 1.66666666666666667e-1f); // 1/6
	q = std::fmaf(q, r, 0.5f);
	q = std::fmaf(q, r, 1.0f);
	q = std::fmaf(q, r, 1.0f);
	int32_t n = (int32_t)fn;
	return u2f(f2u(q) + ((uint32_t)n << 23));
}

float det_sinf(float x) {
	float z = x * x;
	float p = 1.0f / 6227020800.0f;
	p = std::fmaf(p, z, -1.0f / 39916800.0f);
	p = std::fmaf(p, z, 1.0f / 362880.0f);
	p = std::fmaf(p, z, -1.0f / 5040.0f);
	p = std::fmaf(p, z, 1.0f / 120.0f);
	p = std::fmaf(p, z, -1.0f / 6.0f);
	p = std::fmaf(p, z, 1.0f);
	return x * p;
}

static float det_asin_small(float s) {
	float z = s * s;
	float p = 2027025.0f / 175472640.0f;
	p = std::fmaf(p, z, 135135.0f / 9676800.0f);
	p = std::fmaf(p, z, 10395.0f / 599040.0f);
	p = std::fmaf(p, z, 945.0f / 42240.0f);
	p = std::fmaf(p, z, 105.0f / 3456.0f);
	p = std::fmaf(p, z, 15.0f / 336.0f);
	p = std::fmaf(p, z, 3.0f / 40.0f);
	p = std::fmaf(p, z, 1.0f / 6.0f);
	p = std::fmaf(p, z, 1.0f);
	return s * p;
}

float det_acosf(float c) {
	if (c > 0.5f) return 2.0f * det_asin_small(std::sqrt((1.0f - c) * 0.5f));
	return 1.57079632679489662f - det_asin_small(c);
}

} // namespace orc

extern "C" {
void orc_set_math_mode(int mode) { orc::g_math_mode = mode; }
float orc_det_logf(float x) { return orc::det_logf(x); }
float orc_det_expf(float x) { return orc::det_expf(x); }
float orc_det_sinf(float x) { return orc::det_sinf(x); }
float orc_det_acosf(float x) { return orc::det_acosf(x); }
}
