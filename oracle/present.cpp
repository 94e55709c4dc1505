// oracle/present.cpp — TEST INFRASTRUCTURE ONLY (see oracle.h). CPU restatement of the reference's present stage:
// the full-screen blit with the FXAA fragment shader, the premultiplied blend over the clear colour, the 8-bit
// framebuffer and the glReadPixels + vertical flip that feed `output-%03d.png`.
//
// Follows (trxe/SyNeRFgine @ b842828):
//   scripts/virtual_desc/main.frag:24-38   unwarp (piecewise-quadratic foveation inverse)
//   scripts/virtual_desc/main.frag:51-95   fxaa
//   scripts/virtual_desc/main.frag:98-107  main: y flip, unwarp, fxaa(syn_rgba, tex_coords, full_resolution)
//   src/synerfgine/display.cu:104-109      full-screen triangle, UVs = fragment centre / viewport
//   src/synerfgine/display.cu:224-229      sampler state of syn_rgba: GL_REPEAT, GL_LINEAR (present passes GL_LINEAR, :285)
//   src/synerfgine/display.cu:275-281      clear to clear_color, blend ONE / ONE_MINUS_SRC_ALPHA
//   src/synerfgine/display.cu:305-322      glReadPixels GL_RGB / UNSIGNED_BYTE, flipped vertically on write
//
// Parity status: UNPINNED against GL (graphics stack unavailable on this box). What GL leaves to the implementation is
// fixed here and in the CUDA kernel identically: bilinear weights in full fp32 (hardware uses 8-bit fractions), the
// weighted sum in the order of the GL spec's equation (t00, t10, t01, t11), unorm8 = floor(clamp(c) * 255 + 0.5);
// min / max are IEEE fmin / fmax (a NaN operand is ignored, so a NaN colour packs to 0), as CUDA's fminf / fmaxf are.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>

namespace {

struct Warp { float al, bl, cl, am, bm, ar, br, cr, switch_left, switch_right, inv_switch_left, inv_switch_right; };
struct V3 { float x, y, z; };
struct V4 { float x, y, z, w; };

float unwarp1(const Warp& w, float y) {   // main.frag:24-34
	y = std::fmin(std::fmax(y, 0.0f), 1.0f);
	if (y < w.inv_switch_left) return (std::sqrt(-4.0f * w.al * w.cl + 4.0f * w.al * y + w.bl * w.bl) - w.bl) / (2.0f * w.al);
	else if (y > w.inv_switch_right) return (std::sqrt(-4.0f * w.ar * w.cr + 4.0f * w.ar * y + w.br * w.br) - w.br) / (2.0f * w.ar);
	else return (y - w.bm) / w.am;
}

int wrap(int i, int n) { i %= n; return i < 0 ? i + n : i; }   // GL_REPEAT

struct Tex { const float* rgba; int w, h; };

V4 texel(const Tex& t, int i, int j) {
	const float* p = t.rgba + ((size_t)wrap(j, t.h) * t.w + wrap(i, t.w)) * 4;
	return {p[0], p[1], p[2], p[3]};
}

// GL_LINEAR sample at normalised (u, v) with an integer texel offset (textureProjOffset with w = 1).
V4 sample(const Tex& t, float u, float v, int ox, int oy) {
	const float x = u * (float)t.w - 0.5f, y = v * (float)t.h - 0.5f;
	const float fx = std::floor(x), fy = std::floor(y);
	const float a = x - fx, b = y - fy;
	const int i0 = (int)fx + ox, j0 = (int)fy + oy;
	const V4 t00 = texel(t, i0, j0), t10 = texel(t, i0 + 1, j0), t01 = texel(t, i0, j0 + 1), t11 = texel(t, i0 + 1, j0 + 1);
	const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
	V4 r;
	r.x = w00 * t00.x + w10 * t10.x + w01 * t01.x + w11 * t11.x;
	r.y = w00 * t00.y + w10 * t10.y + w01 * t01.y + w11 * t11.y;
	r.z = w00 * t00.z + w10 * t10.z + w01 * t01.z + w11 * t11.z;
	r.w = w00 * t00.w + w10 * t10.w + w01 * t01.w + w11 * t11.w;
	return r;
}

float luma_of(V4 c) { return c.x * 0.299f + c.y * 0.587f + c.z * 0.114f; }   // dot(rgb, luma)

V4 fxaa(const Tex& t, float u, float v, float res_x, float res_y) {   // main.frag:51-95
	const float ivx = 1.0f / res_x, ivy = 1.0f / res_y;
	const V4 nw = sample(t, u, v, -1, 1), ne = sample(t, u, v, 1, 1), sw = sample(t, u, v, -1, -1), se = sample(t, u, v, 1, -1);
	const V4 m = sample(t, u, v, 0, 0);
	const float lnw = luma_of(nw), lne = luma_of(ne), lsw = luma_of(sw), lse = luma_of(se), lm = luma_of(m);
	const float lmin = std::fmin(lm, std::fmin(std::fmin(lnw, lne), std::fmin(lsw, lse)));
	const float lmax = std::fmax(lm, std::fmax(std::fmax(lnw, lne), std::fmax(lsw, lse)));
	float dx = -((lnw + lne) - (lsw + lse));
	float dy = ((lnw + lsw) - (lne + lse));
	const float reduce = std::fmax((lnw + lne + lsw + lse) * (0.25f * (1.0f / 8.0f)), 1.0f / 128.0f);
	const float rcp = 1.0f / (std::fmin(std::fabs(dx), std::fabs(dy)) + reduce);
	dx = std::fmin(8.0f, std::fmax(-8.0f, dx * rcp)) * ivx;
	dy = std::fmin(8.0f, std::fmax(-8.0f, dy * rcp)) * ivy;
	const float k1 = 1.0f / 3.0f - 0.5f, k2 = 2.0f / 3.0f - 0.5f;
	const V4 a1 = sample(t, u + dx * k1, v + dy * k1, 0, 0), a2 = sample(t, u + dx * k2, v + dy * k2, 0, 0);
	const V4 b1 = sample(t, u + dx * -0.5f, v + dy * -0.5f, 0, 0), b2 = sample(t, u + dx * 0.5f, v + dy * 0.5f, 0, 0);
	V4 A, B;
	A.x = 0.5f * (a1.x + a2.x); A.y = 0.5f * (a1.y + a2.y); A.z = 0.5f * (a1.z + a2.z); A.w = m.w;
	B.x = A.x * 0.5f + 0.25f * (b1.x + b2.x); B.y = A.y * 0.5f + 0.25f * (b1.y + b2.y); B.z = A.z * 0.5f + 0.25f * (b1.z + b2.z); B.w = m.w;
	const float lb = luma_of(B);
	return (lb < lmin || lb > lmax) ? A : B;
}

uint8_t unorm8(float c) { c = std::fmin(std::fmax(c, 0.0f), 1.0f); return (uint8_t)std::floor(c * 255.0f + 0.5f); }

} // namespace

extern "C" {

// rgba: in_w x in_h float4, row 0 = top image row (the buffer Engine::frame uploads as the syn_rgba texture).
// out: out_w x out_h x 3 bytes, row 0 = top (what the PNG holds after stbi's vertical flip of the bottom-up read-back).
// warp_x / warp_y: 12 floats each in the order of main.frag's FoveationWarp, or nullptr for the identity.
void orc_present(const float* rgba, int in_w, int in_h, int out_w, int out_h, const float* clear_rgb, const float* warp_x, const float* warp_y, uint8_t* out) {
	const Warp ident = {0, 0, 0, 1, 0, 0, 0, 0, 0, 1, 0, 1};
	Warp wx = ident, wy = ident;
	if (warp_x) std::memcpy(&wx, warp_x, sizeof(Warp));
	if (warp_y) std::memcpy(&wy, warp_y, sizeof(Warp));
	const Tex t = {rgba, in_w, in_h};
	for (int r = 0; r < out_h; ++r)
		for (int x = 0; x < out_w; ++x) {
			const int y_gl = out_h - 1 - r;                                   // fragment row, bottom-up
			const float uv_x = ((float)x + 0.5f) / (float)out_w, uv_y = ((float)y_gl + 0.5f) / (float)out_h;
			const float u = unwarp1(wx, uv_x), v = unwarp1(wy, 1.0f - uv_y);
			const V4 c = fxaa(t, u, v, (float)out_w, (float)out_h);
			const float ia = 1.0f - c.w;                                      // ONE, ONE_MINUS_SRC_ALPHA over the clear colour
			uint8_t* o = out + ((size_t)r * out_w + x) * 3;
			o[0] = unorm8(c.x + clear_rgb[0] * ia); o[1] = unorm8(c.y + clear_rgb[1] * ia); o[2] = unorm8(c.z + clear_rgb[2] * ia);
		}
}

} // extern "C"
