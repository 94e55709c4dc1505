// synerfgine_b200 — present stage: the reference's GL blit + FXAA fragment shader + 8-bit read-back,
// as one CUDA kernel writing a headless RGB8 buffer (top-down rows, what `output-%03d.png` holds).
//
// Replaces (reference): scripts/virtual_desc/main.frag:24-38 (unwarp), :51-95 (fxaa), :98-107 (main);
// src/synerfgine/display.cu:104-109 (full-screen triangle UVs), :224-229 + :285 (syn_rgba sampler: GL_REPEAT, GL_LINEAR),
// :275-281 (clear + ONE / ONE_MINUS_SRC_ALPHA blend), :305-322 (glReadPixels RGB8 + vertical flip -> PNG).
//
// One thread per output pixel, 9 bilinear taps (36 float4 texel reads, L1/L2 hits: neighbouring threads share them), 3 bytes
// out. HBM-bound by construction: 16 B in + 3 B out per pixel when the window has the frame's resolution.
// Implementation-defined GL behaviour is fixed as in oracle/present.cpp (fp32 bilinear weights, spec-order weighted sum,
// unorm8 = floor(clamp(c) * 255 + 0.5)); arithmetic order is identical, so the bytes match the oracle exactly.
#pragma once
#include "snb_device.cuh"

namespace snb {

struct FovWarp { float al, bl, cl, am, bm, ar, br, cr, switch_left, switch_right, inv_switch_left, inv_switch_right; };

SNB_D float fov_unwarp(const FovWarp& w, float y) {
	y = fminf(fmaxf(y, 0.0f), 1.0f);
	if (y < w.inv_switch_left) return (sqrtf(-4.0f * w.al * w.cl + 4.0f * w.al * y + w.bl * w.bl) - w.bl) / (2.0f * w.al);
	else if (y > w.inv_switch_right) return (sqrtf(-4.0f * w.ar * w.cr + 4.0f * w.ar * y + w.br * w.br) - w.br) / (2.0f * w.ar);
	else return (y - w.bm) / w.am;
}

SNB_D int wrap_repeat(int i, int n) {   // GL_REPEAT; in range (the common case) costs one compare instead of an integer division
	if ((unsigned)i < (unsigned)n) return i;
	i %= n;
	return i < 0 ? i + n : i;
}

SNB_D float4 tex_linear(const float4* __restrict__ tex, int w, int h, float u, float v, int ox, int oy) {
	const float x = u * (float)w - 0.5f, y = v * (float)h - 0.5f;
	const float fx = floorf(x), fy = floorf(y);
	const float a = x - fx, b = y - fy;
	const int i0 = wrap_repeat((int)fx + ox, w), j0 = wrap_repeat((int)fy + oy, h);
	const int i1 = i0 + 1 == w ? 0 : i0 + 1, j1 = j0 + 1 == h ? 0 : j0 + 1;
	const float4 t00 = __ldg(tex + (size_t)j0 * w + i0), t10 = __ldg(tex + (size_t)j0 * w + i1);
	const float4 t01 = __ldg(tex + (size_t)j1 * w + i0), t11 = __ldg(tex + (size_t)j1 * w + i1);
	const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
	float4 r;
	r.x = w00 * t00.x + w10 * t10.x + w01 * t01.x + w11 * t11.x;
	r.y = w00 * t00.y + w10 * t10.y + w01 * t01.y + w11 * t11.y;
	r.z = w00 * t00.z + w10 * t10.z + w01 * t01.z + w11 * t11.z;
	r.w = w00 * t00.w + w10 * t10.w + w01 * t01.w + w11 * t11.w;
	return r;
}

SNB_D float luma_of(float4 c) { return c.x * 0.299f + c.y * 0.587f + c.z * 0.114f; }
SNB_D unsigned char to_unorm8(float c) { c = fminf(fmaxf(c, 0.0f), 1.0f); return (unsigned char)floorf(c * 255.0f + 0.5f); }

// Block = 4 warps, each warp = 32 consecutive pixels of one output row (texel reads of a tap are 512 contiguous bytes when the
// window has the frame's resolution); the warp's 96 output bytes are staged in shared memory and leave as 24 aligned 32-bit
// stores when the row pitch allows it.
__global__ void __launch_bounds__(128) k_present(const float4* __restrict__ tex, int in_w, int in_h, int out_w, int out_h, float3 clear, FovWarp wx, FovWarp wy,
                                                unsigned char* __restrict__ out) {
	__shared__ __align__(16) unsigned char s_out[4][96];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int x0 = blockIdx.x * 32, x = x0 + lane, r = blockIdx.y * 4 + warp;
	if (r >= out_h) return;                                   // whole warp
	unsigned char px[3] = {0, 0, 0};
	if (x < out_w) {
		const int y_gl = out_h - 1 - r;
		const float uv_x = ((float)x + 0.5f) / (float)out_w, uv_y = ((float)y_gl + 0.5f) / (float)out_h;
		const float u = fov_unwarp(wx, uv_x), v = fov_unwarp(wy, 1.0f - uv_y);
		const float ivx = 1.0f / (float)out_w, ivy = 1.0f / (float)out_h;
		const float4 nw = tex_linear(tex, in_w, in_h, u, v, -1, 1), ne = tex_linear(tex, in_w, in_h, u, v, 1, 1);
		const float4 sw = tex_linear(tex, in_w, in_h, u, v, -1, -1), se = tex_linear(tex, in_w, in_h, u, v, 1, -1);
		const float4 m = tex_linear(tex, in_w, in_h, u, v, 0, 0);
		const float lnw = luma_of(nw), lne = luma_of(ne), lsw = luma_of(sw), lse = luma_of(se), lm = luma_of(m);
		const float lmin = fminf(lm, fminf(fminf(lnw, lne), fminf(lsw, lse)));
		const float lmax = fmaxf(lm, fmaxf(fmaxf(lnw, lne), fmaxf(lsw, lse)));
		float dx = -((lnw + lne) - (lsw + lse));
		float dy = ((lnw + lsw) - (lne + lse));
		const float reduce = fmaxf((lnw + lne + lsw + lse) * (0.25f * (1.0f / 8.0f)), 1.0f / 128.0f);
		const float rcp = 1.0f / (fminf(fabsf(dx), fabsf(dy)) + reduce);
		dx = fminf(8.0f, fmaxf(-8.0f, dx * rcp)) * ivx;
		dy = fminf(8.0f, fmaxf(-8.0f, dy * rcp)) * ivy;
		const float k1 = 1.0f / 3.0f - 0.5f, k2 = 2.0f / 3.0f - 0.5f;
		const float4 a1 = tex_linear(tex, in_w, in_h, u + dx * k1, v + dy * k1, 0, 0), a2 = tex_linear(tex, in_w, in_h, u + dx * k2, v + dy * k2, 0, 0);
		const float4 b1 = tex_linear(tex, in_w, in_h, u + dx * -0.5f, v + dy * -0.5f, 0, 0), b2 = tex_linear(tex, in_w, in_h, u + dx * 0.5f, v + dy * 0.5f, 0, 0);
		float3 A, B;
		A.x = 0.5f * (a1.x + a2.x); A.y = 0.5f * (a1.y + a2.y); A.z = 0.5f * (a1.z + a2.z);
		B.x = A.x * 0.5f + 0.25f * (b1.x + b2.x); B.y = A.y * 0.5f + 0.25f * (b1.y + b2.y); B.z = A.z * 0.5f + 0.25f * (b1.z + b2.z);
		const float lb = B.x * 0.299f + B.y * 0.587f + B.z * 0.114f;
		const float3 c = (lb < lmin || lb > lmax) ? A : B;
		const float ia = 1.0f - m.w;
		px[0] = to_unorm8(c.x + clear.x * ia); px[1] = to_unorm8(c.y + clear.y * ia); px[2] = to_unorm8(c.z + clear.z * ia);
	}
	unsigned char* row_out = out + ((size_t)r * out_w + x0) * 3;
	const int n_px = min(32, out_w - x0);
	if (n_px == 32 && ((reinterpret_cast<size_t>(row_out) & 3) == 0)) {
		s_out[warp][lane * 3] = px[0]; s_out[warp][lane * 3 + 1] = px[1]; s_out[warp][lane * 3 + 2] = px[2];
		__syncwarp();
		if (lane < 24) reinterpret_cast<uint32_t*>(row_out)[lane] = reinterpret_cast<const uint32_t*>(s_out[warp])[lane];
	} else if (x < out_w) {
		row_out[lane * 3] = px[0]; row_out[lane * 3 + 1] = px[1]; row_out[lane * 3 + 2] = px[2];
	}
}

} // namespace snb
