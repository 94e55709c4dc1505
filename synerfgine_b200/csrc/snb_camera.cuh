// synerfgine_b200 — camera model: pinhole uv -> ray, and the per-pixel rolling-shutter / motion-blur transform.
//
// Replaces (reference): uv_to_ray pinhole path (include/neural-graphics-primitives/common_device.cuh:403-470),
// get_xform_given_rolling_shutter (:361-368) with the quaternion helpers of tcnn vec.h (:1056-1173), and the
// NeRF payload initialisation of init_rays_with_payload_kernel_nerf (src/testbed_nerf.cu:1855-1970).
// In the sng path camera1 == camera0 and rolling_shutter == (0,0,0,1) (src/testbed_nerf.cu:2512-2513,
// include/neural-graphics-primitives/testbed.h:1042), but the reference still runs every pixel's rotation
// through matrix -> quaternion -> slerp(q, q, t) -> normalize -> matrix, which perturbs it by an ulp or two;
// that round trip is kept so ray directions are identical.
#pragma once
#include "snb_device.cuh"

namespace snb {

struct quat { float w, x, y, z; };

SNB_HD quat quat_from_rot(const m3& m) {
	quat q;
	const float tr = m.c0.x + m.c1.y + m.c2.z;
	if (tr > 0.0f) {
		const float S = sqrtf(tr + 1.0f) * 2.0f;
		q.w = 0.25f * S; q.x = (m.c1.z - m.c2.y) / S; q.y = (m.c2.x - m.c0.z) / S; q.z = (m.c0.y - m.c1.x) / S;
	} else if (m.c0.x > m.c1.y && m.c0.x > m.c2.z) {
		const float S = sqrtf(1.0f + m.c0.x - m.c1.y - m.c2.z) * 2.0f;
		q.w = (m.c1.z - m.c2.y) / S; q.x = 0.25f * S; q.y = (m.c1.x + m.c0.y) / S; q.z = (m.c2.x + m.c0.z) / S;
	} else if (m.c1.y > m.c2.z) {
		const float S = sqrtf(1.0f + m.c1.y - m.c0.x - m.c2.z) * 2.0f;
		q.w = (m.c2.x - m.c0.z) / S; q.x = (m.c1.x + m.c0.y) / S; q.y = 0.25f * S; q.z = (m.c2.y + m.c1.z) / S;
	} else {
		const float S = sqrtf(1.0f + m.c2.z - m.c0.x - m.c1.y) * 2.0f;
		q.w = (m.c0.y - m.c1.x) / S; q.x = (m.c2.x + m.c0.z) / S; q.y = (m.c2.y + m.c1.z) / S; q.z = 0.25f * S;
	}
	return q;
}
SNB_HD float qdot(quat a, quat b) { return (a.w * b.w + a.x * b.x) + (a.y * b.y + a.z * b.z); }
SNB_HD quat qmul_r(quat a, float s) { quat r; r.w = a.w * s; r.x = a.x * s; r.y = a.y * s; r.z = a.z * s; return r; }
SNB_HD quat qmul_l(float s, quat a) { quat r; r.w = s * a.w; r.x = s * a.x; r.y = s * a.y; r.z = s * a.z; return r; }
SNB_HD quat qadd(quat a, quat b) { quat r; r.w = a.w + b.w; r.x = a.x + b.x; r.y = a.y + b.y; r.z = a.z + b.z; return r; }
SNB_HD quat qdiv(quat a, float s) { quat r; r.w = a.w / s; r.x = a.x / s; r.y = a.y / s; r.z = a.z / s; return r; }

SNB_HD quat qslerp(quat x, quat y, float t) {
	quat z = y;
	float cos_theta = qdot(x, y);
	if (cos_theta < 0.0f) { z.w = -y.w; z.x = -y.x; z.y = -y.y; z.z = -y.z; cos_theta = -cos_theta; }
	if (cos_theta > 1.0f - 1.1920928955078125e-7f) return qadd(qmul_r(x, 1.0f - t), qmul_r(z, t));
	const float angle = det_acosf(cos_theta);
	return qdiv(qadd(qmul_l(det_sinf((1.0f - t) * angle), x), qmul_l(det_sinf(t * angle), z)), det_sinf(angle));
}
// tcnn's quaternion normalize divides by length2()'s square root, i.e. the 4th root of dot(a,a) (vec.h:1099-1110).
SNB_HD quat qnormalize(quat a) {
	const float len = sqrtf(sqrtf(qdot(a, a)));
	if (len <= 0.0f) { quat r; r.w = 1.0f; r.x = r.y = r.z = 0.0f; return r; }
	return qdiv(a, len);
}
SNB_HD m3 quat_to_rot(quat q) {
	const float qxx = q.x * q.x, qyy = q.y * q.y, qzz = q.z * q.z;
	const float qxz = q.x * q.z, qxy = q.x * q.y, qyz = q.y * q.z;
	const float qwx = q.w * q.x, qwy = q.w * q.y, qwz = q.w * q.z;
	m3 r;
	r.c0 = mk3(1.0f - 2.0f * (qyy + qzz), 2.0f * (qxy + qwz), 2.0f * (qxz - qwy));
	r.c1 = mk3(2.0f * (qxy - qwz), 1.0f - 2.0f * (qxx + qzz), 2.0f * (qyz + qwx));
	r.c2 = mk3(2.0f * (qxz + qwy), 2.0f * (qyz - qwx), 1.0f - 2.0f * (qxx + qyy));
	return r;
}

// Camera of one pixel for start == end and rolling shutter (0,0,0,1).
SNB_HD m4x3 pixel_camera(const m4x3& cam, float motionblur_time, v2 uv) {
	const float pixel_t = 0.0f + 0.0f * uv.x + 0.0f * uv.y + 1.0f * motionblur_time;
	const v3 pos = cam.c3 + (cam.c3 - cam.c3) * pixel_t;
	const quat q = quat_from_rot(rot_of(cam));
	const m3 rot = quat_to_rot(qnormalize(qslerp(q, q, pixel_t)));
	m4x3 r; r.c0 = rot.c0; r.c1 = rot.c1; r.c2 = rot.c2; r.c3 = pos;
	return r;
}

// Pinhole uv -> (origin, unnormalised direction); no lens distortion, foveation, parallax shift or aperture.
SNB_HD void pinhole_ray(v2 uv, int res_x, int res_y, v2 focal, const m4x3& cam, v2 center, v3& o, v3& d) {
	v3 dir = mk3((uv.x - center.x) * (float)res_x / focal.x, (uv.y - center.y) * (float)res_y / focal.y, 1.0f);
	const v3 head = mk3(0.0f, 0.0f, 0.0f);
	dir = dir - head * 0.0f;
	dir = mul(rot_of(cam), dir);
	v3 origin = mul(rot_of(cam), head) + cam.c3;
	origin = origin + dir * 0.0f;
	o = origin; d = dir;
}

// ---- the general camera model (CameraEx) ---------------------------------------------------------------------------------------------
// get_xform_given_rolling_shutter (common_device.cuh:361-368) for two different cameras.
SNB_HD m4x3 pixel_camera_ex(const m4x3& cam0, const CameraEx& ex, float motionblur_time, v2 uv) {
	const float pixel_t = ex.rolling[0] + ex.rolling[1] * uv.x + ex.rolling[2] * uv.y + ex.rolling[3] * motionblur_time;
	const v3 pos = cam0.c3 + (ex.cam1.c3 - cam0.c3) * pixel_t;
	const m3 rot = quat_to_rot(qnormalize(qslerp(quat_from_rot(rot_of(cam0)), quat_from_rot(rot_of(ex.cam1)), pixel_t)));
	m4x3 r; r.c0 = rot.c0; r.c1 = rot.c1; r.c2 = rot.c2; r.c3 = pos;
	return r;
}

// opencv_lens_distortion_delta / opencv_fisheye_lens_distortion_delta (common_device.cuh:249-288)
SNB_HD void lens_delta(int mode, const float* p, float u, float v, float& du, float& dv) {
	if (mode == kLensOpenCV) {
		const float k1 = p[0], k2 = p[1], p1 = p[2], p2 = p[3];
		const float u2 = u * u, uv = u * v, v2 = v * v, r2 = u2 + v2;
		const float radial = k1 * r2 + k2 * r2 * r2;
		du = u * radial + 2.0f * p1 * uv + p2 * (r2 + 2.0f * u2);
		dv = v * radial + 2.0f * p2 * uv + p1 * (r2 + 2.0f * v2);
	} else {
		const float r = sqrtf(u * u + v * v);
		if (r > 2.220446049250313e-16f) {
			const float theta = atanf(r), t2 = theta * theta, t4 = t2 * t2, t6 = t4 * t2, t8 = t4 * t4;
			const float thetad = theta * (1.0f + p[0] * t2 + p[1] * t4 + p[2] * t6 + p[3] * t8);
			du = u * thetad / r - u; dv = v * thetad / r - v;
		} else { du = 0.0f; dv = 0.0f; }
	}
}
// iterative_lens_undistortion (common_device.cuh:290-329): Newton iteration with central differences
SNB_HD void lens_undistort(int mode, const float* p, float& u, float& v) {
	const float eps = 1.1920928955078125e-7f;
	const float x0 = u, y0 = v;
	float x = u, y = v;
	for (int i = 0; i < 100; ++i) {
		const float s0 = fmaxf(eps, fabsf(1e-6f * x)), s1 = fmaxf(eps, fabsf(1e-6f * y));
		float dx, dy, b0x, b0y, f0x, f0y, b1x, b1y, f1x, f1y;
		lens_delta(mode, p, x, y, dx, dy);
		lens_delta(mode, p, x - s0, y, b0x, b0y); lens_delta(mode, p, x + s0, y, f0x, f0y);
		lens_delta(mode, p, x, y - s1, b1x, b1y); lens_delta(mode, p, x, y + s1, f1x, f1y);
		const float j00 = 1.0f + (f0x - b0x) / (2.0f * s0), j10 = (f1x - b1x) / (2.0f * s1), j01 = (f0y - b0y) / (2.0f * s0), j11 = 1.0f + (f1y - b1y) / (2.0f * s1);
		// inverse(J) * (x + dx - x0), J column-major {{j00, j01}, {j10, j11}}
		const float rx = x + dx - x0, ry = y + dy - y0;
		const float det = j00 * j11 - j10 * j01;
		const float sx = (j11 * rx - j10 * ry) / det, sy = (-j01 * rx + j00 * ry) / det;
		x -= sx; y -= sy;
		if (sx * sx + sy * sy < 1e-10f) break;
	}
	u = x; v = y;
}
// square2disk_shirley (random_val.cuh:112-128)
SNB_HD v2 square_to_disk(v2 sq) {
	float phi, r;
	const float a = sq.x, b = sq.y;
	if (a * a > b * b) { r = a; phi = (kPi / 4.0f) * (b / a); }
	else { r = b; phi = (kPi / 2.0f) - (kPi / 4.0f) * (a / b); }
	v2 o; o.x = r * cosf(phi); o.y = r * sinf(phi);
	return o;
}

// uv_to_ray with every option (common_device.cuh:403-470). Returns false for an invalid ray (f-theta outside its domain).
SNB_HD bool general_ray(uint32_t spp, v2 uv, int res_x, int res_y, v2 focal, const m4x3& cam, v2 center, const CameraEx& ex, v3& o, v3& d) {
	v3 dir;
	if (ex.lens_mode == kLensFTheta) {
		const float xpix = (uv.x - center.x) * ex.lens[5], ypix = (uv.y - center.y) * ex.lens[6];
		const float norm = sqrtf(xpix * xpix + ypix * ypix);
		const float alpha = ex.lens[0] + norm * (ex.lens[1] + norm * (ex.lens[2] + norm * (ex.lens[3] + norm * ex.lens[4])));
		const float sa = sinf(alpha), ca = cosf(alpha);
		if (ca <= 1.17549435e-38f || norm == 0.0f) return false;
		const float k = sa * (1.0f / norm);
		dir = mk3(k * xpix, k * ypix, ca);
		if (dir.x == 0.0f && dir.y == 0.0f && dir.z == 0.0f) return false;
	} else if (ex.lens_mode == kLensLatLong) {
		const float theta = (uv.y - 0.5f) * kPi, phi = (uv.x - 0.5f) * kPi * 2.0f;
		dir = mk3(sinf(phi) * cosf(theta), sinf(theta), cosf(phi) * cosf(theta));
	} else if (ex.lens_mode == kLensEquirectangular) {
		const float ct = (uv.y - 0.5f) * 2.0f, st = sqrtf(fmaxf(1.0f - ct * ct, 0.0f)), phi = (uv.x - 0.5f) * kPi * 2.0f;
		dir = mk3(sinf(phi) * st, ct, cosf(phi) * st);
	} else {
		dir = mk3((uv.x - center.x) * (float)res_x / focal.x, (uv.y - center.y) * (float)res_y / focal.y, 1.0f);
		if (ex.lens_mode == kLensOpenCV || ex.lens_mode == kLensOpenCVFisheye) lens_undistort(ex.lens_mode, ex.lens, dir.x, dir.y);
	}
	const v3 head = mk3(ex.parallax.x, ex.parallax.y, 0.0f);
	dir = dir - head * ex.parallax.z;
	dir = mul(rot_of(cam), dir);
	v3 origin = mul(rot_of(cam), head) + cam.c3;
	if (ex.aperture != 0.0f) {
		const v3 lookat = origin + dir * ex.plane_z;
		const int px = (int)(uv.x * (float)res_x), py = (int)(uv.y * (float)res_y);
		const uint32_t seed = (uint32_t)px * 19349663u + (uint32_t)py * 96925573u;
		v2 sq; sq.x = ld_value(spp, seed, 0) * 2.0f - 1.0f; sq.y = ld_value(spp, seed, 1) * 2.0f - 1.0f;
		const v2 disk = square_to_disk(sq);
		origin = origin + cam.c0 * (ex.aperture * disk.x) + cam.c1 * (ex.aperture * disk.y);   // mat2x3(camera) * blur, column by column like mul()
		dir = (lookat - origin) / ex.plane_z;
	}
	origin = origin + dir * ex.near_distance;
	o = origin; d = dir;
	return true;
}

// read_envmap (include/neural-graphics-primitives/envmap.cuh:24-50): bilinear lookup by direction, wrapping in x, clamping in y.
#ifdef __CUDACC__
SNB_D float4 envmap_lookup(const CameraEx& ex, v3 dir) {
	const v3 dd = mk3(dir.z, -dir.x, dir.y);
	const float theta = acosf(fminf(fmaxf(dd.z, -1.0f), 1.0f)), phi = atan2f(dd.y, dd.x);
	const float cx = theta / kPi, cy = phi / (2.0f * kPi) + 0.5f;
	const float fx = cy * (float)(ex.env_w - 1), fy = cx * (float)(ex.env_h - 1);
	const int tx = (int)fx, ty = (int)fy;
	const float wx = fx - (float)tx, wy = fy - (float)ty;
	auto rd = [&](int x, int y) {
		if (x < 0) x += ex.env_w; else if (x >= ex.env_w) x -= ex.env_w;
		y = max(min(y, ex.env_h - 1), 0);
		return __ldg(ex.envmap + (size_t)y * ex.env_w + x);
	};
	const float4 a = rd(tx, ty), b = rd(tx + 1, ty), c = rd(tx, ty + 1), e = rd(tx + 1, ty + 1);
	const float w00 = (1 - wx) * (1 - wy), w10 = wx * (1 - wy), w01 = (1 - wx) * wy, w11 = wx * wy;
	return make_float4(((w00 * a.x + w10 * b.x) + w01 * c.x) + w11 * e.x, ((w00 * a.y + w10 * b.y) + w01 * c.y) + w11 * e.y,
	                   ((w00 * a.z + w10 * b.z) + w01 * c.z) + w11 * e.z, ((w00 * a.w + w10 * b.w) + w01 * c.w) + w11 * e.w);
}
#endif

SNB_HD bool primary_ray_ex(uint32_t idx, v2 uv, const DevView& view, v3& o, v3& d) {
	CameraEx ex = view.ex;
	if (ex.plane_z < 0.0f) ex.aperture = 0.0f;
	const m4x3 cam = pixel_camera_ex(view.cam, ex, ld_value(view.spp, idx * 72239731u), uv);
	return general_ray(view.spp, uv, view.width, view.height, view.focal, cam, view.center, ex, o, d);
}

// NeRF primary ray of pixel (x, y): returns false when the ray misses the render box. `valid_dir` (optional) tells whether o / d hold a
// ray at all (a general-model ray can be invalid); the environment map is looked up for every valid ray, hit or miss. EX = false compiles
// the extended camera out (the kernels are instantiated both ways, so the pinhole instance keeps its registers).
template <bool EX = false>
SNB_HD bool nerf_primary_ray(int x, int y, const DevView& view, const MarchEnv& env, v3& o, v3& d, float& t, bool* valid_dir = nullptr) {
	const uint32_t idx = (uint32_t)x + (uint32_t)view.width * (uint32_t)y;
	const v2 off = view.pix_off;
	v2 uv; uv.x = ((float)x + off.x) / (float)view.width; uv.y = ((float)y + off.y) / (float)view.height;
	if (valid_dir) *valid_dir = true;
	if (!EX || !view.ex.enabled) {
		const m4x3 cam = pixel_camera(view.cam, ld_value(view.spp, idx * 72239731u), uv);
		pinhole_ray(uv, view.width, view.height, view.focal, cam, view.center, o, d);
	} else if (!primary_ray_ex(idx, uv, view, o, d)) {
		if (valid_dir) *valid_dir = false;
		o = mk3(0.0f, 0.0f, 0.0f); d = o; t = 0.0f;
		return false;
	}
	d = normalize(d);
	t = fmaxf(env.render_aabb.ray_intersect(mul(env.to_local, o), mul(env.to_local, d)).x, 0.0f) + 1e-6f;
	return env.render_aabb.contains(mul(env.to_local, o + d * t));
}

} // namespace snb
