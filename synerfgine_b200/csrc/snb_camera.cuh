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

// NeRF primary ray of pixel (x, y): returns false when the ray misses the render box.
SNB_HD bool nerf_primary_ray(int x, int y, const DevView& view, const MarchEnv& env, v3& o, v3& d, float& t) {
	const uint32_t idx = (uint32_t)x + (uint32_t)view.width * (uint32_t)y;
	const v2 off = view.pix_off;
	v2 uv; uv.x = ((float)x + off.x) / (float)view.width; uv.y = ((float)y + off.y) / (float)view.height;
	const m4x3 cam = pixel_camera(view.cam, ld_value(view.spp, idx * 72239731u), uv);
	pinhole_ray(uv, view.width, view.height, view.focal, cam, view.center, o, d);
	d = normalize(d);
	t = fmaxf(env.render_aabb.ray_intersect(mul(env.to_local, o), mul(env.to_local, d)).x, 0.0f) + 1e-6f;
	return env.render_aabb.contains(mul(env.to_local, o + d * t));
}

} // namespace snb
