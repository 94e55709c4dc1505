// synerfgine_b200 — host-side scene layer: virtual-scene JSON, OBJ meshes, BVH build, camera keyframes, animation.
//
// Replaces (reference): Engine::set_virtual_world + the "rendering" block of Engine::init
// (src/synerfgine/engine.cu:21-78,150-228), VirtualObject ctor / next_frame (src/synerfgine/virtual_object.cu:7-88;
// include/synerfgine/virtual_object.cuh:51-64), Material / Light JSON ctors and Light::next_frame
// (include/synerfgine/material.cuh:29-55; light.cuh:17-49), CamPath (include/synerfgine/cam_path.cuh:95-142),
// TriangleBvhWithBranchingFactor<2>::build (src/triangle_bvh.cu:615-718), the Testbed camera setters
// (src/testbed.cu:405-425,470-492) and the subset of tinyobjloader the reference relies on
// (vertex positions, faces with v / v/vt / v//vn / v/vt/vn and negative indices, quads split along the shorter diagonal).
#include <cstring>
#include <cmath>
#include <algorithm>
#include <fstream>
#include <sstream>
#include <stack>
#include <string>
#include <vector>
#include "../../include/snb.h"
#include "snb_json.h"
#include "snb_math.cuh"

using namespace snb;

namespace {

struct Tri9 { v3 a, b, c; };
static_assert(sizeof(Tri9) == 36, "triangle layout");
struct Node8 { v3 lo, hi; int left, right; };
static_assert(sizeof(Node8) == 32, "node layout");

v3 centroid(const Tri9& t) { return (t.a + t.b + t.c) / 3.0f; }
float centroid_axis(const Tri9& t, int axis) { return (comp(t.a, axis) + comp(t.b, axis) + comp(t.c, axis)) / 3; }
v3 vmin3(v3 a, v3 b) { return mk3(std::min(a.x, b.x), std::min(a.y, b.y), std::min(a.z, b.z)); }
v3 vmax3(v3 a, v3 b) { return mk3(std::max(a.x, b.x), std::max(a.y, b.y), std::max(a.z, b.z)); }

void bounds(const Tri9* begin, const Tri9* end, v3& lo, v3& hi) {
	lo = hi = begin->a;
	for (const Tri9* it = begin; it != end; ++it) {
		lo = vmin3(lo, it->a); hi = vmax3(hi, it->a);
		lo = vmin3(lo, it->b); hi = vmax3(hi, it->b);
		lo = vmin3(lo, it->c); hi = vmax3(hi, it->c);
	}
}

// Binary BVH: median split (nth_element) along the axis of largest centroid variance, leaves of <= prims_per_leaf
// triangles encoded as negative [begin, end) ranges; children of a node are adjacent (left_idx, left_idx + 1).
void build_bvh(std::vector<Tri9>& tris, int prims_per_leaf, std::vector<Node8>& nodes) {
	nodes.clear();
	nodes.emplace_back();
	bounds(tris.data(), tris.data() + tris.size(), nodes[0].lo, nodes[0].hi);
	struct Job { int node; size_t begin, end; };
	std::stack<Job> jobs;
	jobs.push({0, 0, tris.size()});
	while (!jobs.empty()) {
		const Job cur = jobs.top(); jobs.pop();
		const size_t n = cur.end - cur.begin;
		v3 mean = mk3(0, 0, 0);
		for (size_t i = cur.begin; i < cur.end; ++i) mean = mean + centroid(tris[i]);
		mean = mean / (float)n;
		v3 var = mk3(0, 0, 0);
		for (size_t i = cur.begin; i < cur.end; ++i) { const v3 d = centroid(tris[i]) - mean; var = var + d * d; }
		var = var / (float)n;
		const float max_val = vmax(var);
		const int axis = var.x == max_val ? 0 : (var.y == max_val ? 1 : 2);
		const size_t mid = cur.begin + n / 2;
		std::nth_element(tris.begin() + cur.begin, tris.begin() + mid, tris.begin() + cur.end,
		                 [axis](const Tri9& t1, const Tri9& t2) { return centroid_axis(t1, axis) < centroid_axis(t2, axis); });
		const Job kids[2] = {{0, cur.begin, mid}, {0, mid, cur.end}};
		nodes[cur.node].left = (int)nodes.size();
		for (int k = 0; k < 2; ++k) {
			Job child = kids[k];
			child.node = (int)nodes.size();
			nodes.emplace_back();
			Node8& nd = nodes.back();
			if (child.begin == child.end) { nd.lo = nd.hi = mk3(0, 0, 0); nd.left = nd.right = -(int)child.begin - 1; continue; }   // only for n == 1
			bounds(tris.data() + child.begin, tris.data() + child.end, nd.lo, nd.hi);
			if ((int)(child.end - child.begin) <= prims_per_leaf) { nd.left = -(int)child.begin - 1; nd.right = -(int)child.end - 1; }
			else jobs.push(child);
		}
		nodes[cur.node].right = (int)nodes.size();
	}
}

// ---- OBJ ---------------------------------------------------------------------------------------------------------------
bool load_obj(const std::string& path, std::vector<Tri9>& out, std::string& err) {
	std::ifstream f(path, std::ios::in | std::ios::binary);
	if (!f) { err = "Error loading file: " + path; return false; }
	std::vector<float> v;
	std::string line;
	auto fix = [&](long idx) -> long { const long nv = (long)(v.size() / 3); if (idx > 0) return idx - 1; if (idx < 0) return nv + idx; return -1; };
	auto vert = [&](long i) { return mk3(v[(size_t)i * 3], v[(size_t)i * 3 + 1], v[(size_t)i * 3 + 2]); };
	while (std::getline(f, line)) {
		size_t s = 0;
		while (s < line.size() && (line[s] == ' ' || line[s] == '\t')) ++s;
		if (s >= line.size() || line[s] == '#') continue;
		if (line[s] == 'v' && s + 1 < line.size() && (line[s + 1] == ' ' || line[s + 1] == '\t')) {
			const char* p = line.c_str() + s + 1; char* e;
			float xyz[3] = {0, 0, 0};
			for (int k = 0; k < 3; ++k) { xyz[k] = (float)std::strtod(p, &e); p = e; }
			v.push_back(xyz[0]); v.push_back(xyz[1]); v.push_back(xyz[2]);
		} else if (line[s] == 'f' && s + 1 < line.size() && (line[s + 1] == ' ' || line[s + 1] == '\t')) {
			std::vector<long> idx;
			const char* p = line.c_str() + s + 1;
			while (*p) {
				while (*p == ' ' || *p == '\t' || *p == '\r') ++p;
				if (!*p) break;
				char* e; const long vi = std::strtol(p, &e, 10);
				if (e == p) break;
				p = e;
				while (*p && *p != ' ' && *p != '\t' && *p != '\r') ++p;   // skip /vt/vn
				idx.push_back(fix(vi));
			}
			const long nv = (long)(v.size() / 3);
			bool ok = idx.size() >= 3;
			for (long i : idx) if (i < 0 || i >= nv) ok = false;
			if (!ok) continue;
			if (idx.size() == 3) { out.push_back({vert(idx[0]), vert(idx[1]), vert(idx[2])}); }
			else if (idx.size() == 4) {
				const v3 p0 = vert(idx[0]), p1 = vert(idx[1]), p2 = vert(idx[2]), p3 = vert(idx[3]);
				const v3 e02 = p2 - p0, e13 = p3 - p1;
				const float sqr02 = e02.x * e02.x + e02.y * e02.y + e02.z * e02.z, sqr13 = e13.x * e13.x + e13.y * e13.y + e13.z * e13.z;
				if (sqr02 < sqr13) { out.push_back({p0, p1, p2}); out.push_back({p0, p2, p3}); }
				else { out.push_back({p0, p1, p3}); out.push_back({p1, p2, p3}); }
			} else {
				for (size_t k = 1; k + 1 < idx.size(); ++k) out.push_back({vert(idx[0]), vert(idx[k]), vert(idx[k + 1])});   // fan (tinyobj ear-clips; convex faces agree)
			}
		}
	}
	if (out.empty()) { err = "no triangles in " + path; return false; }
	return true;
}

struct Keyframe { v3 view, at; float zoom; };

struct Object {
	std::vector<Tri9> tris; std::vector<Node8> nodes;
	m3 rot; v3 pos; float scale; int mat_id;
	v3 anim_centre, anim_axis; float anim_angle;
};
struct LightAnim { v3 start, end; float ratio, step; };

} // namespace

struct snb_scene {
	std::vector<Object> objects;
	std::vector<snb_material> materials;
	std::vector<snb_light> lights;
	std::vector<LightAnim> light_anim;
	snb_rt_params rt;
	// camera block
	v3 view_dir, at_pos; float zoom; int vo_scale; float anim_speed; bool animations; bool playing; int res_factor;
	std::vector<Keyframe> keyframes; int total_time_ms, fps, total_frames, frames_between;
};

static v3 jv3(const snbjson::Value& a) { return mk3(a.at((size_t)0).f(), a.at((size_t)1).f(), a.at((size_t)2).f()); }

static void set_err(char* err, size_t n, const std::string& m) { if (err && n) { std::strncpy(err, m.c_str(), n - 1); err[n - 1] = 0; } }

static int buffer_type_of(const std::string& n) {
	// sng::img_buffer_type_names, include/synerfgine/raytracer.cuh:33-44
	static const char* names[] = {"Final", "Next Origin", "Src Origin", "Next Direction", "Src Direction", "Normal", "Depth", "NerfShadow"};
	for (int i = 0; i < 8; ++i) if (n == names[i]) return i;
	return 0;
}

static void default_rt(snb_rt_params& rt) {
	// RayTracer header defaults (include/synerfgine/raytracer.cuh:156-167) + Engine defaults (engine.cuh:112-120) +
	// Testbed::sng_position_kernel_size (testbed.h:686), exposure 0, tonemap Identity.
	rt.n_steps = 8; rt.samples = 2; rt.bounces = 2; rt.shadow_iters = 4;
	rt.lens_angle_constant = 0.009f; rt.depth_offset = 0.1f; rt.syn_shadow_factor = 1.0f;
	rt.view_nerf_shadow = 1; rt.buffer_type = 0; rt.nerf_shadow_kernel = 1; rt.nerf_shadow_intensity = 2.0f;
	rt.nerf_on_nerf_shadow_threshold = 0.3f; rt.show_shadow_on_nerf = 1; rt.show_nerf = 1; rt.syn_px_scale = 1;
	rt.exposure = 0.0f; rt.tonemap_curve = 0;
}

extern "C" {

int snb_scene_load_json(const char* json_text, const char* base_dir, snb_scene** out, char* err, size_t err_len) {
	if (!json_text || !out) return SNB_ERR_ARG;
	*out = nullptr;
	std::unique_ptr<snb_scene> sc(new snb_scene());
	default_rt(sc->rt);
	sc->view_dir = mk3(0, 0, 0); sc->at_pos = mk3(0, 0, 0); sc->zoom = 1.0f; sc->vo_scale = 4; sc->anim_speed = 1.0f; sc->animations = false; sc->playing = false; sc->res_factor = 64;   // Testbed::m_fixed_res_factor default (engine.cuh:122)
	sc->total_time_ms = 10000; sc->fps = 24; sc->total_frames = 0; sc->frames_between = 1;
	try {
		snbjson::Ptr root = snbjson::parse(json_text);
		const snbjson::Value& cfg = *root;
		if (cfg.has("camera")) {
			const snbjson::Value& cam = cfg.at("camera");
			if (cam.has("view")) sc->view_dir = jv3(cam.at("view"));
			if (cam.has("at")) sc->at_pos = jv3(cam.at("at"));
			if (cam.has("zoom")) sc->zoom = cam.at("zoom").f();
			if (cam.has("vo_scale")) sc->vo_scale = cam.at("vo_scale").i();
			if (cam.has("animation_speed")) { sc->anim_speed = cam.at("animation_speed").f(); sc->animations = sc->anim_speed > 0.0f; }
			if (cam.has("path")) {
				// CamPath(config): "frames" first (accessed unconditionally by the reference; tolerated if absent here), then "path"
				if (cam.has("frames")) for (size_t i = 0; i < cam.at("frames").size(); ++i) { const auto& k = cam.at("frames").at(i); sc->keyframes.push_back({jv3(k.at("view")), jv3(k.at("at")), k.at("zoom").f()}); }
				if (cam.has("total_time_ms")) sc->total_time_ms = cam.at("total_time_ms").i();
				if (cam.has("fps")) sc->fps = cam.at("fps").i();
				if (cam.has("move_on_start")) sc->playing = cam.at("move_on_start").boolean();
				sc->total_frames = sc->total_time_ms * sc->fps / 1000;
				for (size_t i = 0; i < cam.at("path").size(); ++i) { const auto& k = cam.at("path").at(i); sc->keyframes.push_back({jv3(k.at("view")), jv3(k.at("at")), k.at("zoom").f()}); }
				sc->frames_between = sc->total_frames / std::max((int)sc->keyframes.size() - 1, 1);
			}
		}
		if (cfg.has("rendering")) {
			const snbjson::Value& r = cfg.at("rendering");
			snb_rt_params& rt = sc->rt;
			if (r.has("exposure")) rt.exposure = r.at("exposure").f();
			if (r.has("path_trace_depth")) rt.bounces = (uint32_t)r.at("path_trace_depth").i();
			if (r.has("light_samples")) rt.samples = (uint32_t)r.at("light_samples").i();
			if (r.has("nerf_shadow_samples")) rt.nerf_shadow_kernel = r.at("nerf_shadow_samples").i();
			if (r.has("nerf_shadow_intensity")) rt.nerf_shadow_intensity = r.at("nerf_shadow_intensity").f();
			if (r.has("syn_shadow_samples")) rt.shadow_iters = (uint32_t)r.at("syn_shadow_samples").i();
			if (r.has("syn_shadow_intensity")) rt.syn_shadow_factor = r.at("syn_shadow_intensity").f();
			if (r.has("lens_size")) rt.lens_angle_constant = r.at("lens_size").f();
			if (r.has("nerf_on_nerf_shadow_threshold")) rt.nerf_on_nerf_shadow_threshold = r.at("nerf_on_nerf_shadow_threshold").f();
			if (r.has("shadow_on_virtual_obj")) rt.view_nerf_shadow = r.at("shadow_on_virtual_obj").boolean();
			if (r.has("shadow_on_nerf")) rt.show_shadow_on_nerf = r.at("shadow_on_nerf").boolean();
			if (r.has("show_nerf")) rt.show_nerf = r.at("show_nerf").boolean();
			if (r.has("syn_filter")) rt.buffer_type = buffer_type_of(r.at("syn_filter").s());
			if (r.has("depth_offset")) rt.depth_offset = r.at("depth_offset").f();
			if (r.has("res_factor")) sc->res_factor = std::max(1, (int)r.at("res_factor").i());
			// clear_color, smooth_threshold, attenuation, show_light_pos, max_shadow_variance, nerf_ao_intensity,
			// show_virtual_obj, nerf_filter: read by the reference but without effect on this path (or GUI-only).
		}
		const snbjson::Value& mats = cfg.at("materials");
		for (size_t i = 0; i < mats.size(); ++i) {
			const snbjson::Value& m = mats.at(i);
			snb_material o; memset(&o, 0, sizeof o);
			const std::string type = m.at("type").s();
			const v3 kd = jv3(m.at("kd"));
			o.kd[0] = kd.x; o.kd[1] = kd.y; o.kd[2] = kd.z;
			o.ks[0] = o.ks[1] = o.ks[2] = 1.0f;
			if (m.has("ka")) { const v3 ka = jv3(m.at("ka")); o.ka[0] = ka.x; o.ka[1] = ka.y; o.ka[2] = ka.z; }
			if (m.has("ks")) { const v3 ks = jv3(m.at("ks")); o.ks[0] = ks.x; o.ks[1] = ks.y; o.ks[2] = ks.z; }
			o.n = m.at("n").f();
			o.rg = m.has("rg") ? m.at("rg").f() : 0.0f;
			if (type == "lambertian") { o.type = 0; o.spec_angle = 0.0f; }
			else if (type == "glossy") { o.type = 1; o.spec_angle = m.has("spec_angle") ? m.at("spec_angle").f() : 0.001f; }
			else throw std::runtime_error("Material type " + type + " not supported");
			sc->materials.push_back(o);
		}
		const snbjson::Value& objs = cfg.at("objfile");
		const std::string base = base_dir ? std::string(base_dir) : std::string();
		for (size_t i = 0; i < objs.size(); ++i) {
			const snbjson::Value& oc = objs.at(i);
			Object ob;
			ob.rot = identity3(); ob.pos = mk3(0, 0, 0);
			ob.scale = oc.has("scale") ? oc.at("scale").f() : 1.0f;
			const int ppl = oc.has("primitives-per-leaf") ? oc.at("primitives-per-leaf").i() : 4;
			if (oc.has("pos")) ob.pos = jv3(oc.at("pos"));
			if (oc.has("rot")) { const auto& a = oc.at("rot"); float r[9]; for (size_t k = 0; k < 9; ++k) r[k] = a.at(k).f(); ob.rot.c0 = mk3(r[0], r[1], r[2]); ob.rot.c1 = mk3(r[3], r[4], r[5]); ob.rot.c2 = mk3(r[6], r[7], r[8]); }
			ob.anim_centre = mk3(0, 0, 0); ob.anim_axis = mk3(0, 0, 0); ob.anim_angle = 0.0f;
			if (oc.has("anim")) { ob.anim_centre = jv3(oc.at("anim").at("rot_center")); ob.anim_axis = jv3(oc.at("anim").at("rot_axis")); ob.anim_angle = oc.at("anim").at("rot_angle").f(); }
			ob.mat_id = oc.at("material").i();
			if (ob.mat_id < 0 || (size_t)ob.mat_id >= sc->materials.size()) throw std::runtime_error("object material index out of range");
			std::string path = oc.at("file").s();
			if (!path.empty() && path[0] != '/' && !base.empty()) path = base + "/" + path;
			std::string e;
			if (!load_obj(path, ob.tris, e)) throw std::runtime_error(e);
			build_bvh(ob.tris, ppl, ob.nodes);
			sc->objects.push_back(std::move(ob));
		}
		const snbjson::Value& ls = cfg.at("lights");
		for (size_t i = 0; i < ls.size(); ++i) {
			const snbjson::Value& l = ls.at(i);
			snb_light o; memset(&o, 0, sizeof o);
			const v3 p = jv3(l.at("pos"));
			o.pos[0] = p.x; o.pos[1] = p.y; o.pos[2] = p.z;
			o.intensity = l.at("intensity").f(); o.size = l.at("size").f(); o.type = 0;
			if (l.has("type")) { const std::string t = l.at("type").s(); if (t == "point") o.type = 0; else if (t == "directional") o.type = 1; else throw std::runtime_error(t + " light not recognized"); }
			LightAnim an; an.start = p; an.end = p; an.ratio = 0.0f; an.step = 0.0f;
			if (l.has("anim")) { an.end = jv3(l.at("anim").at("end")); an.step = l.at("anim").at("step").f(); }
			sc->lights.push_back(o); sc->light_anim.push_back(an);
		}
	} catch (const std::exception& e) {
		set_err(err, err_len, e.what());
		return SNB_ERR_PARSE;
	}
	sc->rt.syn_px_scale = 1;
	*out = sc.release();
	return SNB_OK;
}

void snb_scene_free(snb_scene* s) { delete s; }

int snb_scene_counts(const snb_scene* s, uint32_t* n_objects, uint32_t* n_materials, uint32_t* n_lights) {
	if (!s) return SNB_ERR_ARG;
	if (n_objects) *n_objects = (uint32_t)s->objects.size();
	if (n_materials) *n_materials = (uint32_t)s->materials.size();
	if (n_lights) *n_lights = (uint32_t)s->lights.size();
	return SNB_OK;
}

int snb_scene_get(const snb_scene* s, snb_object* objects, snb_material* materials, snb_light* lights, snb_rt_params* rt) {
	if (!s) return SNB_ERR_ARG;
	if (objects) for (size_t i = 0; i < s->objects.size(); ++i) {
		const Object& o = s->objects[i]; snb_object& d = objects[i];
		d.nodes = reinterpret_cast<const float*>(o.nodes.data()); d.tris = reinterpret_cast<const float*>(o.tris.data());
		d.n_nodes = (int32_t)o.nodes.size(); d.n_tris = (int32_t)o.tris.size();
		const v3 cols[3] = {o.rot.c0, o.rot.c1, o.rot.c2};
		for (int k = 0; k < 3; ++k) { d.rot[k * 3] = cols[k].x; d.rot[k * 3 + 1] = cols[k].y; d.rot[k * 3 + 2] = cols[k].z; }
		d.pos[0] = o.pos.x; d.pos[1] = o.pos.y; d.pos[2] = o.pos.z; d.scale = o.scale; d.mat_id = o.mat_id;
	}
	if (materials) for (size_t i = 0; i < s->materials.size(); ++i) materials[i] = s->materials[i];
	if (lights) for (size_t i = 0; i < s->lights.size(); ++i) lights[i] = s->lights[i];
	if (rt) *rt = s->rt;
	return SNB_OK;
}

int snb_scene_camera(const snb_scene* s, float* view3, float* at3, float* zoom, float* vo_scale, uint32_t* n_keyframes) {
	if (!s) return SNB_ERR_ARG;
	if (view3) { view3[0] = s->view_dir.x; view3[1] = s->view_dir.y; view3[2] = s->view_dir.z; }
	if (at3) { at3[0] = s->at_pos.x; at3[1] = s->at_pos.y; at3[2] = s->at_pos.z; }
	if (zoom) *zoom = s->zoom;
	if (vo_scale) *vo_scale = (float)s->vo_scale;
	if (n_keyframes) *n_keyframes = (uint32_t)s->keyframes.size();
	return SNB_OK;
}

// CamPath::set_to_frame + CamKeyframe::interpolate for an explicit frame index; total_frames overrides the JSON's
// total_time_ms * fps / 1000 when non-zero (the benchmark resamples the same keyframes to 24 / 64 frames).
int snb_scene_camera_at(const snb_scene* s, uint32_t frame, uint32_t total_frames, float* view3, float* at3, float* zoom) {
	if (!s || s->keyframes.size() < 2) return SNB_ERR_STATE;
	const int total = total_frames ? (int)total_frames : s->total_frames;
	const int between = std::max(1, total / std::max((int)s->keyframes.size() - 1, 1));
	int cur_frame = (int)frame;
	int cur_key = cur_frame / between, next_key = cur_key + 1;
	if (next_key >= (int)s->keyframes.size()) { cur_frame = 0; cur_key = 0; next_key = 1; }
	const Keyframe& a = s->keyframes[cur_key]; const Keyframe& b = s->keyframes[next_key];
	const float k = (float)(cur_frame % between) / (float)between;
	const float invk = (float)(1.0 - k);
	const v3 v = invk * a.view + k * b.view, at = invk * a.at + k * b.at;
	if (view3) { view3[0] = v.x; view3[1] = v.y; view3[2] = v.z; }
	if (at3) { at3[0] = at.x; at3[1] = at.y; at3[2] = at.z; }
	if (zoom) *zoom = invk * a.zoom + k * b.zoom;
	return SNB_OK;
}

// Engine::resize (engine.cu:236-255): the NeRF layer renders at window * min(1, 8 / res_factor) (clamped to [window / 16, window]), the mesh
// layer at min(nerf_res * vo_scale, window); syn_px_scale = mesh width / NeRF width. res_factor 8 (every BASELINE config) gives equal
// resolutions and scale 1.
int snb_scene_resolution(const snb_scene* s, int window_w, int window_h, int* nerf_w, int* nerf_h, int* syn_w, int* syn_h, int* syn_px_scale) {
	if (!s || window_w <= 0 || window_h <= 0) return SNB_ERR_ARG;
	const float factor = std::min(1.0f, 8.0f / (float)s->res_factor);
	auto clampi = [](int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); };
	const int nw = clampi((int)((float)window_w * factor), window_w / 16, window_w), nh = clampi((int)((float)window_h * factor), window_h / 16, window_h);
	const float vs = (float)s->vo_scale;
	const int rw = std::min(clampi((int)((float)nw * vs), nw / 16, nw * 16), window_w), rh = std::min(clampi((int)((float)nh * vs), nh / 16, nh * 16), window_h);
	if (nerf_w) *nerf_w = nw; if (nerf_h) *nerf_h = nh; if (syn_w) *syn_w = rw; if (syn_h) *syn_h = rh;
	if (syn_px_scale) *syn_px_scale = std::max(1, rw / std::max(1, nw));
	return SNB_OK;
}

// CamPath / animation switches of the scene: total_frames = total_time_ms * fps / 1000, playing = "move_on_start", animations =
// "animation_speed" > 0 (engine.cu:43-45; cam_path.cuh:98-115).
int snb_scene_path_info(const snb_scene* s, uint32_t* total_frames, int32_t* playing, int32_t* animations, int32_t* res_factor) {
	if (!s) return SNB_ERR_ARG;
	if (total_frames) *total_frames = (uint32_t)std::max(0, s->total_frames);
	if (playing) *playing = s->playing ? 1 : 0;
	if (animations) *animations = s->animations ? 1 : 0;
	if (res_factor) *res_factor = s->res_factor;
	return SNB_OK;
}

int snb_scene_advance(snb_scene* s) {
	if (!s) return SNB_ERR_ARG;
	if (!s->animations) return SNB_OK;
	const float speed = s->anim_speed;
	for (Object& o : s->objects) {
		if (o.anim_angle == 0.0f) continue;
		const v3 ax = o.anim_axis;
		const float cost = std::cos(o.anim_angle * speed), sint = std::sin(o.anim_angle * speed);
		m3 R;   // columns as listed in the reference's mat3 initialiser
		R.c0 = mk3(cost + ax.x * ax.x * (1.0f - cost), ax.x * ax.y * (1.0f - cost) - ax.z * sint, ax.x * ax.z * (1.0f - cost) + ax.y * sint);
		R.c1 = mk3(ax.x * ax.y * (1.0f - cost) + ax.z * sint, cost + ax.y * ax.y * (1.0f - cost), ax.y * ax.z * (1.0f - cost) - ax.x * sint);
		R.c2 = mk3(ax.z * ax.y * (1.0f - cost) - ax.y * sint, ax.z * ax.y * (1.0f - cost) + ax.x * sint, cost + ax.z * ax.z * (1.0f - cost));
		o.pos = mul(R, mul(o.rot, o.pos - o.anim_centre)) + o.anim_centre;
	}
	for (size_t i = 0; i < s->lights.size(); ++i) {
		LightAnim& a = s->light_anim[i];
		if (a.step == 0.0f) continue;
		float next = a.ratio + a.step;
		if (next > 1.0 || next < 0.0) { a.step = -a.step; next = a.ratio + a.step; }
		a.ratio = next;
		const v3 p = (1.0f - a.ratio) * a.start + a.ratio * a.end;
		s->lights[i].pos[0] = p.x; s->lights[i].pos[1] = p.y; s->lights[i].pos[2] = p.z;
	}
	return SNB_OK;
}

int snb_bvh_build(float* tris, int32_t n_tris, int32_t prims_per_leaf, float* nodes_out, int32_t nodes_cap, int32_t* n_nodes_out) {
	if (!tris || n_tris <= 0 || prims_per_leaf <= 0 || !n_nodes_out) return SNB_ERR_ARG;
	std::vector<Tri9> t(reinterpret_cast<Tri9*>(tris), reinterpret_cast<Tri9*>(tris) + n_tris);
	std::vector<Node8> nodes;
	build_bvh(t, prims_per_leaf, nodes);
	*n_nodes_out = (int32_t)nodes.size();
	if (!nodes_out) return SNB_OK;
	if ((int32_t)nodes.size() > nodes_cap) return SNB_ERR_ARG;
	memcpy(tris, t.data(), (size_t)n_tris * sizeof(Tri9));
	memcpy(nodes_out, nodes.data(), nodes.size() * sizeof(Node8));
	return SNB_OK;
}

int snb_obj_load(const char* path, float* tris_out, int32_t tris_cap, int32_t* n_tris_out) {
	if (!path || !n_tris_out) return SNB_ERR_ARG;
	std::vector<Tri9> t; std::string e;
	if (!load_obj(path, t, e)) return SNB_ERR_IO;
	*n_tris_out = (int32_t)t.size();
	if (!tris_out) return SNB_OK;
	if ((int32_t)t.size() > tris_cap) return SNB_ERR_ARG;
	memcpy(tris_out, t.data(), t.size() * sizeof(Tri9));
	return SNB_OK;
}

// Testbed camera state after reset_camera() (fov 50.625, scale 1.5) followed by set_view_dir(view), set_look_at(at),
// set_scale(zoom) — what Engine::init does with the JSON camera block (engine.cu:148-151). out12 = mat4x3 column-major.
int snb_camera_from_view(const float* view3, const float* at3, float zoom, float* out12) {
	if (!view3 || !at3 || !out12) return SNB_ERR_ARG;
	float scale = 1.5f;
	v3 c0 = mk3(1, 0, 0), c1 = mk3(0, -1, 0), c2 = mk3(0, 0, -1), c3 = mk3(0.5f, 0.5f, 0.5f);
	c3 = c3 - scale * c2;
	const v3 up = mk3(0, 1, 0);
	const v3 dir = mk3(view3[0], view3[1], view3[2]);
	if (dot(dir, dir) != 0.0f) {
		const v3 old_look_at = c3 + c2 * scale;
		c0 = normalize(cross(dir, up));
		c1 = normalize(cross(dir, c0));
		c2 = normalize(dir);
		c3 = c3 + (old_look_at - (c3 + c2 * scale));                       // set_look_at(old_look_at)
		c3 = c3 + (mk3(at3[0], at3[1], at3[2]) - (c3 + c2 * scale));       // set_look_at(at)
		const v3 prev = c3 + c2 * scale;                                    // set_scale(zoom)
		c3 = (c3 - prev) * (zoom / scale) + prev;
		scale = zoom;
	}
	const v3 cols[4] = {c0, c1, c2, c3};
	for (int k = 0; k < 4; ++k) { out12[k * 3] = cols[k].x; out12[k * 3 + 1] = cols[k].y; out12[k * 3 + 2] = cols[k].z; }
	return SNB_OK;
}

// ---- camera_log_lerp (include/neural-graphics-primitives/common_device.cuh:658-660): exp(log(B * A^-1) * t) * A on 4x4 matrices, fp32,
// with tcnn's matrix functions (vec.h:880-958): Denman-Beavers square roots until the matrix is within 1e-5 of the identity, the
// Mercator series for the logarithm, scaling-and-squaring with a (6,6) Pade approximant for the exponential.
} // extern "C"

namespace {
struct M4 { float m[4][4]; };   // m[column][row]
M4 m4_identity() { M4 r{}; for (int i = 0; i < 4; ++i) r.m[i][i] = 1.0f; return r; }
M4 m4_mul(const M4& a, const M4& b) {
	M4 r{};
	for (int c = 0; c < 4; ++c) for (int row = 0; row < 4; ++row) { float acc = 0.0f; for (int k = 0; k < 4; ++k) acc += a.m[k][row] * b.m[c][k]; r.m[c][row] = acc; }
	return r;
}
M4 m4_add(const M4& a, const M4& b) { M4 r; for (int c = 0; c < 4; ++c) for (int k = 0; k < 4; ++k) r.m[c][k] = a.m[c][k] + b.m[c][k]; return r; }
M4 m4_sub(const M4& a, const M4& b) { M4 r; for (int c = 0; c < 4; ++c) for (int k = 0; k < 4; ++k) r.m[c][k] = a.m[c][k] - b.m[c][k]; return r; }
M4 m4_scale(float s, const M4& a) { M4 r; for (int c = 0; c < 4; ++c) for (int k = 0; k < 4; ++k) r.m[c][k] = s * a.m[c][k]; return r; }
float m4_frobenius(const M4& a) { float acc = 0.0f; for (int c = 0; c < 4; ++c) for (int k = 0; k < 4; ++k) acc += a.m[c][k] * a.m[c][k]; return std::sqrt(acc); }
M4 m4_inverse(const M4& a) {   // Gauss-Jordan with partial pivoting
	float w[4][8];
	for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) { w[r][c] = a.m[c][r]; w[r][4 + c] = r == c ? 1.0f : 0.0f; }
	for (int col = 0; col < 4; ++col) {
		int piv = col;
		for (int r = col + 1; r < 4; ++r) if (std::fabs(w[r][col]) > std::fabs(w[piv][col])) piv = r;
		if (piv != col) for (int c = 0; c < 8; ++c) std::swap(w[piv][c], w[col][c]);
		const float d = w[col][col];
		for (int c = 0; c < 8; ++c) w[col][c] /= d;
		for (int r = 0; r < 4; ++r) if (r != col) { const float f = w[r][col]; if (f != 0.0f) for (int c = 0; c < 8; ++c) w[r][c] -= f * w[col][c]; }
	}
	M4 out;
	for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) out.m[c][r] = w[r][4 + c];
	return out;
}
M4 m4_sqrt(const M4& m) {
	M4 X = m, Y = m4_identity();
	for (int i = 0; i < 32; ++i) {
		if (m4_frobenius(m4_sub(m4_mul(X, X), m)) < 1e-10f) return X;
		const M4 iX = m4_inverse(X);
		X = m4_scale(0.5f, m4_add(X, m4_inverse(Y)));
		Y = m4_scale(0.5f, m4_add(Y, iX));
	}
	return X;
}
M4 m4_log(const M4& m) {
	M4 r = m;
	int j = 0;
	for (; j < 32; ++j) {
		if (m4_frobenius(m4_sub(r, m4_identity())) < 1e-5f) break;
		r = m4_sqrt(r);
	}
	const M4 A = m4_sub(r, m4_identity());
	M4 Z = A, X = A;
	for (int i = 2; i < 32; ++i) {
		if (m4_frobenius(Z) < 1e-10f) break;
		Z = m4_mul(Z, A);
		X = m4_add(X, m4_scale(1.0f / (float)i, Z));
	}
	return m4_scale(std::ldexp(1.0f, j), X);
}
M4 m4_exp(const M4& m) {
	const float norm = m4_frobenius(m);
	const int squarings = norm > 0.0f ? std::max(0, 1 + (int)std::floor(std::log2(norm))) : 0;
	const M4 a = m4_scale(std::ldexp(1.0f, -squarings), m);
	M4 mX = m4_identity(), mD = m4_identity(), mN = m4_identity();
	float c = 1.0f, sgn = -1.0f;
	const int q = 6;
	for (int k = 1; k <= q; ++k) {
		c = c * (float)(q - k + 1) / (float)(k * (2 * q - k + 1));
		mX = m4_mul(a, mX);
		const M4 cmX = m4_scale(c, mX);
		mN = m4_add(mN, cmX);
		mD = m4_add(mD, m4_scale(sgn, cmX));
		sgn = -sgn;
	}
	M4 r = m4_mul(m4_inverse(mD), mN);
	for (int i = 0; i < squarings; ++i) r = m4_mul(r, r);
	return r;
}
M4 m4_from_camera(const float* c12) { M4 r = m4_identity(); for (int c = 0; c < 4; ++c) for (int k = 0; k < 3; ++k) r.m[c][k] = c12[c * 3 + k]; return r; }
}  // namespace

extern "C" {

// ≙ camera_log_lerp: the end-of-shutter camera of one motion-blur sample (src/python_api.cu:161: camera_log_lerp(start, end, shutter_fraction)).
int snb_camera_log_lerp(const float* a12, const float* b12, float t, float* out12) {
	if (!a12 || !b12 || !out12) return SNB_ERR_ARG;
	const M4 A = m4_from_camera(a12), B = m4_from_camera(b12);
	const M4 r = m4_mul(m4_exp(m4_scale(t, m4_log(m4_mul(B, m4_inverse(A))))), A);
	for (int c = 0; c < 4; ++c) for (int k = 0; k < 3; ++k) out12[c * 3 + k] = r.m[c][k];
	return SNB_OK;
}

} // extern "C"

