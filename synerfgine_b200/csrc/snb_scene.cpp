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
	v3 view_dir, at_pos; float zoom; int vo_scale; float anim_speed; bool animations;
	std::vector<Keyframe> keyframes; int total_time_ms, fps, total_frames, frames_between;
};

static v3 jv3(const snbjson::Value& a) { return mk3(a.at((size_t)0).f(), a.at((size_t)1).f(), a.at((size_t)2).f()); }

static void set_err(char* err, size_t n, const std::string& m) { if (err && n) { std::strncpy(err, m.c_str(), n - 1); err[n - 1] = 0; } }

static int buffer_type_of(const std::string& n) {
	// sng::buffer_names, include/synerfgine/raytracer.cuh:33-43
	static const char* names[] = {"Final", "Next Origin", "Src Origin", "Next Direction", "Src Direction", "Normal", "Depth", "Nerf Shadow"};
	for (int i = 0; i < 8; ++i) if (n == names[i]) return i;
	return 0;
}
