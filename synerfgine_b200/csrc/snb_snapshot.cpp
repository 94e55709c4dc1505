// synerfgine_b200 — on-disk model format: instant-ngp / SyNeRFgine snapshots (`.ingp` = zlib- or
// gzip-compressed msgpack, `.msgpack` = plain), reader and writer. Host C++ only.
//
// Replaces (reference): Testbed::save_snapshot / load_snapshot (src/testbed.cu:4812-4876, 4878-4992): the msgpack image of
// m_network_config with the "snapshot" sub-object; tcnn Trainer::serialize / deserialize
// (dependencies/tiny-cuda-nn/include/tiny-cuda-nn/trainer.h:274-312: n_params, params_type, params_binary);
// json_binding.h:27-35 (BoundingBox = {"min", "max"}) and tcnn vec_json.h:40-73 (vectors = arrays, matrices = arrays of ROWS);
// network set-up constants derived at load time: reset_network (testbed.cu:3705-3741: base_resolution, per_level_scale) and
// load_nerf_post (testbed_nerf.cu:3078-3085: max_cascade, cone_angle_constant from aabb_scale).
// Third-party formats restated: MessagePack (spec 2.0 type bytes), as nlohmann::json::to_msgpack emits it (json.hpp 3.x: floats
// as float32 when lossless else float64, binaries as bin8/16/32, integers in the smallest type); zlib / gzip via the system zlib.
#include <zlib.h>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <new>
#include <stdexcept>
#include <string>
#include <vector>
#include "../../include/snb.h"

namespace {

struct Node {
	enum Kind { NIL, BOOL, INT, UINT, FLOAT, STR, BIN, ARR, MAP } kind = NIL;
	bool b = false; int64_t i = 0; uint64_t u = 0; double f = 0.0;
	std::string s;                         // STR
	const uint8_t* bin = nullptr; size_t bin_len = 0;   // BIN: view into the file image
	std::vector<Node> arr;
	std::vector<std::pair<std::string, Node>> map;
	const Node* get(const char* key) const {
		if (kind != MAP) return nullptr;
		for (auto& kv : map) if (kv.first == key) return &kv.second;
		return nullptr;
	}
	bool number(double& out) const {
		switch (kind) { case INT: out = (double)i; return true; case UINT: out = (double)u; return true; case FLOAT: out = f; return true; case BOOL: out = b; return true; default: return false; }
	}
};

struct Reader {
	const uint8_t* p; const uint8_t* end; std::string err; int depth = 0;
	bool need(size_t n) { if ((size_t)(end - p) < n) { err = "truncated msgpack"; return false; } return true; }
	uint64_t be(int n) { uint64_t v = 0; for (int k = 0; k < n; ++k) v = (v << 8) | *p++; return v; }
	bool str(size_t n, Node& o) { if (!need(n)) return false; o.kind = Node::STR; o.s.assign((const char*)p, n); p += n; return true; }
	bool bin(size_t n, Node& o) { if (!need(n)) return false; o.kind = Node::BIN; o.bin = p; o.bin_len = n; p += n; return true; }
	// every element takes at least one byte (a map entry two): a count beyond the remaining bytes is a malformed file, not an allocation request
	bool arr(size_t n, Node& o) { if (!need(n)) return false; o.kind = Node::ARR; o.arr.resize(n); for (auto& e : o.arr) if (!value(e)) return false; return true; }
	bool map(size_t n, Node& o) {
		if (n > (size_t)(end - p) / 2) { err = "truncated msgpack"; return false; }
		o.kind = Node::MAP; o.map.resize(n);
		for (auto& kv : o.map) { Node k; if (!value(k)) return false; if (k.kind != Node::STR) { err = "non-string map key"; return false; } kv.first = k.s; if (!value(kv.second)) return false; }
		return true;
	}
	bool value(Node& o) {
		if (++depth > 64) { err = "msgpack nesting too deep"; return false; }
		bool ok = value1(o); --depth; return ok;
	}
	bool value1(Node& o) {
		if (!need(1)) return false;
		const uint8_t t = *p++;
		if (t <= 0x7f) { o.kind = Node::UINT; o.u = t; return true; }
		if (t >= 0xe0) { o.kind = Node::INT; o.i = (int8_t)t; return true; }
		if ((t & 0xf0) == 0x80) return map(t & 0x0f, o);
		if ((t & 0xf0) == 0x90) return arr(t & 0x0f, o);
		if ((t & 0xe0) == 0xa0) return str(t & 0x1f, o);
		switch (t) {
			case 0xc0: o.kind = Node::NIL; return true;
			case 0xc2: case 0xc3: o.kind = Node::BOOL; o.b = t == 0xc3; return true;
			case 0xc4: if (!need(1)) return false; return bin((size_t)be(1), o);
			case 0xc5: if (!need(2)) return false; return bin((size_t)be(2), o);
			case 0xc6: if (!need(4)) return false; return bin((size_t)be(4), o);
			case 0xc7: { if (!need(2)) return false; size_t n = (size_t)be(1); ++p; return bin(n, o); }      // ext8: keep the payload
			case 0xc8: { if (!need(3)) return false; size_t n = (size_t)be(2); ++p; return bin(n, o); }
			case 0xc9: { if (!need(5)) return false; size_t n = (size_t)be(4); ++p; return bin(n, o); }
			case 0xca: { if (!need(4)) return false; uint32_t v = (uint32_t)be(4); float f; std::memcpy(&f, &v, 4); o.kind = Node::FLOAT; o.f = f; return true; }
			case 0xcb: { if (!need(8)) return false; uint64_t v = be(8); double f; std::memcpy(&f, &v, 8); o.kind = Node::FLOAT; o.f = f; return true; }
			case 0xcc: if (!need(1)) return false; o.kind = Node::UINT; o.u = be(1); return true;
			case 0xcd: if (!need(2)) return false; o.kind = Node::UINT; o.u = be(2); return true;
			case 0xce: if (!need(4)) return false; o.kind = Node::UINT; o.u = be(4); return true;
			case 0xcf: if (!need(8)) return false; o.kind = Node::UINT; o.u = be(8); return true;
			case 0xd0: if (!need(1)) return false; o.kind = Node::INT; o.i = (int8_t)be(1); return true;
			case 0xd1: if (!need(2)) return false; o.kind = Node::INT; o.i = (int16_t)be(2); return true;
			case 0xd2: if (!need(4)) return false; o.kind = Node::INT; o.i = (int32_t)be(4); return true;
			case 0xd3: if (!need(8)) return false; o.kind = Node::INT; o.i = (int64_t)be(8); return true;
			case 0xd4: case 0xd5: case 0xd6: case 0xd7: case 0xd8: { const size_t n = (size_t)1 << (t - 0xd4); if (!need(1 + n)) return false; ++p; return bin(n, o); }   // fixext
			case 0xd9: if (!need(1)) return false; return str((size_t)be(1), o);
			case 0xda: if (!need(2)) return false; return str((size_t)be(2), o);
			case 0xdb: if (!need(4)) return false; return str((size_t)be(4), o);
			case 0xdc: if (!need(2)) return false; return arr((size_t)be(2), o);
			case 0xdd: if (!need(4)) return false; return arr((size_t)be(4), o);
			case 0xde: if (!need(2)) return false; return map((size_t)be(2), o);
			case 0xdf: if (!need(4)) return false; return map((size_t)be(4), o);
		}
		err = "unknown msgpack type byte";
		return false;
	}
};

struct Writer {
	std::vector<uint8_t> out;
	void be(uint64_t v, int n) { for (int k = n - 1; k >= 0; --k) out.push_back((uint8_t)(v >> (8 * k))); }
	void map(size_t n) { if (n < 16) out.push_back(0x80 | (uint8_t)n); else { out.push_back(0xde); be(n, 2); } }
	void arr(size_t n) { if (n < 16) out.push_back(0x90 | (uint8_t)n); else { out.push_back(0xdc); be(n, 2); } }
	void str(const std::string& s) { if (s.size() < 32) out.push_back(0xa0 | (uint8_t)s.size()); else { out.push_back(0xd9); be(s.size(), 1); } out.insert(out.end(), s.begin(), s.end()); }
	void uint(uint64_t v) { if (v < 128) out.push_back((uint8_t)v); else if (v < 256) { out.push_back(0xcc); be(v, 1); } else if (v < 65536) { out.push_back(0xcd); be(v, 2); } else if (v < (1ull << 32)) { out.push_back(0xce); be(v, 4); } else { out.push_back(0xcf); be(v, 8); } }
	void f32(float f) { uint32_t v; std::memcpy(&v, &f, 4); out.push_back(0xca); be(v, 4); }
	void boolean(bool b) { out.push_back(b ? 0xc3 : 0xc2); }
	void bin(const void* p, size_t n) { if (n < 256) { out.push_back(0xc4); be(n, 1); } else if (n < 65536) { out.push_back(0xc5); be(n, 2); } else { out.push_back(0xc6); be(n, 4); } const uint8_t* b = (const uint8_t*)p; out.insert(out.end(), b, b + n); }
	void key(const char* k) { str(k); }
	void vec(const float* v, int n) { arr(n); for (int i = 0; i < n; ++i) f32(v[i]); }
};

bool read_file(const char* path, std::vector<uint8_t>& data) {
	FILE* f = std::fopen(path, "rb");
	if (!f) return false;
	std::fseek(f, 0, SEEK_END); const long n = std::ftell(f); std::fseek(f, 0, SEEK_SET);
	data.resize(n > 0 ? (size_t)n : 0);
	const bool ok = n >= 0 && std::fread(data.data(), 1, data.size(), f) == data.size();
	std::fclose(f);
	return ok;
}

bool inflate_all(const std::vector<uint8_t>& in, std::vector<uint8_t>& out, std::string& err) {
	z_stream zs; std::memset(&zs, 0, sizeof zs);
	if (inflateInit2(&zs, 15 + 32) != Z_OK) { err = "inflateInit2 failed"; return false; }   // +32: zlib or gzip header, detected
	zs.next_in = const_cast<Bytef*>(in.data()); zs.avail_in = (uInt)in.size();
	out.resize(std::max<size_t>(in.size() * 4, 1 << 16));
	int rc = Z_OK;
	while (rc != Z_STREAM_END) {
		if (zs.total_out == out.size()) out.resize(out.size() * 2);
		zs.next_out = out.data() + zs.total_out; zs.avail_out = (uInt)std::min<size_t>(out.size() - zs.total_out, 1u << 30);
		rc = inflate(&zs, Z_NO_FLUSH);
		if (rc != Z_OK && rc != Z_STREAM_END) { inflateEnd(&zs); err = "corrupt compressed snapshot"; return false; }
		if (rc == Z_OK && zs.avail_in == 0 && zs.avail_out != 0) { inflateEnd(&zs); err = "truncated compressed snapshot"; return false; }
	}
	out.resize(zs.total_out);
	inflateEnd(&zs);
	return true;
}

bool vecN(const Node* n, float* out, int count) {
	if (!n || n->kind != Node::ARR || (int)n->arr.size() != count) return false;
	for (int i = 0; i < count; ++i) { double d; if (!n->arr[i].number(d)) return false; out[i] = (float)d; }
	return true;
}
// matrices are arrays of ROWS (vec_json.h:40-48); `out` is column-major [col][row]
bool mat(const Node* n, float* out, int cols, int rows) {
	if (!n || n->kind != Node::ARR || (int)n->arr.size() != rows) return false;
	for (int r = 0; r < rows; ++r) {
		const Node& row = n->arr[r];
		if (row.kind != Node::ARR || (int)row.arr.size() != cols) return false;
		for (int c = 0; c < cols; ++c) { double d; if (!row.arr[c].number(d)) return false; out[c * rows + r] = (float)d; }
	}
	return true;
}
uint32_t uval(const Node* n, uint32_t dflt) { double d; return n && n->number(d) ? (uint32_t)d : dflt; }
float fval(const Node* n, float dflt) { double d; return n && n->number(d) ? (float)d : dflt; }
int activation_of(const Node* n, int dflt) {
	if (!n || n->kind != Node::STR) return dflt;
	if (n->s == "None") return SNB_ACT_NONE;
	if (n->s == "ReLU") return SNB_ACT_RELU;
	if (n->s == "Logistic") return SNB_ACT_LOGISTIC;
	if (n->s == "Exponential") return SNB_ACT_EXPONENTIAL;
	return dflt;
}

} // namespace

struct snb_snapshot {
	std::vector<uint8_t> image;            // decompressed msgpack; BIN nodes point into it
	Node root;
	snb_model_desc model;
	const uint16_t* density = nullptr; uint64_t n_density = 0;
	float camera[12]; bool has_camera = false;
	uint32_t aabb_scale = 1, version = 0;
	std::string error;
};

static void set_err(char* err, size_t n, const std::string& s) { if (err && n) { std::snprintf(err, n, "%s", s.c_str()); } }

extern "C" {

static int snapshot_load_impl(const char* path, snb_snapshot** out, char* err, size_t err_len);
int snb_snapshot_load(const char* path, snb_snapshot** out, char* err, size_t err_len) {
	if (!path || !out) return SNB_ERR_ARG;
	*out = nullptr;
	try { return snapshot_load_impl(path, out, err, err_len); }     // nothing may unwind through the C boundary
	catch (const std::bad_alloc&) { set_err(err, err_len, "out of memory while reading the snapshot"); return SNB_ERR_PARSE; }
	catch (const std::exception& e) { set_err(err, err_len, std::string("snapshot: ") + e.what()); return SNB_ERR_PARSE; }
}
static int snapshot_load_impl(const char* path, snb_snapshot** out, char* err, size_t err_len) {
	std::vector<uint8_t> file;
	if (!read_file(path, file)) { set_err(err, err_len, std::string("cannot read ") + path); return SNB_ERR_IO; }
	std::unique_ptr<snb_snapshot> s(new snb_snapshot());
	std::string e;
	const bool compressed = file.size() >= 2 && ((file[0] == 0x1f && file[1] == 0x8b) || (file[0] == 0x78 && ((file[0] << 8 | file[1]) % 31) == 0));
	if (compressed) { if (!inflate_all(file, s->image, e)) { set_err(err, err_len, e); return SNB_ERR_PARSE; } }
	else s->image.swap(file);
	Reader rd{s->image.data(), s->image.data() + s->image.size(), std::string()};
	if (!rd.value(s->root) || s->root.kind != Node::MAP) { set_err(err, err_len, rd.err.empty() ? "snapshot root is not a map" : rd.err); return SNB_ERR_PARSE; }
	const Node* snap = s->root.get("snapshot");
	if (!snap || snap->kind != Node::MAP) { set_err(err, err_len, "file has no 'snapshot' object (File::load_snapshot, testbed.cu:4879-4882)"); return SNB_ERR_PARSE; }
	const Node* mode = snap->get("mode");
	if (mode && mode->kind == Node::STR && mode->s != "nerf" && mode->s != "Nerf") { set_err(err, err_len, "snapshot mode is '" + mode->s + "', only NeRF snapshots are in scope"); return SNB_ERR_UNSUPPORTED; }
	s->version = uval(snap->get("version"), 0);
	if (s->version < 1) { set_err(err, err_len, "Snapshot uses an old format and can not be loaded."); return SNB_ERR_UNSUPPORTED; }   // testbed.cu:4880-4882 (SNAPSHOT_FORMAT_VERSION 1)
	snb_model_desc& m = s->model; std::memset(&m, 0, sizeof m);
	// ---- encoding (reset_network, testbed.cu:3705-3741) ----
	const Node* enc = s->root.get("encoding");
	if (!enc || enc->kind != Node::MAP) { set_err(err, err_len, "no 'encoding' object"); return SNB_ERR_PARSE; }
	{   // the path implements tcnn's hash grid (otype HashGrid, or Grid with type Hash) and 64-wide fully fused MLPs with 1 + 2 hidden layers
		const Node* ot = enc->get("otype"); const Node* ty = enc->get("type");
		const std::string o = ot && ot->kind == Node::STR ? ot->s : "HashGrid", t = ty && ty->kind == Node::STR ? ty->s : "Hash";
		if (!((o == "HashGrid") || (o == "Grid" && t == "Hash"))) { set_err(err, err_len, "encoding '" + o + "' is not supported (expected HashGrid)"); return SNB_ERR_UNSUPPORTED; }
		const Node* net = s->root.get("network"); const Node* rgb = s->root.get("rgb_network");
		if (net && net->kind == Node::MAP && (uval(net->get("n_neurons"), 64) != 64 || uval(net->get("n_hidden_layers"), 1) != 1)) { set_err(err, err_len, "density network must be 64 neurons x 1 hidden layer"); return SNB_ERR_UNSUPPORTED; }
		if (rgb && rgb->kind == Node::MAP && (uval(rgb->get("n_neurons"), 64) != 64 || uval(rgb->get("n_hidden_layers"), 2) != 2)) { set_err(err, err_len, "rgb network must be 64 neurons x 2 hidden layers"); return SNB_ERR_UNSUPPORTED; }
	}
	m.n_features_per_level = uval(enc->get("n_features_per_level"), 2);
	m.n_levels = enc->get("n_features") ? uval(enc->get("n_features"), 32) / std::max(1u, m.n_features_per_level) : uval(enc->get("n_levels"), 16);
	m.log2_hashmap_size = uval(enc->get("log2_hashmap_size"), 15);
	m.base_resolution = uval(enc->get("base_resolution"), 0);
	if (!m.base_resolution) m.base_resolution = 1u << (m.log2_hashmap_size / 3);
	const Node* nerf = snap->get("nerf");
	s->aabb_scale = uval(nerf ? nerf->get("aabb_scale") : nullptr, 1);
	if (s->aabb_scale < 1) s->aabb_scale = 1;
	m.per_level_scale = fval(enc->get("per_level_scale"), 0.0f);
	if (m.per_level_scale <= 0.0f && m.n_levels > 1)
		m.per_level_scale = std::exp(std::log(2048.0f * (float)s->aabb_scale / (float)m.base_resolution) / (float)(m.n_levels - 1));
	// ---- parameters (Trainer::deserialize) ----
	const Node* pt = snap->get("params_type");
	if (pt && pt->kind == Node::STR && pt->s != "__half") { set_err(err, err_len, "params_type '" + pt->s + "' is not supported (expected __half, the inference type)"); return SNB_ERR_UNSUPPORTED; }
	const Node* pb = snap->get("params_binary");
	if (!pb || pb->kind != Node::BIN || pb->bin_len % 2) { set_err(err, err_len, "no params_binary"); return SNB_ERR_PARSE; }
	m.params = reinterpret_cast<const uint16_t*>(pb->bin); m.n_params = pb->bin_len / 2;
	if (snap->get("n_params") && uval(snap->get("n_params"), 0) != m.n_params) { set_err(err, err_len, "n_params does not match params_binary"); return SNB_ERR_PARSE; }
	// ---- boxes, stepping constants (load_nerf_post, testbed_nerf.cu:3078-3085) ----
	const float s_ = (float)s->aabb_scale;
	for (int k = 0; k < 3; ++k) { m.aabb_min[k] = 0.5f - 0.5f * s_; m.aabb_max[k] = 0.5f + 0.5f * s_; }
	if (const Node* bx = snap->get("aabb")) { if (!vecN(bx->get("min"), m.aabb_min, 3) || !vecN(bx->get("max"), m.aabb_max, 3)) { set_err(err, err_len, "bad aabb"); return SNB_ERR_PARSE; } }
	std::memcpy(m.render_aabb_min, m.aabb_min, 12); std::memcpy(m.render_aabb_max, m.aabb_max, 12);
	if (const Node* bx = snap->get("render_aabb")) { if (!vecN(bx->get("min"), m.render_aabb_min, 3) || !vecN(bx->get("max"), m.render_aabb_max, 3)) { set_err(err, err_len, "bad render_aabb"); return SNB_ERR_PARSE; } }
	const float ident[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
	std::memcpy(m.render_aabb_to_local, ident, sizeof ident);
	if (const Node* r = snap->get("render_aabb_to_local")) { if (!mat(r, m.render_aabb_to_local, 3, 3)) { set_err(err, err_len, "bad render_aabb_to_local"); return SNB_ERR_PARSE; } }
	m.max_cascade = 0; while ((1u << m.max_cascade) < s->aabb_scale) ++m.max_cascade;
	m.cone_angle_constant = s->aabb_scale <= 1 ? 0.0f : 1.0f / 256.0f;
	m.rgb_activation = activation_of(nerf ? nerf->get("rgb_activation") : nullptr, SNB_ACT_LOGISTIC);
	m.density_activation = activation_of(nerf ? nerf->get("density_activation") : nullptr, SNB_ACT_EXPONENTIAL);
	m.min_transmittance = 0.01f;
	// ---- density grid (fp16; the bitfield is rebuilt from it: snb_density_to_bitfield) ----
	if (uval(snap->get("density_grid_size"), 128) != 128) { set_err(err, err_len, "density_grid_size != 128"); return SNB_ERR_UNSUPPORTED; }
	if (const Node* dg = snap->get("density_grid_binary")) {
		if (dg->kind != Node::BIN || dg->bin_len % 2) { set_err(err, err_len, "bad density_grid_binary"); return SNB_ERR_PARSE; }
		s->density = reinterpret_cast<const uint16_t*>(dg->bin); s->n_density = dg->bin_len / 2;
		const uint64_t cells = 128ull * 128 * 128;
		if (s->n_density != 0 && s->n_density != cells * (m.max_cascade + 1)) { set_err(err, err_len, "Incompatible number of grid cascades."); return SNB_ERR_PARSE; }   // testbed.cu:4933
	}
	if (const Node* cam = snap->get("camera")) s->has_camera = mat(cam->get("matrix"), s->camera, 4, 3);
	*out = s.release();
	return SNB_OK;
}

void snb_snapshot_free(snb_snapshot* s) { delete s; }

int snb_snapshot_get(const snb_snapshot* s, snb_model_desc* model, const uint16_t** density_fp16, uint64_t* n_density, uint32_t* aabb_scale, float* camera12) {
	if (!s) return SNB_ERR_ARG;
	if (model) *model = s->model;
	if (density_fp16) *density_fp16 = s->density;
	if (n_density) *n_density = s->n_density;
	if (aabb_scale) *aabb_scale = s->aabb_scale;
	if (camera12) { if (!s->has_camera) return SNB_ERR_STATE; std::memcpy(camera12, s->camera, sizeof s->camera); }
	return SNB_OK;
}

int snb_snapshot_save(const char* path, const snb_model_desc* m, const uint16_t* density_fp16, uint64_t n_density, uint32_t aabb_scale, const float* camera12, int compress) {
	if (!path || !m || !m->params) return SNB_ERR_ARG;
	if (density_fp16 == nullptr) n_density = 0;
	Writer w;
	auto mlp = [&](const char* name, unsigned hidden) {
		w.key(name); w.map(5);
		w.key("otype"); w.str("FullyFusedMLP"); w.key("activation"); w.str("ReLU"); w.key("output_activation"); w.str("None");
		w.key("n_neurons"); w.uint(64); w.key("n_hidden_layers"); w.uint(hidden);
	};
	// root: the network description Testbed::reset_network reads back (configs/nerf/base.json defaults) + the snapshot object
	w.map(7);
	mlp("network", 1); mlp("rgb_network", 2);
	w.key("dir_encoding"); w.map(2); w.key("otype"); w.str("Composite"); w.key("nested"); w.arr(2);
	w.map(3); w.key("n_dims_to_encode"); w.uint(3); w.key("otype"); w.str("SphericalHarmonics"); w.key("degree"); w.uint(4);
	w.map(1); w.key("otype"); w.str("Identity");
	w.key("loss"); w.map(1); w.key("otype"); w.str("Huber");
	w.key("optimizer"); w.map(3); w.key("otype"); w.str("Ema"); w.key("decay"); w.f32(0.95f); w.key("nested"); w.map(5);
	w.key("otype"); w.str("ExponentialDecay"); w.key("decay_start"); w.uint(20000); w.key("decay_interval"); w.uint(10000); w.key("decay_base"); w.f32(0.33f); w.key("nested"); w.map(6);
	w.key("otype"); w.str("Adam"); w.key("learning_rate"); w.f32(1e-2f); w.key("beta1"); w.f32(0.9f); w.key("beta2"); w.f32(0.99f); w.key("epsilon"); w.f32(1e-15f); w.key("l2_reg"); w.f32(1e-6f);
	w.key("encoding"); w.map(6);
	w.key("otype"); w.str("HashGrid");
	w.key("n_levels"); w.uint(m->n_levels);
	w.key("n_features_per_level"); w.uint(m->n_features_per_level);
	w.key("log2_hashmap_size"); w.uint(m->log2_hashmap_size);
	w.key("base_resolution"); w.uint(m->base_resolution);
	w.key("per_level_scale"); w.f32(m->per_level_scale);
	w.key("snapshot"); w.map(13 + (camera12 ? 1 : 0));
	w.key("version"); w.uint(1);
	w.key("training_step"); w.uint(0);      // read unconditionally by load_snapshot (testbed.cu:4972-4973)
	w.key("loss"); w.f32(0.0f);
	w.key("mode"); w.str("nerf");
	w.key("n_params"); w.uint(m->n_params);
	w.key("params_type"); w.str("__half");
	w.key("params_binary"); w.bin(m->params, (size_t)m->n_params * 2);
	w.key("density_grid_size"); w.uint(128);
	w.key("density_grid_binary"); w.bin(density_fp16, density_fp16 ? (size_t)n_density * 2 : 0);
	w.key("nerf"); w.map(4);
	w.key("rgb"); w.map(3);                  // training counters, read unconditionally (testbed.cu:4901-4903); Testbed defaults (testbed.h:509-512)
	w.key("rays_per_batch"); w.uint(1u << 12); w.key("measured_batch_size"); w.uint(0); w.key("measured_batch_size_before_compaction"); w.uint(0);
	w.key("aabb_scale"); w.uint(aabb_scale);
	static const char* act[] = {"None", "ReLU", "Logistic", "Exponential"};
	w.key("rgb_activation"); w.str(act[m->rgb_activation & 3]);
	w.key("density_activation"); w.str(act[m->density_activation & 3]);
	w.key("aabb"); w.map(2); w.key("min"); w.vec(m->aabb_min, 3); w.key("max"); w.vec(m->aabb_max, 3);
	w.key("render_aabb"); w.map(2); w.key("min"); w.vec(m->render_aabb_min, 3); w.key("max"); w.vec(m->render_aabb_max, 3);
	w.key("render_aabb_to_local"); w.arr(3);
	for (int r = 0; r < 3; ++r) { w.arr(3); for (int c = 0; c < 3; ++c) w.f32(m->render_aabb_to_local[c * 3 + r]); }
	if (camera12) {
		w.key("camera"); w.map(1); w.key("matrix"); w.arr(3);
		for (int r = 0; r < 3; ++r) { w.arr(4); for (int c = 0; c < 4; ++c) w.f32(camera12[c * 3 + r]); }
	}
	std::vector<uint8_t> file;
	if (compress) {
		uLongf cap = compressBound((uLong)w.out.size());
		file.resize(cap);
		if (compress2(file.data(), &cap, w.out.data(), (uLong)w.out.size(), 1) != Z_OK) return SNB_ERR_IO;   // zlib stream, as zstr::ofstream writes
		file.resize(cap);
	} else file.swap(w.out);
	FILE* f = std::fopen(path, "wb");
	if (!f) return SNB_ERR_IO;
	const bool ok = std::fwrite(file.data(), 1, file.size(), f) == file.size();
	return (std::fclose(f) == 0 && ok) ? SNB_OK : SNB_ERR_IO;
}

} // extern "C"
