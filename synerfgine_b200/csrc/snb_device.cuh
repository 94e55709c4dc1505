// synerfgine_b200 — device-side views of the model, camera and virtual scene (plain structs passed by value
// or through constant/global memory). Layout in HBM is described in DESIGN.md §"Data layout".
#pragma once
#include "snb_march.cuh"

namespace snb {

constexpr int kMaxLevels = 32;
constexpr int kFeat = 32;          // encoded width = n_levels * n_features_per_level (nerf_network.h:82, alignment 16)
constexpr int kHidden = 64;        // base.json network width
constexpr int kMlpHalves = 64 * 32 + 16 * 64 + 64 * 32 + 64 * 64 + 16 * 64; // 10240: dW1 dW2 rW1 rW2 rW3

// Offsets (in halves) of each weight matrix inside the params blob (row-major [out x in],
// reference: dependencies/tiny-cuda-nn/src/fully_fused_mlp.cu:659-671; order nerf_network.h:356-371).
constexpr int kOffDW1 = 0;
constexpr int kOffDW2 = kOffDW1 + 64 * 32;
constexpr int kOffRW1 = kOffDW2 + 16 * 64;
constexpr int kOffRW2 = kOffRW1 + 64 * 32;
constexpr int kOffRW3 = kOffRW2 + 64 * 64;

struct GridDesc {
	uint32_t n_levels, n_feat;
	float scale[kMaxLevels];
	uint32_t res[kMaxLevels];
	uint32_t offset[kMaxLevels + 1];   // in grid entries (each n_feat halves)
};

struct DevModel {
	GridDesc grid;
	const __half* grid_params;   // hash-table features, levels concatenated
	const __half* mlp_rowmajor;  // the 10240 MLP halves as they sit in the snapshot
	const uint32_t* mlp_frag;    // same weights re-ordered into mma.m16n8k16 B-fragment order (snb_network.cuh)
	const __half* mlp_umma;      // same weights in the tcgen05 K-major core-matrix layout (snb_mlp_tc.cuh)
	Aabb train_aabb;
	int rgb_act, density_act;
	float min_transmittance;
	MarchEnv env;
};

// The rest of Testbed::render_nerf's camera model (uv_to_ray, include/neural-graphics-primitives/common_device.cuh:403-470, and the
// arguments of init_rays_with_payload_kernel_nerf, src/testbed_nerf.cu:1855-1970): end-of-shutter camera + rolling shutter (motion blur),
// parallax shift, near plane, thin-lens aperture focused at plane_z, lens models, environment map. enabled == 0: the plain pinhole path.
enum { kLensPerspective = 0, kLensOpenCV = 1, kLensFTheta = 2, kLensLatLong = 3, kLensOpenCVFisheye = 4, kLensEquirectangular = 5 };
struct CameraEx {
	int enabled;
	m4x3 cam1;
	float rolling[4];
	v3 parallax;
	float near_distance, plane_z, aperture;
	int lens_mode;
	float lens[7];
	const float4* envmap; int env_w, env_h;
};

struct DevView {
	m4x3 cam;
	v2 focal, center;
	int width, height;           // full view resolution
	int tx0, ty0, tw, th;        // tile rendered by this context
	uint32_t spp;
	int snap;
	v2 pix_off;                  // ld_pixel_offset(snap ? 0 : spp), evaluated once on the host
	CameraEx ex;
};

struct DevObject {               // sng::ObjectTransform with precomputed world->object map
	const float* nodes;          // 8 floats per node
	const float* tris;           // 9 floats per triangle
	int n_nodes, n_tris;
	m3 rot;
	m3 inv_sr;                   // (I/scale) * inverse(rot), evaluated once per update (reference: per ray, triangle_bvh.cu:313-319)
	v3 pos;
	float scale;
	int mat_id;
};

struct DevMaterial { v3 ka, kd, ks; float n, rg; int type; float spec_angle; };
struct DevLight { v3 pos; int type; float intensity, size; };

struct DevScene {
	const DevObject* objects; uint32_t n_objects;
	const DevMaterial* materials; uint32_t n_materials;
	const DevLight* lights; uint32_t n_lights;
};

struct RayRec {                  // 64 B: one queue entry produced by ray generation
	float ox, oy, oz, t;
	float dx, dy, dz;
	uint32_t idx;                // global pixel index x + width * y
	uint32_t sh[8];              // SH-4 encoding of the direction (16 halves): computed once per ray where the ray is made, so
	                             // the march's refill path is a 32-byte copy instead of ~150 instructions in its hot loop
};

} // namespace snb
