// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h). C interface loaded by tests via ctypes.
#pragma once
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

// Activation codes follow ENerfActivation (include/neural-graphics-primitives/common.h): None, ReLU, Logistic, Exponential.
enum { ORC_ACT_NONE = 0, ORC_ACT_RELU = 1, ORC_ACT_LOGISTIC = 2, ORC_ACT_EXPONENTIAL = 3 };
// Render modes: integer values of ERenderMode (common.h:56-68).
enum { ORC_MODE_AO = 0, ORC_MODE_SHADE = 1, ORC_MODE_NORMALS = 2, ORC_MODE_POSITIONS = 3, ORC_MODE_DEPTH = 4, ORC_MODE_SHADOWDEPTH = 8 };

typedef struct {
	uint32_t n_levels, n_features_per_level, log2_hashmap_size, base_resolution;
	float per_level_scale;
	const uint16_t* params;     // fp16 bits: density MLP | rgb MLP | hash grid (nerf_network.h:356-371)
	uint64_t n_params;
	const uint8_t* bitfield;    // NERF_CASCADES * 128^3 / 8 bytes, Morton order; NULL = occupied everywhere
	float aabb_min[3], aabb_max[3];               // training aabb (m_aabb)
	float render_aabb_min[3], render_aabb_max[3];
	float render_aabb_to_local[9];                // column-major mat3
	uint32_t max_cascade;
	float cone_angle_constant;
	int32_t rgb_activation, density_activation;
	float min_transmittance;
} orc_model;

typedef struct {
	float camera[12];        // mat4x3 column-major: right, up(down), forward, origin
	float focal_length[2];   // pixels (calc_focal_length result)
	float screen_center[2];
	int32_t width, height;
	uint32_t spp;            // sample index (render_buffer.spp)
	int32_t snap_to_pixel_centers;
} orc_view;

// The remaining arguments of init_rays_with_payload_kernel_nerf (src/testbed_nerf.cu:1855-1880) for the stock path. Installed with
// orc_set_camera_ex (a checker-wide override, like orc_set_grid_scales): NULL = pinhole, camera1 == camera0, no envmap.
typedef struct {
	float camera1[12];
	float rolling_shutter[4];
	float parallax_shift[3];
	float near_distance, plane_z, aperture_size;
	int32_t lens_mode;       // ELensMode (common.h:118-126): 0 Perspective, 1 OpenCV, 2 FTheta, 3 LatLong, 4 OpenCVFisheye, 5 Equirectangular
	float lens_params[7];
} orc_camera_ex;

typedef struct {   // ObjectTransform (include/synerfgine/common.cuh:137-147) with host arrays
	const float* nodes;   // n_nodes * 8 floats: bb.min[3], bb.max[3], left_idx(int), right_idx(int)
	const float* tris;    // n_tris * 9 floats
	int32_t n_nodes, n_tris;
	float rot[9];         // column-major
	float pos[3];
	float scale;
	int32_t mat_id;
} orc_object;

typedef struct {   // sng::Material (material.cuh:125-134)
	float ka[3], kd[3], ks[3];
	float n, rg;
	int32_t type;        // 0 lambertian, 1 glossy
	float spec_angle;
} orc_material;

typedef struct {   // sng::Light (light.cuh:79-91)
	float pos[3];
	int32_t type;        // 0 point, 1 directional
	float intensity, size;
} orc_light;

typedef struct {   // RayTracer settings (raytracer.cuh:156-167) + Engine shadow settings (engine.cuh:112-120)
	uint32_t n_steps;            // m_n_steps (8)
	uint32_t samples;            // m_samples (light_samples)
	uint32_t bounces;            // m_ray_iters (path_trace_depth)
	uint32_t shadow_iters;       // m_shadow_iters (syn_shadow_samples)
	float lens_angle_constant;   // lens_size
	float depth_offset;
	float syn_shadow_factor;     // syn_shadow_intensity
	int32_t view_nerf_shadow;    // shadow_on_virtual_obj
	int32_t buffer_type;         // ImgBufferType, 0 = Final
	int32_t nerf_shadow_kernel;  // sng_position_kernel_size = rendering.nerf_shadow_samples
	float nerf_shadow_intensity;
	float nerf_on_nerf_shadow_threshold;
	int32_t show_shadow_on_nerf; // shadow_on_nerf
	int32_t show_nerf;
	int32_t syn_px_scale;
	float exposure;
	int32_t tonemap_curve;       // ETonemapCurve: 0 Identity, 1 ACES, 2 Hable, 3 Reinhard
} orc_rt_params;

void orc_set_math_mode(int mode);   // 0 = libm (reference host behaviour), 1 = deterministic (GPU-comparable)

#ifdef __cplusplus
}
#endif
