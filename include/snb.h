/* snb.h — C ABI of synerfgine_b200 (libsnb.so): the SyNeRFgine render hot path on B200 (sm_100a).
 *
 * Drop-in boundary. Each entry point names the reference interface it replaces; paths are
 * relative to the reference checkout (trxe/SyNeRFgine @ b842828).
 *
 * Conventions: every function returns 0 on success and a negative snb_status otherwise; the message is
 * available from snb_last_error(). No exceptions cross this boundary. Pointers named *_dev are CUDA device
 * pointers on the context's device, *_host are host pointers; nothing else is dereferenced. All work is
 * enqueued on the context's stream; calls that hand results to the host synchronise that stream, all others
 * are asynchronous until snb_sync(). One context per host thread. There is no CPU fallback: without a CUDA
 * device snb_create fails.
 */
#ifndef SNB_H
#define SNB_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct snb_ctx snb_ctx;

enum snb_status { SNB_OK = 0, SNB_ERR_ARG = -1, SNB_ERR_CUDA = -2, SNB_ERR_STATE = -3, SNB_ERR_IO = -4, SNB_ERR_PARSE = -5, SNB_ERR_UNSUPPORTED = -6 };

/* ENerfActivation, include/neural-graphics-primitives/common.h */
enum { SNB_ACT_NONE = 0, SNB_ACT_RELU = 1, SNB_ACT_LOGISTIC = 2, SNB_ACT_EXPONENTIAL = 3 };
/* ERenderMode, include/neural-graphics-primitives/common.h:56-68 */
enum { SNB_MODE_AO = 0, SNB_MODE_SHADE = 1, SNB_MODE_NORMALS = 2, SNB_MODE_POSITIONS = 3, SNB_MODE_DEPTH = 4, SNB_MODE_SHADOWDEPTH = 8 };
/* ETonemapCurve */
enum { SNB_TONEMAP_IDENTITY = 0, SNB_TONEMAP_ACES = 1, SNB_TONEMAP_HABLE = 2, SNB_TONEMAP_REINHARD = 3 };
/* sng::ImgBufferType, include/synerfgine/raytracer.cuh:20-31 */
enum { SNB_BUF_FINAL = 0, SNB_BUF_NEXT_ORIGIN = 1, SNB_BUF_SRC_ORIGIN = 2, SNB_BUF_NEXT_DIRECTION = 3, SNB_BUF_SRC_DIRECTION = 4,
       SNB_BUF_NORMAL = 5, SNB_BUF_DEPTH = 6, SNB_BUF_NERF_SHADOW = 7 };

/* A NeRF model: what Testbed holds after load_snapshot (src/testbed.cu:4878-4992) — NerfNetwork parameters
 * (include/neural-graphics-primitives/nerf_network.h:356-371 order: density MLP, rgb MLP, hash grid; fp16),
 * the occupancy bitfield (src/testbed_nerf.cu:3212-3229), boxes and activations. Host pointers; copied. */
typedef struct {
	uint32_t n_levels, n_features_per_level, log2_hashmap_size, base_resolution;
	float per_level_scale;
	const uint16_t* params;      /* fp16 bits */
	uint64_t n_params;
	const uint8_t* bitfield;     /* 8 * 128^3 / 8 bytes, Morton order; NULL = occupied everywhere */
	float aabb_min[3], aabb_max[3];
	float render_aabb_min[3], render_aabb_max[3];
	float render_aabb_to_local[9];   /* column-major mat3 */
	uint32_t max_cascade;
	float cone_angle_constant;
	int32_t rgb_activation, density_activation;
	float min_transmittance;
} snb_model_desc;

/* Camera + image of one render call: CudaRenderBufferView (render_buffer.h:169-178) + the camera arguments
 * of render_nerf (testbed.h:358-371). */
typedef struct {
	float camera[12];            /* mat4x3, column-major */
	float focal_length[2];       /* pixels: calc_focal_length (src/testbed.cu:4133-4135) */
	float screen_center[2];
	int32_t width, height;
	uint32_t spp;                /* sample index of the accumulation */
	int32_t snap_to_pixel_centers;
} snb_view;

/* The rest of the stock render's camera model: the arguments Testbed::render_nerf passes to NerfTracer::init_rays_from_camera beyond the
 * pinhole camera (src/testbed_nerf.cu:2679-2740 -> init_rays_with_payload_kernel_nerf :1855-1970 -> uv_to_ray,
 * include/neural-graphics-primitives/common_device.cuh:403-470). Applies to snb_render_nerf (and the snb_march_trace tap) only: the
 * sng path's ray generation takes none of them (src/synerfgine/raytracer.cu:59-98). Foveation, hidden-area mask and the learned
 * distortion grid are VR / training features and stay out (SURVEY.md 8 "out of scope"). */
enum { SNB_LENS_PERSPECTIVE = 0, SNB_LENS_OPENCV = 1, SNB_LENS_FTHETA = 2, SNB_LENS_LATLONG = 3, SNB_LENS_OPENCV_FISHEYE = 4, SNB_LENS_EQUIRECTANGULAR = 5 }; /* ELensMode, common.h:118-126 */
typedef struct {
	float camera1[12];           /* camera at the end of the shutter interval (column-major 4x3); snb_view.camera is camera0 */
	float rolling_shutter[4];    /* pixel time = a + b*u + c*v + d*motionblur_time; the windowless render passes {0,0,0,1} */
	float parallax_shift[3];     /* m_parallax_shift */
	float near_distance;         /* m_render_near_distance */
	float plane_z;               /* m_slice_plane_z + m_scale: focus distance; negative = slice plane (rays die, depth = -plane_z) */
	float aperture_size;         /* m_aperture_size: 0 = pinhole */
	int32_t lens_mode;           /* SNB_LENS_* (m_nerf.render_lens.mode when render_with_lens_distortion) */
	float lens_params[7];        /* Lens::params */
} snb_camera_ex;

/* Pixel tile of a view rendered by this context (multi-GPU sharding). Pixel indices used for
 * sampling stay global. A zero-sized tile means the whole view. */
typedef struct { int32_t x0, y0, w, h; } snb_tile;

typedef struct {                 /* sng::ObjectTransform (include/synerfgine/common.cuh:137-147), host arrays */
	const float* nodes;          /* n_nodes * 8 floats: bb.min, bb.max, left_idx, right_idx (ints) */
	const float* tris;           /* n_tris * 9 floats */
	int32_t n_nodes, n_tris;
	float rot[9];                /* column-major */
	float pos[3];
	float scale;
	int32_t mat_id;
} snb_object;

typedef struct {                 /* sng::Material (include/synerfgine/material.cuh:125-134) */
	float ka[3], kd[3], ks[3];
	float n, rg;
	int32_t type;                /* 0 lambertian, 1 glossy */
	float spec_angle;
} snb_material;

typedef struct {                 /* sng::Light (include/synerfgine/light.cuh:79-91) */
	float pos[3];
	int32_t type;                /* 0 point, 1 directional */
	float intensity, size;
} snb_light;

/* RayTracer members (raytracer.cuh:156-167) + Engine shadow settings (engine.cuh:112-120) = "rendering" JSON. */
typedef struct {
	uint32_t n_steps;            /* shadow DDA steps (8) */
	uint32_t samples;            /* light_samples */
	uint32_t bounces;            /* path_trace_depth */
	uint32_t shadow_iters;       /* syn_shadow_samples */
	float lens_angle_constant;   /* lens_size */
	float depth_offset;
	float syn_shadow_factor;     /* syn_shadow_intensity */
	int32_t view_nerf_shadow;    /* shadow_on_virtual_obj */
	int32_t buffer_type;
	int32_t nerf_shadow_kernel;  /* nerf_shadow_samples */
	float nerf_shadow_intensity;
	float nerf_on_nerf_shadow_threshold;
	int32_t show_shadow_on_nerf; /* shadow_on_nerf */
	int32_t show_nerf;
	int32_t syn_px_scale;
	float exposure;
	int32_t tonemap_curve;
} snb_rt_params;

/* ---- present stage ---------------------------------------------------------------------
 * Replaces the GL blit of Display::present / transfer_texture (src/synerfgine/display.cu:153-264,266-303) with the FXAA
 * fragment shader scripts/virtual_desc/main.frag:51-117, and Display::save_image's glReadPixels + stbi_write_png
 * (display.cu:305-322): final float RGBA frame (row 0 = top) -> RGB8 window image (row 0 = top), headless.
 * Foveation warp parameters are main.frag's FoveationWarp uniforms (display.cu:199-218); NULL = identity. */
typedef struct snb_fov_warp {
	float al, bl, cl, am, bm, ar, br, cr, switch_left, switch_right, inv_switch_left, inv_switch_right;
} snb_fov_warp;
int snb_present(snb_ctx* ctx, const float* rgba_dev, int in_w, int in_h, int out_w, int out_h, const float* clear_rgb /* 3 floats or NULL */,
                const snb_fov_warp* warp_x, const snb_fov_warp* warp_y, uint8_t* rgb8_dev /* out_w*out_h*3 */);
int snb_present_host(snb_ctx* ctx, const float* rgba_dev, int in_w, int in_h, int out_w, int out_h, const float* clear_rgb,
                     const snb_fov_warp* warp_x, const snb_fov_warp* warp_y, uint8_t* rgb8_host);
/* 8-bit RGB PNG of a host image (rows top-down), as Display::save_image writes `output-%03d.png` (display.cu:318-319). Pure host code. */
int snb_write_png(const char* path, const uint8_t* rgb8_host, int width, int height);

/* ---- on-disk model format ------------------------------------------------------------------
 * instant-ngp / SyNeRFgine snapshots: `.ingp` (zlib- or gzip-compressed msgpack) and `.msgpack`, as written by
 * Testbed::save_snapshot and read by Testbed::load_snapshot (src/testbed.cu:4812-4992) with tcnn Trainer::serialize
 * (tiny-cuda-nn/trainer.h:274-312). Pure host code. The model descriptor returned by snb_snapshot_get points INTO the
 * snapshot object (params) and has bitfield == NULL: build it from the fp16 density grid with snb_density_to_bitfield
 * (what load_snapshot does through update_density_grid_mean_and_bitfield, testbed.cu:4922-4931). */
typedef struct snb_snapshot snb_snapshot;
int snb_snapshot_load(const char* path, snb_snapshot** out, char* err, size_t err_len);
void snb_snapshot_free(snb_snapshot* s);
int snb_snapshot_get(const snb_snapshot* s, snb_model_desc* model, const uint16_t** density_fp16, uint64_t* n_density, uint32_t* aabb_scale,
                     float* camera12 /* column-major mat4x3 or NULL */);
int snb_snapshot_save(const char* path, const snb_model_desc* model, const uint16_t* density_fp16, uint64_t n_density, uint32_t aabb_scale,
                      const float* camera12 /* or NULL */, int compress);

/* ---- lifetime ----------------------------------------------------------------------------------------- */
int snb_create(int device, void* cuda_stream /* cudaStream_t or NULL */, snb_ctx** out);
void snb_destroy(snb_ctx* ctx);
const char* snb_last_error(const snb_ctx* ctx);
int snb_sync(snb_ctx* ctx);
const char* snb_version(void);
/* Options. Semantics (these change what is computed; the defaults are the reference's behaviour):
 *   "alt_schedule" = "auto" (default) | "external" | "fixed" | "lockstep"   batch lengths of the sng compositor, see snb_get_alt_status below
 *                    (auto: a still frame re-uses the verified sequence of the last one; a frame whose camera or tile changed marches
 *                    the batches shorter than 8 in lock step, one per launch, from parked ray states: exact in one pass. lockstep: always.
 *                    "alt_auto_lockstep" = "0" keeps auto on verified speculation only)
 *   "alt_batch"    = "1".."8"        batch length of alt_schedule = fixed
 *   "dup_reuse"    = "1" (default) | "0"   sng path: the compositor's rewind (composite_kernel_nerf_alt, src/testbed_nerf.cu:574) makes the first
 *                    sample of every batch the last sample of the previous one again — the reference gathers and evaluates it a second time
 *                    at a position that differs by the rounding of the rewind. 1: its network output is re-used (position, dt, depth and
 *                    termination are computed as always): a quarter fewer network evaluations on a 1080p frame, results within the
 *                    frame tolerance (>= 50 dB, counted pixels over 1e-2); 0: every sample goes through the network (exact counts)
 *   "defer_srgb"   = "1" (default) | "0"   sng path, lane march: extract_from_payload's srgb_to_linear (src/testbed_nerf.cu:1614-1640) runs as one pass
 *                    over the tile after the march (1) or per finished ray inside the march loop (0); bit-identical frames either way
 *   "alt_target_queries" = "2097152" target_n_queries of NerfTracer::trace_alt (src/testbed_nerf.cu:2188); tests lower it to exercise short batches on small frames
 * Implementation selectors (A/B measurements; frames are bit-identical unless noted):
 *   "march"      = "lane" (default: one ray per lane, MLPs on warp-level mma.sync) | "ws" (warp-specialised pipeline, MLPs on tcgen05 tensor
 *                  cores with TMEM accumulators; measured slower on C2); same samples, outputs equal within fp32 summation order
 *   "mlp"        = "mma" | "tc"   evaluator of the snb_network_inference parity tap (outputs equal within fp16 rounding)
 *   "overlap"    = "1" | "0"      mesh layer on a second stream
 *   "dead_terms" = "1" | "0"      skip NeRF shadow terms that no written pixel can observe (identical frame and RNG states)
 *   "mesh_priority" = "high" (default) | "default" | "low"   CUDA priority of the mesh layer's stream (set before the first frame)
 *   "lane_warps" = "0" (default: by tile size) | "10" | "12"   CTA shape of the lane march
 *   "sng_block"  = "128" (default) | "64" | "32" | "0"   k_raytrace threads per block, a warp = 8 x 4 pixel patch; 0 = 32 pixels of a row
 *   "terms_block" = "0" (default) | "32" | "64" | "128"   the same for k_nerf_shadow_terms
 * Environment SNB_<OPTION> (upper case) sets the default of an option. */
int snb_set_option(snb_ctx* ctx, const char* key, const char* value);

/* Batch schedule of the sng compositor. NerfTracer::trace_alt marches all alive rays in lock step, S_j = clamp(2^21 / n_alive_j, 1, 8)
 * samples in iteration j (src/testbed_nerf.cu:2189-2190), and composite_kernel_nerf_alt rewinds payload.t after every iteration (:574):
 * the last sample of a batch is composited again as the first of the next, so the image depends on S_0, S_1, ... Rays run
 * independently here; the sequence is an input of the march and the march counts n_alive_j.
 *   auto:     the sequence lives on the device from frame to frame; a pass whose counts contradict the sequence it used is repeated
 *             inside the same snb_frame / snb_render_nerf_with_buffers call (at most 3 passes, conditional launches, no host sync).
 *             Exact once `converged` is 1 (always, unless a cold start needs more than 3 passes or more than 64 batches are short).
 *   external: one pass per call with the sequence given by snb_set_alt_schedule; the caller sums the counts (over GPUs when a frame
 *             is tiled — n_alive_j is a whole-frame quantity) and sets the sequence they imply.
 *   fixed:    S_j = alt_batch for all j: a deliberate deviation (= the reference for frames with <= 262144 alive rays when 8).
 * out132: [0] passes of the last call, [1] converged, [2] first batch whose length was wrong (0xFFFFFFFF none), [3] target_n_queries,
 * [4..67] n_alive_j of the last pass, [68..131] batch lengths the next pass will use (0 = unknown -> S_0). Synchronises. */
int snb_get_alt_status(snb_ctx* ctx, uint32_t* out132_host);
int snb_set_alt_schedule(snb_ctx* ctx, const uint8_t* steps_host, uint32_t n /* <= 64; later batches are 8; n = 0 forgets the sequence */);

/* ---- inputs ------------------------------------------------------------------------------------------- */
/* ≙ Testbed::load_snapshot result handed over as arrays (src/testbed.cu:4878-4992). */
int snb_set_model(snb_ctx* ctx, const snb_model_desc* model);
/* The same with model->params and model->bitfield in DEVICE memory of the context's GPU — the state the reference holds after training or
 * load_snapshot (tcnn's params buffer, Testbed::Nerf::density_grid_bitfield; SURVEY.md 8b "no hidden host copies"): hash table and bitfield
 * are copied device to device, only the 20 KB of MLP weights are read back to be laid out in fragment order. */
int snb_set_model_dev(snb_ctx* ctx, const snb_model_desc* model);
/* ≙ Engine::set_virtual_world after parsing (src/synerfgine/engine.cu:21-78): uploads world, materials, lights. */
int snb_set_scene(snb_ctx* ctx, const snb_object* objects, uint32_t n_objects, const snb_material* materials, uint32_t n_materials,
                  const snb_light* lights, uint32_t n_lights);
/* ≙ Engine::update_world_objects (engine.cu:80-127): per-frame transforms / lights without re-uploading meshes. */
int snb_update_scene(snb_ctx* ctx, const snb_object* objects, uint32_t n_objects, const snb_material* materials, uint32_t n_materials,
                     const snb_light* lights, uint32_t n_lights);
/* ≙ sng::init_rand_state (src/synerfgine/common.cu:22-26) for both state arrays; seed 1999 in the reference. */
int snb_seed_rng(snb_ctx* ctx, uint64_t seed);
/* Clears the ray tracer's temporal blend buffer (≙ RayTracer::m_reset_color_buffer, raytracer.cu:327-336). */
int snb_reset_accumulation(snb_ctx* ctx);

/* ---- the hot path ------------------------------------------------------------------------------------- */
/* ≙ Testbed::render_nerf (testbed.h:358-371; src/testbed_nerf.cu:2679-2837): stock NeRF frame into fb/depth. */
int snb_render_nerf(snb_ctx* ctx, const snb_view* view, const snb_tile* tile, int render_mode, float* fb_rgba_dev, float* depth_dev);
/* Extended camera of the following snb_render_nerf calls (NULL: back to the pinhole path, bit-identical to never having called this). */
int snb_set_camera_ex(snb_ctx* ctx, const snb_camera_ex* ex);
/* ≙ Testbed::m_envmap's inference view (include/neural-graphics-primitives/envmap.cuh:24-50): w x h RGBA fp32 host texels, row-major;
 * background of every valid primary ray of snb_render_nerf, composited under the NeRF colour. NULL or w*h == 0 removes it. */
int snb_set_envmap(snb_ctx* ctx, const float* rgba_host, int32_t w, int32_t h);
/* ≙ RayTracer::render (raytracer.cuh:116-128; src/synerfgine/raytracer.cu:312-370): mesh layer into the
 * context's syn buffers; optional copies out. */
int snb_render_syn(snb_ctx* ctx, const snb_view* view, const snb_tile* tile, const snb_rt_params* rt, float* syn_rgba_dev, float* syn_depth_dev);
/* ≙ Testbed::render sng overload → render_nerf_with_buffers + shade_nerf_shadows (testbed.h:323-357,401-418;
 * src/testbed.cu:4353-4404; src/testbed_nerf.cu:2467-2677). positions/normals may be NULL (context scratch). */
int snb_render_nerf_with_buffers(snb_ctx* ctx, const snb_view* view, const snb_tile* tile, const snb_rt_params* rt, int render_mode,
                                 float* fb_rgba_dev, float* depth_dev, float* positions_dev, float* normals_dev);
/* ≙ RayTracer::overlay (raytracer.cuh:130-131; raytracer.cu:372-392). */
int snb_overlay(snb_ctx* ctx, int syn_w, int syn_h, const snb_rt_params* rt, const float* syn_rgba_dev, const float* syn_depth_dev,
                const float* nerf_rgba_dev, const float* nerf_depth_dev, float* out_rgba_dev, float* out_depth_dev);
/* ≙ render_frame_epilogue → CudaRenderBuffer::accumulate + tonemap (src/testbed.cu:4550-4627; render_buffer.cu:671-735). */
int snb_accumulate_tonemap(snb_ctx* ctx, int n_pixels, const float* frame_rgba_dev, float* accumulate_rgba_dev, uint32_t spp, int color_space_srgb,
                           int output_srgb, float exposure, const float* background_rgba4, int tonemap_curve, float* out_rgba_dev);
/* ≙ Engine::frame minus GL (src/synerfgine/engine.cu:352-433): syn layer, NeRF layer with both shadow
 * directions, overlay. Output on the device … */
int snb_frame(snb_ctx* ctx, const snb_view* view, const snb_tile* tile, const snb_rt_params* rt, float* out_rgba_dev, float* out_depth_dev);
/* … or copied into host buffers (tile-sized; what a host-language caller over cgo/JNI/ctypes uses). */
int snb_frame_host(snb_ctx* ctx, const snb_view* view, const snb_tile* tile, const snb_rt_params* rt, float* out_rgba_host, float* out_depth_host);
/* The same for callers that stream frames (the reference's main loop presents frame after frame): returns once the frame and its
 * device-to-host copy are enqueued; the copy runs on its own stream while the next frame renders. slot = 0 | 1 names the double buffer
 * (device and host side); read a slot's host buffers after snb_frame_host_wait(ctx, slot), reuse them only after it. The host buffers
 * should be pinned, else the copy is staged synchronously. */
int snb_frame_host_async(snb_ctx* ctx, const snb_view* view, const snb_tile* tile, const snb_rt_params* rt, float* out_rgba_host, float* out_depth_host, int32_t slot);
int snb_frame_host_wait(snb_ctx* ctx, int32_t slot);

/* ---- parity taps (the reference's debug views: positions, depth, hit indices — SURVEY.md §4) -------------- */
/* Per-ray march sample sequence: t, dt, cell = mip << 24 | morton (0xFFFFFFFF padding), n per ray. */
int snb_march_trace(snb_ctx* ctx, const snb_view* view, uint32_t max_samples, float* t_dev, float* dt_dev, uint32_t* cell_dev, uint32_t* n_dev,
                    float* origin_dir_dev /* n x 6 or NULL */);
/* The same sequence as recorded by the FUSED MARCH KERNEL itself (not a separate trace kernel): after this call snb_render_nerf (stock
 * path, march=lane) writes, per tile pixel, t and cell of every sample it evaluates (first `cap` per ray; n x cap arrays) and the number
 * of samples (n array, must be zeroed by the caller). NULL arrays switch the tap off. */
int snb_set_sample_tap(snb_ctx* ctx, float* t_dev, uint32_t* cell_dev, uint32_t* n_dev, uint32_t cap);
/* ≙ NerfNetwork::inference_mixed_precision (nerf_network.h:105-139): coords n x 7 fp32 → n x 4 fp16 (rgb, sigma);
 * optional features n x 32 fp16 and hash indices n x (n_levels * 8). */
int snb_network_inference(snb_ctx* ctx, const float* coords_dev, uint64_t n, uint16_t* out_dev, uint16_t* features_dev, uint32_t* hash_idx_dev);
/* ≙ depth_test_world closest hit (src/synerfgine/common.cu:36-67): rays n x 6 → object id, triangle id, t. */
int snb_trace_rays(snb_ctx* ctx, const float* rays_dev, uint32_t n, int32_t* obj_id_dev, int32_t* tri_id_dev, float* t_dev);
/* ≙ sng::init_rand_state + curand_uniform (src/synerfgine/common.cu:22-34): XORWOW state words {v0..v4, d} of subsequences 0..n-1
 * as the render kernels hold them, and the first n_uniform uniforms each would draw. Synchronises. */
int snb_rng_tap(snb_ctx* ctx, uint64_t seed, uint32_t n, uint32_t n_uniform, uint32_t* states6_dev, float* uniforms_dev);
/* Per-level grid scales of the current model as the device evaluates tcnn's grid_scale (exp2f = MUFU.EX2, one FFMA; common_device.h:841-845).
 * Returns the number of levels written or a negative status. */
int snb_get_grid_scales(snb_ctx* ctx, float* out_host, uint32_t cap);
/* Counters of the last NeRF render: [0] network samples evaluated, [1] rays launched, [2] rays that hit, [3] samples composited from a
 * re-used network output (option dup_reuse; [0] + [3] = samples composited = what the reference evaluates). */
int snb_get_counters(snb_ctx* ctx, uint64_t* out4_host);
/* Scheduling statistics of the last NeRF march (instrumentation; SURVEY.md 8d "device counters"): [0] pool iterations summed
 * over CTAs, [1] empty-space DDA steps, [2] search passes (warp level), [3] CTAs of the persistent grid. */
int snb_get_march_stats(snb_ctx* ctx, uint64_t* out4_host);
/* Pipeline balance of the last "ws" march (SM cycles summed over warps): [0] producer warps' lifetime, [1] waiting for a free operand stage,
 * [2] in the group barrier, [3] empty-space search, [4] sample set-up + hash-grid gather + publish, [5] tiles published (per producer warp,
 * summed), [6] epilogue warps' lifetime, [7] epilogue warps waiting for the tensor core. */
int snb_get_ws_profile(snb_ctx* ctx, uint64_t* out8_host);
/* Device time (ms, CUDA events on the context's stream) of the stages of the last frame: [0] NeRF ray generation,
 * [1] fused NeRF march kernel, [2] normals + mesh shadows on the NeRF, [3] mesh layer, [4] overlay. Synchronises. */
int snb_get_stage_ms(snb_ctx* ctx, float* out5_host);
/* Number of kernels this context has launched so far. */
uint64_t snb_launch_count(const snb_ctx* ctx);

/* ---- host-side scene layer (C++ in the reference: Engine / VirtualObject / CamPath / TriangleBvh::build) ---- */
typedef struct snb_scene snb_scene;
/* ≙ Engine::set_virtual_world + Engine::init "rendering" block (engine.cu:21-78,129-234). obj paths resolve against base_dir. */
int snb_scene_load_json(const char* json_text, const char* base_dir, snb_scene** out, char* err, size_t err_len);
void snb_scene_free(snb_scene* s);
int snb_scene_counts(const snb_scene* s, uint32_t* n_objects, uint32_t* n_materials, uint32_t* n_lights);
int snb_scene_get(const snb_scene* s, snb_object* objects, snb_material* materials, snb_light* lights, snb_rt_params* rt);
/* camera block: view dir, look-at, zoom, vo_scale + keyframe path; frame index advances CamPath (cam_path.cuh:116-142). */
int snb_scene_camera(const snb_scene* s, float* view3, float* at3, float* zoom, float* vo_scale, uint32_t* n_keyframes);
int snb_scene_camera_at(const snb_scene* s, uint32_t frame, uint32_t total_frames, float* view3, float* at3, float* zoom);
/* ≙ Engine::resize (engine.cu:236-255): resolutions of the NeRF layer and the mesh layer for a window, and syn_px_scale, from the scene's
 * "res_factor" and "vo_scale". */
int snb_scene_resolution(const snb_scene* s, int window_w, int window_h, int* nerf_w, int* nerf_h, int* syn_w, int* syn_h, int* syn_px_scale);
/* CamPath / animation switches: frames of the path, "move_on_start", "animation_speed" > 0, "res_factor". */
int snb_scene_path_info(const snb_scene* s, uint32_t* total_frames, int32_t* playing, int32_t* animations, int32_t* res_factor);
/* ≙ Testbed::reset_camera + set_view_dir + set_look_at + set_scale (src/testbed.cu:405-425,470-492) as Engine::init
 * applies the JSON camera block (engine.cu:148-151): camera matrix (mat4x3, column-major) out. */
int snb_camera_from_view(const float* view3, const float* at3, float zoom, float* out12);
/* ≙ camera_log_lerp (include/neural-graphics-primitives/common_device.cuh:658-660): exp(log(b * a^-1) * t) * a, the end-of-shutter
 * camera of a motion-blur sample (src/python_api.cu:157-189). Matrices are mat4x3 column-major. */
int snb_camera_log_lerp(const float* a12, const float* b12, float t, float* out12);
/* ≙ VirtualObject::next_frame / Light::next_frame (virtual_object.cuh:53-64; light.cuh:40-49). */
int snb_scene_advance(snb_scene* s);
/* ≙ TriangleBvhWithBranchingFactor<2>::build (src/triangle_bvh.cu:615-718); reorders tris in place. */
int snb_bvh_build(float* tris, int32_t n_tris, int32_t prims_per_leaf, float* nodes_out, int32_t nodes_cap, int32_t* n_nodes_out);
/* The same product built on the DEVICE: a linear BVH (Morton sort + radix tree over clusters of prims_per_leaf triangles +
 * bottom-up bounds) in the node format above, for arrays the caller owns. tris_dev n x 9 -> tris_out_dev (Morton order, must not alias),
 * nodes_out_dev (2 * ceil(n / prims_per_leaf) - 1 nodes of 8 floats), order_out_dev (source index of each output triangle, or NULL).
 * max_depth_out (or NULL; giving it synchronises) = depth of the deepest leaf: the traversal stack holds 32 entries. Triangle ids differ
 * from the host builder's, closest distances do not. snb_set_option(ctx, "bvh", "lbvh") makes snb_set_scene build this way. */
int snb_bvh_build_dev(snb_ctx* ctx, const float* tris_dev, int32_t n_tris, int32_t prims_per_leaf, float* tris_out_dev, float* nodes_out_dev, int32_t nodes_cap,
                      int32_t* order_out_dev, int32_t* n_nodes_out, int32_t* max_depth_out);
/* Bounds of every node recomputed bottom-up from the triangles as they are now (vertex animation); topology and ids stay. Works on any
 * tree in the node format, host-built ones included. */
int snb_bvh_refit_dev(snb_ctx* ctx, const float* tris_dev, float* nodes_dev, int32_t n_nodes);
/* Allocation-free variant of snb_bvh_refit_dev. NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). Same inputs, outputs and result as
 * snb_bvh_refit_dev, but the scratch (8 B per node) lives in a fixed device array inside the library (2 MiB per GPU, held until exit), so the
 * call allocates and frees nothing. n_nodes > SNB_REFIT_NOALLOC_MAX_NODES is rejected with SNB_ERR_ARG and a message. All calls on a GPU
 * share that one array: calls whose kernels overlap on the GPU (different streams / contexts) race; serialise them. */
#define SNB_REFIT_NOALLOC_MAX_NODES 262144
int snb_bvh_refit_dev_noalloc(snb_ctx* ctx, const float* tris_dev, float* nodes_dev, int32_t n_nodes);
/* Deforming mesh of the current scene: new vertices of object `index` (n_tris x 9 floats, in the order snb_set_scene received them), then refit. */
int snb_update_object_triangles(snb_ctx* ctx, uint32_t index, const float* tris_host);
/* ≙ tinyobj::LoadObj as used by VirtualObject (src/synerfgine/virtual_object.cu:38-75): triangles n x 9. */
int snb_obj_load(const char* path, float* tris_out, int32_t tris_cap, int32_t* n_tris_out);
/* ≙ grid_to_bitfield + bitfield_max_pool (src/testbed_nerf.cu:285-332,3212-3229), device kernels. density: fp32
 * [(max_cascade+1) * 128^3] Morton-ordered host array; out: 8 * 128^3 / 8 bytes host. */
int snb_density_to_bitfield(snb_ctx* ctx, const float* density_host, uint32_t max_cascade, uint8_t* bitfield_out_host);

#ifdef __cplusplus
}
#endif
#endif /* SNB_H */
