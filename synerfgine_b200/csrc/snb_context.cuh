// synerfgine_b200 — host-side context behind the C ABI: owns the device copies of the model and the scene, the
// persistent scratch (ray queue, RNG states, layer buffers) and the stream all kernels are enqueued on.
// Mirrors the state the reference spreads over Testbed (network, bitfield, render buffers), sng::Engine
// (world / material / light arrays, NeRF RNG states, normals, positions) and sng::RayTracer (ray SoA, RNG
// states, temporal accumulation) — include/synerfgine/engine.cuh:79-120, raytracer.cuh:150-182.
#pragma once
#include <string>
#include <vector>
#include <cstdio>
#include <cuda_runtime.h>
#include "../../include/snb.h"
#include "snb_device.cuh"
#include "snb_nerf_kernels.cuh"
#include "snb_mlp_tc.cuh"
#include "snb_march_ws.cuh"
#include "snb_present.cuh"
#include "snb_sng_kernels.cuh"

namespace snb {

// Grow-only device buffer (the reference re-acquires arena scratch every frame, src/testbed_nerf.cu:2403-2435).
struct DevBuf {
	void* p = nullptr; size_t cap = 0;
	cudaError_t reserve(size_t bytes) {
		if (bytes <= cap) return cudaSuccess;
		if (p) cudaFree(p);
		p = nullptr; cap = 0;
		cudaError_t e = cudaMalloc(&p, bytes);
		if (e == cudaSuccess) cap = bytes;
		return e;
	}
	void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
	template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct ObjectStore { DevBuf nodes, tris, order; int n_nodes = 0, n_tris = 0; bool permuted = false; };   // order: source index of every device triangle (device-built BVHs)

} // namespace snb

#include "snb_bvh_dev.cuh"   // needs DevBuf

struct snb_ctx {
	int device = 0;
	cudaStream_t stream = nullptr;
	bool own_stream = false;
	std::string err;
	int sm_count = 148;
	uint64_t launches = 0;

	// model
	bool has_model = false;
	snb::DevModel model;
	uint32_t feat_per_level = 2;
	snb::DevBuf d_grid, d_mlp, d_mlp_frag, d_mlp_umma, d_bitfield, d_bits_xyz, d_coarse, d_present;

	// scene
	std::vector<snb::ObjectStore> obj_store;
	snb::DevBuf d_objects, d_materials, d_lights;
	snb::DevScene scene{nullptr, 0, nullptr, 0, nullptr, 0};

	// scratch
	snb::DevBuf d_queue, d_qctl, d_counters;
	snb::DevBuf d_rng_nerf, d_rng_syn;
	int rng_w = 0, rng_h = 0, rng_tx0 = -1, rng_ty0 = -1, rng_tw = 0, rng_th = 0, rng_scale = 0;
	uint64_t rng_seed = 1999;          // PT_SEED, include/synerfgine/common.cuh:20
	bool rng_dirty = true;
	snb::DevBuf d_syn_origins, d_syn_dirs, d_syn_rgba, d_syn_depth;
	bool syn_reset = true;
	int syn_w = 0, syn_h = 0;
	snb::DevBuf d_nerf_fb, d_nerf_depth, d_positions, d_normals, d_terms, d_out_rgba, d_out_depth;
	int fused_blocks_per_sm = 0;
	int lane_blocks_per_sm[32] = {};   // resident CTAs per SM of each k_nerf_march_fused instance launched by THIS context (0 = not set up yet)
	int mlp_variant = 0;               // 0 = warp-level mma.sync evaluator, 1 = CTA-level tcgen05/TMEM evaluator (snb_mlp_tc.cuh)
	int march_variant = 3;             // 3 = one ray per lane + mma.sync MLPs (k_nerf_march_fused: the faster one, default), 6 = warp-specialised tcgen05 pipeline (k_nerf_march_ws)
	bool ws_ready[8] = {};             // dynamic shared-memory opt-in done for each k_nerf_march_ws instance on this context's device
	// batch schedule of the sng compositor (AltState in snb_nerf_kernels.cuh): 0 = "auto" (device-resident sequence, verified and
	// repeated inside the frame; moving frames: lock-step prefix), 1 = "external" (one pass; the caller reads the counts and sets the
	// sequence: multi-GPU), 2 = "fixed", 3 = "lockstep" (always the lock-step prefix)
	int alt_mode = 0;
	int alt_batch = 8;                 // batch length of alt_mode 2
	bool dup_reuse = true;             // sng path: re-use the network output of the sample the compositor's rewind evaluates twice (option dup_reuse)
	bool defer_srgb = true;            // sng path, lane kernel: the march stores sRGB colours, one pass over the tile converts them (option defer_srgb)
	bool alt_auto_lockstep = true;     // alt_mode 0: frames whose camera or tile changed march the lock-step prefix (option alt_auto_lockstep); the march never reads the scene
	float alt_last_cam[12] = {};       // camera and tile size of the last sng NeRF layer
	int alt_last_tw = 0, alt_last_th = 0, alt_last_tx0 = 0, alt_last_ty0 = 0;
	float alt_last_focal = 0.0f;
	snb::DevBuf d_park, d_lock_lists;  // lock-step prefix: RayPark per queued ray, two lists of queue positions
	uint32_t alt_target_queries = 2u * 1024u * 1024u;   // target_n_queries of trace_alt (src/testbed_nerf.cu:2188)
	snb::DevBuf d_alt;
	// BVH construction (option "bvh"): 0 = the nodes the caller passes (reference builder, parity mode), 1 = linear BVH built on the device
	int bvh_builder = 0;
	int bvh_leaf = 4;                  // triangles per leaf of the device builder ("primitives-per-leaf" default, src/synerfgine/virtual_object.cu:14)
	int bvh_max_depth = 0;             // deepest leaf of the last device builds (the traversal stack holds 32 entries)
	snb::LbvhScratch lbvh;
	snb::DevBuf d_tris_in, d_bvh_depth;
	// parity tap of the fused march on the stock path (snb_set_sample_tap): caller-owned device arrays
	float* tap_t = nullptr; uint32_t* tap_cell = nullptr; uint32_t* tap_n = nullptr; uint32_t tap_cap = 0;
	// stock-path camera extras (snb_set_camera_ex / snb_set_envmap); cam_ex.enabled == 0 and no envmap: plain pinhole ray generation
	snb::CameraEx cam_ex = {};
	bool has_cam_ex = false;
	snb::DevBuf d_envmap; int env_w = 0, env_h = 0;

	// per-stage device timing of the last frame (CUDA events on the context's stream; see snb_get_stage_ms)
	static const int kStages = 5;   // 0 nerf raygen, 1 nerf fused march, 2 nerf normals+shadows, 3 mesh layer, 4 overlay
	cudaEvent_t ev[kStages][2] = {};
	bool ev_set[kStages] = {};
	void stage_begin(int s, cudaStream_t st = nullptr) { if (!ev[s][0]) { cudaEventCreate(&ev[s][0]); cudaEventCreate(&ev[s][1]); } cudaEventRecord(ev[s][0], st ? st : stream); ev_set[s] = false; }
	void stage_end(int s, cudaStream_t st = nullptr) { cudaEventRecord(ev[s][1], st ? st : stream); ev_set[s] = true; }

	// second stream for the mesh layer of snb_frame: it does not depend on the NeRF layer until the overlay, and both are
	// latency-bound, so they run concurrently (the reference serialises them with host syncs, engine.cu:386-406)
	cudaStream_t stream2 = nullptr;   // mesh layer
	cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
	// snb_frame_host_async: copy stream + double-buffered device outputs
	cudaStream_t copy_stream = nullptr;
	cudaEvent_t ev_frame_done[2] = {}, ev_copy_done[2] = {};
	snb::DevBuf d_pipe_rgba[2], d_pipe_depth[2];
	bool pipe_used[2] = {false, false};
	bool overlap_layers = true;
	int mesh_priority = 1;         // stream priority of the mesh layer's stream: +1 high (default: the mesh layer finishes under ray generation and the march runs undisturbed), 0 default, -1 low (option "mesh_priority")
	int lane_warps = 0;            // lane march CTA shape: 0 = chosen by tile size, 10 | 12 = forced (option "lane_warps")
	int sng_block = 128;           // k_raytrace: threads per block with a warp = 8 x 4 pixels; 0 = row-linear mapping (option "sng_block")
	int terms_block = 0;           // k_nerf_shadow_terms: same choice (option "terms_block"); row-linear measured faster there
	bool skip_dead_terms = true;   // k_nerf_shadow_terms: skip terms no written pixel can observe (option "dead_terms")

	int fail(int code, const std::string& msg) { err = msg; return code; }
	int cuda_fail(cudaError_t e, const char* what) {
		char buf[256]; snprintf(buf, sizeof buf, "%s: %s", what, cudaGetErrorString(e));
		err = buf; return SNB_ERR_CUDA;
	}
};

#define SNB_CUDA(ctx, call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return (ctx)->cuda_fail(e__, #call); } while (0)
#define SNB_LAUNCHED(ctx) do { ++(ctx)->launches; cudaError_t e__ = cudaGetLastError(); if (e__ != cudaSuccess) return (ctx)->cuda_fail(e__, "kernel launch"); } while (0)
