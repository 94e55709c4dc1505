"""ctypes binding of oracle/liboracle.so — TEST INFRASTRUCTURE ONLY (never imported by synerfgine_b200/)."""
import ctypes as C
import os
import subprocess

import numpy as np

from synerfgine_b200._lib import CameraEx, Light, Material, ModelDesc, Object, RtParams, View

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "liboracle.so")
pf, pu32, pu16, pi32 = C.POINTER(C.c_float), C.POINTER(C.c_uint32), C.POINTER(C.c_uint16), C.POINTER(C.c_int32)
u32, u64, f32 = C.c_uint32, C.c_uint64, C.c_float

_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    srcs = [os.path.join(ORACLE_DIR, f) for f in os.listdir(ORACLE_DIR) if f.endswith((".cpp", ".h"))]
    if not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in srcs):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "liboracle.so"], stdout=subprocess.DEVNULL)
    lib = C.CDLL(LIB)
    M, V = C.POINTER(ModelDesc), C.POINTER(View)
    sig = {
        "orc_set_math_mode": (None, [C.c_int]),
        "orc_det_logf": (f32, [f32]), "orc_det_expf": (f32, [f32]), "orc_det_sinf": (f32, [f32]), "orc_det_acosf": (f32, [f32]),
        "orc_march_trace": (None, [M, V, u32, pf, pf, pu32, pu32, pf]),
        "orc_ld_random_val": (f32, [u32, u32]),
        "orc_ld_random_pixel_offset": (None, [u32, pf]),
        "orc_calc_dt": (f32, [f32, f32]),
        "orc_advance_n_steps": (f32, [f32, f32, f32]),
        "orc_cascaded_grid_idx_at": (u32, [pf, u32]),
        "orc_mip_from_pos": (u32, [pf]),
        "orc_march_to_occupied": (f32, [M, pf, pf, f32]),
        "orc_init_ray": (None, [V, M, u32, u32, pf]),
        "orc_set_grid_scales": (None, [pf, u32]),
        "orc_set_camera_ex": (None, [C.POINTER(CameraEx)]),
        "orc_set_envmap": (None, [pf, C.c_int32, C.c_int32]),
        "orc_n_params": (u64, [u32, u32, u32, u32, f32]),
        "orc_grid_layout": (None, [u32, u32, u32, u32, f32, pf, pu32, pu32]),
        "orc_network_inference": (None, [M, pf, u64, pu16, pu16, pu32]),
        "orc_render_nerf": (u64, [M, V, C.c_int, C.c_int, u32, pf, pf, pf, pu16]),
        "orc_render_nerf_wavefront": (u64, [M, V, C.c_int, C.c_int, u32, pf, pf, pf, pu32, pu32, u32]),
        "orc_render_nerf_sched": (u64, [M, V, C.c_int, C.c_int, pu32, u32, pf, pf, pf]),
        "orc_write_normals": (None, [C.c_int, C.c_int, pf, pf]),
        "orc_accumulate_tonemap": (None, [C.c_int, pf, pf, u32, C.c_int, C.c_int, f32, pf, C.c_int, pf]),
        "orc_tonemap_curve": (None, [pf, C.c_int]),
        "orc_xorwow_states": (None, [u64, u32, pu32]),
        "orc_xorwow_jump_matrix": (None, [pu32]),
        "orc_xorwow_uniforms": (None, [u64, u32, u32, pf]),
        "orc_trace_world": (None, [C.POINTER(Object), u32, pf, u32, pi32, pi32, pf]),
        "orc_syn_init_rays": (None, [V, pf, pf, pf, pf, pf, pf]),
        "orc_syn_raytrace": (None, [u32, pf, pf, V, M, C.POINTER(Object), u32, C.POINTER(Material), u32, C.POINTER(Light), u32, C.POINTER(RtParams), pu32, pf, pf]),
        "orc_shade_nerf_shadows": (None, [C.c_int, C.c_int, M, pf, pf, C.POINTER(Object), u32, C.POINTER(Light), u32, C.POINTER(RtParams), pu32, C.c_int, pf, pf]),
        "orc_overlay": (None, [C.c_int, C.c_int, C.c_int, pf, pf, pf, pf, C.c_int, C.c_int, C.c_int, f32, f32, pf, pf]),
        "orc_present": (None, [pf, C.c_int, C.c_int, C.c_int, C.c_int, pf, pf, pf, C.POINTER(C.c_uint8)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    lib.orc_set_math_mode(1)   # deterministic transcendentals: the mode the CUDA product is bit-comparable with
    _lib = lib
    return lib


def P(a, t=pf):
    return a.ctypes.data_as(t)


def set_camera_ex(ex=None, envmap=None):
    """Stock-path camera extras of the following oracle renders (None, None: pinhole, no environment map)."""
    lib = load()
    lib.orc_set_camera_ex(C.byref(ex) if ex is not None else None)
    if envmap is None:
        lib.orc_set_envmap(None, 0, 0)
    else:
        e = np.ascontiguousarray(envmap, np.float32)
        lib.orc_set_envmap(P(e), e.shape[1], e.shape[0])


def render_nerf(model_desc, view, alt, render_mode=1, steps_per_batch=8, positions=False):
    """Oracle NeRF layer. Returns fb [h,w,4], depth [h,w], positions [h,w,3] or None, n_samples."""
    lib = load()
    n = view.width * view.height
    fb = np.zeros((n, 4), np.float32); depth = np.zeros(n, np.float32)
    pos = np.zeros((n, 3), np.float32) if positions else None
    ns = lib.orc_render_nerf(C.byref(model_desc), C.byref(view), int(alt), render_mode, steps_per_batch, P(fb), P(depth), P(pos) if positions else None, None)
    h, w = view.height, view.width
    return fb.reshape(h, w, 4), depth.reshape(h, w), (pos.reshape(h, w, 3) if positions else None), ns


def render_nerf_wavefront(model_desc, view, alt=1, render_mode=1, target_queries=2 * 1024 * 1024, positions=False, cap=256):
    """Replay of the reference's lock-step march (trace / trace_alt). Returns fb, depth, positions, schedule (S per iteration), alive (n_alive per iteration)."""
    lib = load()
    n = view.width * view.height
    fb = np.zeros((n, 4), np.float32); depth = np.zeros(n, np.float32)
    pos = np.zeros((n, 3), np.float32) if positions else None
    sched = np.zeros(cap, np.uint32); alive = np.zeros(cap, np.uint32)
    lib.orc_render_nerf_wavefront(C.byref(model_desc), C.byref(view), int(alt), render_mode, target_queries, P(fb), P(depth), P(pos) if positions else None,
                                  P(sched, pu32), P(alive, pu32), cap)
    k = int(np.count_nonzero(sched))
    h, w = view.height, view.width
    return fb.reshape(h, w, 4), depth.reshape(h, w), (pos.reshape(h, w, 3) if positions else None), sched[:k].copy(), alive[:k].copy()


def render_nerf_sched(model_desc, view, sched, alt=1, render_mode=1, positions=False):
    """Per-ray march with an explicit batch-length sequence."""
    lib = load()
    n = view.width * view.height
    fb = np.zeros((n, 4), np.float32); depth = np.zeros(n, np.float32)
    pos = np.zeros((n, 3), np.float32) if positions else None
    s = np.ascontiguousarray(sched, np.uint32)
    lib.orc_render_nerf_sched(C.byref(model_desc), C.byref(view), int(alt), render_mode, P(s, pu32), s.size, P(fb), P(depth), P(pos) if positions else None)
    h, w = view.height, view.width
    return fb.reshape(h, w, 4), depth.reshape(h, w), (pos.reshape(h, w, 3) if positions else None)


def frame(model_desc, view, scene, rt, rng_seed=1999, n_frames=1):
    """Oracle of Engine::frame (syn_px_scale 1): returns dict of every intermediate buffer of the last frame."""
    lib = load()
    w, h = view.width, view.height
    n = w * h
    origins = np.zeros((n, 3), np.float32); dirs = np.zeros((n, 3), np.float32)
    acc_rgba = np.zeros((n, 4), np.float32); acc_depth = np.zeros(n, np.float32)
    lib.orc_syn_init_rays(C.byref(view), P(origins), P(dirs), None, None, P(acc_rgba), P(acc_depth))
    rng_syn = np.zeros((n, 6), np.uint32); lib.orc_xorwow_states(rng_seed, n, P(rng_syn, pu32))
    rng_nerf = rng_syn.copy()
    out = {}
    for _ in range(n_frames):
        lib.orc_syn_raytrace(n, P(origins), P(dirs), C.byref(view), C.byref(model_desc), scene.objects, scene.n_objects, scene.materials, scene.n_materials,
                             scene.lights, scene.n_lights, C.byref(rt), P(rng_syn, pu32), P(acc_rgba), P(acc_depth))
        fb = np.zeros((n, 4), np.float32); depth = np.zeros(n, np.float32); pos = np.zeros((n, 3), np.float32); normals = np.zeros((n, 3), np.float32)
        shadow = np.ones(n, np.float32)
        if rt.show_nerf:
            lib.orc_render_nerf(C.byref(model_desc), C.byref(view), 1, 1, 8, P(fb), P(depth), P(pos), None)
            lib.orc_write_normals(w, h, P(pos), P(normals))
            out["nerf_rgba_unshadowed"] = fb.copy().reshape(h, w, 4)
            if rt.show_shadow_on_nerf:
                lib.orc_shade_nerf_shadows(w, h, C.byref(model_desc), P(pos), P(normals), scene.objects, scene.n_objects, scene.lights, scene.n_lights,
                                           C.byref(rt), P(rng_nerf, pu32), 1, P(fb), P(shadow))
        final = np.zeros((n, 4), np.float32); final_depth = np.zeros(n, np.float32)
        lib.orc_overlay(w, h, 1, P(acc_rgba), P(acc_depth), P(fb), P(depth), rt.show_nerf, 1, rt.tonemap_curve, rt.exposure, rt.depth_offset, P(final), P(final_depth))
        out.update(syn_rgba=acc_rgba.copy().reshape(h, w, 4), syn_depth=acc_depth.copy().reshape(h, w), nerf_rgba=fb.reshape(h, w, 4), nerf_depth=depth.reshape(h, w),
                   positions=pos.reshape(h, w, 3), normals=normals.reshape(h, w, 3), shadow=shadow.reshape(h, w), final=final.reshape(h, w, 4),
                   final_depth=final_depth.reshape(h, w))
    return out


class FrameLoop:
    """Oracle of consecutive Engine::frame calls (syn_px_scale 1) with a moving camera and world: generator states carry from frame to
    frame (Engine::resize seeds them once), the mesh layer's rays and temporal blend restart whenever `reset` is set
    (RayTracer::reset_accumulation: a camera move, engine.cu:366-369, or Engine::update_world_objects, :80-127)."""

    def __init__(self, w, h, rng_seed=1999):
        self.lib = load()
        self.w, self.h, self.n = w, h, w * h
        n = self.n
        self.origins = np.zeros((n, 3), np.float32); self.dirs = np.zeros((n, 3), np.float32)
        self.acc_rgba = np.zeros((n, 4), np.float32); self.acc_depth = np.zeros(n, np.float32)
        self.rng_syn = np.zeros((n, 6), np.uint32); self.lib.orc_xorwow_states(rng_seed, n, P(self.rng_syn, pu32))
        self.rng_nerf = self.rng_syn.copy()
        self.fresh = True

    def step(self, model_desc, view, scene, rt, reset):
        lib, n, w, h = self.lib, self.n, self.w, self.h
        if reset or self.fresh:
            lib.orc_syn_init_rays(C.byref(view), P(self.origins), P(self.dirs), None, None, P(self.acc_rgba), P(self.acc_depth))
            self.fresh = False
        lib.orc_syn_raytrace(n, P(self.origins), P(self.dirs), C.byref(view), C.byref(model_desc), scene.objects, scene.n_objects, scene.materials, scene.n_materials,
                             scene.lights, scene.n_lights, C.byref(rt), P(self.rng_syn, pu32), P(self.acc_rgba), P(self.acc_depth))
        fb = np.zeros((n, 4), np.float32); depth = np.zeros(n, np.float32); pos = np.zeros((n, 3), np.float32); normals = np.zeros((n, 3), np.float32)
        shadow = np.ones(n, np.float32)
        if rt.show_nerf:
            lib.orc_render_nerf(C.byref(model_desc), C.byref(view), 1, 1, 8, P(fb), P(depth), P(pos), None)
            lib.orc_write_normals(w, h, P(pos), P(normals))
            if rt.show_shadow_on_nerf:
                lib.orc_shade_nerf_shadows(w, h, C.byref(model_desc), P(pos), P(normals), scene.objects, scene.n_objects, scene.lights, scene.n_lights,
                                           C.byref(rt), P(self.rng_nerf, pu32), 1, P(fb), P(shadow))
        final = np.zeros((n, 4), np.float32); final_depth = np.zeros(n, np.float32)
        lib.orc_overlay(w, h, 1, P(self.acc_rgba), P(self.acc_depth), P(fb), P(depth), rt.show_nerf, 1, rt.tonemap_curve, rt.exposure, rt.depth_offset, P(final), P(final_depth))
        return {"final": final.reshape(h, w, 4), "final_depth": final_depth.reshape(h, w), "syn_depth": self.acc_depth.copy().reshape(h, w),
                "nerf_rgba": fb.reshape(h, w, 4), "shadow": shadow.reshape(h, w)}


def present(rgba, out_w, out_h, clear=(0.0, 0.0, 0.0), warp_x=None, warp_y=None):
    """Oracle present stage (FXAA blit + blend + RGB8): rgba [h, w, 4] float32 -> uint8 [out_h, out_w, 3], rows top-down."""
    lib = load()
    rgba = np.ascontiguousarray(rgba, np.float32)
    out = np.zeros((out_h, out_w, 3), np.uint8)
    cl = np.asarray(clear, np.float32)
    wx = np.asarray(warp_x, np.float32) if warp_x is not None else None
    wy = np.asarray(warp_y, np.float32) if warp_y is not None else None
    lib.orc_present(P(rgba), rgba.shape[1], rgba.shape[0], out_w, out_h, P(cl), P(wx) if wx is not None else None, P(wy) if wy is not None else None,
                    out.ctypes.data_as(C.POINTER(C.c_uint8)))
    return out
