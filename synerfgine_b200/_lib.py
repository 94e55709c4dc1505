"""ctypes binding of libsnb.so (include/snb.h) — the thin reference-side stub described in INTEGRATION.md.

The library is built in-tree by ``__graft_entry__.build()`` / ``make -C synerfgine_b200/csrc``. There is no
fallback: if the shared object is missing this module raises, and every render call needs a CUDA device.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SNB_LIB") or os.path.join(_HERE, "libsnb.so")   # SNB_LIB: alternative build of the same library (tuning runs)

f32, u32, i32, u64, u16, u8 = C.c_float, C.c_uint32, C.c_int32, C.c_uint64, C.c_uint16, C.c_uint8
pf32 = C.POINTER(f32)


class ModelDesc(C.Structure):
    _fields_ = [("n_levels", u32), ("n_features_per_level", u32), ("log2_hashmap_size", u32), ("base_resolution", u32),
                ("per_level_scale", f32), ("params", C.POINTER(u16)), ("n_params", u64), ("bitfield", C.POINTER(u8)),
                ("aabb_min", f32 * 3), ("aabb_max", f32 * 3), ("render_aabb_min", f32 * 3), ("render_aabb_max", f32 * 3),
                ("render_aabb_to_local", f32 * 9), ("max_cascade", u32), ("cone_angle_constant", f32),
                ("rgb_activation", i32), ("density_activation", i32), ("min_transmittance", f32)]


class View(C.Structure):
    _fields_ = [("camera", f32 * 12), ("focal_length", f32 * 2), ("screen_center", f32 * 2), ("width", i32), ("height", i32),
                ("spp", u32), ("snap_to_pixel_centers", i32)]


class CameraEx(C.Structure):
    """snb_camera_ex: the stock render's camera beyond the pinhole (end-of-shutter camera, rolling shutter, parallax, near plane, DoF, lens)."""
    _fields_ = [("camera1", f32 * 12), ("rolling_shutter", f32 * 4), ("parallax_shift", f32 * 3), ("near_distance", f32), ("plane_z", f32),
                ("aperture_size", f32), ("lens_mode", i32), ("lens_params", f32 * 7)]


LENS_MODES = {"perspective": 0, "opencv": 1, "ftheta": 2, "latlong": 3, "opencv_fisheye": 4, "equirectangular": 5}


class Tile(C.Structure):
    _fields_ = [("x0", i32), ("y0", i32), ("w", i32), ("h", i32)]


class FovWarp(C.Structure):
    _fields_ = [(n, f32) for n in ("al", "bl", "cl", "am", "bm", "ar", "br", "cr", "switch_left", "switch_right", "inv_switch_left", "inv_switch_right")]


class Object(C.Structure):
    _fields_ = [("nodes", pf32), ("tris", pf32), ("n_nodes", i32), ("n_tris", i32), ("rot", f32 * 9), ("pos", f32 * 3),
                ("scale", f32), ("mat_id", i32)]


class Material(C.Structure):
    _fields_ = [("ka", f32 * 3), ("kd", f32 * 3), ("ks", f32 * 3), ("n", f32), ("rg", f32), ("type", i32), ("spec_angle", f32)]


class Light(C.Structure):
    _fields_ = [("pos", f32 * 3), ("type", i32), ("intensity", f32), ("size", f32)]


class RtParams(C.Structure):
    _fields_ = [("n_steps", u32), ("samples", u32), ("bounces", u32), ("shadow_iters", u32), ("lens_angle_constant", f32),
                ("depth_offset", f32), ("syn_shadow_factor", f32), ("view_nerf_shadow", i32), ("buffer_type", i32),
                ("nerf_shadow_kernel", i32), ("nerf_shadow_intensity", f32), ("nerf_on_nerf_shadow_threshold", f32),
                ("show_shadow_on_nerf", i32), ("show_nerf", i32), ("syn_px_scale", i32), ("exposure", f32), ("tonemap_curve", i32)]

    def copy(self):
        r = RtParams()
        C.memmove(C.byref(r), C.byref(self), C.sizeof(RtParams))
        return r


vp = C.c_void_p
# name -> (restype, argtypes); every symbol declared in include/snb.h
SIGNATURES = {
    "snb_create": (C.c_int, [C.c_int, vp, C.POINTER(vp)]),
    "snb_destroy": (None, [vp]),
    "snb_last_error": (C.c_char_p, [vp]),
    "snb_sync": (C.c_int, [vp]),
    "snb_version": (C.c_char_p, []),
    "snb_set_option": (C.c_int, [vp, C.c_char_p, C.c_char_p]),
    "snb_set_model": (C.c_int, [vp, C.POINTER(ModelDesc)]),
    "snb_set_model_dev": (C.c_int, [vp, C.POINTER(ModelDesc)]),
    "snb_set_scene": (C.c_int, [vp, C.POINTER(Object), u32, C.POINTER(Material), u32, C.POINTER(Light), u32]),
    "snb_update_scene": (C.c_int, [vp, C.POINTER(Object), u32, C.POINTER(Material), u32, C.POINTER(Light), u32]),
    "snb_seed_rng": (C.c_int, [vp, u64]),
    "snb_reset_accumulation": (C.c_int, [vp]),
    "snb_render_nerf": (C.c_int, [vp, C.POINTER(View), C.POINTER(Tile), C.c_int, vp, vp]),
    "snb_render_syn": (C.c_int, [vp, C.POINTER(View), C.POINTER(Tile), C.POINTER(RtParams), vp, vp]),
    "snb_render_nerf_with_buffers": (C.c_int, [vp, C.POINTER(View), C.POINTER(Tile), C.POINTER(RtParams), C.c_int, vp, vp, vp, vp]),
    "snb_overlay": (C.c_int, [vp, C.c_int, C.c_int, C.POINTER(RtParams), vp, vp, vp, vp, vp, vp]),
    "snb_accumulate_tonemap": (C.c_int, [vp, C.c_int, vp, vp, u32, C.c_int, C.c_int, f32, pf32, C.c_int, vp]),
    "snb_frame": (C.c_int, [vp, C.POINTER(View), C.POINTER(Tile), C.POINTER(RtParams), vp, vp]),
    "snb_frame_host": (C.c_int, [vp, C.POINTER(View), C.POINTER(Tile), C.POINTER(RtParams), vp, vp]),
    "snb_frame_host_async": (C.c_int, [vp, C.POINTER(View), C.POINTER(Tile), C.POINTER(RtParams), vp, vp, i32]),
    "snb_frame_host_wait": (C.c_int, [vp, i32]),
    "snb_march_trace": (C.c_int, [vp, C.POINTER(View), u32, vp, vp, vp, vp, vp]),
    "snb_network_inference": (C.c_int, [vp, vp, u64, vp, vp, vp]),
    "snb_trace_rays": (C.c_int, [vp, vp, u32, vp, vp, vp]),
    "snb_get_counters": (C.c_int, [vp, C.POINTER(u64)]),
    "snb_get_march_stats": (C.c_int, [vp, C.POINTER(u64)]),
    "snb_get_stage_ms": (C.c_int, [vp, pf32]),
    "snb_get_ws_profile": (C.c_int, [vp, C.POINTER(u64)]),
    "snb_get_grid_scales": (C.c_int, [vp, pf32, u32]),
    "snb_rng_tap": (C.c_int, [vp, u64, u32, u32, vp, vp]),
    "snb_get_alt_status": (C.c_int, [vp, C.POINTER(u32)]),
    "snb_set_alt_schedule": (C.c_int, [vp, C.POINTER(u8), u32]),
    "snb_bvh_build_dev": (C.c_int, [vp, vp, i32, i32, vp, vp, i32, vp, C.POINTER(i32), C.POINTER(i32)]),
    "snb_bvh_refit_dev": (C.c_int, [vp, vp, vp, i32]),
    "snb_bvh_refit_dev_noalloc": (C.c_int, [vp, vp, vp, i32]),
    "snb_update_object_triangles": (C.c_int, [vp, u32, pf32]),
    "snb_set_sample_tap": (C.c_int, [vp, vp, vp, vp, u32]),
    "snb_set_camera_ex": (C.c_int, [vp, C.POINTER(CameraEx)]),
    "snb_set_envmap": (C.c_int, [vp, pf32, i32, i32]),
    "snb_camera_log_lerp": (C.c_int, [pf32, pf32, f32, pf32]),
    "snb_launch_count": (u64, [vp]),
    "snb_scene_load_json": (C.c_int, [C.c_char_p, C.c_char_p, C.POINTER(vp), C.c_char_p, C.c_size_t]),
    "snb_scene_free": (None, [vp]),
    "snb_scene_counts": (C.c_int, [vp, C.POINTER(u32), C.POINTER(u32), C.POINTER(u32)]),
    "snb_scene_get": (C.c_int, [vp, C.POINTER(Object), C.POINTER(Material), C.POINTER(Light), C.POINTER(RtParams)]),
    "snb_scene_camera": (C.c_int, [vp, pf32, pf32, pf32, pf32, C.POINTER(u32)]),
    "snb_scene_camera_at": (C.c_int, [vp, u32, u32, pf32, pf32, pf32]),
    "snb_camera_from_view": (C.c_int, [pf32, pf32, f32, pf32]),
    "snb_scene_advance": (C.c_int, [vp]),
    "snb_scene_resolution": (C.c_int, [vp, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "snb_scene_path_info": (C.c_int, [vp, C.POINTER(u32), C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]),
    "snb_bvh_build": (C.c_int, [pf32, i32, i32, pf32, i32, C.POINTER(i32)]),
    "snb_obj_load": (C.c_int, [C.c_char_p, pf32, i32, C.POINTER(i32)]),
    "snb_density_to_bitfield": (C.c_int, [vp, pf32, u32, C.POINTER(u8)]),
    "snb_present": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, pf32, C.POINTER(FovWarp), C.POINTER(FovWarp), vp]),
    "snb_present_host": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, pf32, C.POINTER(FovWarp), C.POINTER(FovWarp), vp]),
    "snb_write_png": (C.c_int, [C.c_char_p, vp, C.c_int, C.c_int]),
    "snb_snapshot_load": (C.c_int, [C.c_char_p, C.POINTER(vp), C.c_char_p, C.c_size_t]),
    "snb_snapshot_free": (None, [vp]),
    "snb_snapshot_get": (C.c_int, [vp, C.POINTER(ModelDesc), C.POINTER(C.POINTER(C.c_uint16)), C.POINTER(u64), C.POINTER(u32), pf32]),
    "snb_snapshot_save": (C.c_int, [C.c_char_p, C.POINTER(ModelDesc), C.POINTER(C.c_uint16), u64, u32, pf32, C.c_int]),
}

_lib = None


def load():
    """Load libsnb.so once and attach the signatures. Raises if the extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU or PyTorch fallback for the render path)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class SnbError(RuntimeError):
    pass


def check(lib, ctx, rc, what):
    if rc != 0:
        msg = lib.snb_last_error(ctx).decode() if ctx else ""
        raise SnbError(f"{what} failed ({rc}): {msg}")
