"""Host-side virtual scene (no torch, no CUDA): the parsed JSON with per-object BVHs, materials, lights, rendering settings
and camera path — ≙ what sng::Engine::set_virtual_world builds (reference src/synerfgine/engine.cu:21-78)."""
import ctypes as C
import os

import numpy as np

from . import _lib
from ._lib import Light, Material, Object, RtParams


class Scene:
    """Parsed virtual-scene JSON (host side): objects with BVHs, materials, lights, rendering settings, camera path."""

    def __init__(self, json_text, base_dir=None):
        self.lib = _lib.load()
        self.handle = C.c_void_p()
        err = C.create_string_buffer(512)
        rc = self.lib.snb_scene_load_json(json_text.encode(), base_dir.encode() if base_dir else None, C.byref(self.handle), err, 512)
        if rc != 0:
            raise _lib.SnbError(f"scene JSON: {err.value.decode()}")
        self.refresh()
        view = (C.c_float * 3)(); at = (C.c_float * 3)(); zoom = C.c_float(); vo = C.c_float(); nk = C.c_uint32()
        self.lib.snb_scene_camera(self.handle, view, at, C.byref(zoom), C.byref(vo), C.byref(nk))
        self.view_dir, self.at_pos, self.zoom, self.vo_scale, self.n_keyframes = list(view), list(at), zoom.value, vo.value, nk.value
        tf, pl, an, rf = C.c_uint32(), C.c_int32(), C.c_int32(), C.c_int32()
        self.lib.snb_scene_path_info(self.handle, C.byref(tf), C.byref(pl), C.byref(an), C.byref(rf))
        self.total_frames, self.move_on_start, self.animations, self.res_factor = tf.value, bool(pl.value), bool(an.value), rf.value

    def resolution(self, window_w, window_h):
        """≙ Engine::resize: (nerf_w, nerf_h, syn_w, syn_h, syn_px_scale) for a window, from "res_factor" and "vo_scale"."""
        o = [C.c_int() for _ in range(5)]
        rc = self.lib.snb_scene_resolution(self.handle, window_w, window_h, *[C.byref(x) for x in o])
        if rc != 0:
            raise _lib.SnbError("snb_scene_resolution")
        return tuple(x.value for x in o)

    @classmethod
    def from_file(cls, path, base_dir=None):
        with open(path) as f:
            return cls(f.read(), base_dir if base_dir is not None else os.path.dirname(os.path.abspath(path)))

    def refresh(self):
        no, nm, nl = C.c_uint32(), C.c_uint32(), C.c_uint32()
        self.lib.snb_scene_counts(self.handle, C.byref(no), C.byref(nm), C.byref(nl))
        self.objects = (Object * max(1, no.value))(); self.materials = (Material * max(1, nm.value))(); self.lights = (Light * max(1, nl.value))()
        self.n_objects, self.n_materials, self.n_lights = no.value, nm.value, nl.value
        self.rt = RtParams()
        self.lib.snb_scene_get(self.handle, self.objects, self.materials, self.lights, C.byref(self.rt))

    def camera_at(self, frame, total_frames=0):
        view = (C.c_float * 3)(); at = (C.c_float * 3)(); zoom = C.c_float()
        rc = self.lib.snb_scene_camera_at(self.handle, frame, total_frames, view, at, C.byref(zoom))
        if rc != 0:
            raise _lib.SnbError("scene has no camera path")
        return list(view), list(at), zoom.value

    def advance(self):
        self.lib.snb_scene_advance(self.handle)
        self.refresh()

    def __del__(self):
        try:
            if self.handle:
                self.lib.snb_scene_free(self.handle)
                self.handle = C.c_void_p()
        except Exception:
            pass

    def camera_matrix(self):
        """Camera matrix (12 floats, column-major mat4x3) Engine::init derives from the JSON camera block."""
        out = np.zeros(12, dtype=np.float32)
        v = np.asarray(self.view_dir, dtype=np.float32); a = np.asarray(self.at_pos, dtype=np.float32)
        self.lib.snb_camera_from_view(v.ctypes.data_as(_lib.pf32), a.ctypes.data_as(_lib.pf32), float(self.zoom), out.ctypes.data_as(_lib.pf32))
        return out

