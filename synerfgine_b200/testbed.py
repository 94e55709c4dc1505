"""Python surface mirroring the reference's ``pyngp.Testbed`` render entry (src/python_api.cu:125-205,415-560) and the
``sng::Engine`` object (include/synerfgine/engine.cuh:22-33), on top of the C ABI in include/snb.h.

torch is used only to own device memory and streams; all rendering happens in libsnb.so.
"""
import ctypes as C
import math
import os

import numpy as np
import torch

from . import _lib
from ._lib import LENS_MODES, CameraEx, RtParams, Tile, View  # noqa: F401
from .scene import Scene  # noqa: F401


def _fp(a):
    return a.ctypes.data_as(_lib.pf32)


class Testbed:
    """NeRF-only render entry: ``render(width, height, spp, linear)`` -> float32 [H, W, 4] like pyngp.Testbed.render."""

    def __init__(self, device=0, stream=None):
        """stream: optional torch.cuda.Stream the context enqueues on (so torch events can time it); default: own stream."""
        if not torch.cuda.is_available():
            raise RuntimeError("synerfgine_b200 needs a CUDA device (no CPU fallback)")
        self.lib = _lib.load()
        self.device = torch.device("cuda", device)
        torch.cuda.set_device(self.device)
        self.ctx = C.c_void_p()
        self.stream = stream
        rc = self.lib.snb_create(device, C.c_void_p(stream.cuda_stream) if stream is not None else None, C.byref(self.ctx))
        if rc != 0:
            raise _lib.SnbError(f"snb_create failed ({rc})")
        # Testbed::reset_camera defaults (src/testbed.cu:470-492)
        self.fov_axis = 1
        self.zoom = 1.0
        self.screen_center = (0.5, 0.5)
        self.relative_focal_length = 0.5 / math.tan(0.5 * 50.625 * math.pi / 180.0)
        self.scale = 1.5
        self.camera_matrix = np.array([[1, 0, 0], [0, -1, 0], [0, 0, -1], [0.5, 0.5, 0.5 + 1.5]], dtype=np.float32)  # columns
        self.exposure = 0.0
        self.background_color = (0.0, 0.0, 0.0, 0.0)
        self.render_mode = 1        # ERenderMode::Shade
        self.tonemap_curve = 0
        self.snap_to_pixel_centers = True
        # the rest of the stock render's camera (testbed.h: m_aperture_size, m_slice_plane_z, m_parallax_shift, m_render_near_distance,
        # m_nerf.render_with_lens_distortion / render_lens, m_envmap); all off by default = the pinhole path
        self.aperture_size = 0.0
        self.slice_plane_z = 0.0
        self.parallax_shift = (0.0, 0.0, 0.0)
        self.render_near_distance = 0.0
        self.render_with_lens_distortion = False
        self.render_lens_mode = "perspective"
        self.render_lens_params = (0.0,) * 7
        self.envmap = None            # float32 [h, w, 4] or None
        self._envmap_uploaded = None
        # camera path of render(start_t, end_t): callable t in [0, 1] -> camera matrix [4, 3] (columns). The reference evaluates its
        # CameraPath spline here (set_camera_from_time, src/testbed.cu:2203-2216); the spline editor itself is GUI state (out of scope),
        # so the path is whatever the caller supplies, e.g. camera_path_from_keyframes(). None = the camera stays (as with an empty path).
        self.camera_path = None
        self.smoothed_camera = None
        self.model = None
        self._accum = None
        self._spp = 0

    def close(self):
        if self.ctx:
            self.lib.snb_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        _lib.check(self.lib, self.ctx, rc, what)

    # ---- inputs --------------------------------------------------------------------------------------------------
    def set_model(self, model):
        """model: synthetic.SyntheticModel (or anything with .desc()); ≙ the state load_snapshot leaves in Testbed."""
        self.model = model
        d = model.desc()
        self._check(self.lib.snb_set_model(self.ctx, C.byref(d)), "snb_set_model")

    def load_snapshot(self, path):
        """≙ Testbed::load_snapshot (src/testbed.cu:4878-4992): read the `.ingp` / `.msgpack` file, rebuild the occupancy
        bitfield from the stored fp16 density grid on the GPU (update_density_grid_mean_and_bitfield), upload the model and
        adopt the stored camera."""
        h = C.c_void_p()
        err = C.create_string_buffer(512)
        rc = self.lib.snb_snapshot_load(str(path).encode(), C.byref(h), err, 512)
        if rc != 0:
            raise RuntimeError(f"load_snapshot({path}): {err.value.decode()} (status {rc})")
        try:
            d = _lib.ModelDesc(); dens = C.POINTER(C.c_uint16)(); n = C.c_uint64(); scale = C.c_uint32(); cam = (C.c_float * 12)()
            rc = self.lib.snb_snapshot_get(h, C.byref(d), C.byref(dens), C.byref(n), C.byref(scale), cam)
            has_camera = rc == 0
            if rc not in (0, -3):
                raise RuntimeError(f"snb_snapshot_get: status {rc}")
            bits = None
            if n.value:
                density = np.ctypeslib.as_array(dens, shape=(n.value,)).view(np.float16).astype(np.float32)
                bits = np.zeros(8 * 128 ** 3 // 8, dtype=np.uint8)
                self._check(self.lib.snb_density_to_bitfield(self.ctx, _fp(density), d.max_cascade, bits.ctypes.data_as(C.POINTER(C.c_uint8))), "snb_density_to_bitfield")
                d.bitfield = bits.ctypes.data_as(C.POINTER(C.c_uint8))
            self._check(self.lib.snb_set_model(self.ctx, C.byref(d)), "snb_set_model")      # copies params and bitfield
            self.model = None
            self.aabb_scale = scale.value
            if has_camera:
                self.camera_matrix = np.array(list(cam), dtype=np.float32).reshape(4, 3)
        finally:
            self.lib.snb_snapshot_free(h)

    def save_snapshot(self, path, model, compress=True):
        """≙ Testbed::save_snapshot (src/testbed.cu:4812-4876) for a model object with .desc(), .density, .aabb_scale."""
        d = model.desc()
        dens = np.ascontiguousarray(model.density.astype(np.float16)).view(np.uint16)
        cam = (C.c_float * 12)(*[float(x) for x in np.asarray(self.camera_matrix, dtype=np.float32).reshape(-1)])
        self._check(self.lib.snb_snapshot_save(str(path).encode(), C.byref(d), dens.ctypes.data_as(C.POINTER(C.c_uint16)), dens.size, int(model.aabb_scale), cam, int(bool(compress))),
                    "snb_snapshot_save")

    def set_camera_from_view(self, view, at, zoom):
        """≙ Engine::init applying the JSON camera block: set_view_dir, set_look_at, set_scale."""
        out = np.zeros(12, dtype=np.float32)
        v = np.asarray(view, dtype=np.float32); a = np.asarray(at, dtype=np.float32)
        rc = self.lib.snb_camera_from_view(_fp(v), _fp(a), float(zoom), _fp(out))
        if rc != 0:
            raise _lib.SnbError("snb_camera_from_view failed")
        self.camera_matrix = out.reshape(4, 3).copy()
        self.scale = float(zoom)

    def calc_focal_length(self, width, height):
        res = (width, height)[self.fov_axis]
        f = np.float32(self.relative_focal_length) * np.float32(res) * np.float32(self.zoom)
        return (float(f), float(f))

    def make_view(self, width, height, spp=0):
        v = View()
        v.camera[:] = [float(x) for x in np.asarray(self.camera_matrix, dtype=np.float32).reshape(-1)]
        v.focal_length[:] = self.calc_focal_length(width, height)
        sc = [(0.5 - c) * self.zoom + 0.5 for c in self.screen_center]   # render_screen_center, src/testbed.cu:4137-4140
        v.screen_center[:] = sc
        v.width, v.height, v.spp = width, height, spp
        v.snap_to_pixel_centers = 1 if self.snap_to_pixel_centers else 0
        return v

    # ---- render ---------------------------------------------------------------------------------------------------
    def render_nerf_raw(self, width, height, spp_index=0, render_mode=None, tile=None):
        """One NeRF sample frame (render_nerf): returns (rgba [h,w,4], depth [h,w]) device tensors, linear colours."""
        v = self.make_view(width, height, spp_index)
        w, h = (tile.w, tile.h) if tile is not None else (width, height)
        fb = torch.empty((h, w, 4), dtype=torch.float32, device=self.device)
        depth = torch.empty((h, w), dtype=torch.float32, device=self.device)
        mode = self.render_mode if render_mode is None else render_mode
        self._check(self.lib.snb_render_nerf(self.ctx, C.byref(v), C.byref(tile) if tile is not None else None, mode, fb.data_ptr(), depth.data_ptr()),
                    "snb_render_nerf")
        self._check(self.lib.snb_sync(self.ctx), "snb_sync")
        return fb, depth

    def _camera_extras(self, camera0, camera1):
        """snb_camera_ex of one sample, or None when every extra is off and the shutter does not move (-> pinhole path)."""
        lens = self.render_with_lens_distortion and self.render_lens_mode != "perspective"
        moving = not np.array_equal(np.asarray(camera0, np.float32), np.asarray(camera1, np.float32))
        if not (lens or moving or self.aperture_size != 0.0 or self.render_near_distance != 0.0 or any(self.parallax_shift)):
            return None
        ex = CameraEx()
        ex.camera1[:] = [float(x) for x in np.asarray(camera1, np.float32).reshape(-1)]
        ex.rolling_shutter[:] = [0.0, 0.0, 0.0, 1.0]            # python_api.cu:181
        ex.parallax_shift[:] = [float(x) for x in self.parallax_shift]
        ex.near_distance = float(self.render_near_distance)
        ex.plane_z = float(self.slice_plane_z) + float(self.scale)  # testbed_nerf.cu:2693 (the Slice view's negated plane is a GUI debug mode)
        ex.aperture_size = float(self.aperture_size)
        ex.lens_mode = LENS_MODES[self.render_lens_mode] if lens else 0
        ex.lens_params[:] = [float(x) for x in self.render_lens_params] if lens else [0.0] * 7
        return ex

    def _sync_envmap(self):
        if self.envmap is self._envmap_uploaded:
            return
        if self.envmap is None:
            self._check(self.lib.snb_set_envmap(self.ctx, None, 0, 0), "snb_set_envmap")
        else:
            e = np.ascontiguousarray(self.envmap, dtype=np.float32)
            if e.ndim != 3 or e.shape[2] != 4:
                raise ValueError("envmap must be float32 [h, w, 4]")
            self._check(self.lib.snb_set_envmap(self.ctx, _fp(e), e.shape[1], e.shape[0]), "snb_set_envmap")
        self._envmap_uploaded = self.envmap

    def camera_log_lerp(self, a, b, t):
        """≙ camera_log_lerp (common_device.cuh:658-660) evaluated by the host library."""
        a = np.ascontiguousarray(a, np.float32).reshape(-1); b = np.ascontiguousarray(b, np.float32).reshape(-1)
        out = np.zeros(12, np.float32)
        if self.lib.snb_camera_log_lerp(_fp(a), _fp(b), float(t), _fp(out)) != 0:
            raise _lib.SnbError("snb_camera_log_lerp failed")
        return out.reshape(4, 3)

    def set_camera_from_time(self, t):
        """≙ Testbed::set_camera_from_time: no path, no change."""
        if self.camera_path is not None:
            self.camera_matrix = np.asarray(self.camera_path(float(t)), dtype=np.float32).reshape(4, 3).copy()

    def render_sample(self, width, height, spp_index, camera0, camera1, tile=None):
        """One sample of the stock render (render_frame -> render_nerf) with the shutter open from camera0 to camera1."""
        self._sync_envmap()
        ex = self._camera_extras(camera0, camera1)
        self._check(self.lib.snb_set_camera_ex(self.ctx, C.byref(ex) if ex is not None else None), "snb_set_camera_ex")
        saved = self.camera_matrix
        self.camera_matrix = np.asarray(camera0, dtype=np.float32).reshape(4, 3)
        try:
            return self.render_nerf_raw(width, height, spp_index=spp_index, tile=tile)
        finally:
            self.camera_matrix = saved

    def render(self, width=1920, height=1080, spp=1, linear=True, start_t=-1.0, end_t=-1.0, fps=30.0, shutter_fraction=1.0):
        """≙ pyngp.Testbed.render (src/python_api.cu:105-205): accumulate ``spp`` samples, tonemap, return float32 ndarray [H, W, 4].
        With start_t >= 0 the camera follows ``camera_path`` from start_t to end_t while the shutter is open for ``shutter_fraction`` of
        the frame: every sample's rays interpolate per pixel between the start camera and camera_log_lerp(start, end, shutter_fraction)."""
        if end_t < 0.0:
            end_t = start_t
        path_animation = start_t >= 0.0
        if not path_animation or self.smoothed_camera is None:
            self.smoothed_camera = np.asarray(self.camera_matrix, np.float32).copy()
        if start_t == 0.0:
            self.set_camera_from_time(start_t)
            self.smoothed_camera = self.camera_matrix.copy()
        start_cam = self.smoothed_camera.copy()
        if path_animation:
            self.set_camera_from_time(end_t)
            self.smoothed_camera = self.camera_matrix.copy()     # apply_camera_smoothing with m_camera_smoothing off (the default)
        end_cam = self.smoothed_camera.copy()
        n = width * height
        accum = torch.zeros((n, 4), dtype=torch.float32, device=self.device)
        out = torch.empty((n, 4), dtype=torch.float32, device=self.device)
        torch.cuda.current_stream(self.device).synchronize()   # the zero fill runs on torch's stream, the context on its own
        bg = np.asarray(self.background_color, dtype=np.float32)
        moving = not np.array_equal(start_cam, end_cam)
        sample_end = self.camera_log_lerp(start_cam, end_cam, shutter_fraction) if moving else start_cam
        for i in range(spp):
            if path_animation:
                a0, a1 = i / spp * shutter_fraction, (i + 1.0) / spp * shutter_fraction
                self.set_camera_from_time(start_t + (end_t - start_t) * (a0 + a1) / 2.0)
            fb, _ = self.render_sample(width, height, i, start_cam, sample_end)
            self._check(self.lib.snb_accumulate_tonemap(self.ctx, n, fb.data_ptr(), accum.data_ptr(), i, 0, 0 if linear else 1, float(self.exposure),
                                                        _fp(bg), self.tonemap_curve, out.data_ptr()), "snb_accumulate_tonemap")
        self._check(self.lib.snb_sync(self.ctx), "snb_sync")
        self._check(self.lib.snb_set_camera_ex(self.ctx, None), "snb_set_camera_ex")
        self.smoothed_camera = end_cam
        return out.reshape(height, width, 4).cpu().numpy()

    def counters(self):
        out = (C.c_uint64 * 4)()
        self._check(self.lib.snb_get_counters(self.ctx, out), "snb_get_counters")
        return {"samples": out[0], "rays": out[1], "hits": out[2], "reused": out[3]}

    def march_stats(self):
        out = (C.c_uint64 * 4)()
        self._check(self.lib.snb_get_march_stats(self.ctx, out), "snb_get_march_stats")
        return {"pool_iterations": out[0], "dda_steps": out[1], "search_passes": out[2], "ctas": out[3]}

    def ws_profile(self):
        """Pipeline balance of the last "ws" march as fractions of the warps' lifetime (snb_get_ws_profile)."""
        o = (C.c_uint64 * 8)()
        self._check(self.lib.snb_get_ws_profile(self.ctx, o), "snb_get_ws_profile")
        p, e = max(1, o[0]), max(1, o[6])
        return {"producer_wait_stage": o[1] / p, "producer_group_barrier": o[2] / p, "producer_search": o[3] / p, "producer_encode_publish": o[4] / p,
                "tiles_per_producer_warp": o[5] / 16.0, "epilogue_wait_tensor_core": o[7] / e, "producer_cycles_per_tile": o[0] / max(1, o[5])}

    def stage_ms(self):
        """Device time of the last frame's stages (synchronises): nerf_raygen, nerf_march, nerf_post, mesh, overlay."""
        out = (C.c_float * 5)()
        self._check(self.lib.snb_get_stage_ms(self.ctx, out), "snb_get_stage_ms")
        return dict(zip(("nerf_raygen", "nerf_march", "nerf_post", "mesh", "overlay"), [float(x) for x in out]))

    def launch_count(self):
        return int(self.lib.snb_launch_count(self.ctx))


def camera_path_from_keyframes(cameras):
    """Piecewise path through camera matrices: t in [0, 1] -> camera_log_lerp between the two neighbouring keyframes (the constant-speed
    stand-in for the reference's CameraPath spline, src/camera_path.cu, whose editor is GUI state)."""
    cams = [np.asarray(c, np.float32).reshape(4, 3) for c in cameras]
    lib = _lib.load()

    def path(t):
        if len(cams) == 1:
            return cams[0]
        x = min(max(float(t), 0.0), 1.0) * (len(cams) - 1)
        i = min(int(x), len(cams) - 2)
        out = np.zeros(12, np.float32)
        a = np.ascontiguousarray(cams[i]).reshape(-1); b = np.ascontiguousarray(cams[i + 1]).reshape(-1)
        if lib.snb_camera_log_lerp(_fp(a), _fp(b), x - i, _fp(out)) != 0:
            raise _lib.SnbError("snb_camera_log_lerp failed")
        return out.reshape(4, 3)
    return path


class Engine:
    """≙ sng::Engine minus GLFW/ImGui/GL: init(w, h, testbed), set_virtual_world(path), frame() -> final RGBA + depth."""

    def __init__(self):
        self.testbed = None
        self.scene = None
        self.rt = None
        self.width = self.height = 0
        self.frame_index = 0
        self.tile = None
        # CamPath state (include/synerfgine/cam_path.cuh:116-152): the camera follows the keyframe path while `is_playing`
        self.is_playing = False
        self.current_frame = 0
        self.path_frames = 0          # 0 = the scene's total_time_ms * fps / 1000; bench.py resamples the path to 24 / 64 frames
        self.enable_animations = False

    def set_virtual_world(self, config_fp, base_dir=None):
        self.scene = Scene.from_file(config_fp, base_dir)
        self.rt = self.scene.rt.copy()
        self.is_playing = self.scene.move_on_start and self.scene.n_keyframes >= 2
        self.enable_animations = self.scene.animations

    def set_virtual_world_json(self, json_text, base_dir=None):
        self.scene = Scene(json_text, base_dir)
        self.rt = self.scene.rt.copy()
        self.is_playing = self.scene.move_on_start and self.scene.n_keyframes >= 2
        self.enable_animations = self.scene.animations

    def init(self, res_width, res_height, testbed, tile=None):
        self.testbed, self.width, self.height, self.tile = testbed, res_width, res_height, tile
        tb, sc = testbed, self.scene
        if sc is None:
            raise RuntimeError("set_virtual_world first")
        if any(x != 0.0 for x in sc.view_dir):
            tb.set_camera_from_view(sc.view_dir, sc.at_pos, sc.zoom)
        tb.exposure = self.rt.exposure
        self.rt.tonemap_curve = tb.tonemap_curve
        tb._check(tb.lib.snb_set_scene(tb.ctx, sc.objects, sc.n_objects, sc.materials, sc.n_materials, sc.lights, sc.n_lights), "snb_set_scene")
        tb._check(tb.lib.snb_seed_rng(tb.ctx, 1999), "snb_seed_rng")
        # full-frame capacity once; a tile (which may change from frame to frame, tiles.StripBalancer) renders into a view of it
        self._rgba_store = torch.empty(res_width * res_height * 4, dtype=torch.float32, device=tb.device)
        self._depth_store = torch.empty(res_width * res_height, dtype=torch.float32, device=tb.device)
        self._bind_outputs(tile)
        if tile is not None:
            # size every grow-only scratch buffer of the context for the FULL frame once, so that moving the tile later
            # (tiles.StripBalancer) never reallocates (cudaFree is a device-wide sync and would stall in-flight NCCL traffic)
            keep, self.tile = self.tile, None
            self._bind_outputs(None)
            self.frame(sync=True)
            self.frame_index -= 1
            self.tile = keep
            self._bind_outputs(keep)
            tb._check(tb.lib.snb_seed_rng(tb.ctx, 1999), "snb_seed_rng")

    def _bind_outputs(self, tile):
        w, h = (tile.w, tile.h) if tile is not None else (self.width, self.height)
        self.out_rgba = self._rgba_store[:h * w * 4].view(h, w, 4)
        self.out_depth = self._depth_store[:h * w].view(h, w)

    def present(self, window_width=None, window_height=None, clear_color=(0.0, 0.0, 0.0), host=True):
        """≙ Display::present minus the window (display.cu:266-303): FXAA blit of the last frame into an RGB8 window image,
        rows top-down. Returns a uint8 numpy array [H, W, 3] (host=True) or a torch uint8 tensor on the device."""
        tb = self.testbed
        h, w = self.out_rgba.shape[0], self.out_rgba.shape[1]
        ow, oh = window_width or w, window_height or h
        cl = (C.c_float * 3)(*clear_color)
        if host:
            out = np.empty((oh, ow, 3), dtype=np.uint8)
            tb._check(tb.lib.snb_present_host(tb.ctx, self.out_rgba.data_ptr(), w, h, ow, oh, cl, None, None, out.ctypes.data), "snb_present_host")
            return out
        out = torch.empty((oh, ow, 3), dtype=torch.uint8, device=tb.device)
        tb._check(tb.lib.snb_present(tb.ctx, self.out_rgba.data_ptr(), w, h, ow, oh, cl, None, None, out.data_ptr()), "snb_present")
        return out

    def save_image(self, path, window_width=None, window_height=None):
        """≙ Display::save_image (display.cu:305-322): the presented window as an 8-bit RGB PNG."""
        img = self.present(window_width, window_height, host=True)
        tb = self.testbed
        tb._check(tb.lib.snb_write_png(str(path).encode(), img.ctypes.data, img.shape[1], img.shape[0]), "snb_write_png")
        return img

    def set_syn_samples(self, n):
        self.rt.shadow_iters = n

    def set_nerf_samples(self, n):
        self.rt.nerf_shadow_kernel = n

    def update_world_objects(self):
        """≙ Engine::update_world_objects (engine.cu:80-127): advance animations, re-upload the POD arrays."""
        tb, sc = self.testbed, self.scene
        sc.advance()
        tb._check(tb.lib.snb_update_scene(tb.ctx, sc.objects, sc.n_objects, sc.materials, sc.n_materials, sc.lights, sc.n_lights), "snb_update_scene")

    def advance_camera_path(self):
        """≙ CamPath::update -> advance_frame -> set_to_frame (cam_path.cuh:116-142), called by every Engine::frame (engine.cu:365):
        the next frame of the keyframe path, wrapping to frame 0 past the last keyframe; moves the testbed's camera."""
        sc, tb = self.scene, self.testbed
        if not self.is_playing or sc.n_keyframes < 2:
            return False
        total = self.path_frames or sc.total_frames
        between = max(1, total // max(sc.n_keyframes - 1, 1))
        self.current_frame += 1
        if self.current_frame // between + 1 >= sc.n_keyframes:
            self.current_frame = 0
        view, at, zoom = sc.camera_at(self.current_frame, total)
        tb.set_camera_from_view(view, at, zoom)
        tb._check(tb.lib.snb_reset_accumulation(tb.ctx), "snb_reset_accumulation")     # m_syn_camera_reset -> RayTracer::reset_accumulation (engine.cu:366-369)
        return True

    def skip_frames(self, n=1):
        """Advance the camera-path position and the animations by `n` frames without rendering or touching the device (frame-batch
        sharding: a rank walks past the frames other ranks render). The next advance_camera_path / update_world_objects pick up from there."""
        sc = self.scene
        if self.is_playing and sc.n_keyframes >= 2:
            total = self.path_frames or sc.total_frames
            between = max(1, total // max(sc.n_keyframes - 1, 1))
            for _ in range(n):
                self.current_frame += 1
                if self.current_frame // between + 1 >= sc.n_keyframes:
                    self.current_frame = 0
        if self.enable_animations:
            for _ in range(n):
                sc.lib.snb_scene_advance(sc.handle)        # host state only; Scene.refresh() runs with the next rendered frame's advance()

    def step(self, **kw):
        """One iteration of the reference's main loop, `while (engine.frame())` (src/main.cu:224): camera path, animations, render."""
        self.advance_camera_path()
        if self.enable_animations:
            self.update_world_objects()
        return self.frame(**kw)

    def frame(self, sync=True, host=False, tile=None):
        """Render with the current camera and world (Engine::frame from RayTracer::render on, engine.cu:374-406). `tile` (an _lib.Tile)
        replaces the tile given to init() from this frame on. The per-frame camera-path / animation update is step()."""
        tb = self.testbed
        if tile is not None and (self.tile is None or (tile.x0, tile.y0, tile.w, tile.h) != (self.tile.x0, self.tile.y0, self.tile.w, self.tile.h)):
            self.tile = tile
            self._bind_outputs(tile)
        v = tb.make_view(self.width, self.height, self.frame_index)
        tile_ref = C.byref(self.tile) if self.tile is not None else None
        if host:
            rgba = np.empty(tuple(self.out_rgba.shape), dtype=np.float32); depth = np.empty(tuple(self.out_depth.shape), dtype=np.float32)
            tb._check(tb.lib.snb_frame_host(tb.ctx, C.byref(v), tile_ref, C.byref(self.rt), rgba.ctypes.data, depth.ctypes.data), "snb_frame_host")
            self.frame_index += 1
            return rgba, depth
        tb._check(tb.lib.snb_frame(tb.ctx, C.byref(v), tile_ref, C.byref(self.rt), self.out_rgba.data_ptr(), self.out_depth.data_ptr()), "snb_frame")
        if sync:
            tb._check(tb.lib.snb_sync(tb.ctx), "snb_sync")
        self.frame_index += 1
        return self.out_rgba, self.out_depth
