"""ctypes binding of oracle/_ref/libref_dev.so — TEST INFRASTRUCTURE ONLY (never imported by synerfgine_b200/).

The library is the REFERENCE's own CUDA code (tiny-cuda-nn + the kernels of src/testbed_nerf.cu, src/synerfgine/*.cu and
src/triangle_bvh.cu) compiled for sm_100a by oracle/Makefile.refdev from the sources where they lie under /root/reference;
see oracle/ref_dev_shim.cu. It needs a GPU (libcuda) to load; `available()` says whether it can be used here.
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "_ref", "libref_dev.so")
pf, pu32, pu16, pi32, pu8 = C.POINTER(C.c_float), C.POINTER(C.c_uint32), C.POINTER(C.c_uint16), C.POINTER(C.c_int32), C.POINTER(C.c_uint8)
u32, u64, f32, vp = C.c_uint32, C.c_uint64, C.c_float, C.c_void_p

_lib = None
_why = None


class RefView(C.Structure):
    _fields_ = [("camera", f32 * 12), ("focal_length", f32 * 2), ("screen_center", f32 * 2), ("spp", u32), ("snap_to_pixel_centers", C.c_int)]


class RefRt(C.Structure):
    _fields_ = [("n_steps", u32), ("samples", u32), ("bounces", u32), ("shadow_iters", u32), ("lens_angle_constant", f32), ("depth_offset", f32),
                ("syn_shadow_factor", f32), ("view_nerf_shadow", C.c_int), ("buffer_type", C.c_int), ("nerf_shadow_kernel", C.c_int),
                ("nerf_shadow_intensity", f32), ("nerf_on_nerf_shadow_threshold", f32), ("show_shadow_on_nerf", C.c_int), ("show_nerf", C.c_int),
                ("exposure", f32), ("tonemap_curve", C.c_int)]


def available():
    """(ok, reason). The library exists only where oracle/Makefile.refdev was run (this container) and loads only next to a CUDA driver."""
    try:
        load()
        return True, ""
    except Exception as e:  # noqa: BLE001
        return False, str(e)


def load():
    global _lib, _why
    if _lib is not None:
        return _lib
    if _why is not None:
        raise RuntimeError(_why)
    if not os.path.exists(LIB):
        _why = "oracle/_ref/libref_dev.so not built (make -C oracle -f Makefile.refdev, needs /root/reference)"
        raise RuntimeError(_why)
    try:
        lib = C.CDLL(LIB)
    except OSError as e:
        _why = f"libref_dev.so does not load: {e}"
        raise RuntimeError(_why)
    sig = {
        "refdev_last_error": (C.c_char_p, []),
        "refdev_init": (C.c_int, [C.c_int]),
        "refdev_model_create": (vp, [u32, u32, u32, u32, f32, pu16, u64, pu8, pf, pf, pf, u32, f32, C.c_int, C.c_int, f32]),
        "refdev_model_free": (None, [vp]),
        "refdev_padded_output_width": (u32, [vp]),
        "refdev_network_inference": (C.c_int, [vp, vp, u32, vp]),
        "refdev_grid_encode": (C.c_int, [vp, vp, u32, vp]),
        "refdev_time_network": (f32, [vp, vp, u32, vp, C.c_int, C.c_int]),
        "refdev_density_to_bitfield": (C.c_int, [pf, u32, pu8]),
        "refdev_scene_create": (vp, [u32, C.POINTER(pf), pi32, C.POINTER(pf), pi32, pf, pf, pf, pi32, u32, pf, u32, pf]),
        "refdev_scene_free": (None, [vp]),
        "refdev_engine_create": (vp, [C.c_int, C.c_int, C.c_int]),
        "refdev_engine_free": (None, [vp]),
        "refdev_engine_reset_accumulation": (None, [vp]),
        "refdev_engine_reseed": (C.c_int, [vp]),
        "refdev_frame": (C.c_int, [vp, vp, vp, C.POINTER(RefView), C.POINTER(RefRt), pf]),
        "refdev_render_nerf": (C.c_int, [vp, vp, vp, C.POINTER(RefView), C.POINTER(RefRt), C.c_int, C.c_int, pf]),
        "refdev_render_syn": (C.c_int, [vp, vp, vp, C.POINTER(RefView), C.POINTER(RefRt)]),
        "refdev_engine_read": (C.c_int, [vp, C.c_int, pf]),
        "refdev_engine_last_n_hit": (u32, [vp]),
        "refdev_init_rays": (C.c_int, [vp, C.POINTER(RefView), C.c_int, C.c_int, pf]),
        "refdev_init_rays_ex": (C.c_int, [vp, C.POINTER(RefView), C.c_int, C.c_int, vp, pf]),
        "refdev_engine_set_camera_ex": (C.c_int, [vp, vp, pf, C.c_int, C.c_int]),
        "refdev_march_samples": (C.c_int, [vp, pf, u32, u32, pf, pu32, pf]),
        "refdev_rand_states": (C.c_int, [u32, u32, pu32, pf]),
        "refdev_trace_rays": (C.c_int, [vp, pf, u32, pi32, pi32, pf]),
        "refdev_shadow_dda": (C.c_int, [vp, pf, u32, u32, pf]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    if lib.refdev_init(0) != 0:
        _why = "refdev_init: " + lib.refdev_last_error().decode()
        raise RuntimeError(_why)
    _lib = lib
    return lib


def _check(rc):
    if rc != 0:
        raise RuntimeError("reference device library: " + load().refdev_last_error().decode())


def P(a, t=pf):
    return a.ctypes.data_as(t)


class Model:
    """The reference's NerfNetwork<__half> + bitfield for a synthetic model (synerfgine_b200.synthetic.SyntheticModel)."""

    def __init__(self, m):
        lib = load()
        self.m = m
        params = np.ascontiguousarray(m.params.view(np.uint16))
        aabb = np.asarray(list(m.aabb_min) + list(m.aabb_max), np.float32)
        raabb = np.asarray(list(m.render_aabb_min) + list(m.render_aabb_max), np.float32)
        tl = np.asarray(m.render_aabb_to_local, np.float32)
        bits = np.ascontiguousarray(m.bitfield) if m.bitfield is not None else None
        self.h = lib.refdev_model_create(m.n_levels, m.n_feat, m.log2_T, m.base_res, m.per_level_scale, P(params, pu16), params.size,
                                         P(bits, pu8) if bits is not None else None, P(aabb), P(raabb), P(tl), m.max_cascade, m.cone_angle,
                                         m.rgb_activation, m.density_activation, m.min_transmittance)
        if not self.h:
            raise RuntimeError("refdev_model_create: " + lib.refdev_last_error().decode())
        self.out_width = lib.refdev_padded_output_width(self.h)

    def close(self):
        if self.h:
            load().refdev_model_free(self.h); self.h = None

    def network_inference(self, coords):
        """coords n x 7 float32 (host) -> rgb sigma n x 4 float16 and the full [16 x n] output."""
        import torch
        n = coords.shape[0]
        npad = (n + 255) // 256 * 256
        c = np.zeros((npad, 7), np.float32); c[:n] = coords
        dc = torch.from_numpy(c).cuda()
        out = torch.zeros((self.out_width, npad), dtype=torch.float16, device="cuda")
        _check(load().refdev_network_inference(self.h, dc.data_ptr(), npad, out.data_ptr()))
        o = out.cpu().numpy()
        return np.ascontiguousarray(o[:4, :n].T), o[:, :n]

    def grid_encode(self, coords):
        """coords n x 7 -> features n x 32 float16 (level-major, feature-minor like the product's AoS row)."""
        import torch
        n = coords.shape[0]
        npad = (n + 255) // 256 * 256
        c = np.zeros((npad, 7), np.float32); c[:n] = coords
        dc = torch.from_numpy(c).cuda()
        out = torch.zeros((32, npad), dtype=torch.float16, device="cuda")
        _check(load().refdev_grid_encode(self.h, dc.data_ptr(), npad, out.data_ptr()))
        return np.ascontiguousarray(out.cpu().numpy()[:, :n].T)

    def time_network(self, coords_dev, n, out_dev, reps=5, grid_only=False):
        ms = load().refdev_time_network(self.h, coords_dev, n, out_dev, reps, int(grid_only))
        if ms < 0:
            raise RuntimeError("refdev_time_network: " + load().refdev_last_error().decode())
        return ms

    def init_rays(self, view, w, h, ex=None):
        """Payloads after init_rays_from_camera: n x {origin3, dir3, t, alive}. view: RefView; ex: synerfgine_b200._lib.CameraEx (same layout) or None."""
        out = np.zeros((w * h, 8), np.float32)
        if ex is None:
            _check(load().refdev_init_rays(self.h, C.byref(view), w, h, P(out)))
        else:
            _check(load().refdev_init_rays_ex(self.h, C.byref(view), w, h, C.byref(ex), P(out)))
        return out

    def march_samples(self, rays7, n_steps):
        rays7 = np.ascontiguousarray(rays7, np.float32)
        n = rays7.shape[0]
        coords = np.zeros((n, n_steps, 7), np.float32); steps = np.zeros(n, np.uint32); t = np.zeros(n, np.float32)
        _check(load().refdev_march_samples(self.h, P(rays7), n, n_steps, P(coords), P(steps, pu32), P(t)))
        return coords, steps, t

    def shadow_dda(self, src_dst, n_steps=8):
        sd = np.ascontiguousarray(src_dst, np.float32)
        out = np.zeros(sd.shape[0], np.float32)
        _check(load().refdev_shadow_dda(self.h, P(sd), sd.shape[0], n_steps, P(out)))
        return out


class Scene:
    """Reference-side copy of a host Scene (synerfgine_b200.scene.Scene): ObjectTransform / Material / Light arrays on the device."""

    def __init__(self, sc):
        lib = load()
        n = sc.n_objects
        nodes = (pf * max(1, n))(); tris = (pf * max(1, n))()
        n_nodes = np.zeros(max(1, n), np.int32); n_tris = np.zeros(max(1, n), np.int32)
        rot = np.zeros((max(1, n), 9), np.float32); pos = np.zeros((max(1, n), 3), np.float32); scale = np.ones(max(1, n), np.float32); mat = np.zeros(max(1, n), np.int32)
        for i in range(n):
            o = sc.objects[i]
            nodes[i] = o.nodes; tris[i] = o.tris; n_nodes[i] = o.n_nodes; n_tris[i] = o.n_tris
            rot[i] = list(o.rot); pos[i] = list(o.pos); scale[i] = o.scale; mat[i] = o.mat_id
        mats = np.zeros((max(1, sc.n_materials), 13), np.float32)
        for i in range(sc.n_materials):
            m = sc.materials[i]
            mats[i] = list(m.ka) + list(m.kd) + list(m.ks) + [m.n, m.rg, float(m.type), m.spec_angle]
        lights = np.zeros((max(1, sc.n_lights), 6), np.float32)
        for i in range(sc.n_lights):
            l = sc.lights[i]
            lights[i] = list(l.pos) + [float(l.type), l.intensity, l.size]
        self.h = lib.refdev_scene_create(n, nodes, P(n_nodes, pi32), tris, P(n_tris, pi32), P(rot), P(pos), P(scale), P(mat, pi32), sc.n_materials, P(mats), sc.n_lights, P(lights))
        if not self.h:
            raise RuntimeError("refdev_scene_create: " + lib.refdev_last_error().decode())

    def close(self):
        if self.h:
            load().refdev_scene_free(self.h); self.h = None

    def trace_rays(self, rays6):
        r = np.ascontiguousarray(rays6, np.float32)
        n = r.shape[0]
        oid = np.zeros(n, np.int32); tid = np.zeros(n, np.int32); t = np.zeros(n, np.float32)
        _check(load().refdev_trace_rays(self.h, P(r), n, P(oid, pi32), P(tid, pi32), P(t)))
        return oid, tid, t


def ref_view(v):
    """synerfgine_b200._lib.View -> RefView."""
    r = RefView()
    r.camera[:] = list(v.camera); r.focal_length[:] = list(v.focal_length); r.screen_center[:] = list(v.screen_center)
    r.spp, r.snap_to_pixel_centers = v.spp, v.snap_to_pixel_centers
    return r


def ref_rt(rt):
    r = RefRt()
    for name, _ in RefRt._fields_:
        setattr(r, name, getattr(rt, name))
    return r


class Engine:
    """Per-resolution state of the reference's Engine + RayTracer (buffers, curand states, temporal blend)."""

    BUFS = {"final": (0, 4, True), "final_depth": (1, 1, True), "nerf_rgba": (2, 4, False), "nerf_depth": (3, 1, False), "positions": (4, 3, False),
            "normals": (5, 3, False), "syn_rgba": (6, 4, True), "syn_depth": (7, 1, True)}

    def __init__(self, w, h, syn_px_scale=1):
        self.w, self.h, self.scale = w, h, syn_px_scale
        self.h_ = load().refdev_engine_create(w, h, syn_px_scale)
        if not self.h_:
            raise RuntimeError("refdev_engine_create: " + load().refdev_last_error().decode())

    def close(self):
        if self.h_:
            load().refdev_engine_free(self.h_); self.h_ = None

    def reset(self, reseed=True):
        load().refdev_engine_reset_accumulation(self.h_)
        if reseed:
            _check(load().refdev_engine_reseed(self.h_))

    def frame(self, model, scene, view, rt):
        """Engine::frame minus GL. Returns stage ms [mesh, nerf, overlay, frame]."""
        ms = (f32 * 4)()
        rv, rr = ref_view(view), ref_rt(rt)
        _check(load().refdev_frame(self.h_, model.h, scene.h, C.byref(rv), C.byref(rr), ms))
        return list(ms)

    def render_nerf(self, model, view, scene=None, rt=None, alt=True, render_mode=1):
        ms = f32()
        rv = ref_view(view)
        rr = ref_rt(rt) if rt is not None else None
        _check(load().refdev_render_nerf(self.h_, model.h, scene.h if scene is not None else None, C.byref(rv), C.byref(rr) if rr is not None else None,
                                         int(alt), render_mode, C.byref(ms)))
        return ms.value

    def set_camera_ex(self, ex=None, envmap=None):
        """Camera extras (CameraEx or None) and environment map (float32 [h, w, 4] or None) of the following stock-path renders."""
        e = np.ascontiguousarray(envmap, np.float32) if envmap is not None else None
        _check(load().refdev_engine_set_camera_ex(self.h_, C.byref(ex) if ex is not None else None, P(e) if e is not None else None,
                                                  e.shape[1] if e is not None else 0, e.shape[0] if e is not None else 0))

    def read(self, name):
        which, ch, syn = self.BUFS[name]
        w, h = (self.w * self.scale, self.h * self.scale) if syn else (self.w, self.h)
        out = np.zeros((h, w, ch) if ch > 1 else (h, w), np.float32)
        _check(load().refdev_engine_read(self.h_, which, P(out)))
        return out

    def n_hit(self):
        return load().refdev_engine_last_n_hit(self.h_)


def rand_states(n, n_uniform=8):
    st = np.zeros((n, 6), np.uint32); un = np.zeros((n, max(1, n_uniform)), np.float32)
    _check(load().refdev_rand_states(n, n_uniform, P(st, pu32), P(un)))
    return st, un


def density_to_bitfield(density, max_cascade):
    d = np.ascontiguousarray(density, np.float32)
    out = np.zeros(8 * 128 ** 3 // 8, np.uint8)
    _check(load().refdev_density_to_bitfield(P(d), max_cascade, P(out, pu8)))
    return out
