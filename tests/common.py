"""Shared helpers for the parity tests."""
import ctypes as C
import math

import numpy as np

from synerfgine_b200 import _lib
from synerfgine_b200._lib import View
from synerfgine_b200.synthetic import SyntheticModel

from synerfgine_b200 import workloads
from synerfgine_b200.workloads import camera_matrix, make_view, model  # noqa: F401  (the tests' names for the shared workload helpers)

_models = {}


def cropped_model():
    """small16x2 with the GUI's crop box applied: a smaller render_aabb and a rotated render_aabb_to_local (the general
    m_render_aabb.contains(m_render_aabb_to_local * pos) path; every other model here has the identity)."""
    import copy
    import math
    if "cropped" not in _models:
        m = copy.copy(model("small16x2"))
        a = math.radians(25.0)
        c, s_ = np.float32(math.cos(a)), np.float32(math.sin(a))
        m.render_aabb_to_local = [float(c), float(s_), 0.0, float(-s_), float(c), 0.0, 0.0, 0.0, 1.0]      # column-major rotation about z
        m.render_aabb_min, m.render_aabb_max = [-0.2, -0.4, -0.3], [1.0, 0.75, 1.2]
        _models["cropped"] = m
    return _models["cropped"]


def psnr(a, b):
    mse = float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2))
    return 99.0 if mse == 0 else 10.0 * math.log10(1.0 / mse)


def default_rt(**kw):
    """RayTracer / Engine header defaults; keyword overrides."""
    from synerfgine_b200._lib import RtParams
    rt = RtParams()
    rt.n_steps, rt.samples, rt.bounces, rt.shadow_iters = 8, 2, 2, 4
    rt.lens_angle_constant, rt.depth_offset, rt.syn_shadow_factor = 0.009, 0.1, 1.0
    rt.view_nerf_shadow, rt.buffer_type, rt.nerf_shadow_kernel = 1, 0, 1
    rt.nerf_shadow_intensity, rt.nerf_on_nerf_shadow_threshold = 2.0, 0.3
    rt.show_shadow_on_nerf, rt.show_nerf, rt.syn_px_scale, rt.exposure, rt.tonemap_curve = 1, 1, 1, 0.0, 0
    for k, v in kw.items():
        assert hasattr(rt, k), k
        setattr(rt, k, v)
    return rt


def use_device_grid_scales(snb, ctx):
    """tcnn's kernel_grid evaluates grid_scale on the device (exp2f = MUFU.EX2, one FFMA); the product does the same once per model
    (k_grid_scales). The oracle cannot restate a hardware function, so it is handed those <= 32 floats for bit-exact feature comparisons;
    their correctness is pinned by the reference's own kernel_grid (tests/test_gpu_refdev.py::test_hash_grid_features_bit_exact)."""
    import oracle_lib
    sc = (C.c_float * 32)()
    n = snb.snb_get_grid_scales(ctx, sc, 32)
    assert n > 0, snb.snb_last_error(ctx)
    arr = np.asarray(list(sc)[:n], np.float32)
    oracle_lib.load().orc_set_grid_scales(arr.ctypes.data_as(C.POINTER(C.c_float)), n)
    return arr


# ---- the stock render's camera extras (snb_camera_ex / orc_camera_ex / the reference's uv_to_ray arguments) ----------------------------
def camera_ex(camera0, camera1=None, rolling=(0.0, 0.0, 0.0, 1.0), parallax=(0.0, 0.0, 0.0), near=0.0, plane_z=1.5, aperture=0.0, lens="perspective", lens_params=()):
    ex = _lib.CameraEx()
    c1 = camera0 if camera1 is None else camera1
    ex.camera1[:] = [float(x) for x in np.asarray(c1, np.float32).reshape(-1)]
    ex.rolling_shutter[:] = [float(x) for x in rolling]
    ex.parallax_shift[:] = [float(x) for x in parallax]
    ex.near_distance, ex.plane_z, ex.aperture_size = float(near), float(plane_z), float(aperture)
    ex.lens_mode = _lib.LENS_MODES[lens]
    p = list(lens_params) + [0.0] * (7 - len(lens_params))
    ex.lens_params[:] = [float(x) for x in p]
    return ex


def camera_ex_cases(w, h):
    """name -> CameraEx for a w x h view from camera_matrix(): every branch of uv_to_ray the stock render can take."""
    cam0 = camera_matrix()
    cam1 = camera_matrix(view=(-0.27, -0.46, -0.85), at=(0.52, 0.43, 0.5), zoom=1.02)
    return {
        "identity": camera_ex(cam0),
        "dof": camera_ex(cam0, aperture=0.02, plane_z=1.1),
        "motion_blur": camera_ex(cam0, cam1),
        "rolling_shutter": camera_ex(cam0, cam1, rolling=(0.1, 0.2, 0.6, 0.1)),
        "parallax_near": camera_ex(cam0, parallax=(0.03, -0.02, 0.8), near=0.25),
        "opencv": camera_ex(cam0, lens="opencv", lens_params=(-0.12, 0.03, 0.002, -0.001)),
        "opencv_fisheye": camera_ex(cam0, lens="opencv_fisheye", lens_params=(0.05, -0.01, 0.002, 0.0)),
        # f-theta: angle = polynomial in the pixel radius of a params[5] x params[6] sensor
        "ftheta": camera_ex(cam0, lens="ftheta", lens_params=(0.0, 3.0e-3, 1.0e-8, 0.0, 0.0, 1000.0, 700.0)),   # corners beyond 90 degrees: invalid rays
        "latlong": camera_ex(cam0, lens="latlong"),
        "equirectangular": camera_ex(cam0, lens="equirectangular"),
        "slice_plane": camera_ex(cam0, plane_z=-1.2, aperture=0.05),
        "everything": camera_ex(cam0, cam1, parallax=(0.01, 0.01, 0.5), near=0.1, plane_z=1.2, aperture=0.01, lens="opencv", lens_params=(-0.08, 0.01, 0.001, 0.001)),
    }


def test_envmap(w=64, h=32, seed=3):
    """Smooth RGBA environment map, float32 [h, w, 4], alpha 1."""
    rng = np.random.default_rng(seed)
    yy, xx = np.meshgrid(np.linspace(0, 1, h, dtype=np.float32), np.linspace(0, 1, w, dtype=np.float32), indexing="ij")
    e = np.stack([0.5 + 0.5 * np.sin(6.0 * xx + 1.0), 0.3 + 0.7 * yy, 0.5 + 0.5 * np.cos(9.0 * xx * yy), np.ones_like(xx)], axis=-1).astype(np.float32)
    return np.ascontiguousarray(e + rng.uniform(0, 0.05, e.shape).astype(np.float32) * np.array([1, 1, 1, 0], np.float32))


test_envmap.__test__ = False
