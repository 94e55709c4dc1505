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
