"""Seeded workloads of the BASELINE.json configurations (SURVEY.md §8d "Concrete configs"): models, views, scenes.

Used by bench.py, the tests and __graft_entry__.smoke(); nothing here touches CUDA.

    C1  256x256, NeRF only, stock render_nerf path (runs in the reference's own build; the CPU oracle replays it)
    C2  1920x1080, armadillo lambertian + 1 point light, K = nerf_shadow_samples 4, S = syn_shadow_samples 2   <- the headline
    C3  3840x2160, 3 meshes, glossy + lambertian, point + directional light, depth 2, light_samples 4, lens_size > 0
    C4  1920x1080, aabb_scale 16 with a 2^22 table, 3 animated objects, 24-frame camera path (mean and p95 ms/frame)
    C5  7680x4320 single frame (c5-8k) and 64 x 1920x1080 frames (c5-batch), both shadow directions, sharded over the GPUs
"""
import math

import numpy as np

from .synthetic import SyntheticModel

MODEL_PRESETS = {
    "fox16x2": dict(n_levels=16, n_feat=2, log2_T=19, aabb_scale=4),       # BASELINE.json's architecture (C1-C3, C5)
    "fox8x4": dict(n_levels=8, n_feat=4, log2_T=19, aabb_scale=4),         # this fork's configs/nerf/base.json
    "unbounded22": dict(n_levels=16, n_feat=2, log2_T=22, aabb_scale=16),  # C4: 2^22 table (268 MB of features, larger than the L2; measured: the touched slice still fits it, L2 hit 96 %)
    "small16x2": dict(n_levels=16, n_feat=2, log2_T=15, aabb_scale=4),
    "unit16x2": dict(n_levels=16, n_feat=2, log2_T=15, aabb_scale=1),
    "big16x2": dict(n_levels=16, n_feat=2, log2_T=16, aabb_scale=16),
    "f1": dict(n_levels=32, n_feat=1, log2_T=14, aabb_scale=2),
    "f8": dict(n_levels=4, n_feat=8, log2_T=14, aabb_scale=2),
    "noocc": dict(n_levels=16, n_feat=2, log2_T=14, aabb_scale=1, occupancy=False, density_gain=12.0),
}
_models = {}


def model(kind="fox16x2"):
    """Cached synthetic model of a preset."""
    if kind not in _models:
        _models[kind] = SyntheticModel(**MODEL_PRESETS[kind])
    return _models[kind]


def camera_matrix(view=(-0.30, -0.45, -0.84), at=(0.5, 0.42, 0.5), zoom=1.0):
    """Testbed::reset_camera + set_view_dir + set_look_at + set_scale as Engine::init applies the JSON camera block."""
    from . import _lib
    lib = _lib.load()
    out = np.zeros(12, np.float32)
    v = np.asarray(view, np.float32); a = np.asarray(at, np.float32)
    assert lib.snb_camera_from_view(v.ctypes.data_as(_lib.pf32), a.ctypes.data_as(_lib.pf32), float(zoom), out.ctypes.data_as(_lib.pf32)) == 0
    return out


def make_view(w, h, cam=None, spp=0, snap=1, fov=50.625):
    """snb_view of a w x h frame: focal length = calc_focal_length(resolution, relative_focal_length (1, 1), fov_axis 1, zoom 1)."""
    from ._lib import View
    v = View()
    cam = camera_matrix() if cam is None else cam
    v.camera[:] = [float(x) for x in cam]
    f = float(np.float32(0.5 / math.tan(0.5 * fov * math.pi / 180.0)) * np.float32(h))
    v.focal_length[:] = [f, f]
    v.screen_center[:] = [0.5, 0.5]
    v.width, v.height, v.spp, v.snap_to_pixel_centers = w, h, spp, snap
    return v


# name -> what bench.py runs. `frames` > 1: the camera follows the scene's keyframe path and animated objects advance every frame
# (Engine::frame, src/synerfgine/engine.cu:365,372). `zoom`: camera distance for the unbounded model.
CONFIGS = {
    "c1": dict(label="C1", model="fox16x2", scene=None, width=256, height=256, frames=1, stock=True,
               workload="C1: fox-architecture NeRF (16x2 hash grid, T=2^19, 64-wide MLPs) 256x256, no virtual objects, 1 spp (stock render_nerf path)"),
    "c2": dict(label="C2", model="fox16x2", scene="c2", width=1920, height=1080, frames=1,
               workload="C2: 1920x1080 fox-architecture NeRF (16x2 hash grid, T=2^19, 64-wide MLPs) + armadillo.obj lambertian + 1 point light, "
                        "nerf_shadow_samples=4, syn_shadow_samples=2"),
    "c3": dict(label="C3", model="fox16x2", scene="c3", width=3840, height=2160, frames=1,
               workload="C3: 3840x2160, armadillo + bunny + monkey, glossy + lambertian, point + directional light, path_trace_depth=2, light_samples=4, lens_size=0.003"),
    "c4": dict(label="C4", model="unbounded22", scene="c4", width=1920, height=1080, frames=24, zoom=2.5,
               workload="C4: 1920x1080, unbounded scene aabb_scale=16 with a 2^22 hash table, 5 occupancy cascades, 3 animated OBJ objects, 24-frame camera path"),
    "c5-8k": dict(label="C5 (8K frame)", model="fox16x2", scene="c3", width=7680, height=4320, frames=1,
                  workload="C5: one 7680x4320 frame tile-sharded over the GPUs, NeRF-on-mesh and mesh-on-NeRF shadows"),
    "c5-batch": dict(label="C5 (frame batch)", model="fox16x2", scene="c4", width=1920, height=1080, frames=64, batch=True,
                     workload="C5: 64-frame 1920x1080 camera-path batch, frames sharded over the GPUs, NeRF-on-mesh and mesh-on-NeRF shadows"),
}
