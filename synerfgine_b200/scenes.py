"""Benchmark / test scenes for the BASELINE.json configs (SURVEY.md §8d "Concrete configs"), authored for the synthetic
model of synthetic.py in the reference's virtual-scene JSON schema (camera, rendering, materials, objfile, lights).

Meshes come from tests/data/meshes.npz (the reference's data/obj inputs converted by tests/data/make_meshes.py) and are
written back to OBJ files in a scratch directory so that the JSON -> OBJ loader -> BVH builder path is the one exercised.
"""
import json
import os
import tempfile

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MESHES_NPZ = os.path.join(_ROOT, "tests", "data", "meshes.npz")

_mesh_dir = None


def mesh_dir():
    """Scratch directory holding <name>.obj for every mesh of the fixture."""
    global _mesh_dir
    if _mesh_dir is not None:
        return _mesh_dir
    d = tempfile.mkdtemp(prefix="snb_meshes_")
    with np.load(MESHES_NPZ) as z:
        for name in z.files:
            write_obj(os.path.join(d, name + ".obj"), z[name])
    _mesh_dir = d
    return d


def write_obj(path, tris):
    """Triangle soup -> OBJ. %.9g round-trips float32 exactly."""
    with open(path, "w") as f:
        v = tris.reshape(-1, 3)
        for p in v:
            f.write("v %.9g %.9g %.9g\n" % (p[0], p[1], p[2]))
        for i in range(tris.shape[0]):
            f.write("f %d %d %d\n" % (3 * i + 1, 3 * i + 2, 3 * i + 3))


CAMERA = {"view": [-0.30, -0.45, -0.84], "at": [0.5, 0.42, 0.5], "zoom": 1.0}
ORBIT = [  # 10-keyframe orbit around the scene centre (same construction as dmrf-compare-abm.json's path)
    [-0.30, -0.45, -0.84], [-0.62, -0.45, -0.64], [-0.88, -0.45, -0.15], [-0.80, -0.45, 0.40], [-0.35, -0.45, 0.82],
    [0.25, -0.45, 0.86], [0.72, -0.45, 0.53], [0.89, -0.45, -0.05], [0.45, -0.45, -0.77], [-0.30, -0.45, -0.84]]
POINT_LIGHT = {"pos": [0.2, 1.6, 1.2], "size": 0.02, "intensity": 1.0, "type": "point"}
DIR_LIGHT = {"pos": [2.43, 1.465, 0.7], "size": 0.3, "intensity": 0.545, "type": "directional"}
LAMBERT = {"id": 0, "type": "lambertian", "n": 64.0, "rg": 0.5, "kd": [0.95, 0.3176, 0.1176]}
GLOSSY = {"id": 1, "type": "glossy", "n": 28.0, "rg": 0.2, "spec_angle": 0.001, "kd": [0.0, 1.0, 0.8]}


def _obj(name, pos, scale, material, rot=None, anim=None):
    o = {"file": name + ".obj", "pos": pos, "scale": scale, "material": material}
    if rot is not None:
        o["rot"] = rot
    if anim is not None:
        o["anim"] = anim
    return o


def scene_c2():
    """C2: one lambertian armadillo, one point light, nerf_shadow_samples 4, syn_shadow_samples 2, lens 0."""
    return {
        "camera": dict(CAMERA),
        "rendering": {"res_factor": 8, "exposure": 0.0, "path_trace_depth": 2, "light_samples": 2, "syn_shadow_samples": 2, "nerf_shadow_samples": 4,
                      "lens_size": 0.0, "nerf_on_nerf_shadow_threshold": 0.002, "nerf_shadow_intensity": 0.6, "syn_shadow_intensity": 1.0,
                      "shadow_on_nerf": True, "show_nerf": True, "shadow_on_virtual_obj": True, "syn_filter": "Final"},
        "materials": [dict(LAMBERT)],
        "objfile": [_obj("armadillo", [0.62, 0.44, 0.74], 0.6, 0, rot=[0.996, 0.0, -0.087, 0.0, 1.0, 0.0, 0.087, 0.0, 0.996])],
        "lights": [dict(POINT_LIGHT)],
    }


def scene_c3(animated=False):
    """C3/C4: armadillo + bunny + monkey, lambertian + glossy, point + directional light, depth 2, light_samples 4, lens > 0."""
    anim = (lambda: {"rot_center": [0.5, 0.4, 0.5], "rot_axis": [0.0, 1.0, 0.0], "rot_angle": 0.1}) if animated else (lambda: None)
    cam = dict(CAMERA)
    cam.update({"fps": 24, "total_time_ms": 1000, "animation_speed": 1.0 if animated else 0.0, "move_on_start": bool(animated),
                "path": [{"view": v, "at": CAMERA["at"], "zoom": CAMERA["zoom"]} for v in ORBIT]})
    return {
        "camera": cam,
        "rendering": {"res_factor": 8, "exposure": 0.002, "path_trace_depth": 2, "lens_size": 0.003, "light_samples": 4, "syn_shadow_samples": 2,
                      "nerf_shadow_samples": 4, "nerf_on_nerf_shadow_threshold": 0.002, "nerf_shadow_intensity": 0.6, "syn_shadow_intensity": 1.0,
                      "shadow_on_nerf": True, "show_nerf": True, "shadow_on_virtual_obj": True, "syn_filter": "Final"},
        "materials": [dict(LAMBERT), dict(GLOSSY)],
        "objfile": [
            _obj("armadillo", [0.62, 0.44, 0.74], 0.6, 0, rot=[0.996, 0.0, -0.087, 0.0, 1.0, 0.0, 0.087, 0.0, 0.996], anim=anim()),
            _obj("bunny", [0.30, 0.41, 0.30], 0.603, 1, rot=[1.0, 0.0, 0.008, 0.0, 1.0, 0.0, -0.008, 0.0, 1.0], anim=anim()),
            _obj("monkey", [0.78, 0.40, 0.20], 0.035, 0, rot=[-0.067, 0.0, -0.998, 0.0, 1.0, 0.0, 0.998, 0.0, -0.067], anim=anim()),
        ],
        "lights": [dict(POINT_LIGHT), dict(DIR_LIGHT)],
    }


def scene_json(name):
    d = {"c2": scene_c2, "c3": scene_c3, "c4": lambda: scene_c3(animated=True)}[name]()
    for o in d["objfile"]:
        for k in [k for k, v in o.items() if v is None]:
            del o[k]
    return json.dumps(d, indent=1)
