"""Converts the reference's committed OBJ inputs (data/obj/*.obj) into tests/data/meshes.npz (float32 triangle soups,
n x 9, in file order, quads split like tinyobjloader does). Run in the build container where /root/reference exists:
    python tests/data/make_meshes.py
The GPU box has no /root/reference; tests and bench.py read the committed npz and write OBJ files from it on demand."""
import ctypes as C
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from synerfgine_b200 import _lib  # noqa: E402

REF = "/root/reference/data/obj"
NAMES = ["armadillo", "bunny", "monkey", "rock", "box", "shortbox", "smallbox", "plane", "aabb"]


def load_obj(lib, path):
    n = C.c_int32()
    assert lib.snb_obj_load(path.encode(), None, 0, C.byref(n)) == 0, path
    tris = np.zeros((n.value, 9), dtype=np.float32)
    assert lib.snb_obj_load(path.encode(), tris.ctypes.data_as(_lib.pf32), n.value, C.byref(n)) == 0
    return tris


if __name__ == "__main__":
    lib = _lib.load()
    out = {}
    for name in NAMES:
        t = load_obj(lib, os.path.join(REF, name + ".obj"))
        out[name] = t
        v = t.reshape(-1, 3)
        print(f"{name}: {t.shape[0]} tris, bounds {v.min(0)} .. {v.max(0)}")
    np.savez_compressed(os.path.join(ROOT, "tests", "data", "meshes.npz"), **out)
