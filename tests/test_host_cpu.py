"""CPU tests (no GPU): the C-ABI library loads and exports everything include/snb.h declares, the host scene layer (JSON,
OBJ, BVH build, camera, animation) behaves like the reference's Engine / VirtualObject / CamPath, and the product refuses to
run without a CUDA device (no CPU fallback)."""
import ctypes as C
import json
import os
import re

import numpy as np
import pytest

import common
import oracle_lib as O
from synerfgine_b200 import _lib, scenes
from synerfgine_b200.scene import Scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ---- C ABI -----------------------------------------------------------------------------------------------------------
def declared_symbols():
    with open(os.path.join(ROOT, "include", "snb.h")) as f:
        text = re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)
    return sorted(set(re.findall(r"\b(snb_[a-z0-9_]+)\s*\(", text)))


def test_abi_exports_every_declared_symbol(snb):
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(snb, n), f"libsnb.so does not export {n} declared in include/snb.h"
    assert sorted(_lib.SIGNATURES) == names, "synerfgine_b200/_lib.py signatures and include/snb.h disagree"
    assert b"sm_100a" in snb.snb_version()


def test_struct_layouts_match_header():
    # sizes the C side static-asserts implicitly by use: a drift here corrupts every call
    assert C.sizeof(_lib.View) == 12 * 4 + 2 * 4 + 2 * 4 + 4 * 4
    assert C.sizeof(_lib.Tile) == 16
    assert C.sizeof(_lib.Material) == 13 * 4
    assert C.sizeof(_lib.Light) == 6 * 4
    assert C.sizeof(_lib.RtParams) == 17 * 4
    assert C.sizeof(_lib.Object) == 8 + 8 + 4 + 4 + 9 * 4 + 3 * 4 + 4 + 4
    assert C.sizeof(_lib.ModelDesc) == C.sizeof(O.ModelDesc)


def test_no_cpu_fallback(snb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    ctx = C.c_void_p()
    rc = snb.snb_create(0, None, C.byref(ctx))
    assert rc == -2 and not ctx.value, "without a CUDA device snb_create must fail with SNB_ERR_CUDA"
    from synerfgine_b200 import Testbed
    with pytest.raises(RuntimeError):
        Testbed(0)


def test_null_arguments_are_rejected(snb):
    assert snb.snb_create(0, None, None) == -1
    assert snb.snb_sync(None) == -1
    assert snb.snb_last_error(None) == b"null context"
    assert snb.snb_launch_count(None) == 0
    err = C.create_string_buffer(256)
    h = C.c_void_p()
    assert snb.snb_scene_load_json(None, None, C.byref(h), err, 256) != 0


# ---- scene JSON (Engine::set_virtual_world / Engine::init "rendering", engine.cu:21-78,129-234) ---------------------------
def test_scene_c2_fields():
    sc = Scene(scenes.scene_json("c2"), scenes.mesh_dir())
    assert (sc.n_objects, sc.n_materials, sc.n_lights) == (1, 1, 1)
    rt = sc.rt
    assert (rt.samples, rt.bounces, rt.shadow_iters, rt.nerf_shadow_kernel) == (2, 2, 2, 4)
    assert rt.n_steps == 8 and rt.lens_angle_constant == 0.0 and rt.syn_px_scale == 1           # res_factor 8 -> scale 1
    assert rt.show_nerf == 1 and rt.show_shadow_on_nerf == 1 and rt.view_nerf_shadow == 1
    assert abs(rt.nerf_shadow_intensity - 0.6) < 1e-7 and abs(rt.nerf_on_nerf_shadow_threshold - 0.002) < 1e-9
    m = sc.materials[0]
    assert m.type == 0 and list(m.kd) == pytest.approx([0.95, 0.3176, 0.1176]) and m.n == 64.0 and m.rg == 0.5
    li = sc.lights[0]
    assert li.type == 0 and list(li.pos) == pytest.approx([0.2, 1.6, 1.2]) and li.size == pytest.approx(0.02)
    o = sc.objects[0]
    assert o.n_tris > 500 and o.n_nodes > o.n_tris // 4 and o.mat_id == 0 and o.scale == pytest.approx(0.6)
    assert sc.view_dir == pytest.approx([-0.30, -0.45, -0.84]) and sc.zoom == 1.0


def test_scene_c3_materials_lights_and_path():
    sc = Scene(scenes.scene_json("c3"), scenes.mesh_dir())
    assert (sc.n_objects, sc.n_materials, sc.n_lights) == (3, 2, 2)
    assert sc.materials[1].type == 1 and sc.materials[1].spec_angle == pytest.approx(0.001)
    assert sc.lights[1].type == 1 and sc.lights[1].intensity == pytest.approx(0.545)
    assert sc.rt.samples == 4 and sc.rt.lens_angle_constant == pytest.approx(0.003)
    assert sc.n_keyframes == 10
    v0, a0, z0 = sc.camera_at(0, 24)
    assert v0 == pytest.approx(scenes.ORBIT[0]) and a0 == pytest.approx(scenes.CAMERA["at"])


@pytest.mark.parametrize("bad,frag", [
    ("{", "expected"), ('{"materials": [{"type": "lambertian", "kd": [1, 1, 1], "n": 1.0}], "objfile": [{"file": "nope.obj", "material": 0}], "lights": []}', "nope.obj"),
])
def test_scene_errors(snb, bad, frag):
    with pytest.raises(_lib.SnbError) as e:
        Scene(bad, scenes.mesh_dir())
    assert frag.lower() in str(e.value).lower()


def test_scene_bad_material_index_is_rejected():
    d = json.loads(scenes.scene_json("c2"))
    d["objfile"][0]["material"] = 7
    with pytest.raises(_lib.SnbError):
        Scene(json.dumps(d), scenes.mesh_dir())


# ---- OBJ loader + BVH builder (virtual_object.cu:38-88, triangle_bvh.cu:615-718) ----------------------------------------
def load_bvh(snb, name, per_leaf=4):
    path = os.path.join(scenes.mesh_dir(), name + ".obj").encode()
    n = C.c_int32()
    tris = np.zeros((20000, 9), np.float32)
    assert snb.snb_obj_load(path, tris.ctypes.data_as(_lib.pf32), 20000, C.byref(n)) == 0
    tris = np.ascontiguousarray(tris[:n.value])
    nodes = np.zeros((4 * n.value + 8, 8), np.float32)
    nn = C.c_int32()
    assert snb.snb_bvh_build(tris.ctypes.data_as(_lib.pf32), n.value, per_leaf, nodes.ctypes.data_as(_lib.pf32), nodes.shape[0], C.byref(nn)) == 0
    return tris, nodes[:nn.value]


@pytest.mark.parametrize("name", ["armadillo", "bunny", "monkey"])
def test_obj_loader_and_bvh_invariants(snb, name):
    with np.load(scenes.MESHES_NPZ) as z:
        src = z[name]
    tris, nodes = load_bvh(snb, name)
    assert tris.shape[0] == src.shape[0]
    # the builder only permutes triangles
    key = lambda t: sorted(map(bytes, np.ascontiguousarray(t.reshape(-1, 9))))
    assert key(tris) == key(src.astype(np.float32))
    idx = nodes[:, 6:8].copy().view(np.int32)
    covered = np.zeros(len(tris), np.int32)
    stack = [0]
    while stack:
        i = stack.pop()
        lo, hi = nodes[i, 0:3], nodes[i, 3:6]
        left, right = int(idx[i, 0]), int(idx[i, 1])
        if left < 0:                                     # leaf: triangles [-left-1, -right-1)
            a, b = -left - 1, -right - 1
            assert 0 < b - a <= 4
            covered[a:b] += 1
            pts = tris[a:b].reshape(-1, 3)
            assert (pts >= lo - 1e-6).all() and (pts <= hi + 1e-6).all()
        else:
            for c in range(left, right):
                assert (nodes[c, 0:3] >= lo - 1e-6).all() and (nodes[c, 3:6] <= hi + 1e-6).all()
                stack.append(c)
    assert (covered == 1).all(), "every triangle must sit in exactly one leaf"


def test_bvh_closest_hit_equals_brute_force(snb, orc):
    """BVH traversal (oracle restatement of ray_intersect_nodes) == brute force over all triangles with the reference's
    triangle test (pinned in test_oracle_pinned.py)."""
    tris, nodes = load_bvh(snb, "bunny")
    rng = np.random.default_rng(11)
    n = 400
    ro = rng.uniform(-1.5, 1.5, (n, 3)).astype(np.float32)
    tgt = tris.reshape(-1, 3)[rng.integers(0, len(tris) * 3, n)] + rng.normal(0, 0.05, (n, 3)).astype(np.float32)
    rd = (tgt - ro).astype(np.float32); rd /= np.linalg.norm(rd, axis=1, keepdims=True)
    rays = np.ascontiguousarray(np.concatenate([ro, rd.astype(np.float32)], axis=1), np.float32)
    obj = _lib.Object()
    obj.nodes = nodes.ctypes.data_as(_lib.pf32); obj.tris = tris.ctypes.data_as(_lib.pf32); obj.n_nodes, obj.n_tris = len(nodes), len(tris)
    obj.rot[:] = [1, 0, 0, 0, 1, 0, 0, 0, 1]; obj.pos[:] = [0, 0, 0]; obj.scale = 1.0; obj.mat_id = 0
    oid = np.zeros(n, np.int32); tid = np.zeros(n, np.int32); t = np.zeros(n, np.float32)
    orc.orc_trace_world((_lib.Object * 1)(obj), 1, O.P(rays), n, O.P(oid, O.pi32), O.P(tid, O.pi32), O.P(t))
    fn = orc.orc_triangle_ray_intersect; fn.restype = C.c_float; fn.argtypes = [O.pf, O.pf, O.pf]
    hits = 0
    for i in range(n):
        # depth_test_world offsets the origin by MIN_DEPTH along the ray (common.cu:36-67)
        o = (rays[i, :3] + rays[i, 3:] * np.float32(1e-5)).astype(np.float32)
        best, arg = np.float32(3.4e38), -1
        for k in range(len(tris)):
            tk = fn(O.P(tris[k]), O.P(o), O.P(np.ascontiguousarray(rays[i, 3:])))
            if tk < best:
                best, arg = tk, k
        if arg >= 0 and best < 16384.0:
            hits += 1
            assert tid[i] == arg or abs(float(t[i]) - float(best)) < 1e-6, (i, tid[i], arg, t[i], best)
        else:
            assert oid[i] < 0
        if hits > 60 and i > 120:
            break
    assert hits > 30


# ---- camera / animation ----------------------------------------------------------------------------------------------
def test_camera_from_view_is_orthonormal_and_looks_at_target():
    cam = common.camera_matrix(view=(0.3, -0.2, 0.9), at=(0.5, 0.4, 0.6), zoom=2.0).reshape(4, 3)
    R = cam[:3]
    np.testing.assert_allclose(R @ R.T, np.eye(3), atol=1e-5)
    fwd = np.array([0.3, -0.2, 0.9], np.float32); fwd /= np.linalg.norm(fwd)
    np.testing.assert_allclose(R[2], fwd, atol=1e-5)
    np.testing.assert_allclose(cam[3] + R[2] * 2.0, [0.5, 0.4, 0.6], atol=1e-5)      # origin = look_at - forward * scale


def test_camera_path_interpolates_keyframes():
    sc = Scene(scenes.scene_json("c4"), scenes.mesh_dir())
    v_first, _, _ = sc.camera_at(0, 24)
    v_last, _, _ = sc.camera_at(23, 24)
    assert v_first == pytest.approx(scenes.ORBIT[0], abs=1e-6)
    mids = [sc.camera_at(f, 24)[0] for f in range(24)]
    steps = [np.linalg.norm(np.subtract(mids[i + 1], mids[i])) for i in range(23)]
    assert max(steps) < 0.6 and len({tuple(np.round(m, 5)) for m in mids}) >= 10           # a path, not a jump; and it moves
    assert np.linalg.norm(np.subtract(v_last, scenes.ORBIT[0])) < 0.6


def test_object_animation_rotates_about_axis_and_light_ping_pongs():
    d = json.loads(scenes.scene_json("c4"))
    d["lights"][0]["anim"] = {"end": [0.4, 1.6, 1.2], "step": 0.25}
    sc = Scene(json.dumps(d), scenes.mesh_dir())
    p0 = np.array(sc.objects[0].pos[:]); r0 = np.array(sc.objects[0].rot[:]).reshape(3, 3)
    l0 = np.array(sc.lights[0].pos[:])
    sc.advance()
    p1 = np.array(sc.objects[0].pos[:]); r1 = np.array(sc.objects[0].rot[:]).reshape(3, 3)
    c = np.array([0.5, 0.4, 0.5])
    # rotation about +y through rot_center; the reference also applies the object's own (3-digit, not exactly orthonormal)
    # rot to the offset every frame (virtual_object.cuh:53-64), hence the loose radius tolerance
    assert abs(np.linalg.norm((p1 - c)[[0, 2]]) - np.linalg.norm((p0 - c)[[0, 2]])) < 1e-3
    assert abs(p1[1] - p0[1]) < 1e-6 and np.linalg.norm(p1 - p0) > 1e-3
    np.testing.assert_array_equal(r0, r1)        # next_frame moves pos only; the object's own rot is untouched (reference behaviour)
    seen = [l0[0]]
    for _ in range(9):
        sc.advance(); seen.append(sc.lights[0].pos[0])
    assert min(seen) >= 0.2 - 1e-5 and max(seen) <= 0.4 + 1e-5 and max(seen) > 0.35              # moves towards `end` and comes back
    assert any(seen[i + 1] < seen[i] for i in range(9))


# ---- model container ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kind", ["small16x2", "fox8x4", "f1", "f8", "big16x2"])
def test_param_count_matches_oracle_layout(orc, kind):
    m = common.model(kind)
    assert m.desc().n_params == orc.orc_n_params(m.n_levels, m.n_feat, m.log2_T, 16, C.c_float(m.per_level_scale))


# ---- present stage: oracle properties + the PNG writer (host code, no GPU) ---------------------------
def decode_png(path):
    """Minimal PNG reader for the test (8-bit RGB, filter types 0-4)."""
    import struct
    import zlib
    with open(path, "rb") as f:
        data = f.read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(data):
        n, typ = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        crc, = struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])
        assert zlib.crc32(typ + body) & 0xFFFFFFFF == crc, "chunk CRC"
        if typ == b"IHDR":
            w, h, depth, ctype, comp, flt, lace = struct.unpack(">IIBBBBB", body)
            assert (depth, ctype, comp, flt, lace) == (8, 2, 0, 0, 0)
        elif typ == b"IDAT":
            idat += body
        pos += 12 + n
    raw = zlib.decompress(idat)
    assert len(raw) == h * (w * 3 + 1)
    rows = np.frombuffer(raw, np.uint8).reshape(h, w * 3 + 1)
    assert (rows[:, 0] == 0).all()
    return rows[:, 1:].reshape(h, w, 3).copy()


@pytest.mark.parametrize("w,h", [(1, 1), (7, 5), (640, 360), (301, 233)])
def test_write_png_round_trip(snb, tmp_path, w, h):
    rng = np.random.default_rng(w * 1000 + h)
    img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    path = str(tmp_path / "output-001.png")
    assert snb.snb_write_png(path.encode(), img.ctypes.data, w, h) == 0
    np.testing.assert_array_equal(decode_png(path), img)
    assert snb.snb_write_png(path.encode(), None, w, h) == -1
    assert snb.snb_write_png(b"/nonexistent-dir/x.png", img.ctypes.data, w, h) == -4


def test_oracle_present_properties():
    rng = np.random.default_rng(11)
    h, w = 24, 40
    # (1) a constant image stays constant: every tap reads the same colour, blend with alpha 1 ignores the clear colour
    flat = np.empty((h, w, 4), np.float32); flat[...] = (0.25, 0.5, 0.75, 1.0)
    out = O.present(flat, w, h, clear=(1.0, 0.0, 0.0))
    assert (out == np.array([64, 128, 191], np.uint8)).all()            # floor(c * 255 + 0.5)
    # (2) alpha 0 everywhere -> premultiplied rgb 0 + clear colour
    clear = np.zeros((h, w, 4), np.float32)
    out = O.present(clear, w, h, clear=(0.2, 0.4, 0.6))
    assert (out == np.array([51, 102, 153], np.uint8)).all()
    # (3) a smooth gradient has no edges: FXAA's blend stays within one 8-bit step of the plain bilinear blit, rows are top-down
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    grad = np.stack([xx / w, yy / h, 0.5 * np.ones_like(xx), np.ones_like(xx)], axis=2).astype(np.float32)
    out = O.present(grad, w, h)
    core = out[2:-2, 2:-2].astype(int)                                  # GL_REPEAT wraps at the border
    expect = np.floor(grad[2:-2, 2:-2, :3] * 255 + 0.5).astype(int)
    assert np.abs(core - expect).max() <= 1
    assert out[2, 5, 1] < out[h - 3, 5, 1]                              # green grows downwards: row 0 is the top
    # (4) a straight axis-aligned edge is left alone (FXAA samples ALONG the edge), a diagonal staircase is softened, and only there
    edge = np.zeros((h, w, 4), np.float32); edge[..., 3] = 1.0; edge[:, w // 2:, :3] = 1.0
    out = O.present(edge, w, h)
    assert set(np.unique(out[3:-3, 3:-3])) == {0, 255}
    stair = np.zeros((h, w, 4), np.float32); stair[..., 3] = 1.0; stair[(xx + yy) > (w + h) / 2, :3] = 1.0
    out = O.present(stair, w, h)[3:-3, 3:-3, 0].astype(int)
    dist = np.abs((xx + yy) - (w + h) / 2)[3:-3, 3:-3]
    soft = (out > 0) & (out < 255)
    assert soft.any() and dist[soft].max() <= 3.0
    assert (out[dist > 4][(xx + yy)[3:-3, 3:-3][dist > 4] < (w + h) / 2] == 0).all()
    # (5) window larger than the frame: bilinear upscale keeps the range and the orientation
    img = rng.random((h, w, 4)).astype(np.float32); img[..., 3] = 1.0
    up = O.present(img, 2 * w, 2 * h)
    assert up.shape == (2 * h, 2 * w, 3) and up.min() >= 0
    # (6) identity warp passed explicitly == default
    ident = [0, 0, 0, 1, 0, 0, 0, 0, 0, 1, 0, 1]
    np.testing.assert_array_equal(O.present(img, w, h, warp_x=ident, warp_y=ident), O.present(img, w, h))


def test_engine_skip_frames_matches_walking_the_path():
    """Engine.skip_frames(n) (frame-batch sharding: a rank walks past the frames other ranks render) leaves the camera-path position and the
    animated objects where n plain iterations would, without touching a device."""
    from synerfgine_b200 import scenes
    from synerfgine_b200.testbed import Engine
    a, b = Engine(), Engine()
    for e in (a, b):
        e.set_virtual_world_json(scenes.scene_json("c4"), scenes.mesh_dir()); e.path_frames = 24
    assert a.is_playing and a.enable_animations
    a.skip_frames(7); a.skip_frames(20)
    sc = b.scene
    between = max(1, 24 // max(sc.n_keyframes - 1, 1))
    for _ in range(27):
        b.current_frame += 1
        if b.current_frame // between + 1 >= sc.n_keyframes:
            b.current_frame = 0
        sc.advance()
    a.scene.refresh()
    assert a.current_frame == b.current_frame
    for i in range(sc.n_objects):
        assert list(a.scene.objects[i].pos) == list(b.scene.objects[i].pos)
