"""SURVEY.md §8 f3 — BVH construction and refit on the device (synerfgine_b200/csrc/snb_bvh_dev.cuh).

The host builder (snb_bvh_build ≙ TriangleBvhWithBranchingFactor<2>::build, src/triangle_bvh.cu:615-718) stays the parity mode. The
device-built tree is validated (a) structurally, (b) by tracing it with the CPU oracle's traversal — ids and distances bit for bit, the
kernels read the same node format —, (c) against the host-built tree: closest distances equal, triangles equal through the build's
permutation (modulo exact ties on shared edges), and (d) through whole frames. Refit is checked against bounds recomputed in numpy and
by rendering a baked (vertex-animated) copy of C4's three animated objects against the rigid-transform animation the reference does."""
import ctypes as C

import numpy as np
import pytest

import common
import oracle_lib as O
from synerfgine_b200 import _lib, scenes

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


@pytest.fixture()
def ctx(snb):
    c = C.c_void_p()
    assert snb.snb_create(0, None, C.byref(c)) == 0
    yield c
    snb.snb_destroy(c)


def dev(shape, dtype=None):
    t = torch.zeros(shape, dtype=dtype or torch.float32, device="cuda")
    torch.cuda.synchronize()
    return t


def finite_psnr(a, b):
    """PSNR over the values finite on both sides, and the fraction of values NaN on one side only. (Under a directional light the reference's
    pow(negative mean, intensity) in shade_with_shadow yields NaN, src/testbed_nerf.cu:1675-1678,1778 — kept for parity; c3 / c4 have one.)"""
    ok = np.isfinite(a) & np.isfinite(b)
    return common.psnr(a[ok], b[ok]), float((np.isnan(a) != np.isnan(b)).mean())


def scene_of(name):
    from synerfgine_b200.testbed import Scene
    return Scene(scenes.scene_json(name), scenes.mesh_dir())


def object_arrays(ob):
    tris = np.ctypeslib.as_array(ob.tris, shape=(ob.n_tris, 9)).copy()
    nodes = np.ctypeslib.as_array(ob.nodes, shape=(ob.n_nodes, 8)).copy()
    return tris, nodes


def build_on_device(snb, ctx, tris, leaf=4):
    n = tris.shape[0]
    d_in = torch.from_numpy(tris).cuda(); torch.cuda.synchronize()
    cap = 2 * ((n + leaf - 1) // leaf) - 1
    d_out, d_nodes, d_order = dev((n, 9)), dev((cap, 8)), dev((n,), torch.int32)
    n_nodes, depth = C.c_int32(), C.c_int32()
    rc = snb.snb_bvh_build_dev(ctx, d_in.data_ptr(), n, leaf, d_out.data_ptr(), d_nodes.data_ptr(), cap, d_order.data_ptr(), C.byref(n_nodes), C.byref(depth))
    assert rc == 0, snb.snb_last_error(ctx)
    assert n_nodes.value == cap
    return d_out.cpu().numpy(), d_nodes.cpu().numpy(), d_order.cpu().numpy(), depth.value


def check_tree(tris, nodes, leaf):
    """Every triangle in exactly one leaf, at most `leaf` per leaf, every box the tight bound of what is below it."""
    ints = nodes.view(np.int32)
    covered = np.zeros(tris.shape[0], np.int32)
    v = tris.reshape(-1, 3, 3)

    def walk(i, depth):
        left, right = int(ints[i, 6]), int(ints[i, 7])
        if left < 0:
            b, e = -left - 1, -right - 1
            assert 0 < e - b <= leaf
            covered[b:e] += 1
            lo, hi = v[b:e].reshape(-1, 3).min(0), v[b:e].reshape(-1, 3).max(0)
            d = depth
        else:
            assert right == left + 2
            lo0, hi0, d0 = walk(left, depth + 1); lo1, hi1, d1 = walk(left + 1, depth + 1)
            lo, hi, d = np.minimum(lo0, lo1), np.maximum(hi0, hi1), max(d0, d1)
        np.testing.assert_array_equal(nodes[i, :3], lo); np.testing.assert_array_equal(nodes[i, 3:6], hi)
        return lo, hi, d

    import sys
    sys.setrecursionlimit(10000)
    _, _, depth = walk(0, 0)
    assert (covered == 1).all()
    return depth


def rays_at(tris, n, seed=7):
    rng = np.random.default_rng(seed)
    lo, hi = tris.reshape(-1, 3).min(0), tris.reshape(-1, 3).max(0)
    c, r = (lo + hi) / 2, np.linalg.norm(hi - lo) / 2
    o = c + rng.normal(size=(n, 3)) * r * 1.5
    tgt = tris.reshape(-1, 3, 3)[rng.integers(0, tris.shape[0], n)].mean(1) + rng.normal(size=(n, 3)) * r * 0.02
    d = tgt - o
    return np.concatenate([o, d / np.linalg.norm(d, axis=1, keepdims=True)], 1).astype(np.float32)


def as_object(tris, nodes, mat_id=0):
    ob = _lib.Object()
    ob.nodes = nodes.ctypes.data_as(_lib.pf32); ob.tris = tris.ctypes.data_as(_lib.pf32)
    ob.n_nodes, ob.n_tris = nodes.shape[0], tris.shape[0]
    ob.rot[:] = [1, 0, 0, 0, 1, 0, 0, 0, 1]; ob.pos[:] = [0, 0, 0]; ob.scale = 1.0; ob.mat_id = mat_id
    return ob


def trace(snb, ctx, sc, obs, rays):
    arr = (_lib.Object * len(obs))(*obs)
    assert snb.snb_set_option(ctx, b"bvh", b"reference") == 0
    assert snb.snb_set_scene(ctx, arr, len(obs), sc.materials, sc.n_materials, sc.lights, sc.n_lights) == 0, snb.snb_last_error(ctx)
    n = rays.shape[0]
    dr = torch.from_numpy(rays).cuda(); torch.cuda.synchronize()
    oid, tid, t = dev((n,), torch.int32), dev((n,), torch.int32), dev((n,))
    assert snb.snb_trace_rays(ctx, dr.data_ptr(), n, oid.data_ptr(), tid.data_ptr(), t.data_ptr()) == 0 and snb.snb_sync(ctx) == 0
    return oid.cpu().numpy(), tid.cpu().numpy(), t.cpu().numpy(), arr


@pytest.mark.parametrize("name,index,leaf", [("c2", 0, 4), ("c3", 1, 4), ("c3", 2, 1), ("c3", 0, 8)])
def test_device_built_tree_is_a_valid_bvh_and_traces_like_the_host_built_one(snb, orc, ctx, name, index, leaf):
    sc = scene_of(name)
    tris, nodes = object_arrays(sc.objects[index])
    d_tris, d_nodes, order, depth = build_on_device(snb, ctx, tris, leaf)
    # (a) structure
    assert sorted(order.tolist()) == list(range(tris.shape[0]))
    np.testing.assert_array_equal(d_tris, tris[order])
    assert check_tree(d_tris, d_nodes, leaf) == depth and depth <= 31
    # (b) the device tree traced by the kernels vs by the CPU oracle's traversal: bit for bit
    rays = rays_at(tris, 20000)
    oid, tid, t, arr = trace(snb, ctx, sc, [as_object(d_tris, d_nodes)], rays)
    n = rays.shape[0]
    roid, rtid, rt_ = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.float32)
    orc.orc_trace_world(arr, 1, O.P(rays), n, O.P(roid, O.pi32), O.P(rtid, O.pi32), O.P(rt_))
    np.testing.assert_array_equal(oid, roid); np.testing.assert_array_equal(tid, rtid); np.testing.assert_array_equal(t.view(np.uint32), rt_.view(np.uint32))
    # (c) vs the reference builder's tree: same closest distance; same triangle through the permutation, modulo exact ties
    hoid, htid, ht, _ = trace(snb, ctx, sc, [as_object(tris, nodes)], rays)
    assert (hoid >= 0).mean() > 0.5
    np.testing.assert_array_equal(oid, hoid)
    np.testing.assert_array_equal(t.view(np.uint32), ht.view(np.uint32))
    hit = hoid >= 0
    same = order[tid[hit]] == htid[hit]
    assert same.mean() > 0.995, same.mean()      # the rest: a ray through a shared edge, both triangles at the same t


def test_refit_follows_deforming_vertices(snb, orc, ctx):
    """snb_bvh_refit_dev on the reference builder's tree: after a non-rigid deformation every box is again the tight bound of what is
    below it (bit-equal to numpy's), and the refitted tree traces the deformed mesh like a tree built for it from scratch."""
    sc = scene_of("c3")
    tris, nodes = object_arrays(sc.objects[1])
    v = tris.reshape(-1, 3)
    bent = (v + 0.15 * np.sin(7.0 * v[:, [1, 2, 0]]) * np.float32([1.0, 0.5, 0.25])).astype(np.float32).reshape(-1, 9)
    d_tris = torch.from_numpy(bent).cuda(); d_nodes = torch.from_numpy(nodes).cuda(); torch.cuda.synchronize()
    assert snb.snb_bvh_refit_dev(ctx, d_tris.data_ptr(), d_nodes.data_ptr(), nodes.shape[0]) == 0 and snb.snb_sync(ctx) == 0
    refit = d_nodes.cpu().numpy()
    np.testing.assert_array_equal(refit.view(np.int32)[:, 6:], nodes.view(np.int32)[:, 6:])      # topology untouched
    check_tree(bent, refit, 4)
    assert not np.array_equal(refit[:, :6], nodes[:, :6])
    # a fresh host build for the deformed mesh (reorders its copy): same closest distances
    fresh_tris = bent.copy()
    cap = 2 * bent.shape[0]
    fresh_nodes = np.zeros((cap, 8), np.float32); nn = C.c_int32()
    assert snb.snb_bvh_build(O.P(fresh_tris), bent.shape[0], 4, O.P(fresh_nodes), cap, C.byref(nn)) == 0
    rays = rays_at(bent, 20000, seed=9)
    _, _, t_refit, _ = trace(snb, ctx, sc, [as_object(bent, refit)], rays)
    oid, _, t_fresh, _ = trace(snb, ctx, sc, [as_object(fresh_tris, np.ascontiguousarray(fresh_nodes[:nn.value]))], rays)
    assert (oid >= 0).mean() > 0.5
    np.testing.assert_array_equal(t_refit.view(np.uint32), t_fresh.view(np.uint32))


def _engine(name, w, h, lbvh):
    from synerfgine_b200.testbed import Engine, Testbed
    tb = Testbed()
    tb.set_model(common.model("small16x2"))
    assert tb.lib.snb_set_option(tb.ctx, b"bvh", b"lbvh" if lbvh else b"reference") == 0
    eng = Engine()
    eng.set_virtual_world_json(scenes.scene_json(name), scenes.mesh_dir())
    eng.init(w, h, tb)
    return tb, eng


@pytest.mark.parametrize("name", ["c2", "c3"])
def test_frames_with_device_built_trees_match_the_reference_trees(snb, name):
    """Whole frames (mesh layer, both shadow directions, overlay) with option bvh=lbvh vs the reference builder's trees: the closest hits
    are the same, so the frames are — up to rays through shared edges, where the other triangle's normal is taken (counted)."""
    w, h = 192, 128
    out = []
    for lbvh in (False, True):
        tb, eng = _engine(name, w, h, lbvh)
        try:
            for _ in range(2):
                rgba, depth = eng.frame()
            out.append((rgba.cpu().numpy().copy(), depth.cpu().numpy().copy()))
        finally:
            tb.close()
    (a, da), (b, db) = out
    diff = np.nan_to_num(np.abs(a - b), nan=0.0).max(-1)
    mesh = (da < 100) | (db < 100)
    assert mesh.mean() > 0.02
    ps, nan_mismatch = finite_psnr(a, b)
    assert ps >= 50.0 and nan_mismatch < 1e-3 and (diff > 1e-3).mean() < 2e-3, (ps, nan_mismatch, (diff > 1e-3).mean())
    np.testing.assert_array_equal(da, db)        # same closest hits -> the same depth, bit for bit (background included)


def test_vertex_animation_through_refit_matches_the_rigid_animation(snb):
    """C4's three animated objects. The reference animates the rigid transform (VirtualObject::next_frame, virtual_object.cuh:53-64) and
    takes rays into object space; here the same motion is BAKED into world-space vertices, uploaded every frame with
    snb_update_object_triangles and followed by the device refit. Frames agree up to the rounding of baking (world-space vs object-space
    intersection arithmetic): >= 40 dB, depth within 1e-3 on all but a counted fraction of the mesh's silhouette pixels."""
    w, h, n_frames = 192, 128, 4
    tb_a, eng_a = _engine("c4", w, h, False)
    tb_b, eng_b = _engine("c4", w, h, True)
    try:
        sc = eng_b.scene
        obj_tris = [object_arrays(sc.objects[i])[0] for i in range(sc.n_objects)]

        def baked(i):
            ob = sc.objects[i]
            rot = np.float32(list(ob.rot)).reshape(3, 3).T          # column-major
            p = obj_tris[i].reshape(-1, 3) @ rot.T * np.float32(ob.scale) + np.float32(list(ob.pos))
            return np.ascontiguousarray(p.reshape(-1, 9), np.float32)

        def upload_baked(first):
            obs = []
            keep = []
            for i in range(sc.n_objects):
                t = baked(i); keep.append(t)
                if first:
                    o = as_object(t, np.zeros((1, 8), np.float32), sc.objects[i].mat_id); obs.append(o)
                else:
                    assert snb.snb_update_object_triangles(tb_b.ctx, i, O.P(t)) == 0, snb.snb_last_error(tb_b.ctx)
            if first:
                arr = (_lib.Object * len(obs))(*obs)
                assert snb.snb_set_scene(tb_b.ctx, arr, len(obs), sc.materials, sc.n_materials, sc.lights, sc.n_lights) == 0, snb.snb_last_error(tb_b.ctx)
                assert snb.snb_seed_rng(tb_b.ctx, 1999) == 0

        upload_baked(True)
        worst = 1e9; moved = 0.0; prev = None; stats = []
        for f in range(n_frames):
            if f:
                eng_a.update_world_objects()
                sc.advance()                       # same animation step for the baked copy
                upload_baked(False)
            a, da = eng_a.frame(); b, db = eng_b.frame()
            a, da, b, db = a.cpu().numpy(), da.cpu().numpy(), b.cpu().numpy(), db.cpu().numpy()
            ps, nan_mismatch = finite_psnr(a, b)
            worst = min(worst, ps)
            assert nan_mismatch < 5e-3, (f, nan_mismatch)
            near = (da < 100) | (db < 100)
            edge = (np.abs(da - db)[near] > 1e-3).sum() / near.sum()
            stats.append((f, round(ps, 1), int(near.sum()), float(edge)))
            assert edge < 0.1, stats          # silhouette pixels whose hit / miss flips under the rounding of baking
            if prev is not None:
                moved = max(moved, float(np.nanmean(np.abs(a - prev))))
            prev = a.copy()
        assert moved > 1e-4, "the objects did not move"
        assert worst >= 40.0, stats
    finally:
        tb_a.close(); tb_b.close()
