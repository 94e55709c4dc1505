"""Pins the CPU oracle (oracle/, test infrastructure) against golden vectors produced by the REFERENCE's own
host-callable code: tests/golden/ref_host.npz was generated in the build container by tests/golden/make_ref_golden.py
through oracle/_ref/libref_shim.so, which is compiled from the headers under /root/reference (nerf_device.cuh,
random_val.cuh, common_device.cuh, triangle.cuh, bounding_box.cuh, tcnn common_device.h). The reference ships no tests
or golden vectors of its own, so these are "outputs of the reference itself run here".

Everything is compared bit for bit in math mode 0 (libm = what the reference headers do on the host). CPU only.
"""
import ctypes as C

import numpy as np
import pytest

import common
import oracle_lib as O

f32, u32 = C.c_float, C.c_uint32
pf = O.pf


@pytest.fixture(scope="module")
def g():
    import os
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_host.npz"))


@pytest.fixture(scope="module")
def lib():
    lib = O.load()
    for name, res, args in [
        ("orc_to_stepping_space", f32, [f32, f32]), ("orc_from_stepping_space", f32, [f32, f32]), ("orc_morton3D", u32, [u32, u32, u32]),
        ("orc_srgb_to_linear", f32, [f32]), ("orc_linear_to_srgb", f32, [f32]), ("orc_triangle_ray_intersect", f32, [pf, pf, pf]),
        ("orc_triangle_normal", None, [pf, pf]), ("orc_aabb_ray_intersect", None, [pf, pf, pf, pf, pf]),
    ]:
        fn = getattr(lib, name); fn.restype, fn.argtypes = res, args
    lib.orc_set_math_mode(0)
    yield lib
    lib.orc_set_math_mode(1)


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def P(a):
    return np.ascontiguousarray(a, np.float32).ctypes.data_as(pf)


@pytest.mark.parametrize("tag,cone", [("c", 1.0 / 256.0), ("z", 0.0)])
def test_stepping_functions(lib, g, tag, cone):
    t = g[f"step_t_{tag}"]
    np.testing.assert_array_equal(bits([lib.orc_calc_dt(float(x), cone) for x in t]), bits(g[f"calc_dt_{tag}"]))
    np.testing.assert_array_equal(bits([lib.orc_to_stepping_space(float(x), cone) for x in t]), bits(g[f"to_step_{tag}"]))
    np.testing.assert_array_equal(bits([lib.orc_from_stepping_space(float(x), cone) for x in g[f"from_n_{tag}"]]), bits(g[f"from_step_{tag}"]))
    np.testing.assert_array_equal(bits([lib.orc_advance_n_steps(float(x), cone, float(k)) for x, k in zip(t, g[f"adv_n_{tag}"])]), bits(g[f"advance_{tag}"]))


def test_cascade_indexing_and_morton(lib, g):
    pos, mips = g["cell_pos"], g["cell_mip"]
    got = np.array([lib.orc_cascaded_grid_idx_at(P(p), int(m)) for p, m in zip(pos, mips)], np.uint32)
    np.testing.assert_array_equal(got, g["cell_idx"])
    np.testing.assert_array_equal(np.array([lib.orc_mip_from_pos(P(p)) for p in pos], np.uint32), g["mip_from_pos"])
    np.testing.assert_array_equal(np.array([lib.orc_morton3D(int(a), int(b), int(c)) for a, b, c in g["morton_xyz"]], np.uint32), g["morton"])


def test_low_discrepancy_sampler(lib, g):
    np.testing.assert_array_equal(bits([lib.orc_ld_random_val(int(i), int(s)) for i, s in zip(g["ld_idx"], g["ld_seed"])]), bits(g["ld_val"]))
    off = np.zeros((64, 2), np.float32)
    for s in range(64):
        lib.orc_ld_random_pixel_offset(s, off[s].ctypes.data_as(pf))
    np.testing.assert_array_equal(bits(off), bits(g["ld_pixel_offset"]))


def test_srgb_transfer(lib, g):
    x = g["srgb_x"]
    np.testing.assert_array_equal(bits([lib.orc_srgb_to_linear(float(v)) for v in x]), bits(g["srgb_to_linear"]))
    np.testing.assert_array_equal(bits([lib.orc_linear_to_srgb(float(v)) for v in x]), bits(g["linear_to_srgb"]))


@pytest.mark.parametrize("kind", ["small16x2", "big16x2"])
def test_march_to_first_occupied_voxel(lib, g, kind):
    """if_unoccupied_advance_to_next_occupied_voxel through the synthetic occupancy fields (aabb 4 and aabb 16 cascades)."""
    m = common.model(kind)
    d = m.desc()
    o, dirs, t0 = g[f"march_{kind}_o"], g[f"march_{kind}_d"], g[f"march_{kind}_t0"]
    got = np.array([lib.orc_march_to_occupied(C.byref(d), P(a), P(b), float(c)) for a, b, c in zip(o, dirs, t0)], np.float32)
    ref = g[f"march_{kind}_t"]
    np.testing.assert_array_equal(bits(got), bits(ref))
    assert (ref < 16384.0).sum() > 100 and (ref >= 16384.0).sum() > 10      # both outcomes are exercised


def test_ray_generation(lib, g):
    """init_rays_with_payload_kernel_nerf's pinhole set-up: origin, direction, entry t, alive."""
    cams, meta, ref = g["ray_cams"], g["ray_meta"], g["ray_out"]
    m = common.model("small16x2")
    d = m.desc()
    d.render_aabb_min[:] = [-1.5] * 3; d.render_aabb_max[:] = [2.5] * 3
    at = 0
    for ci, w, h, spp, snap in meta:
        v = common.make_view(int(w), int(h), cam=cams[ci], spp=int(spp), snap=int(snap))
        f = float(np.float32(h * 1.0572))
        v.focal_length[:] = [f, f]
        out = np.zeros((h, w, 8), np.float32)
        for y in range(h):
            for x in range(w):
                lib.orc_init_ray(C.byref(v), C.byref(d), x, y, out[y, x].ctypes.data_as(pf))
        np.testing.assert_array_equal(bits(out.reshape(-1, 8)), bits(ref[at:at + w * h]))
        at += w * h
    assert at == len(ref)


def test_triangle_and_box_intersection(lib, g):
    tris, ro, rd = g["tri"], g["tri_ro"], g["tri_rd"]
    got = np.array([lib.orc_triangle_ray_intersect(P(a), P(b), P(c)) for a, b, c in zip(tris, ro, rd)], np.float32)
    np.testing.assert_array_equal(bits(got), bits(g["tri_t"]))
    assert (g["tri_t"] < 3e38).sum() > 100
    nrm = np.zeros((len(tris), 3), np.float32)
    for i in range(len(tris)):
        lib.orc_triangle_normal(P(tris[i]), nrm[i].ctypes.data_as(pf))
    np.testing.assert_array_equal(bits(nrm), bits(g["tri_normal"]))
    bb = np.zeros((len(tris), 2), np.float32)
    for i in range(len(tris)):
        lib.orc_aabb_ray_intersect(P(g["bb_lo"][i]), P(g["bb_hi"][i]), P(ro[i]), P(g["bb_rd"][i]), bb[i].ctypes.data_as(pf))
    np.testing.assert_array_equal(bits(bb), bits(g["bb_t"]))


def test_deterministic_math_mode_stays_close_to_libm(lib):
    """Mode 1 (the GPU-comparable transcendentals) must agree with libm to a few ulp on the march's operating range."""
    x = np.exp(np.random.default_rng(1).uniform(np.log(1e-4), np.log(60.0), 2000)).astype(np.float32)
    a = np.array([lib.orc_det_logf(float(v)) for v in x], np.float32)
    np.testing.assert_allclose(a, np.log(x.astype(np.float64)), rtol=3e-7, atol=2e-7)
    y = np.random.default_rng(2).uniform(-20, 6, 2000).astype(np.float32)
    b = np.array([lib.orc_det_expf(float(v)) for v in y], np.float32)
    np.testing.assert_allclose(b, np.exp(y.astype(np.float64)), rtol=4e-7)
