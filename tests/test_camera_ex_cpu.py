"""Host logic and oracle properties of the stock render's camera extras: camera_log_lerp, the general uv_to_ray
restatement (lens models, thin-lens aperture, parallax, near plane, rolling shutter) and the environment-map lookup. No GPU."""
import ctypes as C

import numpy as np
import pytest

import common
import oracle_lib as O
from synerfgine_b200 import _lib


def _m4(c):
    m = np.eye(4); m[:3, :] = np.asarray(c, np.float64).reshape(4, 3).T
    return m


def test_camera_log_lerp_matches_matrix_log_exp(snb):
    """snb_camera_log_lerp = exp(log(B A^-1) t) A (common_device.cuh:658-660) against scipy's logm / expm in float64."""
    sl = pytest.importorskip("scipy.linalg")
    a = common.camera_matrix(); b = common.camera_matrix(view=(-0.2, -0.5, -0.8), at=(0.52, 0.4, 0.5), zoom=1.2)
    out = np.zeros(12, np.float32)
    for t in (0.0, 0.25, 0.5, 1.0):
        assert snb.snb_camera_log_lerp(O.P(a), O.P(b), t, O.P(out)) == 0
        ref = sl.expm(sl.logm(_m4(b) @ np.linalg.inv(_m4(a))) * t) @ _m4(a)
        np.testing.assert_allclose(out.reshape(4, 3).T, ref[:3, :].real, atol=5e-6)
    assert snb.snb_camera_log_lerp(O.P(a), O.P(a), 0.7, O.P(out)) == 0
    np.testing.assert_allclose(out, a, atol=1e-6)
    assert snb.snb_camera_log_lerp(None, O.P(b), 0.5, O.P(out)) != 0


def _rays(orc, m, v, ex=None, envmap=None):
    O.set_camera_ex(ex, envmap)
    try:
        d = m.desc()
        out = np.zeros((v.height, v.width, 8), np.float32)
        for y in range(v.height):
            for x in range(v.width):
                orc.orc_init_ray(C.byref(v), C.byref(d), x, y, O.P(out[y, x]))
        return out
    finally:
        O.set_camera_ex(None, None)


def test_default_extras_are_the_pinhole_camera(orc):
    """camera1 == camera0, no lens, no aperture: uv_to_ray's general path reduces to the pinhole path bit for bit."""
    m = common.model("small16x2"); v = common.make_view(24, 16, spp=3, snap=0)
    a = _rays(orc, m, v); b = _rays(orc, m, v, common.camera_ex(common.camera_matrix()))
    assert np.array_equal(a, b)


def test_aperture_rays_meet_on_the_focus_plane(orc):
    """Thin lens (common_device.cuh:455-461): origin moves on the lens disk, origin + dir * plane_z stays the pinhole ray's focus point."""
    m = common.model("small16x2"); v = common.make_view(24, 16)
    pz, ap = 1.1, 0.05
    pin = _rays(orc, m, v)
    dof = _rays(orc, m, v, common.camera_ex(common.camera_matrix(), aperture=ap, plane_z=pz))
    cam = common.camera_matrix().reshape(4, 3)
    fwd = cam[2]
    # un-normalised pinhole direction has forward component 1: the focus point is origin + dir / dot(dir, fwd) * plane_z
    focus = pin[..., :3] + pin[..., 3:6] / (pin[..., 3:6] @ fwd)[..., None] * pz
    s = ((focus - dof[..., :3]) @ fwd) / (dof[..., 3:6] @ fwd)
    np.testing.assert_allclose(dof[..., :3] + dof[..., 3:6] * s[..., None], focus, atol=2e-5)
    off = dof[..., :3] - pin[..., :3]
    r = np.linalg.norm(off, axis=-1)
    assert r.max() <= ap * 1.0001 and r.max() > 0.5 * ap and np.abs(off @ fwd).max() < 1e-6     # on the lens disk, in the lens plane
    # the slice plane (negative plane_z) closes the aperture and kills every ray
    sl = _rays(orc, m, v, common.camera_ex(common.camera_matrix(), aperture=ap, plane_z=-pz))
    assert not sl[..., 7].any()
    np.testing.assert_allclose(sl[..., :6], pin[..., :6], atol=1e-7)


def test_panoramic_lenses_cover_the_sphere(orc):
    m = common.model("small16x2"); v = common.make_view(32, 16)
    cam = common.camera_matrix().reshape(4, 3)
    for lens in ("latlong", "equirectangular"):
        r = _rays(orc, m, v, common.camera_ex(common.camera_matrix(), lens=lens))
        d = r[..., 3:6]
        np.testing.assert_allclose(np.linalg.norm(d, axis=-1), 1.0, atol=1e-5)
        local = d @ cam[:3].T                  # components along right / down / forward
        assert local[..., 2].min() < -0.9 and local[..., 2].max() > 0.9 and local[..., 1].min() < -0.9 and local[..., 1].max() > 0.9
        np.testing.assert_allclose(r[..., :3], np.broadcast_to(cam[3], r[..., :3].shape), atol=1e-7)


def test_opencv_undistortion_inverts_the_distortion(orc):
    """iterative_opencv_lens_undistortion (common_device.cuh:290-334): distorting the undistorted direction gives the pixel back."""
    m = common.model("small16x2"); w, h = 24, 16
    v = common.make_view(w, h)
    k1, k2, p1, p2 = -0.12, 0.03, 0.002, -0.001
    r = _rays(orc, m, v, common.camera_ex(common.camera_matrix(), lens="opencv", lens_params=(k1, k2, p1, p2)))
    cam = common.camera_matrix().reshape(4, 3)
    local = r[..., 3:6] @ cam[:3].T
    u, vv = local[..., 0] / local[..., 2], local[..., 1] / local[..., 2]
    r2 = u * u + vv * vv
    radial = k1 * r2 + k2 * r2 * r2
    ud = u + u * radial + 2 * p1 * u * vv + p2 * (r2 + 2 * u * u)
    vd = vv + vv * radial + 2 * p2 * u * vv + p1 * (r2 + 2 * vv * vv)
    xs = (np.arange(w, dtype=np.float32) + 0.5) / w; ys = (np.arange(h, dtype=np.float32) + 0.5) / h
    f = v.focal_length[0]
    np.testing.assert_allclose(ud, np.broadcast_to((xs - 0.5) * w / f, ud.shape), atol=2e-5)
    np.testing.assert_allclose(vd, np.broadcast_to(((ys - 0.5) * h / f)[:, None], vd.shape), atol=2e-5)


def test_motion_blur_cameras_lie_between_the_shutter_ends(orc):
    m = common.model("small16x2"); v = common.make_view(24, 16, spp=2)
    cases = common.camera_ex_cases(24, 16)
    r = _rays(orc, m, v, cases["motion_blur"])
    c0 = common.camera_matrix().reshape(4, 3)[3]; c1 = np.asarray(list(cases["motion_blur"].camera1), np.float32).reshape(4, 3)[3]
    s = ((r[..., :3] - c0) @ (c1 - c0)) / ((c1 - c0) @ (c1 - c0))
    assert s.min() >= -1e-5 and s.max() <= 1 + 1e-5 and s.std() > 0.2      # a different shutter time per pixel (ld_random_val(spp, idx * 72239731))
    np.testing.assert_allclose(r[..., :3], c0 + s[..., None] * (c1 - c0), atol=1e-6)


def test_envmap_is_the_background_of_the_stock_render(orc):
    """init_rays_with_payload_kernel_nerf writes read_envmap(dir) for every valid ray; shade_kernel_nerf composites the NeRF over it."""
    m = common.model("small16x2"); w, h = 48, 32
    v = common.make_view(w, h)
    env = common.test_envmap()
    plain = O.render_nerf(m.desc(), v, alt=0)[0]
    O.set_camera_ex(None, env)
    try:
        withenv = O.render_nerf(m.desc(), v, alt=0)[0]
    finally:
        O.set_camera_ex(None, None)
    miss = plain[..., 3] == 0
    assert miss.any() and (~miss).any()
    assert np.abs(withenv[miss][:, 3] - 1.0).max() < 1e-6 and withenv[miss][:, :3].min() >= 0.0 and withenv[miss][:, :3].max() <= 1.06
    # over hit pixels: fb = nerf + envmap * (1 - alpha)
    a = plain[..., 3:4]
    bg = (withenv - plain) / np.maximum(1.0 - a, 1e-6)
    part = (~miss) & (plain[..., 3] < 0.9)
    if part.any():
        assert bg[part][:, :3].min() > -1e-3 and bg[part][:, :3].max() < 1.07
    # texel lookup: a constant map gives that constant everywhere
    const = np.zeros((4, 8, 4), np.float32); const[...] = (0.25, 0.5, 0.75, 1.0)
    O.set_camera_ex(None, const)
    try:
        c = O.render_nerf(m.desc(), v, alt=0)[0]
    finally:
        O.set_camera_ex(None, None)
    np.testing.assert_allclose(c[miss], np.broadcast_to(np.float32([0.25, 0.5, 0.75, 1.0]), c[miss].shape), atol=1e-6)


def test_abi_rejects_bad_camera_extras():
    lib = _lib.load()
    assert lib.snb_set_camera_ex(None, None) != 0 and lib.snb_set_envmap(None, None, 0, 0) != 0
