"""SURVEY.md §8 f4 — the rest of the stock render's camera (DoF, lens models, parallax, near plane, motion blur / rolling shutter,
environment map): CUDA product vs the CPU oracle. The extended path evaluates sinf / cosf / atanf / acosf / atan2f with CUDA's libm where
the oracle uses the host's, so rays agree to float rounding instead of bit for bit; the thresholds below say how closely."""
import ctypes as C

import numpy as np
import pytest

import common
import oracle_lib as O

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

W, H = 64, 48
CASES = list(common.camera_ex_cases(W, H).keys())


@pytest.fixture(scope="module")
def ctx(snb):
    c = C.c_void_p()
    assert snb.snb_create(0, None, C.byref(c)) == 0
    yield c
    snb.snb_destroy(c)


def dev(shape, dtype=None):
    t = torch.zeros(shape, dtype=dtype or torch.float32, device="cuda")
    torch.cuda.synchronize()   # the fill runs on torch's stream, the context renders on its own: finish the fill first
    return t


def _product_rays(snb, ctx, v, ex):
    n = v.width * v.height
    S = 1
    t, dt, cell, cnt, od = dev((n, S)), dev((n, S)), dev((n, S), torch.int32), dev((n,), torch.int32), dev((n, 6))
    assert snb.snb_set_camera_ex(ctx, C.byref(ex) if ex is not None else None) == 0
    try:
        assert snb.snb_march_trace(ctx, C.byref(v), S, t.data_ptr(), dt.data_ptr(), cell.data_ptr(), cnt.data_ptr(), od.data_ptr()) == 0 and snb.snb_sync(ctx) == 0
    finally:
        assert snb.snb_set_camera_ex(ctx, None) == 0
    return od.cpu().numpy()


def _oracle_rays(orc, m, v, ex):
    O.set_camera_ex(ex, None)
    try:
        d = m.desc()
        out = np.zeros((v.width * v.height, 8), np.float32)
        for y in range(v.height):
            for x in range(v.width):
                orc.orc_init_ray(C.byref(v), C.byref(d), x, y, O.P(out[x + v.width * y]))
        return out
    finally:
        O.set_camera_ex(None, None)


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("spp", [0, 5])
def test_primary_rays_match_oracle(snb, orc, ctx, case, spp):
    m = common.model("small16x2")
    d = m.desc(); assert snb.snb_set_model(ctx, C.byref(d)) == 0
    common.use_device_grid_scales(snb, ctx)
    v = common.make_view(W, H, spp=spp, snap=0 if spp else 1)
    ex = common.camera_ex_cases(W, H)[case]
    ours = _product_rays(snb, ctx, v, ex)
    ref = _oracle_rays(orc, m, v, ex)
    if case == "identity":
        assert np.array_equal(ours, ref[:, :6])          # reduces to the pinhole path: bit-exact
    else:
        np.testing.assert_allclose(ours, ref[:, :6], rtol=0, atol=3e-6)
    if case == "ftheta":
        inv = np.abs(ref[:, 3:6]).sum(-1) == 0                      # beyond 90 degrees: Ray::invalid()
        assert 0 < inv.sum() < inv.size // 2 and np.array_equal(inv, np.abs(ours[:, 3:6]).sum(-1) == 0)
    assert np.isfinite(ours).all()


def test_identity_extras_render_bit_identical_to_the_pinhole_path(snb, ctx):
    m = common.model("small16x2")
    d = m.desc(); assert snb.snb_set_model(ctx, C.byref(d)) == 0
    v = common.make_view(128, 96, spp=2, snap=0)
    a, da, b, db = dev((96, 128, 4)), dev((96, 128)), dev((96, 128, 4)), dev((96, 128))
    assert snb.snb_render_nerf(ctx, C.byref(v), None, 1, a.data_ptr(), da.data_ptr()) == 0
    ex = common.camera_ex(common.camera_matrix())
    assert snb.snb_set_camera_ex(ctx, C.byref(ex)) == 0
    assert snb.snb_render_nerf(ctx, C.byref(v), None, 1, b.data_ptr(), db.data_ptr()) == 0
    assert snb.snb_set_camera_ex(ctx, None) == 0 and snb.snb_sync(ctx) == 0
    assert torch.equal(a, b) and torch.equal(da, db)
    assert float(a[..., 3].max()) > 0.5


@pytest.mark.parametrize("case", ["dof", "motion_blur", "opencv", "latlong", "parallax_near", "slice_plane", "everything"])
@pytest.mark.parametrize("envmap", [False, True])
def test_stock_frame_with_extras_matches_oracle(snb, orc, ctx, case, envmap):
    """snb_render_nerf with the extended camera vs the oracle's render_nerf (stock path): frame >= 50 dB, at most 0.5 % of the pixels
    over 1e-2 (rays whose direction differs in the last bit can take another cell at a sample), depth likewise."""
    m = common.model("small16x2")
    d = m.desc(); assert snb.snb_set_model(ctx, C.byref(d)) == 0
    common.use_device_grid_scales(snb, ctx)
    w, h = 96, 64
    v = common.make_view(w, h, spp=1, snap=0)
    ex = common.camera_ex_cases(w, h)[case]
    env = common.test_envmap() if envmap else None
    O.set_camera_ex(ex, env)
    try:
        rfb, rdepth, _, _ = O.render_nerf(m.desc(), v, alt=0)
    finally:
        O.set_camera_ex(None, None)
    fb, depth = dev((h, w, 4)), dev((h, w))
    assert snb.snb_set_camera_ex(ctx, C.byref(ex)) == 0
    assert snb.snb_set_envmap(ctx, O.P(env) if envmap else None, env.shape[1] if envmap else 0, env.shape[0] if envmap else 0) == 0
    try:
        assert snb.snb_render_nerf(ctx, C.byref(v), None, 1, fb.data_ptr(), depth.data_ptr()) == 0 and snb.snb_sync(ctx) == 0
    finally:
        assert snb.snb_set_camera_ex(ctx, None) == 0 and snb.snb_set_envmap(ctx, None, 0, 0) == 0
    fb = fb.cpu().numpy(); depth = depth.cpu().numpy()
    err = np.abs(fb - rfb).max(-1)
    over = int((err > 1e-2).sum())
    assert common.psnr(fb, rfb) >= 50.0, (common.psnr(fb, rfb), over)
    assert over <= max(2, int(5e-3 * w * h)), over
    assert (np.abs(depth - rdepth) > 0.05).mean() < 5e-3
    if case == "slice_plane":
        assert np.allclose(depth, 1.2) and (fb[..., :3] == 0).all() if not envmap else np.allclose(depth, 1.2)
    else:
        assert (rfb[..., 3] > 0.5).mean() > (0.01 if case == "latlong" else 0.03)     # the panorama shows the scene small
    if envmap and case != "slice_plane":
        assert (fb[..., 3] > 0.99).all()         # every valid ray sees the map behind the NeRF


def test_testbed_render_motion_blur_and_dof(snb):
    """pyngp.Testbed.render(spp, start_t, end_t, shutter_fraction) (src/python_api.cu:105-205): a moving camera blurs the frame, a closed
    shutter (shutter_fraction -> 0) gives the still frame back, and an open aperture changes it; envmap fills the background."""
    from synerfgine_b200.testbed import Testbed, camera_path_from_keyframes
    tb = Testbed()
    try:
        tb.set_model(common.model("small16x2"))
        cam0 = common.camera_matrix().reshape(4, 3)
        cam1 = common.camera_matrix(view=(-0.22, -0.47, -0.86), at=(0.55, 0.43, 0.5)).reshape(4, 3)
        tb.camera_matrix = cam0.copy()
        still = tb.render(96, 64, spp=4)
        tb.camera_path = camera_path_from_keyframes([cam0, cam1])
        tb.smoothed_camera = None
        blur = tb.render(96, 64, spp=4, start_t=0.0, end_t=1.0, shutter_fraction=1.0)
        assert np.abs(blur - still).mean() > 1e-3
        tb.smoothed_camera = None
        closed = tb.render(96, 64, spp=4, start_t=0.0, end_t=1.0, shutter_fraction=1e-6)
        assert common.psnr(closed, still) > 45.0
        np.testing.assert_allclose(tb.smoothed_camera, cam1, atol=1e-5)     # the next frame starts where this one ended
        # depth of field
        tb.camera_path = None; tb.camera_matrix = cam0.copy()
        tb.aperture_size = 0.05
        dof = tb.render(96, 64, spp=4)
        assert np.abs(dof - still).mean() > 1e-3
        tb.aperture_size = 0.0
        again = tb.render(96, 64, spp=4)
        assert np.array_equal(again, still)
        # environment map: linear output, background pixels take the map
        tb.envmap = common.test_envmap()
        env = tb.render(96, 64, spp=1)
        assert (env[..., 3] > 0.99).all() and (still[..., 3] < 0.5).any()
        tb.envmap = None
        assert np.array_equal(tb.render(96, 64, spp=4), still)
    finally:
        tb.close()
