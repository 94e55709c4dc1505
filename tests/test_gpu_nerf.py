"""GPU parity: NeRF layer of libsnb.so against the CPU oracle (through the C ABI, same seeded inputs).

Bars (BASELINE.json north_star): occupancy traversal sample indices and hash indices bit-exact; framebuffer
>= 50 dB PSNR and max per-pixel abs error <= 1e-2 for the fp16 MLP output.
"""
import ctypes as C

import numpy as np
import pytest

import common
import oracle_lib as O

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def ctx(snb):
    c = C.c_void_p()
    assert snb.snb_create(0, None, C.byref(c)) == 0
    yield c
    snb.snb_destroy(c)


def set_model(snb, ctx, m):
    d = m.desc()
    rc = snb.snb_set_model(ctx, C.byref(d))
    assert rc == 0, snb.snb_last_error(ctx)
    common.use_device_grid_scales(snb, ctx)
    return d


def dev(shape, dtype):
    t = torch.zeros(shape, dtype=dtype, device="cuda")
    torch.cuda.synchronize()   # the fill runs on torch's stream, the context renders on its own: finish the fill first
    return t


MARCHES = ["lane", "ws"]   # march kernels selectable with snb_set_option(ctx, "march", ...); the first one is the default


@pytest.mark.parametrize("kind,w,h,spp,snap", [("small16x2", 96, 64, 0, 1), ("small16x2", 64, 48, 3, 0), ("unit16x2", 64, 64, 0, 1), ("big16x2", 64, 64, 1, 0),
                                                 ("noocc", 32, 32, 0, 1), ("cropped", 96, 64, 0, 1), ("cropped", 64, 48, 2, 0)])
def test_march_samples_bit_exact(snb, orc, ctx, kind, w, h, spp, snap):
    m = common.cropped_model() if kind == "cropped" else common.model(kind)
    d = set_model(snb, ctx, m)
    cam = common.camera_matrix(zoom=2.5) if kind == "big16x2" else None
    v = common.make_view(w, h, cam=cam, spp=spp, snap=snap)
    S = 48
    n = w * h
    t, dt, cell, cnt, od = dev((n, S), torch.float32), dev((n, S), torch.float32), dev((n, S), torch.int32), dev((n,), torch.int32), dev((n, 6), torch.float32)
    assert snb.snb_march_trace(ctx, C.byref(v), S, t.data_ptr(), dt.data_ptr(), cell.data_ptr(), cnt.data_ptr(), od.data_ptr()) == 0, snb.snb_last_error(ctx)
    assert snb.snb_sync(ctx) == 0
    rt, rdt, rcell, rcnt, rod = np.zeros((n, S), np.float32), np.zeros((n, S), np.float32), np.zeros((n, S), np.uint32), np.zeros(n, np.uint32), np.zeros((n, 6), np.float32)
    orc.orc_march_trace(C.byref(d), C.byref(v), S, O.P(rt), O.P(rdt), O.P(rcell, O.pu32), O.P(rcnt, O.pu32), O.P(rod))
    assert rcnt.sum() > 0, "degenerate test: no samples"
    np.testing.assert_array_equal(od.cpu().numpy().view(np.uint32), rod.view(np.uint32))          # ray origins / directions
    np.testing.assert_array_equal(cnt.cpu().numpy().view(np.uint32), rcnt)                         # samples per ray
    np.testing.assert_array_equal(cell.cpu().numpy().view(np.uint32), rcell)                       # mip << 24 | morton
    np.testing.assert_array_equal(t.cpu().numpy().view(np.uint32), rt.view(np.uint32))             # t bit pattern
    np.testing.assert_array_equal(dt.cpu().numpy().view(np.uint32), rdt.view(np.uint32))


@pytest.mark.parametrize("mlp", ["mma", "tc"])
@pytest.mark.parametrize("kind", ["small16x2", "fox8x4", "f1", "f8", "big16x2"])
def test_network_hash_indices_and_outputs(snb, orc, ctx, kind, mlp):
    """mlp = "mma": warp-level mma.sync evaluator; "tc": CTA-level tcgen05/TMEM evaluator (128-row tiles, ragged last tile)."""
    m = common.model(kind)
    d = set_model(snb, ctx, m)
    assert snb.snb_set_option(ctx, b"mlp", mlp.encode()) == 0
    n = 3000 + 7   # ragged: not a multiple of the 32-row tile
    rng = np.random.default_rng(5)
    coords = np.zeros((n, 7), np.float32)
    coords[:, :3] = rng.uniform(0.0, 1.0, (n, 3)); coords[:, 3] = 0.01; coords[:, 4:] = rng.uniform(0.0, 1.0, (n, 3))
    coords[0, :3] = 0.0; coords[1, :3] = 1.0    # box corners
    L = m.n_levels
    dc = torch.from_numpy(coords).cuda()
    out, feat, hidx = dev((n, 4), torch.float16), dev((n, 32), torch.float16), dev((n, L * 8), torch.int32)
    rc = snb.snb_network_inference(ctx, dc.data_ptr(), n, out.data_ptr(), feat.data_ptr(), hidx.data_ptr())
    snb.snb_set_option(ctx, b"mlp", b"mma")
    assert rc == 0, snb.snb_last_error(ctx)
    assert snb.snb_sync(ctx) == 0, snb.snb_last_error(ctx)
    rout, rfeat, rh = np.zeros((n, 4), np.uint16), np.zeros((n, 32), np.uint16), np.zeros((n, L * 8), np.uint32)
    orc.orc_network_inference(C.byref(d), O.P(coords), n, O.P(rout, O.pu16), O.P(rfeat, O.pu16), O.P(rh, O.pu32))
    np.testing.assert_array_equal(hidx.cpu().numpy().view(np.uint32), rh)                                   # hash indices: bit-exact
    np.testing.assert_array_equal(feat.cpu().numpy().view(np.uint16), rfeat)                                # fp16 features: same fma order -> bit-exact
    a = out.cpu().numpy().astype(np.float32); b = rout.view(np.float16).astype(np.float32)
    # MLP: tensor-core fp32 accumulation order differs from the scalar oracle -> compare within fp16 rounding of the outputs
    tol = 2e-3 * np.maximum(1.0, np.abs(b)) * 4
    assert np.all(np.abs(a - b) <= tol), f"max |diff| {np.abs(a - b).max()}"
    assert np.abs(b).max() > 0.1


@pytest.mark.parametrize("kind,alt,w,h", [("small16x2", 0, 128, 96), ("small16x2", 1, 128, 96), ("fox8x4", 1, 64, 64), ("unit16x2", 0, 64, 64), ("big16x2", 1, 64, 64),
                                          ("cropped", 1, 96, 64)])
@pytest.mark.parametrize("march", MARCHES)
def test_render_nerf_framebuffer(snb, orc, ctx, kind, alt, w, h, march):
    """Every march kernel the library offers (MARCHES) against the oracle."""
    m = common.cropped_model() if kind == "cropped" else common.model(kind)
    d = set_model(snb, ctx, m)
    assert snb.snb_set_option(ctx, b"march", march.encode()) == 0
    cam = common.camera_matrix(zoom=2.5) if kind == "big16x2" else None
    v = common.make_view(w, h, cam=cam)
    fb, depth = dev((h, w, 4), torch.float32), dev((h, w), torch.float32)
    rt = common.default_rt(show_shadow_on_nerf=0)
    if alt:
        pos, nrm = dev((h, w, 3), torch.float32), dev((h, w, 3), torch.float32)
        rc = snb.snb_render_nerf_with_buffers(ctx, C.byref(v), None, C.byref(rt), 1, fb.data_ptr(), depth.data_ptr(), pos.data_ptr(), nrm.data_ptr())
    else:
        rc = snb.snb_render_nerf(ctx, C.byref(v), None, 1, fb.data_ptr(), depth.data_ptr())
    assert rc == 0, snb.snb_last_error(ctx)
    rc = snb.snb_sync(ctx)
    snb.snb_set_option(ctx, b"march", MARCHES[0].encode())
    assert rc == 0, snb.snb_last_error(ctx)
    rfb, rdepth, rpos, ns = O.render_nerf(d, v, alt, positions=bool(alt))
    a = fb.cpu().numpy()
    assert (rfb[..., 3] > 0.5).mean() > 0.05, "degenerate test: nothing visible"
    err = np.abs(a - rfb)
    assert common.psnr(a, rfb) >= 50.0, common.psnr(a, rfb)
    assert err.max() <= 1e-2, err.max()
    # depth: same samples -> agree wherever both wrote it; allow the rare ray whose termination step differs by one sample
    dd = np.abs(depth.cpu().numpy() - rdepth)
    assert (dd > 0.05).mean() < 2e-3, (dd > 0.05).mean()
    cnt = (C.c_uint64 * 4)()
    assert snb.snb_get_counters(ctx, cnt) == 0
    # network samples: the oracle (like the reference) evaluates whole 8-step batches and discards what follows a saturated
    # sample; the fused kernel composites step by step and never evaluates those, so it does the same or less work.
    assert 0.85 * ns <= int(cnt[0]) + int(cnt[3]) <= ns + 8, (cnt[0], cnt[3], ns)   # counted: samples composited = evaluated + composited from a re-used output (dup_reuse)
    if alt:
        pd = np.abs(pos.cpu().numpy() - rpos)
        assert (pd.max(-1) > 0.05).mean() < 2e-3
        rn = np.zeros((h * w, 3), np.float32)
        gp = pos.cpu().numpy().reshape(-1, 3).copy()
        orc.orc_write_normals(w, h, O.P(gp), O.P(rn))                          # normals from the product's own positions: exact stencil parity
        np.testing.assert_allclose(nrm.cpu().numpy().reshape(-1, 3), rn, atol=1e-6)


def test_tile_equals_crop_of_full_frame(snb, ctx):
    from synerfgine_b200._lib import Tile
    m = common.model("small16x2")
    set_model(snb, ctx, m)
    w, h = 96, 64
    v = common.make_view(w, h, spp=2, snap=0)
    full, fdepth = dev((h, w, 4), torch.float32), dev((h, w), torch.float32)
    assert snb.snb_render_nerf(ctx, C.byref(v), None, 1, full.data_ptr(), fdepth.data_ptr()) == 0
    t = Tile(40, 16, 48, 32)
    part, pdepth = dev((t.h, t.w, 4), torch.float32), dev((t.h, t.w), torch.float32)
    assert snb.snb_render_nerf(ctx, C.byref(v), C.byref(t), 1, part.data_ptr(), pdepth.data_ptr()) == 0
    assert snb.snb_sync(ctx) == 0
    assert torch.equal(part, full[t.y0:t.y0 + t.h, t.x0:t.x0 + t.w])
    assert torch.equal(pdepth, fdepth[t.y0:t.y0 + t.h, t.x0:t.x0 + t.w])


def test_bitfield_builder_matches_host_twin(snb, ctx):
    from synerfgine_b200 import synthetic
    m = common.model("small16x2")
    out = np.zeros(synthetic.CASCADES * synthetic.CELLS // 8, np.uint8)
    dens = np.ascontiguousarray(m.density)
    assert snb.snb_density_to_bitfield(ctx, dens.ctypes.data_as(C.POINTER(C.c_float)), m.max_cascade, out.ctypes.data_as(C.POINTER(C.c_uint8))) == 0
    np.testing.assert_array_equal(out, m.bitfield)


def test_errors_are_reported_not_thrown(snb, ctx):
    v = common.make_view(16, 16)
    assert snb.snb_render_nerf(ctx, C.byref(v), None, 1, None, None) != 0
    from synerfgine_b200._lib import ModelDesc
    bad = common.model("small16x2").desc(); bad.n_levels = 5
    assert snb.snb_set_model(ctx, C.byref(bad)) != 0
    assert b"n_levels" in snb.snb_last_error(ctx)


# ---- BASELINE.json's full size (C2 model, 1920x1080) through size-independent properties ---------------------------------
def _render_full(snb, ctx, v, tile=None, march=MARCHES[0].encode(), alt_schedule=b"fixed"):
    """sng NeRF layer (positions + normals, no shadows). alt_schedule "fixed" (8 samples per batch) makes a tile and the frame it
    belongs to march identically whatever their ray counts are — the self-consistency tests below are about tiling, not about
    the reference's count-dependent batch lengths (those: test_alt_schedule_*)."""
    from synerfgine_b200._lib import Tile
    assert snb.snb_set_option(ctx, b"march", march) == 0
    assert snb.snb_set_option(ctx, b"alt_schedule", alt_schedule) == 0
    w, h = (tile.w, tile.h) if tile is not None else (v.width, v.height)
    fb, depth = dev((h, w, 4), torch.float32), dev((h, w), torch.float32)
    pos, nrm = dev((h, w, 3), torch.float32), dev((h, w, 3), torch.float32)
    rt = common.default_rt(show_shadow_on_nerf=0)
    rc = snb.snb_render_nerf_with_buffers(ctx, C.byref(v), C.byref(tile) if tile is not None else None, C.byref(rt), 1, fb.data_ptr(), depth.data_ptr(),
                                          pos.data_ptr(), nrm.data_ptr())
    assert rc == 0, snb.snb_last_error(ctx)
    assert snb.snb_sync(ctx) == 0, snb.snb_last_error(ctx)
    cnt = (C.c_uint64 * 4)()
    assert snb.snb_get_counters(ctx, cnt) == 0
    snb.snb_set_option(ctx, b"march", MARCHES[0].encode())
    snb.snb_set_option(ctx, b"alt_schedule", b"auto")
    return fb, depth, pos, (int(cnt[0]), int(cnt[1]), int(cnt[2]))


def test_full_size_properties(snb, ctx):
    """1080p frame of the benchmark model: (1) two runs are bit-identical (the queue order is not, the result must be);
    (2) every other march kernel evaluates the same samples (counters within a sample per ray) and agrees within the fp16 tolerance;
    (3) four column strips with the multi-GPU halo reproduce the full frame bit for bit, and their sample counts add up to at
    least the full frame's."""
    from synerfgine_b200._lib import Tile
    from synerfgine_b200 import tiles
    m = common.model("fox16x2")
    set_model(snb, ctx, m)
    W, H = 1920, 1080
    v = common.make_view(W, H, spp=1, snap=0)
    a_fb, a_depth, a_pos, a_cnt = _render_full(snb, ctx, v)
    b_fb, b_depth, b_pos, b_cnt = _render_full(snb, ctx, v)
    assert torch.equal(a_fb, b_fb) and torch.equal(a_depth, b_depth) and torch.equal(a_pos, b_pos) and a_cnt == b_cnt   # counters = samples COMPOSITED ("ws" also evaluates, and ignores, the one or two in flight when a ray saturates)
    assert a_cnt[0] > 5_000_000 and (a_fb[..., 3] > 0.5).float().mean() > 0.05
    # the colour conversion of finished rays, inside the march loop (extract_from_payload's place) or as one pass over the tile afterwards
    # (option defer_srgb, the default): the same function of the same value, so the frames are the same bits
    assert snb.snb_set_option(ctx, b"defer_srgb", b"0") == 0
    i_fb, i_depth, i_pos, i_cnt = _render_full(snb, ctx, v)
    assert snb.snb_set_option(ctx, b"defer_srgb", b"1") == 0
    assert torch.equal(i_fb, a_fb) and torch.equal(i_depth, a_depth) and torch.equal(i_pos, a_pos) and i_cnt == a_cnt
    # the other march kernels evaluate every sample through the network: compare them with the lane kernel doing the same (dup_reuse off)
    assert snb.snb_set_option(ctx, b"dup_reuse", b"0") == 0
    x_fb, x_depth, x_pos, x_cnt = _render_full(snb, ctx, v)
    assert snb.snb_set_option(ctx, b"dup_reuse", b"1") == 0
    over = int(((x_fb - a_fb).abs().amax(-1) > 1e-2).sum())
    print(f"[dup_reuse at 1080p] network evaluations {a_cnt[0]} + {x_cnt[0] - a_cnt[0]} re-used; psnr {common.psnr(x_fb.cpu().numpy(), a_fb.cpu().numpy()):.1f} dB, pixels over 1e-2: {over}")
    assert common.psnr(x_fb.cpu().numpy(), a_fb.cpu().numpy()) >= 50.0 and over <= 5e-4 * W * H
    for other in MARCHES[1:]:
        t_fb, t_depth, t_pos, t_cnt = _render_full(snb, ctx, v, march=other.encode())
        err = (t_fb - x_fb).abs()
        assert float(err.max()) <= 1e-2 and common.psnr(t_fb.cpu().numpy(), x_fb.cpu().numpy()) >= 50.0
        assert abs(t_cnt[0] - x_cnt[0]) <= 1e-3 * x_cnt[0] and t_cnt[1] == x_cnt[1]     # a ray may saturate one sample earlier or later
    halo, total = tiles.halo_for(4), 0
    for r in range(4):
        s = tiles.strip_for(r, 4, W, halo)
        fb, depth, pos, cnt = _render_full(snb, ctx, v, tile=Tile(s.hx0, 0, s.render_width, H))
        assert torch.equal(fb[:, s.crop], a_fb[:, s.x0:s.x1]) and torch.equal(depth[:, s.crop], a_depth[:, s.x0:s.x1])
        assert torch.equal(pos[:, s.crop], a_pos[:, s.x0:s.x1])
        total += cnt[0]
    assert a_cnt[0] <= total <= 1.05 * a_cnt[0]                                      # the halo columns are marched twice


@pytest.mark.parametrize("w,h", [(1, 1), (3, 2), (33, 1), (1, 37)])
def test_degenerate_frame_sizes(snb, orc, ctx, w, h):
    """Ragged inputs: frames smaller than a warp's 8x4 patch, single rows and columns."""
    m = common.model("small16x2")
    d = set_model(snb, ctx, m)
    v = common.make_view(w, h)
    fb, depth = dev((h, w, 4), torch.float32), dev((h, w), torch.float32)
    assert snb.snb_render_nerf(ctx, C.byref(v), None, 1, fb.data_ptr(), depth.data_ptr()) == 0, snb.snb_last_error(ctx)
    assert snb.snb_sync(ctx) == 0
    rfb, rdepth, _, _ = O.render_nerf(d, v, 0)
    assert np.abs(fb.cpu().numpy() - rfb).max() <= 1e-2


def test_camera_that_sees_nothing(snb, ctx):
    """Empty input: every ray misses the volume -> zero rays queued, zero samples, a cleared frame, no hang."""
    m = common.model("small16x2")
    set_model(snb, ctx, m)
    v = common.make_view(64, 48)
    for i in (6, 7, 8):
        v.camera[i] = -v.camera[i]                                                     # forward column negated: the eye looks away from the box
    fb, depth = dev((48, 64, 4), torch.float32), dev((48, 64), torch.float32)
    fb.fill_(7.0)
    assert snb.snb_render_nerf(ctx, C.byref(v), None, 1, fb.data_ptr(), depth.data_ptr()) == 0, snb.snb_last_error(ctx)
    assert snb.snb_sync(ctx) == 0
    cnt = (C.c_uint64 * 4)()
    assert snb.snb_get_counters(ctx, cnt) == 0
    assert int(cnt[0]) == 0 and float(fb.abs().max()) == 0.0


def _alt_status(snb, ctx):
    st = (C.c_uint32 * 132)()
    assert snb.snb_get_alt_status(ctx, st) == 0
    st = list(st)
    return {"passes": st[0], "converged": st[1], "mismatch": st[2], "alive": np.asarray(st[4:68], np.uint32), "sched": np.asarray(st[68:132], np.uint32)}


@pytest.mark.parametrize("kind,w,h,target", [("small16x2", 96, 64, 7800), ("small16x2", 128, 96, 17500), ("fox8x4", 64, 64, 7606)])
@pytest.mark.parametrize("march", MARCHES)
def test_alt_schedule_matches_reference_wavefront(snb, orc, ctx, kind, w, h, target, march):
    """The sng compositor's result depends on the batch lengths S_j = clamp(target / n_alive_j, 1, 8) of the reference's lock-step
    march (src/testbed_nerf.cu:2189-2190,574). target_n_queries is lowered so that a frame the CPU oracle can replay has short
    batches at the start, like a 1080p frame has under the real 2^21. After the in-frame passes the library reports the counts
    n_alive_j and the sequence: both must EQUAL the oracle's replay of the wavefront loop, and the image must meet the bar."""
    m = common.model(kind)
    d = set_model(snb, ctx, m)
    assert snb.snb_set_option(ctx, b"march", march.encode()) == 0
    assert snb.snb_set_option(ctx, b"alt_schedule", b"auto") == 0
    assert snb.snb_set_option(ctx, b"dup_reuse", b"0") == 0                               # exact counts: every sample through the network, like the oracle
    assert snb.snb_set_option(ctx, b"alt_target_queries", str(target).encode()) == 0
    assert snb.snb_set_alt_schedule(ctx, None, 0) == 0                                   # cold start
    v = common.make_view(w, h)
    fb, depth = dev((h, w, 4), torch.float32), dev((h, w), torch.float32)
    pos, nrm = dev((h, w, 3), torch.float32), dev((h, w, 3), torch.float32)
    rt = common.default_rt(show_shadow_on_nerf=0)
    rfb, rdepth, rpos, rsched, ralive = O.render_nerf_wavefront(d, v, alt=1, target_queries=target, positions=True)
    assert rsched.min() < 8 and len(set(rsched.tolist())) >= 3, f"degenerate test: schedule {rsched}"
    try:
        for attempt in range(4):                                                       # a cold start may need more than one call
            rc = snb.snb_render_nerf_with_buffers(ctx, C.byref(v), None, C.byref(rt), 1, fb.data_ptr(), depth.data_ptr(), pos.data_ptr(), nrm.data_ptr())
            assert rc == 0, snb.snb_last_error(ctx)
            st = _alt_status(snb, ctx)
            if st["converged"]:
                break
        assert st["converged"] == 1, st
        k = min(len(rsched), 64)
        np.testing.assert_array_equal(st["alive"][:k], ralive[:k])                       # n_alive of every lock-step iteration: exact
        np.testing.assert_array_equal(st["sched"][:k], rsched[:k])
        a = fb.cpu().numpy()
        assert common.psnr(a, rfb) >= 50.0 and np.abs(a - rfb).max() <= 1e-2, (common.psnr(a, rfb), np.abs(a - rfb).max())
        # steady state: the next frame needs one pass
        rc = snb.snb_render_nerf_with_buffers(ctx, C.byref(v), None, C.byref(rt), 1, fb.data_ptr(), depth.data_ptr(), pos.data_ptr(), nrm.data_ptr())
        assert rc == 0
        st2 = _alt_status(snb, ctx)
        assert st2["passes"] == 1 and st2["converged"] == 1
        assert np.abs(fb.cpu().numpy() - rfb).max() <= 1e-2
        # and a fixed batch length is a different image (the deviation the option exists to make explicit)
        assert snb.snb_set_option(ctx, b"alt_schedule", b"fixed") == 0
        rc = snb.snb_render_nerf_with_buffers(ctx, C.byref(v), None, C.byref(rt), 1, fb.data_ptr(), depth.data_ptr(), pos.data_ptr(), nrm.data_ptr())
        assert rc == 0 and snb.snb_sync(ctx) == 0
        r8, _, _, _ = O.render_nerf(d, v, 1, positions=True)
        assert np.abs(fb.cpu().numpy() - r8).max() <= 1e-2
        assert np.abs(r8 - rfb).max() > 2e-2, "degenerate test: the schedule does not matter for this frame"
    finally:
        snb.snb_set_option(ctx, b"alt_schedule", b"auto"); snb.snb_set_option(ctx, b"alt_target_queries", b"2097152"); snb.snb_set_option(ctx, b"march", MARCHES[0].encode())
        snb.snb_set_option(ctx, b"dup_reuse", b"1")
        snb.snb_set_alt_schedule(ctx, None, 0)


def test_alt_schedule_external_tiles_equal_full_frame(snb, ctx):
    """Multi-GPU form: every tile marches with the sequence of the WHOLE frame (n_alive_j is a whole-frame quantity) and the tiles'
    counts add up to the frame's; then tiles are bit-identical crops of the frame although their own ray counts are smaller."""
    from synerfgine_b200._lib import Tile
    from synerfgine_b200 import tiles
    m = common.model("small16x2")
    set_model(snb, ctx, m)
    W, H, target = 160, 96, 22000
    v = common.make_view(W, H, spp=1, snap=0)
    assert snb.snb_set_option(ctx, b"alt_target_queries", str(target).encode()) == 0
    try:
        assert snb.snb_set_alt_schedule(ctx, None, 0) == 0
        for _ in range(4):
            full = _render_full(snb, ctx, v, alt_schedule=b"auto")
            st = _alt_status(snb, ctx)
            if st["converged"]:
                break
        assert st["converged"] == 1 and st["sched"][:8].min() < 8
        sched = st["sched"].astype(np.uint8)
        total = np.zeros(64, np.uint64)
        for r in range(3):
            s = tiles.strip_for(r, 3, W, 2)
            assert snb.snb_set_alt_schedule(ctx, sched.ctypes.data_as(C.POINTER(C.c_uint8)), 64) == 0
            fb, depth, pos, cnt = _render_full(snb, ctx, v, tile=Tile(s.hx0, 0, s.render_width, H), alt_schedule=b"external")
            assert torch.equal(fb[:, s.crop], full[0][:, s.x0:s.x1]) and torch.equal(depth[:, s.crop], full[1][:, s.x0:s.x1])
            own = Tile(s.x0, 0, s.x1 - s.x0, H)                                          # counts of the owned columns only (no halo)
            assert snb.snb_set_alt_schedule(ctx, sched.ctypes.data_as(C.POINTER(C.c_uint8)), 64) == 0
            _render_full(snb, ctx, v, tile=own, alt_schedule=b"external")
            total += _alt_status(snb, ctx)["alive"].astype(np.uint64)
        np.testing.assert_array_equal(total, st["alive"].astype(np.uint64))
    finally:
        snb.snb_set_option(ctx, b"alt_target_queries", b"2097152"); snb.snb_set_alt_schedule(ctx, None, 0)


def test_snapshot_file_round_trip_renders_the_same_frame(snb, tmp_path):
    """save_snapshot -> .ingp -> load_snapshot (bitfield rebuilt on the GPU from the stored fp16 density grid) renders
    the same bits as the model set from memory, and adopts the stored camera."""
    from synerfgine_b200 import Testbed
    m = common.model("small16x2")
    a = Testbed(0)
    a.set_model(m)
    a.camera_matrix = np.asarray(common.camera_matrix(), dtype=np.float32).reshape(4, 3)
    ra, da = a.render_nerf_raw(96, 64)
    ra, da = ra.cpu().numpy().copy(), da.cpu().numpy().copy()
    path = tmp_path / "scene.ingp"
    a.save_snapshot(path, m)
    a.close()
    b = Testbed(0)
    b.load_snapshot(path)
    np.testing.assert_array_equal(np.asarray(b.camera_matrix, np.float32).reshape(-1), np.asarray(common.camera_matrix(), np.float32).reshape(-1))
    rb, db = b.render_nerf_raw(96, 64)
    # the stored density grid is fp16: the bitfield rebuilt from it equals the one built from fp32 densities except where the
    # rounding crosses the occupancy threshold, so compare against a model whose densities went through fp16 as well
    m16 = common.model("small16x2")
    from synerfgine_b200 import synthetic
    m16.density = m.density.astype(np.float16).astype(np.float32)
    m16.bitfield = synthetic.density_to_bitfield(m16.density, m16.max_cascade)
    c = Testbed(0)
    c.set_model(m16)
    c.camera_matrix = b.camera_matrix
    rc_, dc = c.render_nerf_raw(96, 64)
    assert torch.equal(rb, rc_) and torch.equal(db, dc)
    assert np.abs(rb.cpu().numpy() - ra).max() <= 1e-2 or (np.abs(rb.cpu().numpy() - ra) > 1e-2).mean() < 0.01
    assert (rb[..., 3] > 0.5).float().mean() > 0.05
    with pytest.raises(RuntimeError):
        b.load_snapshot(tmp_path / "missing.ingp")
    b.close(); c.close()


def test_large_frame_4k(snb, ctx):
    """Maximum size the reference's GUI offers (4K): buffers grow, 32-bit indexing holds (8.3 M pixels x 16 B rows, queue of
    64 B records), and a tile in the far corner still equals the crop of the full frame."""
    from synerfgine_b200._lib import Tile
    m = common.model("small16x2")
    set_model(snb, ctx, m)
    W, H = 3840, 2160
    v = common.make_view(W, H, spp=0, snap=1)
    fb, depth, pos, cnt = _render_full(snb, ctx, v)
    assert bool(torch.isfinite(fb).all()) and cnt[1] > 100_000 and (fb[..., 3] > 0.5).float().mean() > 0.05
    t = Tile(W - 517, H - 301, 517, 301)
    pfb, pdepth, ppos, _ = _render_full(snb, ctx, v, tile=t)
    assert torch.equal(pfb, fb[t.y0:, t.x0:]) and torch.equal(pdepth, depth[t.y0:, t.x0:]) and torch.equal(ppos, pos[t.y0:, t.x0:])


def test_set_model_from_device_memory_renders_identically(snb):
    """snb_set_model_dev: parameters and bitfield already on the GPU (SURVEY.md 8b "no hidden host copies") -> the same frame, bit for bit."""
    m = common.model("small16x2")
    v = common.make_view(128, 96)
    frames = []
    for on_device in (False, True):
        c = C.c_void_p(); assert snb.snb_create(0, None, C.byref(c)) == 0
        try:
            d = m.desc()
            if on_device:
                n = int(d.n_params)
                params = torch.from_numpy(np.ctypeslib.as_array(d.params, shape=(n,)).astype(np.int16, copy=True)).cuda()
                bits = torch.from_numpy(np.ctypeslib.as_array(d.bitfield, shape=(8 * 128 ** 3 // 8,)).copy()).cuda()
                torch.cuda.synchronize()
                d.params = C.cast(C.c_void_p(params.data_ptr()), C.POINTER(C.c_uint16)); d.bitfield = C.cast(C.c_void_p(bits.data_ptr()), C.POINTER(C.c_uint8))
                assert snb.snb_set_model_dev(c, C.byref(d)) == 0, snb.snb_last_error(c)
            else:
                assert snb.snb_set_model(c, C.byref(d)) == 0
            fb, depth = dev((96, 128, 4), torch.float32), dev((96, 128), torch.float32)
            torch.cuda.synchronize()
            assert snb.snb_render_nerf(c, C.byref(v), None, 1, fb.data_ptr(), depth.data_ptr()) == 0 and snb.snb_sync(c) == 0
            frames.append((fb.cpu().numpy(), depth.cpu().numpy()))
        finally:
            snb.snb_destroy(c)
    assert frames[0][0][..., 3].max() > 0.5
    assert np.array_equal(frames[0][0], frames[1][0]) and np.array_equal(frames[0][1], frames[1][1])


@pytest.mark.parametrize("kind,w,h,target", [("small16x2", 128, 96, 17500), ("fox8x4", 64, 64, 7606), ("small16x2", 96, 64, 2097152)])
def test_lockstep_prefix_equals_the_speculative_schedule(snb, orc, ctx, kind, w, h, target):
    """alt_schedule=lockstep (what "auto" uses for frames whose camera or world moved): the batches shorter than 8 are marched one per
    launch (up to 16) with the rays parked in between, so the sequence is exact in ONE pass whatever the previous frame was. Every ray does the same
    arithmetic as under the speculative schedule once that has converged: frames, positions, counts and sequence are bit-identical —
    for a cold start and along a moving camera."""
    m = common.model(kind)
    set_model(snb, ctx, m)
    assert snb.snb_set_option(ctx, b"march", b"lane") == 0
    assert snb.snb_set_option(ctx, b"alt_target_queries", str(target).encode()) == 0
    rt = common.default_rt(show_shadow_on_nerf=0)
    cams = [common.camera_matrix(), common.camera_matrix(view=(-0.22, -0.47, -0.86), at=(0.55, 0.43, 0.5)), common.camera_matrix(view=(0.3, -0.4, -0.8), zoom=0.8),
            common.camera_matrix(view=(0.3, -0.4, -0.8), zoom=0.8)]
    bufs = lambda: (dev((h, w, 4), torch.float32), dev((h, w), torch.float32), dev((h, w, 3), torch.float32), dev((h, w, 3), torch.float32))

    def render(v, b):
        rc = snb.snb_render_nerf_with_buffers(ctx, C.byref(v), None, C.byref(rt), 1, b[0].data_ptr(), b[1].data_ptr(), b[2].data_ptr(), b[3].data_ptr())
        assert rc == 0, snb.snb_last_error(ctx)
        return _alt_status(snb, ctx)
    try:
        # reference frames: speculation only, repeated until converged
        assert snb.snb_set_option(ctx, b"alt_schedule", b"auto") == 0 and snb.snb_set_option(ctx, b"alt_auto_lockstep", b"0") == 0
        want = []
        for cam in cams:
            v = common.make_view(w, h, cam=cam)
            b = bufs()
            for _ in range(6):
                st = render(v, b)
                if st["converged"]:
                    break
            assert st["converged"] == 1
            want.append(([x.cpu().numpy() for x in b[:3]], st))
        multi = sum(1 for _, st in want if st["passes"] > 1)
        # lock step, cold start, then along the same camera sequence: one pass each, identical results
        assert snb.snb_set_alt_schedule(ctx, None, 0) == 0
        assert snb.snb_set_option(ctx, b"alt_schedule", b"lockstep") == 0
        for cam, (ref, rst) in zip(cams, want):
            v = common.make_view(w, h, cam=cam)
            b = bufs()
            st = render(v, b)
            prefix = int((rst["sched"] < 8).sum())
            if prefix <= 16:                      # kLockAttempts: a longer prefix falls back to verified speculation (more passes, same frame)
                assert st["passes"] == 1 and st["converged"] == 1, st
            for _ in range(6):
                if st["converged"]:
                    break
                st = render(v, b)
            assert st["converged"] == 1, st
            k = 24
            np.testing.assert_array_equal(st["sched"][:k], rst["sched"][:k]); np.testing.assert_array_equal(st["alive"][:k], rst["alive"][:k])
            for got, exp in zip(b[:3], ref):
                assert np.array_equal(got.cpu().numpy(), exp)
        if target < 2097152:
            assert min(want[0][1]["sched"][:4]) < 8, "degenerate test: no prefix"
        # "auto" picks the lock-step prefix when the camera moved and plain speculation when it did not
        assert snb.snb_set_option(ctx, b"alt_schedule", b"auto") == 0 and snb.snb_set_option(ctx, b"alt_auto_lockstep", b"1") == 0
        for cam, (ref, rst) in zip(cams, want):
            b = bufs()
            v = common.make_view(w, h, cam=cam)
            st = render(v, b)
            for _ in range(6):
                if st["converged"]:
                    break
                st = render(v, b)
            assert st["converged"] == 1 and (st["passes"] == 1 or int((rst["sched"] < 8).sum()) > 16), st
            assert np.array_equal(b[0].cpu().numpy(), ref[0])
        print(f"[lockstep {kind} {w}x{h}] frames that needed several speculative passes: {multi} of {len(cams)}")
    finally:
        snb.snb_set_option(ctx, b"alt_schedule", b"auto"); snb.snb_set_option(ctx, b"alt_auto_lockstep", b"1"); snb.snb_set_option(ctx, b"alt_target_queries", b"2097152")
        snb.snb_set_alt_schedule(ctx, None, 0)


def test_two_contexts_on_two_devices_in_one_process(snb):
    """launch set-up (dynamic shared-memory opt-in, occupancy) is cached per CONTEXT, not per process, so a second context on
    another device renders like the first. Needs two GPUs (skipped on a single-GPU box)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    m = common.model("small16x2")
    v = common.make_view(128, 96)
    rt = common.default_rt(show_shadow_on_nerf=0)
    frames = []
    ctxs = []
    try:
        for devi in (0, 1):
            c = C.c_void_p(); assert snb.snb_create(devi, None, C.byref(c)) == 0
            ctxs.append(c)
            d = m.desc(); assert snb.snb_set_model(c, C.byref(d)) == 0
        for devi, c in enumerate(ctxs):                      # both contexts alive at the same time
            with torch.cuda.device(devi):
                fb = torch.zeros((96, 128, 4), device=f"cuda:{devi}"); depth = torch.zeros((96, 128), device=f"cuda:{devi}")
                pos = torch.zeros((96, 128, 3), device=f"cuda:{devi}"); nrm = torch.zeros((96, 128, 3), device=f"cuda:{devi}")
                torch.cuda.synchronize(devi)
                for _ in range(2):
                    rc = snb.snb_render_nerf_with_buffers(c, C.byref(v), None, C.byref(rt), 1, fb.data_ptr(), depth.data_ptr(), pos.data_ptr(), nrm.data_ptr())
                    assert rc == 0, snb.snb_last_error(c)
                assert snb.snb_sync(c) == 0
                sfb = torch.zeros((96, 128, 4), device=f"cuda:{devi}"); sd = torch.zeros((96, 128), device=f"cuda:{devi}")
                torch.cuda.synchronize(devi)
                assert snb.snb_render_nerf(c, C.byref(v), None, 1, sfb.data_ptr(), sd.data_ptr()) == 0 and snb.snb_sync(c) == 0
                frames.append((fb.cpu().numpy(), sfb.cpu().numpy()))
    finally:
        for c in ctxs:
            snb.snb_destroy(c)
    assert frames[0][0][..., 3].max() > 0.5
    assert np.array_equal(frames[0][0], frames[1][0]) and np.array_equal(frames[0][1], frames[1][1])


@pytest.mark.parametrize("kind,w,h", [("small16x2", 128, 96), ("big16x2", 96, 64)])
def test_fused_march_kernel_samples_equal_the_trace(snb, orc, ctx, kind, w, h):
    """k_nerf_march_fused ITSELF (not the separate trace kernel) reports the samples it evaluates (snb_set_sample_tap):
    per ray they are a prefix of the untruncated sequence of snb_march_trace — t bit for bit, same cascade cell — which
    test_march_samples_bit_exact pins to the oracle; the prefix ends where the compositor terminated the ray."""
    m = common.model(kind)
    set_model(snb, ctx, m)
    assert snb.snb_set_option(ctx, b"march", b"lane") == 0
    cam = common.camera_matrix(zoom=2.5) if kind == "big16x2" else None
    v = common.make_view(w, h, cam=cam, spp=2, snap=0)
    n, S = w * h, 48
    t, dt, cell, cnt = dev((n, S), torch.float32), dev((n, S), torch.float32), dev((n, S), torch.int32), dev((n,), torch.int32)
    assert snb.snb_march_trace(ctx, C.byref(v), S, t.data_ptr(), dt.data_ptr(), cell.data_ptr(), cnt.data_ptr(), None) == 0 and snb.snb_sync(ctx) == 0
    ft, fc, fn = dev((n, S), torch.float32), dev((n, S), torch.int32), dev((n,), torch.int32)
    fb, depth = dev((h, w, 4), torch.float32), dev((h, w), torch.float32)
    assert snb.snb_set_sample_tap(ctx, ft.data_ptr(), fc.data_ptr(), fn.data_ptr(), S) == 0
    try:
        assert snb.snb_render_nerf(ctx, C.byref(v), None, 1, fb.data_ptr(), depth.data_ptr()) == 0 and snb.snb_sync(ctx) == 0
    finally:
        assert snb.snb_set_sample_tap(ctx, None, None, None, 0) == 0
    t, cell, cnt = t.cpu().numpy(), cell.cpu().numpy(), cnt.cpu().numpy()
    ft, fc, fn = ft.cpu().numpy(), fc.cpu().numpy(), fn.cpu().numpy()
    assert (fn > 0).sum() > 200 and np.array_equal(fn > 0, cnt > 0)
    k = np.minimum(np.minimum(fn, cnt), S)
    assert (fn <= np.where(cnt < S, cnt, fn)).all()                   # never more samples than the sequence holds
    mask = np.arange(S)[None, :] < k[:, None]
    assert np.array_equal(ft.view(np.uint32)[mask], t.view(np.uint32)[mask])
    assert np.array_equal(fc[mask], cell[mask])
    assert (fn < cnt).sum() > 50, "degenerate test: no ray was terminated by the compositor"
    # the tap is off again: the next render leaves the arrays alone
    before = ft.copy()
    assert snb.snb_render_nerf(ctx, C.byref(v), None, 1, fb.data_ptr(), depth.data_ptr()) == 0 and snb.snb_sync(ctx) == 0
    assert np.array_equal(before, ft)
