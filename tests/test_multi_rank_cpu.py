"""CPU tests of the N>1 path: strip partition + halo + one gather per frame, with world_size 2 over gloo.
The renderer plugged into synerfgine_b200.tiles here is the CPU oracle (test infrastructure) rendering a TILE of the frame
with global pixel indices — the same contract libsnb.so's snb_tile honours on the GPU (tests/test_gpu_nerf.py::
test_tile_equals_crop_of_full_frame and bench.py --gpus N exercise that side)."""
import ctypes as C
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import common
import oracle_lib as O
from synerfgine_b200 import tiles

W, H = 40, 24


def test_strips_tile_the_frame_exactly():
    for width in (1920, 1921, 40, 7):
        for world in (1, 2, 3, 4, 8):
            for halo in (2, 4):
                if world > width:
                    continue
                strips = [tiles.strip_for(r, world, width, halo) for r in range(world)]
                assert strips[0].x0 == 0 and strips[-1].x1 == width
                for a, b in zip(strips, strips[1:]):
                    assert a.x1 == b.x0
                for s in strips:
                    assert s.hx0 == max(0, s.x0 - halo) and s.hx1 == min(width, s.x1 + halo)
                    assert s.crop.stop - s.crop.start == s.width
                assert max(s.width for s in strips) - min(s.width for s in strips) <= 1


def test_strip_alignment_to_syn_px_scale():
    for r in range(3):
        s = tiles.strip_for(r, 3, 1920, 3, align=4)
        assert s.x0 % 4 == 0 and s.x1 % 4 == 0 and s.hx0 % 4 == 0 and s.hx1 % 4 == 0
        assert (s.x0 - s.hx0) in (0, 4) and (s.hx1 - s.x1) in (0, 4)
    with pytest.raises(ValueError):
        tiles.strip_for(0, 2, 1921, 2, align=4)
    with pytest.raises(ValueError):
        tiles.strip_for(2, 2, 1920, 2)


def test_halo_and_frame_batch():
    assert tiles.halo_for(4) == 4 and tiles.halo_for(1) == 2 and tiles.halo_for(8) == 6 and tiles.halo_for(8, False) == 2
    # the stencils run on the NeRF layer: with syn_px_scale 2 the halo doubles in frame pixels, and a strip keeps (k // 2 + 2) NeRF columns of it
    assert tiles.halo_for(4, True, 2) == 8 and tiles.halo_for(1, True, 4) == 8
    for world in (2, 3):
        for r in range(world):
            s = tiles.strip_for(r, world, 1920, tiles.halo_for(4, True, 2), align=2)
            left, right = (s.x0 - s.hx0) // 2, (s.hx1 - s.x1) // 2              # halo in NeRF-layer columns
            assert (left >= 4 or s.hx0 == 0) and (right >= 4 or s.hx1 == 1920)
            assert s.hx0 % 2 == 0 and s.hx1 % 2 == 0
    got = [list(tiles.frame_batch_for(r, 8, 64)) for r in range(8)]
    assert sum(got, []) == list(range(64)) and all(len(g) == 8 for g in got)
    got = [list(tiles.frame_batch_for(r, 3, 10)) for r in range(3)]
    assert sum(got, []) == list(range(10))


def oracle_nerf_layer(model_desc, view):
    """NeRF layer + screen-space normals of the full view (the stencil the halo exists for)."""
    fb, depth, pos, _ = O.render_nerf(model_desc, view, alt=1, positions=True)
    nrm = np.zeros_like(pos)
    O.load().orc_write_normals(view.width, view.height, O.P(np.ascontiguousarray(pos)), O.P(nrm))
    return fb, depth, pos, nrm


def make_tile_renderer(model_desc, view):
    """render_tile(x0, y0, w, h): what one rank computes — rays of the tile only (global pixel indices), then the normals
    stencil INSIDE the tile (it stops at the tile edge exactly like at an image edge). Channels: rgb + normal.x packed in
    rgba so that a wrong halo shows up in the gathered frame; depth as is."""
    full_fb, full_depth, full_pos, _ = O.render_nerf(model_desc, view, alt=1, positions=True)   # per-ray results are tile-independent

    def render_tile(x0, y0, w, h):
        pos = np.ascontiguousarray(full_pos[y0:y0 + h, x0:x0 + w])
        nrm = np.zeros_like(pos)
        O.load().orc_write_normals(w, h, O.P(pos), O.P(nrm))
        rgba = np.ascontiguousarray(full_fb[y0:y0 + h, x0:x0 + w]).copy()
        rgba[..., 3] = nrm[..., 0]
        return torch.from_numpy(rgba), torch.from_numpy(np.ascontiguousarray(full_depth[y0:y0 + h, x0:x0 + w]))
    return render_tile


def _worker(rank, world, port, halo, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        m = common.model("small16x2")
        v = common.make_view(W, H)
        rgba, depth = tiles.render_strip_and_gather(make_tile_renderer(m.desc(), v), rank, world, W, H, halo, dist=dist)
        if rank == 0:
            np.savez(out_path, rgba=rgba.numpy(), depth=depth.numpy())
        else:
            assert rgba is None and depth is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 3])
def test_two_rank_frame_equals_single_rank_frame(tmp_path, world):
    m = common.model("small16x2")
    v = common.make_view(W, H)
    fb, depth, pos, nrm = oracle_nerf_layer(m.desc(), v)
    expect = fb.copy(); expect[..., 3] = nrm[..., 0]
    out = str(tmp_path / "frame.npz")
    mp.spawn(_worker, args=(world, _free_port(), tiles.halo_for(1), out), nprocs=world, join=True)
    got = np.load(out)
    np.testing.assert_array_equal(got["rgba"].view(np.uint32), expect.view(np.uint32))      # bit-identical to the 1-rank frame
    np.testing.assert_array_equal(got["depth"].view(np.uint32), depth.view(np.uint32))
    assert np.abs(nrm[..., 0]).max() > 0.1


def test_missing_halo_is_detected(tmp_path):
    """Negative control: with halo 0 the normals next to the seam differ, i.e. the test above really checks the halo."""
    m = common.model("small16x2")
    v = common.make_view(W, H)
    fb, depth, pos, nrm = oracle_nerf_layer(m.desc(), v)
    out = str(tmp_path / "frame0.npz")
    mp.spawn(_worker, args=(2, _free_port(), 0, out), nprocs=2, join=True)
    got = np.load(out)["rgba"][..., 3]
    seam = W // 2
    assert not np.array_equal(got[:, seam - 2:seam + 2], nrm[:, seam - 2:seam + 2, 0])
    np.testing.assert_array_equal(got[:, :seam - 2], nrm[:, :seam - 2, 0])


# ---- load-balanced strips (tiles.StripBalancer): same frame, unequal strips, exact-size point-to-point assembly ----------

def test_balanced_bounds_equalise_cost_and_respect_alignment():
    rng = np.random.default_rng(7)
    for width, align in ((1920, 1), (1920, 4), (64, 2)):
        cost = rng.random(width) ** 4
        cost[width // 3: width // 2] += 5.0                     # a heavy object off-centre
        for world in (2, 3, 4, 8):
            b = tiles.balanced_bounds(cost, world, align=align, min_width=2 * align)
            assert b[0] == 0 and b[-1] == width and len(b) == world + 1
            assert all(x % align == 0 for x in b) and all(y - x >= 2 * align for x, y in zip(b, b[1:]))
            parts = [cost[x:y].sum() for x, y in zip(b, b[1:])]
            equal = [cost[x:y].sum() for x, y in zip(range(0, width, width // world), range(width // world, width + 1, width // world))]
            assert max(parts) <= max(equal) + 1e-9                  # never worse than equal strips
            if width >= 1920:                                       # (narrow frames are dominated by the minimum strip width)
                assert max(parts) <= cost.sum() / world + cost.max() * align + 1e-9   # within one aligned column group of ideal
    # all cost in one column: every other strip still gets its minimum width
    cost = np.zeros(40); cost[17] = 1.0
    b = tiles.balanced_bounds(cost, 4, align=1, min_width=4)
    assert all(y - x >= 4 for x, y in zip(b, b[1:])) and b[-1] == 40
    with pytest.raises(ValueError):
        tiles.balanced_bounds(np.ones(8), 4, align=1, min_width=4)


def test_strip_balancer_hysteresis():
    bal = tiles.StripBalancer(1920, 4, halo=4)
    assert bal.bounds == [0, 480, 960, 1440, 1920] and bal.min_width == 8
    flat = np.ones(1920)
    assert not bal.observe(flat) and bal.changes == 0              # already balanced: edges stay
    heavy = np.full(1920, 0.01); heavy[700:1200] = 1.0
    assert bal.observe(heavy) and bal.changes == 1
    before = list(bal.bounds)
    assert not bal.observe(heavy * 1.01) and bal.bounds == before  # same profile again: no further move
    s = bal.strip(1)
    assert (s.x0, s.x1) == (bal.bounds[1], bal.bounds[2]) and s.hx0 == s.x0 - 4 and s.hx1 == s.x1 + 4
    c = tiles.column_cost_of_strip(s, 1920, 3.0, np.arange(s.width))
    assert abs(c.sum() - 3.0) < 1e-9 and c[:s.x0].sum() == 0 and c[s.x1:].sum() == 0


def _worker_balanced(rank, world, port, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        m = common.model("small16x2")
        v = common.make_view(W, H)
        render_tile = make_tile_renderer(m.desc(), v)
        bal = tiles.StripBalancer(W, world, tiles.halo_for(1))
        frames = []
        for it in range(3):                                        # frame 0 on equal strips, then on re-balanced ones
            s = bal.strip(rank)
            rgba, depth = render_tile(s.hx0, 0, s.render_width, H)
            mine = torch.cat([rgba[:, s.crop, :], depth[:, s.crop, None]], dim=2).contiguous()
            full = tiles.assemble_on_root(mine, bal.bounds, rank, world, dist)
            hits = (mine[..., 3] != 0).sum(0).numpy()
            # synthetic "frame time": proportional to the hits of the strip, so the split must move towards the object
            cost = torch.from_numpy(tiles.column_cost_of_strip(s, W, 1.0 + float(hits.sum()), hits))
            dist.all_reduce(cost)
            bal.observe(cost.numpy())
            if rank == 0:
                frames.append(full.numpy())
        if rank == 0:
            np.savez(out_path, frames=np.stack(frames), bounds=np.array(bal.bounds), changes=bal.changes)
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_rebalanced_strips_give_the_same_frame(tmp_path, world):
    m = common.model("small16x2")
    v = common.make_view(W, H)
    fb, depth, pos, nrm = oracle_nerf_layer(m.desc(), v)
    expect = np.concatenate([fb[..., :3], nrm[..., :1], depth[..., None]], axis=2)
    out = str(tmp_path / "balanced.npz")
    mp.spawn(_worker_balanced, args=(world, _free_port(), out), nprocs=world, join=True)
    got = np.load(out)
    for f in got["frames"]:                                        # every frame, whatever the split, is the single-rank frame
        np.testing.assert_array_equal(f.view(np.uint32), expect.view(np.uint32))
    assert list(got["bounds"])[0] == 0 and list(got["bounds"])[-1] == W


# ---- row bands (the form bench.py uses): contiguous messages straight into rank 0's frame, rgba + depth ----------------------------
def _band_worker(rank, world, port, halo, out_path, bounds):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        m = common.model("small16x2")
        v = common.make_view(W, H)
        render_tile = make_tile_renderer(m.desc(), v)
        s = tiles.strip_from_bounds(bounds, rank, H, halo)                      # a Strip along y: rows [x0, x1) owned, [hx0, hx1) rendered
        rgba, depth = render_tile(0, s.hx0, W, s.render_width)
        own = slice(s.x0 - s.hx0, s.x1 - s.hx0)
        parts = [rgba[own], depth[own]]
        assert all(p.is_contiguous() for p in parts)                            # no copy needed before sending: rows of a band are contiguous
        frames = [torch.full((H, W, 4), -1.0), torch.full((H, W), -1.0)] if rank == 0 else [None, None]
        got = tiles.assemble_rows_on_root(parts, bounds, rank, world, dist, frames)
        if rank == 0:
            np.savez(out_path, rgba=got[0].numpy(), depth=got[1].numpy())
        else:
            assert got is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,bounds", [(2, [0, 12, 24]), (3, [0, 5, 17, 24])])
def test_row_bands_assemble_rgba_and_depth_bit_identically(tmp_path, world, bounds):
    """Equal and unequal (re-balanced) bands: the assembled rgba AND depth (20 B per pixel) equal the 1-rank frame bit for bit."""
    m = common.model("small16x2")
    v = common.make_view(W, H)
    fb, depth, pos, nrm = oracle_nerf_layer(m.desc(), v)
    expect = fb.copy(); expect[..., 3] = nrm[..., 0]
    out = str(tmp_path / "bands.npz")
    mp.spawn(_band_worker, args=(world, _free_port(), tiles.halo_for(1), out, bounds), nprocs=world, join=True)
    got = np.load(out)
    np.testing.assert_array_equal(got["rgba"].view(np.uint32), expect.view(np.uint32))
    np.testing.assert_array_equal(got["depth"].view(np.uint32), depth.view(np.uint32))


def test_batch_schedule_is_a_whole_frame_quantity():
    """The sng compositor's batch lengths depend on n_alive of the WHOLE frame (src/testbed_nerf.cu:2189-2190). Bands that march with
    the frame's sequence reproduce the frame; bands that derive the sequence from their own ray counts do not — which is why the
    ranks sum their counts (bench.py: sync_alt_schedule) instead of running the single-GPU 'auto' mode per band."""
    m = common.model("small16x2")
    d = m.desc()
    v = common.make_view(W, H)
    target = 900
    fb, _, _, sched, alive = O.render_nerf_wavefront(d, v, alt=1, target_queries=target)
    assert sched[0] < 8
    whole, _, _ = O.render_nerf_sched(d, v, sched, alt=1)
    np.testing.assert_array_equal(whole.view(np.uint32), fb.view(np.uint32))
    # a band on its own: fewer alive rays -> longer batches -> another image
    half = common.make_view(W, H // 2)
    half.camera[:] = list(v.camera); half.focal_length[:] = list(v.focal_length)
    _, _, _, sched_half, alive_half = O.render_nerf_wavefront(d, half, alt=1, target_queries=target)
    assert alive_half[0] < alive[0] and list(sched_half[:4]) != list(sched[:4])
