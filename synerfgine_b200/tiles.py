"""Multi-GPU frame partition: one process per GPU, each renders a column strip of the frame plus the halo
its screen-space stencils read, and rank 0 assembles the strips with one gather per frame. No collective inside the march.

The reference has no multi-GPU split of a frame (its only multi-GPU mode is one VR view per GPU with cudaMemcpyPeerAsync,
src/testbed.cu:5045-5132); rays are independent, so the partition is pure host logic:

* pixel indices used for sampling (ld_random_val(spp, idx * 786433), curand subsequence) stay GLOBAL, so a strip's pixels are
  bit-identical to the same pixels of a single-GPU frame;
* write_normals_to_buffer reads positions at +-1/+-2 px (src/testbed_nerf.cu:1542-1552) and shade_with_shadow a window of
  +-(nerf_shadow_samples // 2) px (:1750-1765): each rank therefore marches `halo_for(...)` extra columns on both sides and
  crops them away before the gather (recompute, no exchange);
* strip edges are multiples of syn_px_scale so overlay_nerf's nearest-neighbour nc = sc / syn_px_scale never straddles ranks.

This module is torch.distributed-only plumbing (NCCL over NVLink on the GPU box, gloo in the CPU tests); the renderer is
passed in as a callable so the same code path is exercised on both.
"""
from dataclasses import dataclass


def halo_for(nerf_shadow_kernel, show_shadow_on_nerf=True, syn_px_scale=1):
    """Columns (or rows) a strip must recompute beyond its own, in FRAME (mesh-layer) pixels: the normals stencil (2) + the shadow
    window (k // 2) run on the NeRF layer, whose pixels are `syn_px_scale` frame pixels wide -> (k // 2 + 2) * syn_px_scale."""
    return max(2, (int(nerf_shadow_kernel) // 2 if show_shadow_on_nerf else 0) + 2) * max(1, int(syn_px_scale))


@dataclass(frozen=True)
class Strip:
    x0: int       # first owned column
    x1: int       # one past the last owned column
    hx0: int      # first rendered column (owned - halo, clamped)
    hx1: int      # one past the last rendered column

    @property
    def width(self):
        return self.x1 - self.x0

    @property
    def render_width(self):
        return self.hx1 - self.hx0

    @property
    def crop(self):
        """Slice of the rendered strip that is owned."""
        return slice(self.x0 - self.hx0, self.x1 - self.hx0)


def strip_for(rank, world, width, halo, align=1):
    """Column strip of `rank` out of `world` for a `width`-pixel frame. Owned ranges tile [0, width) exactly; edges are
    multiples of `align` (syn_px_scale); the halo is clamped at the image border (where the single-GPU stencils stop too)."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    if width % align:
        raise ValueError("width must be a multiple of the alignment")
    units = width // align
    x0 = (units * rank // world) * align
    x1 = (units * (rank + 1) // world) * align
    h = -(-halo // align) * align          # halo rounded up to the alignment
    return Strip(x0, x1, max(0, x0 - h), min(width, x1 + h))


def frame_batch_for(rank, world, n_frames):
    """Frame-batch mode (config C5, 64 frames on 8 GPUs): contiguous frames per rank, no halo."""
    return range(n_frames * rank // world, n_frames * (rank + 1) // world)


def render_strip_and_gather(render_tile, rank, world, width, height, halo, align=1, dist=None, device=None):
    """Render this rank's strip with `render_tile(x0, y0, w, h) -> (rgba [h,w,4], depth [h,w])` (torch tensors) and gather the
    owned columns on rank 0. Returns (rgba [H,W,4], depth [H,W]) on rank 0, (None, None) elsewhere. With world == 1 no
    collective is issued."""
    import torch
    s = strip_for(rank, world, width, halo, align)
    rgba, depth = render_tile(s.hx0, 0, s.render_width, height)
    mine = torch.cat([rgba[:, s.crop, :], depth[:, s.crop, None]], dim=2).contiguous()     # [H, w_owned, 5]: one message per rank
    if world == 1:
        return mine[..., :4], mine[..., 4]
    # dist.gather needs equally sized messages: pad every strip to the widest one, crop on rank 0
    widths = [strip_for(r, world, width, halo, align).width for r in range(world)]
    wmax = max(widths)
    if mine.shape[1] < wmax:
        mine = torch.cat([mine, mine.new_zeros((height, wmax - mine.shape[1], 5))], dim=1).contiguous()
    bufs = [torch.empty_like(mine) for _ in range(world)] if rank == 0 else None
    dist.gather(mine, bufs, dst=0)
    if rank != 0:
        return None, None
    full = torch.cat([b[:, :w, :] for b, w in zip(bufs, widths)], dim=1)
    return full[..., :4], full[..., 4]


# ---- load-balanced strips ------------------------------------------------------------------------------------------------
# Equal-width strips leave most GPUs idle when the NeRF sits in the middle of the frame (N=4, C2: the two outer strips finish
# in 1.8 ms, the two inner ones in 4.1 ms). The partition below moves the strip edges so that every rank gets the same share
# of the measured cost: each rank contributes `its frame time x (per-column hit profile of its strip)`, one all_reduce of a
# `width`-float vector (7.7 KB at 1080p) makes the profile global, and every rank derives the same edges from it.

def balanced_bounds(col_cost, world, align=1, min_width=None):
    """Edges b[0..world] (b[0] = 0, b[world] = width, multiples of `align`, strips at least `min_width` wide) such that the
    cost of every strip is as close to total / world as the alignment allows. `col_cost`: non-negative cost per column."""
    import numpy as np
    c = np.asarray(col_cost, dtype=np.float64)
    width = c.shape[0]
    if width % align:
        raise ValueError("width must be a multiple of the alignment")
    units = width // align
    min_units = max(1, -(-(min_width or align) // align))
    if min_units * world > units:
        raise ValueError("frame too narrow for this many strips")
    cu = c.reshape(units, align).sum(axis=1)
    cu = cu + max(cu.sum(), 1e-30) * 1e-9                      # strictly increasing prefix sums, even over empty columns
    prefix = np.concatenate([[0.0], np.cumsum(cu)])
    bounds = [0]
    for r in range(1, world):
        target = prefix[-1] * r / world
        u = int(np.searchsorted(prefix, target))                # first edge with prefix >= target
        if u > 0 and target - prefix[u - 1] < prefix[u] - target:
            u -= 1
        u = max(u, bounds[-1] + min_units)                      # keep this strip wide enough ...
        u = min(u, units - min_units * (world - r))             # ... and room for the ones to come
        bounds.append(u)
    bounds.append(units)
    return [b * align for b in bounds]


def strip_from_bounds(bounds, rank, width, halo, align=1):
    h = -(-halo // align) * align
    return Strip(bounds[rank], bounds[rank + 1], max(0, bounds[rank] - h), min(width, bounds[rank + 1] + h))


class StripBalancer:
    """Per-frame feedback on the strip edges. `observe(col_cost)` takes the GLOBAL per-column cost (already all-reduced) and
    moves the edges only when the predicted slowest strip gets at least `gain` (default 8 %; twice that once the split has
    settled) faster. A move is cheap — generator states are per global pixel and scratch is sized for the full frame, so nothing
    is re-seeded or reallocated — but it restarts the mesh layer's temporal blend, so the frame that follows a move is not used
    as a measurement (`due()` is False for it)."""

    def __init__(self, width, world, halo, align=1, gain=0.08):
        self.width, self.world, self.halo, self.align, self.gain = width, world, halo, align, gain
        self.min_width = max(align, 2 * (-(-halo // align) * align))
        self.bounds = [strip_for(r, world, width, halo, align).x0 for r in range(world)] + [width]
        self.frames = 0
        self.changes = 0
        self.cooldown = 0

    def strip(self, rank):
        return strip_from_bounds(self.bounds, rank, self.width, self.halo, self.align)

    def due(self):
        """Balance on each of the first 16 frames (except right after a move), then every 8th."""
        return self.world > 1 and self.cooldown == 0 and (self.frames < 16 or self.frames % 8 == 0)

    def tick(self):
        self.frames += 1
        if self.cooldown:
            self.cooldown -= 1

    @staticmethod
    def _slowest(col_cost, bounds):
        return max(float(col_cost[a:b].sum()) for a, b in zip(bounds, bounds[1:]))

    def observe(self, col_cost):
        new = balanced_bounds(col_cost, self.world, self.align, self.min_width)
        gain = self.gain if self.frames < 16 else 2.0 * self.gain
        if new != self.bounds and self._slowest(col_cost, new) < (1.0 - gain) * self._slowest(col_cost, self.bounds):
            self.bounds = new
            self.changes += 1
            self.cooldown = 2          # tick() of this frame takes one off: the NEXT frame is the one skipped
            return True
        return False


def column_cost_of_strip(strip, width, frame_ms, hits_per_column, floor=0.05):
    """This rank's contribution to the global cost vector: its measured frame time spread over its OWNED columns in
    proportion to `floor * max + hits` (hits = pixels of the column that any layer covered); zero elsewhere."""
    import numpy as np
    h = np.asarray(hits_per_column, dtype=np.float64)
    w = h + floor * max(float(h.max()) if h.size else 0.0, 1.0)
    out = np.zeros(width, dtype=np.float64)
    out[strip.x0:strip.x1] = frame_ms * w / w.sum()
    return out


def assemble_on_root(mine, bounds, rank, world, dist, frame=None):
    """Owned columns of every rank -> one [H, W, C] frame on rank 0, with exact-size point-to-point messages (strips are not
    equally wide any more, so dist.gather does not apply). Returns the frame on rank 0, None elsewhere."""
    import torch
    if world == 1:
        return mine
    batched = dist.get_backend() == "nccl"      # one grouped NCCL launch instead of world-1 serialised p2p ops
    if rank != 0:
        if batched:
            for req in dist.batch_isend_irecv([dist.P2POp(dist.isend, mine, 0)]):
                req.wait()
        else:
            dist.send(mine, dst=0)
        return None
    h, c = mine.shape[0], mine.shape[2]
    if frame is None:
        frame = torch.empty((h, bounds[-1], c), dtype=mine.dtype, device=mine.device)
    parts = [torch.empty((h, bounds[r + 1] - bounds[r], c), dtype=mine.dtype, device=mine.device) for r in range(1, world)]
    if batched:
        reqs = dist.batch_isend_irecv([dist.P2POp(dist.irecv, p, r) for r, p in zip(range(1, world), parts)])
    else:
        reqs = [dist.irecv(p, src=r) for r, p in zip(range(1, world), parts)]
    frame[:, bounds[0]:bounds[1]] = mine
    for req in reqs:
        req.wait()
    for r, p in enumerate(parts, start=1):
        frame[:, bounds[r]:bounds[r + 1]] = p
    return frame


def assemble_rows_on_root(parts, bounds, rank, world, dist, frames):
    """Row bands: every rank owns rows [bounds[r], bounds[r+1]) of the frame, which are CONTIGUOUS in a row-major frame, so each
    buffer of a rank (rgba [h, W, 4], depth [h, W]: 20 B per pixel) travels as one exact-size point-to-point message
    straight into its place in rank 0's frame — no .contiguous() copy on the sender, no staging tensor and no second copy on the
    receiver. `parts`: this rank's owned rows of each buffer (contiguous views); `frames`: the full-size buffers on rank 0 (ignored
    elsewhere). All messages of a frame go out in one grouped NCCL launch. Returns `frames` on rank 0, None elsewhere."""
    if world == 1:
        for f, p in zip(frames, parts):
            if f.data_ptr() != p.data_ptr():
                f[bounds[0]:bounds[1]].copy_(p)
        return frames
    for p in parts:
        assert p.is_contiguous(), "row bands are contiguous: slice rows, not columns"
    batched = dist.get_backend() == "nccl"
    if rank != 0:
        if batched:
            for req in dist.batch_isend_irecv([dist.P2POp(dist.isend, p, 0) for p in parts]):
                req.wait()
        else:
            for p in parts:
                dist.send(p, dst=0)
        return None
    views = [[f[bounds[r]:bounds[r + 1]] for f in frames] for r in range(world)]
    if batched:
        reqs = dist.batch_isend_irecv([dist.P2POp(dist.irecv, v, r) for r in range(1, world) for v in views[r]])
    else:
        reqs = [dist.irecv(v, src=r) for r in range(1, world) for v in views[r]]
    for v, p in zip(views[0], parts):
        if v.data_ptr() != p.data_ptr():
            v.copy_(p)
    for req in reqs:
        req.wait()
    return frames
