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


def halo_for(nerf_shadow_kernel, show_shadow_on_nerf=True):
    """Columns a strip must recompute beyond its own: normals stencil (2) + shadow window (k // 2) read normals -> k // 2 + 2."""
    return max(2, (int(nerf_shadow_kernel) // 2 if show_shadow_on_nerf else 0) + 2)


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
