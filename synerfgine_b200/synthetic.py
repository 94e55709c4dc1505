"""Seeded synthetic NeRF models of the base.json architecture family (SURVEY.md §8d "Synthetic inputs").

No trained snapshot ships with the reference and there is no network access, so benchmarks and parity tests use
random-init weights of the exact architecture plus an authored occupancy field:
  * density MLP 32->64->16, rgb MLP 32->64->64->16(3 used), Xavier-uniform, fp16
    (reference init: dependencies/tiny-cuda-nn/src/fully_fused_mlp.cu:866-890);
  * hash-grid features U(-0.5, 0.5) (the reference's +-1e-4 init gives a constant field);
  * analytic density (three soft spheres + a ground slab) sampled on the 128^3 x (max_cascade+1) Morton grid
    and thresholded / max-pooled by the rules of grid_to_bitfield / bitfield_max_pool
    (reference src/testbed_nerf.cu:285-332,3212-3229). The numpy implementation here is the host-side twin of
    the CUDA kernels behind snb_density_to_bitfield; tests compare the two.
"""
import math
import numpy as np

GRID = 128
CELLS = GRID ** 3
CASCADES = 8
MLP_HALVES = 64 * 32 + 16 * 64 + 64 * 32 + 64 * 64 + 16 * 64


def grid_layout(n_levels, n_feat, log2_T, base_res, per_level_scale):
    """Level scales / resolutions / offsets exactly as tcnn computes them (grid.h:692-723)."""
    log2_pls = np.float32(np.log2(np.float32(per_level_scale)))
    scales, ress, offs = [], [], [0]
    for i in range(n_levels):
        scale = np.float32(np.exp2(np.float32(i) * log2_pls) * np.float32(base_res) - np.float32(1.0))
        res = int(math.ceil(float(scale))) + 1
        n = min(res ** 3, 0xFFFFFFFF // 2)
        n = (n + 7) // 8 * 8
        n = min(n, 1 << log2_T)
        scales.append(scale); ress.append(res); offs.append(offs[-1] + n)
    return scales, ress, offs


def per_level_scale_for(aabb_scale, base_res, n_levels, desired=2048):
    """reset_network (reference src/testbed.cu:3739): exp(log(desired_res * aabb_scale / base_res) / (L - 1))."""
    return float(np.exp(np.log(np.float32(desired * aabb_scale) / np.float32(base_res)) / np.float32(n_levels - 1)).astype(np.float32))


def _compact3(x):
    x = x & 0x49249249
    x = (x | (x >> 2)) & 0xC30C30C3
    x = (x | (x >> 4)) & 0x0F00F00F
    x = (x | (x >> 8)) & 0xFF0000FF
    x = (x | (x >> 16)) & 0x0000FFFF
    return x


def _spread3(v):
    v = (v * 0x00010001) & 0xFF0000FF
    v = (v * 0x00000101) & 0x0F00F00F
    v = (v * 0x00000011) & 0xC30C30C3
    v = (v * 0x00000005) & 0x49249249
    return v


def morton3(x, y, z):
    return _spread3(x.astype(np.uint64)) | (_spread3(y.astype(np.uint64)) << 1) | (_spread3(z.astype(np.uint64)) << 2)


def analytic_density(p):
    """Density field in world units: three soft spheres + a ground slab inside the unit cube."""
    spheres = [((0.5, 0.52, 0.5), 0.16), ((0.33, 0.45, 0.62), 0.09), ((0.68, 0.42, 0.38), 0.11)]
    d = np.zeros(p.shape[0], dtype=np.float32)
    for c, r in spheres:
        dist = np.sqrt(((p - np.asarray(c, dtype=np.float32)) ** 2).sum(-1))
        d = np.maximum(d, np.clip((r - dist) / 0.02, 0.0, 1.0).astype(np.float32))
    slab = (p[:, 1] > 0.30) & (p[:, 1] < 0.335) & (np.abs(p[:, 0] - 0.5) < 0.42) & (np.abs(p[:, 2] - 0.5) < 0.42)
    d = np.maximum(d, slab.astype(np.float32))
    return d.astype(np.float32)


def density_grid(max_cascade, field=analytic_density):
    """fp32 density per cell, Morton order, cascades 0..max_cascade concatenated (what the snapshot stores as fp16)."""
    idx = np.arange(CELLS, dtype=np.uint64)
    x = _compact3(idx).astype(np.float32); y = _compact3(idx >> 1).astype(np.float32); z = _compact3(idx >> 2).astype(np.float32)
    unit = (np.stack([x, y, z], -1) + np.float32(0.5)) / np.float32(GRID) - np.float32(0.5)
    out = np.empty(CELLS * (max_cascade + 1), dtype=np.float32)
    for m in range(max_cascade + 1):
        out[m * CELLS:(m + 1) * CELLS] = field((unit * np.float32(2.0 ** m) + np.float32(0.5)).astype(np.float32))
    return out


def density_to_bitfield(density, max_cascade):
    """Host twin of grid_to_bitfield + bitfield_max_pool. Returns CASCADES * 128^3 / 8 bytes."""
    mean = np.float32(np.maximum(density[:CELLS], 0).astype(np.float64).sum() / CELLS)
    thresh = min(np.float32(0.01), mean)
    bits = np.zeros(CASCADES * CELLS // 8, dtype=np.uint8)
    n = CELLS * (max_cascade + 1)
    bits[: n // 8] = np.packbits(density[:n] > thresh, bitorder="little")
    c = np.arange(CELLS // 8, dtype=np.uint64)   # previous-level byte index == Morton index of the 2x2x2 block
    dst = morton3(_compact3(c) + 32, _compact3(c >> 1) + 32, _compact3(c >> 2) + 32)
    for level in range(1, CASCADES):
        prev = bits[(level - 1) * CELLS // 8: level * CELLS // 8]
        nxt = np.unpackbits(bits[level * CELLS // 8:(level + 1) * CELLS // 8], bitorder="little")
        nxt[dst] |= (prev > 0).astype(np.uint8)
        bits[level * CELLS // 8:(level + 1) * CELLS // 8] = np.packbits(nxt, bitorder="little")
    return bits


class SyntheticModel:
    """Arrays + metadata of one model; ``desc()`` fills the ctypes struct shared by libsnb.so and the oracle."""

    def __init__(self, n_levels=16, n_feat=2, log2_T=19, base_res=16, aabb_scale=4, seed=1337, density_gain=8.0, rgb_gain=8.0, feature_amp=0.5, feature_mean=0.0,
                 occupancy=True, rgb_activation=2, density_activation=3, min_transmittance=0.01):
        assert n_levels * n_feat == 32
        self.n_levels, self.n_feat, self.log2_T, self.base_res, self.aabb_scale = n_levels, n_feat, log2_T, base_res, aabb_scale
        self.per_level_scale = per_level_scale_for(aabb_scale, base_res, n_levels)
        self.scales, self.ress, self.offsets = grid_layout(n_levels, n_feat, log2_T, base_res, self.per_level_scale)
        rng = np.random.default_rng(seed)

        def xavier(n_out, n_in):
            s = math.sqrt(6.0 / (n_in + n_out))
            return rng.uniform(-s, s, size=(n_out, n_in)).astype(np.float32)
        dw1, dw2, rw1, rw2, rw3 = xavier(64, 32), xavier(16, 64), xavier(64, 32), xavier(64, 64), xavier(16, 64)
        dw2[0] = np.abs(dw2[0]) * density_gain   # positive, amplified log-density head so rays saturate inside occupied cells
        rw3[:3] *= rgb_gain              # spread the colours away from logistic(0)
        rw3[3:] = 0.0                    # padded outputs of the rgb head (3 live rows)
        grid = (rng.uniform(-feature_amp, feature_amp, size=self.offsets[-1] * n_feat) + feature_mean).astype(np.float32)
        self.params = np.concatenate([w.reshape(-1) for w in (dw1, dw2, rw1, rw2, rw3)] + [grid]).astype(np.float16)
        assert self.params.size == MLP_HALVES + self.offsets[-1] * n_feat
        self.max_cascade = int(round(math.log2(aabb_scale))) if aabb_scale > 1 else 0
        self.cone_angle = 0.0 if aabb_scale <= 1 else 1.0 / 256.0        # load_nerf_post, src/testbed_nerf.cu:3001-3088
        half = 0.5 * aabb_scale
        self.aabb_min = [0.5 - half] * 3; self.aabb_max = [0.5 + half] * 3
        self.render_aabb_min = list(self.aabb_min); self.render_aabb_max = list(self.aabb_max)
        self.render_aabb_to_local = [1, 0, 0, 0, 1, 0, 0, 0, 1]
        self.rgb_activation, self.density_activation, self.min_transmittance = rgb_activation, density_activation, min_transmittance
        self.density = density_grid(self.max_cascade) if occupancy else None
        self.bitfield = density_to_bitfield(self.density, self.max_cascade) if occupancy else None

    def desc(self):
        import ctypes as C
        from ._lib import ModelDesc
        d = ModelDesc()
        d.n_levels, d.n_features_per_level, d.log2_hashmap_size, d.base_resolution = self.n_levels, self.n_feat, self.log2_T, self.base_res
        d.per_level_scale = self.per_level_scale
        d.params = self.params.view(np.uint16).ctypes.data_as(C.POINTER(C.c_uint16))
        d.n_params = self.params.size
        d.bitfield = self.bitfield.ctypes.data_as(C.POINTER(C.c_uint8)) if self.bitfield is not None else None
        d.aabb_min[:] = self.aabb_min; d.aabb_max[:] = self.aabb_max
        d.render_aabb_min[:] = self.render_aabb_min; d.render_aabb_max[:] = self.render_aabb_max
        d.render_aabb_to_local[:] = self.render_aabb_to_local
        d.max_cascade = self.max_cascade
        d.cone_angle_constant = self.cone_angle
        d.rgb_activation, d.density_activation, d.min_transmittance = self.rgb_activation, self.density_activation, self.min_transmittance
        return d
