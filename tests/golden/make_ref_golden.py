"""Generates tests/golden/ref_host.npz by running the REFERENCE's own host-callable code (oracle/_ref/libref_shim.so,
built from the headers under /root/reference by `make -C oracle ref`) on seeded inputs. Run in the build container:
    python tests/golden/make_ref_golden.py
The npz is committed; tests/test_oracle_pinned.py checks the oracle restatement (math mode LIBM = what these headers
do on the host) against it bit for bit, here and on the GPU box where /root/reference does not exist."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import common  # noqa: E402

pf = C.POINTER(C.c_float)
f32, u32 = C.c_float, C.c_uint32


def P(a):
    return a.ctypes.data_as(pf)


def main():
    lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_shim.so"))
    for name in ["ref_calc_dt", "ref_advance_n_steps", "ref_to_stepping_space", "ref_from_stepping_space", "ref_ld_random_val", "ref_srgb_to_linear",
                 "ref_linear_to_srgb", "ref_march_to_occupied", "ref_triangle_ray_intersect"]:
        getattr(lib, name).restype = f32
    lib.ref_calc_dt.argtypes = [f32, f32]; lib.ref_advance_n_steps.argtypes = [f32, f32, f32]
    lib.ref_to_stepping_space.argtypes = [f32, f32]; lib.ref_from_stepping_space.argtypes = [f32, f32]
    lib.ref_ld_random_val.argtypes = [u32, u32]; lib.ref_srgb_to_linear.argtypes = [f32]; lib.ref_linear_to_srgb.argtypes = [f32]
    lib.ref_cascaded_grid_idx_at.restype = u32; lib.ref_cascaded_grid_idx_at.argtypes = [pf, u32]
    lib.ref_mip_from_pos.restype = u32; lib.ref_mip_from_pos.argtypes = [pf]
    lib.ref_morton3D.restype = u32; lib.ref_morton3D.argtypes = [u32, u32, u32]
    lib.ref_ld_random_pixel_offset.argtypes = [u32, pf]
    lib.ref_march_to_occupied.argtypes = [C.POINTER(C.c_uint8), u32, pf, pf, f32, pf, pf, f32]
    lib.ref_init_ray.argtypes = [pf, pf, pf, C.c_int, C.c_int, u32, C.c_int, pf, pf, u32, u32, pf]
    lib.ref_triangle_ray_intersect.argtypes = [pf, pf, pf]
    lib.ref_triangle_normal.argtypes = [pf, pf]; lib.ref_aabb_ray_intersect.argtypes = [pf, pf, pf, pf, pf]

    rng = np.random.default_rng(2024)
    g = {}
    # stepping functions: t over 7 decades, both cone settings
    t = np.exp(rng.uniform(np.log(1e-4), np.log(60.0), 4000)).astype(np.float32)
    for cone, tag in ((1.0 / 256.0, "c"), (0.0, "z")):
        g[f"step_t_{tag}"] = t
        g[f"calc_dt_{tag}"] = np.array([lib.ref_calc_dt(float(x), cone) for x in t], np.float32)
        g[f"to_step_{tag}"] = np.array([lib.ref_to_stepping_space(float(x), cone) for x in t], np.float32)
        nn = rng.uniform(-50, 2500, 4000).astype(np.float32)
        g[f"from_n_{tag}"] = nn
        g[f"from_step_{tag}"] = np.array([lib.ref_from_stepping_space(float(x), cone) for x in nn], np.float32)
        fr = rng.uniform(0, 1, 4000).astype(np.float32)
        g[f"adv_n_{tag}"] = fr
        g[f"advance_{tag}"] = np.array([lib.ref_advance_n_steps(float(x), cone, float(k)) for x, k in zip(t, fr)], np.float32)
    # occupancy cascade indexing
    pos = rng.uniform(-7.5, 8.5, (6000, 3)).astype(np.float32)
    pos[:2000] = rng.uniform(0, 1, (2000, 3)).astype(np.float32)
    mips = rng.integers(0, 8, 6000).astype(np.uint32)
    g["cell_pos"], g["cell_mip"] = pos, mips
    g["cell_idx"] = np.array([lib.ref_cascaded_grid_idx_at(P(np.ascontiguousarray(p)), int(m)) for p, m in zip(pos, mips)], np.uint32)
    g["mip_from_pos"] = np.array([lib.ref_mip_from_pos(P(np.ascontiguousarray(p))) for p in pos], np.uint32)
    xyz = rng.integers(0, 1024, (3000, 3)).astype(np.uint32)
    g["morton_xyz"] = xyz
    g["morton"] = np.array([lib.ref_morton3D(int(a), int(b), int(c)) for a, b, c in xyz], np.uint32)
    # LD sampler
    idx = rng.integers(0, 2 ** 32, 3000, dtype=np.uint64).astype(np.uint32); seed = rng.integers(0, 2 ** 32, 3000, dtype=np.uint64).astype(np.uint32)
    idx[:64] = np.arange(64)
    g["ld_idx"], g["ld_seed"] = idx, seed
    g["ld_val"] = np.array([lib.ref_ld_random_val(int(i), int(s)) for i, s in zip(idx, seed)], np.float32)
    off = np.zeros((64, 2), np.float32)
    for s in range(64):
        lib.ref_ld_random_pixel_offset(s, P(off[s]))
    g["ld_pixel_offset"] = off
    # sRGB
    x = np.concatenate([rng.uniform(0, 1, 2000), rng.uniform(0, 0.01, 500), [0.0, 0.04045, 0.0031308, 1.0]]).astype(np.float32)
    g["srgb_x"] = x
    g["srgb_to_linear"] = np.array([lib.ref_srgb_to_linear(float(v)) for v in x], np.float32)
    g["linear_to_srgb"] = np.array([lib.ref_linear_to_srgb(float(v)) for v in x], np.float32)
    # march through the synthetic occupancy fields (aabb 4 and aabb 16 cascades)
    for kind in ("small16x2", "big16x2"):
        m = common.model(kind)
        n = 3000
        o = rng.uniform(m.aabb_min[0], m.aabb_max[0], (n, 3)).astype(np.float32)
        tgt = rng.uniform(0.2, 0.8, (n, 3)).astype(np.float32)
        d = tgt - o; d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
        t0 = rng.uniform(0, 0.5, n).astype(np.float32)
        mn = np.asarray(m.render_aabb_min, np.float32); mx = np.asarray(m.render_aabb_max, np.float32)
        bf = m.bitfield.ctypes.data_as(C.POINTER(C.c_uint8))
        res = np.array([lib.ref_march_to_occupied(bf, m.max_cascade, P(mn), P(mx), m.cone_angle, P(np.ascontiguousarray(a)), P(np.ascontiguousarray(b)), float(c))
                        for a, b, c in zip(o, d, t0)], np.float32)
        g[f"march_{kind}_o"], g[f"march_{kind}_d"], g[f"march_{kind}_t0"], g[f"march_{kind}_t"] = o, d, t0, res
    # ray generation for three cameras x (snap, spp) settings
    rays = []
    cams = [common.camera_matrix(), common.camera_matrix(view=(0.7, -0.2, 0.3), at=(0.4, 0.5, 0.6), zoom=2.0), common.camera_matrix(view=(0.0, -1.0, 0.001), at=(0.5, 0.3, 0.5), zoom=1.2)]
    mn = np.array([-1.5] * 3, np.float32); mx = np.array([2.5] * 3, np.float32)
    g["ray_cams"] = np.stack(cams)
    meta = []
    for ci, cam in enumerate(cams):
        for (w, h, spp, snap) in ((48, 32, 0, 1), (33, 17, 5, 0)):
            focal = np.array([h * 1.0572, h * 1.0572], np.float32); center = np.array([0.5, 0.5], np.float32)
            out = np.zeros((h, w, 8), np.float32)
            for y in range(h):
                for x_ in range(w):
                    lib.ref_init_ray(P(cam), P(focal), P(center), w, h, spp, snap, P(mn), P(mx), x_, y, P(out[y, x_]))
            rays.append(out.reshape(-1, 8)); meta.append([ci, w, h, spp, snap])
    g["ray_meta"] = np.array(meta, np.int32)
    g["ray_out"] = np.concatenate(rays)
    # triangles / boxes
    n = 4000
    tris = rng.uniform(-1, 1, (n, 9)).astype(np.float32)
    ro = rng.uniform(-2, 2, (n, 3)).astype(np.float32)
    rd = (tris.reshape(n, 3, 3).mean(1) + rng.normal(0, 0.3, (n, 3)) - ro).astype(np.float32)
    g["tri"], g["tri_ro"], g["tri_rd"] = tris, ro, rd
    g["tri_t"] = np.array([lib.ref_triangle_ray_intersect(P(np.ascontiguousarray(a)), P(np.ascontiguousarray(b)), P(np.ascontiguousarray(c))) for a, b, c in zip(tris, ro, rd)], np.float32)
    nrm = np.zeros((n, 3), np.float32)
    for i in range(n):
        lib.ref_triangle_normal(P(np.ascontiguousarray(tris[i])), P(nrm[i]))
    g["tri_normal"] = nrm
    lo = rng.uniform(-1, 0, (n, 3)).astype(np.float32); hi = (lo + rng.uniform(0.1, 1.5, (n, 3))).astype(np.float32)
    rd2 = rd.copy(); rd2[:50, 0] = 0.0; rd2[50:100, 1] = 0.0
    bb = np.zeros((n, 2), np.float32)
    for i in range(n):
        lib.ref_aabb_ray_intersect(P(np.ascontiguousarray(lo[i])), P(np.ascontiguousarray(hi[i])), P(np.ascontiguousarray(ro[i])), P(np.ascontiguousarray(rd2[i])), P(bb[i]))
    g["bb_lo"], g["bb_hi"], g["bb_rd"], g["bb_t"] = lo, hi, rd2, bb
    out_path = os.path.join(ROOT, "tests", "golden", "ref_host.npz")
    np.savez_compressed(out_path, **g)
    print("wrote", out_path, os.path.getsize(out_path), "bytes")


if __name__ == "__main__":
    main()
