"""CPU tests of the on-disk model format: `.ingp` / `.msgpack` snapshots.

The files read here are authored by an INDEPENDENT writer — Python's msgpack + zlib/gzip modules, laid out as
Testbed::save_snapshot does (src/testbed.cu:4812-4876, tcnn trainer.h:274-290: nlohmann::json::to_msgpack of the network
config with the "snapshot" object) — so the C++ msgpack reader is checked against the format, not against its own writer;
the C++ writer is then checked by reading its files back with Python's msgpack."""
import ctypes as C
import gzip
import zlib

import numpy as np
import pytest

import common
from synerfgine_b200 import _lib

msgpack = pytest.importorskip("msgpack")


def rows(mat_colmajor, cols, nrows):
    """tcnn vec_json.h:40-48: a matrix is an array of ROWS."""
    return [[float(mat_colmajor[c * nrows + r]) for c in range(cols)] for r in range(nrows)]


def author_snapshot(m, camera=None, per_level_scale=True, extra=None):
    d = m.desc()
    enc = {"otype": "HashGrid", "n_levels": m.n_levels, "n_features_per_level": m.n_feat, "log2_hashmap_size": m.log2_T, "base_resolution": m.base_res}
    if per_level_scale:
        enc["per_level_scale"] = float(np.float32(m.per_level_scale))
    snap = {
        "version": 1, "mode": "nerf", "n_params": int(m.params.size), "params_type": "__half", "params_binary": m.params.tobytes(),
        "density_grid_size": 128, "density_grid_binary": m.density.astype(np.float16).tobytes(),
        "nerf": {"aabb_scale": m.aabb_scale, "rgb": {"rays_per_batch": 4096, "measured_batch_size": 262144}},
        "training_step": 35000, "loss": 0.00123,
        "aabb": {"min": [float(x) for x in d.aabb_min], "max": [float(x) for x in d.aabb_max]},
        "render_aabb": {"min": [float(x) for x in d.render_aabb_min], "max": [float(x) for x in d.render_aabb_max]},
        "render_aabb_to_local": rows(list(d.render_aabb_to_local), 3, 3),
        "up_dir": [0.0, 1.0, 0.0], "exposure": 0.0, "background_color": [0.0, 0.0, 0.0, 1.0],
    }
    if camera is not None:
        snap["camera"] = {"matrix": rows(list(camera), 4, 3), "fov_axis": 1, "zoom": 1.0, "autofocus": False}
    root = {"loss": {"otype": "Huber"}, "optimizer": {"otype": "Ema", "nested": {"otype": "Adam", "learning_rate": 1e-2}},
            "encoding": enc, "network": {"otype": "FullyFusedMLP", "n_neurons": 64, "n_hidden_layers": 1, "activation": "ReLU"},
            "rgb_network": {"otype": "FullyFusedMLP", "n_neurons": 64, "n_hidden_layers": 2}, "dir_encoding": {"otype": "Composite", "nested": [{"n_dims_to_encode": 3, "otype": "SphericalHarmonics", "degree": 4}]},
            "snapshot": snap}
    if extra:
        root.update(extra)
    return msgpack.packb(root, use_bin_type=True, use_single_float=False)


def load(snb, path):
    h = C.c_void_p()
    err = C.create_string_buffer(256)
    rc = snb.snb_snapshot_load(str(path).encode(), C.byref(h), err, 256)
    return rc, h, err.value.decode()


def get(snb, h):
    d = _lib.ModelDesc(); dens = C.POINTER(C.c_uint16)(); n = C.c_uint64(); scale = C.c_uint32(); cam = (C.c_float * 12)()
    rc = snb.snb_snapshot_get(h, C.byref(d), C.byref(dens), C.byref(n), C.byref(scale), cam)
    return rc, d, dens, n.value, scale.value, list(cam)


def same_model(d, m, ref):
    assert (d.n_levels, d.n_features_per_level, d.log2_hashmap_size, d.base_resolution) == (m.n_levels, m.n_feat, m.log2_T, m.base_res)
    assert np.float32(d.per_level_scale) == np.float32(ref.per_level_scale)
    assert d.n_params == m.params.size
    got = np.ctypeslib.as_array(d.params, shape=(d.n_params,))
    np.testing.assert_array_equal(got, m.params.view(np.uint16))
    for a in ("aabb_min", "aabb_max", "render_aabb_min", "render_aabb_max", "render_aabb_to_local"):
        assert list(getattr(d, a)) == list(getattr(ref, a)), a
    assert (d.max_cascade, d.rgb_activation, d.density_activation) == (ref.max_cascade, ref.rgb_activation, ref.density_activation)
    assert np.float32(d.cone_angle_constant) == np.float32(ref.cone_angle_constant) and np.float32(d.min_transmittance) == np.float32(0.01)
    assert not d.bitfield


@pytest.mark.parametrize("container", ["msgpack", "ingp-zlib", "ingp-gzip"])
@pytest.mark.parametrize("kind", ["small16x2", "fox8x4"])
def test_reads_snapshots_written_by_an_independent_writer(snb, tmp_path, container, kind):
    m = common.model(kind)
    cam = [float(x) for x in common.camera_matrix()]
    blob = author_snapshot(m, camera=cam)
    path = tmp_path / ("model.msgpack" if container == "msgpack" else "model.ingp")
    path.write_bytes(blob if container == "msgpack" else zlib.compress(blob, 1) if container == "ingp-zlib" else gzip.compress(blob, 1))
    rc, h, err = load(snb, path)
    assert rc == 0, err
    rc, d, dens, n, scale, got_cam = get(snb, h)
    assert rc == 0 and scale == m.aabb_scale
    same_model(d, m, m.desc())
    assert n == m.density.size
    np.testing.assert_array_equal(np.ctypeslib.as_array(dens, shape=(n,)), m.density.astype(np.float16).view(np.uint16))
    assert [np.float32(x) for x in got_cam] == [np.float32(x) for x in cam]
    snb.snb_snapshot_free(h)


def test_derived_constants_when_the_config_omits_them(snb, tmp_path):
    """reset_network computes per_level_scale when the config has none (testbed.cu:3737-3741)."""
    m = common.model("small16x2")
    path = tmp_path / "m.msgpack"
    path.write_bytes(author_snapshot(m, per_level_scale=False))
    rc, h, err = load(snb, path)
    assert rc == 0, err
    rc, d, *_ = get(snb, h)
    assert rc == -3                                                  # no camera in this file: SNB_ERR_STATE for the camera, the rest is filled
    assert abs(d.per_level_scale - m.per_level_scale) <= 2e-7 * m.per_level_scale
    snb.snb_snapshot_free(h)


def test_writer_round_trip_and_python_can_read_it(snb, tmp_path):
    m = common.model("small16x2")
    d = m.desc()
    dens = m.density.astype(np.float16).view(np.uint16)
    cam = (C.c_float * 12)(*[float(x) for x in common.camera_matrix()])
    for compress in (0, 1):
        path = tmp_path / ("out.ingp" if compress else "out.msgpack")
        rc = snb.snb_snapshot_save(str(path).encode(), C.byref(d), dens.ctypes.data_as(C.POINTER(C.c_uint16)), dens.size, m.aabb_scale, cam, compress)
        assert rc == 0
        raw = path.read_bytes()
        tree = msgpack.unpackb(zlib.decompress(raw) if compress else raw, raw=False)
        assert tree["snapshot"]["params_type"] == "__half" and tree["snapshot"]["n_params"] == m.params.size
        assert tree["snapshot"]["params_binary"] == m.params.tobytes() and tree["snapshot"]["density_grid_binary"] == dens.tobytes()
        assert tree["encoding"]["n_levels"] == m.n_levels and tree["snapshot"]["nerf"]["aabb_scale"] == m.aabb_scale
        assert tree["snapshot"]["camera"]["matrix"] == rows(list(cam), 4, 3)
        # every key Testbed::load_snapshot / reset_network reads without a default (src/testbed.cu:4878-4992): the reference can load the file
        snap = tree["snapshot"]
        assert snap["version"] >= 1 and snap["training_step"] == 0 and snap["loss"] == 0.0 and snap["density_grid_size"] == 128
        assert set(snap["nerf"]["rgb"]) == {"rays_per_batch", "measured_batch_size", "measured_batch_size_before_compaction"}
        assert tree["network"]["n_neurons"] == 64 and tree["network"]["n_hidden_layers"] == 1 and tree["rgb_network"]["n_hidden_layers"] == 2
        assert tree["dir_encoding"]["nested"][0]["degree"] == 4 and tree["loss"]["otype"] == "Huber"
        assert tree["optimizer"]["nested"]["nested"]["otype"] == "Adam" and tree["encoding"]["otype"] == "HashGrid"
        rc, h, err = load(snb, path)
        assert rc == 0, err
        rc, d2, dp, n, scale, got_cam = get(snb, h)
        assert rc == 0
        same_model(d2, m, d)
        assert n == dens.size and scale == m.aabb_scale
        snb.snb_snapshot_free(h)


def test_malformed_files_are_rejected_with_a_message(snb, tmp_path):
    m = common.model("small16x2")
    blob = author_snapshot(m)
    cases = {
        "truncated.msgpack": blob[: len(blob) // 2],
        "truncated.ingp": zlib.compress(blob, 1)[:-100],
        "garbage.ingp": b"\x78\x9c" + bytes(range(256)) * 4,
        "no_snapshot.msgpack": msgpack.packb({"encoding": {"n_levels": 16}}),
        "not_a_map.msgpack": msgpack.packb([1, 2, 3]),
        "sdf.msgpack": msgpack.packb({"encoding": {}, "snapshot": {"mode": "sdf", "params_binary": b"", "params_type": "__half"}}, use_bin_type=True),
        "float_params.msgpack": msgpack.packb({"encoding": {}, "snapshot": {"mode": "nerf", "params_type": "float", "params_binary": b"\0" * 8}}, use_bin_type=True),
        "empty.msgpack": b"",
        # the format version is enforced, other encodings / MLP shapes are refused, a huge element count is a parse error
        "old_version.msgpack": msgpack.packb({**msgpack.unpackb(blob, raw=False), "snapshot": {**msgpack.unpackb(blob, raw=False)["snapshot"], "version": 0}}, use_bin_type=True),
        "dense_grid.msgpack": msgpack.packb({**msgpack.unpackb(blob, raw=False), "encoding": {**msgpack.unpackb(blob, raw=False)["encoding"], "otype": "DenseGrid"}}, use_bin_type=True),
        "wide_mlp.msgpack": msgpack.packb({**msgpack.unpackb(blob, raw=False), "network": {"otype": "FullyFusedMLP", "n_neurons": 128, "n_hidden_layers": 1}}, use_bin_type=True),
        "huge_array.msgpack": b"\x82\xa8encoding\xdd\xff\xff\xff\xf0\x00",
        "huge_map.msgpack": b"\xdf\xff\xff\xff\xf0\xa1a\x00",
    }
    for name, data in cases.items():
        p = tmp_path / name
        p.write_bytes(data)
        rc, h, err = load(snb, p)
        assert rc in (-5, -6) and not h.value and err, (name, rc, err)
    rc, h, err = load(snb, tmp_path / "does-not-exist.ingp")
    assert rc == -4 and "cannot read" in err
    # wrong number of cascades for the aabb_scale (testbed.cu:4933 "Incompatible number of grid cascades.")
    tree = msgpack.unpackb(blob, raw=False)
    tree["snapshot"]["density_grid_binary"] = tree["snapshot"]["density_grid_binary"][: 128 ** 3 * 2]
    p = tmp_path / "cascades.msgpack"
    p.write_bytes(msgpack.packb(tree, use_bin_type=True))
    rc, h, err = load(snb, p)
    assert (rc == -5 and "cascades" in err) or m.max_cascade == 0
