"""snb_bvh_refit_dev_noalloc — the allocation-free variant of snb_bvh_refit_dev (no counterpart in the reference; added
for callers that cannot call cudaMalloc). Its scratch is a fixed `static __device__` array inside the library that
every call shares and that keeps the previous call's contents, so the test makes TWO calls in a row on different trees and compares each
with the existing entry point, bit for bit (the refit is min / max of fp32 values: exact). Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

from synerfgine_b200 import _lib, scenes

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

MAX_NODES = 262144      # SNB_REFIT_NOALLOC_MAX_NODES, include/snb.h


def _object(sc, i):
    ob = sc.objects[i]
    tris = np.ctypeslib.as_array(ob.tris, shape=(ob.n_tris, 9)).copy()
    nodes = np.ctypeslib.as_array(ob.nodes, shape=(ob.n_nodes, 8)).copy()
    return tris, nodes


def _bend(tris, k):
    v = tris.reshape(-1, 3)
    return np.ascontiguousarray((v + 0.1 * np.sin((5.0 + k) * v[:, [2, 0, 1]])).astype(np.float32).reshape(-1, 9))


def test_refit_without_allocation_matches_the_allocating_entry_point():
    from synerfgine_b200.testbed import Scene
    lib = _lib.load()
    ctx = C.c_void_p()
    assert lib.snb_create(0, None, C.byref(ctx)) == 0
    try:
        sc = Scene(scenes.scene_json("c3"), scenes.mesh_dir())
        cases = []
        for k, i in enumerate((0, 2)):                       # two trees of different size, one after the other
            tris, nodes = _object(sc, i)
            if k:                                            # the second tree: finer leaves -> another node count than the first
                cap = 2 * tris.shape[0]
                nodes = np.zeros((cap, 8), np.float32); nn = C.c_int32()
                assert lib.snb_bvh_build(tris.ctypes.data_as(_lib.pf32), tris.shape[0], 1, nodes.ctypes.data_as(_lib.pf32), cap, C.byref(nn)) == 0
                nodes = np.ascontiguousarray(nodes[:nn.value])
            assert nodes.shape[0] <= MAX_NODES
            bent = _bend(tris, k)
            d_tris = torch.from_numpy(bent).cuda()
            cases.append((bent, nodes, d_tris, torch.from_numpy(nodes).cuda(), torch.from_numpy(nodes).cuda()))
        assert cases[0][1].shape[0] != cases[1][1].shape[0]
        torch.cuda.synchronize()
        # expected: the existing entry point (context scratch, may allocate)
        for bent, nodes, d_tris, d_ref, _ in cases:
            assert lib.snb_bvh_refit_dev(ctx, d_tris.data_ptr(), d_ref.data_ptr(), nodes.shape[0]) == 0, lib.snb_last_error(ctx)
        assert lib.snb_sync(ctx) == 0
        # a first call loads the module's static array; after it the two calls under test must not take device memory
        warm = cases[0][4].clone(); torch.cuda.synchronize()
        assert lib.snb_bvh_refit_dev_noalloc(ctx, cases[0][2].data_ptr(), warm.data_ptr(), cases[0][1].shape[0]) == 0, lib.snb_last_error(ctx)
        assert lib.snb_sync(ctx) == 0
        free_before = torch.cuda.mem_get_info()[0]
        for bent, nodes, d_tris, _, d_out in cases:          # two calls in a row, same stream: ordered, the second sees the first one's leftovers
            assert lib.snb_bvh_refit_dev_noalloc(ctx, d_tris.data_ptr(), d_out.data_ptr(), nodes.shape[0]) == 0, lib.snb_last_error(ctx)
        assert lib.snb_sync(ctx) == 0
        assert torch.cuda.mem_get_info()[0] >= free_before
        for bent, nodes, _, d_ref, d_out in cases:
            ref, out = d_ref.cpu().numpy(), d_out.cpu().numpy()
            assert not np.array_equal(ref[:, :6], nodes[:, :6])                      # the refit did something
            np.testing.assert_array_equal(out.view(np.uint32), ref.view(np.uint32))  # bit for bit, child pointers included
            np.testing.assert_array_equal(out[0, :3], bent.reshape(-1, 3).min(0)); np.testing.assert_array_equal(out[0, 3:6], bent.reshape(-1, 3).max(0))
        # and once more on the first tree after the other one used the array
        again = torch.from_numpy(cases[0][1]).cuda(); torch.cuda.synchronize()
        assert lib.snb_bvh_refit_dev_noalloc(ctx, cases[0][2].data_ptr(), again.data_ptr(), cases[0][1].shape[0]) == 0 and lib.snb_sync(ctx) == 0
        np.testing.assert_array_equal(again.cpu().numpy().view(np.uint32), cases[0][3].cpu().numpy().view(np.uint32))
        # beyond the fixed scratch: rejected with a clear error, nothing launched
        rc = lib.snb_bvh_refit_dev_noalloc(ctx, cases[0][2].data_ptr(), again.data_ptr(), MAX_NODES + 1)
        assert rc != 0 and b"exceed the fixed scratch" in lib.snb_last_error(ctx)
        assert lib.snb_bvh_refit_dev_noalloc(ctx, None, again.data_ptr(), 8) != 0
    finally:
        lib.snb_destroy(ctx)
