"""CPU: the oracle's two forms of the sng march agree — the replay of the reference's lock-step wavefront loop
(trace_alt, src/testbed_nerf.cu:2128-2277: S_j = clamp(target / n_alive_j, 1, 8), payload.t rewound after every batch, :574)
and the per-ray loop given that loop's batch-length sequence. The product's march is the second form (rays run independently);
this is the property that makes it equal to the first."""
import numpy as np

import common
import oracle_lib as O


def test_per_ray_march_with_the_wavefront_sequence_is_the_wavefront_march():
    m = common.model("small16x2")
    d = m.desc()
    v = common.make_view(64, 48)
    fb, depth, pos, sched, alive = O.render_nerf_wavefront(d, v, alt=1, target_queries=3900, positions=True)
    assert sched[0] < 8 and len(set(sched.tolist())) >= 3 and sched[-1] == 8, sched
    assert np.all(np.diff(alive.astype(np.int64)) <= 0)                                  # n_alive never grows
    np.testing.assert_array_equal(sched, np.clip(3900 // alive, 1, 8))
    fb2, depth2, pos2 = O.render_nerf_sched(d, v, sched, alt=1, positions=True)
    np.testing.assert_array_equal(fb.view(np.uint32), fb2.view(np.uint32))
    np.testing.assert_array_equal(depth.view(np.uint32), depth2.view(np.uint32))
    np.testing.assert_array_equal(pos.view(np.uint32), pos2.view(np.uint32))
    # the sequence matters: 8 samples per batch throughout is a visibly different image ...
    fb8, _, _, _ = O.render_nerf(d, v, 1)
    assert np.abs(fb8 - fb).max() > 2e-2
    # ... except on the stock path (trace + composite_kernel_nerf never rewinds), where any sequence gives the same bits
    a, _, _ = O.render_nerf_sched(d, v, sched, alt=0)
    b, _, _, _ = O.render_nerf(d, v, 0)
    np.testing.assert_array_equal(a.view(np.uint32), b.view(np.uint32))


def test_single_sample_batches_stall_the_ray_on_its_first_sample():
    """S = 1 (more than target alive rays — 4K frames under the reference's 2^21): every iteration re-composites the sample the
    previous one ended on, so the march only advances once enough rays have saturated. Recorded here because it is the
    reference's behaviour, not a defect of the restatement."""
    m = common.model("small16x2")
    d = m.desc()
    v = common.make_view(48, 32)
    fb, _, _, sched, alive = O.render_nerf_wavefront(d, v, alt=1, target_queries=int(0.9 * 600))
    assert sched[0] == 1
    n1 = int(np.argmax(sched > 1)) if (sched > 1).any() else len(sched)
    assert n1 >= 10                                                                      # tens of iterations without progress
