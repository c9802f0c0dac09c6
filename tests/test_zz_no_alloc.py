"""GPU: b2fwn_calc_from_vector_device_noalloc, the point-vector entry point for hosts in which this library may not allocate
device memory -- an application that reserved the GPU's memory for its own arena when it started (an allocation made later by a
plug-in fails there, or crowds the application out), or a loop on a latency budget that cannot take the device-wide stall an
allocation brings.  The entry point keeps the Morton sort's index arrays in a fixed array inside the library instead of borrowing
them from the handle's memory pool.

What is held here, through the C ABI: its winding numbers are the ones b2fwn_calc_from_vector_device returns, bit for bit
(the tolerance of the other device-against-device comparisons in this suite: none -- the sort changes the order in which the
queries are walked, not the arithmetic of any of them), on two calls in a row.  The second call is shorter than the first and has
other points, so anything it wrongly took over from the first call's leftovers in the shared array would show.  A vector beyond
the fixed capacity must be refused with a message that names the limit, and the refusal must leave the array usable.

This file is meant to be collected last (pytest -x stops at the first failure; a failure here must not hide other tests) and
passes alone in a fresh process as well as after the rest of the suite."""
import numpy as np
import pytest

import meshlib_b200 as mb
from meshlib_b200 import meshgen

pytestmark = pytest.mark.gpu

MAX_QUERIES = 1 << 20  # B2FWN_NOALLOC_MAX_QUERIES of include/b2fwn.h


def _both(fwn, q, **kw):
    import torch
    d_q = torch.from_numpy(np.ascontiguousarray(q, np.float32)).cuda()
    n = len(q)
    # different fill patterns: a call that wrote nothing cannot pass by accident
    want = torch.full((n,), 7.0, dtype=torch.float32, device="cuda")
    got = torch.full((n,), -7.0, dtype=torch.float32, device="cuda")
    fwn.calc_from_vector_device(d_q.data_ptr(), n, want.data_ptr(), **kw)
    fwn.calc_from_vector_device_noalloc(d_q.data_ptr(), n, got.data_ptr(), **kw)
    torch.cuda.synchronize()
    return got.cpu().numpy(), want.cpu().numpy()


def test_noalloc_point_vector_equals_the_allocating_entry_point_twice_in_a_row():
    import torch
    pts, tris = meshgen.make_torus(2, 1, 96, 48)
    fwn = mb.FastWindingNumber(pts, tris)
    try:
        # first call: large enough for the sorted path (>= 32768 queries), points inside, outside and far from the torus
        q1 = meshgen.query_points(200_000, -4.0, 4.0, seed=31)
        got, want = _both(fwn, q1)
        assert got.tobytes() == want.tobytes()
        assert np.isfinite(got).all() and (np.abs(got) > 0.4).any() and (np.abs(got) < 0.1).any()  # both sides of the surface
        # second call straight after it: fewer, different points, other beta, one face skipped
        q2 = meshgen.query_points(70_001, -2.5, 2.5, seed=32)
        got, want = _both(fwn, q2, beta=3.0, skip_face=5)
        assert got.tobytes() == want.tobytes()
        # below the sort threshold the scratch is not touched at all; the answers are the same again
        got, want = _both(fwn, q2[:1000])
        assert got.tobytes() == want.tobytes()
        # exactly the capacity is accepted ...
        qmax = meshgen.query_points(MAX_QUERIES, -3.5, 3.5, seed=33)
        got, want = _both(fwn, qmax)
        assert got.tobytes() == want.tobytes()
        # ... one more is refused, in words, before anything runs
        d_q = torch.zeros((MAX_QUERIES + 1, 3), dtype=torch.float32, device="cuda")
        d_o = torch.full((MAX_QUERIES + 1,), 5.0, dtype=torch.float32, device="cuda")
        with pytest.raises(mb.B2Error, match="B2FWN_NOALLOC_MAX_QUERIES"):
            fwn.calc_from_vector_device_noalloc(d_q.data_ptr(), MAX_QUERIES + 1, d_o.data_ptr())
        assert float(d_o.min()) == 5.0 and float(d_o.max()) == 5.0
        # and the array still serves the next call
        got, want = _both(fwn, q1[:50_000])
        assert got.tobytes() == want.tobytes()
        # empty vector: nothing to do, no error
        fwn.calc_from_vector_device_noalloc(None, 0, None)
    finally:
        fwn.close()


def test_noalloc_entry_point_serves_a_second_handle_on_the_same_device():
    """two meshes, two handles, one device: both use the one array of that device, one after the other"""
    a = mb.FastWindingNumber(*meshgen.make_uv_sphere(1, 64, 32))
    b = mb.FastWindingNumber(*meshgen.make_torus(2, 1, 64, 32))
    try:
        q = meshgen.query_points(40_000, -3.0, 3.0, seed=34)
        for fwn in (a, b, a):
            got, want = _both(fwn, q)
            assert got.tobytes() == want.tobytes()
    finally:
        a.close()
        b.close()
