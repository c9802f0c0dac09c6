"""Marching-cubes oracle (oracle/port_mc.c) pinned against the compiled reference (oracle/_ref, built from
/root/reference/source/MRVoxels/MRMarchingCubes.cpp) and against the reference's own known-answer test
(source/MRTest/MRMarchingCubesTests.cpp:9-50).  CPU only."""
import os
import re

import numpy as np
import pytest

from oracle import portlib, reflib

needs_ref = pytest.mark.skipif(not reflib.available(), reason="oracle/_ref not built")
REF_MC = "/root/reference/source/MRVoxels/MRMarchingCubes.cpp"


def ball_volume():
    z, y, x = np.meshgrid(*[np.arange(10) - 4.5] * 3, indexing="ij")
    return ((x * x + y * y + z * z) < 16).astype(np.float32).ravel()


def mc_cases():
    rng = np.random.default_rng(7)
    yield "ball", ball_volume(), (10, 10, 10), 1.0, (0, 0, 0), 0.5, False
    for t in range(6):
        dims = tuple(int(v) for v in rng.integers(2, 20, 3))
        vol = rng.standard_normal(dims[0] * dims[1] * dims[2]).astype(np.float32)
        if t % 2:
            vol[rng.random(vol.size) < 0.15 * t] = np.nan
        yield f"rand{t}", vol, dims, (0.3, 0.5, 0.7), (1, -2, 3), 0.1, t % 3 == 0
    n = 36
    g = (np.arange(n) + 0.5) / n * 3 - 1.5
    Z, Y, X = np.meshgrid(g, g, g, indexing="ij")
    sdf = (np.sqrt(X * X + Y * Y + Z * Z) - 1).astype(np.float32)
    sdf[np.abs(sdf) > 0.12] = np.nan  # narrow band with NaN outside, as mcOffsetMesh produces
    yield "band", sdf.ravel(), (n, n, n), 3 / n, (-1.5, -1.5, -1.5), 0.02, True


def test_known_answer_ball_is_closed_midpoint_mesh():
    # MRMarchingCubesTests.cpp: a 0/1 ball volume at iso 0.5 equals the binary-volume mesh, whose vertices are edge midpoints
    v, t = portlib.marching_cubes(ball_volume(), (10, 10, 10), 1.0, (0, 0, 0), 0.5)
    assert len(t) > 0 and len(v) - 3 * len(t) // 2 + len(t) == 2  # closed genus-0 surface
    frac = v - np.floor(v)
    assert np.all(np.isin(frac, [0.0, 0.5])) and np.all((frac == 0.0).sum(1) == 1)
    assert t.min() == 0 and t.max() == len(v) - 1


def test_two_slices_required_and_empty():
    with pytest.raises(RuntimeError):
        portlib.marching_cubes(np.zeros(9, np.float32), (3, 3, 1), 1.0, (0, 0, 0), 0.5)
    v, t = portlib.marching_cubes(np.zeros(0, np.float32), (0, 3, 3), 1.0, (0, 0, 0), 0.5)
    assert len(v) == 0 and len(t) == 0


@needs_ref
@pytest.mark.parametrize("case", list(mc_cases()), ids=lambda c: c[0])
def test_port_equals_reference_bit_exact(case):
    _, vol, dims, vs, org, iso, li = case
    a = reflib.marching_cubes(vol, dims, vs, org, iso, li)
    b = portlib.marching_cubes(vol, dims, vs, org, iso, li)
    assert a[0].shape == b[0].shape and a[1].shape == b[1].shape
    assert np.array_equal(a[0].view(np.uint32), b[0].view(np.uint32))
    assert np.array_equal(a[1], b[1])


@pytest.mark.skipif(not os.path.exists(REF_MC), reason="reference sources not present")
def test_case_table_equals_reference_table():
    src = open(REF_MC).read()
    i = src.index("cTriangleTable = {")
    rows = [[int(x) for x in re.findall(r"-?\d+", m)] for m in re.findall(r"TriangulationPlan\{([^}]*)\}", src[i:src.index("};", i)])]
    tab = portlib.mc_table()
    assert len(rows) == 256
    for k in range(256):
        assert [int(x) for x in tab[k] if x >= 0] == rows[k], k
