"""GPU parity tests at the sizes BASELINE.json names (C2 .. C5) and at the places only large inputs reach: trees of 24 levels,
slivers at the poles of a 10 M-face UV sphere, voxel ids beyond 2^31, the heavy-first brick order, coordinates far from the
origin.  The CUDA path through the C ABI against the unmodified reference (oracle/_ref/libmrref.so, all host threads) on
samples the reference finishes in seconds."""
import numpy as np
import pytest

import meshlib_b200 as mb
from meshlib_b200 import meshgen
from oracle import checker

pytestmark = pytest.mark.gpu


def _oracle(pts, tris):
    orc = checker.best_oracle(pts, tris)
    if orc.kind != "reference":
        pytest.skip("needs oracle/_ref/libmrref.so (the C restatement is single-threaded: minutes at these sizes)")
    return orc


@pytest.fixture(scope="module")
def c5():
    """C5: 9,998,244-face noisy UV sphere, 2048^3 grid, band of 3 voxels (bench.py's c5_band)"""
    pts, tris = meshgen.make_uv_sphere(1, 3162, 1581)
    pts = meshgen.add_noise(pts, 1e-3, seed=42)
    orc = _oracle(pts, tris)
    fwn = mb.FastWindingNumber(pts, orc.tris)
    dims, origin, voxel = meshgen.grid_for_bbox(pts, 2048, 2.0)
    yield pts, orc, fwn, dims, mb.make_basis_xf(origin, voxel), float(voxel)
    fwn.close()


def _check_layers(orc, fwn_vol_of, dims, xf, o, layers):
    layer = dims[0] * dims[1]
    for z in layers:
        want = orc.calc_grid_with_distances_layers(dims, xf, z, z + 1, **o)
        got = fwn_vol_of(z)
        assert got.shape == want.shape
        diff = np.flatnonzero(~np.isnan(got) & ~np.isnan(want) & (np.signbit(got) != np.signbit(want)))
        w = None
        if len(diff):  # excused only where the reference's own winding number is within 1e-3 of the threshold
            idx = diff.astype(np.int64)
            pos = np.stack([idx % dims[0], idx // dims[0], np.full(len(idx), z)], 1).astype(np.float32)
            x = np.asarray(xf, np.float32)
            centres = (pos @ x[:9].reshape(3, 3).T + x[9:]).astype(np.float32)
            w = np.full(layer, 10.0, np.float32)
            w[diff] = orc.calc_from_vector(centres, beta=o["beta"])
        rep = checker.compare_sdf(got, want, w, threshold=o["threshold"])
        assert rep["ok"] and rep["dist_bit_exact_frac"] == 1.0, (z, rep)


def test_c5_tree_depth_and_top_slab_beyond_2_to_31_voxels(c5):
    """layers 2040..2047 of the 2048^3 grid: flat voxel ids up to 8.59e9 (64-bit index arithmetic), the pole of the sphere with
    its slivers inside the band, host output through the chunked call"""
    pts, orc, fwn, dims, xf, voxel = c5
    st = fwn.stats()
    assert st.num_leaves == 9998244 and st.num_nodes == 2 * 9998244 - 1
    o = dict(min_dist_sq=0.0, max_dist_sq=float(np.float32((3.0 * voxel) ** 2)), null_outside=False, threshold=0.5, beta=2.0)
    opts = mb.DistOptions(o["min_dist_sq"], o["max_dist_sq"], False, 0.5, 2.0)
    layer = dims[0] * dims[1]
    slab = fwn.calc_from_grid_with_distances_layers(dims, xf, 2040, 2048, opts)
    assert (2047 * layer + layer - 1) > 2 ** 31
    _check_layers(orc, lambda z: slab[(z - 2040) * layer:(z - 2039) * layer], dims, xf, o, [2040, 2045, 2047])
    # the same layers out of a device-resident slab that starts elsewhere (bricks are relative to the launch)
    import torch
    d = torch.empty(12 * layer, dtype=torch.float32, device="cuda")
    fwn.calc_from_grid_with_distances_device(dims, xf, d.data_ptr(), opts, z_begin=2036, z_end=2048)
    assert d[4 * layer:].cpu().numpy().tobytes() == slab.tobytes()


def test_c5_middle_and_pole_layers_default_options(c5):
    """default DistanceToMeshOptions (exact distance everywhere): 24-level tree, full-range closest-point search"""
    pts, orc, fwn, dims, xf, voxel = c5
    o = dict(min_dist_sq=0.0, max_dist_sq=mb.FLT_MAX, null_outside=True, threshold=0.5, beta=2.0)
    layer = dims[0] * dims[1]
    # one quarter of a layer keeps the reference's share of this test near ten seconds: sub-grid with the same voxel centres
    sub = (512, 512, dims[2])
    x = np.asarray(xf, np.float32).copy()
    shift = np.float32(768)
    x[9] = np.float32(x[9] + x[0] * shift)  # both sides get the same sub-grid transform; it need not reproduce the big grid's centres
    x[10] = np.float32(x[10] + x[4] * shift)
    for z in (6, 1024):
        got = fwn.calc_from_grid_with_distances_layers(sub, x, z, z + 1, mb.DistOptions())
        _check_layers(orc, lambda _z: got, sub, x, o, [z])


def test_c2_ten_million_points_against_a_million_triangles():
    pts, tris = meshgen.make_uv_sphere(1, 1000, 500)
    orc = _oracle(pts, tris)
    fwn = mb.FastWindingNumber(pts, orc.tris)
    q = meshgen.query_points(10_000_000)
    got = fwn.calc_from_vector(q)
    idx = np.arange(0, len(q), 50)
    want = orc.calc_from_vector(q[idx])
    rep = checker.compare_w(got[idx], want, atol=2e-6)
    assert rep["ok"], rep
    inside = np.linalg.norm(q.astype(np.float64), axis=1) < 0.99
    outside = np.linalg.norm(q.astype(np.float64), axis=1) > 1.01
    # the dipole far field at beta = 2 is an approximation (a few per cent right under the surface), the same one on both sides
    assert np.abs(got[inside] - 1).max() < 0.15 and np.abs(got[outside]).max() < 0.15
    fwn.close()


def test_c4_five_million_faces_self_intersections():
    pts, tris = meshgen.two_tori(2.0, 0.7, 1600, 800)
    assert len(tris) == 5_120_000
    orc = _oracle(pts, tris)
    fwn = mb.FastWindingNumber(pts, orc.tris)
    import torch
    d_bits = torch.zeros((len(tris) + 63) // 64, dtype=torch.int64, device="cuda")
    d_w = torch.empty(len(tris), dtype=torch.float32, device="cuda")
    fwn.calc_self_intersections_device(d_bits.data_ptr(), d_w.data_ptr())
    got = fwn.calc_self_intersections()
    assert got.tobytes() == d_bits.cpu().numpy().tobytes()
    # the reference on ALL faces (MRFastWindingNumber.cpp:40-52, TBB); faces whose winding number lies within 1e-3 of 0 or 1 are
    # excused (our own w decides the band: it is within 2e-6 of the reference's, test_calc_from_vector)
    want = orc.calc_self_intersections()
    rep = checker.compare_bits(got, want, d_w.cpu().numpy(), len(tris))
    assert rep["ok"] and rep["mismatch_in_band"] < 1000, rep
    bits = np.unpackbits(got.view(np.uint8), bitorder="little")[:len(tris)].astype(bool)
    assert 0.05 < bits.mean() < 0.6  # the interpenetrating parts, not everything and not nothing
    fwn.close()


def test_c3_offset_mesh_of_two_million_faces_at_512_cubed():
    """generalOffsetMesh( Standard, HoleWindingRule ) incl. marching cubes at the C3 size: grid identical, and the TriMesh identical
    to the reference's marching cubes of the reference's own SDF (counts, triangles, vertices within the distance tolerance)"""
    pts, tris = meshgen.make_uv_sphere(1, 1414, 707)
    pts = meshgen.add_noise(pts, 1e-3, seed=42)
    orc = _oracle(pts, tris)
    fwn = mb.FastWindingNumber(pts, orc.tris)
    voxel, offset = 2.04 / 508.0, 0.02
    v, t, info = fwn.mc_offset_mesh(offset, voxel, close_holes=True)
    wv, wt, wsdf, worigin, wdims = checker.ref_mc_offset_mesh(orc, pts, orc.tris, offset, voxel, close_holes=True)
    assert info["dims"] == wdims and np.array_equal(info["origin"], worigin)
    assert min(wdims) >= 500
    assert (len(v), len(t)) == (len(wv), len(wt)) and len(t) > 1_000_000
    assert np.array_equal(t, wt)
    assert np.abs(v - wv).max() < 1e-5
    fwn.close()


def test_mesh_far_from_the_origin_with_a_fine_voxel():
    """a 6 cm torus at (1000, 1000, 1000) with a 1 mm voxel: coordinates carry 6e-5 of rounding, 1/16 of a voxel -- the FP32 reject
    filter of the distance walk (its allowance scales with |p|) and the padding of the brick cull must still never drop the
    reference's closest triangle"""
    p, t = meshgen.make_torus(0.02, 0.01, 96, 48)
    pts = (p.astype(np.float64) + 1000.0).astype(np.float32)
    orc = _oracle(pts, t)
    fwn = mb.FastWindingNumber(pts, orc.tris)
    dims = (72, 72, 40)
    origin = np.array([1000 - 0.036, 1000 - 0.036, 1000 - 0.02], np.float32)
    xf = mb.make_basis_xf(origin, 1e-3)
    w = orc.calc_from_grid(dims, xf)
    for o in (dict(), dict(min_dist_sq=0.0, max_dist_sq=float(np.float32(0.003 ** 2)), null_outside=False),
              dict(min_dist_sq=0.0, max_dist_sq=float(np.float32(0.003 ** 2)), null_outside=True)):
        want = orc.calc_from_grid_with_distances(dims, xf, **o)
        opts = mb.DistOptions(o.get("min_dist_sq", 0.0), o.get("max_dist_sq", mb.FLT_MAX), o.get("null_outside", True))
        got = fwn.calc_from_grid_with_distances(dims, xf, opts)
        rep = checker.compare_sdf(got, want, w)
        assert rep["ok"] and rep["dist_bit_exact_frac"] == 1.0, (o, rep)
    nodes, dip = fwn.get_tree()
    assert nodes.tobytes() == orc.nodes().tobytes() and dip.tobytes() == orc.dipoles().tobytes()
    q = (meshgen.query_points(50000, -0.04, 0.04) + np.float32(1000)).astype(np.float32)
    rep = checker.compare_w(fwn.calc_from_vector(q), orc.calc_from_vector(q), atol=2e-6)
    assert rep["ok"], rep
    fwn.close()
