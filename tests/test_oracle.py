"""CPU tests: the C restatement (oracle/port.c) against the unmodified reference (oracle/_ref, when built), the
reference's own known-answer tests for this path, and the committed golden fixtures."""
import os

import numpy as np
import pytest

from meshlib_b200 import make_basis_xf, meshgen
from oracle import portlib, reflib

GOLD = os.path.join(os.path.dirname(__file__), "golden")
needs_ref = pytest.mark.skipif(not reflib.available(), reason="oracle/_ref not built (needs /root/reference)")


def test_triangle_solid_angle_known_answers():
    # /root/reference/source/MRTest/MRDipoleTests.cpp:12-51
    tri = [[0, 0, 0], [1, 0, 0], [0, 1, 0]]
    x = portlib.triangle_solid_angle([1 / 3, 1 / 3, 1e-6], tri)
    assert abs(x + 2 * np.pi) < 1e-3 or abs(x - 2 * np.pi) < 1e-3
    y = portlib.triangle_solid_angle([1 / 3, 1 / 3, -1e-6], tri)
    assert abs(x + y) < 1e-3 and x * y < 0
    # in the plane of the triangle, outside it: exactly zero
    for p in ([-1, -1, 0], [2, 2, 0], [0.5, -1, 0]):
        assert portlib.triangle_solid_angle(p, tri) == 0.0
    # exactly at a vertex: zero
    for p in tri:
        assert portlib.triangle_solid_angle(p, tri) == 0.0


def test_tree_known_answers():
    # MRAABBTreeTests.cpp:10-27: node count = 2F-1, root box = mesh bbox (insignificantly expanded)
    pts, tris = meshgen.make_uv_sphere(1, 8, 8)
    m = portlib.PortMesh(pts, tris)
    nodes = m.nodes()
    assert len(nodes) == 2 * len(tris) - 1
    assert (nodes["bmin"][0] <= pts.min(0)).all() and (nodes["bmax"][0] >= pts.max(0)).all()
    assert np.allclose(nodes["bmin"][0], pts.min(0), atol=1e-6) and np.allclose(nodes["bmax"][0], pts.max(0), atol=1e-6)
    leaves = nodes[nodes["r"] < 0]
    assert sorted(leaves["l"].tolist()) == list(range(len(tris)))
    inner = np.nonzero(nodes["r"] >= 0)[0]
    assert (nodes["l"][inner] == inner + 1).all() and (nodes["r"][inner] > inner + 1).all()


def test_chunk_iterator_known_answers():
    # /root/reference/source/MRMesh/MRChunkIterator.cpp:52-80 (TEST(MRMesh, ChunkIterator))
    assert portlib.split_by_chunks(0, 4, 0) == []
    assert portlib.split_by_chunks(8, 0, 0) == []
    assert portlib.split_by_chunks(8, 4, 4) == []
    assert portlib.split_by_chunks(8, 4, 0) == [(0, 4), (4, 4)]
    assert portlib.split_by_chunks(9, 4, 0) == [(0, 4), (4, 4), (8, 1)]
    assert portlib.split_by_chunks(10, 4, 1) == [(0, 4), (3, 4), (6, 4)]
    assert portlib.split_by_chunks(11, 4, 1) == [(0, 4), (3, 4), (6, 4), (9, 2)]


def test_closed_mesh_winding_inside_outside():
    # MRDipoleTests.cpp:35-39 convention: closed outward mesh => w ~ 1 inside, ~ 0 outside
    pts, tris = meshgen.make_uv_sphere(1, 32, 32)
    m = portlib.PortMesh(pts, tris)
    w = m.calc_from_vector([[0, 0, 0], [0.5, 0.2, -0.3], [2, 0, 0], [0, -3, 1]])
    # the beta=2 dipole far-field is a few-percent approximation on a coarse mesh
    assert abs(w[0] - 1) < 0.05 and abs(w[1] - 1) < 0.05 and abs(w[2]) < 0.05 and abs(w[3]) < 0.05
    d = m.find_projection_dist_sq([[2, 0, 0]])
    assert abs(np.sqrt(d[0]) - 1) < 1e-2


@needs_ref
@pytest.mark.parametrize("mesh", ["torus", "sphere", "noisy", "holes"])
def test_port_equals_reference_bit_exact(mesh):
    if mesh == "torus":
        ref = reflib.RefMesh.torus(2, 1, 48, 24)
        pts, tris = ref.points, ref.tris
    elif mesh == "sphere":
        ref = reflib.RefMesh.uv_sphere(1, 40, 20)
        pts, tris = ref.points, ref.tris
    else:
        pts, tris = meshgen.make_uv_sphere(1, 48, 24)
        pts = meshgen.add_noise(pts, 5e-3)
        if mesh == "holes":
            tris = tris[np.arange(len(tris)) % 7 != 3]
        ref = reflib.RefMesh.from_arrays(pts, tris)
        tris = ref.tris
    port = portlib.PortMesh(pts, tris)
    assert ref.nodes().tobytes() == port.nodes().tobytes()
    assert ref.dipoles().tobytes() == port.dipoles().tobytes()
    q = meshgen.query_points(5000, -2.5, 2.5, seed=7)
    assert ref.calc_from_vector(q).tobytes() == port.calc_from_vector(q).tobytes()
    assert ref.calc_from_vector(q[:500], beta=3.0, skip_face=11).tobytes() == port.calc_from_vector(q[:500], beta=3.0, skip_face=11).tobytes()
    assert ref.find_projection_dist_sq(q).tobytes() == port.find_projection_dist_sq(q).tobytes()
    assert ref.find_projection_dist_sq(q, up=0.25, lo=0.01).tobytes() == port.find_projection_dist_sq(q, up=0.25, lo=0.01).tobytes()
    assert np.array_equal(ref.calc_self_intersections(), port.calc_self_intersections())
    dims, xf = (20, 18, 16), make_basis_xf((-2.6, -2.4, -1.4), (0.26, 0.27, 0.18))
    assert ref.calc_from_grid(dims, xf).tobytes() == port.calc_from_grid(dims, xf).tobytes()
    assert ref.calc_from_grid_with_distances(dims, xf).tobytes() == port.calc_from_grid_with_distances(dims, xf).tobytes()
    a = ref.calc_from_grid_with_distances(dims, xf, min_dist_sq=0.01, max_dist_sq=0.5, null_outside=True)
    b = port.calc_from_grid_with_distances(dims, xf, min_dist_sq=0.01, max_dist_sq=0.5, null_outside=True)
    assert a.tobytes() == b.tobytes() and np.isnan(a).any() and not np.isnan(a).all()
    a = ref.calc_from_grid_with_distances(dims, xf, min_dist_sq=0.01, max_dist_sq=0.5, null_outside=False, threshold=0.3)
    b = port.calc_from_grid_with_distances(dims, xf, min_dist_sq=0.01, max_dist_sq=0.5, null_outside=False, threshold=0.3)
    assert a.tobytes() == b.tobytes()


def test_port_against_golden_fixture():
    # tests/golden/ref_golden.npz was produced by the UNMODIFIED reference (tests/golden/make_golden.py)
    g = np.load(os.path.join(GOLD, "ref_golden.npz"))
    for name in ("torus", "noisy_sphere"):
        pts, tris = g[f"{name}_points"], g[f"{name}_tris"]
        port = portlib.PortMesh(pts, tris)
        assert port.nodes().tobytes() == g[f"{name}_nodes"].tobytes()
        assert port.dipoles().tobytes() == g[f"{name}_dipoles"].tobytes()
        assert port.calc_from_vector(g[f"{name}_q"]).tobytes() == g[f"{name}_w"].tobytes()
        dims, xf = tuple(int(x) for x in g[f"{name}_dims"]), g[f"{name}_xf"]
        assert port.calc_from_grid_with_distances(dims, xf).tobytes() == g[f"{name}_sdf"].tobytes()
        o = g[f"{name}_band_opts"]
        got = port.calc_from_grid_with_distances(dims, xf, min_dist_sq=float(o[0]), max_dist_sq=float(o[1]), null_outside=True)
        assert got.tobytes() == g[f"{name}_sdf_band"].tobytes()
        assert np.array_equal(port.calc_self_intersections(), g[f"{name}_selfint"])


def test_projection_restatement_bit_exact_vs_reference():
    """full MeshProjectionResult (face, point, barycentrics, distSq) of the C restatement against
    MR::PointsToMeshProjector::findProjections, with limits and with point / non-rigid tree transforms"""
    from oracle import reflib
    if not reflib.available():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    pts, tris = meshgen.make_torus(2, 1, 48, 24)
    q = meshgen.query_points(4000, -3.5, 3.5, seed=3)
    m, r = portlib.PortMesh(pts, tris), reflib.RefMesh.from_arrays(pts, tris)
    pxf = np.array([0.8, 0.6, 0, -0.6, 0.8, 0, 0, 0, 1, 0.1, -0.2, 0.3], np.float32)
    txf = np.array([1.2, 0.1, 0, 0, 0.9, 0.2, 0.05, 0, 1.1, 0.3, 0.1, -0.2], np.float32)
    for kw in (dict(), dict(up=0.25), dict(lo=0.05), dict(point_xf=pxf), dict(point_xf=pxf, tree_xf=txf), dict(tree_xf=txf, up=1.0)):
        a, b = m.find_projections(q, **kw), r.find_projections(q, **kw)
        for f in ("face", "point", "a", "b", "dist_sq"):
            assert np.array_equal(a[f].view(np.uint32), b[f].view(np.uint32)), (kw.keys(), f)
        assert np.array_equal(b["edge"] >= 0, b["face"] >= 0)


@needs_ref
@pytest.mark.parametrize("name", ["torus", "noisy", "holes"])
def test_projection_normal_arithmetic_of_the_device_header_vs_reference(name):
    """SURVEY 8f N2 (ProjectionNormal): the header the device kernels are built from (b2fwn_normals_math.h), run on the host
    like the kernels run it, against the reference's topology-based pseudonormals and signedDistanceToMesh: face and edge
    pseudonormals bit for bit, vertex pseudonormals up to the ring's summation order, voxel centres bit for bit, and the
    signed distances of query points identical wherever the reference's dot product is not within rounding of zero"""
    from oracle import normalslib
    if not normalslib.available():
        pytest.skip("oracle/liboracle_normals.so not built")
    if name == "torus":
        pts, tris = meshgen.make_torus(2, 1, 96, 48)
    else:
        pts, tris = meshgen.make_uv_sphere(1, 80, 40)
        pts = meshgen.add_noise(pts, 4e-3)
        if name == "holes":
            tris = tris[np.arange(len(tris)) % 9 != 4]
    ref = reflib.RefMesh.from_arrays(pts, tris)
    rp, rt = ref.points, ref.tris
    fn, vn, adj = tabs = normalslib.tables(rp, rt)
    rfn, rvn, ren = ref.pseudonormals(len(rt), len(rp))
    assert fn.tobytes() == rfn.tobytes()
    assert np.abs(vn - rvn).max() < 2e-6
    assert (adj < 0).any() == (name == "holes")
    other = np.where(adj >= 0, adj, np.arange(len(rt))[:, None])
    s = fn[:, None, :] + fn[other]
    length = np.sqrt((s[..., 0] * s[..., 0] + s[..., 1] * s[..., 1]) + s[..., 2] * s[..., 2], dtype=np.float32)
    with np.errstate(divide="ignore", invalid="ignore"):
        en = np.where((adj >= 0)[..., None], (np.float32(1) / length)[..., None] * s, fn[:, None, :]).astype(np.float32)
    assert en.tobytes() == ren.tobytes()
    # signs: the reference's own projections (ours are bit-identical, N1) through the header's sign rule
    q = meshgen.query_points(30000, float(rp.min()) - 0.6, float(rp.max()) + 0.6, seed=11)
    q = np.concatenate([q, rp[::7] + np.float32(0.05) * rvn[::7], rp[::5] - np.float32(0.03) * rvn[::5]]).astype(np.float32)
    for kw in (dict(null_outside=False), dict(min_dist_sq=0.01, max_dist_sq=0.49, null_outside=True)):
        want, dots = ref.signed_distance_proj_normal(q, **kw)
        lo, up = (kw["min_dist_sq"], kw["max_dist_sq"]) if kw["null_outside"] else (0.0, reflib.FLT_MAX)
        rec = ref.find_projections(q, up=up, lo=lo)
        got = normalslib.sign(rec, q, rt, tabs, **kw)
        assert np.array_equal(np.isnan(got), np.isnan(want))
        ok = ~np.isnan(want)
        assert ok.any() and (kw["null_outside"] is False or (~ok).any())
        assert np.abs(got[ok]).tobytes() == np.abs(want[ok]).tobytes()
        clear = ok & (np.abs(dots) > 1e-6 * np.abs(want))
        assert clear.sum() > 0.99 * ok.sum()
        assert np.array_equal(np.signbit(got[clear]), np.signbit(want[clear]))
        assert np.signbit(want[ok]).any() and not np.signbit(want[ok]).all()
    # on-vertex and on-edge closest points must actually occur in the sample
    a, b = rec["a"], rec["b"]
    eps = 10 * np.finfo(np.float32).eps
    assert ((a <= eps) & (b <= eps)).any() and ((b <= eps) & (a > eps) & (np.float32(1) - a - b > eps)).any()
    # voxel centres of meshToDistanceFunctionVolume
    args = ((-1.37, 0.21, 3.3), (0.013, 0.07, 0.11), (17, 9, 5))
    assert normalslib.voxel_centres(*args).tobytes() == reflib.voxel_centres(*args).tobytes()


# ---- golden fixture of the SURVEY 8f rows (tests/golden/ref_golden_next.npz) ----------------------------------------------

class _PortAndHeader:
    """the CPU restatements: oracle/port.c for projections, winding numbers and the refit, the product's normals header run on
    the host (oracle/normals_host.cpp) for the ProjectionNormal sign"""

    def __init__(self, pts, tris):
        from oracle import normalslib
        self.m, self.pts, self.tris = portlib.PortMesh(pts, tris), pts, tris
        self.tabs = normalslib.tables(pts, tris) if normalslib.available() else None

    def find_projections(self, q, up, lo):
        return self.m.find_projections(q, up=up, lo=lo)

    def signed_distance_pn(self, q, min_dist_sq, max_dist_sq, null_outside):
        from oracle import normalslib
        if self.tabs is None:
            return None
        lo, up = (min_dist_sq, max_dist_sq) if null_outside else (0.0, reflib.FLT_MAX)
        rec = self.m.find_projections(q, up=up, lo=lo)
        return normalslib.sign(rec, q, self.tris, self.tabs, min_dist_sq, max_dist_sq, null_outside)

    def make_signed(self, unsigned, dims, xf, threshold, beta):
        from oracle import checker
        return checker.ref_make_signed_by_winding_number(self.m, unsigned, dims, xf, threshold, beta)[0]

    def refit(self, new_points):
        self.m.update_points(new_points)
        return self.m.nodes(), self.m.dipoles(), self.m.calc_from_vector(self.q)


class _Reference:
    def __init__(self, pts, tris):
        self.m = reflib.RefMesh.from_arrays(pts, tris)
        assert self.m.tris.tobytes() == np.ascontiguousarray(tris, np.int32).tobytes()  # fixture holds getTriangulation's order

    def find_projections(self, q, up, lo):
        return self.m.find_projections(q, up=up, lo=lo)

    def signed_distance_pn(self, q, min_dist_sq, max_dist_sq, null_outside):
        return self.m.signed_distance_proj_normal(q, min_dist_sq, max_dist_sq, null_outside)[0]

    def make_signed(self, unsigned, dims, xf, threshold, beta):
        from oracle import checker
        return checker.ref_make_signed_by_winding_number(self.m, unsigned, dims, xf, threshold, beta)[0]

    def refit(self, new_points):
        self.m.update_points(new_points)
        return self.m.nodes(), self.m.dipoles(), self.m.calc_from_vector(self.q)


@pytest.mark.parametrize("name", ["torus", "noisy_sphere"])
def test_restatements_against_golden_next_fixture(name):
    """projection records, ProjectionNormal signed distances and the makeSignedByWindingNumber loop of the CPU restatements
    against vectors generated from the unmodified reference (runs without /root/reference)"""
    import golden_next_check as gc
    g = gc.load()

    def make(pts, tris):
        m = _PortAndHeader(pts, tris)
        m.q = g[f"{name}_q"]
        return m
    ran = gc.check(make, g, name, w_atol=0.0)
    assert f"{name}_proj" in ran and f"{name}_proj_limits" in ran and f"{name}_signed" in ran and f"{name}_refit" in ran
    from oracle import normalslib
    assert (f"{name}_pn" in ran and f"{name}_pn_band" in ran) or not normalslib.available()
    # face normals of the header bit for bit, vertex pseudonormals up to the ring order
    if normalslib.available():
        fn, vn, _adj = normalslib.tables(g[f"{name}_points"], g[f"{name}_tris"])
        assert fn.tobytes() == g[f"{name}_face_normals"].tobytes()
        assert np.abs(vn - g[f"{name}_vert_normals"]).max() < 2e-6


@needs_ref
@pytest.mark.parametrize("name", ["torus", "noisy_sphere"])
def test_golden_next_fixture_is_what_the_reference_gives(name):
    import golden_next_check as gc
    g = gc.load()

    def make(pts, tris):
        r = _Reference(pts, tris)
        r.q = g[f"{name}_q"]
        return r
    assert len(gc.check(make, g, name, w_atol=0.0)) == 6
