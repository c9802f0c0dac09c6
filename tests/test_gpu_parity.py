"""GPU parity tests (run on the B200 box): the CUDA path through the C ABI against the oracle (the unmodified reference
in oracle/_ref when it travelled with the snapshot, else the C restatement) on the same seeded inputs."""
import os

import numpy as np
import pytest

import meshlib_b200 as mb
from meshlib_b200 import meshgen
from oracle import checker, portlib

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _mesh(name):
    if name == "torus":
        return meshgen.make_torus(2, 1, 96, 48)
    if name == "sphere":
        return meshgen.make_uv_sphere(1, 64, 32)
    p, t = meshgen.make_uv_sphere(1, 80, 40)
    p = meshgen.add_noise(p, 4e-3)
    if name == "holes":
        t = t[np.arange(len(t)) % 9 != 4]
    if name == "deleted":
        t = t.copy()
        t[::13] = -1  # deleted faces keep their slot (MeshTopology::getTriangulation)
    return p, t


@pytest.fixture(scope="module", params=["torus", "sphere", "noisy", "holes"])
def pair(request):
    pts, tris = _mesh(request.param)
    orc = checker.best_oracle(pts, tris)
    if orc.kind == "reference":
        tris = orc.tris  # the reference's own per-face vertex order
    fwn = mb.FastWindingNumber(pts, tris)
    yield request.param, pts, tris, orc, fwn
    fwn.close()


def test_tree_and_dipoles_bit_exact(pair):
    _n, _p, _t, orc, fwn = pair
    nodes, dip = fwn.get_tree()
    assert nodes.tobytes() == orc.nodes().tobytes()
    assert dip.tobytes() == orc.dipoles().tobytes()


def test_calc_from_vector(pair):
    _n, pts, _t, orc, fwn = pair
    q = meshgen.query_points(40000, float(pts.min()) - 0.7, float(pts.max()) + 0.7, seed=99)
    got = fwn.calc_from_vector(q)
    want = orc.calc_from_vector(q)
    rep = checker.compare_w(got, want, atol=2e-6)
    assert rep["ok"], rep
    got = fwn.calc_from_vector(q[:3000], beta=3.0, skip_face=5)
    rep = checker.compare_w(got, orc.calc_from_vector(q[:3000], beta=3.0, skip_face=5), atol=2e-6)
    assert rep["ok"], rep
    assert len(fwn.calc_from_vector(np.zeros((0, 3), np.float32))) == 0


def test_calc_from_grid(pair):
    _n, pts, _t, orc, fwn = pair
    dims = (37, 29, 23)  # ragged: not multiples of the brick
    lo, hi = pts.min(0) - 0.3, pts.max(0) + 0.3
    xf = mb.make_basis_xf(lo, (hi - lo) / np.asarray(dims, np.float32))
    rep = checker.compare_w(fwn.calc_from_grid(dims, xf), orc.calc_from_grid(dims, xf), atol=checker.W_GRID_ATOL)
    assert rep["ok"], rep


def test_calc_from_grid_with_distances(pair):
    _n, pts, _t, orc, fwn = pair
    dims = (40, 33, 26)
    lo, hi = pts.min(0) - 0.3, pts.max(0) + 0.3
    xf = mb.make_basis_xf(lo, (hi - lo) / np.asarray(dims, np.float32))
    w = orc.calc_from_grid(dims, xf)
    got = fwn.calc_from_grid_with_distances(dims, xf)
    rep = checker.compare_sdf(got, orc.calc_from_grid_with_distances(dims, xf), w)
    assert rep["ok"], rep
    assert rep["dist_bit_exact_frac"] == 1.0, rep  # the closest point is evaluated in double exactly like the reference
    # narrow band with NaN outside (mcOffsetMesh's call, MROffset.cpp:176-183); minDistSq > 0 makes the reference's
    # early exit order dependent, so in-band distances only have to agree where the reference did not stop early
    lo2, hi2 = 0.02 ** 2, 0.35 ** 2
    got = fwn.calc_from_grid_with_distances(dims, xf, mb.DistOptions(lo2, hi2, True))
    want = orc.calc_from_grid_with_distances(dims, xf, min_dist_sq=lo2, max_dist_sq=hi2, null_outside=True)
    rep = checker.compare_sdf(got, want, w)
    assert rep["ok"], rep
    assert np.isnan(want).any() and not np.isnan(want).all()
    # no NaN, custom threshold and beta
    got = fwn.calc_from_grid_with_distances(dims, xf, mb.DistOptions(0.0, hi2, False, 0.3, 2.5))
    want = orc.calc_from_grid_with_distances(dims, xf, min_dist_sq=0.0, max_dist_sq=hi2, null_outside=False, threshold=0.3, beta=2.5)
    w2 = orc.calc_from_grid(dims, xf, beta=2.5)
    rep = checker.compare_sdf(got, want, w2, threshold=0.3)
    assert rep["ok"], rep


def test_self_intersections(pair):
    _n, _p, tris, orc, fwn = pair
    got = fwn.calc_self_intersections()
    port = portlib.PortMesh(orc.points if hasattr(orc, "points") else _p, tris)
    want, w = port.calc_self_intersections(return_w=True)
    rep = checker.compare_bits(got, want, w, len(tris))
    assert rep["ok"], rep
    if orc.kind == "reference":
        rep = checker.compare_bits(got, orc.calc_self_intersections(), w, len(tris))
        assert rep["ok"], rep


def test_two_tori_self_intersections_flagged():
    pts, tris = meshgen.two_tori(2, 0.7, 96, 48)
    fwn = mb.FastWindingNumber(pts, tris)
    port = portlib.PortMesh(pts, tris)
    got = fwn.calc_self_intersections()
    want, w = port.calc_self_intersections(return_w=True)
    rep = checker.compare_bits(got, want, w, len(tris))
    assert rep["ok"], rep
    assert rep["set"] > 100  # faces of each torus inside the other have w ~ 2


def test_deleted_faces_and_empty_mesh():
    pts, tris = _mesh("deleted")
    orc = portlib.PortMesh(pts, tris)
    fwn = mb.FastWindingNumber(pts, tris)
    nodes, dip = fwn.get_tree()
    assert nodes.tobytes() == orc.nodes().tobytes() and dip.tobytes() == orc.dipoles().tobytes()
    q = meshgen.query_points(5000, -1.5, 1.5, seed=3)
    assert checker.compare_w(fwn.calc_from_vector(q), orc.calc_from_vector(q))["ok"]
    got = fwn.calc_self_intersections()
    want, w = orc.calc_self_intersections(return_w=True)
    assert checker.compare_bits(got, want, w, len(tris))["ok"]
    # empty mesh: no tree; winding numbers are 0, distances stay at maxDistSq => NaN with the default options
    e = mb.FastWindingNumber(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.int32))
    assert (e.calc_from_vector(q[:10]) == 0).all()
    assert len(e.calc_self_intersections()) == 0
    # zero-sized grids
    assert len(fwn.calc_from_grid((0, 5, 5), mb.make_basis_xf((0, 0, 0), 1))) == 0
    assert len(fwn.calc_from_grid_with_distances((4, 4, 0), mb.make_basis_xf((0, 0, 0), 1))) == 0


def test_golden_fixture_from_reference():
    g = np.load(os.path.join(GOLD, "ref_golden.npz"))
    for name in ("torus", "noisy_sphere"):
        fwn = mb.FastWindingNumber(g[f"{name}_points"], g[f"{name}_tris"])
        nodes, dip = fwn.get_tree()
        assert nodes.tobytes() == g[f"{name}_nodes"].tobytes()
        assert dip.tobytes() == g[f"{name}_dipoles"].tobytes()
        assert checker.compare_w(fwn.calc_from_vector(g[f"{name}_q"]), g[f"{name}_w"])["ok"]
        dims, xf = tuple(int(x) for x in g[f"{name}_dims"]), g[f"{name}_xf"]
        assert checker.compare_w(fwn.calc_from_grid(dims, xf), g[f"{name}_wgrid"], atol=checker.W_GRID_ATOL)["ok"]
        rep = checker.compare_sdf(fwn.calc_from_grid_with_distances(dims, xf), g[f"{name}_sdf"], g[f"{name}_wgrid"])
        assert rep["ok"], rep
        o = g[f"{name}_band_opts"]
        got = fwn.calc_from_grid_with_distances(dims, xf, mb.DistOptions(float(o[0]), float(o[1]), True))
        rep = checker.compare_sdf(got, g[f"{name}_sdf_band"], g[f"{name}_wgrid"])
        assert rep["ok"], rep


def test_by_parts_contract():
    pts, tris = _mesh("torus")
    fwn = mb.FastWindingNumber(pts, tris)
    dims = (24, 20, 19)
    xf = mb.make_basis_xf((-3.2, -3.2, -1.3), (0.27, 0.32, 0.14))
    full = fwn.calc_from_grid_with_distances(dims, xf)
    layer = dims[0] * dims[1]
    for overlap, per_part in ((0, 4), (1, 4), (1, 19), (2, 5), (1, 2)):
        seen = []
        vol = np.full(layer * dims[2], np.float32(7e33))

        def part(data, d, z0):
            assert d[0] == dims[0] and d[1] == dims[1] and len(data) == layer * d[2]
            seen.append((z0, d[2]))
            vol[z0 * layer:(z0 + d[2]) * layer] = data
            return True

        fwn.calc_from_grid_with_distances_by_parts(part, dims, xf, layer_overlap=overlap, max_layers_per_part=per_part)
        want = [(o, s) for o, s in portlib.split_by_chunks(dims[2], per_part, overlap)]  # MRChunkIterator.cpp:25-49
        assert seen == want, (seen, want)
        assert vol.tobytes() == full.tobytes()
    w = fwn.calc_from_grid(dims, xf)
    got = np.empty_like(w)

    def part_w(data, d, z0):
        got[z0 * layer:(z0 + d[2]) * layer] = data
        return True

    fwn.calc_from_grid_by_parts(part_w, dims, xf, layer_overlap=1, max_layers_per_part=6)
    assert got.tobytes() == w.tobytes()
    # consumer failure and cancellation surface as errors, reference wording for cancel (MRExpected.h:42-51)
    with pytest.raises(mb.B2Error):
        fwn.calc_from_grid_by_parts(lambda *a: False, dims, xf, max_layers_per_part=4)
    with pytest.raises(mb.B2Error) as e:
        fwn.calc_from_grid_with_distances(dims, xf, cb=lambda v: False)
    assert str(e.value) == "Operation was canceled"
    prog = []
    fwn.calc_from_grid(dims, xf, cb=lambda v: prog.append(v) or True)
    assert prog[0] == 0.0 and prog[-1] == 1.0 and prog == sorted(prog)


def test_z_slab_device_form_matches_full_grid():
    import torch
    pts, tris = _mesh("sphere")
    fwn = mb.FastWindingNumber(pts, tris)
    dims = (32, 32, 30)
    xf = mb.make_basis_xf((-1.3, -1.3, -1.3), (2.6 / 32, 2.6 / 32, 2.6 / 30))
    full = fwn.calc_from_grid_with_distances(dims, xf)
    out = torch.empty(dims[0] * dims[1] * dims[2], dtype=torch.float32, device="cuda")
    layer = dims[0] * dims[1]
    for z0, z1 in ((0, 7), (7, 8), (8, 30)):
        fwn.calc_from_grid_with_distances_device(dims, xf, out.data_ptr() + 4 * z0 * layer, z_begin=z0, z_end=z1)
    assert out.cpu().numpy().tobytes() == full.tobytes()


@pytest.mark.parametrize("world,chunk,stages,dz", [(2, 4, 2, 30), (3, 8, 1, 50), (8, 4, 4, 64)])
def test_z_interleaved_chunks_match_full_grid(world, chunk, stages, dz):
    """every rank's interleaved share (slabs.ChunkPlan -> ..._device_chunks), written in place, gives the full-grid volume
    bit for bit; band limits on, so the cull / heavy-first machinery runs on the chunked layout too"""
    import torch
    from meshlib_b200 import slabs
    pts, tris = _mesh("noisy")
    fwn = mb.FastWindingNumber(pts, tris)
    dims = (40, 36, dz)
    xf = mb.make_basis_xf((-1.3, -1.3, -1.3), (2.6 / 40, 2.6 / 36, 2.6 / dz))
    for opts in (mb.DistOptions(), mb.DistOptions(0.0, 0.04, False, 0.5, 2.0), mb.DistOptions(0.0004, 0.04, True, 0.5, 2.0)):
        full = fwn.calc_from_grid_with_distances(dims, xf, opts)
        out = torch.full((dims[0] * dims[1] * dz,), 123.0, dtype=torch.float32, device="cuda")
        layer = dims[0] * dims[1]
        plan = slabs.ChunkPlan(world, dz, chunk, stages)
        for r in range(world):
            for s in range(plan.stages):
                a = plan.call_args(r, s)
                if a:
                    zb, ze, zc, zs = a
                    fwn.calc_from_grid_with_distances_device_chunks(dims, xf, out.data_ptr() + 4 * zb * layer, opts, z_begin=zb,
                                                                    z_end=ze, z_chunk=zc, z_stride=zs)
        assert out.cpu().numpy().tobytes() == full.tobytes()
    with pytest.raises(Exception):
        fwn.calc_from_grid_with_distances_device_chunks(dims, xf, out.data_ptr(), z_chunk=6, z_stride=12)


@pytest.mark.parametrize("dims", [(36, 40, 48), (40, 36, 48)])
def test_fused_gather_stores_every_result_into_the_peer_volumes(dims):
    """the _p2p form (multi-GPU gather fused into the kernel): two further volumes stand in for the other ranks' memory;
    after all ranks' calls each of them equals the full-grid result.  dx = 36: rows of 16-byte pieces (float4 stores); dx = 40:
    rows of whole 32-byte sectors, the peers receive one 256-bit store per lane"""
    import torch
    from meshlib_b200 import slabs
    pts, tris = _mesh("noisy")
    fwn = mb.FastWindingNumber(pts, tris)
    world = 3
    xf = mb.make_basis_xf((-1.3, -1.3, -1.3), (2.6 / dims[0], 2.6 / dims[1], 2.6 / dims[2]))
    opts = mb.DistOptions(0.0, 0.04, False, 0.5, 2.0)
    full = fwn.calc_from_grid_with_distances(dims, xf, opts)
    layer = dims[0] * dims[1]
    vols = [torch.full((layer * dims[2],), 7.0, dtype=torch.float32, device="cuda") for _ in range(world)]
    plan = slabs.ChunkPlan(world, dims[2], 4, 1)
    for r in range(world):
        zb, ze, zc, zs = plan.call_args(r, 0)
        off = 4 * zb * layer
        fwn.calc_from_grid_with_distances_device_chunks_p2p(dims, xf, vols[r].data_ptr() + off,
                                                            [vols[q].data_ptr() + off for q in range(world) if q != r], opts,
                                                            z_begin=zb, z_end=ze, z_chunk=zc, z_stride=zs)
    for v in vols:
        assert v.cpu().numpy().tobytes() == full.tobytes()
    ptr, token = fwn.shared_alloc(1 << 20)
    assert ptr and len(token) == 64
    t = torch.as_tensor(mb.DeviceBuffer(ptr, 1 << 18), device="cuda")
    t.fill_(3.0)
    assert float(t.sum()) == 3.0 * (1 << 18)
    del t
    fwn.shared_free(ptr)


@pytest.mark.parametrize("n", [1, 31, 33, 100000])
def test_vector_sizes(n):
    pts, tris = _mesh("sphere")
    fwn = mb.FastWindingNumber(pts, tris)
    orc = portlib.PortMesh(pts, tris)
    q = meshgen.query_points(n, -1.5, 1.5, seed=n)
    assert checker.compare_w(fwn.calc_from_vector(q), orc.calc_from_vector(q))["ok"]


def test_c1_full_size_properties():
    """BASELINE config 1 at full size (torus 65k faces, 128^3): sampled parity + size-independent properties"""
    pts, tris = meshgen.make_torus(2, 1, 256, 128)
    fwn = mb.FastWindingNumber(pts, tris)
    dims = (128, 128, 128)
    xf = mb.make_basis_xf((-3.2, -3.2, -3.2), 0.05)
    sdf = fwn.calc_from_grid_with_distances(dims, xf)
    assert not np.isnan(sdf).any()
    vol = (sdf < 0).sum() * 0.05 ** 3  # torus volume 2*pi^2*R*r^2 = 39.48
    assert abs(vol - 2 * np.pi ** 2 * 2) / 39.48 < 0.01
    g = sdf.reshape(128, 128, 128)
    # |grad| <= 1: neighbouring voxels differ by at most one voxel size (distance is 1-Lipschitz)
    for ax in range(3):
        assert np.abs(np.diff(g, axis=ax)).max() <= 0.05 * 1.0001
    # mirror symmetry of the input in z
    assert np.abs(np.abs(g) - np.abs(g[::-1])).max() < 1e-4
    port = portlib.PortMesh(pts, tris)
    stride = 97
    want = port.calc_from_grid_with_distances(dims, xf, stride=stride)
    idx = np.arange(0, sdf.size, stride)
    wref = port.calc_from_vector(_voxel_points(idx, dims, xf))
    rep = checker.compare_sdf(sdf[idx], want[idx], wref)
    assert rep["ok"], rep


def _voxel_points(idx, dims, xf):
    x = (idx % dims[0]).astype(np.float32)
    y = ((idx // dims[0]) % dims[1]).astype(np.float32)
    z = (idx // (dims[0] * dims[1])).astype(np.float32)
    xf = np.asarray(xf, np.float32)
    out = np.empty((len(idx), 3), np.float32)
    for r in range(3):
        out[:, r] = (xf[3 * r] * x + xf[3 * r + 1] * y + xf[3 * r + 2] * z) + xf[9 + r]
    return out


# ---- marching cubes (a20) ------------------------------------------------------------------------------------------
def _mc_cases():
    from test_oracle_mc import mc_cases
    return list(mc_cases())


@pytest.mark.gpu
@pytest.mark.parametrize("case", _mc_cases(), ids=lambda c: c[0])
def test_marching_cubes_identical_to_oracle(case):
    from oracle import portlib
    _, vol, dims, vs, org, iso, li = case
    fwn = mb.FastWindingNumber(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.int32))
    v, t = fwn.marching_cubes(vol, dims, org, vs, iso, li)
    wv, wt = portlib.marching_cubes(vol, dims, vs, org, iso, li)
    assert v.shape == wv.shape and t.shape == wt.shape
    assert np.array_equal(v.view(np.uint32), wv.view(np.uint32))
    assert np.array_equal(t, wt)


@pytest.mark.gpu
def test_marching_cubes_of_device_sdf_counts_match_oracle():
    """SDF -> MC; counts equal the oracle's MC of the oracle's SDF wherever the sign fields match (north_star)"""
    from oracle import checker, portlib
    pts, tris = meshgen.make_torus(2, 1, 96, 48)
    dims = (48, 48, 40)
    org = (-3.3, -3.3, -1.4)
    xf = mb.make_basis_xf(org, 0.14)
    fwn = mb.FastWindingNumber(pts, tris)
    got = fwn.calc_from_grid_with_distances(dims, xf)
    orc = checker.best_oracle(pts, tris)
    want = orc.calc_from_grid_with_distances(dims, xf)
    v, t = fwn.marching_cubes(got, dims, org, 0.14, 0.05, True)
    wv, wt = portlib.marching_cubes(want, dims, 0.14, org, 0.05, True)
    if np.array_equal(got < 0.05, want < 0.05):
        assert len(v) == len(wv) and len(t) == len(wt)
    assert len(t) > 1000 and len(v) - 3 * len(t) // 2 + len(t) == 0  # closed torus: Euler characteristic 0


# ---- the C++ adapter behind MeshLib's own interface (8b) ------------------------------------------------------------
@pytest.mark.gpu
def test_cpp_adapter_is_a_drop_in_for_the_reference_interface():
    """oracle/_ref/adapter_check (built where the reference headers exist): MR::B200::FastWindingNumber used through
    std::shared_ptr<IFastWindingNumber> / IFastWindingNumberByParts by the reference's own callers' code paths, against
    MR::FastWindingNumber (CPU) and the reference's MarchingCubesByParts"""
    import subprocess
    exe = os.path.join(os.path.dirname(GOLD), "..", "oracle", "_ref", "adapter_check")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/adapter_check was not built (needs the reference headers)")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    print(r.stdout, r.stderr)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "FAIL" not in r.stdout and "offsetMesh.counts ok" in r.stdout


# ---- callers on the device: meshToDistanceVolume (a18), mcOffsetMesh (a19) --------------------------------------------
def test_mesh_to_distance_volume_matches_reference_caller():
    pts, tris = meshgen.make_torus(2, 1, 64, 32)
    fwn = mb.FastWindingNumber(pts, tris)
    orc = checker.best_oracle(pts, tris)
    origin, voxel, dims = (-3.3, -3.3, -1.3), (0.21, 0.2, 0.13), (33, 34, 21)
    got, lo, hi = fwn.mesh_to_distance_volume(origin, voxel, dims)
    want, wlo, whi = checker.ref_mesh_to_distance_volume(orc, origin, voxel, dims)
    b = np.asarray(origin, np.float32) + np.float32(0.5) * np.asarray(voxel, np.float32)
    xf = np.array([voxel[0], 0, 0, 0, voxel[1], 0, 0, 0, voxel[2], b[0], b[1], b[2]], np.float32)
    rep = checker.compare_sdf(got, want, orc.calc_from_grid(dims, xf))
    assert rep["ok"], rep
    assert lo == pytest.approx(wlo, rel=1e-5) and hi == pytest.approx(whi, rel=1e-5)
    assert lo < 0 < hi


@pytest.mark.parametrize("offset,close_holes", [(0.15, True), (0.15, False), (-0.1, True)])
def test_mc_offset_mesh_matches_reference_pipeline(offset, close_holes):
    """grid dims/origin identical; vertex and face counts identical wherever the sign fields match (north_star)"""
    pts, tris = meshgen.make_torus(2, 1, 96, 48)
    fwn = mb.FastWindingNumber(pts, tris)
    orc = checker.best_oracle(pts, tris)
    voxel = 0.11
    v, t, info = fwn.mc_offset_mesh(offset, voxel, close_holes=close_holes)
    wv, wt, wsdf, worigin, wdims = checker.ref_mc_offset_mesh(orc, pts, tris, offset, voxel, close_holes=close_holes)
    lo, hi = fwn.bounding_box()
    assert np.array_equal(lo, pts.min(0)) and np.array_equal(hi, pts.max(0))
    assert info["dims"] == wdims and np.array_equal(info["origin"], worigin)
    assert len(t) > 1000 and t.min() >= 0 and t.max() < len(v)
    assert (len(v), len(t)) == (len(wv), len(wt))
    assert np.abs(v - wv).max() < 1e-4  # same vertices up to the 1e-5 relative distance tolerance through the interpolation
    assert np.array_equal(t, wt)


# ---- the reference's own CUDA kernels on this GPU (second baseline, 8d): sanity, not a parity gate ---------------------
def test_reference_cuda_kernels_agree_with_ours():
    from oracle import refcudalib
    if not refcudalib.available():
        pytest.skip("oracle/_ref/libmrrefcuda.so was not built (needs the reference sources)")
    pts, tris = meshgen.make_torus(2, 1, 64, 32)
    orc = checker.best_oracle(pts, tris)
    ref = refcudalib.RefCudaFwn(orc.nodes(), orc.dipoles(), pts, tris)
    dims = (24, 24, 16)
    xf = mb.make_basis_xf((-3.3, -3.3, -1.3), (0.275, 0.275, 0.1625))
    ms, want = ref.signed_distance_layers(dims, xf, 0, dims[2])
    got = mb.FastWindingNumber(pts, tris).calc_from_grid_with_distances(dims, xf)
    assert ms > 0
    # the reference GPU path is float + fast-math (MRCudaMath.cuh:193-247): compare loosely
    assert np.array_equal(np.isnan(got), np.isnan(want))
    ok = ~np.isnan(got)
    assert (np.sign(got[ok]) == np.sign(want[ok])).mean() > 0.999
    assert np.abs(np.abs(got[ok]) - np.abs(want[ok])).max() < 1e-3


def test_general_affine_grid_transform():
    """gridToMeshXf with rotation + shear (the non-axis-aligned kernel variants); same parity rules"""
    pts, tris = meshgen.make_torus(2, 1, 48, 24)
    fwn = mb.FastWindingNumber(pts, tris)
    orc = checker.best_oracle(pts, tris)
    c, s_ = np.cos(0.4), np.sin(0.4)
    A = np.array([[c, -s_, 0.05], [s_, c, 0.0], [0.02, 0.1, 1.0]], np.float32) * np.float32(0.23)
    xf = np.concatenate([A.reshape(9), np.array([-3.1, -3.4, -1.6], np.float32)]).astype(np.float32)
    dims = (29, 30, 14)
    rep = checker.compare_w(fwn.calc_from_grid(dims, xf), orc.calc_from_grid(dims, xf), atol=checker.W_GRID_ATOL)
    assert rep["ok"], rep
    rep = checker.compare_sdf(fwn.calc_from_grid_with_distances(dims, xf), orc.calc_from_grid_with_distances(dims, xf),
                              orc.calc_from_grid(dims, xf))
    assert rep["ok"], rep
    band = dict(min_dist_sq=0.01, max_dist_sq=0.25, null_outside=True)
    got = fwn.calc_from_grid_with_distances(dims, xf, mb.DistOptions(0.01, 0.25, True))
    rep = checker.compare_sdf(got, orc.calc_from_grid_with_distances(dims, xf, **band), orc.calc_from_grid(dims, xf))
    assert rep["ok"], rep


# ---- IPointsToMeshProjector back-end -----------------------------------------------------------------------

@pytest.mark.parametrize("name", ["torus", "noisy", "holes", "deleted"])
def test_find_projections_identical_to_oracle(name):
    """face, closest point, barycentrics and distSq of every query bit for bit (same visiting order as
    findProjectionSubtree, closest point in double), with distance limits and with point / non-rigid tree transforms"""
    pts, tris = _mesh(name)
    orc = checker.best_oracle(pts, tris)
    if getattr(orc, "kind", "") == "reference":
        # the reference's Mesh may start a face's vertex loop at another vertex than the input triple did (getTriangulation);
        # which vertex is `a` changes the last bit of a closest point on an edge, and with it the winner of a tie between
        # the two faces of that edge.  The C++ adapter hands getTriangulation() to b2fwn_create; do the same here.
        pts, tris = orc.points, orc.tris
    fwn = mb.FastWindingNumber(pts, tris)
    q = meshgen.query_points(20000, -2.5, 2.5, seed=11)
    q[::7] = pts[(np.arange(len(q[::7])) * 13) % len(pts)]  # points ON the mesh: ties between the faces around a vertex
    pxf = np.array([0.8, 0.6, 0, -0.6, 0.8, 0, 0, 0, 1, 0.1, -0.2, 0.3], np.float32)
    txf = np.array([1.2, 0.1, 0, 0, 0.9, 0.2, 0.05, 0, 1.1, 0.3, 0.1, -0.2], np.float32)
    for kw in (dict(), dict(up=0.09), dict(lo=0.01), dict(point_xf=pxf), dict(point_xf=pxf, tree_xf=txf), dict(tree_xf=txf, up=0.5, lo=0.001)):
        want = orc.find_projections(q, **kw)
        got = fwn.find_projections(q, kw.get("point_xf"), kw.get("tree_xf"), kw.get("up", mb.FLT_MAX), kw.get("lo", 0.0))
        for f, g in (("face", "face"), ("point", "point"), ("a", "a"), ("b", "b"), ("dist_sq", "dist_sq")):
            assert np.array_equal(got[g].view(np.uint32), want[f].view(np.uint32)), (name, sorted(kw), f)


def test_find_projections_edge_cases():
    pts, tris = _mesh("sphere")
    fwn = mb.FastWindingNumber(pts, tris)
    assert len(fwn.find_projections(np.zeros((0, 3), np.float32))) == 0
    far = fwn.find_projections(np.array([[10, 0, 0]], np.float32), up_dist_limit_sq=1.0)
    assert far["face"][0] == -1 and far["dist_sq"][0] == 1.0 and not far["point"].any()
    empty = mb.FastWindingNumber(pts, np.full((4, 3), -1, np.int32))
    r = empty.find_projections(np.array([[0.5, 0, 0]], np.float32))
    assert r["face"][0] == -1 and r["dist_sq"][0] == mb.FLT_MAX
    # device form
    import torch
    q = torch.from_numpy(meshgen.query_points(5000, -1.5, 1.5, seed=5)).cuda()
    out = torch.zeros(5000 * 8, dtype=torch.float32, device="cuda")
    fwn.find_projections_device(q.data_ptr(), 5000, out.data_ptr())
    host = fwn.find_projections(q.cpu().numpy())
    assert out.cpu().numpy().tobytes() == host.tobytes()


def test_brick_order_does_not_change_results():
    """heavy-first dispatch (walk-length estimates, listed bricks first) is scheduling only: the volume of a grid large enough to
    switch it on (>= 65,536 bricks in one launch) is byte-identical to the same grid computed in z-ranges of 32 layers, each of
    which stays below that size and is therefore walked in plain linear brick order.  (The product library reads no environment
    switch for this -- B2FWN_LINEAR exists in `make DEV=1` builds only.)"""
    import torch
    p, t = meshgen.make_uv_sphere(1, 300, 150)
    p = meshgen.add_noise(p, 2e-3)
    f = mb.FastWindingNumber(p, t)
    try:
        dims = (256, 256, 136)
        layer = dims[0] * dims[1]
        xf = mb.make_basis_xf((-1.1, -1.1, -1.1), (2.2 / 256, 2.2 / 256, 2.2 / 136))
        whole = torch.empty(layer * dims[2], dtype=torch.float32, device="cuda")
        parts = torch.empty_like(whole)
        for o in (mb.DistOptions(0.0, 0.002, False, 0.5, 2.0), mb.DistOptions()):
            whole.fill_(7.0)
            parts.fill_(-7.0)
            f.calc_from_grid_with_distances_device(dims, xf, whole.data_ptr(), o)
            for z0 in range(0, dims[2], 32):
                z1 = min(dims[2], z0 + 32)
                f.calc_from_grid_with_distances_device(dims, xf, parts.data_ptr() + 4 * z0 * layer, o, z_begin=z0, z_end=z1)
            assert whole.cpu().numpy().tobytes() == parts.cpu().numpy().tobytes()
        whole.fill_(7.0)
        parts.fill_(-7.0)
        f.calc_from_grid_device(dims, xf, whole.data_ptr())
        for z0 in range(0, dims[2], 32):
            f.calc_from_grid_device(dims, xf, parts.data_ptr() + 4 * z0 * layer, z_begin=z0, z_end=min(dims[2], z0 + 32))
        assert whole.cpu().numpy().tobytes() == parts.cpu().numpy().tobytes()
    finally:
        f.close()


# ---- makeSignedByWindingNumber on a resident / streamed volume ---------------------------------------------

@pytest.mark.parametrize("name", ["torus", "holes"])
def test_make_signed_by_winding_number_matches_reference_loop(name):
    """unsigned distances + the in-place sign call against the reference's loop (calcFromGrid, negate where w > threshold):
    signs identical outside the |w - threshold| < 1e-3 band, magnitudes untouched bit for bit; and against our own
    calc_from_grid bit for bit (same walk, the numbers just never reach memory)"""
    import torch
    pts, tris = _mesh(name)
    orc = checker.best_oracle(pts, tris)
    if orc.kind == "reference":
        tris = orc.tris
    fwn = mb.FastWindingNumber(pts, tris)
    dims = (41, 35, 27)  # ragged in every direction
    lo, hi = pts.min(0) - 0.3, pts.max(0) + 0.3
    xf = mb.make_basis_xf(lo, (hi - lo) / np.asarray(dims, np.float32))
    unsigned = np.abs(fwn.calc_from_grid_with_distances(dims, xf))
    assert (unsigned >= 0).all()
    for thr, beta in ((0.5, 2.0), (0.3, 2.5), (-0.1, 2.0)):
        want, w_ref = checker.ref_make_signed_by_winding_number(orc, unsigned, dims, xf, thr, beta)
        got = fwn.make_signed_by_winding_number(unsigned.copy(), dims, xf, thr, beta)
        assert np.abs(got).tobytes() == unsigned.tobytes()
        outside = np.abs(w_ref.astype(np.float64) - thr) >= checker.W_BAND
        assert (np.signbit(got) == np.signbit(want))[outside].all()
        assert np.signbit(got).any() and (thr < 0 or not np.signbit(got).all())
        w_own = fwn.calc_from_grid(dims, xf, beta)
        assert got.tobytes() == np.where(w_own > np.float32(thr), -unsigned, unsigned).astype(np.float32).tobytes()
        # device-resident form, in two z-ranges
        d = torch.from_numpy(unsigned.copy()).cuda()
        layer = dims[0] * dims[1]
        fwn.make_signed_by_winding_number_device(d.data_ptr(), dims, xf, thr, beta, z_begin=0, z_end=9)
        fwn.make_signed_by_winding_number_device(d.data_ptr() + 4 * 9 * layer, dims, xf, thr, beta, z_begin=9)
        assert d.cpu().numpy().tobytes() == got.tobytes()
    # unsigned distance + sign == the signed distance of calcFromGridWithDistances (same threshold / beta)
    sdf = fwn.calc_from_grid_with_distances(dims, xf)
    assert fwn.make_signed_by_winding_number(unsigned.copy(), dims, xf).tobytes() == sdf.tobytes()
    fwn.close()


def test_make_signed_by_winding_number_edge_cases():
    import torch
    pts, tris = meshgen.make_torus(2, 1, 48, 24)
    fwn = mb.FastWindingNumber(pts, tris)
    # NaN voxels stay NaN, every other voxel is treated as if they were not there
    dims = (24, 20, 12)
    xf = mb.make_basis_xf((-3.3, -3.3, -1.3), (6.6 / 24, 6.6 / 20, 2.6 / 12))
    vol = np.abs(fwn.calc_from_grid_with_distances(dims, xf))
    holes = vol.copy()
    holes[::7] = np.nan
    got = fwn.make_signed_by_winding_number(holes.copy(), dims, xf)
    full = fwn.make_signed_by_winding_number(vol.copy(), dims, xf)
    assert np.isnan(got[::7]).all()
    keep = np.ones(len(vol), bool)
    keep[::7] = False
    assert got[keep].tobytes() == full[keep].tobytes()
    # rotated + sheared grid transform
    c, s_ = np.cos(0.4), np.sin(0.4)
    A = np.array([[c, -s_, 0.05], [s_, c, 0.0], [0.02, 0.1, 1.0]], np.float32) * np.float32(0.23)
    xf2 = np.concatenate([A.reshape(9), np.array([-3.1, -3.4, -1.6], np.float32)]).astype(np.float32)
    ones = np.ones(29 * 30 * 14, np.float32)
    got = fwn.make_signed_by_winding_number(ones.copy(), (29, 30, 14), xf2)
    assert got.tobytes() == np.where(fwn.calc_from_grid((29, 30, 14), xf2) > np.float32(0.5), -ones, ones).tobytes()
    # empty grids and bad arguments
    fwn.make_signed_by_winding_number(np.zeros(0, np.float32), (0, 5, 5), xf)
    fwn.make_signed_by_winding_number_device(0, (4, 4, 0), xf)
    with pytest.raises(mb.B2Error):
        fwn.make_signed_by_winding_number_device(0, (4, 4, 4), xf)
    with pytest.raises(mb.B2Error):
        fwn.make_signed_by_winding_number_device(0, (4, 4, 4), xf, z_begin=3, z_end=2)
    with pytest.raises(ValueError):
        fwn.make_signed_by_winding_number(np.zeros(5, np.float32), (2, 2, 2), xf)
    # a volume larger than one chunk of the host form (chunks of 32 Mi voxels): H2D, walk, D2H per chunk, two streams
    dims = (512, 512, 160)
    xf = mb.make_basis_xf((-3.3, -3.3, -1.3), (6.6 / 512, 6.6 / 512, 2.6 / 160))
    host = np.full(dims[0] * dims[1] * dims[2], 2.0, np.float32)
    prog = []
    fwn.make_signed_by_winding_number(host, dims, xf, cb=lambda v: prog.append(v) or True)
    assert prog[0] == 0.0 and prog[-1] == 1.0 and len(prog) >= 3
    d = torch.full((host.size,), 2.0, dtype=torch.float32, device="cuda")
    fwn.make_signed_by_winding_number_device(d.data_ptr(), dims, xf)
    assert torch.equal(d.cpu(), torch.from_numpy(host))
    inside = float((host < 0).mean())
    assert abs(inside - 2 * np.pi ** 2 * 2 / (6.6 * 6.6 * 2.6)) < 0.02  # volume fraction of the (48 x 24 faceted) torus R 2, r 1
    fwn.close()


# ---- SignDetectionMode::ProjectionNormal -------------------------------------------------------------------

def _proj_normal_agrees(got, want, dots):
    """NaN masks and magnitudes identical; signs identical wherever the reference's dot( proj - pt, pseudonormal ) is not
    within rounding of zero (vertex pseudonormals are summed in face order here, in ring order there)"""
    assert np.array_equal(np.isnan(got), np.isnan(want))
    ok = ~np.isnan(want)
    ag, aw = np.abs(got[ok]), np.abs(want[ok])
    assert (ag.view(np.uint32) == aw.view(np.uint32)).mean() > 0.999 and (np.abs(ag - aw) <= checker.DIST_REL_TOL * aw).all()
    clear = ok & (np.abs(dots) > 1e-5 * np.abs(want))
    assert clear.sum() > 0.99 * ok.sum()
    assert np.array_equal(np.signbit(got[clear]), np.signbit(want[clear]))
    return ok


@pytest.mark.parametrize("name", ["torus", "holes", "deleted"])
def test_signed_distances_by_projection_normal_match_reference(name):
    import torch
    pts, tris = _mesh(name)
    orc = checker.best_oracle(pts, tris)
    if orc.kind != "reference":
        pytest.skip("needs the reference library (pseudonormals come from its half-edge topology)")
    from oracle import reflib
    rp, rt = orc.points, orc.tris
    fwn = mb.FastWindingNumber(rp, rt)
    q = meshgen.query_points(30000, float(rp.min()) - 0.6, float(rp.max()) + 0.6, seed=11)
    _fn, rvn, _en = orc.pseudonormals(len(rt), len(rp))
    q = np.concatenate([q, rp[::7] + np.float32(0.05) * rvn[::7], rp[::5] - np.float32(0.03) * rvn[::5]]).astype(np.float32)
    for opt, kw in ((mb.DistOptions(0.0, mb.FLT_MAX, False), dict(null_outside=False)),
                    (mb.DistOptions(0.3, 0.1, False), dict(null_outside=False)),  # limits are ignored without the NaN rule
                    (mb.DistOptions(0.01, 0.49, True), dict(min_dist_sq=0.01, max_dist_sq=0.49, null_outside=True))):
        want, dots = orc.signed_distance_proj_normal(q, **kw)
        got = fwn.signed_distances_by_projection_normal(q, opt)
        ok = _proj_normal_agrees(got, want, dots)
        assert np.signbit(got[ok]).any() and not np.signbit(got[ok]).all()
        d_q = torch.from_numpy(q).cuda()
        d_out = torch.empty(len(q), dtype=torch.float32, device="cuda")
        fwn.signed_distances_by_projection_normal_device(d_q.data_ptr(), len(q), d_out.data_ptr(), opt)
        assert d_out.cpu().numpy().tobytes() == got.tobytes()
    # the volume form: voxel centres as meshToDistanceFunctionVolume computes them
    origin = (rp.min(0) - 0.4).astype(np.float32)
    dims = (37, 41, 23)
    voxel = ((rp.max(0) + 0.4 - origin) / np.asarray(dims, np.float32)).astype(np.float32)
    centres = reflib.voxel_centres(origin, voxel, dims)
    for opt, kw in ((mb.DistOptions(0.0, mb.FLT_MAX, False), dict(null_outside=False)),
                    (mb.DistOptions(0.0, 0.09, True), dict(min_dist_sq=0.0, max_dist_sq=0.09, null_outside=True)),
                    (mb.DistOptions(0.0004, 0.09, True), dict(min_dist_sq=0.0004, max_dist_sq=0.09, null_outside=True))):
        want, dots = orc.signed_distance_proj_normal(centres, **kw)
        prog = []
        got, lo, hi = fwn.mesh_to_distance_volume_by_projection_normal(origin, voxel, dims, opt, cb=lambda v: prog.append(v) or True)
        ok = _proj_normal_agrees(got, want, dots)
        assert lo == got[ok].min() and hi == got[ok].max() and lo < 0 < hi
        assert prog[0] == 0.0 and prog[-1] == 1.0
    fwn.close()


def test_projection_normal_edge_cases():
    pts, tris = meshgen.make_torus(2, 1, 48, 24)
    fwn = mb.FastWindingNumber(pts, tris)
    assert len(fwn.signed_distances_by_projection_normal(np.zeros((0, 3), np.float32))) == 0
    got, lo, hi = fwn.mesh_to_distance_volume_by_projection_normal((0, 0, 0), 0.1, (0, 4, 4))
    assert len(got) == 0 and lo == mb.FLT_MAX and hi == -mb.FLT_MAX
    # closed surface: the sign by the pseudonormal agrees with the winding number away from the surface
    q = meshgen.query_points(20000, -3.4, 3.4, seed=5)
    d = fwn.signed_distances_by_projection_normal(q, mb.DistOptions(0.0, mb.FLT_MAX, False))
    w = fwn.calc_from_vector(q)
    far = np.abs(d) > 1e-3
    assert np.array_equal(d[far] < 0, w[far] > 0.5)
    # more than one batch of 4 Mi points through the host form, against the device form in one launch
    import torch
    big = meshgen.query_points(5_000_000, -3.4, 3.4, seed=6)
    opt = mb.DistOptions(0.0, 1.0, True)
    host = fwn.signed_distances_by_projection_normal(big, opt)
    d_q = torch.from_numpy(big).cuda()
    d_out = torch.empty(len(big), dtype=torch.float32, device="cuda")
    fwn.signed_distances_by_projection_normal_device(d_q.data_ptr(), len(big), d_out.data_ptr(), opt)
    assert d_out.cpu().numpy().tobytes() == host.tobytes()
    assert np.isnan(host).any() and not np.isnan(host).all()
    # a volume of more than one batch of layers (8 Mi voxels per batch) against the point form on the same voxel centres
    from oracle import normalslib
    if normalslib.available():
        origin, voxel, dims = (-3.3, -3.3, -1.3), (6.6 / 256, 6.6 / 256, 2.6 / 160), (256, 256, 160)
        centres = normalslib.voxel_centres(origin, voxel, dims)
        # lower limit > 0: the volume call runs the per-point search on generated voxel centres -- same records, same bytes
        band = mb.DistOptions(1e-4, 0.01, True)
        vol, lo, hi = fwn.mesh_to_distance_volume_by_projection_normal(origin, voxel, dims, band)
        assert vol.tobytes() == fwn.signed_distances_by_projection_normal(centres, band).tobytes()
        assert lo == np.nanmin(vol) and hi == np.nanmax(vol) and -0.1 <= lo < 0 < hi <= 0.1
        # no lower limit: the brick walk finds the distances (bit-identical) and keeps the closest triangle; where two
        # triangles tie it may keep the other one, which can only matter where the sign decision is itself a rounding tie
        band = mb.DistOptions(0.0, 0.01, True)
        vol, lo, hi = fwn.mesh_to_distance_volume_by_projection_normal(origin, voxel, dims, band)
        ref_pts = fwn.signed_distances_by_projection_normal(centres, band)
        assert np.array_equal(np.isnan(vol), np.isnan(ref_pts))
        ok = ~np.isnan(vol)
        assert ok.sum() > 100000 and (np.abs(vol[ok]).view(np.uint32) == np.abs(ref_pts[ok]).view(np.uint32)).mean() > 0.999
        assert np.allclose(np.abs(vol[ok]), np.abs(ref_pts[ok]), rtol=1e-5, atol=0)
        assert (np.signbit(vol[ok]) != np.signbit(ref_pts[ok])).mean() < 1e-4
        assert lo == np.nanmin(vol) and hi == np.nanmax(vol) and -0.1 <= lo < 0 < hi <= 0.1
    fwn.close()
    # a mesh without faces has no projection anywhere
    empty = mb.FastWindingNumber(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.int32))
    assert np.isnan(empty.signed_distances_by_projection_normal(q[:100])).all()
    empty.close()


# ---- moved vertices, refit instead of rebuild -------------------------------------------------------------

def test_update_points_matches_reference_refit():
    """b2fwn_update_points against Mesh::updateCaches (AABBTree::refit on the cached tree + calcDipoles): identical nodes and
    dipoles, and every query on the refitted tree under the same parity rules as after a build"""
    pts, tris = _mesh("noisy")
    orc = checker.best_oracle(pts, tris)
    if orc.kind != "reference":
        pytest.skip("needs the reference library (the C restatement has no refit)")
    old, rt = orc.points, orc.tris
    fwn = mb.FastWindingNumber(old, rt)
    nodes0, dip0 = fwn.get_tree()
    new = old.copy()
    new[:, 0] = np.float32(1.15) * old[:, 0] + np.float32(0.08) * np.sin(np.float32(3) * old[:, 1])
    new[:, 2] = old[:, 2] + np.float32(0.1) * np.cos(np.float32(2) * old[:, 0])
    new = new.astype(np.float32)
    orc.update_points(new)
    fwn.update_points(new)
    nodes, dip = fwn.get_tree()
    assert nodes.tobytes() == orc.nodes().tobytes() and dip.tobytes() == orc.dipoles().tobytes()
    assert np.array_equal(nodes["l"], nodes0["l"]) and np.array_equal(nodes["r"], nodes0["r"]) and nodes.tobytes() != nodes0.tobytes()
    q = meshgen.query_points(30000, float(new.min()) - 0.5, float(new.max()) + 0.5, seed=17)
    rep = checker.compare_w(fwn.calc_from_vector(q), orc.calc_from_vector(q), atol=2e-6)
    assert rep["ok"], rep
    dims = (40, 33, 26)
    lo, hi = new.min(0) - 0.3, new.max(0) + 0.3
    xf = mb.make_basis_xf(lo, (hi - lo) / np.asarray(dims, np.float32))
    w = orc.calc_from_grid(dims, xf)
    got = fwn.calc_from_grid_with_distances(dims, xf)
    rep = checker.compare_sdf(got, orc.calc_from_grid_with_distances(dims, xf), w)
    assert rep["ok"] and rep["dist_bit_exact_frac"] > 0.999, rep
    a, b = fwn.find_projections(q[:8000]), orc.find_projections(q[:8000])
    for f_ in ("face", "point", "a", "b", "dist_sq"):
        assert np.array_equal(a[f_].view(np.uint32), b[f_].view(np.uint32)), f_
    want, dots = orc.signed_distance_proj_normal(q[:8000], null_outside=False)
    _proj_normal_agrees(fwn.signed_distances_by_projection_normal(q[:8000], mb.DistOptions(0.0, mb.FLT_MAX, False)), want, dots)
    # a fresh build at the new coordinates has another tree but the same distances
    fresh = mb.FastWindingNumber(new, rt)
    assert fresh.get_tree()[0].tobytes() != nodes.tobytes()
    assert np.abs(fresh.calc_from_grid_with_distances(dims, xf)).tobytes() == np.abs(got).tobytes()
    fresh.close()
    # moving the vertices back gives the tree of the build again
    fwn.update_points(old)
    nodes2, dip2 = fwn.get_tree()
    assert nodes2.tobytes() == nodes0.tobytes() and dip2.tobytes() == dip0.tobytes()
    with pytest.raises(mb.B2Error):
        fwn.update_points(old[:-1])
    fwn.close()
