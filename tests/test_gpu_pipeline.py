"""GPU tests of the host side of the grid calls (meshlib_b200/csrc/b2fwn_pipeline.cu): pageable / pinned / device outputs,
caller-owned by-parts buffers, several devices of one process, cancellation, the memory budget.  The numbers themselves are
pinned against the oracle in test_gpu_parity.py; here every variant must deliver the SAME bits as the device-resident call."""
import os

import numpy as np
import pytest

import meshlib_b200 as mb
from meshlib_b200 import meshgen
from oracle import checker

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torus():
    pts, tris = meshgen.make_torus(2, 1, 96, 48)
    fwn = mb.FastWindingNumber(pts, tris)
    yield pts, tris, fwn
    fwn.close()


@pytest.fixture()
def small_chunks(monkeypatch):
    # test switch of the library: chunks of ~20k voxels instead of >= 32 Mi, so that small grids run many chunks
    monkeypatch.setenv("B2FWN_MIN_CHUNK_VOXELS", "20000")


def _device_volume(fwn, dims, xf, opts):
    import torch
    n = dims[0] * dims[1] * dims[2]
    d = torch.empty(n, dtype=torch.float32, device="cuda")
    fwn.calc_from_grid_with_distances_device(dims, xf, d.data_ptr(), opts)
    return d.cpu().numpy()


@pytest.mark.parametrize("band", [False, True])
def test_pageable_pinned_and_device_outputs_are_identical(torus, small_chunks, band):
    import torch
    _p, _t, fwn = torus
    dims = (64, 60, 44)  # x % 4 == 0: vector stores; 168,960 voxels -> 9 chunks, several staging pieces per chunk
    xf = mb.make_basis_xf((-3.3, -3.3, -1.3), (0.103, 0.11, 0.06))
    opts = mb.DistOptions(0.0, 0.09, False) if band else mb.DistOptions()
    want = _device_volume(fwn, dims, xf, opts)
    pageable = np.full(want.size, np.float32(7e33))
    fwn.calc_from_grid_with_distances(dims, xf, opts, out=pageable)
    assert pageable.tobytes() == want.tobytes()
    pinned = torch.full((want.size,), 7e33, dtype=torch.float32).pin_memory()
    fwn.calc_from_grid_with_distances(dims, xf, opts, out=pinned.numpy())
    assert pinned.numpy().tobytes() == want.tobytes()
    # a z-range into the middle of a pageable buffer, and winding numbers alone
    part = np.empty(dims[0] * dims[1] * 17, np.float32)
    fwn.calc_from_grid_with_distances_layers(dims, xf, 9, 26, opts, out=part)
    layer = dims[0] * dims[1]
    assert part.tobytes() == want[9 * layer:26 * layer].tobytes()
    import torch as _t2
    dw = _t2.empty(want.size, dtype=_t2.float32, device="cuda")
    fwn.calc_from_grid_device(dims, xf, dw.data_ptr())
    assert fwn.calc_from_grid(dims, xf).tobytes() == dw.cpu().numpy().tobytes()


def test_ragged_grid_through_the_staging_ring(torus, small_chunks):
    _p, _t, fwn = torus
    dims = (37, 29, 23)
    xf = mb.make_basis_xf((-3.2, -3.2, -1.2), (0.17, 0.22, 0.1))
    want = _device_volume(fwn, dims, xf, mb.DistOptions())
    got = fwn.calc_from_grid_with_distances(dims, xf)
    assert got.tobytes() == want.tobytes()


def test_by_parts_into_caller_owned_buffers(torus, small_chunks):
    _p, _t, fwn = torus
    dims = (40, 36, 31)
    xf = mb.make_basis_xf((-3.2, -3.2, -1.3), (0.16, 0.18, 0.085))
    full = _device_volume(fwn, dims, xf, mb.DistOptions())
    layer = dims[0] * dims[1]
    for overlap, per_part in ((0, 0), (1, 4), (2, 7), (1, 31)):
        seen, delivered, vol = [], [], np.full(full.size, np.float32(7e33))

        def part(arr, d, z0):
            assert arr.size == layer * d[2] and arr.flags.owndata  # the consumer owns the memory it receives
            seen.append(z0)
            delivered.append((z0, d[2]))
            vol[z0 * layer:(z0 + d[2]) * layer] = arr
            return True

        plan, per_part_used = fwn.plan_parts(dims, overlap, per_part)  # b2fwn_plan_parts_for: what the call is about to do
        fwn.calc_from_grid_with_distances_by_parts_into(part, dims, xf, layer_overlap=overlap, max_layers_per_part=per_part)
        assert seen == sorted(seen) and len(set(seen)) == len(seen) and seen[0] == 0
        assert [(b, n) for b, n, _d in plan] == delivered, (overlap, per_part, plan, delivered)
        assert per_part_used == delivered[0][1] and (per_part <= 0 or per_part_used == min(per_part, dims[2]))
        assert vol.tobytes() == full.tobytes()
    # winding numbers alone, consumer failure
    w = fwn.calc_from_grid(dims, xf)
    got = np.empty_like(w)

    def part_w(arr, d, z0):
        got[z0 * layer:(z0 + d[2]) * layer] = arr

    fwn.calc_from_grid_with_distances_by_parts_into(part_w, dims, xf, layer_overlap=1, max_layers_per_part=5, distances=False)
    assert got.tobytes() == w.tobytes()
    with pytest.raises(mb.B2Error):
        fwn.calc_from_grid_with_distances_by_parts_into(lambda *a: False, dims, xf, max_layers_per_part=4)
    # the handle is still usable after an aborted call
    assert fwn.calc_from_grid_with_distances(dims, xf).tobytes() == full.tobytes()


def test_cancellation_midway_and_reuse(torus, small_chunks):
    _p, _t, fwn = torus
    dims = (48, 48, 40)
    xf = mb.make_basis_xf((-3.2, -3.2, -1.3), (0.135, 0.135, 0.066))
    seen = []
    with pytest.raises(mb.B2Error) as e:
        fwn.calc_from_grid_with_distances(dims, xf, cb=lambda v: seen.append(v) or v < 0.3)
    assert str(e.value) == "Operation was canceled" and e.value.code == 2
    assert seen[0] == 0.0 and 0.3 <= seen[-1] < 1.0
    want = _device_volume(fwn, dims, xf, mb.DistOptions())
    assert fwn.calc_from_grid_with_distances(dims, xf).tobytes() == want.tobytes()


def test_min_dist_with_null_outside_false_option_shape(torus, small_chunks):
    """mcOffsetMesh( closeHoles ) asks for minDistSq > 0 with nullOutsideMinMax = false (MROffset.cpp:176-183).  The reference stops
    a voxel's search as soon as a candidate is within minDist (MRMeshProject.cpp:60-63), so what such a voxel reports depends on the
    visiting order: the contract is |d| <= minDist on both sides there, and bit-identical distances everywhere else -- in the
    full-grid call, in a z-slab and in the chunked host call."""
    pts, tris, fwn = torus
    orc = checker.best_oracle(pts, tris)
    dims = (48, 44, 36)
    xf = mb.make_basis_xf((-3.3, -3.3, -1.3), (0.1375, 0.15, 0.0722))
    min_d, max_d = 0.22, 0.5
    o = dict(min_dist_sq=min_d * min_d, max_dist_sq=max_d * max_d, null_outside=False)
    want = orc.calc_from_grid_with_distances(dims, xf, **o)
    opts = mb.DistOptions(o["min_dist_sq"], o["max_dist_sq"], False)
    full = _device_volume(fwn, dims, xf, opts)
    host = fwn.calc_from_grid_with_distances(dims, xf, opts)
    layer = dims[0] * dims[1]
    slab = fwn.calc_from_grid_with_distances_layers(dims, xf, 10, 22, opts)
    lim = np.float32(np.sqrt(np.float32(o["min_dist_sq"]))) * np.float32(1 + 1e-6)
    for got in (full, host):
        near = np.abs(want) <= lim
        assert near.sum() > 1000 and (~near).sum() > 1000
        assert (np.abs(got[near]) <= lim).all()
        assert np.array_equal(np.abs(got[~near]).view(np.uint32), np.abs(want[~near]).view(np.uint32))
    near = np.abs(want[10 * layer:22 * layer]) <= lim
    assert np.array_equal(slab[~near].view(np.uint32), full[10 * layer:22 * layer][~near].view(np.uint32))
    assert (np.abs(slab[near]) <= lim).all()
    w = orc.calc_from_grid(dims, xf)
    rep = checker.compare_sdf(np.where(np.abs(want) <= lim, want, full), want, w)
    assert rep["sign_mismatch"] == 0 and rep["nan_mismatch"] == 0, rep
    signs_near = np.signbit(full) == np.signbit(want)
    assert signs_near[(np.abs(w - 0.5) >= 1e-3)].all()


def test_safe_memory_limit_and_budgeted_parts(torus):
    _p, _t, fwn = torus
    lim = mb.load_library().b2fwn_safe_memory_limit(0)
    import torch
    free, total = torch.cuda.mem_get_info(0)
    assert 0 < lim < free and lim > 0.5 * free  # 80 % of (free - 128 MB), MRCudaBasic.cpp:75-90
    # a caller asking for parts of "as many layers as possible" gets whole layers within the budget
    dims = (32, 32, 24)
    xf = mb.make_basis_xf((-3.2, -3.2, -1.3), (0.2, 0.2, 0.11))
    parts = []
    fwn.calc_from_grid_with_distances_by_parts(lambda a, d, z: parts.append((z, d[2])) or True, dims, xf, max_layers_per_part=1 << 20)
    assert parts == [(0, 24)]


def test_several_devices_of_one_process(small_chunks):
    L = mb.load_library()
    if L.b2fwn_device_count() < 2:
        pytest.skip("one CUDA device")
    pts, tris = meshgen.make_torus(2, 1, 96, 48)
    n = min(L.b2fwn_device_count(), 4)
    one = mb.FastWindingNumber(pts, tris)
    many = mb.FastWindingNumber(pts, tris, devices=list(range(n)))
    dims = (64, 60, 70)
    xf = mb.make_basis_xf((-3.3, -3.3, -1.3), (0.103, 0.11, 0.038))
    for opts in (mb.DistOptions(), mb.DistOptions(0.0, 0.09, False)):
        want = one.calc_from_grid_with_distances(dims, xf, opts)
        assert many.calc_from_grid_with_distances(dims, xf, opts).tobytes() == want.tobytes()
    layer = dims[0] * dims[1]
    vol = np.empty(layer * dims[2], np.float32)
    seen = []

    def part(arr, d, z0):
        seen.append(z0)
        vol[z0 * layer:(z0 + d[2]) * layer] = arr

    many.calc_from_grid_with_distances_by_parts_into(part, dims, xf, layer_overlap=1, max_layers_per_part=6)
    assert seen == sorted(seen) and vol.tobytes() == one.calc_from_grid_with_distances(dims, xf).tobytes()
    assert many.calc_from_grid(dims, xf).tobytes() == one.calc_from_grid(dims, xf).tobytes()
    got, lo, hi = many.mesh_to_distance_volume((-3.3, -3.3, -1.3), (0.103, 0.11, 0.038), dims)
    w1, lo1, hi1 = one.mesh_to_distance_volume((-3.3, -3.3, -1.3), (0.103, 0.11, 0.038), dims)
    assert got.tobytes() == w1.tobytes() and (lo, hi) == (lo1, hi1)
    # moved vertices reach every replica
    moved = (pts * np.float32(1.05)).astype(np.float32)
    one.update_points(moved)
    many.update_points(moved)
    assert many.calc_from_grid_with_distances(dims, xf).tobytes() == one.calc_from_grid_with_distances(dims, xf).tobytes()
    assert many.stats().kernel_launches > one.stats().kernel_launches / 2
    one.close()
    many.close()


# ---- half-edge tables on the device (first slice) -----------------------------------------------------------
def _check_tables(tris, n_verts, he, epv, epf):
    """structure checks that need no reference, then the reference's own MeshTopology::read + checkValidity where it is present"""
    tris = np.asarray(tris, np.int32)
    assert he.shape[0] % 2 == 0 and epf.shape == (len(tris),) and epv.shape == (n_verts,)
    nxt, prv, org, left = he[:, 0], he[:, 1], he[:, 2], he[:, 3]
    e = np.arange(len(he))
    assert (nxt >= 0).all() and (prv >= 0).all() and np.array_equal(prv[nxt], e) and np.array_equal(nxt[prv], e)
    assert np.array_equal(org[nxt], org)  # next stays in the ring of the origin
    # the three half-edges of every face: e0, prev(sym(e0)), prev(sym(e1))
    e0 = epf
    e1 = prv[e0 ^ 1]
    e2 = prv[e1 ^ 1]
    assert np.array_equal(prv[e2 ^ 1], e0)
    f = np.arange(len(tris))
    assert np.array_equal(left[e0], f) and np.array_equal(left[e1], f) and np.array_equal(left[e2], f)
    assert np.array_equal(np.stack([org[e0], org[e1], org[e2]], 1), tris)
    used = np.zeros(n_verts, bool)
    used[tris.ravel()] = True
    assert np.array_equal(epv >= 0, used) and np.array_equal(org[epv[used]], np.flatnonzero(used))
    from oracle import reflib
    if reflib.available():
        got = reflib.topology_from_tables(he, epv, epf)
        assert got is not None, "the reference's MeshTopology::read rejected the tables"
        rtris, n_ue, n_loops, n_vv = got
        assert np.array_equal(rtris, tris) and n_ue == len(he) // 2 and n_vv == int(used.sum())
        vf, ue, bl, vv, _s = reflib.builder_stats(tris)
        assert (vf, ue, bl, vv) == (len(tris), n_ue, n_loops, n_vv)  # the reference's own builder agrees on every count
        return n_loops
    return None


def test_half_edge_tables_of_closed_and_open_meshes(torus):
    pts, tris, fwn = torus
    he, epv, epf = fwn.build_topology(tris, len(pts))
    assert len(he) == 3 * len(tris)  # closed manifold: E = 3F / 2
    loops = _check_tables(tris, len(pts), he, epv, epf)
    assert loops in (None, 0)
    # an open mesh: a sphere with two runs of faces removed has two boundary loops; unused vertices stay invalid
    p2, t2 = meshgen.make_uv_sphere(1, 48, 24)
    t2 = np.delete(t2, np.r_[300:340, 1500:1530], axis=0)
    he, epv, epf = fwn.build_topology(t2, len(p2) + 5)
    loops = _check_tables(t2, len(p2) + 5, he, epv, epf)
    assert loops in (None, 2)
    assert (epv[len(p2):] == -1).all()
    # holes that touch in a vertex leave it with two fans and two gaps: which gap follows which fan is geometry, not topology --
    # refused like every other vertex that is not a disc or a half-disc
    with pytest.raises(mb.B2Error):
        fwn.build_topology(meshgen.make_uv_sphere(1, 48, 24)[1][np.arange(2304) % 11 != 3], len(p2))
    # non-manifold and degenerate input is refused, not mis-built: the caller keeps the CPU builder for it
    bad = np.concatenate([tris, tris[:1]])  # the same face twice: an edge with two faces on one side
    with pytest.raises(mb.B2Error) as e:
        fwn.build_topology(bad, len(pts))
    assert "manifold" in str(e.value)
    flipped = tris.copy()
    flipped[7] = flipped[7][::-1]  # inconsistent orientation
    with pytest.raises(mb.B2Error):
        fwn.build_topology(flipped, len(pts))
    with pytest.raises(mb.B2Error):
        fwn.build_topology(np.array([[0, 0, 1]], np.int32), 3)
    two_fans = np.array([[0, 1, 2], [0, 2, 3], [0, 3, 1], [1, 3, 2], [0, 5, 4], [0, 6, 5], [0, 4, 6], [4, 5, 6]], np.int32)  # two tetrahedra sharing vertex 0
    with pytest.raises(mb.B2Error) as e:
        fwn.build_topology(two_fans, 7)
    assert "fan" in str(e.value)
    assert fwn.build_topology(np.zeros((0, 3), np.int32), 4)[0].shape == (0, 4)


def test_offset_mesh_to_half_edges_without_leaving_the_device(torus):
    """mcOffsetMesh's last step (Mesh::fromTriMesh, MROffset.cpp:222-225): marching cubes result -> MeshTopology tables, both in HBM"""
    _p, _t, fwn = torus
    v, t, info = fwn.mc_offset_mesh(0.15, 0.11)
    he, epv, epf = fwn.build_topology()
    assert len(epf) == len(t) and len(epv) == len(v)
    _check_tables(t, len(v), he, epv, epf)


def test_suggest_voxel_size_and_malformed_triangulation(torus):
    pts, tris, fwn = torus
    f = np.float32
    ext = (pts.max(0) - pts.min(0)).astype(f)
    vol = f(f(f(f(1) * ext[0]) * ext[1]) * ext[2])  # Box::volume, MRBox.h:112-119
    want = np.cbrt(f(vol / f(5e6)))                  # suggestVoxelSize, MROffset.cpp:25-31
    assert abs(fwn.suggest_voxel_size(5e6) - float(want)) <= 1e-7 * float(want)
    # a vertex id beyond the vertex array is an error of the build, not an out-of-bounds read on the device
    bad = tris.copy()
    bad[11, 1] = len(pts) + 3
    with pytest.raises(mb.B2Error) as e:
        mb.FastWindingNumber(pts, bad)
    assert "vertex ids beyond" in str(e.value)
    # the context survives: the next build works
    ok = mb.FastWindingNumber(pts, tris)
    assert ok.stats().num_leaves == len(tris)
    ok.close()
    # device forms refuse a null output
    with pytest.raises(mb.B2Error):
        fwn.calc_from_grid_with_distances_device((8, 8, 8), mb.make_basis_xf((-3, -3, -1), 0.5), 0)


# ---- marching cubes by slabs and over several devices (the reference's MarchingCubesByParts, in HBM) ------------------------
@pytest.mark.parametrize("close_holes", [True, False])
def test_offset_mesh_in_slabs_is_identical_to_one_shot(torus, close_holes):
    """slabs of 1, 2, 3, 5 ... owner layers (each with the two layers after it) must give the one-shot TriMesh bit for bit:
    vertices, their order, triangles, their order -- with NaN outside the band (close_holes False) and without"""
    _p, _t, fwn = torus
    want_v, want_t, info = fwn.mc_offset_mesh(0.15, 0.11, close_holes=close_holes)
    assert len(want_t) > 1000
    for n in (1, 2, 3, 5, 16, info["dims"][2] - 1, info["dims"][2] + 7):
        v, t, i2 = fwn.mc_offset_mesh(0.15, 0.11, close_holes=close_holes, max_layers_per_slab=n)
        assert i2["dims"] == info["dims"] and (i2["n_verts"], i2["n_tris"]) == (len(want_v), len(want_t)), n
        assert v.tobytes() == want_v.tobytes() and t.tobytes() == want_t.tobytes(), n
    # the half-edge tables of a slabbed result
    he, epv, epf = fwn.build_topology()
    _check_tables(want_t, len(want_v), he, epv, epf)


def test_offset_mesh_over_several_devices():
    L = mb.load_library()
    if L.b2fwn_device_count() < 2:
        pytest.skip("one CUDA device")
    pts, tris = meshgen.make_torus(2, 1, 96, 48)
    n = min(L.b2fwn_device_count(), 4)
    one = mb.FastWindingNumber(pts, tris)
    many = mb.FastWindingNumber(pts, tris, devices=list(range(n)))
    for kw in (dict(), dict(max_layers_per_slab=3), dict(close_holes=False)):
        want_v, want_t, info = one.mc_offset_mesh(-0.1, 0.09, **kw)
        v, t, i2 = many.mc_offset_mesh(-0.1, 0.09, **kw)
        assert i2["dims"] == info["dims"]
        assert v.tobytes() == want_v.tobytes() and t.tobytes() == want_t.tobytes(), kw
    he, epv, epf = many.build_topology()  # triangles gathered from every device
    _check_tables(want_t, len(want_v), he, epv, epf)
    # a plain marching-cubes call afterwards forgets the multi-device result
    vol = np.linspace(-1, 1, 8 * 8 * 8, dtype=np.float32)
    v1, t1 = many.marching_cubes(vol, (8, 8, 8), (0, 0, 0), 1.0, 0.0)
    v2, t2 = one.marching_cubes(vol, (8, 8, 8), (0, 0, 0), 1.0, 0.0)
    assert v1.tobytes() == v2.tobytes() and t1.tobytes() == t2.tobytes()
    one.close()
    many.close()


def test_trim_and_reclaim_of_the_handle_caches_when_memory_runs_out():
    """the caches of a handle (scratch slots, mesher workspace, pool) are given back by b2fwn_trim, and by the handle's own
    allocation sites when cudaMalloc reports out-of-memory -- with the device filled up to 96 MB by somebody else, calls that need
    more than that go through, and give the same results as with an empty device"""
    import torch
    pts, tris = meshgen.make_uv_sphere(1, 96, 48)
    dims = (512, 512, 512)
    vs = 2.6 / 512
    xf = mb.make_basis_xf((-1.3, -1.3, -1.3), vs)
    opt = mb.DistOptions(0.0, (3 * vs) ** 2, False)
    ref = mb.FastWindingNumber(pts, tris)
    vol_ref = ref.calc_from_grid_with_distances(dims, xf, opt)
    ref.close()
    torch.zeros(1, device="cuda")
    MB = 1 << 20

    def free_bytes():
        torch.cuda.synchronize()
        torch.cuda.empty_cache()
        return torch.cuda.mem_get_info()[0]

    def fill_device_up_to(leave):
        return torch.empty(free_bytes() - leave, dtype=torch.uint8, device="cuda")

    fwn = mb.FastWindingNumber(pts, tris)
    try:
        q = meshgen.query_points(5000, -1.5, 1.5, seed=3)
        w0 = fwn.calc_from_vector(q)
        free_start = free_bytes()
        v0, t0, _info = fwn.mc_offset_mesh(0.05, 0.0035)
        topo0 = fwn.build_topology()
        held = free_start - free_bytes()
        assert held > 1000 * MB, held  # the band of ~606^3 voxels and the mesher's workspace stay cached
        assert len(t0) * 3 * 32 > 128 * MB  # the tables' sort needs more than what is left free below

        # (1) the half-edge build cannot allocate its sort buffers: every cache of the handle goes, second attempt succeeds
        hog = fill_device_up_to(96 * MB)
        topo1 = fwn.build_topology()
        for a, b in zip(topo0, topo1):
            assert a.tobytes() == b.tobytes()
        del hog

        # (2) the second scratch slot of a host-output call cannot be allocated: the mesher's workspace goes
        v1, t1, _info = fwn.mc_offset_mesh(0.05, 0.0035)  # caches grown again
        assert v1.tobytes() == v0.tobytes() and t1.tobytes() == t0.tobytes()
        hog = fill_device_up_to(96 * MB)
        vol = fwn.calc_from_grid_with_distances(dims, xf, opt)
        # with the device this full, a wish of 512 layers of 1024^2 per part (2 GiB, twice) is cut to what the handle can count on; the
        # plan says so before the call would
        plan, per_part = fwn.plan_parts((1024, 1024, 1024), 1, 512)
        assert 2 <= per_part < 512 and plan[0][:2] == (0, per_part) and plan[-1][0] + plan[-1][1] == 1024
        assert all(b1 == b0 + per_part - 1 for (b0, _n0, _d0), (b1, _n1, _d1) in zip(plan, plan[1:]))
        del hog
        assert vol.tobytes() == vol_ref.tobytes()

        # (3) b2fwn_trim: the device is as free as before the calls, up to the results the handle still holds (mesh + tables)
        before_trim = free_bytes()
        fwn.trim()
        after_trim = free_bytes()
        assert after_trim - before_trim > 800 * MB, (before_trim, after_trim)
        assert free_start - after_trim < 512 * MB, (free_start, after_trim)
        # and the handle works as before
        assert fwn.calc_from_vector(q).tobytes() == w0.tobytes()
        v2, t2, _info = fwn.mc_offset_mesh(0.05, 0.0035)
        assert v2.tobytes() == v0.tobytes() and t2.tobytes() == t0.tobytes()
    finally:
        fwn.close()
