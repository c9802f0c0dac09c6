"""CPU test of the half-edge construction behind b2fwn_build_topology (meshlib_b200/csrc/b2fwn_topology.cu): the same steps in numpy
-- one key per face corner, sort, runs = undirected edges, next[e_k+1] = sym(e_k), prev[sym(e_k)] = e_k+1, boundary half-edges
chained around their origin -- produce tables the reference's own MeshTopology::read accepts (it runs checkValidity) and from which
it reads back the input triangles; every count equals MeshBuilder::fromTriangles.  The CUDA kernels are held against the same
checks on the GPU box (tests/test_gpu_pipeline.py)."""
import numpy as np
import pytest

from meshlib_b200 import meshgen
from oracle import reflib

pytestmark = pytest.mark.skipif(not reflib.available(), reason="needs oracle/_ref/libmrref.so")


def build_tables(tris, n_verts):
    tris = np.asarray(tris, np.int64)
    F = len(tris)
    a = tris.reshape(-1)
    b = np.roll(tris, -1, axis=1).reshape(-1)
    key = (np.minimum(a, b) << 32) | np.maximum(a, b)
    order = np.argsort(key, kind="stable")
    ks = key[order]
    head = np.ones(len(ks), bool)
    head[1:] = ks[1:] != ks[:-1]
    rank = np.cumsum(head) - 1
    he_sorted = 2 * rank + (a[order] > b[order])
    half_of_corner = np.empty(3 * F, np.int64)
    half_of_corner[order] = he_sorted
    n_half = 2 * (rank[-1] + 1)
    nxt, prv, org, left = (np.full(n_half, -1, np.int64) for _ in range(4))
    org[half_of_corner] = a
    left[half_of_corner] = np.repeat(np.arange(F), 3)
    single = np.bincount(rank)[rank] == 1          # boundary edges: one corner only
    hb = he_sorted[single] ^ 1                     # their half-edges without a face
    org[hb] = b[order][single]
    e = half_of_corner.reshape(F, 3)
    for k in range(3):
        ek, en = e[:, k], e[:, (k + 1) % 3]
        nxt[en] = ek ^ 1
        prv[ek ^ 1] = en
    in_at = np.full(n_verts, -1, np.int64)
    in_at[org[hb ^ 1]] = hb                        # boundary half-edge that ends at the vertex
    hin = in_at[org[hb]]
    nxt[hb] = hin ^ 1
    prv[hin ^ 1] = hb
    epv = np.full(n_verts, 2 ** 31 - 1, np.int64)
    np.minimum.at(epv, org, np.arange(n_half))
    epv[epv == 2 ** 31 - 1] = -1
    return np.stack([nxt, prv, org, left], 1).astype(np.int32), epv.astype(np.int32), e[:, 0].astype(np.int32)


@pytest.mark.parametrize("name", ["torus", "sphere_with_two_holes", "offset_surface"])
def test_tables_load_into_the_reference_topology(name):
    if name == "torus":
        pts, tris = meshgen.make_torus(2, 1, 96, 48)
        n_verts, loops = len(pts), 0
    elif name == "sphere_with_two_holes":
        pts, tris = meshgen.make_uv_sphere(1, 48, 24)
        tris = np.delete(tris, np.r_[300:340, 1500:1530], axis=0)
        n_verts, loops = len(pts) + 3, 2           # three vertices no face uses stay invalid
    else:
        # a marching-cubes result of the reference itself: sphere of radius 0.7 in a 24^3 volume
        g = (np.arange(24, dtype=np.float32) + 0.5) / 12 - 1
        z, y, x = np.meshgrid(g, g, g, indexing="ij")
        vol = (np.sqrt(x * x + y * y + z * z) - np.float32(0.7)).astype(np.float32).ravel()
        verts, tris = reflib.marching_cubes(vol, (24, 24, 24), 1 / 12, (-1, -1, -1), 0.0)
        n_verts, loops = len(verts), 0
    he, epv, epf = build_tables(tris, n_verts)
    got = reflib.topology_from_tables(he, epv, epf)
    assert got is not None, "MeshTopology::read rejected the tables"
    rtris, n_ue, n_loops, n_vv = got
    assert np.array_equal(rtris, tris) and n_ue == len(he) // 2 and n_loops == loops
    vf, ue, bl, vv, _s = reflib.builder_stats(tris)
    assert (vf, ue, bl, vv) == (len(tris), n_ue, n_loops, n_vv)


def test_the_reference_rejects_broken_tables():
    pts, tris = meshgen.make_torus(2, 1, 24, 12)
    he, epv, epf = build_tables(tris, len(pts))
    bad = he.copy()
    bad[5, 0] = bad[6, 0]                          # two half-edges share a next: prev(next(e)) != e
    assert reflib.topology_from_tables(bad, epv, epf) is None
