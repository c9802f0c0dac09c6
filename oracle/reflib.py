"""oracle/reflib.py -- TEST INFRASTRUCTURE ONLY (imported by tests/, __graft_entry__.smoke(), bench.py's
cpu_baseline / --impl reference leg; never by meshlib_b200/).

ctypes loader for oracle/_ref/libmrref.so = the UNMODIFIED reference CPU/TBB implementation of the path
(MR::FastWindingNumber, MR::AABBTree, MR::calcDipoles; /root/reference/source/MRMesh) behind our C wrapper
oracle/ref_wrap.cpp.  Built by `make -C oracle ref` in the build container; the .so travels to the GPU box.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "_ref", "libmrref.so")
_lib = None

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_u64p = np.ctypeslib.ndpointer(np.uint64, flags="C_CONTIGUOUS")


def available():
    return os.path.exists(_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(_PATH)
        L.mrref_hardware_threads.restype = C.c_int
        L.mrref_create.restype = C.c_void_p
        L.mrref_create.argtypes = [_f32p, C.c_size_t, _i32p, C.c_size_t]
        L.mrref_make_torus.restype = C.c_void_p
        L.mrref_make_torus.argtypes = [C.c_float, C.c_float, C.c_int, C.c_int]
        L.mrref_make_uv_sphere.restype = C.c_void_p
        L.mrref_make_uv_sphere.argtypes = [C.c_float, C.c_int, C.c_int]
        L.mrref_destroy.argtypes = [C.c_void_p]
        for n in ("mrref_num_verts", "mrref_num_faces", "mrref_num_nodes"):
            getattr(L, n).restype = C.c_size_t
            getattr(L, n).argtypes = [C.c_void_p]
        L.mrref_num_valid_faces.restype = C.c_int
        L.mrref_num_valid_faces.argtypes = [C.c_void_p]
        L.mrref_get_points.argtypes = [C.c_void_p, _f32p]
        L.mrref_get_tris.argtypes = [C.c_void_p, _i32p]
        L.mrref_get_nodes.argtypes = [C.c_void_p, C.c_void_p]
        L.mrref_get_dipoles.argtypes = [C.c_void_p, _f32p]
        L.mrref_calc_from_vector.argtypes = [C.c_void_p, _f32p, C.c_size_t, C.c_float, C.c_int, _f32p]
        L.mrref_calc_self_intersections.argtypes = [C.c_void_p, C.c_float, _u64p]
        L.mrref_calc_from_grid.argtypes = [C.c_void_p, _i32p, _f32p, C.c_float, _f32p]
        L.mrref_calc_from_grid_with_distances.argtypes = [
            C.c_void_p, _i32p, _f32p, C.c_float, C.c_float, C.c_int, C.c_float, C.c_float, _f32p]
        L.mrref_calc_grid_with_distances_layers.argtypes = [
            C.c_void_p, _i32p, _f32p, C.c_float, C.c_float, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, _f32p]
        L.mrref_find_projection_dist_sq.argtypes = [C.c_void_p, _f32p, C.c_size_t, C.c_float, C.c_float, _f32p]
        L.mrref_find_projections.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_float, C.c_float,
                                             C.c_void_p]
        L.mrref_update_points.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.mrref_signed_distance_proj_normal.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_float, C.c_float, C.c_int,
                                                        C.c_void_p, C.c_void_p]
        L.mrref_voxel_centres.argtypes = [_f32p, _f32p, _i32p, C.c_void_p]
        L.mrref_pseudonormals.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.mrref_marching_cubes.argtypes = [_f32p, _i32p, _f32p, _f32p, C.c_float, C.c_int, C.POINTER(C.c_size_t),
                                           C.POINTER(C.c_size_t)]
        L.mrref_marching_cubes_fetch.argtypes = [_f32p, _i32p]
        ip = C.POINTER(C.c_int)
        L.mrref_topology_from_tables.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p,
                                                 ip, ip, ip]
        L.mrref_builder_stats.argtypes = [C.c_void_p, C.c_size_t, ip, ip, ip, ip, C.POINTER(C.c_double)]
        _lib = L
    return _lib


NODE_DTYPE = np.dtype([("bmin", np.float32, 3), ("bmax", np.float32, 3), ("l", np.int32), ("r", np.int32)])
DIPOLE_DTYPE = np.dtype([("pos", np.float32, 3), ("area", np.float32), ("dirArea", np.float32, 3), ("rr", np.float32)])
FLT_MAX = float(np.finfo(np.float32).max)


PROJ_DTYPE = np.dtype([("face", "<i4"), ("point", "<f4", (3,)), ("a", "<f4"), ("b", "<f4"), ("dist_sq", "<f4"), ("edge", "<i4")])

class RefMesh:
    """The reference's MR::Mesh + MR::FastWindingNumber, driven from numpy."""

    def __init__(self, handle):
        self.h = C.c_void_p(handle)

    @classmethod
    def from_arrays(cls, points, tris):
        points = np.ascontiguousarray(points, np.float32)
        tris = np.ascontiguousarray(tris, np.int32)
        return cls(lib().mrref_create(points, len(points), tris, len(tris)))

    @classmethod
    def torus(cls, R, r, p_res, s_res):
        return cls(lib().mrref_make_torus(R, r, p_res, s_res))

    @classmethod
    def uv_sphere(cls, radius, h_res, v_res):
        return cls(lib().mrref_make_uv_sphere(radius, h_res, v_res))

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.mrref_destroy(self.h)
            self.h = None

    @property
    def points(self):
        out = np.empty((lib().mrref_num_verts(self.h), 3), np.float32)
        lib().mrref_get_points(self.h, out)
        return out

    @property
    def tris(self):
        """triangulation in the reference's per-face vertex order (MeshTopology::getTriangulation)"""
        out = np.empty((lib().mrref_num_faces(self.h), 3), np.int32)
        lib().mrref_get_tris(self.h, out)
        return out

    def nodes(self):
        out = np.empty(lib().mrref_num_nodes(self.h), NODE_DTYPE)
        lib().mrref_get_nodes(self.h, out.ctypes.data_as(C.c_void_p))
        return out

    def dipoles(self):
        n = lib().mrref_num_nodes(self.h)
        out = np.empty((n, 8), np.float32)
        lib().mrref_get_dipoles(self.h, out)
        return out.view(DIPOLE_DTYPE).reshape(n)

    def calc_from_vector(self, pts, beta=2.0, skip_face=-1):
        pts = np.ascontiguousarray(pts, np.float32)
        out = np.empty(len(pts), np.float32)
        assert lib().mrref_calc_from_vector(self.h, pts, len(pts), beta, skip_face, out) == 0
        return out

    def calc_self_intersections(self, beta=2.0):
        nf = lib().mrref_num_faces(self.h)
        out = np.zeros((nf + 63) // 64, np.uint64)
        assert lib().mrref_calc_self_intersections(self.h, beta, out) == 0
        return out

    def calc_from_grid(self, dims, xf12, beta=2.0):
        dims = np.asarray(dims, np.int32)
        out = np.empty(int(np.prod(dims.astype(np.int64))), np.float32)
        assert lib().mrref_calc_from_grid(self.h, dims, np.asarray(xf12, np.float32), beta, out) == 0
        return out

    def calc_from_grid_with_distances(self, dims, xf12, min_dist_sq=0.0, max_dist_sq=FLT_MAX, null_outside=True,
                                      threshold=0.5, beta=2.0):
        dims = np.asarray(dims, np.int32)
        out = np.empty(int(np.prod(dims.astype(np.int64))), np.float32)
        assert lib().mrref_calc_from_grid_with_distances(
            self.h, dims, np.asarray(xf12, np.float32), min_dist_sq, max_dist_sq, int(null_outside), threshold, beta,
            out) == 0
        return out

    def calc_grid_with_distances_layers(self, dims, xf12, z0, z1, min_dist_sq=0.0, max_dist_sq=FLT_MAX,
                                        null_outside=True, threshold=0.5, beta=2.0):
        dims = np.asarray(dims, np.int32)
        out = np.empty(int(dims[0]) * int(dims[1]) * (z1 - z0), np.float32)
        assert lib().mrref_calc_grid_with_distances_layers(
            self.h, dims, np.asarray(xf12, np.float32), min_dist_sq, max_dist_sq, int(null_outside), threshold, beta,
            z0, z1, out) == 0
        return out

    def find_projections(self, pts, point_xf=None, tree_xf=None, up=FLT_MAX, lo=0.0):
        """MR::PointsToMeshProjector::findProjections (the reference's CPU class): structured array incl. mtp.e"""
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 3)
        out = np.zeros(len(pts), PROJ_DTYPE)
        px = None if point_xf is None else np.ascontiguousarray(point_xf, np.float32)
        tx = None if tree_xf is None else np.ascontiguousarray(tree_xf, np.float32)
        lib().mrref_find_projections(self.h, pts.ctypes.data, len(pts), None if px is None else px.ctypes.data,
                                     None if tx is None else tx.ctypes.data, up, lo, out.ctypes.data)
        return out

    def update_points(self, pts):
        """Mesh::updateCaches after moving every vertex: the cached AABB tree is refitted (AABBTree::refit), not rebuilt"""
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 3)
        assert lib().mrref_update_points(self.h, pts.ctypes.data, len(pts)) == 0

    def signed_distance_proj_normal(self, pts, min_dist_sq=0.0, max_dist_sq=FLT_MAX, null_outside=True):
        """MR::signedDistanceToMesh with SignDetectionMode::ProjectionNormal (source/MRMesh/MRMeshDistance.cpp:13-38) per point:
        (signed distances with NaN, dot( proj - pt, pseudonormal ) = the margin of each sign decision)"""
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 3)
        dist, dots = np.empty(len(pts), np.float32), np.empty(len(pts), np.float32)
        lib().mrref_signed_distance_proj_normal(self.h, pts.ctypes.data, len(pts), min_dist_sq, max_dist_sq, int(null_outside),
                                                dist.ctypes.data, dots.ctypes.data)
        return dist, dots

    def pseudonormals(self, n_faces, n_verts):
        """(face normals (F,3), vertex pseudonormals (V,3), edge pseudonormals (F,3,3)) of the reference's topology"""
        fn, vn, en = np.zeros((n_faces, 3), np.float32), np.zeros((n_verts, 3), np.float32), np.zeros((n_faces, 3, 3), np.float32)
        lib().mrref_pseudonormals(self.h, fn.ctypes.data, vn.ctypes.data, en.ctypes.data)
        return fn, vn, en

    def find_projection_dist_sq(self, pts, up=FLT_MAX, lo=0.0):
        pts = np.ascontiguousarray(pts, np.float32)
        out = np.empty(len(pts), np.float32)
        lib().mrref_find_projection_dist_sq(self.h, pts, len(pts), up, lo, out)
        return out


def marching_cubes(volume, dims, voxel_size, origin, iso, less_inside=False):
    """MR::marchingCubesAsTriMesh(SimpleVolume) (source/MRVoxels/MRMarchingCubes.cpp:1112-1115): (verts (n,3) f32, tris (m,3) i32)"""
    vol = np.ascontiguousarray(volume, np.float32).reshape(-1)
    d = np.asarray(dims, np.int32)
    assert vol.size == int(d[0]) * int(d[1]) * int(d[2])
    vs = np.ascontiguousarray(np.asarray(voxel_size, np.float32) * np.ones(3, np.float32))
    o = np.ascontiguousarray(origin, np.float32)
    nv, nt = C.c_size_t(), C.c_size_t()
    if lib().mrref_marching_cubes(vol, d, vs, o, iso, int(less_inside), C.byref(nv), C.byref(nt)) != 0:
        raise RuntimeError("reference marching cubes failed")
    verts = np.empty((nv.value, 3), np.float32)
    tris = np.empty((nt.value, 3), np.int32)
    lib().mrref_marching_cubes_fetch(verts, tris)
    return verts, tris


def topology_from_tables(half_edges, edge_per_vertex, edge_per_face):
    """half-edge tables {next, prev, org, left} through the reference's MeshTopology::read (which validates them with
    checkValidity); returns None when the reference rejects them, else (tris read back, undirected edges, boundary loops, valid verts)"""
    he = np.ascontiguousarray(half_edges, np.int32).reshape(-1, 4)
    epv = np.ascontiguousarray(edge_per_vertex, np.int32)
    epf = np.ascontiguousarray(edge_per_face, np.int32)
    tris = np.full((len(epf), 3), -1, np.int32)
    ue, bl, vv = C.c_int(), C.c_int(), C.c_int()
    rc = lib().mrref_topology_from_tables(he.ctypes.data, len(he), epv.ctypes.data, len(epv), epf.ctypes.data, len(epf), tris.ctypes.data,
                                          C.byref(ue), C.byref(bl), C.byref(vv))
    if rc != 0:
        return None
    return tris, ue.value, bl.value, vv.value


def builder_stats(tris):
    """MeshBuilder::fromTriangles of the reference on the same triangles: (valid faces, undirected edges, boundary loops, valid verts, s)"""
    t = np.ascontiguousarray(tris, np.int32).reshape(-1, 3)
    vf, ue, bl, vv, sec = C.c_int(), C.c_int(), C.c_int(), C.c_int(), C.c_double()
    lib().mrref_builder_stats(t.ctypes.data, len(t), C.byref(vf), C.byref(ue), C.byref(bl), C.byref(vv), C.byref(sec))
    return vf.value, ue.value, bl.value, vv.value, sec.value


def hardware_threads():
    return lib().mrref_hardware_threads()


def voxel_centres(origin, voxel_size, dims):
    """voxel centres as meshToDistanceFunctionVolume computes them (source/MRVoxels/MRMeshToDistanceVolume.cpp:83-84), x fastest"""
    d = np.ascontiguousarray(dims, np.int32)
    out = np.empty((int(d[0]) * int(d[1]) * int(d[2]), 3), np.float32)
    vs = np.ascontiguousarray(np.asarray(voxel_size, np.float32) * np.ones(3, np.float32))
    lib().mrref_voxel_centres(np.ascontiguousarray(origin, np.float32), vs, d, out.ctypes.data)
    return out
