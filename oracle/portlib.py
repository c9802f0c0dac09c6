"""oracle/portlib.py -- TEST INFRASTRUCTURE ONLY.  ctypes loader of oracle/liboracle_port.so (oracle/port.c, our plain-C
restatement of the reference path).  Same method names as oracle/reflib.RefMesh so the two are interchangeable."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "liboracle_port.so")
_lib = None
FLT_MAX = float(np.finfo(np.float32).max)
NODE_DTYPE = np.dtype([("bmin", np.float32, 3), ("bmax", np.float32, 3), ("l", np.int32), ("r", np.int32)])
DIPOLE_DTYPE = np.dtype([("pos", np.float32, 3), ("area", np.float32), ("dirArea", np.float32, 3), ("rr", np.float32)])
COUNTER_NAMES = ("fwn_visits", "fwn_accepts", "fwn_leaves", "box_tests", "tri_tests")


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH):
            subprocess.run(["make", "port"], cwd=_HERE, check=True, stdout=subprocess.DEVNULL)
        L = C.CDLL(_PATH)
        vp, sz = C.c_void_p, C.c_size_t
        L.orc_create.restype = vp
        L.orc_create.argtypes = [vp, sz, vp, sz]
        L.orc_destroy.argtypes = [vp]
        L.orc_num_nodes.restype = sz
        L.orc_num_nodes.argtypes = [vp]
        L.orc_get_nodes.argtypes = [vp, vp]
        L.orc_update_points.argtypes = [vp, vp, sz]
        L.orc_get_dipoles.argtypes = [vp, vp]
        L.orc_calc_from_vector.argtypes = [vp, vp, sz, C.c_float, C.c_int32, vp, vp]
        L.orc_calc_self_intersections.argtypes = [vp, C.c_float, vp, vp]
        L.orc_calc_from_grid.argtypes = [vp, vp, vp, C.c_float, C.c_int, C.c_int, vp, vp]
        L.orc_calc_from_grid_with_distances.argtypes = [vp, vp, vp, C.c_float, C.c_float, C.c_int, C.c_float, C.c_float,
                                                        C.c_int, C.c_int, C.c_longlong, vp, vp]
        L.orc_find_projection_dist_sq.argtypes = [vp, vp, sz, C.c_float, C.c_float, vp]
        L.orc_find_projections.argtypes = [vp, vp, sz, vp, vp, C.c_float, C.c_float, vp]
        L.orc_triangle_solid_angle.restype = C.c_float
        L.orc_triangle_solid_angle.argtypes = [vp, vp]
        L.orc_chunk_count.restype = sz
        L.orc_chunk_count.argtypes = [sz, sz, sz]
        L.orc_chunk.argtypes = [sz, sz, sz, sz, C.POINTER(sz), C.POINTER(sz)]
        L.orc_marching_cubes.argtypes = [vp, vp, vp, vp, C.c_float, C.c_int, C.POINTER(sz), C.POINTER(sz)]
        L.orc_marching_cubes_fetch.argtypes = [vp, vp]
        L.orc_mc_table.restype = C.POINTER(C.c_byte)
        _lib = L
    return _lib


def triangle_solid_angle(p, tri):
    p = np.ascontiguousarray(p, np.float32)
    tri = np.ascontiguousarray(tri, np.float32).reshape(9)
    return float(lib().orc_triangle_solid_angle(p.ctypes.data, tri.ctypes.data))


def marching_cubes(volume, dims, voxel_size, origin, iso, less_inside=False):
    """restatement of MR::marchingCubesAsTriMesh(SimpleVolume) (oracle/port_mc.c): (verts (n,3) f32, tris (m,3) i32)"""
    vol = np.ascontiguousarray(volume, np.float32).reshape(-1)
    d = np.ascontiguousarray(dims, np.int32)
    assert vol.size == int(d[0]) * int(d[1]) * int(d[2])
    vs = np.ascontiguousarray(np.asarray(voxel_size, np.float32) * np.ones(3, np.float32))
    o = np.ascontiguousarray(origin, np.float32)
    nv, nt = C.c_size_t(), C.c_size_t()
    if lib().orc_marching_cubes(vol.ctypes.data, d.ctypes.data, vs.ctypes.data, o.ctypes.data, iso, int(less_inside),
                                C.byref(nv), C.byref(nt)) != 0:
        raise RuntimeError("a part must have at least two Z slices")
    verts = np.empty((nv.value, 3), np.float32)
    tris = np.empty((nt.value, 3), np.int32)
    lib().orc_marching_cubes_fetch(verts.ctypes.data, tris.ctypes.data)
    return verts, tris


def mc_table():
    """the 256 x 16 case table the restatement and the CUDA kernels share"""
    return np.ctypeslib.as_array(lib().orc_mc_table(), shape=(256, 16)).copy()


def split_by_chunks(total, chunk, overlap):
    """[(offset, size), ...] as MR::splitByChunks yields them"""
    n = lib().orc_chunk_count(total, chunk, overlap)
    out = []
    for k in range(n):
        o, s = C.c_size_t(), C.c_size_t()
        lib().orc_chunk(total, chunk, overlap, k, C.byref(o), C.byref(s))
        out.append((o.value, s.value))
    return out


PROJ_DTYPE = np.dtype([("face", "<i4"), ("point", "<f4", (3,)), ("a", "<f4"), ("b", "<f4"), ("dist_sq", "<f4"), ("edge", "<i4")])

class PortMesh:
    kind = "port"

    def __init__(self, points, tris):
        self.points = np.ascontiguousarray(points, np.float32).reshape(-1, 3)
        self.tris = np.ascontiguousarray(tris, np.int32).reshape(-1, 3)
        self.h = C.c_void_p(lib().orc_create(self.points.ctypes.data, len(self.points), self.tris.ctypes.data, len(self.tris)))
        self.counters = np.zeros(5, np.uint64)

    @classmethod
    def from_arrays(cls, points, tris):
        return cls(points, tris)

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.orc_destroy(self.h)
            self.h = None

    def reset_counters(self):
        self.counters[:] = 0

    def counter_dict(self):
        return dict(zip(COUNTER_NAMES, (int(x) for x in self.counters)))

    def update_points(self, pts):
        """Mesh::updateCaches with every vertex moved: AABBTree::refit (MRAABBTree.cpp:62-94) + dipoles on the refitted tree"""
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 3)
        assert lib().orc_update_points(self.h, pts.ctypes.data, len(pts)) == 0
        self.points = pts

    def nodes(self):
        out = np.empty(lib().orc_num_nodes(self.h), NODE_DTYPE)
        lib().orc_get_nodes(self.h, out.ctypes.data)
        return out

    def dipoles(self):
        out = np.empty(lib().orc_num_nodes(self.h), DIPOLE_DTYPE)
        lib().orc_get_dipoles(self.h, out.ctypes.data)
        return out

    def calc_from_vector(self, pts, beta=2.0, skip_face=-1):
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 3)
        out = np.empty(len(pts), np.float32)
        lib().orc_calc_from_vector(self.h, pts.ctypes.data, len(pts), beta, skip_face, out.ctypes.data, self.counters.ctypes.data)
        return out

    def calc_self_intersections(self, beta=2.0, return_w=False):
        nf = len(self.tris)
        out = np.zeros((nf + 63) // 64, np.uint64)
        w = np.zeros(nf, np.float32)
        lib().orc_calc_self_intersections(self.h, beta, out.ctypes.data, w.ctypes.data)
        return (out, w) if return_w else out

    def calc_from_grid(self, dims, xf12, beta=2.0, z0=0, z1=None):
        dims = np.asarray(dims, np.int32)
        z1 = int(dims[2]) if z1 is None else z1
        xf = np.ascontiguousarray(xf12, np.float32)
        out = np.empty(int(dims[0]) * int(dims[1]) * (z1 - z0), np.float32)
        lib().orc_calc_from_grid(self.h, dims.ctypes.data, xf.ctypes.data, beta, z0, z1, out.ctypes.data, self.counters.ctypes.data)
        return out

    def calc_from_grid_with_distances(self, dims, xf12, min_dist_sq=0.0, max_dist_sq=FLT_MAX, null_outside=True,
                                      threshold=0.5, beta=2.0, z0=0, z1=None, stride=1):
        dims = np.asarray(dims, np.int32)
        z1 = int(dims[2]) if z1 is None else z1
        xf = np.ascontiguousarray(xf12, np.float32)
        out = np.zeros(int(dims[0]) * int(dims[1]) * (z1 - z0), np.float32)
        lib().orc_calc_from_grid_with_distances(self.h, dims.ctypes.data, xf.ctypes.data, min_dist_sq, max_dist_sq,
                                                int(null_outside), threshold, beta, z0, z1, stride, out.ctypes.data,
                                                self.counters.ctypes.data)
        return out

    def calc_grid_with_distances_layers(self, dims, xf12, z0, z1, **kw):
        return self.calc_from_grid_with_distances(dims, xf12, z0=z0, z1=z1, **kw)

    def find_projections(self, pts, point_xf=None, tree_xf=None, up=FLT_MAX, lo=0.0):
        """findProjections (MRPointsToMeshProjector.cpp:17-33): structured array face/point/a/b/dist_sq (edge = -1)"""
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 3)
        out = np.zeros(len(pts), PROJ_DTYPE)
        px = None if point_xf is None else np.ascontiguousarray(point_xf, np.float32)
        tx = None if tree_xf is None else np.ascontiguousarray(tree_xf, np.float32)
        lib().orc_find_projections(self.h, pts.ctypes.data, len(pts), None if px is None else px.ctypes.data,
                                   None if tx is None else tx.ctypes.data, up, lo, out.ctypes.data)
        return out

    def find_projection_dist_sq(self, pts, up=FLT_MAX, lo=0.0):
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 3)
        out = np.empty(len(pts), np.float32)
        lib().orc_find_projection_dist_sq(self.h, pts.ctypes.data, len(pts), up, lo, out.ctypes.data)
        return out
