"""TEST INFRASTRUCTURE ONLY: ctypes binding of oracle/liboracle_normals.so -- the ProjectionNormal arithmetic header of the
product (meshlib_b200/csrc/b2fwn_normals_math.h) run on the host, step for step like the device kernels, so that tests
without a GPU can hold it against the reference library."""
import ctypes as C
import os

import numpy as np

_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "liboracle_normals.so")
_lib = None


def available():
    return os.path.exists(_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(_PATH)
        vp = C.c_void_p
        L.b2n_host_tables.argtypes = [vp, C.c_int, vp, C.c_int, vp, vp, vp]
        L.b2n_host_sign.argtypes = [vp, vp, C.c_size_t, vp, vp, vp, vp, C.c_float, C.c_float, C.c_int, vp]
        L.b2n_host_voxel_centres.argtypes = [vp, vp, vp, vp]
        _lib = L
    return _lib


def tables(points, tris):
    """(face normals (F,3), vertex pseudonormals (V,3), face across each edge (F,3))"""
    pts = np.ascontiguousarray(points, np.float32)
    t = np.ascontiguousarray(tris, np.int32)
    fn, vn = np.zeros((len(t), 3), np.float32), np.zeros((len(pts), 3), np.float32)
    adj = np.zeros((len(t), 3), np.int32)
    lib().b2n_host_tables(pts.ctypes.data, len(pts), t.ctypes.data, len(t), fn.ctypes.data, vn.ctypes.data, adj.ctypes.data)
    return fn, vn, adj


def sign(records, query_points, tris, tabs, min_dist_sq=0.0, max_dist_sq=float(np.finfo(np.float32).max), null_outside=True):
    """signed distances from 32-byte projection records (face, point, bary a/b, distSq) of the query points"""
    rec = np.ascontiguousarray(records)
    assert rec.dtype.itemsize == 32
    q = np.ascontiguousarray(query_points, np.float32).reshape(-1, 3)
    t = np.ascontiguousarray(tris, np.int32)
    fn, vn, adj = tabs
    out = np.empty(len(q), np.float32)
    lib().b2n_host_sign(rec.ctypes.data, q.ctypes.data, len(q), t.ctypes.data, fn.ctypes.data, vn.ctypes.data, adj.ctypes.data,
                        min_dist_sq, max_dist_sq, int(null_outside), out.ctypes.data)
    return out


def voxel_centres(origin, voxel_size, dims):
    d = np.ascontiguousarray(dims, np.int32)
    o = np.ascontiguousarray(origin, np.float32)
    vs = np.ascontiguousarray(np.asarray(voxel_size, np.float32) * np.ones(3, np.float32))
    out = np.empty((int(d[0]) * int(d[1]) * int(d[2]), 3), np.float32)
    lib().b2n_host_voxel_centres(o.ctypes.data, vs.ctypes.data, d.ctypes.data, out.ctypes.data)
    return out
