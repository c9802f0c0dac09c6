"""oracle/refcudalib.py -- TEST / BENCH INFRASTRUCTURE ONLY.  ctypes loader of oracle/_ref/libmrrefcuda.so: the reference's
UNMODIFIED CUDA kernels (source/MRCuda/MRCudaFastWindingNumber.cu, -arch sm_100 -use_fast_math) behind our driver
oracle/refcuda_wrap.cu, fed the CPU-built tree and dipoles like MR::Cuda::FastWindingNumber::prepareData_ does.
SURVEY.md 8(d): the second, on-box baseline."""
import ctypes as C
import os

import numpy as np

_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libmrrefcuda.so")
_lib = None


def available():
    return os.path.exists(_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(_PATH)
        L.mrrefcuda_create.restype = C.c_void_p
        L.mrrefcuda_create.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
        L.mrrefcuda_destroy.argtypes = [C.c_void_p]
        L.mrrefcuda_signed_distance.restype = C.c_double
        L.mrrefcuda_signed_distance.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_int, C.c_float,
                                                C.c_float, C.c_int, C.c_int, C.c_void_p]
        _lib = L
    return _lib


class RefCudaFwn:
    """nodes/dipoles: the reference CPU tree (oracle mesh .nodes()/.dipoles()); points, tris as uploaded by prepareData_"""

    def __init__(self, nodes, dipoles, points, tris):
        nodes = np.ascontiguousarray(nodes)
        dip = np.ascontiguousarray(dipoles)
        pts = np.ascontiguousarray(points, np.float32)
        t = np.ascontiguousarray(tris, np.int32)
        assert nodes.itemsize == 32 and dip.nbytes == 32 * len(nodes)
        self.h = lib().mrrefcuda_create(nodes.ctypes.data, dip.ctypes.data, len(nodes), pts.ctypes.data, len(pts), t.ctypes.data, len(t))
        if not self.h:
            raise RuntimeError("mrrefcuda_create failed")

    def signed_distance_layers(self, dims, xf12, z0, z1, min_dist_sq=0.0, max_dist_sq=float(np.finfo(np.float32).max),
                               null_outside=True, threshold=0.5, beta=2.0, fetch=True):
        """returns (kernel ms, values or None)"""
        d = np.ascontiguousarray(dims, np.int32)
        xf = np.ascontiguousarray(xf12, np.float32)
        out = np.empty(int(d[0]) * int(d[1]) * (z1 - z0), np.float32) if fetch else None
        ms = lib().mrrefcuda_signed_distance(self.h, d.ctypes.data, xf.ctypes.data, min_dist_sq, max_dist_sq, int(null_outside),
                                             threshold, beta, z0, z1, out.ctypes.data if fetch else None)
        if ms < 0:
            raise RuntimeError("reference CUDA kernel failed")
        return ms, out

    def close(self):
        if self.h:
            lib().mrrefcuda_destroy(self.h)
        self.h = None

    def __del__(self):
        self.close()
