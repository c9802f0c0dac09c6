"""meshlib_b200.api -- ctypes binding of libb2fwn.so (include/b2fwn.h) with the reference's operator names.

`FastWindingNumber` mirrors MR::IFastWindingNumber / MR::IFastWindingNumberByParts
(/root/reference/source/MRMesh/MRFastWindingNumber.h:11-45, 74-107) and MR::Cuda::FastWindingNumber
(source/MRCuda/MRCudaFastWindingNumber.h:17-38): same method names (snake_case), argument meaning, output
conventions and error strings.  It is a binding, not an implementation: every method is one C-ABI call, and the
module raises at import-of-library time if the CUDA library is missing (there is no CPU fallback).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb2fwn.so")
FLT_MAX = float(np.finfo(np.float32).max)

_lib = None

PROGRESS_CB = C.CFUNCTYPE(C.c_int, C.c_float, C.c_void_p)
PART_CB = C.CFUNCTYPE(C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_int32), C.c_int32, C.c_void_p)
PART_BUFFER_CB = C.CFUNCTYPE(C.c_void_p, C.POINTER(C.c_int32), C.c_int32, C.c_void_p)


class B2Error(RuntimeError):
    """the C ABI returned a non-zero status; str(e) is b2fwn_last_error() (reference: Expected<void>'s error string)"""

    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


MAX_DEVICES = 16


class Config(C.Structure):
    """b2fwn_config: one device, or n_devices > 1 devices of this process (the host-output grid calls use them all)"""
    _fields_ = [("device", C.c_int32), ("n_devices", C.c_int32), ("devices", C.c_int32 * MAX_DEVICES), ("reserved", C.c_int32 * 6)]


class DistOptions(C.Structure):
    """MR::DistanceToMeshOptions (source/MRMesh/MRDistanceToMeshOptions.h:10-33)"""
    _fields_ = [("min_dist_sq", C.c_float), ("max_dist_sq", C.c_float), ("null_outside_min_max", C.c_int32),
                ("winding_number_threshold", C.c_float), ("winding_number_beta", C.c_float)]

    def __init__(self, min_dist_sq=0.0, max_dist_sq=FLT_MAX, null_outside_min_max=True, winding_number_threshold=0.5,
                 winding_number_beta=2.0):
        super().__init__(min_dist_sq, max_dist_sq, int(null_outside_min_max), winding_number_threshold,
                         winding_number_beta)


class Stats(C.Structure):
    _fields_ = [("tree_build_ms", C.c_double), ("last_kernel_ms", C.c_double), ("num_nodes", C.c_uint64),
                ("num_leaves", C.c_uint64), ("kernel_launches", C.c_uint64), ("reserved", C.c_uint64 * 3)]


class McParams(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("voxel_size", C.c_float * 3), ("iso", C.c_float),
                ("less_inside", C.c_int32), ("reserved", C.c_int32 * 4)]


class VolumeParams(C.Structure):
    """MR::DistanceVolumeParams (source/MRVoxels/MRDistanceVolumeParams.h:9-22)"""
    _fields_ = [("origin", C.c_float * 3), ("voxel_size", C.c_float * 3), ("dims", C.c_int32 * 3), ("reserved", C.c_int32 * 3)]


class OffsetParams(C.Structure):
    """the fields of MR::OffsetParameters that mcOffsetMesh reads in HoleWindingRule mode (source/MRVoxels/MROffset.h)"""
    _fields_ = [("voxel_size", C.c_float), ("offset", C.c_float), ("close_holes", C.c_int32),
                ("winding_number_threshold", C.c_float), ("winding_number_beta", C.c_float), ("max_layers_per_slab", C.c_int32),
                ("reserved", C.c_int32 * 2)]


# b2fwn_projection (include/b2fwn.h)
PROJECTION_DTYPE = np.dtype([("face", "<i4"), ("point", "<f4", (3,)), ("a", "<f4"), ("b", "<f4"), ("dist_sq", "<f4"),
                             ("reserved", "<i4")])


class DeviceBuffer:
    """a raw device allocation exposed through __cuda_array_interface__ (zero-copy view: torch.as_tensor(buf, device=...))"""

    def __init__(self, ptr, n_floats):
        self.ptr, self.n = ptr, n_floats
        self.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (ptr, False), "version": 2, "strides": None}


class OffsetResult(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("dims", C.c_int32 * 3), ("n_verts", C.c_uint64), ("n_tris", C.c_uint64),
                ("sdf_ms", C.c_double), ("mc_ms", C.c_double)]


EXPORTS = [
    "b2fwn_last_error", "b2fwn_device_count", "b2fwn_create", "b2fwn_destroy", "b2fwn_calc_from_vector",
    "b2fwn_calc_from_vector_device", "b2fwn_calc_self_intersections", "b2fwn_calc_self_intersections_device",
    "b2fwn_calc_from_grid", "b2fwn_calc_from_grid_device", "b2fwn_calc_from_grid_with_distances",
    "b2fwn_calc_from_grid_with_distances_device", "b2fwn_calc_from_grid_with_distances_device_chunks",
    "b2fwn_calc_from_grid_with_distances_device_chunks_p2p", "b2fwn_shared_alloc", "b2fwn_shared_open", "b2fwn_shared_close",
    "b2fwn_shared_free", "b2fwn_find_projections", "b2fwn_find_projections_device", "b2fwn_calc_from_grid_by_parts",
    "b2fwn_calc_from_grid_with_distances_by_parts", "b2fwn_get_tree", "b2fwn_get_stats", "b2fwn_marching_cubes",
    "b2fwn_marching_cubes_fetch", "b2fwn_calc_from_grid_with_distances_layers", "b2fwn_measure_fp32_peak",
    "b2fwn_get_bounding_box", "b2fwn_mesh_to_distance_volume", "b2fwn_mc_offset_mesh",
    "b2fwn_make_signed_by_winding_number", "b2fwn_make_signed_by_winding_number_device",
    "b2fwn_signed_distances_by_projection_normal", "b2fwn_signed_distances_by_projection_normal_device",
    "b2fwn_mesh_to_distance_volume_by_projection_normal", "b2fwn_update_points",
    "b2fwn_calc_from_grid_by_parts_into", "b2fwn_calc_from_grid_with_distances_by_parts_into", "b2fwn_safe_memory_limit",
    "b2fwn_build_topology", "b2fwn_build_topology_fetch", "b2fwn_suggest_voxel_size", "b2fwn_plan_parts",
    "b2fwn_calc_from_vector_device_noalloc", "b2fwn_trim", "b2fwn_plan_parts_for",
]


def load_library():
    """loads libb2fwn.so; raises (never falls back) when it is missing"""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, f32p, i32p, u64p = C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int32), C.POINTER(C.c_uint64)
    L.b2fwn_last_error.restype = C.c_char_p
    L.b2fwn_device_count.restype = C.c_int
    L.b2fwn_create.argtypes = [vp, C.c_size_t, vp, C.c_size_t, C.POINTER(Config), C.POINTER(vp)]
    L.b2fwn_destroy.argtypes = [vp]
    L.b2fwn_destroy.restype = None
    L.b2fwn_calc_from_vector.argtypes = [vp, vp, C.c_size_t, C.c_float, C.c_int32, vp, PROGRESS_CB, vp]
    L.b2fwn_calc_from_vector_device.argtypes = [vp, vp, C.c_size_t, C.c_float, C.c_int32, vp]
    L.b2fwn_trim.argtypes = [vp]
    L.b2fwn_plan_parts_for.argtypes = [vp, i32p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, i32p, i32p, i32p, i32p, i32p]
    L.b2fwn_calc_from_vector_device_noalloc.argtypes = [vp, vp, C.c_size_t, C.c_float, C.c_int32, vp]
    L.b2fwn_calc_self_intersections.argtypes = [vp, C.c_float, vp, PROGRESS_CB, vp]
    L.b2fwn_calc_self_intersections_device.argtypes = [vp, C.c_float, vp, vp]
    L.b2fwn_calc_from_grid.argtypes = [vp, i32p, f32p, C.c_float, vp, PROGRESS_CB, vp]
    L.b2fwn_calc_from_grid_device.argtypes = [vp, i32p, f32p, C.c_float, C.c_int32, C.c_int32, vp]
    L.b2fwn_calc_from_grid_with_distances.argtypes = [vp, i32p, f32p, C.POINTER(DistOptions), vp, PROGRESS_CB, vp]
    L.b2fwn_calc_from_grid_with_distances_device.argtypes = [vp, i32p, f32p, C.POINTER(DistOptions), C.c_int32,
                                                             C.c_int32, vp]
    L.b2fwn_calc_from_grid_with_distances_device_chunks.argtypes = [vp, i32p, f32p, C.POINTER(DistOptions), C.c_int32,
                                                                    C.c_int32, C.c_int32, C.c_int32, vp]
    L.b2fwn_calc_from_grid_with_distances_device_chunks_p2p.argtypes = [vp, i32p, f32p, C.POINTER(DistOptions), C.c_int32,
                                                                        C.c_int32, C.c_int32, C.c_int32, vp, C.c_int32,
                                                                        C.POINTER(vp)]
    L.b2fwn_find_projections.argtypes = [vp, vp, C.c_size_t, vp, vp, C.c_float, C.c_float, vp, PROGRESS_CB, vp]
    L.b2fwn_find_projections_device.argtypes = [vp, vp, C.c_size_t, vp, vp, C.c_float, C.c_float, vp]
    L.b2fwn_shared_alloc.argtypes = [vp, C.c_size_t, C.POINTER(vp), C.c_char_p]
    L.b2fwn_shared_open.argtypes = [vp, C.c_char_p, C.POINTER(vp)]
    L.b2fwn_shared_close.argtypes = [vp, vp]
    L.b2fwn_shared_free.argtypes = [vp, vp]
    L.b2fwn_calc_from_grid_with_distances_layers.argtypes = [vp, i32p, f32p, C.POINTER(DistOptions), C.c_int32, C.c_int32,
                                                             vp, PROGRESS_CB, vp]
    L.b2fwn_make_signed_by_winding_number_device.argtypes = [vp, i32p, f32p, C.c_float, C.c_float, C.c_int32, C.c_int32, vp]
    L.b2fwn_make_signed_by_winding_number.argtypes = [vp, i32p, f32p, C.c_float, C.c_float, vp, PROGRESS_CB, vp]
    L.b2fwn_measure_fp32_peak.argtypes = [C.c_int32, C.POINTER(C.c_double)]
    L.b2fwn_calc_from_grid_by_parts.argtypes = [vp, i32p, f32p, C.c_float, C.c_int32, C.c_int32, PART_CB, vp,
                                                PROGRESS_CB, vp]
    L.b2fwn_calc_from_grid_with_distances_by_parts.argtypes = [vp, i32p, f32p, C.POINTER(DistOptions), C.c_int32,
                                                               C.c_int32, PART_CB, vp, PROGRESS_CB, vp]
    L.b2fwn_calc_from_grid_by_parts_into.argtypes = [vp, i32p, f32p, C.c_float, C.c_int32, C.c_int32, PART_BUFFER_CB, PART_CB, vp,
                                                     PROGRESS_CB, vp]
    L.b2fwn_calc_from_grid_with_distances_by_parts_into.argtypes = [vp, i32p, f32p, C.POINTER(DistOptions), C.c_int32, C.c_int32,
                                                                    PART_BUFFER_CB, PART_CB, vp, PROGRESS_CB, vp]
    L.b2fwn_build_topology.argtypes = [vp, vp, C.c_size_t, C.c_int32, C.c_size_t, u64p, u64p, u64p]
    L.b2fwn_build_topology_fetch.argtypes = [vp, vp, vp, vp]
    L.b2fwn_suggest_voxel_size.argtypes = [vp, C.c_float, f32p]
    L.b2fwn_plan_parts.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, i32p, i32p, i32p, i32p]
    L.b2fwn_safe_memory_limit.argtypes = [C.c_int32]
    L.b2fwn_safe_memory_limit.restype = C.c_size_t
    L.b2fwn_get_tree.argtypes = [vp, vp, vp, C.c_size_t, C.POINTER(C.c_size_t)]
    L.b2fwn_update_points.argtypes = [vp, vp, C.c_size_t]
    L.b2fwn_get_stats.argtypes = [vp, C.POINTER(Stats)]
    L.b2fwn_marching_cubes.argtypes = [vp, vp, C.c_int32, i32p, C.POINTER(McParams), u64p, u64p]
    L.b2fwn_marching_cubes_fetch.argtypes = [vp, vp, vp]
    L.b2fwn_get_bounding_box.argtypes = [vp, f32p]
    L.b2fwn_mesh_to_distance_volume.argtypes = [vp, C.POINTER(VolumeParams), C.POINTER(DistOptions), vp, f32p, f32p,
                                                PROGRESS_CB, vp]
    L.b2fwn_mc_offset_mesh.argtypes = [vp, C.POINTER(OffsetParams), C.POINTER(OffsetResult), PROGRESS_CB, vp]
    L.b2fwn_signed_distances_by_projection_normal.argtypes = [vp, vp, C.c_size_t, C.POINTER(DistOptions), vp, PROGRESS_CB, vp]
    L.b2fwn_signed_distances_by_projection_normal_device.argtypes = [vp, vp, C.c_size_t, C.POINTER(DistOptions), vp]
    L.b2fwn_mesh_to_distance_volume_by_projection_normal.argtypes = [vp, C.POINTER(VolumeParams), C.POINTER(DistOptions), vp,
                                                                     f32p, f32p, PROGRESS_CB, vp]
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise B2Error(rc, load_library().b2fwn_last_error().decode())


def _dims(dims):
    return (C.c_int32 * 3)(*[int(d) for d in dims])


def _xf(xf12):
    a = np.asarray(xf12, np.float32).reshape(12)
    return (C.c_float * 12)(*a.tolist())


def _progress(cb):
    if cb is None:
        return C.cast(None, PROGRESS_CB)
    return PROGRESS_CB(lambda v, _u: 1 if cb(v) else 0)


def make_basis_xf(origin, voxel_size):
    """gridToMeshXf of meshToDistanceVolume: scale(voxelSize), b = origin + 0.5*voxelSize (voxel centres;
    source/MRVoxels/MRMeshToDistanceVolume.cpp:55)"""
    o = np.asarray(origin, np.float32)
    v = np.asarray(voxel_size, np.float32) * np.ones(3, np.float32)
    b = o + np.float32(0.5) * v
    return np.array([v[0], 0, 0, 0, v[1], 0, 0, 0, v[2], b[0], b[1], b[2]], np.float32)


def plan_parts(n_layers, layers_per_part, layer_overlap=0, n_devices=1):
    """[(first layer, layers, device index), ...]: the partition of the by-parts / host-output calls (b2fwn_plan_parts; needs no GPU)"""
    L = load_library()
    n = C.c_int32()
    _check(L.b2fwn_plan_parts(n_layers, layers_per_part, layer_overlap, n_devices, 0, None, None, None, C.byref(n)))
    b, s, d = ((C.c_int32 * max(n.value, 1))() for _ in range(3))
    _check(L.b2fwn_plan_parts(n_layers, layers_per_part, layer_overlap, n_devices, n.value, b, s, d, C.byref(n)))
    return [(b[k], s[k], d[k]) for k in range(n.value)]


def measure_fp32_peak(device=0):
    """measured FP32 FMA throughput of the device in TFLOP/s (roofline denominator of the traversal kernels)"""
    v = C.c_double()
    _check(load_library().b2fwn_measure_fp32_peak(device, C.byref(v)))
    return v.value


NODE_DTYPE = np.dtype([("bmin", np.float32, 3), ("bmax", np.float32, 3), ("l", np.int32), ("r", np.int32)])
DIPOLE_DTYPE = np.dtype([("pos", np.float32, 3), ("area", np.float32), ("dirArea", np.float32, 3), ("rr", np.float32)])


class FastWindingNumber:
    """B200 implementation of MR::IFastWindingNumber + IFastWindingNumberByParts.

    points: (V,3) float32; tris: (F,3) int32 in MeshTopology::getTriangulation() order (negative ids = deleted face).
    The mesh is uploaded and the AABB tree + dipoles are built on the device in the constructor (the reference does it
    lazily at the first call, MRCudaFastWindingNumber.cpp:46-85)."""

    def __init__(self, points, tris, device=0, devices=None):
        """devices: list of CUDA ordinals of THIS process for the host-output grid calls (b2fwn_config.n_devices > 1)"""
        self._h = None
        L = load_library()
        self.points = np.ascontiguousarray(points, np.float32).reshape(-1, 3)
        self.tris = np.ascontiguousarray(tris, np.int32).reshape(-1, 3)
        cfg = Config(device=device)
        if devices is not None and len(devices) > 1:
            cfg.device, cfg.n_devices = int(devices[0]), len(devices)
            for i, d in enumerate(devices):
                cfg.devices[i] = int(d)
            device = int(devices[0])
        h = C.c_void_p()
        _check(L.b2fwn_create(self.points.ctypes.data, len(self.points), self.tris.ctypes.data, len(self.tris),
                              C.byref(cfg), C.byref(h)))
        self._h = h
        self.device = device

    def close(self):
        if self._h is not None and _lib is not None:
            _lib.b2fwn_destroy(self._h)
        self._h = None

    def __del__(self):
        self.close()

    # ---- IFastWindingNumber ----
    def calc_from_vector(self, points, beta=2.0, skip_face=-1, cb=None, out=None):
        pts = np.ascontiguousarray(points, np.float32).reshape(-1, 3)
        res = out if out is not None else np.empty(len(pts), np.float32)
        _check(_lib.b2fwn_calc_from_vector(self._h, pts.ctypes.data, len(pts), beta, skip_face, res.ctypes.data,
                                           _progress(cb), None))
        return res

    def calc_self_intersections(self, beta=2.0, cb=None):
        """returns the FaceBitSet as uint64 words (bit f set iff w<0 or w>1)"""
        res = np.zeros((len(self.tris) + 63) // 64, np.uint64)
        _check(_lib.b2fwn_calc_self_intersections(self._h, beta, res.ctypes.data, _progress(cb), None))
        return res

    def calc_from_grid(self, dims, grid_to_mesh_xf, beta=2.0, cb=None, out=None):
        n = int(np.prod(np.asarray(dims, np.int64)))
        res = out if out is not None else np.empty(n, np.float32)
        _check(_lib.b2fwn_calc_from_grid(self._h, _dims(dims), _xf(grid_to_mesh_xf), beta, res.ctypes.data,
                                         _progress(cb), None))
        return res

    def calc_from_grid_with_distances(self, dims, grid_to_mesh_xf, options=None, cb=None, out=None):
        options = options or DistOptions()
        n = int(np.prod(np.asarray(dims, np.int64)))
        res = out if out is not None else np.empty(n, np.float32)
        _check(_lib.b2fwn_calc_from_grid_with_distances(self._h, _dims(dims), _xf(grid_to_mesh_xf), C.byref(options),
                                                        res.ctypes.data, _progress(cb), None))
        return res

    def calc_from_grid_with_distances_layers(self, dims, grid_to_mesh_xf, z_begin, z_end, options=None, cb=None, out=None):
        """layers [z_begin, z_end) of the grid into host memory (one rank's z-slab)"""
        options = options or DistOptions()
        n = int(dims[0]) * int(dims[1]) * (z_end - z_begin)
        res = out if out is not None else np.empty(n, np.float32)
        _check(_lib.b2fwn_calc_from_grid_with_distances_layers(self._h, _dims(dims), _xf(grid_to_mesh_xf), C.byref(options),
                                                               z_begin, z_end, res.ctypes.data, _progress(cb), None))
        return res

    # ---- makeSignedByWindingNumber (MRVDBConversions.cpp:491-568): sign of a dense unsigned distance volume ----
    def make_signed_by_winding_number(self, volume, dims, grid_to_mesh_xf, winding_number_threshold=0.5,
                                      winding_number_beta=2.0, cb=None):
        """negates, in place, the values of the voxels whose winding number exceeds the threshold; `volume` is a
        C-contiguous float32 host array of dims[0]*dims[1]*dims[2] values (x fastest), streamed through the device"""
        n = int(dims[0]) * int(dims[1]) * int(dims[2])
        if not (isinstance(volume, np.ndarray) and volume.dtype == np.float32 and volume.flags.c_contiguous
                and volume.flags.writeable and volume.size == n):
            raise ValueError("volume must be a writeable C-contiguous float32 array of dims[0]*dims[1]*dims[2] values")
        _check(_lib.b2fwn_make_signed_by_winding_number(self._h, _dims(dims), _xf(grid_to_mesh_xf), winding_number_beta,
                                                        winding_number_threshold, volume.ctypes.data, _progress(cb), None))
        return volume

    def make_signed_by_winding_number_device(self, d_volume_at_z_begin, dims, grid_to_mesh_xf, winding_number_threshold=0.5,
                                             winding_number_beta=2.0, z_begin=0, z_end=None):
        """the same on a volume resident in device memory (pointer to layer z_begin), layers [z_begin, z_end)"""
        z_end = dims[2] if z_end is None else z_end
        _check(_lib.b2fwn_make_signed_by_winding_number_device(self._h, _dims(dims), _xf(grid_to_mesh_xf), winding_number_beta,
                                                               winding_number_threshold, z_begin, z_end, d_volume_at_z_begin))

    # ---- IFastWindingNumberByParts ----
    def _part_cb(self, res_func):
        def tramp(data, dims, z_offset, _u):
            d = (dims[0], dims[1], dims[2])
            arr = np.ctypeslib.as_array(data, shape=(d[0] * d[1] * d[2],))
            try:
                ok = res_func(arr, d, z_offset)
            except Exception:  # exceptions must not cross the C boundary
                return 1
            return 0 if (ok is None or ok) else 1
        return PART_CB(tramp)

    def calc_from_grid_by_parts(self, res_func, dims, grid_to_mesh_xf, beta=2.0, layer_overlap=0, cb=None,
                                max_layers_per_part=0):
        _check(_lib.b2fwn_calc_from_grid_by_parts(self._h, _dims(dims), _xf(grid_to_mesh_xf), beta, layer_overlap,
                                                  max_layers_per_part, self._part_cb(res_func), None, _progress(cb),
                                                  None))

    def calc_from_grid_with_distances_by_parts(self, res_func, dims, grid_to_mesh_xf, options=None, layer_overlap=0,
                                               cb=None, max_layers_per_part=0):
        options = options or DistOptions()
        _check(_lib.b2fwn_calc_from_grid_with_distances_by_parts(
            self._h, _dims(dims), _xf(grid_to_mesh_xf), C.byref(options), layer_overlap, max_layers_per_part,
            self._part_cb(res_func), None, _progress(cb), None))

    def calc_from_grid_with_distances_by_parts_into(self, res_func, dims, grid_to_mesh_xf, options=None, layer_overlap=0, cb=None,
                                                    max_layers_per_part=0, distances=True, beta=2.0):
        """by-parts with caller-owned part memory (b2fwn_*_by_parts_into): every part lands in a fresh pageable numpy array that
        is handed to res_func(arr, dims, z_offset) -- what the C++ adapter does with std::vector<float>"""
        bufs = {}

        def buffer(dims_p, z_offset, _u):
            try:
                a = np.empty(dims_p[0] * dims_p[1] * dims_p[2], np.float32)
            except MemoryError:
                return None
            bufs[z_offset] = a
            return a.ctypes.data

        def part(_data, dims_p, z_offset, _u):
            a = bufs.pop(z_offset)
            try:
                ok = res_func(a, (dims_p[0], dims_p[1], dims_p[2]), z_offset)
            except Exception:
                return 1
            return 0 if (ok is None or ok) else 1

        bcb, pcb = PART_BUFFER_CB(buffer), PART_CB(part)
        if distances:
            options = options or DistOptions()
            _check(_lib.b2fwn_calc_from_grid_with_distances_by_parts_into(
                self._h, _dims(dims), _xf(grid_to_mesh_xf), C.byref(options), layer_overlap, max_layers_per_part, bcb, pcb, None,
                _progress(cb), None))
        else:
            _check(_lib.b2fwn_calc_from_grid_by_parts_into(self._h, _dims(dims), _xf(grid_to_mesh_xf), beta, layer_overlap,
                                                           max_layers_per_part, bcb, pcb, None, _progress(cb), None))

    # ---- device-pointer forms (multi-GPU slabs, benches) ----
    def calc_from_vector_device(self, d_points_ptr, n, d_out_ptr, beta=2.0, skip_face=-1):
        _check(_lib.b2fwn_calc_from_vector_device(self._h, d_points_ptr, n, beta, skip_face, d_out_ptr))

    def plan_parts(self, dims, layer_overlap=0, max_layers_per_part=0, z_begin=0, z_end=None):
        """([(first layer, layers, device index), ...], layers per part): the partition a by-parts / host-output call of this handle
        would use right now (b2fwn_plan_parts_for: the calls' own part-size choice incl. the free-memory cut)"""
        z_end = dims[2] if z_end is None else z_end
        n, lpp = C.c_int32(), C.c_int32()
        _check(_lib.b2fwn_plan_parts_for(self._h, _dims(dims), z_begin, z_end, layer_overlap, max_layers_per_part, 0, None, None, None,
                                         C.byref(n), C.byref(lpp)))
        b, s, d = ((C.c_int32 * max(n.value, 1))() for _ in range(3))
        _check(_lib.b2fwn_plan_parts_for(self._h, _dims(dims), z_begin, z_end, layer_overlap, max_layers_per_part, n.value, b, s, d,
                                         C.byref(n), C.byref(lpp)))
        return [(b[k], s[k], d[k]) for k in range(n.value)], lpp.value

    def trim(self):
        """gives the handle's cached scratch (device slots, mesher workspace, pinned part buffers, idle pool blocks) back"""
        _check(_lib.b2fwn_trim(self._h))

    def calc_from_vector_device_noalloc(self, d_points_ptr, n, d_out_ptr, beta=2.0, skip_face=-1):
        """calc_from_vector_device without any device allocation: the sort's scratch is a fixed array inside the library
        (at most NOALLOC_MAX_QUERIES points per call), for hosts that own all device memory or cannot afford an allocation stall"""
        _check(_lib.b2fwn_calc_from_vector_device_noalloc(self._h, d_points_ptr, n, beta, skip_face, d_out_ptr))

    def calc_self_intersections_device(self, d_bits_ptr, d_w_ptr=None, beta=2.0):
        _check(_lib.b2fwn_calc_self_intersections_device(self._h, beta, d_bits_ptr, d_w_ptr))

    def calc_from_grid_device(self, dims, grid_to_mesh_xf, d_out_ptr, beta=2.0, z_begin=0, z_end=None):
        z_end = dims[2] if z_end is None else z_end
        _check(_lib.b2fwn_calc_from_grid_device(self._h, _dims(dims), _xf(grid_to_mesh_xf), beta, z_begin, z_end,
                                                d_out_ptr))

    def calc_from_grid_with_distances_device(self, dims, grid_to_mesh_xf, d_out_ptr, options=None, z_begin=0,
                                             z_end=None):
        options = options or DistOptions()
        z_end = dims[2] if z_end is None else z_end
        _check(_lib.b2fwn_calc_from_grid_with_distances_device(self._h, _dims(dims), _xf(grid_to_mesh_xf),
                                                               C.byref(options), z_begin, z_end, d_out_ptr))

    def calc_from_grid_with_distances_device_chunks(self, dims, grid_to_mesh_xf, d_volume_at_z_begin, options=None, z_begin=0,
                                                    z_end=None, z_chunk=8, z_stride=8):
        """layers z_begin + c * z_stride + k (k < z_chunk) below z_end, written in place into the volume that starts at
        layer z_begin (one rank's share of a z-interleaved multi-GPU partition)"""
        options = options or DistOptions()
        z_end = dims[2] if z_end is None else z_end
        _check(_lib.b2fwn_calc_from_grid_with_distances_device_chunks(self._h, _dims(dims), _xf(grid_to_mesh_xf), C.byref(options),
                                                                      z_begin, z_end, z_chunk, z_stride, d_volume_at_z_begin))

    def calc_from_grid_with_distances_device_chunks_p2p(self, dims, grid_to_mesh_xf, d_volume_at_z_begin, peer_ptrs_at_z_begin,
                                                        options=None, z_begin=0, z_end=None, z_chunk=8, z_stride=8):
        """as ..._device_chunks, and every result is also stored into the peer volumes (other GPUs' memory opened with
        shared_open): the all-gather of a multi-GPU run fused into the kernel"""
        options = options or DistOptions()
        z_end = dims[2] if z_end is None else z_end
        n = len(peer_ptrs_at_z_begin)
        arr = (C.c_void_p * max(n, 1))(*peer_ptrs_at_z_begin)
        _check(_lib.b2fwn_calc_from_grid_with_distances_device_chunks_p2p(self._h, _dims(dims), _xf(grid_to_mesh_xf), C.byref(options),
                                                                          z_begin, z_end, z_chunk, z_stride, d_volume_at_z_begin, n, arr))

    # ---- IPointsToMeshProjector::findProjections ----
    def find_projections(self, points, point_xf=None, tree_xf=None, up_dist_limit_sq=FLT_MAX, lo_dist_limit_sq=0.0, cb=None):
        """structured array (PROJECTION_DTYPE) of the closest mesh point of every query point: face (-1 = none within the
        limit), point, barycentrics a/b, dist_sq; point_xf / tree_xf = rigidXfPoint / nonRigidXfTree (12 floats, A then b)"""
        pts = np.ascontiguousarray(points, np.float32).reshape(-1, 3)
        out = np.zeros(len(pts), PROJECTION_DTYPE)
        px = None if point_xf is None else np.ascontiguousarray(point_xf, np.float32)
        tx = None if tree_xf is None else np.ascontiguousarray(tree_xf, np.float32)
        _check(_lib.b2fwn_find_projections(self._h, pts.ctypes.data, len(pts), None if px is None else px.ctypes.data,
                                           None if tx is None else tx.ctypes.data, up_dist_limit_sq, lo_dist_limit_sq,
                                           out.ctypes.data, _progress(cb), None))
        return out

    def find_projections_device(self, d_points_ptr, n, d_out_ptr, point_xf=None, tree_xf=None, up_dist_limit_sq=FLT_MAX,
                                lo_dist_limit_sq=0.0):
        px = None if point_xf is None else np.ascontiguousarray(point_xf, np.float32)
        tx = None if tree_xf is None else np.ascontiguousarray(tree_xf, np.float32)
        _check(_lib.b2fwn_find_projections_device(self._h, d_points_ptr, n, None if px is None else px.ctypes.data,
                                                  None if tx is None else tx.ctypes.data, up_dist_limit_sq, lo_dist_limit_sq,
                                                  d_out_ptr))

    # ---- peer-visible device memory (multi-GPU volumes) ----
    def shared_alloc(self, nbytes):
        """-> (device pointer, 64-byte token another process passes to shared_open)"""
        ptr = C.c_void_p()
        token = C.create_string_buffer(64)
        _check(_lib.b2fwn_shared_alloc(self._h, nbytes, C.byref(ptr), token))
        return ptr.value, token.raw

    def shared_open(self, token):
        ptr = C.c_void_p()
        _check(_lib.b2fwn_shared_open(self._h, token, C.byref(ptr)))
        return ptr.value

    def shared_close(self, ptr):
        _check(_lib.b2fwn_shared_close(self._h, ptr))

    def shared_free(self, ptr):
        _check(_lib.b2fwn_shared_free(self._h, ptr))

    def update_points(self, points):
        """same triangulation, moved vertices: AABBTree::refit + calcDipoles on the device (source/MRMesh/MRAABBTree.cpp:62-94)"""
        pts = np.ascontiguousarray(points, np.float32).reshape(-1, 3)
        _check(_lib.b2fwn_update_points(self._h, pts.ctypes.data, len(pts)))

    # ---- parity / diagnostics ----
    def get_tree(self):
        n = C.c_size_t()
        _check(_lib.b2fwn_get_tree(self._h, None, None, 0, C.byref(n)))
        nodes = np.empty(n.value, NODE_DTYPE)
        dip = np.empty(n.value, DIPOLE_DTYPE)
        if n.value:
            _check(_lib.b2fwn_get_tree(self._h, nodes.ctypes.data, dip.ctypes.data, n.value, C.byref(n)))
        return nodes, dip

    def stats(self):
        s = Stats()
        _check(_lib.b2fwn_get_stats(self._h, C.byref(s)))
        return s

    # ---- callers of the interface, on the device ----
    def suggest_voxel_size(self, approx_num_voxels):
        """MR::suggestVoxelSize (source/MRVoxels/MROffset.cpp:25-31)"""
        v = C.c_float()
        _check(_lib.b2fwn_suggest_voxel_size(self._h, float(approx_num_voxels), C.byref(v)))
        return v.value

    def bounding_box(self):
        bb = (C.c_float * 6)()
        _check(_lib.b2fwn_get_bounding_box(self._h, bb))
        return np.array(bb[:3], np.float32), np.array(bb[3:], np.float32)

    def mesh_to_distance_volume(self, origin, voxel_size, dims, options=None, cb=None, out=None):
        """MR::meshToDistanceVolume, HoleWindingRule (source/MRVoxels/MRMeshToDistanceVolume.cpp:43-62):
        returns (data, min, max) = SimpleVolumeMinMax's payload"""
        options = options or DistOptions()
        vp_ = VolumeParams()
        vp_.origin[:] = [float(x) for x in origin]
        vs = np.asarray(voxel_size, np.float32) * np.ones(3, np.float32)
        vp_.voxel_size[:] = [float(x) for x in vs]
        vp_.dims[:] = [int(d) for d in dims]
        n = int(dims[0]) * int(dims[1]) * int(dims[2])
        res = out if out is not None else np.empty(n, np.float32)
        lo, hi = C.c_float(), C.c_float()
        _check(_lib.b2fwn_mesh_to_distance_volume(self._h, C.byref(vp_), C.byref(options), res.ctypes.data, C.byref(lo),
                                                  C.byref(hi), _progress(cb), None))
        return res, lo.value, hi.value

    # ---- SignDetectionMode::ProjectionNormal (signedDistanceToMesh, source/MRMesh/MRMeshDistance.cpp:13-59) ----
    def signed_distances_by_projection_normal(self, points, options=None, cb=None):
        """distance to the closest point of the mesh for each point, negative where the pseudonormal at the closest point
        says inside; NaN outside [min_dist_sq, max_dist_sq) when options.null_outside_min_max"""
        options = options or DistOptions()
        pts = np.ascontiguousarray(points, np.float32).reshape(-1, 3)
        res = np.empty(len(pts), np.float32)
        _check(_lib.b2fwn_signed_distances_by_projection_normal(self._h, pts.ctypes.data, len(pts), C.byref(options),
                                                                res.ctypes.data, _progress(cb), None))
        return res

    def signed_distances_by_projection_normal_device(self, d_points_ptr, n, d_out_ptr, options=None):
        options = options or DistOptions()
        _check(_lib.b2fwn_signed_distances_by_projection_normal_device(self._h, d_points_ptr, n, C.byref(options), d_out_ptr))

    def mesh_to_distance_volume_by_projection_normal(self, origin, voxel_size, dims, options=None, cb=None):
        """MR::meshToDistanceVolume with SignDetectionMode::ProjectionNormal; returns (data, min, max)"""
        options = options or DistOptions()
        vp_ = VolumeParams()
        vp_.origin[:] = [float(x) for x in origin]
        vs = np.asarray(voxel_size, np.float32) * np.ones(3, np.float32)
        vp_.voxel_size[:] = [float(x) for x in vs]
        vp_.dims[:] = [int(d) for d in dims]
        res = np.empty(int(dims[0]) * int(dims[1]) * int(dims[2]), np.float32)
        lo, hi = C.c_float(), C.c_float()
        _check(_lib.b2fwn_mesh_to_distance_volume_by_projection_normal(self._h, C.byref(vp_), C.byref(options), res.ctypes.data,
                                                                       C.byref(lo), C.byref(hi), _progress(cb), None))
        return res, lo.value, hi.value

    def mc_offset_mesh(self, offset, voxel_size, close_holes=True, threshold=0.5, beta=2.0, cb=None, fetch=True, max_layers_per_slab=0):
        """MR::mcOffsetMesh with HoleWindingRule (source/MRVoxels/MROffset.cpp:117-246), distance band and marching cubes
        both on the device.  Returns (verts, tris, info) -- verts/tris None when fetch is False"""
        p = OffsetParams(float(voxel_size), float(offset), int(close_holes), float(threshold), float(beta), int(max_layers_per_slab))
        r = OffsetResult()
        _check(_lib.b2fwn_mc_offset_mesh(self._h, C.byref(p), C.byref(r), _progress(cb), None))
        info = dict(origin=np.array(r.origin[:], np.float32), dims=tuple(r.dims[:]), n_verts=r.n_verts, n_tris=r.n_tris,
                    sdf_ms=r.sdf_ms, mc_ms=r.mc_ms)
        if not fetch:
            return None, None, info
        verts = np.empty((r.n_verts, 3), np.float32)
        tris = np.empty((r.n_tris, 3), np.int32)
        _check(_lib.b2fwn_marching_cubes_fetch(self._h, verts.ctypes.data, tris.ctypes.data))
        return verts, tris, info

    # ---- half-edge tables (Mesh::fromTriMesh on the device) ----
    def build_topology(self, tris=None, n_verts=0, on_device=False, n_tris=0):
        """MeshTopology tables of a triangulation: tris=None -> the handle's last marching-cubes result (still on the device); a
        host (F,3) int32 array; or a device pointer with n_tris.  Returns (half_edges (2E,4) int32 {next, prev, org, left},
        edge_per_vertex (V,), edge_per_face (F,)); raises B2Error for non-manifold input"""
        nh, nv, nf = C.c_uint64(), C.c_uint64(), C.c_uint64()
        if tris is None:
            _check(_lib.b2fwn_build_topology(self._h, None, 0, 0, 0, C.byref(nh), C.byref(nv), C.byref(nf)))
        elif on_device:
            _check(_lib.b2fwn_build_topology(self._h, tris, n_tris, 1, n_verts, C.byref(nh), C.byref(nv), C.byref(nf)))
        else:
            t = np.ascontiguousarray(tris, np.int32).reshape(-1, 3)
            _check(_lib.b2fwn_build_topology(self._h, t.ctypes.data, len(t), 0, n_verts, C.byref(nh), C.byref(nv), C.byref(nf)))
        he = np.empty((nh.value, 4), np.int32)
        epv = np.empty(nv.value, np.int32)
        epf = np.empty(nf.value, np.int32)
        _check(_lib.b2fwn_build_topology_fetch(self._h, he.ctypes.data, epv.ctypes.data, epf.ctypes.data))
        return he, epv, epf

    # ---- marching cubes ----
    def marching_cubes(self, volume, dims, origin, voxel_size, iso, less_inside=False, on_device=False):
        """volume: host float32 array (or a device pointer when on_device); returns (verts (n,3) f32, tris (m,3) i32)"""
        p = McParams()
        p.origin[:] = [float(x) for x in origin]
        vs = np.asarray(voxel_size, np.float32) * np.ones(3, np.float32)
        p.voxel_size[:] = [float(x) for x in vs]
        p.iso = iso
        p.less_inside = int(less_inside)
        nv, nt = C.c_uint64(), C.c_uint64()
        if on_device:
            ptr = volume
        else:
            volume = np.ascontiguousarray(volume, np.float32)
            ptr = volume.ctypes.data
        _check(_lib.b2fwn_marching_cubes(self._h, ptr, int(on_device), _dims(dims), C.byref(p), C.byref(nv),
                                         C.byref(nt)))
        verts = np.empty((nv.value, 3), np.float32)
        tris = np.empty((nt.value, 3), np.int32)
        _check(_lib.b2fwn_marching_cubes_fetch(self._h, verts.ctypes.data, tris.ctypes.data))
        return verts, tris
