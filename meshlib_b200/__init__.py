"""meshlib_b200 -- B200-native (sm_100a) mesh -> signed-distance-volume hot path of MeshLib behind the reference's
operator surface (MR::IFastWindingNumber / IFastWindingNumberByParts, meshToDistanceVolume).  The arithmetic lives in
libb2fwn.so (CUDA, C ABI: include/b2fwn.h); this package is the thin host-side mirror of the reference interface."""
from .api import (B2Error, DeviceBuffer, DistOptions, FastWindingNumber, FLT_MAX, LIB_PATH, EXPORTS, load_library,  # noqa: F401
                  make_basis_xf, measure_fp32_peak, plan_parts, OffsetParams, PROJECTION_DTYPE, OffsetResult, VolumeParams)
