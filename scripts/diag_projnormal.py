#!/usr/bin/env python
"""N2 diagnostics, SignDetectionMode::ProjectionNormal: meshToDistanceVolume of a 2 M-face mesh on a 256^3 grid, full range
(the sign needs the exact closest point everywhere) and with a 3-voxel band + NaN outside; kernel time by CUDA events and
host-to-host wall clock; first call includes building the pseudonormal tables."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import meshlib_b200 as mb
from meshlib_b200 import meshgen

pts, tris = meshgen.make_uv_sphere(1, 1415, 708)
pts = meshgen.add_noise(pts, 1e-4)
fwn = mb.FastWindingNumber(pts, tris)
n = 256
voxel = 2.4 / n
t0 = time.perf_counter()
fwn.signed_distances_by_projection_normal(np.array([[1.5, 0.2, 0.1]], np.float32))
print(f"pseudonormal tables ({len(tris)} faces) + first call: {(time.perf_counter() - t0) * 1e3:.1f} ms")
for label, opt in (("full range", mb.DistOptions(0.0, mb.FLT_MAX, False)), ("band 3 voxels, NaN outside", mb.DistOptions(0.0, (3 * voxel) ** 2, True))):
    for _ in range(2):
        t0 = time.perf_counter()
        vol, lo, hi = fwn.mesh_to_distance_volume_by_projection_normal((-1.2, -1.2, -1.2), voxel, (n, n, n), opt)
        wall = time.perf_counter() - t0
        ms = fwn.stats().last_kernel_ms
    print(f"N2 ProjectionNormal {label}: {n}^3 voxels kernels {ms:.1f} ms = {n ** 3 / ms / 1e3:.1f} Mvox/s, host-to-host {wall * 1e3:.1f} ms; "
          f"min {lo:.4f} max {hi:.4f} NaN {np.isnan(vol).mean():.3f} inside {(vol < 0).mean():.4f}")
