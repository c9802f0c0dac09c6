#!/usr/bin/env python
"""N3 diagnostics: b2fwn_update_points (refit of the device tree + dipoles + leaf triangles) against a full build, 2 M faces"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import meshlib_b200 as mb
from meshlib_b200 import meshgen

pts, tris = meshgen.make_uv_sphere(1, 1415, 708)
pts = meshgen.add_noise(pts, 1e-4)
t0 = time.perf_counter(); fwn = mb.FastWindingNumber(pts, tris); create = time.perf_counter() - t0
build_ms = fwn.stats().tree_build_ms
new = (pts * np.float32(1.05)).astype(np.float32)
for _ in range(3):
    t0 = time.perf_counter(); fwn.update_points(new); wall = time.perf_counter() - t0
    ms = fwn.stats().last_kernel_ms
print(f"N3 refit: {len(tris)} faces: device build {build_ms:.1f} ms (create call {create * 1e3:.1f} ms host-to-host); "
      f"update_points kernels {ms:.2f} ms, call {wall * 1e3:.1f} ms host-to-host (H2D of {pts.nbytes / 1e6:.0f} MB included)")
