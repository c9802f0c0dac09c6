#!/usr/bin/env python
"""N1 diagnostics: findProjections throughput (10 M random points vs a 1 M-tri sphere), kernel time by CUDA events, next to the
reference CPU class on a bounded sample"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import meshlib_b200 as mb
from meshlib_b200 import meshgen

pts, tris = meshgen.make_uv_sphere(1, 1000, 500)
fwn = mb.FastWindingNumber(pts, tris)
n = 10_000_000
qh = meshgen.query_points(n, -1.5, 1.5, seed=7)
q = torch.from_numpy(qh).cuda()
out = torch.empty(n * 8, dtype=torch.float32, device="cuda")
for _ in range(3):
    fwn.find_projections_device(q.data_ptr(), n, out.data_ptr())
    ms = fwn.stats().last_kernel_ms
print(f"N1 findProjections: {len(tris)} tris, {n} points: {ms:.2f} ms = {n / ms / 1e3:.1f} Mqueries/s")
try:
    from oracle import reflib
    if reflib.available():
        r = reflib.RefMesh.from_arrays(pts, tris)
        r.find_projections(qh[:1000])
        k = 400_000
        t0 = time.perf_counter(); r.find_projections(qh[:k]); dt = time.perf_counter() - t0
        print(f"reference CPU PointsToMeshProjector: {k} points {dt:.2f} s = {k / dt / 1e6:.2f} Mqueries/s on {reflib.hardware_threads()} threads")
except Exception as e:
    print("reference sample failed:", e)
