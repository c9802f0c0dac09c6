#!/usr/bin/env python
"""C2 / C4 diagnostics: winding-number queries/s of calcFromVector (10 M random points vs a 1 M-tri sphere) and of
calcSelfIntersections (two interpenetrating tori, ~5 M tris); kernel time by CUDA events (b2fwn_stats.last_kernel_ms)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import meshlib_b200 as mb
from meshlib_b200 import meshgen

pts, tris = meshgen.make_uv_sphere(1, 1000, 500)
fwn = mb.FastWindingNumber(pts, tris)
n = 10_000_000
q = torch.from_numpy(meshgen.query_points(n, -1.5, 1.5, seed=7)).cuda()
out = torch.empty(n, dtype=torch.float32, device="cuda")
for _ in range(3):
    fwn.calc_from_vector_device(q.data_ptr(), n, out.data_ptr())
    ms = fwn.stats().last_kernel_ms
print(f"C2 calcFromVector: {len(tris)} tris, {n} points: {ms:.2f} ms = {n / ms / 1e3:.1f} Mqueries/s, inside frac {float((out > 0.5).float().mean()):.4f}")
fwn.close()
pts, tris = meshgen.two_tori(2.0, 0.7, 1600, 800, (1.0, 0.0, 0.0))
fwn = mb.FastWindingNumber(pts, tris)
bits = torch.zeros((len(tris) + 63) // 64, dtype=torch.int64, device="cuda")
for _ in range(3):
    fwn.calc_self_intersections_device(bits.data_ptr())
    ms = fwn.stats().last_kernel_ms
nb = int(sum(bin(int(x) & (2**64 - 1)).count("1") for x in bits.cpu().numpy()[:: max(1, len(bits) // 4096)]))
print(f"C4 calcSelfIntersections: {len(tris)} tris: {ms:.2f} ms = {len(tris) / ms / 1e3:.1f} Mqueries/s (sampled set bits {nb})")
