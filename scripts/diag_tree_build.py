#!/usr/bin/env python
"""(needs a library built with `make DEV=1`: the product build ignores B2FWN_NO_SUBTREES)
tree + dipole build time on the device (b2fwn_stats.tree_build_ms) for the bench meshes; B2FWN_NO_SUBTREES=1 = global passes only"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import meshlib_b200 as mb
from meshlib_b200 import meshgen
for name, (h, v) in {"1 M": (1000, 500), "2 M": (1414, 707), "10 M": (3162, 1581)}.items():
    pts, tris = meshgen.make_uv_sphere(1, h, v)
    pts = meshgen.add_noise(pts, 1e-3, seed=42)
    ms = []
    for _ in range(4):
        f = mb.FastWindingNumber(pts, tris); ms.append(f.stats().tree_build_ms); n = f.stats().kernel_launches; f.close()
    print(f"{name} faces ({len(tris)}): tree build {min(ms):.1f} ms (runs: {', '.join('%.1f' % x for x in ms)}), {n} launches", flush=True)
