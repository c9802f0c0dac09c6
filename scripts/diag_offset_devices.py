#!/usr/bin/env python
"""The C5-sized offset (10 M faces, voxel 0.001, 2050^3 grid, incl. marching cubes) on 1 .. N devices of one process and in slabs."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import meshlib_b200 as mb
from meshlib_b200 import meshgen
pts, tris = meshgen.make_uv_sphere(1, 3162, 1581)
pts = meshgen.add_noise(pts, 1e-3, seed=42)
nd = mb.load_library().b2fwn_device_count()
ref = None
for devices, slab in [([0], 0), ([0], 256)] + ([(list(range(nd)), 0)] if nd > 1 else []):
    f = mb.FastWindingNumber(pts, tris, devices=devices)
    for rep in range(2):
        t0 = time.perf_counter()
        _v, _t, info = f.mc_offset_mesh(0.005, 2.04 / 2040.0, fetch=False, max_layers_per_slab=slab)
        wall = time.perf_counter() - t0
    print(f"devices {devices} slab {slab or 'whole'}: grid {info['dims']}, sdf {info['sdf_ms']:.1f} ms, mc {info['mc_ms']:.1f} ms, wall {1e3*wall:.1f} ms, "
          f"{info['n_verts']} verts {info['n_tris']} tris")
    v, t, info = f.mc_offset_mesh(0.005, 2.04 / 2040.0, max_layers_per_slab=slab)
    sig = (len(v), len(t), float(np.abs(v[::997]).sum()), int(t[::997].astype(np.int64).sum()))
    if ref is None:
        ref = sig
    print("   identical to the first run:", sig == ref)
    f.close()
