#!/usr/bin/env python
"""(needs a library built with `make DEV=1`: the product build ignores B2FWN_PROJ_BRICKS)
dev diagnostics: ProjectionNormal volume through the brick walk (B2FWN_PROJ_BRICKS=1) against the per-point search and the
reference, on the open-boundary test mesh; prints the voxels whose sign differs"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import meshlib_b200 as mb
from meshlib_b200 import meshgen
from oracle import checker, reflib

p, t = meshgen.make_uv_sphere(1, 80, 40)
p = meshgen.add_noise(p, 4e-3)
t = t[np.arange(len(t)) % 9 != 4]
orc = checker.best_oracle(p, t)
rp, rt = orc.points, orc.tris
fwn = mb.FastWindingNumber(rp, rt)
origin = (rp.min(0) - 0.4).astype(np.float32)
dims = (37, 41, 23)
voxel = ((rp.max(0) + 0.4 - origin) / np.asarray(dims, np.float32)).astype(np.float32)
centres = reflib.voxel_centres(origin, voxel, dims)
for opt, kw in ((mb.DistOptions(0.0, mb.FLT_MAX, False), dict(null_outside=False)),
                (mb.DistOptions(0.0, 0.09, True), dict(min_dist_sq=0.0, max_dist_sq=0.09, null_outside=True))):
    want, dots = orc.signed_distance_proj_normal(centres, **kw)
    os.environ["B2FWN_PROJ_BRICKS"] = "0"
    pt, _, _ = fwn.mesh_to_distance_volume_by_projection_normal(origin, voxel, dims, opt)
    os.environ["B2FWN_PROJ_BRICKS"] = "1"
    bk, _, _ = fwn.mesh_to_distance_volume_by_projection_normal(origin, voxel, dims, opt)
    ok = ~np.isnan(want)
    clear = ok & (np.abs(dots) > 1e-5 * np.abs(want))
    mp = ok & (np.signbit(pt) != np.signbit(want))
    mb_ = ok & (np.signbit(bk) != np.signbit(want))
    print(f"case {kw}: ok {ok.sum()} clear {clear.sum()} | point path sign mismatches {mp.sum()} (clear {(mp & clear).sum()}) | "
          f"brick path {mb_.sum()} (clear {(mb_ & clear).sum()}) | nan masks equal {np.array_equal(np.isnan(bk), np.isnan(want))} "
          f"| abs equal brick/point {(np.abs(bk[ok]) == np.abs(pt[ok])).mean():.6f}")
    idx = np.nonzero(mb_ & clear)[0][:12]
    rec = fwn.find_projections(centres[idx]) if len(idx) else []
    ref = orc.find_projections(centres[idx]) if len(idx) else []
    for k, i in enumerate(idx):
        print(f"  voxel {i} centre {centres[i].tolist()} want {want[i]:.7g} dot {dots[i]:.4g} brick {bk[i]:.7g} point {pt[i]:.7g} | "
              f"ours face {rec['face'][k]} a {rec['a'][k]:.9g} b {rec['b'][k]:.9g} | ref face {ref['face'][k]} a {ref['a'][k]:.9g} b {ref['b'][k]:.9g}")
