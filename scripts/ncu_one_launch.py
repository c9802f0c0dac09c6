#!/usr/bin/env python
"""One launch sequence of a bench workload's device-resident step (whole grid, one GPU) and nothing else: the program `ncu
--replay-mode application` re-runs once per counter pass, so that a kernel that writes a 34-GB volume can be captured with ALL
passes (kernel replay would save / restore that volume ~40 times and died half way in r02).
usage: ncu_one_launch.py [workload] [launches]   -- the mesh is cached in /tmp between the runs of one capture"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import meshlib_b200 as mb  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c5_band"
launches = int(sys.argv[2]) if len(sys.argv) > 2 else 1
cache = f"/tmp/ncu_one_launch_{name}.npz"
if not os.path.exists(cache):
    wl = bench.build_workload(name)
    np.savez(cache, points=wl["points"], tris=wl["tris"], dims=np.array(wl["dims"]), xf=np.asarray(wl["xf"]),
             opts=np.array([wl["opts"]["min_dist_sq"], wl["opts"]["max_dist_sq"], float(wl["opts"]["null_outside"])], np.float64))
z = np.load(cache)
import torch  # noqa: E402
dims = tuple(int(v) for v in z["dims"])
o = z["opts"]
opt = mb.DistOptions(float(o[0]), float(np.float32(o[1])), bool(o[2]))
fwn = mb.FastWindingNumber(z["points"], z["tris"])
out = torch.empty(dims[0] * dims[1] * dims[2], dtype=torch.float32, device="cuda")
for _ in range(launches):
    fwn.calc_from_grid_with_distances_device(dims, z["xf"], out.data_ptr(), opt)
torch.cuda.synchronize()
print(name, dims, "kernel ms of the last call:", fwn.stats().last_kernel_ms)
fwn.close()
