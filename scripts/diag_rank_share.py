#!/usr/bin/env python
"""Where do the N > 1 ranks lose their time?  On ONE GPU: every rank's z-interleaved share of c5_band (the call bench.py makes at
N ranks, without the peer stores) timed on its own.  If a share takes 1/N of the whole-grid time, the loss of the multi-GPU run
is in the peer stores; what it takes beyond 1/N is per-call work that does not shrink with the share.
usage: diag_rank_share.py [workload] [world ...]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import meshlib_b200 as mb  # noqa: E402
from meshlib_b200 import slabs  # noqa: E402
import torch  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c5_band"
worlds = [int(v) for v in sys.argv[2:]] or [2, 8]
wl = bench.build_workload(name)
dims, xf = wl["dims"], wl["xf"]
o = wl["opts"]
opt = mb.DistOptions(o["min_dist_sq"], o["max_dist_sq"], o["null_outside"], o["threshold"], o["beta"])
fwn = mb.FastWindingNumber(wl["points"], wl["tris"])
layer = dims[0] * dims[1]
vol = torch.empty(layer * dims[2], dtype=torch.float32, device="cuda")


def timed(call, reps=3):
    ms = []
    for _ in range(reps):
        call()
        ms.append(fwn.stats().last_kernel_ms)
    return min(ms)


whole = timed(lambda: fwn.calc_from_grid_with_distances_device(dims, xf, vol.data_ptr(), opt))
print(f"{name}: whole grid {whole:.1f} ms")
for world in worlds:
    plan = slabs.ChunkPlan(world, dims[2], 8, 1)
    for rank in sorted({0, world // 2, world - 1}):
        zb, ze, zc, zs = plan.call_args(rank, 0)
        t = timed(lambda: fwn.calc_from_grid_with_distances_device_chunks(dims, xf, vol.data_ptr() + 4 * zb * layer, opt, z_begin=zb, z_end=ze,
                                                                         z_chunk=zc, z_stride=zs))
        print(f"  world {world} rank {rank}: {t:.1f} ms  (1/N of the whole grid: {whole / world:.1f} ms, ratio {t / (whole / world):.3f})")
    # contiguous slab of the same size for comparison (rank 0's 1/N of the layers, through the middle of the grid)
    n = dims[2] // world
    z0 = (dims[2] - n) // 2
    t = timed(lambda: fwn.calc_from_grid_with_distances_device(dims, xf, vol.data_ptr() + 4 * z0 * layer, opt, z_begin=z0, z_end=z0 + n))
    print(f"  world {world} contiguous middle slab of {n} layers: {t:.1f} ms")
fwn.close()
