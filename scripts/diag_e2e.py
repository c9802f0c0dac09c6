#!/usr/bin/env python
"""where the host-to-host time of the bench workload goes: create (H2D + tree build), the chunked grid call, kernels inside it"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import meshlib_b200 as mb
import bench
wl = bench.build_workload(sys.argv[1] if len(sys.argv) > 1 else "c5_band")
o = wl["opts"]
opts = mb.DistOptions(o["min_dist_sq"], o["max_dist_sq"], o["null_outside"], o["threshold"], o["beta"])
dims = wl["dims"]; total = dims[0] * dims[1] * dims[2]
host = torch.empty(total, dtype=torch.float32, pin_memory=True); hn = host.numpy()
for it in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    f = mb.FastWindingNumber(wl["points"], wl["tris"]); t1 = time.perf_counter()
    st = f.stats()
    f.calc_from_grid_with_distances_layers(dims, np.array(wl["xf"], np.float32), 0, dims[2], opts, out=hn); t2 = time.perf_counter()
    k = f.stats().last_kernel_ms
    f.close(); t3 = time.perf_counter()
    print(f"iter {it}: create {1e3*(t1-t0):.1f} ms (tree build {st.tree_build_ms:.1f}), grid call {1e3*(t2-t1):.1f} ms (kernels {k:.1f}), close {1e3*(t3-t2):.1f} ms, total {1e3*(t3-t0):.1f} ms")
