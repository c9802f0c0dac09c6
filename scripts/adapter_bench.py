#!/usr/bin/env python
"""End to end through the C++ adapter (MR::B200::FastWindingNumber over the reference's headers, oracle/_ref/adapter_check --bench):
a bench workload's mesh is written to /tmp, the adapter process builds a MeshLib Mesh from it and calls
calcFromGridWithDistances( std::vector<float>& ) -- the number a MeshLib caller sees.  usage: adapter_bench.py [workload] [devices] [byparts]"""
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c3_band"
devices = sys.argv[2] if len(sys.argv) > 2 else "0"
wl = bench.build_workload(name)
exe = os.path.join(ROOT, "oracle", "_ref", "adapter_check")
with tempfile.TemporaryDirectory() as d:
    wl["points"].tofile(os.path.join(d, "p.bin"))
    wl["tris"].tofile(os.path.join(d, "t.bin"))
    xf, o = wl["xf"], wl["opts"]
    voxel = float(xf[0])
    origin = [float(xf[9 + k]) - 0.5 * voxel for k in range(3)]  # the adapter_check re-adds half a voxel the same way bench does
    cmd = [exe, "--bench", os.path.join(d, "p.bin"), os.path.join(d, "t.bin")] + [str(x) for x in wl["dims"]] + [repr(x) for x in origin] + \
          [repr(voxel), repr(o["max_dist_sq"]), "1" if o["null_outside"] else "0", devices] + sys.argv[3:4]
    print(wl["desc"], "| devices", devices, flush=True)
    r = subprocess.run(cmd, text=True)
    sys.exit(r.returncode)
