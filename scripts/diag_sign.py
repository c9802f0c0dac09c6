#!/usr/bin/env python
"""N2 diagnostics: makeSignedByWindingNumber on a 512^3 unsigned distance volume of a 2 M-face mesh.
  (a) resident volume, one in-place call (kernel time by CUDA events);
  (b) host volume streamed through the device (H2D, walk, D2H per chunk), wall clock;
  (c) the reference's shape of the same job on OUR winding numbers: calcFromGrid to the host, then the flip loop on the CPU
      (numpy, one thread) -- what MRVDBConversions.cpp:552-565 does with any IFastWindingNumber."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import meshlib_b200 as mb
from meshlib_b200 import meshgen

pts, tris = meshgen.make_uv_sphere(1, 1415, 708)
pts = meshgen.add_noise(pts, 1e-4)
fwn = mb.FastWindingNumber(pts, tris)
n = 512
dims = (n, n, n)
xf = mb.make_basis_xf((-1.2, -1.2, -1.2), 2.4 / n)
vox = n ** 3
d = torch.empty(vox, dtype=torch.float32, device="cuda")
fwn.calc_from_grid_with_distances_device(dims, xf, d.data_ptr())
unsigned = d.abs()
host = unsigned.cpu().numpy()

for _ in range(3):
    v = unsigned.clone()
    fwn.make_signed_by_winding_number_device(v.data_ptr(), dims, xf)
    ms_a = fwn.stats().last_kernel_ms
ref_signed = v.cpu().numpy()
print(f"N2 (a) resident: {len(tris)} faces, {n}^3 voxels: {ms_a:.2f} ms = {vox / ms_a / 1e6:.2f} Gvox/s")

best_b = 1e9
for _ in range(3):
    h = host.copy()
    t0 = time.perf_counter(); fwn.make_signed_by_winding_number(h, dims, xf); best_b = min(best_b, time.perf_counter() - t0)
assert h.tobytes() == ref_signed.tobytes()
print(f"N2 (b) host volume streamed (pageable memory): {best_b * 1e3:.1f} ms = {vox / best_b / 1e9:.2f} Gvox/s")

best_c = 1e9
for _ in range(2):
    h = host.copy()
    t0 = time.perf_counter()
    w = fwn.calc_from_grid(dims, xf)
    flip = w > np.float32(0.5)
    h[flip] = -h[flip]
    best_c = min(best_c, time.perf_counter() - t0)
assert h.tobytes() == ref_signed.tobytes()
print(f"N2 (c) calcFromGrid to the host + flip loop on the CPU: {best_c * 1e3:.1f} ms = {vox / best_c / 1e9:.2f} Gvox/s")

