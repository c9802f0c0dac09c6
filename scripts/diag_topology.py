#!/usr/bin/env python
"""Half-edge tables of the C3 offset mesh on the device (b2fwn_build_topology after b2fwn_mc_offset_mesh) against the reference's
MeshBuilder::fromTriangles on the host cores: time and counts."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import meshlib_b200 as mb
from meshlib_b200 import meshgen
from oracle import reflib
pts, tris = meshgen.make_uv_sphere(1, 1414, 707)
pts = meshgen.add_noise(pts, 1e-3, seed=42)
fwn = mb.FastWindingNumber(pts, tris)
v, t, info = fwn.mc_offset_mesh(0.02, 2.04 / 508.0)
for _ in range(2):
    t0 = time.perf_counter(); he, epv, epf = fwn.build_topology(); wall = time.perf_counter() - t0
    print(f"device: {len(t)} triangles, {len(v)} vertices -> {len(he)} half-edges, kernels {fwn.stats().last_kernel_ms:.2f} ms, with fetch {1e3*wall:.1f} ms")
if reflib.available():
    r = reflib.topology_from_tables(he, epv, epf)
    print("reference MeshTopology::read + checkValidity:", "accepted" if r is not None else "REJECTED", None if r is None else (bool(np.array_equal(r[0], t)), r[1:]))
    vf, ue, bl, vv, sec = reflib.builder_stats(t)
    print(f"reference MeshBuilder::fromTriangles on {reflib.hardware_threads()} threads: {1e3*sec:.1f} ms, faces {vf}, undirected edges {ue}, boundary loops {bl}, vertices {vv}")
