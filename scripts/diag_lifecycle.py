#!/usr/bin/env python
"""where the host-to-host time of the small workloads goes: create / call / fetch / close, five rounds each (C3 offset, C4, C2)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import meshlib_b200 as mb
from meshlib_b200 import meshgen

def rounds(name, pts, tris, call):
    for it in range(5):
        t0 = time.perf_counter(); f = mb.FastWindingNumber(pts, tris); t1 = time.perf_counter()
        call(f); t2 = time.perf_counter()
        f.close(); t3 = time.perf_counter()
        print(f"{name} round {it}: create {1e3*(t1-t0):.1f} ms, call {1e3*(t2-t1):.1f} ms, close {1e3*(t3-t2):.1f} ms, total {1e3*(t3-t0):.1f} ms", flush=True)

pts, tris = meshgen.make_uv_sphere(1, 1414, 707); pts = meshgen.add_noise(pts, 1e-3, seed=42)
rounds("c3_offset", pts, tris, lambda f: f.mc_offset_mesh(0.02, 2.04 / 508.0))
pts, tris = meshgen.two_tori(2.0, 0.7, 1600, 800)
rounds("c4", pts, tris, lambda f: f.calc_self_intersections())
pts, tris = meshgen.make_uv_sphere(1, 1000, 500)
q = meshgen.query_points(10_000_000)
rounds("c2", pts, tris, lambda f: f.calc_from_vector(q))
