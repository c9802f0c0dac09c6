"""Generates tests/golden/ref_golden.npz from the UNMODIFIED reference (oracle/_ref/libmrref.so, compiled from
/root/reference/source by oracle/Makefile).  Run in the build container: python tests/golden/make_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from meshlib_b200 import make_basis_xf, meshgen  # noqa: E402
from oracle import reflib  # noqa: E402

out = {}


def add(name, ref, dims, xf, band):
    pts, tris = ref.points, ref.tris
    q = meshgen.query_points(3000, float(pts.min()) - 0.5, float(pts.max()) + 0.5, seed=123)
    out[f"{name}_points"], out[f"{name}_tris"] = pts, tris
    out[f"{name}_nodes"], out[f"{name}_dipoles"] = ref.nodes(), ref.dipoles()
    out[f"{name}_q"], out[f"{name}_w"] = q, ref.calc_from_vector(q)
    out[f"{name}_dims"], out[f"{name}_xf"] = np.asarray(dims, np.int32), np.asarray(xf, np.float32)
    out[f"{name}_wgrid"] = ref.calc_from_grid(dims, xf)
    out[f"{name}_sdf"] = ref.calc_from_grid_with_distances(dims, xf)
    out[f"{name}_band_opts"] = np.asarray(band, np.float32)
    out[f"{name}_sdf_band"] = ref.calc_from_grid_with_distances(dims, xf, min_dist_sq=band[0], max_dist_sq=band[1], null_outside=True)
    out[f"{name}_selfint"] = ref.calc_self_intersections()


add("torus", reflib.RefMesh.torus(2, 1, 32, 16), (24, 24, 12), make_basis_xf((-3.3, -3.3, -1.3), (0.275, 0.275, 0.22)), (0.0025, 0.36))
p, t = meshgen.make_uv_sphere(1, 30, 15)
p = meshgen.add_noise(p, 0.01)
t = t[np.arange(len(t)) % 11 != 5]  # holes
add("noisy_sphere", reflib.RefMesh.from_arrays(p, t), (20, 20, 20), make_basis_xf((-1.3, -1.3, -1.3), 0.13), (0.001, 0.09))
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_golden.npz"), **out)
print("wrote", {k: v.shape for k, v in out.items()})
