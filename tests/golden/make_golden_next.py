"""Generates tests/golden/ref_golden_next.npz from the UNMODIFIED reference (oracle/_ref/libmrref.so) for the SURVEY 8f rows:
N1 closest-point records (PointsToMeshProjector::findProjections), N2 makeSignedByWindingNumber's loop and
signedDistanceToMesh with SignDetectionMode::ProjectionNormal (+ pseudonormals), N3 Mesh::updateCaches (AABBTree::refit) +
calcDipoles.  Run in the build container: python tests/golden/make_golden_next.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from meshlib_b200 import make_basis_xf, meshgen  # noqa: E402
from oracle import checker, reflib  # noqa: E402

out = {}


def deform(p):
    """the vertex motion of the refit fixture (float32 throughout)"""
    q = p.copy()
    q[:, 0] = np.float32(1.15) * p[:, 0] + np.float32(0.08) * np.sin(np.float32(3) * p[:, 1])
    q[:, 2] = p[:, 2] + np.float32(0.1) * np.cos(np.float32(2) * p[:, 0])
    return q.astype(np.float32)


def add(name, ref, dims, xf):
    pts, tris = ref.points, ref.tris
    out[f"{name}_points"], out[f"{name}_tris"] = pts, tris
    q = meshgen.query_points(2000, float(pts.min()) - 0.5, float(pts.max()) + 0.5, seed=321)
    q[::9] = pts[(np.arange(len(q[::9])) * 7) % len(pts)]  # points ON the mesh: ties around a vertex
    out[f"{name}_q"] = q
    # N1
    out[f"{name}_proj"] = ref.find_projections(q)
    out[f"{name}_proj_limits"] = ref.find_projections(q, up=0.25, lo=0.01)
    # N2 (b)
    fn, vn, en = ref.pseudonormals(len(tris), len(pts))
    out[f"{name}_face_normals"], out[f"{name}_vert_normals"], out[f"{name}_edge_normals"] = fn, vn, en
    out[f"{name}_pn"], out[f"{name}_pn_dots"] = ref.signed_distance_proj_normal(q, null_outside=False)
    out[f"{name}_pn_band"], out[f"{name}_pn_band_dots"] = ref.signed_distance_proj_normal(q, min_dist_sq=0.0, max_dist_sq=0.25, null_outside=True)
    # N2 (a): unsigned distances of a grid, then the reference's sign loop at threshold 0.4, beta 2.5
    out[f"{name}_dims"], out[f"{name}_xf"] = np.asarray(dims, np.int32), np.asarray(xf, np.float32)
    unsigned = np.abs(ref.calc_from_grid_with_distances(dims, xf))
    signed, w = checker.ref_make_signed_by_winding_number(ref, unsigned, dims, xf, 0.4, 2.5)
    out[f"{name}_unsigned"], out[f"{name}_signed"], out[f"{name}_sign_w"] = unsigned, signed, w
    # N3: every vertex moved, tree refitted in place (this changes `ref`: keep it last)
    new = deform(pts)
    ref.update_points(new)
    out[f"{name}_moved_points"] = new
    out[f"{name}_refit_nodes"], out[f"{name}_refit_dipoles"] = ref.nodes(), ref.dipoles()
    out[f"{name}_refit_w"] = ref.calc_from_vector(q)


if __name__ == "__main__":
    add("torus", reflib.RefMesh.torus(2, 1, 32, 16), (24, 24, 12), make_basis_xf((-3.3, -3.3, -1.3), (0.275, 0.275, 0.22)))
    p, t = meshgen.make_uv_sphere(1, 30, 15)
    p = meshgen.add_noise(p, 0.01)
    t = t[np.arange(len(t)) % 11 != 5]  # holes
    add("noisy_sphere", reflib.RefMesh.from_arrays(p, t), (20, 20, 20), make_basis_xf((-1.3, -1.3, -1.3), 0.13))
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_golden_next.npz"), **out)
    print("wrote", {k: v.shape for k, v in out.items()})
